"""GPU: ReplayBuffer.save / load in the reference's chunk format, against chunk files written by the
reference (tests/golden/chunks_ref) and round trips through the device ring."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
CH = os.path.join(GOLDEN, "chunks_ref")


def same(batch, d):
    obs, act, rew, nxt, nd = batch
    return (torch.equal(obs.cpu(), torch.from_numpy(d["out_obs"]).float()) and torch.equal(nxt.cpu(), torch.from_numpy(d["out_next"]).float())
            and torch.equal(act.cpu(), torch.from_numpy(d["out_act"])) and torch.equal(rew.cpu(), torch.from_numpy(d["out_rew"]))
            and torch.equal(nd.cpu(), torch.from_numpy(d["out_nd"])))


def test_load_reference_chunks_then_incremental_save(golden, tmp_path):
    from augcl_b200 import utils as U
    d = golden("chunks_ref")
    S, A, cap = int(d["S"]), int(d["A"]), int(d["cap"])
    buf = U.ReplayBuffer((9, S, S), (A,), cap, 9)
    buf.IO_BLOCK = 4                                   # several staging blocks per chunk
    buf.load(CH)
    assert buf.idx == 25 and buf.last_save == 25 and buf.last_requeue_save == 25 and not buf.full
    assert same(buf.sample(idxs=d["idxs"]), d)
    # end_step stops after the chunk that ends there (src/utils.py:199-200)
    b2 = U.ReplayBuffer((9, S, S), (A,), cap, 9)
    b2.load(CH, end_step=10, is_requeue_load=False)
    assert b2.idx == 10 and b2.last_save == 10 and b2.last_requeue_save == 0
    # incremental save: same chunk names as the reference, then a new transition and a second chunk
    buf.last_save = 0
    buf.save(str(tmp_path))
    assert sorted(os.listdir(tmp_path)) == ["0_25.tar"] and buf.last_save == 25
    buf.save(str(tmp_path))                            # nothing new: no file (src/utils.py:157-158)
    assert sorted(os.listdir(tmp_path)) == ["0_25.tar"]
    fr = np.random.RandomState(0).randint(0, 256, (4, 3, S, S)).astype(np.uint8)
    buf.add(U.LazyFrames(list(fr[0:3])), np.ones(A, np.float32), 2.5, U.LazyFrames(list(fr[1:4])), True)
    buf.save(str(tmp_path))
    assert sorted(os.listdir(tmp_path)) == ["0_25.tar", "25_26.tar"]
    b3 = U.ReplayBuffer((9, S, S), (A,), cap, 9)
    b3.load(str(tmp_path))
    assert b3.idx == 26 and same(b3.sample(idxs=d["idxs"]), d)
    o, a, r, x, nd = b3.sample(idxs=np.array([25]))
    assert torch.equal(o.cpu()[0], torch.from_numpy(np.concatenate(list(fr[0:3]))).float())
    assert torch.equal(x.cpu()[0], torch.from_numpy(np.concatenate(list(fr[1:4]))).float())
    assert float(r) == 2.5 and float(nd) == 0.0 and torch.equal(a.cpu()[0], torch.ones(A))
    with pytest.raises(AssertionError):                # contiguity assert of the reference (src/utils.py:188-190)
        b3.load(str(tmp_path))
