"""The sample entry points are CUDA-graph capturable when every draw is device-resident: they enqueue one kernel on the
given stream, allocate nothing themselves, never synchronise, and the fused sampler's ticket counters re-arm at kernel end
(SURVEY.md 7.2 "offer CUDA-graph capture").  A captured update replays with new indices / offsets written into the same
device tensors and equals the eager call."""
import numpy as np
import pytest
import torch

from conftest import TOL_NORM

pytestmark = pytest.mark.gpu


def test_captured_samplers_replay_with_new_draws():
    from augcl_b200 import _lib as L, augmentations as A, utils as U
    cap, B, S = 5000, 128, 84
    buf = U.ReplayBuffer((9, S, S), (6,), cap, B)
    buf.fill_synthetic(seed=2)
    A.set_background_bank(None)
    A.seed(3)
    A.load_dataloader(B, S)
    g = torch.Generator().manual_seed(1)
    idx = torch.randint(0, cap, (B,), generator=g).cuda()
    off = torch.randint(0, 9, (B, 2), generator=g, dtype=torch.int32).cuda()
    off2 = torch.randint(0, 9, (B, 2), generator=g, dtype=torch.int32).cuda()
    rows = torch.randint(0, A.SYNTHETIC_BANK_LEN, (B,), generator=g).cuda()
    wt = torch.randn((B, 3, 3, 3, 3), generator=g).cuda()

    def update():
        drq = buf.sample_drq(idxs=idx, offsets={"obs": off, "next_obs": off2})
        aug = buf.sample_augcl(idxs=idx, offsets={"obs": off}, bank_idx=rows, weights=wt)
        return drq, aug

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    try:
        with torch.cuda.stream(side):
            update()                                   # warm-up outside the capture (lazy init, func attributes)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        n0 = L.launch_count()
        with torch.cuda.graph(graph):
            drq_g, aug_g = update()
        assert L.launch_count() - n0 == 2              # two kernels in the graph
        for trial in range(3):
            idx.copy_(torch.randint(0, cap, (B,), generator=g))
            off.copy_(torch.randint(0, 9, (B, 2), generator=g, dtype=torch.int32))
            graph.replay()
            drq_e, aug_e = update()
            torch.cuda.synchronize()
            for a, b in zip(drq_g, drq_e):
                assert torch.equal(a, b)
            assert torch.equal(aug_g[0], aug_e[0]) and torch.equal(aug_g[4], aug_e[4])
            assert (aug_g[1] - aug_e[1]).abs().max().item() / 255.0 <= TOL_NORM
    finally:
        A.set_background_bank(None)
