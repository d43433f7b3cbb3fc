"""GPU: the EXECUTED drop-in.  The reference's own, unmodified algorithm classes
(/root/reference/src/algorithms/{sac,rad,drq,svea,soda,curl,augcl,pad,non_naive_rad}.py) are imported from oracle/_ref
(the reference compiled to bytecode by oracle/build_ref.py; the live tree when it is mounted) with augcl_b200's shim
modules ``augmentations`` / ``utils`` ahead of them on sys.path (augcl_b200/dropin), and each runs ``update()`` on a
device ring filled through ``add(LazyFrames, ...)`` -- the calls src/train.py:251,256 make.  Checked: no exception,
finite losses, parameters finite and changed, replay tensors and the ring unmutated, kwargs-by-name dispatch of
``RAD.apply_aug`` (src/algorithms/rad.py:18-22), ``train.py``'s poke of ``last_save`` (src/train.py:184)."""
import json
import math
import os
import sys
import types

import numpy as np
import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _ref_dir():
    from oracle import ref_harness as R
    return R.ref_dir()


class _Log:
    """Stands in for logger.Logger (src/logger.py needs termcolor / tensorboard): records what update() logs."""

    def __init__(self):
        self.rows = {}

    def log(self, key, value, step, n=1):
        self.rows[key] = float(value)


def _args(algorithm, data_aug="identity", aug_params=None, **kw):
    # defaults of src/arguments.py:11-106
    a = dict(algorithm=algorithm, discount=0.99, batch_size=16, hidden_dim=64, actor_lr=1e-3, actor_beta=0.9,
             actor_log_std_min=-10, actor_log_std_max=2, actor_update_freq=2, critic_lr=1e-3, critic_beta=0.9,
             critic_tau=0.01, critic_target_update_freq=2, num_shared_layers=4, num_head_layers=0, num_filters=8,
             projection_dim=20, encoder_tau=0.05, init_temperature=0.1, alpha_lr=1e-4, alpha_beta=0.5, aux_lr=1e-3,
             aux_beta=0.9, aux_update_freq=2, soda_batch_size=24, soda_tau=0.005, svea_alpha=0.5, svea_beta=0.5,
             drq_k=2, drq_m=2, data_aug=data_aug, aug_params=json.dumps(aug_params) if aug_params else None)
    a.update(kw)
    return types.SimpleNamespace(**a)


@pytest.fixture(scope="module")
def dropin():
    ref = _ref_dir()
    if ref is None:
        pytest.skip("no reference available (neither /root/reference nor oracle/_ref)")
    import augcl_b200.dropin as D
    from oracle import ref_harness as R
    saved_path = list(sys.path)
    saved_mods = {k: sys.modules.get(k) for k in ("utils", "augmentations", "logger")}
    if ref == R.REF_BUILT:          # compiled reference (GPU box): resolved by name through the harness's finder
        R.install_ref_importer()
        D.install(None)
    else:
        D.install(ref)
    lg = types.ModuleType("logger")
    lg.Logger = _Log
    sys.modules["logger"] = lg
    import augmentations
    import utils
    from algorithms.factory import make_agent
    assert utils.ReplayBuffer.__module__ == "augcl_b200.utils" and augmentations.random_shift.__module__ == "augcl_b200.augmentations"
    assert sys.modules["algorithms.sac"].__spec__.origin.startswith(ref), "algorithm classes must be the reference's"
    assert utils.soft_update_params.__module__ == "_augcl_reference_utils"
    yield types.SimpleNamespace(utils=utils, aug=augmentations, make_agent=make_agent)
    D.uninstall()
    R.remove_ref_importer()
    sys.path[:] = saved_path
    for k, v in saved_mods.items():
        if v is None:
            sys.modules.pop(k, None)
        else:
            sys.modules[k] = v


def _fill(utils, S, B, A=4, cap=96, k=3, seed=0):
    """A ring filled the way src/train.py:256 does: add(LazyFrames, action, reward, LazyFrames, done) with the frame
    sharing of FrameStack (src/env/wrappers.py:183-203); frames are DMC-like (bright flat background + saturated blobs)
    so splice keeps some pixels and replaces others."""
    rng = np.random.RandomState(seed)
    frames = np.full((cap + k, 3, S, S), 200, dtype=np.uint8)
    for f in frames:
        y, x = rng.randint(0, S - 20, 2)
        f[:, y:y + 20, x:x + 20] = rng.randint(0, 256, (3, 1, 1))
        f += rng.randint(0, 8, f.shape).astype(np.uint8)
    buf = utils.ReplayBuffer(obs_shape=(3 * k, S, S), action_shape=(A,), capacity=cap, batch_size=B)
    fl = [frames[i] for i in range(cap + k)]
    for t in range(cap - 7):
        buf.add(utils.LazyFrames(fl[t:t + k]), rng.uniform(-1, 1, A).astype(np.float32), float(rng.rand()),
                utils.LazyFrames(fl[t + 1:t + 1 + k]), bool(rng.rand() < 0.05))
    return buf


CASES = [
    ("sac", 84, {}),
    ("rad", 84, dict(data_aug="shift")),
    ("rad", 100, dict(data_aug="crop", aug_params={"crop": {"size": 84}})),
    ("rad", 84, dict(data_aug="splice-conv", aug_params={"splice": {"val_thres": 0.4}})),
    ("non_naive_rad", 84, dict(data_aug="overlay")),
    ("drq", 84, {}),
    ("svea", 84, dict(data_aug="overlay")),
    ("svea", 84, dict(data_aug="conv")),
    ("soda", 100, dict(data_aug="overlay")),
    ("curl", 100, {}),
    ("pad", 84, {}),
    ("augcl", 84, dict(data_aug="splice-conv")),
    ("augcl", 84, dict(data_aug="CS_splice")),
]


@pytest.mark.parametrize("algo,S,kw", CASES, ids=[f"{a}-{k.get('data_aug', 'none')}-{s}" for a, s, k in CASES])
def test_reference_update_runs_on_the_dropin(dropin, algo, S, kw):
    torch.manual_seed(0)
    np.random.seed(0)
    dropin.aug.seed(3)
    dropin.aug.set_background_bank(None)          # a synthetic bank of the right size is made on first use
    args = _args(algo, **kw)
    buf = _fill(dropin.utils, S, args.batch_size)
    agent = dropin.make_agent(obs_shape=(9, 84, 84), action_shape=(4,), args=args)
    before = [p.detach().clone() for p in agent.critic.parameters()]
    ring_before = buf.sample(idxs=np.arange(8))[0].clone()
    buf.last_save = 5                              # src/train.py:184 writes the attribute from outside
    log = _Log()
    for step in range(2):                          # step 0 exercises the actor / target / aux branches (freq 2), step 1 the rest
        agent.update(buf, log, step)
    torch.cuda.synchronize()
    assert log.rows, "update() logged nothing"
    for k, v in log.rows.items():
        assert math.isfinite(v), f"{algo}: {k} = {v}"
    after = list(agent.critic.parameters())
    assert all(torch.isfinite(p).all() for p in after)
    assert any(not torch.equal(a, b) for a, b in zip(after, before)), "critic did not train"
    assert torch.equal(buf.sample(idxs=np.arange(8))[0], ring_before), "the ring was mutated by update()"
    assert buf.last_save == 5 and buf.idx == 96 - 7 and not buf.full
    dropin.aug.set_background_bank(None)


@pytest.mark.parametrize("algo,kw", [("augcl", dict(data_aug="splice-conv")), ("drq", {}), ("svea", dict(data_aug="overlay"))])
def test_train_loop_pattern_add_then_update(dropin, algo, kw):
    """src/train.py:251-256 in miniature: every step adds one transition (frames shared with the previous step, like
    FrameStack) and runs one update() of the unmodified reference agent on the drop-in buffer -- 40 steps across the
    staging flushes, the wrap-around of the write cursor, the recycled output sets and back-to-back sample launches."""
    torch.manual_seed(1)
    np.random.seed(1)
    dropin.aug.seed(9)
    dropin.aug.set_background_bank(None)
    args = _args(algo, **kw)
    buf = _fill(dropin.utils, 84, args.batch_size, cap=256)
    start = buf.idx
    agent = dropin.make_agent(obs_shape=(9, 84, 84), action_shape=(4,), args=args)
    rng = np.random.RandomState(5)
    frames = [rng.randint(0, 256, (3, 84, 84)).astype(np.uint8) for _ in range(3)]
    log = _Log()
    for step in range(40):
        frames.append(rng.randint(0, 256, (3, 84, 84)).astype(np.uint8))
        buf.add(dropin.utils.LazyFrames(frames[-4:-1]), rng.uniform(-1, 1, 4).astype(np.float32), float(rng.rand()),
                dropin.utils.LazyFrames(frames[-3:]), False)
        agent.update(buf, log, step)
    torch.cuda.synchronize()
    assert buf.idx == (start + 40) % 256 and buf.full          # 249 + 40: the write cursor wrapped around the ring
    for k, v in log.rows.items():
        assert math.isfinite(v), f"{algo}: {k} = {v}"
    # the last transition added is what the ring returns for its slot
    obs, _, _, nxt, _ = buf.sample(idxs=np.array([buf.idx - 1]))
    assert torch.equal(obs[0].cpu(), torch.from_numpy(np.concatenate(frames[-4:-1])).float())
    assert torch.equal(nxt[0].cpu(), torch.from_numpy(np.concatenate(frames[-3:])).float())
    dropin.aug.set_background_bank(None)


def test_apply_aug_kwargs_by_name_and_input_aliasing(dropin):
    """RAD.apply_aug passes --aug_params by keyword (rad.py:18-22); SVEA augments obs.clone() (svea.py:49) and AugCL
    reuses obs after random_shift (augcl.py:70-71): neither call may mutate its input."""
    dropin.aug.seed(5)
    dropin.aug.set_background_bank(None)
    args = _args("rad", data_aug="shift-splice", aug_params={"shift": {"pad": 2}, "splice": {"hue_thres": 0, "sat_thres": 0, "val_thres": 0.4}})
    agent = dropin.make_agent(obs_shape=(9, 84, 84), action_shape=(4,), args=args)
    assert list(agent.aug_funcs) == ["shift", "splice"] and agent.aug_funcs["shift"]["params"] == {"pad": 2}
    buf = _fill(dropin.utils, 84, 16)
    obs = buf.sample()[0]
    keep = obs.clone()
    out = agent.apply_aug(obs)
    assert out.shape == obs.shape and out.dtype == torch.float32 and out.is_cuda and out.is_contiguous()
    assert out.data_ptr() != obs.data_ptr() and torch.equal(obs, keep)
    sh = dropin.aug.random_shift(obs)
    assert torch.equal(obs, keep) and sh.data_ptr() != obs.data_ptr()
    with pytest.raises(TypeError):                 # an unknown keyword fails like it does against the reference's functions
        dropin.aug.random_shift(obs, padding=3)
    dropin.aug.set_background_bank(None)


def test_consecutive_samples_do_not_alias(dropin):
    """The reference returns fresh tensors from every sample() (src/utils.py:206-214): two consecutive calls of the
    drop-in must not share storage while both results are alive."""
    buf = _fill(dropin.utils, 84, 16)
    a = buf.sample_drq()
    b = buf.sample_drq()
    for x, y in zip(a, b):
        assert x.data_ptr() != y.data_ptr()
    a2 = buf.sample()
    assert all(x.data_ptr() != y.data_ptr() for x, y in zip(a2, b))
