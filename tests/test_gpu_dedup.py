"""GPU parity of the de-duplicated ring ("v2" layout: augcl_ring_create_dedup / put_frames / put_refs,
ReplayBuffer(dedup=True)) against the CPU oracle's HostRing and against the full-stack ring: frames are stored once,
shared by object identity like the reference's LazyFrames (src/utils.py:283-322) coming out of FrameStack
(src/env/wrappers.py:183-203), and every sample entry point reads through the reference table."""
import collections

import numpy as np
import pytest
import torch

from conftest import TOL_NORM

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def U():
    from augcl_b200 import utils
    return utils


@pytest.fixture(scope="module")
def A():
    from augcl_b200 import augmentations
    return augmentations


@pytest.fixture(scope="module")
def O():
    from oracle import augcl_oracle
    return augcl_oracle


def frame_stack_stream(rng, steps, k, S, episode=17):
    """(obs_frames, next_frames, action, reward, done) like FrameStack: a deque of k frame OBJECTS, the reset frame
    repeated k times, one new frame per step, an episode boundary every `episode` steps."""
    dq = collections.deque(maxlen=k)
    t_ep = episode
    for t in range(steps):
        if t_ep == episode:
            f0 = rng.randint(0, 256, (3, S, S)).astype(np.uint8)
            for _ in range(k):
                dq.append(f0)
            t_ep = 0
        obs = list(dq)
        dq.append(rng.randint(0, 256, (3, S, S)).astype(np.uint8))
        t_ep += 1
        yield obs, list(dq), rng.randn(4).astype(np.float32), float(rng.randn()), t_ep == episode


@pytest.mark.parametrize("S,k", [(84, 3), (20, 2), (100, 3)])
def test_dedup_ring_matches_oracle_and_shares_frames(U, O, S, k):
    rng = np.random.RandomState(S + k)
    cap, B, steps, pad = 50, 16, 130, 4 if S > 30 else 2
    buf = U.ReplayBuffer((3 * k, S, S), (4,), cap, B, dedup=True)
    ref = U.ReplayBuffer((3 * k, S, S), (4,), cap, B)
    ora = O.HostRing((3 * k, S, S), (4,), cap, B)
    allocated = 0
    for obs, nxt, a, r, d in frame_stack_stream(rng, steps, k, S):
        buf.add(U.LazyFrames(obs), a, r, U.LazyFrames(nxt), d)
        ref.add(U.LazyFrames(obs), a, r, U.LazyFrames(nxt), d)
        ora.add(O.LazyFrames(np.stack(obs)), a, r, O.LazyFrames(np.stack(nxt)), d)
    # one new frame per step + one per reset (8 episodes): the pool cursor wrapped once (pool = 62+12+16)
    assert buf._adds == steps and (buf.idx, buf.full) == (ora.idx, ora.full)
    assert buf.ring_bytes() < 0.45 * ref.ring_bytes()
    for _ in range(3):
        idxs = rng.randint(0, cap, size=B)
        o_obs, o_act, o_rew, o_next, o_nd = ora.sample(idxs)
        obs, act, rew, nxt, nd = buf.sample(idxs=idxs)
        assert torch.equal(obs.cpu(), o_obs) and torch.equal(nxt.cpu(), o_next)
        assert torch.equal(act.cpu(), o_act) and torch.equal(rew.cpu(), o_rew) and torch.equal(nd.cpu(), o_nd)
        off = {"obs": rng.randint(0, 2 * pad + 1, (B, 2)), "next_obs": rng.randint(0, 2 * pad + 1, (B, 2))}
        a1 = buf.sample_drq(pad=pad, idxs=idxs, offsets=off)
        assert torch.equal(a1[0].cpu(), O.random_shift_int(o_obs, off["obs"], pad))
        assert torch.equal(a1[3].cpu(), O.random_shift_int(o_next, off["next_obs"], pad))
        size = S - 8
        co = {key: rng.randint(0, 8, (B, 2)) for key in ("pos", "obs", "next_obs")}
        c1 = buf.sample_curl(idxs=idxs, size=size, offsets=co)
        assert torch.equal(c1[0].cpu(), O.random_crop(o_obs, size, co["obs"][:, 0], co["obs"][:, 1]))
        assert torch.equal(c1[5].cpu(), O.random_crop(o_obs, size, co["pos"][:, 0], co["pos"][:, 1]))
    from augcl_b200 import _lib as L
    L.set_option("staging", "generic")          # the unstaged kernel reads through the same table
    try:
        assert torch.equal(buf.sample(idxs=idxs)[0].cpu(), o_obs)
    finally:
        L.set_option("staging", "tma")


def test_dedup_ring_fused_sampler_and_unshared_frames(U, A, O):
    rng = np.random.RandomState(3)
    S, k, cap, B, pad = 84, 3, 40, 12, 4
    buf = U.ReplayBuffer((3 * k, S, S), (4,), cap, B, dedup=True)
    ref = U.ReplayBuffer((3 * k, S, S), (4,), cap, B)
    for obs, nxt, a, r, d in frame_stack_stream(rng, 70, k, S, episode=11):
        buf.add(U.LazyFrames(obs), a, r, U.LazyFrames(nxt), d)
        ref.add(U.LazyFrames(obs), a, r, U.LazyFrames(nxt), d)
    bank = torch.from_numpy(rng.randint(0, 256, (6, 3, S, S)).astype(np.uint8)).cuda()
    A.set_background_bank(bank)
    try:
        idxs = rng.randint(0, cap, size=B)
        kw = dict(idxs=idxs, offsets={"obs": rng.randint(0, 9, (B, 2))}, bank_idx=rng.randint(0, 6, size=B),
                  weights=torch.from_numpy(rng.randn(B, 3, 3, 3, 3).astype(np.float32)))
        for fn in ("sample_augcl", "sample_svea_aug"):
            x, y = getattr(buf, fn)(**kw), getattr(ref, fn)(**kw)
            for a, b in zip(x, y):
                assert torch.equal(a, b), fn
    finally:
        A.set_background_bank(None)
    # plain (C,H,W) arrays are never shared: 2k frames per add, and the pool refuses to overwrite live frames
    small = U.ReplayBuffer((3 * k, 20, 20), (4,), 50, 4, dedup=True)
    x = rng.randint(0, 256, (3 * k, 20, 20)).astype(np.uint8)
    with pytest.raises(RuntimeError, match="pool"):
        for _ in range(40):
            small.add(x.copy(), np.zeros(4, np.float32), 0.0, x.copy(), False)
    ok = U.ReplayBuffer((3 * k, 20, 20), (4,), 50, 4, dedup=True, pool_factor=2 * k + 1)
    hst = O.HostRing((3 * k, 20, 20), (4,), 50, 4)
    for t in range(120):
        a, b = rng.randint(0, 256, (2, 3 * k, 20, 20)).astype(np.uint8)
        ok.add(a, np.zeros(4, np.float32), float(t), b, False)
        hst.add(a, np.zeros(4, np.float32), float(t), b, False)
    idxs = rng.randint(0, 50, size=9)
    assert torch.equal(ok.sample(idxs=idxs)[0].cpu(), hst.sample(idxs)[0])
    assert torch.equal(ok.sample(idxs=idxs)[3].cpu(), hst.sample(idxs)[3])


def test_dedup_ring_persistence_roundtrip(U, tmp_path):
    rng = np.random.RandomState(5)
    S, k, cap, B = 84, 3, 64, 8
    src = U.ReplayBuffer((3 * k, S, S), (4,), cap, B, dedup=True)
    stream = list(frame_stack_stream(rng, 45, k, S, episode=13))
    for i, (obs, nxt, a, r, d) in enumerate(stream):
        src.add(U.LazyFrames(obs), a, r, U.LazyFrames(nxt), d)
        if i in (19, 44):
            src.save(str(tmp_path))
    idxs = np.arange(45)
    want = src.sample(idxs=idxs)
    for dedup in (True, False):
        dst = U.ReplayBuffer((3 * k, S, S), (4,), cap, B, dedup=dedup)
        dst.load(str(tmp_path))
        assert dst.idx == 45
        got = dst.sample(idxs=idxs)
        for a, b in zip(want, got):
            assert torch.equal(a, b)
        if dedup:       # the chunk's shared LazyFrames frames were de-duplicated again: ~1 frame per transition
            assert dst._pool_cur <= 45 + 4 * k + 3


def test_dedup_ring_full_size_synthetic_window(U):
    """BASELINE.json configs[1] size on the v2 layout: slot i holds frames i..i+k-1 / i+1..i+k of the pool, so the last
    k-1 frames of obs ARE the first k-1 frames of next_obs, and 100k transitions take ~2.1 GB instead of 12.7 GB."""
    cap, B = 100_000, 128
    buf = U.ReplayBuffer((9, 84, 84), (6,), cap, B, dedup=True)
    buf.fill_synthetic(seed=1)
    assert buf.ring_bytes() < 2.8e9
    g = torch.Generator().manual_seed(0)
    idxs = torch.randint(0, cap, (B,), generator=g).numpy()
    obs, act, rew, nxt, nd = buf.sample(idxs=idxs)
    assert torch.equal(obs[:, 3:], nxt[:, :6])
    assert obs.min() >= 0 and obs.max() <= 255 and obs.std() > 50
    o2 = buf.sample(idxs=(idxs + 1) % cap)[0]
    keep = torch.from_numpy(idxs + 1 < cap)
    assert torch.equal(o2[keep], nxt[keep])
    off = torch.randint(0, 9, (B, 2), generator=g).numpy()
    sh = buf.sample_drq(idxs=idxs, offsets={"obs": off, "next_obs": off})
    assert torch.equal(sh[0][:, 3:], sh[3][:, :6])
    with pytest.raises(RuntimeError):
        buf.add(np.zeros((9, 84, 84), np.uint8), np.zeros(6, np.float32), 0.0, np.zeros((9, 84, 84), np.uint8), False)
