"""GPU parity of the device-resident replay ring and its fused sample kernel
(augcl_b200.utils.ReplayBuffer -> augcl_ring_* C ABI) against the reference's recorded outputs
(tests/golden/replay_*.npz), the CPU oracle, and size-independent properties at BASELINE.json's
full size (100k-transition ring)."""
import numpy as np
import pytest
import torch

from conftest import REPLAY_CASES

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def U():
    from augcl_b200 import utils
    return utils


@pytest.fixture(scope="module")
def O():
    from oracle import augcl_oracle
    return augcl_oracle


def replay_into(d, buf, LF):
    k, fr = int(d["k"]), d["frames"]
    for t in range(len(d["actions"])):
        buf.add(LF(list(fr[t:t + k])), d["actions"][t], d["rewards"][t], LF(list(fr[t + 1:t + 1 + k])), bool(d["dones"][t]))


@pytest.mark.parametrize("name", REPLAY_CASES)
def test_add_sample_against_reference_golden(U, golden, name):
    d = golden(name)
    S, k = d["frames"].shape[-1], int(d["k"])
    buf = U.ReplayBuffer((3 * k, S, S), d["actions"].shape[1:], int(d["cap"]), len(d["idxs"]))
    replay_into(d, buf, U.LazyFrames)
    assert buf.idx == int(d["end_idx"]) and buf.full == bool(d["end_full"])
    obs, act, rew, nxt, nd = buf.sample(idxs=d["idxs"])
    assert obs.dtype == torch.float32 and obs.is_cuda and obs.is_contiguous()
    assert torch.equal(obs.cpu(), torch.from_numpy(d["out_obs"]).float())
    assert torch.equal(nxt.cpu(), torch.from_numpy(d["out_next"]).float())
    assert torch.equal(act.cpu(), torch.from_numpy(d["out_act"]))
    assert torch.equal(rew.cpu(), torch.from_numpy(d["out_rew"])) and rew.shape == (len(d["idxs"]), 1)
    assert torch.equal(nd.cpu(), torch.from_numpy(d["out_nd"]))
    o2, n2, a2, r2, nd2 = buf.tensor_buffer_samples(d["idxs"])      # reference's other return order
    assert torch.equal(o2, obs) and torch.equal(n2, nxt) and torch.equal(a2, act)


@pytest.mark.parametrize("S,size,pad", [(84, 64, 4), (100, 84, 4), (21, 16, 3), (18, 12, 2)])
def test_fused_samplers_against_oracle(U, O, S, size, pad):
    """sample_drq / sample_curl / sample_svea / sample_obs == oracle gather followed by oracle aug."""
    rng = np.random.RandomState(S)
    cap, k, A_, B = 50, 3, 5, 13
    buf = U.ReplayBuffer((3 * k, S, S), (A_,), cap, B)
    ora = O.HostRing((3 * k, S, S), (A_,), cap, B)
    frames = rng.randint(0, 256, (70 + k, 3, S, S)).astype(np.uint8)
    for t in range(70):     # wraps the ring (cap 50)
        a, r, dn = rng.randn(A_).astype(np.float32), float(rng.randn()), bool(rng.rand() < 0.2)
        buf.add(U.LazyFrames(list(frames[t:t + k])), a, r, U.LazyFrames(list(frames[t + 1:t + 1 + k])), dn)
        ora.add(O.LazyFrames(frames[t:t + k]), a, r, O.LazyFrames(frames[t + 1:t + 1 + k]), dn)
    assert (buf.idx, buf.full) == (ora.idx, ora.full)
    idxs = rng.randint(0, cap, size=B)
    o_obs, o_act, o_rew, o_next, o_nd = ora.sample(idxs)

    off = {"obs": rng.randint(0, 2 * pad + 1, (B, 2)), "next_obs": rng.randint(0, 2 * pad + 1, (B, 2))}
    obs, act, rew, nxt, nd = buf.sample_drq(pad=pad, idxs=idxs, offsets=off)
    assert torch.equal(obs.cpu(), O.random_shift_int(o_obs, off["obs"], pad))
    assert torch.equal(nxt.cpu(), O.random_shift_int(o_next, off["next_obs"], pad))
    assert torch.equal(act.cpu(), o_act) and torch.equal(rew.cpu(), o_rew) and torch.equal(nd.cpu(), o_nd)

    obs, _, _, nxt, _ = buf.sample_svea(pad=pad, idxs=idxs, offsets=off["obs"])
    assert torch.equal(obs.cpu(), O.random_shift_int(o_obs, off["obs"], pad)) and torch.equal(nxt.cpu(), o_next)

    co = {key: rng.randint(0, S - size, (B, 2)) for key in ("pos", "obs", "next_obs")}
    co["obs"][0] = (0, S - size - 1)
    obs, act, rew, nxt, nd, pos = buf.sample_curl(idxs=idxs, size=size, offsets=co)
    assert obs.shape == (B, 3 * k, size, size)
    assert torch.equal(obs.cpu(), O.random_crop(o_obs, size, co["obs"][:, 0], co["obs"][:, 1]))
    assert torch.equal(nxt.cpu(), O.random_crop(o_next, size, co["next_obs"][:, 0], co["next_obs"][:, 1]))
    assert torch.equal(pos.cpu(), O.random_crop(o_obs, size, co["pos"][:, 0], co["pos"][:, 1]))

    np.random.seed(11)
    want_idx = np.random.randint(0, cap, size=B)
    np.random.seed(11)
    assert torch.equal(buf.sample_obs().cpu(), ora.sample(want_idx)[0])      # same numpy index stream as the reference

    # staging A/B: the generic (unstaged) kernel returns the same tensors
    from augcl_b200 import _lib as L
    a = buf.sample_drq(pad=pad, idxs=idxs, offsets=off)
    L.set_option("staging", "generic")
    try:
        b = buf.sample_drq(pad=pad, idxs=idxs, offsets=off)
    finally:
        L.set_option("staging", "tma")
    assert all(torch.equal(p, q) for p, q in zip(a, b))


def test_full_size_ring_properties(U):
    """BASELINE.json config 2 at full size: 100k-transition ring, k=3, 84x84, batch 128 (and 4096)."""
    from augcl_b200 import augmentations as A
    cap = 100_000
    buf = U.ReplayBuffer((9, 84, 84), (6,), cap, 128)
    buf.fill_synthetic(seed=7)
    assert buf.full and buf.idx == 0
    assert buf.ring_bytes() >= 2 * cap * 9 * 84 * 84
    g = torch.Generator().manual_seed(0)
    for B in (128, 4096):
        idxs = torch.randint(0, cap, (B,), generator=g)
        idxs[0], idxs[1] = 0, cap - 1                       # extreme slots
        obs, act, rew, nxt, nd = buf.sample(idxs=idxs.numpy())
        # integer-valued pixels in 0..255, exact
        assert torch.equal(obs, obs.round()) and float(obs.min()) >= 0 and float(obs.max()) <= 255
        # idempotence / determinism: same indices -> same bytes
        obs2 = buf.sample(idxs=idxs.numpy())[0]
        assert torch.equal(obs, obs2)
        # duplicated indices give duplicated rows
        dup = torch.cat([idxs[:5], idxs[:5]])
        od = buf.sample(idxs=dup.numpy())[0]
        assert torch.equal(od[:5], od[5:]) and torch.equal(od[:5], obs[:5])
        # fused shift == stand-alone shift kernel applied to the plain sample (two independent kernels)
        off_o, off_n = torch.randint(0, 9, (B, 2), generator=g), torch.randint(0, 9, (B, 2), generator=g)
        so, _, _, sn, _ = buf.sample_drq(idxs=idxs.numpy(), offsets={"obs": off_o, "next_obs": off_n})
        assert torch.equal(so, A.random_shift(obs, 4, offsets=off_o)) and torch.equal(sn, A.random_shift(nxt, 4, offsets=off_n))
        # centre offsets are the identity
        ident = torch.full((B, 2), 4)
        io = buf.sample_drq(idxs=idxs.numpy(), offsets={"obs": ident, "next_obs": ident})[0]
        assert torch.equal(io, obs)
        # checksum of checksums against the raw bytes read back through ring_get
        raw = torch.empty((8, 9, 84, 84), dtype=torch.uint8)
        from augcl_b200 import _lib as L
        s0 = int(idxs[2]) if int(idxs[2]) + 8 <= cap else 0
        L.call("augcl_ring_get", buf._ring, s0, 8, raw.data_ptr(), None, None, None, None, None)
        torch.cuda.synchronize()
        got = buf.sample(idxs=np.arange(s0, s0 + 8))[0]
        assert torch.equal(got.cpu(), raw.float())
        assert act.shape == (B, 6) and rew.shape == (B, 1) and nd.shape == (B, 1)
        assert torch.equal(act.cpu(), torch.from_numpy(buf.actions[idxs.numpy()]))
        assert torch.equal(nd.cpu(), torch.from_numpy(buf.not_dones[idxs.numpy()]))


def test_normalize_flag_and_empty_requests(U):
    buf = U.ReplayBuffer((9, 84, 84), (2,), 64, 8)
    buf.fill_synthetic(seed=1)
    idxs = np.arange(8)
    o = buf._sample(idxs)[0]
    on = buf._sample(idxs, normalize=True)[0]
    assert torch.equal(on.cpu(), o.cpu() / 255.0)   # IEEE division == NormalizeImg (modules.py:85-90) on the CPU oracle of record
    only_next = buf._sample(idxs, obs=False, scalars=False)
    assert only_next[0] is None and only_next[1] is None and only_next[3] is not None
    from augcl_b200._lib import AugclError
    with pytest.raises(AugclError):
        buf._sample(idxs, obs=False, next_obs=False)
    with pytest.raises(AugclError):
        buf._sample(np.zeros(0, dtype=np.int64))


@pytest.mark.parametrize("S,aug,B", [(84, "shift", 128), (84, "none", 1), (84, "shift", 1000), (100, "crop", 77),
                                      (100, "none", 130), (84, "svea", 5)])
def test_band_kernel_equals_generic_kernel_full_ring(U, S, aug, B):
    """Size-independent property: every staging variant (band / plane TMA pipeline, with and without PDL,
    odd stage and CTA counts) returns bit-identical tensors to the no-staging one-thread-per-element kernel,
    on a ring filled on the device, including out-of-range offsets at the replicate border."""
    from augcl_b200 import _lib as L
    cap = 3000
    buf = U.ReplayBuffer((9, S, S), (4,), cap, B)
    buf.fill_synthetic(seed=S + B)
    rng = np.random.RandomState(B)
    idxs = rng.randint(0, cap, size=B)
    idxs[0], idxs[-1] = 0, cap - 1
    pad = 4

    def draw():
        if aug == "shift":
            o = {"obs": rng.randint(0, 2 * pad + 1, (B, 2)), "next_obs": rng.randint(0, 2 * pad + 1, (B, 2))}
            o["obs"][0] = (0, 0); o["next_obs"][0] = (2 * pad, 2 * pad); o["obs"][-1] = (2 * pad, 0)
            return lambda: buf.sample_drq(pad=pad, idxs=idxs, offsets=o)
        if aug == "svea":
            o = rng.randint(0, 2 * pad + 1, (B, 2))
            return lambda: buf.sample_svea(pad=pad, idxs=idxs, offsets=o)
        if aug == "crop":
            o = {k: rng.randint(0, S - 84, (B, 2)) for k in ("pos", "obs", "next_obs")}
            o["obs"][0] = (0, 0); o["next_obs"][0] = (S - 85, S - 85)
            return lambda: buf.sample_curl(idxs=idxs, size=84, offsets=o)
        return lambda: buf.sample(idxs=idxs)

    call = draw()
    try:
        L.set_option("staging", "generic")
        want = [t.clone() for t in call() if t is not None]
        L.set_option("staging", "tma")
        for band, pdl, stages, cps in [(1, 0, 3, 6), (1, 1, 3, 6), (1, 1, 2, 7), (1, 0, 4, 3), (1, 0, 1, 1), (0, 0, 3, 6), (0, 1, 2, 5)]:
            for k, v in (("sample_band", band), ("sample_pdl", pdl), ("sample_stages", stages), ("sample_ctas_per_sm", cps)):
                L.set_option(k, str(v))
            for rep in range(2):       # back-to-back launches exercise the PDL overlap
                got = [t for t in call() if t is not None]
            for w, g in zip(want, got):
                assert torch.equal(w, g), (band, pdl, stages, cps)
    finally:
        for k, v in (("staging", "tma"), ("sample_band", 1), ("sample_pdl", 1), ("sample_stages", 0), ("sample_ctas_per_sm", 6)):
            L.set_option(k, str(v))


@pytest.mark.parametrize("S,B", [(84, 128), (84, 300), (100, 64), (36, 10)])
def test_fused_philox_draws_equal_explicit_draws(U, O, S, B):
    """Offsets / indices drawn INSIDE the sample kernel are the same numbers augcl_rng_fill_* produce for the same
    (seed, subsequence) -- and the same the oracle's Philox restatement produces -- so a fused-draw call equals the
    call with those draws supplied explicitly (inline by-value path for B <= 256, staged path above)."""
    from augcl_b200 import _lib as L
    from augcl_b200 import augmentations as A
    cap, pad = 500, 4
    buf = U.ReplayBuffer((9, S, S), (3,), cap, B)
    buf.fill_synthetic(seed=7)
    A.seed(4242)
    sub = A._calls                      # the next _sample reserves (sub, sub+1, sub+2)
    got = buf.sample_drq(pad=pad, device_indices=True)
    idx = O.rng_randint(B, 0, cap, 4242, sub)
    o1 = O.rng_randint(2 * B, 0, 2 * pad + 1, 4242, sub + 1).reshape(B, 2)
    o2 = O.rng_randint(2 * B, 0, 2 * pad + 1, 4242, sub + 2).reshape(B, 2)
    want = buf.sample_drq(pad=pad, idxs=idx, offsets={"obs": o1, "next_obs": o2})
    for g, w in zip(got, want):
        assert torch.equal(g, w)
    # SVEA pattern: obs drawn, next_obs untouched
    sub = A._calls
    got = buf.sample_svea(pad=pad, idxs=idx)
    o1 = O.rng_randint(2 * B, 0, 2 * pad + 1, 4242, sub + 1).reshape(B, 2)
    want = buf.sample_svea(pad=pad, idxs=idx, offsets=o1)
    plain = buf.sample(idxs=idx)
    assert torch.equal(got[0], want[0]) and torch.equal(got[3], plain[3]) and torch.equal(want[3], plain[3])
    # same call with DEVICE tensors for indices and offsets
    dev = buf.sample_svea(pad=pad, idxs=torch.from_numpy(idx).cuda(), offsets=torch.from_numpy(o1).cuda())
    assert torch.equal(dev[0], want[0]) and torch.equal(dev[1], want[1])
    if S > 84:      # CURL: crop draws have the reference's exclusive upper bound
        sub = A._calls
        got = buf.sample_curl(idxs=idx, size=84)
        cm = S - 84
        oo = O.rng_randint(2 * B, 0, cm, 4242, sub + 1).reshape(B, 2)
        on = O.rng_randint(2 * B, 0, cm, 4242, sub + 2).reshape(B, 2)
        op = O.rng_randint(2 * B, 0, cm, 4242, sub + 3).reshape(B, 2)     # one launch: (indices, obs, next_obs, pos)
        want = buf.sample_curl(idxs=idx, size=84, offsets={"obs": oo, "next_obs": on, "pos": op})
        for g, w in zip(got, want):
            assert torch.equal(g, w)


def test_host_draws_are_validated(U):
    from augcl_b200._lib import AugclError
    buf = U.ReplayBuffer((9, 84, 84), (2,), 50, 8)
    buf.fill_synthetic(seed=1)
    with pytest.raises(AugclError):
        buf.sample(idxs=np.array([0, 50]))           # the reference raises IndexError here
    with pytest.raises(AugclError):
        buf.sample(idxs=np.array([-1, 3]))
    with pytest.raises(AugclError):
        buf.sample_drq(pad=4, idxs=np.array([1, 2]), offsets={"obs": np.array([[0, 9], [0, 0]]), "next_obs": np.zeros((2, 2))})
    with pytest.raises(ValueError):
        buf.sample_drq(pad=4, idxs=np.array([1, 2]), offsets={"obs": np.zeros((3, 2)), "next_obs": np.zeros((2, 2))})


def test_sweep_sizes_and_edge_batches(U):
    """BASELINE.json configs[4] sizes: the batch-16384 launch (the sweep's maximum) equals the same draws taken 128 at a
    time; batch 1, repeated indices and the ring's first / last slot behave like any other."""
    cap, S = 100_000, 84
    buf = U.ReplayBuffer((9, S, S), (6,), cap, 128)
    buf.fill_synthetic(seed=9)
    g = torch.Generator().manual_seed(4)
    B = 16384
    idxs = torch.randint(0, cap, (B,), generator=g).numpy()
    idxs[:4] = (0, cap - 1, 0, cap - 1)
    idxs[100:164] = idxs[100]                                   # 64 copies of one transition
    off = {"obs": torch.randint(0, 9, (B, 2), generator=g).numpy(), "next_obs": torch.randint(0, 9, (B, 2), generator=g).numpy()}
    big = buf.sample_drq(idxs=idxs, offsets=off)
    assert big[0].shape == (B, 9, S, S) and big[1].shape == (B, 6) and big[2].shape == (B, 1)
    assert torch.equal(big[0][100], big[0][101]) is (bool((off["obs"][100] == off["obs"][101]).all()))
    for lo in (0, 128, 8192, B - 128):
        sl = slice(lo, lo + 128)
        part = buf.sample_drq(idxs=idxs[sl], offsets={"obs": off["obs"][sl], "next_obs": off["next_obs"][sl]})
        for a, b in zip(big, part):
            assert torch.equal(a[sl], b)
    one = buf.sample_drq(idxs=idxs[:1], offsets={"obs": off["obs"][:1], "next_obs": off["next_obs"][:1]})
    assert one[0].shape == (1, 9, S, S) and torch.equal(one[0][0], big[0][0]) and torch.equal(one[3][0], big[3][0])
    plain = buf.sample(idxs=idxs[:300])                          # > 256: the staged (not by-value) host-draw path
    ident = buf.sample_drq(idxs=idxs[:300], offsets={"obs": np.full((300, 2), 4), "next_obs": np.full((300, 2), 4)})
    assert torch.equal(plain[0], ident[0]) and torch.equal(plain[3], ident[3])      # offset == pad is the identity
    dev = buf.sample(idxs=torch.from_numpy(idxs[:300]).cuda())   # device-resident indices
    assert torch.equal(plain[0], dev[0]) and torch.equal(plain[2], dev[2])
    with pytest.raises(U.L.AugclError):
        buf.sample(idxs=np.array([0, cap]))
    with pytest.raises(U.L.AugclError):
        buf.sample(idxs=np.array([-1]))


def test_sample_curl_is_one_launch_with_three_independent_crops(U, O):
    """sample_curl (src/utils.py:247-258): obs, next_obs and pos come out of ONE kernel launch; offsets that are not
    supplied are drawn in the kernel from three distinct Philox subsequences (the same values augcl_rng_fill_offsets gives)."""
    from augcl_b200 import _lib as L, augmentations as A
    rng = np.random.RandomState(11)
    S, size, k, cap, B = 100, 84, 3, 32, 10
    buf = U.ReplayBuffer((3 * k, S, S), (2,), cap, B)
    ora = O.HostRing((3 * k, S, S), (2,), cap, B)
    frames = rng.randint(0, 256, (cap + k, 3, S, S)).astype(np.uint8)
    for t in range(cap):
        a = rng.randn(2).astype(np.float32)
        buf.add(U.LazyFrames(list(frames[t:t + k])), a, 1.0, U.LazyFrames(list(frames[t + 1:t + 1 + k])), False)
        ora.add(O.LazyFrames(frames[t:t + k]), a, 1.0, O.LazyFrames(frames[t + 1:t + 1 + k]), False)
    idxs = rng.randint(0, cap, size=B)
    o_obs, _, _, o_next, _ = ora.sample(idxs)
    A.seed(77)
    n0 = L.launch_count()
    obs, act, rew, nxt, nd, pos = buf.sample_curl(idxs=idxs, size=size)
    assert L.launch_count() - n0 == 1
    draws = [O.rng_randint(2 * B, 0, S - size, 77, sub).reshape(B, 2) for sub in (1, 2, 3)]
    assert torch.equal(obs.cpu(), O.random_crop(o_obs, size, draws[0][:, 0], draws[0][:, 1]))
    assert torch.equal(nxt.cpu(), O.random_crop(o_next, size, draws[1][:, 0], draws[1][:, 1]))
    assert torch.equal(pos.cpu(), O.random_crop(o_obs, size, draws[2][:, 0], draws[2][:, 1]))
    assert not np.array_equal(draws[0], draws[2])
    # mixed: one supplied (host) offset array, the other two drawn
    A.seed(78)
    mine = rng.randint(0, S - size, (B, 2))
    obs2, _, _, nxt2, _, pos2 = buf.sample_curl(idxs=idxs, size=size, offsets={"pos": mine})
    assert torch.equal(pos2.cpu(), O.random_crop(o_obs, size, mine[:, 0], mine[:, 1]))
    d1 = O.rng_randint(2 * B, 0, S - size, 78, 1).reshape(B, 2)
    assert torch.equal(obs2.cpu(), O.random_crop(o_obs, size, d1[:, 0], d1[:, 1]))


@pytest.mark.parametrize("S,B", [(84, 24), (84, 300), (20, 7)])
def test_sample_drq_km_one_launch(U, O, S, B):
    """DrQ's K shifted copies of obs and M of next_obs (src/algorithms/drq.py:67-70) from one launch: each view equals
    random_shift of the gathered tensor with that view's offsets; offsets not supplied come from the view's own Philox
    subsequence (seed, sub + 3 + view)."""
    from augcl_b200 import _lib as L, augmentations as A
    rng = np.random.RandomState(S + B)
    cap, pad, k = 60, 4 if S > 30 else 2, 3
    buf = U.ReplayBuffer((3 * k, S, S), (3,), cap, B)
    ora = O.HostRing((3 * k, S, S), (3,), cap, B)
    frames = rng.randint(0, 256, (cap + k, 3, S, S)).astype(np.uint8)
    for t in range(cap):
        a = rng.randn(3).astype(np.float32)
        buf.add(U.LazyFrames(list(frames[t:t + k])), a, float(t), U.LazyFrames(list(frames[t + 1:t + 1 + k])), t % 7 == 0)
        ora.add(O.LazyFrames(frames[t:t + k]), a, float(t), O.LazyFrames(frames[t + 1:t + 1 + k]), t % 7 == 0)
    idxs = rng.randint(0, cap, size=B)
    o_obs, o_act, o_rew, o_next, o_nd = ora.sample(idxs)
    offs = {"obs": [rng.randint(0, 2 * pad + 1, (B, 2)) for _ in range(2)],
            "next_obs": [rng.randint(0, 2 * pad + 1, (B, 2)) for _ in range(3)]}
    n0 = L.launch_count()
    obs_l, act, rew, next_l, nd = buf.sample_drq_km(pad=pad, k=2, m=3, idxs=idxs, offsets=offs)
    assert L.launch_count() - n0 == 1 and len(obs_l) == 2 and len(next_l) == 3
    for v, o in zip(obs_l, offs["obs"]):
        assert torch.equal(v.cpu(), O.random_shift_int(o_obs, o, pad))
    for v, o in zip(next_l, offs["next_obs"]):
        assert torch.equal(v.cpu(), O.random_shift_int(o_next, o, pad))
    assert torch.equal(act.cpu(), o_act) and torch.equal(rew.cpu(), o_rew) and torch.equal(nd.cpu(), o_nd)
    A.seed(31)
    obs_l, _, _, next_l, _ = buf.sample_drq_km(pad=pad, k=2, m=2, idxs=idxs)
    draws = [O.rng_randint(2 * B, 0, 2 * pad + 1, 31, 3 + v).reshape(B, 2) for v in range(4)]
    assert torch.equal(obs_l[0].cpu(), O.random_shift_int(o_obs, draws[0], pad))
    assert torch.equal(obs_l[1].cpu(), O.random_shift_int(o_obs, draws[1], pad))
    assert torch.equal(next_l[0].cpu(), O.random_shift_int(o_next, draws[2], pad))
    assert torch.equal(next_l[1].cpu(), O.random_shift_int(o_next, draws[3], pad))
    assert not torch.equal(obs_l[0], obs_l[1])


@pytest.mark.parametrize("S,size", [(100, 84), (84, 84), (100, None), (20, 12)])
def test_sample_encoder_input_is_centercrop_plus_normalize(U, S, size):
    """NormalizeImg(CenterCrop(size)(obs)) of src/algorithms/modules.py:69-90 out of the sample kernel,
    bit-equal to slicing and dividing the plain sample() with torch (x[:, :, p:-p, p:-p] / 255.0)."""
    cap, B = 60, 10
    buf = U.ReplayBuffer((9, S, S), (3,), cap, B)
    buf.fill_synthetic(seed=S)
    idxs = np.random.RandomState(S).randint(0, cap, size=B)
    plain = buf.sample(idxs=idxs)
    got = buf.sample_encoder_input(idxs=idxs, center_crop=size)
    p = 0 if size is None else (S - size) // 2
    for i in (0, 3):
        x = plain[i].cpu()
        want = (x[:, :, p:S - p, p:S - p] if p else x) / 255.0
        assert torch.equal(got[i].cpu(), want)
    for i in (1, 2, 4):
        assert torch.equal(got[i], plain[i])
    if S == 20:
        with pytest.raises(ValueError):
            buf.sample_encoder_input(idxs=idxs, center_crop=13)


def test_recycled_output_sets_behave_like_fresh_tensors(U):
    """ReplayBuffer._outputs hands a set of output tensors out again only when nothing outside the pool can see it: the
    reference's contract (fresh tensors from every sample(), src/utils.py:206-214) as far as a caller can observe."""
    buf = U.ReplayBuffer((9, 84, 84), (3,), 256, 8)
    buf.fill_synthetic(seed=2)
    rng = np.random.RandomState(0)

    def draw():
        return buf.sample(idxs=rng.randint(0, 256, 8))

    a = draw()
    a_copy = [t.clone() for t in a]
    b = draw()
    assert all(x.data_ptr() != y.data_ptr() for x, y in zip(a, b))          # both alive: never the same memory
    for _ in range(5):
        draw()                                                               # results dropped at once: sets are recycled
    assert all(torch.equal(x, y) for x, y in zip(a, a_copy)), "a live result was overwritten"
    assert len(buf._out_pool[(8, 84, 84)]) <= U.ReplayBuffer.POOL_SETS
    # only a VIEW of a result stays alive
    view = draw()[0][2:4, 3]
    view_copy = view.clone()
    for _ in range(6):
        draw()
    assert torch.equal(view, view_copy), "memory behind a live view was recycled"
    # autograd holds a result (saved for backward) after the caller dropped it
    w = torch.ones(8, 9, 84, 84, device="cuda", requires_grad=True)
    loss = (w * draw()[0]).sum()
    expect = None
    for _ in range(6):
        draw()
    loss.backward()
    expect = w.grad.clone()
    obs = buf.sample(idxs=np.arange(8))[0]
    w.grad = None
    (w * obs).sum().backward()
    assert w.grad is not None and expect is not None and torch.isfinite(expect).all()
    # every set busy -> plain allocation, still correct
    held = [draw() for _ in range(U.ReplayBuffer.POOL_SETS + 2)]
    ptrs = {h[0].data_ptr() for h in held}
    assert len(ptrs) == len(held)
    # recycling really happens when results are dropped (otherwise the pool is dead weight)
    p1 = draw()[0].data_ptr()
    del held
    seen = {draw()[0].data_ptr() for _ in range(8)}
    assert len(seen) <= U.ReplayBuffer.POOL_SETS
    # pool off == plain allocations, same values
    plain = U.ReplayBuffer((9, 84, 84), (3,), 256, 8, pool_outputs=False)
    plain.fill_synthetic(seed=2)
    i = rng.randint(0, 256, 8)
    assert all(torch.equal(x, y) for x, y in zip(buf.sample(idxs=i), plain.sample(idxs=i)))


def test_device_draws_written_just_before_the_launch_are_seen(U):
    """The sample kernels read their draws and start their first bulk copies BEFORE griddepcontrol.wait (programmatic launch).
    That is only allowed to overlap kernels that released their successor early -- this library's own kernels, which never
    write draws or ring data.  A torch kernel that writes the index / offset tensors, or the ring fill, immediately before the
    launch (no synchronize in between) must be fully visible: same result as the host-draw call."""
    S, k, cap = 84, 3, 4000
    for B in (128, 512):           # plane kernel / band-of-frames kernel
        buf = U.ReplayBuffer((3 * k, S, S), (6,), cap, B)
        for rep in range(4):
            buf.fill_synthetic(seed=100 + rep)                                      # kernels, not synchronized
            base = torch.arange(B, device="cuda", dtype=torch.int64)
            idx_dev = (base * 7919 + 13 * rep) % cap                                  # written by torch kernels on the same stream
            off_dev = ((base.view(B, 1) * 3 + torch.tensor([rep, 2 * rep + 1], device="cuda")) % 9).to(torch.int32)
            got = buf.sample_drq(pad=4, idxs=idx_dev, offsets={"obs": off_dev, "next_obs": off_dev})
            idx_h, off_h = idx_dev.cpu().numpy(), off_dev.cpu().numpy()
            want = buf.sample_drq(pad=4, idxs=idx_h, offsets={"obs": off_h, "next_obs": off_h})
            for g, w in zip(got, want):
                assert torch.equal(g, w), (B, rep)
