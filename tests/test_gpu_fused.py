"""GPU parity of the fused per-algorithm sampler (augcl_ring_sample_fused, ReplayBuffer.sample_fused /
sample_augcl / sample_svea_aug) against the CPU oracle: oracle gather followed by the oracle's
random_shift / splice / CS_splice / random_overlay / random_conv in the reference's order
(src/algorithms/augcl.py:69-72, svea.py:66-70 + :45-52, rad.py:18-22).
Bit-exact for the copies and the splice select; <= 1e-5 on the /255 scale for overlay and conv."""
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


def dmc_like_frames(rng, n, S):
    """Frames with flat bright low-saturation background and saturated blobs, so splice keeps 20-50 % of the pixels."""
    fr = np.empty((n, 3, S, S), dtype=np.uint8)
    for i in range(n):
        base = rng.randint(150, 256)
        fr[i] = base
        fr[i, :, : S // 3] = rng.randint(0, 120)                       # dark band: below the value threshold
        for _ in range(4):
            y, x = rng.randint(0, S - 6, 2)
            hh, ww = rng.randint(3, S // 2, 2)
            fr[i, :, y:y + hh, x:x + ww] = rng.randint(0, 256, (3, 1, 1))
        noise = rng.rand(S, S) < 0.15
        fr[i][:, noise] = rng.randint(0, 256, (3, int(noise.sum())))
    return fr


def build_pair(U, O, rng, S, k, cap, B, A_=5, steps=None):
    steps = steps or cap + 7
    buf = U.ReplayBuffer((3 * k, S, S), (A_,), cap, B)
    ora = O.HostRing((3 * k, S, S), (A_,), cap, B)
    frames = dmc_like_frames(rng, steps + k, S)
    for t in range(steps):
        a, r, dn = rng.randn(A_).astype(np.float32), float(rng.randn()), bool(rng.rand() < 0.2)
        buf.add(U.LazyFrames(list(frames[t:t + k])), a, r, U.LazyFrames(list(frames[t + 1:t + 1 + k])), dn)
        ora.add(O.LazyFrames(frames[t:t + k]), a, r, O.LazyFrames(frames[t + 1:t + 1 + k]), dn)
    return buf, ora


def oracle_chain(O, x, keys, imgs, wt, params):
    for key in keys:
        if key == "splice":
            x = O.splice(x, imgs, **params.get("splice", {}))
        elif key == "CS_splice":
            x = O.cs_splice(x, imgs, **params.get("CS_splice", {}))
        elif key == "overlay":
            x = O.random_overlay(x, imgs, params.get("overlay", {}).get("alpha", 0.5))
        elif key == "conv":
            x = O.random_conv(x, wt)
    return x


CHAINS = ["splice-conv", "splice", "CS_splice", "overlay", "conv", "overlay-conv", "CS_splice-conv"]


@pytest.mark.parametrize("S,k,pad", [(84, 3, 4), (100, 3, 4), (20, 2, 3), (84, 1, 4)])
@pytest.mark.parametrize("aug_input", ["plain", "shift"])
def test_fused_chains_against_oracle(U, A, O, S, k, pad, aug_input):
    rng = np.random.RandomState(S * 10 + k)
    cap, B = 40, 11
    buf, ora = build_pair(U, O, rng, S, k, cap, B)
    bank = torch.from_numpy(rng.randint(0, 256, (9, 3, S, S)).astype(np.uint8))
    A.set_background_bank(bank.cuda())
    try:
        for chain in CHAINS:
            keys = chain.split("-")
            idxs = rng.randint(0, cap, size=B)
            off = {"obs": rng.randint(0, 2 * pad + 1, (B, 2)), "next_obs": rng.randint(0, 2 * pad + 1, (B, 2))}
            off["obs"][0] = (0, 0); off["obs"][1] = (2 * pad, 2 * pad); off["obs"][2] = (0, 2 * pad)
            rows = rng.randint(0, 9, size=B)
            wt = torch.from_numpy(rng.randn(B, 3, 3, 3, 3).astype(np.float32))
            params = {"splice": {"val_thres": 0.4}} if chain == "splice" else {}
            out = buf.sample_fused(idxs=idxs, pad=pad, data_aug=chain, aug_params=params, aug_input=aug_input,
                                   want=("obs", "obs_shift", "obs_aug", "next_obs", "next_obs_shift"), offsets=off,
                                   bank_idx=rows, weights=wt)
            o_obs, o_act, o_rew, o_next, o_nd = ora.sample(idxs)
            o_shift = O.random_shift_int(o_obs, off["obs"], pad)
            assert torch.equal(out["obs"].cpu(), o_obs), chain
            assert torch.equal(out["obs_shift"].cpu(), o_shift), chain
            assert torch.equal(out["next_obs"].cpu(), o_next), chain
            assert torch.equal(out["next_obs_shift"].cpu(), O.random_shift_int(o_next, off["next_obs"], pad)), chain
            assert torch.equal(out["action"].cpu(), o_act) and torch.equal(out["reward"].cpu(), o_rew)
            assert torch.equal(out["not_done"].cpu(), o_nd)
            imgs = O.bank_rows_f32(bank, rows)
            want = oracle_chain(O, o_shift if aug_input == "shift" else o_obs, keys, imgs, wt, params)
            got = out["obs_aug"].cpu()
            if chain in ("splice", "CS_splice"):
                assert torch.equal(got, want), f"{chain}/{aug_input}: select must be bit-exact"
            else:
                err = (got - want).abs().max().item() / 255.0
                assert err <= TOL_NORM, f"{chain}/{aug_input}: {err}"
    finally:
        A.set_background_bank(None)


def test_fused_subset_outputs_normalize_and_wrappers(U, A, O):
    rng = np.random.RandomState(7)
    S, k, cap, B, pad = 84, 3, 64, 16, 4
    buf, ora = build_pair(U, O, rng, S, k, cap, B)
    bank = torch.from_numpy(rng.randint(0, 256, (5, 3, S, S)).astype(np.uint8))
    A.set_background_bank(bank.cuda())
    try:
        idxs = rng.randint(0, cap, size=B)
        off = rng.randint(0, 2 * pad + 1, (B, 2))
        rows = rng.randint(0, 5, size=B)
        wt = torch.from_numpy(rng.randn(B, 3, 3, 3, 3).astype(np.float32))
        o_obs, o_act, o_rew, o_next, o_nd = ora.sample(idxs)
        imgs = O.bank_rows_f32(bank, rows)
        # AugCL wrapper (src/algorithms/augcl.py:69-72)
        sh, aug, act, rew, nxt, nd = buf.sample_augcl(idxs=idxs, offsets={"obs": off}, bank_idx=rows, weights=wt)
        assert torch.equal(sh.cpu(), O.random_shift_int(o_obs, off, pad)) and torch.equal(nxt.cpu(), o_next)
        want = O.random_conv(O.splice(o_obs, imgs), wt)
        assert (aug.cpu() - want).abs().max().item() / 255.0 <= TOL_NORM
        assert torch.equal(act.cpu(), o_act) and torch.equal(rew.cpu(), o_rew) and torch.equal(nd.cpu(), o_nd)
        # the unfused product path gives the same tensors as the fused one
        plain = buf.sample(idxs=idxs)[0]
        chain = A.random_conv(A.splice(plain, bank_idx=rows), weights=wt)
        assert (aug - chain).abs().max().item() / 255.0 <= TOL_NORM
        # SVEA wrapper (src/algorithms/svea.py:66-70): overlay of the SHIFTED obs
        sh2, aug2, _, _, nxt2, _ = buf.sample_svea_aug(idxs=idxs, offsets={"obs": off}, bank_idx=rows)
        assert torch.equal(sh2, sh) and torch.equal(nxt2, nxt)
        want2 = O.random_overlay(O.random_shift_int(o_obs, off, pad), imgs)
        assert (aug2.cpu() - want2).abs().max().item() / 255.0 <= TOL_NORM
        # subsets + normalize: every output divided by 255 (IEEE)
        out = buf.sample_fused(idxs=idxs, data_aug="splice", want=("obs_aug",), bank_idx=rows, normalize=True)
        assert set(out) == {"obs_aug", "action", "reward", "not_done"}
        assert torch.equal(out["obs_aug"].cpu(), O.splice(o_obs, imgs) / 255.0)
        out = buf.sample_fused(idxs=idxs, want=("next_obs",))
        assert torch.equal(out["next_obs"].cpu(), o_next)
        # device-side draws: same Philox streams as sample_svea / augcl_rng_fill_offsets
        A.seed(123)
        a1 = buf.sample_fused(idxs=idxs, want=("obs_shift", "next_obs"))
        A.seed(123)
        b1 = buf.sample_svea(idxs=idxs)
        assert torch.equal(a1["obs_shift"], b1[0]) and torch.equal(a1["next_obs"], b1[3])
        A.seed(5)
        a2 = buf.sample_fused(n=B, want=("obs", "next_obs"), device_indices=True)
        A.seed(5)
        b2 = buf.sample(n=B, device_indices=True)
        assert torch.equal(a2["obs"], b2[0]) and torch.equal(a2["next_obs"], b2[3]) and torch.equal(a2["reward"], b2[2])
        with pytest.raises(NotImplementedError):
            buf.sample_fused(idxs=idxs, data_aug="conv-splice", want=("obs_aug",))
        with pytest.raises(U.L.AugclError):
            buf.sample_fused(idxs=np.array([cap] * B), want=("obs",))
    finally:
        A.set_background_bank(None)


def test_fused_full_size_properties(U, A):
    """BASELINE.json configs[3] size (batch 128, 100k ring): fused == unfused product chain on the same draws."""
    cap, B, S, k = 100_000, 128, 84, 3
    buf = U.ReplayBuffer((3 * k, S, S), (6,), cap, B)
    buf.fill_synthetic(seed=3)
    A.set_background_bank(None)
    A.seed(11)
    A.load_dataloader(B, S)
    g = torch.Generator().manual_seed(0)
    idxs = torch.randint(0, cap, (B,), generator=g).numpy()
    off = torch.randint(0, 9, (B, 2), generator=g).numpy()
    rows = torch.randint(0, A.SYNTHETIC_BANK_LEN, (B,), generator=g).numpy()
    wt = torch.randn((B, 3, 3, 3, 3), generator=g)
    try:
        sh, aug, act, rew, nxt, nd = buf.sample_augcl(idxs=idxs, offsets={"obs": off}, bank_idx=rows, weights=wt)
        obs, act2, rew2, nxt2, nd2 = buf.sample(idxs=idxs)
        assert torch.equal(nxt, nxt2) and torch.equal(act, act2) and torch.equal(rew, rew2) and torch.equal(nd, nd2)
        assert torch.equal(sh, A.random_shift(obs, 4, offsets=off))
        chain = A.random_conv(A.splice(obs, bank_idx=rows), weights=wt)
        assert (aug - chain).abs().max().item() / 255.0 <= TOL_NORM
        sp = buf.sample_fused(idxs=idxs, data_aug="splice", want=("obs_aug",), bank_idx=rows)["obs_aug"]
        assert torch.equal(sp, A.splice(obs, bank_idx=rows))
        # batch 1024 exercises > 2 units per CTA on every CTA
        idx2 = torch.randint(0, cap, (1024,), generator=g).numpy()
        rows2 = torch.randint(0, A.SYNTHETIC_BANK_LEN, (1024,), generator=g).numpy()
        o2 = buf.sample_fused(idxs=idx2, data_aug="overlay", aug_input="shift", want=("obs_shift", "obs_aug", "next_obs"),
                              bank_idx=rows2, offsets={"obs": np.full((1024, 2), 3)})
        assert (o2["obs_aug"] - A.random_overlay(o2["obs_shift"], bank_idx=rows2)).abs().max().item() / 255.0 <= TOL_NORM
        assert torch.equal(o2["next_obs"], buf.sample(idxs=idx2)[3])
    finally:
        A.set_background_bank(None)


def test_fast_splice_mask_equals_float_hsv_on_all_colours(U, A):
    """The default-threshold splice mask of the fused kernel runs on raw bytes (fused.cu, kPreSpliceFast); it must
    select exactly like the float kornia-HSV path (stand-alone augcl_splice, bit-exact vs the oracle) on every one of
    the 2^24 u8 colours, for the reference's val_thres 0.6 (default) and 0.4, and like the general fused path."""
    from augcl_b200 import _lib as L
    S, P = 84, 84 * 84
    n = (1 << 24 + 0) // P + 1
    col = np.arange(n * P, dtype=np.int64) & 0xFFFFFF
    frames = np.stack([(col >> 16) & 255, (col >> 8) & 255, col & 255], axis=0).astype(np.uint8)   # [3, n*P]
    frames = np.ascontiguousarray(frames.reshape(3, n, S, S).transpose(1, 0, 2, 3))                 # [n,3,S,S]
    buf = U.ReplayBuffer((3, S, S), (1,), n, 64)
    z = np.zeros((n, 1), dtype=np.float32)
    L.call("augcl_ring_put", buf._ring, 0, n, frames.ctypes.data, frames.ctypes.data, z.ctypes.data, z.ctypes.data,
           z.ctypes.data, None)
    torch.cuda.synchronize()
    buf.idx, buf.full = 0, True
    rng = np.random.RandomState(0)
    bank = torch.from_numpy(rng.randint(0, 256, (4, 3, S, S)).astype(np.uint8)).cuda()
    A.set_background_bank(bank)
    try:
        idxs = np.arange(n)
        rows = rng.randint(0, 4, size=n)
        plain = buf.sample(idxs=idxs)[0]
        assert torch.equal(plain.cpu().to(torch.uint8), torch.from_numpy(frames))
        for vt in (0.6, 0.4):
            prm = {"splice": {"val_thres": vt}}
            fast = buf.sample_fused(idxs=idxs, data_aug="splice", aug_params=prm, want=("obs_aug",), bank_idx=rows)["obs_aug"]
            ref = A.splice(plain, val_thres=vt, bank_idx=rows)
            assert torch.equal(fast, ref), f"val_thres {vt}: {(fast != ref).sum().item()} elements differ"
            L.set_option("fused_fast_mask", "0")
            try:
                slow = buf.sample_fused(idxs=idxs, data_aug="splice", aug_params=prm, want=("obs_aug",), bank_idx=rows)["obs_aug"]
            finally:
                L.set_option("fused_fast_mask", "1")
            assert torch.equal(slow, ref)
    finally:
        A.set_background_bank(None)


def _ring_stacks(buf, idxs):
    """u8 (obs, next_obs) stacks of the sampled slots, read back slot by slot with augcl_ring_get (host tensors)."""
    from augcl_b200 import _lib as L
    c, h, w = buf.obs_shape
    obs = torch.empty((len(idxs), c, h, w), dtype=torch.uint8, device="cuda")
    nxt = torch.empty_like(obs)
    st = torch.cuda.current_stream().cuda_stream
    for j, i in enumerate(idxs):
        L.call("augcl_ring_get", buf._ring, int(i), 1, obs[j].data_ptr(), nxt[j].data_ptr(), None, None, None, st)
    torch.cuda.synchronize()
    return obs.cpu().float(), nxt.cpu().float()


@pytest.mark.parametrize("chain,aug_input,B", [("overlay", "shift", 256), ("splice-conv", "plain", 128)])
def test_fused_against_oracle_at_baseline_sizes(U, A, O, chain, aug_input, B):
    """BASELINE.json configs[2] (SVEA shift + overlay, batch 256) and configs[3] (AugCL splice-conv, batch 128) at
    full size -- 100k ring, 9x84x84 -- against the ORACLE chain (not the unfused product), fed with the u8 stacks read
    back from the sampled ring slots."""
    cap, S, k, pad = 100_000, 84, 3, 4
    buf = U.ReplayBuffer((3 * k, S, S), (6,), cap, B)
    buf.fill_synthetic(seed=17)
    g = torch.Generator().manual_seed(1)
    bank = torch.randint(0, 256, (64, 3, S, S), generator=g, dtype=torch.uint8)
    A.set_background_bank(bank.cuda())
    try:
        idxs = torch.randint(0, cap, (B,), generator=g).numpy()
        off = {"obs": torch.randint(0, 2 * pad + 1, (B, 2), generator=g).numpy()}
        rows = torch.randint(0, 64, (B,), generator=g).numpy()
        wt = torch.randn((B, 3, 3, 3, 3), generator=g)
        out = buf.sample_fused(idxs=idxs, pad=pad, data_aug=chain, aug_input=aug_input, offsets=off, bank_idx=rows, weights=wt,
                               want=("obs_shift", "obs_aug", "next_obs"))
        o_obs, o_next = _ring_stacks(buf, idxs)
        o_shift = O.random_shift_int(o_obs, off["obs"], pad)
        assert torch.equal(out["obs_shift"].cpu(), o_shift)
        assert torch.equal(out["next_obs"].cpu(), o_next)
        imgs = O.bank_rows_f32(bank, rows)
        want = oracle_chain(O, o_shift if aug_input == "shift" else o_obs, chain.split("-"), imgs, wt, {})
        err = (out["obs_aug"].cpu() - want).abs().max().item() / 255.0
        assert err <= TOL_NORM, f"{chain}: {err}"
    finally:
        A.set_background_bank(None)


@pytest.mark.parametrize("chain,B", [("splice-conv", 1024), ("overlay", 640), ("conv", 300), ("", 2048)])
def test_fused_repeated_launches_are_bit_identical(U, A, chain, B):
    """The warp-specialised kernel hands stages and tiles over through mbarriers only; units are claimed dynamically, so every
    launch runs a different schedule.  Same draws -> bit-identical outputs, launch after launch (a missed hand-over shows up as
    a stale stage or tile), and equal to the unfused product chain."""
    cap, S, k = 20_000, 84, 3
    buf = U.ReplayBuffer((3 * k, S, S), (6,), cap, B)
    buf.fill_synthetic(seed=23)
    A.set_background_bank(None)
    A.seed(5)
    A.load_dataloader(64, S)
    g = torch.Generator().manual_seed(3)
    idxs = torch.randint(0, cap, (B,), generator=g).numpy()
    off = {"obs": torch.randint(0, 9, (B, 2), generator=g).numpy(), "next_obs": torch.randint(0, 9, (B, 2), generator=g).numpy()}
    rows = torch.randint(0, A.SYNTHETIC_BANK_LEN, (B,), generator=g).numpy()
    wt = torch.randn((B, 3, 3, 3, 3), generator=g)
    want = ("obs_shift", "obs_aug", "next_obs", "next_obs_shift") if chain else ("obs_shift", "next_obs_shift")
    try:
        first = None
        for rep in range(12):
            out = buf.sample_fused(idxs=idxs, data_aug=chain, aug_input="plain", offsets=off, bank_idx=rows, weights=wt, want=want)
            cur = {key: out[key].clone() for key in want}
            if first is None:
                first = cur
                continue
            for key in want:
                assert torch.equal(cur[key], first[key]), (chain, rep, key)
        obs, _, _, nxt, _ = buf.sample(idxs=idxs)
        assert torch.equal(first["obs_shift"], A.random_shift(obs, 4, offsets=off["obs"]))
        assert torch.equal(first["next_obs_shift"], A.random_shift(nxt, 4, offsets=off["next_obs"]))
    finally:
        A.set_background_bank(None)


def test_two_streams_sample_one_ring_concurrently(U, A):
    """One set of schedule counters per (ring, stream): launches of the fused sampler that overlap on two streams each
    process exactly their own units (round 1 had ONE counter per ring and documented the race instead)."""
    cap, S, k, B = 20_000, 84, 3, 1024
    buf = U.ReplayBuffer((3 * k, S, S), (6,), cap, B)
    buf.fill_synthetic(seed=29)
    g = torch.Generator().manual_seed(4)
    draws = [(torch.randint(0, cap, (B,), generator=g).numpy(), {"obs": torch.randint(0, 9, (B, 2), generator=g).numpy()}) for _ in range(2)]
    ref = [buf.sample_fused(idxs=i, offsets=o, data_aug="", want=("obs_shift", "next_obs")) for i, o in draws]
    ref = [{key: r[key].clone() for key in ("obs_shift", "next_obs")} for r in ref]
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    for rep in range(6):
        outs = []
        for st, (i, o) in zip(streams, draws):
            with torch.cuda.stream(st):
                outs.append(buf.sample_fused(idxs=i, offsets=o, data_aug="", want=("obs_shift", "next_obs")))
        torch.cuda.synchronize()
        for got, want in zip(outs, ref):
            assert torch.equal(got["obs_shift"], want["obs_shift"]) and torch.equal(got["next_obs"], want["next_obs"]), rep


@pytest.mark.parametrize("H,W,k,pad,B", [(8, 8, 1, 2, 5), (12, 16, 2, 3, 7), (64, 48, 2, 4, 40), (128, 128, 1, 4, 9), (16, 128, 3, 2, 12),
                                         (52, 8, 2, 1, 6)])
def test_fused_geometries_equal_the_unfused_chain(U, A, H, W, k, pad, B):
    """The generic-width instantiations: non-square frames, the narrowest / widest supported rows, bands shorter than a
    tile, on every chain: all five image outputs against the unfused product kernels (themselves oracle-checked)."""
    buf = U.ReplayBuffer((3 * k, H, W), (4,), 300, B)
    buf.fill_synthetic(seed=H * 7 + W)
    A.set_background_bank(torch.randint(0, 256, (13, 3, H, W), dtype=torch.uint8, device="cuda"))
    try:
        g = torch.Generator().manual_seed(H + W)
        idxs = torch.randint(0, 300, (B,), generator=g).numpy()
        off = {"obs": torch.randint(0, 2 * pad + 1, (B, 2), generator=g).numpy(),
               "next_obs": torch.randint(0, 2 * pad + 1, (B, 2), generator=g).numpy()}
        rows = torch.randint(0, 13, (B,), generator=g).numpy()
        wt = torch.randn((B, 3, 3, 3, 3), generator=g)
        obs, _, _, nxt, _ = buf.sample(idxs=idxs)
        sh = A.random_shift(obs, pad, offsets=off["obs"])
        nsh = A.random_shift(nxt, pad, offsets=off["next_obs"])
        fns = {"splice": lambda x: A.splice(x, bank_idx=rows), "CS_splice": lambda x: A.CS_splice(x, bank_idx=rows),
               "overlay": lambda x: A.random_overlay(x, bank_idx=rows), "conv": lambda x: A.random_conv(x, weights=wt)}
        for chain in ["splice-conv", "overlay", "conv", "CS_splice-conv", "splice", "overlay-conv"]:
            for aug_input in ["plain", "shift"]:
                out = buf.sample_fused(idxs=idxs, pad=pad, data_aug=chain, aug_input=aug_input, offsets=off, bank_idx=rows, weights=wt,
                                       want=("obs", "obs_shift", "obs_aug", "next_obs", "next_obs_shift"))
                assert torch.equal(out["obs"], obs) and torch.equal(out["next_obs"], nxt), (chain, aug_input)
                assert torch.equal(out["obs_shift"], sh) and torch.equal(out["next_obs_shift"], nsh), (chain, aug_input)
                x = sh if aug_input == "shift" else obs
                for key in chain.split("-"):
                    x = fns[key](x)
                assert (out["obs_aug"] - x).abs().max().item() / 255.0 <= TOL_NORM, (chain, aug_input)
    finally:
        A.set_background_bank(None)


def test_in_kernel_conv_weights_equal_explicit_weights(U, A):
    """desc.conv_weights == NULL: the 81 N(0,1) weights per sample are drawn inside the fused kernel (by the B-warps) and
    are the values augcl_rng_fill_normal(out, B*81, seed, conv_subsequence) writes -- AugCL's update is ONE launch."""
    from augcl_b200 import _lib as L
    S, cap, B = 84, 90, 23
    buf = U.ReplayBuffer((9, S, S), (4,), cap, B)
    buf.fill_synthetic(seed=5)
    bank = torch.randint(0, 256, (11, 3, S, S), dtype=torch.uint8, device="cuda")
    A.set_background_bank(bank)
    try:
        idxs = np.random.RandomState(8).randint(0, cap, size=B)
        off = {"obs": np.random.RandomState(9).randint(0, 9, (B, 2))}
        rows = torch.from_numpy(np.random.RandomState(10).randint(0, 11, size=B)).cuda()
        for chain, fn in (("splice-conv", "sample_augcl"), ("conv", "sample_svea_aug")):
            A.seed(33)
            n0 = L.lib().augcl_launch_count()
            got = getattr(buf, fn)(idxs=idxs, data_aug=chain, offsets=off, bank_idx=rows)
            assert L.lib().augcl_launch_count() - n0 == 1, "one launch per update"
            # the call reserves subsequences 0..2 for its own draws; explicit bank rows take none; the weights take 3
            wt = torch.empty((B, 3, 3, 3, 3), device="cuda")
            L.call("augcl_rng_fill_normal", wt.data_ptr(), wt.numel(), 33, 3, torch.cuda.current_stream().cuda_stream)
            A.seed(33)
            want = getattr(buf, fn)(idxs=idxs, data_aug=chain, offsets=off, bank_idx=rows, weights=wt)
            for a, b in zip(got, want):
                assert torch.equal(a, b), chain
            assert float(wt.std()) > 0.8
    finally:
        A.set_background_bank(None)


@pytest.mark.parametrize("chain,fn", [("overlay", "sample_svea_aug"), ("splice", "sample_augcl")])
def test_in_kernel_bank_rows_equal_explicit_rows(U, A, chain, fn):
    """desc.bank_idx == NULL: the background rows are drawn inside the fused kernel and are the values
    augcl_rng_fill_indices(out, B, 0, bank_len, seed, subsequence) writes (no extra launch)."""
    from augcl_b200 import _lib as L
    S, cap, B = 84, 120, 20
    buf = U.ReplayBuffer((9, S, S), (4,), cap, B)
    buf.fill_synthetic(seed=9)
    bank = torch.randint(0, 256, (37, 3, S, S), dtype=torch.uint8, device="cuda")
    A.set_background_bank(bank)
    try:
        idxs = np.random.RandomState(4).randint(0, cap, size=B)
        off = {"obs": np.random.RandomState(5).randint(0, 9, (B, 2))}
        A.seed(21)
        got = getattr(buf, fn)(idxs=idxs, data_aug=chain, offsets=off)
        # the call reserves subsequences 0..2 for its own draws, then takes the next one (3) for the bank rows
        rows = torch.empty(B, dtype=torch.int64, device="cuda")
        L.call("augcl_rng_fill_indices", rows.data_ptr(), B, 0, bank.shape[0], 21, 3, torch.cuda.current_stream().cuda_stream)
        A.seed(21)
        want = getattr(buf, fn)(idxs=idxs, data_aug=chain, offsets=off, bank_idx=rows)
        for a, b in zip(got, want):
            assert torch.equal(a, b)
        assert len(set(rows.tolist())) > 5
    finally:
        A.set_background_bank(None)
