"""TEST INFRASTRUCTURE ONLY — writes tests/golden/*.npz by running the REFERENCE's own
functions (through oracle/ref_harness.py) on seeded inputs.  Run in the build
container, where /root/reference is mounted:

    python oracle/make_golden.py

The fixtures pin oracle/augcl_oracle.py (CPU tests) and the CUDA kernels (GPU
tests) to outputs of the reference itself.  Inputs are stored as uint8 where the
reference path would see buffer-fresh (integer-valued) pixels.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as R  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def dmc_like(g, n, c, h, w):
    """Frames with flat bright background + saturated blobs so splice keeps a real fraction."""
    x = torch.empty(n, c, h, w, dtype=torch.uint8)
    for i in range(n):
        for f in range(c // 3):
            bg = torch.randint(150, 256, (1,), generator=g).item()
            fr = torch.full((3, h, w), bg, dtype=torch.uint8)
            for _ in range(4):
                y0 = torch.randint(0, h - 4, (1,), generator=g).item()
                x0 = torch.randint(0, w - 4, (1,), generator=g).item()
                hh = torch.randint(3, max(4, h // 2), (1,), generator=g).item()
                ww = torch.randint(3, max(4, w // 2), (1,), generator=g).item()
                col = torch.randint(0, 256, (3, 1, 1), generator=g, dtype=torch.uint8)
                fr[:, y0:y0 + hh, x0:x0 + ww] = col
            noise = torch.randint(0, 256, (3, h, w), generator=g, dtype=torch.uint8)
            sel = torch.rand(h, w, generator=g) < 0.15
            fr[:, sel] = noise[:, sel]
            x[i, 3 * f:3 * f + 3] = fr
    return x


def case(name, seed, n, c, h, w, pad=4, crop_from=None):
    g = torch.Generator().manual_seed(seed)
    d = {}
    x_u8 = dmc_like(g, n, c, h, w)
    x = x_u8.float()
    d["x_u8"] = x_u8.numpy()
    # integer shift (reference random_shift, kornia "slice" crop) incl. the extreme offsets
    off = torch.randint(0, 2 * pad + 1, (n, 2), generator=g).numpy().astype(np.int32)
    off[0] = (0, 0)
    if n > 1:
        off[1] = (2 * pad, 2 * pad)
    if n > 2:
        off[2] = (0, 2 * pad)
    d["shift_off"] = off
    d["shift_out"] = R.ref_random_shift(x, off, pad, "slice").numpy()
    d["pad"] = np.int32(pad)
    # conv
    wts = torch.randn(n, 3, 3, 3, 3, generator=g)
    d["conv_w"] = wts.numpy()
    d["conv_out"] = R.ref_random_conv(x, wts).numpy()
    # overlay / splice with a u8 background batch (ToTensor == u8/255)
    bank = torch.randint(0, 256, (n, 3, h, w), generator=g, dtype=torch.uint8)
    imgs = bank.float() / 255.0
    d["bank_u8"] = bank.numpy()
    d["overlay_out"] = R.ref_random_overlay(x, imgs).numpy()
    d["splice_out_06"] = R.ref_splice(x, imgs, 0, 0, 0.6).numpy()
    d["splice_out_04"] = R.ref_splice(x, imgs, 0, 0, 0.4).numpy()
    d["splice_out_hs"] = R.ref_splice(x, imgs, 1.0, 0.2, 0.3).numpy()
    d["cs_splice_out"] = R.ref_cs_splice(x, imgs).numpy()
    # chained: splice -> conv (AugCL strong phase, --data_aug splice-conv); conv -> splice hits
    # the non-integer input path of the HSV mask
    d["splice_conv_out"] = R.ref_random_conv(torch.from_numpy(d["splice_out_06"]), wts).numpy()
    d["conv_splice_out"] = R.ref_splice(torch.from_numpy(d["conv_out"]), imgs, 0, 0, 0.6).numpy()
    # crop (square inputs only, like the reference)
    if crop_from is not None:
        S, size = crop_from
        xc_u8 = torch.randint(0, 256, (n, c, S, S), generator=g, dtype=torch.uint8)
        w1 = torch.randint(0, S - size, (n,), generator=g)
        h1 = torch.randint(0, S - size, (n,), generator=g)
        w1[0], h1[0] = 0, S - size - 1
        d["crop_x_u8"] = xc_u8.numpy()
        d["crop_w1"], d["crop_h1"] = w1.numpy().astype(np.int32), h1.numpy().astype(np.int32)
        d["crop_size"] = np.int32(size)
        d["crop_out"] = R.ref_random_crop(xc_u8.float(), size, w1, h1).numpy()
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **d)
    print(name, {k: v.shape for k, v in d.items() if hasattr(v, "shape") and v.ndim}, os.path.getsize(path))


def replay_case(name, seed, cap, k, S, A, n_add, B):
    """Reference ReplayBuffer.add x n_add (with wrap-around) then sample(idxs)."""
    g = torch.Generator().manual_seed(seed)
    rng = np.random.RandomState(seed)
    buf = R.ref_replay_buffer((3 * k, S, S), (A,), cap, B)
    frames = torch.randint(0, 256, (n_add + k, 3, S, S), generator=g, dtype=torch.uint8).numpy()
    actions = rng.randn(n_add, A).astype(np.float32)
    rewards = rng.randn(n_add).astype(np.float32)
    dones = rng.rand(n_add) < 0.1
    for t in range(n_add):
        obs = R.ref_lazy_frames(frames[t:t + k])
        nxt = R.ref_lazy_frames(frames[t + 1:t + 1 + k])
        buf.add(obs, actions[t], rewards[t], nxt, bool(dones[t]))
    hi = cap if buf.full else buf.idx
    idxs = rng.randint(0, hi, size=B)
    obs, act, rew, nobs, nd = R.ref_sample(buf, idxs)
    d = dict(frames=frames, actions=actions, rewards=rewards, dones=dones, idxs=idxs.astype(np.int64),
             cap=np.int32(cap), k=np.int32(k), out_obs=obs.numpy().astype(np.uint8), out_next=nobs.numpy().astype(np.uint8),
             out_act=act.numpy(), out_rew=rew.numpy(), out_nd=nd.numpy(),
             end_idx=np.int32(buf.idx), end_full=np.bool_(buf.full))
    assert (obs == obs.round()).all()
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **d)
    print(name, os.path.getsize(path))


def chunk_case(name, seed, cap, k, S, A, saves, B):
    """Reference ReplayBuffer: add transitions (two episodes: reset repeats the first frame k times,
    src/env/wrappers.py:193-197), save() at the steps in `saves` -> chunk files written BY THE REFERENCE,
    then sample(idxs) from a fresh reference buffer that load()ed them."""
    import shutil
    rng = np.random.RandomState(seed)
    n_add = saves[-1]
    out_dir = os.path.join(OUT, name)
    shutil.rmtree(out_dir, ignore_errors=True)
    os.makedirs(out_dir)
    buf = R.ref_replay_buffer((3 * k, S, S), (A,), cap, B)
    raw = rng.randint(0, 256, (n_add + 2, 3, S, S)).astype(np.uint8)
    actions = rng.randn(n_add, A).astype(np.float32)
    rewards = rng.randn(n_add).astype(np.float32)
    ep_len = n_add // 2 + 1
    window, dones = None, []
    for t in range(n_add):
        if t % ep_len == 0:
            window = [raw[t]] * k                      # reset: first frame k times (same ndarray object)
        obs = R.ref_lazy_frames(list(window))
        window = window[1:] + [raw[t + 1]]
        nxt = R.ref_lazy_frames(list(window))
        done = (t + 1) % ep_len == 0
        dones.append(done)
        buf.add(obs, actions[t], rewards[t], nxt, done)
        if t + 1 in saves:
            R.ref_buffer_save(buf, out_dir)
    fresh = R.ref_replay_buffer((3 * k, S, S), (A,), cap, B)
    R.ref_buffer_load(fresh, out_dir)
    assert fresh.idx == n_add and fresh.last_save == n_add
    idxs = rng.randint(0, n_add, size=B)
    obs, act, rew, nobs, nd = R.ref_sample(fresh, idxs)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), idxs=idxs.astype(np.int64), cap=np.int32(cap), k=np.int32(k),
                        S=np.int32(S), A=np.int32(A), n_add=np.int32(n_add),
                        out_obs=obs.numpy().astype(np.uint8), out_next=nobs.numpy().astype(np.uint8),
                        out_act=act.numpy(), out_rew=rew.numpy(), out_nd=nd.numpy())
    print(name, sorted(os.listdir(out_dir)), [os.path.getsize(os.path.join(out_dir, f)) for f in sorted(os.listdir(out_dir))])


def warp_cases(names):
    """random_shift as the pinned-era environment runs it (setup/conda.yaml:12,22 -> kornia 0.5.x: RandomCrop applies the
    crop as a bilinear warp, align_corners=True): the reference's random_shift executed with the harness's crop stand-in
    in "warp" mode on the inputs / offsets of the existing cases.  Kept in its own small file so the other fixtures stay
    byte-identical."""
    d = {}
    for name in names:
        c = dict(np.load(os.path.join(OUT, name + ".npz")))
        x = torch.from_numpy(c["x_u8"]).float()
        d[name] = R.ref_random_shift(x, c["shift_off"], int(c["pad"]), "warp").numpy()
    path = os.path.join(OUT, "shift_warp.npz")
    np.savez_compressed(path, **d)
    print("shift_warp", {k: v.shape for k, v in d.items()}, os.path.getsize(path))


if __name__ == "__main__":
    assert R.available(), "needs /root/reference"
    os.makedirs(OUT, exist_ok=True)
    if "--warp-only" in sys.argv:
        warp_cases(["aug_84", "aug_small", "aug_rect", "aug_odd"])
        sys.exit(0)
    case("aug_84", seed=1, n=2, c=9, h=84, w=84, pad=4, crop_from=(100, 84))
    case("aug_small", seed=2, n=5, c=6, h=20, w=20, pad=3, crop_from=(24, 20))
    case("aug_rect", seed=3, n=3, c=3, h=12, w=28, pad=2)          # non-square, w%4==0
    case("aug_odd", seed=4, n=4, c=9, h=13, w=11, pad=4, crop_from=(17, 10))   # widths not multiple of 4
    replay_case("replay_small", seed=5, cap=24, k=3, S=16, A=6, n_add=37, B=10)
    replay_case("replay_84", seed=6, cap=16, k=3, S=84, A=6, n_add=12, B=4)
    chunk_case("chunks_ref", seed=7, cap=40, k=3, S=12, A=4, saves=[10, 25], B=9)
    warp_cases(["aug_84", "aug_small", "aug_rect", "aug_odd"])
