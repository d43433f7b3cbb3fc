"""TEST INFRASTRUCTURE ONLY — CPU oracle for the replay-gather + augmentation hot path.

A CPU restatement (numpy / torch-CPU) of the reference algorithms the CUDA
kernels replace.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import this
module; the product package ``augcl_b200`` never does and has no CPU fallback.

Pinning: the GT-RIPL/AugCL reference has NO golden vectors or assertion tests
for this path.  This oracle is pinned instead against
outputs of the reference's own functions executed in the build container
(``oracle/ref_harness.py`` -> ``oracle/make_golden.py`` -> ``tests/golden/``),
and ``tests/test_oracle_vs_reference.py`` re-checks it live whenever
/root/reference is present.  The two kornia call sites (``RandomCrop``,
``rgb_to_hsv``; kornia is unpinned and absent) are restated from the published
kornia 0.5.x algorithm: at that boundary parity is UNPINNED by the reference.

Every function takes the random draws (indices, offsets, conv weights,
background rows) as explicit arguments: parity is defined "given identical
indices / offsets / weights" (BASELINE.json north_star).
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

HSV_EPS = 1e-6  # kornia 0.5.x rgb_to_hsv default eps
TWO_PI_F32 = np.float32(2 * math.pi)


# ----------------------------------------------------------------------------
# Replay storage + gather  (reference: src/utils.py:136-154, 216-240, 272-322)
# ----------------------------------------------------------------------------
class LazyFrames:
    """src/utils.py:283-322 — k frames concatenated along channels on demand."""

    def __init__(self, frames):
        self._frames = list(frames)

    def __array__(self, dtype=None, copy=None):
        out = np.concatenate(self._frames, axis=0)
        return out if dtype is None else out.astype(dtype)

    def __len__(self):
        return len(self._frames)


class HostRing:
    """The reference's host ring: list of (obs, next_obs) + numpy scalars.

    add:      src/utils.py:216-227      _get_idxs: src/utils.py:229-232
    gather:   src/utils.py:234-240      sample:    src/utils.py:206-214, 272-280
    """

    def __init__(self, obs_shape, action_shape, capacity, batch_size):
        self.capacity, self.batch_size = capacity, batch_size
        self._obses = []
        self.actions = np.empty((capacity, *action_shape), dtype=np.float32)
        self.rewards = np.empty((capacity, 1), dtype=np.float32)
        self.not_dones = np.empty((capacity, 1), dtype=np.float32)
        self.idx, self.full = 0, False

    def add(self, obs, action, reward, next_obs, done):
        if self.idx >= len(self._obses):
            self._obses.append((obs, next_obs))
        else:
            self._obses[self.idx] = (obs, next_obs)
        np.copyto(self.actions[self.idx], action)
        np.copyto(self.rewards[self.idx], reward)
        np.copyto(self.not_dones[self.idx], not done)
        self.idx = (self.idx + 1) % self.capacity
        self.full = self.full or self.idx == 0

    def get_idxs(self, n=None):
        n = self.batch_size if n is None else n
        return np.random.randint(0, self.capacity if self.full else self.idx, size=n)

    def encode_obses(self, idxs):
        obses, nexts = [], []
        for i in idxs:
            o, no = self._obses[i]
            obses.append(np.asarray(o))
            nexts.append(np.asarray(no))
        return np.array(obses), np.array(nexts)

    def sample(self, idxs):
        """-> (obs, actions, rewards, next_obs, not_dones), f32 CPU tensors, obs in 0..255."""
        o, no = self.encode_obses(idxs)
        return (torch.as_tensor(o).float(), torch.as_tensor(self.actions[idxs]),
                torch.as_tensor(self.rewards[idxs]), torch.as_tensor(no).float(),
                torch.as_tensor(self.not_dones[idxs]))


def gather_u8(store_u8: np.ndarray, idxs) -> torch.Tensor:
    """Dense-array form of encode_obses + ``.float()``: store (cap,C,H,W) u8 -> (B,C,H,W) f32."""
    return torch.from_numpy(store_u8[np.asarray(idxs)]).float()


# ----------------------------------------------------------------------------
# random_shift  (src/augmentations.py:378-382)
# ----------------------------------------------------------------------------
def random_shift_int(x: torch.Tensor, offsets, pad: int = 4) -> torch.Tensor:
    """Replicate-pad by ``pad`` then crop (h,w) at offsets[i]=(row,col) in [0,2*pad]."""
    n, c, h, w = x.shape
    xp = F.pad(x, (pad, pad, pad, pad), mode="replicate")
    off = np.asarray(offsets)
    return torch.stack([xp[i, :, off[i, 0]:off[i, 0] + h, off[i, 1]:off[i, 1] + w] for i in range(n)])


def drq_base_grid(h: int, pad: int) -> torch.Tensor:
    """DrQ-v2 ``arange`` row: linspace(-1+eps, 1-eps, h+2p)[:h], eps=1/(h+2p), fp32."""
    eps = 1.0 / (h + 2 * pad)
    return torch.linspace(-1.0 + eps, 1.0 - eps, h + 2 * pad, dtype=torch.float32)[:h]


def random_shift_bilinear(x: torch.Tensor, shifts, pad: int = 4) -> torch.Tensor:
    """DrQ-v2 RandomShiftsAug (the "DrQ grid_sample semantics" of the north star).

    shifts[i] = (sx, sy) integers in [0, 2*pad] — NOTE grid order: column shift first,
    as in ``grid = base_grid + shift`` where grid[..., 0] is x.  Square images only."""
    n, c, h, w = x.shape
    assert h == w
    xp = F.pad(x, (pad, pad, pad, pad), mode="replicate")
    arange = drq_base_grid(h, pad).unsqueeze(0).repeat(h, 1).unsqueeze(2)
    base_grid = torch.cat([arange, arange.transpose(1, 0)], dim=2).unsqueeze(0).repeat(n, 1, 1, 1)
    shift = torch.as_tensor(np.asarray(shifts)).to(torch.float32).reshape(n, 1, 1, 2).clone()
    shift *= 2.0 / (h + 2 * pad)
    grid = base_grid + shift
    return F.grid_sample(xp, grid, padding_mode="zeros", align_corners=False)


# ----------------------------------------------------------------------------
# random_crop  (src/augmentations.py:385-433)
# ----------------------------------------------------------------------------
def random_crop(x: torch.Tensor, size: int, w1, h1) -> torch.Tensor:
    """out[i] = x[i, :, w1_i:w1_i+size, h1_i:h1_i+size]  (w1 = ROW offset, fact 6)."""
    n = x.shape[0]
    if x.shape[-1] - size <= 0:
        return x
    w1, h1 = np.asarray(w1), np.asarray(h1)
    return torch.stack([x[i, :, w1[i]:w1[i] + size, h1[i]:h1[i] + size] for i in range(n)]).contiguous()


# ----------------------------------------------------------------------------
# random_conv  (src/augmentations.py:265-274)
# ----------------------------------------------------------------------------
def random_conv(x: torch.Tensor, weights: torch.Tensor) -> torch.Tensor:
    """weights (n,3,3,3,3) = per-sample (out,in,kh,kw); same weights for the k frames."""
    n, c, h, w = x.shape
    outs = []
    for i in range(n):
        t = x[i:i + 1].reshape(-1, 3, h, w) / 255.0
        t = F.pad(t, pad=[1] * 4, mode="replicate")
        outs.append(torch.sigmoid(F.conv2d(t, weights[i])) * 255.0)
    return torch.cat(outs, 0).reshape(n, c, h, w)


# ----------------------------------------------------------------------------
# random_overlay  (src/augmentations.py:253-262)
# ----------------------------------------------------------------------------
def random_overlay(x: torch.Tensor, imgs: torch.Tensor, alpha: float = 0.5) -> torch.Tensor:
    """imgs (n,3,h,w) f32 in [0,1]; tiled over channels (frame f uses the same image)."""
    imgs = imgs.repeat(1, x.size(1) // 3, 1, 1)
    return ((1 - alpha) * (x / 255.0) + (alpha) * imgs) * 255.0


def bank_rows_f32(bank_u8: torch.Tensor, bank_idx) -> torch.Tensor:
    """Background batch as the reference's loader yields it: ToTensor() == u8/255 in f32."""
    return bank_u8[torch.as_tensor(np.asarray(bank_idx)).long()].float() / 255.0


# ----------------------------------------------------------------------------
# splice / CS_splice  (src/augmentations.py:277-284, 295-324) + kornia.color.rgb_to_hsv
# ----------------------------------------------------------------------------
def rgb_to_hsv(x: torch.Tensor, eps: float = HSV_EPS) -> torch.Tensor:
    """kornia 0.5.x rgb_to_hsv restated per pixel: (N,3,H,W) in [0,1] -> [h in [0,2pi), s, v].

    Written as the straight-line per-pixel arithmetic the CUDA kernel performs
    (first-argmax channel selection), not as kornia's tensor program."""
    r, g, b = x[:, 0], x[:, 1], x[:, 2]
    maxc = torch.maximum(torch.maximum(r, g), b)
    minc = torch.minimum(torch.minimum(r, g), b)
    v = maxc
    delta = maxc - minc
    s = delta / (v + eps)
    d = torch.where(delta == 0, torch.ones_like(delta), delta)
    rc, gc, bc = maxc - r, maxc - g, maxc - b
    h_r = bc - gc
    h_g = 2.0 * d + rc - bc
    h_b = 4.0 * d + gc - rc
    hsel = torch.where(r == maxc, h_r, torch.where(g == maxc, h_g, h_b))
    hh = (hsel / d) / 6.0
    hh = torch.where(hh < 0, hh + 1.0, hh)       # == torch.remainder(hh, 1.0) for |hh| < 1 (may round to 1.0, as torch does)
    return torch.stack([TWO_PI_F32.item() * hh, s, v], dim=1)


def hsv_mask(x_rgb, hue_thres, sat_thres, val_thres, eps: float = HSV_EPS) -> torch.Tensor:
    """(N,H,W) bool: all(hsv > thresholds) — src/augmentations.py:277-284."""
    hsv = rgb_to_hsv(x_rgb, eps)
    thr = torch.tensor([hue_thres, sat_thres, val_thres], dtype=torch.float32)
    return (hsv[:, 0] > thr[0]) & (hsv[:, 1] > thr[1]) & (hsv[:, 2] > thr[2])


def splice(x, imgs, hue_thres=0, sat_thres=0, val_thres=0.6) -> torch.Tensor:
    """imgs (n,3,h,w) f32 [0,1]; row j=(sample j//k, frame j%k) takes background j mod n (fact 7)."""
    n, c, h, w = x.shape
    k = c // 3
    overlay = imgs.repeat(k, 1, 1, 1)
    x_rgb = x.reshape(-1, 3, h, w) / 255.0
    keep = hsv_mask(x_rgb, hue_thres, sat_thres, val_thres).unsqueeze(1)
    out = torch.where(keep, x_rgb, overlay)
    return out.reshape(n, c, h, w) * 255.0


def cs_splice(x, imgs, hue_thres=3.5, sat_thres=0, val_thres=0) -> torch.Tensor:
    """Cartpole variant: overlay[mask2]=x[mask2]; overlay[~mask]=x[~mask] (src/augmentations.py:309-324)."""
    n, c, h, w = x.shape
    k = c // 3
    overlay = imgs.repeat(k, 1, 1, 1)
    x_rgb = x.reshape(-1, 3, h, w) / 255.0
    mask = hsv_mask(x_rgb, hue_thres, sat_thres, val_thres).unsqueeze(1)
    mask2 = hsv_mask(x_rgb, 0, sat_thres, 0.6).unsqueeze(1)
    out = torch.where(mask2 | ~mask, x_rgb, overlay)
    return out.reshape(n, c, h, w) * 255.0


# ----------------------------------------------------------------------------
# Device RNG restated: Philox4x32-10 (Salmon et al. 2011), the generator the
# library uses when the caller does not supply offsets / indices / weights.
# (No reference counterpart: the reference draws from torch/numpy global RNGs,
#  src/augmentations.py:269,404-405, src/utils.py:232.)
# ----------------------------------------------------------------------------
_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)


def philox4x32_10(counter: np.ndarray, key) -> np.ndarray:
    """counter (N,4) uint32, key (2,) uint32 -> (N,4) uint32."""
    c = counter.astype(np.uint32).copy()
    k0, k1 = np.uint32(key[0]), np.uint32(key[1])
    mask = np.uint64(0xFFFFFFFF)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = _M0 * c[:, 0].astype(np.uint64)
            p1 = _M1 * c[:, 2].astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), (p0 & mask).astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), (p1 & mask).astype(np.uint32)
            c = np.stack([hi1 ^ c[:, 1] ^ k0, lo1, hi0 ^ c[:, 3] ^ k1, lo0], axis=1)
            k0 = np.uint32(k0 + _W0)
            k1 = np.uint32(k1 + _W1)
    return c


def _philox_words(count: int, seed: int, subsequence: int) -> np.ndarray:
    """First ``count`` 32-bit words of stream (seed, subsequence): block b = counter (b, 0, sub_lo, sub_hi)."""
    nblk = (count + 3) // 4
    ctr = np.zeros((nblk, 4), dtype=np.uint32)
    b = np.arange(nblk, dtype=np.uint64)
    ctr[:, 0] = (b & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    ctr[:, 1] = (b >> np.uint64(32)).astype(np.uint32)
    ctr[:, 2] = np.uint32(subsequence & 0xFFFFFFFF)
    ctr[:, 3] = np.uint32((subsequence >> 32) & 0xFFFFFFFF)
    key = (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    return philox4x32_10(ctr, key).reshape(-1)[:count]


def rng_randint(count: int, lo: int, hi: int, seed: int, subsequence: int) -> np.ndarray:
    """Uniform integers in [lo, hi): lo + ((u32 * (hi-lo)) >> 32) (multiply-shift, no rejection)."""
    u = _philox_words(count, seed, subsequence).astype(np.uint64)
    return (lo + ((u * np.uint64(hi - lo)) >> np.uint64(32)).astype(np.int64))


def rng_normal(count: int, seed: int, subsequence: int) -> np.ndarray:
    """Standard normals via Box-Muller on word pairs (u1,u2): element 2j -> r*cos, 2j+1 -> r*sin."""
    npair = (count + 1) // 2
    wds = _philox_words(2 * npair, seed, subsequence).astype(np.float64)
    u1 = (wds[0::2] + 1.0) * (1.0 / 4294967296.0)   # (0,1]
    u2 = wds[1::2] * (1.0 / 4294967296.0)           # [0,1)
    r = np.sqrt(-2.0 * np.log(u1))
    out = np.empty(2 * npair, dtype=np.float64)
    out[0::2] = r * np.cos(2 * math.pi * u2)
    out[1::2] = r * np.sin(2 * math.pi * u2)
    return out[:count].astype(np.float32)


# ----------------------------------------------------------------------------
# The remaining aug_to_func entries.  Draws are explicit arguments.
# NOTE these reference functions were inherited from RAD, where images are in [0,1]; the
# reference applies them to 0..255 tensors as they are (e.g. grayscale's x*255 -> uint8 wraps
# modulo 256 on the CPU).  The restatements follow the reference's arithmetic literally.
# ----------------------------------------------------------------------------
def random_grayscale(x: torch.Tensor, mask) -> torch.Tensor:
    """src/augmentations.py:88-136.  mask (n,) bool = (np.random.uniform <= p)."""
    n, c, h, w = x.shape
    k = c // 3
    u8 = ((x * 255.0).to(torch.int64) % 256).to(torch.uint8)          # .type(torch.uint8): truncation, wraps mod 256
    v = u8.view(n, k, 3, h, w)
    gray = v[:, :, 0] * 0.2989 + v[:, :, 1] * 0.587 + v[:, :, 2] * 0.114
    gray = gray.type(torch.uint8).float()[:, :, None].expand(n, k, 3, h, w)
    m = torch.as_tensor(np.asarray(mask)).bool().view(n, 1, 1, 1, 1)
    out = torch.where(m, gray, v.float())
    return out.reshape(n, c, h, w) / 255.0


def random_flip(x: torch.Tensor, mask) -> torch.Tensor:
    """src/augmentations.py:139-168.  Horizontal flip of the samples where mask is set."""
    m = torch.as_tensor(np.asarray(mask)).bool().view(-1, 1, 1, 1)
    return torch.where(m, x.flip([3]), x)


def random_rotation(x: torch.Tensor, rot) -> torch.Tensor:
    """src/augmentations.py:171-213.  rot (n,) in {0,1,2,3} = rnd_rot * (rnd <= p): quarter turns."""
    rot = np.asarray(rot)
    return torch.stack([torch.rot90(x[i], int(rot[i]), [1, 2]) for i in range(x.shape[0])])


def random_cutout_color(x: torch.Tensor, w1, h1, colors) -> torch.Tensor:
    """src/augmentations.py:327-350.  Box rows h1:2*h1, cols w1:2*w1 (start == size, a reference quirk)
    filled per channel with colors (n,c) = randint(0,255)/255.0 (float64, cast to the image dtype)."""
    out = x.clone()
    w1, h1 = np.asarray(w1), np.asarray(h1)
    col = torch.as_tensor(np.asarray(colors, dtype=np.float64)).to(x.dtype)
    for i in range(x.shape[0]):
        out[i, :, h1[i]:2 * h1[i], w1[i]:2 * w1[i]] = col[i].view(-1, 1, 1)
    return out
