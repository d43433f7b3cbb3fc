"""Drop-in for the reference's ``augmentations`` module (/root/reference/src/augmentations.py).

Same function names, positional arguments and error behaviour as the reference, so
``src/algorithms/*`` and ``RAD.apply_aug`` (src/algorithms/rad.py:9-22) call it unchanged:
tensors in, tensors out, CUDA f32 NCHW in the 0..255 range, inputs never mutated.
Every function enqueues ONE hand-written sm_100a kernel through the C ABI
(include/augcl_b200.h) on torch's current stream.  There is no CPU path.

Random draws.  The reference draws on the host (kornia crop parameters :382,
``LongTensor.random_`` :404-405, ``torch.randn`` :269, the DataLoader's shuffled batch :76-86).
Here every draw can be SUPPLIED through a keyword-only argument (``offsets=``, ``weights=``,
``bank_idx=``) — that is how parity "given identical draws" is tested — and otherwise comes from
the device Philox stream (seed with :func:`seed`; stream = (seed, call counter)).

Backgrounds.  Places365 / COCO are not available offline; ``random_overlay`` / ``splice`` read a
device-resident background bank ([M,3,h,w] u8 or f32-in-[0,1]) registered with
:func:`set_background_bank`.  ``load_dataloader`` keeps the reference's name and dataset check
(:240-250) and, when no bank has been registered, creates a synthetic one.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib as L

__all__ = ["random_shift", "random_crop", "random_conv", "random_overlay", "splice", "CS_splice",
           "random_grayscale", "random_flip", "random_rotation", "random_cutout_color",
           "identity", "aug_to_func", "seed", "set_background_bank", "load_dataloader", "create_hsv_mask",
           "view_as_windows_cuda", "color_jitter", "batch_from_obs", "prepare_pad_batch"]

# --------------------------------------------------------------------------------------------
# process-global state (the reference keeps `dataloader` / `data_iter` globals, :11-12)
# --------------------------------------------------------------------------------------------
_bank = None            # device tensor [M,3,h,w], uint8 or float32 in [0,1]
_seed = 0
_calls = 0              # Philox subsequence of the next draw
_base_grids = {}        # (h, pad, device) -> DrQ-v2 base grid on the device
SYNTHETIC_BANK_LEN = 4096


def seed(s: int) -> None:
    """Seed the device draws of this module (the reference uses torch/numpy global RNGs)."""
    global _seed, _calls
    _seed, _calls = int(s) & (2**64 - 1), 0


def _next_subsequence() -> int:
    global _calls
    _calls += 1
    return _calls - 1


def _reserve_subsequences(k: int) -> int:
    """First of k consecutive Philox subsequences (the fused sample kernel uses 3: indices, obs and
    next_obs offsets)."""
    global _calls
    _calls += k
    return _calls - k


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _prep(x: torch.Tensor, who: str) -> torch.Tensor:
    if not isinstance(x, torch.Tensor) or not x.is_cuda:
        raise AssertionError(f"{who}: input must be a CUDA tensor (augcl_b200 has no CPU path)")
    if x.dim() != 4:
        raise ValueError(f"{who}: expected (n,c,h,w), got {tuple(x.shape)}")
    if x.dtype != torch.float32:
        x = x.float()
    return x.contiguous()


def _dev_i32(v, n: int, device, who: str) -> torch.Tensor:
    t = torch.as_tensor(v)
    if t.shape != (n, 2):
        raise ValueError(f"{who}: offsets must have shape ({n}, 2), got {tuple(t.shape)}")
    return t.to(device=device, dtype=torch.int32, non_blocking=True).contiguous()


def _draw_offsets(n: int, lo: int, hi: int, device) -> torch.Tensor:
    out = torch.empty((n, 2), dtype=torch.int32, device=device)
    L.call("augcl_rng_fill_offsets", out.data_ptr(), 2 * n, lo, hi, _seed, _next_subsequence(), _stream(out))
    return out


def identity(x):
    """src/augmentations.py:374."""
    return x


# --------------------------------------------------------------------------------------------
# random_shift  (src/augmentations.py:378-382)
# --------------------------------------------------------------------------------------------
def random_shift(imgs, pad=4, *, offsets=None, mode="int"):
    """Replicate-pad by ``pad`` and crop back to (h,w) at a per-sample random origin.

    offsets: optional (n,2) integers (row, col) in [0, 2*pad] (the kornia RandomCrop draw).
    mode "int" = exact integer crop (bit-exact); "bilinear" = the DrQ-v2 grid_sample formulation.
    """
    x = _prep(imgs, "random_shift")
    n, c, h, w = x.shape
    with torch.cuda.device(x.device):
        off = _draw_offsets(n, 0, 2 * pad + 1, x.device) if offsets is None else _dev_i32(offsets, n, x.device, "random_shift")
        out = torch.empty_like(x)
        if mode == "int":
            L.call("augcl_random_shift", x.data_ptr(), out.data_ptr(), n, c, h, w, int(pad), off.data_ptr(),
                   L.SHIFT_INT, None, _stream(x))
        elif mode == "bilinear":
            L.call("augcl_random_shift", x.data_ptr(), out.data_ptr(), n, c, h, w, int(pad), off.data_ptr(),
                   L.SHIFT_BILINEAR, _base_grid(h, int(pad), x.device).data_ptr(), _stream(x))
        else:
            raise ValueError(f"random_shift: unknown mode {mode!r}")
    return out


def _base_grid(h: int, pad: int, device) -> torch.Tensor:
    key = (h, pad, str(device))
    g = _base_grids.get(key)
    if g is None:
        buf = (C.c_float * h)()
        L.call("augcl_drq_base_grid", h, pad, buf)
        g = torch.tensor(list(buf), dtype=torch.float32).to(device)
        _base_grids[key] = g
    return g


# --------------------------------------------------------------------------------------------
# random_crop  (src/augmentations.py:385-413)
# --------------------------------------------------------------------------------------------
def random_crop(x, size=84, w1=None, h1=None, return_w1_h1=False):
    """Vectorized CUDA implementation of random crop; ``w1`` indexes rows (reference quirk)."""
    assert (w1 is None and h1 is None) or (w1 is not None and h1 is not None), \
        "must either specify both w1 and h1 or neither of them"
    assert isinstance(x, torch.Tensor) and x.is_cuda, "input must be cuda tensor"
    xin = x
    x = _prep(x, "random_crop")
    n, c, H, W = x.shape
    crop_max = W - size
    if crop_max <= 0:                      # :396-399 passthrough
        return (xin, None, None) if return_w1_h1 else xin
    with torch.cuda.device(x.device):
        if w1 is None:
            off = _draw_offsets(n, 0, crop_max, x.device)       # random_(0, crop_max): upper bound exclusive
        else:
            w1t, h1t = torch.as_tensor(w1).reshape(n), torch.as_tensor(h1).reshape(n)
            for name, t in (("w1", w1t), ("h1", h1t)):
                # the reference's windows[arange, w1, h1] raises IndexError outside [0, crop_max); host draws are checked
                # here, device draws are trusted (checking them would cost a device->host sync; the kernel clamps)
                if not t.is_cuda and t.numel() and (int(t.min()) < -crop_max or int(t.max()) >= crop_max):
                    raise IndexError(f"random_crop: {name} outside [0, {crop_max}) (window positions of view_as_windows_cuda)")
            if not w1t.is_cuda:     # negative host indices wrap like torch advanced indexing does
                w1t, h1t = w1t % crop_max, h1t % crop_max
            off = torch.stack([w1t, h1t], dim=1)
            off = off.to(device=x.device, dtype=torch.int32).contiguous()
        out = torch.empty((n, c, size, size), dtype=torch.float32, device=x.device)
        L.call("augcl_random_crop", x.data_ptr(), out.data_ptr(), n, c, H, W, int(size), off.data_ptr(), _stream(x))
        if xin.dtype != torch.float32:       # the reference's window view keeps the input dtype (uint8 from prepare_pad_batch)
            out = out.to(xin.dtype)
    if return_w1_h1:
        return out, off[:, 0].long(), off[:, 1].long()
    return out


# --------------------------------------------------------------------------------------------
# random_conv  (src/augmentations.py:265-274)
# --------------------------------------------------------------------------------------------
def random_conv(x, *, weights=None):
    """Per-sample random 3x3 conv (3->3 channels) on replicate-padded x/255, sigmoid, *255.

    weights: optional (n,3,3,3,3) f32 = the reference's per-sample ``torch.randn(3,3,3,3)``."""
    x = _prep(x, "random_conv")
    n, c, h, w = x.shape
    with torch.cuda.device(x.device):
        if weights is None:
            wt = torch.empty((n, 3, 3, 3, 3), dtype=torch.float32, device=x.device)
            L.call("augcl_rng_fill_normal", wt.data_ptr(), wt.numel(), _seed, _next_subsequence(), _stream(x))
        else:
            wt = torch.as_tensor(weights).to(device=x.device, dtype=torch.float32).contiguous()
            if wt.shape != (n, 3, 3, 3, 3):
                raise ValueError(f"random_conv: weights must be ({n},3,3,3,3), got {tuple(wt.shape)}")
        out = torch.empty_like(x)
        L.call("augcl_random_conv", x.data_ptr(), out.data_ptr(), n, c, h, w, wt.data_ptr(), _stream(x))
    return out


# --------------------------------------------------------------------------------------------
# background bank (stands in for the Places365 / COCO DataLoader, src/augmentations.py:19-86)
# --------------------------------------------------------------------------------------------
def set_background_bank(bank) -> None:
    """Register the device-resident background bank: [M,3,h,w] uint8, or float32 in [0,1]."""
    global _bank
    if bank is None:
        _bank = None
        return
    if not (isinstance(bank, torch.Tensor) and bank.is_cuda and bank.dim() == 4 and bank.shape[1] == 3):
        raise ValueError("background bank must be a CUDA tensor of shape [M,3,h,w]")
    if bank.dtype not in (torch.uint8, torch.float32):
        raise ValueError("background bank must be uint8 or float32")
    _bank = bank.contiguous()


def get_background_bank():
    return _bank


def load_dataloader(batch_size, image_size, dataset="coco", *, device=None):
    """Reference name and dataset check (src/augmentations.py:240-250).  The image folders are not
    available offline: when no bank is registered, a synthetic u8 bank of SYNTHETIC_BANK_LEN rows is
    generated on the device (Philox bytes)."""
    global _bank
    if dataset not in ("places365_standard", "coco"):
        raise NotImplementedError(f'overlay has not been implemented for dataset "{dataset}"')
    if _bank is None:
        device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        bank = torch.empty((SYNTHETIC_BANK_LEN, 3, image_size, image_size), dtype=torch.uint8, device=device)
        with torch.cuda.device(device):
            L.call("augcl_rng_fill_bytes", bank.data_ptr(), bank.numel(), _seed ^ 0xBA4C, 0, _stream(bank))
        _bank = bank


def _bank_for(x: torch.Tensor, who: str, dataset: str):
    n, c, h, w = x.shape
    load_dataloader(batch_size=n, image_size=h, dataset=dataset, device=x.device)
    b = _bank
    if b.device != x.device:
        raise ValueError(f"{who}: background bank lives on {b.device}, input on {x.device}")
    if tuple(b.shape[2:]) != (h, w):
        raise ValueError(f"{who}: background bank is {tuple(b.shape[2:])}, input images are {(h, w)}")
    return b, (L.BANK_U8 if b.dtype == torch.uint8 else L.BANK_F32)


def _bank_rows(n: int, bank_len: int, bank_idx, device) -> torch.Tensor:
    if bank_idx is None:
        out = torch.empty(n, dtype=torch.int64, device=device)
        L.call("augcl_rng_fill_indices", out.data_ptr(), n, 0, bank_len, _seed, _next_subsequence(), _stream(out))
        return out
    t = torch.as_tensor(bank_idx).reshape(-1).to(device=device, dtype=torch.int64).contiguous()
    if t.numel() != n:
        raise ValueError(f"bank_idx must have {n} entries, got {t.numel()}")
    return t


# --------------------------------------------------------------------------------------------
# random_overlay  (src/augmentations.py:253-262)
# --------------------------------------------------------------------------------------------
def random_overlay(x, dataset="coco", *, bank_idx=None, alpha=0.5):
    """((1-alpha)*(x/255) + alpha*img)*255 with img = bank[bank_idx[i]], same image for every frame."""
    x = _prep(x, "random_overlay")
    n, c, h, w = x.shape
    with torch.cuda.device(x.device):
        bank, bdt = _bank_for(x, "random_overlay", dataset)
        rows = _bank_rows(n, bank.shape[0], bank_idx, x.device)
        out = torch.empty_like(x)
        L.call("augcl_random_overlay", x.data_ptr(), out.data_ptr(), n, c, h, w, bank.data_ptr(), bdt,
               bank.shape[0], rows.data_ptr(), float(alpha), _stream(x))
    return out


# --------------------------------------------------------------------------------------------
# splice / CS_splice  (src/augmentations.py:295-306, 309-324; mask :277-284)
# --------------------------------------------------------------------------------------------
def _splice(x, thr, variant, who, bank_idx, dataset):
    x = _prep(x, who)
    n, c, h, w = x.shape
    with torch.cuda.device(x.device):
        bank, bdt = _bank_for(x, who, dataset)
        rows = _bank_rows(n, bank.shape[0], bank_idx, x.device)
        out = torch.empty_like(x)
        L.call("augcl_splice", x.data_ptr(), out.data_ptr(), n, c, h, w, bank.data_ptr(), bdt, bank.shape[0],
               rows.data_ptr(), float(thr[0]), float(thr[1]), float(thr[2]), variant, _stream(x))
    return out


def create_hsv_mask(x_rgb, hue_thres, sat_thres, val_thres):
    """src/augmentations.py:277-284: boolean mask (n,3,h,w) of the pixels of an RGB batch IN [0,1] whose kornia-HSV hue,
    saturation and value all exceed the thresholds (splice evaluates the same test inside its kernel)."""
    x = _prep(x_rgb, "create_hsv_mask")
    n, c, h, w = x.shape
    if c != 3:
        raise ValueError(f"create_hsv_mask: expected (n,3,h,w), got {tuple(x.shape)}")
    with torch.cuda.device(x.device):
        m = torch.empty((n, h, w), dtype=torch.uint8, device=x.device)
        L.call("augcl_hsv_mask", x.data_ptr(), m.data_ptr(), n, h, w, float(hue_thres), float(sat_thres), float(val_thres),
               _stream(x))
    return m.bool().unsqueeze(1).repeat(1, 3, 1, 1)


def view_as_windows_cuda(x, window_shape):
    """src/augmentations.py:416-433: strided view of every window (pure metadata; random_crop itself is a kernel)."""
    assert isinstance(window_shape, tuple) and len(window_shape) == len(x.shape), \
        "window_shape must be a tuple with same number of dimensions as x"
    win_indices_shape = [x.size(0), x.size(1) - int(window_shape[1]), x.size(2) - int(window_shape[2]), x.size(3)]
    return x.as_strided(tuple(win_indices_shape + list(window_shape)), tuple(list(x.stride()) + list(x.stride())))


def splice(x, hue_thres=0, sat_thres=0, val_thres=0.6, *, bank_idx=None, dataset="coco"):
    """Keep pixels whose HSV exceeds the thresholds, replace the rest by the background;
    frame f of sample s takes background row (s*k+f) mod n (the reference's batch tiling)."""
    return _splice(x, (hue_thres, sat_thres, val_thres), L.SPLICE, "splice", bank_idx, dataset)


def CS_splice(x, hue_thres=3.5, sat_thres=0, val_thres=0, *, bank_idx=None, dataset="coco"):
    return _splice(x, (hue_thres, sat_thres, val_thres), L.CS_SPLICE, "CS_splice", bank_idx, dataset)


# --------------------------------------------------------------------------------------------
# the remaining aug_to_func entries (src/augmentations.py:111-213, 327-350).  Their per-sample decisions are
# drawn with numpy's global RNG in the reference's own order, so a seeded run makes the same decisions;
# every draw can also be supplied.
# --------------------------------------------------------------------------------------------
def _dev(a, dtype, device):
    return torch.as_tensor(np.ascontiguousarray(a)).to(device=device, dtype=dtype).contiguous()


def random_grayscale(images, p=0.3, *, mask=None):
    """src/augmentations.py:111-136.  mask: optional (n,) bools = the reference's ``rnd <= p``."""
    x = _prep(images, "random_grayscale")
    n, c, h, w = x.shape
    if mask is None:
        mask = np.random.uniform(0.0, 1.0, size=(n,)) <= p
    with torch.cuda.device(x.device):
        m = _dev(np.asarray(mask).astype(np.uint8), torch.uint8, x.device)
        out = torch.empty_like(x)
        L.call("augcl_random_grayscale", x.data_ptr(), out.data_ptr(), n, c, h, w, m.data_ptr(), _stream(x))
    return out


def random_flip(images, p=0.2, *, mask=None):
    """src/augmentations.py:139-168."""
    x = _prep(images, "random_flip")
    n, c, h, w = x.shape
    if mask is None:
        mask = np.random.uniform(0.0, 1.0, size=(n,)) <= p
    with torch.cuda.device(x.device):
        m = _dev(np.asarray(mask).astype(np.int32), torch.int32, x.device)
        out = torch.empty_like(x)
        L.call("augcl_random_flip", x.data_ptr(), out.data_ptr(), n, c, h, w, m.data_ptr(), _stream(x))
    return out


def random_rotation(images, p=0.3, *, rot=None):
    """src/augmentations.py:171-213.  rot: optional (n,) quarter turns in {0,1,2,3}."""
    x = _prep(images, "random_rotation")
    n, c, h, w = x.shape
    if rot is None:
        rnd = np.random.uniform(0.0, 1.0, size=(n,))
        rnd_rot = np.random.randint(1, 4, size=(n,))
        rot = rnd_rot * (rnd <= p)
    with torch.cuda.device(x.device):
        r = _dev(np.asarray(rot).astype(np.int32), torch.int32, x.device)
        out = torch.empty_like(x)
        L.call("augcl_random_rotation", x.data_ptr(), out.data_ptr(), n, c, h, w, r.data_ptr(), _stream(x))
    return out


def random_cutout_color(imgs, min_cut=10, max_cut=30, *, w1=None, h1=None, colors=None):
    """src/augmentations.py:327-350 (the reference round-trips through numpy and returns a CPU tensor; this
    stays on the device).  colors: optional (n,c) floats = the reference's randint(0,255)/255.0."""
    x = _prep(imgs, "random_cutout_color")
    n, c, h, w = x.shape
    if w1 is None:
        w1 = np.random.randint(min_cut, max_cut, n)
    if h1 is None:
        h1 = np.random.randint(min_cut, max_cut, n)
    if colors is None:
        colors = np.random.randint(0, 255, size=(n, c)) / 255.0
    box = np.stack([np.asarray(w1).reshape(n), np.asarray(h1).reshape(n)], axis=1).astype(np.int32)
    with torch.cuda.device(x.device):
        b = _dev(box, torch.int32, x.device)
        col = _dev(np.asarray(colors, dtype=np.float64).reshape(n, c), torch.float32, x.device)
        out = torch.empty_like(x)
        L.call("augcl_random_cutout_color", x.data_ptr(), out.data_ptr(), n, c, h, w, b.data_ptr(), col.data_ptr(),
               _stream(x))
    return out


def color_jitter(imgs, p=1, *, sampled_idxs=None, params=None):
    """src/augmentations.py:225-237: torchvision ``ColorJitter(0, (0,2), (0,3), (-0.5,0.5))`` on a random subset of
    the frames, IN PLACE like the reference (one kernel: augcl_color_jitter, the torchvision op sequence restated).
    The draws are the reference's: ``np.random.randint(0, b, p*b*(c/3))`` frame indices (its quirk: only the first b of
    the b*c/3 frames are candidates) and ``ColorJitter.get_params`` (torch CPU RNG: randperm(4), then the contrast,
    saturation and hue factors).  ``sampled_idxs`` / ``params=(fn_idx, brightness, contrast, saturation, hue)``
    supply them explicitly; a ``None`` factor turns its op off, as in torchvision."""
    if not (isinstance(imgs, torch.Tensor) and imgs.is_cuda):
        raise AssertionError("color_jitter: input must be a CUDA tensor (augcl_b200 has no CPU path)")
    if imgs.dtype != torch.float32 or not imgs.is_contiguous():
        raise ValueError("color_jitter: expected a contiguous float32 tensor (the op is in place)")
    b, c, h, w = imgs.shape
    num_samples = int(p * b * int(c / 3))
    if sampled_idxs is None:
        sampled_idxs = np.random.randint(0, b, num_samples)
    if params is None:
        # torchvision ColorJitter.get_params(None, [0,2], [0,3], [-0.5,0.5]) restated: same calls on torch's CPU RNG, in order
        fn_idx = torch.randperm(4)
        cf = float(torch.empty(1).uniform_(0.0, 2.0))
        sf = float(torch.empty(1).uniform_(0.0, 3.0))
        hf = float(torch.empty(1).uniform_(-0.5, 0.5))
        params = (fn_idx, None, cf, sf, hf)
    fn_idx, bf, cf, sf, hf = params
    sel_np = np.unique(np.asarray(sampled_idxs, dtype=np.int64))
    if sel_np.size and (sel_np[0] < 0 or sel_np[-1] >= b * (c // 3)):
        raise IndexError("color_jitter: frame index out of range")
    sel = torch.from_numpy(sel_np.astype(np.int32)).to(imgs.device)
    order = (C.c_int32 * 4)(*[int(v) for v in fn_idx])
    enabled = sum(1 << i for i, f in enumerate((bf, cf, sf, hf)) if f is not None)
    L.call("augcl_color_jitter", imgs.data_ptr(), b * (c // 3), h, w, sel.data_ptr(), int(sel_np.size), order,
           float(bf or 0.0), float(cf or 0.0), float(sf or 0.0), float(hf or 0.0), enabled, _stream(imgs))
    return imgs.view(-1, 3, h, w).view(b, c, h, w)


def batch_from_obs(obs, batch_size=32):
    """src/augmentations.py:353-362: one observation repeated along a new batch axis (tensor in, tensor out; array in,
    array out)."""
    if isinstance(obs, torch.Tensor):
        return (obs.unsqueeze(0) if obs.dim() == 3 else obs).repeat(batch_size, 1, 1, 1)
    obs = np.asarray(obs)
    return np.repeat(obs[None] if obs.ndim == 3 else obs, repeats=batch_size, axis=0)


def prepare_pad_batch(obs, next_obs, action, batch_size=32):
    """src/augmentations.py:365-371 (PAD's test-time batch, called from src/eval.py:36): the observation pair repeated
    batch_size times on the device, each copy randomly cropped by the random_crop kernel."""
    dev = torch.device("cuda", torch.cuda.current_device())
    batch_obs = batch_from_obs(torch.from_numpy(np.asarray(obs)).to(dev), batch_size)
    batch_next_obs = batch_from_obs(torch.from_numpy(np.asarray(next_obs)).to(dev), batch_size)
    batch_action = torch.from_numpy(np.asarray(action)).to(dev).unsqueeze(0).repeat(batch_size, 1)
    return random_crop(batch_obs), random_crop(batch_next_obs), batch_action


# same 12 keys as src/augmentations.py:436-449
aug_to_func = {
    "grayscale": random_grayscale,
    "flip": random_flip,
    "rotate": random_rotation,
    "conv": random_conv,
    "shift": random_shift,
    "color_jitter": color_jitter,
    "identity": identity,
    "splice": splice,
    "crop": random_crop,
    "cutout_color": random_cutout_color,
    "CS_splice": CS_splice,
    "overlay": random_overlay,
}
