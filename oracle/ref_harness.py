"""TEST INFRASTRUCTURE ONLY — never imported by the product package.

Runs the reference's own Python code (GT-RIPL/AugCL, mounted read-only at
/root/reference) on CPU so that (1) the restatement in ``augcl_oracle.py`` can
be pinned against it and (2) golden vectors can be generated for
``tests/golden/`` (``make_golden.py``).  /root/reference does not exist on the GPU
box; there the harness falls back to ``oracle/_ref/`` — the same modules compiled to
bytecode by the committed recipe ``oracle/build_ref.py`` (a built artefact that ships
with the snapshot like the ``.so``; nothing reads /root/reference at run time).

The reference does not import in this container as-is:
  * ``src/augmentations.py:6`` imports kornia (absent, unpinned) -> stub module;
  * ``src/utils.py:238-239`` uses ``np.array(copy=False)`` (raises on numpy 2);
  * ``.cuda()`` / ``is_cuda`` / ``get_device()`` are hard-coded
    (``src/utils.py:208-212``, ``src/augmentations.py:86,280-282,390``).
The shims below neutralise exactly those three things and nothing else; all
arithmetic executed is the reference's.  Determinism is injected by replacing
the random draws (kornia RandomCrop parameters, ``torch.randn`` in
``random_conv``, the background batch) with supplied tensors.
"""
from __future__ import annotations

import contextlib
import importlib.abc
import importlib.machinery
import importlib.util
import math
import os
import sys
import types

import numpy as np
import torch

REF_LIVE = "/root/reference/src"                                             # build container only
REF_BUILT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")  # oracle/build_ref.py output (ships to the GPU box)


def ref_dir():
    """Directory the reference's modules are imported from: the live tree when it is mounted, else the compiled copy."""
    if os.path.isfile(os.path.join(REF_LIVE, "augmentations.py")):
        return REF_LIVE
    if os.path.isfile(os.path.join(REF_BUILT, "augmentations.bin")):
        return REF_BUILT
    return None


class RefFinder(importlib.abc.MetaPathFinder):
    """Imports the reference's modules from oracle/_ref: ``<root>/<pkg>/<mod>.bin`` holds the bytecode
    oracle/build_ref.py compiled from /root/reference/src/<pkg>/<mod>.py; directories are namespace packages (the
    reference's ``algorithms`` has no __init__.py either)."""

    def __init__(self, root=REF_BUILT):
        self.root = root

    def find_spec(self, fullname, path=None, target=None):
        rel = os.path.join(self.root, *fullname.split("."))
        if os.path.isfile(rel + ".bin"):
            loader = importlib.machinery.SourcelessFileLoader(fullname, rel + ".bin")
            return importlib.util.spec_from_loader(fullname, loader, origin=rel + ".bin")
        if os.path.isdir(rel):
            spec = importlib.machinery.ModuleSpec(fullname, None, is_package=True)
            spec.submodule_search_locations = [rel]
            return spec
        return None


def install_ref_importer(front=False):
    """Make the reference's modules importable by name on a machine without /root/reference.  front=False: only names
    nothing else provides (``algorithms.*``; ``utils`` / ``augmentations`` when no shim shadows them)."""
    for f in sys.meta_path:
        if isinstance(f, RefFinder):
            return f
    f = RefFinder()
    if front:
        sys.meta_path.insert(0, f)
    else:
        sys.meta_path.append(f)
    return f


def remove_ref_importer():
    sys.meta_path[:] = [f for f in sys.meta_path if not isinstance(f, RefFinder)]


REF_SRC = ref_dir() or REF_LIVE


def available() -> bool:
    return ref_dir() is not None


# --------------------------------------------------------------------------
# kornia stand-in.  kornia is a third-party dependency absent from
# /root/reference (setup/conda.yaml:22, unpinned; the contemporaries torch==1.7.1 /
# numpy==1.19.5 resolve to a 0.5.x release).  The two call sites are restated
# from the published algorithm.  "parity unpinned":
# the reference holds no test vectors at this boundary.
# --------------------------------------------------------------------------
HSV_EPS = 1e-6  # kornia 0.5.x rgb_to_hsv default


def kornia_rgb_to_hsv(image: torch.Tensor, eps: float = HSV_EPS) -> torch.Tensor:
    maxc, _ = image.max(-3)
    maxc_mask = image == maxc.unsqueeze(-3)
    _, max_indices = ((maxc_mask.cumsum(-3) == 1) & maxc_mask).max(-3)
    minc = image.min(-3)[0]
    v = maxc
    deltac = maxc - minc
    s = deltac / (v + eps)
    deltac = torch.where(deltac == 0, torch.ones_like(deltac), deltac)
    maxc_tmp = maxc.unsqueeze(-3) - image
    rc, gc, bc = maxc_tmp[..., 0, :, :], maxc_tmp[..., 1, :, :], maxc_tmp[..., 2, :, :]
    h = torch.stack([bc - gc, 2.0 * deltac + rc - bc, 4.0 * deltac + gc - rc], dim=-3)
    h = torch.gather(h, dim=-3, index=max_indices[..., None, :, :]).squeeze(-3)
    h = h / deltac
    h = (h / 6.0) % 1.0
    h = 2 * math.pi * h
    return torch.stack([h, s, v], dim=-3)


class _RandomCrop:
    """kornia.augmentation.RandomCrop((h, w)) with injected parameters.

    ``_next_offsets`` (n,2) int (row, col) must be set by the harness before the
    call; mode "slice" (kornia>=0.6 default) or "warp" (0.5.x: bilinear
    warp_affine, align_corners=True)."""

    _next_offsets = None
    _mode = "slice"

    def __init__(self, size):
        self.size = size

    def __call__(self, x):
        h, w = self.size
        off = type(self)._next_offsets
        assert off is not None, "harness must inject crop offsets"
        n = x.shape[0]
        if type(self)._mode == "slice":
            return torch.stack(
                [x[i, :, int(off[i, 0]):int(off[i, 0]) + h, int(off[i, 1]):int(off[i, 1]) + w] for i in range(n)]
            )
        # warp mode: translation-only affine sampled with bilinear/zeros/align_corners=True
        H, W = x.shape[-2:]
        ys = torch.arange(h, dtype=torch.float32)
        xs = torch.arange(w, dtype=torch.float32)
        out = []
        for i in range(n):
            sy = ys + float(off[i, 0])
            sx = xs + float(off[i, 1])
            gy = sy * 2.0 / (H - 1) - 1.0
            gx = sx * 2.0 / (W - 1) - 1.0
            grid = torch.stack(torch.meshgrid(gx, gy, indexing="xy"), dim=-1)[None]
            out.append(torch.nn.functional.grid_sample(
                x[i:i + 1], grid, mode="bilinear", padding_mode="zeros", align_corners=True))
        return torch.cat(out)


def _install_kornia_stub():
    if "kornia" in sys.modules and not getattr(sys.modules["kornia"], "_augcl_stub", False):
        return
    k = types.ModuleType("kornia")
    k._augcl_stub = True
    k.color = types.ModuleType("kornia.color")
    k.color.rgb_to_hsv = kornia_rgb_to_hsv
    k.augmentation = types.ModuleType("kornia.augmentation")
    k.augmentation.RandomCrop = _RandomCrop
    sys.modules["kornia"] = k
    sys.modules["kornia.color"] = k.color
    sys.modules["kornia.augmentation"] = k.augmentation


_ref_mods = None


def load_reference():
    """Import the reference's ``augmentations`` and ``utils`` modules (cached)."""
    global _ref_mods
    if _ref_mods is not None:
        return _ref_mods
    src = ref_dir()
    if src is None:
        raise RuntimeError("neither /root/reference nor oracle/_ref (python oracle/build_ref.py) is present on this machine")
    _install_kornia_stub()
    saved = {k: sys.modules.pop(k) for k in ("augmentations", "utils") if k in sys.modules}
    built = src == REF_BUILT
    finder = RefFinder() if built else None
    if built:
        sys.meta_path.insert(0, finder)       # ahead of the path finder: the reference's modules, whatever sys.path holds
    else:
        sys.path.insert(0, src)
    try:
        import augmentations as ref_aug  # noqa: E402
        import utils as ref_utils  # noqa: E402
    finally:
        if built:
            sys.meta_path.remove(finder)
        else:
            sys.path.remove(src)
        # keep the reference modules out of the global namespace
        sys.modules.pop("augmentations", None)
        sys.modules.pop("utils", None)
        sys.modules.update(saved)
    _ref_mods = (ref_aug, ref_utils)
    return _ref_mods


@contextlib.contextmanager
def numpy2_shim():
    """``np.array(obj, copy=False)`` (src/utils.py:238-239) raises on numpy >= 2 when a copy is needed; numpy 1.19, which
    the reference pins, copied silently.  Restores that behaviour for the duration of a call."""
    orig_np_array = np.array

    def np_array(obj, *a, copy=True, **kw):
        if copy is False:
            return np.asarray(obj, *a, **kw)
        return orig_np_array(obj, *a, copy=copy, **kw)

    np.array = np_array
    try:
        yield
    finally:
        np.array = orig_np_array


@contextlib.contextmanager
def cpu_shims():
    """Neutralise .cuda()/is_cuda/get_device() and numpy-2 ``copy=False``."""
    orig_cuda = torch.Tensor.cuda
    orig_get_device = torch.Tensor.get_device
    orig_is_cuda = torch.Tensor.is_cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.Tensor.get_device = lambda self: "cpu"
    torch.Tensor.is_cuda = property(lambda self: True)
    try:
        with numpy2_shim():
            yield
    finally:
        torch.Tensor.cuda = orig_cuda
        torch.Tensor.get_device = orig_get_device
        torch.Tensor.is_cuda = orig_is_cuda


# -------------------------- reference entry points --------------------------

def ref_random_shift(x: torch.Tensor, offsets, pad: int = 4, mode: str = "slice") -> torch.Tensor:
    """src/augmentations.py:378-382 with kornia's draw replaced by ``offsets`` (n,2)=(row,col)."""
    aug, _ = load_reference()
    _RandomCrop._next_offsets = np.asarray(offsets)
    _RandomCrop._mode = mode
    try:
        with cpu_shims():
            return aug.random_shift(x, pad)
    finally:
        _RandomCrop._next_offsets = None


def ref_random_crop(x: torch.Tensor, size: int, w1, h1) -> torch.Tensor:
    """src/augmentations.py:385-413 with supplied offsets (w1 indexes rows)."""
    aug, _ = load_reference()
    with cpu_shims():
        return aug.random_crop(x, size, w1=torch.as_tensor(w1).long(), h1=torch.as_tensor(h1).long())


def ref_random_conv(x: torch.Tensor, weights: torch.Tensor) -> torch.Tensor:
    """src/augmentations.py:265-274 with ``torch.randn(3,3,3,3)`` replaced by weights[i]."""
    aug, _ = load_reference()
    it = iter(weights)
    orig = aug.torch.randn
    aug.torch.randn = lambda *shape, **kw: next(it).clone()
    try:
        with cpu_shims():
            return aug.random_conv(x)
    finally:
        aug.torch.randn = orig


@contextlib.contextmanager
def _bank_batch(aug, imgs):
    orig_load, orig_get = aug.load_dataloader, aug._get_data_batch
    aug.load_dataloader = lambda *a, **k: None
    aug._get_data_batch = lambda batch_size: imgs.clone()
    try:
        yield
    finally:
        aug.load_dataloader, aug._get_data_batch = orig_load, orig_get


def ref_random_overlay(x: torch.Tensor, imgs: torch.Tensor) -> torch.Tensor:
    """src/augmentations.py:253-262; ``imgs`` (n,3,h,w) f32 in [0,1] replaces the Places/COCO batch."""
    aug, _ = load_reference()
    with cpu_shims(), _bank_batch(aug, imgs):
        return aug.random_overlay(x)


def ref_splice(x, imgs, hue_thres=0, sat_thres=0, val_thres=0.6) -> torch.Tensor:
    """src/augmentations.py:295-306 (+277-284)."""
    aug, _ = load_reference()
    with cpu_shims(), _bank_batch(aug, imgs):
        return aug.splice(x.clone(), hue_thres=hue_thres, sat_thres=sat_thres, val_thres=val_thres)


def ref_cs_splice(x, imgs, hue_thres=3.5, sat_thres=0, val_thres=0) -> torch.Tensor:
    """src/augmentations.py:309-324."""
    aug, _ = load_reference()
    with cpu_shims(), _bank_batch(aug, imgs):
        return aug.CS_splice(x.clone(), hue_thres=hue_thres, sat_thres=sat_thres, val_thres=val_thres)


def ref_replay_buffer(obs_shape, action_shape, capacity, batch_size):
    """A reference ``utils.ReplayBuffer`` (src/utils.py:136-154), prefill off."""
    _, ref_utils = load_reference()
    return ref_utils.ReplayBuffer(obs_shape, action_shape, capacity, batch_size, prefill=False)


def ref_lazy_frames(frames):
    _, ref_utils = load_reference()
    return ref_utils.LazyFrames(list(frames))


def ref_sample(buf, idxs):
    """src/utils.py:272-280 with explicit idxs, run on CPU."""
    with cpu_shims():
        return buf.sample(idxs=np.asarray(idxs))


# -------------------------- persistence (src/utils.py:156-204) --------------------------

@contextlib.contextmanager
def _as_utils_module():
    """Pickle resolves ``utils.LazyFrames`` through sys.modules; the reference's module is kept out of it
    (load_reference), so install it under its own name for the duration of a save / load.  torch >= 2.6
    defaults torch.load to weights_only=True, which the reference's torch 1.7.1 did not have: restore the
    old default for the call (third shim of the same kind as numpy's copy=False)."""
    import functools
    _, ref_utils = load_reference()
    saved = sys.modules.get("utils")
    sys.modules["utils"] = ref_utils
    orig_load = ref_utils.torch.load
    ref_utils.torch.load = functools.partial(orig_load, weights_only=False)
    try:
        yield ref_utils
    finally:
        ref_utils.torch.load = orig_load
        if saved is None:
            sys.modules.pop("utils", None)
        else:
            sys.modules["utils"] = saved


def ref_buffer_save(buf, save_dir):
    """ReplayBuffer.save (src/utils.py:169-170) of a reference buffer."""
    with _as_utils_module():
        buf.save(save_dir)


def ref_buffer_load(buf, save_dir, **kw):
    """ReplayBuffer.load (src/utils.py:177-204) into a reference buffer."""
    with _as_utils_module(), cpu_shims():
        buf.load(save_dir, **kw)


# -------------------------- remaining aug_to_func entries --------------------------

@contextlib.contextmanager
def _np_random(aug, uniform=None, randint=None):
    """Replace the np.random draws inside the reference module with queued arrays."""
    class _R:
        def __init__(self):
            self.u, self.r = list(uniform or []), list(randint or [])
        def uniform(self, *a, **k):
            return np.asarray(self.u.pop(0), dtype=np.float64)
        def randint(self, *a, **k):
            return np.asarray(self.r.pop(0))
    orig = aug.np
    shim = types.SimpleNamespace(**{k: getattr(orig, k) for k in dir(orig) if not k.startswith("__")})
    shim.random = _R()
    aug.np = shim
    try:
        yield
    finally:
        aug.np = orig


def ref_random_grayscale(x, mask, p=0.3):
    """src/augmentations.py:111-136 with rnd := 0 where mask else 1."""
    aug, _ = load_reference()
    u = np.where(np.asarray(mask), 0.0, 1.0)
    with cpu_shims(), _np_random(aug, uniform=[u]):
        return aug.random_grayscale(x.clone(), p)


def ref_random_flip(x, mask, p=0.2):
    aug, _ = load_reference()
    u = np.where(np.asarray(mask), 0.0, 1.0)
    with cpu_shims(), _np_random(aug, uniform=[u]):
        return aug.random_flip(x.clone(), p)


def ref_random_rotation(x, rot, p=0.3):
    """rot (n,) in {0..3}: 0 = not rotated (rnd > p), else rnd_rot."""
    aug, _ = load_reference()
    rot = np.asarray(rot)
    u = np.where(rot > 0, 0.0, 1.0)
    rr = np.where(rot > 0, rot, 1)
    with cpu_shims(), _np_random(aug, uniform=[u], randint=[rr]):
        return aug.random_rotation(x.clone(), p)


def ref_random_cutout_color(x, w1, h1, color_ints):
    """color_ints (n,c) = the reference's randint(0,255) draw (it divides by 255.0 itself)."""
    aug, _ = load_reference()
    with cpu_shims(), _np_random(aug, randint=[np.asarray(w1), np.asarray(h1), np.asarray(color_ints)]):
        return aug.random_cutout_color(x.clone())


# -------------------------- the reference on a GPU (bench.py's torch_cuda_reference leg) --------------------------
# The reference's functions are written for CUDA (.cuda() / is_cuda are hard-coded); on a GPU box they run as they are,
# with only the kornia stand-in, the numpy-2 shim and the injected draws.

def gpu_sample(buf, idxs):
    """src/utils.py:272-280 on the GPU: host gather (_encode_obses), pageable H2D, .float()."""
    with numpy2_shim():
        return buf.sample(idxs=np.asarray(idxs))


def gpu_random_shift(x, offsets, pad=4, mode="slice"):
    """src/augmentations.py:378-382 on CUDA tensors (F.pad + the kornia crop stand-in)."""
    aug, _ = load_reference()
    _RandomCrop._next_offsets = np.asarray(offsets)
    _RandomCrop._mode = mode
    try:
        return aug.random_shift(x, pad)
    finally:
        _RandomCrop._next_offsets = None


def gpu_splice_conv(x, imgs, weights):
    """random_conv(splice(x)) of src/augmentations.py:295-306, :265-274 on CUDA tensors: the AugCL strong phase
    (--data_aug splice-conv through RAD.apply_aug, src/algorithms/rad.py:18-22)."""
    aug, _ = load_reference()
    it = iter(weights)
    orig = aug.torch.randn
    aug.torch.randn = lambda *shape, **kw: next(it)
    try:
        with _bank_batch(aug, imgs):
            return aug.random_conv(aug.splice(x))
    finally:
        aug.torch.randn = orig
