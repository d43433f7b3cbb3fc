"""CPU, build container only: oracle restatement vs the reference's own code run live
(skipped where /root/reference is absent, e.g. on the GPU box)."""
import numpy as np
import pytest
import torch

from oracle import augcl_oracle as O
from oracle import ref_harness as R

pytestmark = pytest.mark.skipif(not R.available(), reason="/root/reference not mounted")


@pytest.mark.parametrize("shape,pad", [((3, 9, 84, 84), 4), ((4, 6, 17, 23), 3), ((2, 3, 8, 8), 1)])
def test_live_parity(shape, pad):
    g = torch.Generator().manual_seed(sum(shape))
    n, c, h, w = shape
    x = torch.randint(0, 256, shape, generator=g, dtype=torch.uint8).float()
    xf = torch.rand(shape, generator=g) * 255.0
    off = torch.randint(0, 2 * pad + 1, (n, 2), generator=g).numpy()
    assert torch.equal(R.ref_random_shift(x, off, pad), O.random_shift_int(x, off, pad))
    wt = torch.randn(n, 3, 3, 3, 3, generator=g)
    imgs = torch.randint(0, 256, (n, 3, h, w), generator=g, dtype=torch.uint8).float() / 255
    for xin in (x, xf):
        assert torch.equal(R.ref_random_conv(xin, wt), O.random_conv(xin, wt))
        assert torch.equal(R.ref_random_overlay(xin, imgs), O.random_overlay(xin, imgs))
        for thr in [(0, 0, 0.6), (0, 0, 0.4), (2.0, 0.3, 0.2)]:
            assert torch.equal(R.ref_splice(xin, imgs, *thr), O.splice(xin, imgs, *thr))
        assert torch.equal(R.ref_cs_splice(xin, imgs), O.cs_splice(xin, imgs))


def test_live_crop_and_hsv():
    g = torch.Generator().manual_seed(7)
    x = torch.randint(0, 256, (5, 9, 100, 100), generator=g, dtype=torch.uint8).float()
    w1 = torch.randint(0, 16, (5,), generator=g)
    h1 = torch.randint(0, 16, (5,), generator=g)
    assert torch.equal(R.ref_random_crop(x, 84, w1, h1), O.random_crop(x, 84, w1, h1))
    xr = torch.rand(16, 3, 32, 32, generator=g)
    assert torch.equal(R.kornia_rgb_to_hsv(xr), O.rgb_to_hsv(xr))
    # all 2^24 uint8 colours would take minutes; a 2^18 random subset + ties
    cols = torch.randint(0, 256, (1, 3, 512, 512), generator=g).float() / 255
    cols[0, :, :16] = cols[0, 0:1, :16]                     # greys (max==min)
    cols[0, 1, 16:32] = cols[0, 0, 16:32]                   # r==g ties
    assert torch.equal(R.kornia_rgb_to_hsv(cols), O.rgb_to_hsv(cols))


def test_live_replay_gather():
    rng = np.random.RandomState(0)
    k, S, A, cap = 3, 12, 4, 9
    ref = R.ref_replay_buffer((3 * k, S, S), (A,), cap, 5)
    ora = O.HostRing((3 * k, S, S), (A,), cap, 5)
    frames = rng.randint(0, 256, (20 + k, 3, S, S)).astype(np.uint8)
    for t in range(20):
        a, r, dn = rng.randn(A).astype(np.float32), float(rng.randn()), bool(rng.rand() < 0.2)
        ref.add(R.ref_lazy_frames(frames[t:t + k]), a, r, R.ref_lazy_frames(frames[t + 1:t + 1 + k]), dn)
        ora.add(O.LazyFrames(frames[t:t + k]), a, r, O.LazyFrames(frames[t + 1:t + 1 + k]), dn)
    assert (ref.idx, ref.full) == (ora.idx, ora.full)
    idxs = rng.randint(0, cap, size=7)
    for a, b in zip(R.ref_sample(ref, idxs), ora.sample(idxs)):
        assert torch.equal(a, b)


@pytest.mark.parametrize("unit_range", [False, True])
def test_live_remaining_augs(unit_range):
    """grayscale / flip / rotate / cutout_color: oracle restatement == the reference's functions with the
    np.random draws injected."""
    g = torch.Generator().manual_seed(3)
    rng = np.random.RandomState(3)
    n = 7
    x = torch.rand(n, 9, 24, 24, generator=g) * (1.0 if unit_range else 255.0)
    mask = rng.rand(n) < 0.5
    assert torch.equal(R.ref_random_grayscale(x, mask), O.random_grayscale(x, mask))
    assert torch.equal(R.ref_random_flip(x, mask), O.random_flip(x, mask))
    rot = rng.randint(0, 4, n)
    assert torch.equal(R.ref_random_rotation(x, rot), O.random_rotation(x, rot))
    w1, h1, ci = rng.randint(1, 15, n), rng.randint(1, 15, n), rng.randint(0, 255, (n, 9))
    assert torch.equal(R.ref_random_cutout_color(x, w1, h1, ci), O.random_cutout_color(x, w1, h1, ci / 255.0))


def test_live_view_as_windows_and_hsv_mask():
    """The product's pure-metadata view_as_windows_cuda equals the reference's (src/augmentations.py:416-433); the oracle's
    hsv_mask equals the reference's create_hsv_mask (:277-284) run with the kornia restatement of ref_harness."""
    from augcl_b200 import augmentations as A
    aug, _ = R.load_reference()
    g = torch.Generator().manual_seed(9)
    y = torch.rand(3, 20, 20, 9, generator=g)
    assert torch.equal(aug.view_as_windows_cuda(y, (1, 12, 12, 1)), A.view_as_windows_cuda(y, (1, 12, 12, 1)))
    x = torch.randint(0, 256, (4, 3, 16, 16), generator=g).float() / 255.0
    with R.cpu_shims():
        for thr in ((0, 0, 0.6), (3.5, 0, 0), (1.0, 0.3, 0.2)):
            want = aug.create_hsv_mask(x, *thr)
            assert torch.equal(O.hsv_mask(x, *thr).unsqueeze(1).repeat(1, 3, 1, 1), want), thr
