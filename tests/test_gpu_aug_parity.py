"""GPU parity: the CUDA augmentation kernels (through the C ABI, via augcl_b200.augmentations) against
(1) the REFERENCE's own outputs stored in tests/golden and (2) the CPU oracle on seeded inputs.

Tolerance contract: shift(int) / crop / splice selection bit-exact; conv, overlay,
bilinear shift <= 1e-5 absolute on the /255-normalised scale."""
import numpy as np
import pytest
import torch

from conftest import AUG_CASES, TOL_NORM

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def A():
    from augcl_b200 import augmentations
    return augmentations


@pytest.fixture(scope="module")
def O():
    from oracle import augcl_oracle
    return augcl_oracle


def dev(a, dtype=None):
    t = torch.from_numpy(np.asarray(a)).cuda()
    return t if dtype is None else t.to(dtype)


def norm_err(got, want):
    return (got.detach().cpu().double() - torch.as_tensor(want).double()).abs().max().item() / 255.0


@pytest.mark.parametrize("name", AUG_CASES)
def test_against_reference_golden(A, golden, name):
    d = golden(name)
    x = dev(d["x_u8"]).float()
    n = x.shape[0]
    x0 = x.clone()
    bank = dev(d["bank_u8"])
    rows = torch.arange(n)
    A.set_background_bank(bank)
    try:
        out = A.random_shift(x, int(d["pad"]), offsets=d["shift_off"])
        assert torch.equal(out.cpu(), torch.from_numpy(d["shift_out"]))
        # the same call as the pinned-era environment runs it (kornia 0.5.x: RandomCrop is a bilinear warp,
        # /root/reference/src/augmentations.py:382): reference output recorded in warp mode, float tolerance
        assert norm_err(out, golden("shift_warp")[name]) <= TOL_NORM
        assert norm_err(A.random_conv(x, weights=d["conv_w"]), d["conv_out"]) <= TOL_NORM
        assert norm_err(A.random_overlay(x, bank_idx=rows), d["overlay_out"]) <= TOL_NORM
        for thr, key in [((0, 0, 0.6), "splice_out_06"), ((0, 0, 0.4), "splice_out_04"), ((1.0, 0.2, 0.3), "splice_out_hs")]:
            got = A.splice(x, *thr, bank_idx=rows)
            assert torch.equal(got.cpu(), torch.from_numpy(d[key])), key
        assert torch.equal(A.CS_splice(x, bank_idx=rows).cpu(), torch.from_numpy(d["cs_splice_out"]))
        # chains: splice -> conv (AugCL strong phase) and conv -> splice (non-integer pixels into the HSV mask)
        sc = A.random_conv(A.splice(x, 0, 0, 0.6, bank_idx=rows), weights=d["conv_w"])
        assert norm_err(sc, d["splice_conv_out"]) <= TOL_NORM
        cs = A.splice(dev(d["conv_out"]), 0, 0, 0.6, bank_idx=rows)
        assert torch.equal(cs.cpu(), torch.from_numpy(d["conv_splice_out"]))
        # f32 bank in [0,1] (what the reference's loader yields) gives the same results as the u8 bank
        A.set_background_bank((bank.cpu().float() / 255.0).cuda())    # IEEE division on the host, like ToTensor
        assert torch.equal(A.splice(x, 0, 0, 0.6, bank_idx=rows).cpu(), torch.from_numpy(d["splice_out_06"]))
        assert norm_err(A.random_overlay(x, bank_idx=rows), d["overlay_out"]) <= TOL_NORM
        if "crop_out" in d:
            xc = dev(d["crop_x_u8"]).float()
            got = A.random_crop(xc, int(d["crop_size"]), w1=torch.from_numpy(d["crop_w1"]), h1=torch.from_numpy(d["crop_h1"]))
            assert torch.equal(got.cpu(), torch.from_numpy(d["crop_out"]))
        assert torch.equal(x, x0), "inputs must not be mutated"
    finally:
        A.set_background_bank(None)


SHAPES = [((128, 9, 84, 84), 4), ((7, 6, 33, 40), 3), ((5, 3, 17, 23), 2), ((1, 9, 84, 84), 4), ((3, 12, 100, 100), 4)]


@pytest.mark.parametrize("shape,pad", SHAPES)
@pytest.mark.parametrize("integer_pixels", [True, False])
def test_against_oracle(A, O, shape, pad, integer_pixels):
    g = torch.Generator().manual_seed(sum(shape) + pad)
    n, c, h, w = shape
    if integer_pixels:
        x = torch.randint(0, 256, shape, generator=g, dtype=torch.uint8).float()
    else:
        x = torch.rand(shape, generator=g) * 255.0
    xd = x.cuda()
    off = torch.randint(0, 2 * pad + 1, (n, 2), generator=g)
    assert torch.equal(A.random_shift(xd, pad, offsets=off).cpu(), O.random_shift_int(x, off.numpy(), pad))
    size = h - 5
    if h == w:
        w1 = torch.randint(0, w - size, (n,), generator=g)
        h1 = torch.randint(0, w - size, (n,), generator=g)
        got, gw1, gh1 = A.random_crop(xd, size, w1=w1, h1=h1, return_w1_h1=True)
        assert torch.equal(got.cpu(), O.random_crop(x, size, w1.numpy(), h1.numpy()))
        assert torch.equal(gw1.cpu(), w1) and torch.equal(gh1.cpu(), h1)
        assert norm_err(A.random_shift(xd, pad, offsets=off, mode="bilinear"),
                        O.random_shift_bilinear(x, off.flip(1).numpy(), pad)) <= TOL_NORM
    wt = torch.randn(n, 3, 3, 3, 3, generator=g)
    assert norm_err(A.random_conv(xd, weights=wt), O.random_conv(x, wt)) <= TOL_NORM
    M = 11
    bank = torch.randint(0, 256, (M, 3, h, w), generator=g, dtype=torch.uint8)
    rows = torch.randint(0, M, (n,), generator=g)
    imgs = O.bank_rows_f32(bank, rows.numpy())
    A.set_background_bank(bank.cuda())
    try:
        ov = A.random_overlay(xd, bank_idx=rows)
        assert norm_err(ov, O.random_overlay(x, imgs)) <= TOL_NORM
        assert torch.equal(ov.cpu(), O.random_overlay(x, imgs)), "overlay restates the reference op order exactly"
        for thr in [(0, 0, 0.6), (0, 0, 0.4), (2.0, 0.3, 0.2)]:
            assert torch.equal(A.splice(xd, *thr, bank_idx=rows).cpu(), O.splice(x, imgs, *thr))
        assert torch.equal(A.CS_splice(xd, bank_idx=rows).cpu(), O.cs_splice(x, imgs))
    finally:
        A.set_background_bank(None)


@pytest.mark.parametrize("S", [64, 85, 150, 252, 300, 258])
def test_bilinear_shift_kernel_variants(A, O, S):
    """Planes up to 96 KB take the shared-memory staged kernel (85: scalar fill, ragged row chunks), wider multiples of 4
    the 4-pixel one, anything else the generic one.  The pixel coordinate's ulp grows with S, and with it the weight a
    floor() flip moves between neighbouring taps: the tolerance scales accordingly above the 84-pixel case."""
    g = torch.Generator().manual_seed(S)
    x = torch.rand((3, 4, S, S), generator=g) * 255.0
    for pad in (4, 11):
        off = torch.randint(0, 2 * pad + 1, (3, 2), generator=g)
        assert norm_err(A.random_shift(x.cuda(), pad, offsets=off, mode="bilinear"),
                        O.random_shift_bilinear(x, off.flip(1).numpy(), pad)) <= TOL_NORM * max(1.0, S / 84)


@pytest.mark.parametrize("shape", [(3, 3, 8, 8), (2, 6, 16, 16), (5, 9, 84, 84), (2, 3, 33, 40), (3, 6, 100, 100),
                                   (1, 3, 50, 200)])
def test_random_conv_kernel_variants(A, O, shape):
    """The part-of-frame kernel (any part count, including the ragged last part and tiny images whose CTA is smaller
    than the 81 weight loaders) and the 12-row band kernels give the oracle's result."""
    from augcl_b200 import _lib as L
    g = torch.Generator().manual_seed(sum(shape))
    x = torch.rand(shape, generator=g) * 255.0
    wt = torch.randn(shape[0], 3, 3, 3, 3, generator=g)
    want = O.random_conv(x, wt)
    try:
        for parts in (1, 0, 2, 3, 4, 6):
            L.set_option("conv_part", str(parts))
            assert norm_err(A.random_conv(x.cuda(), weights=wt), want) <= TOL_NORM, parts
    finally:
        L.set_option("conv_part", "1")


def test_known_answers(A):
    # zero weights -> sigmoid(0)*255 == 127.5 everywhere
    x = torch.rand(2, 6, 10, 12, device="cuda") * 255
    # (255 / (1 + 2^acc) is evaluated as rcp(fma(2^acc, 1/255, 1/255)): 127.5 to one ulp of fl(2/255), not bit-exact)
    assert (A.random_conv(x, weights=torch.zeros(2, 3, 3, 3, 3)) - 127.5).abs().max().item() / 255.0 <= 1e-6
    # centre offset == identity; crop passthrough when size >= width
    assert torch.equal(A.random_shift(x, 3, offsets=torch.full((2, 2), 3)), x)
    assert A.random_crop(x, 12) is x
    out, w1, h1 = A.random_crop(x, 12, return_w1_h1=True)
    assert out is x and w1 is None and h1 is None
    # overlay fixed points
    xe = torch.tensor([0.0, 255.0], device="cuda").reshape(1, 1, 1, 2).repeat(1, 3, 2, 1).contiguous()
    A.set_background_bank(torch.tensor([0, 255], dtype=torch.uint8, device="cuda").reshape(1, 1, 1, 2).repeat(1, 3, 2, 1).contiguous())
    try:
        assert torch.equal(A.random_overlay(xe, bank_idx=[0]), xe)
    finally:
        A.set_background_bank(None)


def test_device_draws_match_oracle_philox(A, O):
    """When the caller supplies no draws they come from Philox (seed, call counter)."""
    A.seed(123)
    x = torch.randint(0, 256, (16, 9, 84, 84), dtype=torch.uint8).float()
    got = A.random_shift(x.cuda(), 4)                         # call 0: offsets
    off = O.rng_randint(32, 0, 9, 123, 0).reshape(16, 2)
    assert torch.equal(got.cpu(), O.random_shift_int(x, off, 4))
    got = A.random_conv(x.cuda())                             # call 1: normals
    wt = torch.from_numpy(O.rng_normal(16 * 81, 123, 1)).reshape(16, 3, 3, 3, 3)
    assert norm_err(got, O.random_conv(x, wt)) <= 5e-5        # device sin/cos/log vs numpy f64 in the weights
    xc = torch.randint(0, 256, (4, 9, 100, 100), dtype=torch.uint8).float()
    got, w1, h1 = A.random_crop(xc.cuda(), 84, return_w1_h1=True)   # call 2
    off = O.rng_randint(8, 0, 16, 123, 2).reshape(4, 2)
    assert w1.cpu().tolist() == off[:, 0].tolist() and h1.cpu().tolist() == off[:, 1].tolist()
    assert int(w1.max()) <= 15
    assert torch.equal(got.cpu(), O.random_crop(xc, 84, off[:, 0], off[:, 1]))


def test_staging_switch_gives_identical_results(A):
    from augcl_b200 import _lib as L
    g = torch.Generator().manual_seed(5)
    x = (torch.rand(9, 9, 84, 84, generator=g) * 255).cuda()
    off = torch.randint(0, 9, (9, 2), generator=g)
    a = A.random_shift(x, 4, offsets=off)
    L.set_option("staging", "generic")
    try:
        b = A.random_shift(x, 4, offsets=off)
    finally:
        L.set_option("staging", "tma")
    assert torch.equal(a, b)


def test_errors(A):
    from augcl_b200._lib import AugclError
    x = torch.zeros(2, 9, 8, 8, device="cuda")
    with pytest.raises(AssertionError):
        A.random_crop(x.cpu())
    with pytest.raises(AssertionError):
        A.random_crop(x, 4, w1=torch.zeros(2))
    with pytest.raises(AugclError):
        A.random_conv(torch.zeros(2, 4, 8, 8, device="cuda"))       # c not a multiple of 3
    with pytest.raises(NotImplementedError):
        A.random_overlay(x, dataset="imagenet")
    with pytest.raises(ValueError):
        A.random_shift(x, 4, offsets=torch.zeros(3, 2))
    assert set(A.aug_to_func) == {"grayscale", "flip", "rotate", "conv", "shift", "color_jitter", "identity",
                                  "splice", "crop", "cutout_color", "CS_splice", "overlay"}
