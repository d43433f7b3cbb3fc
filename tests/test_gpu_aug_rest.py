"""GPU parity of the remaining aug_to_func entries (grayscale, flip, rotate, cutout_color) against the CPU
oracle (itself checked live against the reference's functions in tests/test_oracle_vs_reference.py): all four are
bit-exact.  color_jitter is checked against torchvision's own CPU ops (the reference calls torchvision directly)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(16, 9, 84, 84), (5, 3, 17, 17), (3, 6, 40, 40), (1, 9, 100, 100)])
@pytest.mark.parametrize("unit_range", [False, True])
def test_rest_against_oracle(shape, unit_range):
    from augcl_b200 import augmentations as A
    from oracle import augcl_oracle as O
    n, c, h, w = shape
    g = torch.Generator().manual_seed(sum(shape))
    rng = np.random.RandomState(sum(shape))
    x = torch.rand(shape, generator=g) * (1.0 if unit_range else 255.0)
    if not unit_range:
        x[0] = torch.randint(0, 256, shape[1:], generator=g).float()      # buffer-fresh integer pixels
    xd = x.cuda()
    mask = rng.rand(n) < 0.5
    mask[0] = True
    assert torch.equal(A.random_grayscale(xd, mask=mask).cpu(), O.random_grayscale(x, mask))
    assert torch.equal(A.random_flip(xd, mask=mask).cpu(), O.random_flip(x, mask))
    rot = rng.randint(0, 4, n)
    assert torch.equal(A.random_rotation(xd, rot=rot).cpu(), O.random_rotation(x, rot))
    w1, h1 = rng.randint(1, max(2, w // 2 + 3), n), rng.randint(1, max(2, h // 2 + 3), n)   # boxes clipped at the border too
    col = rng.randint(0, 255, (n, c)) / 255.0
    assert torch.equal(A.random_cutout_color(xd, w1=w1, h1=h1, colors=col).cpu(), O.random_cutout_color(x, w1, h1, col))
    assert torch.equal(xd.cpu(), x), "inputs must not be mutated"


def test_rest_draws_follow_numpy_like_the_reference():
    """Without supplied draws the decisions come from numpy's global RNG in the reference's order."""
    from augcl_b200 import augmentations as A
    from oracle import augcl_oracle as O
    x = torch.rand(8, 9, 84, 84) * 255
    np.random.seed(11)
    got = A.random_rotation(x.cuda(), p=0.6)
    np.random.seed(11)
    rnd = np.random.uniform(0.0, 1.0, size=(8,))
    rr = np.random.randint(1, 4, size=(8,))
    assert torch.equal(got.cpu(), O.random_rotation(x, rr * (rnd <= 0.6)))
    np.random.seed(12)
    got = A.random_cutout_color(x.cuda())
    np.random.seed(12)
    w1, h1 = np.random.randint(10, 30, 8), np.random.randint(10, 30, 8)
    col = np.random.randint(0, 255, size=(8, 9)) / 255.0
    assert torch.equal(got.cpu(), O.random_cutout_color(x, w1, h1, col))
    assert A.aug_to_func["rotate"] is A.random_rotation and A.aug_to_func["cutout_color"] is A.random_cutout_color
    with pytest.raises(Exception):
        A.random_rotation(torch.zeros(2, 3, 8, 10, device="cuda"))      # non-square


def test_create_hsv_mask_and_view_as_windows():
    """The two helpers splice / random_crop are built from in the reference (src/augmentations.py:277-284, :416-433)."""
    from augcl_b200 import augmentations as A
    from oracle import augcl_oracle as O
    g = torch.Generator().manual_seed(5)
    u8 = torch.randint(0, 256, (6, 3, 84, 84), generator=g, dtype=torch.uint8)
    u8[0, :, :10] = u8[0, 0:1, :10]                       # grey rows: saturation 0
    u8[1, 1, :20] = u8[1, 2, :20]                         # g == b where r may be the maximum: hue 0
    x = u8.float() / 255.0
    for thr in ((0, 0, 0.6), (0, 0, 0.4), (3.5, 0, 0), (1.0, 0.3, 0.2)):
        got = A.create_hsv_mask(x.cuda(), *thr)
        want = O.hsv_mask(x, *thr)
        assert got.dtype == torch.bool and got.shape == (6, 3, 84, 84)
        assert torch.equal(got.cpu(), want.unsqueeze(1).repeat(1, 3, 1, 1)), thr
    # used the way the reference's random_crop does (:401-408): NHWC view, windows[arange(n), w1, h1]
    y = torch.randint(0, 256, (4, 9, 20, 20), generator=g).float()
    w1, h1 = torch.tensor([0, 3, 7, 2]), torch.tensor([7, 0, 1, 5])
    win = A.view_as_windows_cuda(y.cuda().permute(0, 2, 3, 1), (1, 12, 12, 1))[..., 0, :, :, 0]
    assert win.shape == (4, 8, 8, 9, 12, 12)
    assert torch.equal(win[torch.arange(4), w1, h1].cpu(), O.random_crop(y, 12, w1.numpy(), h1.numpy()))


def _tv_jitter(flat, fn_idx, bf, cf, sf, hf):
    """torchvision's ColorJitter.forward with given parameters, on CPU (the reference's op: src/augmentations.py:233)."""
    import torchvision.transforms.functional as F
    for fn_id in fn_idx:
        if fn_id == 0 and bf is not None:
            flat = F.adjust_brightness(flat, bf)
        elif fn_id == 1 and cf is not None:
            flat = F.adjust_contrast(flat, cf)
        elif fn_id == 2 and sf is not None:
            flat = F.adjust_saturation(flat, sf)
        elif fn_id == 3 and hf is not None:
            flat = F.adjust_hue(flat, hf)
    return flat


@pytest.mark.parametrize("shape", [(6, 9, 84, 84), (4, 3, 17, 23), (3, 6, 100, 100)])
@pytest.mark.parametrize("unit_range", [True, False])
def test_color_jitter_kernel_against_torchvision(shape, unit_range):
    """color_jitter (src/augmentations.py:225-237) as one kernel vs torchvision's own CPU ops with the same order and
    factors: every permutation class of the four ops, factors at and inside the range ends, on [0,1] images (where the ops
    are meaningful) and on the reference's 0..255 images (where _blend's clamp to [0,1] saturates almost everything).
    Elementwise ops are restated exactly; only contrast's mean is summed in another order, a few-ulp difference that a
    following hue adjustment (divisions by the channel range) can amplify: <= 5e-5 absolute on [0,1] outputs, and exact
    whenever contrast is off."""
    import itertools
    from augcl_b200 import augmentations as A
    g = torch.Generator().manual_seed(sum(shape))
    b, c, h, w = shape
    rng = np.random.RandomState(b)
    perms = list(itertools.permutations(range(4)))
    cases = [(perms[i], None, float(rng.uniform(0, 2)), float(rng.uniform(0, 3)), float(rng.uniform(-0.5, 0.5)))
             for i in rng.choice(24, 6, replace=False)]
    cases += [((0, 1, 2, 3), 1.3, 0.0, 3.0, 0.5), ((3, 2, 1, 0), None, 2.0, 0.0, -0.5), ((1, 3, 0, 2), None, None, 1.0, None),
              ((2, 0, 3, 1), None, 1.0, None, 0.0)]
    for fn_idx, bf, cf, sf, hf in cases:
        x = torch.rand(shape, generator=g) * (1.0 if unit_range else 255.0)
        if unit_range:
            x[0, :3, :4] = 0.5                       # grey pixels: max == min branches of _rgb2hsv
            x[0, :3, 4:6] = torch.tensor([1.0, 0.0, 0.0]).view(3, 1, 1)
        idxs = rng.randint(0, b, size=b * (c // 3))
        want = x.clone().view(-1, 3, h, w)
        sel = torch.from_numpy(idxs)
        want[sel] = _tv_jitter(want[sel], fn_idx, bf, cf, sf, hf)
        xd = x.cuda()
        got = A.color_jitter(xd, sampled_idxs=idxs, params=(torch.tensor(fn_idx), bf, cf, sf, hf))
        assert got.data_ptr() == xd.data_ptr() and got.shape == x.shape      # in place, like the reference
        err = (got.cpu().view(-1, 3, h, w) - want).abs().max().item()
        assert err <= (5e-5 if cf is not None else 0.0), (fn_idx, bf, cf, sf, hf, err)
        untouched = np.setdiff1d(np.arange(b * (c // 3)), idxs)
        assert torch.equal(got.cpu().view(-1, 3, h, w)[untouched], x.view(-1, 3, h, w)[untouched])


def test_color_jitter_draws_follow_the_reference():
    """Without supplied draws: numpy's global RNG picks the frames, torch's CPU RNG the order and factors, consumed
    exactly as `model(flat[sampled_idxs])` with a fresh ColorJitter does (src/augmentations.py:229-233)."""
    import torchvision.transforms as TF
    from augcl_b200 import augmentations as A
    x = torch.rand(8, 9, 84, 84)
    np.random.seed(5)
    torch.manual_seed(5)
    got = A.color_jitter(x.clone().cuda(), p=0.5)
    np.random.seed(5)
    torch.manual_seed(5)
    flat = x.clone().view(-1, 3, 84, 84)
    idxs = torch.from_numpy(np.random.randint(0, 8, int(0.5 * 8 * 3)))
    flat[idxs] = TF.ColorJitter(0, (0, 2), (0, 3), (-0.5, 0.5))(flat[idxs])
    assert (got.cpu() - flat.view(8, 9, 84, 84)).abs().max().item() <= 5e-5
    assert A.aug_to_func["color_jitter"] is A.color_jitter
    with pytest.raises(Exception):
        A.color_jitter(x.cuda(), params=(torch.tensor([0, 1, 2, 2]), None, 1.0, 1.0, 0.0))    # not a permutation
    with pytest.raises(Exception):
        A.color_jitter(x.cuda(), params=(torch.tensor([0, 1, 2, 3]), None, 1.0, 1.0, 0.7))    # hue outside [-0.5, 0.5]
