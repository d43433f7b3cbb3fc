"""GPU parity of the remaining aug_to_func entries (grayscale, flip, rotate, cutout_color) against the CPU
oracle (itself checked live against the reference's functions in tests/test_oracle_vs_reference.py).
All four are bit-exact."""
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
