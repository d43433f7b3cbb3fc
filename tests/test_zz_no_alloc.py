"""augcl_ring_sample_noalloc (include/augcl_b200.h): the HOST-draw sample call whose staging scratch is a fixed file-scope
device array instead of stream-ordered allocations.  No counterpart in the reference.
Collected last on purpose.

Checked through the C ABI against the existing entry point (augcl_ring_sample_ex with the same HOST draws), on two calls in
a row with different draws -- the array keeps the previous call's contents -- and at both sides of the batch threshold that
switches kernels.  Gather + integer shift are bit-exact in this repository's tests, so the comparison is torch.equal."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

S, K, PAD, CAP, A_DIM = 84, 3, 4, 3000, 6
MAX_B = 16384      # AUGCL_NOALLOC_MAX_BATCH


def _ring():
    from augcl_b200 import utils as U
    buf = U.ReplayBuffer((3 * K, S, S), (A_DIM,), CAP, 128)
    buf.fill_synthetic(seed=41)
    return buf


def _outputs(B):
    dev = torch.device("cuda", 0)
    return (torch.empty((B, 3 * K, S, S), device=dev), torch.empty((B, 3 * K, S, S), device=dev),
            torch.empty((B, A_DIM), device=dev), torch.empty((B, 1), device=dev), torch.empty((B, 1), device=dev))


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.mark.parametrize("B", [128, 700])      # the plane kernel / the band-of-frames kernel
def test_noalloc_equals_sample_ex_on_two_calls_in_a_row(B):
    from augcl_b200 import _lib as L
    buf = _ring()
    st = torch.cuda.current_stream().cuda_stream
    rng = np.random.RandomState(B)
    for call in range(2):
        idx = np.ascontiguousarray(rng.randint(0, CAP, size=B).astype(np.int64))
        off_o = np.ascontiguousarray(rng.randint(0, 2 * PAD + 1, size=(B, 2)).astype(np.int32))
        off_n = np.ascontiguousarray(rng.randint(0, 2 * PAD + 1, size=(B, 2)).astype(np.int32))
        for mode, on in ((L.AUG_SHIFT, off_n), (L.AUG_NONE, None), (L.AUG_SHIFT_OBS, None)):
            got, want = _outputs(B), _outputs(B)
            L.call("augcl_ring_sample_noalloc", buf._ring, _ptr(idx), B, got[0].data_ptr(), got[1].data_ptr(), got[2].data_ptr(),
                   got[3].data_ptr(), got[4].data_ptr(), mode, _ptr(off_o), None if on is None else _ptr(on), PAD, 0, st)
            L.call("augcl_ring_sample_ex", buf._ring, L.SRC_HOST, _ptr(idx), 0, B, want[0].data_ptr(), want[1].data_ptr(),
                   want[2].data_ptr(), want[3].data_ptr(), want[4].data_ptr(), mode, L.SRC_HOST, _ptr(off_o),
                   None if on is None else _ptr(on), PAD, 0, 0, 0, st)
            torch.cuda.synchronize()
            for g, w, name in zip(got, want, ("obs", "next_obs", "action", "reward", "not_done")):
                assert torch.equal(g, w), (call, mode, name)
            assert float(got[0].abs().sum()) > 0.0


def test_noalloc_rejects_what_does_not_fit_and_bad_draws():
    from augcl_b200 import _lib as L
    buf = _ring()
    st = torch.cuda.current_stream().cuda_stream
    o = _outputs(4)
    args = lambda idx, B, off: (buf._ring, _ptr(idx), B, o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(), o[3].data_ptr(),
                                o[4].data_ptr(), L.AUG_SHIFT, _ptr(off), None, PAD, 0, st)
    big = np.zeros(MAX_B + 1, dtype=np.int64)
    with pytest.raises(L.AugclError, match="exceeds the fixed scratch"):
        L.call("augcl_ring_sample_noalloc", *args(big, MAX_B + 1, np.zeros((MAX_B + 1, 2), dtype=np.int32)))
    with pytest.raises(L.AugclError, match="out of range"):
        L.call("augcl_ring_sample_noalloc", *args(np.array([0, 1, CAP, 2], dtype=np.int64), 4, np.zeros((4, 2), dtype=np.int32)))
    with pytest.raises(L.AugclError, match="offset"):
        L.call("augcl_ring_sample_noalloc", *args(np.arange(4, dtype=np.int64), 4, np.full((4, 2), 2 * PAD + 1, dtype=np.int32)))
    torch.cuda.synchronize()
