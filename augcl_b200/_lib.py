"""ctypes binding of libaugcl_b200.so (the C ABI declared in include/augcl_b200.h).

The product path has NO CPU fallback: if the shared library is missing or a call fails,
an exception is raised.  Nothing under ``oracle/`` is ever imported from here.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libaugcl_b200.so")

# enums of include/augcl_b200.h
AUG_NONE, AUG_SHIFT, AUG_CROP, AUG_SHIFT_OBS = 0, 1, 2, 3
SHIFT_INT, SHIFT_BILINEAR = 0, 1
BANK_U8, BANK_F32 = 0, 1
SPLICE, CS_SPLICE = 0, 1
SRC_DEVICE, SRC_HOST, SRC_PHILOX = 0, 1, 2
OP_SPLICE, OP_CS_SPLICE, OP_OVERLAY, OP_CONV = 1, 2, 4, 8
AUG_INPUT_PLAIN, AUG_INPUT_SHIFTED = 0, 1

_p, _i, _i64, _u64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_float



class FusedDesc(C.Structure):
    """augcl_fused_desc of include/augcl_b200.h (field for field)."""
    _fields_ = [
        ("struct_size", C.c_int32), ("B", C.c_int32), ("idx_src", C.c_int32), ("off_src", C.c_int32),
        ("idx", _p), ("idx_hi", _i64), ("off_obs", _p), ("off_next", _p),
        ("pad", C.c_int32), ("normalize", C.c_int32), ("seed", _u64), ("subsequence", _u64),
        ("out_obs", _p), ("out_obs_shift", _p), ("out_obs_aug", _p), ("out_next", _p), ("out_next_shift", _p),
        ("out_act", _p), ("out_rew", _p), ("out_nd", _p),
        ("aug_input", C.c_int32), ("aug_ops", C.c_int32),
        ("bank", _p), ("bank_len", _i64), ("bank_idx", _p), ("conv_weights", _p),
        ("hue_thres", _f), ("sat_thres", _f), ("val_thres", _f), ("alpha", _f),
        ("bank_subsequence", _u64),
        ("conv_subsequence", _u64),
    ]


# name -> (restype, argtypes); every symbol include/augcl_b200.h declares
SIGNATURES = {
    "augcl_version": (C.c_char_p, []),
    "augcl_last_error": (C.c_char_p, []),
    "augcl_launch_count": (_i64, []),
    "augcl_set_option": (_i, [C.c_char_p, C.c_char_p]),
    "augcl_stream_delay": (_i, [_u64, _p]),
    "augcl_ring_create": (_i, [C.POINTER(_p), _i, _i64, _i, _i, _i, _i, _i]),
    "augcl_ring_create_dedup": (_i, [C.POINTER(_p), _i, _i64, _i, _i, _i, _i, _i, _i64]),
    "augcl_ring_put_frames": (_i, [_p, _i64, _i64, _p, _p]),
    "augcl_ring_put_refs": (_i, [_p, _i64, _i64, _p, _p, _p, _p, _p]),
    "augcl_ring_destroy": (_i, [_p]),
    "augcl_ring_bytes": (_i64, [_p]),
    "augcl_ring_put": (_i, [_p, _i64, _i64, _p, _p, _p, _p, _p, _p]),
    "augcl_ring_get": (_i, [_p, _i64, _i64, _p, _p, _p, _p, _p, _p]),
    "augcl_ring_fill_synthetic": (_i, [_p, _i64, _u64, _p]),
    "augcl_ring_sample": (_i, [_p, _p, _i, _p, _p, _p, _p, _p, _i, _p, _p, _i, _i, _p]),
    "augcl_ring_sample_noalloc": (_i, [_p, _p, _i, _p, _p, _p, _p, _p, _i, _p, _p, _i, _i, _p]),
    "augcl_ring_sample_ex": (_i, [_p, _i, _p, _i64, _i, _p, _p, _p, _p, _p, _i, _i, _p, _p, _i, _i, _u64, _u64, _p]),
    "augcl_ring_sample_curl": (_i, [_p, _i, _p, _i64, _i, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p, _i, _i, _u64, _u64, _p]),
    "augcl_ring_sample_views": (_i, [_p, _i, _p, _i64, _i, _i, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _u64, _u64, _p]),
    "augcl_ring_sample_fused": (_i, [_p, C.POINTER(FusedDesc), _p]),
    "augcl_random_shift": (_i, [_p, _p, _i, _i, _i, _i, _i, _p, _i, _p, _p]),
    "augcl_drq_base_grid": (_i, [_i, _i, C.POINTER(_f)]),
    "augcl_random_crop": (_i, [_p, _p, _i, _i, _i, _i, _i, _p, _p]),
    "augcl_random_conv": (_i, [_p, _p, _i, _i, _i, _i, _p, _p]),
    "augcl_random_overlay": (_i, [_p, _p, _i, _i, _i, _i, _p, _i, _i64, _p, _f, _p]),
    "augcl_splice": (_i, [_p, _p, _i, _i, _i, _i, _p, _i, _i64, _p, _f, _f, _f, _i, _p]),
    "augcl_hsv_mask": (_i, [_p, _p, _i, _i, _i, _f, _f, _f, _p]),
    "augcl_color_jitter": (_i, [_p, _i, _i, _i, _p, _i, C.POINTER(C.c_int32), C.c_double, C.c_double, C.c_double,
                                C.c_double, _i, _p]),
    "augcl_random_grayscale": (_i, [_p, _p, _i, _i, _i, _i, _p, _p]),
    "augcl_random_flip": (_i, [_p, _p, _i, _i, _i, _i, _p, _p]),
    "augcl_random_rotation": (_i, [_p, _p, _i, _i, _i, _i, _p, _p]),
    "augcl_random_cutout_color": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _p]),
    "augcl_rng_fill_offsets": (_i, [_p, _i64, C.c_int32, C.c_int32, _u64, _u64, _p]),
    "augcl_rng_fill_indices": (_i, [_p, _i64, _i64, _i64, _u64, _u64, _p]),
    "augcl_rng_fill_normal": (_i, [_p, _i64, _u64, _u64, _p]),
    "augcl_rng_fill_bytes": (_i, [_p, _i64, _u64, _u64, _p]),
}


class AugclError(RuntimeError):
    """A C-ABI call returned a negative augcl_status."""

    def __init__(self, fn: str, status: int, msg: str):
        super().__init__(f"{fn} failed with status {status}: {msg}")
        self.status = status


_lib = None


def lib() -> C.CDLL:
    """Loads the library once; raises (never falls back) when it is absent."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build it with `python -m augcl_b200.build` "
                "(augcl_b200 has no CPU fallback)")
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)
            fn.restype, fn.argtypes = res, args
        _lib = h
    return _lib


def call(name: str, *args) -> None:
    """Invoke a status-returning entry point; raise AugclError on a negative status."""
    h = lib()
    rc = getattr(h, name)(*args)
    if rc != 0:
        raise AugclError(name, rc, h.augcl_last_error().decode(errors="replace"))


def launch_count() -> int:
    return int(lib().augcl_launch_count())


def set_option(key: str, value: str) -> None:
    call("augcl_set_option", key.encode(), value.encode())


def version() -> str:
    return lib().augcl_version().decode()
