"""``import augmentations`` of the reference (src/augmentations.py) resolved to augcl_b200's kernels (see __init__.py)."""
from augcl_b200.augmentations import *          # noqa: F401,F403
from augcl_b200.augmentations import (          # noqa: F401  names a star import skips or callers reach by attribute
    aug_to_func, get_background_bank, _bank_rows, _next_subsequence, _reserve_subsequences)
import augcl_b200.augmentations as _impl


def __getattr__(name):          # module state (seed, call counter, bank) lives in one place: augcl_b200.augmentations
    return getattr(_impl, name)
