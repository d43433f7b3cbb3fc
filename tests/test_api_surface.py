"""CPU: augcl_b200 keeps the reference's call surface for the path.  tests/golden/api_surface.json
was recorded from the reference's own modules by oracle/make_api_golden.py (parameter names, order, defaults, the
aug_to_func table); every reference parameter must exist here with the same name, position and default — extra
parameters are allowed only as keyword-only additions (the supplied draws)."""
import inspect
import json
import os

import pytest

from conftest import GOLDEN

SURFACE = json.load(open(os.path.join(GOLDEN, "api_surface.json")))


def check(ref_params, fn, who):
    ours = list(inspect.signature(fn).parameters.values())
    for pos, rp in enumerate(ref_params):
        assert pos < len(ours), f"{who}: parameter {rp['name']} missing"
        op = ours[pos]
        assert op.name == rp["name"], f"{who}: parameter {pos} is {op.name!r}, the reference calls it {rp['name']!r}"
        assert op.kind.name == rp["kind"], f"{who}.{op.name}: kind {op.kind.name} != {rp['kind']}"
        od = None if op.default is inspect.Parameter.empty else repr(op.default)
        assert od == rp["default"], f"{who}.{op.name}: default {od} != reference {rp['default']}"
    for op in ours[len(ref_params):]:
        assert op.kind in (inspect.Parameter.KEYWORD_ONLY, inspect.Parameter.VAR_KEYWORD), \
            f"{who}: extra parameter {op.name!r} must be keyword-only"


@pytest.mark.parametrize("name", sorted(SURFACE["augmentations"]))
def test_augmentation_signatures(name):
    from augcl_b200 import augmentations as A
    assert hasattr(A, name), f"augmentations.{name} is missing"
    check(SURFACE["augmentations"][name], getattr(A, name), f"augmentations.{name}")


def test_aug_to_func_table():
    from augcl_b200 import augmentations as A
    assert {k: v.__name__ for k, v in A.aug_to_func.items()} == SURFACE["aug_to_func"]


@pytest.mark.parametrize("name", sorted(SURFACE["ReplayBuffer"]))
def test_replay_buffer_signatures(name):
    from augcl_b200 import utils as U
    assert hasattr(U.ReplayBuffer, name), f"ReplayBuffer.{name} is missing"
    check(SURFACE["ReplayBuffer"][name], getattr(U.ReplayBuffer, name), f"ReplayBuffer.{name}")


def test_lazy_frames_surface():
    from augcl_b200 import utils as U
    for name, params in SURFACE["LazyFrames"].items():
        check(params, getattr(U.LazyFrames, name), f"LazyFrames.{name}")
    import numpy as np
    fr = [np.full((3, 4, 4), i, dtype=np.uint8) for i in range(3)]
    lf = U.LazyFrames(fr)
    assert np.asarray(lf).shape == (9, 4, 4) and lf.count() == 3 and lf.frame(1).max() == 1 and len(lf) == 3
    assert U.LazyFrames.__module__ == "utils"          # chunks pickled here load under the reference
