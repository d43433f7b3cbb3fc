"""TEST INFRASTRUCTURE ONLY — records the call surface of the reference's ``augmentations`` and ``utils.ReplayBuffer``
(parameter names, order, defaults; the aug_to_func keys) into tests/golden/api_surface.json, by importing the
reference's own modules in the build container:

    python oracle/make_api_golden.py

tests/test_api_surface.py then checks, on any machine, that augcl_b200 keeps that surface."""
from __future__ import annotations

import inspect
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as R  # noqa: E402

AUG_FUNCS = ["random_shift", "random_crop", "random_conv", "random_overlay", "splice", "CS_splice", "random_grayscale",
             "random_flip", "random_rotation", "random_cutout_color", "color_jitter", "identity", "load_dataloader",
             "view_as_windows_cuda", "create_hsv_mask", "batch_from_obs", "prepare_pad_batch"]
RB_METHODS = ["__init__", "add", "sample", "sample_obs", "sample_curl", "sample_drq", "save", "requeue_save", "load",
              "tensor_buffer_samples", "_get_idxs"]


def sig(fn):
    out = []
    for name, p in inspect.signature(fn).parameters.items():
        d = None if p.default is inspect.Parameter.empty else repr(p.default)
        out.append({"name": name, "kind": p.kind.name, "default": d})
    return out


def main():
    aug, utils = R.load_reference()
    doc = {"augmentations": {f: sig(getattr(aug, f)) for f in AUG_FUNCS if hasattr(aug, f)},
           "aug_to_func": {k: v.__name__ for k, v in aug.aug_to_func.items()},
           "ReplayBuffer": {m: sig(getattr(utils.ReplayBuffer, m)) for m in RB_METHODS if hasattr(utils.ReplayBuffer, m)},
           "LazyFrames": {m: sig(getattr(utils.LazyFrames, m)) for m in ("__init__", "frame", "count")}}
    path = os.path.join(ROOT, "tests", "golden", "api_surface.json")
    json.dump(doc, open(path, "w"), indent=1, sort_keys=True)
    print("wrote", path, {k: len(v) for k, v in doc.items()})


if __name__ == "__main__":
    main()
