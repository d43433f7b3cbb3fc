"""Module drop-in for the reference tree: two shim modules named like the reference's own
(``augmentations``, ``utils``) that ``src/train.py`` and ``src/algorithms/*`` pick up instead of ``src/augmentations.py``
and ``src/utils.py`` when this directory comes first on ``sys.path``:

    PYTHONPATH=<repo>/augcl_b200/dropin:<reference>/src  python <reference>/src/train.py ...
    # or, inside a process:
    import augcl_b200.dropin as dropin; dropin.install("<reference>/src")

No reference file is edited.  ``augmentations`` is augcl_b200's module; ``utils`` is the reference's own ``utils`` (found
further down ``sys.path``: config / seeding / soft-update helpers stay the reference's) with ``ReplayBuffer`` and
``LazyFrames`` replaced by the device-resident ones.  tests/test_gpu_dropin.py runs one ``update()`` of every unmodified
algorithm class of the reference (src/algorithms/*.py) on top of this.
"""
from __future__ import annotations

import os
import sys

SHIM_DIR = os.path.dirname(os.path.abspath(__file__))


def install(reference_src=None) -> None:
    """Put the shim modules ahead of ``reference_src`` on sys.path (both are added if missing) and drop already-imported
    ``utils`` / ``augmentations`` / ``algorithms`` modules so the next import resolves through the shim.
    reference_src=None: the reference is provided by some other importer (sys.meta_path); only the shim is installed."""
    dirs = [SHIM_DIR] + ([os.path.abspath(reference_src)] if reference_src else [])
    for p in dirs:
        while p in sys.path:
            sys.path.remove(p)
    sys.path[:0] = dirs
    for name in [m for m in sys.modules if m in ("utils", "augmentations", "algorithms") or m.startswith("algorithms.")]:
        del sys.modules[name]


def uninstall() -> None:
    while SHIM_DIR in sys.path:
        sys.path.remove(SHIM_DIR)
    for name in [m for m in sys.modules if m in ("utils", "augmentations", "algorithms", "_augcl_reference_utils")
                 or m.startswith("algorithms.")]:
        del sys.modules[name]
