"""TEST / BASELINE INFRASTRUCTURE ONLY — never imported by the product package.

Builds ``oracle/_ref/``: the reference's own Python modules for the hot path and its callers, COMPILED to bytecode
from the sources where they lie under /root/reference (``py_compile``; no source text is copied into the repo).

    python oracle/build_ref.py            # also run by __graft_entry__.build() when /root/reference exists

``oracle/_ref/`` is git-ignored (it stays out of history, like the built ``.so``) but NOT gpurun-ignored, so it
travels to the GPU box, where /root/reference does not exist.  There it serves as
  * the checker of record: ``oracle/ref_harness.py`` falls back to it, so ``bench.py --impl reference`` and the
    ``cpu_baseline`` leg time the reference's own functions (``kind: "reference"``), not the restatement;
  * the caller side of the executed drop-in test (tests/test_gpu_dropin.py): the unmodified algorithm classes
    (src/algorithms/*.py) run one ``update()`` each on top of augcl_b200's modules;
  * the "reference on the same GPU" column of bench.py (the reference's torch op sequence on CUDA).

Modules compiled (all under /root/reference/src): utils.py, augmentations.py, algorithms/*.py.
The files are "unchecked hash" CPython bytecode (the .pyc format) stored as ``<module>.bin`` -- snapshot tools tend to
drop ``*.pyc`` -- and imported through the meta-path finder in oracle/ref_harness.py (``RefFinder``); they are bound to
this image's CPython minor version (the GPU box runs the same image).
"""
from __future__ import annotations

import glob
import json
import os
import py_compile
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/src"
OUT = os.path.join(HERE, "_ref")
MODULES = ["utils.py", "augmentations.py"]     # + algorithms/*.py


def sources():
    rel = list(MODULES)
    rel += sorted(os.path.relpath(p, REF_SRC) for p in glob.glob(os.path.join(REF_SRC, "algorithms", "*.py")))
    return rel


def up_to_date() -> bool:
    stamp = os.path.join(OUT, "BUILD.json")
    if not os.path.isfile(stamp):
        return False
    try:
        meta = json.load(open(stamp))
    except Exception:
        return False
    if meta.get("python") != list(sys.version_info[:2]):
        return False
    for rel in sources():
        src = os.path.join(REF_SRC, rel)
        dst = os.path.join(OUT, rel[:-3] + ".bin")
        if not os.path.isfile(dst) or meta.get("mtimes", {}).get(rel) != os.path.getmtime(src):
            return False
    return True


def build(force: bool = False) -> str | None:
    """Compiles the reference modules into oracle/_ref/.  Returns the directory, or None when /root/reference is absent
    (GPU box: the prebuilt directory that travelled with the snapshot is used as it is)."""
    if not os.path.isdir(REF_SRC):
        return OUT if os.path.isdir(OUT) else None
    if not force and up_to_date():
        return OUT
    mtimes = {}
    for rel in sources():
        src = os.path.join(REF_SRC, rel)
        dst = os.path.join(OUT, rel[:-3] + ".bin")     # utils.py -> utils.bin (bytecode; see the module docstring)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        py_compile.compile(src, cfile=dst, dfile="<reference>/src/" + rel, doraise=True,
                           invalidation_mode=py_compile.PycInvalidationMode.UNCHECKED_HASH)
        mtimes[rel] = os.path.getmtime(src)
    json.dump({"python": list(sys.version_info[:2]), "from": REF_SRC, "modules": sources(), "mtimes": mtimes},
              open(os.path.join(OUT, "BUILD.json"), "w"), indent=1)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
