"""``import utils`` of the reference (src/utils.py) with the replay buffer replaced (see __init__.py): every public name
of the reference's own ``utils`` module -- found as the next ``utils.py`` / ``utils.pyc`` after this directory on
``sys.path`` -- is re-exported unchanged; ``ReplayBuffer`` and ``LazyFrames`` (src/utils.py:136-322) are augcl_b200's."""
import importlib.machinery
import importlib.util
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))


def _load_reference_utils():
    for d in sys.path:
        d = os.path.abspath(d or ".")
        if d == _HERE:
            continue
        for fn, loader in (("utils.py", importlib.machinery.SourceFileLoader),
                           ("utils.pyc", importlib.machinery.SourcelessFileLoader)):
            path = os.path.join(d, fn)
            # the reference's module is recognised by what it defines, not by where it lies
            if os.path.isfile(path) and os.path.isfile(os.path.join(d, "augmentations" + fn[5:])):
                spec = importlib.util.spec_from_loader("_augcl_reference_utils", loader("_augcl_reference_utils", path))
                mod = importlib.util.module_from_spec(spec)
                sys.modules["_augcl_reference_utils"] = mod
                spec.loader.exec_module(mod)      # its `import augmentations` resolves to the shim next to this file
                return mod
    # not a plain directory of sources: ask the import system's other finders (e.g. a finder for a compiled reference)
    for finder in sys.meta_path:
        find = getattr(finder, "find_spec", None)
        spec = find("utils", None) if find else None
        if spec is None or spec.loader is None or not spec.origin or os.path.dirname(os.path.abspath(spec.origin)) == _HERE:
            continue
        loader = type(spec.loader)("_augcl_reference_utils", spec.origin)
        spec = importlib.util.spec_from_loader("_augcl_reference_utils", loader, origin=spec.origin)
        mod = importlib.util.module_from_spec(spec)
        sys.modules["_augcl_reference_utils"] = mod
        spec.loader.exec_module(mod)
        return mod
    raise ImportError("augcl_b200.dropin.utils: the reference's src/ directory (utils.py + augmentations.py) must come "
                      "after augcl_b200/dropin on sys.path")


_ref = sys.modules.get("_augcl_reference_utils") or _load_reference_utils()
globals().update({k: v for k, v in vars(_ref).items() if not k.startswith("__")})

from augcl_b200.utils import ReplayBuffer, LazyFrames   # noqa: E402,F401  (override the reference's classes)
