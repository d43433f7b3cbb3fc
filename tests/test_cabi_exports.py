"""CPU: the C-ABI library builds, loads without a GPU, and exports every symbol include/augcl_b200.h declares;
the ctypes table binds exactly that set; argument validation that needs no device works."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "augcl_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(augcl_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as entry
    entry.build()
    from augcl_b200 import _lib
    return _lib


def test_header_symbols_are_exported_and_bound(lib):
    names = declared_symbols()
    assert len(names) >= 20
    h = ctypes.CDLL(lib.LIB_PATH)
    for n in names:
        assert hasattr(h, n), f"{n} declared in the header but not exported"
    assert sorted(lib.SIGNATURES) == names, "ctypes table and header disagree"


def test_version_and_option_validation(lib):
    assert "sm_100a" in lib.version()
    lib.set_option("staging", "generic")
    lib.set_option("staging", "tma")
    with pytest.raises(lib.AugclError) as e:
        lib.set_option("staging", "cpu")
    assert e.value.status == -1 and "staging" in str(e.value)
    with pytest.raises(lib.AugclError):
        lib.set_option("no_such_key", "1")
    assert lib.launch_count() >= 0


def test_argument_validation_without_device(lib):
    # bad arguments are rejected before any CUDA call
    with pytest.raises(lib.AugclError) as e:
        lib.call("augcl_random_conv", None, None, 1, 3, 8, 8, None, None)
    assert e.value.status == -1
    with pytest.raises(lib.AugclError):
        lib.call("augcl_ring_sample", None, None, 4, None, None, None, None, None, 0, None, None, 0, 0, None)
    buf = (ctypes.c_float * 84)()
    lib.call("augcl_drq_base_grid", 84, 4, buf)      # host helper: linspace(-1+1/92, 1-1/92, 92)[:84]
    import torch
    eps = 1.0 / 92
    want = torch.linspace(-1.0 + eps, 1.0 - eps, 92, dtype=torch.float32)[:84]
    assert torch.equal(torch.tensor(list(buf), dtype=torch.float32), want)


def test_fused_desc_abi_and_validation_without_device(lib):
    """The ctypes mirror of augcl_fused_desc has the C struct's size (struct_size guards the ABI), and the fused /
    de-duplicated entry points reject bad arguments before touching a device."""
    d = lib.FusedDesc()
    d.struct_size = ctypes.sizeof(lib.FusedDesc)
    h = lib.lib()
    assert h.augcl_ring_sample_fused(None, ctypes.byref(d), None) == -1
    assert "ring is NULL" in h.augcl_last_error().decode()          # the size check passed, the ring check failed
    d.struct_size -= 8
    assert h.augcl_ring_sample_fused(None, ctypes.byref(d), None) == -1
    assert "ABI mismatch" in h.augcl_last_error().decode()
    assert h.augcl_ring_sample_fused(None, None, None) == -1
    with pytest.raises(lib.AugclError):
        lib.call("augcl_ring_put_frames", None, 0, 1, None, None)
    with pytest.raises(lib.AugclError):
        lib.call("augcl_ring_put_refs", None, 0, 1, None, None, None, None, None)
    out = ctypes.c_void_p()
    with pytest.raises(lib.AugclError) as e:
        lib.call("augcl_ring_create_dedup", ctypes.byref(out), 0, 10, 3, 3, 84, 84, 6, 2)     # pool <= k
    assert e.value.status == -1 and "pool_frames" in str(e.value)
    for key in ("fused_stages", "fused_ctas_per_sm", "fused_cost_obs", "fused_cost_next"):
        lib.set_option(key, "0")
    lib.set_option("fused_fast_mask", "1")
    with pytest.raises(lib.AugclError):
        lib.set_option("fused_stages", "9")


def test_product_has_no_cpu_path_and_never_imports_oracle():
    import torch
    from augcl_b200 import augmentations as A
    with pytest.raises(AssertionError):
        A.random_shift(torch.zeros(1, 3, 8, 8))
    for f in ("_lib.py", "augmentations.py", "utils.py", "build.py", "__init__.py"):
        src = open(os.path.join(ROOT, "augcl_b200", f)).read()
        assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
    if not torch.cuda.is_available():
        from augcl_b200 import utils as U
        with pytest.raises(RuntimeError):
            U.ReplayBuffer((9, 84, 84), (6,), 10, 2)
