"""The C-ABI library loads and exports every symbol include/gs_b200.h declares (no compute calls,
so this runs without a GPU), and the ctypes table binds exactly that set."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "gs_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gs_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported_and_bound():
    from taichi_3d_gaussian_splatting_b200 import _capi
    names = _declared()
    assert len(names) >= 12
    lib = ctypes.CDLL(_capi.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in gs_b200.h but not exported"
    assert sorted(_capi.SIGNATURES) == names


def test_abi_version_and_status_strings():
    from taichi_3d_gaussian_splatting_b200 import _capi
    assert _capi.lib.gs_abi_version() == _capi.ABI_VERSION
    assert _capi.lib.gs_status_string(0) == b"ok"
    assert b"invalid" in _capi.lib.gs_status_string(-1)
    assert _capi.lib.gs_project_workspace_bytes(1000) >= 8 * 4
    assert _capi.lib.gs_sort_workspace_bytes(10000) > 0


def test_library_is_sm100a_only():
    """No multi-arch dispatch: the fat binary holds sm_100a code only."""
    import shutil
    import subprocess
    from taichi_3d_gaussian_splatting_b200 import _capi
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        return
    out = subprocess.run([cuobjdump, "-lelf", _capi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs
