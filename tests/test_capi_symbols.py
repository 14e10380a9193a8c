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


def test_struct_layouts_match_the_header():
    """gs_view_t and gs_exchange_t as ctypes declares them have the size and field offsets the C compiler
    gives the header's structs (they cross the ABI by pointer / by value in kernel arguments)."""
    import subprocess
    import tempfile
    from taichi_3d_gaussian_splatting_b200 import _capi
    fields = {"gs_view_t": ["slot_of_point", "point_in_camera", "splat_grads", "intrinsics", "q_cam_pc",
                            "t_pc_cam", "reserved"],
              "gs_exchange_t": ["multicast_base", "local_base", "flags", "rank", "world", "chunks",
                                "reducer_ctas", "epoch", "reserved", "scratch"]}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "gs_b200.h"', 'int main(void) {']
    for name, fs in fields.items():
        lines.append(f'printf("{name} %zu", sizeof({name}));')
        for f in fs:
            lines.append(f'printf(" %zu", offsetof({name}, {f}));')
        lines.append('printf("\\n");')
    lines += ['return 0;', '}']
    with tempfile.TemporaryDirectory() as tmp:
        src, exe = os.path.join(tmp, "layout.c"), os.path.join(tmp, "layout")
        open(src, "w").write("\n".join(lines))
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), src, "-o", exe], check=True)
        out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split("\n")
    for line in out:
        if not line.strip():
            continue
        name, size, *offsets = line.split()
        ct = getattr(_capi, name)
        assert ctypes.sizeof(ct) == int(size), name
        assert [getattr(ct, f).offset for f in fields[name]] == [int(o) for o in offsets], name


def test_exchange_sizes_without_a_gpu():
    from taichi_3d_gaussian_splatting_b200 import _capi
    assert _capi.lib.gs_exchange_flag_words(4, 8) == 2 * 4 * 8
    assert _capi.lib.gs_exchange_scratch_bytes(4) == 4 * (2 + 2 * 4)
    assert _capi.lib.gs_exchange_flag_words(0, 8) < 0 and _capi.lib.gs_exchange_scratch_bytes(0) < 0
