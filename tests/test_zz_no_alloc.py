"""gs_project_cull_sh_static (include/gs_b200.h): the projection pass with its look-back scratch in a
file-scope device array of the library instead of a caller-allocated workspace.  No counterpart in the
reference; added for callers that cannot allocate.

Called through the C ABI (ctypes) and compared with gs_project_cull_sh on the same inputs: the two run the
same kernel, so every output must have the same BITS (the forward's tolerance in this repository is zero).
Two calls in a row on different scenes, the second smaller than the first, so that a call that read what
the previous one left in the shared array would show.  This file is named to be collected last.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

NEAR, FAR, KEY_SCALE = 0.8, 1000.0, 100.0
OUTPUTS = (  # name, columns, dtype; every per-point array is sized for n rows, the first M are defined
    ("ids", 1, torch.int32), ("point_uv", 2, torch.float32), ("point_in_camera", 3, torch.float32),
    ("conic", 4, torch.float32), ("alpha", 1, torch.float32), ("color", 3, torch.float32),
    ("radii", 1, torch.float32), ("num_tiles", 1, torch.int32), ("tile_offsets", 1, torch.int64),
    ("records", 12, torch.float32), ("bbox", 4, torch.uint16), ("depth_key", 1, torch.int32),
    ("slot_of_point", 1, torch.int32),   # the one array written for EVERY point (position or -1)
)


def _inputs(scene, dev):
    from taichi_3d_gaussian_splatting_b200 import _capi
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    q_pc, t_pc = t(scene.q_pointcloud_camera), t(scene.t_pointcloud_camera)
    q_cam, t_cam = torch.empty_like(q_pc), torch.empty_like(t_pc)
    _capi.check(_capi.lib.gs_inverse_pose(q_pc.data_ptr(), t_pc.data_ptr(), q_pc.shape[0], q_cam.data_ptr(),
                                          t_cam.data_ptr(), None), "gs_inverse_pose")
    return dict(xyz=t(scene.point_cloud), features=t(scene.point_cloud_features),
                invalid=t(scene.point_invalid_mask), object_id=t(scene.point_object_id),
                intrinsics=t(scene.camera_intrinsics), q_cam=q_cam, t_cam=t_cam, n_obj=int(q_pc.shape[0]))


def _outputs(n, dev):
    # poisoned, so that rows a call did not write cannot compare equal by accident of a fresh allocation
    out = {}
    for name, cols, dtype in OUTPUTS:
        raw = torch.full((max(n, 1) * cols * torch.empty((), dtype=dtype).element_size(),), 0xA5,
                         dtype=torch.uint8, device=dev)
        out[name] = raw.view(dtype)
    out["totals"] = torch.full((2,), -7, dtype=torch.int64, device=dev)
    return out


def _call(entry, scene, inp, features, out, workspace):
    """One call of either entry point on torch's current stream.  `features` is this call's own copy: the
    pass normalises the quaternion columns in place (R.py:264-266)."""
    from taichi_3d_gaussian_splatting_b200 import _capi
    args = [inp["xyz"].data_ptr(), features.data_ptr(), inp["invalid"].data_ptr(), inp["object_id"].data_ptr(),
            scene.point_cloud.shape[0], inp["intrinsics"].data_ptr(), inp["q_cam"].data_ptr(),
            inp["t_cam"].data_ptr(), inp["n_obj"], NEAR, FAR, scene.width, scene.height, KEY_SCALE]
    args += [out[name].data_ptr() for name, _, _ in OUTPUTS]
    args.append(out["totals"].data_ptr())
    if workspace is not None:
        args.append(workspace.data_ptr())
    args.append(torch.cuda.current_stream().cuda_stream)
    return getattr(_capi.lib, entry)(*args)


def _compare(scene, dev, check_memory):
    from taichi_3d_gaussian_splatting_b200 import _capi
    n = scene.point_cloud.shape[0]
    inp = _inputs(scene, dev)
    ref, got = _outputs(n, dev), _outputs(n, dev)
    ws = torch.empty(max(int(_capi.lib.gs_project_workspace_bytes(n)), 16), dtype=torch.uint8, device=dev)
    ref_features, got_features = inp["features"].clone(), inp["features"].clone()
    _capi.check(_call("gs_project_cull_sh", scene, inp, ref_features, ref, ws), "gs_project_cull_sh")
    torch.cuda.synchronize()
    # every tensor of this comparison exists by now: nothing between the two readings allocates through torch
    free_before = torch.cuda.mem_get_info()[0]
    _capi.check(_call("gs_project_cull_sh_static", scene, inp, got_features, got, None),
                "gs_project_cull_sh_static")
    torch.cuda.synchronize()
    if check_memory:  # the first call of a process may load the kernel's module lazily; later ones take nothing
        assert torch.cuda.mem_get_info()[0] == free_before
    m, k = (int(v) for v in ref["totals"].tolist())
    assert got["totals"].tolist() == [m, k]
    assert 0 < m < n and k > 0
    assert torch.equal(ref["slot_of_point"], got["slot_of_point"]) and int((ref["slot_of_point"] >= 0).sum()) == m
    for name, cols, _ in OUTPUTS:
        if name == "slot_of_point":
            continue
        a, b = ref[name][: m * cols].view(torch.uint8), got[name][: m * cols].view(torch.uint8)
        assert torch.equal(a, b), f"{name} differs between gs_project_cull_sh and its _static variant"
        # neither writes past the M rows it reports
        assert torch.equal(ref[name][m * cols:].view(torch.uint8), got[name][m * cols:].view(torch.uint8)), name
    assert torch.equal(ref_features.view(torch.int32), got_features.view(torch.int32))
    return m, k


def test_static_scratch_variant_equals_the_workspace_variant_on_two_calls_in_a_row():
    from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    first = make_scene(60_000, 976, 544, seed=4101, num_objects=3)
    second = make_scene(7_013, 320, 240, seed=4102)   # fewer blocks: most of the array keeps call 1's words
    m1, _ = _compare(first, dev, check_memory=False)
    m2, _ = _compare(second, dev, check_memory=True)
    m1_again, _ = _compare(first, dev, check_memory=True)        # and back up in size
    assert m1_again == m1 and m2 < m1


def test_static_scratch_variant_rejects_sizes_beyond_its_array():
    """n > GS_PROJECT_STATIC_MAX_POINTS (2^24) is refused before anything is read or launched."""
    from taichi_3d_gaussian_splatting_b200 import _capi
    limit = 1 << 24
    args = [None, None, None, None, limit + 1, None, None, None, 1, NEAR, FAR, 976, 544, KEY_SCALE]
    args += [None] * (len(OUTPUTS) + 1) + [None]
    status = _capi.lib.gs_project_cull_sh_static(*args)
    assert status == -3
    assert b"static" in _capi.lib.gs_status_string(status)
    with pytest.raises(_capi.GsError, match="fixed scratch"):
        _capi.check(status, "gs_project_cull_sh_static")
    torch.cuda.synchronize()
