#!/usr/bin/env python
"""Generate golden fixtures by EXECUTING the reference's own Python source.

    python tests/golden/gen_golden.py            # writes tests/golden/ref_*.npz

Runs in the build container only (needs /root/reference).  Taichi cannot be installed there, so
the reference package is imported against tests/golden/ti_shim (a numpy-float32 stand-in for the
`taichi` module, see its docstring for the semantics it assumes) and the reference's
`GaussianPointCloudRasterisation._module_function.forward / backward` (reference
GaussianPointCloudRasterisation.py:831-1163) are called directly on CPU tensors with a stand-in
autograd context.  Every kernel call is traced, so the fixtures hold every intermediate buffer:
cull mask, per-point attributes, tile counts, unsorted and sorted keys, payload, tile ranges,
image / depth / alpha / last index / counts, the gradients and the hook payload.

Two deliberate pins where the reference leaves behaviour to its runtime:
  * `Tensor.sort()` (R.py:947) is called with stable=True: ties keep emission order, which is
    what the CUDA radix sort does in practice and what the oracle and the CUDA path fix;
  * `ti.exp` is the oracle's gso_expf (see ti_shim/taichi/__init__.py).

The fixtures are small (the emulation is pure Python): the committed .npz files are what the
tests read; this script is the record of how they were made.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(HERE, "ti_shim"))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, ROOT)

import taichi as ti  # noqa: E402  (the shim)
import taichi_3d_gaussian_splatting.GaussianPointCloudRasterisation as R  # noqa: E402
from taichi_3d_gaussian_splatting.Camera import CameraInfo  # noqa: E402
from taichi_3d_gaussian_splatting_b200.synthetic import make_scene  # noqa: E402

CASES = {
    # name: (make_scene kwargs, (near, far, key scale), colour band, rgb_only)
    "ref_a_300pts_48x32": (dict(num_points=300, width=48, height=32, seed=2001, scale_factor=0.06,
                                num_objects=2), (0.8, 1000.0, 100.0), 3, False),
    "ref_b_dense_120pts_32x32": (dict(num_points=120, width=32, height=32, seed=2002, scale_factor=0.25,
                                      invalid_fraction=0.05), (0.8, 1000.0, 10.0), 1, False),
    # one tile whose list is longer than one 256-entry group of the reference's blend loops
    "ref_c_deep_700pts_16x16": (dict(num_points=700, width=16, height=16, seed=2003, scale_factor=0.12,
                                     z_range=(2.0, 30.0)), (0.8, 1000.0, 100.0), 2, False),
}


class _Ctx:
    """What the autograd context offers the reference's forward / backward."""
    needs_input_grad = (True, True, False, False, False, False, False, False)

    def save_for_backward(self, *tensors):
        self.saved_tensors = tensors


def _stable_sort(self, *args, **kwargs):
    kwargs.setdefault("stable", True)
    return _ORIG_SORT(self, *args, **kwargs)


_ORIG_SORT = torch.Tensor.sort


def run_case(name, scene_kwargs, cfg, band, rgb_only):
    near, far, scale = cfg
    scene = make_scene(**scene_kwargs)
    hooked = {}
    module = R.GaussianPointCloudRasterisation(
        config=R.GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig(
            near_plane=near, far_plane=far, depth_to_sort_key_scale=scale, rgb_only=rgb_only),
        backward_valid_point_hook=lambda h: hooked.update(h=h))
    fn = module._module_function
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a).copy())
    xyz, feat = t(scene.point_cloud), t(scene.point_cloud_features)
    cam = CameraInfo(camera_intrinsics=t(scene.camera_intrinsics), camera_height=scene.height,
                     camera_width=scene.width, camera_id=0)
    ctx = _Ctx()
    ti.KERNEL_TRACE.clear()
    ti.TRACE["on"] = True
    torch.Tensor.sort = _stable_sort
    try:
        t0 = time.time()
        image, depth, count = fn.forward(ctx, xyz, feat, t(scene.point_invalid_mask), t(scene.point_object_id),
                                         t(scene.q_pointcloud_camera), t(scene.t_pointcloud_camera), cam, band)
        t_fwd = time.time() - t0
        g = np.random.default_rng(77).uniform(0, 1, (scene.height, scene.width, 3)).astype(np.float32)
        t0 = time.time()
        grads = fn.backward(ctx, t(g), None, None)
        t_bwd = time.time() - t0
    finally:
        torch.Tensor.sort = _ORIG_SORT
        ti.TRACE["on"] = False
    trace = {}
    for kname, args in ti.KERNEL_TRACE:
        trace.setdefault(kname, args)
    out = dict(
        # inputs (the quaternion columns of features_out were normalised in place, R.py:264-266)
        point_cloud=scene.point_cloud, point_cloud_features=scene.point_cloud_features,
        point_invalid_mask=scene.point_invalid_mask, point_object_id=scene.point_object_id,
        camera_intrinsics=scene.camera_intrinsics, q_pointcloud_camera=scene.q_pointcloud_camera,
        t_pointcloud_camera=scene.t_pointcloud_camera,
        width=np.int32(scene.width), height=np.int32(scene.height),
        near=np.float32(near), far=np.float32(far), key_scale=np.float32(scale),
        band=np.int32(band), rgb_only=np.int32(rgb_only), image_grad=g,
        features_out=feat.numpy(),
    )
    f = trace["filter_point_in_camera"]
    out["point_in_camera_mask"] = f["point_in_camera_mask"]
    out["q_camera_pointcloud"], out["t_camera_pointcloud"] = f["q_camera_pointcloud"], f["t_camera_pointcloud"]
    a = trace["generate_point_attributes_in_camera_plane"]
    for key in ("point_uv", "point_in_camera", "point_uv_conic_and_rescale",
                "point_alpha_after_activation", "point_color", "point_radii"):
        out[key] = a[key]
    out["point_id_in_camera_list"] = a["point_id_list"]
    out["num_overlap_tiles"] = trace["generate_num_overlap_tiles"]["num_overlap_tiles"]
    k = trace["generate_point_sort_key_by_num_overlap_tiles"]
    out["accumulated_num_overlap_tiles"] = k["accumulated_num_overlap_tiles"]
    out["unsorted_keys"], out["unsorted_payload"] = k["point_in_camera_sort_key"], k["point_offset_with_sort_key"]
    r = trace["find_tile_start_and_end"]
    out["point_in_camera_sort_key"] = r["point_in_camera_sort_key"]
    out["tile_points_start"], out["tile_points_end"] = r["tile_points_start"], r["tile_points_end"]
    b = trace["gaussian_point_rasterisation"]
    out["point_offset_with_sort_key"] = b["point_offset_with_sort_key"]
    for key in ("rasterized_image", "rasterized_depth", "pixel_accumulated_alpha",
                "pixel_offset_of_last_effective_point", "pixel_valid_point_count"):
        out[key] = b[key]
    assert np.array_equal(out["rasterized_image"], image.numpy())
    out["grad_pointcloud"], out["grad_pointcloud_features"] = grads[0].numpy(), grads[1].numpy()
    h = hooked["h"]
    for key in ("point_id_in_camera_list", "grad_point_in_camera", "grad_pointfeatures_in_camera",
                "grad_viewspace", "magnitude_grad_viewspace", "magnitude_grad_viewspace_on_image",
                "num_overlap_tiles", "num_affected_pixels", "point_depth", "point_uv_in_camera"):
        out["hook_" + key] = getattr(h, key).numpy()
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print(f"{name}: M={out['point_id_in_camera_list'].shape[0]} K={out['unsorted_keys'].shape[0]} "
          f"fwd {t_fwd:.1f}s bwd {t_bwd:.1f}s -> {path} ({os.path.getsize(path)} B)", flush=True)


if __name__ == "__main__":
    only = sys.argv[1:]
    for name, (kw, cfg, band, rgb_only) in CASES.items():
        if only and name not in only:
            continue
        run_case(name, kw, cfg, band, rgb_only)
