"""Fixtures produced by EXECUTING the reference's own Python source (tests/golden/gen_golden.py:
the reference package imported against a numpy-float32 Taichi stand-in, its autograd Function's
forward / backward called on CPU tensors, every kernel call traced) against

  * the CPU oracle -- every forward buffer value-exact, gradients and hook payload at 1e-5
    (the oracle sums in double, the reference's atomics in float);                 [CPU test]
  * the CUDA path through the C ABI -- every forward buffer value-exact, gradients at the
    north_star tolerance of 1e-4.                                                   [GPU test]

This is the pin of the oracle to the reference's source line by line (reference
GaussianPointCloudRasterisation.py:31-772, 831-1163, GaussianPoint3D.py, SphericalHarmonics.py,
utils.py:257-359, 396-432); what stays unpinned is Taichi's own code generation."""
import glob
import os

import numpy as np
import pytest

import util

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_*.npz")))


def _load(path):
    return {k: v for k, v in np.load(path).items()}


def _scene(fx):
    from taichi_3d_gaussian_splatting_b200.synthetic import SyntheticScene
    return SyntheticScene(fx["point_cloud"], fx["point_cloud_features"], fx["point_invalid_mask"],
                          fx["point_object_id"], fx["camera_intrinsics"], fx["q_pointcloud_camera"],
                          fx["t_pointcloud_camera"], int(fx["width"]), int(fx["height"]))


def test_fixtures_are_present():
    assert len(GOLDEN) >= 3


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_reproduces_the_reference_source(path, oracle32):
    fx = _load(path)
    scene = _scene(fx)
    f = util.oracle_forward(oracle32, scene, float(fx["near"]), float(fx["far"]), float(fx["key_scale"]),
                            bool(fx["rgb_only"]))
    assert np.array_equal(f["point_in_camera_mask"].astype(np.int8), fx["point_in_camera_mask"])
    report = util.mismatch_report(fx, f)
    assert not report, "\n".join(report)
    assert fx["point_in_camera_sort_key"].shape[0] > 100 and (fx["pixel_valid_point_count"] > 0).any()
    b = oracle32.backward(f, fx["image_grad"], int(fx["band"]))
    for key in ("grad_pointcloud", "grad_pointcloud_features"):
        l2, mx = util.grad_error(fx[key], b[key])
        assert l2 <= 1e-5 and mx <= 1e-5, (key, l2, mx)
    ids = fx["hook_point_id_in_camera_list"]
    assert np.array_equal(ids, f["point_id_in_camera_list"])
    assert np.array_equal(fx["hook_num_overlap_tiles"], f["num_overlap_tiles"])
    assert np.array_equal(fx["hook_num_affected_pixels"], b["num_affected_pixels"])
    assert np.array_equal(fx["hook_point_depth"], f["point_in_camera"][:, 2])
    assert np.array_equal(fx["hook_point_uv_in_camera"], f["point_uv"])
    for key, got in (("hook_grad_viewspace", b["grad_viewspace"][ids]),
                     ("hook_magnitude_grad_viewspace", b["magnitude_grad_viewspace"][ids]),
                     ("hook_magnitude_grad_viewspace_on_image", b["magnitude_grad_viewspace_on_image"]),
                     ("hook_grad_point_in_camera", b["grad_pointcloud"][ids]),
                     ("hook_grad_pointfeatures_in_camera", b["grad_pointcloud_features"][ids])):
        l2, mx = util.grad_error(fx[key], np.asarray(got))
        assert l2 <= 1e-5 and mx <= 1e-5, (key, l2, mx)


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["depth_first", "binned", "literal"])
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_cuda_path_reproduces_the_reference_source(path, pipeline):
    import torch
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation as Ras
    fx = _load(path)
    scene = _scene(fx)
    near, far, scale = float(fx["near"]), float(fx["far"]), float(fx["key_scale"])
    got = util.cuda_forward_stages(scene, near, far, scale, bool(fx["rgb_only"]), pipeline=pipeline)
    report = util.mismatch_report(fx, got)
    assert not report, "\n".join(report)
    # backward + hook through the public module
    payload = {}
    ras = Ras(Ras.GaussianPointCloudRasterisationConfig(near_plane=near, far_plane=far,
                                                        depth_to_sort_key_scale=scale),
              backward_valid_point_hook=lambda h: payload.update(h=h))
    t = util.scene_tensors(scene, requires_grad=True)
    cam = CameraInfo(t["camera_intrinsics"], scene.height, scene.width, 0)
    image, depth, count = ras(Ras.GaussianPointCloudRasterisationInput(
        point_cloud=t["point_cloud"], point_cloud_features=t["point_cloud_features"],
        point_object_id=t["point_object_id"], point_invalid_mask=t["point_invalid_mask"],
        camera_info=cam, q_pointcloud_camera=t["q_pointcloud_camera"],
        t_pointcloud_camera=t["t_pointcloud_camera"], color_max_sh_band=int(fx["band"])))
    assert np.array_equal(image.detach().cpu().numpy(), fx["rasterized_image"])
    assert np.array_equal(depth.detach().cpu().numpy(), fx["rasterized_depth"])
    assert np.array_equal(count.cpu().numpy(), fx["pixel_valid_point_count"])
    (image * torch.from_numpy(fx["image_grad"]).cuda()).sum().backward()
    for key, g in (("grad_pointcloud", t["point_cloud"].grad), ("grad_pointcloud_features",
                                                                 t["point_cloud_features"].grad)):
        l2, mx = util.grad_error(fx[key], g.cpu().numpy())
        assert l2 <= 1e-4 and mx <= 1e-4, (key, l2, mx)
    h = payload["h"]
    assert np.array_equal(h.point_id_in_camera_list.cpu().numpy(), fx["hook_point_id_in_camera_list"])
    assert np.array_equal(h.num_overlap_tiles.cpu().numpy(), fx["hook_num_overlap_tiles"])
    assert np.array_equal(h.num_affected_pixels.cpu().numpy(), fx["hook_num_affected_pixels"])
    assert np.array_equal(h.point_depth.cpu().numpy(), fx["hook_point_depth"])
    assert np.array_equal(h.point_uv_in_camera.cpu().numpy(), fx["hook_point_uv_in_camera"])
    for key, got_t in (("hook_grad_viewspace", h.grad_viewspace),
                       ("hook_magnitude_grad_viewspace", h.magnitude_grad_viewspace),
                       ("hook_magnitude_grad_viewspace_on_image", h.magnitude_grad_viewspace_on_image),
                       ("hook_grad_point_in_camera", h.grad_point_in_camera),
                       ("hook_grad_pointfeatures_in_camera", h.grad_pointfeatures_in_camera)):
        l2, mx = util.grad_error(fx[key], got_t.cpu().numpy())
        assert l2 <= 1e-4 and mx <= 2e-4, (key, l2, mx)
