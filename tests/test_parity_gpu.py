"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on identical
seeded inputs.  Forward stages are compared value-exact (every buffer of the reference's
forward, R.py:831-1023); gradients at 1e-4 relative (atomic order differs, north_star)."""
import numpy as np
import pytest
import torch

from taichi_3d_gaussian_splatting_b200.synthetic import make_scene

import util

pytestmark = pytest.mark.gpu

GRAD_RTOL = 1e-4  # north_star: gradients within 1e-4 relative

SCENES = {
    # name: (make_scene kwargs, rasteriser config (near, far, key scale))
    "c1_1k_64x64": (dict(num_points=1000, width=64, height=64, seed=1001, scale_factor=0.05),
                    (0.8, 1000.0, 100.0)),
    "multi_object_96x64": (dict(num_points=3000, width=96, height=64, seed=1002, scale_factor=0.04,
                                num_objects=3), (0.8, 1000.0, 100.0)),
    "dense_deep_lists_32x32": (dict(num_points=6000, width=32, height=32, seed=1003,
                                    scale_factor=0.15), (0.8, 1000.0, 100.0)),
    "mid_20k_256x160": (dict(num_points=20000, width=256, height=160, seed=1004,
                             scale_factor=0.01), (0.8, 1000.0, 100.0)),
    "truck_cfg_100k_976x544": (dict(num_points=100000, width=976, height=544, seed=1005,
                                    scale_factor=0.004), (0.4, 2000.0, 10.0)),
}


# all three key pipelines must give the reference's sorted keys, payload order and tile ranges
CASES = [(name, "depth_first") for name in SCENES] + [(name, "binned") for name in SCENES] + \
        [(name, "literal") for name in ("c1_1k_64x64", "dense_deep_lists_32x32", "truck_cfg_100k_976x544")]


@pytest.fixture(scope="module", params=CASES, ids=lambda c: f"{c[0]}-{c[1]}")
def case(request, oracle32):
    name, pipeline = request.param
    kwargs, (near, far, scale) = SCENES[name]
    scene = make_scene(**kwargs)
    ref = util.oracle_forward(oracle32, scene, near, far, scale)
    got = util.cuda_forward_stages(scene, near, far, scale, pipeline=pipeline)
    assert got["_state"].info["key_pipeline"] == pipeline
    return dict(name=f"{name}-{pipeline}", scene=scene, ref=ref, got=got, cfg=(near, far, scale),
                pipeline=pipeline)


def test_forward_every_stage_exact(case):
    ref, got = case["ref"], case["got"]
    assert len(ref["point_id_in_camera_list"]) > 0 and len(ref["point_in_camera_sort_key"]) > 0
    report = util.mismatch_report(ref, got)
    assert not report, f"{case['name']}:\n" + "\n".join(report)


def test_sorted_keys_are_sorted_and_ranges_consistent(case):
    got = case["got"]
    keys = got["point_in_camera_sort_key"]
    assert np.all(keys[1:] >= keys[:-1])
    tiles = (keys >> 32).astype(np.int64)
    start, end = got["tile_points_start"], got["tile_points_end"]
    for t in np.unique(tiles):
        idx = np.nonzero(tiles == t)[0]
        assert start[t] == idx[0] and end[t] == idx[-1] + 1
    assert int((end - start).sum()) == keys.shape[0]


def _backward_oracles(case, oracle32, oracle64, band):
    scene, ref = case["scene"], case["ref"]
    g = np.random.default_rng(77).uniform(0, 1, (scene.height, scene.width, 3)).astype(np.float32)
    b32 = oracle32.backward(ref, g, band)
    ref64 = {k: (v.astype(np.float64) if isinstance(v, np.ndarray) and v.dtype == np.float32 else v)
             for k, v in ref.items()}
    b64 = oracle64.backward(ref64, g.astype(np.float64), band)
    return g, b32, b64


@pytest.mark.parametrize("band", [3, 1])
def test_backward_gradients(case, oracle32, oracle64, band):
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation
    scene = case["scene"]
    near, far, scale = case["cfg"]
    g, b32, b64 = _backward_oracles(case, oracle32, oracle64, band)
    t = util.scene_tensors(scene, requires_grad=True)
    payload = {}
    ras = GaussianPointCloudRasterisation(
        GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig(
            near_plane=near, far_plane=far, depth_to_sort_key_scale=scale),
        backward_valid_point_hook=lambda h: payload.update(h=h))
    cam = CameraInfo(t["camera_intrinsics"], scene.height, scene.width, 0)
    image, depth, count = ras(GaussianPointCloudRasterisation.GaussianPointCloudRasterisationInput(
        point_cloud=t["point_cloud"], point_cloud_features=t["point_cloud_features"],
        point_object_id=t["point_object_id"], point_invalid_mask=t["point_invalid_mask"],
        camera_info=cam, q_pointcloud_camera=t["q_pointcloud_camera"],
        t_pointcloud_camera=t["t_pointcloud_camera"], color_max_sh_band=band))
    (image * torch.from_numpy(g).cuda()).sum().backward()
    torch.cuda.synchronize()
    gx = t["point_cloud"].grad.cpu().numpy()
    gf = t["point_cloud_features"].grad.cpu().numpy()
    assert np.isfinite(gx).all() and np.isfinite(gf).all()

    lines = []
    for name, cuda, o32, o64 in (("grad_pointcloud", gx, b32["grad_pointcloud"], b64["grad_pointcloud"]),
                                 ("grad_pointcloud_features", gf, b32["grad_pointcloud_features"],
                                  b64["grad_pointcloud_features"])):
        l2, mx = util.grad_error(o64, cuda)
        l2o, mxo = util.grad_error(o64, o32)
        rows, rows_o = util.row_error_summary(o64, cuda), util.row_error_summary(o64, o32)
        lines.append(f"{name}: cuda vs fp64 oracle rel-l2 {l2:.2e} max-col-norm {mx:.2e} "
                     f"per-row p50/p99.9/max {rows['p50']:.1e}/{rows['p999']:.1e}/{rows['max']:.1e}; "
                     f"fp32 oracle vs fp64 {l2o:.2e} {mxo:.2e} per-row {rows_o['p50']:.1e}/"
                     f"{rows_o['p999']:.1e}/{rows_o['max']:.1e}")
        assert l2 <= GRAD_RTOL and mx <= GRAD_RTOL, "\n".join(lines)
        # every row on its own (a row that is small against its column's maximum must be right too):
        # 99.9 % of the rows within 1e-4 of their own norm; the rest are rows whose terms cancel, where
        # the reference's own fp32 arithmetic (the literal oracle) is no closer to the fp64 value
        assert rows["p999"] <= max(GRAD_RTOL, 2.0 * rows_o["p999"]), "\n".join(lines)
    print("\n".join(lines))
    # band mask: coefficients above the band have exactly zero gradient (R.py:1167-1182)
    keep = {0: 1, 1: 4, 2: 9, 3: 16}[band]
    for base in (8, 24, 40):
        assert np.all(gf[:, base + keep:base + 16] == 0)

    # hook payload (R.py:1127-1142): M rows, list order, same values as the reference buffers
    h = payload["h"]
    ids = case["ref"]["point_id_in_camera_list"]
    m = ids.shape[0]
    assert h.point_id_in_camera_list.shape == (m,) and h.grad_point_in_camera.shape == (m, 3)
    assert h.grad_pointfeatures_in_camera.shape == (m, 56) and h.grad_viewspace.shape == (m, 2)
    assert h.magnitude_grad_viewspace.shape == (m,) and h.num_overlap_tiles.shape == (m,)
    assert h.num_affected_pixels.shape == (m,) and h.point_depth.shape == (m,)
    assert h.point_uv_in_camera.shape == (m, 2)
    assert h.magnitude_grad_viewspace_on_image.shape == (scene.height, scene.width, 2)
    assert np.array_equal(h.point_id_in_camera_list.cpu().numpy(), ids)
    assert np.array_equal(h.num_overlap_tiles.cpu().numpy(), case["ref"]["num_overlap_tiles"])
    assert np.array_equal(h.point_uv_in_camera.cpu().numpy(), case["ref"]["point_uv"])
    assert np.array_equal(h.point_depth.cpu().numpy(), case["ref"]["point_in_camera"][:, 2])
    assert np.array_equal(h.grad_point_in_camera.cpu().numpy(), gx[ids])
    assert np.array_equal(h.grad_pointfeatures_in_camera.cpu().numpy(), gf[ids])
    # which pairs take part is the reference's fp32 decision (R.py:634): the backward re-decides pairs on
    # the alpha = 1/255 edge with the literal arithmetic, so the counts are EQUAL to the literal oracle's
    # (and to the fp64 arbiter's, which takes the same fp32 decisions)
    aff = h.num_affected_pixels.cpu().numpy()
    assert np.array_equal(aff, b32["num_affected_pixels"])
    assert np.array_equal(aff, b64["num_affected_pixels"])
    for name, cuda, o64 in (("grad_viewspace", h.grad_viewspace.cpu().numpy(), b64["grad_uv"][ids]),
                            ("magnitude", h.magnitude_grad_viewspace.cpu().numpy()[:, None],
                             b64["magnitude_grad_viewspace"][ids][:, None])):
        l2, mx = util.grad_error(o64, cuda)
        assert l2 <= GRAD_RTOL and mx <= GRAD_RTOL, f"{name}: {l2:.2e} {mx:.2e}"
    # the per-pixel map is a sum over one pixel's pairs only; with the same pairs taking part no
    # pixel is off by a whole pair any more
    cuda = h.magnitude_grad_viewspace_on_image.cpu().numpy().reshape(-1, 2)
    o64 = b64["magnitude_grad_viewspace_on_image"].reshape(-1, 2)
    l2, mx = util.grad_error(o64, cuda)
    assert l2 <= GRAD_RTOL and mx <= GRAD_RTOL, f"on_image: {l2:.2e} {mx:.2e}"
    assert util.outlier_fraction(o64, cuda, GRAD_RTOL) == 0.0


def test_forward_is_deterministic(case):
    near, far, scale = case["cfg"]
    again = util.cuda_forward_stages(case["scene"], near, far, scale, pipeline=case["pipeline"])
    assert not util.mismatch_report(case["got"], again)


def test_quaternions_normalised_in_place(case):
    """R.py:264-266: visible rows get unit quaternions, all other rows/columns are untouched."""
    scene, got = case["scene"], case["got"]
    ids = got["point_id_in_camera_list"]
    out = got["features_out"]
    assert np.allclose(np.linalg.norm(out[ids, :4], axis=1), 1.0, atol=1e-6)
    untouched = np.ones(scene.point_cloud.shape[0], bool)
    untouched[ids] = False
    assert np.array_equal(out[untouched], scene.point_cloud_features[untouched])
    assert np.array_equal(out[:, 4:], scene.point_cloud_features[:, 4:])
