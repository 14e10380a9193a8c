"""Helpers shared by the parity tests: run the CUDA path stage by stage and name every buffer
like the oracle (oracle/gs_oracle.py) does, so the two dicts can be diffed key by key."""
from __future__ import annotations

import numpy as np
import torch


def scene_tensors(scene, device="cuda", requires_grad=False):
    t = dict(
        point_cloud=torch.from_numpy(scene.point_cloud).to(device),
        point_cloud_features=torch.from_numpy(scene.point_cloud_features.copy()).to(device),
        point_invalid_mask=torch.from_numpy(scene.point_invalid_mask).to(device),
        point_object_id=torch.from_numpy(scene.point_object_id).to(device),
        camera_intrinsics=torch.from_numpy(scene.camera_intrinsics).to(device),
        q_pointcloud_camera=torch.from_numpy(scene.q_pointcloud_camera).to(device),
        t_pointcloud_camera=torch.from_numpy(scene.t_pointcloud_camera).to(device),
    )
    if requires_grad:
        t["point_cloud"].requires_grad_(True)
        t["point_cloud_features"].requires_grad_(True)
    return t


def oracle_forward(oracle, scene, near=0.8, far=1000.0, scale=100.0, rgb_only=False):
    return oracle.forward(scene.point_cloud, scene.point_cloud_features, scene.point_invalid_mask,
                          scene.point_object_id, scene.camera_intrinsics,
                          scene.q_pointcloud_camera, scene.t_pointcloud_camera, scene.width,
                          scene.height, near, far, scale, rgb_only)


def tile_offsets_of(st) -> torch.Tensor:
    """accumulated_num_overlap_tiles (R.py:913-922) of a forward state: the kernel's own scan if it was
    asked for, else (inputs beyond the packed scan's capacity) the exclusive cumsum of the tile counts."""
    buf = st.buffers
    if buf.has("tile_offsets"):
        return buf.view("tile_offsets", st.m)
    counts = buf.view("num_tiles", st.m).to(torch.int64)
    return torch.cumsum(counts, 0) - counts


def literal_unsorted_keys(st, width):
    """The reference's unsorted keys / payload in emission order (R.py:131-172), from the literal
    emission entry point run on the forward's per-point buffers."""
    from taichi_3d_gaussian_splatting_b200 import _capi
    dev = st.payload.device
    keys = torch.empty(max(st.k, 1), dtype=torch.int64, device=dev)
    pay = torch.empty(max(st.k, 1), dtype=torch.int32, device=dev)
    if st.k > 0:
        buf = st.buffers
        offsets = tile_offsets_of(st).contiguous()
        _capi.check(_capi.lib.gs_emit_keys(buf.ptr("bbox"), buf.ptr("depth_key"),
                                           offsets.data_ptr(), st.m, st.k, width,
                                           keys.data_ptr(), pay.data_ptr(),
                                           torch.cuda.current_stream(dev).cuda_stream), "gs_emit_keys")
    return keys[:st.k], pay[:st.k]


def _scan_fits(scene) -> bool:
    """gs_project_cull_sh produces the tile-offset scan only while n x tiles-per-image < 2^35."""
    n = scene.point_cloud.shape[0]
    return n < (1 << 27) and n * (scene.width // 16) * (scene.height // 16) < (1 << 35)


def cuda_forward_state(scene, near=0.8, far=1000.0, scale=100.0):
    """The operator's forward state with every buffer left on the device (large cases)."""
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    t = scene_tensors(scene)
    cfg = impl.GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig(
        near_plane=near, far_plane=far, depth_to_sort_key_scale=scale)
    cam = CameraInfo(camera_intrinsics=t["camera_intrinsics"], camera_height=scene.height,
                     camera_width=scene.width, camera_id=0)
    st = impl._forward_impl(cfg, t["point_cloud"], t["point_cloud_features"], t["point_invalid_mask"],
                            t["point_object_id"], t["q_pointcloud_camera"], t["t_pointcloud_camera"], cam,
                            debug_outputs=True, offsets_output=_scan_fits(scene))
    torch.cuda.synchronize()
    return st


def cuda_forward_stages(scene, near=0.8, far=1000.0, scale=100.0, rgb_only=False, pipeline=None):
    """Run the product forward through its C ABI and return every stage as numpy arrays.
    pipeline: None (the module default), "binned", "depth_first" or "literal"."""
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation

    t = scene_tensors(scene)
    cfg = GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig(
        near_plane=near, far_plane=far, depth_to_sort_key_scale=scale, rgb_only=rgb_only)
    cam = CameraInfo(camera_intrinsics=t["camera_intrinsics"], camera_height=scene.height,
                     camera_width=scene.width, camera_id=0)
    saved = impl.KEY_PIPELINE
    if pipeline is not None:
        impl.KEY_PIPELINE = pipeline
    try:
        st = impl._forward_impl(cfg, t["point_cloud"], t["point_cloud_features"],
                                t["point_invalid_mask"], t["point_object_id"],
                                t["q_pointcloud_camera"], t["t_pointcloud_camera"], cam,
                                keep_unsorted=True, debug_outputs=True, offsets_output=_scan_fits(scene))
    finally:
        impl.KEY_PIPELINE = saved
    if st.unsorted_keys is None and st.k > 0:
        st.unsorted_keys, st.unsorted_payload = literal_unsorted_keys(st, scene.width)
    torch.cuda.synchronize()
    m, k, buf = st.m, st.k, st.buffers
    np_ = lambda x: x.detach().cpu().numpy()
    out = dict(
        point_id_in_camera_list=np_(buf.view("ids", m)),
        point_uv=np_(buf.view("point_uv", m)),
        point_in_camera=np_(buf.view("point_in_camera", m)),
        point_uv_conic_and_rescale=np_(buf.view("conic", m)),
        point_alpha_after_activation=np_(buf.view("alpha", m)),
        point_color=np_(buf.view("color", m)),
        point_radii=np_(buf.view("radii", m)),
        num_overlap_tiles=np_(buf.view("num_tiles", m)),
        accumulated_num_overlap_tiles=np_(tile_offsets_of(st)),
        records=np_(buf.view("records", m)),
        point_in_camera_sort_key=np_(st.sorted_keys()),
        point_offset_with_sort_key=np_(st.payload),
        tile_points_start=np_(st.tile_start),
        tile_points_end=np_(st.tile_end),
        rasterized_image=np_(st.image),
        rasterized_depth=np_(st.depth),
        pixel_accumulated_alpha=np_(st.acc_alpha),
        pixel_offset_of_last_effective_point=np_(st.last),
        pixel_valid_point_count=np_(st.count),
        q_camera_pointcloud=np_(st.q_cam),
        t_camera_pointcloud=np_(st.t_cam),
        features_out=np_(t["point_cloud_features"]),
    )
    if st.unsorted_keys is not None:
        out["unsorted_keys"] = np_(st.unsorted_keys)
        out["unsorted_payload"] = np_(st.unsorted_payload)
    else:
        out["unsorted_keys"] = np.zeros(0, np.int64)
        out["unsorted_payload"] = np.zeros(0, np.int32)
    out["_state"] = st
    out["_tensors"] = t
    return out


FORWARD_KEYS_EXACT = [
    "q_camera_pointcloud", "t_camera_pointcloud", "point_id_in_camera_list", "features_out",
    "point_uv", "point_in_camera", "point_uv_conic_and_rescale", "point_alpha_after_activation",
    "point_color", "point_radii", "num_overlap_tiles", "accumulated_num_overlap_tiles",
    "unsorted_keys", "unsorted_payload", "point_in_camera_sort_key", "point_offset_with_sort_key",
    "tile_points_start", "tile_points_end", "rasterized_image", "rasterized_depth",
    "pixel_accumulated_alpha", "pixel_offset_of_last_effective_point", "pixel_valid_point_count",
]


def mismatch_report(ref: dict, got: dict, keys=FORWARD_KEYS_EXACT) -> list:
    """Value-exact comparison (NaN == NaN, -0 == +0); returns one line per differing buffer."""
    lines = []
    for key in keys:
        a, b = np.asarray(ref[key]), np.asarray(got[key])
        if a.shape != b.shape:
            lines.append(f"{key}: shape {a.shape} vs {b.shape}")
            continue
        if a.size == 0:
            continue
        if np.issubdtype(a.dtype, np.floating):
            bad = ~((a == b) | (np.isnan(a) & np.isnan(b)))
        else:
            bad = a != b
        nb = int(bad.sum())
        if nb:
            idx = np.argwhere(bad)[0]
            if np.issubdtype(a.dtype, np.floating):
                err = float(np.nanmax(np.abs(a.astype(np.float64) - b.astype(np.float64))[bad]))
            else:
                err = float(np.abs(a.astype(np.int64) - b.astype(np.int64))[bad].max())
            lines.append(f"{key}: {nb}/{a.size} differ, first at {tuple(idx)}: "
                         f"oracle {a[tuple(idx)]!r} cuda {b[tuple(idx)]!r}, max abs {err:.3e}")
    return lines


def grad_error(ref: np.ndarray, got: np.ndarray):
    """(relative L2 error, max elementwise error normalised by the column's max magnitude)."""
    ref = np.asarray(ref, np.float64)
    got = np.asarray(got, np.float64)
    denom = np.linalg.norm(ref)
    rel_l2 = float(np.linalg.norm(ref - got) / denom) if denom > 0 else float(np.abs(got).max())
    r2 = ref.reshape(ref.shape[0], -1)
    g2 = got.reshape(got.shape[0], -1)
    col = np.abs(r2).max(axis=0)
    col[col == 0] = 1.0
    max_norm = float((np.abs(r2 - g2) / col).max()) if r2.size else 0.0
    return rel_l2, max_norm


def outlier_fraction(ref: np.ndarray, got: np.ndarray, tol: float) -> float:
    """Fraction of elements whose error, normalised by the column's max magnitude, exceeds tol."""
    r2 = np.asarray(ref, np.float64).reshape(ref.shape[0], -1)
    g2 = np.asarray(got, np.float64).reshape(got.shape[0], -1)
    col = np.abs(r2).max(axis=0)
    col[col == 0] = 1.0
    return float(((np.abs(r2 - g2) / col) > tol).mean()) if r2.size else 0.0


def row_relative_errors(ref: np.ndarray, got: np.ndarray, floor: float = 1e-3) -> np.ndarray:
    """Per-row relative error ||got_r - ref_r|| / ||ref_r|| over the rows whose reference norm exceeds
    `floor` x the median norm of the non-zero rows (rows below that are rounding noise of a few pairs).
    A row that is 100 % wrong shows up here whatever the magnitude of the other rows."""
    r2 = np.asarray(ref, np.float64).reshape(ref.shape[0], -1)
    g2 = np.asarray(got, np.float64).reshape(got.shape[0], -1)
    norms = np.linalg.norm(r2, axis=1)
    nz = norms[norms > 0]
    if nz.size == 0:
        return np.zeros(0)
    keep = norms > floor * np.median(nz)
    return np.linalg.norm(g2[keep] - r2[keep], axis=1) / norms[keep]


def row_error_summary(ref: np.ndarray, got: np.ndarray) -> dict:
    e = row_relative_errors(ref, got)
    if e.size == 0:
        return dict(rows=0, p50=0.0, p99=0.0, p999=0.0, max=0.0)
    return dict(rows=int(e.size), p50=float(np.percentile(e, 50)), p99=float(np.percentile(e, 99)),
                p999=float(np.percentile(e, 99.9)), max=float(e.max()))


def decision_band_log2(conic_and_rescale: np.ndarray, alpha: np.ndarray) -> np.ndarray:
    """The 12th float of a splat record as gs_project_cull_sh writes it (include/gs_b200.h, gs_blend_bwd):
    s = log2(255 peak / (1 + band)), band = min(3e-6 + 6e-7 tau (1 + rho) / (1 - rho), 0.02); 0 for records
    that are irregular or can never reach alpha = 1/255."""
    c = np.asarray(conic_and_rescale, np.float64)
    a, b, cc = c[:, 0], c[:, 1], c[:, 2]
    peak = c[:, 3] * np.asarray(alpha, np.float64)
    with np.errstate(all="ignore"):
        ok = (peak > 0) & (a > 0) & (cc > 0) & (a * cc > b * b) & np.isfinite(a + b + cc + peak)
        rho = np.minimum(np.sqrt(b * b / (a * cc)), 0.9999)
        tau = np.maximum(2.0 * np.log(255.0 * peak), 0.0)
        band = np.minimum(3e-6 + 6e-7 * tau * (1 + rho) / (1 - rho), 0.02)
        s = np.log2(255.0 * peak / (1.0 + band))
    return np.where(ok, s, 0.0).astype(np.float32)
