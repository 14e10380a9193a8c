"""ctypes/numpy front end of the CPU oracle (oracle/gs_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(taichi_3d_gaussian_splatting_b200) never imports this module.

`forward()` / `backward()` chain the per-stage C functions in exactly the order
of the reference's autograd Function (reference
taichi_3d_gaussian_splatting/GaussianPointCloudRasterisation.py:831-1023 and
:1026-1163) and return every intermediate buffer, so a test can diff any stage.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

GRAD_FACTORS = dict(q=1.0, s=0.5, alpha=20.0, color=5.0, high_order_color=1.0)  # R.py:782-786


def build(force: bool = False) -> None:
    """Compile libgs_oracle.so / libgs_oracle64.so with oracle/Makefile."""
    outs = [os.path.join(_HERE, n) for n in ("libgs_oracle.so", "libgs_oracle64.so")]
    src = os.path.join(_HERE, "gs_oracle.c")
    fresh = all(os.path.exists(o) and os.path.getmtime(o) >= os.path.getmtime(src) for o in outs)
    if fresh and not force:
        return
    subprocess.run(["make", "-C", _HERE, "-B" if force else "-s"], check=True,
                   stdout=subprocess.DEVNULL)


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleLib:
    """One build of the oracle: float (literal fp32 profile) or double."""

    def __init__(self, double: bool = False):
        name = "libgs_oracle64.so" if double else "libgs_oracle.so"
        path = os.path.join(_HERE, name)
        if not os.path.exists(path):
            build()
        self.lib = C.CDLL(path)
        self.double = double
        self.real = np.float64 if double else np.float32
        self.creal = C.c_double if double else C.c_float
        self.prefix = "gso64_" if double else "gso_"

    def fn(self, name, restype=None):
        f = getattr(self.lib, self.prefix + name)
        f.restype = restype
        return f

    def arr(self, a, dtype=None):
        return np.ascontiguousarray(a, dtype=self.real if dtype is None else dtype)

    # -- individual stages -------------------------------------------------
    def num_threads(self) -> int:
        return int(self.fn("num_threads", C.c_int)())

    def set_num_threads(self, threads: int) -> None:
        """OpenMP threads of the following calls (a launcher may have exported OMP_NUM_THREADS=1)."""
        self.fn("set_num_threads")(C.c_int(int(threads)))

    def set_tile_row_band(self, begin: int = -1, end: int = -1) -> None:
        """Restrict both blend loops to tile rows [begin, end) (bench.py's bounded CPU sample)."""
        self.fn("set_tile_row_band")(C.c_int(begin), C.c_int(end))

    def inverse_pose(self, q_pc_cam, t_pc_cam):
        q = self.arr(q_pc_cam).reshape(-1, 4)
        t = self.arr(t_pc_cam).reshape(-1, 3)
        qo, to = np.empty_like(q), np.empty_like(t)
        self.fn("inverse_pose")(_ptr(q), _ptr(t), C.c_int(q.shape[0]), _ptr(qo), _ptr(to))
        return qo, to

    def cull(self, xyz, invalid_mask, object_id, K, q_cam_pc, t_cam_pc, near, far, W, H):
        n = xyz.shape[0]
        mask = np.zeros(n, np.int8)
        self.fn("cull")(_ptr(xyz), _ptr(invalid_mask), _ptr(object_id), C.c_int64(n), _ptr(K),
                        _ptr(q_cam_pc), _ptr(t_cam_pc), self.creal(near), self.creal(far),
                        C.c_int(W), C.c_int(H), _ptr(mask))
        return mask

    def compact(self, mask):
        ids = np.empty(mask.shape[0], np.int32)
        m = self.fn("compact", C.c_int64)(_ptr(mask), C.c_int64(mask.shape[0]), _ptr(ids))
        return ids[:m].copy()

    def project(self, xyz, features, object_id, K, q_cam_pc, t_cam_pc, ids):
        m = ids.shape[0]
        out = dict(
            point_uv=np.empty((m, 2), self.real),
            point_in_camera=np.empty((m, 3), self.real),
            point_uv_conic_and_rescale=np.empty((m, 4), self.real),
            point_alpha_after_activation=np.empty(m, self.real),
            point_color=np.empty((m, 3), self.real),
            point_radii=np.empty(m, self.real),
        )
        self.fn("project")(_ptr(xyz), _ptr(features), _ptr(object_id), _ptr(K), _ptr(q_cam_pc),
                           _ptr(t_cam_pc), _ptr(ids), C.c_int64(m), _ptr(out["point_uv"]),
                           _ptr(out["point_in_camera"]), _ptr(out["point_uv_conic_and_rescale"]),
                           _ptr(out["point_alpha_after_activation"]), _ptr(out["point_color"]),
                           _ptr(out["point_radii"]))
        return out

    def tile_counts(self, point_uv, radii, W, H):
        m = radii.shape[0]
        counts = np.empty(m, np.int32)
        self.fn("tile_counts")(_ptr(point_uv), _ptr(radii), C.c_int64(m), C.c_int(W), C.c_int(H),
                               _ptr(counts))
        return counts

    def exclusive_scan(self, counts):
        offsets = np.empty(counts.shape[0], np.int64)
        total = self.fn("exclusive_scan", C.c_int64)(_ptr(counts), C.c_int64(counts.shape[0]),
                                                     _ptr(offsets))
        return offsets, int(total)

    def emit_keys(self, point_uv, point_in_camera, radii, offsets, total, W, H, scale):
        keys = np.empty(total, np.int64)
        payload = np.empty(total, np.int32)
        self.fn("emit_keys")(_ptr(point_uv), _ptr(point_in_camera), _ptr(radii), _ptr(offsets),
                             C.c_int64(radii.shape[0]), C.c_int(W), C.c_int(H), self.creal(scale),
                             _ptr(keys), _ptr(payload))
        return keys, payload

    def sort_pairs(self, keys, payload):
        keys = np.ascontiguousarray(keys, np.int64).copy()
        payload = np.ascontiguousarray(payload, np.int32).copy()
        self.fn("sort_pairs")(_ptr(keys), _ptr(payload), C.c_int64(keys.shape[0]))
        return keys, payload

    def tile_ranges(self, sorted_keys, num_tiles):
        start = np.zeros(num_tiles, np.int32)
        end = np.zeros(num_tiles, np.int32)
        keys = np.ascontiguousarray(sorted_keys, np.int64)
        self.fn("tile_ranges")(_ptr(keys), C.c_int64(keys.shape[0]), _ptr(start), _ptr(end))
        return start, end

    def blend_fwd(self, H, W, start, end, payload, point_uv, point_in_camera, conic, alpha, color,
                  rgb_only=False):
        image = np.zeros((H, W, 3), self.real)
        depth = np.zeros((H, W), self.real)
        acc_alpha = np.zeros((H, W), self.real)
        last = np.zeros((H, W), np.int32)
        count = np.zeros((H, W), np.int32)
        visited = C.c_int64(0)
        self.fn("blend_fwd")(C.c_int(H), C.c_int(W), _ptr(start), _ptr(end), _ptr(payload),
                             _ptr(point_uv), _ptr(point_in_camera), _ptr(conic), _ptr(alpha),
                             _ptr(color), _ptr(image), _ptr(depth), _ptr(acc_alpha), _ptr(last),
                             _ptr(count), C.c_int(int(rgb_only)), C.byref(visited))
        return dict(rasterized_image=image, rasterized_depth=depth,
                    pixel_accumulated_alpha=acc_alpha,
                    pixel_offset_of_last_effective_point=last, pixel_valid_point_count=count,
                    pairs_fwd=int(visited.value))

    def blend_bwd(self, H, W, n, start, end, payload, ids, image_grad, acc_alpha, last, point_uv,
                  conic, alpha, color):
        m = ids.shape[0]
        grad_uv = np.zeros((n, 2), np.float64)
        grad_logit = np.zeros(n, np.float64)
        magnitude = np.zeros(n, np.float64)
        cov_buf = np.zeros((m, 3), np.float64)
        color_buf = np.zeros((m, 3), np.float64)
        affected = np.zeros(m, np.int32)
        on_image = np.zeros((H, W, 2), self.real)
        visited = C.c_int64(0)
        self.fn("blend_bwd")(C.c_int(H), C.c_int(W), _ptr(start), _ptr(end), _ptr(payload),
                             _ptr(ids), _ptr(image_grad), _ptr(acc_alpha), _ptr(last),
                             _ptr(point_uv), _ptr(conic), _ptr(alpha), _ptr(color), _ptr(grad_uv),
                             _ptr(grad_logit), _ptr(magnitude), _ptr(cov_buf), _ptr(color_buf),
                             _ptr(affected), _ptr(on_image), C.byref(visited))
        return dict(grad_uv=grad_uv, grad_logit=grad_logit, magnitude_grad_viewspace=magnitude,
                    cov_buf=cov_buf, color_buf=color_buf, num_affected_pixels=affected,
                    magnitude_grad_viewspace_on_image=on_image, pairs_bwd=int(visited.value))

    def point_bwd(self, xyz, features, object_id, K, q_cam_pc, t_cam_pc, t_pc_cam, ids,
                  point_in_camera, grad_uv, cov_buf, color_buf, grad_xyz, grad_features):
        self.fn("point_bwd")(_ptr(xyz), _ptr(features), _ptr(object_id), _ptr(K), _ptr(q_cam_pc),
                             _ptr(t_cam_pc), _ptr(t_pc_cam), _ptr(ids), C.c_int64(ids.shape[0]),
                             _ptr(point_in_camera), _ptr(grad_uv), _ptr(cov_buf), _ptr(color_buf),
                             _ptr(grad_xyz), _ptr(grad_features))

    def grad_epilogue(self, grad_features, band, factors=GRAD_FACTORS):
        self.fn("grad_epilogue")(_ptr(grad_features), C.c_int64(grad_features.shape[0]),
                                 C.c_int(int(band)), self.creal(factors["q"]),
                                 self.creal(factors["s"]), self.creal(factors["alpha"]),
                                 self.creal(factors["color"]),
                                 self.creal(factors["high_order_color"]))

    # -- whole path ----------------------------------------------------------
    def forward(self, point_cloud, point_cloud_features, point_invalid_mask, point_object_id,
                camera_intrinsics, q_pointcloud_camera, t_pointcloud_camera, width, height,
                near_plane=0.8, far_plane=1000.0, depth_to_sort_key_scale=100.0,
                rgb_only=False) -> Dict[str, np.ndarray]:
        """Reference forward (R.py:831-1023).  `features_out` is the in-place-normalised copy."""
        assert width % 16 == 0 and height % 16 == 0  # R.py:1193-1194
        xyz = self.arr(point_cloud)
        feat = self.arr(point_cloud_features).copy()
        inv = np.ascontiguousarray(point_invalid_mask, np.int8)
        oid = np.ascontiguousarray(point_object_id, np.int32)
        K = self.arr(camera_intrinsics)
        q_pc = self.arr(q_pointcloud_camera).reshape(-1, 4)
        t_pc = self.arr(t_pointcloud_camera).reshape(-1, 3)
        q_cam, t_cam = self.inverse_pose(q_pc, t_pc)
        mask = self.cull(xyz, inv, oid, K, q_cam, t_cam, near_plane, far_plane, width, height)
        ids = self.compact(mask)
        out = self.project(xyz, feat, oid, K, q_cam, t_cam, ids)
        counts = self.tile_counts(out["point_uv"], out["point_radii"], width, height)
        offsets, total = self.exclusive_scan(counts)
        keys, payload = self.emit_keys(out["point_uv"], out["point_in_camera"], out["point_radii"],
                                       offsets, total, width, height, depth_to_sort_key_scale)
        skeys, spayload = self.sort_pairs(keys, payload)
        num_tiles = (width // 16) * (height // 16)
        start, end = self.tile_ranges(skeys, num_tiles)
        out.update(self.blend_fwd(height, width, start, end, spayload, out["point_uv"],
                                  out["point_in_camera"], out["point_uv_conic_and_rescale"],
                                  out["point_alpha_after_activation"], out["point_color"],
                                  rgb_only))
        out.update(point_in_camera_mask=mask, point_id_in_camera_list=ids,
                   num_overlap_tiles=counts, accumulated_num_overlap_tiles=offsets,
                   unsorted_keys=keys, unsorted_payload=payload,
                   point_in_camera_sort_key=skeys, point_offset_with_sort_key=spayload,
                   tile_points_start=start, tile_points_end=end,
                   q_camera_pointcloud=q_cam, t_camera_pointcloud=t_cam,
                   features_out=feat, xyz=xyz, object_id=oid, K=K, t_pointcloud_camera=t_pc,
                   width=width, height=height)
        return out

    def backward(self, fwd: Dict[str, np.ndarray], grad_rasterized_image, color_max_sh_band=2,
                 factors=GRAD_FACTORS) -> Dict[str, np.ndarray]:
        """Reference backward (R.py:1026-1163) on the buffers `forward` returned."""
        H, W = fwd["height"], fwd["width"]
        xyz, feat = fwd["xyz"], fwd["features_out"]
        n = xyz.shape[0]
        ids = fwd["point_id_in_camera_list"]
        g_img = self.arr(grad_rasterized_image)
        b = self.blend_bwd(H, W, n, fwd["tile_points_start"], fwd["tile_points_end"],
                           fwd["point_offset_with_sort_key"], ids, g_img,
                           fwd["pixel_accumulated_alpha"],
                           fwd["pixel_offset_of_last_effective_point"], fwd["point_uv"],
                           fwd["point_uv_conic_and_rescale"], fwd["point_alpha_after_activation"],
                           fwd["point_color"])
        grad_xyz = np.zeros((n, 3), self.real)
        grad_feat = np.zeros((n, 56), self.real)
        grad_feat[:, 7] = b["grad_logit"].astype(self.real)
        self.point_bwd(xyz, feat, fwd["object_id"], fwd["K"], fwd["q_camera_pointcloud"],
                       fwd["t_camera_pointcloud"], fwd["t_pointcloud_camera"], ids,
                       fwd["point_in_camera"], self.arr(b["grad_uv"]), self.arr(b["cov_buf"]),
                       self.arr(b["color_buf"]), grad_xyz, grad_feat)
        self.grad_epilogue(grad_feat, color_max_sh_band, factors)
        b.update(grad_pointcloud=grad_xyz, grad_pointcloud_features=grad_feat,
                 grad_viewspace=self.arr(b["grad_uv"]))
        return b


_LIBS: Dict[bool, OracleLib] = {}


def get(double: bool = False) -> OracleLib:
    if double not in _LIBS:
        _LIBS[double] = OracleLib(double)
    return _LIBS[double]
