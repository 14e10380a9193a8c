#!/usr/bin/env python
"""Experiment: how the forward's culling bits are distributed at C2 (per 8x8 quarter and list entry:
0 = not reached / culled, 1 or 2 = one 8x4 half reachable, 3 = both)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl  # noqa: E402
from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo  # noqa: E402
from taichi_3d_gaussian_splatting_b200.synthetic import make_scene  # noqa: E402
import util  # noqa: E402

n, w, h, sf, seed = bench.WORKLOADS["c2_1m_1920x1088"]
scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
t = util.scene_tensors(scene)
cfg = impl.GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig()
cam = CameraInfo(camera_intrinsics=t["camera_intrinsics"], camera_height=h, camera_width=w, camera_id=0)
real_empty = torch.empty


def zero_u8(*a, **k):
    return torch.zeros(*a, **k) if k.get("dtype") == torch.uint8 else real_empty(*a, **k)


torch.empty = zero_u8
st = impl._forward_impl(cfg, t["point_cloud"], t["point_cloud_features"], t["point_invalid_mask"],
                        t["point_object_id"], t["q_pointcloud_camera"], t["t_pointcloud_camera"], cam,
                        keep_cell_masks=True)
torch.empty = real_empty
torch.cuda.synchronize()
hist = torch.bincount(st.cell_masks.to(torch.int64), minlength=4).cpu().numpy()
tot = hist.sum()
print("K =", st.k, "bytes =", tot)
print({v: f"{hist[v] / tot:.3f}" for v in range(4)})
vis = hist[1] + hist[2] + hist[3]
print("of the reachable (quarter, entry) pairs: one half only", f"{(hist[1] + hist[2]) / vis:.3f}", "both", f"{hist[3] / vis:.3f}")
