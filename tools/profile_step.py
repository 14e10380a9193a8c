#!/usr/bin/env python
"""A few forward+backward steps of the C2 workload (or --workload), for ncu / compute-sanitizer.

    python tools/profile_step.py [--steps 3] [--workload c2_1m_1920x1088]
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo  # noqa: E402
from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import \
    GaussianPointCloudRasterisation as Ras  # noqa: E402
from taichi_3d_gaussian_splatting_b200.synthetic import make_scene  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--workload", default="c2_1m_1920x1088", choices=list(bench.WORKLOADS))
    ap.add_argument("--batch", type=int, default=0,
                    help="after the single-view steps, one gradient batch of this many views (the step's "
                         "per-point pass: point_bwd_batch_kernel)")
    args = ap.parse_args()
    n, w, h, sf, seed = bench.WORKLOADS[args.workload]
    dev = torch.device("cuda:0")
    scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
    xyz = torch.nn.Parameter(torch.from_numpy(scene.point_cloud).to(dev))
    feat = torch.nn.Parameter(torch.from_numpy(scene.point_cloud_features).to(dev))
    inv = torch.from_numpy(scene.point_invalid_mask).to(dev)
    oid = torch.from_numpy(scene.point_object_id).to(dev)
    cam = CameraInfo(torch.from_numpy(scene.camera_intrinsics).to(dev), h, w, 0)
    q = torch.from_numpy(scene.q_pointcloud_camera).to(dev)
    t = torch.from_numpy(scene.t_pointcloud_camera).to(dev)
    wt = torch.from_numpy(np.random.default_rng(3).uniform(0, 1, (h, w, 3)).astype(np.float32)).to(dev)
    ras = Ras(Ras.GaussianPointCloudRasterisationConfig())
    for _ in range(args.steps):
        xyz.grad = feat.grad = None
        image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(
            point_cloud=xyz, point_cloud_features=feat, point_object_id=oid, point_invalid_mask=inv,
            camera_info=cam, q_pointcloud_camera=q, t_pointcloud_camera=t, color_max_sh_band=3))
        (image * wt).sum().backward()
    if args.batch > 0:
        gf = torch.empty(n, 56, device=dev)
        gx = torch.empty(n, 3, device=dev)
        qs, ts = bench.make_views(args.batch)
        ras.begin_gradient_batch(gx, gf)
        for v in range(args.batch):
            image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(
                point_cloud=xyz, point_cloud_features=feat, point_object_id=oid, point_invalid_mask=inv,
                camera_info=cam, q_pointcloud_camera=torch.from_numpy(qs[v:v + 1]).to(dev),
                t_pointcloud_camera=torch.from_numpy(ts[v:v + 1]).to(dev), color_max_sh_band=3))
            (image * wt).sum().backward()
        ras.end_gradient_batch()
    torch.cuda.synchronize()
    print("ok", ras.last_forward_info)


if __name__ == "__main__":
    main()
