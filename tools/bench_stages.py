#!/usr/bin/env python
"""CUDA-event time of every forward stage of one view (single stream), median over calls: the quick A/B
for the small kernels (projection, point sort, binning) at a bench workload.

    [GS_B200_LIB_NAME=variant.so] python tools/bench_stages.py [--workload c2_1m_1920x1088] [--iters 30]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import util  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2_1m_1920x1088", choices=list(bench.WORKLOADS))
    ap.add_argument("--iters", type=int, default=30)
    args = ap.parse_args()
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    from taichi_3d_gaussian_splatting_b200.profiling import StageProfiler
    from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
    n, w, h, sf, seed = bench.WORKLOADS[args.workload]
    scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
    t = util.scene_tensors(scene)
    cfg = impl.GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig()
    cam = CameraInfo(camera_intrinsics=t["camera_intrinsics"], camera_height=h, camera_width=w, camera_id=0)
    prof = StageProfiler()
    per_stage = {}
    st = None
    for it in range(args.iters + 3):
        impl.STAGE_PROFILER = prof if it >= 3 else None
        prof.reset()
        st = impl._forward_impl(cfg, t["point_cloud"], t["point_cloud_features"], t["point_invalid_mask"],
                                t["point_object_id"], t["q_pointcloud_camera"], t["t_pointcloud_camera"], cam,
                                keep_cell_masks=True)
        if it >= 3:
            for name, rec in prof.summary().items():
                per_stage.setdefault(name, []).append(rec["ms"])
    impl.STAGE_PROFILER = None
    out = {"workload": args.workload, "m": st.m, "k": st.k, "lib": os.environ.get("GS_B200_LIB_NAME", "libgs_b200.so")}
    for name, ts in per_stage.items():
        out[name] = round(float(np.median(ts)), 4)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
