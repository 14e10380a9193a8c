#!/usr/bin/env python
"""Experiment: a batch of views on one stream against the same batch spread over L streams
(view v on stream v mod L; the gradients are summed into one buffer by RED either way).

    python tools/exp_two_streams.py [--lanes 2] [--views 8] [--steps 10]
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
    ap.add_argument("--lanes", type=int, default=2)
    ap.add_argument("--views", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--workload", default="c2_1m_1920x1088")
    args = ap.parse_args()
    n, w, h, sf, seed = bench.WORKLOADS[args.workload]
    dev = torch.device("cuda:0")
    scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
    t = lambda a: torch.from_numpy(a).to(dev)
    xyz = torch.nn.Parameter(t(scene.point_cloud))
    feat = torch.nn.Parameter(t(scene.point_cloud_features))
    inv, oid = t(scene.point_invalid_mask), t(scene.point_object_id)
    cam = CameraInfo(t(scene.camera_intrinsics), h, w, 0)
    q, tr = t(scene.q_pointcloud_camera), t(scene.t_pointcloud_camera)
    wt = t(np.random.default_rng(3).uniform(0, 1, (h, w, 3)).astype(np.float32))
    ras = Ras(Ras.GaussianPointCloudRasterisationConfig())
    flat = torch.zeros(n * 59, device=dev)
    gx, gf = flat[:3 * n].view(n, 3), flat[3 * n:].view(n, 56)

    def one_view():
        image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(
            point_cloud=xyz, point_cloud_features=feat, point_object_id=oid, point_invalid_mask=inv,
            camera_info=cam, q_pointcloud_camera=q, t_pointcloud_camera=tr, color_max_sh_band=3))
        (image * wt).sum().backward()

    def step(lanes):
        main_stream = torch.cuda.current_stream(dev)
        flat.zero_()
        ras.accumulate_gradients_into(gx, gf)
        if len(lanes) == 1:
            for _ in range(args.views):
                one_view()
        else:
            start = torch.cuda.Event()
            start.record(main_stream)
            for s in lanes:
                s.wait_event(start)
            for v in range(args.views):
                with torch.cuda.stream(lanes[v % len(lanes)]):
                    one_view()
            for s in lanes:
                e = torch.cuda.Event()
                e.record(s)
                main_stream.wait_event(e)
        ras.accumulate_gradients_into(None, None)

    def timed(lanes):
        for _ in range(3):
            step(lanes)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.steps):
            step(lanes)
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / (args.steps * args.views)

    ms1 = timed([torch.cuda.current_stream(dev)])
    ref = flat.clone()
    msl = timed([torch.cuda.Stream(dev) for _ in range(args.lanes)])
    rel = float((flat - ref).norm() / ref.norm())
    print(f"1 stream: {ms1:.4f} ms/view ({1e3 / ms1:.1f} it/s)   {args.lanes} streams: {msl:.4f} ms/view "
          f"({1e3 / msl:.1f} it/s)   gradient rel diff {rel:.2e}")


if __name__ == "__main__":
    main()
