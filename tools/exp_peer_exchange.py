#!/usr/bin/env python
"""The fused per-point pass + exchange (gs_point_bwd_batch_exchange over NVLink multicast) against the pass
followed by NCCL's all_reduce: same gradient, and the time of either on the bench workload.

    timeout 300 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 \
        tools/exp_peer_exchange.py [--points 1000000] [--views 8] [--chunks 16] [--reducers 0]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=1_000_000)
    ap.add_argument("--width", type=int, default=976)
    ap.add_argument("--height", type=int, default=544)
    ap.add_argument("--views", type=int, default=8)
    ap.add_argument("--chunks", type=int, default=4)
    ap.add_argument("--reducers", type=int, default=0)
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import bench
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    from taichi_3d_gaussian_splatting_b200.distributed import PeerExchange
    from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
    Ras = impl.GaussianPointCloudRasterisation
    n, w, h = args.points, args.width, args.height
    scene = make_scene(n, w, h, seed=11, scale_factor=0.004)
    xyz = torch.nn.Parameter(torch.from_numpy(scene.point_cloud).to(dev))
    feat = torch.nn.Parameter(torch.from_numpy(scene.point_cloud_features).to(dev))
    inv = torch.from_numpy(scene.point_invalid_mask).to(dev)
    oid = torch.from_numpy(scene.point_object_id).to(dev)
    cam = CameraInfo(torch.from_numpy(scene.camera_intrinsics).to(dev), h, w, 0)
    q_np, t_np = bench.make_views_spread(args.views * world, seed=3)
    qs, ts = torch.from_numpy(q_np).to(dev), torch.from_numpy(t_np).to(dev)
    wt = torch.from_numpy(np.random.default_rng(3).uniform(0, 1, (h, w, 3)).astype(np.float32)).to(dev)
    ras = Ras(Ras.GaussianPointCloudRasterisationConfig())
    mine = list(range(rank, args.views * world, world))

    def views():
        for v in mine:
            image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(
                point_cloud=xyz, point_cloud_features=feat, point_object_id=oid, point_invalid_mask=inv,
                camera_info=cam, q_pointcloud_camera=qs[v:v + 1], t_pointcloud_camera=ts[v:v + 1],
                color_max_sh_band=3))
            (image * wt).sum().backward()

    out = {"world": world, "points": n, "views_per_rank": args.views}
    flat = torch.empty(59 * n, device=dev)
    gf, gx = flat[:56 * n].view(n, 56), flat[56 * n:].view(n, 3)

    def step_nccl():
        ras.begin_gradient_batch(gx, gf)
        views()
        ras.end_gradient_batch()
        dist.all_reduce(flat)

    peer = PeerExchange(n, dev, chunks=args.chunks, reducer_ctas=args.reducers)
    out["chunks"] = peer.chunks

    def step_fused():
        ras.begin_gradient_batch(peer.grad_xyz, peer.grad_features)
        views()
        ras.end_gradient_batch(exchange=peer)

    step_nccl()
    torch.cuda.synchronize()
    want = flat.clone()
    step_fused()
    torch.cuda.synchronize()
    got = peer.flat
    step_nccl()
    torch.cuda.synchronize()
    again = flat.clone()

    def rows_off(a, b):
        a2 = torch.cat([a[:56 * n].view(n, 56), a[56 * n:].view(n, 3)], 1).double()
        b2 = torch.cat([b[:56 * n].view(n, 56), b[56 * n:].view(n, 3)], 1).double()
        err = (a2 - b2).norm(dim=1) / (b2.norm(dim=1) + 1e-3 * b2.norm(dim=1).median())
        return {"rows_over_1e-4": int((err > 1e-4).sum()), "max_row_err": float(err.max()),
                "worst_row": int(err.argmax())}
    out["rows_fused_vs_nccl"] = rows_off(got, want)
    out["rows_nccl_vs_nccl"] = rows_off(again, want)
    out["rel_l2_nccl_vs_nccl"] = float((again.double() - want.double()).norm() / want.double().norm())
    out["rel_l2"] = float((got.double() - want.double()).norm() / want.double().norm())
    out["finite"] = bool(torch.isfinite(got).all())
    out["exchange_error"] = peer.error()

    def timed(fn):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        ms = torch.tensor([a.elapsed_time(b) / args.iters], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    out["ms_step_nccl"] = timed(step_nccl)
    out["ms_step_fused"] = timed(step_fused)
    out["exchange_error_after"] = peer.error()

    # the tail alone: park once, then time only the pass (+ exchange)
    def tail(fused):
        def run():
            if fused:
                ras.begin_gradient_batch(peer.grad_xyz, peer.grad_features)
            else:
                ras.begin_gradient_batch(gx, gf)
            views()
            torch.cuda.synchronize()
            dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            if fused:
                ras.end_gradient_batch(exchange=peer)
            else:
                ras.end_gradient_batch()
                dist.all_reduce(flat)
            b.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b)
        run()
        ms = torch.tensor([float(np.median([run() for _ in range(5)]))], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    out["ms_tail_nccl"] = tail(False)
    out["ms_tail_fused"] = tail(True)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
