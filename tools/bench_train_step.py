#!/usr/bin/env python
"""One full training iteration at a bench workload (default C2: 1M points, 1920x1088), as the
reference trainer runs it (GaussianPointTrainer.py:146-184): zero_grad, rasterise, clamp + permute,
loss (L1 + SSIM + scale regulariser), backward (with the densification hook), two Adam steps.

Two arms on the same GPU, same rasteriser:
  fused : LossFunction.forward_rasterised + FusedAdam + AdaptiveControllerAccumulators (this repo)
  torch : the torch calls the reference makes for those three steps (clamp/permute, conv2d SSIM with
          autograd, torch.optim.Adam, index-add statistics)
Prints one JSON line (ms per iteration for each arm).
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench  # noqa: E402
from bench_loss import WINDOW, torch_step  # noqa: E402
from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo  # noqa: E402
from taichi_3d_gaussian_splatting_b200.GaussianPointAdaptiveController import AdaptiveControllerAccumulators  # noqa: E402
from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation as Ras  # noqa: E402
from taichi_3d_gaussian_splatting_b200.LossFunction import LossFunction  # noqa: E402
from taichi_3d_gaussian_splatting_b200.optim import FusedAdam  # noqa: E402
from taichi_3d_gaussian_splatting_b200.synthetic import make_scene  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="c2_1m_1920x1088", choices=list(bench.WORKLOADS))
    ap.add_argument("--arms", default="fused,torch")
    args = ap.parse_args()
    n, w, h, sf, seed = bench.WORKLOADS[args.workload]
    dev = torch.device("cuda:0")
    scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
    t = lambda a: torch.from_numpy(a).to(dev)
    inv, oid = t(scene.point_invalid_mask), t(scene.point_object_id)
    cam = CameraInfo(t(scene.camera_intrinsics), h, w, 0)
    q, tr = t(scene.q_pointcloud_camera), t(scene.t_pointcloud_camera)
    gt = torch.from_numpy(np.random.default_rng(3).uniform(0, 1, (3, h, w)).astype(np.float32)).to(dev)
    win = torch.tensor(WINDOW, device=dev).repeat(3, 1, 1, 1)
    out = {"what": "training iteration", "workload": args.workload, "steps": args.steps}

    for arm in args.arms.split(","):
        xyz = torch.nn.Parameter(t(scene.point_cloud))
        feat = torch.nn.Parameter(t(scene.point_cloud_features))
        if arm == "fused":
            acc = AdaptiveControllerAccumulators(n, dev)
            hook = acc.update
            o_feat, o_xyz = FusedAdam([feat], lr=5e-3), FusedAdam([xyz], lr=1e-5)
            lf = LossFunction(LossFunction.LossFunctionConfig(), fold_regulariser_gradient=True)
        else:
            stats = {k: torch.zeros(n, device=dev) for k in "abcd"}
            stats["p"] = torch.zeros(n, 3, device=dev)

            def hook(hk):
                ids = hk.point_id_in_camera_list.long()
                stats["a"][ids] += 1
                stats["b"][ids] += hk.num_affected_pixels
                stats["c"][ids] += hk.magnitude_grad_viewspace
                avg = hk.magnitude_grad_viewspace / hk.num_affected_pixels
                avg[torch.isnan(avg)] = 0
                stats["d"][ids] += avg
                stats["p"][ids] += hk.grad_point_in_camera
                stats["a"][ids] += hk.grad_point_in_camera.norm(dim=1)
            o_feat = torch.optim.Adam([feat], lr=5e-3, betas=(0.9, 0.999))
            o_xyz = torch.optim.Adam([xyz], lr=1e-5, betas=(0.9, 0.999))
        ras = Ras(Ras.GaussianPointCloudRasterisationConfig(), backward_valid_point_hook=hook)

        def iteration():
            o_feat.zero_grad()
            o_xyz.zero_grad()
            image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(
                point_cloud=xyz, point_cloud_features=feat, point_object_id=oid, point_invalid_mask=inv,
                camera_info=cam, q_pointcloud_camera=q, t_pointcloud_camera=tr, color_max_sh_band=3))
            if arm == "fused":
                loss, _, _ = lf.forward_rasterised(image, gt, inv, feat)
                loss.backward()
            else:
                s = feat[inv == 0, 4:7]
                reg = torch.norm(torch.exp(s), dim=1).mean()
                pred = torch.clamp(image, min=0, max=1).permute(2, 0, 1).unsqueeze(0)
                y = gt.unsqueeze(0)
                import torch.nn.functional as F

                def gf(a):
                    o = F.conv2d(a, weight=win.transpose(2, -1), stride=1, padding=0, groups=3)
                    return F.conv2d(o, weight=win, stride=1, padding=0, groups=3)
                l1 = torch.abs(pred - y).mean()
                mu1, mu2 = gf(pred), gf(y)
                mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
                s1, s2, s12 = gf(pred * pred) - mu1_sq, gf(y * y) - mu2_sq, gf(pred * y) - mu1_mu2
                cs = (2 * s12 + 0.03 ** 2) / (s1 + s2 + 0.03 ** 2)
                sm = ((2 * mu1_mu2 + 0.01 ** 2) / (mu1_sq + mu2_sq + 0.01 ** 2)) * cs
                loss = 0.8 * l1 + 0.2 * (1 - torch.flatten(sm, 2).mean(-1).mean()) + 2 * reg
                loss.backward()
            o_feat.step()
            o_xyz.step()
            return loss

        for _ in range(args.warmup):
            iteration()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.steps):
            last = iteration()
        b.record()
        torch.cuda.synchronize()
        out[f"{arm}_ms_per_iteration"] = round(a.elapsed_time(b) / args.steps, 4)
        out[f"{arm}_loss"] = float(last)
        del xyz, feat, o_feat, o_xyz, ras
        torch.cuda.empty_cache()
    if "torch_ms_per_iteration" in out and "fused_ms_per_iteration" in out:
        out["speedup"] = round(out["torch_ms_per_iteration"] / out["fused_ms_per_iteration"], 2)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
