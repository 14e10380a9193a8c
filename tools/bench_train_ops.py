#!/usr/bin/env python
"""Times gs_adam_step (FusedAdam) over the C2 parameter set (N x 3 + N x 56 floats) against
torch.optim.Adam on the same GPU, and gs_controller_accumulate against the reference's torch
statements (GaussianPointAdaptiveController.py:130-146).  Prints one JSON line.

Algorithmic bytes: Adam 28 B per element (read p, g, m, v; write p, m, v).
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from taichi_3d_gaussian_splatting_b200.GaussianPointAdaptiveController import AdaptiveControllerAccumulators  # noqa: E402
from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation as R  # noqa: E402
from taichi_3d_gaussian_splatting_b200.optim import FusedAdam  # noqa: E402


def timed(fn, steps, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=1_000_000)
    ap.add_argument("--steps", type=int, default=50)
    args = ap.parse_args()
    n, dev = args.points, torch.device("cuda:0")
    out = {"what": "post-backward steps", "points": n}
    for name, make in (("fused", lambda ps, lr: FusedAdam(ps, lr=lr)),
                       ("torch", lambda ps, lr: torch.optim.Adam(ps, lr=lr, betas=(0.9, 0.999)))):
        xyz = torch.nn.Parameter(torch.randn(n, 3, device=dev))
        feat = torch.nn.Parameter(torch.randn(n, 56, device=dev))
        xyz.grad, feat.grad = torch.randn_like(xyz), torch.randn_like(feat)
        o1, o2 = make([feat], 5e-3), make([xyz], 1e-4)

        def step():
            o1.step()
            o2.step()
        out[f"adam_{name}_ms"] = round(timed(step, args.steps), 4)
        del xyz, feat, o1, o2
    out["adam_algorithmic_bytes"] = 28 * n * 59
    out["adam_fused_gb_s"] = round(28 * n * 59 / (out["adam_fused_ms"] * 1e-3) / 1e9, 1)

    m = int(0.92 * n)
    ids = torch.sort(torch.randperm(n, device=dev)[:m]).values.to(torch.int32)
    px = torch.randint(0, 100, (m,), device=dev, dtype=torch.int32)
    mag = torch.rand(m, device=dev)
    gx = torch.randn(m, 3, device=dev)
    fields = {f: None for f in R.BackwardValidPointHookInput.__dataclass_fields__}
    fields.update(point_id_in_camera_list=ids, num_affected_pixels=px, magnitude_grad_viewspace=mag,
                  grad_point_in_camera=gx)
    h = R.BackwardValidPointHookInput(**fields)
    acc = AdaptiveControllerAccumulators(n, dev)
    out["controller_fused_ms"] = round(timed(lambda: acc.update(h), args.steps), 4)
    t = {k: torch.zeros_like(getattr(acc, k)) for k, _, _ in acc.FIELDS}
    idl = ids.long()

    def torch_update():
        t["accumulated_num_in_camera"][idl] += 1
        t["accumulated_num_pixels"][idl] += px
        t["accumulated_view_space_position_gradients"][idl] += mag
        avg = mag / px
        avg[torch.isnan(avg)] = 0
        t["accumulated_view_space_position_gradients_avg"][idl] += avg
        t["accumulated_position_gradients"][idl] += gx
        t["accumulated_position_gradients_norm"][idl] += gx.norm(dim=1)
    out["controller_torch_ms"] = round(timed(torch_update, args.steps), 4)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
