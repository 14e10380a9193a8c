#!/usr/bin/env python
"""Times the fused loss front-end (csrc/loss.cu) at the C2 frame size and, beside it, the same
step written with the torch calls the reference makes (clamp, permute, abs().mean(), the grouped
conv2d SSIM of pytorch_msssim, autograd backward) on the same GPU.

    python tools/bench_loss.py [--height 1088 --width 1920 --steps 50]

Prints one JSON line.  Algorithmic bytes per step: forward reads image + target
(24 B/px), writes 9 maps ((H-10)(W-10) x 36 B); backward reads the maps, image and target and
writes the gradient (12 B/px).
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from taichi_3d_gaussian_splatting_b200.LossFunction import LossFunction  # noqa: E402

WINDOW = [0.0010283803567290306, 0.0075987582094967365, 0.036000773310661316, 0.10936068743467331,
          0.21300552785396576, 0.26601171493530273, 0.21300552785396576, 0.10936068743467331,
          0.036000773310661316, 0.0075987582094967365, 0.0010283803567290306]


def torch_step(img, gt, lam, win):
    pred = torch.clamp(img, min=0, max=1).permute(2, 0, 1).unsqueeze(0)
    y = gt.unsqueeze(0)

    def gf(a):
        out = F.conv2d(a, weight=win.transpose(2, -1), stride=1, padding=0, groups=3)
        return F.conv2d(out, weight=win, stride=1, padding=0, groups=3)

    l1 = torch.abs(pred - y).mean()
    mu1, mu2 = gf(pred), gf(y)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1, s2, s12 = gf(pred * pred) - mu1_sq, gf(y * y) - mu2_sq, gf(pred * y) - mu1_mu2
    cs = (2 * s12 + 0.03 ** 2) / (s1 + s2 + 0.03 ** 2)
    ssim_map = ((2 * mu1_mu2 + 0.01 ** 2) / (mu1_sq + mu2_sq + 0.01 ** 2)) * cs
    loss = (1 - lam) * l1 + lam * (1 - torch.flatten(ssim_map, 2).mean(-1).mean())
    loss.backward()
    return loss


def timed(fn, steps, warmup):
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
    ap.add_argument("--height", type=int, default=1088)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    args = ap.parse_args()
    h, w = args.height, args.width
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(1)
    gt = torch.from_numpy(rng.uniform(0, 1, (3, h, w)).astype(np.float32)).to(dev)
    img = (gt.permute(1, 2, 0) + 0.1 * torch.randn(h, w, 3, device=dev)).contiguous().requires_grad_(True)
    lf = LossFunction(LossFunction.LossFunctionConfig(enable_regularization=False))
    win = torch.tensor(WINDOW, device=dev).repeat(3, 1, 1, 1)

    def fused():
        img.grad = None
        loss, _, _ = lf.forward_rasterised(img, gt)
        loss.backward()
        return loss

    def reference_calls():
        img.grad = None
        return torch_step(img, gt, 0.2, win)

    lf_val, ref_val = fused().item(), reference_calls().item()
    ms_fused = timed(fused, args.steps, args.warmup)
    ms_torch = timed(reference_calls, args.steps, args.warmup)
    px, opx = h * w, (h - 10) * (w - 10)
    bytes_fwd, bytes_bwd = 24 * px + 36 * opx, 36 * opx + 24 * px + 12 * px
    print(json.dumps({
        "what": "loss front-end fwd+bwd", "height": h, "width": w, "steps": args.steps,
        "fused_ms": round(ms_fused, 4), "torch_reference_calls_ms": round(ms_torch, 4),
        "speedup": round(ms_torch / ms_fused, 2), "loss_fused": lf_val, "loss_torch": ref_val,
        "algorithmic_bytes": bytes_fwd + bytes_bwd,
        "gb_s": round((bytes_fwd + bytes_bwd) / (ms_fused * 1e-3) / 1e9, 1)}))


if __name__ == "__main__":
    main()
