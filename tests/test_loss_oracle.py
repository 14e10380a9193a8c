"""Pins oracle/loss_oracle.py: the numpy restatement of the loss front-end against the same
algorithm written with the torch calls pytorch_msssim and the reference make
(F.conv2d with groups=C, torch.clamp, permute, abs().mean(), autograd), in fp64 on the CPU.

Reference: LossFunction.py:20-52, GaussianPointTrainer.py:172-182; pytorch_msssim ssim.py.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import loss_oracle as LO


def _torch_ssim(x, y):
    """pytorch_msssim.ssim(X, Y, data_range=1, size_average=True) on [1,C,H,W] tensors."""
    c = x.shape[1]
    win = torch.from_numpy(LO.gaussian_window()).to(x.dtype).repeat(c, 1, 1, 1)  # [C,1,1,11]

    def gf(a):
        out = a
        for i, s in enumerate(a.shape[2:]):
            assert s >= win.shape[-1]
            out = F.conv2d(out, weight=win.transpose(2 + i, -1), stride=1, padding=0, groups=c)
        return out

    c1, c2 = 0.01 ** 2, 0.03 ** 2
    mu1, mu2 = gf(x), gf(y)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = gf(x * x) - mu1_sq
    s2 = gf(y * y) - mu2_sq
    s12 = gf(x * y) - mu1_mu2
    cs_map = (2 * s12 + c2) / (s1 + s2 + c2)
    ssim_map = ((2 * mu1_mu2 + c1) / (mu1_sq + mu2_sq + c1)) * cs_map
    return torch.flatten(ssim_map, 2).mean(-1).mean()


def _torch_loss(image_hwc, gt_chw, lam, features=None, mask=None, weight=2.0):
    pred = torch.clamp(image_hwc, min=0, max=1).permute(2, 0, 1).unsqueeze(0)
    gt = gt_chw.unsqueeze(0)
    l1 = torch.abs(pred - gt).mean()
    ld = 1 - _torch_ssim(pred, gt)
    loss = (1 - lam) * l1 + lam * ld
    if features is not None:
        s = features[mask == 0, 4:7]
        loss = loss + weight * torch.norm(torch.exp(s), dim=1).mean()
    return loss, l1, ld


def _images(h, w, seed):
    rng = np.random.default_rng(seed)
    gt = rng.uniform(0, 1, (3, h, w))
    # a prediction that resembles the target, leaves [0,1] in places and hits 0 exactly in others
    img = gt.transpose(1, 2, 0) + rng.normal(0, 0.15, (h, w, 3))
    img[rng.uniform(size=(h, w, 3)) < 0.02] = 0.0
    return img.astype(np.float32), gt.astype(np.float32)


def test_window_matches_torch():
    coords = torch.arange(11, dtype=torch.float) - 11 // 2
    g = torch.exp(-(coords ** 2) / (2 * 1.5 ** 2))
    g /= g.sum()
    assert np.array_equal(LO.gaussian_window(), g.numpy())


@pytest.mark.parametrize("h,w,lam", [(16, 32, 0.2), (48, 32, 0.2), (32, 64, 0.7), (11, 11, 0.2)])
def test_image_loss_matches_torch_autograd(h, w, lam):
    img, gt = _images(h, w, 10 + h + w)
    o = LO.image_loss(img, gt, lam)
    t_img = torch.from_numpy(img).double().requires_grad_(True)
    loss, l1, ld = _torch_loss(t_img, torch.from_numpy(gt).double(), lam)
    loss.backward()
    assert abs(o["loss"] - loss.item()) < 1e-12
    assert abs(o["l1"] - l1.item()) < 1e-12
    assert abs(o["ld_ssim"] - ld.item()) < 1e-12
    g = t_img.grad.numpy()
    assert np.abs(o["grad_image_hwc"] - g).max() <= 1e-12 * max(1.0, np.abs(g).max())
    # the clamp blocks the gradient outside [0,1] and passes it on the boundary
    assert np.all(o["grad_image_hwc"][(img < 0) | (img > 1)] == 0)


def test_fp32_oracle_is_close_to_fp64():
    img, gt = _images(32, 48, 5)
    a, b = LO.image_loss(img, gt, 0.2), LO.image_loss(img, gt, 0.2, dtype=np.float32)
    assert abs(a["loss"] - b["loss"]) < 1e-6
    assert np.abs(a["grad_image_hwc"] - b["grad_image_hwc"]).max() < 1e-4 * np.abs(a["grad_image_hwc"]).max()


def test_total_loss_with_regulariser_matches_torch():
    img, gt = _images(16, 16, 3)
    rng = np.random.default_rng(4)
    feat = rng.normal(0, 1, (50, 56)).astype(np.float32)
    mask = (rng.uniform(size=50) < 0.2).astype(np.int8)
    o = LO.total_loss(img, gt, feat, mask, 0.2, True, 2.0)
    t_img = torch.from_numpy(img).double().requires_grad_(True)
    t_feat = torch.from_numpy(feat).double().requires_grad_(True)
    loss, _, _ = _torch_loss(t_img, torch.from_numpy(gt).double(), 0.2, t_feat, torch.from_numpy(mask), 2.0)
    loss.backward()
    assert abs(o["loss"] - loss.item()) < 1e-12
    assert np.abs(o["grad_scale"] - t_feat.grad.numpy()[:, 4:7]).max() < 1e-12
    assert np.abs(t_feat.grad.numpy()[:, :4]).max() == 0 and np.abs(t_feat.grad.numpy()[:, 7:]).max() == 0
