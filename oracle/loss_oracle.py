"""CPU oracle of the loss front-end.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module; the
product path (taichi_3d_gaussian_splatting_b200/LossFunction.py) never does.

What it restates, in numpy (fp64 by default, fp32 on request):

  * the trainer's glue between the rasteriser and the loss, reference GaussianPointTrainer.py:172-175:
    `clamp(image, 0, 1)` then HWC -> CHW;
  * reference LossFunction.py:20-38: `L = (1 - lambda) * L1 + lambda * (1 - SSIM)` (+ the scale
    regularisation of LossFunction.py:40-52 when features are given);
  * `pytorch_msssim.ssim(X, Y, data_range=1, size_average=True)`, a THIRD-PARTY dependency that is
    not vendored under /root/reference and not pinned (requirements.txt:4 just names it).  Its
    published algorithm (pytorch_msssim 1.0.0, ssim.py: `_fspecial_gauss_1d`, `gaussian_filter`,
    `_ssim`, `ssim`) is restated here: 11-tap Gaussian window, sigma 1.5, normalised in fp32;
    separable VALID (unpadded) filtering, H first then W, of X, Y, X*X, Y*Y, X*Y; K1 = 0.01,
    K2 = 0.03; `ssim_map = (2 mu1 mu2 + C1)/(mu1^2 + mu2^2 + C1) * (2 s12 + C2)/(s1 + s2 + C2)`;
    mean over the (H-10)x(W-10) map per channel, then over channels.
    Parity of this part is pinned by tests/test_loss_oracle.py against the same algorithm written
    with torch.nn.functional.conv2d + autograd (the calls pytorch_msssim itself makes); the
    package is absent here, so parity with the package's own binaries is "unpinned".

The gradient is analytic (the adjoint of the valid filter is the full correlation of the three
derivative maps) and is what the CUDA kernels are compared with.
"""
from __future__ import annotations

import numpy as np

WIN_SIZE = 11
WIN_SIGMA = 1.5
K1, K2 = 0.01, 0.03


# `_fspecial_gauss_1d(11, 1.5)` as torch evaluates it in fp32 (exp, then division by the fp32 sum);
# tests/test_loss_oracle.py::test_window_matches_torch re-derives the table with torch.
_WINDOW_HEX = ("0x1.0d9570p-10", "0x1.f1fe02p-8", "0x1.26eb18p-5", "0x1.bff0fep-4", "0x1.b43c3ep-3",
               "0x1.106560p-2", "0x1.b43c3ep-3", "0x1.bff0fep-4", "0x1.26eb18p-5", "0x1.f1fe02p-8",
               "0x1.0d9570p-10")


def gaussian_window() -> np.ndarray:
    """The 11 fp32 taps of pytorch_msssim's window (sigma 1.5)."""
    return np.array([float.fromhex(v) for v in _WINDOW_HEX], dtype=np.float32)


def _filter_valid(a: np.ndarray, w: np.ndarray) -> np.ndarray:
    """[C,H,W] -> [C,H-10,W-10]: `gaussian_filter`, along H first, then along W, no padding."""
    n = w.shape[0]
    h = a.shape[1] - n + 1
    out = np.zeros((a.shape[0], h, a.shape[2]), a.dtype)
    for k in range(n):
        out += w[k] * a[:, k:k + h, :]
    wd = a.shape[2] - n + 1
    out2 = np.zeros((a.shape[0], h, wd), a.dtype)
    for k in range(n):
        out2 += w[k] * out[:, :, k:k + wd]
    return out2


def _filter_adjoint(m: np.ndarray, w: np.ndarray, height: int, width: int) -> np.ndarray:
    """Adjoint of _filter_valid: [C,H-10,W-10] -> [C,H,W]."""
    n = w.shape[0]
    tmp = np.zeros((m.shape[0], m.shape[1], width), m.dtype)
    for k in range(n):
        tmp[:, :, k:k + m.shape[2]] += w[k] * m
    out = np.zeros((m.shape[0], height, width), m.dtype)
    for k in range(n):
        out[:, k:k + m.shape[1], :] += w[k] * tmp
    return out


def ssim_maps(x: np.ndarray, y: np.ndarray, dtype=np.float64):
    """ssim_map and its three derivative maps.

    Returns (ssim_map, d_mu, d_xx, d_xy): the derivative of ssim_map with respect to the filtered
    X (holding the filtered X*X and X*Y fixed), the filtered X*X and the filtered X*Y.
    """
    w = gaussian_window().astype(dtype)
    x = x.astype(dtype)
    y = y.astype(dtype)
    c1, c2 = dtype(K1 ** 2), dtype(K2 ** 2)
    mu1, mu2 = _filter_valid(x, w), _filter_valid(y, w)
    s1 = _filter_valid(x * x, w) - mu1 * mu1
    s2 = _filter_valid(y * y, w) - mu2 * mu2
    s12 = _filter_valid(x * y, w) - mu1 * mu2
    a1, a2 = 2 * mu1 * mu2 + c1, 2 * s12 + c2
    b1, b2 = mu1 * mu1 + mu2 * mu2 + c1, s1 + s2 + c2
    ssim_map = (a1 / b1) * (a2 / b2)
    d_s1 = -(a1 * a2) / (b1 * b2 * b2)
    d_s12 = 2 * a1 / (b1 * b2)
    d_mu1 = 2 * mu2 * a2 / (b1 * b2) - 2 * mu1 * a1 * a2 / (b1 * b1 * b2)
    d_mu = d_mu1 - 2 * mu1 * d_s1 - mu2 * d_s12
    return ssim_map, d_mu, d_s1, d_s12


def image_loss(image_hwc: np.ndarray, gt_chw: np.ndarray, lambda_value: float = 0.2,
               dtype=np.float64):
    """Trainer glue + LossFunction.forward without the regulariser.

    image_hwc: the rasteriser's H x W x 3 output, NOT clamped.  gt_chw: 3 x H x W.
    Returns dict(loss, l1, ld_ssim, grad_image_hwc) with the gradient of `loss` with respect to
    image_hwc (clamp included: it passes where 0 <= value <= 1, as torch.clamp's does).
    """
    raw = image_hwc.astype(dtype).transpose(2, 0, 1)
    x = np.clip(raw, 0, 1)
    y = gt_chw.astype(dtype)
    c, h, wd = x.shape
    lam = dtype(lambda_value)
    diff = x - y
    l1 = np.abs(diff).mean(dtype=dtype)
    ssim_map, d_mu, d_xx, d_xy = ssim_maps(x, y, dtype)
    ssim = ssim_map.mean(dtype=dtype)
    ld = 1 - ssim
    loss = (1 - lam) * l1 + lam * ld
    w = gaussian_window().astype(dtype)
    g_map = -lam / dtype(ssim_map.size)
    gx = g_map * (_filter_adjoint(d_mu, w, h, wd) + 2 * x * _filter_adjoint(d_xx, w, h, wd)
                  + y * _filter_adjoint(d_xy, w, h, wd))
    gx += (1 - lam) / dtype(x.size) * np.sign(diff)
    gx *= ((raw >= 0) & (raw <= 1))
    return dict(loss=loss, l1=l1, ld_ssim=ld, ssim=ssim,
                grad_image_hwc=np.ascontiguousarray(gx.transpose(1, 2, 0)))


def scale_regularisation(features: np.ndarray, invalid_mask: np.ndarray, dtype=np.float64):
    """LossFunction._regularization_loss (LossFunction.py:40-52): mean over valid points of
    ||exp(s)||_2 with s = features[:, 4:7].  Returns (value, gradient [N, 3] for columns 4:7)."""
    valid = invalid_mask == 0
    e = np.exp(features[:, 4:7].astype(dtype))
    norm = np.sqrt((e * e).sum(axis=1))
    count = int(valid.sum())
    value = norm[valid].mean(dtype=dtype) if count else dtype(np.nan)
    grad = np.zeros((features.shape[0], 3), dtype)
    if count:
        grad[valid] = (e[valid] * e[valid]) / norm[valid, None] / count
    return value, grad


def total_loss(image_hwc, gt_chw, features=None, invalid_mask=None, lambda_value=0.2,
               enable_regularization=True, regularization_weight=2.0, dtype=np.float64):
    """LossFunction.forward (LossFunction.py:20-38) end to end; returns the image_loss dict plus
    `reg` and `grad_scale` ([N,3], already multiplied by the weight) when the regulariser applies."""
    out = image_loss(image_hwc, gt_chw, lambda_value, dtype)
    if features is not None and enable_regularization:
        reg, g = scale_regularisation(features, invalid_mask, dtype)
        out["reg"] = reg
        out["loss"] = out["loss"] + dtype(regularization_weight) * reg
        out["grad_scale"] = dtype(regularization_weight) * g
    return out
