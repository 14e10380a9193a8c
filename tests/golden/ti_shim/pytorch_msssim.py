"""Stand-in for `pytorch_msssim` -- TEST INFRASTRUCTURE ONLY (the package is not installed here or
on the GPU box).  `ssim` restates the published algorithm of pytorch_msssim.ssim with the torch
calls the package makes (separable 11-tap Gaussian window, sigma 1.5, F.conv2d with groups = C,
valid padding, K1 = 0.01, K2 = 0.03) so that the reference's LossFunction / trainer modules
(LossFunction.py:4,30-31, GaussianPointTrainer.py:16,290-291) import and run unchanged.  The same
function pins oracle/loss_oracle.py in tests/test_loss_oracle.py."""
import torch
import torch.nn.functional as F


def _window(size=11, sigma=1.5, dtype=torch.float32, device=None):
    coords = torch.arange(size, dtype=torch.float) - size // 2
    g = torch.exp(-(coords ** 2) / (2 * sigma ** 2))
    g /= g.sum()
    return g.to(dtype=dtype, device=device)


def _filter(a, win):
    c = a.shape[1]
    out = a
    for i, s in enumerate(a.shape[2:]):
        if s >= win.shape[-1]:
            out = F.conv2d(out, weight=win.transpose(2 + i, -1), stride=1, padding=0, groups=c)
    return out


def ssim(X, Y, data_range=255, size_average=True, win_size=11, win_sigma=1.5, win=None, K=(0.01, 0.03),
         nonnegative_ssim=False):
    if X.shape != Y.shape or X.dim() != 4:
        raise ValueError("expected two [N, C, H, W] tensors of the same shape")
    c = X.shape[1]
    w = _window(win_size, win_sigma, X.dtype, X.device).repeat(c, 1, 1, 1)  # [C, 1, 1, size]
    c1, c2 = (K[0] * data_range) ** 2, (K[1] * data_range) ** 2
    mu1, mu2 = _filter(X, w), _filter(Y, w)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = _filter(X * X, w) - mu1_sq
    s2 = _filter(Y * Y, w) - mu2_sq
    s12 = _filter(X * Y, w) - mu1_mu2
    cs_map = (2 * s12 + c2) / (s1 + s2 + c2)
    ssim_map = ((2 * mu1_mu2 + c1) / (mu1_sq + mu2_sq + c1)) * cs_map
    per_channel = torch.flatten(ssim_map, 2).mean(-1)
    if nonnegative_ssim:
        per_channel = torch.relu(per_channel)
    return per_channel.mean() if size_average else per_channel.mean(1)
