"""Parity of the fused loss front-end (csrc/loss.cu through the C ABI and the LossFunction module)
with oracle/loss_oracle.py evaluated in fp64.

Tolerances (fp32 kernels against an fp64 oracle):
  * L, L1, 1-SSIM:  |difference| <= 2e-6  (values are O(0.1 .. 1))
  * dL/d(image):    max |difference| <= 1e-4 * max |oracle gradient|
  * regulariser gradient: relative 1e-5 per element
Reference: LossFunction.py:20-52, GaussianPointTrainer.py:172-182.
"""
import numpy as np
import pytest
import torch

from oracle import loss_oracle as LO

pytestmark = pytest.mark.gpu

LOSS_TOL = 2e-6
GRAD_TOL = 1e-4


def _module(**kw):
    from taichi_3d_gaussian_splatting_b200.LossFunction import LossFunction
    return LossFunction(LossFunction.LossFunctionConfig(**kw))


def _images(h, w, seed, smooth=False):
    rng = np.random.default_rng(seed)
    gt = rng.uniform(0, 1, (3, h, w))
    if smooth:  # low-variance windows: the SSIM denominators get close to C1, C2
        yy, xx = np.mgrid[0:h, 0:w]
        gt = np.stack([0.5 + 0.4 * np.sin(xx / 37.0 + c) * np.cos(yy / 23.0) for c in range(3)])
    img = gt.transpose(1, 2, 0) + rng.normal(0, 0.15, (h, w, 3))
    img[rng.uniform(size=(h, w, 3)) < 0.02] = 0.0
    img[rng.uniform(size=(h, w, 3)) < 0.01] = 1.0
    return img.astype(np.float32), gt.astype(np.float32)


def _check(h, w, lam, seed, smooth=False):
    img, gt = _images(h, w, seed, smooth)
    o = LO.image_loss(img, gt, lam)
    dev = torch.device("cuda")
    t_img = torch.from_numpy(img).to(dev).requires_grad_(True)
    t_gt = torch.from_numpy(gt).to(dev)
    lf = _module(lambda_value=lam)
    loss, l1, ld = lf.forward_rasterised(t_img, t_gt)
    loss.backward()
    assert abs(loss.item() - o["loss"]) <= LOSS_TOL
    assert abs(l1.item() - o["l1"]) <= LOSS_TOL
    assert abs(ld.item() - o["ld_ssim"]) <= LOSS_TOL
    g, go = t_img.grad.cpu().numpy(), o["grad_image_hwc"]
    assert g.shape == go.shape
    err = np.abs(g - go).max()
    assert err <= GRAD_TOL * np.abs(go).max(), (err, np.abs(go).max())
    assert np.all(g[(img < 0) | (img > 1)] == 0)
    return err / np.abs(go).max()


@pytest.mark.parametrize("h,w,lam", [(11, 11, 0.2), (16, 16, 0.2), (32, 48, 0.2), (37, 53, 0.2),
                                     (42, 43, 0.5), (64, 64, 0.0), (75, 33, 1.0), (128, 208, 0.2)])
def test_fused_loss_matches_oracle(h, w, lam):
    _check(h, w, lam, 100 + h * 7 + w)


def test_fused_loss_smooth_images():
    _check(96, 160, 0.2, 7, smooth=True)


def test_fused_loss_full_size():
    """The C2 frame: 1920 x 1088 (2040 tiles forward, 2040 backward)."""
    _check(1088, 1920, 0.2, 8)


def test_reference_call_style_matches_fused():
    """forward(clamp(image).permute(2,0,1), gt): read through the view's strides; torch's clamp
    backward supplies the mask.  Must agree with the fused entry to the last bit in value and
    gradient (same kernels, same inputs after the clamp)."""
    img, gt = _images(48, 80, 21)
    dev = torch.device("cuda")
    lf = _module()
    a = torch.from_numpy(img).to(dev).requires_grad_(True)
    b = torch.from_numpy(img).to(dev).requires_grad_(True)
    t_gt = torch.from_numpy(gt).to(dev)
    la, l1a, lda = lf.forward_rasterised(a, t_gt)
    la.backward()
    pred = torch.clamp(b, min=0, max=1).permute(2, 0, 1)
    lb, l1b, ldb = lf(pred.unsqueeze(0), t_gt.unsqueeze(0))
    lb.backward()
    assert la.item() == lb.item() and l1a.item() == l1b.item() and lda.item() == ldb.item()
    assert torch.equal(a.grad, b.grad)


def test_chw_contiguous_prediction_and_strided_target():
    img, gt = _images(40, 56, 22)
    img = np.clip(img, 0, 1)
    o = LO.image_loss(img, gt, 0.2)
    dev = torch.device("cuda")
    pred = torch.from_numpy(np.ascontiguousarray(img.transpose(2, 0, 1))).to(dev).requires_grad_(True)
    t_gt = torch.from_numpy(np.ascontiguousarray(gt.transpose(1, 2, 0))).to(dev).permute(2, 0, 1)  # HWC storage
    loss, _, _ = _module()(pred, t_gt)
    loss.backward()
    assert abs(loss.item() - o["loss"]) <= LOSS_TOL
    go = o["grad_image_hwc"].transpose(2, 0, 1)
    assert np.abs(pred.grad.cpu().numpy() - go).max() <= GRAD_TOL * np.abs(go).max()


def test_upstream_scale_and_regulariser():
    img, gt = _images(32, 32, 23)
    rng = np.random.default_rng(24)
    n = 5000
    feat = rng.normal(0, 1, (n, 56)).astype(np.float32)
    mask = (rng.uniform(size=n) < 0.1).astype(np.int8)
    o = LO.total_loss(img, gt, feat, mask, 0.2, True, 2.0)
    dev = torch.device("cuda")
    t_img = torch.from_numpy(img).to(dev).requires_grad_(True)
    t_feat = torch.from_numpy(feat).to(dev).requires_grad_(True)
    lf = _module(lambda_value=0.2, enable_regularization=True, regularization_weight=2.0)
    loss, _, _ = lf.forward_rasterised(t_img, torch.from_numpy(gt).to(dev), torch.from_numpy(mask).to(dev), t_feat)
    assert abs(loss.item() - o["loss"]) <= 1e-5 * abs(o["loss"])
    (3.0 * loss).backward()
    g = t_feat.grad.cpu().numpy()
    assert np.all(g[:, :4] == 0) and np.all(g[:, 7:] == 0) and np.all(g[mask != 0] == 0)
    want = 3.0 * o["grad_scale"]
    assert np.allclose(g[:, 4:7], want, rtol=1e-5, atol=0)
    gi, go = t_img.grad.cpu().numpy(), 3.0 * o["grad_image_hwc"]
    assert np.abs(gi - go).max() <= GRAD_TOL * np.abs(go).max()
    # regulariser switched off: features get no gradient and L is the image loss
    t_feat2 = torch.from_numpy(feat).to(dev).requires_grad_(True)
    lf2 = _module(enable_regularization=False)
    l2, _, _ = lf2.forward_rasterised(torch.from_numpy(img).to(dev), torch.from_numpy(gt).to(dev),
                                      torch.from_numpy(mask).to(dev), t_feat2)
    assert abs(l2.item() - LO.image_loss(img, gt, 0.2)["loss"]) <= LOSS_TOL


def test_accumulate_into_existing_feature_gradient():
    from taichi_3d_gaussian_splatting_b200 import _capi
    rng = np.random.default_rng(25)
    n = 1000
    feat = rng.normal(0, 1, (n, 56)).astype(np.float32)
    mask = (rng.uniform(size=n) < 0.1).astype(np.int8)
    _, grad = LO.scale_regularisation(feat, mask)
    dev = torch.device("cuda")
    t_feat, t_mask = torch.from_numpy(feat).to(dev), torch.from_numpy(mask).to(dev)
    scalars = torch.zeros(6, device=dev)
    scalars[4] = float((mask == 0).sum())
    base = torch.from_numpy(rng.normal(0, 1, (n, 56)).astype(np.float32)).to(dev)
    buf = base.clone()
    st = torch.cuda.current_stream().cuda_stream
    _capi.check(_capi.lib.gs_scale_reg_bwd(t_feat.data_ptr(), t_mask.data_ptr(), n, 2.0, scalars.data_ptr(),
                                           None, buf.data_ptr(), 1, st), "gs_scale_reg_bwd")
    want = base.cpu().numpy().astype(np.float64)
    want[:, 4:7] += 2.0 * grad
    assert np.allclose(buf.cpu().numpy(), want, rtol=1e-5, atol=1e-7)


def test_invalid_arguments():
    from taichi_3d_gaussian_splatting_b200 import _capi
    assert _capi.lib.gs_loss_workspace_bytes(10, 64) < 0
    assert _capi.lib.gs_loss_maps_bytes(64, 5) < 0
    dev = torch.device("cuda")
    with pytest.raises(ValueError):
        _module()(torch.zeros(3, 8, 64, device=dev), torch.zeros(3, 8, 64, device=dev))
    with pytest.raises(ValueError):
        _module()(torch.zeros(2, 3, 16, 16, device=dev), torch.zeros(2, 3, 16, 16, device=dev))
    with pytest.raises(ValueError):
        _module()(torch.zeros(3, 16, 16), torch.zeros(3, 16, 16))


def test_folded_regulariser_gradient_matches_the_autograd_one():
    """LossFunction(fold_regulariser_gradient=True): the regulariser's gradient is added in place to
    the gradient arriving at the features from another consumer (here a stand-in for the rasteriser)
    instead of travelling as a second dense tensor; same totals, upstream scale included."""
    img, gt = _images(32, 48, 31)
    rng = np.random.default_rng(32)
    n = 4000
    feat = rng.normal(0, 1, (n, 56)).astype(np.float32)
    mask = (rng.uniform(size=n) < 0.1).astype(np.int8)
    wmat = rng.normal(0, 1, (n, 56)).astype(np.float32)
    dev = torch.device("cuda")
    results = []
    for fold in (False, True):
        from taichi_3d_gaussian_splatting_b200.LossFunction import LossFunction
        lf = LossFunction(LossFunction.LossFunctionConfig(), fold_regulariser_gradient=fold)
        t_img = torch.from_numpy(img).to(dev).requires_grad_(True)
        t_feat = torch.from_numpy(feat).to(dev).requires_grad_(True)
        other = (t_feat * torch.from_numpy(wmat).to(dev)).sum() * 1e-3      # the "rasteriser" path
        loss, _, _ = lf.forward_rasterised(t_img, torch.from_numpy(gt).to(dev), torch.from_numpy(mask).to(dev), t_feat)
        (2.5 * (loss + other)).backward()
        results.append((loss.item(), t_feat.grad.cpu().numpy(), t_img.grad.cpu().numpy()))
    assert results[0][0] == results[1][0]
    assert np.allclose(results[0][1], results[1][1], rtol=1e-6, atol=1e-9)
    assert np.array_equal(results[0][2], results[1][2])
    o = LO.total_loss(img, gt, feat, mask, 0.2, True, 2.0)
    want = 2.5 * (1e-3 * wmat.astype(np.float64))
    want[:, 4:7] += 2.5 * o["grad_scale"]
    assert np.allclose(results[1][1], want, rtol=2e-5, atol=1e-8)


def test_against_the_package_the_reference_imports():
    """The pin the loss row is missing: the reference computes SSIM with
    pytorch_msssim.ssim(pred, gt, data_range=1, size_average=True) (LossFunction.py:4,30-31), a third-party
    package that is not vendored.  Where it is importable this test compares gs_loss_fwd / gs_loss_bwd with
    that exact call + autograd; where it is not (this image: neither here nor on the GPU box, and nothing
    can be installed) it SKIPS -- it never passes vacuously -- and the row stays "parity unpinned"."""
    try:
        import pytorch_msssim
    except ImportError:
        pytest.skip("pytorch_msssim is not installed on this box: loss parity against the package the "
                    "reference imports stays unpinned (oracle = restatement of its published algorithm)")
    if "ti_shim" in (getattr(pytorch_msssim, "__file__", "") or ""):
        pytest.skip("only the test stand-in of pytorch_msssim is importable here, not the package")
    img, gt = _images(96, 128, 77)
    pred = torch.from_numpy(img).cuda().requires_grad_(True)
    target = torch.from_numpy(gt).cuda()
    lam = 0.2
    x = torch.clamp(pred, 0, 1).permute(2, 0, 1)
    l1 = torch.abs(x - target).mean()
    ld = 1 - pytorch_msssim.ssim(x.unsqueeze(0), target.unsqueeze(0), data_range=1, size_average=True)
    want = (1 - lam) * l1 + lam * ld
    want.backward()
    want_grad = pred.grad.clone()
    pred.grad = None
    got, got_l1, got_ld = _module(lambda_value=lam, enable_regularization=False).forward_rasterised(pred, target)
    got.backward()
    print("pytorch_msssim", getattr(pytorch_msssim, "__version__", "?"), "loss", float(want), float(got))
    assert abs(float(got) - float(want)) <= LOSS_TOL and abs(float(got_ld) - float(ld)) <= LOSS_TOL
    assert float((pred.grad - want_grad).abs().max()) <= GRAD_TOL * float(want_grad.abs().max())
