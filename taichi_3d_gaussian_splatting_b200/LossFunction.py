"""Drop-in for the reference `LossFunction` (taichi_3d_gaussian_splatting/LossFunction.py:9-52),
running on the fused sm_100a kernels of csrc/loss.cu through the C ABI (gs_loss_fwd,
gs_loss_bwd, gs_scale_reg_bwd in include/gs_b200.h).  No torch fallback: without libgs_b200.so
this module does not import.

    loss_function = LossFunction(LossFunction.LossFunctionConfig())
    L, L1, LD_SSIM = loss_function(image_pred, image_gt, point_invalid_mask, pointcloud_features)
    L.backward()

`forward` keeps the reference's signature and return tuple.  Two ways to call it:

  * the reference's way (GaussianPointTrainer.py:172-182): `predicted_image` is the clamped,
    permuted (3, H, W) view of the rasteriser's image.  It is read in place through its strides;
    torch's own clamp / permute backward then carry the gradient to the rasteriser.
  * fused: `forward_rasterised(image_hwc, ...)` takes the rasteriser's raw H x W x 3 output; the
    clamp and the layout change happen inside the kernels (three full-image round trips fewer),
    and the gradient comes back directly in the layout gs_blend_bwd reads.

Differences from the reference: one image per call (its trainer never batches,
GaussianPointTrainer.py:125-126); fp32 CUDA tensors only.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn

from . import _capi

try:  # the reference derives the config from dataclass_wizard.YAMLWizard (LossFunction.py:5,11)
    from dataclass_wizard import YAMLWizard
except Exception:  # pragma: no cover
    class YAMLWizard:  # type: ignore
        pass

_lib = _capi.lib


def _chw_strides(t: torch.Tensor):
    """(h, w, c) element strides of a logical (3, H, W) tensor."""
    return t.stride(1), t.stride(2), t.stride(0)


class _FusedLoss(torch.autograd.Function):
    """(predicted (3,H,W) logical view, features or None) -> scalars [6] (see gs_loss_fwd)."""

    @staticmethod
    def forward(ctx, pred_chw, gt_chw, features, invalid_mask, lambda_value, reg_weight, use_reg, folded=None):
        for t, name in ((pred_chw, "predicted_image"), (gt_chw, "ground_truth_image")):
            if not t.is_cuda or t.dtype != torch.float32:
                raise ValueError(f"{name} must be a float32 CUDA tensor")
        if pred_chw.shape != gt_chw.shape or pred_chw.dim() != 3 or pred_chw.shape[0] != 3:
            raise ValueError(f"expected two (3, H, W) images, got {tuple(pred_chw.shape)} and {tuple(gt_chw.shape)}")
        _, h, w = pred_chw.shape
        if h < 11 or w < 11:
            raise ValueError("SSIM's 11x11 window needs H, W >= 11")
        dev = pred_chw.device
        use_reg = bool(use_reg and features is not None)
        if use_reg:
            if invalid_mask is None:
                raise ValueError("point_invalid_mask is required with pointcloud_features")
            if features.dtype != torch.float32 or features.dim() != 2 or features.shape[1] != 56:
                raise ValueError("pointcloud_features must be [N, 56] float32")
            features_c = features if features.is_contiguous() else features.contiguous()
            mask_c = invalid_mask.to(torch.int8).contiguous()
        stream = torch.cuda.current_stream(dev).cuda_stream
        with torch.cuda.device(dev):
            maps = torch.empty(_lib.gs_loss_maps_bytes(h, w) // 4, dtype=torch.float32, device=dev)
            ws = torch.empty(_lib.gs_loss_workspace_bytes(h, w), dtype=torch.uint8, device=dev)
            scalars = torch.empty(6, dtype=torch.float32, device=dev)
            ph, pw, pc = _chw_strides(pred_chw)
            gh, gw, gc = _chw_strides(gt_chw)
            _capi.check(_lib.gs_loss_fwd(
                pred_chw.data_ptr(), ph, pw, pc, gt_chw.data_ptr(), gh, gw, gc, h, w, float(lambda_value),
                features_c.data_ptr() if use_reg else None, mask_c.data_ptr() if use_reg else None,
                features_c.shape[0] if use_reg else 0, float(reg_weight),
                maps.data_ptr(), ws.data_ptr(), scalars.data_ptr(), stream), "gs_loss_fwd")
        ctx.save_for_backward(pred_chw, gt_chw, maps, scalars,
                              features_c if use_reg else None, mask_c if use_reg else None)
        ctx.cfg = (h, w, float(lambda_value), float(reg_weight), use_reg)
        ctx.folded = folded
        if folded is not None and use_reg:   # see LossFunction(fold_regulariser_gradient=True)
            folded.update(features=features_c, mask=mask_c, scalars=scalars, weight=float(reg_weight))
        return scalars

    @staticmethod
    def backward(ctx, grad_scalars):
        pred_chw, gt_chw, maps, scalars, features, mask = ctx.saved_tensors
        h, w, lam, reg_weight, use_reg = ctx.cfg
        dev = pred_chw.device
        # only slot 0 (L) is differentiable; L1 and 1-SSIM are reported values, as in the reference
        # trainer, which calls backward on L alone (GaussianPointTrainer.py:182)
        upstream = grad_scalars[0:1].contiguous()
        if ctx.folded is not None:
            ctx.folded["upstream"] = upstream   # the tensor hook on the features runs after this
        stream = torch.cuda.current_stream(dev).cuda_stream
        grad_pred = grad_feat = None
        with torch.cuda.device(dev):
            if ctx.needs_input_grad[0]:
                grad_hwc = torch.empty((h, w, 3), dtype=torch.float32, device=dev)
                ph, pw, pc = _chw_strides(pred_chw)
                gh, gw, gc = _chw_strides(gt_chw)
                _capi.check(_lib.gs_loss_bwd(
                    pred_chw.data_ptr(), ph, pw, pc, gt_chw.data_ptr(), gh, gw, gc, h, w, lam,
                    maps.data_ptr(), upstream.data_ptr(), grad_hwc.data_ptr(), stream), "gs_loss_bwd")
                grad_pred = grad_hwc.permute(2, 0, 1)
            if use_reg and ctx.needs_input_grad[2]:
                grad_feat = torch.zeros_like(features)
                _capi.check(_lib.gs_scale_reg_bwd(
                    features.data_ptr(), mask.data_ptr(), features.shape[0], reg_weight,
                    scalars.data_ptr(), upstream.data_ptr(), grad_feat.data_ptr(), 0, stream),
                    "gs_scale_reg_bwd")
        return grad_pred, None, grad_feat, None, None, None, None, None


class LossFunction(nn.Module):
    @dataclass
    class LossFunctionConfig(YAMLWizard):
        lambda_value: float = 0.2
        enable_regularization: bool = True
        regularization_weight: float = 2

    def __init__(self, config: "LossFunction.LossFunctionConfig", fold_regulariser_gradient: bool = False):
        """fold_regulariser_gradient: the scale regulariser touches 3 of the 56 feature columns, but as
        an autograd gradient it is a second dense N x 56 tensor (zero fill, strided write, and a full
        add when autograd sums it with the rasteriser's gradient: 0.18 ms at 1M points).  With this
        switch the term's value still enters L, but its gradient is ADDED IN PLACE to the gradient
        that reaches `pointcloud_features` from the rasteriser (a one-shot tensor hook; 0.01 ms).
        Requires that the features of this call do receive a gradient in the same backward pass (they
        do in the trainer: the image depends on them); otherwise leave it off."""
        super().__init__()
        self.config = config
        self.fold_regulariser_gradient = bool(fold_regulariser_gradient)

    @staticmethod
    def _single(image: torch.Tensor, name: str) -> torch.Tensor:
        if image.dim() == 4:
            if image.shape[0] != 1:
                raise ValueError(f"{name}: one image per call (got a batch of {image.shape[0]})")
            image = image[0]
        return image

    def _run(self, pred_chw, gt_chw, point_invalid_mask, pointcloud_features):
        cfg = self.config
        fold = (self.fold_regulariser_gradient and cfg.enable_regularization and pointcloud_features is not None
                and pointcloud_features.requires_grad and torch.is_grad_enabled())
        if not fold:
            scalars = _FusedLoss.apply(pred_chw, gt_chw, pointcloud_features, point_invalid_mask,
                                       cfg.lambda_value, cfg.regularization_weight, cfg.enable_regularization)
            return scalars[0], scalars[1].detach(), scalars[2].detach()
        folded = {}
        scalars = _FusedLoss.apply(pred_chw, gt_chw, pointcloud_features.detach(), point_invalid_mask,
                                   cfg.lambda_value, cfg.regularization_weight, True, folded)
        handle = []

        def add_regulariser_gradient(grad):
            handle[0].remove()
            if "upstream" not in folded:      # L itself was not part of this backward pass
                return grad
            g = grad if grad.is_contiguous() else grad.contiguous()
            with torch.cuda.device(g.device):
                _capi.check(_lib.gs_scale_reg_bwd(
                    folded["features"].data_ptr(), folded["mask"].data_ptr(), g.shape[0], folded["weight"],
                    folded["scalars"].data_ptr(), folded["upstream"].data_ptr(), g.data_ptr(), 1,
                    torch.cuda.current_stream(g.device).cuda_stream), "gs_scale_reg_bwd")
            return g

        handle.append(pointcloud_features.register_hook(add_regulariser_gradient))
        return scalars[0], scalars[1].detach(), scalars[2].detach()

    def forward(self, predicted_image, ground_truth_image, point_invalid_mask=None, pointcloud_features=None):
        """L = (1 - lambda) L1 + lambda (1 - SSIM) [+ weight * scale regulariser]
        (LossFunction.py:20-38).  predicted_image, ground_truth_image: (3, H, W) or (1, 3, H, W).
        Returns (L, L1, LD_SSIM) as 0-dim tensors; only L carries a gradient."""
        pred = self._single(predicted_image, "predicted_image")
        gt = self._single(ground_truth_image, "ground_truth_image")
        return self._run(pred, gt, point_invalid_mask, pointcloud_features)

    def forward_rasterised(self, rasterized_image, ground_truth_image, point_invalid_mask=None,
                           pointcloud_features=None):
        """Same value as forward(torch.clamp(rasterized_image, 0, 1).permute(2, 0, 1), ...) for the
        rasteriser's raw (H, W, 3) image (GaussianPointTrainer.py:172-182), with the clamp and
        the layout change folded into the kernels."""
        if rasterized_image.dim() != 3 or rasterized_image.shape[2] != 3:
            raise ValueError("rasterized_image must be (H, W, 3)")
        gt = self._single(ground_truth_image, "ground_truth_image")
        return self._run(rasterized_image.permute(2, 0, 1), gt, point_invalid_mask, pointcloud_features)
