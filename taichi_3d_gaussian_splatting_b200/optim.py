"""`FusedAdam`: torch.optim.Adam for the trainer's two parameter tensors (reference
GaussianPointTrainer.py:131-134: point_cloud_features and point_cloud, betas (0.9, 0.999),
amsgrad off, no weight decay) as one streaming kernel per tensor (gs_adam_step, csrc/train_ops.cu).

It is a torch.optim.Optimizer, so `zero_grad()`, `param_groups[...]["lr"]` and
torch.optim.lr_scheduler.ExponentialLR (GaussianPointTrainer.py:136-137) work unchanged.
No torch fallback: parameters must be contiguous float32 CUDA tensors.
"""
from __future__ import annotations

import torch

from . import _capi

_lib = _capi.lib


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1):
            raise ValueError("invalid Adam hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        """One update.  `grads`: optional {param: gradient tensor} to read instead of `param.grad`
        (e.g. views of the all-reduced flat gradient buffer of distributed.ViewShardedTrainerStep)."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                g = grads.get(p) if grads is not None else p.grad
                if g is None:
                    continue
                if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()):
                    raise ValueError("FusedAdam needs contiguous float32 CUDA parameters")
                if g.dtype != torch.float32 or g.shape != p.shape or g.device != p.device:
                    raise ValueError("gradient must match its parameter (float32, same shape and device)")
                if not g.is_contiguous():
                    g = g.contiguous()
                state = self.state[p]
                if not state:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                state["step"] += 1
                with torch.cuda.device(p.device):
                    _capi.check(_lib.gs_adam_step(
                        p.data_ptr(), g.data_ptr(), state["exp_avg"].data_ptr(), state["exp_avg_sq"].data_ptr(),
                        p.numel(), float(group["lr"]), float(beta1), float(beta2), float(group["eps"]),
                        int(state["step"]), None, 0, torch.cuda.current_stream(p.device).cuda_stream),
                        "gs_adam_step")
        return loss
