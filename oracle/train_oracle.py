"""CPU oracle of the two post-backward steps.  TEST
INFRASTRUCTURE ONLY: imported by tests/ (and nothing in the product path).

  * adam_step: torch.optim.Adam's single-tensor update (torch/optim/adam.py `_single_tensor_adam`,
    amsgrad off, no weight decay, not maximising), which is what the reference trainer runs for
    its two optimisers (GaussianPointTrainer.py:131-134,183-184).  torch is a third-party
    dependency of the reference; tests/test_train_oracle.py pins this restatement against
    torch.optim.Adam itself on the CPU.
  * controller_update: GaussianPointAdaptiveController.update, reference
    GaussianPointAdaptiveController.py:130-146, statement by statement.
"""
from __future__ import annotations

import numpy as np


def adam_step(p, g, m, v, lr, beta1, beta2, eps, step, dtype=np.float64):
    """Returns the updated (p, m, v); `step` is 1-based."""
    p, g, m, v = (a.astype(dtype) for a in (p, g, m, v))
    m = m + dtype(1 - beta1) * (g - m)                     # exp_avg.lerp_(grad, 1 - beta1)
    v = dtype(beta2) * v + dtype(1 - beta2) * g * g        # mul_(beta2).addcmul_(grad, grad, 1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    step_size = dtype(lr / bc1)
    denom = np.sqrt(v) / dtype(bc2 ** 0.5) + dtype(eps)
    p = p - step_size * (m / denom)                        # addcdiv_(exp_avg, denom, value=-step_size)
    return p, m, v


ACCUMULATORS = (
    ("accumulated_num_pixels", np.int32, 1),
    ("accumulated_num_in_camera", np.int32, 1),
    ("accumulated_view_space_position_gradients", np.float32, 1),
    ("accumulated_view_space_position_gradients_avg", np.float32, 1),
    ("accumulated_position_gradients", np.float32, 3),
    ("accumulated_position_gradients_norm", np.float32, 1),
)


def new_accumulators(n):
    return {name: np.zeros((n,) if cols == 1 else (n, cols), dtype) for name, dtype, cols in ACCUMULATORS}


def controller_update(acc, ids, num_affected_pixels, magnitude_grad_viewspace, grad_point_in_camera):
    """In place, fp32 like the reference's tensors (AC.py:130-146)."""
    ids = ids.astype(np.int64)
    acc["accumulated_num_in_camera"][ids] += 1
    acc["accumulated_num_pixels"][ids] += num_affected_pixels
    acc["accumulated_view_space_position_gradients"][ids] += magnitude_grad_viewspace
    with np.errstate(divide="ignore", invalid="ignore"):
        avg = magnitude_grad_viewspace / num_affected_pixels.astype(np.float32)
    avg[np.isnan(avg)] = 0
    acc["accumulated_view_space_position_gradients_avg"][ids] += avg.astype(np.float32)
    acc["accumulated_position_gradients"][ids] += grad_point_in_camera
    norm = np.sqrt((grad_point_in_camera.astype(np.float64) ** 2).sum(axis=1)).astype(np.float32)
    acc["accumulated_position_gradients_norm"][ids] += norm
    return acc
