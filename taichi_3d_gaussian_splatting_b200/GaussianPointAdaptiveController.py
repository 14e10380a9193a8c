"""The accumulator half of the reference's `GaussianPointAdaptiveController`
(taichi_3d_gaussian_splatting/GaussianPointAdaptiveController.py:113-146 and the resets at
:151-166): the six per-point statistics the densification rules read, kept on the device and
updated from the rasteriser's backward hook by ONE kernel (gs_controller_accumulate,
csrc/train_ops.cu) instead of five gathers and six index-adds.

Only the statistics are here.  The densification / pruning policy that consumes them
(`_find_densify_points`, `_add_densify_points`, floater removal, opacity reset,
AC.py:168-347) is host-side control logic outside the hot path; it
can read these tensors under the same attribute names.

    accumulators = AdaptiveControllerAccumulators(num_points, device)
    rasterisation = GaussianPointCloudRasterisation(config, backward_valid_point_hook=accumulators.update)
"""
from __future__ import annotations

import torch

from . import _capi

_lib = _capi.lib


class AdaptiveControllerAccumulators:
    # attribute name, dtype, columns (AC.py:113-125)
    FIELDS = (
        ("accumulated_num_pixels", torch.int32, 1),
        ("accumulated_num_in_camera", torch.int32, 1),
        ("accumulated_view_space_position_gradients", torch.float32, 1),
        ("accumulated_view_space_position_gradients_avg", torch.float32, 1),
        ("accumulated_position_gradients", torch.float32, 3),
        ("accumulated_position_gradients_norm", torch.float32, 1),
    )

    def __init__(self, num_points: int, device):
        self.num_points = int(num_points)
        self.device = torch.device(device)
        self.iteration_counter = -1
        self.reset()

    def reset(self) -> None:
        """Zero every statistic (what `refinement` does after a densification, AC.py:151-166)."""
        for name, dtype, cols in self.FIELDS:
            shape = (self.num_points,) if cols == 1 else (self.num_points, cols)
            setattr(self, name, torch.zeros(shape, dtype=dtype, device=self.device))

    @torch.no_grad()
    def update(self, input_data) -> None:
        """`input_data`: the BackwardValidPointHookInput of one backward (R.py:806-817)."""
        if self.device.type != "cuda":
            raise ValueError("update() runs a CUDA kernel: the accumulators must live on a CUDA device")
        self.iteration_counter += 1
        ids = input_data.point_id_in_camera_list
        m = int(ids.shape[0])
        if m == 0:
            return
        num_px = input_data.num_affected_pixels
        mag = input_data.magnitude_grad_viewspace
        gxyz = input_data.grad_point_in_camera
        for t, dtype in ((ids, torch.int32), (num_px, torch.int32), (mag, torch.float32), (gxyz, torch.float32)):
            if t.dtype != dtype or not t.is_cuda or not t.is_contiguous():
                raise ValueError("hook payload tensors must be contiguous CUDA tensors of the operator's dtypes")
        with torch.cuda.device(self.device):
            _capi.check(_lib.gs_controller_accumulate(
                ids.data_ptr(), m, num_px.data_ptr(), mag.data_ptr(), gxyz.data_ptr(),
                self.accumulated_num_in_camera.data_ptr(), self.accumulated_num_pixels.data_ptr(),
                self.accumulated_view_space_position_gradients.data_ptr(),
                self.accumulated_view_space_position_gradients_avg.data_ptr(),
                self.accumulated_position_gradients.data_ptr(),
                self.accumulated_position_gradients_norm.data_ptr(),
                torch.cuda.current_stream(self.device).cuda_stream), "gs_controller_accumulate")

    def all_reduce(self, group=None) -> None:
        """View-sharded training (distributed.ViewShardedTrainerStep): every rank has accumulated the
        statistics of ITS views; summing them over the ranks gives what a single process that saw
        every view would hold.  Call it before the densification policy reads the tensors (once per
        densification interval, not per step)."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return
        for name, _, _ in self.FIELDS:
            dist.all_reduce(getattr(self, name), group=group)
