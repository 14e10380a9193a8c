"""Drop-in for the reference operator `GaussianPointCloudRasterisation`.

Same module / nested dataclasses / `torch.autograd.Function` forward and backward
signatures as the reference (taichi_3d_gaussian_splatting/GaussianPointCloudRasterisation.py
:775-1204, "R.py" below); the body of forward/backward launches the hand-written sm_100a
kernels of libgs_b200.so through the C ABI of include/gs_b200.h.  No Taichi, no Triton, no
CPU fallback: without the library this module does not import.

Host-side differences from the reference, none of which change results:
  * one host sync per forward (read M and K) instead of two (R.py:864, :916-931);
  * per-point intermediates live in one buffer sized for N points (only the first M rows
    are meaningful) instead of M-sized tensors allocated after a sync;
  * empty frustum / empty key list returns zero images instead of uninitialised memory
    (R.py:967-980).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Any, Callable, Dict, Optional

import torch

from . import _capi
from .Camera import CameraInfo

try:  # the reference derives its config from dataclass_wizard.YAMLWizard (R.py:20,777)
    from dataclass_wizard import YAMLWizard
except Exception:  # pragma: no cover - dataclass_wizard is optional here
    class YAMLWizard:  # type: ignore
        pass

BOUNDARY_TILES = 3  # R.py:26
TILE_WIDTH = 16     # R.py:27
TILE_HEIGHT = 16    # R.py:28

_lib = _capi.lib
_REC = 12
_GRAD_SLOTS = 12

# "depth_first": sort the points by depth, emit tile ids in that order, sort by tile id only (a third
# of the sort traffic; same payload order and tile ranges).  "literal": build and sort the reference's
# 64-bit keys (R.py:131-172, 947-950); also taken automatically when depth keys can be negative.
# "binned": after the same point sort, split the points over tile columns and then tile rows
# (gs_bin_tiles): no (tile id, payload) pair per list entry is written or sorted; same payload order and
# tile ranges.  Falls back to "depth_first" when the image has more than 256 tile columns or rows.
KEY_PIPELINE = os.environ.get("GS_KEY_PIPELINE", "binned")

# operator calls made through this module since import (acceptance tests check that the reference's
# unchanged callers really ran THIS code)
CALL_COUNTS = {"forward": 0, "backward": 0, "hook": 0}

# A plain backward (no hook, no accumulation target, no batch) stores the gradient rows of all points with
# gs_point_bwd_all_points instead of zero-filling N x 59 floats and scattering the visible rows into them.
# GS_STORE_ALL_ROWS=0 keeps the two-step form (tests compare the two).
STORE_ALL_ROWS = os.environ.get("GS_STORE_ALL_ROWS", "1") != "0"

# set to a profiling.StageProfiler to get one CUDA event per stage (bench.py); None = no events
STAGE_PROFILER = None

# Priority sections.  The two blend kernels fill every SM's register file and are bound by instruction
# issue; the kernels around them (projection, point sort, binning, per-point backward) are latency /
# HBM bound.  When several views are in flight on different streams, the CTAs of a small kernel would
# wait behind the thousands of queued CTAs of another view's blend kernel.  With this switch the small
# kernels of a view run on a HIGH-PRIORITY companion of the caller's stream (chained to it by events, so
# the order of a view's kernels is unchanged): their CTAs are dispatched as soon as blend CTAs retire
# and run in the issue slots the blend kernels leave idle.  Off by default: a single view on a single
# stream gains nothing and pays four event hand-offs.
PRIORITY_SECTIONS = os.environ.get("GS_PRIORITY_SECTIONS", "0") != "0"
_COMPANION = {}


def _companion_stream(cur: "torch.cuda.Stream") -> "torch.cuda.Stream":
    key = (cur.device.index, cur.cuda_stream)
    s = _COMPANION.get(key)
    if s is None:
        s = _COMPANION[key] = torch.cuda.Stream(cur.device, priority=-1)
    return s


def _hand_over(src: "torch.cuda.Stream", dst: "torch.cuda.Stream") -> None:
    """Work enqueued on dst from here on runs after everything enqueued on src so far."""
    if src is not dst:
        ev = torch.cuda.Event()
        ev.record(src)
        dst.wait_event(ev)


def _mark(name: str) -> None:
    if STAGE_PROFILER is not None:
        STAGE_PROFILER.mark(name)


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _align(x: int, a: int = 256) -> int:
    return (x + a - 1) // a * a


class _PointBuffers:
    """Carves the per-visible-point arrays (capacity n rows) out of one byte buffer.

    CORE is what the path itself reads (76 bytes per point); HOOK the two reference-named buffers a
    backward hook shows (R.py:1127-1142); DEBUG the remaining reference-named buffers (R.py:873-884),
    which nothing but tests looks at; OFFSETS the scan the literal key pipeline reads."""

    FIELDS = (  # name, dtype, columns
        ("ids", torch.int32, 1),
        ("point_uv", torch.float32, 2),
        ("point_in_camera", torch.float32, 3),
        ("conic", torch.float32, 4),
        ("alpha", torch.float32, 1),
        ("color", torch.float32, 3),
        ("radii", torch.float32, 1),
        ("num_tiles", torch.int32, 1),
        ("tile_offsets", torch.int64, 1),
        ("records", torch.float32, _REC),
        ("bbox", torch.int16, 4),
        ("depth_key", torch.int32, 1),
        ("slot_of_point", torch.int32, 1),
    )
    CORE = ("ids", "point_in_camera", "records", "bbox", "depth_key")
    HOOK = ("point_uv", "num_tiles")
    DEBUG = ("conic", "alpha", "color", "radii")
    OFFSETS = ("tile_offsets",)
    SLOTS = ("slot_of_point",)   # the inverse of ids, for gs_point_bwd_batch

    def __init__(self, n: int, device: torch.device, hook: bool = False, debug: bool = False,
                 offsets: bool = False, slots: bool = False):
        self.n = n
        wanted = set(self.CORE)
        if hook or debug:
            wanted |= set(self.HOOK)
        if debug:
            wanted |= set(self.DEBUG)
        if offsets:
            wanted |= set(self.OFFSETS)
        if slots:
            wanted |= set(self.SLOTS)
        self.offsets: Dict[str, int] = {}
        off = 0
        for name, dtype, cols in self.FIELDS:
            if name not in wanted:
                continue
            self.offsets[name] = off
            off = _align(off + max(n, 1) * cols * torch.empty((), dtype=dtype).element_size())
        self.storage = torch.empty(off, dtype=torch.uint8, device=device)
        self.base = self.storage.data_ptr()

    def has(self, name: str) -> bool:
        return name in self.offsets

    def ptr(self, name: str) -> Optional[int]:
        """Device address of the array, or None (a NULL pointer: not produced) if it was not asked for."""
        return self.base + self.offsets[name] if name in self.offsets else None

    def view(self, name: str, rows: int) -> torch.Tensor:
        for fname, dtype, cols in self.FIELDS:
            if fname == name:
                nbytes = rows * cols * torch.empty((), dtype=dtype).element_size()
                off = self.offsets[name]
                t = self.storage[off:off + nbytes].view(dtype)
                return t.view(rows, cols) if cols > 1 else t
        raise KeyError(name)


def sort_key_bits(near_plane: float, far_plane: float, depth_to_sort_key_scale: float,
                  num_tiles: int):
    """How many low bits the depth part and how many bits the tile part of a key need.

    depth key = i32(z * scale) with near < z < far (R.py:72-73,159-160).  A negative near plane
    (negative depth keys) or a depth that does not fit 31 bits selects the full signed 64-bit sort.
    """
    max_depth = far_plane * depth_to_sort_key_scale * 1.001 + 2.0
    if near_plane < 0.0 or not (max_depth < 2.0 ** 31):
        return 32, 32
    depth_bits = max(1, int(max_depth).bit_length())
    tile_bits = max(1, (max(num_tiles, 1) - 1).bit_length())
    return depth_bits, tile_bits


def find_tile_start_and_end(point_in_camera_sort_key: torch.Tensor,
                            tile_points_start: torch.Tensor,
                            tile_points_end: torch.Tensor) -> None:
    """Same call as the reference kernel of that name (R.py:175-193) on CUDA tensors."""
    assert point_in_camera_sort_key.dtype == torch.int64 and point_in_camera_sort_key.is_cuda
    assert tile_points_start.dtype == torch.int32 and tile_points_end.dtype == torch.int32
    stream = torch.cuda.current_stream(point_in_camera_sort_key.device).cuda_stream
    _capi.check(_lib.gs_tile_ranges(point_in_camera_sort_key.data_ptr(),
                                    point_in_camera_sort_key.shape[0],
                                    tile_points_start.data_ptr(), tile_points_end.data_ptr(),
                                    tile_points_start.shape[0], stream), "gs_tile_ranges")


class GaussianPointCloudRasterisation(torch.nn.Module):
    @dataclass
    class GaussianPointCloudRasterisationConfig(YAMLWizard):
        near_plane: float = 0.8
        far_plane: float = 1000.
        depth_to_sort_key_scale: float = 100.
        rgb_only: bool = False
        # un-annotated, i.e. class constants, exactly like R.py:782-786
        grad_color_factor = 5.
        grad_high_order_color_factor = 1.
        grad_s_factor = 0.5
        grad_q_factor = 1.
        grad_alpha_factor = 20.

    @dataclass
    class GaussianPointCloudRasterisationInput:
        point_cloud: torch.Tensor  # Nx3
        point_cloud_features: torch.Tensor  # Nx56
        point_object_id: torch.Tensor  # N int32
        point_invalid_mask: torch.Tensor  # N int8
        camera_info: CameraInfo
        q_pointcloud_camera: torch.Tensor  # Kx4 xyzw
        t_pointcloud_camera: torch.Tensor  # Kx3
        color_max_sh_band: int = 2

    @dataclass
    class BackwardValidPointHookInput:
        point_id_in_camera_list: torch.Tensor  # M
        grad_point_in_camera: torch.Tensor  # Mx3
        grad_pointfeatures_in_camera: torch.Tensor  # Mx56
        grad_viewspace: torch.Tensor  # Mx2
        magnitude_grad_viewspace: torch.Tensor  # M
        magnitude_grad_viewspace_on_image: torch.Tensor  # HxWx2
        num_overlap_tiles: torch.Tensor  # M
        num_affected_pixels: torch.Tensor  # M
        point_depth: torch.Tensor  # M
        point_uv_in_camera: torch.Tensor  # Mx2

    def __init__(
        self,
        config: "GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig",
        backward_valid_point_hook: Optional[Callable[["GaussianPointCloudRasterisation.BackwardValidPointHookInput"], None]] = None,
    ):
        super().__init__()
        self.config = config
        self.last_forward_info: Dict[str, object] = {}
        self._accumulate_into = None
        self._batch: Optional[_GradientBatch] = None
        self._has_hook = backward_valid_point_hook is not None
        outer = self

        class _module_function(torch.autograd.Function):

            @staticmethod
            def forward(ctx,
                        pointcloud,
                        pointcloud_features,
                        point_invalid_mask,
                        point_object_id,
                        q_pointcloud_camera,
                        t_pointcloud_camera,
                        camera_info,
                        color_max_sh_band,
                        ):
                cfg = outer.config
                out = _forward_impl(cfg, pointcloud, pointcloud_features, point_invalid_mask,
                                    point_object_id, q_pointcloud_camera, t_pointcloud_camera,
                                    camera_info,
                                    keep_cell_masks=ctx.needs_input_grad[0] or ctx.needs_input_grad[1],
                                    hook_outputs=backward_valid_point_hook is not None,
                                    # a plain backward without a hook stores the gradient rows of ALL
                                    # points in one pass (no zero fill): it needs the slot map too
                                    slots_output=outer._batch is not None or (
                                        STORE_ALL_ROWS and backward_valid_point_hook is None
                                        and outer._accumulate_into is None
                                        and (ctx.needs_input_grad[0] or ctx.needs_input_grad[1])))
                outer.last_forward_info = out.info
                ctx.set_materialize_grads(False)
                ctx.camera_info = camera_info
                ctx.color_max_sh_band = color_max_sh_band
                # where this call's gradients go is fixed NOW: a later accumulate_gradients_into()
                # (or another user of the module) must not redirect a backward already recorded
                ctx.accumulate_into = outer._accumulate_into
                ctx.batch = outer._batch
                # the outputs must not be reachable from ctx (output -> grad_fn -> ctx cycle)
                image, depth, count = out.image, out.depth, out.count
                out.image = out.depth = out.count = None
                ctx.state = out
                ctx.save_for_backward(pointcloud, pointcloud_features, point_object_id,
                                      q_pointcloud_camera, t_pointcloud_camera)
                return image, depth, count

            @staticmethod
            def backward(ctx, grad_rasterized_image, grad_rasterized_depth=None,
                         grad_pixel_valid_point_count=None):
                grad_pointcloud = grad_pointcloud_features = None
                if ctx.needs_input_grad[0] or ctx.needs_input_grad[1]:
                    pointcloud, pointcloud_features, point_object_id, _, t_pointcloud_camera = \
                        ctx.saved_tensors
                    grad_pointcloud, grad_pointcloud_features = _backward_impl(
                        outer.config, ctx.state, pointcloud, pointcloud_features, point_object_id,
                        t_pointcloud_camera, ctx.camera_info, ctx.color_max_sh_band,
                        grad_rasterized_image, backward_valid_point_hook,
                        GaussianPointCloudRasterisation.BackwardValidPointHookInput,
                        accumulate_into=ctx.accumulate_into, batch=ctx.batch)
                return grad_pointcloud, grad_pointcloud_features, None, None, None, None, None, None

        self._module_function = _module_function

    def accumulate_gradients_into(self, grad_point_cloud: Optional[torch.Tensor],
                                  grad_point_cloud_features: Optional[torch.Tensor]) -> None:
        """Extension for multi-view batches (not in the reference, which renders one view per step).

        While set, every backward ADDS its point gradients into the two given float32 buffers
        ([N,3] and [N,56], contiguous, zeroed by the caller once per batch) and returns no gradient
        to autograd for those two inputs -- so a batch of views costs neither a zero fill nor an
        AccumulateGrad pass per view.  Call with (None, None) to restore the reference behaviour.
        The target is captured by each forward call: a backward always adds into the buffers that
        were set when ITS forward ran.  The rows are written as 16-byte vectors, so both buffers
        must start on a 16-byte boundary (a flat N x 59 buffer must hold the N x 56 block first).
        """
        if grad_point_cloud is None or grad_point_cloud_features is None:
            self._accumulate_into = None
            return
        for t, cols in ((grad_point_cloud, 3), (grad_point_cloud_features, 56)):
            if not (t.is_cuda and t.dtype == torch.float32 and t.dim() == 2 and t.shape[1] == cols
                    and t.is_contiguous()):
                raise ValueError("gradient buffers must be contiguous CUDA float32 [N,3] and [N,56]")
        if grad_point_cloud_features.data_ptr() % 16:
            raise ValueError("grad_point_cloud_features must be 16-byte aligned (rows are added with "
                             "red.global.add.v4.f32); put the N x 56 block first in a flat buffer")
        self._accumulate_into = (grad_point_cloud, grad_point_cloud_features)

    def begin_gradient_batch(self, grad_point_cloud: torch.Tensor, grad_point_cloud_features: torch.Tensor) -> None:
        """Extension for multi-view batches, deferred form (not in the reference).  Until
        `end_gradient_batch`, every backward of this module runs its blend backward only and PARKS the
        view's reduced screen-space gradients (48 bytes per visible point); no gradient is returned to
        autograd for the two point inputs.  `end_gradient_batch` then runs the per-point chain rule of
        ALL parked views in one pass over the points (gs_point_bwd_batch) and STORES the summed gradient
        rows into the two buffers ([N,3] and [N,56], float32, contiguous, 16-byte aligned; they need not
        be zeroed): each feature row is read once per batch instead of once per view, and nothing is
        read-modify-written.  Not available with a backward_valid_point_hook (its payload is per view)."""
        if self._has_hook:
            raise ValueError("gradient batches do not produce the per-view hook payload: use "
                             "accumulate_gradients_into with a backward_valid_point_hook")
        self.accumulate_gradients_into(grad_point_cloud, grad_point_cloud_features)   # validates
        self._accumulate_into = None
        self._batch = _GradientBatch(grad_point_cloud, grad_point_cloud_features)

    def end_gradient_batch(self, chunks: int = 1,
                           after_chunk: Optional[Callable[[int, int, int], None]] = None,
                           exchange: Any = None) -> int:
        """Runs the parked views' per-point backward on the current stream (the caller has made it wait
        for the streams the views ran on) and leaves batch mode.  With chunks > 1 the points are cut into
        that many id ranges, one launch each, and `after_chunk(index, first_point, num_points)` is called
        after each launch -- e.g. to start the all-reduce of that range while the next one is computed.
        exchange: a distributed.PeerExchange whose buffers the batch was begun on -- the pass and the sum of
        the gradient over the ranks then run as ONE kernel (gs_point_bwd_batch_exchange: ranges are reduced
        in the NVLink switch while later ranges are computed); every rank must have parked at least one view.
        Returns the number of views that were in the batch."""
        batch, self._batch = self._batch, None
        if batch is None:
            raise RuntimeError("end_gradient_batch without begin_gradient_batch")
        return batch.finish(self.config, max(1, int(chunks)), after_chunk, exchange)

    def forward(self, input_data: "GaussianPointCloudRasterisation.GaussianPointCloudRasterisationInput"):
        camera_info = input_data.camera_info
        assert camera_info.camera_width % TILE_WIDTH == 0
        assert camera_info.camera_height % TILE_HEIGHT == 0
        return self._module_function.apply(
            input_data.point_cloud,
            input_data.point_cloud_features,
            input_data.point_invalid_mask,
            input_data.point_object_id,
            input_data.q_pointcloud_camera,
            input_data.t_pointcloud_camera,
            camera_info,
            input_data.color_max_sh_band,
        )


class _GradientBatch:
    """The views parked between begin_gradient_batch and end_gradient_batch."""

    def __init__(self, grad_point_cloud: torch.Tensor, grad_point_cloud_features: torch.Tensor):
        self.grad_xyz, self.grad_features = grad_point_cloud, grad_point_cloud_features
        self.views = []       # per view: the tensors gs_view_t points into (kept alive until finish)
        self.common = None    # (xyz, features, object_id, colour band) -- the same for every view

    def park(self, pointcloud, pointcloud_features, point_object_id, color_max_sh_band, st, splat_grads,
             t_pc) -> None:
        common = (pointcloud.data_ptr(), pointcloud_features.data_ptr(), point_object_id.data_ptr(),
                  int(color_max_sh_band), pointcloud.shape[0])
        if self.common is None:
            self.common = common
            self.tensors = (pointcloud, pointcloud_features, point_object_id)
        elif common != self.common:
            raise ValueError("the views of a gradient batch must share the point tensors and the colour band")
        buf = st.buffers
        # views of the forward's per-point buffer (80 B per point of capacity stay alive until the step's
        # pass; copying the two arrays out cost 0.02 ms per view at 1M points)
        self.views.append(dict(slot=buf.view("slot_of_point", st.n),
                               pic=buf.view("point_in_camera", max(st.m, 1)),
                               sg=splat_grads, K=st.intrinsics, q_cam=st.q_cam, t_pc=t_pc))

    def finish(self, cfg, chunks: int, after_chunk, exchange=None) -> int:
        n = self.grad_xyz.shape[0]
        dev = self.grad_xyz.device
        cur = torch.cuda.current_stream(dev)
        if exchange is not None:
            if self.common is None:
                raise _capi.GsError("the fused exchange needs at least one view of the batch on every rank")
            if (self.grad_features.data_ptr() != exchange.grad_features.data_ptr()
                    or self.grad_xyz.data_ptr() != exchange.grad_xyz.data_ptr() or exchange.n != n):
                raise ValueError("the gradient batch was not begun on the exchange's buffers")
        if self.common is None:            # no view on this rank: the gradient is zero
            self.grad_xyz.zero_()
            self.grad_features.zero_()
            if after_chunk is not None:
                after_chunk(0, 0, n)
            return 0
        xyz, features, object_id = self.tensors
        rows = []
        for v in self.views:
            rows.append([v["slot"].data_ptr(), v["pic"].data_ptr(), v["sg"].data_ptr(), v["K"].data_ptr(),
                         v["q_cam"].data_ptr(), v["t_pc"].data_ptr(), 0, 0])
            for t in v.values():
                t.record_stream(cur)       # produced on the view's lane, consumed here
        table = torch.tensor(rows, dtype=torch.int64).pin_memory().to(dev, non_blocking=True)
        band = self.common[3]
        if exchange is not None:
            _capi.check(_lib.gs_point_bwd_batch_exchange(
                xyz.detach().contiguous().data_ptr(), features.detach().data_ptr(),
                object_id.contiguous().data_ptr(), n, table.data_ptr(), len(rows), band,
                float(cfg.grad_q_factor), float(cfg.grad_s_factor), float(cfg.grad_alpha_factor),
                float(cfg.grad_color_factor), float(cfg.grad_high_order_color_factor),
                exchange.next_call(), cur.cuda_stream), "gs_point_bwd_batch_exchange")
            _mark("point_bwd_batch_exchange")
            num = len(self.views)
            self.views = []
            return num
        chunks = max(1, min(chunks, (n + 127) // 128))
        per = ((n + chunks - 1) // chunks + 127) // 128 * 128
        first, index = 0, 0
        while first < n:
            count = min(per, n - first)
            _capi.check(_lib.gs_point_bwd_batch(
                xyz.detach().contiguous().data_ptr(), features.detach().data_ptr(),
                object_id.contiguous().data_ptr(), first, count, table.data_ptr(), len(rows), band,
                float(cfg.grad_q_factor), float(cfg.grad_s_factor), float(cfg.grad_alpha_factor),
                float(cfg.grad_color_factor), float(cfg.grad_high_order_color_factor),
                self.grad_xyz.data_ptr(), self.grad_features.data_ptr(), cur.cuda_stream), "gs_point_bwd_batch")
            if after_chunk is not None:
                after_chunk(index, first, count)
            first += count
            index += 1
        _mark("point_bwd_batch")
        num = len(self.views)
        self.views = []
        return num


_READBACK: Dict[int, tuple] = {}  # device index -> (side stream, pinned int64[2])


def _read_totals(totals: torch.Tensor, ready: "torch.cuda.Event"):
    """{M, K} to the host without waiting for work enqueued after `ready` (the point sort): the
    copy runs on a side stream that waits only for the projection kernel."""
    dev = totals.device
    key = dev.index if dev.index is not None else torch.cuda.current_device()
    if key not in _READBACK:
        _READBACK[key] = (torch.cuda.Stream(dev), torch.empty(2, dtype=torch.int64).pin_memory())
    side, host = _READBACK[key]
    with torch.cuda.stream(side):
        side.wait_event(ready)
        host.copy_(totals, non_blocking=True)
        done = torch.cuda.Event()
        done.record(side)
    totals.record_stream(side)
    done.synchronize()
    return int(host[0]), int(host[1])


class _ForwardState:
    """Everything the backward needs, plus handles a test can read every stage from."""
    __slots__ = ("n", "m", "k", "width", "height", "buffers", "payload", "keys", "tile_start",
                 "tile_end", "image", "depth", "count", "acc_alpha", "last", "q_cam", "t_cam",
                 "intrinsics", "info", "unsorted_keys", "unsorted_payload", "tile_ids", "cell_masks",
                 "quarter_flags")

    def sorted_keys(self) -> torch.Tensor:
        """The reference's sorted 64-bit keys (point_in_camera_sort_key, R.py:947) of this forward."""
        if self.keys is not None:
            return self.keys
        keys = torch.empty(self.k, dtype=torch.int64, device=self.payload.device)
        if self.k > 0 and self.tile_ids is None:
            # binned pipeline: the lists are contiguous in tile order, so the ids follow from the ranges
            lengths = (self.tile_end - self.tile_start).to(torch.int64)
            self.tile_ids = torch.repeat_interleave(
                torch.arange(lengths.numel(), dtype=torch.int32, device=lengths.device), lengths)
        if self.k > 0:
            stream = torch.cuda.current_stream(keys.device).cuda_stream
            _capi.check(_lib.gs_sorted_keys(self.tile_ids.data_ptr(), self.payload.data_ptr(),
                                            self.buffers.ptr("depth_key"), self.k, keys.data_ptr(),
                                            stream), "gs_sorted_keys")
        return keys


def _check_inputs(pointcloud, features, invalid_mask, object_id, q, t, intrinsics):
    if not pointcloud.is_cuda:
        raise _capi.GsError("GaussianPointCloudRasterisation needs CUDA tensors (no CPU path exists)")
    n = pointcloud.shape[0]
    if pointcloud.dtype != torch.float32 or pointcloud.dim() != 2 or pointcloud.shape[1] != 3:
        raise ValueError("point_cloud must be float32 [N,3]")
    if features.dtype != torch.float32 or features.shape != (n, 56):
        raise ValueError("point_cloud_features must be float32 [N,56]")
    if not features.is_contiguous():
        raise ValueError("point_cloud_features must be contiguous (it is normalised in place)")
    if invalid_mask.dtype != torch.int8 or invalid_mask.shape[0] != n:
        raise ValueError("point_invalid_mask must be int8 [N]")
    if object_id.dtype != torch.int32 or object_id.shape[0] != n:
        raise ValueError("point_object_id must be int32 [N]")
    if q.dim() != 2 or q.shape[1] != 4 or t.dim() != 2 or t.shape[1] != 3 or q.shape[0] != t.shape[0]:
        raise ValueError("q_pointcloud_camera must be [K,4] and t_pointcloud_camera [K,3]")
    if tuple(intrinsics.shape) != (3, 3):
        raise ValueError("camera_intrinsics must be 3x3")


def _forward_impl(cfg, pointcloud, pointcloud_features, point_invalid_mask, point_object_id,
                  q_pointcloud_camera, t_pointcloud_camera, camera_info,
                  keep_unsorted: bool = False, keep_cell_masks: bool = False,
                  hook_outputs: bool = False, debug_outputs: bool = False,
                  offsets_output: bool = False, slots_output: bool = False) -> _ForwardState:
    """hook_outputs: also produce point_uv / num_overlap_tiles (a backward hook will show them);
    debug_outputs: every reference-named per-point buffer (tests); offsets_output: the exclusive scan
    of the tile counts even when the key pipeline in use does not read it (tests)."""
    CALL_COUNTS["forward"] += 1
    dev = pointcloud.device
    intrinsics = camera_info.camera_intrinsics
    _check_inputs(pointcloud, pointcloud_features, point_invalid_mask, point_object_id,
                  q_pointcloud_camera, t_pointcloud_camera, intrinsics)
    W, H = int(camera_info.camera_width), int(camera_info.camera_height)
    if W % TILE_WIDTH or H % TILE_HEIGHT:
        raise ValueError("camera width and height must be multiples of 16")
    n = pointcloud.shape[0]
    n_obj = q_pointcloud_camera.shape[0]
    xyz = pointcloud.detach().contiguous()
    feat = pointcloud_features.detach()
    inv = point_invalid_mask.contiguous()
    oid = point_object_id.contiguous()
    Kmat = intrinsics.detach().to(device=dev, dtype=torch.float32).contiguous()
    q_pc = q_pointcloud_camera.detach().to(device=dev, dtype=torch.float32).contiguous()
    t_pc = t_pointcloud_camera.detach().to(device=dev, dtype=torch.float32).contiguous()
    s_big = torch.cuda.current_stream(dev)
    s_small = _companion_stream(s_big) if PRIORITY_SECTIONS else s_big
    _hand_over(s_big, s_small)
    stream = s_small.cuda_stream  # everything up to the tile lists; the blend runs on s_big

    st = _ForwardState()
    st.n, st.width, st.height, st.intrinsics = n, W, H, Kmat
    st.unsorted_keys = st.unsorted_payload = None
    _mark("forward_begin")
    st.q_cam = torch.empty_like(q_pc)
    st.t_cam = torch.empty_like(t_pc)
    _capi.check(_lib.gs_inverse_pose(q_pc.data_ptr(), t_pc.data_ptr(), n_obj, st.q_cam.data_ptr(),
                                     st.t_cam.data_ptr(), stream), "gs_inverse_pose")

    num_tiles = (W // TILE_WIDTH) * (H // TILE_HEIGHT)
    depth_bits, tile_bits = sort_key_bits(float(cfg.near_plane), float(cfg.far_plane),
                                          float(cfg.depth_to_sort_key_scale), num_tiles)
    depth_first = KEY_PIPELINE in ("depth_first", "binned") and depth_bits <= 31
    binned = (depth_first and KEY_PIPELINE == "binned" and W // TILE_WIDTH <= 256
              and H // TILE_HEIGHT <= 256)
    # the per-point scan of the tile counts is read by the literal key emission only (binned falls back
    # to depth_first, never to literal)
    buf = _PointBuffers(n, dev, hook=hook_outputs, debug=debug_outputs,
                        offsets=offsets_output or not depth_first, slots=slots_output)
    st.buffers = buf
    totals = torch.empty(2, dtype=torch.int64, device=dev)
    ws = torch.empty(max(int(_lib.gs_project_workspace_bytes(n)), 16), dtype=torch.uint8, device=dev)
    _capi.check(_lib.gs_project_cull_sh(
        xyz.data_ptr(), feat.data_ptr(), inv.data_ptr(), oid.data_ptr(), n, Kmat.data_ptr(),
        st.q_cam.data_ptr(), st.t_cam.data_ptr(), n_obj, float(cfg.near_plane),
        float(cfg.far_plane), W, H, float(cfg.depth_to_sort_key_scale),
        buf.ptr("ids"), buf.ptr("point_uv"), buf.ptr("point_in_camera"), buf.ptr("conic"),
        buf.ptr("alpha"), buf.ptr("color"), buf.ptr("radii"), buf.ptr("num_tiles"),
        buf.ptr("tile_offsets"), buf.ptr("records"), buf.ptr("bbox"), buf.ptr("depth_key"),
        buf.ptr("slot_of_point"), totals.data_ptr(), ws.data_ptr(), stream), "gs_project_cull_sh")
    _mark("project_cull_sh")
    after_project = torch.cuda.Event()
    after_project.record(s_small)
    order = None
    if depth_first and n > 0:
        # The argsort of the visible points by depth needs M only on the device: enqueue it for
        # capacity N now, so the GPU works while the host waits for {M, K} below.
        order = torch.empty(n, dtype=torch.int32, device=dev)
        pts_ws = torch.empty(int(_lib.gs_point_sort_workspace_bytes(n)), dtype=torch.uint8, device=dev)
        _capi.check(_lib.gs_sort_points_by_depth(buf.ptr("depth_key"), n, totals.data_ptr(), depth_bits,
                                                 order.data_ptr(), pts_ws.data_ptr(), stream),
                    "gs_sort_points_by_depth")
        _mark("sort_points")
    # everything that does not depend on {M, K} is allocated while the GPU is still busy
    st.tile_start = torch.empty(num_tiles, dtype=torch.int32, device=dev)
    st.tile_end = torch.empty(num_tiles, dtype=torch.int32, device=dev)
    st.image = torch.empty(H, W, 3, dtype=torch.float32, device=dev)
    rgb_only = bool(cfg.rgb_only)
    if rgb_only:
        # the reference leaves these uninitialised (R.py:478-484); zeros are the safe superset
        st.depth = torch.zeros(H, W, dtype=torch.float32, device=dev)
        st.acc_alpha = torch.zeros(H, W, dtype=torch.float32, device=dev)
        st.last = torch.zeros(H, W, dtype=torch.int32, device=dev)
        st.count = torch.zeros(H, W, dtype=torch.int32, device=dev)
    else:
        st.depth = torch.empty(H, W, dtype=torch.float32, device=dev)
        st.acc_alpha = torch.empty(H, W, dtype=torch.float32, device=dev)
        st.last = torch.empty(H, W, dtype=torch.int32, device=dev)
        st.count = torch.empty(H, W, dtype=torch.int32, device=dev)

    m, k = _read_totals(totals, after_project)  # the one host sync of the forward
    st.m, st.k = m, k

    kk = max(k, 1)
    st.tile_ids = None
    if binned and k < (1 << 30):
        # points sorted by depth -> split over tile columns -> split over tile rows
        st.keys = None
        st.payload = torch.empty(kk, dtype=torch.int32, device=dev)[:k]
        if k > 0:
            entries = torch.empty(int(_lib.gs_bin_entries_capacity(m, k, W)), dtype=torch.int64, device=dev)
            bin_ws = torch.empty(int(_lib.gs_bin_workspace_bytes(m, k, W, H)), dtype=torch.uint8, device=dev)
        _mark("host_sync_and_alloc")
        _capi.check(_lib.gs_bin_tiles(order.data_ptr() if k > 0 else None, buf.ptr("bbox"), m, k, W, H,
                                      entries.data_ptr() if k > 0 else None,
                                      st.payload.data_ptr() if k > 0 else None,
                                      st.tile_start.data_ptr(), st.tile_end.data_ptr(),
                                      bin_ws.data_ptr() if k > 0 else None, stream), "gs_bin_tiles")
        _mark("bin_tiles")
    elif depth_first:
        # points sorted by depth -> tile ids emitted in that order -> stable sort by tile id only
        tid2 = torch.empty(2 * kk, dtype=torch.int32, device=dev)
        pay2 = torch.empty(2 * kk, dtype=torch.int32, device=dev)
        ws_bytes = max(int(_lib.gs_sort_workspace_bytes(k)), int(_lib.gs_emit_workspace_bytes(m)), 16)
        sort_ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        _mark("host_sync_and_alloc")
        st.keys = None
        if k > 0:
            # the emission counts the tile-id digits while it writes them: the pair sort starts
            # from that histogram instead of re-reading the K ids
            digit_hist = torch.empty(4 * 256, dtype=torch.int32, device=dev)
            _capi.check(_lib.gs_emit_tile_keys(order.data_ptr(), buf.ptr("bbox"), m, k, W,
                                               tid2.data_ptr(), pay2.data_ptr(), sort_ws.data_ptr(),
                                               digit_hist.data_ptr(), max(tile_bits, 1), stream),
                        "gs_emit_tile_keys")
            _mark("emit_keys")
            in_alt = C.c_int(0)
            _capi.check(_lib.gs_sort_pairs32(tid2.data_ptr(), pay2.data_ptr(),
                                             tid2.data_ptr() + 4 * kk, pay2.data_ptr() + 4 * kk, k,
                                             tile_bits, sort_ws.data_ptr(), digit_hist.data_ptr(),
                                             C.byref(in_alt), stream),
                        "gs_sort_pairs32")
            _mark("sort_pairs")
            half = slice(kk, kk + k) if in_alt.value else slice(0, k)
            st.tile_ids, st.payload = tid2[half], pay2[half]
        else:
            st.tile_ids, st.payload = tid2[:0], pay2[:0]
        _capi.check(_lib.gs_tile_ranges32(st.tile_ids.data_ptr() if k > 0 else None, k,
                                          st.tile_start.data_ptr(), st.tile_end.data_ptr(), num_tiles,
                                          stream), "gs_tile_ranges32")
        _mark("tile_ranges")
    else:
        # literal pipeline: the reference's 64-bit keys in emission order, full stable sort
        keys2 = torch.empty(2 * kk, dtype=torch.int64, device=dev)
        pay2 = torch.empty(2 * kk, dtype=torch.int32, device=dev)
        keys_a, keys_b = keys2[:kk], keys2[kk:]
        pay_a, pay_b = pay2[:kk], pay2[kk:]
        _mark("host_sync_and_alloc")
        if k > 0:
            _capi.check(_lib.gs_emit_keys(buf.ptr("bbox"), buf.ptr("depth_key"),
                                          buf.ptr("tile_offsets"), m, k, W, keys_a.data_ptr(),
                                          pay_a.data_ptr(), stream), "gs_emit_keys")
            _mark("emit_keys")
            if keep_unsorted:
                st.unsorted_keys, st.unsorted_payload = keys_a[:k].clone(), pay_a[:k].clone()
            sort_ws = torch.empty(int(_lib.gs_sort_workspace_bytes(k)), dtype=torch.uint8, device=dev)
            in_alt = C.c_int(0)
            _capi.check(_lib.gs_sort_pairs(keys_a.data_ptr(), pay_a.data_ptr(), keys_b.data_ptr(),
                                           pay_b.data_ptr(), k, depth_bits, tile_bits,
                                           sort_ws.data_ptr(), C.byref(in_alt), stream),
                        "gs_sort_pairs")
            _mark("sort_pairs")
            st.keys = (keys_b if in_alt.value else keys_a)[:k]
            st.payload = (pay_b if in_alt.value else pay_a)[:k]
        else:
            st.keys, st.payload = keys_a[:0], pay_a[:0]
        _capi.check(_lib.gs_tile_ranges(st.keys.data_ptr() if k > 0 else None, k,
                                        st.tile_start.data_ptr(), st.tile_end.data_ptr(), num_tiles,
                                        stream), "gs_tile_ranges")
        _mark("tile_ranges")
    # what the backward blend takes over from the forward blend instead of redoing the arithmetic: the
    # exact culling bits per (list entry, 8x4 cell) -- four planes (one per 8x8 quarter of a tile) of k
    # bytes -- and one flag per quarter (does it reach an entry with an irregular conic)
    if keep_cell_masks:
        st.cell_masks = torch.empty(max(4 * k, 1), dtype=torch.uint8, device=dev)
        st.quarter_flags = torch.empty(4 * num_tiles, dtype=torch.uint8, device=dev)
    else:
        st.cell_masks = st.quarter_flags = None
    _hand_over(s_small, s_big)
    stream = s_big.cuda_stream
    _capi.check(_lib.gs_blend_fwd(H, W, st.tile_start.data_ptr(), st.tile_end.data_ptr(),
                                  st.payload.data_ptr() if k > 0 else None, buf.ptr("records"),
                                  st.image.data_ptr(), st.depth.data_ptr(), st.acc_alpha.data_ptr(),
                                  st.last.data_ptr(), st.count.data_ptr(), int(rgb_only),
                                  _ptr(st.cell_masks), k, _ptr(st.quarter_flags), stream),
                "gs_blend_fwd")
    _mark("blend_fwd")
    st.info = dict(num_points=n, num_points_in_camera=m, num_keys=k, depth_bits=depth_bits,
                   tile_bits=tile_bits,
                   key_pipeline=("binned" if binned else "depth_first") if depth_first else "literal")
    return st


def _backward_impl(cfg, st: _ForwardState, pointcloud, pointcloud_features, point_object_id,
                   t_pointcloud_camera, camera_info, color_max_sh_band, grad_rasterized_image,
                   hook, hook_input_type, accumulate_into=None, batch=None):
    CALL_COUNTS["backward"] += 1
    store_all = False
    dev = pointcloud.device
    n, m, W, H = st.n, st.m, st.width, st.height
    s_big = torch.cuda.current_stream(dev)
    s_small = _companion_stream(s_big) if PRIORITY_SECTIONS else s_big
    stream = s_big.cuda_stream
    _mark("backward_begin")
    if batch is not None:
        grad_pointcloud = grad_features = None   # written by end_gradient_batch
    elif accumulate_into is not None:
        grad_pointcloud, grad_features = accumulate_into
        if grad_pointcloud.shape[0] != n or grad_features.shape[0] != n:
            raise ValueError("accumulation buffers do not match the point cloud")
    elif (STORE_ALL_ROWS and hook is None and m > 0 and st.buffers.has("slot_of_point")
          and pointcloud_features.data_ptr() % 16 == 0):
        # plain backward (what the unchanged trainer runs): every row is stored by gs_point_bwd_all_points,
        # zeros for the points this view does not see -- no zero fill of the N x 59 floats (R.py:1051-1058)
        store_all = True
        grad_pointcloud = torch.empty_like(pointcloud, memory_format=torch.contiguous_format)
        grad_features = torch.empty_like(pointcloud_features, memory_format=torch.contiguous_format)
    else:
        grad_pointcloud = torch.zeros_like(pointcloud)
        grad_features = torch.zeros_like(pointcloud_features, memory_format=torch.contiguous_format)
    if grad_rasterized_image is None:
        grad_rasterized_image = torch.zeros(H, W, 3, dtype=torch.float32, device=dev)
    g_img = grad_rasterized_image.detach().to(torch.float32).contiguous()
    want_hook = hook is not None
    # |d uv| magnitudes (R.py:700-704) are visible only through the hook; without one the kernel
    # skips them (map_on_image = NULL)
    on_image = torch.empty(H, W, 2, dtype=torch.float32, device=dev) if want_hook else None
    splat_grads = torch.zeros(max(m, 1), _GRAD_SLOTS, dtype=torch.float32, device=dev)
    buf = st.buffers
    if st.cell_masks is None:
        raise _capi.GsError("backward of a forward that ran without gradient tracking (no culling bits kept)")
    scratch = torch.empty(max(int(_lib.gs_blend_bwd_scratch_bytes(H, W)), 1), dtype=torch.uint8, device=dev)
    _mark("zero_fill")
    _capi.check(_lib.gs_blend_bwd(H, W, st.tile_start.data_ptr(), st.tile_end.data_ptr(),
                                  st.payload.data_ptr() if st.k > 0 else None, buf.ptr("records"),
                                  g_img.data_ptr(), st.acc_alpha.data_ptr(), st.last.data_ptr(),
                                  splat_grads.data_ptr(), _ptr(on_image), st.cell_masks.data_ptr(), st.k,
                                  st.quarter_flags.data_ptr(), scratch.data_ptr(), stream),
                "gs_blend_bwd")
    _mark("blend_bwd")
    if batch is not None:
        t_pc = t_pointcloud_camera.detach().to(device=dev, dtype=torch.float32).contiguous()
        batch.park(pointcloud, pointcloud_features, point_object_id, color_max_sh_band, st, splat_grads, t_pc)
        _mark("park_view")
        return None, None
    h_xyz = torch.empty(m, 3, dtype=torch.float32, device=dev) if want_hook else None
    h_feat = torch.empty(m, 56, dtype=torch.float32, device=dev) if want_hook else None
    h_view = torch.empty(m, 2, dtype=torch.float32, device=dev) if want_hook else None
    h_mag = torch.empty(m, dtype=torch.float32, device=dev) if want_hook else None
    h_aff = torch.empty(m, dtype=torch.int32, device=dev) if want_hook else None
    t_pc = t_pointcloud_camera.detach().to(device=dev, dtype=torch.float32).contiguous()
    if store_all:
        view = _capi.gs_view_t()
        view.slot_of_point, view.point_in_camera = buf.ptr("slot_of_point"), buf.ptr("point_in_camera")
        view.splat_grads, view.intrinsics = splat_grads.data_ptr(), st.intrinsics.data_ptr()
        view.q_cam_pc, view.t_pc_cam = st.q_cam.data_ptr(), t_pc.data_ptr()
        _hand_over(s_big, s_small)
        import ctypes as _C
        _capi.check(_lib.gs_point_bwd_all_points(
            pointcloud.detach().contiguous().data_ptr(), pointcloud_features.detach().data_ptr(),
            point_object_id.contiguous().data_ptr(), n, _C.byref(view), int(color_max_sh_band),
            float(cfg.grad_q_factor), float(cfg.grad_s_factor), float(cfg.grad_alpha_factor),
            float(cfg.grad_color_factor), float(cfg.grad_high_order_color_factor),
            grad_pointcloud.data_ptr(), grad_features.data_ptr(), s_small.cuda_stream), "gs_point_bwd_all_points")
        _hand_over(s_small, s_big)
    elif m > 0:
        _hand_over(s_big, s_small)
        stream = s_small.cuda_stream
        _capi.check(_lib.gs_point_bwd(
            pointcloud.detach().contiguous().data_ptr(), pointcloud_features.detach().data_ptr(),
            point_object_id.contiguous().data_ptr(), st.intrinsics.data_ptr(), st.q_cam.data_ptr(),
            st.t_cam.data_ptr(), t_pc.data_ptr(), st.q_cam.shape[0], buf.ptr("ids"), m,
            buf.ptr("point_in_camera"), splat_grads.data_ptr(), int(color_max_sh_band),
            float(cfg.grad_q_factor), float(cfg.grad_s_factor), float(cfg.grad_alpha_factor),
            float(cfg.grad_color_factor), float(cfg.grad_high_order_color_factor),
            grad_pointcloud.data_ptr(), grad_features.data_ptr(), _ptr(h_xyz), _ptr(h_feat),
            _ptr(h_view), _ptr(h_mag), _ptr(h_aff), int(accumulate_into is not None), stream),
            "gs_point_bwd")
        _hand_over(s_small, s_big)
    _mark("point_bwd")
    if want_hook:
        CALL_COUNTS["hook"] += 1
        point_in_camera = buf.view("point_in_camera", m)
        hook(hook_input_type(
            point_id_in_camera_list=buf.view("ids", m),
            grad_point_in_camera=h_xyz,
            grad_pointfeatures_in_camera=h_feat,
            grad_viewspace=h_view,
            magnitude_grad_viewspace=h_mag,
            magnitude_grad_viewspace_on_image=on_image,
            num_overlap_tiles=buf.view("num_tiles", m),
            num_affected_pixels=h_aff,
            point_depth=point_in_camera[:, 2],
            point_uv_in_camera=buf.view("point_uv", m),
        ))
    if accumulate_into is not None:
        return None, None  # already added into the caller's buffers
    return grad_pointcloud, grad_features


def install() -> None:
    """Make the reference package use this operator: GaussianPointTrainer, gaussian_point_render.py,
    visualizer / benchmark scripts and the reference's tests then run unchanged.

        import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as drop_in
        drop_in.install()          # any time: before or after the callers are imported

    What it does, in the already importable reference package (raises ImportError otherwise):
      * `taichi_3d_gaussian_splatting.GaussianPointCloudRasterisation.GaussianPointCloudRasterisation`
        and `.find_tile_start_and_end` become this module's;
      * every loaded module that already did `from ... import GaussianPointCloudRasterisation` (the
        trainer, the renderer, a test module) is re-pointed too, so the call order does not matter;
      * this class ADOPTS the reference's own nested record types (`...Config`, `...Input`,
        `BackwardValidPointHookInput`): a `TrainConfig` built earlier (its `default_factory`,
        GaussianPointTrainer.py:56-57), a YAML round trip through the reference's YAMLWizard base and
        `isinstance` checks in the callers all keep working on the very same classes.
    Idempotent."""
    import importlib
    import sys
    ref = importlib.import_module("taichi_3d_gaussian_splatting.GaussianPointCloudRasterisation")
    old_cls = ref.GaussianPointCloudRasterisation
    old_find = getattr(ref, "find_tile_start_and_end", None)
    if old_cls is GaussianPointCloudRasterisation:
        return
    for name in ("GaussianPointCloudRasterisationConfig", "GaussianPointCloudRasterisationInput",
                 "BackwardValidPointHookInput"):
        setattr(GaussianPointCloudRasterisation, name, getattr(old_cls, name))
    ref.GaussianPointCloudRasterisation = GaussianPointCloudRasterisation
    ref.find_tile_start_and_end = find_tile_start_and_end
    for mod in list(sys.modules.values()):
        names = getattr(mod, "__dict__", None)
        if not isinstance(names, dict) or mod is sys.modules[__name__]:
            continue
        for attr, val in list(names.items()):
            if val is old_cls:
                names[attr] = GaussianPointCloudRasterisation
            elif old_find is not None and val is old_find:
                names[attr] = find_tile_start_and_end
