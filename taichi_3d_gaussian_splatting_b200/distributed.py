"""View-sharded multi-GPU step (new functionality; the reference is single-GPU, README.md:325).

The path shards over camera views: every rank holds the whole point cloud, renders its share of a
batch's views forward+backward with the single-view operator, and the one exchange step is the sum of
the N x 56 + N x 3 gradient over ranks (`all_reduce`, NCCL over NVLink on GPUs; gloo in the CPU
tests).  One process per GPU, launched with torch.distributed.run.

Which rank renders which view: round robin (`views_of_rank`) by default; with per-view costs from a
previous pass (`assign_views_lpt`, longest-processing-time first on K + 3 M: list entries and visible
points) the ranks' loads differ by a few per cent instead of by the spread of the views, and a step
ends when its SLOWEST rank does.

The operator normalises the quaternion columns of the visible points in place (reference
R.py:264-266), and different ranks see different points, so the replicas would drift apart by
rounding in those four columns; `sync_quaternions` re-broadcasts them from rank 0 at the start of a
step (16 B per point, against 236 B per point for the gradient all-reduce).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def views_of_rank(num_views: int, rank: int, world: int) -> List[int]:
    """Indices of the views rank `rank` renders: v = rank (mod world)."""
    return list(range(rank, num_views, world))


def view_cost(num_points_in_camera: float, num_keys: float) -> float:
    """Work of one view forward+backward, in list-entry equivalents: the blend and binning kernels go
    with the K list entries, the projection and the per-point backward with the M visible points (stage
    times at 2M points, 976x544: ~0.07 ns per entry, ~0.23 ns per point)."""
    return float(num_keys) + 3.0 * float(num_points_in_camera)


def assign_views_lpt(costs: Sequence[float], world: int) -> Tuple[List[List[int]], List[float]]:
    """Longest-processing-time-first: views by decreasing cost, each to the least loaded rank.
    Deterministic, so every rank computes the same assignment from the same (all-reduced) costs.
    Returns (views per rank, load per rank)."""
    loads = [0.0] * world
    mine: List[List[int]] = [[] for _ in range(world)]
    for v in sorted(range(len(costs)), key=lambda i: (-costs[i], i)):
        r = min(range(world), key=lambda i: (loads[i], len(mine[i]), i))
        mine[r].append(v)
        loads[r] += costs[v]
    return mine, loads


class PeerExchange:
    """The buffers of the fused per-point pass + exchange (gs_point_bwd_batch_exchange, include/gs_b200.h):
    the flat N x 59 gradient as SYMMETRIC memory (one allocation per rank, mapped into every peer, with an
    NVLink multicast address) and a symmetric flag array, set up with torch.distributed._symmetric_memory
    (plumbing: allocation and the exchange of the mappings; the reduction itself is this package's kernel).
    Collective: every rank of `group` constructs it at the same point.  Raises if the devices have no
    multicast mapping (no NVSwitch, or one GPU): the caller then keeps NCCL's all_reduce.

    grad_features [N,56] / grad_xyz [N,3] are views of `flat` (features first, as everywhere); pass them to
    begin_gradient_batch and the object to end_gradient_batch(exchange=...)."""

    def __init__(self, n: int, device: torch.device, group: Optional[dist.ProcessGroup] = None,
                 chunks: int = 4, reducer_ctas: int = 0):
        import ctypes as C

        import torch.distributed._symmetric_memory as symm_mem

        from . import _capi
        group = group if group is not None else dist.group.WORLD
        self.n, self.group = int(n), group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world < 2 or self.world > 16:
            raise RuntimeError("the fused exchange is for 2..16 ranks")
        blocks = (self.n + 127) // 128
        per = (blocks + max(1, chunks) - 1) // max(1, chunks)
        self.chunks = (blocks + per - 1) // per           # the library's own rule (every range >= one block)
        numel = (59 * self.n + 3) // 4 * 4 + 4            # the position segment's last 16-byte vector may overhang
        self._all = symm_mem.empty(numel, dtype=torch.float32, device=device)
        self._all.zero_()
        self._handle = symm_mem.rendezvous(self._all, group.group_name)
        if not getattr(self._handle, "multicast_ptr", 0):
            raise RuntimeError("no NVLink multicast mapping for symmetric memory on this system")
        words = int(_capi.lib.gs_exchange_flag_words(self.chunks, self.world))
        self._flags = symm_mem.empty(words, dtype=torch.int32, device=device)
        self._flags.zero_()
        self._flag_handle = symm_mem.rendezvous(self._flags, group.group_name)
        self._scratch = torch.zeros(int(_capi.lib.gs_exchange_scratch_bytes(self.chunks)) // 4,
                                    dtype=torch.int32, device=device)
        torch.cuda.synchronize(device)
        dist.barrier(group)                               # every rank's flags are zero before anyone writes
        self.flat = self._all[:59 * self.n]
        self.grad_features = self.flat[:56 * self.n].view(self.n, 56)
        self.grad_xyz = self.flat[56 * self.n:].view(self.n, 3)
        self.epoch = 0
        x = _capi.gs_exchange_t()
        x.multicast_base = int(self._handle.multicast_ptr)
        x.local_base = self._all.data_ptr()
        ptrs = list(self._flag_handle.buffer_ptrs)
        for r in range(self.world):
            x.flags[r] = int(ptrs[r])
        x.rank, x.world, x.chunks, x.reducer_ctas = self.rank, self.world, self.chunks, int(reducer_ctas)
        x.scratch = self._scratch.data_ptr()
        self._struct, self._byref = x, C.byref(x)

    def next_call(self):
        """The gs_exchange_t of the next launch (a new epoch: the flags are never reset)."""
        self.epoch += 1
        self._struct.epoch = self.epoch
        return self._byref

    def error(self) -> int:
        """Non-zero if a launch gave up waiting for a peer (synchronises)."""
        return int(self._scratch[1].item())


@dataclass
class StepResult:
    loss: torch.Tensor            # sum of the per-view losses over ALL ranks (scalar)
    num_local_views: int


class ViewShardedTrainerStep:
    """One data-parallel step over a batch of views.

    render(point_cloud, point_cloud_features, view) -> image  is the single-view operator call
    (default: none -- pass a closure around GaussianPointCloudRasterisation); loss_fn(image, view)
    -> scalar.  After `__call__`, `point_cloud.grad` / `point_cloud_features.grad` hold the gradient
    of the summed loss over every view of the batch, identical on every rank.
    """

    def __init__(self, render: Callable[[torch.Tensor, torch.Tensor, Any], torch.Tensor],
                 loss_fn: Callable[[torch.Tensor, Any], torch.Tensor],
                 group: Optional[dist.ProcessGroup] = None, sync_quaternions: bool = True,
                 rasteriser: Any = None, lanes: int = 4, balance: bool = False, exchange_chunks: int = 1,
                 peer_exchange: bool = True):
        """rasteriser: the GaussianPointCloudRasterisation module `render` calls, if the step should
        use its fused accumulation (`accumulate_gradients_into`): every view's backward then adds
        into one flat N x 59 buffer (no per-view zero fill, no AccumulateGrad pass) and the exchange
        is ONE all-reduce of that buffer.  Gradients that reach the two parameters from loss terms
        that do not pass through the rasteriser (the scale regulariser of LossFunction) arrive through
        autograd as usual and are added into the buffer before the exchange.  (Do not combine the fused
        path with LossFunction(fold_regulariser_gradient=True): its tensor hook rides on the
        rasteriser's gradient, which the fused path never hands to autograd.)
        lanes: with the fused accumulation on a CUDA device, this rank's views are spread over that many
        CUDA streams (view j on lane j mod lanes).  The views are independent (their gradients meet only
        in the RED adds into the flat buffer), the blend kernels are issue-bound and the key pipeline is
        latency / HBM bound, so views in flight fill each other's idle slots (+8 % with 2 lanes, +10 % with 4 at 1M points,
        1920x1088; tools/exp_two_streams.py).  The sum order of the gradient then varies from run to
        run (differences of 1e-7 relative).  A backward_valid_point_hook of the rasteriser runs on its
        view's lane, concurrently with other views' hooks: it must only ADD into shared state with
        atomics (AdaptiveControllerAccumulators.update does); a hook with plain read-modify-write on
        shared tensors needs lanes=1.
        balance: assign the views by longest-processing-time-first on the costs measured in the previous
        call (the rasteriser's `last_forward_info`; needs `rasteriser`) instead of round robin.
        exchange_chunks: a rasteriser WITHOUT a backward hook runs the step as a gradient batch
        (`begin_gradient_batch` / `end_gradient_batch`: the per-point chain rule of all the rank's views in
        one pass over the points, gs_point_bwd_batch).  With exchange_chunks > 1 and more than one rank
        that pass is cut into this many point ranges and the all-reduce of a range starts as soon as the
        range is written, so that the exchange overlaps the computation of the following ranges.  The
        default is ONE all-reduce after the pass: on 8 B200s the pass is 0.3 ms and every extra NCCL
        collective costs more than the part of it that it can hide behind.
        peer_exchange: on CUDA devices with an NVLink multicast mapping, gradient batches end in the fused
        pass + exchange kernel (PeerExchange / gs_point_bwd_batch_exchange) instead of the pass followed by
        NCCL's all_reduce; falls back to NCCL when the mapping cannot be set up, and for a step in which
        some rank has no view."""
        self.lanes = max(1, int(lanes))
        self._streams: List[Any] = []
        self.render = render
        self.loss_fn = loss_fn
        self.group = group
        self.sync_quaternions = sync_quaternions
        self.rasteriser = rasteriser
        self.balance = bool(balance) and rasteriser is not None
        self.exchange_chunks = max(1, int(exchange_chunks))
        self.peer_exchange = bool(peer_exchange)
        self._peer: Optional[PeerExchange] = None
        self._peer_failed = False
        self._drop_peer = False
        self._flat: Optional[torch.Tensor] = None
        self._costs: Optional[List[float]] = None   # per view, from the previous call

    def _render_view(self, point_cloud, point_cloud_features, views, v, costs):
        image = self.render(point_cloud, point_cloud_features, views[v])
        if costs is not None:
            info = self.rasteriser.last_forward_info
            costs[v] = view_cost(info.get("num_points_in_camera", 0), info.get("num_keys", 0))
        loss = self.loss_fn(image, views[v])
        loss.backward()
        return loss.detach()

    def _run_on_lanes(self, point_cloud, point_cloud_features, views, mine, costs) -> torch.Tensor:
        dev = point_cloud.device
        cur = torch.cuda.current_stream(dev)
        while len(self._streams) < self.lanes:
            self._streams.append(torch.cuda.Stream(dev))
        start = torch.cuda.Event()
        start.record(cur)
        totals = []
        for s in self._streams:
            s.wait_event(start)
            with torch.cuda.stream(s):
                totals.append(torch.zeros((), dtype=torch.float32, device=dev))
        for j, v in enumerate(mine):
            lane = j % self.lanes
            with torch.cuda.stream(self._streams[lane]):
                totals[lane] += self._render_view(point_cloud, point_cloud_features, views, v, costs)
        for s in self._streams:
            done = torch.cuda.Event()
            done.record(s)
            cur.wait_event(done)
        return torch.stack(totals).sum()

    def _world(self):
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(self.group), dist.get_world_size(self.group)
        return 0, 1

    def __call__(self, point_cloud: torch.Tensor, point_cloud_features: torch.Tensor,
                 views: Sequence[Any]) -> StepResult:
        rank, world = self._world()
        if world > 1 and self.sync_quaternions:
            with torch.no_grad():
                q = point_cloud_features.detach()[:, 0:4].contiguous()
                dist.broadcast(q, src=dist.get_global_rank(self.group, 0) if self.group else 0,
                               group=self.group)
                point_cloud_features.detach()[:, 0:4] = q
        point_cloud.grad = None
        point_cloud_features.grad = None
        if self.balance and self._costs is not None and len(self._costs) == len(views):
            mine = assign_views_lpt(self._costs, world)[0][rank]
        else:
            mine = views_of_rank(len(views), rank, world)
        costs = [0.0] * len(views) if self.balance else None
        total = torch.zeros((), dtype=torch.float32, device=point_cloud.device)
        fused = self.rasteriser is not None
        batched = False
        if fused:
            n = point_cloud.shape[0]
            if self._drop_peer:
                self._peer, self._peer_failed, self._flat, self._drop_peer = None, True, None, False
            if self._flat is None or self._flat.numel() != n * 59 or self._flat.device != point_cloud.device:
                self._peer = None
                if (self.peer_exchange and not self._peer_failed and world > 1 and point_cloud.is_cuda
                        and not getattr(self.rasteriser, "_has_hook", True)):
                    try:      # collective: every rank takes the same branch (same n, same devices)
                        self._peer = PeerExchange(n, point_cloud.device, self.group)
                    except Exception:   # noqa: BLE001 -- no multicast here: NCCL does the exchange
                        self._peer_failed = True
                self._flat = (self._peer.flat if self._peer is not None else
                              torch.empty(n * 59, dtype=torch.float32, device=point_cloud.device))
            # features first: their rows are added as 16-byte vectors and 224 n is a multiple of 16
            # for every n, 12 n is not
            gf, gx = self._flat[:56 * n].view(n, 56), self._flat[56 * n:].view(n, 3)
            batched = point_cloud.is_cuda and not getattr(self.rasteriser, "_has_hook", True)
            if batched:
                self.rasteriser.begin_gradient_batch(gx, gf)      # every row is stored: no zero fill
            else:
                self._flat.zero_()
                self.rasteriser.accumulate_gradients_into(gx, gf)
        use_lanes = fused and self.lanes > 1 and point_cloud.is_cuda and len(mine) > 1
        try:
            if use_lanes:
                total = total + self._run_on_lanes(point_cloud, point_cloud_features, views, mine, costs)
            else:
                for v in mine:     # autograd accumulates across this rank's views (not fused)
                    total = total + self._render_view(point_cloud, point_cloud_features, views, v, costs)
        except BaseException:
            if batched:
                self.rasteriser._batch = None        # leave batch mode, drop the parked views
            raise
        finally:
            if fused and not batched:
                self.rasteriser.accumulate_gradients_into(None, None)
        if costs is not None:
            shared = torch.tensor(costs, dtype=torch.float64, device=point_cloud.device)
            if world > 1:
                dist.all_reduce(shared, group=self.group)   # every view was rendered by exactly one rank
            self._costs = shared.tolist()
        if fused:
            # whatever reached the parameters through autograd (loss terms outside the rasteriser)
            extra_x, extra_f = point_cloud.grad, point_cloud_features.grad
            fused_exchange = (batched and self._peer is not None and extra_x is None and extra_f is None
                              and len(views) >= world)
            if fused_exchange:
                self.rasteriser.end_gradient_batch(exchange=self._peer)
            elif batched:
                if self._peer is not None and (extra_x is not None or extra_f is not None):
                    # this training loop sends gradients around the rasteriser every step (the same on every
                    # rank: they come from the loss configuration): the fused exchange will never apply, and
                    # NCCL is slow on the symmetric allocation -- from the next step on, a plain buffer
                    self._drop_peer = True
                pending = []

                def after_chunk(index, first, count):
                    rows = slice(first, first + count)
                    if extra_x is not None:
                        gx[rows].add_(extra_x[rows])
                    if extra_f is not None:
                        gf[rows].add_(extra_f[rows])
                    if world > 1 and count == n:
                        pending.append(dist.all_reduce(self._flat, group=self.group, async_op=True))
                    elif world > 1:   # asynchronous: the next range is computed while this one is exchanged
                        pending.append(dist.all_reduce(gf[rows], group=self.group, async_op=True))
                        pending.append(dist.all_reduce(gx[rows], group=self.group, async_op=True))

                try:
                    self.rasteriser.end_gradient_batch(self.exchange_chunks if world > 1 else 1, after_chunk)
                finally:
                    for work in pending:
                        work.wait()
            else:
                if extra_x is not None:
                    gx.add_(extra_x)
                if extra_f is not None:
                    gf.add_(extra_f)
                if world > 1:
                    dist.all_reduce(self._flat, group=self.group)
            if world > 1:
                dist.all_reduce(total, group=self.group)
            point_cloud.grad, point_cloud_features.grad = gx, gf
            return StepResult(loss=total, num_local_views=len(mine))
        # a rank with no view of this batch still takes part in the exchange with zeros
        if point_cloud.grad is None:
            point_cloud.grad = torch.zeros_like(point_cloud)
        if point_cloud_features.grad is None:
            point_cloud_features.grad = torch.zeros_like(point_cloud_features)
        if world > 1:
            dist.all_reduce(point_cloud.grad, group=self.group)
            dist.all_reduce(point_cloud_features.grad, group=self.group)
            dist.all_reduce(total, group=self.group)
        return StepResult(loss=total, num_local_views=len(mine))
