"""View-sharded multi-GPU step (new functionality; the reference is single-GPU, README.md:325).

The path shards over camera views: every rank holds the whole point cloud, renders the views
`v = rank, rank + world, ...` of a batch forward+backward with the single-view operator, and the one
exchange step is the sum of the N x 3 + N x 56 gradient over ranks (`all_reduce`, NCCL over NVLink on
GPUs; gloo in the CPU tests).  One process per GPU, launched with torch.distributed.run.

The operator normalises the quaternion columns of the visible points in place (reference
R.py:264-266), and different ranks see different points, so the replicas would drift apart by
rounding in those four columns; `sync_quaternions` re-broadcasts them from rank 0 at the start of a
step (16 B per point, against 236 B per point for the gradient all-reduce).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable, List, Optional, Sequence

import torch
import torch.distributed as dist


def views_of_rank(num_views: int, rank: int, world: int) -> List[int]:
    """Indices of the views rank `rank` renders: v = rank (mod world)."""
    return list(range(rank, num_views, world))


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
                 rasteriser: Any = None, lanes: int = 4):
        """rasteriser: the GaussianPointCloudRasterisation module `render` calls, if the step should
        use its fused accumulation (`accumulate_gradients_into`): every view's backward then adds
        into one flat N x 59 buffer (no per-view zero fill, no AccumulateGrad pass) and the exchange
        is ONE all-reduce of that buffer.
        lanes: with the fused accumulation on a CUDA device, this rank's views are spread over that many
        CUDA streams (view j on lane j mod lanes).  The views are independent (their gradients meet only
        in the RED adds into the flat buffer), the blend kernels are issue-bound and the key pipeline is
        latency / HBM bound, so views in flight fill each other's idle slots (+8 % with 2 lanes, +10 % with 4 at 1M points,
        1920x1088; tools/exp_two_streams.py).  The sum order of the gradient then varies from run to
        run (differences of 1e-7 relative)."""
        self.lanes = max(1, int(lanes))
        self._streams: List[Any] = []
        self.render = render
        self.loss_fn = loss_fn
        self.group = group
        self.sync_quaternions = sync_quaternions
        self.rasteriser = rasteriser
        self._flat: Optional[torch.Tensor] = None

    def _run_on_lanes(self, point_cloud, point_cloud_features, views, mine) -> torch.Tensor:
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
                image = self.render(point_cloud, point_cloud_features, views[v])
                loss = self.loss_fn(image, views[v])
                loss.backward()
                totals[lane] += loss.detach()
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
        mine = views_of_rank(len(views), rank, world)
        total = torch.zeros((), dtype=torch.float32, device=point_cloud.device)
        fused = self.rasteriser is not None
        if fused:
            n = point_cloud.shape[0]
            if self._flat is None or self._flat.numel() != n * 59 or self._flat.device != point_cloud.device:
                self._flat = torch.empty(n * 59, dtype=torch.float32, device=point_cloud.device)
            self._flat.zero_()
            # features first: their rows are added as 16-byte vectors and 224 n is a multiple of 16
            # for every n, 12 n is not
            gf, gx = self._flat[:56 * n].view(n, 56), self._flat[56 * n:].view(n, 3)
            self.rasteriser.accumulate_gradients_into(gx, gf)
        use_lanes = fused and self.lanes > 1 and point_cloud.is_cuda and len(mine) > 1
        try:
            if use_lanes:
                total = total + self._run_on_lanes(point_cloud, point_cloud_features, views, mine)
            else:
                for v in mine:
                    image = self.render(point_cloud, point_cloud_features, views[v])
                    loss = self.loss_fn(image, views[v])
                    loss.backward()          # accumulates across this rank's views
                    total = total + loss.detach()
        finally:
            if fused:
                self.rasteriser.accumulate_gradients_into(None, None)
        if fused:
            if world > 1:
                dist.all_reduce(self._flat, group=self.group)
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
