"""World-size-2 gloo test (CPU) of the view-sharded step: the N>1 host logic of SURVEY.md section 8(e).
The single-view operator is replaced by a small differentiable torch stand-in (the real one needs a
GPU); what is checked is the sharding, the gradient exchange and the replica consistency."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from taichi_3d_gaussian_splatting_b200.distributed import ViewShardedTrainerStep, views_of_rank


def test_views_of_rank_partition():
    for world in (1, 2, 3, 8):
        for nv in (0, 1, 5, 64):
            seen = sorted(v for r in range(world) for v in views_of_rank(nv, r, world))
            assert seen == list(range(nv))


def test_assign_views_lpt_partitions_and_balances():
    import numpy as np
    from taichi_3d_gaussian_splatting_b200.distributed import assign_views_lpt, view_cost
    rng = np.random.default_rng(0)
    costs = [view_cost(m, k) for m, k in zip(rng.uniform(6e5, 2e6, 64), rng.uniform(4e6, 1e7, 64))]
    for world in (1, 2, 4, 8):
        mine, loads = assign_views_lpt(costs, world)
        assert sorted(v for r in mine for v in r) == list(range(64))
        assert all(abs(loads[r] - sum(costs[v] for v in mine[r])) < 1e-6 * loads[r] for r in range(world))
        rr = [sum(costs[v] for v in views_of_rank(64, r, world)) for r in range(world)]
        assert max(loads) <= max(rr) + 1e-9
        assert max(loads) / (sum(loads) / world) < 1.02          # 64 views over <= 8 ranks: within 2 %
    assert assign_views_lpt(costs, 4) == assign_views_lpt(list(costs), 4)   # deterministic


def _render(xyz, feat, view):
    # differentiable, view dependent, and like the operator it normalises the q columns in place
    # for the points this view "sees"
    vis = ((torch.arange(xyz.shape[0]) + view) % 3) != 0
    with torch.no_grad():
        q = feat.detach()[vis, 0:4]
        feat.detach()[vis, 0:4] = q / q.norm(dim=1, keepdim=True)
    w = torch.sin(torch.arange(56, dtype=torch.float32) * (view + 1))
    return (xyz * (view + 1)).sum(1)[vis].sum() + (feat[vis] * w).sum() + (feat[vis, :4] ** 2).sum()


def _loss(image, view):
    return image * 0.5


def _worker(rank, world, port, n, views, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    xyz = torch.randn(n, 3, requires_grad=True)
    feat = torch.randn(n, 56, requires_grad=True)
    step = ViewShardedTrainerStep(_render, _loss)
    res = step(xyz, feat, views)
    out[rank] = (xyz.grad.clone(), feat.grad.clone(), res.loss.clone(), feat.detach().clone(),
                 res.num_local_views)
    res2 = step(xyz, feat, views)   # replicas must start step 2 from identical parameters
    out[rank + world] = (feat.detach().clone(), res2.loss.clone())
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.timeout(300)
def test_two_rank_step_matches_single_process():
    n, views, world = 200, list(range(5)), 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, _free_port(), n, views, out), nprocs=world, join=True)
        out = dict(out)
    torch.manual_seed(0)   # single-process reference: all views on one "rank"
    xyz = torch.randn(n, 3, requires_grad=True)
    feat = torch.randn(n, 56, requires_grad=True)
    res = ViewShardedTrainerStep(_render, _loss)(xyz, feat, views)
    gx0, gf0, l0, _, nv0 = out[0]
    gx1, gf1, l1, _, nv1 = out[1]
    assert nv0 == 3 and nv1 == 2
    assert torch.equal(gx0, gx1) and torch.equal(gf0, gf1) and torch.equal(l0, l1)
    assert torch.allclose(gx0, xyz.grad, rtol=1e-5, atol=1e-5)
    assert torch.allclose(gf0, feat.grad, rtol=1e-4, atol=1e-4)
    assert torch.allclose(l0, res.loss, rtol=1e-5)
    # replicas drifted in the q columns during step 1 (different visible sets) ...
    assert not torch.equal(out[0][3][:, :4], out[1][3][:, :4])
    # ... and step 2's losses agree exactly because the q columns are re-broadcast first
    assert torch.equal(out[2][1], out[3][1])


class _FusedStub:
    """Stand-in for the rasteriser's fused accumulation (`accumulate_gradients_into`): a view's
    backward adds its gradient into the caller's buffers and hands autograd nothing."""

    def __init__(self):
        self.buffers = None
        stub = self

        class _Fn(torch.autograd.Function):
            @staticmethod
            def forward(ctx, xyz, feat, view):
                ctx.view = view
                ctx.save_for_backward(xyz, feat)
                with torch.enable_grad():
                    pass
                return _render(xyz.detach(), feat.detach(), view)

            @staticmethod
            def backward(ctx, g):
                xyz, feat = ctx.saved_tensors
                with torch.enable_grad():
                    x = xyz.detach().requires_grad_(True)
                    f = feat.detach().clone().requires_grad_(True)
                    gx, gf = torch.autograd.grad(_render(x, f, ctx.view), (x, f), g)
                if stub.buffers is None:
                    return gx, gf, None
                stub.buffers[0].add_(gx)
                stub.buffers[1].add_(gf)
                return None, None, None

        self.fn = _Fn

    def accumulate_gradients_into(self, gx, gf):
        self.buffers = None if gx is None else (gx, gf)

    def render(self, xyz, feat, view):
        return self.fn.apply(xyz, feat, view)


def _fused_worker(rank, world, port, n, views, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    xyz = torch.randn(n, 3, requires_grad=True)
    feat = torch.randn(n, 56, requires_grad=True)
    stub = _FusedStub()
    step = ViewShardedTrainerStep(stub.render, _loss, rasteriser=stub)
    res = step(xyz, feat, views)
    assert stub.buffers is None  # accumulation mode is switched off again after the step
    out[rank] = (xyz.grad.clone(), feat.grad.clone(), res.loss.clone())
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_fused_accumulation_matches_plain_step():
    n, views, world = 150, list(range(6)), 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_fused_worker, args=(world, _free_port(), n, views, out), nprocs=world, join=True)
        out = dict(out)
    torch.manual_seed(0)
    xyz = torch.randn(n, 3, requires_grad=True)
    feat = torch.randn(n, 56, requires_grad=True)
    res = ViewShardedTrainerStep(_render, _loss)(xyz, feat, views)
    for r in range(world):
        gx, gf, loss = out[r]
        assert torch.allclose(gx, xyz.grad, rtol=1e-5, atol=1e-5)
        assert torch.allclose(gf, feat.grad, rtol=1e-4, atol=1e-4)
        assert torch.allclose(loss, res.loss, rtol=1e-5)
    assert torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][1], out[1][1])


def _stats_worker(rank, world, port, n, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from taichi_3d_gaussian_splatting_b200.GaussianPointAdaptiveController import AdaptiveControllerAccumulators
    acc = AdaptiveControllerAccumulators(n, "cpu")   # holding and exchanging needs no GPU; update() does
    g = torch.Generator().manual_seed(100 + rank)
    for name, dtype, _ in acc.FIELDS:
        t = getattr(acc, name)
        if dtype == torch.int32:
            t.copy_(torch.randint(0, 50, t.shape, generator=g, dtype=torch.int32))
        else:
            t.copy_(torch.rand(t.shape, generator=g))
    before = {name: getattr(acc, name).clone() for name, _, _ in acc.FIELDS}
    acc.all_reduce()
    out[rank] = (before, {name: getattr(acc, name).clone() for name, _, _ in acc.FIELDS})
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_densification_statistics_are_summed_over_ranks():
    """SURVEY.md section 8(f) row 2, multi-GPU half: each rank accumulates the statistics of its own
    views; all_reduce() gives every rank the totals."""
    n, world = 300, 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_stats_worker, args=(world, _free_port(), n, out), nprocs=world, join=True)
        out = dict(out)
    for name in out[0][0]:
        total = out[0][0][name] + out[1][0][name]
        assert torch.equal(out[0][1][name], out[1][1][name])
        if total.dtype == torch.int32:
            assert torch.equal(out[0][1][name], total)
        else:
            assert torch.allclose(out[0][1][name], total, rtol=1e-6)
