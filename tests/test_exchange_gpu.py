"""gs_point_bwd_batch_exchange (the per-point pass of a gradient batch fused with the sum over the ranks, in
the NVLink switch) against the pass followed by NCCL's all_reduce, on two GPUs of one box.  Skips where the
box has one GPU or symmetric memory has no multicast mapping (then the product falls back to NCCL too)."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, n, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    try:
        import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
        from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
        from taichi_3d_gaussian_splatting_b200.distributed import PeerExchange
        from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
        Ras = impl.GaussianPointCloudRasterisation
        w, h = 256, 192
        scene = make_scene(n, w, h, seed=31, scale_factor=0.02, num_objects=2)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        xyz = t(scene.point_cloud).requires_grad_(True)
        feat = t(scene.point_cloud_features).requires_grad_(True)
        cam = CameraInfo(t(scene.camera_intrinsics), h, w, 0)
        oid, inv = t(scene.point_object_id), t(scene.point_invalid_mask)   # one tensor each for every view
        q_pc, t_pc = t(scene.q_pointcloud_camera), t(scene.t_pointcloud_camera)
        g = t(np.random.default_rng(1).uniform(0, 1, (h, w, 3)).astype(np.float32))
        ras = Ras(Ras.GaussianPointCloudRasterisationConfig())
        shifts = [(0.0, 0.0, 0.0), (0.05, 0.0, 0.1), (-0.04, 0.03, 0.0), (0.0, -0.05, 0.2)]
        mine = shifts[rank::world]

        def views():
            for s in mine:
                image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(
                    point_cloud=xyz, point_cloud_features=feat, point_object_id=oid,
                    point_invalid_mask=inv, camera_info=cam, q_pointcloud_camera=q_pc,
                    t_pointcloud_camera=t_pc + torch.tensor([list(s)], device=dev),
                    color_max_sh_band=3))
                (image * g).sum().backward()

        try:
            peer = PeerExchange(n, dev)
        except RuntimeError as e:
            out[rank] = ("skip", str(e))
            return
        flat = torch.full((59 * n,), float("nan"), device=dev)
        gf, gx = flat[:56 * n].view(n, 56), flat[56 * n:].view(n, 3)
        views()   # the first forward normalises the quaternions in place: do it before either measurement
        xyz.grad = feat.grad = None
        worst = 0.0
        for step in range(3):   # several epochs on the same flags
            ras.begin_gradient_batch(gx, gf)
            views()
            ras.end_gradient_batch()
            dist.all_reduce(flat)
            peer.flat.fill_(float("nan"))
            ras.begin_gradient_batch(peer.grad_xyz, peer.grad_features)
            views()
            ras.end_gradient_batch(exchange=peer)
            torch.cuda.synchronize()
            assert peer.error() == 0
            assert bool(torch.isfinite(peer.flat).all())
            rel = float((peer.flat.double() - flat.double()).norm() / flat.double().norm())
            worst = max(worst, rel)
        out[rank] = ("ok", worst, peer.chunks)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [5000, 40_013])
def test_fused_pass_and_exchange_equals_pass_then_all_reduce(n):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs on one box")
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(2, port, n, out), nprocs=2, join=True)
    assert set(out.keys()) == {0, 1}
    if out[0][0] == "skip":
        pytest.skip(f"no symmetric-memory multicast here: {out[0][1]}")
    for rank in (0, 1):
        status, worst, chunks = out[rank]
        # the two paths differ by the order of fp32 additions only (atomics inside a view, ranks in the switch)
        assert status == "ok" and worst <= 1e-5, (rank, worst)
        assert 1 <= chunks <= 4


def _step_worker(rank, world, port, out):
    """ViewShardedTrainerStep on two ranks: the fused pass + exchange against NCCL's all_reduce after the pass
    (peer_exchange=False), LPT assignment on, six views."""
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    try:
        import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
        from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
        from taichi_3d_gaussian_splatting_b200.distributed import ViewShardedTrainerStep
        from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
        Ras = impl.GaussianPointCloudRasterisation
        n, w, h = 20_000, 256, 160
        scene = make_scene(n, w, h, seed=23, scale_factor=0.02)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        cam = CameraInfo(t(scene.camera_intrinsics), h, w, 0)
        oid, inv = t(scene.point_object_id), t(scene.point_invalid_mask)
        q_pc, t_pc = t(scene.q_pointcloud_camera), t(scene.t_pointcloud_camera)
        g = t(np.random.default_rng(4).uniform(0, 1, (h, w, 3)).astype(np.float32))
        views = [(0.01 * i, -0.005 * i) for i in range(6)]
        results = {}
        for peer_exchange in (False, True):
            ras = Ras(Ras.GaussianPointCloudRasterisationConfig())
            xyz = t(scene.point_cloud).requires_grad_(True)
            feat = t(scene.point_cloud_features).requires_grad_(True)

            def render(p, f, view, ras=ras):
                image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(
                    point_cloud=p, point_cloud_features=f, point_object_id=oid, point_invalid_mask=inv,
                    camera_info=cam, q_pointcloud_camera=q_pc,
                    t_pointcloud_camera=t_pc + torch.tensor([[view[0], view[1], 0.0]], device=dev),
                    color_max_sh_band=3))
                return image

            step = ViewShardedTrainerStep(render, lambda image, view: (image * g).sum(), rasteriser=ras,
                                          lanes=2, balance=True, peer_exchange=peer_exchange)
            for _ in range(3):   # the second call uses the measured costs (LPT), the third a later epoch
                res = step(xyz, feat, views)
            torch.cuda.synchronize()
            results[peer_exchange] = (xyz.grad.clone(), feat.grad.clone(), float(res.loss),
                                      step._peer is not None)
        gx0, gf0, loss0, used0 = results[False]
        gx1, gf1, loss1, used1 = results[True]
        assert not used0
        rel_f = float((gf1.double() - gf0.double()).norm() / gf0.double().norm())
        rel_x = float((gx1.double() - gx0.double()).norm() / gx0.double().norm())
        out[rank] = (used1, rel_f, rel_x, abs(loss1 - loss0) / abs(loss0))
    finally:
        dist.destroy_process_group()


def test_view_sharded_step_with_the_fused_exchange_equals_the_nccl_step():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs on one box")
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = mp.Manager().dict()
    mp.spawn(_step_worker, args=(2, port, out), nprocs=2, join=True)
    assert set(out.keys()) == {0, 1}
    if not out[0][0]:
        pytest.skip("no symmetric-memory multicast here: the step kept NCCL")
    for rank in (0, 1):
        used, rel_f, rel_x, rel_loss = out[rank]
        assert used and rel_f <= 1e-5 and rel_x <= 1e-5 and rel_loss <= 1e-5, out[rank]
