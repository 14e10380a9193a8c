"""GPU parity of gs_adam_step / FusedAdam and gs_controller_accumulate / the accumulator class
with oracle/train_oracle.py.

Tolerances: Adam parameters after 6 steps within 2e-6 relative of the fp64 oracle (fp32 kernel);
integer statistics exact; fp32 statistics within 1e-6 relative.
Reference: GaussianPointTrainer.py:131-137,183-184; GaussianPointAdaptiveController.py:113-146.
"""
import numpy as np
import pytest
import torch

from oracle import train_oracle as TO

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(1000, 3), (1001, 3), (513, 56), (7,), (4096,)])
def test_fused_adam_matches_oracle(shape):
    from taichi_3d_gaussian_splatting_b200.optim import FusedAdam
    rng = np.random.default_rng(sum(shape))
    p0 = rng.normal(0, 1, shape).astype(np.float32)
    dev = torch.device("cuda")
    p = torch.nn.Parameter(torch.from_numpy(p0).to(dev))
    opt = FusedAdam([p], lr=3e-3, betas=(0.9, 0.999), eps=1e-8)
    sched = torch.optim.lr_scheduler.ExponentialLR(optimizer=opt, gamma=0.9)
    po, m, v = p0.astype(np.float64), np.zeros(shape), np.zeros(shape)
    lr = 3e-3
    for step in range(1, 7):
        g = rng.normal(0, 1, shape).astype(np.float32) * (0.0 if step == 3 else 1.0)
        p.grad = torch.from_numpy(g).to(dev)
        opt.step()
        po, m, v = TO.adam_step(po, g, m, v, lr, 0.9, 0.999, 1e-8, step)
        sched.step()
        lr *= 0.9
        assert np.abs(p.detach().cpu().numpy() - po).max() <= 2e-6 * max(1.0, np.abs(po).max())
    opt.zero_grad()
    assert p.grad is None


def test_adam_unaligned_slice_of_flat_buffer_and_mask():
    """Gradient read in place from a slice of a flat N x 59 buffer whose offset is not 16-byte
    aligned (N odd); optional row mask leaves masked rows untouched."""
    from taichi_3d_gaussian_splatting_b200 import _capi
    n = 1231
    rng = np.random.default_rng(5)
    dev = torch.device("cuda")
    flat = torch.from_numpy(rng.normal(0, 1, n * 59).astype(np.float32)).to(dev)
    gf = flat[3 * n:].view(n, 56)
    assert gf.data_ptr() % 16 != 0
    p0 = rng.normal(0, 1, (n, 56)).astype(np.float32)
    mask = (rng.uniform(size=n) < 0.3).astype(np.int8)
    st = torch.cuda.current_stream().cuda_stream
    for use_mask in (False, True):
        p = torch.from_numpy(p0).to(dev)
        m, v = torch.zeros_like(p), torch.zeros_like(p)
        t_mask = torch.from_numpy(mask).to(dev)
        _capi.check(_capi.lib.gs_adam_step(p.data_ptr(), gf.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(),
                                           1e-2, 0.9, 0.999, 1e-8, 1, t_mask.data_ptr() if use_mask else None,
                                           56 if use_mask else 0, st), "gs_adam_step")
        po, mo, vo = TO.adam_step(p0, gf.cpu().numpy(), np.zeros_like(p0), np.zeros_like(p0), 1e-2, 0.9, 0.999,
                                  1e-8, 1)
        got = p.cpu().numpy()
        keep = mask != 0 if use_mask else np.zeros(n, bool)
        assert np.array_equal(got[keep], p0[keep])
        assert np.abs(got[~keep] - po[~keep]).max() <= 2e-6 * np.abs(po).max()
        assert np.all(m.cpu().numpy()[keep] == 0)
    assert _capi.lib.gs_adam_step(p.data_ptr(), gf.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), 1e-2, 0.9,
                                  0.999, 1e-8, 0, None, 0, st) < 0          # step is 1-based


def test_controller_accumulators_match_oracle():
    from taichi_3d_gaussian_splatting_b200.GaussianPointAdaptiveController import AdaptiveControllerAccumulators
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation
    rng = np.random.default_rng(11)
    n, dev = 20000, torch.device("cuda")
    acc = AdaptiveControllerAccumulators(n, dev)
    want = TO.new_accumulators(n)
    Hook = GaussianPointCloudRasterisation.BackwardValidPointHookInput
    for it in range(4):
        m = int(rng.integers(1, n))
        ids = np.sort(rng.choice(n, m, replace=False)).astype(np.int32)
        px = rng.integers(0, 80, m).astype(np.int32)
        mag = rng.uniform(0, 1, m).astype(np.float32)
        mag[px == 0] = 0.0
        gx = rng.normal(0, 1e-3, (m, 3)).astype(np.float32)
        TO.controller_update(want, ids, px, mag, gx)
        t = lambda a: torch.from_numpy(a).to(dev)
        fields = {f: None for f in Hook.__dataclass_fields__}
        fields.update(point_id_in_camera_list=t(ids), num_affected_pixels=t(px), magnitude_grad_viewspace=t(mag),
                      grad_point_in_camera=t(gx))
        acc.update(Hook(**fields))
    assert acc.iteration_counter == 3
    for name, dtype, _ in TO.ACCUMULATORS:
        got = getattr(acc, name).cpu().numpy()
        if dtype == np.int32:
            assert np.array_equal(got, want[name]), name
        else:
            assert np.allclose(got, want[name], rtol=1e-6, atol=1e-9), name
    acc.reset()
    assert int(acc.accumulated_num_in_camera.sum()) == 0


def test_accumulators_from_a_real_backward_hook():
    """End to end: the operator's own hook payload drives the accumulators."""
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    from taichi_3d_gaussian_splatting_b200.GaussianPointAdaptiveController import AdaptiveControllerAccumulators
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation as R
    from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
    scene = make_scene(3000, 64, 48, seed=31, scale_factor=0.05)
    dev = torch.device("cuda")
    acc = AdaptiveControllerAccumulators(3000, dev)
    seen = {}

    def hook(h):
        seen["h"] = h
        acc.update(h)

    ras = R(R.GaussianPointCloudRasterisationConfig(), backward_valid_point_hook=hook)
    t = lambda a: torch.from_numpy(a).to(dev)
    xyz, feat = torch.nn.Parameter(t(scene.point_cloud)), torch.nn.Parameter(t(scene.point_cloud_features))
    image, _, _ = ras(R.GaussianPointCloudRasterisationInput(
        point_cloud=xyz, point_cloud_features=feat, point_object_id=t(scene.point_object_id),
        point_invalid_mask=t(scene.point_invalid_mask),
        camera_info=CameraInfo(t(scene.camera_intrinsics), scene.height, scene.width, 0),
        q_pointcloud_camera=t(scene.q_pointcloud_camera), t_pointcloud_camera=t(scene.t_pointcloud_camera),
        color_max_sh_band=3))
    image.sum().backward()
    h = seen["h"]
    want = TO.new_accumulators(3000)
    TO.controller_update(want, h.point_id_in_camera_list.cpu().numpy(), h.num_affected_pixels.cpu().numpy(),
                         h.magnitude_grad_viewspace.cpu().numpy(), h.grad_point_in_camera.cpu().numpy())
    assert int(acc.accumulated_num_in_camera.sum()) == h.point_id_in_camera_list.shape[0] > 0
    for name, dtype, _ in TO.ACCUMULATORS:
        got = getattr(acc, name).cpu().numpy()
        assert np.allclose(got, want[name], rtol=1e-6, atol=1e-12), name


def _torch_reference_loss(image, gt, feat, mask, lam=0.2, weight=2.0):
    """The reference's statements: GaussianPointTrainer.py:172-182, LossFunction.py:20-52, with
    pytorch_msssim.ssim written out (see tests/test_loss_oracle.py)."""
    import torch.nn.functional as F
    from oracle import loss_oracle
    win = torch.from_numpy(loss_oracle.gaussian_window()).to(image.device).repeat(3, 1, 1, 1)
    pred = torch.clamp(image, min=0, max=1).permute(2, 0, 1).unsqueeze(0)
    y = gt.unsqueeze(0)

    def gf(a):
        out = F.conv2d(a, weight=win.transpose(2, -1), stride=1, padding=0, groups=3)
        return F.conv2d(out, weight=win, stride=1, padding=0, groups=3)

    l1 = torch.abs(pred - y).mean()
    mu1, mu2 = gf(pred), gf(y)
    s1, s2, s12 = gf(pred * pred) - mu1 * mu1, gf(y * y) - mu2 * mu2, gf(pred * y) - mu1 * mu2
    ssim_map = ((2 * mu1 * mu2 + 1e-4) / (mu1 * mu1 + mu2 * mu2 + 1e-4)) * ((2 * s12 + 9e-4) / (s1 + s2 + 9e-4))
    reg = torch.norm(torch.exp(feat[mask == 0, 4:7]), dim=1).mean()
    return (1 - lam) * l1 + lam * (1 - ssim_map.mean()) + weight * reg


def test_training_iterations_track_the_reference_calls(tmp_path):
    """Whole iteration, both ways, from the same scene file: (a) this package's LossFunction +
    FusedAdam + accumulators, (b) the torch statements the reference trainer runs
    (GaussianPointTrainer.py:131-184).  After 8 iterations the parameters agree to 2e-4 of their
    range, the losses to 1e-5, and the loss has gone down."""
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    from taichi_3d_gaussian_splatting_b200.GaussianPointAdaptiveController import AdaptiveControllerAccumulators
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation as R
    from taichi_3d_gaussian_splatting_b200.GaussianPointCloudScene import GaussianPointCloudScene as Scene
    from taichi_3d_gaussian_splatting_b200.LossFunction import LossFunction
    from taichi_3d_gaussian_splatting_b200.optim import FusedAdam
    from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
    sc = make_scene(6000, 96, 64, seed=51, scale_factor=0.04, invalid_fraction=0.0)
    path = str(tmp_path / "scene.parquet")
    Scene(sc.point_cloud, Scene.PointCloudSceneConfig(), torch.from_numpy(sc.point_cloud_features)).to_parquet(path)
    dev = torch.device("cuda")
    t = lambda a: torch.from_numpy(a).to(dev)
    cam = CameraInfo(t(sc.camera_intrinsics), sc.height, sc.width, 0)
    q, tr = t(sc.q_pointcloud_camera), t(sc.t_pointcloud_camera)
    gt = t(np.random.default_rng(52).uniform(0.2, 0.8, (3, sc.height, sc.width)).astype(np.float32))

    def run(fused):
        scene = Scene.from_parquet(path).to_device(dev)
        xyz, feat, mask = scene.point_cloud, scene.point_cloud_features, scene.point_invalid_mask
        acc = AdaptiveControllerAccumulators(xyz.shape[0], dev)
        ras = R(R.GaussianPointCloudRasterisationConfig(), backward_valid_point_hook=acc.update)
        Opt = FusedAdam if fused else (lambda ps, lr: torch.optim.Adam(ps, lr=lr, betas=(0.9, 0.999)))
        o_feat, o_xyz = Opt([feat], lr=5e-3), Opt([xyz], lr=1e-4)
        sched = torch.optim.lr_scheduler.ExponentialLR(optimizer=o_xyz, gamma=0.95)
        lf = LossFunction(LossFunction.LossFunctionConfig())
        losses = []
        for it in range(8):
            o_feat.zero_grad()
            o_xyz.zero_grad()
            image, _, _ = ras(R.GaussianPointCloudRasterisationInput(
                point_cloud=xyz, point_cloud_features=feat, point_object_id=scene.point_object_id,
                point_invalid_mask=mask, camera_info=cam, q_pointcloud_camera=q, t_pointcloud_camera=tr,
                color_max_sh_band=3))
            if fused:
                loss, _, _ = lf.forward_rasterised(image, gt, mask, feat)
            else:
                loss = _torch_reference_loss(image, gt, feat, mask)
            loss.backward()
            o_feat.step()
            o_xyz.step()
            sched.step()
            losses.append(loss.item())
        return xyz.detach().cpu().numpy(), feat.detach().cpu().numpy(), losses, acc

    xa, fa, la, acc = run(True)
    xb, fb, lb, _ = run(False)
    assert la[-1] < la[0]
    assert np.allclose(la, lb, rtol=0, atol=1e-5), (la, lb)
    assert np.abs(xa - xb).max() <= 2e-4 * np.abs(xb).max()
    assert np.abs(fa - fb).max() <= 2e-4 * np.abs(fb).max()
    assert int(acc.accumulated_num_in_camera.max()) == 8 and acc.iteration_counter == 7
