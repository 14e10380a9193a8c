"""Stand-in for compute-sanitizer memcheck / initcheck (the tool is closed on the GPU pool, see
profiles/r2_sanitizer_unavailable.txt): every buffer the operator allocates sits between two canary
regions and starts out filled with 0xFF bytes (NaN as float, -1 as int).

  * canaries intact after forward + backward  => no out-of-bounds WRITE past either end of any buffer;
  * outputs identical to a run on ordinary buffers => no READ of a byte the path did not write itself
    (such a read would surface as NaN in a result or as a wild index).

Covers the three key pipelines, both launch shapes of the backward blend, the hook path, an empty
frustum and a point count that is not a multiple of the kernels' block sizes.
"""
import os

import numpy as np
import pytest
import torch

import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
from taichi_3d_gaussian_splatting_b200.synthetic import make_scene

import util

pytestmark = pytest.mark.gpu

GUARD = 4096
POISON = 0xFF
CANARY = 0x5C


class GuardedTorch:
    """Stands in for the `torch` name inside the operator module: empty / zeros (and their _like forms)
    come out of canary-fenced, poisoned storage; everything else is torch's."""

    def __init__(self):
        self.blocks = []

    def __getattr__(self, name):
        return getattr(torch, name)

    def _carve(self, shape, dtype, device, zero):
        if isinstance(shape, int):
            shape = (shape,)
        shape = tuple(int(s) for s in shape)
        dtype = dtype or torch.float32
        nbytes = int(np.prod(shape, dtype=np.int64)) * torch.empty((), dtype=dtype).element_size()
        if torch.device(device).type != "cuda":
            return torch.zeros(shape, dtype=dtype, device=device) if zero else torch.empty(shape, dtype=dtype, device=device)
        padded = (nbytes + 255) // 256 * 256
        raw = torch.full((2 * GUARD + padded,), CANARY, dtype=torch.uint8, device=device)
        body = raw[GUARD:GUARD + nbytes]
        body.fill_(0 if zero else POISON)
        self.blocks.append((raw, nbytes))
        return body.view(dtype).view(shape)

    def empty(self, *shape, dtype=None, device=None, **kw):
        shape = shape[0] if len(shape) == 1 and not isinstance(shape[0], int) else shape
        return self._carve(shape, dtype, device or "cpu", zero=False)

    def zeros(self, *shape, dtype=None, device=None, **kw):
        shape = shape[0] if len(shape) == 1 and not isinstance(shape[0], int) else shape
        return self._carve(shape, dtype, device or "cpu", zero=True)

    def empty_like(self, t, **kw):
        return self._carve(tuple(t.shape), t.dtype, t.device, zero=False)

    def zeros_like(self, t, memory_format=None, **kw):
        return self._carve(tuple(t.shape), t.dtype, t.device, zero=True)

    def check(self):
        bad = []
        for raw, nbytes in self.blocks:
            head, tail = raw[:GUARD], raw[GUARD + nbytes:]
            if not bool((head == CANARY).all()) or not bool((tail == CANARY).all()):
                bad.append((nbytes, int((head != CANARY).sum()), int((tail != CANARY).sum())))
        return bad


def _run(scene, pipeline, persistent, hook, guarded, monkeypatch, near=0.8):
    monkeypatch.setenv("GS_BLEND_PERSISTENT", persistent)
    monkeypatch.setattr(impl, "KEY_PIPELINE", pipeline)
    guard = GuardedTorch()
    if guarded:
        monkeypatch.setattr(impl, "torch", guard)
    try:
        t = util.scene_tensors(scene, requires_grad=True)
        seen = {}
        ras = impl.GaussianPointCloudRasterisation(
            impl.GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig(near_plane=near),
            backward_valid_point_hook=(lambda h: seen.update(h=h)) if hook else None)
        cam = CameraInfo(t["camera_intrinsics"], scene.height, scene.width, 0)
        image, depth, count = ras(impl.GaussianPointCloudRasterisation.GaussianPointCloudRasterisationInput(
            point_cloud=t["point_cloud"], point_cloud_features=t["point_cloud_features"],
            point_object_id=t["point_object_id"], point_invalid_mask=t["point_invalid_mask"], camera_info=cam,
            q_pointcloud_camera=t["q_pointcloud_camera"], t_pointcloud_camera=t["t_pointcloud_camera"],
            color_max_sh_band=3))
        g = torch.from_numpy(np.random.default_rng(5).uniform(0, 1, (scene.height, scene.width, 3))
                             .astype(np.float32)).cuda()
        (image * g).sum().backward()
        torch.cuda.synchronize()
        out = dict(image=image.detach().cpu().numpy(), depth=depth.detach().cpu().numpy(),
                   count=count.cpu().numpy(), gx=t["point_cloud"].grad.cpu().numpy(),
                   gf=t["point_cloud_features"].grad.cpu().numpy(), info=dict(ras.last_forward_info))
        if hook and "h" in seen:
            h = seen["h"]
            out.update(aff=h.num_affected_pixels.cpu().numpy(),
                       on_image=h.magnitude_grad_viewspace_on_image.cpu().numpy(),
                       uv=h.point_uv_in_camera.cpu().numpy(), tiles=h.num_overlap_tiles.cpu().numpy())
        return out, guard
    finally:
        if guarded:
            monkeypatch.setattr(impl, "torch", torch)


SCENES = {
    "odd_7013pts_two_objects": dict(num_points=7013, width=128, height=96, seed=3, scale_factor=0.03, num_objects=2),
    "deep_lists": dict(num_points=5000, width=32, height=32, seed=4, scale_factor=0.15),
}


@pytest.mark.parametrize("scene_name", list(SCENES))
@pytest.mark.parametrize("pipeline", ["binned", "depth_first", "literal"])
@pytest.mark.parametrize("persistent", ["1", "0"])
@pytest.mark.parametrize("hook", [False, True])
def test_canaries_intact_and_nothing_uninitialised_is_read(scene_name, pipeline, persistent, hook, monkeypatch):
    scene = make_scene(**SCENES[scene_name])
    plain, _ = _run(scene, pipeline, persistent, hook, False, monkeypatch)
    fenced, guard = _run(scene, pipeline, persistent, hook, True, monkeypatch)
    assert len(guard.blocks) >= 12            # the operator's allocations did go through the fence
    assert guard.check() == []
    assert plain["info"] == fenced["info"] and plain["info"]["key_pipeline"] == pipeline
    for key in ("image", "depth", "count"):   # the forward has no atomics: bit-identical
        assert np.array_equal(plain[key], fenced[key]), key
    for key in ("gx", "gf"):
        assert np.isfinite(fenced[key]).all(), key
        l2, mx = util.grad_error(plain[key], fenced[key])
        assert l2 <= 1e-5 and mx <= 1e-5, (key, l2, mx)
    if hook:
        for key in ("aff", "uv", "tiles"):
            assert np.array_equal(plain[key], fenced[key]), key
        assert np.isfinite(fenced["on_image"]).all()


def test_fenced_buffers_with_empty_frustum_and_negative_depth_keys(monkeypatch):
    scene = make_scene(num_points=3000, width=64, height=64, seed=6, scale_factor=0.05)
    plain, _ = _run(scene, "binned", "1", True, False, monkeypatch, near=-1.0)     # literal 64-bit sort
    fenced, guard = _run(scene, "binned", "1", True, True, monkeypatch, near=-1.0)
    assert guard.check() == [] and plain["info"]["key_pipeline"] == "literal"
    assert np.array_equal(plain["image"], fenced["image"]) and np.array_equal(plain["aff"], fenced["aff"])
    far = make_scene(num_points=2000, width=64, height=64, seed=7, scale_factor=0.05)
    far.point_cloud[:, 2] = -5.0                                                    # behind the camera
    fenced, guard = _run(far, "binned", "1", True, True, monkeypatch)
    assert guard.check() == [] and fenced["info"]["num_points_in_camera"] == 0
    assert not fenced["image"].any() and not fenced["gf"].any()
