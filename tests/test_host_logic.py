"""Host-side logic of the drop-in module that needs no GPU: API surface equal to the reference's
(R.py:775-826), sort-bit selection, buffer carving, and loud failure without CUDA tensors."""
import dataclasses

import pytest
import torch

from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import (
    GaussianPointCloudRasterisation, sort_key_bits)
import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
from taichi_3d_gaussian_splatting_b200.synthetic import make_scene

R = GaussianPointCloudRasterisation


def test_config_fields_and_constants_match_reference():
    cfg = R.GaussianPointCloudRasterisationConfig()
    fields = {f.name: getattr(cfg, f.name) for f in dataclasses.fields(cfg)}
    assert fields == dict(near_plane=0.8, far_plane=1000., depth_to_sort_key_scale=100.,
                          rgb_only=False)                       # R.py:778-781
    assert (cfg.grad_color_factor, cfg.grad_high_order_color_factor, cfg.grad_s_factor,
            cfg.grad_q_factor, cfg.grad_alpha_factor) == (5., 1., 0.5, 1., 20.)  # R.py:782-786
    assert (impl.TILE_WIDTH, impl.TILE_HEIGHT, impl.BOUNDARY_TILES) == (16, 16, 3)


def test_input_and_hook_records_match_reference():
    names = [f.name for f in dataclasses.fields(R.GaussianPointCloudRasterisationInput)]
    assert names == ["point_cloud", "point_cloud_features", "point_object_id", "point_invalid_mask",
                     "camera_info", "q_pointcloud_camera", "t_pointcloud_camera",
                     "color_max_sh_band"]                        # R.py:788-804
    assert R.GaussianPointCloudRasterisationInput.__dataclass_fields__["color_max_sh_band"].default == 2
    hook = [f.name for f in dataclasses.fields(R.BackwardValidPointHookInput)]
    assert hook == ["point_id_in_camera_list", "grad_point_in_camera",
                    "grad_pointfeatures_in_camera", "grad_viewspace", "magnitude_grad_viewspace",
                    "magnitude_grad_viewspace_on_image", "num_overlap_tiles",
                    "num_affected_pixels", "point_depth", "point_uv_in_camera"]  # R.py:806-817


def test_module_is_nn_module_with_function_closure():
    m = R(R.GaussianPointCloudRasterisationConfig())
    assert isinstance(m, torch.nn.Module)
    assert issubclass(m._module_function, torch.autograd.Function)
    m2 = R(R.GaussianPointCloudRasterisationConfig(near_plane=0.1))
    assert m._module_function is not m2._module_function  # one closure per instance, R.py:828


def test_sort_key_bits():
    assert sort_key_bits(0.8, 1000., 100., 8160) == (17, 13)
    assert sort_key_bits(0.4, 2000., 10., 2074) == (15, 12)
    assert sort_key_bits(-1.0, 1000., 100., 8160) == (32, 32)    # negative depths: full signed sort
    assert sort_key_bits(0.8, 1e9, 100., 8160) == (32, 32)
    assert sort_key_bits(0.0, 10., 100., 1) == (10, 1)


def test_point_buffers_hold_only_what_was_asked_for():
    core = impl._PointBuffers(1000, torch.device("cpu"))
    assert set(core.offsets) == set(impl._PointBuffers.CORE)
    assert core.ptr("conic") is None and core.ptr("tile_offsets") is None and core.ptr("records") is not None
    per_point = sum(cols * torch.empty((), dtype=dtype).element_size()
                    for name, dtype, cols in core.FIELDS if name in core.offsets)
    assert per_point == 76   # ids 4 + camera-space xyz 12 + record 48 + bbox 8 + depth key 4
    hook = impl._PointBuffers(1000, torch.device("cpu"), hook=True)
    assert set(hook.offsets) == set(impl._PointBuffers.CORE) | {"point_uv", "num_tiles"}
    literal = impl._PointBuffers(1000, torch.device("cpu"), offsets=True)
    assert set(literal.offsets) == set(impl._PointBuffers.CORE) | {"tile_offsets"}
    batch = impl._PointBuffers(1000, torch.device("cpu"), slots=True)
    assert set(batch.offsets) == set(impl._PointBuffers.CORE) | {"slot_of_point"}


def test_point_buffers_are_aligned_and_disjoint():
    buf = impl._PointBuffers(1000, torch.device("cpu"), debug=True, offsets=True, slots=True)
    spans = []
    for name, dtype, cols in buf.FIELDS:
        size = 1000 * cols * torch.empty((), dtype=dtype).element_size()
        assert buf.offsets[name] % 256 == 0
        spans.append((buf.offsets[name], buf.offsets[name] + size))
        v = buf.view(name, 10)
        assert v.shape[0] == 10 and v.dtype == dtype
    spans.sort()
    assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:]))
    assert spans[-1][1] <= buf.storage.numel()


def test_cpu_tensors_fail_loudly_not_silently():
    scene = make_scene(10, 64, 64, seed=1)
    m = R(R.GaussianPointCloudRasterisationConfig())
    cam = CameraInfo(torch.from_numpy(scene.camera_intrinsics), 64, 64, 0)
    inp = R.GaussianPointCloudRasterisationInput(
        point_cloud=torch.from_numpy(scene.point_cloud),
        point_cloud_features=torch.from_numpy(scene.point_cloud_features),
        point_object_id=torch.from_numpy(scene.point_object_id),
        point_invalid_mask=torch.from_numpy(scene.point_invalid_mask),
        camera_info=cam, q_pointcloud_camera=torch.from_numpy(scene.q_pointcloud_camera),
        t_pointcloud_camera=torch.from_numpy(scene.t_pointcloud_camera))
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(inp)
    cam_bad = CameraInfo(cam.camera_intrinsics, 60, 64, 0)
    inp.camera_info = cam_bad
    with pytest.raises(AssertionError):                          # R.py:1193-1194
        m(inp)


def test_product_package_never_imports_the_oracle():
    import os
    import re
    root = os.path.dirname(os.path.abspath(impl.__file__))
    for dirpath, _, files in os.walk(root):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle|gs_oracle\.(py|so)|libgs_oracle",
                                     text, flags=re.M), f


def test_gradient_batch_api_rejects_misuse_on_the_host():
    """begin_gradient_batch / end_gradient_batch / accumulate_gradients_into validate before anything
    is launched: CPU buffers, wrong shapes, a hook, a missing begin."""
    import pytest
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    Ras = impl.GaussianPointCloudRasterisation
    ras = Ras(Ras.GaussianPointCloudRasterisationConfig())
    gx, gf = torch.zeros(8, 3), torch.zeros(8, 56)
    with pytest.raises(ValueError):
        ras.accumulate_gradients_into(gx, gf)          # CPU tensors: there is no CPU path
    with pytest.raises(ValueError):
        ras.begin_gradient_batch(gx, gf)
    with pytest.raises(RuntimeError):
        ras.end_gradient_batch()                        # nothing was begun
    hooked = Ras(Ras.GaussianPointCloudRasterisationConfig(), backward_valid_point_hook=lambda h: None)
    with pytest.raises(ValueError, match="hook"):
        hooked.begin_gradient_batch(gx, gf)
    ras.accumulate_gradients_into(None, None)           # restoring the reference behaviour always works
    assert ras._accumulate_into is None and ras._batch is None


def test_views_and_lpt_assignment_cover_every_view_once():
    from taichi_3d_gaussian_splatting_b200.distributed import assign_views_lpt, view_cost, views_of_rank
    costs = [view_cost(m, k) for m, k in ((1.0e6, 9.0e6), (0.6e6, 4.4e6), (1.9e6, 9.9e6), (1.2e6, 7.0e6),
                                            (0.9e6, 5.0e6))]
    for world in (1, 2, 3, 5, 8):
        mine, loads = assign_views_lpt(costs, world)
        assert sorted(v for share in mine for v in share) == list(range(len(costs)))
        assert abs(sum(loads) - sum(costs)) < 1e-6 * sum(costs)
        if world <= len(costs):
            assert all(share for share in mine)          # >= world views: nobody is left without one
        assert sorted(v for r in range(world) for v in views_of_rank(len(costs), r, world)) == list(range(len(costs)))
