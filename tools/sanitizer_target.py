#!/usr/bin/env python
"""What compute-sanitizer runs (tools/run_sanitizer.sh): every kernel of the library once or twice on
small inputs -- the three key pipelines, both launch shapes of the backward blend, the hook path, the
literal pass of the backward blend (irregular conics and threshold pairs), the batch extension on two
lanes, the loss front-end, the densification statistics and the fused Adam.

    PYTORCH_NO_CUDA_MEMORY_CACHING=1 compute-sanitizer --tool memcheck python tools/sanitizer_target.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl  # noqa: E402
from taichi_3d_gaussian_splatting_b200 import _capi  # noqa: E402
from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo  # noqa: E402
from taichi_3d_gaussian_splatting_b200.distributed import ViewShardedTrainerStep  # noqa: E402
from taichi_3d_gaussian_splatting_b200.GaussianPointAdaptiveController import \
    AdaptiveControllerAccumulators  # noqa: E402
from taichi_3d_gaussian_splatting_b200.LossFunction import LossFunction  # noqa: E402
from taichi_3d_gaussian_splatting_b200.optim import FusedAdam  # noqa: E402
from taichi_3d_gaussian_splatting_b200.synthetic import make_scene  # noqa: E402

Ras = impl.GaussianPointCloudRasterisation
dev = torch.device("cuda:0")


def tensors(scene):
    t = {k: torch.from_numpy(np.ascontiguousarray(getattr(scene, k))).to(dev)
         for k in ("point_cloud", "point_cloud_features", "point_invalid_mask", "point_object_id",
                   "camera_intrinsics", "q_pointcloud_camera", "t_pointcloud_camera")}
    t["point_cloud"].requires_grad_(True)
    t["point_cloud_features"].requires_grad_(True)
    return t


def step(scene, pipeline, persistent, hook, near=0.8):
    impl.KEY_PIPELINE = pipeline
    os.environ["GS_BLEND_PERSISTENT"] = persistent
    t = tensors(scene)
    seen = {}
    ras = Ras(Ras.GaussianPointCloudRasterisationConfig(near_plane=near),
              backward_valid_point_hook=(lambda h: seen.update(h=h)) if hook else None)
    cam = CameraInfo(t["camera_intrinsics"], scene.height, scene.width, 0)
    image, depth, count = ras(Ras.GaussianPointCloudRasterisationInput(
        point_cloud=t["point_cloud"], point_cloud_features=t["point_cloud_features"],
        point_object_id=t["point_object_id"], point_invalid_mask=t["point_invalid_mask"], camera_info=cam,
        q_pointcloud_camera=t["q_pointcloud_camera"], t_pointcloud_camera=t["t_pointcloud_camera"],
        color_max_sh_band=3))
    (image.sum() + 0 * depth.sum()).backward()
    torch.cuda.synchronize()
    return ras.last_forward_info["key_pipeline"], t, ras


def main():
    scene = make_scene(6000, 128, 96, seed=3, scale_factor=0.03, num_objects=2)
    ran = []
    for pipeline in ("binned", "depth_first", "literal"):
        for persistent in ("1", "0"):
            for hook in (False, True):
                ran.append(step(scene, pipeline, persistent, hook)[0])
    ran.append(step(scene, "binned", "1", True, near=-1.0)[0])     # negative depth keys: full 64-bit sort
    impl.KEY_PIPELINE = "binned"
    os.environ["GS_BLEND_PERSISTENT"] = "1"

    # the literal pass of the backward blend on hand-made records (irregular conics, threshold pairs)
    import test_kernels_gpu as T
    c = T._blend_case_by_hand()
    W, H, m = c["W"], c["H"], c["m"]
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    start, end, payload, records = tt(c["start"]), tt(c["end"]), tt(c["payload"]), tt(c["records"])
    k = payload.shape[0]
    image = torch.empty(H, W, 3, device=dev)
    depth, acc = torch.empty(H, W, device=dev), torch.empty(H, W, device=dev)
    last = torch.empty(H, W, dtype=torch.int32, device=dev)
    count = torch.empty(H, W, dtype=torch.int32, device=dev)
    masks = torch.empty(4 * k, dtype=torch.uint8, device=dev)
    flags = torch.empty(16, dtype=torch.uint8, device=dev)
    scratch = torch.empty(int(_capi.lib.gs_blend_bwd_scratch_bytes(H, W)), dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    _capi.check(_capi.lib.gs_blend_fwd(H, W, start.data_ptr(), end.data_ptr(), payload.data_ptr(),
                                       records.data_ptr(), image.data_ptr(), depth.data_ptr(), acc.data_ptr(),
                                       last.data_ptr(), count.data_ptr(), 0, masks.data_ptr(), k,
                                       flags.data_ptr(), st), "gs_blend_fwd")
    g = torch.rand(H, W, 3, device=dev)
    for magnitude in (True, False):
        grads = torch.zeros(m, 12, device=dev)
        on_image = torch.empty(H, W, 2, device=dev)
        _capi.check(_capi.lib.gs_blend_bwd(H, W, start.data_ptr(), end.data_ptr(), payload.data_ptr(),
                                           records.data_ptr(), g.data_ptr(), acc.data_ptr(), last.data_ptr(),
                                           grads.data_ptr(), on_image.data_ptr() if magnitude else None,
                                           masks.data_ptr(), k, flags.data_ptr(), scratch.data_ptr(), st),
                    "gs_blend_bwd")
    torch.cuda.synchronize()
    fixed = int(scratch.view(torch.int32)[0])

    # batch extension on two lanes with the densification statistics hook, then loss + Adam
    t = tensors(scene)
    acc_stats = AdaptiveControllerAccumulators(t["point_cloud"].shape[0], dev)
    ras = Ras(Ras.GaussianPointCloudRasterisationConfig(), backward_valid_point_hook=acc_stats.update)
    cam = CameraInfo(t["camera_intrinsics"], scene.height, scene.width, 0)
    gt = torch.rand(3, scene.height, scene.width, device=dev)
    loss_fn = LossFunction(LossFunction.LossFunctionConfig())

    def render(xyz, feat, view):
        return ras(Ras.GaussianPointCloudRasterisationInput(
            point_cloud=xyz, point_cloud_features=feat, point_object_id=t["point_object_id"],
            point_invalid_mask=t["point_invalid_mask"], camera_info=cam,
            q_pointcloud_camera=t["q_pointcloud_camera"],
            t_pointcloud_camera=t["t_pointcloud_camera"] + view, color_max_sh_band=3))[0]

    def loss(image, view):
        return loss_fn.forward_rasterised(image, gt, t["point_invalid_mask"], t["point_cloud_features"])[0]

    views = [torch.tensor([[0.01 * i, 0.0, 0.0]], device=dev) for i in range(4)]
    opt = FusedAdam([t["point_cloud"], t["point_cloud_features"]], lr=1e-3)
    sharded = ViewShardedTrainerStep(render, loss, rasteriser=ras, lanes=2)
    for _ in range(2):
        sharded(t["point_cloud"], t["point_cloud_features"], views)
        opt.step()
    torch.cuda.synchronize()
    print("sanitizer target ok:", sorted(set(ran)), "literal-pass pixels", fixed,
          "statistics max", int(acc_stats.accumulated_num_in_camera.max()))


if __name__ == "__main__":
    main()
