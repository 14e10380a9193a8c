#!/usr/bin/env python
"""Host-side cost of one fused training iteration at C2: enqueue time against total time, and a
cProfile of 20 iterations (tools/bench_train_step.py measures the device side)."""
import cProfile, pstats, sys, os, io, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np, torch
import bench
from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
from taichi_3d_gaussian_splatting_b200.GaussianPointAdaptiveController import AdaptiveControllerAccumulators
from taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation import GaussianPointCloudRasterisation as Ras
from taichi_3d_gaussian_splatting_b200.LossFunction import LossFunction
from taichi_3d_gaussian_splatting_b200.optim import FusedAdam
from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
n, w, h, sf, seed = bench.WORKLOADS["c2_1m_1920x1088"]
dev = torch.device("cuda:0")
scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
t = lambda a: torch.from_numpy(a).to(dev)
inv, oid = t(scene.point_invalid_mask), t(scene.point_object_id)
cam = CameraInfo(t(scene.camera_intrinsics), h, w, 0)
q, tr = t(scene.q_pointcloud_camera), t(scene.t_pointcloud_camera)
gt = torch.rand(3, h, w, device=dev)
xyz = torch.nn.Parameter(t(scene.point_cloud)); feat = torch.nn.Parameter(t(scene.point_cloud_features))
acc = AdaptiveControllerAccumulators(n, dev)
o1, o2 = FusedAdam([feat], lr=5e-3), FusedAdam([xyz], lr=1e-5)
lf = LossFunction(LossFunction.LossFunctionConfig(), fold_regulariser_gradient=True)
ras = Ras(Ras.GaussianPointCloudRasterisationConfig(), backward_valid_point_hook=acc.update)
def it():
    o1.zero_grad(); o2.zero_grad()
    image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(point_cloud=xyz, point_cloud_features=feat, point_object_id=oid, point_invalid_mask=inv, camera_info=cam, q_pointcloud_camera=q, t_pointcloud_camera=tr, color_max_sh_band=3))
    loss, _, _ = lf.forward_rasterised(image, gt, inv, feat)
    loss.backward(); o1.step(); o2.step()
for _ in range(5): it()
torch.cuda.synchronize()
# host enqueue time per iteration
ts=[]
for _ in range(20):
    torch.cuda.synchronize(); t0=time.perf_counter(); it(); t1=time.perf_counter(); torch.cuda.synchronize(); t2=time.perf_counter()
    ts.append((t1-t0, t2-t0))
print("host enqueue ms (median)", 1e3*sorted(x[0] for x in ts)[10], "total ms", 1e3*sorted(x[1] for x in ts)[10])
pr=cProfile.Profile(); pr.enable()
for _ in range(20): it()
torch.cuda.synchronize(); pr.disable()
s=io.StringIO(); pstats.Stats(pr,stream=s).sort_stats("tottime").print_stats(18); print(s.getvalue()[:3500])
