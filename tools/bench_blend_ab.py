#!/usr/bin/env python
"""A/B of the backward blend's launch shape at a bench workload: persistent (warps pull (tile, part)
units from a counter; default) against one CTA per tile (GS_BLEND_PERSISTENT=0).  CUDA-event times of
gs_blend_fwd / gs_blend_bwd alone on one forward state (the forward has one launch shape: a persistent
variant measured 0.605 ms against 0.572 ms and was dropped).

    python tools/bench_blend_ab.py [--workload c2_1m_1920x1088] [--iters 30]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import util  # noqa: E402
from taichi_3d_gaussian_splatting_b200 import _capi  # noqa: E402
from taichi_3d_gaussian_splatting_b200.synthetic import make_scene  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2_1m_1920x1088", choices=list(bench.WORKLOADS))
    ap.add_argument("--iters", type=int, default=30)
    args = ap.parse_args()
    n, w, h, sf, seed = bench.WORKLOADS[args.workload]
    scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    t = util.scene_tensors(scene)
    cfg = impl.GaussianPointCloudRasterisation.GaussianPointCloudRasterisationConfig()
    cam = CameraInfo(camera_intrinsics=t["camera_intrinsics"], camera_height=h, camera_width=w, camera_id=0)
    st = impl._forward_impl(cfg, t["point_cloud"], t["point_cloud_features"], t["point_invalid_mask"],
                            t["point_object_id"], t["q_pointcloud_camera"], t["t_pointcloud_camera"], cam,
                            keep_cell_masks=True)
    dev = torch.device("cuda")
    g = torch.from_numpy(np.random.default_rng(3).uniform(0, 1, (h, w, 3)).astype(np.float32)).to(dev)
    grads = torch.zeros(st.m, 12, dtype=torch.float32, device=dev)
    on_image = torch.zeros(h, w, 2, dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    lib = _capi.lib
    scratch = torch.empty(int(lib.gs_blend_bwd_scratch_bytes(h, w)), dtype=torch.uint8, device=dev)

    def fwd():
        _capi.check(lib.gs_blend_fwd(h, w, st.tile_start.data_ptr(), st.tile_end.data_ptr(), st.payload.data_ptr(),
                                     st.buffers.ptr("records"), st.image.data_ptr(), st.depth.data_ptr(),
                                     st.acc_alpha.data_ptr(), st.last.data_ptr(), st.count.data_ptr(), 0,
                                     st.cell_masks.data_ptr(), st.k, st.quarter_flags.data_ptr(), stream),
                    "gs_blend_fwd")

    def bwd(magnitude):
        _capi.check(lib.gs_blend_bwd(h, w, st.tile_start.data_ptr(), st.tile_end.data_ptr(), st.payload.data_ptr(),
                                     st.buffers.ptr("records"), g.data_ptr(), st.acc_alpha.data_ptr(),
                                     st.last.data_ptr(), grads.data_ptr(),
                                     on_image.data_ptr() if magnitude else None, st.cell_masks.data_ptr(), st.k,
                                     st.quarter_flags.data_ptr(), scratch.data_ptr(), stream), "gs_blend_bwd")

    def time_it(fn):
        ts = []
        for it in range(args.iters + 5):
            grads.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            if it >= 5:
                ts.append(e0.elapsed_time(e1))
        return round(float(np.median(ts)), 4)

    out = {"workload": args.workload, "m": st.m, "k": st.k}
    bwd(False)
    torch.cuda.synchronize()
    fix = scratch.view(torch.int32)
    entries = fix[1:1 + int(fix[0])]
    out["pixels_flagged_ambiguous"] = int((entries >= 0).sum())
    out["pixels_flagged_whole_quarter"] = int((entries < 0).sum())
    os.environ["GS_BWD_FIXUP"] = "0"
    out["bwd_without_literal_pass_ms"] = time_it(lambda: bwd(False))
    os.environ.pop("GS_BWD_FIXUP")
    for mode in ("1", "0"):
        os.environ["GS_BLEND_PERSISTENT"] = mode
        tag = "persistent" if mode == "1" else "cta_per_tile"
        out[f"fwd_{tag}_ms"] = time_it(fwd)
        out[f"bwd_{tag}_ms"] = time_it(lambda: bwd(False))
        out[f"bwd_magnitude_{tag}_ms"] = time_it(lambda: bwd(True))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
