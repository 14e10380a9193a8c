#!/usr/bin/env python
"""Host time of one view forward+backward through the module (gradient batch, lanes as in the bench): the
same calls on a cloud so small that the GPU work is negligible, so the wall time per view is what the host
needs to enqueue a view (Python, allocator, launches, the {M, K} read-back round trip).

    python tools/exp_host_overhead.py [--views 64]      # lanes 1, 4 and 8 are all measured
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--views", type=int, default=64)
    args = ap.parse_args()
    import bench
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    dev = torch.device("cuda:0")
    out = {}
    for lanes in (1, 4, 8):
        bench.WORKLOADS["tiny"] = (2000, 64, 64, 0.02, 5)
        rig = bench.Rig(torch, impl, dev, "tiny", lanes)
        q_np, t_np = bench.make_views(8)
        q, t = torch.from_numpy(q_np).to(dev), torch.from_numpy(t_np).to(dev)

        def per_view(j, stream):
            v = j % 8
            return rig.render_loss(q[v:v + 1], t[v:v + 1], rig.wt_dev)

        def step():
            rig.begin_step()
            rig.run_views(args.views, per_view)
            rig.finish_step(None, 1)

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            step()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        out[f"lanes={lanes}"] = {"host_ms_per_view": (t1 - t0) / (5 * args.views) * 1e3,
                                 "drain_ms_after_enqueue": (t2 - t1) * 1e3}
        del rig
    print(json.dumps(out))


if __name__ == "__main__":
    main()
