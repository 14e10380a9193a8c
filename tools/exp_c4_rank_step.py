#!/usr/bin/env python
"""What the ranks of the 8-GPU C4 run do between two exchanges, emulated on ONE GPU: every rank's share of
the 64 spread views (2M points, 976x544) run alone as an 8-view step on 8 lanes, for three assignments --
round robin, longest-processing-time-first on the cost model K + 3 M, and LPT on the MEASURED time of every
view alone -- beside the per-rank share of a 64-view step (what the views cost inside a long step, where
nothing ramps up or drains).  (Earlier sweeps with this tool, in profiles/r2_exp_c4_rank_step.json: lanes
1 / 2 / 4 / 8 and the order of a rank's views; profiles/r2_exp_c4_assignments_one_gpu.json: this output.)

    python tools/exp_c4_rank_step.py [--steps 6]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--views", type=int, default=8)
    args = ap.parse_args()
    import torch

    import bench
    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    from taichi_3d_gaussian_splatting_b200.distributed import assign_views_lpt, view_cost

    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    q_np, t_np = bench.make_views_spread(bench.C4_VIEWS)
    q_dev, t_dev = torch.from_numpy(q_np).to(dev), torch.from_numpy(t_np).to(dev)
    out = {}

    def make_step(rig, mine, stats=None):
        def per_view(j, stream):
            v = mine[j]
            loss = rig.render_loss(q_dev[v:v + 1], t_dev[v:v + 1], rig.wt_dev)
            if stats is not None:
                info = rig.ras.last_forward_info
                stats[v] = view_cost(info["num_points_in_camera"], info["num_keys"])
            return loss

        def step():
            rig.begin_step()
            rig.run_views(len(mine), per_view)
            rig.finish_step(None, 1)
        return step

    def timed(fn, steps):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    rig = bench.Rig(torch, impl, dev, "c4_2m_976x544", 4)
    costs = {}
    make_step(rig, list(range(bench.C4_VIEWS)), costs)()
    torch.cuda.synchronize()
    cost_list = [costs[v] for v in range(bench.C4_VIEWS)]
    mine = assign_views_lpt(cost_list, bench.C4_VIEWS // args.views)[0][0]   # rank 0's share at that world size
    ms_all = timed(make_step(rig, list(range(bench.C4_VIEWS))), 3)
    share = sum(cost_list[v] for v in mine) / sum(cost_list)
    out["ms_64_views_4_lanes"] = ms_all
    out["rank_views"] = mine
    out["no_bubble_bound_ms"] = ms_all * share
    del rig
    # per-view time, alone on one stream (CUDA events around forward + backward of the view)
    rig = bench.Rig(torch, impl, dev, "c4_2m_976x544", 1)
    view_ms = []
    for v in range(bench.C4_VIEWS):
        ts = []
        for rep in range(3):
            rig.begin_step()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rig.render_loss(q_dev[v:v + 1], t_dev[v:v + 1], rig.wt_dev)
            e1.record()
            rig.finish_step(None, 1)
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        view_ms.append(min(ts[1:]))
    del rig
    world = bench.C4_VIEWS // args.views
    assignments = {
        "round_robin": [list(range(r, bench.C4_VIEWS, world)) for r in range(world)],
        "lpt_on_K+3M": assign_views_lpt(cost_list, world)[0],
        "lpt_on_measured_ms": assign_views_lpt(view_ms, world)[0],
    }
    rig = bench.Rig(torch, impl, dev, "c4_2m_976x544", 8)
    corr = float(np.corrcoef(np.array(cost_list), np.array(view_ms))[0, 1])
    out["view_ms_min_max"] = [min(view_ms), max(view_ms)]
    out["corr_model_vs_measured"] = corr
    for name, shares in assignments.items():
        times = [timed(make_step(rig, share), 4) for share in shares]
        out[f"shares {name}"] = {"max_ms": max(times), "mean_ms": sum(times) / len(times),
                                 "per_rank_ms": [round(x, 2) for x in times]}
    del rig
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
