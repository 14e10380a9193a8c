#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 rasteriser (contract in the task statement, section 4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is one forward+backward of the operator over one camera view of the synthetic workload
BASELINE.json's metric is quoted on: configs[1] = 1M random Gaussians, SH degree 3, 1920x1088
(1080p rounded up to whole 16x16 tiles), generated as SURVEY.md section 8(d)
prescribes.  Views cycle through a small set of nearby camera poses.

  value   fwd+bwd iterations/s with every input already resident in HBM (CUDA events)
  e2e     the same through the public module with HOST inputs: per step the camera pose and the
          per-pixel loss weights come from pinned host memory (what the reference's trainer
          uploads per step, Tr.py:149-159: pose + ground-truth image; the Gaussians are
          nn.Parameters resident on the GPU, Tr.py:65-98) and the scalar loss is read back every
          step (asynchronous copy into pinned memory, consumed one step later, as a trainer logs it)
  N > 1   views are sharded over ranks against a replicated cloud; every step ends with an NCCL
          all-reduce of the N x 59 gradient; weak scaling
  --impl reference   the CPU oracle (the reference's Taichi kernels cannot run here: no Taichi)
                     timed on the host cores with OpenMP, same config and metric
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (points, width, height, scale_factor, seed)
    "c2_1m_1920x1088": (1_000_000, 1920, 1088, 0.004, 1002),
    "c3_430k_976x544": (430_000, 976, 544, 0.004, 1003),
    "small_100k_976x544": (100_000, 976, 544, 0.004, 1005),
    "c4_2m_976x544": (2_000_000, 976, 544, 0.004, 1004),
    "c5_5m_3840x2160": (5_000_000, 3840, 2160, 0.004, 1006),
}
METRIC = "fwd+bwd iters/s"  # BASELINE.json metric, first half; render Mpix/s rides along as render_mpix_per_s
NUM_VIEWS = 8
VIEWS_PER_RANK = 8   # views each rank renders per step (C4: 64 views over 8 GPUs); weak scaling
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.5, SURVEY.md section 8(d)
FLOP_PER_PAIR_FWD, FLOP_PER_PAIR_BWD = 28.0, 64.0     # SURVEY.md section 8(d) convention


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2_1m_1920x1088", choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--timeline", default=None,
                    help="also write the kernel timeline (name, stream, start, duration; CUPTI through "
                         "torch.profiler) of two untimed resident steps to this JSON file")
    ap.add_argument("--priority-sections", type=int, default=0,
                    help="with lanes > 1: the small kernels of a view (projection, sorts, binning, per-point "
                         "backward) run on a high-priority companion stream of the lane")
    ap.add_argument("--lanes", type=int, default=4,
                    help="CUDA streams the views of a step are spread over (view j on lane j mod L)")
    return ap.parse_args()


def make_views(num_views: int, seed: int = 5):
    """Nearby camera poses (T_pointcloud_camera as xyzw quaternion + translation)."""
    rng = np.random.default_rng(seed)
    q = np.zeros((num_views, 4), np.float32)
    q[:, 3] = 1.0
    q[1:, :3] = rng.normal(0, 0.01, (num_views - 1, 3))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    t = np.zeros((num_views, 3), np.float32)
    t[1:] = rng.normal(0, 0.05, (num_views - 1, 3))
    return q.astype(np.float32), t.astype(np.float32)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[1]))
                smax.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # median over the upper half of the samples = clocks under load (idle samples at the ends)
        busy = sorted(sm)[len(sm) // 2:]
        return {"sm_mhz": statistics.median(busy), "sm_max_mhz": max(smax), "reasons": sorted(reasons),
                "samples": len(sm)}


def stage_bytes(n, m, k, tiles, hw, info):
    """ALGORITHMIC bytes per call of each stage (stated in DESIGN.md)."""
    depth_bits, tile_bits = int(info["depth_bits"]), int(info["tile_bits"])
    if info.get("key_pipeline") == "binned":
        pm = (depth_bits + 7) // 8
        keys = {
            "sort_points": m * (4 + 16 * pm - 4),
            # order + bbox in, payload + ranges out (the 8-byte column entries in between are the
            # implementation's, not the algorithm's)
            "bin_tiles": 12 * m + 4 * k + 8 * tiles,
        }
    elif info.get("key_pipeline") == "depth_first":
        pm, pk = (depth_bits + 7) // 8, (tile_bits + 7) // 8
        keys = {
            # argsort of M 4-byte depth keys: histogram read + per pass key/value read and write
            "sort_points": m * (4 + 16 * pm - 4),
            "emit_keys": 12 * m + 8 * k,            # order + bbox in, tile id + payload out
            "sort_pairs": k * (16 * pk),             # per pass 8 B in, 8 B out (histogram: from the emission)
            "tile_ranges": 4 * k + 8 * tiles,
        }
    else:
        passes = (depth_bits + tile_bits + 7) // 8
        keys = {
            "emit_keys": 24 * m + 12 * k,
            "sort_pairs": k * (8 + 24 * passes),
            "tile_ranges": 8 * k + 8 * tiles,
        }
    return {
        "project_cull_sh": 17 * n + 296 * m,
        **keys,
        "blend_fwd": 48 * k + 28 * hw,
        "zero_fill": 4 * 59 * n + 48 * m,
        "blend_bwd": 44 * k + 28 * hw + 8 * hw + 44 * m,
        "point_bwd": 280 * m + 232 * m,
    }


def oracle_step(o, scene, q, t, wt, cfg=(0.8, 1000.0, 100.0), band=None):
    """One fwd+bwd of the CPU oracle on one view; returns (seconds, forward dict, backward dict)."""
    if band is not None:
        o.set_tile_row_band(*band)
    t0 = time.perf_counter()
    f = o.forward(scene.point_cloud, scene.point_cloud_features, scene.point_invalid_mask,
                  scene.point_object_id, scene.camera_intrinsics, q, t, scene.width, scene.height,
                  *cfg)
    b = o.backward(f, wt, 3)
    dt = time.perf_counter() - t0
    if band is not None:
        o.set_tile_row_band(-1, -1)
    return dt, f, b


def run_reference(args):
    """--impl reference: the CPU restatement of the reference path on the host cores.

    The reference's own kernels are Taichi; Taichi is not installable here, so kind = "port".
    Under torchrun only rank 0 works."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from oracle import gs_oracle
    from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
    gs_oracle.build()
    o = gs_oracle.get(False)
    n, w, h, sf, seed = WORKLOADS[args.workload]
    scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
    qs, ts = make_views(NUM_VIEWS)
    wt = np.random.default_rng(3).uniform(0, 1, (h, w, 3)).astype(np.float32)
    cores = o.num_threads()
    # probe one full step; if warmup+steps full steps would exceed ~4 minutes, time the blend loops
    # on a band of tile rows and extrapolate them to the frame (per-point stages always run in full)
    probe, f, b = oracle_step(o, scene, qs[0:1], ts[0:1], wt)
    total_steps = args.steps + args.warmup
    rows = h // 16
    band = None
    sample = f"full frame, every stage, {args.steps} timed steps"
    if probe * total_steps > 240.0:
        band_rows = max(1, min(rows, int(rows * 240.0 / (probe * total_steps))))
        lo = (rows - band_rows) // 2
        band = (lo, lo + band_rows)
        sample = (f"per-point stages in full + blend fwd/bwd on tile rows [{band[0]},{band[1]}) of "
                  f"{rows}; blend time scaled by {rows}/{band_rows}")
    times = []
    for i in range(total_steps):
        v = i % NUM_VIEWS
        if band is None:
            dt, _, _ = oracle_step(o, scene, qs[v:v + 1], ts[v:v + 1], wt)
        else:
            t_pts, _, _ = oracle_step(o, scene, qs[v:v + 1], ts[v:v + 1], wt, band=(0, 0))
            t_all, _, _ = oracle_step(o, scene, qs[v:v + 1], ts[v:v + 1], wt, band=band)
            dt = t_pts + (t_all - t_pts) * rows / (band[1] - band[0])
        if i >= args.warmup:
            times.append(dt)
    total = sum(times)
    value = len(times) / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "it/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": args.workload, "points": n, "width": w, "height": h, "sh_degree": 3,
                   "near": 0.8, "far": 1000.0, "depth_to_sort_key_scale": 100.0},
        "cpu_baseline": {"value": value, "unit": "it/s", "cores": cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "it/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "render_mpix_per_s": None,
        "note": "reference's Taichi kernels unavailable on this box (no taichi wheel); CPU oracle "
                "= C restatement of the same path, OpenMP over points/tiles",
    }
    _emit(line)


def run_b200(args):
    import torch
    import torch.distributed as dist

    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
    from taichi_3d_gaussian_splatting_b200.profiling import StageProfiler
    from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
    Ras = impl.GaussianPointCloudRasterisation

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n, w, h, sf, seed = WORKLOADS[args.workload]
    scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
    xyz = torch.nn.Parameter(torch.from_numpy(scene.point_cloud).to(dev))
    feat = torch.nn.Parameter(torch.from_numpy(scene.point_cloud_features).to(dev))
    inv = torch.from_numpy(scene.point_invalid_mask).to(dev)
    oid = torch.from_numpy(scene.point_object_id).to(dev)
    cam = CameraInfo(torch.from_numpy(scene.camera_intrinsics).to(dev), h, w, 0)
    qs_np, ts_np = make_views(NUM_VIEWS)
    wt_np = np.random.default_rng(3).uniform(0, 1, (h, w, 3)).astype(np.float32)
    # host side of the e2e leg: pinned buffers
    qs_host = torch.from_numpy(qs_np).pin_memory()
    ts_host = torch.from_numpy(ts_np).pin_memory()
    wt_host = torch.from_numpy(wt_np).pin_memory()
    # device-resident copies for the kernel-only leg
    qs_dev, ts_dev, wt_dev = qs_host.to(dev), ts_host.to(dev), wt_host.to(dev)
    ras = Ras(Ras.GaussianPointCloudRasterisationConfig())

    # A step = one batch of VIEWS_PER_RANK views on every rank (views sharded
    # over ranks against a replicated cloud): forward+backward of every view, the point gradients
    # summed into one N x 59 buffer by the backward itself, then ONE all-reduce of that buffer.
    flat = torch.zeros(n * 59, dtype=torch.float32, device=dev)
    gf, gx = flat[:56 * n].view(n, 56), flat[56 * n:].view(n, 3)   # 16-byte aligned feature rows for every n
    ras.accumulate_gradients_into(gx, gf)
    copy_stream = torch.cuda.Stream(dev)

    def view_of(step, j):  # global view index of this rank's j-th view of the step
        return ((step * world + rank) * VIEWS_PER_RANK + j) % NUM_VIEWS

    def render_loss(q, t, wt):
        image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(
            point_cloud=xyz, point_cloud_features=feat, point_object_id=oid,
            point_invalid_mask=inv, camera_info=cam, q_pointcloud_camera=q,
            t_pointcloud_camera=t, color_max_sh_band=3))
        loss = (image * wt).sum()
        loss.backward()
        return loss.detach()

    def finish_step():
        if world > 1:  # the one exchange step of the path: sum the N x 59 gradient over ranks
            dist.all_reduce(flat)
        xyz.grad, feat.grad = gx, gf

    lanes = max(1, min(args.lanes, VIEWS_PER_RANK))
    lane_streams = [torch.cuda.Stream(dev) for _ in range(lanes)] if lanes > 1 else []

    def run_views(per_view, use_lanes=True):
        """The step's views, view j on lane j mod L (the views are independent: their gradients are
        summed into `flat` by RED either way).  The hot kernels are issue-bound, the key pipeline is
        latency / HBM bound: two views in flight fill each other's idle slots."""
        cur = torch.cuda.current_stream(dev)
        if not (use_lanes and lane_streams):
            total = torch.zeros((), device=dev)
            for j in range(VIEWS_PER_RANK):
                total += per_view(j, cur)
            return total
        start = torch.cuda.Event()
        start.record(cur)
        totals = []
        for s in lane_streams:
            s.wait_event(start)
            with torch.cuda.stream(s):
                totals.append(torch.zeros((), device=dev))
        for j in range(VIEWS_PER_RANK):
            s = lane_streams[j % lanes]
            with torch.cuda.stream(s):
                totals[j % lanes] += per_view(j, s)
        for s in lane_streams:
            e = torch.cuda.Event()
            e.record(s)
            cur.wait_event(e)
        return torch.stack(totals).sum()

    def resident_view(i):
        def per_view(j, stream):
            v = view_of(i, j)
            return render_loss(qs_dev[v:v + 1], ts_dev[v:v + 1], wt_dev)
        return per_view

    def step_resident(i):
        flat.zero_()
        total = run_views(resident_view(i))
        finish_step()
        return total

    def step_resident_single_stream(i):  # the instrumented pass: per-stage CUDA events need one stream
        flat.zero_()
        total = run_views(resident_view(i), use_lanes=False)
        finish_step()
        return total

    num_slots = 2 * lanes                                      # e2e: staging buffers of the targets
    wt_stage = [torch.empty_like(wt_dev) for _ in range(num_slots)]
    uploads, finished, next_upload = {}, {}, [0]
    loss_host = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ready, loss_log = [None, None], []

    def upload(g):
        """pose + target of the g-th view this rank renders, from pinned host memory, on the copy
        stream, into staging slot g mod num_slots (free once view g - num_slots has been rendered)"""
        v = view_of(g // VIEWS_PER_RANK, g % VIEWS_PER_RANK)
        slot = g % num_slots
        with torch.cuda.stream(copy_stream):
            prev = finished.pop(g - num_slots, None)
            if prev is not None:
                copy_stream.wait_event(prev)
            q = qs_host[v:v + 1].to(dev, non_blocking=True)
            t = ts_host[v:v + 1].to(dev, non_blocking=True)
            wt_stage[slot].copy_(wt_host, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        uploads[g] = (q, t, wt_stage[slot], ev)

    def step_e2e(i):
        flat.zero_()
        if next_upload[0] < i * VIEWS_PER_RANK or not uploads:   # first call: nothing prefetched yet
            copy_stream.wait_stream(torch.cuda.current_stream(dev))
            uploads.clear()
            finished.clear()
            next_upload[0] = i * VIEWS_PER_RANK

        def per_view(j, stream):
            g = i * VIEWS_PER_RANK + j
            # the loader runs `lanes` views ahead of the renderer, across step boundaries too
            while next_upload[0] <= g + lanes:
                upload(next_upload[0])
                next_upload[0] += 1
            q, t, wt, ev = uploads.pop(g)
            stream.wait_event(ev)
            loss = render_loss(q, t, wt)
            done = torch.cuda.Event()
            done.record(stream)
            finished[g] = done
            return loss

        total = run_views(per_view)
        finish_step()
        # device -> host read of the step's result, every step, the way a trainer logs a loss: an
        # asynchronous copy into pinned memory, consumed one step later (the host never stalls the
        # queue; the last read completes before the timed region's closing synchronize)
        slot = i % 2
        loss_host[slot].copy_(total.detach().reshape(1), non_blocking=True)
        loss_ready[slot] = torch.cuda.Event()
        loss_ready[slot].record(torch.cuda.current_stream(dev))
        prev = loss_ready[1 - slot]
        if prev is not None:
            prev.synchronize()
            loss_log.append(float(loss_host[1 - slot][0]))
        return None

    def step_render(i):
        v = view_of(i, 0)
        with torch.no_grad():
            image, _, _ = ras(Ras.GaussianPointCloudRasterisationInput(
                point_cloud=xyz, point_cloud_features=feat, point_object_id=oid,
                point_invalid_mask=inv, camera_info=cam, q_pointcloud_camera=qs_dev[v:v + 1],
                t_pointcloud_camera=ts_dev[v:v + 1], color_max_sh_band=3))
        return image

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, steps, warmup):
        """Device time of exactly `steps` steps: barrier+sync, event pair, barrier+sync; max over ranks."""
        for i in range(warmup):
            step_fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            step_fn(warmup + i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    warmup = max(args.warmup, 3)
    sampler = ClockSampler(local_rank)
    prof = StageProfiler()
    prio = bool(args.priority_sections) and lanes > 1
    impl.PRIORITY_SECTIONS = prio
    timed(step_resident, 0, warmup)            # untimed warm-up (allocator, module load)
    sampler.start()
    ms_resident = timed(step_resident, args.steps, 0)
    impl.PRIORITY_SECTIONS = False
    # per-stage CUDA-event times (stages, roofline): the same steps once more on ONE stream, because
    # events bracket a stage only when nothing else is in flight (all ranks run it: all-reduce inside)
    stage_steps = min(args.steps, 10)
    timed(step_resident_single_stream, 0, 1)
    impl.STAGE_PROFILER = prof if rank == 0 else None
    ms_single_stream = timed(step_resident_single_stream, stage_steps, 0)
    stage_ms = prof.summary() if rank == 0 else {}
    impl.STAGE_PROFILER = None
    impl.PRIORITY_SECTIONS = prio
    ms_e2e = timed(step_e2e, args.steps, warmup)
    impl.PRIORITY_SECTIONS = False
    ms_render = timed(step_render, args.steps, warmup)
    clocks = sampler.stop()
    if args.timeline and rank == 0:
        from torch.profiler import ProfilerActivity, profile
        impl.PRIORITY_SECTIONS = prio
        step_resident(0)
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as tl:
            step_resident(1)
            step_resident(2)
            torch.cuda.synchronize()
        impl.PRIORITY_SECTIONS = False
        tl.export_chrome_trace(args.timeline)

    info = ras.last_forward_info
    m, k = int(info["num_points_in_camera"]), int(info["num_keys"])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    tiles, hw = (w // 16) * (h // 16), w * h
    hbm_peak, peak_src = peaks()
    sbytes = stage_bytes(n, m, k, tiles, hw, info)
    stages = {}
    for name, rec in stage_ms.items():
        per_call = rec["ms"] / max(rec["calls"], 1)
        entry = {"ms": round(per_call, 4), "calls": rec["calls"]}
        if name in sbytes:
            gbs = sbytes[name] / (per_call * 1e-3) / 1e9
            entry.update(algorithmic_bytes=sbytes[name], gb_s=round(gbs, 1),
                         hbm_frac=round(gbs / hbm_peak, 4))
        stages[name] = entry
    kernel_stages = [s for s in stages if s in sbytes]
    dominant = max(kernel_stages, key=lambda s: stages[s]["ms"]) if kernel_stages else None

    cpu_baseline = None
    pairs = {}
    if not args.no_cpu_baseline:
        from oracle import gs_oracle
        gs_oracle.build()
        o = gs_oracle.get(False)
        dt, f, b = oracle_step(o, scene, qs_np[0:1], ts_np[0:1], wt_np)
        pairs = {"pairs_fwd": f["pairs_fwd"], "pairs_bwd": b["pairs_bwd"]}
        cpu_baseline = {"value": 1.0 / dt, "unit": "it/s", "cores": o.num_threads(), "kind": "port",
                        "sample": "one full fwd+bwd step of view 0 (whole frame, every stage), "
                                  f"{dt:.1f} s; oracle/gs_oracle.c with OpenMP"}
        for name, key, flop in (("blend_fwd", "pairs_fwd", FLOP_PER_PAIR_FWD),
                                ("blend_bwd", "pairs_bwd", FLOP_PER_PAIR_BWD)):
            if name in stages:
                tf = pairs[key] * flop / (stages[name]["ms"] * 1e-3) / 1e12
                stages[name].update(pairs=pairs[key], fp32_tflops=round(tf, 2),
                                    fp32_frac=round(tf / FP32_PEAK_TFLOPS, 4))

    roofline = None
    if dominant is not None:
        d = stages[dominant]
        roofline = {"kernel": dominant, "bound": "hbm", "achieved": d["gb_s"], "peak": hbm_peak,
                    "unit": "GB/s", "frac": round(d["gb_s"] / hbm_peak, 4), "traffic": None,
                    "peak_source": peak_src, "ms_per_launch": d["ms"],
                    "algorithmic_bytes": d["algorithmic_bytes"]}
        try:  # measured DRAM bytes per launch of that kernel, from the committed ncu --set full capture
            with open(os.path.join(ROOT, "profiles", "ncu_dram_traffic.json")) as f:
                entry = json.load(f).get(args.workload, {}).get(dominant)
            if entry:
                roofline.update(traffic=entry["bytes"], traffic_source=entry["source"])
        except (OSError, ValueError):
            pass
        if "fp32_tflops" in d:  # the blend kernels are FP32-issue bound, not HBM bound
            roofline.update(alt_bound="fp32", alt_achieved=d["fp32_tflops"],
                            alt_peak=round(FP32_PEAK_TFLOPS, 1), alt_unit="TFLOP/s",
                            alt_frac=d["fp32_frac"])

    depth_bits, tile_bits = int(info["depth_bits"]), int(info["tile_bits"])
    if info.get("key_pipeline") == "binned":
        # pose, project | point sort: hist+scan+passes | bin: count, columns, rows, ranges |
        # blend_fwd, blend_bwd, point_bwd   (memsets are not counted)
        launches_per_step = 2 + (2 + (depth_bits + 7) // 8) + 4 + 3
    elif info.get("key_pipeline") == "depth_first":
        # pose, project | point sort: hist+scan+passes | emit | pair sort: hist+scan+passes |
        # ranges, blend_fwd, blend_bwd, point_bwd
        launches_per_step = 2 + (2 + (depth_bits + 7) // 8) + 1 + (2 + (tile_bits + 7) // 8) + 4
    else:
        launches_per_step = 2 + 1 + (2 + (depth_bits + tile_bits + 7) // 8) + 4
    h2d = VIEWS_PER_RANK * (wt_host.numel() * 4 + 16 + 12)
    iters = world * args.steps * VIEWS_PER_RANK   # view forward+backward passes in the timed region
    line = {
        "metric": METRIC, "value": iters / (ms_resident * 1e-3), "unit": "it/s",
        "n_gpus": world, "steps": args.steps, "warmup": warmup,
        "ms_per_step": ms_resident / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "points": n, "width": w, "height": h, "sh_degree": 3,
                   "near": 0.8, "far": 1000.0, "depth_to_sort_key_scale": 100.0,
                   "visible_points": m, "keys": k, "views": NUM_VIEWS,
                   "step": f"{VIEWS_PER_RANK} views forward+backward per rank, gradients summed in "
                           "place, one all-reduce; an iteration (it) = one view forward+backward",
                   "views_per_rank_per_step": VIEWS_PER_RANK,
                   "lanes": lanes, "priority_sections": prio,
                   "lanes_note": f"view j of a step runs on CUDA stream j mod {lanes}; `stages` / `roofline` are "
                                 f"timed in a separate single-stream pass of {stage_steps} steps "
                                 f"({ms_single_stream / stage_steps:.3f} ms per step there)",
                   "key_pipeline": info.get("key_pipeline"),
                   "parallelism": f"views sharded over {world} rank(s), replicated cloud"
                                  + (", NCCL all-reduce of the Nx59 gradient per step" if world > 1 else ""),
                   "l2": "inputs exceed L2 (236 MB of point data read per step)"},
        "e2e": {"value": iters / (ms_e2e * 1e-3), "unit": "it/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                "ms_per_step": ms_e2e / args.steps,
                "loss_readback": "every step, asynchronous copy into pinned memory, read one step later",
                "losses_read": len(loss_log)},
        "render_mpix_per_s": world * args.steps * hw / (ms_render * 1e-3) / 1e6,
        "render_ms_per_frame": ms_render / args.steps,
        "gpu_launches": launches_per_step * VIEWS_PER_RANK * args.steps,
        "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline, "stages": stages,
    }
    _emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    # stdout carries exactly one JSON line: anything libraries print there meanwhile (NCCL's version
    # banner, for one) is sent to stderr
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        _run(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
    if _LINE:
        print(_LINE[0], flush=True)


_LINE = []


def _emit(line: dict) -> None:
    _LINE.append(json.dumps(line))


def _run(args):
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
