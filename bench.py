#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 rasteriser (contract in the task statement, section 4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

The workload is the configuration BASELINE.json's metric is quoted on: configs[1] = 1M random
Gaussians, SH degree 3, 1920x1088 (1080p rounded up to whole 16x16 tiles),
generated as SURVEY.md section 8(d) prescribes.

  value        fwd+bwd iterations/s with every input already resident in HBM (CUDA events).  A step is
               one batch of 8 views per rank, gradients summed in place into one N x 59 buffer, views
               spread over `lanes` CUDA streams; for N > 1 every step ends with an NCCL all-reduce of
               that buffer (views sharded over ranks against a replicated cloud).
  e2e          the same through the public module with HOST inputs: per view the camera pose and the
               per-pixel target come from pinned host memory (what the reference's trainer uploads per
               step, Tr.py:149-159) and the scalar loss is read back every step.
  single_view  what the UNCHANGED reference trainer sees: plain `module(input)` + `loss.backward()`, one
               view per call, fresh gradient tensors from autograd, one stream, no batch extension.
  render_e2e   the render half of the metric with the reference's own timing method
               (benchmark/inference_benchmark.py:13-14,133-160: long warm-up, then 100 timed forwards under
               no_grad) and what gaussian_point_render.py:100-121 does per frame: pose from the host, the
               frame converted to 8-bit and copied to host memory.
  c4           BASELINE.json configs[3] as written: 64 DISTINCT views x 2M Gaussians at 976x544, the 64
               views of a step split over the ranks (strong scaling), longest-processing-time-first
               assignment from the previous pass's per-view work, one all-reduce per step, and an on-box
               check that the reduced gradient equals rank 0 rendering all 64 views alone.
  --impl reference   the CPU oracle (the reference's Taichi kernels cannot run here: no Taichi)
               timed on ALL host cores with OpenMP, same config and metric.
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
METRIC = "fwd+bwd iters/s"  # BASELINE.json metric, first half; the render half rides along (render_*)
NUM_VIEWS = 8
VIEWS_PER_RANK = int(os.environ.get("GS_BENCH_VIEWS_PER_STEP", "8"))   # views each rank renders per step of the headline leg (weak scaling; the env override is for experiments)
C4_VIEWS = 64        # BASELINE.json configs[3]: 64 camera views per step, split over the ranks
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.5 (theoretical: 148 SMs x 128 lanes x FMA at 1965 MHz)
FLOP_PER_PAIR_FWD, FLOP_PER_PAIR_BWD = 28.0, 64.0     # SURVEY.md section 8(d) convention
RASTER_CFG = (0.8, 1000.0, 100.0)                     # near, far, depth_to_sort_key_scale (R.py:778-780)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2_1m_1920x1088", choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-c4", action="store_true", help="skip the 64-view x 2M-point strong-scaling leg")
    ap.add_argument("--render-frames", type=int, default=100)
    ap.add_argument("--render-warmup", type=int, default=1000)
    ap.add_argument("--timeline", default=None,
                    help="also write the kernel timeline (name, stream, start, duration; CUPTI through "
                         "torch.profiler) of two untimed resident steps to this JSON file")
    ap.add_argument("--priority-sections", type=int, default=0,
                    help="with lanes > 1: the small kernels of a view (projection, sorts, binning, per-point "
                         "backward) run on a high-priority companion stream of the lane")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="N > 1: the step's exchange -- fused into the per-point pass over NVLink multicast "
                         "(falls back to nccl if symmetric memory has no multicast mapping), or NCCL's all_reduce")
    ap.add_argument("--exchange-chunks", type=int, default=1,
                    help="N > 1: point ranges the per-point backward of a step is cut into, each all-reduced as "
                         "soon as it is written (1 = one all-reduce after the pass)")
    ap.add_argument("--c4-assign", default="time", choices=["time", "model"],
                    help="c4 leg: what the longest-processing-time-first assignment balances -- the measured "
                         "time of every view alone in the previous pass, or the cost model K + 3 M")
    ap.add_argument("--c4-lanes", type=int, default=8,
                    help="lanes of the c4 leg (its 976x544 views fill the machine less than a 1920x1088 one: "
                         "8 lanes run a rank's 8 views 10 %% faster than 4, tools/exp_c4_rank_step.py)")
    ap.add_argument("--lanes", type=int, default=4,
                    help="CUDA streams the views of a step are spread over (view j on lane j mod L)")
    return ap.parse_args()


def make_views(num_views: int, seed: int = 5):
    """Nearby camera poses (T_pointcloud_camera as xyzw quaternion + translation): the headline leg
    keeps every view on the configuration the metric is quoted on."""
    rng = np.random.default_rng(seed)
    q = np.zeros((num_views, 4), np.float32)
    q[:, 3] = 1.0
    q[1:, :3] = rng.normal(0, 0.01, (num_views - 1, 3))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    t = np.zeros((num_views, 3), np.float32)
    t[1:] = rng.normal(0, 0.05, (num_views - 1, 3))
    return q.astype(np.float32), t.astype(np.float32)


def make_views_spread(num_views: int, seed: int = 9):
    """DISTINCT poses with a real spread (the C4 leg): the camera turns by up to +-20 degrees about its
    vertical and horizontal axes, rolls by up to +-10, and moves sideways / back and forth inside the
    near part of the cloud, so that the visible points M and the list entries K differ by well over 2x
    between views (printed in the bench line)."""
    rng = np.random.default_rng(seed)
    q = np.zeros((num_views, 4), np.float32)
    t = np.zeros((num_views, 3), np.float32)
    for i in range(num_views):
        yaw, pitch = np.radians(rng.uniform(-20, 20, 2))
        roll = np.radians(rng.uniform(-10, 10))
        shift = [rng.uniform(-0.6, 0.6), rng.uniform(-0.4, 0.4), rng.uniform(-1.5, 3.0)]
        if i == 0:
            yaw = pitch = roll = 0.0
            shift = [0.0, 0.0, 0.0]
        cy, sy, cp, sp, cr, sr = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch), np.cos(roll), np.sin(roll)
        Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
        Rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
        Rz = np.array([[cr, -sr, 0], [sr, cr, 0], [0, 0, 1]])
        R = Ry @ Rx @ Rz
        w = np.sqrt(max(0.0, 1.0 + R[0, 0] + R[1, 1] + R[2, 2])) / 2.0
        q[i] = [(R[2, 1] - R[1, 2]) / (4 * w), (R[0, 2] - R[2, 0]) / (4 * w), (R[1, 0] - R[0, 1]) / (4 * w), w]
        t[i] = shift
    q /= np.linalg.norm(q, axis=1, keepdims=True)
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
        # what the timed region really zeroes: the M x 12 reduced splat gradients (the N x 59 zero fill of
        # R.py:1051-1067 is fused away by the in-place accumulation; it is paid once per STEP, outside)
        "zero_fill": 48 * m,
        "blend_bwd": 44 * k + 28 * hw + 8 * hw + 44 * m,
        "point_bwd": 280 * m + 232 * m,
        # per STEP of V views: every feature row in, every gradient row out, per view the slot map and,
        # per visible point, 48 bytes of reduced gradients + 12 of camera-space position
        "point_bwd_batch": (224 + 236) * n + VIEWS_PER_RANK * (4 * n + 60 * m),
    }


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def oracle_step(o, scene, q, t, wt, cfg=RASTER_CFG, band=None):
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
    Under torchrun only rank 0 works -- on every host core (torchrun exports OMP_NUM_THREADS=1 to its
    workers, so the thread count is set explicitly)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from oracle import gs_oracle
    from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
    gs_oracle.build()
    o = gs_oracle.get(False)
    o.set_num_threads(host_threads())
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
                "= C restatement of the same path, OpenMP over points/tiles, one view per step",
    }
    _emit(line)


class Rig:
    """One workload resident on this rank's GPU: the cloud as nn.Parameters, the operator, one flat
    N x 59 gradient buffer the batch extension sums into, and the lane streams."""

    def __init__(self, torch, impl, dev, workload, lanes, peer_exchange=False):
        from taichi_3d_gaussian_splatting_b200.Camera import CameraInfo
        from taichi_3d_gaussian_splatting_b200.synthetic import make_scene
        self.torch, self.impl, self.dev = torch, impl, dev
        self.Ras = impl.GaussianPointCloudRasterisation
        n, w, h, sf, seed = WORKLOADS[workload]
        self.n, self.w, self.h, self.workload = n, w, h, workload
        self.scene = make_scene(n, w, h, seed=seed, scale_factor=sf)
        self.xyz = torch.nn.Parameter(torch.from_numpy(self.scene.point_cloud).to(dev))
        self.feat = torch.nn.Parameter(torch.from_numpy(self.scene.point_cloud_features).to(dev))
        self.inv = torch.from_numpy(self.scene.point_invalid_mask).to(dev)
        self.oid = torch.from_numpy(self.scene.point_object_id).to(dev)
        self.cam = CameraInfo(torch.from_numpy(self.scene.camera_intrinsics).to(dev), h, w, 0)
        self.wt_np = np.random.default_rng(3).uniform(0, 1, (h, w, 3)).astype(np.float32)
        self.wt_host = torch.from_numpy(self.wt_np).pin_memory()
        self.wt_dev = self.wt_host.to(dev)
        self.ras = self.Ras(self.Ras.GaussianPointCloudRasterisationConfig())
        # features first: their rows are added as 16-byte vectors (224 n is a multiple of 16 for every n)
        # N > 1 on NVSwitch: the buffer is symmetric memory and the step ends in the fused pass + exchange
        # kernel (gs_point_bwd_batch_exchange); otherwise (one GPU, no multicast mapping, --exchange nccl) a
        # plain buffer, the pass, and NCCL's all_reduce
        self.peer, self.peer_note = None, "nccl all_reduce after the per-point pass"
        if peer_exchange:
            try:
                from taichi_3d_gaussian_splatting_b200.distributed import PeerExchange
                self.peer = PeerExchange(n, dev)
                self.peer_note = ("fused: gs_point_bwd_batch_exchange (ranges reduced in the NVLink switch with "
                                  f"multimem.ld_reduce / multimem.st while later ranges are computed, "
                                  f"{self.peer.chunks} ranges)")
            except Exception as e:  # noqa: BLE001
                self.peer_note = f"nccl all_reduce (no symmetric-memory multicast: {type(e).__name__}: {str(e)[:120]})"
        self.flat = self.peer.flat if self.peer is not None else torch.zeros(n * 59, dtype=torch.float32, device=dev)
        self.gf, self.gx = self.flat[:56 * n].view(n, 56), self.flat[56 * n:].view(n, 3)
        self.lanes = max(1, lanes)
        self.lane_streams = [torch.cuda.Stream(dev) for _ in range(self.lanes)] if self.lanes > 1 else []

    def begin_step(self):
        """The views that follow are a gradient batch: their backward parks the reduced screen-space
        gradients, `finish_step` turns all of them into point gradients in one pass over the points."""
        self.ras.begin_gradient_batch(self.gx, self.gf)

    def finish_step(self, dist, world, chunks=1):
        """Per-point backward of the step's views (gs_point_bwd_batch) and, for world > 1, the exchange: one
        all-reduce of the flat N x 59 buffer.  With chunks > 1 the pass is cut into that many point ranges
        and the all-reduce of a range starts as soon as the range is written (measured, not adopted: at
        N = 8 eight ranges cost 14.17 ms per step against 13.3 with one all-reduce -- sixteen small
        collectives cost more than the 0.3 ms of computation they can hide behind)."""
        if world > 1 and self.peer is not None:
            self.ras.end_gradient_batch(exchange=self.peer)
            self.xyz.grad, self.feat.grad = self.gx, self.gf
            return
        pending = []

        def after_chunk(index, first, count):
            if count == self.n:
                pending.append(dist.all_reduce(self.flat, async_op=True))
            else:
                pending.append(dist.all_reduce(self.gf[first:first + count], async_op=True))
                pending.append(dist.all_reduce(self.gx[first:first + count], async_op=True))

        self.ras.end_gradient_batch(chunks if world > 1 else 1, after_chunk if world > 1 else None)
        for work in pending:
            work.wait()
        self.xyz.grad, self.feat.grad = self.gx, self.gf

    def forward(self, q, t):
        Ras = self.Ras
        return self.ras(Ras.GaussianPointCloudRasterisationInput(
            point_cloud=self.xyz, point_cloud_features=self.feat, point_object_id=self.oid,
            point_invalid_mask=self.inv, camera_info=self.cam, q_pointcloud_camera=q,
            t_pointcloud_camera=t, color_max_sh_band=3))

    def render_loss(self, q, t, wt):
        image, _, _ = self.forward(q, t)
        loss = (image * wt).sum()
        loss.backward()
        return loss.detach()

    def run_views(self, count, per_view, use_lanes=True):
        """`count` views, view j on lane j mod L (the views are independent: their gradients are summed
        into `flat` by RED either way).  The hot kernels are issue-bound, the key pipeline is latency /
        HBM bound: views in flight fill each other's idle slots."""
        torch, dev = self.torch, self.dev
        cur = torch.cuda.current_stream(dev)
        if not (use_lanes and self.lane_streams) or count <= 1:
            total = torch.zeros((), device=dev)
            for j in range(count):
                total += per_view(j, cur)
            return total
        start = torch.cuda.Event()
        start.record(cur)
        totals = []
        for s in self.lane_streams:
            s.wait_event(start)
            with torch.cuda.stream(s):
                totals.append(torch.zeros((), device=dev))
        for j in range(count):
            s = self.lane_streams[j % self.lanes]
            with torch.cuda.stream(s):
                totals[j % self.lanes] += per_view(j, s)
        for s in self.lane_streams:
            e = torch.cuda.Event()
            e.record(s)
            cur.wait_event(e)
        return torch.stack(totals).sum()


def run_b200(args):
    import torch
    import torch.distributed as dist

    import taichi_3d_gaussian_splatting_b200.GaussianPointCloudRasterisation as impl
    from taichi_3d_gaussian_splatting_b200.profiling import StageProfiler

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

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

    lanes = max(1, min(args.lanes, VIEWS_PER_RANK))
    rig = Rig(torch, impl, dev, args.workload, lanes, peer_exchange=world > 1 and args.exchange == "fused")
    n, w, h = rig.n, rig.w, rig.h
    xyz, feat, ras, flat, gx, gf = rig.xyz, rig.feat, rig.ras, rig.flat, rig.gx, rig.gf
    qs_np, ts_np = make_views(NUM_VIEWS)
    # host side of the e2e leg: pinned buffers
    qs_host = torch.from_numpy(qs_np).pin_memory()
    ts_host = torch.from_numpy(ts_np).pin_memory()
    wt_host, wt_dev = rig.wt_host, rig.wt_dev
    qs_dev, ts_dev = qs_host.to(dev), ts_host.to(dev)   # device-resident copies for the kernel-only leg
    copy_stream = torch.cuda.Stream(dev)

    def view_of(step, j):  # global view index of this rank's j-th view of the step
        return ((step * world + rank) * VIEWS_PER_RANK + j) % NUM_VIEWS

    def finish_step():
        # per-point backward of the 8 views + the exchange of the N x 59 gradient
        rig.finish_step(dist, world, args.exchange_chunks)

    def resident_view(i):
        def per_view(j, stream):
            v = view_of(i, j)
            return rig.render_loss(qs_dev[v:v + 1], ts_dev[v:v + 1], wt_dev)
        return per_view

    def step_resident(i):
        rig.begin_step()
        total = rig.run_views(VIEWS_PER_RANK, resident_view(i))
        finish_step()
        return total

    def step_resident_single_stream(i):  # the instrumented pass: per-stage CUDA events need one stream
        rig.begin_step()
        total = rig.run_views(VIEWS_PER_RANK, resident_view(i), use_lanes=False)
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
        rig.begin_step()
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
            loss = rig.render_loss(q, t, wt)
            done = torch.cuda.Event()
            done.record(stream)
            finished[g] = done
            return loss

        total = rig.run_views(VIEWS_PER_RANK, per_view)
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

    def step_single_view(i):
        """One iteration as the unchanged reference trainer runs it (Tr.py:143-184 without the optimiser):
        gradients set to None, one module call, loss.backward() -- autograd allocates and fills fresh
        N x 3 / N x 56 gradient tensors, everything on the current stream."""
        xyz.grad = None
        feat.grad = None
        v = view_of(i, 0)
        return rig.render_loss(qs_dev[v:v + 1], ts_dev[v:v + 1], wt_dev)

    def step_render(i):
        v = view_of(i, 0)
        with torch.no_grad():
            image, _, _ = rig.forward(qs_dev[v:v + 1], ts_dev[v:v + 1])
        return image

    frame_host = torch.empty((h, w, 3), dtype=torch.uint8).pin_memory()

    def step_render_e2e(i):
        """gaussian_point_render.py:100-121 per frame: pose from the host, forward under no_grad, the
        frame clamped, converted to 8 bits and brought to host memory (a blocking copy, as `.cpu()`)."""
        v = view_of(i, 0)
        q = qs_host[v:v + 1].to(dev, non_blocking=True)
        t = ts_host[v:v + 1].to(dev, non_blocking=True)
        with torch.no_grad():
            image, _, _ = rig.forward(q, t)
            frame_host.copy_(torch.clamp(image * 255, 0, 255).byte())
        return None

    warmup = max(args.warmup, 3)
    sampler = ClockSampler(local_rank)
    prof = StageProfiler()
    prio = bool(args.priority_sections) and lanes > 1
    impl.PRIORITY_SECTIONS = prio
    # untimed warm-up (allocator, module load); the caching allocator settles after about five steps of
    # 8 views on 4 lanes, so a run asked for fewer warm-up steps gets the difference as extra untimed steps
    timed(step_resident, 0, max(warmup, 5))
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
    info = dict(ras.last_forward_info)
    # the drop-in figure and the two render figures (no batch extension, no lanes)
    single_iters = args.steps * VIEWS_PER_RANK
    ms_single = timed(step_single_view, single_iters, 2 * warmup)
    xyz.grad = feat.grad = None
    ms_render = timed(step_render, args.steps, warmup)
    ms_render_e2e = timed(step_render_e2e, args.render_frames, args.render_warmup)
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

    # release the headline workload before the C4 leg allocates its own
    scene_main = rig.scene
    xyz.grad = feat.grad = None
    rig_exchange_note = rig.peer_note
    exchange_error = rig.peer.error() if rig.peer is not None else 0
    del rig, xyz, feat, ras, flat, gx, gf, wt_stage, uploads, finished
    torch.cuda.empty_cache()
    c4 = None
    if not args.no_c4:
        c4 = run_c4_leg(torch, dist, impl, dev, rank, world, max(1, args.c4_lanes), args, barrier)

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
        o.set_num_threads(host_threads())
        dt, f, b = oracle_step(o, scene_main, qs_np[0:1], ts_np[0:1], wt_host.numpy())
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
        hbm = {"achieved": d["gb_s"], "peak": hbm_peak, "unit": "GB/s", "frac": round(d["gb_s"] / hbm_peak, 4),
               "peak_source": peak_src}
        traffic, traffic_source = None, None
        try:  # measured DRAM bytes per launch of that kernel, from the committed ncu --set full capture
            with open(os.path.join(ROOT, "profiles", "ncu_dram_traffic.json")) as f:
                entry = json.load(f).get(args.workload, {}).get(dominant)
            if entry:
                traffic, traffic_source = entry["bytes"], entry["source"]
        except (OSError, ValueError):
            pass
        if "fp32_tflops" in d:
            # the blend kernels are bound by FP32 instruction issue, not by HBM (ncu: issue slots 75-85 %
            # busy, DRAM 2 %): the binding resource is the primary bound, HBM rides along as alt_*
            flop = FLOP_PER_PAIR_BWD if dominant == "blend_bwd" else FLOP_PER_PAIR_FWD
            roofline = {"kernel": dominant, "bound": "fp32", "achieved": d["fp32_tflops"],
                        "peak": round(FP32_PEAK_TFLOPS, 1), "unit": "TFLOP/s", "frac": d["fp32_frac"],
                        "traffic": traffic, "traffic_source": traffic_source,
                        "peak_source": "THEORETICAL: 148 SM x 128 lanes x 2 flop x 1.965 GHz (MEASURED_PEAKS.json "
                                       f"has no FP32 entry); flops by the SURVEY.md 8(d) convention, {flop:.0f} per "
                                       "VISITED pair -- the kernel skips the pairs that cannot contribute, so this "
                                       "is a useful-work rate, not pipe utilisation",
                        "ms_per_launch": d["ms"], "algorithmic_flop": d["pairs"] * flop,
                        "algorithmic_bytes": d["algorithmic_bytes"],
                        "alt_bound": "hbm", "alt_achieved": hbm["achieved"], "alt_peak": hbm["peak"],
                        "alt_unit": "GB/s", "alt_frac": hbm["frac"], "alt_peak_source": peak_src}
        else:
            roofline = {"kernel": dominant, "bound": "hbm", **hbm, "traffic": traffic,
                        "traffic_source": traffic_source, "ms_per_launch": d["ms"],
                        "algorithmic_bytes": d["algorithmic_bytes"]}

    depth_bits, tile_bits = int(info["depth_bits"]), int(info["tile_bits"])
    if info.get("key_pipeline") == "binned":
        # pose, project | point sort: hist+scan+passes | bin: count, columns, rows, ranges |
        # blend_fwd, blend_bwd + its literal pass   (memsets are not counted; + 1 per step: the batched
        # per-point backward, added below)
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
                   "step": f"{VIEWS_PER_RANK} views forward+backward per rank as one gradient batch (per-point "
                           "chain rule of all views in one pass over the points), then one all-reduce of the "
                           "N x 59 gradient; an iteration (it) = one view forward+backward",
                   "views_per_rank_per_step": VIEWS_PER_RANK,
                   "lanes": lanes, "priority_sections": prio,
                   "lanes_note": f"view j of a step runs on CUDA stream j mod {lanes}; `stages` / `roofline` are "
                                 f"timed in a separate single-stream pass of {stage_steps} steps "
                                 f"({ms_single_stream / stage_steps:.3f} ms per step there)",
                   "key_pipeline": info.get("key_pipeline"),
                   "exchange": rig_exchange_note if world > 1 else "none (one rank)",
                   "exchange_error": exchange_error,
                   "parallelism": f"views sharded over {world} rank(s), replicated cloud"
                                  + (", NCCL all-reduce of the Nx59 gradient per step" if world > 1 else ""),
                   "l2": "inputs exceed L2 (236 MB of point data read per step)"},
        "e2e": {"value": iters / (ms_e2e * 1e-3), "unit": "it/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                "ms_per_step": ms_e2e / args.steps,
                "loss_readback": "every step, asynchronous copy into pinned memory, read one step later",
                "losses_read": len(loss_log)},
        "single_view": {"value": world * single_iters / (ms_single * 1e-3), "unit": "it/s",
                        "ms_per_iteration": ms_single / single_iters, "iterations": single_iters,
                        "how": "the unchanged trainer's call pattern: grads set to None, module(input), "
                               "loss.backward() into fresh autograd-allocated N x 3 / N x 56 tensors, one "
                               "stream, no accumulate_gradients_into, no lanes" + (
                                   "; every rank runs its own replica, no exchange" if world > 1 else "")},
        "render_mpix_per_s": world * args.steps * hw / (ms_render * 1e-3) / 1e6,
        "render_ms_per_frame": ms_render / args.steps,
        "render_e2e": {"mpix_per_s": world * args.render_frames * hw / (ms_render_e2e * 1e-3) / 1e6,
                       "ms_per_frame": ms_render_e2e / args.render_frames, "frames": args.render_frames,
                       "warmup_frames": args.render_warmup, "h2d_bytes_per_frame": 28,
                       "d2h_bytes_per_frame": hw * 3,
                       "how": "benchmark/inference_benchmark.py's method (long warm-up, then one CUDA-event pair "
                              "around the timed forwards under no_grad) with gaussian_point_render.py's frame "
                              "handling: pose uploaded from pinned host memory, clamp(image * 255).byte() copied "
                              "to host memory with a blocking copy every frame"},
        "gpu_launches": (launches_per_step * VIEWS_PER_RANK + 1) * args.steps,
        "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline, "stages": stages,
        "c4": c4,
    }
    _emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_c4_leg(torch, dist, impl, dev, rank, world, lanes, args, barrier):
    """BASELINE.json configs[3]: 64 distinct views x 2M Gaussians, the views of a step split over the
    ranks (STRONG scaling: E(G) = T(1, 64 views) / (G T(G, 64 views)))."""
    name = "c4_2m_976x544"
    rig = Rig(torch, impl, dev, name, lanes, peer_exchange=world > 1 and args.exchange == "fused")
    q_np, t_np = make_views_spread(C4_VIEWS)
    q_dev, t_dev = torch.from_numpy(q_np).to(dev), torch.from_numpy(t_np).to(dev)
    stats = np.zeros((C4_VIEWS, 3), np.float64)   # per view: M, K, ms alone on one stream (recording pass)
    view_events = {}

    def render_views(mine, record=False):
        def per_view(j, stream):
            v = mine[j]
            if record:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            loss = rig.render_loss(q_dev[v:v + 1], t_dev[v:v + 1], rig.wt_dev)
            if record:
                e1.record()
                view_events[v] = (e0, e1)
                info = rig.ras.last_forward_info
                stats[v, 0:2] = (float(info["num_points_in_camera"]), float(info["num_keys"]))
            return loss
        # the recording pass runs the views one after another on one stream, so that the events around a
        # view bracket that view's work alone
        return rig.run_views(len(mine), per_view, use_lanes=not record)

    def step(mine, exchange=True):
        rig.begin_step()
        total = render_views(mine)
        rig.finish_step(dist, world if exchange else 1, args.exchange_chunks)
        return total

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / steps

    # pass 0 (round robin) measures every view's work; from then on the views are dealt out by
    # longest-processing-time-first on it, the way a trainer would use the previous epoch's numbers
    round_robin = [list(range(r, C4_VIEWS, world)) for r in range(world)]
    for _ in range(2):   # the second pass is the one recorded (allocator and caches warm)
        view_events.clear()
        rig.begin_step()
        render_views(round_robin[rank], record=True)
        rig.finish_step(dist, 1)
        torch.cuda.synchronize()
    for v, (e0, e1) in view_events.items():
        stats[v, 2] = e0.elapsed_time(e1)
    if world > 1:
        shared = torch.from_numpy(stats).to(dev)
        dist.all_reduce(shared)
        stats = shared.cpu().numpy()
    m_view, k_view, ms_view = stats[:, 0], stats[:, 1], stats[:, 2]
    from taichi_3d_gaussian_splatting_b200.distributed import assign_views_lpt, view_cost
    model = [view_cost(m, k) for m, k in zip(m_view, k_view)]
    # LPT on the MEASURED time of every view (alone, one stream, previous pass): the cost model K + 3 M
    # correlates 0.8 with it, and the slowest rank is what a step waits for
    costs = [float(x) for x in ms_view] if args.c4_assign == "time" else model
    assignment, loads = assign_views_lpt(costs, world)
    rr_loads = [sum(costs[v] for v in round_robin[r]) for r in range(world)]
    mine = assignment[rank]
    steps = max(3, min(args.steps, 6))
    for _ in range(2):
        step(mine)
    ms_step = timed(lambda: step(mine), steps)
    ms_no_exchange = timed(lambda: step(mine, exchange=False), steps)
    if world > 1:
        step(round_robin[rank])   # other views per rank, other buffer sizes: let the allocator settle first
    ms_round_robin = timed(lambda: step(round_robin[rank]), steps) if world > 1 else ms_step
    ms_exchange_alone = None
    verify = None
    if world > 1:
        # NCCL's all_reduce of a buffer of this size, alone, for comparison (an ordinary allocation: the
        # symmetric buffer of the fused exchange is not one NCCL runs at speed on)
        plain = (torch.zeros(rig.flat.numel(), dtype=torch.float32, device=dev) if rig.peer is not None
                 else rig.flat)
        for _ in range(3):   # NCCL sets its channels up on the first large calls
            dist.all_reduce(plain)
        ms_exchange_alone = timed(lambda: dist.all_reduce(plain), 5)
        del plain
        # on-box check of the exchange: the all-reduced gradient of one step against rank 0 rendering all
        # 64 views by itself into the same kind of buffer (rows differ only by the order of the additions)
        step(mine)
        reduced = rig.flat.clone()
        result = torch.zeros(3, dtype=torch.float64, device=dev)
        if rank == 0:
            step(list(range(C4_VIEWS)), exchange=False)
            diff = (reduced.double() - rig.flat.double())
            result[0] = diff.norm() / rig.flat.double().norm()
            result[1] = diff.abs().max()
            result[2] = rig.flat.double().abs().max()
        dist.broadcast(result, 0)
        verify = {"rel_l2": float(result[0]), "max_abs_diff": float(result[1]), "max_abs": float(result[2]),
                  "tolerance_rel_l2": 1e-5, "ok": bool(float(result[0]) <= 1e-5),
                  "what": f"all-reduced N x 59 gradient of one step ({world} ranks) vs rank 0 rendering all "
                          f"{C4_VIEWS} views alone"}
    barrier()
    out = {
        "workload": name, "points": rig.n, "width": rig.w, "height": rig.h, "views_per_step": C4_VIEWS,
        "scaling": "strong", "n_gpus": world, "steps": steps, "lanes": lanes,
        "value": C4_VIEWS / (ms_step * 1e-3), "unit": "it/s", "ms_per_step": ms_step,
        "ms_per_step_without_exchange": ms_no_exchange,
        "ms_per_step_round_robin_assignment": ms_round_robin,
        "exchange": rig.peer_note if world > 1 else "none (one rank)",
        "exchange_error": rig.peer.error() if rig.peer is not None else 0,
        "exchange_ms_alone": ms_exchange_alone, "exchange_ms_alone_what": "NCCL all_reduce of an ordinary buffer of this size", "exchange_bytes": rig.flat.numel() * 4 if world > 1 else 0,
        "assignment": ("longest-processing-time first on the measured time of every view alone (previous pass)"
                       if args.c4_assign == "time" else
                       "longest-processing-time first on K + 3 M of the previous pass"),
        "view_ms_alone_min_max": [float(ms_view.min()), float(ms_view.max())],
        "cost_model_vs_measured_time_correlation": float(np.corrcoef(np.array(model), ms_view)[0, 1]),
        "rank_load_max_over_mean": max(loads) / (sum(loads) / world),
        "rank_load_max_over_mean_round_robin": max(rr_loads) / (sum(rr_loads) / world),
        "views_per_rank": [len(a) for a in assignment],
        "view_visible_points_min_max": [int(m_view.min()), int(m_view.max())],
        "view_keys_min_max": [int(k_view.min()), int(k_view.max())],
        "verify": verify,
    }
    rig.xyz.grad = rig.feat.grad = None
    del rig
    torch.cuda.empty_cache()
    return out


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
