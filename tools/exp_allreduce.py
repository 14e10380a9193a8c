#!/usr/bin/env python
"""Experiment: the step's one exchange (sum of the N x 59 fp32 gradient over ranks) with NCCL's
all_reduce against torch's symmetric-memory all-reduces (NVLink multicast / peer loads).

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/exp_allreduce.py
"""
import os
import sys
import traceback

import torch
import torch.distributed as dist


def timed(fn, steps=20, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    ms = torch.tensor([a.elapsed_time(b) / steps], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    numel = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000 * 59
    out = {}
    flat = torch.randn(numel, device=dev)
    out["nccl_all_reduce"] = timed(lambda: dist.all_reduce(flat))
    try:
        import torch.distributed._symmetric_memory as symm_mem
        group = dist.group.WORLD
        t = symm_mem.empty(numel, dtype=torch.float32, device=dev)
        symm_mem.rendezvous(t, group.group_name)
        t.copy_(torch.randn(numel, device=dev) * 0.01)
        ref = t.clone()
        dist.all_reduce(ref)
        for name in ("multimem_all_reduce_", "two_shot_all_reduce_"):
            try:
                op = getattr(torch.ops.symm_mem, name)
                probe = t.clone()
                t2 = symm_mem.empty(numel, dtype=torch.float32, device=dev)
                symm_mem.rendezvous(t2, group.group_name)
                t2.copy_(probe)
                dist.barrier()
                op(t2, "sum", group.group_name)
                torch.cuda.synchronize()
                err = float((t2 - ref).abs().max())
                out[name] = timed(lambda: op(t2, "sum", group.group_name))
                out[name + "_max_err"] = err
            except Exception as e:  # noqa: BLE001
                out[name] = f"failed: {type(e).__name__}: {str(e)[:200]}"
    except Exception as e:  # noqa: BLE001
        out["symm_mem"] = f"failed: {type(e).__name__}: {str(e)[:300]}"
        traceback.print_exc()
    if rank == 0:
        print({"world": world, "bytes": numel * 4, **out})
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
