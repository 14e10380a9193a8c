#!/usr/bin/env python
"""Summary of a kernel timeline written by `bench.py --timeline FILE` (chrome trace from
torch.profiler / CUPTI): GPU busy fraction (union of kernel intervals), time with 1 / 2 / 3+ kernels
in flight, per-kernel mean duration (to compare with the stand-alone times of `stages`), and the
longest idle gaps.     python tools/timeline_summary.py FILE [--top 12]"""
import collections
import json
import re
import sys


def main():
    d = json.load(open(sys.argv[1]))
    ev = d["traceEvents"] if isinstance(d, dict) else d
    ks = [e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "dur" in e]
    ks.sort(key=lambda e: e["ts"])
    t0, t1 = ks[0]["ts"], max(e["ts"] + e["dur"] for e in ks)
    pts = []
    for e in ks:
        pts.append((e["ts"], 1))
        pts.append((e["ts"] + e["dur"], -1))
    pts.sort()
    depth, last, hist, gaps = 0, t0, collections.Counter(), []
    for t, d_ in pts:
        if t > last:
            hist[min(depth, 3)] += t - last
            if depth == 0:
                gaps.append((t - last, last - t0))
        depth += d_
        last = t
    span = t1 - t0
    print(f"span {span / 1e3:.3f} ms, {len(ks)} device activities, streams {len(set(e['args'].get('stream') for e in ks))}")
    for k in range(4):
        print(f"  {k}{'+' if k == 3 else ' '} in flight: {hist[k] / 1e3:8.3f} ms  {100 * hist[k] / span:5.1f} %")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for e in ks:
        m = re.search(r"(\w+_kernel|\w+)(?:<|\(|$)", e["name"].replace("void ", "").replace("(anonymous namespace)::", ""))
        name = m.group(1) if m else e["name"][:40]
        agg[name][0] += 1
        agg[name][1] += e["dur"]
    print("kernel, launches, mean us, total ms, share of summed kernel time")
    tot = sum(v[1] for v in agg.values())
    for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:16]:
        print(f"  {name:42s} {n:5d} {t / n:9.1f} {t / 1e3:9.3f} {100 * t / tot:5.1f} %")
    print(f"summed kernel time {tot / 1e3:.3f} ms = {tot / span:.2f} x span")
    gaps.sort(reverse=True)
    print("longest idle gaps (us, at ms):", [(round(g, 1), round(at / 1e3, 3)) for g, at in gaps[:8]])


if __name__ == "__main__":
    main()
