#!/usr/bin/env python
"""Aggregate an ncu report's per-instruction counters by CUDA source line.

    python tools/ncu_by_line.py <report.ncu-rep> <kernel-regex> <cubin> <mangled-substring> [top] [launch]

`launch` picks the n-th matching launch of the report (default 0).

ncu's `--page source --csv` lists SASS instructions in program order with executed counts and
stall samples; `nvdisasm -g -c` lists the same instructions with `//## File ..., line N` marks.
The two are joined by instruction index.  (Run where ncu and nvdisasm exist; no GPU needed.)
"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict


def main():
    rep, kregex, cubin, mangled = sys.argv[1:5]
    top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name",
                          f"regex:{kregex}"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    launch = int(sys.argv[6]) if len(sys.argv) > 6 else 0
    hdr_i = [i for i, r in enumerate(rows) if r and r[0] == "Address"][launch]
    hdr = rows[hdr_i]
    body = []
    for r in rows[hdr_i + 1:]:
        if r and r[0] == "Kernel Name":
            break  # only the first matching launch
        if len(r) == len(hdr):
            body.append(r)
    ci = hdr.index("Instructions Executed")
    si = hdr.index("# Samples")
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    lines, cur, active = [], None, False
    for ln in dis.splitlines():
        if ln.startswith("//---") and ".text." in ln:
            active = mangled in ln
            continue
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
            lines.append((cur, ln.strip()))
    if len(lines) != len(body):
        print(f"warning: {len(lines)} disassembled vs {len(body)} profiled instructions", file=sys.stderr)
    agg = defaultdict(lambda: [0, 0, 0])
    total = total_s = 0
    for (loc, _), r in zip(lines, body):
        n, s = int(float(r[ci] or 0)), int(float(r[si] or 0))
        agg[loc][0] += n
        agg[loc][1] += s
        agg[loc][2] += 1
        total += n
        total_s += s
    print(f"total warp instructions {total}, samples {total_s}, static instructions {len(body)}")
    src_cache = {}
    for loc, (n, s, k) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        text = ""
        if loc:
            try:
                if loc[0] not in src_cache:
                    path = subprocess.run(["find", ".", "-name", loc[0]], capture_output=True,
                                          text=True).stdout.split()[0]
                    src_cache[loc[0]] = open(path).read().splitlines()
                text = src_cache[loc[0]][loc[1] - 1].strip()[:90]
            except Exception:
                pass
        print(f"{100*n/total:5.1f}% inst {100*s/max(total_s,1):5.1f}% smpl {k:4d} sass  {loc}  {text}")


if __name__ == "__main__":
    main()
