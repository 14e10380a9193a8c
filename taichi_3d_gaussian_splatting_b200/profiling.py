"""Per-stage CUDA-event timing of the rasteriser (used by bench.py for the roofline numbers).

Install a StageProfiler in GaussianPointCloudRasterisation.STAGE_PROFILER and every stage of the
forward / backward records one event on the current stream when it has been enqueued; the time
between two consecutive events is charged to the later one's stage.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, List, Tuple

import torch


class StageProfiler:
    def __init__(self):
        self.events: List[Tuple[str, torch.cuda.Event]] = []

    def mark(self, name: str) -> None:
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        self.events.append((name, e))

    def reset(self) -> None:
        self.events = []

    def summary(self) -> "OrderedDict[str, Dict[str, float]]":
        """{stage: {"ms": total, "calls": n}}; stages named *_begin only open an interval."""
        torch.cuda.synchronize()
        out: "OrderedDict[str, Dict[str, float]]" = OrderedDict()
        for (_, e0), (name, e1) in zip(self.events, self.events[1:]):
            if name.endswith("_begin"):
                name = "between_calls"
            rec = out.setdefault(name, {"ms": 0.0, "calls": 0})
            rec["ms"] += e0.elapsed_time(e1)
            rec["calls"] += 1
        return out
