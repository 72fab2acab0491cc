"""How far is the cost-ordered tile schedule from a perfectly packed one? Reads the per-tile cost the megakernel records
(slowest warp of each 16x8 tile, SM cycles) and compares sum(cost) / resident CTA slots with the launch time. One JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import gpu_util  # noqa: E402
from _loader import load_package  # noqa: E402

rt = load_package()
W, H = 1920, 1080
d = gpu_util.DeviceScene(rt, torch, W, H, "poseB", seed=1234, env=True)
out = {}
for ordered in (False, True):
    d.workspace.zero_()
    ms = []
    for k in range(4):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); d.process(k, spp=64, bounces=4, write_last=1, ordered=True); b.record(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
        if not ordered:
            d.workspace[:16].zero_()  # forget the history: the next launch runs in grid order again but still records costs
    ws = d.workspace.view(torch.int32).cpu().numpy().view(np.uint32)
    n = int(ws[2]) if ws[2] else (W // 16) * (H // 8)
    cost = ws[4 + n:4 + 2 * n].astype(np.float64) * 16.0
    slots = 148 * 10
    clock = 1.965e9
    out["ordered" if ordered else "grid order"] = {
        "launch_ms": ms[-1], "tiles": n, "sum_cost_ms_per_slot": cost.sum() / slots / clock * 1e3,
        "max_tile_ms": cost.max() / clock * 1e3, "mean_tile_ms": cost.mean() / clock * 1e3,
        "p99_tile_ms": float(np.percentile(cost, 99)) / clock * 1e3}
print(json.dumps(out))
