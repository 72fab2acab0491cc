"""The cost-ordered tile schedule under a MOVING camera: an orbit of `frames` poses around pose B, one 64-spp
frame per pose, the tile order of every frame taken from the PREVIOUS pose's measured tile costs; against the same frames without
any history (grid order) and against a static camera. One JSON line."""
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import gpu_util  # noqa: E402
from _loader import load_package  # noqa: E402

rt = load_package()
W, H, FRAMES = 1920, 1080, 12
d = gpu_util.DeviceScene(rt, torch, W, H, "poseB", seed=1234, env=True)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def run(poses, ordered):
    d.workspace.zero_()
    ms = []
    for k, pose in enumerate(poses):
        d.set_pose(pose)
        flush.fill_(k & 0xFF)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); d.process(0, spp=64, bounces=4, write_last=1, ordered=ordered); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return ms


def orbit(step_deg):
    out = []
    for k in range(FRAMES):
        a = math.radians(step_deg * k)
        out.append(((32.0 + 6.0 * math.sin(a), -1.0 + 0.5 * math.sin(2 * a), 2.0 + 3.0 * (1 - math.cos(a))), 2.0 * math.sin(a), 180.0 + step_deg * k * 0.5))
    return out


static = [((32.0, -1.0, 2.0), 0.0, 180.0)] * FRAMES
run(static, True)  # warm everything
res = {"frames": FRAMES, "what": "1080p, 64 spp, 4 bounces per frame; ms per frame, first frame of a sequence has no tile-cost history"}
for name, poses in (("static", static), ("orbit_2deg_per_frame", orbit(2.0)), ("orbit_10deg_per_frame", orbit(10.0))):
    a, b = run(poses, True), run(poses, False)
    res[name] = {"cost_ordered_ms": [round(x, 2) for x in a], "grid_order_ms": [round(x, 2) for x in b],
                 "mean_after_first": [round(sum(a[1:]) / (FRAMES - 1), 2), round(sum(b[1:]) / (FRAMES - 1), 2)]}
print(json.dumps(res))
