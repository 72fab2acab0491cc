"""Config 4b alone (10x10 bunny grid, 13.9 M triangles, 1080p, 16 spp, 4 bounces): warm-up launches + timed launches, one JSON line.
For ncu: `-k regex:pathtrace_megakernel -s 2 -c 1` captures the first timed launch.
    python benchmarks/grid_step.py [--steps 3] [--variant 0] [--which b]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import gpu_util  # noqa: E402
import util  # noqa: E402
from _loader import load_package  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--variant", type=int, default=0)
ap.add_argument("--which", default="b")
a = ap.parse_args()
rt = load_package()
W, H = 1920, 1080
scene, mesh, pose, _ = util.scaled_scene(rt, a.which, env=True)
g = gpu_util.DeviceScene(rt, torch, W, H, pose, seed=1234, env=True, scene=scene)
f = lambda: g.process(0, spp=16, bounces=4, write_last=1, variant=a.variant)
for _ in range(2):
    f()
torch.cuda.synchronize()
g.counter.zero_()
ms = []
for k in range(a.steps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); f(); e1.record(); torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1))
print(json.dumps({"config": "4" + a.which, "variant": a.variant, "ms": ms, "mrays_s": g.rays() / sum(ms) / 1e3, "accel": {k: int(v) for k, v in scene.accel_info().items()}}))
