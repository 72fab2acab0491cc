"""Smallest run that touches every kernel once (for compute-sanitizer): primary/rays queries (both variants), RNG init/draw,
path tracing variants 0/1/2/3/10, AO, tone-map, the facade and its RenderToHost."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402

import gpu_util  # noqa: E402
import util  # noqa: E402
from _loader import load_package  # noqa: E402

rt = load_package()
W, H = 72, 40
d = gpu_util.DeviceScene(rt, torch, W, H, "poseB", seed=3, env=True)
for v in (0, 1):
    d.trace_primary(1, v)
rays = np.random.default_rng(0).normal(size=(5000, 6)).astype(np.float32)
rays[:, :3] = rays[:, :3] * 10 + (30, 0, 20)
for v in (0, 1):
    d.trace_rays(rays, True, v)
work = torch.zeros(2, dtype=torch.int64, device="cuda")
from cuda_raytracing_b200.api import process_ex
for v in (0, 1, 10):
    d.process(0, spp=3, bounces=4, variant=v, write_last=1)
    d.process(1, spp=2, bounces=4, variant=v, write_last=1)   # second launch: the tile schedule reads the first one's costs
for v in (2, 3):
    process_ex(d.surface, d.last, W, H, d.pitch, 0, d.gpu, d.accel, spp=1, bounces=4, variant=v, ray_counter=d.counter, work_counters=work)
ao = torch.zeros(W * H, dtype=torch.float32, device="cuda")
rt.check(rt.lib().rt_ambient_occlusion(d.accel, d.gpu, W, H, 4, rt.ptr(ao), rt.ptr(d.counter), None))
out = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda")
rt.check(rt.lib().rt_tonemap_rgba8(rt.ptr(d.surface), W, H, W * 16, 1.0, rt.ptr(out), None))
a = util.assets()
s = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda"); l = torch.zeros_like(s)
tr = rt.Tracer(s, l, W, H, W * 16, bunny_obj=a["obj"], env_dds=a["dds"], seed=5)
host = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
tr.render_to_host(2, host, W * 16)
torch.cuda.synchronize()
print("sanitize case ok", d.rays(), float(host.mean()))
