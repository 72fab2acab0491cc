"""Experiment: cost-binned pixel -> lane assignment. A counting launch records the traversal steps every pixel cost; the next
launches give each warp 32 pixels of similar cost (several groupings), on top of the cost-ordered tile schedule. The image must
stay bit-identical (pixels are independent). One JSON line per grouping.

    python -c "from _loader import load_build; load_build().build(defines=['RT_PT_BINNING=1'], out='libb200rt_binning.so')"
    B200RT_LIB=$PWD/cuda-raytracing_b200/lib/libb200rt_binning.so python benchmarks/pixel_binning.py > gpurun_out/pixel_binning.jsonl
"""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

import gpu_util  # noqa: E402
from _loader import load_package  # noqa: E402

rt = load_package()
W, H, SPP, BLOCK = 1920, 1080, 64, 128
d = gpu_util.DeviceScene(rt, torch, W, H, "poseB", seed=1234, env=True)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
dbg = rt.lib().rt_debug_pixel_binning
dbg.restype = None


def set_debug(perm, cost):
    dbg(C.c_void_p(perm.data_ptr() if perm is not None else 0), C.c_void_p(cost.data_ptr() if cost is not None else 0))


def run(reps=4):
    """fresh accumulation + fresh schedule history; returns (median ms of frames 1.., surface after frame 0)"""
    d.reseed(1234); d.workspace.zero_()
    d.process(0, spp=SPP, bounces=4, write_last=1); torch.cuda.synchronize()
    first = d.surface.clone()
    ms = []
    for k in range(reps):
        flush.fill_(k)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); d.process(k + 1, spp=SPP, bounces=4, write_last=1); b.record(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return sorted(ms)[len(ms) // 2], first


set_debug(None, None)
base_ms, ref = run()
print(json.dumps({"grouping": "raster tiles 16x8 (product)", "ms": base_ms}), flush=True)

cost = torch.zeros(W * H, dtype=torch.int32, device="cuda")
set_debug(None, cost)
count_ms, first = run(reps=2)
assert torch.equal(first, ref)
set_debug(None, None)
c = cost.clone().to(torch.int64)
print(json.dumps({"grouping": "counting launch", "ms": count_ms, "steps_mean": float(c.float().mean()), "steps_max": int(c.max()),
                  "steps_p50": float(c.float().median()), "steps_p99": float(torch.quantile(c.float()[::7], 0.99))}), flush=True)

pix = torch.arange(W * H, device="cuda", dtype=torch.int64)
px, py = pix % W, pix // W
cmax = int(c.max()) + 1


def padded(order):
    n = (len(order) + BLOCK - 1) // BLOCK * BLOCK
    out = torch.full((n,), 0xFFFFFFFF, dtype=torch.int64, device="cuda")
    out[:len(order)] = order
    return out.to(torch.uint32) if hasattr(torch, "uint32") else out.to(torch.int32)


def region_key(rw, rh, levels=None):
    region = (py // rh) * ((W + rw - 1) // rw) + (px // rw)
    cc = c if levels is None else (c * levels // cmax)
    top = cmax if levels is None else levels
    return region * top + (top - 1 - cc)


groupings = {
    "global sort by cost": torch.argsort(-c, stable=True),
    "global, 16 cost levels (raster order inside a level)": torch.argsort(region_key(W, H, 16), stable=True),
    "global, 64 cost levels": torch.argsort(region_key(W, H, 64), stable=True),
    "128x64 regions, sort by cost": torch.argsort(region_key(128, 64), stable=True),
    "64x32 regions, sort by cost": torch.argsort(region_key(64, 32), stable=True),
    "32x16 regions, sort by cost": torch.argsort(region_key(32, 16), stable=True),
    "16x8 regions (= product tiles), sort by cost": torch.argsort(region_key(16, 8), stable=True),
    "64x32 regions, 8 cost levels": torch.argsort(region_key(64, 32, 8), stable=True),
}
for name, order in groupings.items():
    perm = padded(order)
    set_debug(perm, None)
    ms, first = run()
    set_debug(None, None)
    print(json.dumps({"grouping": name, "ms": ms, "bit_identical": bool(torch.equal(first, ref))}), flush=True)
