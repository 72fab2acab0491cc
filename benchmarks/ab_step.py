"""A/B of one build of libb200rt (B200RT_LIB=<path>) on the bench step, in ONE process: median / min of N timed 1080p 64-spp
frames (pose B, 4 bounces; the bench.py protocol: L2 flush between steps, CUDA events) + the frame/RNG hash that says whether the
build is bit-identical + optionally the 16-spp frame of config 4b (10x10 bunny grid). One JSON line.
    B200RT_LIB=$PWD/cuda-raytracing_b200/lib/libb200rt_x.so python benchmarks/ab_step.py [--steps 10] [--grid] [--variant 0]"""
import argparse
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "benchmarks")):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402

import gpu_util  # noqa: E402
from _loader import load_package  # noqa: E402


def timed(fn, reps, flush):
    ms = []
    for k in range(reps):
        flush.fill_(k & 0xFF)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--grid", action="store_true")
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--no-hash", action="store_true")
    ap.add_argument("--counters", action="store_true", help="wide-node steps and triangle tests per ray of the shipped kernel (variant 3, 2 spp)")
    a = ap.parse_args()
    rt = load_package()
    out = {"lib": os.path.basename(os.environ.get("B200RT_LIB", "libb200rt.so")), "variant": a.variant}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    if not a.no_hash:
        d = gpu_util.DeviceScene(rt, torch, 640, 360, "poseB", seed=1234, env=True)
        for k in range(2):
            d.process(k, spp=16, bounces=4, write_last=1, variant=a.variant)
        torch.cuda.synchronize()
        h = hashlib.sha1()
        h.update(d.surface.cpu().numpy().tobytes()); h.update(d.rng.cpu().numpy().tobytes())
        out["hash"] = h.hexdigest()[:16]; out["hash_rays"] = d.rays()
        del d
    W, H = 1920, 1080
    d = gpu_util.DeviceScene(rt, torch, W, H, "poseB", seed=1234, env=True)
    f = lambda: d.process(0, spp=64, bounces=4, write_last=1, variant=a.variant)
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    d.counter.zero_()
    ms = timed(f, a.steps, flush)
    out.update({"ms_median": sorted(ms)[len(ms) // 2], "ms_min": min(ms), "ms_max": max(ms), "mrays_s": d.rays() / sum(ms) / 1e3})
    if a.counters:
        from cuda_raytracing_b200.api import process_ex
        work = torch.zeros(2, dtype=torch.int64, device="cuda")
        d.counter.zero_()
        process_ex(d.surface, d.last, W, H, d.pitch, 0, d.gpu, d.accel, spp=2, bounces=4, blend_mode=0, variant=3, ray_counter=d.counter,
                   stream=torch.cuda.current_stream().cuda_stream, work_counters=work)
        torch.cuda.synchronize()
        rays = int(d.counter.item())
        out.update({"node_steps_per_ray": work[0].item() / rays, "tri_tests_per_ray": work[1].item() / rays})
    if a.grid:
        import scaled_meshes
        del d
        scene, mesh, pose, _ = scaled_meshes.build_scene(rt, "b")
        g = gpu_util.DeviceScene(rt, torch, W, H, pose, seed=1234, env=True, scene=scene)
        fg = lambda: g.process(0, spp=16, bounces=4, write_last=1, variant=a.variant)
        for _ in range(2):
            fg()
        torch.cuda.synchronize()
        g.counter.zero_()
        msg = timed(fg, 5, flush)
        out.update({"grid_ms_median": sorted(msg)[2], "grid_ms_min": min(msg), "grid_mrays_s": g.rays() / sum(msg) / 1e3})
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
