"""Secondary measurements for BASELINE.json's other configs (bench.py stays the one contract line):
primary-ray closest-hit throughput, AO stress (config 3), path tracing at both poses / variants, the reference's own CUDA
kernel (sm_100a build, oracle/_ref/libref_gpu.so) on the same GPUScene. One JSON line per measurement.

    python benchmarks/configs.py [--quick] > gpurun_out/configs.jsonl
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

import gpu_util  # noqa: E402
import util  # noqa: E402
from _loader import load_package  # noqa: E402


def timed(fn, reps, flush):
    fn(); torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return sorted(ms)[len(ms) // 2]


def main():
    quick = "--quick" in sys.argv
    rt = load_package()
    W, H = 1920, 1080
    reps = 3 if quick else 7
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    d = gpu_util.DeviceScene(rt, torch, W, H, "poseB", seed=1234, env=True)
    out = []

    def emit(**kw):
        print(json.dumps(kw), flush=True)
        out.append(kw)

    ids = torch.zeros(W * H, dtype=torch.int32, device="cuda"); ts = torch.zeros(W * H, dtype=torch.float32, device="cuda")
    for pose in ("poseB", "default"):
        d.set_pose(pose)
        for variant in (0, 1):
            f = lambda: rt.check(rt.lib().rt_trace_primary(d.accel, d.gpu, W, H, 1, variant, rt.ptr(ids), rt.ptr(ts), None, None))
            ms = timed(f, reps, flush)
            emit(config="primary closest-hit 1080p (1 ray/pixel, spheres+BVH)", pose=pose,
                 kernel="trace_primary_kernel<%d>" % variant + (" (reference traversal order)" if variant else " (product)"),
                 ms=ms, mrays_s=W * H / ms / 1e3)
    d.set_pose("poseB")
    ao = torch.zeros(W * H, dtype=torch.float32, device="cuda")
    d.counter.zero_()
    f = lambda: rt.check(rt.lib().rt_ambient_occlusion(d.accel, d.gpu, W, H, 16, rt.ptr(ao), rt.ptr(d.counter), None))
    f(); n = d.rays(); d.counter.zero_()
    ms = timed(f, reps, flush)
    emit(config="AO stress 1080p: primary + 16 diffuse-lobe rays per hit (BASELINE config 3)", pose="poseB", kernel="ambient_occlusion_kernel",
         ms=ms, rays=n, mrays_s=n / ms / 1e3)
    for pose in ("poseB", "default"):
        d.set_pose(pose)
        for variant in (0, 1):
            for bounces, spp in ((1, 16), (4, 16)):
                d.counter.zero_()
                g = lambda: d.process(0, spp=spp, bounces=bounces, variant=variant, write_last=1)
                g(); n = d.rays()
                ms = timed(g, reps, flush)
                emit(config=f"path trace 1080p, {spp} spp/launch, max_bounces={bounces}" + (" (= jittered primary rays only)" if bounces == 1 else ""),
                     pose=pose, kernel=f"variant {variant}", ms=ms, rays=n, mrays_s=n / ms / 1e3)
    # the reference's own CUDA kernel, unmodified, sm_100a build: 5 spp x 6 bounces per launch (main_raytracing.cu:115,167)
    import ref_oracle as ro
    if ro.gpu_available():
        g = ro.gpu_lib()
        for pose in ("poseB", "default"):
            d.set_pose(pose)
            d.counter.zero_()
            d.process(0, spp=5, bounces=6, variant=0)  # ray count of the same workload through the product's counter
            n = d.rays()
            mine = timed(lambda: d.process(0, spp=5, bounces=6, variant=0, write_last=1), reps, flush)
            def ref():
                g.raytracing_process(rt.ptr(d.surface), rt.ptr(d.last), W, H, d.pitch, 0, d.gpu)
                d.last.copy_(d.surface)  # RayTracing.cpp:233 - the reference's per-frame D2D copy
            ms = timed(ref, reps, flush)
            emit(config="reference frame: 1080p, 5 spp x 6 bounces (raytracing_process + the D2D copy of RayTracing.cpp:233)", pose=pose,
                 kernel="REFERENCE raytracing_kernel_main (unmodified source, nvcc -O3 sm_100a)", ms=ms, rays_approx=n, mrays_s=n / ms / 1e3,
                 product_ms_same_workload=mine, product_mrays_s=n / mine / 1e3, speedup=ms / mine)


if __name__ == "__main__":
    main()
