"""How many lanes of a warp still have a ray per traversal iteration (and how many stack entries they hold), for the bench
workload. Needs the measurement build:  python cuda-raytracing_b200/build.py -DRT_TW_MODE=8 --out=libb200rt_hist.so
    B200RT_LIB=$PWD/cuda-raytracing_b200/lib/libb200rt_hist.so python benchmarks/lane_histogram.py"""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import util  # noqa: E402


def main():
    import torch
    import gpu_util
    from _loader import load_package
    rt = load_package()
    W, H = 1920, 1080
    d = gpu_util.DeviceScene(rt, torch, W, H, util.POSES["poseB"], seed=1234, env=True)
    lib = C.CDLL(rt.LIB_PATH)
    hist = (C.c_ulonglong * 66)()
    d.process(0, spp=2, bounces=4)   # warm-up
    torch.cuda.synchronize()
    assert lib.rt_debug_tw_hist(hist, 1) == 0
    d.process(1, spp=16, bounces=4)
    torch.cuda.synchronize()
    assert lib.rt_debug_tw_hist(hist, 1) == 0
    it = [hist[n] for n in range(33)]
    pend = [hist[33 + n] for n in range(33)]
    total = sum(it)
    cum = 0
    rows = []
    for n in range(1, 33):
        cum += it[n]
        rows.append({"alive": n, "iterations_frac": it[n] / total, "cumulative": cum / total,
                     "pending_entries_per_alive_lane": (pend[n] / (it[n] * n)) if it[n] else 0.0})
    print(json.dumps({"workload": "bench (1080p pose B, 4 bounces, 16 spp)", "warp_iterations": total,
                      "mean_alive": sum(n * it[n] for n in range(33)) / total, "rows": rows}))


if __name__ == "__main__":
    main()
