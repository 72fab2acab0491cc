"""How much does regrouping incoherent rays buy? Traces the same diffuse-lobe ray set (benchmarks/incoherent.py) in different
orders through rt_trace_rays variant 0: as generated (pixel order), regrouped by direction octant inside blocks of 256 / 1024 /
8192 rays (what a CTA- or wavefront-local regroup can do), and globally sorted by (octant, origin Morton code)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "benchmarks")):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402

import gpu_util  # noqa: E402
from _loader import load_package  # noqa: E402
from incoherent import make_rays  # noqa: E402


def octant(r):
    d = r[:, 3:]
    return ((d[:, 0] < 0).long() | ((d[:, 1] < 0).long() << 1) | ((d[:, 2] < 0).long() << 2))


def morton(r, bits=7):
    p = r[:, :3]
    lo, hi = p.min(0).values, p.max(0).values
    q = ((p - lo) / (hi - lo + 1e-9) * ((1 << bits) - 1)).long()
    code = torch.zeros(len(r), dtype=torch.long, device=r.device)
    for b in range(bits):
        for k in range(3):
            code |= ((q[:, k] >> b) & 1) << (3 * b + k)
    return code


def main():
    rt = load_package()
    W, H = 1920, 1080
    d = gpu_util.DeviceScene(rt, torch, W, H, "poseB", seed=1, env=False)
    rays = make_rays(rt, d, W, H, 8)
    n = rays.shape[0]
    ids = torch.zeros(n, dtype=torch.int32, device="cuda"); ts = torch.zeros(n, dtype=torch.float32, device="cuda")

    def run(r, label):
        r = r.contiguous()
        f = lambda: rt.check(rt.lib().rt_trace_rays(d.accel, d.gpu, rt.ptr(r), n, 1, 0, rt.ptr(ids), rt.ptr(ts), None, None))
        f(); torch.cuda.synchronize()
        ms = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); f(); b.record(); torch.cuda.synchronize(); ms.append(a.elapsed_time(b))
        m = sorted(ms)[2]
        print(json.dumps({"order": label, "rays": n, "ms": m, "mrays_s": n / m / 1e3}), flush=True)

    run(rays, "as generated (pixel order, 8 rays per hit point consecutive)")
    perm = torch.randperm(n, device="cuda")
    run(rays[perm], "random permutation")
    oc = octant(rays)
    for blk in (256, 1024, 8192, 65536):
        key = (torch.arange(n, device="cuda") // blk) * 8 + oc
        run(rays[torch.argsort(key, stable=True)], f"octant-regrouped inside blocks of {blk}")
    run(rays[torch.argsort(oc, stable=True)], "global sort by octant (origins stay in pixel order)")
    run(rays[torch.argsort(oc * (1 << 21) + morton(rays), stable=True)], "global sort by (octant, origin Morton)")
    key = morton(rays) * 8 + oc
    run(rays[torch.argsort(key, stable=True)], "global sort by (origin Morton, octant)")


if __name__ == "__main__":
    main()
