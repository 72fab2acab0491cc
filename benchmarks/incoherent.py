"""Traversal-only measurement on an incoherent ray set: diffuse-lobe rays from the primary hit points of the 1080p pose-B
view (what bounce 1 of the path tracer traces). Times rt_trace_rays variant 0 (product) and 1 (reference order) and checks they agree bit for bit.
    python benchmarks/incoherent.py [rays_per_hit]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402

import gpu_util  # noqa: E402
from _loader import load_package  # noqa: E402


def make_rays(rt, d, W, H, per_hit, seed=5):
    ids, ts = d.trace_primary(1, 0)
    cam = np.frombuffer(bytes(d.scene.gpu_scene().camera), np.float32)
    o, hz, vt, llc = cam[0:3], cam[6:9], cam[9:12], cam[12:15]
    ys, xs = np.divmod(np.arange(W * H), W)
    u = ((xs + 0.5) / W).astype(np.float32)[:, None]; v = ((ys + 0.5) / H).astype(np.float32)[:, None]
    dirs = llc + u * hz + v * vt - o
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    hit = ids != 0xFFFFFFFF
    p = (o + dirs[hit] * ts[hit, None]).astype(np.float32)
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    P = torch.from_numpy(p).cuda().repeat_interleave(per_hit, dim=0)
    D = torch.randn(P.shape, generator=g, device="cuda")
    D = D / D.norm(dim=1, keepdim=True)
    P = P + 0.02 * D
    return torch.cat([P, D], 1).contiguous()


def main():
    per_hit = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    rt = load_package()
    W, H = 1920, 1080
    d = gpu_util.DeviceScene(rt, torch, W, H, "poseB", seed=1, env=False)
    rays = make_rays(rt, d, W, H, per_hit)
    n = rays.shape[0]
    ids = [torch.zeros(n, dtype=torch.int32, device="cuda") for _ in range(3)]
    ts = [torch.zeros(n, dtype=torch.float32, device="cuda") for _ in range(3)]
    res = {}
    for variant in (0, 1):
        f = lambda: rt.check(rt.lib().rt_trace_rays(d.accel, d.gpu, rt.ptr(rays), n, 1, variant, rt.ptr(ids[variant]), rt.ptr(ts[variant]), None, None))
        f(); torch.cuda.synchronize()
        ms = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); f(); b.record(); torch.cuda.synchronize(); ms.append(a.elapsed_time(b))
        res[variant] = sorted(ms)[2]
    same = bool(torch.equal(ids[0], ids[1]) and torch.equal(ts[0].view(torch.int32), ts[1].view(torch.int32)))
    print(json.dumps({"rays": n, "ms_product": res[0], "ms_reference_order": res[1], "mrays_product": n / res[0] / 1e3,
                      "mrays_reference_order": n / res[1] / 1e3, "identical": same}))


if __name__ == "__main__":
    main()
