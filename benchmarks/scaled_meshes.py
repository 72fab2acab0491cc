"""BASELINE config 4: synthetic scaled meshes through the reference's own AddLoadedScene semantics (every face twice: smooth +
flat duplicate) and BVH builder mirror:
  (a) 2x Loop-subdivided bunny: 1 111 216 faces -> 2 222 432 BVH triangles (+ Cornell box, floor, spheres = the reference scene)
  (b) 10x10 grid of bunnies, pitch 30 units in x/z: 6 945 100 faces -> 13 890 200 BVH triangles (working set > L2)
For each: primary-ray parity property at 1080p (product traversal == reference-order traversal, ids and t bit for bit),
path tracing 1080p / 4 bounces, Mrays/s, and the bytes-per-ray roofline from reference-order work counters.
    python benchmarks/scaled_meshes.py [a|b|ab]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import util  # noqa: E402

GRID_POSE = util.GRID_POSE
bunny_transform = util.bunny_transform


def build_scene(rt, which, env=True):
    """The config-4 scenes live in tests/util.py (the -m gpu parity tests build the same ones)."""
    return util.scaled_scene(rt, which, env)


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "ab"
    import torch
    import gpu_util
    from _loader import load_package
    rt = load_package()
    from cuda_raytracing_b200.api import process_ex
    W, H = 1920, 1080
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    for w in which:
        scene, mesh, pose, times = build_scene(rt, w)
        t0 = time.time()
        d = gpu_util.DeviceScene(rt, torch, W, H, pose, seed=1234, env=True, scene=scene)
        torch.cuda.synchronize()
        times["upload_and_relayout_s"] = time.time() - t0
        c = scene.counts()
        info = {"config": "4a: 2x Loop-subdivided bunny in the reference scene" if w == "a" else "4b: 10x10 bunny grid",
                "triangles": c["faces"], "bvh_nodes": c["nodes"], **times}
        i0, t0_ = d.trace_primary(1, 0)
        i1, t1_ = d.trace_primary(1, 1)
        info["primary_hit_fraction"] = float((i0 != 0xFFFFFFFF).mean())
        info["primary_ids_equal_reference_order"] = bool(np.array_equal(i0, i1))
        info["primary_t_bit_equal_reference_order"] = bool(np.array_equal(t0_.view(np.uint32), t1_.view(np.uint32)))
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        ids = torch.zeros(W * H, dtype=torch.int32, device="cuda"); ts = torch.zeros(W * H, dtype=torch.float32, device="cuda")

        def timed(fn, reps=5):
            fn(); torch.cuda.synchronize()
            ms = []
            for _ in range(reps):
                flush.fill_(1)
                a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a_.record(); fn(); b_.record(); torch.cuda.synchronize(); ms.append(a_.elapsed_time(b_))
            return sorted(ms)[len(ms) // 2]
        ms = timed(lambda: rt.check(rt.lib().rt_trace_primary(d.accel, d.gpu, W, H, 1, 0, rt.ptr(ids), rt.ptr(ts), None, None)))
        info["primary_ms"] = ms; info["primary_mrays_s"] = W * H / ms / 1e3
        # work counters (reference order) on the PT workload -> bytes per ray
        # (quarter-resolution frame: the per-ray averages are what matters, and the reference order can be very slow)
        work = torch.zeros(2, dtype=torch.int64, device="cuda")
        d.counter.zero_()
        qs = torch.zeros((H // 4, W // 4, 4), dtype=torch.float32, device="cuda")
        scene.set_camera(pose[0], pose[1], pose[2], W // 4, H // 4); scene.upload(d.rng)
        process_ex(qs, qs, W // 4, H // 4, (W // 4) * 16, 0, d.gpu, d.accel, spp=1, bounces=4, variant=2, ray_counter=d.counter, work_counters=work)
        n = d.rays()
        nodes, tris = work[0].item() / n, work[1].item() / n
        d.set_pose(pose); d.reseed(1234)
        bytes_per_ray = 32 * nodes + 48 * tris + 64
        spp = 16
        d.counter.zero_()
        d.process(0, spp=spp, bounces=4, write_last=1); rays = d.rays()
        ms = timed(lambda: d.process(0, spp=spp, bounces=4, write_last=1), reps=3)
        achieved = (rays * bytes_per_ray + W * H * 80) / (ms * 1e-3) / 1e9
        info.update({"pt_spp_per_launch": spp, "pt_ms": ms, "pt_rays": rays, "pt_mrays_s": rays / ms / 1e3,
                     "nodes_per_ray_reference_order": nodes, "tris_per_ray_reference_order": tris, "bytes_per_ray": bytes_per_ray,
                     "roofline_achieved_gbs": achieved, "roofline_frac_of_measured_hbm": achieved / peak})
        import ctypes as C
        inf = (C.c_uint64 * 8)()
        rt.check(rt.lib().rt_accel_info(d.accel, inf))
        info["accel_device_bytes"] = int(inf[5] | (inf[6] << 32)); info["wide_nodes"] = int(inf[0]); info["fat_leaves"] = int(inf[3])
        print(json.dumps(info), flush=True)
        del d, scene, mesh
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
