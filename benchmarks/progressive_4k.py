"""BASELINE config 5: progressive 3840x2160 render, 1024 spp in total, spp split across the ranks of one node, ONE NCCL reduce
of the fp32 accumulation buffer at the end (strong scaling: total work fixed).

    python benchmarks/progressive_4k.py [--spp 1024]                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        benchmarks/progressive_4k.py [--spp 1024]

Prints one JSON line (rank 0) and saves a 4x4-block-averaged copy of the final mean image to
gpurun_out/progressive_4k_<N>gpu.npy so runs at different N can be compared (RMSE) afterwards with --compare."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--spp-per-launch", type=int, default=64)
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--bounces", type=int, default=4)
    ap.add_argument("--compare", nargs="*", help="block-averaged .npy images to compare against the first one")
    a = ap.parse_args()
    if a.compare:
        ref = np.load(a.compare[0])
        for other in a.compare[1:]:
            img = np.load(other)
            rmse = float(np.sqrt(np.mean((img - ref) ** 2)) / np.sqrt(np.mean(ref ** 2)))
            print(json.dumps({"reference": a.compare[0], "image": other, "relative_rmse_block4x4": rmse}))
        return 0

    import torch
    import torch.distributed as dist
    import gpu_util
    from _loader import load_package
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))
    rt = load_package()
    from cuda_raytracing_b200 import multigpu
    from cuda_raytracing_b200.api import process_ex
    W, H = a.width, a.height
    d = gpu_util.DeviceScene(rt, torch, W, H, "poseB", seed=1234, env=True, device=dev)
    n_states = gpu_util.rng_state_count(W, H)
    t0 = time.perf_counter()
    d.reseed(1234, multigpu.rank_first_subsequence(rank, n_states))
    torch.cuda.synchronize()
    rng_init_s = time.perf_counter() - t0
    my_spp = multigpu.partition_spp(a.spp, world)[rank]
    acc = torch.zeros((H, W, 4), dtype=torch.float32, device=dev)
    # warm-up launch on a scratch copy of the states (excluded)
    keep = d.rng.clone()
    process_ex(acc, d.last, W, H, d.pitch, 0, d.gpu, d.accel, spp=1, bounces=a.bounces, blend_mode=1, ray_counter=d.counter)
    torch.cuda.synchronize()
    d.rng.copy_(keep); acc.zero_(); d.counter.zero_()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    for n in multigpu.launches_for(my_spp, a.spp_per_launch):
        process_ex(acc, d.last, W, H, d.pitch, 0, d.gpu, d.accel, spp=n, bounces=a.bounces, blend_mode=1, ray_counter=d.counter)
    e1.record()
    multigpu.reduce_frame(acc, dist if world > 1 else None)
    e2.record()
    torch.cuda.synchronize()
    render_ms, total_ms = e0.elapsed_time(e1), e0.elapsed_time(e2)
    t = torch.tensor([render_ms, total_ms, float(d.rays())], dtype=torch.float64, device=dev)
    if world > 1:
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        render_ms, total_ms, rays = mx[0].item(), mx[1].item(), sm[2].item()
    else:
        rays = t[2].item()
    if rank == 0:
        mean = multigpu.finalize_mean(acc)
        assert bool((acc[..., 3] == a.spp).all()), "sample count in the reduced buffer"
        blk = mean[..., :3].reshape(H // 4, 4, W // 4, 4, 3).mean(dim=(1, 3)).cpu().numpy()
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        np.save(os.path.join(ROOT, "gpurun_out", f"progressive_4k_{world}gpu.npy"), blk)
        print(json.dumps({"config": f"progressive {W}x{H}, {a.spp} spp total, {a.bounces} bounces, pose B", "n_gpus": world,
                          "spp_per_rank": multigpu.partition_spp(a.spp, world), "render_ms_max_over_ranks": render_ms,
                          "render_plus_reduce_ms": total_ms, "reduce_ms": total_ms - render_ms, "rays": rays,
                          "mrays_s": rays / total_ms / 1e3, "rng_init_s_rank0": rng_init_s, "reduce_bytes": W * H * 16}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
