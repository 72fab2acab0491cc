"""bench.py - BASELINE.json's metric (Mrays/s, ms/frame @ 64 spp) on its config[1]:
stanford-bunny scene, 1920x1080 path tracing, 4 bounces, sunset env map, 64 spp.

    python bench.py [--gpus N --steps K --warmup W] [--impl reference]

One "step" = one 1080p frame of 64 samples per pixel (camera rays -> BVH traversal -> shading -> accumulation).
  value      whole-job Mrays/s (closest-hit queries counted on device / CUDA-event time, max over ranks), inputs resident
  e2e        same metric through the headless CUDARayTracer facade call (RenderToHost): per step the mutable scene
             state goes host->device and the float4 frame comes back into pinned HOST memory, wall clock around the call
  roofline   path-trace kernel against the bytes-per-ray memory roofline of SURVEY.md §8(d)
  cpu_baseline   the reference's own kernel body compiled host-side (oracle/_ref), all host threads, bounded sample
N > 1: every rank renders its own 64 spp of the same frame with disjoint XORWOW subsequences into a sum buffer, then
one NCCL reduce (fp32 sum) to rank 0 per frame - weak scaling in samples (BASELINE config 5's partitioning).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import util  # noqa: E402

METRIC = "Mrays/s (bunny 1080p, 4-bounce PT, 64 spp/frame)"
UNIT = "Mrays/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--spp-per-launch", type=int, default=64)
    ap.add_argument("--bounces", type=int, default=4)
    ap.add_argument("--pose", default="poseB", choices=list(util.POSES))
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def workload_name(a):
    return (f"stanford-bunny scene (138916 tris, 9 spheres), {a.width}x{a.height}, {a.bounces} bounces, sunset env map, "
            f"{a.spp} spp/frame, camera {a.pose}")


# ------------------------------------------------------------------------------------------------
# CPU reference (oracle/_ref): the reference's own kernel body, host-compiled, OpenMP over rows
# ------------------------------------------------------------------------------------------------
class CpuReference:
    def __init__(self, a):
        import ref_oracle as ro
        if not ro.available():
            raise RuntimeError("oracle/_ref/libref_oracle.so missing (built by __graft_entry__.build() where /root/reference exists)")
        self.ro, self.a = ro, a
        try:  # all host cores this process may use (torchrun sets OMP_NUM_THREADS=1 for its ranks)
            ro.lib().ref_set_num_threads(len(os.sched_getaffinity(0)))
        except (AttributeError, OSError):
            pass
        # the reference's own scene assembly (GLM restatement of RayTracing.cpp:33-203 in oracle/ref_shim) and its verbatim BVH.cpp on
        # the committed ingest fixture: this arm never loads libb200rt.so
        m = np.load(os.path.join(util.GOLDEN, "bunny_smooth_mesh.npz"))
        self.sc = ro.build_scene(np.ascontiguousarray(m["positions"]), np.ascontiguousarray(m["normals"]), np.ascontiguousarray(m["indices"]))
        self.nodes, self.fi = ro.bvh_build(self.sc["vertices"], self.sc["faces"])
        p, ax, ay = util.POSES[a.pose]
        self.cam = ro.camera(p, ax, ay, a.width, a.height)
        self.cm = util.cubemap_texels()
        n = ((a.width * a.height + 127) // 128) * 128
        self.states = ro.rng_init(n, a.seed)
        self.surf = np.zeros((a.height, a.width, 4), np.float32)
        self.last = np.zeros_like(self.surf)
        self.row_begin = 0

    def run(self, row_stride, spp):
        a = self.a
        s, rays, threads = self.ro.bench_render(self.sc, self.nodes, self.fi, self.cam, self.cm, self.states, self.surf, self.last,
                                                a.width, a.height, 0, self.row_begin % row_stride, row_stride, spp, a.bounces)
        self.row_begin += 1
        return s, rays, threads

    def calibrate(self, target_seconds):
        """Pick (row_stride, spp) so one sample takes about target_seconds."""
        s, rays, threads = self.run(32, 1)
        rate = rays / max(s, 1e-6)
        rays_per_row_spp = rays / (self.a.height / 32.0)
        want = rate * target_seconds
        for stride in (8, 4, 2, 1):  # densest row sampling whose sample count still fits the workload's spp
            spp = int(round(want / (rays_per_row_spp * self.a.height / stride)))
            if spp <= self.a.spp:
                break
        return stride, int(max(1, min(self.a.spp, spp)))

    def sample_desc(self, stride, spp):
        a = self.a
        return (f"every {stride}{ {1: 'st', 2: 'nd', 3: 'rd'}.get(stride, 'th')} row of the {a.width}x{a.height} frame ({(a.height + stride - 1) // stride} rows), {spp} spp, "
                f"{a.bounces} bounces, reference raytracing_kernel_main body host-compiled (-O3, OpenMP)")


def run_reference_arm(a):
    """--impl reference: the reference's CPU implementation of the path on the host cores, same metric/config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cpu = CpuReference(a)
    stride, spp = cpu.calibrate(3.0)
    for _ in range(a.warmup):
        cpu.run(stride, spp)
    total_s, total_rays, threads = 0.0, 0, 1
    for _ in range(a.steps):
        s, rays, threads = cpu.run(stride, spp)
        total_s += s; total_rays += rays
    value = total_rays / total_s / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": 1e3 * total_s / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": workload_name(a), "step": "bounded sample: " + cpu.sample_desc(stride, spp)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "reference", "sample": cpu.sample_desc(stride, spp)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print_json(line)
    return 0


# ------------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md "clocks line")
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.path = os.path.join(tempfile.gettempdir(), f"b200rt_clocks_{os.getpid()}.csv")
        self.proc = None
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(index)],
                                         stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        for line in open(self.path):
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.remove(self.path)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# own arm
# ------------------------------------------------------------------------------------------------
def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def run_b200_arm(a):
    import torch
    import torch.distributed as dist
    from _loader import load_package
    import gpu_util
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: there is no CPU fallback")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))
    rt = load_package()
    W, H = a.width, a.height
    launches_per_step = (a.spp + a.spp_per_launch - 1) // a.spp_per_launch
    spp_launch = [min(a.spp_per_launch, a.spp - i * a.spp_per_launch) for i in range(launches_per_step)]

    d = gpu_util.DeviceScene(rt, torch, W, H, a.pose, seed=a.seed, env=True, device=dev)
    n_states = gpu_util.rng_state_count(W, H)
    from cuda_raytracing_b200 import multigpu
    if world > 1:  # rank r owns cuRAND subsequences [r*n, (r+1)*n): disjoint streams, rank 0 == the 1-GPU stream
        d.reseed(a.seed, multigpu.rank_first_subsequence(rank, n_states))
    stream = torch.cuda.current_stream().cuda_stream
    from cuda_raytracing_b200.api import process_ex

    def frame(blend, surface, variant=a.variant, work=None):
        for i, s in enumerate(spp_launch):
            process_ex(surface, d.last, W, H, d.pitch, i if blend == 0 else 0, d.gpu, d.accel, spp=s, bounces=a.bounces,
                       blend_mode=blend, variant=variant, write_last=1 if blend == 0 else 0, ray_counter=d.counter,
                       stream=stream, work_counters=work, workspace=d.workspace)

    # ---- algorithmic bytes per ray: reference-order work counters on this exact workload (2 spp is plenty) ----
    work = torch.zeros(2, dtype=torch.int64, device=dev)
    rng_keep = d.rng.clone()
    d.counter.zero_()
    process_ex(d.surface, d.last, W, H, d.pitch, 0, d.gpu, d.accel, spp=2, bounces=a.bounces, blend_mode=0, variant=2,
               ray_counter=d.counter, stream=stream, work_counters=work)
    torch.cuda.synchronize()
    cnt_rays = int(d.counter.item()); nodes_per_ray = work[0].item() / cnt_rays; tris_per_ray = work[1].item() / cnt_rays
    bytes_per_ray = 32.0 * nodes_per_ray + 48.0 * tris_per_ray + 64.0
    d.rng.copy_(rng_keep)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    blend = 0 if world == 1 else 1
    sum_buf = d.surface

    def step():
        if blend == 1:
            sum_buf.zero_()
        frame(blend, sum_buf)
        multigpu.reduce_frame(sum_buf, dist if world > 1 else None)

    for _ in range(max(a.warmup, 0)):
        flush.fill_(1); step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    d.counter.zero_()
    uuid = str(torch.cuda.get_device_properties(local).uuid)
    sampler = ClockSampler(uuid if uuid.startswith("GPU-") else "GPU-" + uuid) if rank == 0 else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    wall0 = time.perf_counter()
    for k in range(a.steps):
        flush.fill_(k & 0xFF)  # L2 flush between timed steps (outside the event pairs)
        ev[k][0].record()
        if blend == 1:
            sum_buf.zero_()
        frame(blend, sum_buf)
        ev[k][1].record()      # kernel-only span (roofline)
        multigpu.reduce_frame(sum_buf, dist if world > 1 else None)
        ev[k][2].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop() if sampler else None
    step_ms = sum(e[0].elapsed_time(e[2]) for e in ev)
    kern_ms = sum(e[0].elapsed_time(e[1]) for e in ev)
    rays = int(d.counter.item())
    t = torch.tensor([step_ms, kern_ms, float(rays)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        step_ms, kern_ms, rays_all = tmax[0].item(), tmax[1].item(), tsum[2].item()
    else:
        rays_all = float(rays)
    value = rays_all / (step_ms * 1e-3) / 1e6

    # ---- e2e: the headless facade call with HOST buffers ----
    e2e = None
    if not a.no_e2e:
        e2e = run_e2e(a, rt, torch, dev, world, rank)

    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peaks()
    traffic = None  # dram bytes per launch from the committed ncu capture of this exact launch shape (never measured under bench)
    issue = None    # what actually bounds the kernel: warp instructions per launch (same capture) against the SMs' issue rate
    try:
        if a.variant == 0 and (a.width, a.height, a.spp_per_launch, a.bounces, a.pose) == (1920, 1080, 64, 4, "poseB"):
            cap = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["pathtrace_megakernel"]
            traffic = cap["dram_bytes_per_launch"]
            props = torch.cuda.get_device_properties(dev)
            mhz = (clocks or {}).get("sm_mhz") or 1965.0
            peak_ginst = props.multi_processor_count * 4 * mhz * 1e-3  # 4 schedulers per SM, 1 warp instruction per clock each
            ach_ginst = cap["warp_instructions_per_launch"] / ((kern_ms / launches_per_step / a.steps) * 1e-3) / 1e9
            issue = {"achieved": ach_ginst, "peak": peak_ginst, "unit": "G warp-inst/s", "frac": ach_ginst / peak_ginst,
                     "threads_per_instruction": cap["threads_per_instruction"],
                     "source": "instruction count from the committed ncu capture (profiles/traffic.json), time and SM clock live"}
    except Exception:
        traffic = None
    launches = a.steps * launches_per_step
    px = W * H
    alg_bytes_per_launch = (rays / launches) * bytes_per_ray + px * 80.0
    achieved = alg_bytes_per_launch / ((kern_ms / launches) * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": step_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(a), "spp_per_launch": a.spp_per_launch, "kernel_variant": a.variant,
                   "l2": "flushed between timed steps (256 MiB fill, outside the event pairs); scene working set itself is L2-resident by nature",
                   "parallelism": "1 GPU" if world == 1 else f"spp split: {world} ranks x {a.spp} spp, disjoint XORWOW subsequences, 1 NCCL reduce(sum fp32) of the frame per step",
                   "rays_per_step": rays_all / a.steps, "ms_per_frame_64spp": step_ms / a.steps, "wall_s": wall},
        "clocks": clocks, "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "kernel": "pathtrace_megakernel", "peak_source": peak_src, "issue": issue,
                     "bytes_per_ray": bytes_per_ray, "nodes_per_ray_reference_order": nodes_per_ray, "tris_per_ray_reference_order": tris_per_ray,
                     "note": "algorithmic bytes = rays x (32 B x nodes + 48 B x tris under the REFERENCE traversal order + 64) + pixels x 80; "
                             "the bunny scene is L2-resident, so frac > 1 means cache/culling amplification, not HBM traffic (SURVEY §8d); `traffic` (ncu, same protocol) includes the write-back of the L2-flush buffer's dirty lines evicted during the launch (~230 MB) on top of surfaces + RNG states + first-touch scene"},
    }
    if e2e:
        line["e2e"] = e2e
    if world == 1 and not a.no_cpu_baseline:
        try:
            cpu = CpuReference(a)
            stride, spp = cpu.calibrate(12.0)  # ~10-30 s of CPU work (all host threads)
            s, r, threads = cpu.run(stride, spp)
            line["cpu_baseline"] = {"value": r / s / 1e6, "unit": UNIT, "cores": threads, "kind": "reference", "sample": cpu.sample_desc(stride, spp),
                                    "seconds": s}
        except Exception as e:  # the checker is optional on a box without the prebuilt oracle
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {e}"}
    print_json(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_e2e(a, rt, torch, dev, world, rank):
    """The call a headless user makes: CUDARayTracer::RenderToHost (rt_host_tracer_render_to_host). Per step the mutable
    scene state (spheres, materials; the camera rides in the by-value GPUScene) goes host->device, `spp/spp_per_launch`
    Process() frames run, and the float4 surface is copied into pinned HOST memory. Wall clock around the synchronous call."""
    import torch.distributed as dist
    W, H = a.width, a.height
    assets = util.assets()
    surface = torch.zeros((H, W, 4), dtype=torch.float32, device=dev)
    last = torch.zeros_like(surface)
    host = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
    tr = rt.Tracer(surface, last, W, H, W * 16, bunny_obj=assets["obj"], env_dds=assets["dds"], samples_per_launch=a.spp_per_launch,
                   max_bounces=a.bounces, seed=a.seed + 17 * rank + 1, kernel_variant=a.variant)
    p, ax, ay = util.POSES[a.pose]
    tr.scene.set_camera(p, ax, ay, W, H)
    frames = max(1, a.spp // a.spp_per_launch)
    tr.render_to_host(frames, host, W * 16)  # first call uploads the scene + builds the re-layout (cold path)
    for _ in range(max(a.warmup - 1, 0)):
        tr.render_to_host(frames, host, W * 16)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    r0 = tr.ray_count
    h2d = 0
    t0 = time.perf_counter()
    for _ in range(a.steps):
        h2d += tr.render_to_host(frames, host, W * 16)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    rays = tr.ray_count - r0
    assert float(host[..., 3].min()) == 1.0 and bool(torch.isfinite(host).all())
    t = torch.tensor([dt, float(rays)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dt, rays = tmax[0].item(), tsum[1].item()
    return {"value": rays / dt / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d // a.steps, "d2h_bytes_per_step": W * H * 16,
            "ms_per_step": 1e3 * dt / a.steps,
            "call": "rt_host_tracer_render_to_host (CUDARayTracer::RenderToHost): pinned host frame out, scene state in"}


def main():
    a = parse_args()
    # The contract is ONE JSON line on stdout. Libraries (NCCL prints its version banner, torchrun its OMP notice) write to
    # fd 1 too, so everything but the final line is routed to stderr: fd 1 is pointed at stderr for the run and the JSON line
    # is written to the saved real stdout.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    global print_json
    def print_json(line):
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if a.impl == "reference":
        return run_reference_arm(a)
    return run_b200_arm(a)


if __name__ == "__main__":
    sys.exit(main())
