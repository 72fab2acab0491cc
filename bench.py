"""bench.py - BASELINE.json's metric (Mrays/s, ms/frame @ 64 spp) on its config[1]:
stanford-bunny scene, 1920x1080 path tracing, 4 bounces, sunset env map, 64 spp.

    python bench.py [--gpus N --steps K --warmup W] [--impl reference]

One "step" = one 1080p frame of 64 samples per pixel (camera rays -> BVH traversal -> shading -> accumulation).
  value      whole-job Mrays/s (closest-hit queries counted on device / CUDA-event time, max over ranks), inputs resident
  e2e        same metric through the headless CUDARayTracer facade call (RenderToHost): per step the mutable scene
             state goes host->device and the float4 frame comes back into pinned HOST memory, wall clock around the call
  roofline   path-trace kernel against the bound that binds it: warp-instruction issue (scene is L2-resident); the bytes-per-ray
             figures of SURVEY.md §8(d) (reference order and product order), DRAM and L2 traffic ride along as flat keys
  configs    bounded side measurements, same event protocol: primary-only / bounce-only Mrays/s, AO (config 3), 10x10 grid (config 4b),
             the reference's own CUDA kernel (unmodified, sm_100a) on the same frame
  strong     one timed step of BASELINE config 5 (4K, 1024 spp total split over the ranks, one reduce): the strong-scaling curve
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
    ap.add_argument("--no-flush", action="store_true", help="profiling runs only: no L2-flush fill between steps")
    ap.add_argument("--no-strong", action="store_true", help="skip the config-5 strong-scaling step (4K, 1024 spp)")
    ap.add_argument("--no-configs", action="store_true", help="skip the side measurements (primary / bounce / AO / 4b / reference CUDA kernel)")
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
            cmd = ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"]
            if index is not None:   # one GPU, or a comma-separated list (multi-rank runs: the step time is the slowest rank's)
                cmd += ["-i", str(index)]
            self.proc = subprocess.Popen(cmd, stdout=self.f, stderr=subprocess.DEVNULL)
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
        sm, mx, reasons, per_gpu = [], [], set(), {}
        for line in open(self.path):
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
                per_gpu.setdefault(c[0], ([], []))[0].append(float(c[1])); per_gpu[c[0]][1].append(float(c[3]))
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
        out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
        if len(per_gpu) > 1:
            out["per_gpu"] = {k: {"sm_mhz": float(np.median(v[0])), "sm_mhz_min": float(min(v[0])), "power_w": float(np.median(v[1]))} for k, v in sorted(per_gpu.items())}
        return out


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


def csrc_hash():
    """sha1 over the sources the megakernel is compiled from: profiles/traffic.json records it, so a stale ncu capture is never
    applied to another kernel."""
    import hashlib
    h = hashlib.sha1()
    d = os.path.join(ROOT, "cuda-raytracing_b200", "csrc")
    # what the megakernel is compiled from (the tree it walks is tied to a capture by its measured steps per ray, load_capture)
    for name in ("accel.h", "kernels.h", "pathtrace.cu", "pathtrace_common.cuh", "rng.cuh", "shade.cuh", "traverse.cuh"):
        h.update(name.encode()); h.update(open(os.path.join(d, name), "rb").read())
    return h.hexdigest()[:16]


def load_capture(key, node_steps_per_ray=None, tri_tests_per_ray=None):
    """One entry of profiles/traffic.json (written by benchmarks/update_traffic.py from an ncu --set full report), or None when it
    was captured on other device sources than the ones in this source tree, or on another BVH: the capture records the wide-node
    steps and triangle tests per ray that the captured launch's tree gave (variant 3 counters of the same bench command), and this
    run's live counters must agree - a different tree means a different instruction count per ray."""
    try:
        cap = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[key]
    except Exception:
        return None, "no capture in profiles/traffic.json"
    if cap.get("csrc_hash") != csrc_hash():
        return None, f"capture in profiles/traffic.json was taken on csrc {cap.get('csrc_hash')}, tree is {csrc_hash()}: re-run benchmarks/update_traffic.py"
    for name, live in (("node_steps_per_ray", node_steps_per_ray), ("tri_tests_per_ray", tri_tests_per_ray)):
        if live is not None and (cap.get(name) is None or abs(cap[name] - live) > 1e-4 * live):
            return None, f"capture in profiles/traffic.json was taken on a BVH with {name} = {cap.get(name)}, this run's has {live}: re-run benchmarks/update_traffic.py"
    return cap, None


def timed_median(torch, fn, reps, flush):
    """median CUDA-event time of fn() in ms: one untimed call first, L2 flushed before every timed call (outside the events)"""
    fn(); torch.cuda.synchronize()
    ms = []
    for k in range(reps):
        flush.fill_(k & 0xFF)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return sorted(ms)[len(ms) // 2]


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
    host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))
        host_group = dist.new_group(backend="gloo")  # host-side waits: ranks that idle while rank 0 drives every GPU must not park a NCCL kernel on theirs
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
    # ... and what the shipped kernel actually executes (variant 3: the megakernel with two counters): wide-node steps x 64 B +
    # triangle tests x 48 B + 64 B per ray
    work.zero_(); d.counter.zero_()
    process_ex(d.surface, d.last, W, H, d.pitch, 0, d.gpu, d.accel, spp=2, bounces=a.bounces, blend_mode=0, variant=3,
               ray_counter=d.counter, stream=stream, work_counters=work)
    torch.cuda.synchronize()
    cnt3 = int(d.counter.item()); steps_per_ray = work[0].item() / cnt3; tests_per_ray = work[1].item() / cnt3
    bytes_per_ray_product = 64.0 * steps_per_ray + 48.0 * tests_per_ray + 64.0
    d.rng.copy_(rng_keep)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    blend = 0 if world == 1 else 1
    sum_buf = d.surface

    def step():
        if blend == 1:
            sum_buf.zero_()
        frame(blend, sum_buf)
        multigpu.reduce_frame(sum_buf, dist if world > 1 else None)

    first_launch_ms = None
    for w in range(max(a.warmup, 0)):
        if not a.no_flush:
            flush.fill_(1)
        if w == 0:  # the one launch without tile-cost history (grid order): what a camera jump or resize costs
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record(); step(); f1.record(); torch.cuda.synchronize()
            first_launch_ms = f0.elapsed_time(f1)
        else:
            step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    d.counter.zero_()
    uuid = str(torch.cuda.get_device_properties(local).uuid)
    uuid = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
    if world > 1:   # the GPUs of THIS job (idle GPUs of the node sit at 120 MHz and must not enter the median)
        uuids = [None] * world
        dist.all_gather_object(uuids, uuid, group=host_group)
        uuid = ",".join(uuids)
    sampler = ClockSampler(uuid) if rank == 0 else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    wall0 = time.perf_counter()
    for k in range(a.steps):
        if not a.no_flush:
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
        per_rank = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(per_rank, t)
        per_rank_ms = [round(x[0].item() / a.steps, 3) for x in per_rank]
        step_ms, kern_ms, rays_all = tmax[0].item(), tmax[1].item(), tsum[2].item()
    else:
        rays_all = float(rays); per_rank_ms = None
    value = rays_all / (step_ms * 1e-3) / 1e6

    # ---- e2e: the headless facade call with HOST buffers ----
    e2e = None
    if not a.no_e2e:
        e2e = run_e2e(a, rt, torch, dev, world, rank) if world == 1 else run_e2e_multi(a, rt, torch, dist, host_group, world, rank)

    strong = None
    if not a.no_strong:
        strong = run_strong_step(a, rt, torch, dist if world > 1 else None, dev, world, rank)
    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0
    configs = None
    if world == 1 and not a.no_configs:
        configs = run_configs_block(a, rt, torch, d, flush)

    peak, peak_src = measured_peaks()
    launches = a.steps * launches_per_step
    px = W * H
    launch_s = (kern_ms / launches) * 1e-3
    rays_per_launch = rays / launches
    props = torch.cuda.get_device_properties(dev)
    mhz = (clocks or {}).get("sm_mhz") or 1965.0
    peak_ginst = props.multi_processor_count * 4 * mhz * 1e-3  # 4 schedulers per SM, one warp instruction per clock each
    # The scene is L2-resident (22 MB), so HBM does not bound this kernel: instruction issue does. Warp-instruction count, DRAM and
    # L2 bytes per launch come from the committed ncu capture of this exact launch (never measured under bench); time and clock are live.
    cap, cap_note = (None, "not the captured launch shape")
    if a.variant == 0 and (a.width, a.height, a.spp_per_launch, a.bounces, a.pose) == (1920, 1080, 64, 4, "poseB"):
        cap, cap_note = load_capture("pathtrace_megakernel", steps_per_ray, tests_per_ray)
    alg_ref = rays_per_launch * bytes_per_ray + px * 80.0
    alg_prod = rays_per_launch * bytes_per_ray_product + px * 80.0
    roofline = {"bound": "issue", "achieved": None, "peak": peak_ginst, "unit": "G warp-inst/s", "frac": None, "traffic": None,
                "kernel": "pathtrace_megakernel<0,5>", "threads_per_instruction": None, "lane_frac": None,
                "warp_instructions_per_ray": None, "dram_bytes": None, "dram_gbs": None, "dram_frac": None, "l2_gbs": None,
                "hbm_peak_gbs": peak, "hbm_peak_source": peak_src,
                "peak_source": f"{props.multi_processor_count} SMs x 4 schedulers x {mhz:.0f} MHz (median SM clock sampled during the timed region)",
                "capture": cap["source"] if cap else cap_note,
                "bytes_per_ray_reference_order": bytes_per_ray, "nodes_per_ray_reference_order": nodes_per_ray, "tris_per_ray_reference_order": tris_per_ray,
                "hbm_algorithmic_gbs": alg_ref / launch_s / 1e9,
                "bytes_per_ray_product_order": bytes_per_ray_product, "node_steps_per_ray_product_order": steps_per_ray,
                "tri_tests_per_ray_product_order": tests_per_ray, "product_order_gbs": alg_prod / launch_s / 1e9,
                "note": "bound = warp-instruction issue: the 22 MB scene is L2-resident and DRAM carries only surfaces + RNG states (dram_frac). "
                        "hbm_algorithmic_gbs = SURVEY 8(d)'s bytes per ray under the REFERENCE traversal order over the kernel time (informational: "
                        "it exceeds the HBM peak because near-first order + the fat-leaf sub-tree remove most of that work and none of it goes to "
                        "HBM); product_order_gbs = the bytes the shipped kernel's own node steps and triangle tests read, served by L1 / L2"}
    if cap:
        ach = cap["warp_instructions_per_launch"] / launch_s / 1e9
        roofline.update({"achieved": ach, "frac": ach / peak_ginst, "traffic": cap["dram_bytes_per_launch"],
                         "threads_per_instruction": cap["threads_per_instruction"], "lane_frac": cap["threads_per_instruction"] / 32.0,
                         "warp_instructions_per_ray": cap["warp_instructions_per_launch"] / rays_per_launch,
                         "dram_bytes": cap["dram_bytes_per_launch"], "dram_gbs": cap["dram_bytes_per_launch"] / launch_s / 1e9,
                         "dram_frac": cap["dram_bytes_per_launch"] / launch_s / 1e9 / peak,
                         "dram_bytes_algorithmic": px * 80.0 + 22e6,
                         "l2_gbs": cap.get("l2_bytes_per_launch", 0) / launch_s / 1e9 if cap.get("l2_bytes_per_launch") else None})
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": step_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(a), "spp_per_launch": a.spp_per_launch, "kernel_variant": a.variant,
                   "l2": "flushed between timed steps (256 MiB fill, outside the event pairs); scene working set itself is L2-resident by nature",
                   "schedule": "timed steps are history-warmed: the cost-ordered tile schedule uses the previous launch's measured tile costs (static camera); "
                               "the first launch after a resize or camera jump runs in grid order (see config.first_launch_ms)",
                   "first_launch_ms": first_launch_ms,
                   "parallelism": "1 GPU" if world == 1 else f"spp split: {world} ranks x {a.spp} spp, disjoint XORWOW subsequences, 1 NCCL reduce(sum fp32) of the frame per step",
                   "rays_per_step": rays_all / a.steps, "ms_per_frame_64spp": step_ms / a.steps, "wall_s": wall,
                   **({"ms_per_step_per_rank": per_rank_ms} if per_rank_ms else {})},
        "clocks": clocks, "gpu_launches": launches * (2 if a.variant == 0 else 1), "roofline": roofline,
    }
    if strong:
        line["strong"] = strong
    if configs:
        line["configs"] = configs
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


def run_e2e_multi(a, rt, torch, dist, host_group, world, rank):
    """N > 1: the partitioned job through the repo's own multi-GPU call (rt_host_multi_tracer_render, host/MultiGPU.h): ONE process
    (rank 0) drives all N devices with one host thread each - spp split, ncclReduce inside the library, resolve, float4 frame
    into pinned HOST memory. The other ranks wait on the host (gloo) with their GPUs idle. Weak scaling like `value`: every device
    renders a.spp samples per step, the frame holds N * a.spp."""
    torch.cuda.synchronize()
    dist.barrier(group=host_group)
    out = None
    if rank == 0:
        W, H = a.width, a.height
        assets = util.assets()
        host = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
        m = rt.MultiTracer(W, H, bunny_obj=assets["obj"], env_dds=assets["dds"], max_bounces=a.bounces, seed=a.seed, devices=list(range(world)))
        m.set_camera(*util.POSES[a.pose])
        for _ in range(max(a.warmup, 1)):
            m.render(a.spp * world, a.spp_per_launch, host)
        r0 = m.ray_count
        t0 = time.perf_counter()
        tm = None
        for _ in range(a.steps):
            tm = m.render(a.spp * world, a.spp_per_launch, host)
        dt = time.perf_counter() - t0
        rays = m.ray_count - r0
        assert float(host[..., 3].min()) == 1.0 and bool(torch.isfinite(host).all())
        out = {"value": rays / dt / 1e6, "unit": UNIT, "h2d_bytes_per_step": world * 136 * ((a.spp + a.spp_per_launch - 1) // a.spp_per_launch),
               "d2h_bytes_per_step": W * H * 16, "ms_per_step": 1e3 * dt / a.steps, "last_step": tm,
               "call": f"rt_host_multi_tracer_render (MultiGPUTracer::Render): {world} devices x {a.spp} spp, one ncclReduce, pinned host frame out"}
        del m
    dist.barrier(group=host_group)
    return out


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


def run_configs_block(a, rt, torch, d, flush):
    """Bounded side measurements (about 15 s), the bench step's own protocol (L2 flushed, CUDA events, median of 3): the north-star
    targets (primary >= 2 Grays/s, diffuse bounce >= 0.5 Grays/s), BASELINE configs 3 and 4b, and the reference's own CUDA kernel."""
    import ref_oracle as ro
    W, H = a.width, a.height
    out = {}

    def pt(dev, spp, bounces):
        dev.counter.zero_()
        f = lambda: dev.process(0, spp=spp, bounces=bounces, write_last=1)
        f(); n = dev.rays()
        return n, timed_median(torch, f, 3, flush)

    n1, ms1 = pt(d, 16, 1)
    n4, ms4 = pt(d, 16, a.bounces)
    out["primary_only"] = {"what": f"{W}x{H}, 16 spp, max_bounces=1 (jittered camera rays only), camera {a.pose}", "rays": n1, "ms": ms1, "mrays_s": n1 / ms1 / 1e3}
    out["bounce_only"] = {"what": f"rays of depth >= 1: the {a.bounces}-bounce 16-spp launch ({n4} rays, {ms4:.3f} ms) minus the primary-only launch",
                          "rays": n4 - n1, "ms": ms4 - ms1, "mrays_s": (n4 - n1) / (ms4 - ms1) / 1e3}
    ao = torch.zeros(W * H, dtype=torch.float32, device=d.device)
    d.counter.zero_()
    f = lambda: rt.check(rt.lib().rt_ambient_occlusion(d.accel, d.gpu, W, H, 16, rt.ptr(ao), rt.ptr(d.counter), None))
    f(); n = d.rays()
    ms = timed_median(torch, f, 3, flush)
    out["ao_config3"] = {"what": f"{W}x{H}: pixel-centre primary ray + 16 diffuse-lobe rays per hit", "rays": n, "ms": ms, "mrays_s": n / ms / 1e3}
    # the reference's own CUDA kernel (unmodified main_raytracing.cu / Random.cu, nvcc -O3 sm_100a) on the same GPUScene: its frame is
    # 5 spp x 6 bounces + the caller's D2D copy (RayTracing.cpp:232-233); ray count of that workload from the product's counter
    if ro.gpu_available():
        g = ro.gpu_lib()
        nref, mine = pt(d, 5, 6)

        def ref_frame():
            g.raytracing_process(rt.ptr(d.surface), rt.ptr(d.last), W, H, d.pitch, 0, d.gpu)
            d.last.copy_(d.surface)
        ms = timed_median(torch, ref_frame, 3, flush)
        out["reference_cuda_kernel"] = {"what": f"UNMODIFIED reference raytracing_kernel_main (oracle/_ref/libref_gpu.so), {W}x{H}, 5 spp x 6 bounces + its D2D copy, camera {a.pose}",
                                        "ms": ms, "rays_approx": nref, "mrays_s": nref / ms / 1e3,
                                        "product_ms_same_frame": mine, "product_mrays_s": nref / mine / 1e3, "speedup": ms / mine}
    else:
        out["reference_cuda_kernel"] = {"unavailable": "oracle/_ref/libref_gpu.so not built"}
    # config 4b: the only config whose working set exceeds L2
    t0 = time.perf_counter()
    scene, mesh, pose, times = util.scaled_scene(rt, "b", env=True)
    import gpu_util
    g4 = gpu_util.DeviceScene(rt, torch, W, H, pose, seed=a.seed, env=True, scene=scene, device=d.device)
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0
    n, ms = pt(g4, 16, a.bounces)
    out["grid_config4b"] = {"what": f"10x10 bunny grid, {scene.counts()['faces']} triangles, {W}x{H}, 16 spp, {a.bounces} bounces",
                            "rays": n, "ms": ms, "mrays_s": n / ms / 1e3, "scene_build_upload_s": build_s,
                            "accel_bytes": int(scene.accel_info()["device_bytes"])}
    del g4, scene, mesh
    torch.cuda.empty_cache()
    return out


def run_strong_step(a, rt, torch, dist, dev, world, rank):
    """BASELINE config 5 as ONE timed step: 3840x2160, 1024 spp in total split evenly over the ranks (64-spp launches into a sum
    buffer, disjoint XORWOW subsequences), one reduce(sum, fp32) of the 132.7 MB frame to rank 0. RNG init is outside the timed
    region (reported). Time = max over ranks, CUDA events."""
    import gpu_util
    from cuda_raytracing_b200 import multigpu
    W, H, TOTAL = 3840, 2160, 1024
    if TOTAL % world:
        return {"unavailable": f"1024 spp do not split evenly over {world} ranks"}
    spp_rank = TOTAL // world
    d = gpu_util.DeviceScene(rt, torch, W, H, a.pose, seed=a.seed, env=True, device=dev)
    r0, r1, e1, e2, e3 = (torch.cuda.Event(enable_timing=True) for _ in range(5))
    n_states = gpu_util.rng_state_count(W, H)
    r0.record(); d.reseed(a.seed, multigpu.rank_first_subsequence(rank, n_states)); r1.record()
    acc = d.surface
    d.process(0, spp=4, bounces=a.bounces, blend_mode=1, surface=acc)  # warm-up launch: gives the tile schedule its cost history
    d.reseed(a.seed, multigpu.rank_first_subsequence(rank, n_states))
    acc.zero_(); d.counter.zero_()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e1.record()
    launches = 0
    left = spp_rank
    while left > 0:
        s = min(64, left)
        d.process(launches, spp=s, bounces=a.bounces, blend_mode=1, surface=acc)
        left -= s; launches += 1
    e2.record()
    multigpu.reduce_frame(acc, dist if world > 1 else None)
    e3.record()
    torch.cuda.synchronize()
    t = torch.tensor([e1.elapsed_time(e3), e1.elapsed_time(e2), e2.elapsed_time(e3), float(d.rays())], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, render_ms, reduce_ms, rays = tmax[0].item(), tmax[1].item(), tmax[2].item(), tsum[3].item()
    else:
        ms, render_ms, reduce_ms, rays = t[0].item(), t[1].item(), t[2].item(), t[3].item()
    ok = True
    if rank == 0:
        img = acc[..., :3] / acc[..., 3:4]
        ok = bool(torch.isfinite(img).all()) and float(acc[..., 3].min()) == float(TOTAL) == float(acc[..., 3].max())
    del d
    torch.cuda.empty_cache()
    return {"width": W, "height": H, "spp_total": TOTAL, "spp_per_rank": spp_rank, "launches_per_rank": launches, "ms": ms,
            "render_ms": render_ms, "reduce_ms": reduce_ms, "rng_init_ms_untimed": r0.elapsed_time(r1), "reduce_bytes": W * H * 16 if world > 1 else 0, "rays": rays,
            "mrays_s": rays / ms / 1e3, "alpha_equals_spp_total": ok,
            "limiter": "per-rank render time (the slowest rank's launch tail); the reduce is reduce_ms of ms"}


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
