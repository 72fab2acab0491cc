"""CPU tests of bench.py's host-side bookkeeping (no GPU): which ncu capture may feed the roofline of a bench line."""
import importlib.util
import json
import os

import util


def load_bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(util.ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_capture_is_tied_to_kernel_sources_and_to_the_tree():
    """profiles/traffic.json holds the warp-instruction count of ONE captured launch. bench.py may apply it to a live run only when
    the kernel sources hash to what the capture recorded AND the tree the live run walks gives the wide-node steps and triangle
    tests per ray the captured launch's tree gave (a different BVH = a different instruction count per ray)."""
    bench = load_bench()
    cap = json.load(open(os.path.join(util.ROOT, "profiles", "traffic.json")))["pathtrace_megakernel"]
    assert cap["csrc_hash"] == bench.csrc_hash(), "profiles/traffic.json is stale: re-capture (benchmarks/gpu_bench_and_profile.sh <tag> full, update_traffic.py)"
    assert cap["node_steps_per_ray"] and cap["tri_tests_per_ray"] and cap["warp_instructions_per_launch"] > 1e9
    got, note = bench.load_capture("pathtrace_megakernel", cap["node_steps_per_ray"], cap["tri_tests_per_ray"])
    assert got is not None and note is None
    got, note = bench.load_capture("pathtrace_megakernel", cap["node_steps_per_ray"] * 1.01, cap["tri_tests_per_ray"])
    assert got is None and "BVH" in note
    got, note = bench.load_capture("no_such_kernel")
    assert got is None
