"""Host BVH build time on the scaled meshes of benchmarks/scaled_meshes.py, no GPU needed:
    B200RT_BVH_TIMING=1 OMP_NUM_THREADS=<n> python benchmarks/bvh_build_time.py [a|b|ab]
Topology identity with the reference's BVH.cpp is a CPU test (tests/test_layout_and_scene.py)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "benchmarks")]
from _loader import load_package  # noqa: E402

rt = load_package()
import scaled_meshes as sm  # noqa: E402

for w in (sys.argv[1] if len(sys.argv) > 1 else "ab"):
    scene, mesh, pose, t = sm.build_scene(rt, w, env=False)
    c = scene.counts()
    print(json.dumps({"config": w, "threads": os.environ.get("OMP_NUM_THREADS", "all"), "triangles": c["faces"], "nodes": c["nodes"], **t}), flush=True)
