"""Build recipe for libb200rt.so (sm_100a only) - explicit nvcc, in-tree output.

    python cuda-raytracing_b200/build.py [--force]

Device translation units are compiled with -fmad=false and host code with -ffp-contract=off: the
parity oracle (the reference compiled host-side) rounds every multiply and add separately, and
bit-exact hit ids hang on doing the same (DESIGN.md "contraction policy"). FMAs the kernels want are
written explicitly as fmaf().
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_DIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIB_DIR, "libb200rt.so")

SOURCES = [
    "csrc/kernels.cu", "csrc/pathtrace.cu", "csrc/pathtrace_regroup.cu", "csrc/pathtrace_wavefront.cu", "csrc/capi.cu", "csrc/capi_render.cu",
    "csrc/accel_build.cpp", "csrc/dropin.cpp", "csrc/dds.cpp", "csrc/rng_host.cpp",
    "host/Math.cpp", "host/BVH.cpp", "host/MeshLoader.cpp", "host/Subdivide.cpp", "host/Scene.cpp",
    "host/SceneSetup.cpp", "host/SceneUpload.cpp", "host/RayTracing.cpp", "host/MultiGPU.cpp", "host/host_capi.cpp",
]
HEADERS_DIRS = ["csrc", "host", os.path.join(ROOT, "include")]


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: libb200rt.so cannot be built (there is no CPU fallback)")
    return exe


def _newest_source_mtime():
    newest = os.path.getmtime(os.path.abspath(__file__))
    for d in HEADERS_DIRS:
        d = d if os.path.isabs(d) else os.path.join(HERE, d)
        for name in os.listdir(d):
            if name.endswith((".cu", ".cuh", ".cpp", ".h")):
                newest = max(newest, os.path.getmtime(os.path.join(d, name)))
    return newest


def needs_build():
    return not os.path.exists(LIB) or os.path.getmtime(LIB) < _newest_source_mtime()


def build(force=False, verbose=False, defines=(), out=None):
    """Compile libb200rt.so for sm_100a. Returns the library path. `defines`/`out`: A/B builds of tuning switches
    (e.g. defines=["RT_TW_FMA=1"], out="libb200rt_fma.so"; select at run time with B200RT_LIB=<path>)."""
    global LIB
    if out:
        force = True
    if not force and not needs_build():
        return LIB
    os.makedirs(LIB_DIR, exist_ok=True)
    obj_dir = os.path.join(HERE, "build")
    os.makedirs(obj_dir, exist_ok=True)
    common = [
        _nvcc(), "-std=c++20", "-O3", "-lineinfo", "-fmad=false",
        "-gencode", "arch=compute_100a,code=sm_100a",
        "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=default,-Wall,-Wno-unused-function,-fopenmp",
        "-I", os.path.join(ROOT, "include"), "-I", os.path.join(HERE, "csrc"), "-I", os.path.join(HERE, "host"),
    ]
    if verbose:
        common += ["-Xptxas", "-v"]
    common += ["-D" + d for d in defines]
    target = os.path.join(LIB_DIR, out) if out else LIB
    if out:
        obj_dir = os.path.join(HERE, "build", out.replace(".so", ""))
        os.makedirs(obj_dir, exist_ok=True)
    procs = []
    objects = []
    for src in SOURCES:
        obj = os.path.join(obj_dir, src.replace("/", "_") + ".o")
        objects.append(obj)
        cmd = common + ["-x", "cu", "-c", os.path.join(HERE, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write(f"--- {src} ---\n{out}\n")
        elif verbose or "warning" in out:
            sys.stderr.write(f"--- {src} ---\n{out}\n")
    if failed:
        raise RuntimeError("nvcc failed")
    link = [_nvcc(), "-shared", "-o", target] + objects + ["-cudart", "static", "-Xcompiler", "-fopenmp", "-lgomp", "-ldl"]
    subprocess.run(link, check=True)
    return target


REDO_LIB = os.path.join(LIB_DIR, "libb200rt_redo.so")


def build_redo_variant(force=False):
    """Test build (tests/test_gpu_redo_build.py): RT_TW_FORCE_REDO=7 sends every ray whose winning slot is a multiple of 8 through
    the redo path of trace_wide (per-candidate exact leaf-box check), which a grazing ray takes about never; RT_STACK_SIZE=32 (64 in
    the product) makes the builder's depth limiter - SAH splits replaced by halving near the end of the traversal stack - work on
    the bunny scene, whose SAH tree is 21 levels deep with an 8-level fat-leaf sub-tree below."""
    if not force and os.path.exists(REDO_LIB) and os.path.getmtime(REDO_LIB) >= _newest_source_mtime():
        return REDO_LIB
    return build(defines=["RT_TW_FORCE_REDO=7", "RT_STACK_SIZE=32"], out="libb200rt_redo.so")


if __name__ == "__main__":
    defs = [a[2:] for a in sys.argv[1:] if a.startswith("-D")]
    outs = [a[6:] for a in sys.argv[1:] if a.startswith("--out=")]
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, defines=defs, out=outs[0] if outs else None))
