"""CPU tests: data-model layout, host scene assembly and BVH topology against the reference
(oracle/_ref = the reference's GPUScene.h / BVH.cpp / GLM compiled host-side) and the golden hashes."""
import ctypes as C
import hashlib
import json
import os

import numpy as np

import util


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_library_exports_every_declared_symbol(rt):
    lib = rt.lib()
    import re
    declared = set()
    for h in ("b200rt.h", "b200rt_host.h"):
        text = open(os.path.join(util.ROOT, "include", h)).read()
        declared |= set(re.findall(r"\b((?:rt_|raytracing_process|init_rng)\w*)\s*\(", text))
    declared -= {"rt_last_error()"}
    assert {"raytracing_process", "init_rng", "rt_process_ex", "rt_accel_create"} <= declared
    for name in sorted(declared):
        assert hasattr(lib, name), name
    assert declared <= set(rt.SYMBOLS), declared - set(rt.SYMBOLS)


def test_shipped_sass_has_no_contracted_packed_fma(rt):
    """Contraction policy guard: the device code is built with -fmad=false so that every a*b+c rounds
    twice like the host-compiled reference. nvcc nevertheless fuses a packed __fmul2_rn feeding a packed __fadd2_rn into FFMA2; the
    product's packed code is written so that this never happens (the FMA slab form, RT_PT_FMA, is off). Scalar FFMA may only appear
    where fmaf() is written (conservative box culls) and inside the IEEE division / sqrt sequences. The only FFMA2 the source
    writes is the conservative wide-node test of the centre + half-extent form (traverse.cuh center_slabs: half * (-/+)|1/d| + mid);
    each of those carries the |.| operand modifier, which no contracted multiply-add of the triangle / shading arithmetic has."""
    import shutil
    import subprocess
    from _loader import load_build
    tool = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(tool):
        import pytest
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([tool, "-sass", load_build().LIB], capture_output=True, text=True).stdout
    assert "pathtrace_megakernel" in sass and "FMUL2" in sass and "FADD2" in sass  # the packed paths are really there
    per_function, name = {}, None
    for l in sass.splitlines():
        if "Function :" in l:
            name = l.split("Function :")[1].strip()
        elif "FFMA2" in l:
            per_function.setdefault(name, []).append(l)
    assert per_function
    for name, lines in per_function.items():
        # six per wide-node test; a kernel holds the test twice (the deferred-check loop and the __noinline__ redo copy)
        assert len(lines) == 12, (name, len(lines))
    mega = [v for k, v in per_function.items() if "pathtrace_megakernelILi0" in k][0]
    assert all("|R" in l for l in mega), [l for l in mega if "|R" not in l][:3]


def test_pod_layout_matches_reference_structs(rt, ref):
    s = ref.lib().ref_sizeof
    assert [s(i) for i in range(8)] == [32, 16, 32, 64, 32, 60, 136, 48]
    assert (s(8), s(9), s(10)) == (72, 56, 64)          # GPUScene: camera, rng_state, env tex
    assert (s(11), s(12), s(13)) == (4, 8, 24)          # GPUFace v2, v1 ; GPUBVHNode first_index
    assert (s(14), s(15), s(16), s(17)) == (48, 16, 24, 48)
    assert C.sizeof(rt.GPUScene) == 136 and rt.CAMERA.itemsize == 60


def test_compute_entry_points_fail_loudly_without_device(rt):
    if rt.lib().rt_device_count() > 0:
        return
    scene = rt.Scene()
    scene.add_triangle((0, 0, 1), (1, 0, 1), (0, 1, 1), scene.add_material())
    try:
        scene.upload(None)
    except rt.RtError as e:
        assert "no CPU fallback" in str(e) or "CUDA" in str(e)
    else:
        raise AssertionError("upload must fail without a device")


def test_scene_assembly_bitwise_vs_glm_build(rt, ref, bunny):
    a = bunny["arrays"]
    pos, nrm, idx, _ = bunny["mesh"].arrays()
    r = ref.build_scene(pos, nrm, idx)
    assert np.array_equal(util.raw(a["vertices"]), r["vertices"])
    assert np.array_equal(util.raw(a["faces"]), r["faces"])
    assert np.array_equal(util.raw(a["spheres"])[:, :20], r["spheres"][:, :20])
    assert np.array_equal(util.raw(a["materials"])[:, :60], r["materials"][:, :60])
    # Appendix A quirks 1/2: ids interleave smooth/flat, Cornell 0..11, floor last two
    f = a["faces"]
    assert len(f) == 138916 and (f["material"][:12:2] == np.arange(6)).all()
    k = 12 + 2 * 1000
    assert f["v2"][k + 1] == f["v0"][k + 1] + 1 and f["v1"][k + 1] == f["v0"][k + 1] + 2


def test_smooth_mesh_fixture_is_the_ingest_output(rt, ref, bunny):
    """tests/golden/bunny_smooth_mesh.npz (what bench.py's reference arm assembles its scene from, without libb200rt.so) is the
    product's ingest output, and the reference-side assembly of it is the scene every parity test runs on."""
    import os
    m = np.load(os.path.join(util.GOLDEN, "bunny_smooth_mesh.npz"))
    pos, nrm, idx, _ = bunny["mesh"].arrays()
    assert np.array_equal(m["positions"], pos) and np.array_equal(m["normals"], nrm) and np.array_equal(m["indices"], idx)
    r = ref.build_scene(np.ascontiguousarray(m["positions"]), np.ascontiguousarray(m["normals"]), np.ascontiguousarray(m["indices"]))
    assert np.array_equal(util.raw(bunny["arrays"]["vertices"]), r["vertices"]) and np.array_equal(util.raw(bunny["arrays"]["faces"]), r["faces"])


def test_bvh_topology_identical_to_reference_builder(rt, ref, bunny):
    a = bunny["arrays"]
    nodes, face_indices = ref.bvh_build(util.raw(a["vertices"]), util.raw(a["faces"]))
    assert len(nodes) == 138545
    assert np.array_equal(util.raw(a["nodes"]), nodes)
    assert np.array_equal(a["face_indices"], face_indices)
    leaves = a["nodes"][a["nodes"]["prim_count"] > 0]
    assert len(leaves) == 69273 and leaves["prim_count"].max() == 345  # the fat leaf


def test_small_scenes_bvh_vs_reference(rt, ref):
    rng = np.random.default_rng(3)
    for n in (1, 2, 3, 17, 200):
        scene = rt.Scene()
        m = scene.add_material()
        for _ in range(n):
            c = rng.uniform(-5, 5, 3)
            tri = (c + rng.normal(size=(3, 3))).astype(np.float32)
            scene.add_triangle(tri[0], tri[1], tri[2], m)
        scene.build_bvh()
        a = scene.arrays()
        nodes, fi = ref.bvh_build(util.raw(a["vertices"]), util.raw(a["faces"]))
        assert np.array_equal(util.raw(a["nodes"]), nodes) and np.array_equal(a["face_indices"], fi)


def test_parallel_bvh_build_adversarial_soup_vs_reference(rt, ref):
    """The parallel builder (level-synchronous top tree + per-thread subtrees + block renumbering, host/BVH.cpp) against the
    reference's recursive BVH.cpp on a soup that exercises every branch: several top-tree levels (120 k triangles > 7 subtrees),
    an unsplittable leaf LARGER than a subtree (20 k identical triangles stay one top-tree leaf), clusters of different density,
    triangles much larger than their neighbours, and axis fallbacks (flat clusters)."""
    rng = np.random.default_rng(11)
    scene = rt.Scene()
    m = scene.add_material()
    tris = []
    for centre, spread, n in (((0, 0, 0), 20.0, 60000), ((50, 5, -30), 2.0, 30000), ((-40, -10, 10), 0.5, 10000)):
        c = rng.normal(size=(n, 1, 3)) * spread + np.asarray(centre)
        tris.append(c + rng.normal(size=(n, 3, 3)) * 0.3)
    flat = rng.uniform(-30, 30, size=(2000, 1, 3)); flat[..., 1] = 7.0           # coplanar sheet: y extent 0 for its nodes
    tris.append(flat + rng.normal(size=(2000, 3, 3)) * np.asarray([0.5, 0.0, 0.5]))
    tris.append(np.repeat(np.asarray([[[3.0, 3.0, 3.0], [4.0, 3.0, 3.0], [3.0, 4.0, 3.0]]]), 20000, axis=0))  # 20 k identical
    tris.append(np.asarray([[[-200.0, -50.0, -200.0], [200.0, -50.0, -200.0], [0.0, -50.0, 300.0]]]))         # one giant triangle
    soup = np.concatenate(tris).astype(np.float32)
    soup = soup[rng.permutation(len(soup))]
    for t in soup:
        scene.add_triangle(t[0], t[1], t[2], m)
    scene.build_bvh()
    a = scene.arrays()
    nodes, fi = ref.bvh_build(util.raw(a["vertices"]), util.raw(a["faces"]))
    assert np.array_equal(a["face_indices"], fi)
    assert np.array_equal(util.raw(a["nodes"]), nodes)
    assert a["nodes"]["prim_count"].max() >= 20000


def test_accel_relayout_structure_on_the_host(rt, bunny):
    """The B200 re-layout (one arena, byte-offset references, fat-leaf sub-trees) built and walked on the host, no device: every BVH slot
    reachable exactly once from the root, references aligned and in range, leaf bookkeeping consistent, and every wide-node child box
    (centre + half extent) holding, in exact arithmetic, the reference leaf boxes / triangles below it (what makes the conservative node
    test safe), plus a replay of the kernel's own fp32 node test on sampled rays that hit those bounds by an exact double-precision slab
    test: none may be rejected (with the builder's and the kernel's margins removed the replay does find rejected hits) - bunny scene, every fat-leaf threshold, and small random soups through the reference-identical BVH builder."""
    def check(a, threshold):
        info = (C.c_uint64 * 4)()
        r = rt.lib().rt_accel_check_host(rt.ptr(a["nodes"]), len(a["nodes"]), rt.ptr(a["face_indices"]), rt.ptr(a["faces"]), len(a["faces"]),
                                         rt.ptr(a["vertices"]), len(a["vertices"]), threshold, info)
        assert r == 0, rt.lib().rt_last_error()
        return [int(v) for v in info]
    a = bunny["arrays"]
    inner, leaves, depth, arena = check(a, 0)
    assert arena - len(a["faces"]) * 48 == inner * 64 and leaves >= 69273 and depth < 62
    for threshold in (1, 2, 4, 16):
        i2, l2, d2, _ = check(a, threshold)
        assert i2 + 1 == l2 or i2 == l2 - 1  # a binary tree
    rng = np.random.default_rng(5)
    for n in (1, 2, 3, 17, 400):
        scene = rt.Scene(None)
        for _ in range(n):
            c = rng.uniform(-10, 10, 3); e = rng.normal(size=(3, 3))
            scene.add_triangle(*(c + e[k] for k in range(3)), 0)
        scene.build_bvh()
        check(scene.arrays(), 0)


def test_own_topology_over_the_reference_leaves(rt, bunny, monkeypatch):
    """accel_build.cpp builds its own SAH tree (+ one insertion-based optimisation pass) over the reference's leaves; B200RT_TOPOLOGY=ref
    keeps BVH.cpp's midpoint tree. Both are binary trees over the same leaves and pass the same host check (containment of the exact
    reference leaf boxes, replay of the kernel's node test); a build with a 32-entry traversal stack (the test build) has the builder halve by count where SAH
    splits would leave the stack, and still passes."""
    def check(a):
        info = (C.c_uint64 * 4)()
        r = rt.lib().rt_accel_check_host(rt.ptr(a["nodes"]), len(a["nodes"]), rt.ptr(a["face_indices"]), rt.ptr(a["faces"]), len(a["faces"]),
                                         rt.ptr(a["vertices"]), len(a["vertices"]), 0, info)
        assert r == 0, rt.lib().rt_last_error()
        return [int(v) for v in info]
    a = bunny["arrays"]
    own = check(a)
    monkeypatch.setenv("B200RT_REINSERT", "0")
    sah_only = check(a)
    monkeypatch.setenv("B200RT_TOPOLOGY", "ref")
    ref_tree = check(a)
    monkeypatch.delenv("B200RT_TOPOLOGY"); monkeypatch.delenv("B200RT_REINSERT")
    assert own[:2] == sah_only[:2] == ref_tree[:2] and own[3] == ref_tree[3]   # same leaves, same number of inner nodes, same arena
    assert own[2] < 40 and ref_tree[2] < 40
    # the depth limiter: the test build has a 32-entry traversal stack (build.py build_redo_variant)
    from _loader import load_build
    small = C.CDLL(load_build().REDO_LIB)
    small.rt_accel_check_host.restype = C.c_int
    small.rt_accel_check_host.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64)]
    info = (C.c_uint64 * 4)()
    r = small.rt_accel_check_host(rt.ptr(a["nodes"]), len(a["nodes"]), rt.ptr(a["face_indices"]), rt.ptr(a["faces"]), len(a["faces"]),
                                  rt.ptr(a["vertices"]), len(a["vertices"]), 0, info)
    assert r == 0
    assert list(info)[:2] == own[:2] and info[2] + 2 <= 32
    assert info[2] < own[2]   # (18 against 21 levels: the limiter did have work to do, and the reinsertion pass, whose result was 24 deep, was undone)


def test_relayout_fuzz_through_the_host_check(rt):
    """120 random soups through BVH.cpp's builder, the re-layout (own SAH tree, fat-leaf sub-trees) and the host check: 1 to 3 000
    triangles, scene scales 1e-3 to 1e4, triangle sizes 1e-4 to 2 scene units, and the degenerate families - identical centroids (one
    fat leaf), collinear triangles, everything in one axis-aligned plane (zero-thickness boxes), every triangle four times, a scene
    1 000 times longer than wide."""
    rng = np.random.default_rng(11)
    for case in range(120):
        n = int(rng.choice([1, 2, 3, 5, 9, 17, 40, 100, 300, 1000, 3000]))
        kind = case % 6
        scale = float(rng.choice([1e-3, 0.1, 1.0, 30.0, 1e4]))
        size = float(rng.choice([1e-4, 0.01, 0.3, 2.0])) * scale
        c = rng.uniform(-scale, scale, (n, 3)); e = rng.normal(size=(n, 3, 3)) * size
        if kind == 1:
            c[:] = c[0]
        if kind == 2:
            e[::2, 1] = e[::2, 0] * 2; e[::2, 2] = e[::2, 0] * 3   # every other triangle has its three vertices on one line
        if kind == 3:
            c[:, 1] = 0; e[:, :, 1] = 0
        if kind == 4:
            c = np.repeat(c[: max(1, n // 4)], 4, axis=0)[:n]; e = np.repeat(e[: max(1, n // 4)], 4, axis=0)[:n]
        if kind == 5:
            c[:, 0] *= 1e3
        scene = rt.Scene(None)
        for i in range(len(c)):
            scene.add_triangle(c[i] + e[i, 0], c[i] + e[i, 1], c[i] + e[i, 2], 0)
        scene.build_bvh()
        a = scene.arrays()
        for threshold in (0, 2):
            info = (C.c_uint64 * 4)()
            r = rt.lib().rt_accel_check_host(rt.ptr(a["nodes"]), len(a["nodes"]), rt.ptr(a["face_indices"]), rt.ptr(a["faces"]), len(a["faces"]),
                                             rt.ptr(a["vertices"]), len(a["vertices"]), threshold, info)
            assert r == 0, (case, n, kind, scale, size, threshold, rt.lib().rt_last_error())


def test_relayout_of_scenes_with_nan_and_infinite_vertices(rt):
    """BVH.cpp never splits what it cannot order, so NaN / infinite vertices end up in fat leaves with NaN / infinite boxes. The re-layout
    must come back (it used to recurse without end in the fat leaf's SAH sweep when no split had a finite cost) and pass the host check:
    infinite bounds are stored as +-1.7e38, NaN bounds as an empty box (no ray passes a NaN box in the reference either)."""
    rng = np.random.default_rng(3)
    for mode in ("nan", "inf", "both"):
        scene = rt.Scene(None)
        for i in range(3000):
            c = rng.uniform(-10, 10, 3); e = rng.normal(size=(3, 3))
            p = [c + e[k] for k in range(3)]
            if mode in ("nan", "both") and i % 97 == 0:
                p[1] = np.array([np.nan, 1.0, 2.0])
            if mode in ("inf", "both") and i % 101 == 0:
                p[2] = np.array([np.inf, 1.0, -np.inf])
            scene.add_triangle(*p, 0)
        scene.build_bvh()
        a = scene.arrays()
        info = (C.c_uint64 * 4)()
        r = rt.lib().rt_accel_check_host(rt.ptr(a["nodes"]), len(a["nodes"]), rt.ptr(a["face_indices"]), rt.ptr(a["faces"]), len(a["faces"]),
                                         rt.ptr(a["vertices"]), len(a["vertices"]), 0, info)
        assert r == 0, (mode, rt.lib().rt_last_error())
        assert info[2] + 2 <= 64


def test_golden_hashes(rt, bunny):
    """Runs everywhere (no /root/reference needed): arrays hash to what the GLM + BVH.cpp build produced."""
    h = json.load(open(os.path.join(util.GOLDEN, "scene_hashes.json")))
    a = bunny["arrays"]
    assert sha(util.raw(a["vertices"])) == h["vertices"]
    assert sha(util.raw(a["faces"])) == h["faces"]
    assert sha(util.raw(a["nodes"])) == h["nodes"]
    assert sha(a["face_indices"]) == h["face_indices"]
    assert sha(util.raw(a["spheres"])[:, :20]) == h["spheres20"]
    assert sha(util.raw(a["materials"])[:, :60]) == h["materials60"]
    for name, (p, ax, ay) in util.POSES.items():
        bunny["scene"].set_camera(p, ax, ay, 320, 240)
        cam, _ = bunny["scene"].camera()
        assert sha(cam.view(np.uint8)) == h["cameras"][name]


def test_camera_vs_glm(rt, ref):
    scene = rt.Scene()
    for p, ax, ay, w, h in [((0, 0, 0), 0, 180, 640, 480), ((32, -1, 2), 0, 180, 1920, 1080), ((3, 4, -5), 17.5, -63.25, 3840, 2160)]:
        scene.set_camera(p, ax, ay, w, h)
        cam, _ = scene.camera()
        assert np.array_equal(cam.view(np.uint8).ravel(), ref.camera(p, ax, ay, w, h))


def test_dds_reader_roundtrip(rt):
    tex = util.cubemap_texels()
    out = np.zeros_like(tex)
    size = rt.lib().rt_dds_read_level0(util.assets()["dds"].encode(), rt.ptr(out), out.size)
    assert size == 128 and np.array_equal(out, tex)
    assert rt.lib().rt_dds_read_level0(b"/nonexistent.dds", None, 0) == -4
