"""-m gpu parity tests: the sm_100a kernels, called through the C ABI (include/b200rt.h), against
  * the committed golden vectors produced by the reference's own code compiled host-side
    (tests/golden/make_fixtures.py, oracle/_ref), and
  * oracle/_ref itself on fresh seeded inputs when the prebuilt library travelled to the box.
Bars (BASELINE.json north_star): primary closest-hit ids bit-exact (exact ties -> lowest id), t within 1e-5 relative
(observed: bit-exact), RNG stream bit-exact, path-traced images by RMSE (the hardware cubemap filter and device
sin/cos differ from the host restatement in the last ulps)."""
import ctypes as C
import os

import numpy as np
import pytest

import gpu_util
import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev(rt, torch_cuda):
    return gpu_util.DeviceScene(rt, torch_cuda, 320, 240, "default", seed=1234)


def golden(name):
    return np.load(os.path.join(util.GOLDEN, name))


def check_ids(ids, ts, gid_low, gt, gid_first=None):
    """Bit-exact bar: t equals the reference's t bit for bit, ids equal the golden ids under BASELINE.json's rule (min t,
    exact ties -> lowest id; produced by oracle/rt_port.c after it was pinned to the unmodified reference functions).
    Where the reference's own first-visited id differs, it can only be a HIGHER id at the same t."""
    assert np.array_equal(ts.view(np.uint32), gt.view(np.uint32)), "t differs: %d rays" % int((ts.view(np.uint32) != gt.view(np.uint32)).sum())
    assert np.array_equal(ids, gid_low), "ids differ: %d rays" % int((ids != gid_low).sum())
    if gid_first is not None:
        moved = ids != gid_first
        assert (ids[moved] < gid_first[moved]).all()
        return int(moved.sum())
    return 0


# ------------------------------------------------------------------ RNG
def test_rng_known_answers_bit_exact(rt, torch_cuda):
    t = torch_cuda
    g = golden("rng_kat.npz")
    for i, seed in enumerate(g["seeds"]):
        for j, sub in enumerate(g["subsequences"]):
            st = t.zeros(48, dtype=t.uint8, device="cuda")
            out = t.zeros(8, dtype=t.int32, device="cuda")
            rt.check(rt.lib().rt_init_rng(rt.ptr(st), 1, int(seed), int(sub), None))
            rt.check(rt.lib().rt_rng_draw(rt.ptr(st), 1, 8, rt.ptr(out), None))
            assert np.array_equal(out.cpu().numpy().view(np.uint32), g["draws"][i, j]), (seed, sub)


def test_rng_init_matches_curand_init_for_every_pixel_stream(rt, torch_cuda, ref):
    """init_rng (the reference-named entry) over a whole padded frame == host curand_init for each subsequence."""
    t = torch_cuda
    n = gpu_util.rng_state_count(320, 240)
    st = t.zeros(n * 48, dtype=t.uint8, device="cuda")
    rt.lib().init_rng(n // 128, 128, rt.ptr(st), 0xDEADBEEF)
    t.cuda.synchronize()
    mine = st.cpu().numpy().reshape(n, 48)
    want = ref.rng_init(n, 0xDEADBEEF)
    assert np.array_equal(mine[:, :24], want[:, :24])  # d, v[5]
    assert not mine[:, 24:36].any()                    # Box-Muller flags/extra zeroed like curand_init


def test_rng_uniform_and_state_roundtrip(rt, torch_cuda):
    t = torch_cuda
    g = golden("rng_kat.npz")
    st = t.zeros(48 * 4, dtype=t.uint8, device="cuda")
    rt.check(rt.lib().rt_init_rng(rt.ptr(st), 4, 1234, 0, None))
    a = t.zeros(4 * 16, dtype=t.int32, device="cuda")
    # two draws of 8 == one draw of 16 (state persists in the 48-B curandState between launches)
    b = t.zeros(4 * 8, dtype=t.int32, device="cuda"); c = t.zeros(4 * 8, dtype=t.int32, device="cuda")
    st2 = st.clone()
    rt.check(rt.lib().rt_rng_draw(rt.ptr(st), 4, 16, rt.ptr(a), None))
    rt.check(rt.lib().rt_rng_draw(rt.ptr(st2), 4, 8, rt.ptr(b), None))
    rt.check(rt.lib().rt_rng_draw(rt.ptr(st2), 4, 8, rt.ptr(c), None))
    a = a.cpu().numpy().view(np.uint32).reshape(4, 16)
    assert np.array_equal(a[:, :8], b.cpu().numpy().view(np.uint32).reshape(4, 8))
    assert np.array_equal(a[:, 8:], c.cpu().numpy().view(np.uint32).reshape(4, 8))
    # curand_uniform = x * 2^-32 + 2^-33 (curand_uniform.h:69-72)
    uni = (a[2].astype(np.float32) * np.float32(2.3283064e-10) + np.float32(2.3283064e-10 / 2)).astype(np.float32)
    assert np.array_equal(uni, g["uniform_seed1234_sub2"])


# ------------------------------------------------------------------ closest hit
@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("pose", ["default", "poseB"])
@pytest.mark.parametrize("with_spheres", [0, 1])
def test_primary_hits_vs_reference_golden(dev, pose, with_spheres, variant):
    g = golden("primary_320x240.npz")
    dev.set_pose(pose)
    ids, ts = dev.trace_primary(with_spheres, variant)
    low = golden("primary_320x240_lowest_id.npz")
    check_ids(ids, ts, low[f"{pose}_s{with_spheres}_id"], g[f"{pose}_s{with_spheres}_t"], g[f"{pose}_s{with_spheres}_id"])
    hit = ids != 0xFFFFFFFF
    assert 0.3 < hit.mean() <= 1.0


@pytest.mark.parametrize("variant", [0, 1])
def test_incoherent_rays_vs_reference_golden(dev, variant):
    g = golden("rays_kat.npz")
    ids, ts = dev.trace_rays(g["rays"], True, variant)
    check_ids(ids, ts, golden("primary_320x240_lowest_id.npz")["rays_id"], g["t"], g["ids"])


def test_fresh_random_rays_vs_live_oracle(dev, ref):
    """Seeded rays that are NOT in the fixtures, through oracle/_ref's unmodified GetRayHit on the box's host (t, first-visited
    ids) and through the pinned plain-C port for the lowest-id rule."""
    import port_oracle as po
    rng = np.random.default_rng(20261104)
    n = 20000
    o = rng.uniform([15, -12, 5], [45, 12, 35], (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    rays = np.concatenate([o, d], 1)
    # axis-aligned directions (zero components: division by zero in the slab test, Math.h:50-61)
    rays[:64, 3:] = 0; rays[:64, 3 + (np.arange(64) % 3)] = np.where(np.arange(64) % 2, 1.0, -1.0)
    a = dev.scene.arrays()
    gid, gt = ref.trace_rays(util.raw(a["vertices"]), util.raw(a["faces"]), util.raw(a["nodes"]), a["face_indices"],
                             util.raw(a["spheres"]), rays, True)
    ps = po.PortScene({k: util.raw(a[k]) for k in ("vertices", "faces", "spheres", "materials")}, util.raw(a["nodes"]), a["face_indices"])
    pid0, pt0 = ps.trace_rays(rays, True, 0)
    assert np.array_equal(pid0, gid) and np.array_equal(pt0.view(np.uint32), gt.view(np.uint32))  # port pinned on these rays too
    low, _ = ps.trace_rays(rays, True, 1)
    for variant in (0, 1):
        ids, ts = dev.trace_rays(rays, True, variant)
        check_ids(ids, ts, low, gt, gid)


@pytest.mark.parametrize("count,size", [(1, 1.0), (2, 1.0), (3, 4.0), (60, 0.3), (3000, 1.0), (3000, 6.0), (40000, 0.2)])
def test_random_soups_vs_live_oracle(rt, torch_cuda, ref, count, size):
    """The product's own tree (SAH + reinsertion over the reference's leaves, with fat-leaf sub-trees where BVH.cpp could not split) on
    scenes that are not the bunny: triangle soups from one triangle to 40 000, sparse to everything-overlaps-everything. Product
    traversal and reference-order traversal against the unmodified reference's GetRayHit on the host: t bit for bit on every ray, ids
    only ever lower at an exact tie."""
    rng = np.random.default_rng(1000 + count + int(size * 10))
    scene = rt.Scene(None)
    m = scene.add_material(albedo=(0.5, 0.5, 0.5))
    c = rng.uniform(-10, 10, (count, 3)); e = rng.normal(size=(count, 3, 3)) * size
    for i in range(count):
        scene.add_triangle(c[i] + e[i, 0], c[i] + e[i, 1], c[i] + e[i, 2], m)
    scene.build_bvh()
    d = gpu_util.DeviceScene(rt, torch_cuda, 64, 48, ((0, 0, -30), 0.0, 0.0), env=False, scene=scene)
    n = 30000
    o = rng.uniform(-16, 16, (n, 3)).astype(np.float32)
    target = (c[rng.integers(0, count, n)] + rng.normal(size=(n, 3)) * size * 0.5).astype(np.float32)   # towards the triangles
    dirs = np.where(rng.uniform(size=(n, 1)) < 0.7, target - o, rng.normal(size=(n, 3))).astype(np.float32)
    dirs[::3] /= np.linalg.norm(dirs[::3], axis=1, keepdims=True)
    rays = np.concatenate([o, dirs], 1)
    rays[:64, 3:] = 0; rays[:64, 3 + (np.arange(64) % 3)] = np.where(np.arange(64) % 2, 1.0, -1.0)   # axis-parallel
    a = scene.arrays()
    gid, gt = ref.trace_rays(util.raw(a["vertices"]), util.raw(a["faces"]), util.raw(a["nodes"]), a["face_indices"], util.raw(a["spheres"]), rays, False)
    assert (gid != 0xFFFFFFFF).mean() > 0.03   # (4 % for the one-triangle scene, 98 % for the dense soup)
    for variant in (0, 1):
        ids, ts = d.trace_rays(rays, False, variant)
        assert np.array_equal(ts.view(np.uint32), gt.view(np.uint32)), (variant, int((ts.view(np.uint32) != gt.view(np.uint32)).sum()))
        moved = ids != gid
        assert (ids[moved] < gid[moved]).all() and moved.mean() < 0.02, (variant, float(moved.mean()))


@pytest.mark.parametrize("kind", [1, 2, 3, 4, 5])
def test_degenerate_soups_product_vs_reference_order(rt, torch_cuda, ref, kind):
    """The degenerate families of tests/test_layout_and_scene.py::test_relayout_fuzz_through_the_host_check on the device: identical
    centroids (one fat leaf), collinear triangles, everything in one axis-aligned plane (zero-thickness boxes - with rays that run
    exactly inside that plane, where the reference's box test divides 0 by 0), every triangle four times (exact ties), a scene 1 000
    times longer than wide. Product traversal against the reference-order traversal (device semantics of the reference's box test)
    on every ray, and against the unmodified reference on the host on the rays whose box tests stay free of NaN."""
    rng = np.random.default_rng(77 + kind)
    for n, scale, size in ((300, 1.0, 0.3), (3000, 30.0, 0.6), (40, 0.1, 0.03)):
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
        m = scene.add_material(albedo=(0.5, 0.5, 0.5))
        for i in range(len(c)):
            scene.add_triangle(c[i] + e[i, 0], c[i] + e[i, 1], c[i] + e[i, 2], m)
        scene.build_bvh()
        d = gpu_util.DeviceScene(rt, torch_cuda, 64, 48, ((0, 0, -30), 0.0, 0.0), env=False, scene=scene)
        k = 20000
        span = np.abs(c).max(0) + 2 * size + 1e-3
        o = (rng.uniform(-1.5, 1.5, (k, 3)) * span).astype(np.float32)
        target = (c[rng.integers(0, len(c), k)] + rng.normal(size=(k, 3)) * size * 0.3).astype(np.float32)
        dirs = (target - o).astype(np.float32)
        dirs[::2] /= np.maximum(np.linalg.norm(dirs[::2], axis=1, keepdims=True), 1e-30)
        rays = np.concatenate([o, dirs], 1)
        in_plane = np.zeros(k, bool)
        if kind == 3:   # rays inside the plane y = 0 of every box of this scene
            rays[:2000, 1] = 0; rays[:2000, 4] = 0; in_plane[:2000] = True
        i0, t0 = d.trace_rays(rays, False, 0)
        i1, t1 = d.trace_rays(rays, False, 1)
        a = scene.arrays()
        gid, gt = ref.trace_rays(util.raw(a["vertices"]), util.raw(a["faces"]), util.raw(a["nodes"]), a["face_indices"], util.raw(a["spheres"]), rays, False)
        ok = ~in_plane
        assert (gid[ok] != 0xFFFFFFFF).mean() > 0.02
        if kind != 3:
            assert np.array_equal(t1[ok].view(np.uint32), gt[ok].view(np.uint32))
            assert np.array_equal(t0.view(np.uint32), t1.view(np.uint32)), (n, int((t0.view(np.uint32) != t1.view(np.uint32)).sum()))
            moved = i0 != i1
            assert (i0[moved] < i1[moved]).all()
        else:
            # Overlapping COPLANAR triangles: a ray meets several at the same point, their computed t differ in the last ulps, and the
            # reference keeps whichever its own visiting order lets through its strict `tmin < best` cull (order
            # dependence on adversarial rays). The reference-order kernel of rt_trace_rays enters a box whose entry EQUALS the best (it needs
            # to, for the lowest-id tie rule) and the product returns the smallest t among all candidates: same hit / miss, each never
            # farther than the one before, all within 1e-5.
            assert ((i0 == 0xFFFFFFFF) == (i1 == 0xFFFFFFFF)).all() and ((i1[ok] == 0xFFFFFFFF) == (gid[ok] == 0xFFFFFFFF)).all()
            assert (t0 <= t1).all() and (t1[ok] <= gt[ok]).all()
            rel = np.abs(t0[ok] - gt[ok]) / np.maximum(gt[ok], 1e-30)
            assert rel.max() <= 1e-5, float(rel.max())
            assert (rel == 0).mean() > 0.7, float((rel == 0).mean())   # measured 0.84 on the densest of the three scenes
            assert (i1[in_plane] == 0xFFFFFFFF).all() and (i0[in_plane] == 0xFFFFFFFF).all()   # rays inside the plane see every triangle edge-on: det == 0


@pytest.mark.parametrize("mode", ["nan", "inf", "both"])
def test_soup_with_nan_and_infinite_vertices(rt, torch_cuda, mode):
    """Degenerate input as the domain has it: a soup in which some vertices are NaN or infinite. BVH.cpp's boxes then hold infinite
    bounds and need not nest; the re-layout keeps the reference's own topology for such a scene (accel_build.cpp) and stores an axis
    with a bound at infinity as unconstrained. The judge is the reference-order traversal on the device (variant 1: the reference's
    nodes, its box test with the DEVICE's fminf / fmaxf, which is what utils/CUDAHelper.h:32 gives the reference's own kernel; the
    host build of the same header uses std::min, so the host oracle answers differently as soon as a NaN enters a box test and is
    not consulted here): same t bit for bit, same hit / miss, ids only ever lower at the same t."""
    rng = np.random.default_rng(3)
    scene = rt.Scene(None)
    m = scene.add_material(albedo=(0.5, 0.5, 0.5))
    for i in range(3000):
        c = rng.uniform(-10, 10, 3); e = rng.normal(size=(3, 3))
        p = [c + e[k] for k in range(3)]
        if mode in ("nan", "both") and i % 97 == 0:
            p[1] = np.array([np.nan, 1.0, 2.0])
        if mode in ("inf", "both") and i % 101 == 0:
            p[2] = np.array([np.inf, 1.0, -np.inf])
        scene.add_triangle(*p, m)
    scene.build_bvh()
    d = gpu_util.DeviceScene(rt, torch_cuda, 64, 48, ((0, 0, -30), 0.0, 0.0), env=False, scene=scene)
    n = 20000
    o = rng.uniform(-14, 14, (n, 3)).astype(np.float32)
    dirs = rng.normal(size=(n, 3)).astype(np.float32)
    rays = np.concatenate([o, dirs], 1)
    rays[:64, 3:] = 0; rays[:64, 3 + (np.arange(64) % 3)] = np.where(np.arange(64) % 2, 1.0, -1.0)   # axis-parallel: 1 / 0 in the slab test
    i0, t0 = d.trace_rays(rays, False, 0)
    i1, t1 = d.trace_rays(rays, False, 1)
    assert (i1 != 0xFFFFFFFF).mean() > 0.05
    assert np.array_equal(t0.view(np.uint32), t1.view(np.uint32)), int((t0.view(np.uint32) != t1.view(np.uint32)).sum())
    moved = i0 != i1
    assert (i0[moved] < i1[moved]).all() and moved.mean() < 0.01


def edge_aimed_rays(scene_arrays, n=60000, seed=7):
    """Rays aimed exactly AT bunny vertices and at points ON bunny edges (half of them a hair beside), half unit-length."""
    a = scene_arrays
    rng = np.random.default_rng(seed)
    faces = a["faces"][12:-2:2]  # the smooth copy of every bunny face (Scene.cpp:110-127)
    pick = faces[rng.integers(0, len(faces), n)]
    pos = a["vertices"]["position"]
    p0, p1, p2 = pos[pick["v0"]], pos[pick["v1"]], pos[pick["v2"]]
    w = rng.uniform(0, 1, (n, 1)).astype(np.float32)
    kind = rng.integers(0, 4, n)[:, None]
    target = np.where(kind == 0, p0, np.where(kind == 1, p0 + w * (p1 - p0), np.where(kind == 2, p1 + w * (p2 - p1), p2 + w * (p0 - p2))))
    jitter = np.where(rng.uniform(size=(n, 1)) < 0.5, 0.0, 1.0) * rng.normal(scale=2e-6, size=(n, 3)) * np.abs(target)
    target = (target + jitter).astype(np.float32)
    o = rng.uniform([10, -12, 0], [50, 14, 40], (n, 3)).astype(np.float32)
    rays = np.concatenate([o, (target - o).astype(np.float32)], 1)
    rays[::2, 3:] /= np.linalg.norm(rays[::2, 3:], axis=1, keepdims=True)
    return rays


def test_rays_aimed_at_edges_and_vertices(dev):
    """Adversarial rays for the order-independent hit rule: where several faces meet, their t differ in the
    last ulps and the hit point lies ON leaf-box planes, so whether the reference still enters the second leaf (`tmin < best`,
    Math.h:50-61 against the CURRENT best) depends on its own visiting order; the near-first product order culls against a
    different current best. Both outcomes are the reference algorithm's under some visiting order. They differ in the last ulps
    of t, or - when one of the faces around the vertex is seen edge-on, where Moeller-Trumbore's t is ill-conditioned (error
    ~ eps |dist| |e| / det) and disagrees with its own leaf box's entry distance - by up to ~2e-4 relative.
    Measured on 200 000 such rays (benchmarks/edge_rays.py, profiles/r01g_edge_rays.json): 0.18 % differ in t (97 % of those toward
    the CLOSER hit in the product), 0.11 % same t but another face of the exact tie.
    Bars: same hit/miss; bit-identical t and id on > 99.5 % of even these rays; the north star's 1e-5 relative on > 99.95 %;
    never more than 1e-3."""
    rays = edge_aimed_rays(dev.scene.arrays())
    i0, t0 = dev.trace_rays(rays, True, 0)
    i1, t1 = dev.trace_rays(rays, True, 1)
    assert (i0 != 0xFFFFFFFF).mean() > 0.9
    assert ((i0 == 0xFFFFFFFF) == (i1 == 0xFFFFFFFF)).all()
    same_t = t0.view(np.uint32) == t1.view(np.uint32)
    assert same_t.mean() > 0.995 and (same_t & (i0 == i1)).mean() > 0.995
    rel = np.abs(t0 - t1) / np.maximum(t1, 1e-30)
    assert (rel <= 1e-5).mean() > 0.9995 and rel.max() < 1e-3, float(rel.max())


def test_product_kernel_equals_reference_order_kernel_at_1080p(rt, torch_cuda):
    """Size-independent property at BASELINE's full resolution: the near-first wide traversal returns exactly what
    the reference-order traversal returns (ids and t bit for bit), both poses."""
    d = gpu_util.DeviceScene(rt, torch_cuda, 1920, 1080, "poseB", env=False)
    for pose in ("poseB", "default"):
        d.set_pose(pose)
        i0, t0 = d.trace_primary(1, 0)
        i1, t1 = d.trace_primary(1, 1)
        assert np.array_equal(i0, i1) and np.array_equal(t0.view(np.uint32), t1.view(np.uint32))
    assert d.rays() == 4 * 1920 * 1080


def test_edge_cases(rt, torch_cuda):
    t = torch_cuda
    # one-triangle scene, no spheres, width/height not multiples of the 16x16 tile
    scene = rt.Scene()
    scene.add_triangle((-1, -1, 3), (1, -1, 3), (0, 1, 3), scene.add_material(albedo=(0.5, 0.5, 0.5)))
    scene.build_bvh()
    d = gpu_util.DeviceScene(rt, t, 37, 23, ((0, 0, 0), 0.0, 180.0), env=False, scene=scene)
    for variant in (0, 1):
        ids, ts = d.trace_primary(0, variant)
        assert set(np.unique(ids)) == {0, 0xFFFFFFFF}
        assert (ts[ids == 0] >= 3).all() and (ts[ids == 0] < 6).all() and (ts[ids == 0xFFFFFFFF] == np.float32(1e30)).all()
    # zero rays is a no-op, not an error
    ids, ts = d.trace_rays(np.zeros((0, 6), np.float32), False, 0)
    assert len(ids) == 0
    # argument errors come back as codes, never exits (utils/CUDAHelper.h:7-22 exits; this layer must not)
    assert rt.lib().rt_trace_primary(None, d.gpu, 4, 4, 0, 0, None, None, None, None) == -1
    assert rt.lib().rt_trace_primary(d.accel, d.gpu, 4, 4, 0, 7, rt.ptr(d.counter), rt.ptr(d.counter), None, None) == -1
    assert rt.lib().rt_process_ex(None) == -1
    assert b"bad argument" in rt.lib().rt_last_error()
    # a path-traced frame of the one-triangle scene: finite, alpha 1
    d.process(0)
    s = d.surface.cpu().numpy()
    assert np.isfinite(s).all() and (s[..., 3] == 1).all()


# ------------------------------------------------------------------ path tracing
def rel_rmse(a, b):
    return float(np.sqrt(np.mean((a[..., :3] - b[..., :3]) ** 2)) / max(np.sqrt(np.mean(b[..., :3] ** 2)), 1e-12))


def test_path_trace_frames_vs_reference_golden(rt, torch_cuda):
    """Same seed, same per-pixel stream, same fp32 expression order => the frames track the host-run reference kernel
    body pixel by pixel. Residual differences: device sinf/cosf vs glibc, hardware cubemap filter vs cubemap_cpu.h."""
    g = golden("pt_96x72.npz")
    d = gpu_util.DeviceScene(rt, torch_cuda, 96, 72, "poseB", seed=1234, env=True)
    measured = []
    for f in range(3):
        d.process(f, write_last=1)
        got = d.surface.cpu().numpy()
        want = g["frames"][f]
        assert (got[..., 3] == 1).all()
        close = np.isclose(got[..., :3], want[..., :3], rtol=2e-2, atol=2e-3).all(axis=-1)
        measured.append((f, float(close.mean()), rel_rmse(got, want)))
        # frame 0 measured 0.994 / 0.6 %; a pixel whose path length differs once (device sinf/cosf vs libm at a grazing hit) has
        # another RNG stream from then on, so the per-pixel agreement of the running mean decays a little with every frame
        assert np.array_equal(d.last.cpu().numpy(), got)  # folded D2D copy (RayTracing.cpp:233)
    # the per-pixel XORWOW streams advanced by the same number of draws as the reference's (path lengths equal)
    st = d.rng.cpu().numpy().reshape(-1, 48)[: 96 * 72, :24]
    same = (st == g["rng_v"]).all(axis=1).mean()
    os.makedirs(os.path.join(util.ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(util.ROOT, "gpurun_out", "golden_frames_measured.txt"), "w") as fh:
        fh.write("frame, pixels within 2e-2, rel RMSE: %s; identical RNG streams after 3 frames: %.4f\n" % (measured, same))
    for f, c, r in measured:
        # measured (gpurun_out/golden_frames_measured.txt): 0.9944 / 0.9906 / 0.9893 of the pixels, RMSE 2.9 / 2.5 / 1.9 % (a few
        # bright pixels whose path took another branch carry the RMSE; the per-pixel bar is the sharp one)
        assert c > 0.99 - 0.004 * f and r < 0.045, measured
    assert same > 0.997, same  # measured 0.9991


def test_reference_named_entry_and_accumulation_semantics(rt, torch_cuda):
    t = torch_cuda
    d = gpu_util.DeviceScene(rt, t, 64, 48, "poseB", seed=7, env=True)
    # raytracing_process (main_raytracing.cu:202): 5 spp, 6 bounces, lerp(last, mean, 1/(frame+1)), alpha = 1
    rt.lib().raytracing_process(rt.ptr(d.surface), rt.ptr(d.last), 64, 48, d.pitch, 0, d.gpu)
    t.cuda.synchronize()
    f0 = d.surface.clone()
    assert (f0[..., 3] == 1).all() and t.isfinite(f0).all()
    assert (d.last == 0).all()  # the reference-named entry leaves the last-frame copy to its caller (RayTracing.cpp:233)
    d2 = gpu_util.DeviceScene(rt, t, 64, 48, "poseB", seed=7, env=True)
    d2.process(0, spp=5, bounces=6)
    assert t.equal(d2.surface, f0)  # same seed -> bit-identical frame through either entry
    # frame 1 = lerp(frame0, estimate1, 1/2): run the same estimate into an empty buffer and blend on the host
    d.last.copy_(d.surface)
    rng_before = d.rng.clone()
    rt.lib().raytracing_process(rt.ptr(d.surface), rt.ptr(d.last), 64, 48, d.pitch, 1, d.gpu)
    t.cuda.synchronize()
    d.rng.copy_(rng_before)
    est = t.zeros_like(d.surface)
    d.process(0, surface=est, last=t.zeros_like(est))
    a = np.float32(0.5)
    want = f0.cpu().numpy() * (np.float32(1) - a) + est.cpu().numpy() * a
    want[..., 3] = 1
    assert np.array_equal(d.surface.cpu().numpy(), want)
    # blend_mode 1: sum buffer (+ sample count in alpha) for the multi-GPU split
    d.rng.copy_(rng_before)
    acc = t.zeros_like(d.surface)
    d.process(0, blend_mode=1, surface=acc, last=None)
    got = acc.cpu().numpy()
    assert (got[..., 3] == 5).all()
    assert np.allclose(got[..., :3] / 5, est.cpu().numpy()[..., :3], rtol=1e-6, atol=1e-7)


def test_cost_ordered_tile_schedule_does_not_change_the_frame(rt, torch_cuda):
    """With a workspace the megakernel schedules tiles heaviest-first from the previous launch's measured costs (pathtrace.cu).
    Pixels are independent: three progressive frames with and without it are bit-identical in both surfaces and in every pixel's
    RNG state; the order the third launch used is a permutation of the tiles, and a frame-size change resets the history."""
    import ctypes as C
    W, H = 200, 120
    a = gpu_util.DeviceScene(rt, torch_cuda, W, H, "poseB", seed=77, env=True)
    b = gpu_util.DeviceScene(rt, torch_cuda, W, H, "poseB", seed=77, env=True)
    for k in range(3):
        a.process(k, spp=4, bounces=4, write_last=1, ordered=True)
        b.process(k, spp=4, bounces=4, write_last=1, ordered=False)
    torch_cuda.cuda.synchronize()
    assert torch_cuda.equal(a.surface, b.surface) and torch_cuda.equal(a.last, b.last) and torch_cuda.equal(a.rng, b.rng)
    assert a.rays() == b.rays()
    ws = a.workspace.view(torch_cuda.int32).cpu().numpy().view(np.uint32)
    n = ((W + 15) // 16) * ((H + 7) // 8)
    assert tuple(ws[:3]) == (W, H, n) and ws[3] == 2  # two launches found a cost history
    order, cost = ws[8:8 + n], ws[8 + n:8 + 2 * n]
    assert np.array_equal(np.sort(order), np.arange(n, dtype=np.uint32))
    assert (cost > 0).all()  # every tile ran and reported
    prev = b.workspace  # the other scene's workspace was never used
    assert int(prev.view(torch_cuda.int32).abs().sum().item()) == 0
    assert rt.lib().rt_process_workspace_bytes(W, H, 0) == (8 + 3 * n) * 4


def test_tile_schedule_under_a_moving_camera_does_not_change_the_frames(rt, torch_cuda):
    """After a camera move the schedule dilates the previous pose's tile costs (pathtrace.cu tile_order_kernel): still only an order
    of independent pixels - frames and RNG states equal the unordered run's at every pose, the workspace tracks the camera."""
    W, H = 200, 120
    a = gpu_util.DeviceScene(rt, torch_cuda, W, H, "poseB", seed=5, env=True)
    b = gpu_util.DeviceScene(rt, torch_cuda, W, H, "poseB", seed=5, env=True)
    hashes = set()
    for k, pose in enumerate((util.POSES["poseB"], ((30.0, -1.0, 3.0), 1.0, 184.0), ((27.0, 0.0, 5.0), 3.0, 195.0), ((27.0, 0.0, 5.0), 3.0, 195.0))):
        a.set_pose(pose); b.set_pose(pose)
        a.process(0, spp=4, bounces=4, write_last=1, ordered=True)
        b.process(0, spp=4, bounces=4, write_last=1, ordered=False)
        torch_cuda.cuda.synchronize()
        assert torch_cuda.equal(a.surface, b.surface) and torch_cuda.equal(a.rng, b.rng), k
        ws = a.workspace.view(torch_cuda.int32).cpu().numpy().view(np.uint32)
        n = ((W + 15) // 16) * ((H + 7) // 8)
        assert np.array_equal(np.sort(ws[8:8 + n]), np.arange(n, dtype=np.uint32))  # still a permutation of the tiles
        hashes.add(int(ws[4]))
    assert len(hashes) == 3  # three distinct cameras seen, the repeated pose kept its hash


def test_sample_split_is_stream_continuous(rt, torch_cuda):
    """One launch of 10 spp == two launches of 5 spp (sum buffers): the per-pixel stream is continuous across launches."""
    t = torch_cuda
    a = gpu_util.DeviceScene(rt, t, 64, 48, "poseB", seed=99, env=True)
    b = gpu_util.DeviceScene(rt, t, 64, 48, "poseB", seed=99, env=True)
    sa = t.zeros_like(a.surface); sb = t.zeros_like(a.surface)
    a.process(0, spp=10, bounces=4, blend_mode=1, surface=sa)
    b.process(0, spp=5, bounces=4, blend_mode=1, surface=sb)
    b.process(1, spp=5, bounces=4, blend_mode=1, surface=sb)
    assert t.equal(a.rng, b.rng)
    assert t.allclose(sa, sb, rtol=1e-5, atol=1e-6)
    assert a.rays() == b.rays() > 64 * 48 * 10


def test_tracer_facade_process_and_resize(rt, torch_cuda):
    """CUDARayTracer mirror (RayTracing.h:12-45): Process() advances frame_index, OnResize restarts accumulation."""
    t = torch_cuda
    a = util.assets()
    s = t.zeros((48, 64, 4), dtype=t.float32, device="cuda"); l = t.zeros_like(s)
    tr = rt.Tracer(s, l, 64, 48, 64 * 16, bunny_obj=a["obj"], env_dds=a["dds"], seed=5)
    tr.scene.set_camera(*util.POSES["poseB"], 64, 48)
    tr.process(); tr.process()
    t.cuda.synchronize()
    assert tr.frame_index == 2 and t.equal(s, l) and (s[..., 3] == 1).all()
    assert tr.ray_count > 2 * 5 * 64 * 48
    s2 = t.zeros((24, 32, 4), dtype=t.float32, device="cuda"); l2 = t.zeros_like(s2)
    tr.on_resize(s2, l2, 32, 24, 32 * 16)
    tr.process()
    t.cuda.synchronize()
    assert tr.frame_index == 1 and (s2[..., 3] == 1).all() and float(s2[..., :3].mean()) > 0


def test_checkpoint_resume_continues_bit_for_bit(rt, torch_cuda, tmp_path):
    """{frame index, surfaces, RNG states} saved after 3 frames; a NEW tracer that loads the file and renders 2
    more frames ends on exactly the image (and RNG stream position) of the uninterrupted 5-frame run."""
    t = torch_cuda
    a = util.assets()
    W, H = 80, 56

    def make():
        s = t.zeros((H, W, 4), dtype=t.float32, device="cuda"); l = t.zeros_like(s)
        tr = rt.Tracer(s, l, W, H, W * 16, bunny_obj=a["obj"], env_dds=a["dds"], seed=77, samples_per_launch=3, max_bounces=4)
        tr.scene.set_camera(*util.POSES["poseB"], W, H)
        return tr, s, l

    tr, s, l = make()
    for _ in range(3):
        tr.process()
    path = str(tmp_path / "run.ckpt")
    tr.save_checkpoint(path)
    for _ in range(2):
        tr.process()
    t.cuda.synchronize()
    full = s.clone()
    assert tr.frame_index == 5

    tr2, s2, l2 = make()
    tr2.load_checkpoint(path)
    assert tr2.frame_index == 3
    for _ in range(2):
        tr2.process()
    t.cuda.synchronize()
    assert tr2.frame_index == 5
    assert t.equal(s2, full) and t.equal(l2, full)
    # a third run without the checkpoint differs after 2 frames (the test would otherwise pass vacuously)
    tr3, s3, _ = make()
    tr3.process(); tr3.process()
    t.cuda.synchronize()
    assert not t.equal(s3, full)
    with pytest.raises(RuntimeError):
        tr3.load_checkpoint(str(tmp_path / "missing.ckpt"))


def test_ambient_occlusion_and_tonemap(rt, torch_cuda):
    t = torch_cuda
    d = gpu_util.DeviceScene(rt, t, 160, 120, "poseB", seed=3, env=False)
    ao = t.zeros(160 * 120, dtype=t.float32, device="cuda")
    rt.check(rt.lib().rt_ambient_occlusion(d.accel, d.gpu, 160, 120, 16, rt.ptr(ao), rt.ptr(d.counter), None))
    v = ao.cpu().numpy()
    assert ((v >= 0) & (v <= 1)).all() and 0.05 < v.mean() < 0.99
    ids, _ = d.trace_primary(1, 0)
    hits = int((ids != 0xFFFFFFFF).sum())
    assert d.rays() == 160 * 120 + 16 * hits + 160 * 120  # AO: 1 primary/pixel + 16 per hit; then the id query
    # tonemap: exposure 0.5 + ACES + sRGB + vertical flip (main.cpp:65-94), against numpy
    surf = t.rand((120, 160, 4), dtype=t.float32, device="cuda") * 4
    out = t.zeros((120, 160, 4), dtype=t.uint8, device="cuda")
    rt.check(rt.lib().rt_tonemap_rgba8(rt.ptr(surf), 160, 120, 160 * 16, 1.0, rt.ptr(out), None))
    x = surf.cpu().numpy()[..., :3].astype(np.float64) * 0.5
    y = np.clip((x * (2.51 * x + 0.03)) / (x * (2.43 * x + 0.59) + 0.14), 0, 1)
    y = np.where(y <= 0.0031308, 12.92 * y, 1.055 * y ** (1 / 2.4) - 0.055)
    want = np.floor(y * 255 + 0.5)[::-1]
    assert np.abs(out.cpu().numpy()[..., :3].astype(np.float64) - want).max() <= 1


# ------------------------------------------------------------------ the reference's own CUDA kernel on the same GPU
def test_against_reference_cuda_kernel_on_device(rt, torch_cuda, ref):
    """The reference's UNMODIFIED main_raytracing.cu / Random.cu built for sm_100a (oracle/_ref/libref_gpu.so), fed the SAME
    GPUScene (reference-layout arrays, texture object) and the same seed: init_rng states identical bit for bit; frames agree
    pixel for pixel up to FMA contraction (the reference build contracts, this build does not) - this also pins the jitter
    draw order (x first) and the hardware cubemap fetch."""
    if not ref.gpu_available():
        pytest.skip("oracle/_ref/libref_gpu.so not built")
    t = torch_cuda
    g = ref.gpu_lib()
    W, H = 160, 120
    d = gpu_util.DeviceScene(rt, t, W, H, "poseB", seed=4321, env=True)
    mine = d.rng.clone()
    theirs = t.zeros_like(d.rng)
    g.init_rng(d.rng.numel() // 48 // 128, 128, rt.ptr(theirs), 4321)
    t.cuda.synchronize()
    assert t.equal(mine.view(-1, 48)[:, :24], theirs.view(-1, 48)[:, :24])
    rs = t.zeros((H, W, 4), dtype=t.float32, device="cuda"); rl = t.zeros_like(rs)
    frames = 4
    for f in range(frames):
        d.scene.gpu_scene().rng_state = theirs.data_ptr()
        g.raytracing_process(rt.ptr(rs), rt.ptr(rl), W, H, W * 16, f, d.gpu)
        t.cuda.synchronize()
        rl.copy_(rs)
        d.scene.gpu_scene().rng_state = d.rng.data_ptr()
        d.process(f, write_last=1)
        t.cuda.synchronize()
    a, b = d.surface.cpu().numpy(), rs.cpu().numpy()
    # Per-pixel identity is NOT expected: every bunny face exists twice (smooth + flat duplicate, Scene.cpp:110-127) and
    # which copy wins is decided in the last ulp of t, which FMA contraction changes - the reference's own GPU build and its
    # own host build already disagree there. Non-bunny pixels track closely; the whole image agrees statistically.
    close = np.isclose(a[..., :3], b[..., :3], rtol=2e-2, atol=2e-3).all(axis=-1).mean()
    blk = lambda x: x[..., :3].reshape(H // 8, 8, W // 8, 8, 3).mean(axis=(1, 3))
    block_rmse = float(np.sqrt(np.mean((blk(a) - blk(b)) ** 2)) / np.sqrt(np.mean(blk(b) ** 2)))
    mean_ratio = float(a[..., :3].mean() / b[..., :3].mean())
    same_stream = (d.rng.view(-1, 48)[: W * H, :24] == theirs.view(-1, 48)[: W * H, :24]).all(dim=1).float().mean().item()
    os.makedirs(os.path.join(util.ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(util.ROOT, "gpurun_out", "vs_reference_cuda_kernel.txt"), "w") as f:
        f.write(f"pixels within 2e-2: {close:.4f}  8x8-block rel RMSE: {block_rmse:.4f}  mean ratio: {mean_ratio:.4f}  "
                f"identical RNG streams after {frames} frames: {same_stream:.4f}\n")
    # bars = 3x the measured values (0.73 / 0.0147 / 1.002, profiles/r01a_vs_reference_cuda_kernel.txt)
    assert close > 0.70, close
    assert block_rmse < 0.045, block_rmse
    assert abs(mean_ratio - 1) < 0.01, mean_ratio


@pytest.mark.parametrize("size", [(64, 48), (37, 23), (200, 120)])
def test_regroup_variant_bit_identical_to_megakernel(rt, torch_cuda, size):
    """variant 1 (persistent, parked lanes shaded/refilled together by warp ballot, pixels pulled from a global counter) runs the
    same per-pixel sequence of draws and fp32 operations as variant 0: surfaces, RNG states and ray counts are identical."""
    t = torch_cuda
    W, H = size
    a = gpu_util.DeviceScene(rt, t, W, H, "poseB", seed=21, env=True)
    b = gpu_util.DeviceScene(rt, t, W, H, "poseB", seed=21, env=True)
    for f in range(2):
        a.process(f, spp=7, bounces=4, variant=0, write_last=1)
        b.process(f, spp=7, bounces=4, variant=1, write_last=1)
    t.cuda.synchronize()
    assert t.equal(a.surface, b.surface) and t.equal(a.last, b.last)
    assert t.equal(a.rng, b.rng)
    assert a.rays() == b.rays()
    sa = t.zeros_like(a.surface); sb = t.zeros_like(a.surface)
    a.process(0, spp=3, bounces=6, blend_mode=1, variant=0, surface=sa)
    b.process(0, spp=3, bounces=6, blend_mode=1, variant=1, surface=sb)
    assert t.equal(sa, sb)


@pytest.mark.parametrize("variant", [10])
def test_wavefront_variant_bit_identical_to_megakernel(rt, torch_cuda, variant):
    """variant 10 (queue-based wavefront: path state in HBM, ballot / prefix-sum compaction into camera + octant queues between
    stages, north-star item 3) runs the same per-pixel sequence of draws and fp32 operations as variant 0: surfaces, RNG states
    and ray counts are identical. Frame sizes that do not fill the tiles; running-mean and sum blend."""
    t = torch_cuda
    for (W, H) in ((200, 120), (37, 23)):
        a = gpu_util.DeviceScene(rt, t, W, H, "poseB", seed=21, env=True)
        b = gpu_util.DeviceScene(rt, t, W, H, "poseB", seed=21, env=True)
        for f in range(2):
            a.process(f, spp=7, bounces=4, variant=0, write_last=1)
            b.process(f, spp=7, bounces=4, variant=variant, write_last=1)
        t.cuda.synchronize()
        assert t.equal(a.surface, b.surface) and t.equal(a.last, b.last)
        assert t.equal(a.rng, b.rng)
        assert a.rays() == b.rays()
        sa = t.zeros_like(a.surface); sb = t.zeros_like(a.surface)
        a.process(0, spp=3, bounces=6, blend_mode=1, variant=0, surface=sa)
        b.process(0, spp=3, bounces=6, blend_mode=1, variant=variant, surface=sb)
        assert t.equal(sa, sb)


def test_converged_image_rmse_vs_reference_at_1024spp(rt, torch_cuda):
    """BASELINE.json's image bar: the path-traced image at ~1024 spp has RMSE <= 1 % of the reference's converged image.
    Reference = the pinned plain-C port of raytracing_kernel_main (bit-identical to oracle/_ref) run live on the box's host cores:
    205 progressive frames of 5 spp x 6 bounces (the reference's own frame protocol, RayTracing.cpp:205-235), same seed."""
    import port_oracle as po
    t = torch_cuda
    W, H, FRAMES = 128, 96, 205
    d = gpu_util.DeviceScene(rt, t, W, H, "poseB", seed=2024, env=True)
    for f in range(FRAMES):
        d.process(f, spp=5, bounces=6, write_last=1)
    got = d.surface.cpu().numpy()
    a = d.scene.arrays()
    cam, _ = d.scene.camera()
    ps = po.PortScene({k: util.raw(a[k]) for k in ("vertices", "faces", "spheres", "materials")}, util.raw(a["nodes"]), a["face_indices"],
                      cam.view(np.uint8).ravel().copy(), util.cubemap_texels())
    st = po.rng_init(gpu_util.rng_state_count(W, H), 2024)
    surf = np.zeros((H, W, 4), np.float32); last = np.zeros((H, W, 4), np.float32)
    for f in range(FRAMES):
        ps.render(st, surf, last, W, H, f)
        last[:] = surf
    rmse = rel_rmse(got, surf)
    with open(os.path.join(util.ROOT, "gpurun_out", "rmse_1024spp.txt"), "w") as fh:
        fh.write(f"{W}x{H}, {FRAMES * 5} spp, 6 bounces, pose B: relative RMSE vs reference = {rmse:.5f}\n")
    assert rmse <= 0.01, rmse
    assert (got[..., 3] == 1).all()
