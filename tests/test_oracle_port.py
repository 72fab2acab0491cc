"""CPU tests that PIN the plain-C oracle (oracle/rt_port.c) against the reference's own code:
  * the committed golden vectors (made by oracle/_ref = unmodified BVHRayHit / GetRayHit / raytracing_kernel_main and the
    toolkit's cuRAND header, tests/golden/make_fixtures.py) - these run everywhere;
  * oracle/_ref live on fresh seeded inputs - these run where the prebuilt library exists.
Also checks the order-independent hit rule (min t, exact ties -> lowest id) the B200 kernels implement against the
reference's first-visited rule: t is always bit-identical; ids differ only on exact ties, toward the lower id."""
import os

import numpy as np
import pytest

import gpu_util
import util


@pytest.fixture(scope="module")
def port():
    import port_oracle
    port_oracle.lib()
    return port_oracle


@pytest.fixture(scope="module")
def pscene(port, bunny):
    a = bunny["arrays"]
    arrays = {k: util.raw(a[k]) for k in ("vertices", "faces", "spheres", "materials")}
    return port.PortScene(arrays, util.raw(a["nodes"]), a["face_indices"], None, util.cubemap_texels())


def cam60(bunny, pose, w, h):
    p, ax, ay = util.POSES[pose]
    bunny["scene"].set_camera(p, ax, ay, w, h)
    cam, _ = bunny["scene"].camera()
    return cam.view(np.uint8).ravel().copy()


def golden(name):
    return np.load(os.path.join(util.GOLDEN, name))


def test_port_rng_known_answers(port):
    g = golden("rng_kat.npz")
    st = np.zeros(48, np.uint8)
    for i, seed in enumerate(g["seeds"]):
        for j, sub in enumerate(g["subsequences"]):
            port.lib().port_rng_init(st.ctypes.data, int(seed), int(sub))
            assert np.array_equal(port.rng_draw(st, 8), g["draws"][i, j]), (seed, sub)
    st4 = port.rng_init(4, 1234)
    assert np.array_equal(port.rng_uniform(st4[2], 16), g["uniform_seed1234_sub2"])


def test_port_rng_vs_live_curand(port, ref):
    n = 4096
    assert np.array_equal(port.rng_init(n, 0xDEADBEEF)[:, :24], ref.rng_init(n, 0xDEADBEEF)[:, :24])
    a, b = port.rng_init(1, 77), ref.rng_init(1, 77)
    for _ in range(64):  # GetRandomPointOnSphere (Random.h:23-31), bitwise
        pa, pb = np.zeros(3, np.float32), np.zeros(3, np.float32)
        port.lib().port_rng_sphere(a.ctypes.data, pa.ctypes.data)
        ref.lib().ref_rng_point_on_sphere(b.ctypes.data, pb.ctypes.data)
        assert np.array_equal(pa.view(np.uint32), pb.view(np.uint32))


@pytest.mark.parametrize("pose", ["default", "poseB"])
@pytest.mark.parametrize("with_spheres", [0, 1])
def test_port_primary_hits_equal_reference_golden(bunny, pscene, pose, with_spheres):
    g = golden("primary_320x240.npz")
    pscene.set(cam60(bunny, pose, 320, 240))
    ids, ts = pscene.primary_hits(320, 240, with_spheres, 0)
    assert np.array_equal(ids, g[f"{pose}_s{with_spheres}_id"])
    assert np.array_equal(ts.view(np.uint32), g[f"{pose}_s{with_spheres}_t"].view(np.uint32))
    # order-independent rule: same t everywhere; ids move only to a LOWER id, only on exact ties
    ids1, ts1 = pscene.primary_hits(320, 240, with_spheres, 1)
    assert np.array_equal(ts1.view(np.uint32), ts.view(np.uint32))
    moved = ids1 != ids
    assert (ids1[moved] < ids[moved]).all()
    lo = golden("primary_320x240_lowest_id.npz")
    assert np.array_equal(ids1, lo[f"{pose}_s{with_spheres}_id"])


def test_port_incoherent_rays_equal_reference_golden(pscene):
    g = golden("rays_kat.npz")
    ids, ts = pscene.trace_rays(g["rays"], True, 0)
    assert np.array_equal(ids, g["ids"]) and np.array_equal(ts.view(np.uint32), g["t"].view(np.uint32))
    # lowest-id rule on arbitrary rays (incl. |d| < 1, where the reference's own visiting order decides what is found):
    # same t bit for bit, ids only move down
    ids1, ts1 = pscene.trace_rays(g["rays"], True, 1)
    assert np.array_equal(ts1.view(np.uint32), ts.view(np.uint32)) and (ids1 <= ids).all()
    assert np.array_equal(ids1, golden("primary_320x240_lowest_id.npz")["rays_id"])


def test_port_fresh_rays_vs_live_reference(bunny, pscene, ref):
    rng = np.random.default_rng(11)
    n = 30000
    o = rng.uniform([-12, -12, 0], [45, 12, 40], (n, 3)).astype(np.float32)
    d = (rng.normal(size=(n, 3)) * rng.uniform(0.3, 3, (n, 1))).astype(np.float32)
    rays = np.concatenate([o, d], 1)
    rays[:30, 3:] = 0; rays[:30, 3 + (np.arange(30) % 3)] = np.where(np.arange(30) % 2, 2.0, -0.5)  # zero components
    a = bunny["arrays"]
    for ws in (True, False):
        gid, gt = ref.trace_rays(util.raw(a["vertices"]), util.raw(a["faces"]), util.raw(a["nodes"]), a["face_indices"],
                                 util.raw(a["spheres"]), rays, ws)
        ids, ts = pscene.trace_rays(rays, ws, 0)
        assert np.array_equal(ids, gid) and np.array_equal(ts.view(np.uint32), gt.view(np.uint32))


def test_port_frames_equal_reference_kernel_body(port, bunny, pscene):
    """port_render == raytracing_kernel_main run by oracle/_ref (golden frames): surfaces and RNG states bit for bit."""
    g = golden("pt_96x72.npz")
    W, H = 96, 72
    pscene.set(cam60(bunny, "poseB", W, H))
    st = port.rng_init(gpu_util.rng_state_count(W, H), 1234)
    surf = np.zeros((H, W, 4), np.float32); surf.view(np.uint8)[:] = 1
    last = np.zeros((H, W, 4), np.float32)
    for f in range(3):
        pscene.render(st, surf, last, W, H, f)
        last[:] = surf
        assert np.array_equal(surf.view(np.uint32), g["frames"][f].view(np.uint32)), f
    assert np.array_equal(st[: W * H, :24], g["rng_v"])


def test_port_sum_mode_and_counters(port, bunny, pscene):
    W, H = 48, 36
    pscene.set(cam60(bunny, "poseB", W, H))
    st = port.rng_init(gpu_util.rng_state_count(W, H), 5)
    acc = np.zeros((H, W, 4), np.float32)
    cnt = np.zeros(3, np.uint64)
    pscene.render(st, acc, None, W, H, 0, spp=3, bounces=4, blend_mode=1, counters=cnt)
    assert (acc[..., 3] == 3).all() and cnt[0] >= 3 * W * H and cnt[1] > cnt[0] and cnt[2] > 0
    # empty row range is a no-op
    before = acc.copy()
    pscene.render(st, acc, None, W, H, 0, spp=3, bounces=4, blend_mode=1, row_begin=10, row_end=10)
    assert np.array_equal(before, acc)


def test_ao_port_sequence_and_stream(rt, bunny):
    """oracle/rt_port.c port_ambient_occlusion (BASELINE config 3) is a composition of pinned pieces (GetRayHit, the diffuse lobe,
    XORWOW): check what is its own - one primary query per pixel, 16 more per hit, two draws per AO ray and none on a miss
    (final states = the states 32 draws further), counts in sixteenths."""
    import port_oracle as po
    W, H, N = 64, 48, 16
    a = bunny["arrays"]
    scene = bunny["scene"]
    scene.set_camera(*util.POSES["poseB"], W, H)
    cam, _ = scene.camera()
    ps = po.PortScene({k: util.raw(a[k]) for k in ("vertices", "faces", "spheres", "materials")}, util.raw(a["nodes"]), a["face_indices"],
                      cam.view(np.uint8).ravel().copy())
    st = po.rng_init(W * H, 5)
    before = st.copy()
    ao, ids, rays = ps.ambient_occlusion(st, W, H, N, 0)
    pid, _ = ps.primary_hits(W, H, 1, 0)
    assert np.array_equal(ids, pid)
    hit = ids != 0xFFFFFFFF
    assert rays == W * H + N * int(hit.sum()) and 0.3 < hit.mean() < 1.0
    assert (ao[~hit] == 1).all() and np.array_equal((ao * N).round() / N, ao) and 0.05 < ao[hit].mean() < 0.99
    for i in (0, 7, int(np.flatnonzero(hit)[0]), int(np.flatnonzero(hit)[-1])):
        s = before[i].copy()
        if hit[i]:
            po.rng_draw(s, 2 * N)
        assert np.array_equal(s[:24], st[i, :24])
