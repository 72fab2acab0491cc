"""-m gpu, collected last on purpose: rt_process_ex_static_scratch, the allocation-free variant of
rt_process_ex (include/b200rt.h; no counterpart in the reference). Its tile-schedule
scratch is a file-scope device array in csrc/pathtrace.cu instead of a caller-allocated workspace. Checked through the C ABI against
rt_process_ex on two calls in a row: surfaces, last-frame copies, RNG states and ray counts bit-identical (the tolerance the
variant-vs-variant tests of this suite use: none), device memory untouched, oversized frames rejected with a clear error."""
import ctypes as C

import numpy as np
import pytest

import gpu_util

pytestmark = pytest.mark.gpu


def process_static(rt, d, frame_index, spp, bounces, write_last=1, width=None, height=None):
    from cuda_raytracing_b200 import ProcessParams
    W, H = width or d.width, height or d.height
    p = ProcessParams(rt.ptr(d.surface), rt.ptr(d.last), W, H, W * 16, frame_index, spp, bounces, 0, 0, write_last,
                      d.gpu, d.accel, rt.ptr(d.counter), None, None, 0, None)
    return rt.lib().rt_process_ex_static_scratch(C.byref(p))


def test_static_scratch_entry_matches_rt_process_ex_on_two_calls_in_a_row(rt, torch_cuda):
    t = torch_cuda
    W, H = 200, 120
    a = gpu_util.DeviceScene(rt, t, W, H, "poseB", seed=77, env=True)   # existing entry point, caller-owned workspace
    b = gpu_util.DeviceScene(rt, t, W, H, "poseB", seed=77, env=True)   # new entry point, static scratch
    assert rt.lib().rt_process_static_scratch_bytes() == 1 << 20
    for k in range(2):  # two calls in a row: the second runs on the tile costs the first left in the shared array
        a.process(k, spp=4, bounces=4, write_last=1, ordered=True)
        rt.check(process_static(rt, b, k, 4, 4))
        t.cuda.synchronize()
        assert t.equal(a.surface, b.surface) and t.equal(a.last, b.last) and t.equal(a.rng, b.rng), k
        assert a.rays() == b.rays() > W * H * 4 * (k + 1)
    # a third call bracketed by the driver's free-memory figure (nothing of torch's runs in between): no device allocation
    t.cuda.synchronize()
    free_before, _ = t.cuda.mem_get_info()
    rt.check(process_static(rt, b, 2, 4, 4))
    t.cuda.synchronize()
    free_after, _ = t.cuda.mem_get_info()
    assert free_after == free_before
    a.process(2, spp=4, bounces=4, write_last=1, ordered=True)
    t.cuda.synchronize()
    assert t.equal(a.surface, b.surface) and t.equal(a.rng, b.rng)
    # a second, different frame size right after: the stale contents of the shared array (another size's order / costs) must not leak in
    c = gpu_util.DeviceScene(rt, t, 96, 72, "default", seed=5, env=True)
    e = gpu_util.DeviceScene(rt, t, 96, 72, "default", seed=5, env=True)
    for k in range(2):
        c.process(k, spp=3, bounces=6, write_last=1, ordered=False)
        rt.check(process_static(rt, e, k, 3, 6))
    t.cuda.synchronize()
    assert t.equal(c.surface, e.surface) and t.equal(c.rng, e.rng) and c.rays() == e.rays()
    assert float(e.surface[..., 3].min()) == 1.0 and bool(t.isfinite(e.surface).all())


def test_static_scratch_entry_rejects_what_it_cannot_hold(rt, torch_cuda):
    d = gpu_util.DeviceScene(rt, torch_cuda, 64, 48, "poseB", seed=1, env=False)
    # 16384 x 16384 = 2 097 152 tiles: beyond the 1 MiB array (87 378 tiles); rejected before anything is launched or touched
    assert process_static(rt, d, 0, 1, 1, width=16384, height=16384) == -1
    msg = rt.lib().rt_last_error().decode()
    assert "static array holds 1048576" in msg and "rt_process_ex" in msg
    from cuda_raytracing_b200 import ProcessParams
    p = ProcessParams(rt.ptr(d.surface), rt.ptr(d.last), 64, 48, 64 * 16, 0, 1, 1, 0, 1, 0, d.gpu, d.accel, None, None, None, 0, None)
    assert rt.lib().rt_process_ex_static_scratch(C.byref(p)) == -1  # variant 1 has no scratch of this kind
    assert rt.lib().rt_process_ex_static_scratch(None) == -1
    torch_cuda.cuda.synchronize()
