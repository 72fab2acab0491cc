"""Error-bound check for the clear-pass filter in front of the exact reference-leaf box test (traverse.cuh exact_box,
RT_TW_BOX_PREFILTER, a round-2 candidate that is off in the shipped build): whenever the slab test evaluated with an approximate
reciprocal (any value within 2^-23 of the exact one, as PTX specifies for rcp.approx.f32) passes with the filter's margins, the
reference's own test (Math::IntersectAABB, RayTracing/Math.h:50-61: IEEE divisions) passes too. numpy float32 arithmetic is IEEE,
so the exact side is the reference's bit for bit; the rays are aimed at box corners, edges and faces so that the comparisons sit
on their thresholds, plus axis-parallel and tiny components."""
import numpy as np


def exact_test(lo, hi, o, d, length):
    with np.errstate(all="ignore"):
        t1 = (lo - o) / d
        t2 = (hi - o) / d
        tmin = np.fmax(np.fmax(np.fmin(t1[:, 0], t2[:, 0]), np.fmin(t1[:, 1], t2[:, 1])), np.fmin(t1[:, 2], t2[:, 2]))
        tmax = np.fmin(np.fmin(np.fmax(t1[:, 0], t2[:, 0]), np.fmax(t1[:, 1], t2[:, 1])), np.fmax(t1[:, 2], t2[:, 2]))
        return (tmax >= tmin) & (tmin < length) & (tmax > 0)


def filter_pass(lo, hi, o, d, length, rel):
    f = np.float32
    with np.errstate(all="ignore"):
        r = (f(1.0) / d).astype(np.float64) * (1.0 + rel)          # anything within 2^-23 of the exact reciprocal ...
        r = r.astype(np.float32)                                   # ... as an fp32 value
        r = np.where(np.abs(d) < f(1.1754944e-38), np.where(np.signbit(d), -np.inf, np.inf).astype(np.float32), r)  # .ftz
        a1 = ((lo - o) * r).astype(np.float32)
        a2 = ((hi - o) * r).astype(np.float32)
        tlo = np.fmax(np.fmax(np.fmin(a1[:, 0], a2[:, 0]), np.fmin(a1[:, 1], a2[:, 1])), np.fmin(a1[:, 2], a2[:, 2]))
        thi = np.fmin(np.fmin(np.fmax(a1[:, 0], a2[:, 0]), np.fmax(a1[:, 1], a2[:, 1])), np.fmax(a1[:, 2], a2[:, 2]))
        lo_up = (tlo.astype(np.float64) + 1.0e-6 * np.abs(tlo.astype(np.float64))).astype(np.float32) + f(1e-30)
        hi_dn = (thi.astype(np.float64) - 1.0e-6 * np.abs(thi.astype(np.float64))).astype(np.float32) - f(1e-30)
        return (hi_dn >= lo_up) & (lo_up < length) & (hi_dn > 0)


def test_filter_pass_implies_exact_pass():
    g = np.random.default_rng(11)
    n = 400_000
    total_pass = 0
    for scale in (1.0, 300.0, 1e-3):
        c = g.uniform(-20, 20, (n, 3)) * scale
        h = np.abs(g.normal(size=(n, 3))) * scale * g.choice([1e-3, 0.3, 5.0], (n, 1))
        h[g.random(n) < 0.2, g.integers(0, 3)] = 0.0                       # flat boxes (axis-aligned triangles)
        lo, hi = (c - h).astype(np.float32), (c + h).astype(np.float32)
        # targets on corners / edges / faces / inside
        w = g.choice([0.0, 1.0, 0.5], (n, 3), p=[0.4, 0.4, 0.2])
        inside = g.random(n) < 0.3
        w[inside] = g.random((int(inside.sum()), 3))
        target = lo + (hi - lo) * w.astype(np.float32)
        o = (target + g.normal(size=(n, 3)) * scale * g.choice([0.01, 1.0, 50.0], (n, 1))).astype(np.float32)
        d = (target - o).astype(np.float32)
        d /= np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-30).astype(np.float32)
        d = d.astype(np.float32)
        k = g.random(n)
        d[k < 0.05, g.integers(0, 3)] = 0.0                                # axis-parallel
        d[(k > 0.05) & (k < 0.08), g.integers(0, 3)] = np.float32(1e-41)   # denormal component
        # ray length: far, or right at the entry distance (the `tmin < length` comparison on its threshold)
        dist = np.linalg.norm((target - o).astype(np.float64), axis=1).astype(np.float32)
        length = np.where(g.random(n) < 0.5, np.float32(1e30), dist * g.choice([1.0, 1.0 + 2e-7, 1.0 - 2e-7, 0.5], n).astype(np.float32)).astype(np.float32)
        ex = exact_test(lo, hi, o, d, length)
        for rel in (0.0, 2.0 ** -23, -(2.0 ** -23)):
            fp = filter_pass(lo, hi, o, d, length, rel)
            assert not (fp & ~ex).any(), int((fp & ~ex).sum())
            total_pass += int(fp.sum())
        assert ex.mean() > 0.2                                             # the cases do exercise the passing side
    assert total_pass > 100_000                                            # and the filter fires on a useful share of them
