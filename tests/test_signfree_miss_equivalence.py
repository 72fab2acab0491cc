"""The triangle test's miss decision in its sign-free form (traverse.cuh tri_test, RT_TW_TRI_SIGNFLIP: flip the sign bits of u, v,
u + v by det's and compare against |det|) is the SAME predicate as GLM's two sign cases (glm/gtx/intersect.inl:44-63) for every fp32
input - thresholds, signed zeros, infinities and NaNs included. Bit-level restatement of both in numpy; the CUDA code itself is held
to the reference by the GPU parity tests."""
import numpy as np

EPS = np.float32(1.1920928955078125e-07)


def glm_two_cases(det, u, v, uv):
    pos = det > EPS
    neg = det < -EPS
    miss_pos = (u < 0) | (u > det) | (v < 0) | (uv > det)
    miss_neg = (u > 0) | (u < det) | (v > 0) | (uv < det)
    return np.where(pos, miss_pos, np.where(neg, miss_neg, True))


def sign_free(det, u, v, uv):
    flip = det.view(np.uint32) & np.uint32(0x80000000)
    ad = np.abs(det)
    us, vs, uvs = ((x.view(np.uint32) ^ flip).view(np.float32) for x in (u, v, uv))
    return ~(ad > EPS) | ((us < 0) | (us > ad)) | ((vs < 0) | (uvs > ad))


def test_sign_free_form_is_the_same_predicate():
    g = np.random.default_rng(3)
    special = np.array([0.0, -0.0, EPS, -EPS, np.nextafter(EPS, np.float32(1)), np.nextafter(-EPS, np.float32(-1)), np.nextafter(EPS, np.float32(0)),
                        1.0, -1.0, 1e-45, -1e-45, 3e38, -3e38, np.inf, -np.inf, np.nan], np.float32)
    n = 1_000_000
    with np.errstate(all="ignore"):
        det = np.where(g.random(n) < 0.3, g.choice(special, n), (g.normal(size=n) * g.choice([1e-7, 1e-3, 1.0, 1e6], n))).astype(np.float32)
        # u, v around the thresholds 0 and det, or special
        def around(n):
            base = np.where(g.random(n) < 0.5, 0.0, det.astype(np.float64))
            x = (base + g.choice([0.0, 1.0, -1.0, 1e-3, -1e-3], n) * np.abs(det.astype(np.float64)) * g.choice([0.0, 1e-7, 0.5, 2.0], n)).astype(np.float32)
            return np.where(g.random(n) < 0.15, g.choice(special, n), x).astype(np.float32)
        u, v = around(n), around(n)
        uv = (u + v).astype(np.float32)
        a, b = glm_two_cases(det, u, v, uv), sign_free(det, u, v, uv)
    assert np.array_equal(a, b), int((a != b).sum())
    assert 0.05 < a.mean() < 0.995  # both outcomes occur
