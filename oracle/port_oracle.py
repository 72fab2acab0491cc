"""TEST INFRASTRUCTURE. ctypes binding of oracle/_build/librt_port.so = the plain-C restatement of the reference's
hot path (oracle/rt_port.c). Pinned bit for bit against oracle/_ref (the reference's own code) by
tests/test_oracle_port.py. Used only by tests/, smoke() and bench.py's cpu_baseline leg - never by the product path."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "librt_port.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(os.path.join(HERE, "rt_port.c")):
            subprocess.run(["bash", os.path.join(HERE, "build_port.sh")], check=True)
        h = C.CDLL(LIB_PATH)
        h.port_sizeof_scene.restype = C.c_uint32
        h.port_scene_fill.argtypes = [C.c_void_p] * 6 + [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        h.port_primary_hits.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        h.port_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        h.port_rng_init.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64]
        h.port_rng_init_many.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64]
        h.port_rng_draw.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        h.port_rng_uniform.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        h.port_rng_sphere.argtypes = [C.c_void_p, C.c_void_p]
        h.port_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_int,
                                  C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        h.port_ambient_occlusion.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = h
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class PortScene:
    """Holds the numpy arrays alive and the filled PScene struct. `arrays` = raw uint8 views of the reference-layout
    arrays (vertices, faces, spheres, materials), nodes [n,32] u8, face_indices u32, cam60 u8[60], cubemap f32[6,n,n,4]."""

    def __init__(self, arrays, nodes, face_indices, cam60=None, cubemap=None):
        self.keep = (arrays, nodes, face_indices, cam60, cubemap)
        self.buf = np.zeros(lib().port_sizeof_scene(), np.uint8)
        self.set(cam60)

    def set(self, cam60):
        arrays, nodes, face_indices, _, cubemap = self.keep
        self.cam = None if cam60 is None else np.ascontiguousarray(cam60)
        lib().port_scene_fill(_p(self.buf), _p(arrays["vertices"]), _p(arrays["faces"]), _p(nodes), _p(face_indices),
                              _p(arrays["spheres"]), len(arrays["spheres"]), _p(arrays["materials"]), _p(cubemap),
                              0 if cubemap is None else cubemap.shape[1], _p(self.cam))

    def primary_hits(self, width, height, with_spheres, tie_mode, work=False):
        ids = np.zeros(width * height, np.uint32); ts = np.zeros(width * height, np.float32)
        w = np.zeros((width * height, 2), np.uint32) if work else None
        lib().port_primary_hits(_p(self.buf), width, height, int(with_spheres), tie_mode, _p(ids), _p(ts), _p(w))
        return (ids, ts, w) if work else (ids, ts)

    def trace_rays(self, rays, with_spheres, tie_mode):
        rays = np.ascontiguousarray(rays, np.float32)
        ids = np.zeros(len(rays), np.uint32); ts = np.zeros(len(rays), np.float32)
        lib().port_trace_rays(_p(self.buf), _p(rays), len(rays), int(with_spheres), tie_mode, _p(ids), _p(ts))
        return ids, ts

    def ambient_occlusion(self, rng_states, width, height, ao_rays=16, tie_mode=0):
        """BASELINE config 3 on the host: (unoccluded fraction per pixel, primary hit ids, closest-hit queries issued)."""
        ao = np.zeros(width * height, np.float32); ids = np.zeros(width * height, np.uint32); n = np.zeros(1, np.uint64)
        lib().port_ambient_occlusion(_p(self.buf), _p(rng_states), width, height, ao_rays, tie_mode, _p(ao), _p(ids), _p(n))
        return ao, ids, int(n[0])

    def render(self, rng_states, surface, last, width, height, frame_index, spp=5, bounces=6, blend_mode=0, tie_mode=0,
               row_begin=0, row_end=0, row_stride=1, counters=None):
        lib().port_render(_p(self.buf), _p(rng_states), _p(surface), _p(last), width, height, width * 16, frame_index, spp, bounces,
                          blend_mode, tie_mode, row_begin, row_end, row_stride, _p(counters))


def rng_init(count, seed, first_subsequence=0):
    st = np.zeros((count, 48), np.uint8)
    lib().port_rng_init_many(_p(st), count, seed, first_subsequence)
    return st


def rng_draw(state48, n):
    out = np.zeros(n, np.uint32)
    lib().port_rng_draw(_p(state48), n, _p(out))
    return out


def rng_uniform(state48, n):
    out = np.zeros(n, np.float32)
    lib().port_rng_uniform(_p(state48), n, _p(out))
    return out
