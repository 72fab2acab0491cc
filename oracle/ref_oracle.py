"""TEST INFRASTRUCTURE. ctypes binding of oracle/_ref/libref_oracle.so = the REFERENCE'S OWN code
compiled host-side (oracle/build_ref.sh, oracle/ref_shim/ref_api.cpp). Used only by tests/, smoke() and
bench.py's cpu_baseline / --impl reference legs - never by the product path."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libref_oracle.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        h = C.CDLL(LIB_PATH)
        h.ref_sizeof.restype = C.c_uint32; h.ref_sizeof.argtypes = [C.c_int]
        h.ref_num_threads.restype = C.c_int
        h.ref_scene_new.restype = C.c_void_p
        for name in ("ref_scene_free", "ref_scene_setup_cornell"):
            getattr(h, name).argtypes = [C.c_void_p]; getattr(h, name).restype = None
        h.ref_scene_setup_bunny.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
        h.ref_scene_add_mesh.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int]
        h.ref_scene_counts.argtypes = [C.c_void_p, C.c_void_p]
        h.ref_scene_copy.argtypes = [C.c_void_p] * 5
        h.ref_camera_update.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p]
        h.ref_bvh_build.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        h.ref_primary_hits.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                                       C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        h.ref_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                                     C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
        h.ref_rng_init.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32]
        h.ref_rng_init_one.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64]
        h.ref_rng_draw_u32.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        h.ref_rng_draw_uniform.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        h.ref_rng_point_on_sphere.argtypes = [C.c_void_p, C.c_void_p]
        h.ref_render_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                                       C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_int]
        h.ref_dds_info.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]; h.ref_dds_info.restype = C.c_int
        _lib = h
    return _lib


def dds_info(data, capacity=12 + 3 * 256):
    """The reference's own GetDDSSubresources on a DDS file image: (subresource count or < 0, info array) - layout of rt_dds_info."""
    buf = np.frombuffer(bytes(data), np.uint8).copy()
    info = np.zeros(capacity, np.uint64)
    n = lib().ref_dds_info(_p(buf), len(buf), _p(info), capacity)
    return n, info


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def bvh_build(vertices, faces):
    """Reference BVH::Calculate (verbatim). Returns (nodes[32 B], face_indices)."""
    nf = len(faces)
    nodes = np.zeros((max(2 * nf - 1, 1), 32), np.uint8)
    used = C.c_uint32()
    idx = np.zeros(nf, np.uint32)
    lib().ref_bvh_build(_p(vertices), len(vertices), _p(faces), nf, _p(nodes), C.byref(used), _p(idx))
    return nodes[: used.value].copy(), idx


def build_scene(mesh_pos=None, mesh_nrm=None, mesh_idx=None, cornell=True):
    """GLM-based assembly of the reference's fixed scene. Returns dict of raw byte arrays."""
    h = lib()
    s = h.ref_scene_new()
    try:
        if cornell:
            h.ref_scene_setup_cornell(s)
        if mesh_pos is not None:
            h.ref_scene_setup_bunny(s, _p(mesh_pos), _p(mesh_nrm), len(mesh_pos), _p(mesh_idx), len(mesh_idx))
        c = np.zeros(4, np.uint32)
        h.ref_scene_counts(s, _p(c))
        out = {"vertices": np.zeros((c[0], 32), np.uint8), "faces": np.zeros((c[1], 16), np.uint8),
               "spheres": np.zeros((c[2], 32), np.uint8), "materials": np.zeros((c[3], 64), np.uint8)}
        h.ref_scene_copy(s, _p(out["vertices"]), _p(out["faces"]), _p(out["spheres"]), _p(out["materials"]))
        return out
    finally:
        h.ref_scene_free(s)


def camera(position, angle_x, angle_y, width, height):
    out = np.zeros(60, np.uint8)
    lib().ref_camera_update(_p(np.asarray(position, np.float32)), angle_x, angle_y, float(width), float(height), _p(out))
    return out


def primary_hits(vertices, faces, nodes, face_indices, spheres, cam60, width, height, with_spheres):
    ids = np.zeros(width * height, np.uint32); ts = np.zeros(width * height, np.float32)
    lib().ref_primary_hits(_p(vertices), _p(faces), len(faces), _p(nodes), _p(face_indices), _p(spheres),
                           0 if spheres is None else len(spheres), _p(cam60), width, height, int(with_spheres), _p(ids), _p(ts))
    return ids, ts


def trace_rays(vertices, faces, nodes, face_indices, spheres, rays, with_spheres):
    rays = np.ascontiguousarray(rays, np.float32)
    n = len(rays)
    ids = np.zeros(n, np.uint32); ts = np.zeros(n, np.float32)
    lib().ref_trace_rays(_p(vertices), _p(faces), len(faces), _p(nodes), _p(face_indices), _p(spheres),
                         0 if spheres is None else len(spheres), _p(rays), n, int(with_spheres), _p(ids), _p(ts))
    return ids, ts


def rng_init(count, seed):
    states = np.zeros((count, 48), np.uint8)
    lib().ref_rng_init(_p(states), count, seed)
    return states


def rng_draw_u32(state48, n):
    out = np.zeros(n, np.uint32)
    lib().ref_rng_draw_u32(_p(state48), n, _p(out))
    return out


def render_frame(scene, nodes, face_indices, cam60, cubemap, rng_states, surface, last, width, height, frame_index,
                 row_begin=0, row_end=0):
    """One raytracing_kernel_main pass (5 spp, 6 bounces) on the host; surfaces are float32 [H, W, 4]."""
    lib().ref_render_frame(_p(scene["vertices"]), _p(scene["faces"]), _p(nodes), _p(face_indices), _p(scene["spheres"]),
                           len(scene["spheres"]), _p(scene["materials"]), len(scene["materials"]), _p(cam60),
                           _p(cubemap), 0 if cubemap is None else cubemap.shape[1], _p(rng_states), _p(surface), _p(last),
                           width, height, width * 16, frame_index, row_begin, row_end)


def bench_render(scene, nodes, face_indices, cam60, cubemap, rng_states, surface, last, width, height, frame_index,
                 row_begin, row_stride, spp, bounces):
    """CPU baseline: the reference kernel body (constants made runtime, ref_shim/ref_bench.cpp) over every
    `row_stride`-th row with all host threads. Returns (seconds, rays = GetRayHit calls, threads)."""
    h = lib()
    h.ref_bench_render.restype = C.c_double
    h.ref_bench_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                                   C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.c_void_p, C.c_void_p]
    rays = C.c_ulonglong(); threads = C.c_int()
    s = h.ref_bench_render(_p(scene["vertices"]), _p(scene["faces"]), _p(nodes), _p(face_indices), _p(scene["spheres"]),
                           len(scene["spheres"]), _p(scene["materials"]), len(scene["materials"]), _p(cam60),
                           _p(cubemap), 0 if cubemap is None else cubemap.shape[1], _p(rng_states), _p(surface), _p(last),
                           width, height, width * 16, frame_index, row_begin, row_stride, spp, bounces,
                           C.addressof(rays), C.addressof(threads))
    return s, rays.value, threads.value


GPU_LIB_PATH = os.path.join(HERE, "_ref", "libref_gpu.so")
_gpu_lib = None


def gpu_available():
    return os.path.exists(GPU_LIB_PATH)


def gpu_lib():
    """The reference's UNMODIFIED main_raytracing.cu + Random.cu built for sm_100a (oracle/build_ref.sh): its own
    extern "C" raytracing_process / init_rng. Comparator only (tests, bench 'reference_cuda_kernel' note)."""
    global _gpu_lib
    if _gpu_lib is None:
        h = C.CDLL(GPU_LIB_PATH, mode=os.RTLD_LOCAL | os.RTLD_NOW)
        h.raytracing_process.restype = None
        h.raytracing_process.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_void_p]
        h.init_rng.restype = None
        h.init_rng.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint]
        _gpu_lib = h
    return _gpu_lib
