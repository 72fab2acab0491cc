"""ctypes binding of libb200rt.so (include/b200rt.h, include/b200rt_host.h).

Python is orchestration only (tests, bench.py): every compute call goes through the C ABI into the
sm_100a kernels. There is no CPU fallback - if the library is missing or no device is present the
compute entry points raise.

The directory name has a hyphen, so load this file with ``load_package()`` from ``_loader.py`` at the
repo root (module name ``cuda_raytracing_b200``).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200RT_LIB") or os.path.join(_HERE, "lib", "libb200rt.so")  # B200RT_LIB: A/B builds (build.py --out=)

# numpy views of the reference-layout PODs (include/rt_types.h)
VERTEX = np.dtype([("position", "<f4", 3), ("normal", "<f4", 3), ("uv", "<f4", 2)])
FACE = np.dtype([("v0", "<u4"), ("v2", "<u4"), ("v1", "<u4"), ("material", "<u4")])
BVH_NODE = np.dtype([("min", "<f4", 3), ("max", "<f4", 3), ("first_index", "<u4"), ("prim_count", "<u4")])
SPHERE = np.dtype([("position", "<f4", 3), ("radius", "<f4"), ("material", "<i4"), ("_pad", "<u4", 3)])
MATERIAL = np.dtype([("albedo", "<f4", 4), ("emissive", "<f4", 4), ("specular", "<f4", 4), ("roughness", "<f4"),
                     ("specular_percent", "<f4"), ("IOR", "<f4"), ("_pad", "<u4")])
CAMERA = np.dtype([("origin", "<f4", 3), ("viewport_worldspace_size", "<f4", 2), ("aspect", "<f4"),
                   ("horizontal", "<f4", 3), ("vertical", "<f4", 3), ("lower_left_corner", "<f4", 3)])
RNG_STATE_BYTES = 48
MISS_ID = 0xFFFFFFFF
SPHERE_ID_BIT = 0x80000000
assert VERTEX.itemsize == 32 and FACE.itemsize == 16 and BVH_NODE.itemsize == 32
assert SPHERE.itemsize == 32 and MATERIAL.itemsize == 64 and CAMERA.itemsize == 60


class GPUScene(C.Structure):
    """RtGPUScene / the reference's GPUScene (136 B)."""
    _fields_ = [("gpu_spheres", C.c_void_p), ("gpu_materials", C.c_void_p), ("gpu_bvh_nodes", C.c_void_p),
                ("gpu_bvh_face_indices", C.c_void_p), ("gpu_vertices", C.c_void_p), ("gpu_faces", C.c_void_p),
                ("sphere_count", C.c_int32), ("material_count", C.c_int32), ("rng_state", C.c_void_p),
                ("environment_cubemap_tex", C.c_uint64), ("camera", C.c_uint8 * 60), ("_tail", C.c_uint8 * 4)]


class ProcessParams(C.Structure):
    """RtProcessParams."""
    _fields_ = [("surface", C.c_void_p), ("surface_last_frame", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32),
                ("pitch", C.c_uint64), ("frame_index", C.c_int32), ("samples_per_launch", C.c_int32),
                ("max_bounces", C.c_int32), ("blend_mode", C.c_int32), ("variant", C.c_int32),
                ("write_last_frame", C.c_int32), ("scene", C.c_void_p), ("accel", C.c_void_p),
                ("d_ray_counter", C.c_void_p), ("stream", C.c_void_p), ("workspace", C.c_void_p),
                ("workspace_bytes", C.c_uint64), ("d_work_counters", C.c_void_p)]


class TracerOptions(C.Structure):
    _fields_ = [("bunny_obj", C.c_char_p), ("env_dds", C.c_char_p), ("cornell_box", C.c_int32),
                ("samples_per_launch", C.c_int32), ("max_bounces", C.c_int32), ("seed", C.c_uint32),
                ("kernel_variant", C.c_int32)]


assert C.sizeof(GPUScene) == 136

# every symbol include/*.h declares: name -> (restype, argtypes)
_P, _I, _U32, _U64, _F = C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_float
SYMBOLS = {
    "rt_last_error": (C.c_char_p, []),
    "rt_device_count": (_I, []),
    "raytracing_process": (None, [_P, _P, _I, _I, C.c_size_t, _I, _P]),
    "init_rng": (None, [_U32, _U32, _P, C.c_uint]),
    "rt_init_rng": (_I, [_P, _U64, _U64, _U64, _P]),
    "rt_rng_draw": (_I, [_P, _U64, _U32, _P, _P]),
    "rt_accel_create": (_I, [_P, _U32, _P, _P, _U32, _P, _U32, _U32, C.POINTER(_P)]),
    "rt_accel_destroy": (_I, [_P]),
    "rt_accel_check_host": (_I, [_P, _U32, _P, _P, _U32, _P, _U32, _U32, C.POINTER(_U64)]),
    "rt_accel_info": (_I, [_P, C.POINTER(_U64)]),
    "rt_accel_bind": (_I, [_P, _P]),
    "rt_accel_unbind": (_I, [_P]),
    "rt_envmap_create_from_file": (_I, [C.c_char_p, C.POINTER(_P)]),
    "rt_envmap_create_from_memory": (_I, [_P, C.c_size_t, C.POINTER(_P)]),
    "rt_envmap_texture": (_U64, [_P]),
    "rt_envmap_destroy": (_I, [_P]),
    "rt_dds_read_level0": (_I, [C.c_char_p, _P, C.c_size_t]),
    "rt_trace_primary": (_I, [_P, _P, _I, _I, _I, _I, _P, _P, _P, _P]),
    "rt_trace_rays": (_I, [_P, _P, _P, _U64, _I, _I, _P, _P, _P, _P]),
    "rt_texture_create_from_file": (_I, [C.c_char_p, _I, C.POINTER(_P)]),
    "rt_texture_create_from_memory": (_I, [_P, C.c_size_t, _I, C.POINTER(_P)]),
    "rt_texture_fetch": (_I, [_P, _I, _P, C.c_float, _P, _P]),
    "rt_dds_info": (_I, [_P, C.c_size_t, C.POINTER(_U64), _I]),
    "rt_process_ex": (_I, [_P]),
    "rt_process_workspace_bytes": (_U64, [_I, _I, _I]),
    "rt_process_ex_static_scratch": (_I, [_P]),
    "rt_process_static_scratch_bytes": (_U64, []),
    "rt_ambient_occlusion": (_I, [_P, _P, _I, _I, _I, _P, _P, _P]),
    "rt_resolve_sum": (_I, [_P, _P, _I, _I, _U64, _U64, _P]),
    "rt_tonemap_rgba8": (_I, [_P, _I, _I, _U64, _F, _P, _P]),
    "rt_count_reference_work": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    # b200rt_host.h
    "rt_host_mesh_load_obj": (_I, [C.c_char_p, C.POINTER(_P)]),
    "rt_host_mesh_load_obj_subdivided": (_I, [C.c_char_p, _I, C.POINTER(_P)]),
    "rt_host_mesh_destroy": (None, [_P]),
    "rt_host_mesh_counts": (_I, [_P, C.POINTER(_U32), C.POINTER(_U32)]),
    "rt_host_mesh_copy": (_I, [_P, _P, _P, _P, _P]),
    "rt_host_scene_create": (_P, [C.c_char_p]),
    "rt_host_scene_destroy": (None, [_P]),
    "rt_host_scene_setup_cornell_box": (_I, [_P]),
    "rt_host_scene_setup_stanford_bunny": (_I, [_P, _P]),
    "rt_host_scene_add_material": (_I, [_P, _P]),
    "rt_host_scene_add_sphere": (_I, [_P, _P, _F, _I]),
    "rt_host_scene_add_triangle": (_I, [_P, _P, _P, _P, _I]),
    "rt_host_scene_add_loaded_scene": (_I, [_P, _P, _P, _I]),
    "rt_host_scene_build_bvh": (_I, [_P]),
    "rt_host_scene_counts": (_I, [_P, C.POINTER(_U32)]),
    "rt_host_scene_copy": (_I, [_P, _P, _P, _P, _P, _P, _P]),
    "rt_host_scene_set_camera": (_I, [_P, _P, _F, _F, _F, _F]),
    "rt_host_scene_get_camera": (_I, [_P, _P, _P]),
    "rt_host_scene_add_dirty_flags": (_I, [_P, _U32]),
    "rt_host_scene_upload": (_I, [_P, _P]),
    "rt_host_scene_gpu": (_P, [_P]),
    "rt_host_scene_accel": (_P, [_P]),
    "rt_host_tracer_create": (_I, [_P, _P, _I, _I, _U64, _P, C.POINTER(_P)]),
    "rt_host_tracer_destroy": (None, [_P]),
    "rt_host_tracer_process": (_I, [_P]),
    "rt_host_tracer_on_resize": (_I, [_P, _P, _P, _I, _I, _U64]),
    "rt_host_tracer_render_to_host": (_I, [_P, _I, _P, _U64, C.POINTER(_U64)]),
    "rt_host_tracer_scene": (_P, [_P]),
    "rt_host_tracer_frame_index": (_I, [_P]),
    "rt_host_tracer_set_sampling": (_I, [_P, _I, _I, _I]),
    "rt_host_tracer_ray_count": (C.c_ulonglong, [_P]),
    "rt_host_tracer_save_checkpoint": (_I, [_P, C.c_char_p]),
    "rt_host_tracer_load_checkpoint": (_I, [_P, C.c_char_p]),
    "rt_host_multi_tracer_create": (_I, [_I, _I, _P, _P, _I, C.POINTER(_P)]),
    "rt_host_multi_tracer_destroy": (None, [_P]),
    "rt_host_multi_tracer_device_count": (_I, [_P]),
    "rt_host_multi_tracer_set_camera": (_I, [_P, _P, _F, _F]),
    "rt_host_multi_tracer_reseed": (_I, [_P, _U32]),
    "rt_host_multi_tracer_render": (_I, [_P, _I, _I, _P, _U64, _P]),
    "rt_host_multi_tracer_ray_count": (C.c_ulonglong, [_P]),
    "rt_host_write_ppm": (_I, [C.c_char_p, _P, _I, _I]),
}

_lib = None


class RtError(RuntimeError):
    pass


def lib():
    """The loaded library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtError(f"{LIB_PATH} is missing - run __graft_entry__.build() (there is no CPU fallback)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(code):
    if code < 0:
        raise RtError(f"libb200rt error {code}: {lib().rt_last_error().decode()}")
    return code


def ptr(a):
    """void* of a numpy array (host) or an int device pointer / torch tensor."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    return C.c_void_p(int(a))


from .api import Mesh, Scene, Tracer, MultiTracer, Accel, EnvMap  # noqa: E402,F401
