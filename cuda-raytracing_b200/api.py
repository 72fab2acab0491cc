"""Thin Python handles over the host C wrapper (include/b200rt_host.h). Names follow the reference's
classes (Scene / Camera / CUDARayTracer, RayTracing/Scene.h, RayTracing/RayTracing.h)."""
import ctypes as C
import os

import numpy as np

from . import (BVH_NODE, CAMERA, FACE, MATERIAL, SPHERE, VERTEX, GPUScene, ProcessParams, TracerOptions, check, lib,
               ptr)


def _b(s):
    return None if s is None else str(s).encode()


def _safe_destroy(name, handle):
    """Destructors may run during interpreter shutdown, when module globals are already gone."""
    try:
        getattr(lib(), name)(handle)
    except Exception:
        pass


class Mesh:
    """Loader::LoadedMesh (OBJ + smooth normals) - the stand-in for Loader::ImportScene (utils/AssimpLoader.cpp:29)."""

    def __init__(self, obj_path, subdivide=0):
        h = C.c_void_p()
        check(lib().rt_host_mesh_load_obj_subdivided(_b(obj_path), subdivide, C.byref(h)))
        self.handle = h

    def __del__(self, _safe_destroy=_safe_destroy):   # bound at definition: module globals are None by the time exit-time destructors run
        if getattr(self, "handle", None):
            _safe_destroy("rt_host_mesh_destroy", self.handle)
            self.handle = None

    def counts(self):
        nv, nf = C.c_uint32(), C.c_uint32()
        check(lib().rt_host_mesh_counts(self.handle, C.byref(nv), C.byref(nf)))
        return nv.value, nf.value

    def arrays(self):
        nv, nf = self.counts()
        pos = np.empty((nv, 3), np.float32); nrm = np.empty((nv, 3), np.float32)
        idx = np.empty((nf, 3), np.uint32); xf = np.empty(16, np.float32)
        check(lib().rt_host_mesh_copy(self.handle, ptr(pos), ptr(nrm), ptr(idx), ptr(xf)))
        return pos, nrm, idx, xf


class Scene:
    """RayTracing::Scene. Host assembly works without a device; upload() needs one."""

    def __init__(self, env_dds=None, _borrowed=None):
        self._owned = _borrowed is None
        self.handle = C.c_void_p(lib().rt_host_scene_create(_b(env_dds))) if self._owned else C.c_void_p(_borrowed)
        if not self.handle:
            raise RuntimeError("rt_host_scene_create failed")

    def __del__(self, _safe_destroy=_safe_destroy):   # bound at definition: module globals are None by the time exit-time destructors run
        if getattr(self, "_owned", False) and getattr(self, "handle", None):
            _safe_destroy("rt_host_scene_destroy", self.handle)
            self.handle = None

    def setup_cornell_box(self):
        check(lib().rt_host_scene_setup_cornell_box(self.handle))

    def setup_stanford_bunny(self, mesh):
        check(lib().rt_host_scene_setup_stanford_bunny(self.handle, mesh.handle))

    def add_material(self, albedo=(1, 1, 1), emissive=(0, 0, 0), specular=(0, 0, 0, 0), roughness=0.9, specular_percent=0.0):
        m = np.zeros(1, MATERIAL)
        m["albedo"][0] = (*albedo, 1); m["emissive"][0] = (*emissive, 1); m["specular"][0] = specular
        m["roughness"] = roughness; m["specular_percent"] = specular_percent; m["IOR"] = 1.0
        return check(lib().rt_host_scene_add_material(self.handle, ptr(m)))

    def add_sphere(self, position, radius, material=0):
        check(lib().rt_host_scene_add_sphere(self.handle, ptr(np.asarray(position, np.float32)), radius, material))

    def add_triangle(self, a, b, c, material=0):
        f = lambda v: ptr(np.ascontiguousarray(v, np.float32))
        check(lib().rt_host_scene_add_triangle(self.handle, f(a), f(b), f(c), material))

    def add_loaded_scene(self, mesh, transform=None, material=0):
        t = None if transform is None else np.ascontiguousarray(np.asarray(transform, np.float32).T.reshape(16))
        check(lib().rt_host_scene_add_loaded_scene(self.handle, mesh.handle, ptr(t), material))

    def build_bvh(self):
        check(lib().rt_host_scene_build_bvh(self.handle))

    def counts(self):
        c = (C.c_uint32 * 6)()
        check(lib().rt_host_scene_counts(self.handle, c))
        return dict(zip(("vertices", "faces", "spheres", "materials", "nodes", "face_indices"), list(c)))

    def arrays(self):
        """Host copies of the reference-layout arrays (structured numpy)."""
        n = self.counts()
        out = {"vertices": np.zeros(n["vertices"], VERTEX), "faces": np.zeros(n["faces"], FACE),
               "spheres": np.zeros(n["spheres"], SPHERE), "materials": np.zeros(n["materials"], MATERIAL),
               "nodes": np.zeros(n["nodes"], BVH_NODE), "face_indices": np.zeros(n["face_indices"], np.uint32)}
        check(lib().rt_host_scene_copy(self.handle, ptr(out["vertices"]), ptr(out["faces"]), ptr(out["spheres"]),
                                       ptr(out["materials"]), ptr(out["nodes"]) if n["nodes"] else None,
                                       ptr(out["face_indices"]) if n["face_indices"] else None))
        return out

    def set_camera(self, position, angle_x, angle_y, width, height):
        p = None if position is None else np.asarray(position, np.float32)
        check(lib().rt_host_scene_set_camera(self.handle, ptr(p), angle_x, angle_y, float(width), float(height)))

    def camera(self):
        cam = np.zeros(1, CAMERA); ang = np.zeros(2, np.float32)
        check(lib().rt_host_scene_get_camera(self.handle, ptr(cam), ptr(ang)))
        return cam, ang

    def upload(self, d_rng_state=None):
        check(lib().rt_host_scene_upload(self.handle, ptr(d_rng_state)))

    def gpu_scene_ptr(self):
        return C.c_void_p(lib().rt_host_scene_gpu(self.handle))

    def gpu_scene(self):
        return GPUScene.from_address(lib().rt_host_scene_gpu(self.handle))

    def accel_ptr(self):
        return C.c_void_p(lib().rt_host_scene_accel(self.handle))

    def accel_info(self):
        """rt_accel_info of the re-layout Scene::Upload built (needs upload() first)."""
        return _accel_info(self.accel_ptr())


def _accel_info(handle):
    v = (C.c_uint64 * 8)()
    check(lib().rt_accel_info(handle, v))
    return {"wide_nodes": v[0], "slots": v[1], "ref_leaves": v[2], "fat_leaves": v[3],
            "wide_depth": v[4], "device_bytes": v[5] | (v[6] << 32), "ref_depth": v[7]}


class Accel:
    """RtAccel built directly from host arrays (rt_accel_create)."""

    def __init__(self, nodes, face_indices, faces, vertices, fat_leaf_threshold=0):
        h = C.c_void_p()
        check(lib().rt_accel_create(ptr(nodes), len(nodes), ptr(face_indices), ptr(faces), len(faces), ptr(vertices),
                                    len(vertices), fat_leaf_threshold, C.byref(h)))
        self.handle = h

    def info(self):
        return _accel_info(self.handle)

    def __del__(self, _safe_destroy=_safe_destroy):   # bound at definition: module globals are None by the time exit-time destructors run
        if getattr(self, "handle", None):
            _safe_destroy("rt_accel_destroy", self.handle)
            self.handle = None


class EnvMap:
    def __init__(self, dds_path):
        h = C.c_void_p()
        check(lib().rt_envmap_create_from_file(_b(dds_path), C.byref(h)))
        self.handle = h

    @property
    def texture(self):
        return lib().rt_envmap_texture(self.handle)

    def __del__(self, _safe_destroy=_safe_destroy):   # bound at definition: module globals are None by the time exit-time destructors run
        if getattr(self, "handle", None):
            _safe_destroy("rt_envmap_destroy", self.handle)
            self.handle = None


class Tracer:
    """RayTracing::CUDARayTracer over caller-owned surfaces (device pointers / torch tensors)."""

    def __init__(self, surface, last_frame_surface, width, height, pitch, bunny_obj=None, env_dds=None, cornell_box=True,
                 samples_per_launch=5, max_bounces=6, seed=0, kernel_variant=0):
        self._keep = (surface, last_frame_surface)
        o = TracerOptions(_b(bunny_obj), _b(env_dds), int(cornell_box), samples_per_launch, max_bounces, seed, kernel_variant)
        h = C.c_void_p()
        check(lib().rt_host_tracer_create(ptr(surface), ptr(last_frame_surface), width, height, pitch, C.byref(o), C.byref(h)))
        self.handle = h
        self.scene = Scene(_borrowed=lib().rt_host_tracer_scene(h))
        self.scene._owner = self  # the borrowed RtHostScene lives inside this tracer: whoever keeps the scene keeps the tracer alive

    def process(self):
        check(lib().rt_host_tracer_process(self.handle))

    def render_to_host(self, frames, host_pixels, host_pitch):
        """Headless e2e call: `frames` progressive frames, surface copied into HOST memory. Returns h2d bytes moved."""
        n = C.c_uint64()
        check(lib().rt_host_tracer_render_to_host(self.handle, frames, ptr(host_pixels), host_pitch, C.byref(n)))
        return n.value

    def save_checkpoint(self, path):
        """{frame index, both surfaces, RNG states} -> file (CUDARayTracer::SaveCheckpoint)."""
        check(lib().rt_host_tracer_save_checkpoint(self.handle, os.fsencode(path)))

    def load_checkpoint(self, path):
        check(lib().rt_host_tracer_load_checkpoint(self.handle, os.fsencode(path)))

    def on_resize(self, surface, last_frame_surface, width, height, pitch):
        self._keep = (surface, last_frame_surface)
        check(lib().rt_host_tracer_on_resize(self.handle, ptr(surface), ptr(last_frame_surface), width, height, pitch))

    def set_sampling(self, samples_per_launch, max_bounces, kernel_variant=0):
        check(lib().rt_host_tracer_set_sampling(self.handle, samples_per_launch, max_bounces, kernel_variant))

    @property
    def frame_index(self):
        return lib().rt_host_tracer_frame_index(self.handle)

    @property
    def ray_count(self):
        return lib().rt_host_tracer_ray_count(self.handle)

    def __del__(self, _safe_destroy=_safe_destroy):   # bound at definition: module globals are None by the time exit-time destructors run
        if getattr(self, "handle", None):
            _safe_destroy("rt_host_tracer_destroy", self.handle)
            self.handle = None


class MultiTracer:
    """RayTracing::MultiGPUTracer (host/MultiGPU.h): spp split over the devices of one node, one ncclReduce per frame, frame to host."""

    def __init__(self, width, height, bunny_obj=None, env_dds=None, cornell_box=True, max_bounces=6, seed=0, devices=None):
        o = TracerOptions(_b(bunny_obj), _b(env_dds), int(cornell_box), 0, max_bounces, seed, 0)
        dev = (C.c_int * len(devices))(*devices) if devices else None
        h = C.c_void_p()
        check(lib().rt_host_multi_tracer_create(width, height, C.byref(o), dev, len(devices) if devices else 0, C.byref(h)))
        self.handle = h
        self.width, self.height = width, height

    @property
    def device_count(self):
        return lib().rt_host_multi_tracer_device_count(self.handle)

    def set_camera(self, position, angle_x, angle_y):
        p = None if position is None else np.asarray(position, np.float32)
        check(lib().rt_host_multi_tracer_set_camera(self.handle, ptr(p), angle_x, angle_y))

    def reseed(self, seed):
        check(lib().rt_host_multi_tracer_reseed(self.handle, seed))

    def render(self, total_spp, spp_per_launch, host_pixels, host_pitch=None):
        """-> {render_ms, reduce_ms, total_ms}; host_pixels: numpy float32 [H, W, 4] or a (pinned) torch CPU tensor."""
        t = (C.c_double * 3)()
        check(lib().rt_host_multi_tracer_render(self.handle, total_spp, spp_per_launch, ptr(host_pixels), host_pitch or self.width * 16, t))
        return {"render_ms": t[0], "reduce_ms": t[1], "total_ms": t[2]}

    @property
    def ray_count(self):
        return lib().rt_host_multi_tracer_ray_count(self.handle)

    def __del__(self, _safe_destroy=_safe_destroy):   # bound at definition: module globals are None by the time exit-time destructors run
        if getattr(self, "handle", None):
            _safe_destroy("rt_host_multi_tracer_destroy", self.handle)
            self.handle = None


def process_ex(surface, last, width, height, pitch, frame_index, scene_ptr, accel_ptr, spp=5, bounces=6, blend_mode=0,
               variant=0, write_last=0, ray_counter=None, stream=None, work_counters=None, workspace=None):
    """workspace: zero-filled uint8 device tensor of process_workspace_bytes(width, height) (cost-ordered tile scheduling
    state kept between launches), or None = grid order."""
    ws_bytes = int(workspace.numel() * workspace.element_size()) if workspace is not None else 0
    p = ProcessParams(ptr(surface), ptr(last), width, height, pitch, frame_index, spp, bounces, blend_mode, variant,
                      write_last, scene_ptr, accel_ptr, ptr(ray_counter), ptr(stream), ptr(workspace), ws_bytes, ptr(work_counters))
    check(lib().rt_process_ex(C.byref(p)))


def process_workspace_bytes(width, height, variant=0):
    return int(lib().rt_process_workspace_bytes(width, height, variant))
