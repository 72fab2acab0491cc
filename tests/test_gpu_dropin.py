"""-m gpu: libb200rt.so as a TRUE drop-in for an unmodified host (VERDICT r01 missing #5). The reference's host only ever does what
RayTracing/Scene.cpp:189-231 and RayTracing.cpp:213-233 do: cudaMalloc + cudaMemcpy of its six arrays, init_rng, then per frame
raytracing_process(surface, last, W, H, pitch, frame_index, &scene) and a D2D copy surface -> last. This test does exactly that
through the two reference-named symbols - plain cudaMalloc'ed arrays, a hand-filled 136-byte GPUScene, NO rt_accel_* / Scene::Upload
call - and expects the frames rt_process_ex renders for the same scene and seed, bit for bit. Then the host frees its arrays and
uploads ANOTHER scene (the allocator is free to hand the same addresses back): the library must render the new scene, not a stale
re-layout keyed by the old pointer (buffer-id check, csrc/dropin.cpp)."""
import ctypes as C
import glob
import os

import numpy as np
import pytest

import gpu_util
import util

pytestmark = pytest.mark.gpu


def cudart():
    import torch
    for pat in (os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*"),
                "/usr/local/cuda/lib64/libcudart.so*"):
        hits = sorted(glob.glob(pat))
        if hits:
            h = C.CDLL(hits[0])
            h.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]; h.cudaFree.argtypes = [C.c_void_p]
            h.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
            return h
    pytest.skip("no libcudart to play the host with")


class PlainHost:
    """What the reference's Scene::Upload leaves behind: six cudaMalloc'ed arrays and a GPUScene naming them."""

    def __init__(self, rt, rtlib, arrays, cam60, rng_ptr, env_tex=0, capacity=None):
        self.rtlib, self.ptrs = rtlib, []
        self.scene = rt.GPUScene()
        for field, key in (("gpu_spheres", "spheres"), ("gpu_materials", "materials"), ("gpu_bvh_nodes", "nodes"),
                           ("gpu_bvh_face_indices", "face_indices"), ("gpu_vertices", "vertices"), ("gpu_faces", "faces")):
            a = np.ascontiguousarray(arrays[key])
            p = C.c_void_p()
            # (capacity: a host that sizes its buffers generously - and makes the allocator hand the same addresses back on re-upload)
            assert rtlib.cudaMalloc(C.byref(p), max(a.nbytes, 16, (capacity or {}).get(key, 0))) == 0
            assert rtlib.cudaMemcpy(p, a.ctypes.data_as(C.c_void_p), a.nbytes, 1) == 0
            setattr(self.scene, field, p.value)
            self.ptrs.append(p)
        self.scene.sphere_count = len(arrays["spheres"]); self.scene.material_count = len(arrays["materials"])
        self.scene.rng_state = rng_ptr
        self.scene.environment_cubemap_tex = env_tex
        C.memmove(C.addressof(self.scene) + rt.GPUScene.camera.offset, cam60.ctypes.data, 60)
        self.addresses = [p.value for p in self.ptrs]

    def free(self):
        for p in self.ptrs:
            assert self.rtlib.cudaFree(p) == 0
        self.ptrs = []


def reference_host_frames(rt, t, host, W, H, seed, frames):
    """RayTracing.cpp:213-233 verbatim in spirit: init_rng, then raytracing_process + the caller's D2D copy per frame."""
    n = gpu_util.rng_state_count(W, H)
    surface = t.zeros((H, W, 4), dtype=t.float32, device="cuda"); last = t.zeros_like(surface)
    rt.lib().init_rng(n // 128, 128, C.c_void_p(host.scene.rng_state), seed)
    out = []
    for f in range(frames):
        rt.lib().raytracing_process(rt.ptr(surface), rt.ptr(last), W, H, W * 16, f, C.byref(host.scene))
        last.copy_(surface)  # RayTracing.cpp:233
        t.cuda.synchronize()
        out.append(surface.clone())
    return out


def test_unmodified_host_call_sequence_needs_no_rt_calls(rt, torch_cuda):
    t = torch_cuda
    rtlib = cudart()
    W, H, SEED = 96, 64, 909
    # expected: the same scene through Scene::Upload + rt_process_ex
    d = gpu_util.DeviceScene(rt, t, W, H, "poseB", seed=SEED, env=True)
    want = []
    for f in range(3):
        d.process(f, spp=5, bounces=6, write_last=1, ordered=False)
        t.cuda.synchronize()
        want.append(d.surface.clone())
    arrays = d.scene.arrays()
    cam, _ = d.scene.camera()
    rng = t.zeros(gpu_util.rng_state_count(W, H) * 48, dtype=t.uint8, device="cuda")
    capacity = {k: ((np.ascontiguousarray(v).nbytes >> 21) + 2) << 21 for k, v in arrays.items()}
    host = PlainHost(rt, rtlib, arrays, cam.view(np.uint8).ravel().copy(), rng.data_ptr(), d.scene.gpu_scene().environment_cubemap_tex, capacity)
    got = reference_host_frames(rt, t, host, W, H, SEED, 3)
    for f in range(3):
        assert t.equal(got[f], want[f]), f
    assert t.equal(rng, d.rng)

    # the host drops that scene and uploads another one - the bunny moved and turned, same buffer capacities: freshly cudaMalloc'ed
    # arrays, in practice at the very same addresses
    scene_b = rt.Scene(util.assets()["dds"])
    scene_b.setup_cornell_box()
    scene_b.add_loaded_scene(d.mesh, util.bunny_transform(27, -16, 23), scene_b.add_material(albedo=(0.9, 0.3, 0.2), specular=(1, 1, 1, 0), specular_percent=0.2))
    scene_b.build_bvh()
    db = gpu_util.DeviceScene(rt, t, W, H, "poseB", seed=SEED, env=True, scene=scene_b)
    db.process(0, spp=5, bounces=6, ordered=False)
    t.cuda.synchronize()
    assert not t.equal(db.surface, want[0])
    arrays_b = scene_b.arrays()
    host.free()  # ... and nothing else allocates device memory between the free and the re-upload
    host_b = PlainHost(rt, rtlib, arrays_b, cam.view(np.uint8).ravel().copy(), rng.data_ptr(), db.scene.gpu_scene().environment_cubemap_tex, capacity)
    got_b = reference_host_frames(rt, t, host_b, W, H, SEED, 1)
    assert t.equal(got_b[0], db.surface)
    with open(os.path.join(util.ROOT, "gpurun_out", "dropin_addresses.txt"), "w") as fh:
        fh.write(f"node array of scene A at {host.addresses[2]:#x}, of scene B at {host_b.addresses[2]:#x} "
                 f"({'SAME address: the buffer-id check did the work' if host.addresses[2] == host_b.addresses[2] else 'different address'})\n")
    host_b.free()
