"""Device-side helpers for the -m gpu tests, smoke() and bench.py: torch is used for device memory only."""
import ctypes as C

import numpy as np

import util


def rng_state_count(width, height):
    """RayTracing.cpp:215-218: ceil(W*H/128)*128 states."""
    return ((width * height + 127) // 128) * 128


class DeviceScene:
    """The reference's fixed scene uploaded to cuda:0 through Scene::Upload, plus RNG states and two float4 surfaces."""

    def __init__(self, rt, torch, width, height, pose="poseB", seed=1234, env=True, scene=None, device="cuda:0"):
        self.rt, self.torch, self.width, self.height = rt, torch, width, height
        self.device = device
        if scene is None:
            scene, self.mesh = util.bunny_scene(rt, env=env)
        self.scene = scene
        n = rng_state_count(width, height)
        self.rng = torch.zeros(n * 48, dtype=torch.uint8, device=device)
        self.seed = seed
        self.reseed(seed)
        self.pitch = width * 16
        # main.cpp:638,645: surface memset to bytes 0x01, last-frame to 0
        self.surface = torch.full((height, width, 4), 0, dtype=torch.float32, device=device)
        self.surface.view(torch.uint8).fill_(1)
        self.last = torch.zeros((height, width, 4), dtype=torch.float32, device=device)
        self.counter = torch.zeros(1, dtype=torch.int64, device=device)
        from cuda_raytracing_b200.api import process_workspace_bytes
        self.workspace = torch.zeros(max(process_workspace_bytes(width, height), 4), dtype=torch.uint8, device=device)
        self.set_pose(pose)

    def reseed(self, seed, first_subsequence=0):
        self.rt.check(self.rt.lib().rt_init_rng(self.rt.ptr(self.rng), self.rng.numel() // 48, seed, first_subsequence, None))

    def set_pose(self, pose):
        p, ax, ay = util.POSES[pose] if isinstance(pose, str) else pose
        self.scene.set_camera(p, ax, ay, self.width, self.height)
        self.scene.upload(self.rng)

    @property
    def gpu(self):
        return self.scene.gpu_scene_ptr()

    @property
    def accel(self):
        return self.scene.accel_ptr()

    def trace_primary(self, with_spheres, variant):
        t = self.torch
        ids = t.zeros(self.width * self.height, dtype=t.int32, device=self.device)
        ts = t.zeros(self.width * self.height, dtype=t.float32, device=self.device)
        self.rt.check(self.rt.lib().rt_trace_primary(self.accel, self.gpu, self.width, self.height, int(with_spheres), variant,
                                                     self.rt.ptr(ids), self.rt.ptr(ts), self.rt.ptr(self.counter), None))
        t.cuda.synchronize()
        return ids.cpu().numpy().view(np.uint32), ts.cpu().numpy()

    def trace_rays(self, rays, with_spheres, variant):
        t = self.torch
        d_rays = t.from_numpy(np.ascontiguousarray(rays, np.float32)).to(self.device)
        n = len(rays)
        ids = t.zeros(max(n, 1), dtype=t.int32, device=self.device)
        ts = t.zeros(max(n, 1), dtype=t.float32, device=self.device)
        self.rt.check(self.rt.lib().rt_trace_rays(self.accel, self.gpu, self.rt.ptr(d_rays), n, int(with_spheres), variant,
                                                  self.rt.ptr(ids), self.rt.ptr(ts), None, None))
        t.cuda.synchronize()
        return ids.cpu().numpy().view(np.uint32)[:n], ts.cpu().numpy()[:n]

    def process(self, frame_index, spp=5, bounces=6, blend_mode=0, variant=0, write_last=0, surface=None, last=None, ordered=True):
        from cuda_raytracing_b200.api import process_ex, process_workspace_bytes
        ws = self.workspace if ordered else None
        if variant == 10:  # queue-based wavefront: path states + ray queues live in its (much larger) workspace
            if getattr(self, "wavefront_workspace", None) is None:
                self.wavefront_workspace = self.torch.empty(process_workspace_bytes(self.width, self.height, 10), dtype=self.torch.uint8, device=self.device)
            ws = self.wavefront_workspace
        process_ex(self.surface if surface is None else surface, self.last if last is None else last, self.width, self.height,
                   self.pitch, frame_index, self.gpu, self.accel, spp=spp, bounces=bounces, blend_mode=blend_mode, variant=variant,
                   write_last=write_last, ray_counter=self.counter, workspace=ws)

    def rays(self):
        self.torch.cuda.synchronize()
        return int(self.counter.item())


def oracle_frames(ro, rt, scene, width, height, pose, seed, frames, cubemap):
    """The reference kernel body (oracle/_ref) run on the host for `frames` progressive frames. Returns list of surfaces."""
    a = scene.arrays()
    p, ax, ay = util.POSES[pose] if isinstance(pose, str) else pose
    cam = ro.camera(p, ax, ay, width, height)
    sc = {k: util.raw(a[k]) for k in ("vertices", "faces", "spheres", "materials")}
    st = ro.rng_init(rng_state_count(width, height), seed)
    surf = np.zeros((height, width, 4), np.float32)
    surf.view(np.uint8)[:] = 1
    last = np.zeros((height, width, 4), np.float32)
    out = []
    for f in range(frames):
        ro.render_frame(sc, util.raw(a["nodes"]), a["face_indices"], cam, cubemap, st, surf, last, width, height, f)
        last[:] = surf  # RayTracing.cpp:233
        out.append(surf.copy())
    return out, st
