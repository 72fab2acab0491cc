"""-m gpu tests of the general DDS -> CUDA texture path (utils/CUDATexture.cpp:9-75,112-220): texel values through
the hardware sampler against numpy decodes of the same file bytes, for plain, normalised, sRGB, half-float, 3D, cubemap and
block-compressed (BC1, hardware-decoded) containers with mip chains; and an RGBA16F cubemap as the path tracer's environment map."""
import ctypes as C
import struct

import numpy as np
import pytest

import gpu_util
import util

pytestmark = pytest.mark.gpu


class Texture:
    def __init__(self, rt, torch, data, srgb=0):
        self.rt, self.torch = rt, torch
        self.buf = np.frombuffer(bytes(data), np.uint8).copy()
        self.handle = C.c_void_p()
        rt.check(rt.lib().rt_texture_create_from_memory(rt.ptr(self.buf), len(self.buf), srgb, C.byref(self.handle)))

    def fetch(self, coords, lod=0.0):
        t = self.torch
        c = t.from_numpy(np.ascontiguousarray(coords, np.float32)).cuda()
        out = t.zeros((len(coords), 4), dtype=t.float32, device="cuda")
        self.rt.check(self.rt.lib().rt_texture_fetch(self.handle, len(coords), self.rt.ptr(c), float(lod), self.rt.ptr(out), None))
        t.cuda.synchronize()
        return out.cpu().numpy()

    def __del__(self):
        if getattr(self, "handle", None):
            self.rt.lib().rt_envmap_destroy(self.handle)
            self.handle = None


def centres(w, h):
    ys, xs = np.mgrid[0:h, 0:w]
    return np.stack([(xs.ravel() + 0.5) / w, (ys.ravel() + 0.5) / h, np.zeros(w * h)], 1)


def mip_chain(w, h, channels, rng, dtype):
    levels = []
    while True:
        levels.append(rng.integers(0, 256, (h, w, channels)).astype(dtype))
        if w == 1 and h == 1:
            return levels
        w, h = max(w // 2, 1), max(h // 2, 1)


def test_rgba8_mips_plain_and_srgb(rt, torch_cuda):
    rng = np.random.default_rng(1)
    levels = mip_chain(8, 4, 4, rng, np.uint8)
    data = util.make_dds(8, 4, b"".join(l.tobytes() for l in levels), pixel_format=(0x41, 0, 32, 0xFF, 0xFF00, 0xFF0000, 0xFF000000), mips=len(levels))
    lin = Texture(rt, torch_cuda, data, srgb=0)
    for m, l in enumerate(levels):
        got = lin.fetch(centres(l.shape[1], l.shape[0]), lod=m)
        assert np.array_equal(got, l.reshape(-1, 4).astype(np.float32) / np.float32(255)), m
    srgb = Texture(rt, torch_cuda, data, srgb=1)  # the reference's default (CUDATexture.h: sRGB = true): colour decoded, alpha linear
    got = srgb.fetch(centres(8, 4))
    c = levels[0].reshape(-1, 4).astype(np.float64) / 255
    want = np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4)
    want[:, 3] = c[:, 3]
    assert np.abs(got - want).max() < 5e-3  # the sampler's sRGB decode is a ~9-bit table


def test_half_float_cubemap_and_volume(rt, torch_cuda):
    rng = np.random.default_rng(2)
    faces = rng.uniform(0, 4, (6, 4, 4, 4)).astype(np.float16)
    data = util.make_dds(4, 4, faces.tobytes(), dxgi=10, cube=True)
    tex = Texture(rt, torch_cuda, data)
    # centre of every face: the major-axis direction picks the face (+x, -x, +y, -y, +z, -z), s/t = 0 -> between the 4 centre texels
    dirs = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], np.float32)
    got = tex.fetch(dirs)
    want = faces[:, 1:3, 1:3, :].astype(np.float32).mean(axis=(1, 2))
    assert np.allclose(got, want, rtol=0, atol=4 / 256)  # 1.8 fixed-point filter weights
    vol = rng.uniform(-1, 1, (4, 2, 8)).astype(np.float32)  # depth 4, height 2, width 8
    data = util.make_dds(8, 2, vol.tobytes(), dxgi=41, dimension=4, depth=4, volume=True)
    t3 = Texture(rt, torch_cuda, data)
    zs, ys, xs = np.mgrid[0:4, 0:2, 0:8]
    coords = np.stack([(xs.ravel() + 0.5) / 8, (ys.ravel() + 0.5) / 2, (zs.ravel() + 0.5) / 4], 1)
    got = t3.fetch(coords)
    assert np.array_equal(got[:, 0], vol.ravel()) and (got[:, 1:] == 0).all()  # CUDA fills absent channels with 0 (D3D would give alpha 1)


def bc1_block(c0, c1, indices):
    """c0, c1: RGB565 ints; indices: 16 two-bit selectors, row-major."""
    bits = 0
    for i, s in enumerate(indices):
        bits |= int(s) << (2 * i)
    return struct.pack("<HHI", c0, c1, bits)


def bc1_decode(block):
    c0, c1, bits = struct.unpack("<HHI", block)

    def rgb(c):
        r, g, b = (c >> 11) & 31, (c >> 5) & 63, c & 31
        return np.array([r / 31.0, g / 63.0, b / 31.0, 1.0])
    p0, p1 = rgb(c0), rgb(c1)
    if c0 > c1:
        pal = [p0, p1, (2 * p0 + p1) / 3, (p0 + 2 * p1) / 3]
    else:
        pal = [p0, p1, (p0 + p1) / 2, np.array([0.0, 0.0, 0.0, 0.0])]
    return np.array([pal[(bits >> (2 * i)) & 3] for i in range(16)]).reshape(4, 4, 4)


def test_bc1_with_mips_against_numpy_decode(rt, torch_cuda):
    rng = np.random.default_rng(3)
    w = h = 8
    blocks, images = [], []
    lw, lh = w, h
    while True:
        bw, bh = max(lw // 4, 1), max(lh // 4, 1)
        img = np.zeros((bh * 4, bw * 4, 4))
        for by in range(bh):
            for bx in range(bw):
                b = bc1_block(int(rng.integers(0, 65536)), int(rng.integers(0, 65536)), rng.integers(0, 4, 16))
                blocks.append(b)
                img[by * 4:by * 4 + 4, bx * 4:bx * 4 + 4] = bc1_decode(b)
        images.append(img[:lh, :lw])
        if lw == 1 and lh == 1:
            break
        lw, lh = max(lw // 2, 1), max(lh // 2, 1)
    data = util.make_dds(w, h, b"".join(blocks), pixel_format=(0x4, util.fourcc("DXT1"), 0, 0, 0, 0, 0), mips=len(images))
    tex = Texture(rt, torch_cuda, data, srgb=0)
    for m, img in enumerate(images):
        got = tex.fetch(centres(img.shape[1], img.shape[0]), lod=m)
        assert np.abs(got - img.reshape(-1, 4)).max() < 2.5 / 255, m  # the hardware decoder interpolates the two derived colours with its own rounding (observed <= 1.9/255)


def test_formats_without_a_cuda_array_format_fail_loudly(rt, torch_cuda):
    for dxgi in (26, 67, 87, 24):  # R11G11B10F, R9G9B9E5, B8G8R8A8, R10G10B10A2
        data = util.make_dds(4, 4, b"\0" * 64, dxgi=dxgi)
        buf = np.frombuffer(data, np.uint8).copy()
        h = C.c_void_p()
        assert rt.lib().rt_texture_create_from_memory(rt.ptr(buf), len(buf), 0, C.byref(h)) < 0
        assert b"no CUDA array format" in rt.lib().rt_last_error()


def test_rgba16f_cubemap_as_environment_map(rt, torch_cuda, tmp_path):
    """The same sunset texels stored as RGBA16F (DX10 header) drive the path tracer through rt_envmap_create_*: the frame matches the
    RGBA32F one to half-float rounding of the radiance (2^-11 relative per texel)."""
    texels = util.cubemap_texels()
    half = texels.astype(np.float16)
    size = texels.shape[1]
    p16 = tmp_path / "sunset_rgba16f.dds"
    p16.write_bytes(util.make_dds(size, size, half.tobytes(), dxgi=10, cube=True))
    a = util.assets()
    frames = []
    for env_path in (a["dds"], str(p16)):
        scene = rt.Scene(env_path)
        scene.setup_cornell_box(); scene.setup_stanford_bunny(rt.Mesh(a["obj"])); scene.build_bvh()
        d = gpu_util.DeviceScene(rt, torch_cuda, 160, 120, "poseB", seed=99, env=True, scene=scene)
        d.process(0, spp=8, bounces=4, write_last=1)
        frames.append(d.surface.cpu().numpy()[..., :3].astype(np.float64))
    ref32, got16 = frames
    assert ref32.mean() > 0.05
    rel = np.abs(got16 - ref32) / (np.abs(ref32) + 1e-3)
    assert rel.max() < 4e-3 and np.abs(got16.mean() - ref32.mean()) / ref32.mean() < 1e-3
