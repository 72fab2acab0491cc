"""Shared helpers for tests / bench / smoke: asset materialisation from the committed fixtures, poses."""
import os
import struct
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if os.path.join(ROOT, "oracle") not in sys.path:
    sys.path.insert(0, os.path.join(ROOT, "oracle"))

# SURVEY.md §8(d) config 1 poses: (position, angle_x, angle_y)
POSES = {"default": ((0.0, 0.0, 0.0), 0.0, 180.0), "poseB": ((32.0, -1.0, 2.0), 0.0, 180.0)}

_ASSET_DIR = None


def write_obj(path, positions, indices):
    with open(path, "w") as f:
        f.write("# regenerated from tests/golden/bunny_mesh.npz (values round-trip float32 exactly)\n")
        for p in positions:
            f.write("v %.9g %.9g %.9g\n" % (p[0], p[1], p[2]))
        for t in indices:
            f.write("f %d %d %d\n" % (t[0] + 1, t[1] + 1, t[2] + 1))


def write_dds_cubemap(path, texels):
    """Legacy DDS, FourCC 116 (RGBA32F), six faces, one mip level - the layout of sunset_uncompressed.dds."""
    size = texels.shape[1]
    header = struct.pack("<7I", 124, 0x1 | 0x2 | 0x4 | 0x1000 | 0x20000, size, size, size * 16, 0, 1) + b"\0" * 44
    header += struct.pack("<8I", 32, 0x4, 116, 0, 0, 0, 0, 0)
    header += struct.pack("<5I", 0x1000 | 0x8 | 0x400000, 0xFE00, 0, 0, 0)
    assert len(header) == 124
    with open(path, "wb") as f:
        f.write(b"DDS " + header + np.ascontiguousarray(texels, np.float32).tobytes())


def assets():
    """Paths of a temp OBJ + DDS rebuilt from the fixtures (the GPU box has no /root/reference)."""
    global _ASSET_DIR
    if _ASSET_DIR is None:
        d = tempfile.mkdtemp(prefix="b200rt_assets_")
        mesh = np.load(os.path.join(GOLDEN, "bunny_mesh.npz"))
        write_obj(os.path.join(d, "stanford-bunny.obj"), mesh["positions"], mesh["indices"])
        write_dds_cubemap(os.path.join(d, "sunset_uncompressed.dds"), np.load(os.path.join(GOLDEN, "sunset_level0.npz"))["texels"])
        _ASSET_DIR = d
    return {"obj": os.path.join(_ASSET_DIR, "stanford-bunny.obj"), "dds": os.path.join(_ASSET_DIR, "sunset_uncompressed.dds")}


def cubemap_texels():
    return np.ascontiguousarray(np.load(os.path.join(GOLDEN, "sunset_level0.npz"))["texels"], np.float32)


def bunny_scene(rt, env=False):
    """The reference's fixed scene (Cornell box + bunny + floor + emissive sphere), host side, BVH built."""
    a = assets()
    mesh = rt.Mesh(a["obj"])
    scene = rt.Scene(a["dds"] if env else None)
    scene.setup_cornell_box()
    scene.setup_stanford_bunny(mesh)
    scene.build_bvh()
    return scene, mesh


# ---- BASELINE config 4: synthetic scaled meshes ----------------------------------------------------------
GRID_POSE = ((165.0, 60.0, -40.0), -20.0, 180.0)  # above the 10x10 grid, looking along +z and down: bunnies cover 22 % of the pixels


def bunny_transform(tx, ty, tz):
    """T(tx,ty,tz) * Ry(-pi) * Rx(pi/2) * S(150): the reference's bunny placement (RayTracing.cpp:42-46) moved to (tx,ty,tz)."""
    import math
    c, s = math.cos(-math.pi), math.sin(-math.pi)
    ry = np.array([[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1]], np.float32)
    c2, s2 = math.cos(math.pi / 2), math.sin(math.pi / 2)
    rx = np.array([[1, 0, 0, 0], [0, c2, -s2, 0], [0, s2, c2, 0], [0, 0, 0, 1]], np.float32)
    t = np.eye(4, dtype=np.float32); t[:3, 3] = (tx, ty, tz)
    return t @ ry @ rx @ np.diag([150, 150, 150, 1]).astype(np.float32)


def scaled_scene(rt, which, env=True, grid=10):
    """BASELINE config 4 scenes through the reference's own AddLoadedScene semantics (every face twice, Scene.cpp:75-132):
    'a' = 2x Loop-subdivided bunny in the reference scene (2 222 432 BVH triangles, one 1 079-prim reference leaf);
    'b' = grid x grid bunnies, pitch 30 in x/z, in the Cornell box scene WITHOUT the floor quad (10x10: 13 890 200 triangles;
    with BVH.cpp's midpoint split two grid-sized floor triangles pin ~3 M primitives into one leaf, BVH.cpp:79-107).
    Returns (scene, mesh, pose, timings)."""
    import time
    a = assets()
    t0 = time.time()
    scene = rt.Scene(a["dds"] if env else None)
    scene.setup_cornell_box()
    if which == "a":
        mesh = rt.Mesh(a["obj"], subdivide=2)
        scene.setup_stanford_bunny(mesh)
        pose = POSES["poseB"]
    else:
        mesh = rt.Mesh(a["obj"])
        m = scene.add_material(albedo=(1, 1, 1), specular=(0.3, 1.0, 0.3, 0.0), roughness=0.8, specular_percent=0.5)
        for gx in range(grid):
            for gz in range(grid):
                scene.add_loaded_scene(mesh, bunny_transform(30 + 30 * gx, -18, 20 + 30 * gz), m)
        pose = GRID_POSE
    t1 = time.time()
    scene.build_bvh()
    t2 = time.time()
    return scene, mesh, pose, {"assemble_s": t1 - t0, "bvh_build_s": t2 - t1}


def raw(a):
    """uint8 [n, itemsize] view of a structured array (what oracle/ref_oracle.py takes)."""
    return np.ascontiguousarray(a).view(np.uint8).reshape(len(a), a.dtype.itemsize)


# ---- DDS containers for the general loader tests ------------------------------------------------------------
DDSD = {"caps": 0x1, "height": 0x2, "width": 0x4, "pitch": 0x8, "pixelformat": 0x1000, "mips": 0x20000, "linear": 0x80000, "depth": 0x800000}


def fourcc(s):
    return struct.unpack("<I", s.encode("ascii"))[0]


def make_dds(width, height, payload, pixel_format=None, dxgi=None, depth=0, mips=1, cube=False, array=1, dimension=3, volume=False,
             caps2=None, mip_field=None):
    """A DDS file image. pixel_format = (flags, fourCC, bits, R, G, B, A) for a legacy header, or dxgi = DXGI_FORMAT for a DX10 one."""
    flags = DDSD["caps"] | DDSD["height"] | DDSD["width"] | DDSD["pixelformat"] | (DDSD["mips"] if mips > 1 else 0) | (DDSD["depth"] if volume else 0)
    if dxgi is not None:
        pixel_format = (0x4, fourcc("DX10"), 0, 0, 0, 0, 0)
    header = struct.pack("<7I", 124, flags, height, width, 0, depth, mips if mip_field is None else mip_field) + b"\0" * 44
    header += struct.pack("<8I", 32, *pixel_format)
    if caps2 is None:
        caps2 = (0xFE00 if cube and dxgi is None else 0) | (0x200000 if volume else 0)
    header += struct.pack("<5I", 0x1000 | (0x8 if cube or mips > 1 else 0) | (0x400000 if mips > 1 else 0), caps2, 0, 0, 0)
    assert len(header) == 124
    out = b"DDS " + header
    if dxgi is not None:
        out += struct.pack("<5I", dxgi, dimension, 0x4 if cube else 0, array, 0)
    return out + bytes(payload)


def dds_payload_bytes(width, height, depth, mips, elements, bits_per_pixel, block_bytes):
    """Independent size computation (array element major, mips inside)."""
    total = 0
    for _ in range(elements):
        w, h, d = width, height, max(depth, 1)
        for _ in range(mips):
            if block_bytes >= 8:
                total += max(1, (w + 3) // 4) * max(1, (h + 3) // 4) * block_bytes * d
            elif block_bytes == 4:
                total += ((w + 1) // 2) * 4 * h * d
            else:
                total += ((w * bits_per_pixel + 7) // 8) * h * d
            w, h, d = max(w // 2, 1), max(h // 2, 1), max(d // 2, 1)
    return total
