"""Generates the committed golden fixtures. Runs ONLY in the build container, where /root/reference and
oracle/_ref/libref_oracle.so (the reference's own code compiled host-side) exist:

    python tests/golden/make_fixtures.py

Outputs (tests/golden/):
  bunny_mesh.npz      positions f32[35947,3] + 0-based triangle indices u32[69451,3] parsed from
                      data/stanford-bunny.obj (binary restatement of the asset; the GPU box has no /root/reference)
  sunset_level0.npz   level-0 texels f32[6,128,128,4] of data/sunset_uncompressed.dds (the only level sampled)
  scene_hashes.json   sha256 of the reference-layout arrays + BVH the GLM/BVH.cpp build produces
  primary_320x240.npz reference BVHRayHit/GetRayHit closest-hit ids (first-visited rule) + t, two poses,
                      with and without spheres
  primary_320x240_lowest_id.npz  the same queries under the order-independent rule (min t, ties -> lowest id), from
                      oracle/rt_port.c after its reference-rule mode is asserted equal to the reference's output
  rng_kat.npz         curand_init(seed, subsequence, 0) + curand() known answers from the toolkit's header
  rays_kat.npz        4096 seeded incoherent rays through GetRayHit (ids, t)
  pt_96x72.npz        raytracing_kernel_main run per pixel on the host (5 spp x 6 bounces, seed 1234, pose B,
                      sunset env map through oracle/cubemap_cpu.h): surfaces after frames 0,1,2 + final RNG states
"""
import hashlib
import json
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
REF = "/root/reference"

POSES = {"default": ((0.0, 0.0, 0.0), 0.0, 180.0), "poseB": ((32.0, -1.0, 2.0), 0.0, 180.0)}


def parse_obj(path):
    pos, idx = [], []
    for line in open(path):
        if line.startswith("v "):
            pos.append([np.float32(x) for x in line.split()[1:4]])
        elif line.startswith("f "):
            c = [int(t.split("/")[0]) - 1 for t in line.split()[1:]]
            for k in range(1, len(c) - 1):
                idx.append([c[0], c[k], c[k + 1]])
    return np.asarray(pos, np.float32), np.asarray(idx, np.uint32)


def parse_dds_level0(path):
    b = open(path, "rb").read()
    assert b[:4] == b"DDS " and struct.unpack_from("<I", b, 84)[0] == 116
    size, mips = struct.unpack_from("<I", b, 16)[0], struct.unpack_from("<I", b, 28)[0]
    chain = sum((size >> m) ** 2 * 16 for m in range(mips))
    faces = [np.frombuffer(b, np.float32, size * size * 4, 128 + f * chain).reshape(size, size, 4) for f in range(6)]
    return np.stack(faces)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def seeded_rays(n, seed):
    rng = np.random.default_rng(seed)
    o = rng.uniform([-10, -10, 0], [45, 10, 40], (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d *= rng.uniform(0.5, 3.0, (n, 1)).astype(np.float32)  # unnormalised on purpose
    return np.concatenate([o, d], 1).astype(np.float32)


def main():
    import ref_oracle as ro
    from _loader import load_package
    rt = load_package()

    pos, idx = parse_obj(os.path.join(REF, "data/stanford-bunny.obj"))
    np.savez_compressed(os.path.join(HERE, "bunny_mesh.npz"), positions=pos, indices=idx)
    texels = parse_dds_level0(os.path.join(REF, "data/sunset_uncompressed.dds"))
    np.savez_compressed(os.path.join(HERE, "sunset_level0.npz"), texels=texels)

    # smooth mesh through the product loader (Assimp stand-in), then GLM assembly + verbatim BVH.cpp
    mesh = rt.Mesh(os.path.join(REF, "data/stanford-bunny.obj"))
    mpos, mnrm, midx, _ = mesh.arrays()
    scene = ro.build_scene(mpos, mnrm, midx)
    nodes, face_indices = ro.bvh_build(scene["vertices"], scene["faces"])
    hashes = {"vertices": sha(scene["vertices"]), "faces": sha(scene["faces"]), "spheres20": sha(scene["spheres"][:, :20]),
              "materials60": sha(scene["materials"][:, :60]), "nodes": sha(nodes), "face_indices": sha(face_indices),
              "counts": {k: int(len(v)) for k, v in scene.items()} | {"nodes": int(len(nodes))},
              "cameras": {k: sha(ro.camera(p, ax, ay, 320, 240)) for k, (p, ax, ay) in POSES.items()}}
    json.dump(hashes, open(os.path.join(HERE, "scene_hashes.json"), "w"), indent=1)

    out = {}
    for name, (p, ax, ay) in POSES.items():
        cam = ro.camera(p, ax, ay, 320, 240)
        for ws in (0, 1):
            ids, ts = ro.primary_hits(scene["vertices"], scene["faces"], nodes, face_indices, scene["spheres"], cam, 320, 240, ws)
            out[f"{name}_s{ws}_id"] = ids
            out[f"{name}_s{ws}_t"] = ts
    np.savez_compressed(os.path.join(HERE, "primary_320x240.npz"), **out)

    rays = seeded_rays(4096, 7)
    ids, ts = ro.trace_rays(scene["vertices"], scene["faces"], nodes, face_indices, scene["spheres"], rays, True)
    np.savez_compressed(os.path.join(HERE, "rays_kat.npz"), rays=rays, ids=ids, t=ts)

    kat = {}
    seeds = [0, 1234, 0xDEADBEEF]
    subs = [0, 1, 2, 3, 31, 32, 33, 127, 128, 1000, 65535, 65536, 307199, 2073599, 8294399, (1 << 24) + 12345, (1 << 33) + 7]
    table = np.zeros((len(seeds), len(subs), 8), np.uint32)
    state = np.zeros(48, np.uint8)
    for i, s in enumerate(seeds):
        for j, q in enumerate(subs):
            ro.lib().ref_rng_init_one(state.ctypes.data, s, q, 0)
            table[i, j] = ro.rng_draw_u32(state, 8)
    kat = {"seeds": np.asarray(seeds, np.uint64), "subsequences": np.asarray(subs, np.uint64), "draws": table}
    st = ro.rng_init(4, 1234)
    uni = np.zeros(16, np.float32)
    ro.lib().ref_rng_draw_uniform(st[2].ctypes.data, 16, uni.ctypes.data)
    kat["uniform_seed1234_sub2"] = uni
    np.savez_compressed(os.path.join(HERE, "rng_kat.npz"), **kat)
    # ids under the ORDER-INDEPENDENT rule the product implements (min t, exact ties -> lowest id; BASELINE.json):
    # produced by the plain-C port (oracle/rt_port.c, tie_mode 1) AFTER asserting that its tie_mode 0 reproduces the
    # unmodified reference functions bit for bit on the same rays.
    import port_oracle as po
    ps = po.PortScene(scene, nodes, face_indices, None, None)
    low = {}
    for name, (p, ax, ay) in POSES.items():
        ps.set(ro.camera(p, ax, ay, 320, 240))
        for ws in (0, 1):
            i0, t0 = ps.primary_hits(320, 240, ws, 0)
            assert np.array_equal(i0, out[f"{name}_s{ws}_id"]) and np.array_equal(t0.view(np.uint32), out[f"{name}_s{ws}_t"].view(np.uint32))
            i1, t1 = ps.primary_hits(320, 240, ws, 1)
            assert np.array_equal(t1.view(np.uint32), t0.view(np.uint32)) and (i1 <= i0).all()
            low[f"{name}_s{ws}_id"] = i1
            print(name, ws, "tie swaps:", int((i1 != i0).sum()))
    i0, t0 = ps.trace_rays(rays, True, 0)
    assert np.array_equal(i0, ids) and np.array_equal(t0.view(np.uint32), ts.view(np.uint32))
    low["rays_id"], _ = ps.trace_rays(rays, True, 1)
    np.savez_compressed(os.path.join(HERE, "primary_320x240_lowest_id.npz"), **low)

    # path-traced frames (full kernel body incl. accumulation), small enough to commit
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import gpu_util
    import util
    uscene, _ = util.bunny_scene(rt)
    frames, states = gpu_util.oracle_frames(ro, rt, uscene, 96, 72, "poseB", 1234, 3, np.ascontiguousarray(texels, np.float32))
    np.savez_compressed(os.path.join(HERE, "pt_96x72.npz"), frames=np.stack(frames), rng_v=states[:96 * 72, :24].copy())
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
