"""Headless render of the reference scene to an image file: CUDARayTracer mirror -> float4 surface -> display transform
(exposure 0.5, ACES, sRGB, flip: main.cpp:65-94, as a CUDA kernel) -> PPM (rt_host_write_ppm) and PNG (stdlib zlib).
    python benchmarks/render_image.py [--pose poseB|default] [--width 960 --height 540] [--frames 40] [--out gpurun_out/bunny]"""
import argparse
import os
import struct
import sys
import zlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import util  # noqa: E402


def write_png(path, rgba):
    h, w, _ = rgba.shape
    raw = b"".join(b"\x00" + rgba[y, :, :3].tobytes() for y in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pose", default="poseB")
    ap.add_argument("--width", type=int, default=960)
    ap.add_argument("--height", type=int, default=540)
    ap.add_argument("--frames", type=int, default=40)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--bounces", type=int, default=6)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "bunny"))
    a = ap.parse_args()
    import torch
    from _loader import load_package
    rt = load_package()
    assets = util.assets()
    W, H = a.width, a.height
    surf = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda"); last = torch.zeros_like(surf)
    tr = rt.Tracer(surf, last, W, H, W * 16, bunny_obj=assets["obj"], env_dds=assets["dds"], seed=1234,
                   samples_per_launch=a.spp, max_bounces=a.bounces)
    tr.scene.set_camera(*util.POSES[a.pose], W, H)
    for _ in range(a.frames):
        tr.process()
    rgba = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda")
    rt.check(rt.lib().rt_tonemap_rgba8(rt.ptr(surf), W, H, W * 16, 1.0, rt.ptr(rgba), None))
    torch.cuda.synchronize()
    img = rgba.cpu().numpy()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    rt.check(rt.lib().rt_host_write_ppm((a.out + ".ppm").encode(), rt.ptr(img), W, H))
    write_png(a.out + ".png", img)
    print(f"{a.out}.png  {W}x{H}  {a.frames * a.spp} spp  rays={tr.ray_count}")


if __name__ == "__main__":
    main()
