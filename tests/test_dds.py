"""CPU tests of the general DDS container parser (csrc/dds.cpp ParseDds through rt_dds_info) against the REFERENCE's own parser
(utils/image/DDSLoader.cpp GetDDSSubresources, compiled from /root/reference into oracle/_ref with a Windows.h stand-in) on
generated files covering what the reference accepts and rejects: legacy and DX10 headers, 1D/2D/3D, cubemaps, arrays, mip chains
down to 1x1, block-compressed and packed formats, truncation and bad headers."""
import ctypes as C

import numpy as np
import pytest

import util

RGB, LUM, ALPHA, BUMP, FOURCC = 0x40, 0x20000, 0x2, 0x80000, 0x4

# (name, make_dds kwargs without payload, bits per pixel, block bytes, elements)
CASES = [
    ("legacy RGBA8 2D, full mips", dict(width=16, height=8, pixel_format=(RGB | 1, 0, 32, 0xFF, 0xFF00, 0xFF0000, 0xFF000000), mips=5), 32, 0, 1),
    ("legacy BGRA8", dict(width=5, height=3, pixel_format=(RGB | 1, 0, 32, 0xFF0000, 0xFF00, 0xFF, 0xFF000000)), 32, 0, 1),
    ("legacy B5G6R5, odd size, mips", dict(width=7, height=5, pixel_format=(RGB, 0, 16, 0xF800, 0x07E0, 0x001F, 0), mips=3), 16, 0, 1),
    ("legacy L8", dict(width=9, height=4, pixel_format=(LUM, 0, 8, 0xFF, 0, 0, 0)), 8, 0, 1),
    ("legacy A8L8 written with 8 bits", dict(width=4, height=4, pixel_format=(LUM | 1, 0, 8, 0xFF, 0, 0, 0xFF00)), 16, 0, 1),
    ("legacy A8", dict(width=4, height=2, pixel_format=(ALPHA, 0, 8, 0, 0, 0, 0xFF)), 8, 0, 1),
    ("legacy V8U8", dict(width=4, height=4, pixel_format=(BUMP, 0, 16, 0xFF, 0xFF00, 0, 0)), 16, 0, 1),
    ("legacy DXT1 with mips to 1x1", dict(width=16, height=16, pixel_format=(FOURCC, util.fourcc("DXT1"), 0, 0, 0, 0, 0), mips=5), 4, 8, 1),
    ("legacy DXT5 non-multiple-of-4", dict(width=10, height=6, pixel_format=(FOURCC, util.fourcc("DXT5"), 0, 0, 0, 0, 0), mips=2), 8, 16, 1),
    ("legacy ATI2", dict(width=8, height=8, pixel_format=(FOURCC, util.fourcc("ATI2"), 0, 0, 0, 0, 0)), 8, 16, 1),
    ("legacy RGBG packed", dict(width=5, height=2, pixel_format=(FOURCC, util.fourcc("RGBG"), 0, 0, 0, 0, 0)), 32, 4, 1),
    ("legacy FourCC 113 RGBA16F cubemap, mips", dict(width=8, height=8, pixel_format=(FOURCC, 113, 0, 0, 0, 0, 0), mips=4, cube=True), 64, 0, 6),
    ("legacy FourCC 116 cubemap (the asset's layout)", dict(width=4, height=4, pixel_format=(FOURCC, 116, 0, 0, 0, 0, 0), mips=3, cube=True), 128, 0, 6),
    ("legacy volume R32F", dict(width=4, height=4, depth=4, pixel_format=(FOURCC, 114, 0, 0, 0, 0, 0), mips=3, volume=True), 32, 0, 1),
    ("DX10 RGBA8 sRGB 2D array of 3", dict(width=4, height=4, dxgi=29, array=3, mips=2), 32, 0, 3),
    ("DX10 BC6H cubemap", dict(width=8, height=8, dxgi=95, cube=True, mips=2), 8, 16, 6),
    ("DX10 BC7 sRGB", dict(width=12, height=8, dxgi=99, mips=3), 8, 16, 1),
    ("DX10 two cubemaps", dict(width=4, height=4, dxgi=10, cube=True, array=2), 64, 0, 12),
    ("DX10 1D R16F", dict(width=16, height=1, dxgi=54, dimension=2, mips=5), 16, 0, 1),
    ("DX10 3D RG16 UNORM", dict(width=4, height=2, depth=8, dxgi=35, dimension=4, volume=True, mips=2), 32, 0, 1),
    ("DX10 R11G11B10F (parse only)", dict(width=3, height=3, dxgi=26), 32, 0, 1),
    ("DX10 R9G9B9E5 (parse only)", dict(width=2, height=2, dxgi=67), 32, 0, 1),
    ("mip count field 0 means 1", dict(width=4, height=4, dxgi=28, mip_field=0), 32, 0, 1),
]


def our_info(rt, data, capacity=12 + 3 * 256):
    buf = np.frombuffer(bytes(data), np.uint8).copy()
    info = np.zeros(capacity, np.uint64)
    n = rt.lib().rt_dds_info(rt.ptr(buf), len(buf), info.ctypes.data_as(C.POINTER(C.c_uint64)), capacity)
    return n, info


def build(case):
    name, kw, bpp, block, elements = case
    kw = dict(kw)
    n = util.dds_payload_bytes(kw["width"], kw["height"], kw.get("depth", 0), kw.get("mips", 1), elements, bpp, block)
    payload = (np.arange(n, dtype=np.uint32) * 37 % 251).astype(np.uint8).tobytes()
    return util.make_dds(payload=payload, **kw), n


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_parse_matches_reference_parser(rt, ref, case):
    data, payload_bytes = build(case)
    n, info = our_info(rt, data)
    rn, rinfo = ref.dds_info(data)
    assert rn > 0, "the reference's parser rejects this file: fix the test case"
    assert n == rn
    assert np.array_equal(info[:9], rinfo[:9]), (info[:12], rinfo[:12])
    assert info[11] == rinfo[11] == n
    assert np.array_equal(info[12:12 + 3 * n], rinfo[12:12 + 3 * n])
    # independent of both parsers: the subresources tile the payload exactly
    bpp, block = case[2], case[3]
    assert info[8] == bpp and info[9] == block
    header = len(data) - payload_bytes
    assert info[12] == header
    # one byte short: both reject
    assert our_info(rt, data[:-1])[0] < 0 and ref.dds_info(data[:-1])[0] < 0


def test_parse_without_reference(rt):
    """The same property with no oracle (runs wherever the .so is): offsets start after the header(s), grow monotonically, the last
    subresource ends at the end of the file."""
    for case in CASES:
        data, payload_bytes = build(case)
        n, info = our_info(rt, data)
        assert n > 0, (case[0], rt.lib().rt_last_error())
        offs = info[12:12 + 3 * n:3].astype(np.int64)
        assert offs[0] == len(data) - payload_bytes and (np.diff(offs) > 0).all(), case[0]


BAD = [
    ("bad magic", lambda d: b"DDX " + d[4:]),
    ("header size field", lambda d: d[:4] + (123).to_bytes(4, "little") + d[8:]),
    ("pixel format size field", lambda d: d[:4 + 72] + (31).to_bytes(4, "little") + d[4 + 76:]),
    ("too small", lambda d: d[:100]),
]


@pytest.mark.parametrize("name,mutate", BAD, ids=[b[0] for b in BAD])
def test_rejects_what_the_reference_rejects(rt, ref, name, mutate):
    data, _ = build(CASES[0])
    bad = mutate(data)
    assert our_info(rt, bad)[0] < 0
    assert ref.dds_info(bad)[0] < 0


def test_rejections_of_container_rules(rt, ref):
    px = b"\0" * 4096
    cases = {
        "partial cubemap": util.make_dds(4, 4, px, pixel_format=(FOURCC, 116, 0, 0, 0, 0, 0), cube=True, caps2=0x200 | 0x400 | 0x800),
        "unknown legacy masks": util.make_dds(4, 4, px, pixel_format=(RGB, 0, 24, 0xFF0000, 0xFF00, 0xFF, 0)),
        "17 mips": util.make_dds(4, 4, px, dxgi=28, mips=17),
        "DX10 array size 0": util.make_dds(4, 4, px, dxgi=28, array=0),
        "DX10 unknown format": util.make_dds(4, 4, px, dxgi=200),
        "DX10 palette format": util.make_dds(4, 4, px, dxgi=113),
        "DX10 3D without the volume flag": util.make_dds(4, 4, px, dxgi=28, dimension=4, depth=4),
        "DX10 3D array": util.make_dds(4, 4, px, dxgi=28, dimension=4, depth=4, volume=True, array=2),
        "DX10 1D with height 2": util.make_dds(4, 2, px, dxgi=28, dimension=2),
        "DX10 buffer dimension": util.make_dds(4, 4, px, dxgi=28, dimension=1),
        "2D too wide": util.make_dds(16385, 1, px, dxgi=61),
    }
    for name, data in cases.items():
        assert our_info(rt, data)[0] < 0, name
        assert ref.dds_info(data)[0] < 0, name + " (reference accepts it)"


def test_malformed_cubemap_headers_fail_before_allocating(rt, tmp_path):
    """the RGBA32F cubemap reader must not trust mipMapCount / width before it has checked the file size (a header
    that promises 2^31 mips or a 16384^2 x 6 chain used to allocate tens of GB and throw across the C ABI)."""
    import struct
    texels = np.zeros((6, 4, 4, 4), np.float32)
    path = str(tmp_path / "ok.dds")
    util.write_dds_cubemap(path, texels)
    good = bytearray(open(path, "rb").read())
    assert rt.lib().rt_dds_read_level0(path.encode(), None, 0) == 4
    for name, offset, value in (("mips", 4 + 24, 0x7FFFFFFF), ("mips33", 4 + 24, 33), ("width", 4 + 12, 16384), ("huge", 4 + 12, 1 << 20)):
        bad = bytearray(good)
        struct.pack_into("<I", bad, offset, value)
        if name in ("width", "huge"):
            struct.pack_into("<I", bad, 4 + 8, value)  # height too (faces must be square)
        p = str(tmp_path / (name + ".dds"))
        open(p, "wb").write(bad)
        assert rt.lib().rt_dds_read_level0(p.encode(), None, 0) < 0, name
        assert b"DDS" in rt.lib().rt_last_error()
