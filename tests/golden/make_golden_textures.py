"""Generate tests/golden/textures.npz: image files (baseline and progressive JPEG in every sampling layout, PNG, TGA) and the RGBA8 texels the
REFERENCE gets from each -- its own vendored decoder (/root/reference/include/stb/stb_image.h) behind loadTexture's two
steps (src/Model.cpp:99-118: stbi_load(..., STBI_rgb_alpha), then the mirror along y), compiled where it lies as
oracle/_ref/stb_decode (oracle/Makefile target ref_stb).  host/image_io.cpp is pinned against these bit for bit
(tests/test_host.py::test_host_textures_equal_the_reference_decoder).

Run in the build container (needs /root/reference, and Pillow / OpenCV to WRITE the test files; the test itself needs neither):
    make -C oracle ref_stb && python tests/golden/make_golden_textures.py
"""
import io
import os
import subprocess
import sys
import tempfile

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
STB = os.path.join(ROOT, "oracle", "_ref", "stb_decode")


def smooth(rng, h, w, c):
    a = rng.random((h // 4 + 2, w // 4 + 2, c))
    img = np.kron(a, np.ones((4, 4, 1)))[:h, :w]
    for _ in range(2):
        img = (img + np.roll(img, 1, 0) + np.roll(img, 1, 1) + np.roll(img, -1, 0) + np.roll(img, -1, 1)) / 5
    return (img * 255).astype(np.uint8)


def reference_texels(data, suffix):
    with tempfile.NamedTemporaryFile(suffix=suffix, delete=False) as f:
        f.write(data)
    try:
        out = subprocess.run([STB, f.name], capture_output=True, check=True).stdout
    finally:
        os.unlink(f.name)
    head, body = out.split(b"\n", 1)
    w, h = (int(x) for x in head.split())
    return np.frombuffer(body, np.uint8).reshape(h, w, 4).copy()


def png_bytes(px, ctype, depth, interlace=0, palette=None, trns=None, filt=0):
    """A PNG file of px[h, w, channels] (integers below 2**depth): any colour type / bit depth, optionally Adam7."""
    import struct
    import zlib

    h, w, ch = px.shape

    def chunk(t, body):
        return struct.pack(">I", len(body)) + t + body + struct.pack(">I", zlib.crc32(t + body) & 0xffffffff)

    def rows(sub):
        out = bytearray()
        for r in sub:
            v = r.reshape(-1).astype(np.uint32)
            if depth == 16:
                b = np.stack([v >> 8, v & 255], -1).astype(np.uint8).tobytes()
            elif depth == 8:
                b = v.astype(np.uint8).tobytes()
            else:
                per = 8 // depth
                pad = (-len(v)) % per
                vv = np.concatenate([v, np.zeros(pad, np.uint32)]).reshape(-1, per)
                b = sum(vv[:, i] << (8 - depth * (i + 1)) for i in range(per)).astype(np.uint8).tobytes()
            out += bytes([0]) + b          # filter 0 (the five filters are covered by test_host_texture_formats)
        return bytes(out)

    raw = b""
    if interlace:
        for x0, y0, dx, dy in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
            sub = px[y0::dy, x0::dx]
            if sub.shape[0] and sub.shape[1]:
                raw += rows(sub)
    else:
        raw = rows(px)
    out = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, interlace))
    if palette is not None:
        out += chunk(b"PLTE", bytes(palette))
    if trns is not None:
        out += chunk(b"tRNS", bytes(trns))
    return out + chunk(b"IDAT", zlib.compress(raw)) + chunk(b"IEND", b"")


def main():
    if not os.path.exists(STB):
        sys.exit("oracle/_ref/stb_decode is not built: make -C oracle ref_stb (needs /root/reference)")
    rng = np.random.default_rng(2026)
    files = {}

    def jpeg(name, arr, **kw):
        b = io.BytesIO()
        Image.fromarray(arr).save(b, format="JPEG", **kw)
        files[name + ".jpg"] = b.getvalue()

    k = 0
    for (w, h) in ((16, 16), (37, 23), (8, 8), (50, 41), (1, 1), (3, 17)):
        for sub, tag in ((0, "444"), (1, "422"), (2, "420")):
            for kw, t2 in (({}, "std"), ({"optimize": True}, "opt"), ({"restart_marker_blocks": 3}, "rst")):
                if (w, h) in ((1, 1), (3, 17)) and t2 != "std":
                    continue
                jpeg("c%d_%dx%d_%s_%s" % (k, w, h, tag, t2), smooth(rng, h, w, 3), quality=(92, 75, 40)[k % 3], subsampling=sub, **kw)
                k += 1
    jpeg("noise_420", rng.integers(0, 256, (24, 40, 3), dtype=np.uint8), quality=95, subsampling=2)  # clamping, large AC
    jpeg("grey_33x20", smooth(rng, 20, 33, 1)[..., 0], quality=90)
    jpeg("grey_9x9_rst", smooth(rng, 9, 9, 1)[..., 0], quality=60, restart_marker_blocks=1)
    # 4:4:0 (luma 1x2: the vertical-only filter) and 4:1:1 (luma 4x1: the replicating path) through OpenCV's encoder
    import cv2

    for tag, flag, (w, h) in (("440", cv2.IMWRITE_JPEG_SAMPLING_FACTOR_440, (20, 30)), ("411", cv2.IMWRITE_JPEG_SAMPLING_FACTOR_411, (41, 18)),
                              ("440", cv2.IMWRITE_JPEG_SAMPLING_FACTOR_440, (5, 9)), ("411", cv2.IMWRITE_JPEG_SAMPLING_FACTOR_411, (7, 3))):
        ok, buf = cv2.imencode(".jpg", smooth(rng, h, w, 3), [cv2.IMWRITE_JPEG_QUALITY, 85, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, flag])
        assert ok
        files["cv_%s_%dx%d.jpg" % (tag, w, h)] = buf.tobytes()
    # progressive (SOF2): spectral selection + successive approximation, DC scans interleaved, AC scans per component
    k = 0
    for (w, h) in ((16, 16), (37, 23), (50, 41), (1, 1), (9, 3)):
        for sub, tag in ((0, "444"), (1, "422"), (2, "420")):
            for kw, t2 in (({}, "std"), ({"optimize": True}, "opt"), ({"restart_marker_blocks": 2}, "rst")):
                if (w, h) in ((1, 1), (9, 3)) and t2 != "std":
                    continue
                jpeg("p%d_%dx%d_%s_%s" % (k, w, h, tag, t2), smooth(rng, h, w, 3), quality=(92, 75, 40)[k % 3], subsampling=sub,
                     progressive=True, **kw)
                k += 1
    jpeg("p_noise_420", rng.integers(0, 256, (24, 40, 3), dtype=np.uint8), quality=95, subsampling=2, progressive=True)
    jpeg("p_grey_33x20", smooth(rng, 20, 33, 1)[..., 0], quality=90, progressive=True)
    ok, buf = cv2.imencode(".jpg", smooth(rng, 21, 30, 3), [cv2.IMWRITE_JPEG_QUALITY, 80, cv2.IMWRITE_JPEG_PROGRESSIVE, 1,
                                                            cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_440])
    assert ok
    files["p_cv_440_30x21.jpg"] = buf.tobytes()
    b = io.BytesIO()
    Image.fromarray(smooth(rng, 11, 13, 3)).save(b, format="PNG")
    files["rgb_13x11.png"] = b.getvalue()
    b = io.BytesIO()
    Image.fromarray(np.concatenate([smooth(rng, 6, 9, 3), smooth(rng, 6, 9, 1)], -1), "RGBA").save(b, format="PNG")
    files["rgba_9x6.png"] = b.getvalue()
    # PNG: every colour type x bit depth, plain and Adam7 (sizes below 8 leave some passes empty), tRNS on every type
    for (w, h) in ((13, 9), (3, 2), (1, 1), (8, 8)):
        for il in (0, 1):
            tag = "%dx%d_%s" % (w, h, "adam7" if il else "plain")
            for depth in (1, 2, 4, 8, 16):
                top = 1 << depth
                files["g%d_%s.png" % (depth, tag)] = png_bytes(rng.integers(0, top, (h, w, 1)), 0, depth, il)
                if depth <= 8:
                    pal = rng.integers(0, 256, 3 * min(top, 256)).astype(np.uint8)
                    files["p%d_%s.png" % (depth, tag)] = png_bytes(rng.integers(0, min(top, 256), (h, w, 1)), 3, depth, il, palette=pal,
                                                                   trns=rng.integers(0, 256, min(top, 256) // 2 + 1).astype(np.uint8))
                if depth >= 8:
                    files["rgb%d_%s.png" % (depth, tag)] = png_bytes(rng.integers(0, top, (h, w, 3)), 2, depth, il)
                    files["ga%d_%s.png" % (depth, tag)] = png_bytes(rng.integers(0, top, (h, w, 2)), 4, depth, il)
                    files["rgba%d_%s.png" % (depth, tag)] = png_bytes(rng.integers(0, top, (h, w, 4)), 6, depth, il)
    few = rng.integers(0, 3, (9, 13, 3))
    files["rgb8_trns.png"] = png_bytes(few * 100, 2, 8, 0, trns=[0, 100, 0, 200, 0, 0])
    files["rgb16_trns.png"] = png_bytes(few * 30000, 2, 16, 1, trns=[0x75, 0x30, 0xea, 0x60, 0, 0])
    for depth in (1, 2, 4, 8, 16):
        v = rng.integers(0, min(1 << depth, 4), (9, 13, 1))
        files["g%d_trns.png" % depth] = png_bytes(v, 0, depth, depth & 1, trns=[0, 1])
    # TGA: 15 / 16-bit colour, grey + alpha, colour-mapped (24 / 32 / 16-bit palettes, 8- and 16-bit indices, an index
    # past the palette), run-length encoded forms, top-down, an image id field
    import struct

    def tga(name, type_, w, h, bits, data, desc=0, cmap=None, cmap_bits=24, idlen=0, cmap_first=0):
        n_pal = 0 if cmap is None else len(cmap) // ((cmap_bits + 7) // 8)
        hdr = bytes([idlen, 1 if cmap is not None else 0, type_]) + struct.pack("<HHB", cmap_first, n_pal, cmap_bits if cmap is not None else 0)
        hdr += struct.pack("<HHHHBB", 0, 0, w, h, bits, desc)
        files[name + ".tga"] = hdr + bytes(idlen) + (bytes(cmap) if cmap is not None else b"") + bytes(data)

    def rle(units, size):
        out, i = bytearray(), 0
        while i < len(units):
            n = min(3, len(units) - i)
            out += bytes([n - 1]) + b"".join(units[i:i + n])
            i += n
            if i < len(units):
                n = min(5, len(units) - i)
                out += bytes([128 | (n - 1)]) + units[i]
                i += n
        return out

    w, h = 7, 5
    tga("t_rgb16", 2, w, h, 16, rng.integers(0, 65536, w * h, dtype=np.uint16).tobytes())
    tga("t_rgb15", 2, w, h, 15, rng.integers(0, 65536, w * h, dtype=np.uint16).tobytes())
    tga("t_ga16", 3, w, h, 16, rng.integers(0, 256, w * h * 2, dtype=np.uint8).tobytes())
    tga("t_g8_top_id", 3, w, h, 8, rng.integers(0, 256, w * h, dtype=np.uint8).tobytes(), desc=0x20, idlen=9)
    for cb in (24, 32, 16, 15):
        pal = rng.integers(0, 256, ((cb + 7) // 8) * 16, dtype=np.uint8)
        tga("t_cm8_pal%d" % cb, 1, w, h, 8, rng.integers(0, 20, w * h, dtype=np.uint8).tobytes(), cmap=pal, cmap_bits=cb)   # 16..19: past the palette
        idx = [bytes([int(v)]) for v in rng.integers(0, 16, w * h)]
        tga("t_cm8_rle_pal%d" % cb, 9, w, h, 8, rle(idx, 1), cmap=pal, cmap_bits=cb)
    pal = rng.integers(0, 256, 3 * 300, dtype=np.uint8)
    tga("t_cm16_pal24", 1, w, h, 16, rng.integers(0, 300, w * h).astype("<u2").tobytes(), cmap=pal, cmap_bits=24)
    tga("t_rgba32_rle_top", 10, w, h, 32, rle([rng.integers(0, 256, 4, dtype=np.uint8).tobytes() for _ in range(w * h)], 4), desc=0x28)
    tga("t_rgb16_rle", 10, w, h, 16, rle([rng.integers(0, 256, 2, dtype=np.uint8).tobytes() for _ in range(w * h)], 2))
    tga("t_g8_rle", 11, w, h, 8, rle([bytes([int(v)]) for v in rng.integers(0, 256, w * h)], 1))
    # BMP: OS/2 (12-byte) and Windows (40 / 56 / 108 / 124-byte) headers; 4 / 8-bit palette, 16-bit 5-5-5 and 5-6-5 and odd
    # masks, 24-bit, 32-bit with alpha, with all-zero alpha, with explicit masks up to bit 31; bottom-up and top-down;
    # row padding (widths 5, 6, 7)
    def bmp(name, w, h, bpp, rows, hsz=40, compress=0, masks=(), pal=None, top_down=False):
        pal_bytes = b"" if pal is None else bytes(pal)
        extra = b"".join(struct.pack("<I", m) for m in masks)
        if hsz == 12:
            info = struct.pack("<IHHHH", 12, w, h, 1, bpp)
        else:
            info = struct.pack("<IiiHHIIiiII", hsz, w, -h if top_down else h, 1, bpp, compress, 0, 2835, 2835, 0, 0)
            if hsz == 40:
                info += extra[:12] if compress == 3 else b""
            elif hsz == 56:
                info += struct.pack("<IIII", *(list(masks) + [0] * 4)[:4])
            else:
                info += struct.pack("<IIII", *(list(masks) + [0] * 4)[:4]) + bytes(hsz - 40 - 16)
        offset = 14 + len(info) + len(pal_bytes)
        body = b"".join(r + bytes((-len(r)) & 3) for r in rows)
        files[name + ".bmp"] = b"BM" + struct.pack("<IHHI", offset + len(body), 0, 0, offset) + info + pal_bytes + body

    def packed(vals, bpp):
        if bpp == 8:
            return bytes(int(v) for v in vals)
        v = list(vals) + [0]
        return bytes((int(v[i]) << 4) | int(v[i + 1]) for i in range(0, len(vals), 2))

    for w in (5, 6, 7):
        h = 4
        pal3 = rng.integers(0, 256, 3 * 16, dtype=np.uint8)
        pal4 = rng.integers(0, 256, 4 * 16, dtype=np.uint8)
        for bpp in (4, 8):
            idx = rng.integers(0, 16, (h, w))
            bmp("b_os2_p%d_w%d" % (bpp, w), w, h, bpp, [packed(r, bpp) for r in idx], hsz=12, pal=pal3)
            bmp("b_p%d_w%d" % (bpp, w), w, h, bpp, [packed(r, bpp) for r in idx], pal=pal4, top_down=(w == 6))
        px16 = rng.integers(0, 65536, (h, w)).astype("<u2")
        bmp("b_555_w%d" % w, w, h, 16, [r.tobytes() for r in px16])
        bmp("b_565_w%d" % w, w, h, 16, [r.tobytes() for r in px16], compress=3, masks=(0xf800, 0x07e0, 0x001f))
        bmp("b_odd16_w%d" % w, w, h, 16, [r.tobytes() for r in px16], compress=3, masks=(0x7000, 0x0f80, 0x0003))
        px24 = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        bmp("b_24_w%d" % w, w, h, 24, [r.tobytes() for r in px24], top_down=(w == 5))
        bmp("b_os2_24_w%d" % w, w, h, 24, [r.tobytes() for r in px24], hsz=12)
        px32 = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        bmp("b_32_w%d" % w, w, h, 32, [r.tobytes() for r in px32])
        zero_a = px32.copy()
        zero_a[..., 3] = 0
        bmp("b_32_zero_alpha_w%d" % w, w, h, 32, [r.tobytes() for r in zero_a])
        raw32 = rng.integers(0, 2 ** 32, (h, w), dtype=np.uint64).astype("<u4")
        bmp("b_32_masks_w%d" % w, w, h, 32, [r.tobytes() for r in raw32], compress=3, masks=(0xffc00000, 0x003ff000, 0x00000ffc))
        bmp("b_v4_w%d" % w, w, h, 32, [r.tobytes() for r in raw32], hsz=108, compress=3, masks=(0x00ff0000, 0x0000ff00, 0x000000ff, 0xff000000))
        bmp("b_v5_w%d" % w, w, h, 32, [r.tobytes() for r in raw32], hsz=124, compress=3, masks=(0x000000ff, 0x0000ff00, 0x00ff0000, 0xf0000000))
        bmp("b_v3_56_w%d" % w, w, h, 16, [r.tobytes() for r in px16], hsz=56)
    b = io.BytesIO()
    Image.fromarray(smooth(rng, 7, 12, 3)).save(b, format="BMP")
    files["pil_rgb.bmp"] = b.getvalue()
    b = io.BytesIO()
    Image.fromarray(smooth(rng, 7, 12, 3)).save(b, format="TGA")
    files["rgb_12x7.tga"] = b.getvalue()
    names = sorted(files)
    out = {"names": np.array(names)}
    for i, n in enumerate(names):
        out["file_%d" % i] = np.frombuffer(files[n], np.uint8)
        out["rgba_%d" % i] = reference_texels(files[n], os.path.splitext(n)[1])
    np.savez_compressed(os.path.join(HERE, "textures.npz"), **out)
    print("textures.npz: %d files, %d bytes of image files" % (len(names), sum(len(v) for v in files.values())))


if __name__ == "__main__":
    main()
