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
