"""Golden texels for JPEG texture files: small JPEG files written by Pillow, decoded by the REFERENCE's own
image reader (oracle/_ref: DefaultImageReader::read -> stbi_load(..., 3) -> byte / 255.0f, core/image.cpp:105-121).
Run in the build container (needs /root/reference through oracle/_ref and Pillow); the files and the texels are
committed as tests/golden/reference_vectors_jpeg.npz so that the tests need neither.

    python tests/golden/make_golden_jpeg.py
"""
import ctypes as C
import io
import os
import tempfile

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libakari_ref.so")


def picture(rng, h, w):
    """Smooth gradients + a hard-edged pattern + noise: exercises DC prediction, long AC runs, EOB runs and clamping."""
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.stack([128 + 120 * np.sin(x / 5.0) * np.cos(y / 7.0), 255 * ((x // 4 + y // 3) % 2), 40 + 200 * (x + y) / max(1, w + h - 2)], -1)
    img += rng.normal(0, 12, img.shape)
    img[: h // 3, : w // 3] = (255, 0, 255)  # saturated block: IDCT overshoot is clamped
    return np.clip(img, 0, 255).astype(np.uint8)


def jpeg_files(rng):
    out = {}

    def add(name, arr, mode="RGB", **kw):
        b = io.BytesIO()
        Image.fromarray(arr, mode).save(b, "JPEG", **kw)
        out[name] = b.getvalue()

    add("base444_37x29", picture(rng, 29, 37), quality=90, subsampling=0)
    add("base420_37x29", picture(rng, 29, 37), quality=85, subsampling=2)
    add("base420_64x48_q30", picture(rng, 48, 64), quality=30, subsampling=2)
    add("base422_50x21", picture(rng, 21, 50), quality=80, subsampling=1)
    add("base411_45x19", picture(rng, 19, 45), quality=80, subsampling="4:1:1")
    add("base420_1x9", picture(rng, 9, 1), quality=90, subsampling=2)       # one chroma sample per row: no interpolation
    add("base420_2x2", picture(rng, 2, 2), quality=90, subsampling=2)
    add("base420_17x1", picture(rng, 1, 17), quality=90, subsampling=2)
    add("base444_q100_16x16", picture(rng, 16, 16), quality=100, subsampling=0)
    add("optimized420_40x33", picture(rng, 33, 40), quality=75, subsampling=2, optimize=True)  # file-specific Huffman tables
    add("restart420_61x43", picture(rng, 43, 61), quality=75, subsampling=2, restart_marker_blocks=3)
    add("restart444_rows_33x31", picture(rng, 31, 33), quality=75, subsampling=0, restart_marker_rows=1)
    add("prog420_61x43", picture(rng, 43, 61), quality=80, subsampling=2, progressive=True)
    add("prog444_35x20", picture(rng, 20, 35), quality=92, subsampling=0, progressive=True)
    add("prog420_restart_52x37", picture(rng, 37, 52), quality=60, subsampling=2, progressive=True, restart_marker_blocks=2)
    add("prog_grey_29x23", picture(rng, 23, 29)[..., 0].copy(), "L", quality=85, progressive=True)
    add("grey_33x17", picture(rng, 17, 33)[..., 1].copy(), "L", quality=85)
    cmyk = np.concatenate([picture(rng, 14, 22), picture(rng, 14, 22)[..., :1]], -1)
    add("cmyk_22x14", cmyk, "CMYK", quality=85)
    return out


def main():
    rng = np.random.default_rng(20260211)
    lib = C.CDLL(REF_SO)
    out = {}
    for name, data in jpeg_files(rng).items():
        with tempfile.TemporaryDirectory() as td:
            path = os.path.join(td, name + ".jpg")
            open(path, "wb").write(data)
            ww, hh = C.c_int(), C.c_int()
            assert lib.akref_read_image(path.encode(), C.byref(ww), C.byref(hh), None, C.c_size_t(0)) == 0, name
            tex = np.zeros((hh.value, ww.value, 4), np.float32)
            assert lib.akref_read_image(path.encode(), C.byref(ww), C.byref(hh), tex.ctypes.data_as(C.c_void_p), C.c_size_t(tex.size)) == 0
        out["jpeg_" + name + "_file"] = np.frombuffer(data, np.uint8)
        out["jpeg_" + name + "_rgba"] = tex
        print(name, len(data), "bytes ->", tex.shape, "mean", tex[..., :3].mean())
    np.savez_compressed(os.path.join(HERE, "reference_vectors_jpeg.npz"), **out)


if __name__ == "__main__":
    main()
