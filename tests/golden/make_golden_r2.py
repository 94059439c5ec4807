"""Generates tests/golden/reference_vectors_r2.npz by running the REFERENCE's own code (oracle/_ref/libakari_ref.so):

  * a multi-mesh scene (4 overlapping placed copies of a small M1, per-mesh material tables; plus a variant with a
    fifth mesh coincident with the first): closest hits with geom_id / prim_id, any-hit flags, a Path film
    (kernel/bvh-accelerator.h:651-678, kernel/scene.h:70-90) -- BASELINE config C4's shape;
  * a many-light scene with 10 000 emissive triangles (config C5's light count; 14-step CDF search), no Russian
    roulette (the reference has none): Path film, depth 16;
  * the byte stage: Film::write_image -> GammaCorrection -> DefaultImageWriter (core/film.h:97-113,
    core/image.cpp:38-71) of golden films, PNG decoded back to RGB8 with the reference's own reader.

Run in the build container only:  make -C oracle ref && python tests/golden/make_golden_r2.py
"""
import ctypes as C
import importlib
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("akarirender-1_b200")
from oracle_libs import REF_SO, Reference  # noqa: E402
from make_golden import rays_for, secondary  # noqa: E402


def r2_scenes():
    s = pkg.scenes
    return {
        "multi": s.multi_mesh_scene(n=20, resolution=(64, 36), spp=4, max_depth=5),
        "multi_co": s.multi_mesh_scene(n=20, resolution=(64, 36), spp=4, max_depth=5, coincident=True),
        "lights10k": s.many_light_scene(n=48, resolution=(48, 27), spp=4, max_depth=16, n_quads=5000, rr_start_depth=0),
    }


def boundary_rays(rng, n):
    """Rays skimming the strips where neighbouring copies overlap (x or z near 0), in both directions."""
    o = np.stack([rng.uniform(-12, 12, n), rng.uniform(-2, 6, n), rng.uniform(-95, 95, n)], -1)
    d = np.stack([rng.normal(0, 1, n), rng.normal(0, 0.15, n), rng.normal(0, 1, n)], -1)
    swap = rng.random(n) < 0.5
    o[swap] = o[swap][:, [2, 1, 0]]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    org = np.concatenate([o, np.full((n, 1), 1e-3)], 1).astype(np.float32)
    dr = np.concatenate([d, np.full((n, 1), np.inf)], 1).astype(np.float32)
    return org, dr


def srgb8_via_reference(film):
    """Film::write_image on (sums, weights) -> PNG -> RGB8, all with the reference's own code."""
    lib = C.CDLL(REF_SO)
    h, w = film.shape[:2]
    f = np.ascontiguousarray(film, np.float32)
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "out.png").encode()
        assert lib.akref_write_image(f.ctypes.data_as(C.c_void_p), w, h, path) == 0
        ww, hh = C.c_int(), C.c_int()
        out = np.zeros((h, w, 4), np.float32)
        assert lib.akref_read_image(path, C.byref(ww), C.byref(hh), out.ctypes.data_as(C.c_void_p), C.c_size_t(out.size)) == 0
        assert (ww.value, hh.value) == (w, h)
    b = np.rint(out[..., :3] * 255.0)
    assert np.array_equal((b / 255.0).astype(np.float32), out[..., :3])  # reader stores byte / 255.0f
    return b.astype(np.uint8)


def hdr_files(rng):
    """Radiance .hdr files as bytes: flat pixels (width < 8), new-style RLE scanlines with runs and dumps, and the case
    stb treats specially - a width in the RLE range whose data is not run-length encoded."""
    def header(w, h, ident=b"#?RADIANCE"):
        return ident + b"\n# made by make_golden_r2.py\nFORMAT=32-bit_rle_rgbe\nEXPOSURE=1.0\n\n-Y %d +X %d\n" % (h, w)

    def rgbe(w, h):
        px = rng.integers(0, 256, (h, w, 4)).astype(np.uint8)
        px[..., 3] = rng.integers(118, 140, (h, w))  # exponents around 1.0
        px[0, 0, 3] = 0                              # e = 0 -> black
        px[..., :3][px[..., :3].max(-1) < 128] |= 128  # a valid RGBE pixel has one mantissa >= 128
        return px

    files = {}
    p = rgbe(5, 4)
    files["flat_5x4"] = header(5, 4) + p.tobytes()
    p = rgbe(19, 6)
    p[2, 3:15] = p[2, 3]  # a long run
    body = b""
    for row in p:
        body += bytes([2, 2, 0, 19])
        for k in range(4):
            ch, i = row[:, k], 0
            while i < 19:
                j = i
                while j + 1 < 19 and ch[j + 1] == ch[i] and j - i < 126:
                    j += 1
                if j - i >= 2:
                    body += bytes([128 + (j - i + 1), ch[i]])
                    i = j + 1
                else:
                    n = min(19 - i, int(rng.integers(1, 6)))
                    body += bytes([n]) + ch[i:i + n].tobytes()
                    i += n
    files["rle_19x6"] = header(19, 6, b"#?RGBE") + body
    p = rgbe(9, 3)
    files["flatdata_9x3"] = header(9, 3) + p.tobytes()
    return files


def main():
    out = {}
    rng = np.random.default_rng(20261105)
    S = r2_scenes()
    for name in ("multi", "multi_co"):
        ref = Reference(S[name])
        xy, u2, org, dr = rays_for(ref, 64, 36, 1, rng)
        hits = ref.trace(org, dr)
        o2, d2 = secondary(org, dr, hits, rng)
        o3, d3 = boundary_rays(rng, 6000)
        org_all, dir_all = np.concatenate([org, o2, o3]), np.concatenate([dr, d2, d3])
        da = dir_all.copy()
        da[:, 3] = 60.0
        out.update({f"{name}_org": org_all, f"{name}_dir": dir_all, f"{name}_hits": ref.trace(org_all, dir_all),
                    f"{name}_dir_any": da, f"{name}_occ": ref.trace(org_all, da, any_hit=True)})
        h = out[f"{name}_hits"]
        print(name, "hit fraction", (h["prim_id"] >= 0).mean(), "geoms hit", np.unique(h["geom_id"]))
    out["multi_film"] = Reference(S["multi"]).render(S["multi"].integrator.params())
    lref = Reference(S["lights10k"])
    assert lref.n_lights() == 10000
    out["lights10k_film"] = lref.render(S["lights10k"].integrator.params())
    # ---- C1 at its full size (Cornell 512 x 512, 16 spp, depth 5): the reference's film against the brute-force oracle.
    # 4 MB of floats are not committed; what is: a hash of the brute-force oracle's film, the pixels where the
    # reference's film differs from it, and for each of those the one ray that explains it with both answers (an
    # exact-t tie between two triangles, or a hit the reference's unpadded slab test culled: DESIGN.md section 2).
    import hashlib
    from oracle_libs import Oracle
    c1 = pkg.scenes.cornell_box((512, 512), 16, 5)
    p1 = c1.integrator.params()
    ob, rf = Oracle(c1, brute_force=True), Reference(c1)
    fb, fr = ob.render(p1), rf.render(p1)
    diff = np.argwhere((fb.view(np.uint32) != fr.view(np.uint32)).any(-1))
    rays_o, rays_d, ans_b, ans_r = [], [], [], []
    for y, x in diff:
        anyhit, org, dr = ob.pixel_rays(p1, x, y)
        cl = ~anyhit
        hb, hr = ob.trace(org[cl], dr[cl]), rf.trace(org[cl], dr[cl])
        bad = np.flatnonzero(hb != hr)
        assert len(bad) >= 1, "difference not explained by a closest-hit ray"
        i = bad[0]
        rays_o.append(org[cl][i]), rays_d.append(dr[cl][i]), ans_b.append(hb[i]), ans_r.append(hr[i])
    out.update(c1_full_brute_sha256=np.frombuffer(hashlib.sha256(fb.tobytes()).digest(), np.uint8),
               c1_full_ref_diff_yx=diff.astype(np.int32), c1_full_ref_diff_vals=np.stack([fr[y, x] for y, x in diff]),
               c1_full_diff_org=np.stack(rays_o), c1_full_diff_dir=np.stack(rays_d),
               c1_full_diff_hit_brute=np.array(ans_b), c1_full_diff_hit_ref=np.array(ans_r))
    print("C1 full frame: reference differs from the brute-force oracle in", len(diff), "pixels:", diff.tolist())
    for b, r in zip(ans_b, ans_r):
        print("   brute", b, "reference", r)
    # ---- .hdr image textures: the reference's reader (stbi_loadf, core/image.cpp:85-103) on small generated files
    lib = C.CDLL(REF_SO)
    for name, data in hdr_files(rng).items():
        with tempfile.TemporaryDirectory() as td:
            path = os.path.join(td, name + ".hdr")
            open(path, "wb").write(data)
            ww, hh = C.c_int(), C.c_int()
            assert lib.akref_read_image(path.encode(), C.byref(ww), C.byref(hh), None, C.c_size_t(0)) == 0
            tex = np.zeros((hh.value, ww.value, 4), np.float32)
            assert lib.akref_read_image(path.encode(), C.byref(ww), C.byref(hh), tex.ctypes.data_as(C.c_void_p), C.c_size_t(tex.size)) == 0
        out["hdr_" + name + "_file"] = np.frombuffer(data, np.uint8)
        out["hdr_" + name + "_rgba"] = tex
    # ---- byte stage: the reference's own writer on golden films (incl. values <0.0031308, >1 and weight 0)
    G = np.load(os.path.join(HERE, "reference_vectors.npz"))
    for key in ("cornell32_film", "mixed32_film", "terrain_film"):
        out["srgb8_" + key] = srgb8_via_reference(G[key])
    ramp = np.zeros((64, 64, 4), np.float32)
    v = np.concatenate([np.linspace(0, 0.01, 1024), np.linspace(0.01, 1.2, 3072)]).astype(np.float32).reshape(64, 64)
    ramp[..., 0], ramp[..., 1], ramp[..., 2] = v * 3, v * 7, v * 5
    ramp[..., 3] = np.array([3, 7, 5, 0], np.float32)[np.arange(64) % 4][None, :]
    ramp[..., :3][ramp[..., 3] == 0] = v[ramp[..., 3] == 0][:, None]
    out["srgb8_ramp_in"], out["srgb8_ramp"] = ramp, srgb8_via_reference(ramp)
    np.savez_compressed(os.path.join(HERE, "reference_vectors_r2.npz"), **out)
    print("wrote reference_vectors_r2.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
