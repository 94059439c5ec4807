"""Generates tests/golden/reference_vectors.npz by running the REFERENCE's own code (oracle/_ref/libakari_ref.so, built
from /root/reference by oracle/Makefile) on small fixed inputs. Run in the build container only:

    make -C oracle ref && python tests/golden/make_golden.py

The committed .npz is what pins the oracle (and, through it, the CUDA path) on machines where /root/reference does
not exist. Inputs are stored next to the outputs so the tests do not depend on scene-generation code staying fixed.
"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("akarirender-1_b200")
from oracle_libs import Reference  # noqa: E402


def rays_for(ref, w, h, step, rng):
    ys, xs = np.mgrid[0:h:step, 0:w:step]
    xy = np.stack([xs.ravel(), ys.ravel()], -1).astype(np.int32)
    u2 = rng.random((xy.shape[0], 2), dtype=np.float32)
    org, dr = ref.generate_rays(xy, u2)
    return xy, u2, org, dr


def secondary(org, dr, hits, rng):
    """Incoherent rays leaving the primary hit points (misses get random origins)."""
    n = len(hits)
    t = np.where(hits["prim_id"] >= 0, hits["t"], 5.0).astype(np.float32)
    p = org[:, :3] + dr[:, :3] * t[:, None] * np.float32(0.999)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o2 = np.concatenate([p, np.full((n, 1), 1e-3, np.float32)], 1).astype(np.float32)
    d2 = np.concatenate([d, np.full((n, 1), np.inf, np.float32)], 1).astype(np.float32)
    return o2, d2


def main():
    out = {}
    rng = np.random.default_rng(20261104)
    # --- sampler KATs (kernel/sampler.h)
    sc = pkg.scenes.cornell_box((1024, 1024), 16, 5)
    ref = Reference(sc)
    for kind, name in ((0, "lcg"), (1, "pcg")):
        for seed in (0, 7, 123456, 0xFFFFFFF0):
            d, n2 = ref.sampler_kat(kind, seed, 8)
            out[f"kat_{name}_{seed}_draws"], out[f"kat_{name}_{seed}_next2d"] = d, n2
    # --- Cornell 1024^2: camera rays, closest hits, any-hit, secondary rays
    xy, u2, org, dr = rays_for(ref, 1024, 1024, 32, rng)
    hits = ref.trace(org, dr)
    o2, d2 = secondary(org, dr, hits, rng)
    d2a = d2.copy()
    d2a[:, 3] = 1.5
    out.update(cornell_xy=xy, cornell_u2=u2, cornell_org=org, cornell_dir=dr, cornell_hits=hits,
               cornell_org2=o2, cornell_dir2=d2, cornell_hits2=ref.trace(o2, d2),
               cornell_dir2_any=d2a, cornell_occ2=ref.trace(o2, d2a, any_hit=True))
    # --- Cornell 32x32 16 spp film (Path, depth 5, LCG) and the survey's two 1024^2 pixel sums
    sc32 = pkg.scenes.cornell_box((32, 32), 16, 5)
    out["cornell32_film"] = Reference(sc32).render(sc32.integrator.params())
    film = ref.render(sc.integrator.params(), rect=(512, 512, 513, 513))
    out["cornell1024_px_512_512"] = film[512, 512]
    film = ref.render(sc.integrator.params(), rect=(256, 768, 257, 769))
    out["cornell1024_px_256_768"] = film[768, 256]
    # --- every reference material kind (Diffuse, Glossy, nested Mix, Emissive)
    mb = pkg.scenes.mixed_box((32, 32), 8, 5)
    out["mixed32_film"] = Reference(mb).render(mb.integrator.params())
    mbp = pkg.scenes.mixed_box((24, 24), 4, 3, sampler="pcg")
    out["mixed24_pcg_film"] = Reference(mbp).render(mbp.integrator.params())
    # --- small terrain (M1 family, 4418 triangles): rays + film
    ts = pkg.scenes.terrain_scene(n=48, resolution=(64, 36), spp=4, max_depth=5, materials="c3ref")
    tref = Reference(ts)
    xy, u2, org, dr = rays_for(tref, 64, 36, 1, rng)
    hits = tref.trace(org, dr)
    o2, d2 = secondary(org, dr, hits, rng)
    d2[:, 1] = np.abs(d2[:, 1]) * np.float32(0.05)  # grazing
    d2a = d2.copy()
    d2a[:, 3] = 50.0
    out.update(terrain_xy=xy, terrain_u2=u2, terrain_org=org, terrain_dir=dr, terrain_hits=hits,
               terrain_org2=o2, terrain_dir2=d2, terrain_hits2=tref.trace(o2, d2),
               terrain_dir2_any=d2a, terrain_occ2=tref.trace(o2, d2a, any_hit=True),
               terrain_film=tref.render(ts.integrator.params()))
    np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **out)
    print("wrote reference_vectors.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
