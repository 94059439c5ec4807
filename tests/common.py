"""Shared helpers for the parity tests."""
import importlib
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("akarirender-1_b200")
scenes, abi, nodes = pkg.scenes, pkg.abi, pkg.nodes
GOLDEN = os.path.join(ROOT, "tests", "golden", "reference_vectors.npz")


def golden():
    return np.load(GOLDEN)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def same_bits(a, b):
    return np.array_equal(bits(a), bits(b))


def hit_mismatch(a, b):
    """Closest-hit comparison per the contract: prim/geom ids bit-exact except exact-t ties; t/u/v bit-exact off ties.
    Returns (n_id_mismatch_not_tie, n_ties, n_t_bit_mismatch)."""
    id_bad = (a["prim_id"] != b["prim_id"]) | (a["geom_id"] != b["geom_id"])
    t_bad = bits(a["t"]) != bits(b["t"])
    ties = id_bad & ~t_bad
    return int((id_bad & t_bad).sum()), int(ties.sum()), int(t_bad.sum())


def golden_scenes():
    return {
        "cornell": scenes.cornell_box((1024, 1024), 16, 5),
        "cornell32": scenes.cornell_box((32, 32), 16, 5),
        "mixed32": scenes.mixed_box((32, 32), 8, 5),
        "mixed24_pcg": scenes.mixed_box((24, 24), 4, 3, sampler="pcg"),
        "terrain": scenes.terrain_scene(n=48, resolution=(64, 36), spp=4, max_depth=5, materials="c3ref"),
    }


def rel_rmse(img, ref):
    """sqrt(mean((I-R)^2 / (R^2 + 1e-2))) over RGB of radiance/weight."""
    a = img[..., :3] / np.maximum(img[..., 3:4], 1)
    b = ref[..., :3] / np.maximum(ref[..., 3:4], 1)
    return float(np.sqrt(np.mean((a - b) ** 2 / (b ** 2 + 1e-2))))


def random_rays(rng, n, lo, hi):
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    org = np.concatenate([o, np.full((n, 1), 1e-3, np.float32)], 1).astype(np.float32)
    dr = np.concatenate([d, np.full((n, 1), np.inf, np.float32)], 1).astype(np.float32)
    return org, dr
