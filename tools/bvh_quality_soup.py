"""Tree quality of the two builders on a non-terrain scene: clustered triangle soup, sizes over three decades
(development aid). Prints build time, node count and closest-hit / any-hit rates for random rays."""
import importlib
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")
abi, scenes = pkg.abi, pkg.scenes


def main():
    nt = int(os.environ.get("QP_NT", "1000000"))
    rng = np.random.default_rng(7)
    centres = rng.normal(0, 20, (200, 3))
    c = centres[rng.integers(0, 200, nt)] + rng.normal(0, 2.0, (nt, 3))
    size = 10 ** rng.uniform(-2.5, 0.3, (nt, 1, 1))
    v = (c[:, None, :] + rng.normal(0, 1, (nt, 3, 3)) * size).astype(np.float32)
    mesh = scenes.AkariMesh(materials=[scenes.DiffuseMaterial(0.5)], vertices=v.reshape(-1, 3),
                            indices=np.arange(3 * nt, dtype=np.int32), material_indices=np.zeros(nt, np.int32))
    sc = scenes.Scene(scenes.PerspectiveCamera(fov=60, resolution=(16, 16)), scenes.Path(spp=1, ray_clamp=0.0), [mesh])
    n = 1 << 22
    o = rng.uniform(-50, 50, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    org = torch.from_numpy(np.concatenate([o, np.full((n, 1), 1e-3, np.float32)], 1)).cuda()
    dr = torch.from_numpy(np.concatenate([d, np.full((n, 1), np.inf, np.float32)], 1)).cuda()
    hits = torch.zeros(n * 5, dtype=torch.int32, device="cuda")
    occ = torch.zeros(n, dtype=torch.int32, device="cuda")
    ref = None
    for name, kind in (("host SAH", abi.BVH_HOST_SAH), ("device PLOC", abi.BVH_DEVICE)):
        t = time.time()
        be = pkg.Backend(sc, 0, bvh_build=kind)
        info = be.info()
        for _ in range(3):
            ms_c = be.trace_device(org.data_ptr(), dr.data_ptr(), n, 0, hits.data_ptr(), occ.data_ptr())
        h = hits.clone()
        for _ in range(3):
            ms_a = be.trace_device(org.data_ptr(), dr.data_ptr(), n, 1, hits.data_ptr(), occ.data_ptr())
        print("%-12s BVH %.3f s  nodes %d depth %d  closest %.1f Mrays/s  any-hit %.1f Mrays/s  hit fraction %.2f"
              % (name, info["build_us"] / 1e6, info["n_bvh_nodes"], info["bvh_depth"], n / ms_c / 1e3, n / ms_a / 1e3,
                 float((h.view(n, 5)[:, 3] >= 0).float().mean())), flush=True)
        if ref is None:
            ref = h
        else:
            print("hits identical:", bool((ref == h).all()))
        be.close()


if __name__ == "__main__":
    main()
