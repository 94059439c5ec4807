"""Ad-hoc timing of the hot path on one GPU (development aid; bench.py is the contract)."""
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
    n = int(os.environ.get("QP_N", "708"))
    res = (1920, 1080)
    spp = int(os.environ.get("QP_SPP", "16"))
    t = time.time()
    sc = scenes.terrain_scene(n=n, resolution=res, spp=spp, max_depth=8, materials="c3")
    be = pkg.Backend(sc, 0)
    print("scene build %.2fs" % (time.time() - t), be.info(), flush=True)
    # primary + secondary rays on device
    W, H = res
    ys, xs = np.mgrid[0:H, 0:W]
    xy = np.stack([xs.ravel(), ys.ravel()], -1).astype(np.int32)
    org, dr = be.generate_rays(xy, np.full((xy.shape[0], 2), 0.5, np.float32))
    d_org, d_dir = torch.from_numpy(org).cuda(), torch.from_numpy(dr).cuda()
    nr = org.shape[0]
    d_hits = torch.empty(nr * 5, dtype=torch.int32, device="cuda")
    d_occ = torch.empty(nr, dtype=torch.int32, device="cuda")
    for rep in range(3):
        ms = be.trace_device(d_org.data_ptr(), d_dir.data_ptr(), nr, 0, d_hits.data_ptr(), d_occ.data_ptr())
    print("coherent closest: %.3f ms  %.1f Mrays/s" % (ms, nr / ms / 1e3), flush=True)
    hits = d_hits.cpu().numpy().view(pkg.HIT_DTYPE)
    rng = np.random.default_rng(1)
    hit = hits["prim_id"] >= 0
    p = org[:, :3] + dr[:, :3] * np.where(hit, hits["t"], 10)[:, None]
    d = rng.normal(size=(nr, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:, 1] = np.abs(d[:, 1])
    o2 = np.concatenate([p + np.array([0, 1e-2, 0]), np.full((nr, 1), 1e-3)], 1).astype(np.float32)
    d2 = np.concatenate([d, np.full((nr, 1), np.inf)], 1).astype(np.float32)
    d_org2, d_dir2 = torch.from_numpy(o2).cuda(), torch.from_numpy(d2).cuda()
    for rep in range(3):
        ms = be.trace_device(d_org2.data_ptr(), d_dir2.data_ptr(), nr, 0, d_hits.data_ptr(), d_occ.data_ptr())
    print("incoherent closest: %.3f ms  %.1f Mrays/s" % (ms, nr / ms / 1e3), flush=True)
    for rep in range(3):
        ms = be.trace_device(d_org2.data_ptr(), d_dir2.data_ptr(), nr, 1, d_hits.data_ptr(), d_occ.data_ptr())
    print("incoherent any-hit: %.3f ms  %.1f Mrays/s" % (ms, nr / ms / 1e3), flush=True)
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    for mode, sort in ((abi.MODE_WAVEFRONT, 1), (abi.MODE_WAVEFRONT, 0), (abi.MODE_MEGAKERNEL, 0)):
        p = sc.integrator.params()
        p.mode, p.sort_materials = mode, sort
        for rep in range(2):
            st = be.render_device(film.data_ptr(), p)
        rays = st.n_closest_rays + st.n_shadow_rays
        print("render mode=%d sort=%d: %.2f ms  %.1f Msamples/s  %.1f Mrays/s  rays/sample %.2f launches %d extend %.2f shade %.2f"
              % (mode, sort, st.ms_total, st.n_samples / st.ms_total / 1e3, rays / st.ms_total / 1e3, rays / max(1, st.n_samples),
                 st.n_kernel_launches, st.ms_extend, st.ms_shade), flush=True)
    # cornell
    sc = scenes.cornell_box((512, 512), 16, 5)
    be2 = pkg.Backend(sc, 0)
    film = torch.zeros(512 * 512 * 4, dtype=torch.float32, device="cuda")
    for mode in (abi.MODE_WAVEFRONT, abi.MODE_MEGAKERNEL):
        p = sc.integrator.params()
        p.mode = mode
        for rep in range(2):
            st = be2.render_device(film.data_ptr(), p)
        rays = st.n_closest_rays + st.n_shadow_rays
        print("cornell mode=%d: %.2f ms  %.1f Msamples/s  %.1f Mrays/s launches %d" % (mode, st.ms_total, st.n_samples / st.ms_total / 1e3,
                                                                                   rays / st.ms_total / 1e3, st.n_kernel_launches), flush=True)


if __name__ == "__main__":
    main()
