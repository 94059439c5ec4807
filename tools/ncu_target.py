"""Small fixed workload for ncu captures: C3 scene (999 700 triangles, 1080p) at QP_SPP spp through the wavefront, plus
one coherent and one incoherent closest-hit batch. Run it plain first, then under ncu (B200_PROFILING.md)."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")


def main():
    spp = int(os.environ.get("QP_SPP", "4"))
    sc = pkg.scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=spp, max_depth=8, materials="c3")
    be = pkg.Backend(sc, 0)
    W, H = 1920, 1080
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    p = sc.integrator.params()
    p.mode = pkg.abi.MODE_MEGAKERNEL if os.environ.get("QP_MODE") == "mega" else pkg.abi.MODE_WAVEFRONT
    for _ in range(2):
        st = be.render_device(film.data_ptr(), p)
    print("render: %.2f ms, %d samples, %d+%d rays, %d launches" % (st.ms_total, st.n_samples, st.n_closest_rays, st.n_shadow_rays,
                                                                  st.n_kernel_launches))
    ys, xs = np.mgrid[0:H, 0:W]
    xy = np.stack([xs.ravel(), ys.ravel()], -1).astype(np.int32)
    org, dr = be.generate_rays(xy, np.full((xy.shape[0], 2), 0.5, np.float32))
    n = org.shape[0]
    d_o, d_d = torch.from_numpy(org).cuda(), torch.from_numpy(dr).cuda()
    d_h = torch.empty(n * 5, dtype=torch.int32, device="cuda")
    d_occ = torch.empty(n, dtype=torch.int32, device="cuda")
    ms = be.trace_device(d_o.data_ptr(), d_d.data_ptr(), n, 0, d_h.data_ptr(), d_occ.data_ptr())
    hits = d_h.cpu().numpy().view(pkg.HIT_DTYPE)
    rng = np.random.default_rng(1)
    hit = hits["prim_id"] >= 0
    pnt = org[:, :3] + dr[:, :3] * np.where(hit, hits["t"], 10)[:, None]
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:, 1] = np.abs(d[:, 1])
    o2 = np.concatenate([pnt + np.array([0, 1e-2, 0]), np.full((n, 1), 1e-3)], 1).astype(np.float32)
    d2 = np.concatenate([d, np.full((n, 1), np.inf)], 1).astype(np.float32)
    d_o2, d_d2 = torch.from_numpy(o2).cuda(), torch.from_numpy(d2).cuda()
    ms2 = be.trace_device(d_o2.data_ptr(), d_d2.data_ptr(), n, 0, d_h.data_ptr(), d_occ.data_ptr())
    print("trace: coherent %.3f ms, incoherent %.3f ms (%d rays)" % (ms, ms2, n))


if __name__ == "__main__":
    main()
