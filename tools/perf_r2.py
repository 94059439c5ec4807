"""Development aid: stage times of the wavefront on C3 (and optionally C1/C5) for A/B runs of library variants.

    AKB_LIB=/root/repo/akarirender-1_b200/libakari_b200_X.so python tools/perf_r2.py [c3] [c1] [c5] [rays]

Prints one line per workload: whole-frame ms (best of 3, normal scheduling), then the per-stage device times of one
profiled frame (akb_set_option profile_stages: one pipeline, no graphs), and the ray microbench."""
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
tag = os.path.basename(os.environ.get("AKB_LIB", "default")).replace("libakari_b200_", "").replace(".so", "")


def frame(sc, name, sort=0):
    be = pkg.Backend(sc, 0)
    W, H = sc.camera.resolution
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    p = sc.integrator.params()
    p.sort_materials = sort
    best = 1e9
    for _ in range(4):
        st = be.render_device(film.data_ptr(), p)
        best = min(best, st.ms_total)
    pkg.set_option("profile_stages", 1)
    be.render_device(film.data_ptr(), p)
    sp = be.render_device(film.data_ptr(), p)
    pkg.set_option("profile_stages", 0)
    h = int(torch.sum(film.view(torch.int32).long() & 0xFFFF).item())
    print("[%s] %s sort=%d: frame %.2f ms (%.1f Msamples/s) | profiled: total %.2f extend %.2f shade %.2f | rays/sample %.2f | film-hash %d"
          % (tag, name, sort, best, st.n_samples / best / 1e3, sp.ms_total, sp.ms_extend, sp.ms_shade,
             (st.n_closest_rays + st.n_shadow_rays) / max(1, st.n_samples), h), flush=True)
    return be


def rays(be, res):
    W, H = res
    ys, xs = np.mgrid[0:H, 0:W]
    xy = np.stack([xs.ravel(), ys.ravel()], -1).astype(np.int32)
    org, dr = be.generate_rays(xy, np.full((xy.shape[0], 2), 0.5, np.float32))
    d_org, d_dir = torch.from_numpy(org).cuda(), torch.from_numpy(dr).cuda()
    nr = org.shape[0]
    d_hits = torch.empty(nr * 5, dtype=torch.int32, device="cuda")
    d_occ = torch.empty(nr, dtype=torch.int32, device="cuda")

    def run(o, d, any_hit):
        return min(be.trace_device(o.data_ptr(), d.data_ptr(), nr, any_hit, d_hits.data_ptr(), d_occ.data_ptr()) for _ in range(4))
    ms_c = run(d_org, d_dir, 0)
    hits = d_hits.cpu().numpy().view(pkg.HIT_DTYPE)
    rng = np.random.default_rng(1)
    hit = hits["prim_id"] >= 0
    p = org[:, :3] + dr[:, :3] * np.where(hit, hits["t"], 10)[:, None]
    d = rng.normal(size=(nr, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:, 1] = np.abs(d[:, 1])
    o2 = np.concatenate([p + np.array([0, 1e-2, 0]), np.full((nr, 1), 1e-3)], 1).astype(np.float32)
    d2 = np.concatenate([d, np.full((nr, 1), np.inf)], 1).astype(np.float32)
    d_o2, d_d2 = torch.from_numpy(o2).cuda(), torch.from_numpy(d2).cuda()
    ms_i = run(d_o2, d_d2, 0)
    hsum = int(d_hits.long().sum().item())
    ms_a = run(d_o2, d_d2, 1)
    print("[%s] rays: coherent closest %.1f, incoherent closest %.1f, incoherent any-hit %.1f Mrays/s | hit-hash %d"
          % (tag, nr / ms_c / 1e3, nr / ms_i / 1e3, nr / ms_a / 1e3, hsum), flush=True)


def main():
    what = sys.argv[1:] or ["c3", "rays"]
    spp = int(os.environ.get("QP_SPP", "16"))
    be = None
    if "c3" in what or "rays" in what:
        sc = scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=spp, max_depth=8, materials="c3")
        if "c3" in what:
            be = frame(sc, "C3 %d spp" % spp)
        else:
            be = pkg.Backend(sc, 0)
        if "c3sort" in what:
            frame(sc, "C3 %d spp" % spp, sort=1).close()
            frame(sc, "C3 %d spp" % spp, sort=2).close()
        if "rays" in what:
            rays(be, (1920, 1080))
        be.close()
    if "c1" in what:
        frame(scenes.cornell_box((512, 512), 16, 5), "C1").close()
    if "c5" in what:
        frame(scenes.many_light_scene(n=708, resolution=(1920, 1080), spp=spp, max_depth=16, n_quads=5000), "C5 %d spp" % spp).close()


if __name__ == "__main__":
    main()
