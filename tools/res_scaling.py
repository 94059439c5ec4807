"""Development aid: the same scene and the same number of samples per frame at three resolutions (paths per launch x 1/4,
x 1, x 4): how much of a frame is per-bounce cost (launch tails, the slowest ray of each launch) rather than work."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")

for (W, H, spp) in ((960, 540, 256), (1920, 1080, 64), (3840, 2160, 16), (7680, 4320, 4)):
    sc = pkg.scenes.terrain_scene(n=708, resolution=(W, H), spp=spp, max_depth=8, materials="c3")
    be = pkg.Backend(sc, 0)
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    p = sc.integrator.params()
    for mode in (pkg.abi.MODE_WAVEFRONT, pkg.abi.MODE_MEGAKERNEL):
        p.mode = mode
        best = 1e9
        for _ in range(3):
            st = be.render_device(film.data_ptr(), p)
            best = min(best, st.ms_total)
        print("%dx%d %d spp mode %d: %.2f ms, %.1f Msamples/s, rays/sample %.3f, launches %d" % (
            W, H, spp, mode, best, st.n_samples / best / 1e3, (st.n_closest_rays + st.n_shadow_rays) / st.n_samples,
            st.n_kernel_launches), flush=True)
    be.close()
    del film
