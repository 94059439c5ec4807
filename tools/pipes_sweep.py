"""Development aid: frame time of C3 / C1 / C5 for the wavefront's pipeline options (count, tile interleave)."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")

spp = int(os.environ.get("QP_SPP", "16"))
for wl in sys.argv[1:] or ["c3"]:
    if wl == "c1":
        sc = pkg.scenes.cornell_box((512, 512), 16, 5)
    elif wl == "c3_4k":
        sc = pkg.scenes.terrain_scene(n=708, resolution=(3840, 2160), spp=spp, max_depth=8, materials="c3")
    elif wl == "c5":
        sc = pkg.scenes.many_light_scene(n=708, resolution=(1920, 1080), spp=spp, max_depth=16, n_quads=5000, rr_start_depth=3)
    else:
        sc = pkg.scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=spp, max_depth=8, materials="c3")
    W, H = sc.camera.resolution
    be = pkg.Backend(sc, 0)
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    p = sc.integrator.params()
    ref = None
    for pipes, inter in [tuple(int(v) for v in a.split(":")) for a in os.environ.get("QP_SETS", "2:0,2:1,3:1,4:1,4:0,1:1").split(",")]:
        pkg.set_option("pipelines", pipes)
        pkg.set_option("pipeline_interleave", inter)
        best = 1e9
        for _ in range(4):
            st = be.render_device(film.data_ptr(), p)
            best = min(best, st.ms_total)
        h = int(torch.sum(film.view(torch.int32).long() & 0xFFFFFF).item())
        ref = h if ref is None else ref
        print("%s %d spp: pipelines %d interleave %d: %.3f ms (%.1f Msamples/s) launches %d film %s" % (
            wl, sc.integrator.spp, pipes, inter, best, st.n_samples / best / 1e3, st.n_kernel_launches, "same" if h == ref else "DIFFERENT"), flush=True)
    be.close()
