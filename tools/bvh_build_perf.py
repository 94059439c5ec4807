"""Host SAH build vs device PLOC build: build time, tree size and traversal speed on the same scene.
Development aid; run on the GPU box: python tools/bvh_build_perf.py [c3|c4]"""
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
    which = sys.argv[1] if len(sys.argv) > 1 else "c3"
    spp = int(os.environ.get("QP_SPP", "16"))
    if which == "c4":
        sc = scenes.instanced_scene(resolution=(3840, 2160), spp=spp, max_depth=8)
    else:
        sc = scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=spp, max_depth=8, materials="c3")
    W, H = sc.camera.resolution
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    films = []
    for name, kind in (("host SAH", abi.BVH_HOST_SAH), ("device PLOC", abi.BVH_DEVICE), ("device PLOC (2nd)", abi.BVH_DEVICE)):
        t = time.time()
        be = pkg.Backend(sc, 0, bvh_build=kind)
        create_s = time.time() - t
        info = be.info()
        p = sc.integrator.params()
        for _ in range(2):
            st = be.render_device(film.data_ptr(), p)
        films.append(film.clone())
        print("%-18s create %.3f s (BVH %.3f s, built_by %d)  nodes %d depth %d  render %.2f ms  %.1f Msamples/s"
              % (name, create_s, info["build_us"] / 1e6, info["bvh_build"], info["n_bvh_nodes"], info["bvh_depth"], st.ms_total,
                 st.n_samples / st.ms_total / 1e3), flush=True)
        be.close()
    for i, j in ((0, 1), (1, 2)):
        a, b = films[i].view(torch.int32).view(-1, 4), films[j].view(torch.int32).view(-1, 4)
        bad = (a != b).any(1).nonzero().flatten()
        print("films %d vs %d: %d pixels differ" % (i, j, bad.numel()))
        for k in bad[:6].tolist():
            print("   pixel", k % W, k // W, films[i].view(-1, 4)[k].tolist(), films[j].view(-1, 4)[k].tolist())


if __name__ == "__main__":
    main()
