"""AO integrator on the C3 mesh, wavefront vs one-thread-per-pixel kernel (development aid)."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")
abi, scenes, nodes = pkg.abi, pkg.scenes, pkg.nodes


def main():
    spp = int(os.environ.get("QP_SPP", "16"))
    sc = scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=1, max_depth=8, materials="c3")
    sc.integrator = nodes.AO(spp=spp, occlude=float(os.environ.get("QP_OCCLUDE", "inf")))
    be = pkg.Backend(sc, 0)
    film = torch.zeros(1080 * 1920 * 4, dtype=torch.float32, device="cuda")
    out = []
    for name, mode in (("wavefront", abi.MODE_WAVEFRONT), ("megakernel", abi.MODE_MEGAKERNEL)):
        p = sc.integrator.params()
        p.mode = mode
        for _ in range(3):
            st = be.render_device(film.data_ptr(), p)
        out.append(film.clone())
        rays = st.n_closest_rays + st.n_shadow_rays
        print("AO %-10s %.2f ms  %.1f Msamples/s  %.1f Mrays/s  launches %d" % (name, st.ms_total, st.n_samples / st.ms_total / 1e3, rays / st.ms_total / 1e3,
                                                                            st.n_kernel_launches))
    print("bit-identical:", bool((out[0].view(torch.int32) == out[1].view(torch.int32)).all()))


if __name__ == "__main__":
    main()
