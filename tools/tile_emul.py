"""Development aid: one GPU renders rank 0's share of an N-GPU tile split (C3 by default) under different scheduling
options - the per-GPU time of the bit-identical tile split without needing N GPUs.

    python tools/tile_emul.py [world=8] [name=value ...]     e.g.  python tools/tile_emul.py 8 pipelines=1 check_every=8
"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")
abi, scenes = pkg.abi, pkg.scenes


def main():
    args = sys.argv[1:]
    world = int(args[0]) if args and args[0].isdigit() else 8
    opts = dict(a.split("=", 1) for a in args if "=" in a)
    mode = opts.pop("mode", "wavefront")
    wl = opts.pop("workload", "c3")
    spp = int(opts.pop("spp", "64"))
    for k, v in opts.items():
        pkg.set_option(k, v)
    if wl == "c3":
        sc = scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=spp, max_depth=8, materials="c3")
    else:
        sc = scenes.cornell_box((512, 512), 16, 5)
    be = pkg.Backend(sc, 0)
    W, H = sc.camera.resolution
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    out = []
    for rank in (0, world // 2):
        p = sc.integrator.params(rank=rank, world_size=world, shard_tile=32)
        p.mode = abi.MODE_MEGAKERNEL if mode == "megakernel" else abi.MODE_WAVEFRONT
        best = 1e9
        for _ in range(4):
            st = be.render_device(film.data_ptr(), p)
            best = min(best, st.ms_total)
        if opts.get("profile_stages") == "1":
            print("   stages: total %.2f extend %.2f shade %.2f ms, launches %d" % (st.ms_total, st.ms_extend, st.ms_shade, st.n_kernel_launches))
        h = int(torch.sum(film.view(torch.int32).long() & 0xFFFFF).item())
        out.append("rank %d: %.2f ms (%d launches, %d+%d rays, film-hash %d)" % (rank, best, st.n_kernel_launches, st.n_closest_rays, st.n_shadow_rays, h))
    ideal = None
    print("tile split 1/%d %s %s: %s" % (world, mode, opts, "; ".join(out)), flush=True)
    be.close()


if __name__ == "__main__":
    main()
