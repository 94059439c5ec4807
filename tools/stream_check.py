"""Development aid: the streaming wavefront (option stream = 1) against the bounce-synchronous one: films bit for bit,
ray counts, device time.   python tools/stream_check.py [quick|c3|c1|c5|all]"""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")
scenes = pkg.scenes


def run(sc, name, reps=3):
    W, H = sc.camera.resolution
    be = pkg.Backend(sc, 0)
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    p = sc.integrator.params()
    out = {}
    for stream in (0, 1):
        pkg.set_option("stream", stream)
        best, st = 1e9, None
        for _ in range(reps):
            st = be.render_device(film.data_ptr(), p)
            best = min(best, st.ms_total)
        out[stream] = (film.cpu().numpy().copy(), best, st.n_samples, st.n_closest_rays, st.n_shadow_rays, st.n_kernel_launches)
    pkg.set_option("stream", 0)
    a, b = out[0], out[1]
    same = a[0].view(np.uint32) == b[0].view(np.uint32)
    print("%s: classic %.3f ms (%.1f Msamples/s, %d launches) | stream %.3f ms (%.1f Msamples/s) | film identical: %s (%d of %d words differ) | "
          "samples %d/%d closest %d/%d shadow %d/%d" % (name, a[1], a[2] / a[1] / 1e3, a[5], b[1], b[2] / b[1] / 1e3, bool(same.all()),
                                                       int((~same).sum()), same.size, a[2], b[2], a[3], b[3], a[4], b[4]), flush=True)
    be.close()
    return bool(same.all())


what = sys.argv[1] if len(sys.argv) > 1 else "quick"
ok = True
if what in ("quick", "all"):
    ok &= run(scenes.cornell_box((64, 64), 4, 5), "cornell 64x64 4spp")
    ok &= run(scenes.cornell_box((512, 512), 16, 5), "C1 cornell 512x512 16spp")
    ok &= run(scenes.terrain_scene(n=128, resolution=(256, 256), spp=8, max_depth=8, materials="c3"), "terrain 32k tris 256x256 8spp")
if what in ("c3", "all", "quick"):
    ok &= run(scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=16, max_depth=8, materials="c3"), "C3 1080p 16spp")
if what in ("c3full", "all"):
    ok &= run(scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=64, max_depth=8, materials="c3"), "C3 1080p 64spp")
if what in ("c5", "all"):
    ok &= run(scenes.many_light_scene(n=708, resolution=(1920, 1080), spp=32, max_depth=16, n_quads=5000, rr_start_depth=3), "C5 1080p 32spp")
print("ALL IDENTICAL" if ok else "MISMATCH")
