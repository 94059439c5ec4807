"""Development aid: C3 (16 spp) through the streaming wavefront for several settings of its options."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
if wl == "c1":
    sc = pkg.scenes.cornell_box((512, 512), 16, 5)
else:
    sc = pkg.scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=16, max_depth=8, materials="c3")
W, H = sc.camera.resolution
be = pkg.Backend(sc, 0)
film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
p = sc.integrator.params()


def run(tag):
    best = 1e9
    for _ in range(3):
        st = be.render_device(film.data_ptr(), p)
        best = min(best, st.ms_total)
    print("%-28s %.3f ms  %.1f Msamples/s" % (tag, best, st.n_samples / best / 1e3), flush=True)


pkg.set_option("stream", 0)
run("classic")
pkg.set_option("stream", 1)
pkg.set_option("timeline", 1)
for opts in sys.argv[2:] or ["stream_flex_mask=1", "stream_flex_mask=0", "stream_flex_mask=3", "stream_flex_mask=15"]:
    for kv in opts.split(","):
        k, v = kv.split("=")
        pkg.set_option(k, v)
    run(opts)
