"""Development aid: two C3 frames rendered at the same time (two scenes, two host threads, two streams) against the
same two frames one after the other: is there idle machine left in a frame that finer interleaving could use?"""
import importlib
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")

spp = int(os.environ.get("QP_SPP", "16"))
sc = pkg.scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=spp, max_depth=8, materials="c3")
bes = [pkg.Backend(sc, 0) for _ in range(2)]
films = [torch.zeros(1080 * 1920 * 4, dtype=torch.float32, device="cuda") for _ in range(2)]
streams = [torch.cuda.Stream() for _ in range(2)]
p = sc.integrator.params()


def frames(i, n, out):
    for _ in range(n):
        st = bes[i].render_device(films[i].data_ptr(), p, streams[i].cuda_stream)
    out[i] = st.ms_total


for k in range(2):
    frames(k, 2, [0, 0])
torch.cuda.synchronize()
N = 4
t = time.perf_counter()
o = [0, 0]
frames(0, N, o)
frames(1, N, o)
torch.cuda.synchronize()
seq = time.perf_counter() - t
t = time.perf_counter()
th = [threading.Thread(target=frames, args=(i, N, o)) for i in range(2)]
[x.start() for x in th]
[x.join() for x in th]
torch.cuda.synchronize()
par = time.perf_counter() - t
print("2 x %d frames of C3 at %d spp: one after the other %.1f ms per frame, at the same time %.1f ms per frame (device time of one frame then: %.1f ms)"
      % (N, spp, seq / (2 * N) * 1e3, par / (2 * N) * 1e3, o[0]))
