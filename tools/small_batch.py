"""Development aid: device time of akb_trace_device for small ray batches (the latency floor of one traversal launch)."""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")
sc = pkg.scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=1, max_depth=8, materials="c3")
be = pkg.Backend(sc, 0)
W, H = 1920, 1080
rng = np.random.default_rng(1)
n = 1 << 20
xy = np.stack([rng.integers(0, W, n), rng.integers(0, H, n)], -1).astype(np.int32)
org, dr = be.generate_rays(xy, rng.random((n, 2), dtype=np.float32))
h = be.trace(org, dr)
hit = h["prim_id"] >= 0
p = org[:, :3] + dr[:, :3] * np.where(hit, h["t"], 10)[:, None]
d = rng.normal(size=(n, 3)).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True); d[:, 1] = np.abs(d[:, 1]) * 0.3
o2 = np.concatenate([p + np.array([0, 1e-2, 0]), np.full((n, 1), 1e-3)], 1).astype(np.float32)[hit]
d2 = np.concatenate([d, np.full((n, 1), np.inf)], 1).astype(np.float32)[hit]
d_o, d_d = torch.from_numpy(o2).cuda(), torch.from_numpy(d2).cuda()
d_h = torch.empty(len(o2) * 5, dtype=torch.int32, device="cuda"); d_occ = torch.empty(len(o2), dtype=torch.int32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for m in (1, 32, 1024, 32768, 262144, len(o2)):
    ts = []
    for rep in range(5):
        off = rep * m % max(1, len(o2) - m)
        flush.fill_(rep)  # evict the BVH from L2
        ts.append(be.trace_device(d_o[off:].data_ptr(), d_d[off:].data_ptr(), m, 0, d_h.data_ptr(), d_occ.data_ptr()))
    cold = min(ts)
    ts = [be.trace_device(d_o.data_ptr(), d_d.data_ptr(), m, 0, d_h.data_ptr(), d_occ.data_ptr()) for _ in range(5)]
    print("batch %7d rays: L2 flushed %.1f us, warm %.1f us" % (m, cold * 1e3, min(ts) * 1e3), flush=True)
