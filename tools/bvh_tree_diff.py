"""Development aid: the same rays through a host-built and a device-built tree; reports hits that differ."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")
abi, scenes = pkg.abi, pkg.scenes


def main():
    sc = scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=1, max_depth=8, materials="c3")
    a = pkg.Backend(sc, 0, bvh_build=abi.BVH_HOST_SAH)
    b = pkg.Backend(sc, 0, bvh_build=abi.BVH_DEVICE)
    W, H = 1920, 1080
    ys, xs = np.mgrid[0:H, 0:W]
    xy = np.stack([xs.ravel(), ys.ravel()], -1).astype(np.int32)
    org, dr = a.generate_rays(xy, np.full((xy.shape[0], 2), 0.5, np.float32))
    rng = np.random.default_rng(3)
    sets = [("primary", org, dr)]
    ha = a.trace(org, dr)
    hit = ha["prim_id"] >= 0
    p = org[:, :3] + dr[:, :3] * np.where(hit, ha["t"], 10)[:, None]
    for k in range(4):
        d = rng.normal(size=(org.shape[0], 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        if k < 2:
            d[:, 1] = np.abs(d[:, 1])
        o2 = np.concatenate([p + np.array([0, 1e-2, 0]), np.full((org.shape[0], 1), 1e-3)], 1).astype(np.float32)
        d2 = np.concatenate([d, np.full((org.shape[0], 1), np.inf)], 1).astype(np.float32)
        sets.append(("bounce%d" % k, o2, d2))
    for name, o, d in sets:
        ha, hb = a.trace(o, d), b.trace(o, d)
        diff = np.nonzero((ha["prim_id"] != hb["prim_id"]) | (ha["t"].view(np.int32) != hb["t"].view(np.int32)))[0]
        print(name, "rays", o.shape[0], "differ", diff.size)
        for i in diff[:8]:
            print("   ray", i, "host", ha[i], "device", hb[i], "o", o[i], "d", d[i])
        da = d.copy()
        da[:, 3] = 40.0
        oa, ob = a.trace(o, da, any_hit=True), b.trace(o, da, any_hit=True)
        print(name, "any-hit differ", int((oa != ob).sum()))


if __name__ == "__main__":
    main()
