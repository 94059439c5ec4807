"""Small fixed workload for compute-sanitizer (tools/sanitize.sh): the 64 x 64 smoke scene through the wavefront (both
pipelines, graphs, finish kernel), the megakernel and the AO integrator, a batch of rays through akb_trace, a
device-built BVH, and the byte stage. Exits non-zero if any result differs from the unsanitised expectation."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")
abi, scenes = pkg.abi, pkg.scenes


def main():
    pkg.set_option("pipeline_min_slots", 1)  # two pipelines even at this size
    sc = scenes.cornell_box((64, 64), 8, 5)
    be = pkg.Backend(sc, 0)
    films = []
    for mode in (abi.MODE_WAVEFRONT, abi.MODE_MEGAKERNEL):
        p = sc.integrator.params()
        p.mode = mode
        films.append(be.render(p)[0])
    assert np.array_equal(films[0].view(np.uint32), films[1].view(np.uint32)), "wavefront != megakernel"
    p = sc.integrator.params()
    p.sort_materials = 1
    assert np.array_equal(be.render(p)[0].view(np.uint32), films[0].view(np.uint32)), "material sort changed the film"
    be.film_to_srgb8(films[0])
    be.close()
    ts = scenes.terrain_scene(n=48, resolution=(96, 54), spp=4, max_depth=8, materials="c3", rr_start_depth=3)
    for kind in (abi.BVH_HOST_SAH, abi.BVH_DEVICE):
        be = pkg.Backend(ts, 0, bvh_build=kind)
        f, st = be.render()
        rng = np.random.default_rng(1)
        n = 20000
        o = np.concatenate([rng.uniform(-50, 50, (n, 3)), np.full((n, 1), 1e-3)], 1).astype(np.float32)
        o[:, 1] = rng.uniform(-2, 15, n)
        d = rng.normal(size=(n, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        d = np.concatenate([d, np.full((n, 1), np.inf, np.float32)], 1).astype(np.float32)
        h = be.trace(o, d)
        occ = be.trace(o, d, any_hit=True)
        assert np.array_equal(occ.astype(bool), h["prim_id"] >= 0)
        be.close()
    ao = scenes.Scene(ts.camera, pkg.nodes.AO(spp=2, occlude=5.0), ts.shapes)
    be = pkg.Backend(ao, 0)
    be.render()
    p = ao.integrator.params()
    p.mode = abi.MODE_WAVEFRONT
    be.render(p)
    be.close()
    print("sanitize target: ok")


if __name__ == "__main__":
    main()
