"""akb_scene_create wall time and its phases (AKB_BVH_VERBOSE=1) on C3 / C4 (development aid)."""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("akarirender-1_b200")

which = sys.argv[1] if len(sys.argv) > 1 else "c4"
sc = (pkg.scenes.instanced_scene(resolution=(3840, 2160), spp=1, max_depth=8) if which == "c4"
      else pkg.scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=1, max_depth=8, materials="c3"))
for kind in (0, 0, pkg.abi.BVH_HOST_SAH if which != "c4" else 0):
    t = time.time()
    be = pkg.Backend(sc, 0, bvh_build=kind)
    print("Backend() %.2f s, built_by %d" % (time.time() - t, be.info()["bvh_build"]), flush=True)
    be.close()
