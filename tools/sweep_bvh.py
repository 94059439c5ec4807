"""Development aid: sweep the BVH cost-model constants on the GPU (C3 scene, 16 spp wavefront + ray batches)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for trav in sys.argv[1].split(","):
    for prim in sys.argv[2].split(","):
        env = dict(os.environ, AKB_BVH_TRAV_COST=trav, AKB_BVH_PRIM_COST=prim)
        if len(sys.argv) > 3:
            env["AKB_BVH_MAX_LEAF"] = sys.argv[3]
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "quick_perf.py")], env=env, capture_output=True, text=True)
        keep = [l for l in r.stdout.splitlines() if l.startswith(("scene", "incoherent closest", "coherent", "render mode=0 sort=0"))]
        print("trav=%s prim=%s | %s" % (trav, prim, " | ".join(k[:95] for k in keep)), flush=True)
