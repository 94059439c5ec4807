"""Generates tests/golden/reference_hispp.npz: 4096-spp float films from the REFERENCE's own CPU path tracer
(oracle/_ref/libakari_ref.so) for the image-level criterion of the north star ("<= 1 % relative RMSE at 4096 spp").
Each scene is rendered twice: with the LCG sampler (the seeds the GPU path reproduces) and with the PCG sampler (an
independent random stream, for the different-seeds comparison). Build container only:

    make -C oracle ref && python tests/golden/make_golden_hispp.py
"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("akarirender-1_b200")
from oracle_libs import Reference  # noqa: E402

SPP = 4096


def hispp_cases():
    s = pkg.scenes
    return {"cornell48": s.cornell_box((48, 48), SPP, 5),
            "terrain64": s.terrain_scene(n=48, resolution=(64, 36), spp=SPP, max_depth=5, materials="c3ref")}


def main():
    out = {}
    for name, sc in hispp_cases().items():
        ref = Reference(sc)
        for sampler, kind in (("lcg", pkg.abi.SAMPLER_LCG), ("pcg", pkg.abi.SAMPLER_PCG)):
            p = sc.integrator.params()
            p.sampler = kind
            out["%s_%s" % (name, sampler)] = ref.render(p)
    np.savez_compressed(os.path.join(HERE, "reference_hispp.npz"), **out)
    print({k: (v.shape, float(v[..., :3].sum())) for k, v in out.items()})


if __name__ == "__main__":
    main()
