"""Generates tests/golden/reference_vectors_ao.npz: AO-integrator films from the REFERENCE's own code
(oracle/_ref/libakari_ref.so; the `Li` of cpu::AmbientOcclusion::render driven on the reference's Scene, Triangle,
Frame3f and sampling<C> - see oracle/ref_harness.cpp). Build container only:

    make -C oracle ref && python tests/golden/make_golden_ao.py
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


def ao_cases():
    """name -> scene with an AO integrator node (shared with the tests)."""
    s, n = pkg.scenes, pkg.nodes
    out = {}
    for name, sc, spp, occ in (("ao_cornell32", s.cornell_box((32, 32), 1, 1), 16, float("inf")),
                               ("ao_cornell32_near", s.cornell_box((32, 32), 1, 1), 8, 0.75),
                               ("ao_mixed24_pcg", s.mixed_box((24, 24), 1, 1), 4, 1.5),
                               ("ao_terrain", s.terrain_scene(n=48, resolution=(64, 36), spp=1, materials="c3ref"), 4, 6.0)):
        sc.integrator = n.AO(spp=spp, occlude=occ)
        out[name] = sc
    return out


def main():
    out = {}
    for name, sc in ao_cases().items():
        p = sc.integrator.params()
        if name.endswith("pcg"):
            p.sampler = pkg.abi.SAMPLER_PCG
        out[name] = Reference(sc).render(p)
    np.savez_compressed(os.path.join(HERE, "reference_vectors_ao.npz"), **out)
    print({k: (v.shape, float(v[..., 0].sum())) for k, v in out.items()})


if __name__ == "__main__":
    main()
