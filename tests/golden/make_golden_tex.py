"""Generates tests/golden/reference_vectors_tex.npz: films of the image-textured test scene from the REFERENCE's own
code (oracle/_ref/libakari_ref.so: the reference's ImageTexture / Material / AreaLight over the description's texels).
Build container only:

    make -C oracle ref && python tests/golden/make_golden_tex.py

Note: the emissive test texture holds values above 1 (a PNG cannot), so that case is fed through the C ABI's float
texels; the scene-file route is covered with 8-bit textures in tests/test_textures.py.
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


def tex_cases():
    s = pkg.scenes
    return {"tex40": s.textured_box((40, 40), 8, 5), "tex24_pcg": s.textured_box((24, 24), 4, 3, sampler="pcg")}


def main():
    out = {name: Reference(sc).render(sc.integrator.params()) for name, sc in tex_cases().items()}
    np.savez_compressed(os.path.join(HERE, "reference_vectors_tex.npz"), **out)
    print({k: (v.shape, float(v[..., :3].sum())) for k, v in out.items()})


if __name__ == "__main__":
    main()
