"""the reference's own CUDA wavefront path tracer, patched in transit to build for sm_100a
(oracle/Makefile target ref_gpu), runs on the B200 and renders the same image as this backend - statistically: it is
built with --use_fast_math, spends one more sampler draw per bounce and clamps samples, so there is no bit parity to
ask for. Skipped where oracle/_ref/libakari_ref_gpu.so was not built."""
import numpy as np
import pytest

from common import pkg, rel_rmse, scenes
from oracle_libs import ReferenceGPU, have_ref_gpu

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not have_ref_gpu(), reason="oracle/_ref/libakari_ref_gpu.so not built (needs /root/reference)")
def test_reference_cuda_wavefront_renders_the_same_image():
    # Cornell box: Diffuse + Emissive only. (With Mix materials the reference's GPU path selects the branch with a
    # sampler draw of its own, gpu/cuda/integrator.cpp:286-287, and its images differ from its own CPU path's by
    # almost 1 % in the mean - measured on scenes.mixed_box - so Mix is not part of this comparison.)
    sc = scenes.cornell_box((64, 64), 1024, 5)
    p = sc.integrator.params()
    p.ray_clamp = 10.0  # the reference's GPU integrator always clamps; ask the same of ours
    ref = ReferenceGPU(sc)
    films = {}
    for name, wf in (("wavefront", True), ("megakernel", False)):
        films[name], ms = ref.render_gpu(p, tile_size=64, wavefront=wf)
        assert (films[name][..., 3] > 0).all() and ms > 0  # (the megakernel path stores the mean with weight 1)
    ref.close()
    be = pkg.Backend(sc, 0)
    ours, st = be.render(p)
    be.close()
    noise = rel_rmse(films["wavefront"], films["megakernel"])  # two reference renders with different draw sequences
    for name, f in films.items():
        e = rel_rmse(ours, f)
        assert e < 0.03 and e < 2.0 * max(noise, 0.005), (name, e, noise)
        m_ref, m_ours = (f[..., :3] / f[..., 3:4]).mean(), (ours[..., :3] / ours[..., 3:4]).mean()
        assert abs(m_ref - m_ours) / m_ref < 0.005, (name, m_ref, m_ours)
