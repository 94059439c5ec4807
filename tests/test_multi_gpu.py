"""The multi-GPU frame behind the C ABI (run with -m gpu; the >= 2-GPU cases skip on a one-GPU box): ONE process,
akb_init(n_devices), akb_render(params.n_gpus = n): one host thread + stream per GPU inside the library, the scene
replicated from one build, one film collective (ncclReduce, or the library's own peer-memory kernels)."""
import ctypes as C

import numpy as np
import pytest

from common import abi, bits, pkg, same_bits, scenes
from oracle_libs import Oracle

pytestmark = pytest.mark.gpu


def n_gpus():
    return abi.load().akb_device_count()


def test_one_device_through_the_multi_gpu_entry():
    """akb_init(1) + n_gpus = 1 is the one-GPU path (and what a one-GPU host of the section-8b boundary calls)."""
    sc = scenes.mixed_box((100, 70), 3, 4)
    ref, _ = pkg.Backend(sc, 0).render()
    be = pkg.Backend(sc, n_devices=1)
    assert be.info()["n_devices"] == 1 and be.lib.akb_n_devices() == 1
    p = sc.integrator.params()
    p.n_gpus = 1
    film, st = be.render(p)
    assert same_bits(film, ref) and st.n_gpus_used == 1 and st.ms_wall >= st.ms_total > 0
    # asking for more GPUs than the scene lives on is an error, not a silent single-GPU render
    p.n_gpus = 2
    out = np.zeros_like(film)
    assert be.lib.akb_render(be.h, C.byref(p), out.ctypes.data_as(C.c_void_p), None) == abi.INVALID_ARG
    assert b"n_gpus" in be.lib.akb_last_error()
    be.close()
    pkg.init(0)


@pytest.mark.parametrize("combine", ["nccl", "p2p"])
def test_tile_split_is_bit_identical_to_one_gpu(combine):
    n = n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    scs = (scenes.mixed_box((100, 70), 3, 4),  # ragged: not a multiple of the tile
           scenes.terrain_scene(n=96, resolution=(640, 360), spp=4, max_depth=8, materials="c3", rr_start_depth=3))
    refs = []
    for sc in scs:
        be = pkg.Backend(sc, 0)
        refs.append(be.render()[0])
        be.close()
    try:
        pkg.set_option("combine", combine)
        for sc, ref in zip(scs, refs):
            be = pkg.Backend(sc, n_devices=n)
            assert be.info()["n_devices"] == n
            for k in sorted({2, n}):
                for tile in (32, 8):
                    p = sc.integrator.params(shard_tile=tile)
                    p.n_gpus = k
                    film, st = be.render(p)
                    assert same_bits(film, ref), (combine, k, tile)
                    assert st.n_gpus_used == k and st.n_samples == ref.shape[0] * ref.shape[1] * sc.integrator.spp
            # megakernel mode and the AO integrator go through the same split
            p = sc.integrator.params()
            p.n_gpus, p.mode = n, abi.MODE_MEGAKERNEL
            assert same_bits(be.render(p)[0], ref)
            # ... and so does the streaming wavefront (one persistent launch per GPU, its own rings on every replica)
            try:
                pkg.set_option("stream", 1)
                p = sc.integrator.params()
                p.n_gpus = n
                assert same_bits(be.render(p)[0], ref)
            finally:
                pkg.set_option("stream", 0)
            be.close()
    finally:
        pkg.set_option("combine", "nccl")
        pkg.init(0)


@pytest.mark.parametrize("combine", ["nccl", "p2p"])
def test_sample_split_sums_the_per_gpu_oracle_films(combine):
    n = n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    sc = scenes.mixed_box((64, 48), 10, 4)
    o = Oracle(sc)
    parts = [o.render(sc.integrator.params(rank=r, world_size=n, split=abi.SPLIT_SAMPLE)) for r in range(n)]
    try:
        pkg.set_option("combine", combine)
        be = pkg.Backend(sc, n_devices=n)
        p = sc.integrator.params(split=abi.SPLIT_SAMPLE)
        p.n_gpus = n
        film, st = be.render(p)
        assert (film[..., 3] == 10.0).all() and st.n_samples == 64 * 48 * 10
        if combine == "p2p":  # the peer-memory sum adds in rank order
            want = parts[0].copy()
            for q in parts[1:]:
                want = want + q
            assert same_bits(film, want)
        else:  # NCCL picks its own order: float sums agree to rounding
            assert np.allclose(film, sum(parts), rtol=1e-5, atol=1e-6)
        be.close()
    finally:
        pkg.set_option("combine", "nccl")
        pkg.init(0)
