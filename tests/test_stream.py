"""The streaming wavefront (akb_set_option("stream", 1): one persistent launch, ray ring and hit ring between traversal
and shading warps, no barrier between bounces; csrc/kernels.cu). It reorders nothing WITHIN a pixel - shadow result, then
the vertex, samples in sequence - so its films must be the bounce-synchronous kernels' films bit for bit, which are in
turn pinned by the reference-generated goldens and the oracle: both comparisons are made here, through the C ABI."""
import numpy as np
import pytest

from common import bits, golden, golden_scenes, pkg, scenes
from oracle_libs import Oracle

pytestmark = pytest.mark.gpu


def n_diff_px(a, b):
    return int((bits(a) != bits(b)).any(-1).sum())


def render_both(sc, params=None):
    be = pkg.Backend(sc, 0)
    p = params if params is not None else sc.integrator.params()
    out = []
    try:
        for stream in (0, 1):
            pkg.set_option("stream", stream)
            film, st = be.render(p)
            out.append((film, st))
    finally:
        pkg.set_option("stream", 0)
        be.close()
    return out


def test_stream_films_equal_goldens():
    G, S = golden(), golden_scenes()
    for name in ("cornell32", "mixed32", "mixed24_pcg", "terrain"):
        (classic, st0), (stream, st1) = render_both(S[name])
        assert n_diff_px(stream, G[name + "_film"]) == 0, name
        assert n_diff_px(stream, classic) == 0, name
        assert (st1.n_samples, st1.n_closest_rays, st1.n_shadow_rays) == (st0.n_samples, st0.n_closest_rays, st0.n_shadow_rays), name
        assert st1.n_kernel_launches == 4  # init, seed + stream (one persistent launch), resolve


def test_stream_extensions_and_sharding_equal_oracle():
    # Russian roulette, dielectric, depth 16, 1 000 lights: the oracle is the pin
    ml = scenes.many_light_scene(n=48, resolution=(64, 36), spp=8, max_depth=16, n_quads=500)
    want = Oracle(ml).render(ml.integrator.params())
    (classic, _), (stream, _) = render_both(ml)
    assert n_diff_px(stream, want) == 0 and n_diff_px(classic, want) == 0
    # a tile-split share (interleaved 32x32 tiles): non-owned pixels stay 0, owned ones are the 1-GPU film's
    sc = scenes.terrain_scene(n=96, resolution=(160, 96), spp=6, max_depth=8, materials="c3")
    full = render_both(sc)[1][0]
    acc = np.zeros_like(full)
    for rank in range(3):
        share = render_both(sc, sc.integrator.params(rank=rank, world_size=3, shard_tile=32))[1][0]
        assert n_diff_px(acc * (share != 0), 0 * acc) == 0  # disjoint
        acc = acc + share
    assert n_diff_px(acc, full) == 0
    # a frame larger than one wave of the persistent grid, so that both rings wrap many laps of work through the warps
    big = scenes.terrain_scene(n=256, resolution=(640, 360), spp=8, max_depth=8, materials="c3")
    (classic, st0), (stream, st1) = render_both(big)
    assert n_diff_px(stream, classic) == 0
    assert st1.n_closest_rays == st0.n_closest_rays and st1.n_shadow_rays == st0.n_shadow_rays


def test_pipeline_count_and_tile_order_are_neutral():
    """Options pipelines (1..4 groups of the owned tiles, each with its own queues and stream) and pipeline_interleave
    (tiles dealt out to the pipelines in turn instead of contiguous ranges) change which slot a pixel lives in and which
    launches trace it, never its samples: films equal the goldens bit for bit, ragged frames and tile-split shares too."""
    G, S = golden(), golden_scenes()
    ragged = scenes.terrain_scene(n=96, resolution=(150, 70), spp=5, max_depth=8, materials="c3")
    be0 = pkg.Backend(ragged, 0)
    ragged_ref = be0.render()[0]
    share_ref = be0.render(ragged.integrator.params(rank=1, world_size=3, shard_tile=8))[0]
    be0.close()
    try:
        pkg.set_option("pipeline_min_slots", 0)
        for pipes, inter in ((1, 0), (2, 1), (3, 0), (3, 1), (4, 0), (4, 1)):
            pkg.set_option("pipelines", pipes)
            pkg.set_option("pipeline_interleave", inter)
            for name in ("cornell32", "mixed32", "terrain"):
                be = pkg.Backend(S[name], 0)
                assert n_diff_px(be.render()[0], G[name + "_film"]) == 0, (name, pipes, inter)
                be.close()
            be = pkg.Backend(ragged, 0)
            assert n_diff_px(be.render()[0], ragged_ref) == 0, (pipes, inter)
            assert n_diff_px(be.render(ragged.integrator.params(rank=1, world_size=3, shard_tile=8))[0], share_ref) == 0, (pipes, inter)
            be.close()
    finally:
        pkg.set_option("pipeline_min_slots", 128 * 1024)
        pkg.set_option("pipelines", 2)
        pkg.set_option("pipeline_interleave", 0)
