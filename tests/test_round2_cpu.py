"""CPU tests (no GPU) for the round-2 golden vectors (tests/golden/make_golden_r2.py, generated from the reference's own
code): the oracle port and the product's host-compiled traversal/shading (tests/hostcheck) against
  * the multi-mesh scene (BASELINE config C4's shape: geom_id / prim_id, per-mesh material tables, the exact-tie rule),
  * the 10 000-light scene (config C5's light count),
  * the reference-written sRGB bytes (Film::write_image -> DefaultImageWriter);
and the CPU baseline driver's tile scheduling."""
import os
import sys

import numpy as np
import pytest

from common import ROOT, bits, golden, hit_mismatch, same_bits, scenes
from oracle_libs import HostCheck, Oracle

sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from make_golden_r2 import r2_scenes  # noqa: E402


@pytest.fixture(scope="module")
def G2():
    return np.load(os.path.join(ROOT, "tests", "golden", "reference_vectors_r2.npz"))


@pytest.fixture(scope="module")
def S2():
    return r2_scenes()


@pytest.mark.parametrize("which", ["oracle", "oracle_brute", "hostcheck"])
@pytest.mark.parametrize("name", ["multi", "multi_co"])
def test_multi_mesh_hits_golden(G2, S2, name, which):
    sc = S2[name]
    impl = {"oracle": lambda: Oracle(sc), "oracle_brute": lambda: Oracle(sc, brute_force=True),
            "hostcheck": lambda: HostCheck(sc)}[which]()
    g = G2[f"{name}_hits"]
    h = impl.trace(G2[f"{name}_org"], G2[f"{name}_dir"])
    bad, ties, tb = hit_mismatch(g, h)
    assert bad == 0 and tb == 0, (bad, ties, tb)
    assert len(np.unique(g["geom_id"])) >= 4  # the rays do reach several meshes
    # exact-t ties between the coincident meshes 0 and 3: the reference keeps the mesh its TLAS visits first, which is
    # mesh 0 here; the contract picks the lowest flat triangle id = the lowest mesh: the same answer
    assert ties == 0
    if name == "multi_co":
        assert not (h["geom_id"] == 3).any() or (h["prim_id"][h["geom_id"] == 3] == 724).all()
    assert same_bits(h["u"], g["u"]) and same_bits(h["v"], g["v"])
    occ = impl.trace(G2[f"{name}_org"], G2[f"{name}_dir_any"], any_hit=True)
    assert np.array_equal(occ, G2[f"{name}_occ"])


@pytest.mark.parametrize("which", ["oracle", "hostcheck"])
def test_multi_mesh_and_10k_light_films_golden(G2, S2, which):
    for name in ("multi", "lights10k"):
        sc = S2[name]
        impl = Oracle(sc) if which == "oracle" else HostCheck(sc)
        if name == "lights10k":
            assert impl.n_lights() == 10000
        film = impl.render(sc.integrator.params())
        assert same_bits(film, G2[name + "_film"]), name


def test_reference_tlas_loses_meshes_from_five_on():
    """Not a golden: documents why C4 (10 meshes) is pinned by the oracle and the 4-mesh golden instead of a 10-mesh
    reference run - with 5x2 placed copies the oracle (checked against brute force) hits 9 meshes from this camera."""
    sc = scenes.instanced_scene(n=20, resolution=(64, 36), spp=1, max_depth=5)
    o, ob = Oracle(sc), Oracle(sc, brute_force=True)
    ys, xs = np.mgrid[0:36, 0:64]
    xy = np.stack([xs.ravel(), ys.ravel()], -1).astype(np.int32)
    org, dr = o.generate_rays(xy, np.full((xy.shape[0], 2), 0.5, np.float32))
    h, hb = o.trace(org, dr), ob.trace(org, dr)
    assert hit_mismatch(hb, h) == (0, 0, 0)
    assert len(np.unique(h["geom_id"][h["geom_id"] >= 0])) >= 8


def test_srgb8_bytes_golden(G2):
    """The oracle's byte stage against bytes the reference's own writer produced (PNG written by Film::write_image,
    decoded by the reference's reader): bit-exact."""
    G = golden()
    o = Oracle(scenes.cornell_box((8, 8), 1, 1))
    for key in ("cornell32_film", "mixed32_film", "terrain_film"):
        assert np.array_equal(o.film_to_srgb8(G[key]), G2["srgb8_" + key]), key
    assert np.array_equal(o.film_to_srgb8(G2["srgb8_ramp_in"]), G2["srgb8_ramp"])
    assert len(np.unique(G2["srgb8_ramp"])) == 256  # every byte value is exercised


def test_render_rects_is_render():
    """The CPU-baseline driver (16x16 tiles of several rectangles in one work list) writes what per-rectangle calls
    write, and keeps its workers busy."""
    sc = scenes.mixed_box((70, 50), 3, 4)
    o = Oracle(sc)
    p = sc.integrator.params()
    rects = [(0, 3, 70, 7), (5, 20, 33, 41), (0, 49, 70, 50)]
    a = o.render_rects(p, rects, threads=4)
    b = np.zeros_like(a)
    for r in rects:
        f = o.render(p, rect=r, threads=2)
        x0, y0, x1, y1 = r
        b[y0:y1, x0:x1] = f[y0:y1, x0:x1]
    assert same_bits(a, b)
    assert o.busy[0] > 0 and o.busy[1] > 0
    full = o.render(p)
    assert same_bits(a[20:41, 5:33], full[20:41, 5:33])


def test_c1_full_size_differences_from_the_reference_are_the_documented_ones(G2):
    """C1 at full size: the product's host-compiled traversal against the committed rays where the reference's own film
    departs from the brute-force oracle (exact ties / reference false misses)."""
    sc = scenes.cornell_box((512, 512), 16, 5)
    hc = HostCheck(sc)
    h = hc.trace(G2["c1_full_diff_org"], G2["c1_full_diff_dir"])
    assert h.tobytes() == G2["c1_full_diff_hit_brute"].tobytes()
    assert len(G2["c1_full_ref_diff_yx"]) == 4
