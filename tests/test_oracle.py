"""CPU tests (no GPU): pin the oracle (oracle/akari_oracle.c) against the golden vectors generated from the reference's
own code (tests/golden/make_golden.py) and, where oracle/_ref is built, against the reference itself; check the
product's host logic (BVH8 build, flattening, slot map) through tests/hostcheck."""
import numpy as np
import pytest

from common import bits, golden, golden_scenes, hit_mismatch, random_rays, same_bits, scenes
from oracle_libs import HostCheck, Oracle, Reference, have_ref


@pytest.fixture(scope="module")
def G():
    return golden()


@pytest.fixture(scope="module")
def S():
    return golden_scenes()


def test_sampler_kat_survey_values():
    # LCG seed 7, g++ right-to-left next2d
    o = Oracle(scenes.cornell_box((8, 8), 1, 1))
    d, n2 = o.sampler_kat(0, 7, 4)
    assert np.allclose(d, [0.798528016, 0.149632409, 0.165837333, 0.345986694], rtol=0, atol=1e-9)
    assert n2[0] == d[1] and n2[1] == d[0]


def test_sampler_kat_golden(G, S):
    o = Oracle(S["cornell32"])
    for kind, name in ((0, "lcg"), (1, "pcg")):
        for seed in (0, 7, 123456, 0xFFFFFFF0):
            d, n2 = o.sampler_kat(kind, seed, 8)
            assert same_bits(d, G[f"kat_{name}_{seed}_draws"])
            assert same_bits(n2, G[f"kat_{name}_{seed}_next2d"])


def test_lcg_can_return_one():
    # kernel/sampler.h:60-63: (float)0xFFFFFFFF == 2^32, so seeds >= 0xFFFFFF80 give exactly 1.0
    o = Oracle(scenes.cornell_box((8, 8), 1, 1))
    inv = pow(1103515245, -1, 2 ** 32)
    seed = ((0xFFFFFFFF - 12345) * inv) % 2 ** 32
    d, _ = o.sampler_kat(0, seed, 1)
    assert d[0] == 1.0


@pytest.mark.parametrize("which", ["oracle", "oracle_brute", "hostcheck"])
@pytest.mark.parametrize("name", ["cornell", "terrain"])
def test_rays_and_hits_golden(G, S, name, which):
    sc = S[name]
    impl = {"oracle": lambda: Oracle(sc), "oracle_brute": lambda: Oracle(sc, brute_force=True),
            "hostcheck": lambda: HostCheck(sc)}[which]()
    org, dr = impl.generate_rays(G[f"{name}_xy"], G[f"{name}_u2"])
    assert same_bits(org, G[f"{name}_org"]) and same_bits(dr, G[f"{name}_dir"])
    for sfx in ("", "2"):
        h = impl.trace(G[f"{name}_org{sfx}"], G[f"{name}_dir{sfx}"])
        bad, ties, tb = hit_mismatch(G[f"{name}_hits{sfx}"], h)
        assert bad == 0 and tb == 0, (name, sfx, bad, ties, tb)
        ok = G[f"{name}_hits{sfx}"]["prim_id"] == h["prim_id"]
        assert same_bits(h["u"][ok], G[f"{name}_hits{sfx}"]["u"][ok]) and same_bits(h["v"][ok], G[f"{name}_hits{sfx}"]["v"][ok])
    occ = impl.trace(G[f"{name}_org2"], G[f"{name}_dir2_any"], any_hit=True)
    assert np.array_equal(occ, G[f"{name}_occ2"])


@pytest.mark.parametrize("post_limit", [-1, 0, 2])
@pytest.mark.parametrize("name", ["cornell", "terrain"])
def test_if_if_step_same_hits(G, S, name, post_limit):
    """trav_step_ii (the step the persistent warps run on real trees) returns exactly what trav_step does: default
    parking limit, no parking at all (0) and a nearly full stack (2)."""
    hc = HostCheck(S[name])
    for sfx in ("", "2"):
        a = hc.trace(G[f"{name}_org{sfx}"], G[f"{name}_dir{sfx}"])
        b = hc.trace_ii(G[f"{name}_org{sfx}"], G[f"{name}_dir{sfx}"], post_limit=post_limit)
        assert a.tobytes() == b.tobytes()
    a = hc.trace(G[f"{name}_org2"], G[f"{name}_dir2_any"], any_hit=True)
    b = hc.trace_ii(G[f"{name}_org2"], G[f"{name}_dir2_any"], any_hit=True, post_limit=post_limit)
    assert np.array_equal(a, b) and np.array_equal(b, G[f"{name}_occ2"])


def test_survey_cornell_kats(S):
    # camera rays with u2 = (0.5, 0.5) on the shipped 1024^2 Cornell scene
    o = Oracle(S["cornell"])
    xy = np.array([[0, 0], [512, 512], [256, 768], [1023, 1023], [512, 10]], np.int32)
    org, dr = o.generate_rays(xy, np.full((5, 2), 0.5, np.float32))
    h = o.trace(org, dr)
    assert list(h["prim_id"]) == [-1, 30, 25, -1, -1] and list(h["geom_id"]) == [-1, 0, 0, -1, -1]
    assert np.allclose(h["t"][[1, 2]], [9.07773304, 9.14707279], rtol=1e-7)
    assert np.allclose(dr[1, :3], [0.000127106905, -0.000127106905, -1], atol=1e-9)
    assert np.isinf(h["t"][0])


@pytest.mark.parametrize("which", ["oracle", "hostcheck"])
def test_films_golden(G, S, which):
    for name, key in (("cornell32", "cornell32_film"), ("mixed32", "mixed32_film"), ("mixed24_pcg", "mixed24_pcg_film"),
                      ("terrain", "terrain_film")):
        sc = S[name]
        impl = Oracle(sc) if which == "oracle" else HostCheck(sc)
        film = impl.render(sc.integrator.params())
        assert same_bits(film, G[key]), (name, int((bits(film) != bits(G[key])).any(-1).sum()))


def test_survey_pixel_sums(G, S):
    # sum of 16 path samples of two pixels of the shipped 1024^2 scene
    o = Oracle(S["cornell"])
    p = S["cornell"].integrator.params()
    f = o.render(p, rect=(512, 512, 513, 513))
    assert same_bits(f[512, 512], G["cornell1024_px_512_512"])
    assert np.allclose(f[512, 512, :3], [2.32806635, 1.63713145, 0.457783341], rtol=1e-6)
    f = o.render(p, rect=(256, 768, 257, 769))
    assert same_bits(f[768, 256], G["cornell1024_px_256_768"])
    assert np.allclose(f[768, 256, :3], [1.36838162, 0.412676632, 0.11681854], rtol=1e-6)
    srgb = o.film_to_srgb8(f)
    assert tuple(srgb[768, 256]) == (83, 45, 21)


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_vs_reference_live():
    """Fresh inputs (not the committed vectors): oracle and hostcheck against the reference's own code."""
    rng = np.random.default_rng(5)
    sc = scenes.terrain_scene(n=96, resolution=(80, 45), spp=3, max_depth=6, materials="c3ref")
    ref, o, h = Reference(sc), Oracle(sc), HostCheck(sc)
    org, dr = random_rays(rng, 20000, (-50, 0, -50), (50, 12, 50))
    hr = ref.trace(org, dr)
    for impl in (o, h):
        bad, ties, tb = hit_mismatch(hr, impl.trace(org, dr))
        assert bad == 0 and tb == 0
    dra = dr.copy()
    dra[:, 3] = 30
    occ = ref.trace(org, dra, any_hit=True)
    assert np.array_equal(occ, o.trace(org, dra, any_hit=True)) and np.array_equal(occ, h.trace(org, dra, any_hit=True))
    p = sc.integrator.params()
    fr = ref.render(p)
    assert same_bits(fr, o.render(p)) and same_bits(fr, h.render(p))


def test_extensions_oracle_vs_hostcheck():
    """Dielectric + Russian roulette are extensions (parity unpinned by the reference): the oracle's CPU statement and
    the product's device functions (compiled for the host) must agree bit for bit."""
    sc = scenes.terrain_scene(n=40, resolution=(48, 27), spp=4, max_depth=8, materials="c3", rr_start_depth=3)
    p = sc.integrator.params()
    assert same_bits(Oracle(sc).render(p), HostCheck(sc).render(p))
    sc = scenes.many_light_scene(n=32, resolution=(40, 24), spp=2, max_depth=6, n_quads=200)
    p = sc.integrator.params()
    assert same_bits(Oracle(sc).render(p), HostCheck(sc).render(p))


def test_ray_clamp_and_sharding_hostcheck():
    sc = scenes.cornell_box((40, 24), 4, 5)
    sc.integrator.ray_clamp = 0.5
    p = sc.integrator.params()
    full = HostCheck(sc).render(p)
    assert same_bits(full, Oracle(sc).render(p))
    assert full[..., :3].max() <= 0.5 * 4 + 1e-6
    # interleaved 8x8 tiles over 3 ranks: disjoint cover, sum == single-rank film
    parts = [HostCheck(sc).render(sc.integrator.params(rank=r, world_size=3, shard_tile=8)) for r in range(3)]
    owned = np.stack([(q[..., 3] > 0) for q in parts])
    assert (owned.sum(0) == 1).all()
    assert same_bits(sum(parts), full)


def test_bvh8_shape():
    sc = scenes.terrain_scene(n=128, resolution=(8, 8), spp=1)
    info = HostCheck(sc).info()
    assert info["n_triangles"] == 2 * 127 * 127 + 2
    assert info["bvh_bytes"] == 80 * info["n_bvh_nodes"] and info["triangle_bytes"] == 48 * info["n_triangles"]
    assert info["bvh_depth"] <= 30
    # wide nodes should be well filled: < 1 node per 3 triangles
    assert info["n_bvh_nodes"] * 3 < info["n_triangles"]


def test_degenerate_scenes_hostcheck():
    # empty scene, a scene with no lights, coincident triangles (exact ties -> lowest id wins, like brute force)
    empty = scenes.Scene(scenes.PerspectiveCamera(resolution=(8, 8)), scenes.Path(spp=1, ray_clamp=0.0), [])
    p = empty.integrator.params()
    f = HostCheck(empty).render(p)
    assert (f[..., :3] == 0).all() and (f[..., 3] == 1).all() and same_bits(f, Oracle(empty).render(p))
    v = np.array([[-1, -1, -3], [1, -1, -3], [0, 1, -3]], np.float32)
    idx = np.array([[0, 1, 2], [0, 1, 2], [0, 1, 2]], np.int32)
    mesh = scenes.AkariMesh(materials=[scenes.DiffuseMaterial(0.5)], vertices=v, indices=idx, material_indices=np.zeros(3, np.int32))
    sc = scenes.Scene(scenes.PerspectiveCamera(fov=60, resolution=(16, 16)), scenes.Path(spp=2, ray_clamp=0.0), [mesh])
    h, ob = HostCheck(sc), Oracle(sc, brute_force=True)
    org = np.array([[0, 0, 0, 1e-3]], np.float32)
    dr = np.array([[0, 0, -1, np.inf]], np.float32)
    hh, ho = h.trace(org, dr), ob.trace(org, dr)
    assert hh["prim_id"][0] == 0 == ho["prim_id"][0] and hh["t"][0] == 3.0
    assert same_bits(h.render(sc.integrator.params()), ob.render(sc.integrator.params()))


def test_device_trig_restates_glibc():
    """csrc/shade.h sincosf_glibc (what the kernels call) against this host's libm sinf/cosf."""
    import ctypes as C
    from oracle_libs import HOSTCHECK_SO, build_hostcheck
    build_hostcheck()
    lib = C.CDLL(HOSTCHECK_SO)
    lib.hc_trig_mismatches.restype = C.c_long
    assert lib.hc_trig_mismatches(C.c_long(5_000_000)) <= 1
