"""Scene-file front-end: `.akari` parser, `.mesh` reader, node flattening, PNG/PFM writers, CLI.

CPU tests: the C++ front-end must flatten a scene file to the very bytes the node mirror (validated against the
reference through the golden vectors) produces, accept/reject what the reference's parser accepts/rejects
(core/parser.cpp), and - in the build container, where /root/reference exists - read the reference's own shipped
Cornell scene. GPU tests: a scene file rendered through the front-end (library and `akari-b200` command) gives the
same film / PNG as the direct C-ABI route.
"""
import ctypes as C
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

from common import ROOT, abi, nodes, pkg, scenes

REF_SCENE = "/root/reference/resources/data/cornell_box/scene.akari"
CLI = os.path.join(ROOT, "akarirender-1_b200", "bin", "akari-b200")


def struct_bytes(s):
    return bytes(C.string_at(C.addressof(s), C.sizeof(s)))


def mesh_arrays(m):
    nv, nt = int(m.n_vertices), int(m.n_triangles)
    out = []
    for ptr, n in ((m.vertices, 3 * nv), (m.normals, 9 * nt), (m.texcoords, 6 * nt), (m.indices, 3 * nt),
                   (m.material_indices, nt), (m.materials, int(m.n_mesh_materials))):
        out.append(bytes(C.string_at(ptr, 4 * n)) if n else b"")
    return out


def assert_same_desc(a, b):
    assert a.abi_version == b.abi_version and a.n_meshes == b.n_meshes and a.n_materials == b.n_materials
    assert struct_bytes(a.camera) == struct_bytes(b.camera)
    for i in range(a.n_materials):
        assert struct_bytes(a.materials[i]) == struct_bytes(b.materials[i]), "material %d" % i
    for i in range(a.n_meshes):
        assert mesh_arrays(a.meshes[i]) == mesh_arrays(b.meshes[i]), "mesh %d" % i


ROUND_TRIP = {
    "cornell": lambda: scenes.cornell_box((1024, 1024), 16, 5),
    "mixed_pcg": lambda: scenes.mixed_box((24, 24), 4, 3, sampler="pcg", sort_materials=True),
    "terrain_c3": lambda: scenes.terrain_scene(n=24, resolution=(64, 36), spp=4, max_depth=8, materials="c3"),
    "instanced": lambda: scenes.instanced_scene(n=12, resolution=(48, 27), spp=2, copies=(2, 2)),
    "many_light": lambda: scenes.many_light_scene(n=16, resolution=(32, 18), spp=2, n_quads=200),
    "megakernel": lambda: _with_clamp(scenes.cornell_box((16, 16), 2, 3, megakernel=True), 10.0),
}


def _with_clamp(sc, clamp):
    sc.integrator.ray_clamp = clamp
    return sc


@pytest.mark.parametrize("name", sorted(ROUND_TRIP))
def test_scene_file_flattens_like_the_node_mirror(tmp_path, name):
    sc = ROUND_TRIP[name]()
    path = nodes.write_akari(sc, str(tmp_path))
    sf = nodes.SceneFile(path)
    want, keep = sc.flatten()
    assert_same_desc(sf.desc, want)
    assert struct_bytes(sf.params()) == struct_bytes(sc.integrator.params())
    assert sf.output == sc.output
    sf.close()


def test_ao_node_and_defaults(tmp_path):
    sc = scenes.cornell_box((16, 16), 2, 3)
    sc.integrator = nodes.AO(spp=7, occlude=2.5)
    sf = nodes.SceneFile(nodes.write_akari(sc, str(tmp_path)))
    p = sf.params()
    assert (p.integrator, p.spp, p.ao_occlude) == (abi.INTEGRATOR_AO, 7, 2.5)
    # node defaults: core/nodes/integrator.cpp:30-32, :53-57; core/nodes/camera.cpp:30-31
    (tmp_path / "d.akari").write_text(
        'export m = DiffuseMaterial { color: 0.5 }\nexport s = AkariMesh { path: "shape0.mesh", materials: [$m,$m,$m,$m,$m,$m,$m,$m] }\n'
        "export scene = Scene { camera: PerspectiveCamera {}, integrator: Path {}, shapes: [$s] }\n")
    sf = nodes.SceneFile(str(tmp_path / "d.akari"))
    p = sf.params()
    assert (p.integrator, p.spp, p.max_depth, p.ray_clamp, p.mode, p.sampler) == (abi.INTEGRATOR_PATH, 16, 5, 10.0,
                                                                                   abi.MODE_WAVEFRONT, abi.SAMPLER_LCG)
    assert sf.resolution == (512, 512) and sf.output == "out.png"
    want = nodes.PerspectiveCamera().desc()
    assert struct_bytes(sf.desc.camera) == struct_bytes(want)
    assert sf.desc.n_materials == 1 and list(sf.desc.materials[0].color) == [0.5, 0.5, 0.5]  # one shared node
    (tmp_path / "ao.akari").write_text(
        'export s = AkariMesh { path: "shape0.mesh", materials: [] }\n'
        "export scene = Scene { camera: PerspectiveCamera {}, integrator: AO {}, shapes: [$s] }\n")
    p = nodes.SceneFile(str(tmp_path / "ao.akari")).params()
    assert p.integrator == abi.INTEGRATOR_AO and p.spp == 16 and np.isinf(p.ao_occlude)


def _load_error(path):
    lib = abi.load()
    h = C.c_void_p()
    rc = lib.akb_scene_file_load(str(path).encode(), C.byref(h))
    if rc == abi.OK:
        lib.akb_scene_file_destroy(h)
        return None
    assert not h.value
    return lib.akb_last_error().decode()


def test_parser_grammar(tmp_path):
    sc = scenes.cornell_box((8, 8), 1, 1)
    nodes.write_akari(sc, str(tmp_path))  # leaves shape0.mesh
    (tmp_path / "lib").mkdir()
    nodes.write_mesh(str(tmp_path / "lib" / "m.mesh"), sc.shapes[0])
    (tmp_path / "lib" / "inner.akari").write_text(
        "// nested module; mesh path is relative to THIS file\n"
        'export white = DiffuseMaterial { color: [0.5, 0.25, 0.125], }\n'
        'export box = AkariMesh { path: "m.mesh", materials: [$white, $white, $white, $white, $white, $white, $white, $white,], }\n')
    (tmp_path / "mid.akari").write_text('import "lib/inner.akari" as inner\nexport fwd = $inner.box\n')
    (tmp_path / "top.akari").write_text(
        'import "mid.akari" as mid   // trailing comment\n'
        "let fov = 33.5\n"
        "export fov = 99\n"
        'let name = "a \\"quoted\\" \\\\ name"\n'
        "export scene = Scene {\n"
        "  // comment inside an object\n"
        "  camera: PerspectiveCamera { fov: $fov, position: [-1.5, 0, 2.25], resolution: [8, 4] },\n"
        "  integrator : Path { spp : 3 , unknown_field: [1, [2, 3]], wavefront: false },\n"
        "  output: $name,\n"
        "  shapes: [ $mid.inner.box, $mid.fwd ]\n"
        "}")  # no trailing newline
    sf = nodes.SceneFile(str(tmp_path / "top.akari"))
    assert sf.output == 'a "quoted" \\ name'
    assert sf.desc.n_meshes == 2 and sf.desc.n_materials == 1  # the same node through two reference paths
    assert list(sf.desc.materials[0].color) == [0.5, 0.25, 0.125]
    p = sf.params()
    assert p.spp == 3 and p.mode == abi.MODE_MEGAKERNEL
    want = nodes.PerspectiveCamera(fov=33.5, position=(-1.5, 0, 2.25), resolution=(8, 4)).desc()  # the local wins
    assert struct_bytes(sf.desc.camera) == struct_bytes(want)

    bad = {
        "export scene = Nope { }\n": "unknown node type Nope",
        "export a = 1\nexport a = 2\n": "already defined",
        "export a = DiffuseMaterial { color: 1, color: 2 }\n": "redefined",
        "export a = $missing\n": "no exported variable",
        "export a = $m.x\n": "no submodule",
        'import "nowhere.akari" as n\n': "not found",
        "export a = 1 export b = 2\n": "newline expected",
        "export a = 1e5\n": "newline expected",  # no exponent syntax (core/parser.cpp:338-363)
        "export a = .5\n": "stray token",
        "frobnicate\n": "stray token",
        'export a = "x\\q"\n': "escape",
        "export a = [1, 2\n": "expected",
        "export a = 1\n": "'scene' not found",
        "export scene = DiffuseMaterial { color: 1 }\n": "not a Scene",
        "export scene = Scene { camera: PerspectiveCamera {}, integrator: Path {}, shapes: [ DiffuseMaterial { color: 1 } ] }\n": "AkariMesh",
        'export scene = Scene { camera: PerspectiveCamera {}, integrator: Path {}, shapes: [ AkariMesh { path: "none.mesh" } ] }\n': "none.mesh",
        'export scene = Scene { camera: PerspectiveCamera {}, integrator: Path {}, shapes: [ AkariMesh { path: "shape0.mesh", materials: [ DiffuseMaterial { color: "tex.png" } ] } ] }\n': "tex.png: cannot open",
    }
    for text, needle in bad.items():
        (tmp_path / "bad.akari").write_text(text)
        err = _load_error(tmp_path / "bad.akari")
        assert err is not None and needle in err, (text, err)
    (tmp_path / "trunc.mesh").write_bytes(open(tmp_path / "shape0.mesh", "rb").read()[:200])
    (tmp_path / "bad.akari").write_text(
        'export scene = Scene { camera: PerspectiveCamera {}, integrator: Path {}, shapes: [ AkariMesh { path: "trunc.mesh" } ] }\n')
    assert "truncated" in _load_error(tmp_path / "bad.akari")
    assert "cannot open" in _load_error(tmp_path / "does_not_exist.akari")


def test_number_literals_read_like_the_reference(tmp_path):
    """core/parser.cpp:338-363 reads n + frac / 10^k in double; every float32 the serializer emits must come back."""
    rng = np.random.default_rng(7)
    vals = np.concatenate([rng.uniform(-1e3, 1e3, 200), rng.uniform(0, 1, 200) ** 8, [0.0, 1 / 3, 1e-7, 16777217.0]])
    vals = vals.astype(np.float32)
    sc = scenes.cornell_box((8, 8), 1, 1)
    nodes.write_akari(sc, str(tmp_path))
    mats = "\n".join("export m%d = DiffuseMaterial { color: [%s, 0, 0] }" % (i, nodes.sdl_number(v)) for i, v in enumerate(vals))
    refs = ", ".join("$m%d" % i for i in range(len(vals)))
    (tmp_path / "n.akari").write_text(
        mats + '\nexport s = AkariMesh { path: "shape0.mesh", materials: [%s] }\n' % refs
        + "export scene = Scene { camera: PerspectiveCamera {}, integrator: Path {}, shapes: [$s] }\n")
    sf = nodes.SceneFile(str(tmp_path / "n.akari"))
    got = np.array([sf.desc.materials[i].color[0] for i in range(len(vals))], np.float32)
    assert np.array_equal(got.view(np.uint32), vals.view(np.uint32))


@pytest.mark.skipif(not os.path.exists(REF_SCENE), reason="reference tree not present (GPU box)")
def test_reads_the_reference_cornell_scene():
    """The reference's own shipped scene (scene.akari -> cornell_box.akari -> CornellBox-Original.obj.mesh)."""
    sf = nodes.SceneFile(REF_SCENE)
    sc = scenes.cornell_box((1024, 1024), 16, 5)
    want, keep = sc.flatten()
    assert_same_desc(sf.desc, want)
    p = sf.params()
    assert (p.integrator, p.spp, p.max_depth, p.ray_clamp, p.mode) == (abi.INTEGRATOR_PATH, 16, 5, 10.0, abi.MODE_WAVEFRONT)
    assert sf.output == "out.png"


def decode_png(path):
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    o, idat, w = 8, b"", None
    while o < len(b):
        n, tag = struct.unpack(">I4s", b[o:o + 8])
        body = b[o + 8:o + 8 + n]
        assert struct.unpack(">I", b[o + 8 + n:o + 12 + n])[0] == zlib.crc32(tag + body)
        if tag == b"IHDR":
            w, h, depth, ctype, comp, flt, lace = struct.unpack(">IIBBBBB", body)
            assert (depth, ctype, comp, flt, lace) == (8, 2, 0, 0, 0)
        elif tag == b"IDAT":
            idat += body
        o += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 3 * w + 1)
    assert not raw[:, 0].any()
    return raw[:, 1:].reshape(h, w, 3)


def test_png_and_pfm_writers(tmp_path):
    lib = abi.load()
    rng = np.random.default_rng(3)
    for (w, h) in ((1, 1), (37, 11), (300, 120)):  # the last one needs more than one stored block
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        p = str(tmp_path / ("t%d.png" % w))
        assert lib.akb_write_png(p.encode(), img.ctypes.data_as(C.c_void_p), w, h) == abi.OK
        assert np.array_equal(decode_png(p), img)
    film = rng.random((5, 7, 4)).astype(np.float32)
    film[..., 3] = 4.0
    film[0, 0, 3] = 0.0
    p = str(tmp_path / "t.pfm")
    assert lib.akb_write_pfm(p.encode(), film.ctypes.data_as(C.c_void_p), 7, 5) == abi.OK
    b = open(p, "rb").read()
    head = b"PF\n7 5\n-1.0\n"
    assert b.startswith(head)
    px = np.frombuffer(b[len(head):], "<f4").reshape(5, 7, 3)[::-1]
    want = film[..., :3] / np.where(film[..., 3:] != 0, film[..., 3:], 1)
    assert np.array_equal(px, want.astype(np.float32))
    assert lib.akb_write_png(b"/nonexistent_dir/x.png", film.ctypes.data_as(C.c_void_p), 1, 1) == abi.INVALID_ARG


def test_cli_fails_loudly_without_a_device(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    assert os.path.exists(CLI), "akari-b200 not built"
    path = nodes.write_akari(scenes.cornell_box((8, 8), 1, 1), str(tmp_path))
    r = subprocess.run([CLI, path], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode != 0 and "no CUDA device" in r.stderr
    assert not os.path.exists(tmp_path / "out.png")


@pytest.mark.gpu
def test_scene_file_renders_like_the_direct_route(tmp_path):
    sc = scenes.mixed_box((64, 48), 8, 5)
    path = nodes.write_akari(sc, str(tmp_path))
    sf = nodes.SceneFile(path)
    a = pkg.Backend(sf, 0)
    b = pkg.Backend(sc, 0)
    fa, sa = a.render(sf.params())
    fb, sb = b.render(sc.integrator.params())
    assert np.array_equal(fa.view(np.uint32), fb.view(np.uint32))
    assert sa.n_closest_rays == sb.n_closest_rays and sa.n_shadow_rays == sb.n_shadow_rays
    # the command-line driver: same pixels in the PNG it writes as Film::write_image's post-process gives
    r = subprocess.run([CLI, "--gpu", path, "--pfm", "out.pfm"], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr
    assert "render done" in r.stdout
    assert np.array_equal(decode_png(str(tmp_path / "out.png")), b.film_to_srgb8(fb))
    raw = open(tmp_path / "out.pfm", "rb").read()
    px = np.frombuffer(raw[len(b"PF\n64 48\n-1.0\n"):], "<f4").reshape(48, 64, 3)[::-1]
    assert np.array_equal(px, (fb[..., :3] / fb[..., 3:]).astype(np.float32))
    # --bvh device: the tree is built on the GPU, the image is the same
    r = subprocess.run([CLI, path, "--bvh", "device", "-o", "dev.png"], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 0 and "device build" in r.stdout, r.stdout + r.stderr
    assert np.array_equal(decode_png(str(tmp_path / "dev.png")), b.film_to_srgb8(fb))
    a.close(), b.close()
