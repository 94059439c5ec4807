"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/akari_b200.h declares,
its structs have the layout the ctypes mirror assumes, and compute calls fail loudly without a GPU."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from common import ROOT, abi, pkg, scenes


def test_library_exports_every_declared_symbol():
    lib = abi.load()
    hdr = open(os.path.join(ROOT, "include", "akari_b200.h")).read()
    declared = set(re.findall(r"\b(akb_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(abi.SYMBOLS), declared ^ set(abi.SYMBOLS)
    for s in declared:
        assert hasattr(lib, s), s


def test_struct_layouts_match_header(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "akari_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(akb_material),sizeof(akb_mesh),sizeof(akb_camera),sizeof(akb_scene_desc),"
                   "sizeof(akb_render_params),sizeof(akb_render_stats),sizeof(akb_hit),"
                   "(size_t)&((akb_scene_desc*)0)->camera);return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(abi.Material), C.sizeof(abi.Mesh), C.sizeof(abi.Camera), C.sizeof(abi.SceneDesc),
            C.sizeof(abi.RenderParams), C.sizeof(abi.RenderStats), C.sizeof(abi.Hit), abi.SceneDesc.camera.offset]
    assert got == want


def test_no_cpu_fallback_without_a_device():
    lib = abi.load()
    if lib.akb_device_count() > 0:
        pytest.skip("a CUDA device is present")
    assert lib.akb_init(0) == abi.NO_DEVICE
    assert b"no CPU fallback" in lib.akb_last_error()
    assert lib.akb_init_device(0) == abi.NO_DEVICE and lib.akb_n_devices() == 0
    sc = scenes.cornell_box((8, 8), 1, 1)
    desc, keep = sc.flatten()
    h = C.c_void_p()
    assert lib.akb_scene_create(C.byref(desc), C.byref(h)) == abi.NO_DEVICE and not h.value
    film = np.zeros((8, 8, 4), np.float32)
    p = sc.integrator.params()
    assert lib.akb_render(None, C.byref(p), film.ctypes.data_as(C.c_void_p), None) == abi.NO_DEVICE
    with pytest.raises(pkg.AkariError):
        pkg._initialised = None
        pkg.Backend(sc)


def test_product_package_never_touches_the_oracle():
    """The oracle is a checker: nothing under the package (Python or csrc) may name it."""
    pdir = os.path.join(ROOT, "akarirender-1_b200")
    for dp, _, files in os.walk(pdir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, f)).read()
                assert "libakari_oracle" not in txt and "oracle/" not in txt.replace("oracle/akari_oracle.c", "") \
                    and "hostcheck" not in txt.replace("tests/hostcheck", ""), f


def test_options_are_explicit_not_environment():
    """Tuning goes through akb_set_option (no device needed to set one): unknown names and bad values are refused,
    and the library's sources do not read the environment."""
    lib = abi.load()
    assert lib.akb_set_option(b"finish_div", b"8") == abi.OK
    assert lib.akb_set_option(b"finish_div", b"4") == abi.OK
    assert lib.akb_set_option(b"combine", b"p2p") == abi.OK and lib.akb_set_option(b"combine", b"nccl") == abi.OK
    assert lib.akb_set_option(b"no_such_switch", b"1") == abi.INVALID_ARG
    assert b"unknown option" in lib.akb_last_error()
    assert lib.akb_set_option(b"pipelines", b"7") == abi.INVALID_ARG
    assert lib.akb_set_option(b"bvh_build", b"maybe") == abi.INVALID_ARG
    assert lib.akb_set_option(None, b"1") == abi.INVALID_ARG
    for d in ("csrc", "host"):
        for f in os.listdir(os.path.join(ROOT, "akarirender-1_b200", d)):
            assert "getenv" not in open(os.path.join(ROOT, "akarirender-1_b200", d, f)).read(), f
