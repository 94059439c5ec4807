"""ctypes loaders for the two CPU checkers (TEST INFRASTRUCTURE — never imported by the product package):

  * `Oracle`   : oracle/libakari_oracle.so  — the plain-C restatement (oracle/akari_oracle.c)
  * `Reference`: oracle/_ref/libakari_ref.so — the reference's own code behind oracle/ref_harness.cpp

Both take the same `akb_scene_desc` the product library takes.
"""
import ctypes as C
import importlib
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("akarirender-1_b200")
abi = pkg.abi

ORACLE_SO = os.path.join(ROOT, "oracle", "libakari_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libakari_ref.so")


def build_oracle():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True)


def have_ref():
    return os.path.exists(REF_SO)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class _Base:
    prefix = None

    def __init__(self, scene, lib, create_args=()):
        self.lib = lib
        self.desc, self._keep = scene.flatten()
        self.width, self.height = scene.camera.resolution
        create = getattr(lib, self.prefix + "_scene_create")
        create.restype = C.c_void_p
        self.h = create(C.byref(self.desc), *create_args)
        if not self.h:
            raise RuntimeError(f"{self.prefix}_scene_create failed")
        self.h = C.c_void_p(self.h)

    def close(self):
        if self.h:
            getattr(self.lib, self.prefix + "_scene_destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def n_lights(self):
        f = getattr(self.lib, self.prefix + "_n_lights")
        f.restype = C.c_uint64
        return int(f(self.h))

    def trace(self, org_tmin, dir_tmax, any_hit=False, threads=0):
        org = np.ascontiguousarray(org_tmin, np.float32)
        dr = np.ascontiguousarray(dir_tmax, np.float32)
        n = org.shape[0]
        hits = np.zeros(n, dtype=HIT_DTYPE)
        occ = np.zeros(n, np.int32)
        f = getattr(self.lib, self.prefix + "_trace")
        rc = f(self.h, _ptr(org), _ptr(dr), C.c_size_t(n), C.c_int(int(any_hit)), _ptr(hits), _ptr(occ),
               C.c_int(threads or os.cpu_count()))
        assert rc == 0
        return occ if any_hit else hits

    def generate_rays(self, xy, u2):
        xy = np.ascontiguousarray(xy, np.int32)
        u2 = np.ascontiguousarray(u2, np.float32)
        n = xy.shape[0]
        org = np.zeros((n, 4), np.float32)
        dr = np.zeros((n, 4), np.float32)
        rc = getattr(self.lib, self.prefix + "_generate_rays")(self.h, _ptr(xy), _ptr(u2), C.c_size_t(n), _ptr(org),
                                                                _ptr(dr))
        assert rc == 0
        return org, dr

    def sampler_kat(self, kind, seed, n):
        out = np.zeros(n, np.float32)
        n2 = np.zeros(2, np.float32)
        getattr(self.lib, self.prefix + "_sampler_kat")(C.c_int(kind), C.c_uint64(seed), C.c_int(n), _ptr(out), _ptr(n2))
        return out, n2


HIT_DTYPE = np.dtype([("t", np.float32), ("u", np.float32), ("v", np.float32), ("prim_id", np.int32),
                      ("geom_id", np.int32)])


class Oracle(_Base):
    prefix = "akor"

    def __init__(self, scene, brute_force=False):
        if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(
                os.path.join(ROOT, "oracle", "akari_oracle.c")):
            build_oracle()
        super().__init__(scene, C.CDLL(ORACLE_SO), (C.c_int(int(brute_force)),))

    def render(self, params, rect=None, threads=0):
        x0, y0, x1, y1 = rect or (0, 0, self.width, self.height)
        film = np.zeros((self.height, self.width, 4), np.float32)
        counts = (C.c_uint64 * 3)()
        rc = self.lib.akor_render(self.h, C.byref(params), x0, y0, x1, y1, _ptr(film),
                                  C.c_int(threads or os.cpu_count()), counts)
        assert rc == 0
        self.counts = tuple(int(c) for c in counts)
        return film

    def render_rects(self, params, rects, threads=0):
        """All rectangles in ONE call (one work list of 16x16 tiles over every host thread); sets self.busy =
        (CPU seconds summed over the workers, wall seconds)."""
        film = np.zeros((self.height, self.width, 4), np.float32)
        r = np.ascontiguousarray(rects, np.int32).reshape(-1, 4)
        counts = (C.c_uint64 * 3)()
        busy = (C.c_double * 2)()
        rc = self.lib.akor_render_rects(self.h, C.byref(params), _ptr(r), C.c_int(r.shape[0]), _ptr(film),
                                        C.c_int(threads or os.cpu_count()), counts, busy)
        assert rc == 0
        self.counts = tuple(int(c) for c in counts)
        self.busy = (float(busy[0]), float(busy[1]))
        return film

    def pixel_rays(self, params, x, y, cap=65536):
        """Every ray the Path integrator traces for pixel (x, y), in order: (is_any_hit[n], org_tmin[n,4], dir_tmax[n,4])."""
        buf = np.zeros((cap, 9), np.float32)
        n = self.lib.akor_pixel_rays(self.h, C.byref(params), int(x), int(y), _ptr(buf), cap)
        r = buf[:n]
        return (r[:, 0] != 0, np.ascontiguousarray(np.concatenate([r[:, 1:4], r[:, 7:8]], 1)),
                np.ascontiguousarray(np.concatenate([r[:, 4:7], r[:, 8:9]], 1)))

    def film_to_srgb8(self, film):
        h, w = film.shape[:2]
        out = np.zeros((h, w, 3), np.uint8)
        self.lib.akor_film_to_srgb8(_ptr(np.ascontiguousarray(film, np.float32)), w, h, _ptr(out))
        return out


class Reference(_Base):
    prefix = "akref"

    def __init__(self, scene):
        super().__init__(scene, C.CDLL(REF_SO))

    def render(self, params, rect=None, threads=0):
        x0, y0, x1, y1 = rect or (0, 0, self.width, self.height)
        film = np.zeros((self.height, self.width, 4), np.float32)
        rc = self.lib.akref_render(self.h, C.byref(params), x0, y0, x1, y1, _ptr(film),
                                   C.c_int(threads or os.cpu_count()))
        assert rc == 0
        return film


    def render_rects(self, params, rects, threads=0):
        """All rectangles in ONE call: the reference's scheduling (16x16 tiles, one shared work list, all host threads);
        sets self.busy = (CPU seconds summed over the workers, wall seconds)."""
        film = np.zeros((self.height, self.width, 4), np.float32)
        r = np.ascontiguousarray(rects, np.int32).reshape(-1, 4)
        busy = (C.c_double * 2)()
        rc = self.lib.akref_render_rects(self.h, C.byref(params), _ptr(r), C.c_int(r.shape[0]), _ptr(film),
                                         C.c_int(threads or os.cpu_count()), busy)
        assert rc == 0
        self.busy = (float(busy[0]), float(busy[1]))
        return film


REF_GPU_SO = os.path.join(ROOT, "oracle", "_ref", "libakari_ref_gpu.so")


def have_ref_gpu():
    return os.path.exists(REF_GPU_SO)


class ReferenceGPU(_Base):
    """The reference's OWN CUDA wavefront path tracer (gpu::PathTracer, kernel/integrators/gpu/cuda/integrator.cpp),
    patched in transit to compile for sm_100a (oracle/Makefile target ref_gpu). A comparison point
    on the same box, not a parity oracle: it is built with --use_fast_math, draws one more random number per bounce and
    clamps samples (gpu/cuda/integrator.cpp:286-287,397-398)."""
    prefix = "akref"

    def __init__(self, scene):
        super().__init__(scene, C.CDLL(REF_GPU_SO))

    def render_gpu(self, params, tile_size=256, wavefront=True):
        """-> (film, milliseconds of render() + device sync)."""
        film = np.zeros((self.height, self.width, 4), np.float32)
        ms = C.c_double(0)
        rc = self.lib.akref_gpu_render(self.h, C.byref(params), C.c_int(tile_size), C.c_int(int(wavefront)), _ptr(film),
                                       C.byref(ms))
        if rc != 0:
            self.lib.akref_last_error.restype = C.c_char_p
            raise RuntimeError(self.lib.akref_last_error().decode(errors="replace"))
        return film, ms.value


HOSTCHECK_DIR = os.path.join(ROOT, "tests", "hostcheck")
HOSTCHECK_SO = os.path.join(HOSTCHECK_DIR, "libakari_hostcheck.so")
CSRC = os.path.join(ROOT, "akarirender-1_b200", "csrc")


def build_hostcheck():
    srcs = [os.path.join(HOSTCHECK_DIR, "hostcheck.cpp"), os.path.join(CSRC, "scene_host.cpp"),
            os.path.join(CSRC, "bvh8.cpp")]
    deps = srcs + [os.path.join(CSRC, f) for f in ("scene_host.h", "bvh8.h", "shade.h", "traverse.h", "vecmath.h")]
    if os.path.exists(HOSTCHECK_SO) and all(os.path.getmtime(HOSTCHECK_SO) >= os.path.getmtime(d) for d in deps):
        return
    subprocess.run(["g++", "-std=c++17", "-O2", "-mavx2", "-mfma", "-ffp-contract=off", "-fPIC", "-shared", "-pthread",
                    "-o", HOSTCHECK_SO] + srcs, check=True)


class HostCheck(_Base):
    """The product's host-side scene build + the kernels' AKB_HD device functions compiled for the CPU
    (tests/hostcheck/hostcheck.cpp). Test infrastructure: checks host logic without a GPU."""
    prefix = "hc"

    def __init__(self, scene, threads=0):
        build_hostcheck()
        super().__init__(scene, C.CDLL(HOSTCHECK_SO), (C.c_int(threads or os.cpu_count()),))
        self.lib.hc_info.restype = C.c_uint64

    def info(self):
        keys = ["n_triangles", "n_bvh_nodes", "bvh_bytes", "n_lights", "triangle_bytes", "bvh_depth", "build_us"]
        return {k: int(self.lib.hc_info(self.h, i)) for i, k in enumerate(keys)}

    def trace_ii(self, org_tmin, dir_tmax, any_hit=False, post_limit=-1):
        """The same rays through the if-if traversal step (traverse.h trav_step_ii)."""
        org = np.ascontiguousarray(org_tmin, np.float32)
        dr = np.ascontiguousarray(dir_tmax, np.float32)
        n = org.shape[0]
        hits = np.zeros(n, dtype=HIT_DTYPE)
        occ = np.zeros(n, np.int32)
        rc = self.lib.hc_trace_ii(self.h, _ptr(org), _ptr(dr), C.c_size_t(n), C.c_int(int(any_hit)), C.c_int(post_limit),
                                  _ptr(hits), _ptr(occ))
        assert rc == 0
        return occ if any_hit else hits

    def render(self, params, threads=0):
        film = np.zeros((self.height, self.width, 4), np.float32)
        rc = self.lib.hc_render(self.h, C.byref(params), _ptr(film), C.c_int(threads or os.cpu_count()))
        assert rc == 0
        return film
