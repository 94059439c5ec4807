"""akarirender-1_b200 — B200-native backend for AkariRender's Path-integrator hot path.

The product is libakari_b200.so (csrc/, hand-written sm_100a CUDA behind the C ABI of include/akari_b200.h).
This package is the thin host-side mirror used by tests and bench.py: the reference's node vocabulary
(`nodes`), synthetic scenes of the BASELINE shapes (`scenes`) and `Backend`, which drives the C ABI through
ctypes. There is no CPU fallback: everything here fails loudly if the CUDA library is missing.

(The directory name is not a Python identifier; import it with importlib.import_module("akarirender-1_b200").)
"""
import ctypes as C

import numpy as np

from . import abi, nodes, scenes  # noqa: F401

# `distributed` (torch.distributed plumbing for the multi-GPU frame split) is imported lazily: it needs torch.

HIT_DTYPE = np.dtype([("t", np.float32), ("u", np.float32), ("v", np.float32), ("prim_id", np.int32),
                      ("geom_id", np.int32)])


class AkariError(RuntimeError):
    pass


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


_initialised = None

# Development aid of THIS Python layer (the library itself never reads the environment): AKB_<OPTION>=value variables
# are forwarded once to akb_set_option, so A/B scripts can flip a switch without editing code. tests use set_option().
_ENV_OPTIONS = ("pipelines", "pipeline_min_slots", "graphs", "check_every", "poll_lag", "finish_div", "finish_max",
                "no_park", "shade_partition", "tail_share", "tail_share_min", "refill_below", "fetch_chunk", "profile_stages", "timeline", "trace_chunk", "trav_ifif",
                "bvh_build", "bvh_verbose", "bvh_trav_cost", "bvh_prim_cost", "bvh_max_leaf", "bvh_device_algo",
                "ploc_radius", "combine", "nvtx", "stream", "stream_flex_mask", "stream_watchdog_ms")
_env_forwarded = False


def _forward_env(lib):
    global _env_forwarded
    if _env_forwarded:
        return
    _env_forwarded = True
    import os
    for name in _ENV_OPTIONS:
        v = os.environ.get("AKB_" + name.upper())
        if v is not None:
            _check(lib, lib.akb_set_option(name.encode(), v.encode()))


def set_option(name, value):
    """akb_set_option: a tuning switch of the library (csrc/options.h)."""
    lib = abi.load()
    _check(lib, lib.akb_set_option(str(name).encode(), str(value).encode()))


def init(device_index=0):
    """akb_init_device: bind this process to one GPU (the one-process-per-GPU arrangement)."""
    global _initialised
    lib = abi.load()
    _forward_env(lib)
    if _initialised != ("device", device_index):
        _check(lib, lib.akb_init_device(device_index))
        _initialised = ("device", device_index)
    return lib


def init_multi(n_devices=0):
    """akb_init(n_devices): bind this process to the first n GPUs (0 = all); scenes are replicated on all of them and
    akb_render_params.n_gpus splits a frame over them."""
    global _initialised
    lib = abi.load()
    _forward_env(lib)
    if _initialised != ("multi", n_devices):
        _check(lib, lib.akb_init(n_devices))
        _initialised = ("multi", n_devices)
    return lib


def _check(lib, rc):
    if rc != abi.OK:
        raise AkariError(f"akb status {rc}: {lib.akb_last_error().decode(errors='replace')}")


class Backend:
    """A committed scene on the device: the replacement for `scene.accel = BVHAccelerator; scene.commit()` +
    `gpu::Integrator::render` (core/nodes/scene.cpp:119-162)."""

    def __init__(self, scene, device_index=0, bvh_build=0, n_devices=None):
        """bvh_build: abi.BVH_DEFAULT / BVH_HOST_SAH / BVH_DEVICE (include/akari_b200.h akb_bvh_build).
        n_devices: None = one process per GPU on `device_index`; a number = this process drives that many GPUs
        (0 = all), the scene is replicated on each and render(params with n_gpus) splits the frame."""
        self.lib = init(device_index) if n_devices is None else init_multi(n_devices)
        self.scene = scene
        self.width, self.height = scene.camera.resolution
        desc, keep = scene.flatten()
        if bvh_build:
            desc.bvh_build = bvh_build
        h = C.c_void_p()
        _check(self.lib, self.lib.akb_scene_create(C.byref(desc), C.byref(h)))
        self.h = h
        del keep  # the library deep-copies

    def close(self):
        if getattr(self, "h", None):
            self.lib.akb_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        keys = ["n_triangles", "n_bvh_nodes", "bvh_bytes", "n_lights", "triangle_bytes", "bvh_depth", "build_us", "bvh_build",
                "n_devices", "upload_us"]
        return {k: int(self.lib.akb_scene_info(self.h, i)) for i, k in enumerate(keys)}

    # ---- Scene::intersect / occlude over host ray batches
    def trace(self, org_tmin, dir_tmax, any_hit=False, out=None):
        """`out`: a caller-owned result array (HIT_DTYPE, or int32 for any_hit) to fill instead of a new one."""
        org = np.ascontiguousarray(org_tmin, np.float32)
        dr = np.ascontiguousarray(dir_tmax, np.float32)
        n = org.shape[0]
        hits = out if (out is not None and not any_hit) else np.zeros(0 if any_hit else n, HIT_DTYPE)
        occ = out if (out is not None and any_hit) else np.zeros(n if any_hit else 0, np.int32)
        assert (occ if any_hit else hits).shape[0] == n
        _check(self.lib, self.lib.akb_trace(self.h, _ptr(org), _ptr(dr), n, int(any_hit), _ptr(hits), _ptr(occ)))
        return occ if any_hit else hits

    def trace_device(self, d_org, d_dir, n, any_hit, d_hits, d_occ, stream=0):
        """Device pointers (ints). Returns the traversal kernel's device time in ms."""
        ms = C.c_float(0)
        _check(self.lib, self.lib.akb_trace_device(self.h, d_org, d_dir, n, int(any_hit), d_hits, d_occ, stream,
                                                   C.byref(ms)))
        return ms.value

    def generate_rays(self, xy, u2):
        xy = np.ascontiguousarray(xy, np.int32)
        u2 = np.ascontiguousarray(u2, np.float32)
        n = xy.shape[0]
        org = np.zeros((n, 4), np.float32)
        dr = np.zeros((n, 4), np.float32)
        _check(self.lib, self.lib.akb_generate_rays(self.h, _ptr(xy), _ptr(u2), n, _ptr(org), _ptr(dr)))
        return org, dr

    # ---- Integrator::render
    def render(self, params=None, out=None):
        """Host film (H, W, 4) float32 = (sum R, sum G, sum B, sum weight). Returns (film, stats)."""
        p = params if params is not None else self.scene.integrator.params()
        film = out if out is not None else np.zeros((self.height, self.width, 4), np.float32)
        st = abi.RenderStats()
        _check(self.lib, self.lib.akb_render(self.h, C.byref(p), _ptr(film), C.byref(st)))
        return film, st

    def render_device(self, d_film, params=None, stream=0):
        p = params if params is not None else self.scene.integrator.params()
        st = abi.RenderStats()
        _check(self.lib, self.lib.akb_render_device(self.h, C.byref(p), d_film, stream, C.byref(st)))
        return st

    def film_to_srgb8(self, film):
        h, w = film.shape[:2]
        out = np.zeros((h, w, 3), np.uint8)
        _check(self.lib, self.lib.akb_film_to_srgb8(_ptr(np.ascontiguousarray(film, np.float32)), w, h, _ptr(out)))
        return out

    def film_to_srgb8_noalloc(self, film):
        """akb_film_to_srgb8_noalloc: same bytes, static device scratch, no allocation (include/akari_b200.h)."""
        h, w = film.shape[:2]
        out = np.zeros((h, w, 3), np.uint8)
        _check(self.lib, self.lib.akb_film_to_srgb8_noalloc(_ptr(np.ascontiguousarray(film, np.float32)), w, h, _ptr(out)))
        return out
