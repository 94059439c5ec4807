"""akb_film_to_srgb8_noalloc (include/akari_b200.h): the byte stage with its device staging in a static device array, for
callers that may not allocate. It has no counterpart in the reference; the check is against the existing entry point
akb_film_to_srgb8 through the C ABI - byte work, so bit-exact like the other byte-stage tests - on two calls in a row
(the second call finds the first call's contents in the shared array). Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

from common import abi, pkg, scenes

pytestmark = pytest.mark.gpu


def film_like(h, w, seed):
    """Sums-and-weight film: radiance from dark to over-range, some negative, some zero-weight pixels."""
    rng = np.random.default_rng(seed)
    f = np.empty((h, w, 4), np.float32)
    wgt = rng.integers(0, 5, size=(h, w)).astype(np.float32)
    f[..., :3] = (rng.random((h, w, 3), dtype=np.float32) ** 3 * 1.5 - 0.05) * np.maximum(wgt, 1)[..., None]
    f[..., 3] = wgt
    return f


def test_noalloc_byte_stage_equals_allocating_entry_point():
    be = pkg.Backend(scenes.cornell_box((8, 8), 1, 1), 0)
    first = film_like(1080, 1920, 1)  # the largest film of the BASELINE configs that fits the static scratch
    second = film_like(37, 53, 2)     # then a small ragged one: most of the array still holds the first call's data
    for film in (first, second, first):
        want = be.film_to_srgb8(film)
        got = be.film_to_srgb8_noalloc(film)
        assert got.shape == want.shape and np.array_equal(got, want)
        assert want.any() and (want != want.flat[0]).any()  # a real image, not a constant
    be.close()


def test_noalloc_rejects_films_beyond_the_static_scratch():
    be = pkg.Backend(scenes.cornell_box((8, 8), 1, 1), 0)
    lib = be.lib
    max_px = 1 << 21  # AKB_NOALLOC_MAX_PIXELS
    film = np.zeros((1, max_px + 1, 4), np.float32)
    out = np.zeros((1, max_px + 1, 3), np.uint8)
    rc = lib.akb_film_to_srgb8_noalloc(film.ctypes.data_as(C.c_void_p), max_px + 1, 1, out.ctypes.data_as(C.c_void_p))
    assert rc == abi.INVALID_ARG
    assert b"AKB_NOALLOC_MAX_PIXELS" in lib.akb_last_error()
    # 4K (C4's resolution) is beyond it too; the exact maximum is accepted
    assert lib.akb_film_to_srgb8_noalloc(film.ctypes.data_as(C.c_void_p), 3840, 2160, out.ctypes.data_as(C.c_void_p)) == abi.INVALID_ARG
    ok = film_like(1, max_px, 3)
    assert np.array_equal(be.film_to_srgb8_noalloc(ok), be.film_to_srgb8(ok))
    be.close()
