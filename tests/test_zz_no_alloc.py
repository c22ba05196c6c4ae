"""np_apply_noalloc (include/nodalproj.h): the operator hook without any device allocation, for callers that cannot call
cudaMalloc.  No counterpart in the reference.  Collected LAST on purpose.
Checked through the C ABI against the existing entry point np_apply on two calls in a
row (the scratch array is shared between calls and keeps the last call's contents), at two MG levels, and for the size
limit.  Tolerance: the kernel-level bar of tests/test_kernels.py (1e-13 relative L-inf); the two are in fact bit-identical."""
import ctypes as C

import numpy as np
import pytest

import nphelp
from incflo_b200 import capi


def _apply(drv, fn, l, x):
    fx, fy = drv.nfab(l, x), drv.nfab(l)
    rc = fn(drv.h, l, C.byref(fx.c()), C.byref(fy.c()))
    return rc, fy.numpy_valid()[0]


@pytest.mark.parametrize("variable_sigma", [False, True])
def test_apply_noalloc_matches_apply_twice_in_a_row(backend, variable_sigma):
    rng = np.random.default_rng(7)
    ncell = (48, 32, 16)
    drv = nphelp.Driver(ncell, (0, 0, 0), (1.5, 1.0, 0.5), (0, 2, 1), (0, 2, 2), **backend)
    L = drv.L
    drv.set_sigma(rng.uniform(0.5, 2.0, ncell[::-1]) if variable_sigma else np.full(ncell[::-1], 0.75))
    for call, l in enumerate((0, 0, 1, 0)):                 # back to back; level 1 in between leaves other data in the scratch
        n = drv.P.level_ncell(l)
        x = rng.standard_normal((n[2] + 1, n[1] + 1, n[0] + 1))
        rc0, y0 = _apply(drv, L.np_apply, l, x)
        rc1, y1 = _apply(drv, L.np_apply_noalloc, l, x)
        assert rc0 == 0 and rc1 == 0, (call, rc0, rc1)
        assert np.abs(y0).max() > 0
        assert np.abs(y1 - y0).max() <= 1e-13 * np.abs(y0).max(), call


def test_apply_noalloc_rejects_what_it_cannot_hold(backend):
    ncell = (96, 96, 96)                                    # 100 x 99 x 99 doubles per array: two do not fit 2^20
    drv = nphelp.Driver(ncell, **backend)
    L = drv.L
    drv.set_sigma(np.ones(ncell[::-1]))
    rc, _ = _apply(drv, L.np_apply_noalloc, 0, np.zeros((97, 97, 97)))
    assert rc == capi.NP_ERR_TOO_LARGE
    assert b"too large" in L.np_strerror(rc)
    rc, _ = _apply(drv, L.np_apply_noalloc, 1, np.ones((49, 49, 49)))      # the next level fits
    assert rc == 0
    if backend["like"] != "numpy":                          # CUDA build: host fabs would need a staging allocation
        hx = nphelp.Fab.zeros(drv.P.level_ncell(1), 1, 1, True, "numpy", None)
        hy = nphelp.Fab.zeros(drv.P.level_ncell(1), 1, 1, True, "numpy", None)
        assert L.np_apply_noalloc(drv.h, 1, C.byref(hx.c()), C.byref(hy.c())) == capi.NP_ERR_INVALID
