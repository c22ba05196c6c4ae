"""Small smoother / operator cases for `compute-sanitizer --tool racecheck|memcheck python tools/sanitize_case.py`:
the nodal two-sweep smoother (constant and variable sigma, a partial tile in x and y: 70 x 36 x 8 cells, Neumann + periodic),
the nodal operator, and the MAC one-pass red-black smoother on 128 x 32 x 8 cells (whole tiles), through the kernel hooks."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from incflo_b200 import capi                                                        # noqa: E402
from incflo_b200.mac_projector import MacProjector, face_fab_from_array             # noqa: E402
from incflo_b200.projector import Fab, Geometry, LPInfo, NodalProjector             # noqa: E402

L = capi.load()
dev = "cuda:0"
torch.cuda.set_device(0)
rng = np.random.default_rng(0)
ncell = (70, 36, 8)
for var in (False, True):
    vel = Fab.zeros(ncell, 3, 1, False, "torch", dev)
    sig = Fab.from_array(rng.uniform(0.5, 2.0, ncell[::-1]) if var else np.ones(ncell[::-1]), 0, False, "torch", dev)
    P = NodalProjector(vel, sig, Geometry(ncell, (0, 0, 0), (1.0, 0.5, 0.1), [False, True, True]), LPInfo(), lib=L)
    P.setDomainBC((2, 0, 0), (2, 0, 0))
    capi.check(L, L.np_set_sigma(P.h, C.byref(sig.c())))
    x = Fab.from_array(rng.standard_normal((9, 37, 71)), 1, True, "torch", dev)
    r = Fab.from_array(rng.standard_normal((9, 37, 71)), 1, True, "torch", dev)
    y = Fab.zeros(ncell, 1, 1, True, "torch", dev)
    capi.check(L, L.np_smooth(P.h, 0, C.byref(x.c()), C.byref(r.c()), 2))
    capi.check(L, L.np_apply(P.h, 0, C.byref(x.c()), C.byref(y.c())))
    print("nodal", "var" if var else "const", float(x.data.abs().max()), float(y.data.abs().max()))
    P.close()
L.np_cache_clear()
mc = (128, 32, 8)
from oracle.mac import face_shapes                                                  # noqa: E402  (shapes only)
for const in (True, False):
    beta = [np.full(s, 0.7) if const else rng.uniform(0.5, 2.0, s) for s in face_shapes(mc)]
    for d, ax in ((0, 2), (1, 1), (2, 0)):
        sl_hi = [slice(None)] * 3; sl_lo = [slice(None)] * 3; sl_hi[ax], sl_lo[ax] = -1, 0
        beta[d][tuple(sl_hi)] = beta[d][tuple(sl_lo)]
    M = MacProjector(Geometry(mc, (0, 0, 0), (4.0, 1.0, 0.25), [True] * 3), lib=L)
    M.setDomainBC((0, 0, 0), (0, 0, 0))
    M.initProjector(LPInfo().setMaxCoarseningLevel(100), [face_fab_from_array(b, 0, "torch", dev) for b in beta], fused_rb_min_n=2, fused_rb_const=1)
    fx = Fab.from_array(rng.standard_normal(mc[::-1]), 0, False, "torch", dev)
    fb = Fab.from_array(rng.standard_normal(mc[::-1]), 0, False, "torch", dev)
    capi.check(L, L.mp_smooth(M.h, 0, C.byref(fx.c()), C.byref(fb.c()), 2))
    print("mac", "const" if const else "var", float(fx.data.abs().max()))
    M.close()
print("sanitize_case: done")
