"""Test helpers: drive the C ABI (CUDA build on the GPU box, emulation build on CPU) with global numpy
arrays, so kernel-level results can be compared 1:1 with the oracle's."""
import ctypes as C
import os
import subprocess
import numpy as np

from incflo_b200 import capi
from incflo_b200.projector import Fab, Geometry, LPInfo, NodalProjector

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU_LIB = os.path.join(ROOT, "tests", "_emu", "libnodalproj_emu.so")


def emu_lib():
    """TEST-ONLY emulation build of the product sources (tests/emu): kernels run on the host."""
    dirs = [os.path.join(ROOT, "incflo_b200", "csrc"), os.path.join(ROOT, "tests", "emu"), os.path.join(ROOT, "include")]
    srcs = [os.path.join(d, f) for d in dirs for f in os.listdir(d)]
    if not os.path.exists(EMU_LIB) or any(os.path.getmtime(s) > os.path.getmtime(EMU_LIB) for s in srcs if os.path.isfile(s)):
        subprocess.check_call([os.path.join(ROOT, "tests", "emu", "build_emu.sh")])
    return capi.load(EMU_LIB)


class Driver:
    """Wraps a NodalProjector handle and exposes the kernel hooks on global arrays [k,j,i]."""

    def __init__(self, ncell, prob_lo=(0, 0, 0), prob_hi=(1, 1, 1), bclo=(0, 0, 0), bchi=(0, 0, 0), lib=None,
                 like="numpy", device="cuda", **opts):
        self.like, self.device = like, device
        self.ncell = tuple(ncell)
        geom = Geometry(ncell, prob_lo, prob_hi, [b == 0 for b in bclo])
        self.vel = Fab.zeros(ncell, 3, 1, False, like, device)
        self.sig = Fab.zeros(ncell, 1, 0, False, like, device)
        self.P = NodalProjector(self.vel, self.sig, geom, LPInfo(), lib=lib, **opts)
        self.P.setDomainBC(bclo, bchi)
        self.L, self.h = self.P.L, self.P.h

    def nfab(self, l, arr=None):
        n = self.P.level_ncell(l)
        f = Fab.zeros(n, 1, 1, True, self.like, self.device)
        if arr is not None:
            a = np.ascontiguousarray(arr, dtype=np.float64)[None]
            if self.like == "numpy":
                f.valid()[...] = a
            else:
                import torch
                f.valid().copy_(torch.from_numpy(a))
        return f

    def cfab(self, arr, ng):
        return Fab.from_array(arr, ng, False, self.like, self.device)

    def set_sigma(self, sigma):
        f = self.cfab(sigma, 0)
        capi.check(self.L, self.L.np_set_sigma(self.h, C.byref(f.c())))

    def get_sigma(self, l):
        n = self.P.level_ncell(l)
        f = Fab.zeros(n, 1, 0, False, self.like, self.device)
        capi.check(self.L, self.L.np_get_sigma(self.h, l, C.byref(f.c())))
        return f.numpy_valid()[0]

    def apply(self, l, x):
        fx, fy = self.nfab(l, x), self.nfab(l)
        capi.check(self.L, self.L.np_apply(self.h, l, C.byref(fx.c()), C.byref(fy.c())))
        return fy.numpy_valid()[0]

    def smooth(self, l, x, rhs, ncalls=1):
        fx, fr = self.nfab(l, x), self.nfab(l, rhs)
        capi.check(self.L, self.L.np_smooth(self.h, l, C.byref(fx.c()), C.byref(fr.c()), ncalls))
        return fx.numpy_valid()[0]

    def restrict(self, lc, fine):
        ff, fc = self.nfab(lc - 1, fine), self.nfab(lc)
        capi.check(self.L, self.L.np_restrict(self.h, lc, C.byref(ff.c()), C.byref(fc.c())))
        return fc.numpy_valid()[0]

    def resid_restrict(self, l, cor, res):
        fc, fr, fo = self.nfab(l, cor), self.nfab(l, res), self.nfab(l + 1)
        capi.check(self.L, self.L.np_resid_restrict(self.h, l, C.byref(fc.c()), C.byref(fr.c()), C.byref(fo.c())))
        return fo.numpy_valid()[0]

    def interp_add(self, lf, fine, crse):
        ff, fc = self.nfab(lf, fine), self.nfab(lf + 1, crse)
        capi.check(self.L, self.L.np_interp_add(self.h, lf, C.byref(ff.c()), C.byref(fc.c())))
        return ff.numpy_valid()[0]

    def divu(self, vel, ghost_setter=None):
        fv = self.cfab(vel, 1)
        if ghost_setter:
            ghost_setter(fv)
        fr = self.nfab(0)
        capi.check(self.L, self.L.np_divu(self.h, C.byref(fv.c()), C.byref(fr.c())))
        return fr.numpy_valid()[0]

    def mknewu(self, phi, vel):
        fp, fv = self.nfab(0, phi), self.cfab(vel, 1)
        fg = Fab.zeros(self.ncell, 3, 1, False, self.like, self.device)
        capi.check(self.L, self.L.np_mknewu(self.h, C.byref(fp.c()), C.byref(fv.c()), C.byref(fg.c())))
        return fv.numpy_valid(), fg.numpy_valid()

    def bottom_solve(self, b):
        l = self.P.num_levels - 1
        fb, fx = self.nfab(l, b), self.nfab(l)
        it = C.c_int(0)
        capi.check(self.L, self.L.np_bottom_solve(self.h, C.byref(fb.c()), C.byref(fx.c()), C.byref(it)))
        return fx.numpy_valid()[0], it.value

    def vcycle(self, res0):
        fr, fc = self.nfab(0, res0), self.nfab(0)
        capi.check(self.L, self.L.np_vcycle(self.h, C.byref(fr.c()), C.byref(fc.c())))
        return fc.numpy_valid()[0]

    def dot(self, l, x, y):
        fx, fy = self.nfab(l, x), self.nfab(l, y)
        out = C.c_double(0)
        capi.check(self.L, self.L.np_dot(self.h, l, C.byref(fx.c()), C.byref(fy.c()), C.byref(out)))
        return out.value

    def project(self, vel, sigma, rtol=1e-11, atol=1e-14, ghost_setter=None, raise_on_fail=True):
        """Full projection on global arrays; returns (stats, vel, phi, gradphi)."""
        fv, fs = self.cfab(vel, 1), self.cfab(sigma, 0)
        if ghost_setter:
            ghost_setter(fv)
        self.P.vel, self.P.sigma = fv, fs
        st = self.P.project(rtol, atol, raise_on_fail=raise_on_fail)
        phi = self.P.getPhi().numpy_valid()[0]
        gp = self.P.getGradPhi().numpy_valid()
        return st, fv.numpy_valid(), phi, gp


def assign(dst, arr):
    """dst[...] = arr for a numpy view or a torch (CPU / CUDA) view."""
    if hasattr(dst, "copy_"):
        import torch
        dst.copy_(torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float64)))
    else:
        dst[...] = arr


def periodic_consistent(x, bclo):
    """Make a full nodal array consistent on periodic images (high face = low face)."""
    x = np.array(x)
    for d, ax in ((0, 2), (1, 1), (2, 0)):
        if bclo[d] == 0:
            sl_hi = [slice(None)] * 3
            sl_lo = [slice(None)] * 3
            sl_hi[ax], sl_lo[ax] = -1, 0
            x[tuple(sl_hi)] = x[tuple(sl_lo)]
    return x


def zero_dirichlet(x, bclo, bchi):
    x = np.array(x)
    for d, ax in ((0, 2), (1, 1), (2, 0)):
        sl = [slice(None)] * 3
        if bclo[d] == 1:
            sl[ax] = 0
            x[tuple(sl)] = 0
        if bchi[d] == 1:
            sl[ax] = -1
            x[tuple(sl)] = 0
    return x
