"""ORACLE — TEST INFRASTRUCTURE ONLY (parity unpinned upstream; see oracle/README.md).

ctypes binding of oracle/_build/libmacoracle.so, the CPU restatement of the reference's MAC projection
(amrex::MacProjector / MLMG / MLABecLaplacian as called from
/root/reference/src/convection/incflo_compute_MAC_projected_velocities.cpp:36-46,115-131).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
Arrays: numpy C order [k, j, i]; face arrays bx/u (nz, ny, nx+1), by/v (nz, ny+1, nx), bz/w (nz+1, ny, nx).
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libmacoracle.so")
PERIODIC, DIRICHLET, NEUMANN = 0, 1, 2


class Options(C.Structure):
    _fields_ = [("max_coarsening_level", C.c_int), ("maxiter", C.c_int), ("bottom_maxiter", C.c_int), ("nu1", C.c_int),
                ("nu2", C.c_int), ("nuf", C.c_int), ("nub", C.c_int), ("bottom_solver", C.c_int), ("verbose", C.c_int),
                ("maxorder", C.c_int), ("nthreads", C.c_int), ("bottom_rtol", C.c_double), ("bottom_atol", C.c_double)]


class Stats(C.Structure):
    _fields_ = [("n_vcycles", C.c_int), ("bottom_iters_total", C.c_int), ("status", C.c_int), ("n_levels", C.c_int),
                ("rhs_norm0", C.c_double), ("res_norm0", C.c_double), ("res_norm_final", C.c_double),
                ("res_history", C.c_double * 128), ("t_setup", C.c_double), ("t_solve", C.c_double), ("t_total", C.c_double)]


def build():
    src = os.path.join(_HERE, "mac_oracle.cpp")
    if not os.path.exists(_LIB) or os.path.getmtime(src) > os.path.getmtime(_LIB):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
        _lib.mor_create.restype = C.c_void_p
    return _lib


def _p(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def face_shapes(n):
    nx, ny, nz = n
    return (nz, ny, nx + 1), (nz, ny + 1, nx), (nz + 1, ny, nx)


class MacOracle:
    def __init__(self, ncell, prob_lo=(0, 0, 0), prob_hi=(1, 1, 1), bclo=(0, 0, 0), bchi=(0, 0, 0), **opts):
        L = lib()
        o = Options()
        L.mor_default_options(C.byref(o))
        for k, v in opts.items():
            setattr(o, k, v)
        self.L, self.ncell = L, tuple(ncell)
        self.h = C.c_void_p(L.mor_create((C.c_int * 3)(*ncell), (C.c_double * 3)(*prob_lo), (C.c_double * 3)(*prob_hi),
                                         (C.c_int * 3)(*bclo), (C.c_int * 3)(*bchi), C.byref(o)))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.mor_destroy(self.h)
            self.h = None

    @property
    def num_levels(self):
        return self.L.mor_num_levels(self.h)

    def level_ncell(self, l):
        n = (C.c_int * 3)()
        self.L.mor_level_ncell(self.h, l, n)
        return tuple(n)

    def _cell(self, l):
        n = self.level_ncell(l)
        return (n[2], n[1], n[0])

    def set_beta(self, bx, by, bz):
        a = [np.ascontiguousarray(x, dtype=np.float64) for x in (bx, by, bz)]
        assert tuple(x.shape for x in a) == face_shapes(self.ncell)
        self.L.mor_set_beta(self.h, _p(a[0]), _p(a[1]), _p(a[2]))

    def get_beta(self, l):
        out = [np.zeros(s) for s in face_shapes(self.level_ncell(l))]
        self.L.mor_get_beta(self.h, l, _p(out[0]), _p(out[1]), _p(out[2]))
        return out

    def set_umac(self, u, v, w):
        a = [np.ascontiguousarray(x, dtype=np.float64) for x in (u, v, w)]
        assert tuple(x.shape for x in a) == face_shapes(self.ncell)
        self.L.mor_set_umac(self.h, _p(a[0]), _p(a[1]), _p(a[2]))

    def get_umac(self):
        out = [np.zeros(s) for s in face_shapes(self.ncell)]
        self.L.mor_get_umac(self.h, _p(out[0]), _p(out[1]), _p(out[2]))
        return out

    def set_phi(self, phi):
        a = np.ascontiguousarray(phi, dtype=np.float64)
        self.L.mor_set_phi(self.h, _p(a))

    def get_phi(self):
        out = np.zeros(self._cell(0))
        self.L.mor_get_phi(self.h, _p(out))
        return out

    def project(self, rtol, atol, have_guess=False):
        st = Stats()
        self.L.mor_project(self.h, C.c_double(rtol), C.c_double(atol), int(have_guess), C.byref(st))
        return st

    def apply(self, l, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.zeros_like(x)
        self.L.mor_apply(self.h, l, _p(x), _p(y))
        return y

    def smooth(self, l, x, b, ncalls=1):
        x = np.array(x, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        self.L.mor_smooth(self.h, l, _p(x), _p(b), ncalls)
        return x

    def restrict(self, lc, fine):
        f = np.ascontiguousarray(fine, dtype=np.float64)
        c = np.zeros(self._cell(lc))
        self.L.mor_restrict(self.h, lc, _p(f), _p(c))
        return c

    def interp_add(self, lf, fine, crse):
        f = np.array(fine, dtype=np.float64)
        c = np.ascontiguousarray(crse, dtype=np.float64)
        self.L.mor_interp_add(self.h, lf, _p(f), _p(c))
        return f

    def divergence(self):
        d = np.zeros(self._cell(0))
        self.L.mor_divergence(self.h, _p(d))
        return d

    def bottom_solve(self, b):
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros_like(b)
        it = self.L.mor_bottom_solve(self.h, _p(b), _p(x))
        return x, it

    # ---- general ABecLaplacian solve (scalar diffusion; SURVEY.md 8(f) row 3 groundwork, CPU only) ----
    def set_scalars(self, alpha, beta):
        self.L.mor_set_scalars(self.h, C.c_double(alpha), C.c_double(beta))

    def set_acoef(self, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        self.L.mor_set_acoef(self.h, _p(a))

    def solve_abec(self, rhs, rtol, atol, bcvals=None):
        """(alpha a - beta div b grad) phi = rhs; phi starts from set_phi() (or 0); bcvals: dict face -> 2-D array of Dirichlet
        values, faces 'xlo','ylo','zlo','xhi','yhi','zhi', arrays indexed [higher transverse dir][lower transverse dir]."""
        r = np.ascontiguousarray(rhs, dtype=np.float64)
        self.L.mor_set_rhs(self.h, _p(r))
        keep, ptrs = [], []
        for f in ("xlo", "ylo", "zlo", "xhi", "yhi", "zhi"):
            v = (bcvals or {}).get(f)
            if v is None:
                ptrs.append(None)
            else:
                a = np.ascontiguousarray(v, dtype=np.float64)
                keep.append(a)
                ptrs.append(_p(a))
        st = Stats()
        self.L.mor_solve_abec(self.h, C.c_double(rtol), C.c_double(atol), *ptrs, C.byref(st))
        return st
