"""ORACLE — TEST INFRASTRUCTURE ONLY (parity unpinned upstream; see oracle/README.md).

ctypes binding of oracle/_build/liboracle.so, the CPU restatement of the reference's nodal
projection (amrex::NodalProjector / MLMG / MLNodeLaplacian as called from
/root/reference/src/projection/incflo_apply_nodal_projection.cpp:148-166).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
Array conventions: numpy C-order arrays indexed [k, j, i] (x fastest); velocity [3, k, j, i].
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "liboracle.so")

PERIODIC, DIRICHLET, NEUMANN = 0, 1, 2
LEX, TILED = 0, 1
BICGCG, BICGSTAB, CG, SMOOTHER = 0, 1, 2, 3


class Options(C.Structure):
    _fields_ = [("max_coarsening_level", C.c_int), ("maxiter", C.c_int), ("bottom_maxiter", C.c_int),
                ("nu1", C.c_int), ("nu2", C.c_int), ("nuf", C.c_int), ("nub", C.c_int),
                ("smooth_sweeps", C.c_int), ("bottom_solver", C.c_int),
                ("bottom_rtol", C.c_double), ("bottom_atol", C.c_double),
                ("verbose", C.c_int), ("smoother", C.c_int), ("tile", C.c_int * 3),
                ("interp_sigma_weighted", C.c_int), ("bottom_normalize", C.c_int),
                ("agg_level", C.c_int), ("nthreads", C.c_int), ("tile_variant", C.c_int), ("tile_overlap", C.c_int), ("tile_z_target", C.c_int),
                ("tile_owned_only", C.c_int)]


class Stats(C.Structure):
    _fields_ = [("n_vcycles", C.c_int), ("bottom_iters_total", C.c_int), ("status", C.c_int), ("n_levels", C.c_int),
                ("rhs_norm0", C.c_double), ("res_norm0", C.c_double), ("res_norm_final", C.c_double),
                ("res_history", C.c_double * 128),
                ("t_setup", C.c_double), ("t_solve", C.c_double), ("t_total", C.c_double)]


def build(force=False):
    if force or not os.path.exists(_LIB) or any(
            os.path.getmtime(os.path.join(_HERE, f)) > os.path.getmtime(_LIB)
            for f in ("nodal_oracle.cpp", "stencil.hpp")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        # On the GPU box the prebuilt .so travels with the snapshot; rebuild only if missing/stale.
        build()
        _lib = C.CDLL(_LIB)
        _lib.orc_create.restype = C.c_void_p
        _lib.orc_dot.restype = C.c_double
    return _lib


def _p(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def chop_boxes(ncell, max_grid):
    """Cut the domain into boxes of at most max_grid cells per side (AMReX max_grid_size)."""
    mg = [max_grid] * 3 if np.isscalar(max_grid) else list(max_grid)
    boxes = []
    for k0 in range(0, ncell[2], mg[2]):
        for j0 in range(0, ncell[1], mg[1]):
            for i0 in range(0, ncell[0], mg[0]):
                boxes.append([i0, j0, k0, min(i0 + mg[0], ncell[0]) - 1, min(j0 + mg[1], ncell[1]) - 1,
                              min(k0 + mg[2], ncell[2]) - 1])
    return boxes


class Oracle:
    def __init__(self, ncell, prob_lo=(0, 0, 0), prob_hi=(1, 1, 1), bclo=(0, 0, 0), bchi=(0, 0, 0),
                 boxes=None, **opts):
        L = lib()
        self.ncell = tuple(int(n) for n in ncell)
        o = Options()
        L.orc_default_options(C.byref(o))
        for k, v in opts.items():
            if k == "tile":
                o.tile[0], o.tile[1], o.tile[2] = [int(t) for t in v]
            else:
                if not hasattr(o, k):
                    raise KeyError(k)
                setattr(o, k, v)
        self.opts = o
        if boxes is None:
            boxes = [[0, 0, 0, self.ncell[0] - 1, self.ncell[1] - 1, self.ncell[2] - 1]]
        self.boxes = boxes
        bx = np.asarray(boxes, dtype=np.int32).reshape(-1, 6)
        a3i = lambda v: (C.c_int * 3)(*[int(x) for x in v])
        a3d = lambda v: (C.c_double * 3)(*[float(x) for x in v])
        self.h = C.c_void_p(L.orc_create(a3i(ncell), a3d(prob_lo), a3d(prob_hi), a3i(bclo), a3i(bchi),
                                         bx.shape[0], bx.ctypes.data_as(C.POINTER(C.c_int)), C.byref(o)))
        self.L = L

    def close(self):
        if self.h:
            self.L.orc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- shapes --
    @property
    def num_levels(self):
        return self.L.orc_num_levels(self.h)

    def level_ncell(self, l):
        n = (C.c_int * 3)()
        self.L.orc_level_ncell(self.h, l, n)
        return tuple(n)

    def nshape(self, l=0):
        n = self.level_ncell(l)
        return (n[2] + 1, n[1] + 1, n[0] + 1)

    def cshape(self, l=0):
        n = self.level_ncell(l)
        return (n[2], n[1], n[0])

    # -- full projection --
    def set_vel(self, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        assert u.shape == (3,) + self.cshape(0)
        self.L.orc_set_vel(self.h, _p(u))

    def set_vel_ghost(self, i, j, k, n, v):
        self.L.orc_set_vel_ghost(self.h, int(i), int(j), int(k), int(n), C.c_double(v))

    def set_sigma(self, s):
        s = np.ascontiguousarray(s, dtype=np.float64)
        assert s.shape == self.cshape(0)
        self.L.orc_set_sigma(self.h, _p(s))

    def prepare_sigma(self):
        self.L.orc_prepare_sigma(self.h)

    def get_sigma(self, l):
        s = np.empty(self.cshape(l))
        self.L.orc_get_sigma(self.h, l, _p(s))
        return s

    def project(self, rtol=1e-11, atol=1e-14):
        st = Stats()
        self.L.orc_project(self.h, C.c_double(rtol), C.c_double(atol), C.byref(st))
        return st

    def get_vel(self):
        u = np.empty((3,) + self.cshape(0))
        self.L.orc_get_vel(self.h, _p(u))
        return u

    def get_gradphi(self):
        g = np.empty((3,) + self.cshape(0))
        self.L.orc_get_gradphi(self.h, _p(g))
        return g

    def get_phi(self):
        p = np.empty(self.nshape(0))
        self.L.orc_get_phi(self.h, _p(p))
        return p

    def get_rhs(self):
        p = np.empty(self.nshape(0))
        self.L.orc_get_rhs(self.h, _p(p))
        return p

    # -- kernel-level hooks --
    def divu(self):
        r = np.empty(self.nshape(0))
        self.L.orc_divu(self.h, _p(r))
        return r

    def apply(self, l, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.zeros(self.nshape(l))
        self.L.orc_apply(self.h, l, _p(x), _p(y))
        return y

    def smooth(self, l, x, rhs, ncalls=1):
        x = np.array(x, dtype=np.float64, order="C")
        rhs = np.ascontiguousarray(rhs, dtype=np.float64)
        self.L.orc_smooth(self.h, l, _p(x), _p(rhs), ncalls)
        return x

    def restrict(self, lc, fine):
        fine = np.ascontiguousarray(fine, dtype=np.float64)
        c = np.zeros(self.nshape(lc))
        self.L.orc_restrict(self.h, lc, _p(fine), _p(c))
        return c

    def interp_add(self, lf, fine, crse):
        fine = np.array(fine, dtype=np.float64, order="C")
        crse = np.ascontiguousarray(crse, dtype=np.float64)
        self.L.orc_interp_add(self.h, lf, _p(fine), _p(crse))
        return fine

    def mknewu(self, phi, u):
        phi = np.ascontiguousarray(phi, dtype=np.float64)
        u = np.array(u, dtype=np.float64, order="C")
        g = np.empty_like(u)
        self.L.orc_mknewu(self.h, _p(phi), _p(u), _p(g))
        return u, g

    def dot(self, l, x, y):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.float64)
        return self.L.orc_dot(self.h, l, _p(x), _p(y))

    def bottom_solve(self, b):
        l = self.num_levels - 1
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros(self.nshape(l))
        it = C.c_int(0)
        self.L.orc_bottom_solve(self.h, _p(b), _p(x), C.byref(it))
        return x, it.value

    def vcycle(self, res0):
        res0 = np.ascontiguousarray(res0, dtype=np.float64)
        cor = np.zeros(self.nshape(0))
        self.L.orc_vcycle(self.h, _p(res0), _p(cor))
        return cor
