"""Host-side mirror of the reference's implicit scalar diffusion solve over the C ABI of include/macproj.h.

    DiffusionScalarOp::diffuse_scalar(tracer, density, eta, dt)      /root/reference/src/diffusion/DiffusionScalarOp.cpp:139-260
        (rho - dt div(eta grad)) tracer_new = rho tracer_old,   one component at a time
        m_reg_scal_solve_op->setScalars(1.0, dt)            :179        -> mp_set_scalars
        setACoeffs(lev, density)                            :181        -> mp_set_acoeffs
        setBCoeffs(lev, average_scalar_eta_to_faces(..))    :196-197    -> mp_update_beta (the face averaging stays with the caller,
                                                                           incflo::average_scalar_eta_to_faces, incflo_diffusion.cpp:262-284)
        rhs = rho * tracer                                  :207-217    -> folded into mp_solve (rho_or_null)
        setLevelBC(lev, &phi)                               :227        -> the ghost cells of the tracer fab (Dirichlet values)
        mlmg.solve(phi, rhs, m_mg_rtol, m_mg_atol)          :250        -> mp_solve
    DiffusionScalarOp::diffuse_vel_components (:262-380) is the same solve per velocity component with eta = mu.
Domain BCs: incflo::get_diffuse_scalar_bc (incflo_diffusion.cpp:204-233): Dirichlet at mass inflow, Neumann at every other
non-periodic face.  Solver options: the "scalar_diffusion.*" ParmParse keys (mg_rtol 1e-11, mg_atol 1e-14, mg_max_iter 100,
bottom solver bicgstab: DiffusionScalarOp.cpp readParameters)."""
import ctypes as C

import numpy as np

from . import capi
from .mac_projector import MacProjector, face_fab_from_array
from .projector import LPInfo


def average_cellcenter_to_face(eta_g):
    """amrex::average_cellcenter_to_face + incflo::fixup_eta_on_domain_faces for one box: eta_g is [k, j, i] WITH one ghost
    layer (periodic images or the boundary values); returns (bx, by, bz) face arrays without ghosts.  Interior faces: mean
    of the two cells; faces on a non-periodic domain boundary: the ghost value (fixup: the boundary value itself) - pass
    ghosts equal to the first interior cell for a homogeneous wall."""
    e = np.asarray(eta_g, dtype=np.float64)
    c = e[1:-1, 1:-1, 1:-1]
    bx = 0.5 * (e[1:-1, 1:-1, :-1] + e[1:-1, 1:-1, 1:])
    by = 0.5 * (e[1:-1, :-1, 1:-1] + e[1:-1, 1:, 1:-1])
    bz = 0.5 * (e[:-1, 1:-1, 1:-1] + e[1:, 1:-1, 1:-1])
    assert bx.shape[2] == c.shape[2] + 1 and by.shape[1] == c.shape[1] + 1 and bz.shape[0] == c.shape[0] + 1
    return np.ascontiguousarray(bx), np.ascontiguousarray(by), np.ascontiguousarray(bz)


class DiffusionScalarOp:
    """One AMR level, one box, one rank.  diffuse_scalar(tracer, density, eta_faces, dt) updates every component of the tracer fab
    in place; the tracer fab carries >= 1 ghost layer whose cells next to Dirichlet (mass inflow) faces hold the inflow values."""

    def __init__(self, geom, bclo, bchi, lib=None, like="numpy", device="cuda", **options):
        self.L = lib or capi.load()
        self.geom, self.like, self.device = geom, like, device
        options.setdefault("maxiter", 100)
        options.setdefault("bottom_maxiter", 100)
        options.setdefault("bottom_solver", 1)          # bicgstab ("The default bottom solver is BiCG", DiffusionScalarOp.cpp:234)
        self.options = options
        self.P = MacProjector(geom, lib=self.L)
        self.P.setDomainBC(bclo, bchi)
        self.stats = []

    def close(self):
        self.P.close()

    def _init(self, beta):
        if self.P.needInitialization():
            self.P.initProjector(LPInfo().setMaxCoarseningLevel(100), beta, **self.options)
        else:
            self.P.updateBeta(beta)

    def diffuse_scalar(self, tracer, density, eta_faces, dt, rtol=1e-11, atol=1e-14, stream=None, raise_on_fail=True):
        """tracer: cell Fab, ncomp components, >= 1 ghost; density: cell Fab (1 comp); eta_faces: per component a triple of face
        Fabs (b = eta on faces).  Returns the list of np_stats (one per component)."""
        from .projector import Fab
        L, P = self.L, self.P
        ncomp = tracer.data.shape[0]
        self.stats = []
        for comp in range(ncomp):
            self._init(eta_faces[comp])
            capi.check(L, L.mp_set_scalars(P.h, 1.0, float(dt)))
            a = density.c()
            capi.check(L, L.mp_set_acoeffs(P.h, C.byref(a)))
            phi = Fab(tracer.data[comp:comp + 1], tracer.ngrow)         # alias of one component (amrex::make_alias, :203)
            p, r, rho = phi.c(), phi.c(), density.c()
            st = capi.np_stats()
            rc = L.mp_solve(P.h, C.byref(p), C.byref(r), C.byref(rho), rtol, atol, C.c_void_p(stream) if stream else None, C.byref(st))
            self.stats.append(st)
            if rc and raise_on_fail:
                capi.check(L, rc)
        return self.stats

    def solve(self, phi, rhs, alpha, beta_scalar, a, b_faces, rtol=1e-11, atol=1e-14, raise_on_fail=True):
        """The bare MLMG::solve on (alpha a - beta div(b grad)) phi = rhs (tests / kernels hooks)."""
        L, P = self.L, self.P
        self._init(b_faces)
        capi.check(L, L.mp_set_scalars(P.h, float(alpha), float(beta_scalar)))
        if a is not None:
            ac = a.c()
            capi.check(L, L.mp_set_acoeffs(P.h, C.byref(ac)))
        p, r = phi.c(), rhs.c()
        st = capi.np_stats()
        rc = L.mp_solve(P.h, C.byref(p), C.byref(r), None, rtol, atol, None, C.byref(st))
        if rc and raise_on_fail:
            capi.check(L, rc)
        return st
