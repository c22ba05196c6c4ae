"""Host-side mirror of amrex::MacProjector as incflo drives it, over the C ABI of include/macproj.h.

Call order of the reference (/root/reference/src/convection/incflo_compute_MAC_projected_velocities.cpp):

    macproj = MacProjector(geom)                                  # incflo.cpp: macproj.reset(new MacProjector(Geom(0,finest_level)))
    macproj.initProjector(lp_info, inv_rho)   or   macproj.updateBeta(inv_rho)           # :39-46
    macproj.setDomainBC(bclo, bchi)                                                       # :44
    macproj.setUMAC(mac_vec)                                                              # :115
    macproj.project(rtol, atol)   or   macproj.project(mac_phi, rtol, atol)               # :127-131
    macproj.getFluxes(fluxes, mac_phi)        # macproj->getLinOp().getFluxes(...)          :63

Face Fabs (one box, FArrayBox layout): direction d has shape (1, nz+[d==2]+2g, ny+[d==1]+2g, nx+[d==0]+2g).
"""
import ctypes as C

from . import capi
from .projector import Fab, LPInfo


def face_fab_zeros(ncell, d, ngrow=0, like="numpy", device="cuda"):
    n = [int(x) + 2 * ngrow for x in ncell]
    n[d] += 1
    return Fab(Fab._alloc((1, n[2], n[1], n[0]), like, device), ngrow, nodal=False)


def face_fab_from_array(arr, ngrow=0, like="numpy", device="cuda"):
    """arr: [k, j, i] face values of one direction (no ghosts)."""
    import numpy as np
    a = np.asarray(arr, dtype=np.float64)
    nz, ny, nx = a.shape
    f = Fab(Fab._alloc((1, nz + 2 * ngrow, ny + 2 * ngrow, nx + 2 * ngrow), like, device), ngrow, nodal=False)
    v = f.valid()
    if like == "numpy":
        v[...] = a[None]
    else:
        import torch
        v.copy_(torch.from_numpy(np.ascontiguousarray(a[None])))
    return f


class MacProjector:
    """Drop-in for amrex::MacProjector on one AMR level, one rank.

    boxes: list of (lo, hi) cell boxes of the level's BoxArray (default: one box covering the domain).  With several
    boxes every per-direction argument (beta[d], umac[d], fluxes[d]) and mac_phi is a LIST of Fabs in BoxArray order."""

    def __init__(self, geom, lib=None, boxes=None, rank=0, nranks=1, nccl_comm=None):
        """nranks > 1 (one GPU each): `geom` is the whole domain, the level is cut into nranks equal z-slabs, this rank
        holds slab `rank` (its fabs have ncell[2] / nranks planes), nccl_comm: void* from capi np_nccl_comm_create."""
        self.L = lib or capi.load()
        self.geom = geom
        self.rank, self.nranks, self.nccl_comm = rank, nranks, nccl_comm
        if nranks > 1:
            nzl = geom.ncell[2] // nranks
            self.all_boxes = [((0, 0, r * nzl), (geom.ncell[0] - 1, geom.ncell[1] - 1, (r + 1) * nzl - 1)) for r in range(nranks)]
            self.owner = list(range(nranks))
            self.boxes = [self.all_boxes[rank]]
        else:
            self.boxes = boxes or [((0, 0, 0), tuple(n - 1 for n in geom.ncell))]
            self.all_boxes, self.owner = self.boxes, [0] * len(self.boxes)
        self.h = None
        self.umac = None
        self.stats = None
        self._bc = None

    def needInitialization(self):
        return self.h is None

    def initProjector(self, lp_info, beta, **options):
        info = lp_info or LPInfo()
        options.setdefault("max_coarsening_level", min(info.max_coarsening_level, 100))
        self.opt = capi.mac_default_options(self.L, **options)
        nb = len(self.all_boxes)
        bx = (capi.np_box * nb)()
        for b, (lo, hi) in enumerate(self.all_boxes):
            for d in range(3):
                bx[b].lo[d], bx[b].hi[d] = int(lo[d]), int(hi[d])
        owner = (C.c_int * nb)(*self.owner)
        h = C.c_void_p()
        g = self.geom.c()
        capi.check(self.L, self.L.mp_create(C.byref(h), C.byref(g), nb, bx, owner, self.rank, self.nranks,
                                            C.c_void_p(self.nccl_comm) if self.nccl_comm else None, C.byref(self.opt)))
        self.h = h
        if self._bc:
            self.setDomainBC(*self._bc)
        self.updateBeta(beta)

    def close(self):
        if getattr(self, "h", None):
            self.L.mp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def setDomainBC(self, lo, hi):
        self._bc = (tuple(lo), tuple(hi))
        if self.h is not None:
            capi.check(self.L, self.L.mp_set_bc(self.h, (C.c_int * 3)(*[int(x) for x in lo]), (C.c_int * 3)(*[int(x) for x in hi])))

    def _arr(self, fabs):
        return (capi.np_fab * len(fabs))(*[f.c() for f in fabs])

    def updateBeta(self, beta):
        if len(self.boxes) > 1:
            capi.check(self.L, self.L.mp_update_beta_boxes(self.h, len(self.boxes), *[self._arr(beta[d]) for d in range(3)]))
            return
        b = (capi.np_fab * 3)(*[f.c() for f in beta])
        capi.check(self.L, self.L.mp_update_beta(self.h, b))

    def setUMAC(self, umac):
        self.umac = list(umac)

    def project(self, *args, stream=None, raise_on_fail=True):
        """project(rtol, atol) or project(mac_phi, rtol, atol) - the two overloads the reference uses."""
        mac_phi = None
        if len(args) == 3:
            mac_phi, rtol, atol = args
        else:
            rtol, atol = args
        st = capi.np_stats()
        if len(self.boxes) > 1:
            rc = self.L.mp_project_boxes(self.h, len(self.boxes), *[self._arr(self.umac[d]) for d in range(3)],
                                         self._arr(mac_phi) if mac_phi is not None else None, rtol, atol,
                                         C.c_void_p(stream) if stream else None, C.byref(st))
        else:
            u = (capi.np_fab * 3)(*[f.c() for f in self.umac])
            phi = mac_phi.c() if mac_phi is not None else None
            rc = self.L.mp_project(self.h, u, C.byref(phi) if phi is not None else None, rtol, atol,
                                   C.c_void_p(stream) if stream else None, C.byref(st))
        self.stats = st
        if rc and raise_on_fail:
            capi.check(self.L, rc)
        return st

    def getFluxes(self, fluxes, mac_phi):
        if len(self.boxes) > 1:
            capi.check(self.L, self.L.mp_get_fluxes_boxes(self.h, len(self.boxes), *[self._arr(fluxes[d]) for d in range(3)],
                                                          self._arr(mac_phi)))
            return
        f = (capi.np_fab * 3)(*[x.c() for x in fluxes])
        p = mac_phi.c()
        capi.check(self.L, self.L.mp_get_fluxes(self.h, f, C.byref(p)))

    def copyPhi(self, dst):
        d = dst.c()
        capi.check(self.L, self.L.mp_copy_phi(self.h, C.byref(d)))

    @property
    def num_levels(self):
        return self.L.mp_num_levels(self.h)

    def level_ncell(self, l):
        n = (C.c_int * 3)()
        capi.check(self.L, self.L.mp_level_ncell(self.h, l, n))
        return tuple(n)
