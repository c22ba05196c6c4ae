"""Host-side mirror of the reference's projector interface, over the C ABI.

Names and call order follow what incflo::ApplyProjection does with amrex::NodalProjector
(/root/reference/src/projection/incflo_apply_nodal_projection.cpp:148-166):

    info = LPInfo(); info.setMaxCoarseningLevel(n)
    proj = NodalProjector(vel, sigma, geom, info)
    proj.setDomainBC(bclo, bchi)
    proj.project(rtol, atol)
    phi, gradphi = proj.getPhi(), proj.getGradPhi()

`vel` / `sigma` are Fab objects: one box of a MultiFab in AMReX's FArrayBox layout, i.e. an array of
shape (ncomp, nz+2g, ny+2g, nx+2g) in C order (x fastest, component slowest).  The storage is a torch
tensor (CUDA: zero-copy; CPU/pinned: staged inside np_project) or a numpy array.
"""
import ctypes as C
import numpy as np

from . import capi


class Fab:
    """One box of a MultiFab.  nodal=True -> valid range has n+1 points per direction."""

    def __init__(self, data, ngrow, nodal=False):
        self.data = data
        self.ngrow = int(ngrow)
        self.nodal = bool(nodal)
        shp = tuple(data.shape)
        assert len(shp) == 4, "Fab storage is (ncomp, nz+2g, ny+2g, nx+2g)"
        self.ncomp = shp[0]
        if hasattr(data, "is_contiguous"):
            assert data.is_contiguous() and str(data.dtype) == "torch.float64"
        else:
            assert data.flags["C_CONTIGUOUS"] and data.dtype == np.float64

    @property
    def ptr(self):
        d = self.data
        return d.data_ptr() if hasattr(d, "data_ptr") else d.ctypes.data

    def c(self):
        return capi.np_fab(C.c_void_p(self.ptr), self.ngrow, self.ncomp)

    def valid(self):
        """View of the valid region (no ghosts): (ncomp, nz, ny, nx)."""
        g = self.ngrow
        if g == 0:
            return self.data
        return self.data[:, g:-g, g:-g, g:-g]

    # ---- construction helpers -------------------------------------------------------------
    @staticmethod
    def _alloc(shape, like, device):
        if like == "numpy":
            return np.zeros(shape, dtype=np.float64)
        import torch
        if like == "pinned":
            return torch.zeros(shape, dtype=torch.float64).pin_memory()
        return torch.zeros(shape, dtype=torch.float64, device=device)

    @classmethod
    def zeros(cls, ncell, ncomp, ngrow, nodal=False, like="numpy", device="cuda"):
        n = [int(x) + (1 if nodal else 0) + 2 * ngrow for x in ncell]
        return cls(cls._alloc((ncomp, n[2], n[1], n[0]), like, device), ngrow, nodal)

    @classmethod
    def from_array(cls, arr, ngrow, nodal=False, like="numpy", device="cuda"):
        """arr: (ncomp, nz, ny, nx) or (nz, ny, nx) valid data (numpy); ghosts are zero."""
        a = np.asarray(arr, dtype=np.float64)
        if a.ndim == 3:
            a = a[None]
        ncomp, nz, ny, nx = a.shape
        off = 1 if nodal else 0
        f = cls.zeros((nx - off, ny - off, nz - off), ncomp, ngrow, nodal, like, device)
        if like == "numpy":
            f.valid()[...] = a
        else:
            import torch
            f.valid().copy_(torch.from_numpy(np.ascontiguousarray(a)))
        return f

    def numpy_valid(self):
        v = self.valid()
        if hasattr(v, "cpu"):
            return v.cpu().numpy().copy()
        return np.array(v)


class Geometry:
    """amrex::Geometry subset: domain cells, physical box, periodicity."""

    def __init__(self, ncell, prob_lo=(0, 0, 0), prob_hi=(1, 1, 1), periodic=(1, 1, 1)):
        self.ncell = tuple(int(n) for n in ncell)
        self.prob_lo = tuple(float(x) for x in prob_lo)
        self.prob_hi = tuple(float(x) for x in prob_hi)
        self.periodic = tuple(int(bool(p)) for p in periodic)

    def isPeriodic(self, d):
        return bool(self.periodic[d])

    def c(self):
        g = capi.np_geom()
        for d in range(3):
            g.ncell[d] = self.ncell[d]
            g.prob_lo[d] = self.prob_lo[d]
            g.prob_hi[d] = self.prob_hi[d]
            g.periodic[d] = self.periodic[d]
        return g


class LPInfo:
    """amrex::LPInfo subset used by the reference (…nodal_projection.cpp:148-149)."""

    def __init__(self):
        self.max_coarsening_level = 30

    def setMaxCoarseningLevel(self, n):
        self.max_coarsening_level = int(n)
        return self


class NodalProjector:
    """Drop-in for amrex::NodalProjector on one AMR level.

    boxes: list of (lo, hi) cell boxes of the level-0 BoxArray, default one box covering the domain; owner: the
    DistributionMapping (rank of every box), default box b -> rank b.  The boxes of rank r must tile the r-th of
    nranks equal z-slabs.  With several boxes on this rank `vel` / `sigma` are lists of Fabs in BoxArray order (what
    MFIter visits) and getPhi / getGradPhi return lists.  nccl_comm: void* from capi np_nccl_comm_create when nranks > 1.
    Extra keyword options are the `nodal_proj.*` ParmParse keys / implementation knobs of np_options.
    """

    def __init__(self, vel, sigma, geom, info=None, boxes=None, rank=0, nranks=1, nccl_comm=None, lib=None,
                 owner=None, **options):
        self.L = lib or capi.load()
        self.vel, self.sigma, self.geom = vel, sigma, geom
        info = info or LPInfo()
        self.opt = capi.default_options(self.L, max_coarsening_level=info.max_coarsening_level, **options)
        if boxes is None:
            boxes = [((0, 0, 0), tuple(n - 1 for n in geom.ncell))]
        bx = (capi.np_box * len(boxes))()
        for b, (lo, hi) in enumerate(boxes):
            for d in range(3):
                bx[b].lo[d], bx[b].hi[d] = int(lo[d]), int(hi[d])
        owner = list(range(len(boxes))) if owner is None else [int(r) for r in owner]
        self.local_boxes = [(tuple(lo), tuple(hi)) for (lo, hi), r in zip(boxes, owner) if r == rank]
        owner = (C.c_int * len(boxes))(*owner)
        self.h = C.c_void_p()
        g = geom.c()
        capi.check(self.L, self.L.np_create(C.byref(self.h), C.byref(g), len(boxes), bx, owner, rank, nranks,
                                            C.c_void_p(nccl_comm) if nccl_comm else None, C.byref(self.opt)))
        self.rank, self.nranks = rank, nranks
        self.nloc = (geom.ncell[0], geom.ncell[1], geom.ncell[2] // nranks)
        self.stats = None

    def close(self):
        if getattr(self, "h", None):
            self.L.np_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- the reference's five calls ----
    def setDomainBC(self, lo, hi):
        a = (C.c_int * 3)(*[int(x) for x in lo])
        b = (C.c_int * 3)(*[int(x) for x in hi])
        capi.check(self.L, self.L.np_set_bc(self.h, a, b))

    def project(self, rtol, atol, stream=None, raise_on_fail=True):
        st = capi.np_stats()
        if isinstance(self.vel, (list, tuple)):
            nloc = len(self.vel)
            v = (capi.np_fab * nloc)(*[f.c() for f in self.vel])
            s = (capi.np_fab * nloc)(*[f.c() for f in self.sigma])
            rc = self.L.np_project(self.h, nloc, v, s, rtol, atol, C.c_void_p(stream) if stream else None, C.byref(st))
        else:
            v, s = self.vel.c(), self.sigma.c()
            rc = self.L.np_project(self.h, 1, C.byref(v), C.byref(s), rtol, atol, C.c_void_p(stream) if stream else None,
                                   C.byref(st))
        self.stats = st
        if rc and raise_on_fail:
            capi.check(self.L, rc)      # reference behaviour: amrex::Abort("MLMG failed")
        return st

    def _borrow(self, getter, ncell, ncomp, nodal, box=0):
        f = capi.np_fab()
        capi.check(self.L, getter(self.h, box, C.byref(f)))
        n = [x + (1 if nodal else 0) + 2 * f.ngrow for x in ncell]
        shape = (ncomp, n[2], n[1], n[0])
        if not hasattr(self.L, "np_emu_set_comm"):
            # the CUDA library hands back DEVICE pointers whatever the inputs were: zero-copy alias of the
            # projector-owned buffer (valid until the next project()/close())
            import torch
            v0 = self.vel[0] if isinstance(self.vel, (list, tuple)) else self.vel
            dev = v0.data.device if getattr(v0.data, "is_cuda", False) else torch.device("cuda", torch.cuda.current_device())
            return Fab(_alias_cuda(f.p, shape, dev), f.ngrow, nodal)
        arr = np.ctypeslib.as_array(C.cast(f.p, C.POINTER(C.c_double)), shape=shape)
        return Fab(arr, f.ngrow, nodal)

    def _results(self, getter, ncomp, nodal):
        if not isinstance(self.vel, (list, tuple)):
            return self._borrow(getter, self.nloc, ncomp, nodal)
        return [self._borrow(getter, tuple(h - l + 1 for l, h in zip(lo, hi)), ncomp, nodal, b)
                for b, (lo, hi) in enumerate(self.local_boxes)]

    def getPhi(self):
        return self._results(self.L.np_get_phi, 1, True)

    def getGradPhi(self):
        return self._results(self.L.np_get_gradphi, 3, False)

    # ---- introspection ----
    @property
    def num_levels(self):
        return self.L.np_num_levels(self.h)

    def level_ncell(self, l):
        n = (C.c_int * 3)()
        capi.check(self.L, self.L.np_level_ncell(self.h, l, n))
        return tuple(n)


def _alias_cuda(ptr, shape, device):
    """torch tensor aliasing device memory owned by the library (through __cuda_array_interface__)."""
    import torch

    class _Holder:
        pass

    h = _Holder()
    h.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f8", "data": (int(ptr), False), "version": 2,
                                  "strides": None}
    return torch.as_tensor(h, device=device)
