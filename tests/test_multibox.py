"""Several boxes per rank (np_project with nlocal > 1): the MultiFab layout incflo really hands to
amrex::NodalProjector (amr.max_grid_size chops the domain; /root/reference/src/projection/
incflo_apply_nodal_projection.cpp:148-153 passes whatever BoxArray the level has).

The boxes are gathered into the rank's slab, so the result must be BIT-IDENTICAL to the one-box call on the same
global arrays, and with that equal to the oracle's (tests/test_project.py pins the one-box call to it).
"""
import ctypes as C

import numpy as np
import pytest

import nphelp
from incflo_b200 import capi, problems
from incflo_b200.projector import Fab, Geometry, LPInfo, NodalProjector
from oracle.oracle import chop_boxes as _chop


def chop_boxes(ncell, max_grid):
    return [((b[0], b[1], b[2]), (b[3], b[4], b[5])) for b in _chop(ncell, max_grid)]


def _box_fabs(arr, boxes, ng, like, device, ghost_src=None):
    """Cut a global [comp,k,j,i] array into per-box Fabs with ng ghosts.  ghost_src: global array WITH one ghost
    layer whose out-of-domain cells are copied into the matching ghost cells of the box fabs (inflow)."""
    fabs = []
    nz, ny, nx = arr.shape[1:]
    for lo, hi in boxes:
        sub = arr[:, lo[2]:hi[2] + 1, lo[1]:hi[1] + 1, lo[0]:hi[0] + 1]
        f = Fab.from_array(sub, ng, False, "numpy")
        if ghost_src is not None:
            # junk in ghost cells that lie over other boxes (AMReX leaves stale data there until FillBoundary) ...
            f.data[...] = 777.0
            f.valid()[...] = sub
            # ... and the physical ghost cells (outside the domain) from ghost_src
            for kk in range(lo[2] - 1, hi[2] + 2):
                for jj in range(lo[1] - 1, hi[1] + 2):
                    for ii in range(lo[0] - 1, hi[0] + 2):
                        if not (0 <= ii < nx and 0 <= jj < ny and 0 <= kk < nz):
                            f.data[:, kk - lo[2] + ng, jj - lo[1] + ng, ii - lo[0] + ng] = ghost_src[:, kk + 1, jj + 1, ii + 1]
        if like != "numpy":
            import torch
            f = Fab(torch.from_numpy(f.data).to(device), ng, False)
        fabs.append(f)
    return fabs


def _run_boxes(ncell, prob_lo, prob_hi, bclo, bchi, vel, sigma, boxes, backend, ghost_src=None, vel_ng=2, like=None):
    like = like or backend["like"]
    device = backend.get("device", "cuda")
    geom = Geometry(ncell, prob_lo, prob_hi, [b == 0 for b in bclo])
    fv = _box_fabs(vel, boxes, vel_ng, like, device, ghost_src)
    fs = _box_fabs(sigma[None], boxes, 0, like, device)
    P = NodalProjector(fv, fs, geom, LPInfo(), boxes=boxes, owner=[0] * len(boxes), lib=backend["lib"])
    P.setDomainBC(bclo, bchi)
    st = P.project(1e-11, 1e-14)
    phis, gps = P.getPhi(), P.getGradPhi()
    u = np.zeros_like(vel)
    phi = np.zeros((ncell[2] + 1, ncell[1] + 1, ncell[0] + 1))
    gp = np.zeros_like(vel)
    for b, (lo, hi) in enumerate(boxes):
        u[:, lo[2]:hi[2] + 1, lo[1]:hi[1] + 1, lo[0]:hi[0] + 1] = fv[b].numpy_valid()
        gp[:, lo[2]:hi[2] + 1, lo[1]:hi[1] + 1, lo[0]:hi[0] + 1] = gps[b].numpy_valid()
        pb = phis[b].numpy_valid()[0]
        cur = phi[lo[2]:hi[2] + 2, lo[1]:hi[1] + 2, lo[0]:hi[0] + 2]
        # nodes shared by two boxes must carry the same value in both (nodalSync)
        seen = cur != 0
        assert np.array_equal(cur[seen], pb[seen])
        phi[lo[2]:hi[2] + 2, lo[1]:hi[1] + 2, lo[0]:hi[0] + 2] = pb
    # ghost layers of the per-box results = the neighbouring boxes' values (what FillBoundary leaves)
    lo, hi = boxes[0]
    g0 = gps[0].data
    g0 = g0.cpu().numpy() if hasattr(g0, "cpu") else g0
    if hi[0] + 1 < ncell[0]:
        assert np.array_equal(g0[:, 1:-1, 1:-1, -1], gp[:, lo[2]:hi[2] + 1, lo[1]:hi[1] + 1, hi[0] + 1])
    P.close()
    return st, u, phi, gp


@pytest.mark.parametrize("name,ncell,maxgrid", [("taylor_green", (32, 32, 32), (16, 16, 16)),
                                                ("rayleigh_taylor", (16, 16, 32), (8, 16, 8)),
                                                ("double_shear", (32, 32, 16), (32, 8, 16))])
def test_multibox_equals_onebox(backend, name, ncell, maxgrid):
    kw = dict(zwave=True, prob_hi=(1.0, 1.0, 0.5)) if name == "double_shear" else {}
    prob = problems.PROBLEMS[name](ncell, **kw)
    drv = nphelp.Driver(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"], **backend)
    st1, u1, phi1, gp1 = drv.project(prob["vel"], prob["sigma"], 1e-11, 1e-14)
    boxes = chop_boxes(ncell, maxgrid)
    assert len(boxes) > 1
    st, u, phi, gp = _run_boxes(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"], prob["vel"],
                                prob["sigma"], boxes, backend)
    assert st.status == 0 and st.n_vcycles == st1.n_vcycles
    assert np.array_equal(u, u1) and np.array_equal(phi, phi1) and np.array_equal(gp, gp1)


@pytest.mark.parametrize("host", [False, pytest.param(True, marks=pytest.mark.gpu)])
def test_multibox_inflow_ghosts(backend, host):
    """Inflow ghost cells of the boxes that touch the x-lo face must reach the divergence; ghost cells over other boxes
    (stale in AMReX until FillBoundary) must not.  host=True on the GPU: host fabs in, staged per box."""
    if host and backend["like"] == "numpy":
        pytest.skip("host fabs are a case of the CUDA build only (the emulation has one kind of memory)")
    rng = np.random.default_rng(5)
    ncell = (32, 16, 8)
    bclo, bchi = (2, 2, 0), (1, 2, 0)
    sigma = rng.uniform(0.5, 2.0, ncell[::-1])
    vel = 0.1 * rng.standard_normal((3,) + ncell[::-1])
    vel[0] += 1.0
    y = (np.arange(16) + 0.5) / 16
    uin = 6 * y * (1 - y)
    gsrc = np.zeros((3, 10, 18, 34))
    gsrc[0, 1:-1, 1:-1, 0] = uin[None, :]

    def ghosts(fv):
        nphelp.assign(fv.data[0, 1:-1, 1:-1, 0], np.broadcast_to(uin[None, :], (8, 16)))

    drv = nphelp.Driver(ncell, (0, 0, 0), (2.0, 1.0, 0.5), bclo, bchi, **backend)
    st1, u1, phi1, gp1 = drv.project(vel, sigma, 1e-11, 1e-14, ghost_setter=ghosts)
    boxes = chop_boxes(ncell, (8, 8, 8))
    st, u, phi, gp = _run_boxes(ncell, (0, 0, 0), (2.0, 1.0, 0.5), bclo, bchi, vel, sigma, boxes, backend, ghost_src=gsrc,
                                vel_ng=1, like="numpy" if host else None)
    assert st.status == 0 and st.n_vcycles == st1.n_vcycles
    assert np.array_equal(u, u1) and np.array_equal(phi, phi1) and np.array_equal(gp, gp1)


def test_multibox_layout_checks(backend):
    """Layouts np_create must refuse: overlapping boxes, holes, a box outside its owner's slab."""
    L = backend["lib"]
    geom = Geometry((16, 16, 16))

    def create(boxes, owner, nranks=1):
        bx = (capi.np_box * len(boxes))()
        for b, (lo, hi) in enumerate(boxes):
            for d in range(3):
                bx[b].lo[d], bx[b].hi[d] = lo[d], hi[d]
        h = C.c_void_p()
        g = geom.c()
        opt = capi.default_options(L)
        rc = L.np_create(C.byref(h), C.byref(g), len(boxes), bx, (C.c_int * len(boxes))(*owner), 0, nranks, None,
                         C.byref(opt))
        if rc == 0:
            L.np_destroy(h)
        return rc

    good = chop_boxes((16, 16, 16), (8, 8, 8))
    assert create(good, [0] * 8) == 0
    assert create(good[:-1], [0] * 7) == 1                                   # hole
    bad = list(good)
    bad[0] = ((0, 0, 0), (8, 7, 7))                                          # overlaps its x neighbour
    bad[1] = ((9, 0, 0), (15, 7, 7))
    assert create(bad, [0] * 8) == 0                                         # still a tiling (9 + 7 cells)
    bad[1] = ((8, 0, 0), (15, 7, 7))
    assert create(bad, [0] * 8) == 1                                         # overlap
