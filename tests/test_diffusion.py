"""Implicit scalar diffusion solve (DiffusionScalarOp::diffuse_scalar,
/root/reference/src/diffusion/DiffusionScalarOp.cpp:139-260) through the C ABI of include/macproj.h (mp_set_scalars,
mp_set_acoeffs, mp_update_beta, mp_solve) against the oracle's general MLABecLaplacian solve (oracle/mac_oracle.cpp:
solve_abec, pinned by tests/test_mac_oracle_pins.py: quadratic reproduced exactly, sparse direct solve).  Red-black
Gauss-Seidel does not depend on the order inside a colour, so the CUDA path follows the reference's algorithm exactly:
same V-cycle and bottom-iteration counts, fields to 1e-10 relative L2 (the north-star tolerance), operator and smoother
hooks to 1e-13."""
import ctypes as C

import numpy as np
import pytest

import nphelp  # noqa: F401  (emu build)
from incflo_b200 import capi
from incflo_b200.diffusion import DiffusionScalarOp, average_cellcenter_to_face
from incflo_b200.mac_projector import face_fab_from_array
from incflo_b200.projector import Fab, Geometry
from oracle.mac import MacOracle, face_shapes

FACES = ("xlo", "ylo", "zlo", "xhi", "yhi", "zhi")


def l2rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def random_beta(rng, ncell, bclo, const=False):
    b = [np.full(s, 0.7) if const else rng.uniform(0.5, 2.0, s) for s in face_shapes(ncell)]
    if bclo[0] == 0: b[0][:, :, -1] = b[0][:, :, 0]
    if bclo[1] == 0: b[1][:, -1, :] = b[1][:, 0, :]
    if bclo[2] == 0: b[2][-1, :, :] = b[2][0, :, :]
    return b


def ghost_face_values(phi_g, bclo, bchi):
    """the wall values setLevelBC reads: the first ghost layer of a [k, j, i] array with one ghost layer, per Dirichlet face"""
    v = {}
    if bclo[0] == 1: v["xlo"] = phi_g[1:-1, 1:-1, 0]
    if bchi[0] == 1: v["xhi"] = phi_g[1:-1, 1:-1, -1]
    if bclo[1] == 1: v["ylo"] = phi_g[1:-1, 0, 1:-1]
    if bchi[1] == 1: v["yhi"] = phi_g[1:-1, -1, 1:-1]
    if bclo[2] == 1: v["zlo"] = phi_g[0, 1:-1, 1:-1]
    if bchi[2] == 1: v["zhi"] = phi_g[-1, 1:-1, 1:-1]
    return v


BCS = [((0, 0, 0), (0, 0, 0)), ((1, 2, 0), (2, 2, 0)), ((2, 1, 2), (1, 2, 1))]


@pytest.mark.parametrize("bclo,bchi", BCS)
@pytest.mark.parametrize("const_b", [False, True])
def test_abec_operator_and_smoother_hooks(backend, bclo, bchi, const_b):
    """A x and two smooths of (alpha a - beta div b grad) on two MG levels against the oracle."""
    rng = np.random.default_rng(3)
    ncell, hi = (16, 12, 8), (1.0, 0.6, 0.5)
    like, dev = backend["like"], backend.get("device", "cuda")
    beta = random_beta(rng, ncell, bclo, const_b)
    a = rng.uniform(0.5, 2.0, ncell[::-1])
    alpha, bscal = 1.3, 0.02
    op = DiffusionScalarOp(Geometry(ncell, (0, 0, 0), hi, [b == 0 for b in bclo]), bclo, bchi, lib=backend["lib"], like=like, device=dev)
    op._init([face_fab_from_array(b, 0, like, dev) for b in beta])
    L, h = op.L, op.P.h
    capi.check(L, L.mp_set_scalars(h, alpha, bscal))
    fa = Fab.from_array(a, 0, False, like, dev)              # (kept alive: the C struct only borrows the pointer)
    capi.check(L, L.mp_set_acoeffs(h, C.byref(fa.c())))
    o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
    o.set_scalars(alpha, bscal); o.set_acoef(a); o.set_beta(*beta)      # set_beta averages a and b down the hierarchy
    for l in (0, 1):
        n = op.P.level_ncell(l)
        x, b = rng.standard_normal(n[::-1]), rng.standard_normal(n[::-1])
        fx, fy = Fab.from_array(x, 0, False, like, dev), Fab.zeros(n, 1, 0, False, like, dev)
        capi.check(L, L.mp_apply(h, l, C.byref(fx.c()), C.byref(fy.c())))
        yo = o.apply(l, x)
        assert np.abs(fy.numpy_valid()[0] - yo).max() <= 1e-13 * np.abs(yo).max()
        fb = Fab.from_array(b, 0, False, like, dev)
        capi.check(L, L.mp_smooth(h, l, C.byref(fx.c()), C.byref(fb.c()), 2))
        so = o.smooth(l, x, b, 2)
        assert np.abs(fx.numpy_valid()[0] - so).max() <= 1e-13 * np.abs(so).max()
    op.close()


@pytest.mark.parametrize("bclo,bchi", BCS)
@pytest.mark.parametrize("const_b", [False, True])
@pytest.mark.parametrize("ncell", [(128, 32, 8), (64, 16, 16)])
def test_abec_one_pass_smoother(backend, bclo, bchi, const_b, ncell):
    """The one-pass two-colour smoother with the alpha a term (k_mac_rb3<., true>: levels made of whole 64 x 16 tiles; forced
    onto these small grids with fused_rb_min_n = 2): smooth hook and a whole solve with Dirichlet data against the oracle."""
    if backend["like"] == "numpy" and (ncell != (64, 16, 16) or (bclo, bchi) == BCS[1]):
        pytest.skip("CPU emulation: the small grid and two BC sets keep the CPU suite short; every case runs on the GPU")
    rng = np.random.default_rng(13)
    hi = (2.0, 0.5, 0.25)
    like, dev = backend["like"], backend.get("device", "cuda")
    beta = random_beta(rng, ncell, bclo, const_b)
    a = rng.uniform(0.5, 2.0, ncell[::-1])
    alpha, bscal = 1.1, 0.03
    op = DiffusionScalarOp(Geometry(ncell, (0, 0, 0), hi, [b == 0 for b in bclo]), bclo, bchi, lib=backend["lib"], like=like, device=dev,
                           bottom_solver=0, fused_rb_min_n=2)
    op._init([face_fab_from_array(b, 0, like, dev) for b in beta])
    L, h = op.L, op.P.h
    capi.check(L, L.mp_set_scalars(h, alpha, bscal))
    fa = Fab.from_array(a, 0, False, like, dev)
    capi.check(L, L.mp_set_acoeffs(h, C.byref(fa.c())))
    o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
    o.set_scalars(alpha, bscal); o.set_acoef(a); o.set_beta(*beta)
    x, b = rng.standard_normal(ncell[::-1]), rng.standard_normal(ncell[::-1])
    fx, fb = Fab.from_array(x, 0, False, like, dev), Fab.from_array(b, 0, False, like, dev)
    capi.check(L, L.mp_smooth(h, 0, C.byref(fx.c()), C.byref(fb.c()), 2))
    so = o.smooth(0, x, b, 2)
    assert np.abs(fx.numpy_valid()[0] - so).max() <= 1e-13 * np.abs(so).max()
    rhs = rng.standard_normal(ncell[::-1])
    phi_g = rng.standard_normal(tuple(n + 2 for n in ncell[::-1]))
    fphi = Fab.zeros(ncell, 1, 1, False, like, dev)
    nphelp.assign(fphi.data, phi_g[None])
    st = op.solve(fphi, Fab.from_array(rhs, 0, False, like, dev), alpha, bscal, fa, [face_fab_from_array(b, 0, like, dev) for b in beta], 1e-11, 1e-14)
    o.set_phi(phi_g[1:-1, 1:-1, 1:-1])
    s2 = o.solve_abec(rhs, 1e-11, 1e-14, ghost_face_values(phi_g, bclo, bchi))
    assert st.status == 0 and s2.status == 0
    assert st.n_vcycles == s2.n_vcycles and st.bottom_iters_total == s2.bottom_iters_total
    assert l2rel(fphi.numpy_valid()[0], o.get_phi()) <= 1e-10
    op.close()


@pytest.mark.parametrize("bclo,bchi", BCS)
def test_abec_solve_matches_oracle(backend, bclo, bchi):
    """mp_solve with variable a and b, an initial guess and inhomogeneous Dirichlet values in the ghost cells."""
    rng = np.random.default_rng(5)
    ncell, hi = (32, 16, 16), (2.0, 1.0, 1.0)
    like, dev = backend["like"], backend.get("device", "cuda")
    beta = random_beta(rng, ncell, bclo)
    a = rng.uniform(0.5, 2.0, ncell[::-1])
    alpha, bscal = 1.0, 0.05
    rhs = rng.standard_normal(ncell[::-1])
    phi_g = rng.standard_normal(tuple(n + 2 for n in ncell[::-1]))        # guess in the valid cells, wall values in the ghosts
    op = DiffusionScalarOp(Geometry(ncell, (0, 0, 0), hi, [b == 0 for b in bclo]), bclo, bchi, lib=backend["lib"], like=like, device=dev,
                           bottom_solver=0)
    fphi = Fab.zeros(ncell, 1, 1, False, like, dev)
    nphelp.assign(fphi.data, phi_g[None])
    st = op.solve(fphi, Fab.from_array(rhs, 0, False, like, dev), alpha, bscal, Fab.from_array(a, 0, False, like, dev),
                  [face_fab_from_array(b, 0, like, dev) for b in beta], 1e-11, 1e-14)
    o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
    o.set_beta(*beta); o.set_scalars(alpha, bscal); o.set_acoef(a); o.set_phi(phi_g[1:-1, 1:-1, 1:-1])
    so = o.solve_abec(rhs, 1e-11, 1e-14, ghost_face_values(phi_g, bclo, bchi))
    assert st.status == 0 and so.status == 0
    assert st.n_vcycles == so.n_vcycles and st.bottom_iters_total == so.bottom_iters_total and st.n_levels == so.n_levels
    for i in range(so.n_vcycles):
        assert abs(st.res_history[i] - so.res_history[i]) <= 1e-6 * so.res_history[i] + 1e-14 * so.rhs_norm0
    assert l2rel(fphi.numpy_valid()[0], o.get_phi()) <= 1e-10
    op.close()


def test_diffuse_scalar_two_tracers(backend):
    """DiffusionScalarOp::diffuse_scalar end to end: two tracer components, rhs = rho * tracer formed inside mp_solve, mass
    inflow (Dirichlet) at x-lo with the inflow values in the ghost cells, Neumann elsewhere in x / y, periodic z; against the
    same sequence of calls on the oracle."""
    rng = np.random.default_rng(9)
    ncell, hi = (32, 16, 8), (2.0, 1.0, 0.5)
    bclo, bchi = (1, 2, 0), (2, 2, 0)
    like, dev = backend["like"], backend.get("device", "cuda")
    dt = 0.01
    rho = rng.uniform(0.5, 2.0, ncell[::-1])
    tr_g = rng.uniform(0.0, 1.0, (2,) + tuple(n + 2 for n in ncell[::-1]))
    tr_g[0][:, :, 0] = 1.0                                  # inflow values of component 0 / 1 in the x-lo ghost layer
    tr_g[1][:, :, 0] = 0.25
    eta = []
    for comp in range(2):
        e = rng.uniform(0.01, 0.05, tuple(n + 2 for n in ncell[::-1]))
        e[0], e[-1] = e[-2], e[1]                           # periodic images in z
        eta.append(average_cellcenter_to_face(e))
    op = DiffusionScalarOp(Geometry(ncell, (0, 0, 0), hi, [b == 0 for b in bclo]), bclo, bchi, lib=backend["lib"], like=like, device=dev,
                           bottom_solver=0)
    ftr = Fab.zeros(ncell, 2, 1, False, like, dev)
    nphelp.assign(ftr.data, tr_g)
    stats = op.diffuse_scalar(ftr, Fab.from_array(rho, 0, False, like, dev),
                              [[face_fab_from_array(b, 0, like, dev) for b in eta[c]] for c in range(2)], dt)
    out = ftr.numpy_valid()
    for comp in range(2):
        o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
        o.set_beta(*eta[comp]); o.set_scalars(1.0, dt); o.set_acoef(rho); o.set_phi(tr_g[comp][1:-1, 1:-1, 1:-1])
        so = o.solve_abec(rho * tr_g[comp][1:-1, 1:-1, 1:-1], 1e-11, 1e-14, ghost_face_values(tr_g[comp], bclo, bchi))
        assert stats[comp].status == 0 and so.status == 0 and stats[comp].n_vcycles == so.n_vcycles
        assert l2rel(out[comp], o.get_phi()) <= 1e-10
        # the implicit step keeps the tracer within the bounds of its data (M-matrix): a sanity check on the physics
        assert out[comp].min() >= -1e-12 and out[comp].max() <= 1.0 + 1e-12
    op.close()


def test_abec_quadratic_is_reproduced(backend):
    """Known-answer pin through the CUDA path itself: phi = x^2 with Dirichlet walls in x is solved exactly by the second-order
    scheme with the maxorder-3 boundary stencil (the same pin the oracle passes in tests/test_mac_oracle_pins.py)."""
    ncell, hi = (16, 8, 8), (1.0, 0.5, 0.5)
    bclo, bchi = (1, 0, 0), (1, 0, 0)
    like, dev = backend["like"], backend.get("device", "cuda")
    alpha, bscal, b0 = 1.0, 0.05, 0.7
    x = (np.arange(ncell[0]) + 0.5) * hi[0] / ncell[0]
    exact = np.broadcast_to(x ** 2, ncell[::-1]).copy()
    a = 1.0 + 0.5 * np.cos(3.0 * exact)
    rhs = alpha * a * exact - bscal * b0 * 2.0
    phi_g = np.zeros(tuple(n + 2 for n in ncell[::-1]))
    phi_g[:, :, 0], phi_g[:, :, -1] = 0.0, hi[0] ** 2      # wall values
    op = DiffusionScalarOp(Geometry(ncell, (0, 0, 0), hi, [b == 0 for b in bclo]), bclo, bchi, lib=backend["lib"], like=like, device=dev)
    fphi = Fab.zeros(ncell, 1, 1, False, like, dev)
    nphelp.assign(fphi.data, phi_g[None])
    st = op.solve(fphi, Fab.from_array(rhs, 0, False, like, dev), alpha, bscal, Fab.from_array(a, 0, False, like, dev),
                  [face_fab_from_array(np.full(s, b0), 0, like, dev) for s in face_shapes(ncell)], 1e-12, 0.0)
    assert st.status == 0
    assert np.abs(fphi.numpy_valid()[0] - exact).max() <= 1e-10
    op.close()


@pytest.mark.gpu
def test_diffusion_iteration_graph_matches_kernel_by_kernel_path():
    """mp_options.use_graph on the diffusion solve (alpha a term, Dirichlet data, initial guess): bitwise equal to the
    kernel-by-kernel path; the second tracer component replays the graph captured for the first one's successor key."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    rng = np.random.default_rng(9)
    ncell, hi = (64, 32, 32), (2.0, 1.0, 1.0)
    bclo, bchi = (1, 2, 0), (2, 2, 0)
    rho = rng.uniform(0.5, 2.0, ncell[::-1])
    tr_g = rng.uniform(0.0, 1.0, (2,) + tuple(n + 2 for n in ncell[::-1]))
    eta = [average_cellcenter_to_face(rng.uniform(0.01, 0.05, tuple(n + 2 for n in ncell[::-1]))) for _ in range(2)]

    def run(use_graph):
        op = DiffusionScalarOp(Geometry(ncell, (0, 0, 0), hi, [b == 0 for b in bclo]), bclo, bchi, lib=capi.load(), like="torch", device="cuda",
                               use_graph=use_graph)
        ftr = Fab.zeros(ncell, 2, 1, False, "torch", "cuda")
        nphelp.assign(ftr.data, tr_g)
        stats = op.diffuse_scalar(ftr, Fab.from_array(rho, 0, False, "torch", "cuda"),
                                  [[face_fab_from_array(b, 0, "torch", "cuda") for b in eta[c]] for c in range(2)], 0.01)
        out = ftr.numpy_valid().copy()
        res = [(s.status, s.n_vcycles, s.bottom_iters_total, s.graph_replays) for s in stats]
        op.close()
        return out, res

    (a, ra), (b, rb) = run(0), run(1)
    assert np.array_equal(a, b)
    for x, y in zip(ra, rb):
        assert x[:3] == y[:3] and x[0] == 0 and x[3] == 0 and y[3] == y[1]
