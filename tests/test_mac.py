"""MAC projection through the C ABI of include/macproj.h against the oracle (oracle/mac_oracle.cpp),
on the CPU emulation build and on the GPU.  Red-black Gauss-Seidel does not depend on the order inside a colour, so the
CUDA path follows the reference's algorithm exactly and iterates agree to round-off: same V-cycle count, residual
history to 1e-6 relative, umac / phi within 1e-10 relative L2 (the north-star tolerance)."""
import ctypes as C

import numpy as np
import pytest

import nphelp  # noqa: F401  (emu build)
from incflo_b200 import capi, problems
from incflo_b200.mac_projector import MacProjector, face_fab_from_array, face_fab_zeros
from incflo_b200.projector import Fab, Geometry, LPInfo
from oracle.mac import MacOracle, face_shapes

BCS = [((0, 0, 0), (0, 0, 0)), ((2, 2, 0), (2, 2, 0)), ((2, 0, 2), (1, 0, 2)), ((1, 2, 2), (1, 2, 1))]


def l2rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def random_beta(rng, ncell, bclo, const=False):
    b = [np.full(s, 0.7) if const else rng.uniform(0.5, 2.0, s) for s in face_shapes(ncell)]
    if bclo[0] == 0: b[0][:, :, -1] = b[0][:, :, 0]
    if bclo[1] == 0: b[1][:, -1, :] = b[1][:, 0, :]
    if bclo[2] == 0: b[2][-1, :, :] = b[2][0, :, :]
    return b


class Drv:
    """MacProjector handle + the kernel hooks on global numpy arrays [k, j, i]."""

    def __init__(self, ncell, prob_hi, bclo, bchi, beta, backend, **opts):
        self.like, self.device = backend["like"], backend.get("device", "cuda")
        self.ncell = tuple(ncell)
        geom = Geometry(ncell, (0, 0, 0), prob_hi, [b == 0 for b in bclo])
        self.P = MacProjector(geom, lib=backend["lib"])
        self.P.setDomainBC(bclo, bchi)
        self.P.initProjector(LPInfo().setMaxCoarseningLevel(100), [self.face(b) for b in beta], **opts)
        self.L, self.h = self.P.L, self.P.h

    def face(self, a, ng=0):
        return face_fab_from_array(a, ng, self.like, self.device)

    def cell(self, l, arr=None, ng=0):
        n = self.P.level_ncell(l)
        if arr is None:
            return Fab.zeros(n, 1, ng, False, self.like, self.device)
        return Fab.from_array(arr, ng, False, self.like, self.device)

    def apply(self, l, x):
        fx, fy = self.cell(l, x), self.cell(l)
        capi.check(self.L, self.L.mp_apply(self.h, l, C.byref(fx.c()), C.byref(fy.c())))
        return fy.numpy_valid()[0]

    def smooth(self, l, x, b, n=1):
        fx, fb = self.cell(l, x), self.cell(l, b)
        capi.check(self.L, self.L.mp_smooth(self.h, l, C.byref(fx.c()), C.byref(fb.c()), n))
        return fx.numpy_valid()[0]

    def restrict(self, lc, fine):
        ff, fc = self.cell(lc - 1, fine), self.cell(lc)
        capi.check(self.L, self.L.mp_restrict(self.h, lc, C.byref(ff.c()), C.byref(fc.c())))
        return fc.numpy_valid()[0]

    def interp_add(self, lf, fine, crse):
        ff, fc = self.cell(lf, fine), self.cell(lf + 1, crse)
        capi.check(self.L, self.L.mp_interp_add(self.h, lf, C.byref(ff.c()), C.byref(fc.c())))
        return ff.numpy_valid()[0]

    def get_beta(self, l):
        n = self.P.level_ncell(l)
        fb = [face_fab_zeros(n, d, 0, self.like, self.device) for d in range(3)]
        arr = (capi.np_fab * 3)(*[f.c() for f in fb])
        capi.check(self.L, self.L.mp_get_beta(self.h, l, arr))
        return [f.numpy_valid()[0] for f in fb]

    def bottom_solve(self, b):
        l = self.P.num_levels - 1
        fb, fx = self.cell(l, b), self.cell(l)
        it = C.c_int(0)
        capi.check(self.L, self.L.mp_bottom_solve(self.h, C.byref(fb.c()), C.byref(fx.c()), C.byref(it)))
        return fx.numpy_valid()[0], it.value

    def project(self, umac, rtol=1e-11, atol=1e-14, phi_guess=None, ng=0, raise_on_fail=True):
        fu = [self.face(u, ng) for u in umac]
        self.P.setUMAC(fu)
        if phi_guess is not None:
            fp = self.cell(0, phi_guess, 1)
            st = self.P.project(fp, rtol, atol, raise_on_fail=raise_on_fail)
            phi = fp.numpy_valid()[0]
        else:
            st = self.P.project(rtol, atol, raise_on_fail=raise_on_fail)
            fp = self.cell(0)
            self.P.copyPhi(fp)
            phi = fp.numpy_valid()[0]
        return st, [f.numpy_valid()[0] for f in fu], phi


@pytest.mark.parametrize("bclo,bchi", BCS)
@pytest.mark.parametrize("const", [False, True])
@pytest.mark.parametrize("fused", [128, 2])
def test_mac_kernels(backend, bclo, bchi, const, fused):
    """fused = 2: every level that qualifies runs both colours of a smooth in one pass (k_mac_rb2) - tiles, z-chunks and
    halos of 2 on tiny levels; 128 (the default): one launch per colour here."""
    rng = np.random.default_rng(7)
    ncell, hi = (16, 8, 12), (1.6, 1.0, 0.9)
    beta = random_beta(rng, ncell, bclo, const)
    drv = Drv(ncell, hi, bclo, bchi, beta, backend, fused_rb_min_n=fused, fused_rb_const=1)
    o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
    o.set_beta(*beta)
    assert drv.P.num_levels == o.num_levels == 2 + 1
    for l in range(o.num_levels):
        n = o.level_ncell(l)
        assert drv.P.level_ncell(l) == n
        for a, b in zip(drv.get_beta(l), o.get_beta(l)):
            assert np.abs(a - b).max() <= 1e-15
        x = rng.standard_normal(n[::-1]); b = rng.standard_normal(n[::-1])
        y, yo = drv.apply(l, x), o.apply(l, x)
        assert np.abs(y - yo).max() <= 1e-13 * np.abs(yo).max()
        s, so = drv.smooth(l, x, b, 2), o.smooth(l, x, b, 2)
        assert np.abs(s - so).max() <= 1e-13 * np.abs(so).max()
        if l > 0:
            nf = o.level_ncell(l - 1)
            f = rng.standard_normal(nf[::-1])
            assert np.abs(drv.restrict(l, f) - o.restrict(l, f)).max() <= 1e-15
            assert np.array_equal(drv.interp_add(l - 1, f, x), o.interp_add(l - 1, f, x))


@pytest.mark.parametrize("bclo,bchi", BCS[:3])
def test_mac_bottom_solve(backend, bclo, bchi):
    rng = np.random.default_rng(8)
    ncell = (16, 16, 16)
    beta = random_beta(rng, ncell, bclo)
    drv = Drv(ncell, (1, 1, 1), bclo, bchi, beta, backend, max_coarsening_level=1)      # bottom level 8^3
    o = MacOracle(ncell, (0, 0, 0), (1, 1, 1), bclo, bchi, max_coarsening_level=1)
    o.set_beta(*beta)
    b = rng.standard_normal((8, 8, 8))
    x, it = drv.bottom_solve(b)
    xo, ito = o.bottom_solve(b)
    assert it == ito and it > 2
    # same iteration count; the iterates differ by the summation order of the dot products, amplified by the Krylov
    # recurrences (the solve itself stops at bottom_rtol = 1e-4)
    assert np.abs(x - xo).max() <= 1e-7 * np.abs(xo).max()


CASES = [("double_shear", (32, 32, 16), dict(prob_hi=(1.0, 1.0, 0.5))), ("rayleigh_taylor", (16, 16, 32), {}),
         ("rayleigh_taylor", (64, 16, 32), dict(prob_hi=(2.0, 0.5, 1.0)))]      # whole 64 x 16 tiles: k_mac_rb3 on level 0


@pytest.mark.parametrize("name,ncell,kw", CASES)
@pytest.mark.parametrize("fused", [128, 2])
def test_mac_project_matches_oracle(backend, name, ncell, kw, fused):
    if ncell[0] == 64 and fused != 2:
        pytest.skip("the whole-tile case is there for the one-pass smoother")
    prob = problems.MAC_PROBLEMS[name](ncell, **kw)
    drv = Drv(ncell, prob["prob_hi"], prob["bclo"], prob["bchi"], prob["beta"], backend, fused_rb_min_n=fused, fused_rb_const=1)
    st, u, phi = drv.project(prob["umac"], ng=1)
    o = MacOracle(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"])
    o.set_beta(*prob["beta"]); o.set_umac(*prob["umac"])
    div0 = o.divergence()
    so = o.project(1e-11, 1e-14)
    assert st.status == 0 and so.status == 0
    assert st.n_vcycles == so.n_vcycles and st.n_levels == so.n_levels
    assert st.bottom_iters_total == so.bottom_iters_total
    for i in range(st.n_vcycles):
        assert abs(st.res_history[i] - so.res_history[i]) <= 1e-6 * so.res_history[i] + 1e-14 * so.rhs_norm0
    for a, b in zip(u, o.get_umac()):
        assert l2rel(a, b) < 1e-12
    po = o.get_phi()
    assert l2rel(phi - phi.mean(), po - po.mean()) < 1e-10
    assert st.sigma_is_const == (1 if name == "double_shear" else 0)
    # the projected field is discretely divergence-free
    o.set_umac(*u)
    assert np.abs(o.divergence()).max() <= 2e-11 * np.abs(div0).max()


def test_mac_project_dirichlet_and_guess(backend):
    """Outflow (Dirichlet) on x-hi, inflow/walls (Neumann) elsewhere in x/y, periodic z; then the mac_phi overload:
    the converged phi as initial guess needs no V-cycle and returns phi unchanged."""
    rng = np.random.default_rng(9)
    ncell, hi = (32, 16, 8), (2.0, 1.0, 0.5)
    bclo, bchi = (2, 2, 0), (1, 2, 0)
    beta = random_beta(rng, ncell, bclo)
    umac = [rng.standard_normal(s) for s in face_shapes(ncell)]
    umac[0][:, :, 0] = 1.0                      # inflow face values stay as given
    umac[1][:, 0, :] = 0.0; umac[1][:, -1, :] = 0.0
    umac[2][-1] = umac[2][0]
    drv = Drv(ncell, hi, bclo, bchi, beta, backend, fused_rb_min_n=2, fused_rb_const=1)
    st, u, phi = drv.project(umac)
    o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
    o.set_beta(*beta); o.set_umac(*umac)
    so = o.project(1e-11, 1e-14)
    assert st.status == 0 and st.n_vcycles == so.n_vcycles
    for a, b in zip(u, o.get_umac()):
        assert l2rel(a, b) < 1e-12
    assert l2rel(phi, o.get_phi()) < 1e-10
    assert np.array_equal(u[0][:, :, 0], umac[0][:, :, 0])          # Neumann face: no correction
    st2, u2, phi2 = drv.project(umac, phi_guess=phi)
    assert st2.n_vcycles == 0 and np.array_equal(phi2, phi)
    for a, b in zip(u2, u):
        assert l2rel(a, b) < 1e-9
    # getFluxes(mac_phi): u - u* = flux
    fl = [face_fab_zeros(ncell, d, 0, drv.like, drv.device) for d in range(3)]
    drv.P.getFluxes(fl, drv.cell(0, phi, 1))
    for d in range(3):
        f = fl[d].numpy_valid()[0]
        assert np.abs(f - (u[d] - umac[d])).max() <= 1e-9 * np.abs(f).max()


def test_mac_not_converged_and_layout_checks(backend):
    prob = problems.mac_double_shear((16, 16, 16))
    drv = Drv((16, 16, 16), (1, 1, 1), prob["bclo"], prob["bchi"], prob["beta"], backend, maxiter=1)
    st, *_ = drv.project(prob["umac"], raise_on_fail=False)
    assert st.status == 5 and st.n_vcycles == 1
    with pytest.raises(Exception):
        drv.project(prob["umac"])


def test_mac_nan_input_is_reported(backend):
    """A NaN face velocity must not come back as status 0 with zero V-cycles (AMReX's MLMG fails on it)."""
    prob = problems.mac_double_shear((16, 16, 16))
    drv = Drv((16, 16, 16), (1, 1, 1), prob["bclo"], prob["bchi"], prob["beta"], backend)
    umac = [u.copy() for u in prob["umac"]]
    umac[0][3, 4, 5] = np.nan
    st, *_ = drv.project(umac, raise_on_fail=False)
    assert st.status == 6


@pytest.mark.gpu
def test_mac_host_fabs_on_gpu():
    """A CPU build of the caller hands host fabs: staged inside the library."""
    import torch
    assert torch.cuda.is_available()
    backend = dict(lib=capi.load(), like="numpy")
    ncell = (32, 32, 16)
    prob = problems.mac_rayleigh_taylor(ncell, prob_hi=(0.5, 0.5, 0.25))
    drv = Drv(ncell, prob["prob_hi"], prob["bclo"], prob["bchi"], prob["beta"], backend)
    st, u, phi = drv.project(prob["umac"], ng=1)
    o = MacOracle(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"])
    o.set_beta(*prob["beta"]); o.set_umac(*prob["umac"])
    so = o.project(1e-11, 1e-14)
    assert st.n_vcycles == so.n_vcycles
    for a, b in zip(u, o.get_umac()):
        assert l2rel(a, b) < 1e-12


@pytest.mark.parametrize("bclo,bchi", [BCS[0], BCS[3], BCS[2]])
@pytest.mark.parametrize("ncell", [(80, 24, 8), (128, 32, 8)])
def test_mac_fused_smoother_tiles(backend, bclo, bchi, ncell):
    """The one-pass smoother across tile and chunk borders: 2 x 2 tiles of 64 x 16 cells, two z-chunks, three smooths in a
    row.  (80, 24, 8): partial tiles -> k_mac_rb2; (128, 32, 8): whole tiles -> k_mac_rb3 (hoisted indices)."""
    rng = np.random.default_rng(11)
    hi = (1.0, 0.3, 0.1)
    beta = random_beta(rng, ncell, bclo)
    drv = Drv(ncell, hi, bclo, bchi, beta, backend, fused_rb_min_n=2, fused_rb_const=1)
    o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
    o.set_beta(*beta)
    x = rng.standard_normal(ncell[::-1]); b = rng.standard_normal(ncell[::-1])
    s, so = drv.smooth(0, x, b, 3), o.smooth(0, x, b, 3)
    assert np.abs(s - so).max() <= 1e-13 * np.abs(so).max()


@pytest.mark.parametrize("host", [False, pytest.param(True, marks=pytest.mark.gpu)])
def test_mac_multibox_equals_onebox(backend, host):
    """amr.max_grid_size chops the level: per-box face fabs (shared faces duplicated, 1 ghost layer of junk) through the
    *_boxes entry points must give bit-identical umac / phi / fluxes to the one-box call.  host=True on the GPU: host fabs."""
    if host and backend["like"] == "numpy":
        pytest.skip("host fabs are a case of the CUDA build only (the emulation has one kind of memory)")
    from oracle.oracle import chop_boxes
    rng = np.random.default_rng(13)
    ncell, hi = (32, 16, 8), (2.0, 1.0, 0.5)
    bclo, bchi = (2, 2, 0), (1, 2, 0)
    beta = random_beta(rng, ncell, bclo)
    umac = [rng.standard_normal(s) for s in face_shapes(ncell)]
    umac[2][-1] = umac[2][0]
    drv = Drv(ncell, hi, bclo, bchi, beta, backend)
    st1, u1, phi1 = drv.project(umac)
    like = "numpy" if host else backend["like"]
    dev = backend.get("device", "cuda")
    boxes = [((b[0], b[1], b[2]), (b[3], b[4], b[5])) for b in chop_boxes(ncell, (16, 8, 4))]

    def cut(a, d, ng, fill=0.0):
        out = []
        for lo, h in boxes:
            e = [h[q] + 1 + (1 if q == d else 0) for q in range(3)]
            sub = a[lo[2]:e[2], lo[1]:e[1], lo[0]:e[0]]
            full = np.full((1,) + tuple(x + 2 * ng for x in sub.shape), fill)
            full[(0,) + tuple(slice(ng, ng + x) for x in sub.shape)] = sub
            f = Fab(full, ng)
            if like != "numpy":
                import torch
                f = Fab(torch.from_numpy(full).to(dev), ng)
            out.append(f)
        return out

    P = MacProjector(Geometry(ncell, (0, 0, 0), hi, [b == 0 for b in bclo]), lib=backend["lib"], boxes=boxes)
    P.setDomainBC(bclo, bchi)
    P.initProjector(LPInfo().setMaxCoarseningLevel(100), [cut(beta[d], d, 0) for d in range(3)])
    fu = [cut(umac[d], d, 1, 777.0) for d in range(3)]
    fphi = cut(np.zeros(ncell[::-1]), -1, 1, 555.0)
    P.setUMAC(fu)
    st = P.project(fphi, 1e-11, 1e-14)
    assert st.status == 0 and st.n_vcycles == st1.n_vcycles

    def paste(fabs, d, shape):
        out = np.full(shape, np.nan)
        for (lo, h), f in zip(boxes, fabs):
            e = [h[q] + 1 + (1 if q == d else 0) for q in range(3)]
            v = f.numpy_valid()[0]
            cur = out[lo[2]:e[2], lo[1]:e[1], lo[0]:e[0]]
            seen = ~np.isnan(cur)
            assert np.array_equal(cur[seen], v[seen])            # a face shared by two boxes agrees in both
            out[lo[2]:e[2], lo[1]:e[1], lo[0]:e[0]] = v
        return out

    for d in range(3):
        assert np.array_equal(paste(fu[d], d, umac[d].shape), u1[d])
        g = fu[d][0].data
        g = g.cpu().numpy() if hasattr(g, "cpu") else g
        assert g[0, 0, 0, 0] == 777.0                              # ghost cells untouched
    assert np.array_equal(paste(fphi, -1, ncell[::-1]), phi1)
    fl = [cut(np.zeros(s), d, 0) for d, s in enumerate(face_shapes(ncell))]
    P.getFluxes(fl, fphi)
    for d in range(3):
        f = paste(fl[d], d, umac[d].shape)
        assert np.abs(f - (u1[d] - umac[d])).max() <= 1e-9 * np.abs(f).max()
    P.close()


@pytest.mark.gpu
@pytest.mark.parametrize("const", [True, False])
def test_mac_one_pass_smoother_is_deterministic_on_gpu(const):
    """Race proxy for k_mac_rb3 (compute-sanitizer is closed on this pool): 20 repeats of two one-pass smooths on whole tiles
    (128 x 32 x 24 cells, three z-chunks) must be bitwise identical."""
    import torch
    torch.cuda.set_device(0)
    rng = np.random.default_rng(8)
    ncell, hi = (128, 32, 24), (4.0, 1.0, 0.75)
    bclo = bchi = (0, 0, 0)
    beta = random_beta(rng, ncell, bclo, const)
    drv = Drv(ncell, hi, bclo, bchi, beta, dict(lib=capi.load(), like="cuda", device="cuda:0"), fused_rb_min_n=2, fused_rb_const=1)
    x, b = rng.standard_normal(ncell[::-1]), rng.standard_normal(ncell[::-1])
    ref = None
    for rep in range(20):
        s = drv.smooth(0, x, b, 2)
        if ref is None:
            ref = s
        assert np.array_equal(s, ref), rep


@pytest.mark.gpu
def test_mac_iteration_graph_matches_kernel_by_kernel_path():
    """mp_options.use_graph: the MLMG iteration replayed from a CUDA graph gives bit-identical umac / phi, counts and residual
    histories to the kernel-by-kernel path (uniform and variable beta, the one-pass and the colour-per-launch smoother, walls
    with a Dirichlet face, an initial guess, updateBeta between projections of one handle); a bottom solve that needs MLMG's
    fallback (bottom_maxiter 1) makes the solve start again without the graph from the same initial guess."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    backend = dict(lib=capi.load(), like="torch", device="cuda")
    rng = np.random.default_rng(41)
    ncell, hi = (128, 32, 32), (4.0, 1.0, 1.0)
    umac0 = [rng.standard_normal(s) for s in face_shapes(ncell)]
    guess = rng.standard_normal(ncell[::-1])

    def run(use_graph, bclo, bchi, **opts):
        out = []
        d = Drv(ncell, hi, bclo, bchi, random_beta(rng_b(), ncell, bclo, const=True), backend, use_graph=use_graph, **opts)
        for const, g in ((True, None), (False, None), (False, guess)):
            d.P.updateBeta([d.face(b) for b in random_beta(rng_b(), ncell, bclo, const=const)])
            um = [u.copy() for u in umac0]
            if bclo[0] == 0: um[0][:, :, -1] = um[0][:, :, 0]
            if bclo[1] == 0: um[1][:, -1, :] = um[1][:, 0, :]
            if bclo[2] == 0: um[2][-1, :, :] = um[2][0, :, :]
            st, u, phi = d.project(um, phi_guess=g, raise_on_fail=False)
            out.append((st.status, st.n_vcycles, st.bottom_iters_total, [st.res_history[i] for i in range(st.n_vcycles)], u, phi,
                        st.graph_replays))
        d.P.close()
        return out

    def rng_b():
        return np.random.default_rng(43)

    for bclo, bchi in (BCS[0], BCS[3]):
        for opts in (dict(fused_rb_min_n=64), dict(fused_rb_min_n=100000)):
            plain, graph = run(0, bclo, bchi, **opts), run(1, bclo, bchi, **opts)
            for p, g in zip(plain, graph):
                assert p[0] == 0 and g[0] == 0 and p[6] == 0 and g[6] == g[1] > 0
                assert p[1] == g[1] and p[2] == g[2] and p[3] == g[3]
                assert all(np.array_equal(a, b) for a, b in zip(p[4], g[4])) and np.array_equal(p[5], g[5])
    pf, gf = run(0, *BCS[3], bottom_maxiter=1), run(1, *BCS[3], bottom_maxiter=1)
    for p, g in zip(pf, gf):
        assert p[0] == g[0] and p[1] == g[1] and g[6] <= 1
        assert all(np.array_equal(a, b) for a, b in zip(p[4], g[4])) and np.array_equal(p[5], g[5])
