"""Whole projection (np_project through the C ABI) vs the oracle, on the CPU emulation build.

Two comparisons per configuration (SURVEY.md 8c "parity definition"):
  * oracle with the SAME smoother ordering (TILED)      -> same V-cycle count, iterates equal to round-off
  * oracle with the reference's ordering (LEX, AMReX CPU) -> u, grad(phi), phi-mean within 1e-10 relative L2
    at a tolerance both reach (1e-13), V-cycle counts within +-1.
"""
import numpy as np
import pytest

import nphelp
from incflo_b200 import capi, problems
from oracle.oracle import Oracle, TILED, LEX

BIG = 1 << 30


def l2rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def run_oracle(ncell, prob, rtol, atol, **kw):
    o = Oracle(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"], **kw)
    o.set_sigma(prob["sigma"])
    o.set_vel(prob["vel"])
    st = o.project(rtol, atol)
    return st, o.get_vel(), o.get_phi(), o.get_gradphi()


CASES = [("taylor_green", (32, 32, 32), {}), ("rayleigh_taylor", (16, 16, 32), {}),
         ("double_shear", (32, 32, 16), dict(zwave=True, prob_hi=(1.0, 1.0, 0.5)))]


@pytest.mark.parametrize("name,ncell,kw", CASES)
def test_project_same_ordering(backend, name, ncell, kw):
    prob = problems.PROBLEMS[name](ncell, **kw)
    drv = nphelp.Driver(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"], **backend)
    st, u, phi, gp = drv.project(prob["vel"], prob["sigma"], 1e-11, 1e-14)
    so, uo, po, go = run_oracle(ncell, prob, 1e-11, 1e-14, smoother=TILED, tile=(63, 31, BIG), tile_z_target=-1)
    assert st.status == 0 and so.status == 0
    assert st.n_vcycles == so.n_vcycles and st.n_levels == so.n_levels
    assert st.bottom_iters_total == so.bottom_iters_total
    for i in range(st.n_vcycles):
        assert abs(st.res_history[i] - so.res_history[i]) <= 1e-6 * so.res_history[i] + 1e-14 * so.rhs_norm0
    assert l2rel(u, uo) < 1e-12 and l2rel(gp, go) < 1e-10
    assert l2rel(phi - phi.mean(), po - po.mean()) < 1e-10
    assert st.sigma_is_const == (1 if name != "rayleigh_taylor" else 0)


@pytest.mark.parametrize("name,ncell,kw", CASES[:2])
def test_project_vs_reference_ordering(backend, name, ncell, kw):
    prob = problems.PROBLEMS[name](ncell, **kw)
    drv = nphelp.Driver(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"], **backend)
    st, u, phi, gp = drv.project(prob["vel"], prob["sigma"], 1e-13, 0.0)
    so, uo, po, go = run_oracle(ncell, prob, 1e-13, 0.0, smoother=LEX)
    assert st.status == 0 and so.status == 0
    assert abs(st.n_vcycles - so.n_vcycles) <= 1
    assert l2rel(u, uo) < 1e-10 and l2rel(gp, go) < 1e-10
    assert l2rel(phi - phi.mean(), po - po.mean()) < 1e-10


def test_project_dirichlet_outflow_with_inflow_ghosts(backend):
    """x: mass inflow (Neumann, ghost cells carry u_in) / pressure outflow (Dirichlet); y walls; z periodic."""
    rng = np.random.default_rng(2)
    ncell = (32, 16, 8)
    bclo, bchi = (2, 2, 0), (1, 2, 0)
    sigma = rng.uniform(0.5, 2.0, ncell[::-1])
    vel = 0.1 * rng.standard_normal((3,) + ncell[::-1])
    vel[0] += 1.0
    y = (np.arange(16) + 0.5) / 16
    uin = 6 * y * (1 - y)                      # prob_set_inflow_velocity, probtype 31

    def ghosts(fv):                            # Fab with 1 ghost: [comp, k, j, i]
        nphelp.assign(fv.data[0, 1:-1, 1:-1, 0], np.broadcast_to(uin[None, :], (8, 16)))

    drv = nphelp.Driver(ncell, (0, 0, 0), (2.0, 1.0, 0.5), bclo, bchi, **backend)
    st, u, phi, gp = drv.project(vel, sigma, 1e-11, 1e-14, ghost_setter=ghosts)
    o = Oracle(ncell, (0, 0, 0), (2.0, 1.0, 0.5), bclo, bchi, smoother=TILED, tile=(63, 31, BIG), tile_z_target=-1)
    o.set_sigma(sigma)
    o.set_vel(vel)
    for k in range(8):
        for j in range(16):
            o.set_vel_ghost(-1, j, k, 0, uin[j])
    so = o.project(1e-11, 1e-14)
    assert st.status == 0 and st.n_vcycles == so.n_vcycles
    assert l2rel(u, o.get_vel()) < 1e-11 and l2rel(gp, o.get_gradphi()) < 1e-10 and l2rel(phi, o.get_phi()) < 1e-10


def test_not_converged_is_reported(backend):
    prob = problems.taylor_green((16, 16, 16))
    drv = nphelp.Driver((16, 16, 16), **backend, maxiter=1)
    st, *_ = drv.project(prob["vel"], prob["sigma"], 1e-11, 1e-14, raise_on_fail=False)
    assert st.status == 5 and st.n_vcycles == 1
    with pytest.raises(Exception):
        drv.project(prob["vel"], prob["sigma"], 1e-11, 1e-14)


@pytest.mark.parametrize("host_out", [False, True])
def test_copy_out_and_host_fabs(backend, host_out):
    """np_copy_phi / np_copy_gradphi into caller fabs of other ghost widths; on the GPU also host (numpy) fabs for
    every argument, i.e. the staged path a CPU build of the caller takes."""
    import ctypes as C
    from incflo_b200 import capi
    from incflo_b200.projector import Fab
    ncell = (16, 16, 16)
    prob = problems.taylor_green(ncell)
    bk = dict(backend)
    if host_out:
        bk["like"] = "numpy"                       # GPU backend: host pointers in and out; emulation: same as before
    drv = nphelp.Driver(ncell, **bk)
    st, u, phi, gp = drv.project(prob["vel"], prob["sigma"], 1e-11, 1e-14)
    so, uo, po, go = run_oracle(ncell, prob, 1e-11, 1e-14, smoother=TILED, tile=(63, 31, BIG), tile_z_target=-1)
    assert st.n_vcycles == so.n_vcycles and l2rel(u, uo) < 1e-12
    for ng in (0, 1, 2):
        fp = Fab.zeros(ncell, 1, ng, True, bk["like"], bk.get("device", "cuda"))
        fg = Fab.zeros(ncell, 3, ng, False, bk["like"], bk.get("device", "cuda"))
        capi.check(drv.L, drv.L.np_copy_phi(drv.h, 0, C.byref(fp.c())))
        capi.check(drv.L, drv.L.np_copy_gradphi(drv.h, 0, C.byref(fg.c())))
        assert np.array_equal(fp.numpy_valid()[0], phi) and np.array_equal(fg.numpy_valid(), gp)


def test_project_nan_input_is_reported(backend):
    """A NaN in u* must come back as NP_ERR_DIVERGED, not as a converged zero-cycle solve (fmax-based norms drop NaNs)."""
    ncell = (16, 16, 16)
    prob = problems.PROBLEMS["taylor_green"](ncell)
    vel = prob["vel"].copy()
    vel[1, 3, 4, 5] = np.nan
    drv = nphelp.Driver(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"], **backend)
    st, *_ = drv.project(vel, prob["sigma"], 1e-11, 1e-14, raise_on_fail=False)
    assert st.status == 6


def test_handle_cache_reuse_is_stateless(backend):
    """np_destroy parks the handle, the next np_create with the same key gets it back reset: same address, and the second
    use (other BC kind, other sigma) gives bit-identical results to a fresh handle (NP_HANDLE_CACHE=0 style: cache cleared)."""
    rng = np.random.default_rng(11)
    ncell = (16, 16, 16)
    L = backend["lib"]
    prob = problems.PROBLEMS["taylor_green"](ncell)
    geom_kw = dict(prob_lo=(0, 0, 0), prob_hi=(1, 1, 1))

    def run(bclo, bchi, vel, sigma):
        drv = nphelp.Driver(ncell, geom_kw["prob_lo"], geom_kw["prob_hi"], bclo, bchi, **backend)
        addr = drv.h.value if hasattr(drv.h, "value") else int(drv.h)
        st, u, phi, gp = drv.project(vel, sigma, 1e-11, 1e-14)
        drv.P.close()
        return addr, st.n_vcycles, u, phi, gp

    sig2 = rng.uniform(0.5, 2.0, ncell[::-1])
    vel2 = rng.standard_normal((3,) + ncell[::-1])
    L.np_cache_clear()
    a1, *_ = run((0, 0, 0), (0, 0, 0), prob["vel"], prob["sigma"])
    a2, vc2, u2, phi2, gp2 = run((0, 0, 0), (0, 0, 0), vel2, sig2)          # same key: the parked handle comes back
    assert a2 == a1
    L.np_cache_clear()
    a3, vc3, u3, phi3, gp3 = run((0, 0, 0), (0, 0, 0), vel2, sig2)          # fresh handle
    assert vc2 == vc3 and np.array_equal(u2, u3) and np.array_equal(phi2, phi3) and np.array_equal(gp2, gp3)
    # same key, other BC kinds than the previous use (non-periodic geometry): the arrays are cleared, results equal a fresh handle's
    def run_np(bclo, bchi):
        drv = nphelp.Driver(ncell, (0, 0, 0), (1, 1, 1), bclo, bchi, **backend)
        st, u, phi, gp = drv.project(vel2, sig2, 1e-11, 1e-14)
        drv.P.close()
        return st.n_vcycles, u, phi, gp
    run_np((2, 2, 2), (1, 2, 2))
    r1 = run_np((1, 2, 2), (2, 1, 2))
    L.np_cache_clear()
    r2 = run_np((1, 2, 2), (2, 1, 2))
    assert r1[0] == r2[0] and all(np.array_equal(a, b) for a, b in zip(r1[1:], r2[1:]))
    # another geometry evicts the parked handle and works
    drv = nphelp.Driver((16, 16, 8), **backend)
    drv.P.close()
    L.np_cache_clear()


@pytest.mark.gpu
def test_iteration_graph_matches_kernel_by_kernel_path():
    """np_options.use_graph: one MLMG iteration replayed from a CUDA graph gives bit-identical fields, counts and residual
    histories to the kernel-by-kernel path; the graph is captured again when what it bakes in changes (constant -> variable
    sigma, other BCs through the handle cache); a bottom solve that needs MLMG's fallback path (here: bottom_maxiter 1)
    restarts the solve without the graph and still agrees."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    backend = dict(lib=capi.load(), like="torch", device="cuda")
    L = backend["lib"]
    rng = np.random.default_rng(17)
    ncell = (32, 32, 64)
    vel = rng.standard_normal((3,) + ncell[::-1])
    sig_c = np.full(ncell[::-1], 0.8)
    sig_v = rng.uniform(0.5, 2.0, ncell[::-1])

    def run(use_graph, cases, **opts):
        out = []
        for bclo, bchi, sigma in cases:
            drv = nphelp.Driver(ncell, (0, 0, 0), (1, 1, 2), bclo, bchi, use_graph=use_graph, **opts, **backend)
            st, u, phi, gp = drv.project(vel, sigma, 1e-11, 1e-14)
            out.append((st.status, st.n_vcycles, st.bottom_iters_total, [st.res_history[i] for i in range(st.n_vcycles)], u, phi, gp,
                        st.kernel_launches))
            drv.P.close()                                   # parks the handle: the next case reuses it, graph included
        L.np_cache_clear()
        return out

    per, walls = ((0, 0, 0), (0, 0, 0)), ((2, 1, 2), (2, 2, 1))
    cases = [(*per, sig_c), (*per, sig_c), (*per, sig_v), (*walls, sig_v), (*walls, sig_c)]
    plain, graph = run(0, cases), run(1, cases)
    for p, g in zip(plain, graph):
        assert p[0] == 0 and g[0] == 0
        assert p[1] == g[1] and p[2] == g[2] and p[3] == g[3] and p[7] == g[7]
        assert all(np.array_equal(a, b) for a, b in zip(p[4:7], g[4:7]))
    # fallback: one bottom iteration cannot converge -> MLMG smooths the bottom level instead; the graph path restarts kernel by kernel
    pf, gf = run(0, cases[:1], bottom_maxiter=1), run(1, cases[:1], bottom_maxiter=1)
    assert pf[0][0] == 0 and gf[0][0] == 0 and pf[0][1] == gf[0][1]
    for a, b in zip(pf[0][4:7], gf[0][4:7]):
        assert np.linalg.norm(a - b) <= 1e-12 * max(np.linalg.norm(a), 1e-300)
