"""Pins for the MAC-projection oracle (oracle/mac_oracle.cpp; parity unpinned upstream).

AMReX is absent, so the restatement is pinned by identities that must hold for the published algorithm:
  * mlabeclap_adotx == -Div diag(b) Grad assembled independently with scipy (1-D difference matrices + kron), including
    the boundary rows: periodic wrap, Neumann (no flux), Dirichlet with the maxorder-3 ghost (-2 phi0 + phi1/3);
  * the projected face velocities are discretely divergence-free to the solver tolerance;
  * phi == sparse direct solve of the assembled operator;
  * average_down_faces / average_down / piecewise-constant interpolation against numpy one-liners;
  * the red-black sweep against a direct numpy restatement of abec_gsrb.
"""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from oracle.mac import MacOracle, face_shapes, PERIODIC, DIRICHLET, NEUMANN

BCS = [((0, 0, 0), (0, 0, 0)), ((2, 2, 0), (2, 2, 0)), ((2, 0, 2), (1, 0, 2)), ((1, 2, 2), (1, 2, 1))]


def grad_1d(n, lo, hi):
    """(n+1) x n matrix: cell values -> face differences (per unit cell), boundary rows per BC."""
    G = sp.lil_matrix((n + 1, n))
    for f in range(1, n):
        G[f, f] = 1.0
        G[f, f - 1] = -1.0
    if lo == PERIODIC:
        G[0, 0] = 1.0; G[0, n - 1] = -1.0
        G[n, 0] = 1.0; G[n, n - 1] = -1.0
    else:
        if lo == DIRICHLET:          # phi0 - ghost, ghost = -2 phi0 + phi1/3
            G[0, 0] = 3.0; G[0, 1] = -1.0 / 3.0
        if hi == DIRICHLET:          # ghost - phi_{n-1}
            G[n, n - 1] = -3.0; G[n, n - 2] = 1.0 / 3.0
    return G.tocsr()


def div_1d(n):
    D = sp.lil_matrix((n, n + 1))
    for c in range(n):
        D[c, c + 1] = 1.0
        D[c, c] = -1.0
    return D.tocsr()


def assemble(ncell, dx, bclo, bchi, beta):
    nx, ny, nz = ncell
    I = [sp.identity(n, format="csr") for n in ncell]
    G = [grad_1d(ncell[d], bclo[d], bchi[d]) for d in range(3)]
    D = [div_1d(n) for n in ncell]
    # C order [k, j, i]: kron(z, y, x)
    Gx, Gy, Gz = sp.kron(I[2], sp.kron(I[1], G[0])), sp.kron(I[2], sp.kron(G[1], I[0])), sp.kron(G[2], sp.kron(I[1], I[0]))
    Dx, Dy, Dz = sp.kron(I[2], sp.kron(I[1], D[0])), sp.kron(I[2], sp.kron(D[1], I[0])), sp.kron(D[2], sp.kron(I[1], I[0]))
    A = -(Dx @ sp.diags(beta[0].ravel()) @ Gx / dx[0] ** 2 + Dy @ sp.diags(beta[1].ravel()) @ Gy / dx[1] ** 2
          + Dz @ sp.diags(beta[2].ravel()) @ Gz / dx[2] ** 2)
    return A.tocsr()


def random_beta(rng, ncell, bclo):
    b = [rng.uniform(0.5, 2.0, s) for s in face_shapes(ncell)]
    if bclo[0] == PERIODIC: b[0][:, :, -1] = b[0][:, :, 0]
    if bclo[1] == PERIODIC: b[1][:, -1, :] = b[1][:, 0, :]
    if bclo[2] == PERIODIC: b[2][-1, :, :] = b[2][0, :, :]
    return b


@pytest.mark.parametrize("bclo,bchi", BCS)
def test_adotx_is_assembled_fv_operator(bclo, bchi):
    rng = np.random.default_rng(0)
    ncell, hi = (8, 6, 4), (0.8 * 3, 0.6 * 5, 0.4 * 7)
    dx = [hi[d] / ncell[d] for d in range(3)]
    o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
    beta = random_beta(rng, ncell, bclo)
    o.set_beta(*beta)
    A = assemble(ncell, dx, bclo, bchi, beta)
    x = rng.standard_normal(ncell[::-1])
    y = o.apply(0, x)
    ref = (A @ x.ravel()).reshape(x.shape)
    assert np.abs(y - ref).max() <= 1e-13 * np.abs(ref).max()


@pytest.mark.parametrize("bclo,bchi", BCS)
def test_projection_is_divergence_free_and_matches_direct_solve(bclo, bchi):
    rng = np.random.default_rng(1)
    ncell, hi = (16, 16, 8), (1.0, 1.0, 0.5)
    dx = [hi[d] / ncell[d] for d in range(3)]
    o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
    beta = random_beta(rng, ncell, bclo)
    o.set_beta(*beta)
    u = [rng.standard_normal(s) for s in face_shapes(ncell)]
    for d, ax in ((0, 2), (1, 1), (2, 0)):       # periodic image faces agree; wall faces carry no flow
        sl_lo = [slice(None)] * 3; sl_hi = [slice(None)] * 3
        sl_lo[ax], sl_hi[ax] = 0, -1
        if bclo[d] == PERIODIC:
            u[d][tuple(sl_hi)] = u[d][tuple(sl_lo)]
        if bclo[d] == NEUMANN:
            u[d][tuple(sl_lo)] = 0
        if bchi[d] == NEUMANN:
            u[d][tuple(sl_hi)] = 0
    o.set_umac(*u)
    div0 = o.divergence()
    st = o.project(1e-11, 1e-14)
    assert st.status == 0 and st.n_vcycles <= 14
    div1 = o.divergence()
    assert np.abs(div1).max() <= 2e-11 * np.abs(div0).max()
    A = assemble(ncell, dx, bclo, bchi, beta)
    rhs = -div0.ravel()
    singular = not any(b == DIRICHLET for b in bclo + bchi)
    if singular:
        rhs = rhs - rhs.mean()
        A = A.tolil(); A[0, :] = 0; A[0, 0] = 1.0; rhs[0] = 0.0; A = A.tocsr()
    ref = spla.spsolve(A.tocsc(), rhs).reshape(div0.shape)
    phi = o.get_phi()
    if singular:
        ref -= ref.mean(); phi = phi - phi.mean()
    assert np.linalg.norm(phi - ref) <= 1e-8 * np.linalg.norm(ref)


def test_transfer_operators():
    rng = np.random.default_rng(2)
    ncell = (8, 8, 4)
    o = MacOracle(ncell)
    beta = random_beta(rng, ncell, (0, 0, 0))
    o.set_beta(*beta)
    bx1, by1, bz1 = o.get_beta(1)
    assert np.allclose(bx1, 0.25 * (beta[0][0::2, 0::2, ::2] + beta[0][1::2, 0::2, ::2] + beta[0][0::2, 1::2, ::2] + beta[0][1::2, 1::2, ::2]), rtol=1e-15)
    assert np.allclose(bz1, 0.25 * (beta[2][::2, 0::2, 0::2] + beta[2][::2, 1::2, 0::2] + beta[2][::2, 0::2, 1::2] + beta[2][::2, 1::2, 1::2]), rtol=1e-15)
    f = rng.standard_normal(ncell[::-1])
    c = o.restrict(1, f)
    assert np.allclose(c, f.reshape(2, 2, 4, 2, 4, 2).mean(axis=(1, 3, 5)), rtol=1e-15)
    g = o.interp_add(0, f, c)
    assert np.array_equal(g, f + np.repeat(np.repeat(np.repeat(c, 2, 0), 2, 1), 2, 2))


def test_gsrb_against_numpy_restatement():
    """abec_gsrb on a Neumann/Dirichlet box, both colours, against explicit loops."""
    rng = np.random.default_rng(3)
    ncell, bclo, bchi = (6, 4, 4), (2, 0, 1), (1, 0, 2)
    o = MacOracle(ncell, (0, 0, 0), (1.0, 1.0, 1.0), bclo, bchi)
    beta = random_beta(rng, ncell, bclo)
    o.set_beta(*beta)
    x = rng.standard_normal(ncell[::-1]); b = rng.standard_normal(ncell[::-1])
    got = o.smooth(0, x, b, 1)
    nx, ny, nz = ncell
    dh = [(ncell[d] / 1.0) ** 2 for d in range(3)]
    fac = {PERIODIC: 0.0, NEUMANN: 1.0, DIRICHLET: -2.0}
    phi = x.copy()
    for rb in range(2):
        g = np.zeros((nz + 2, ny + 2, nx + 2)); g[1:-1, 1:-1, 1:-1] = phi
        g[1:-1, 1:-1, 0] = phi[:, :, 0]; g[1:-1, 1:-1, -1] = -2 * phi[:, :, -1] + phi[:, :, -2] / 3     # x: Neumann / Dirichlet
        g[1:-1, 0, 1:-1] = phi[:, -1, :]; g[1:-1, -1, 1:-1] = phi[:, 0, :]                             # y: periodic
        g[0, 1:-1, 1:-1] = -2 * phi[0] + phi[1] / 3; g[-1, 1:-1, 1:-1] = phi[-1]                       # z: Dirichlet / Neumann
        for k in range(nz):
            for j in range(ny):
                for i in range(nx):
                    if (i + j + k + rb) % 2:
                        continue
                    bm = [beta[0][k, j, i], beta[1][k, j, i], beta[2][k, j, i]]
                    bp = [beta[0][k, j, i + 1], beta[1][k, j + 1, i], beta[2][k + 1, j, i]]
                    idx = (i, j, k)
                    gamma = sum(dh[d] * (bm[d] + bp[d]) for d in range(3))
                    delta = sum(dh[d] * (bm[d] * (fac[bclo[d]] if idx[d] == 0 else 0.0) + bp[d] * (fac[bchi[d]] if idx[d] == ncell[d] - 1 else 0.0))
                                for d in range(3))
                    rho = (dh[0] * (bm[0] * g[k + 1, j + 1, i] + bp[0] * g[k + 1, j + 1, i + 2]) + dh[1] * (bm[1] * g[k + 1, j, i + 1] + bp[1] * g[k + 1, j + 2, i + 1])
                           + dh[2] * (bm[2] * g[k, j + 1, i + 1] + bp[2] * g[k + 2, j + 1, i + 1]))
                    res = b[k, j, i] - (gamma * phi[k, j, i] - rho)
                    phi[k, j, i] += 1.15 / (gamma - delta) * res
    assert np.abs(got - phi).max() <= 1e-13 * np.abs(phi).max()


def test_abec_solve_reproduces_a_quadratic_exactly():
    """General MLABecLaplacian solve (scalar diffusion groundwork): (alpha a - beta b lap) phi = rhs with inhomogeneous
    Dirichlet values.  For phi quadratic in x the 3-point stencil and the maxorder-3 boundary polynomial are exact, so the
    solver must return phi at the cell centres to the solver tolerance - an independent pin of the boundary treatment."""
    ncell, hi = (16, 8, 8), (2.0, 1.0, 1.0)
    bclo, bchi = (DIRICHLET, PERIODIC, NEUMANN), (DIRICHLET, PERIODIC, NEUMANN)
    alpha, beta, b0 = 1.0, 0.05, 0.7
    o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
    o.set_beta(*[np.full(s, b0) for s in face_shapes(ncell)])
    rng = np.random.default_rng(4)
    a = rng.uniform(0.5, 2.0, ncell[::-1])
    x = (np.arange(ncell[0]) + 0.5) * hi[0] / ncell[0]
    f = lambda t: 1.0 + 0.5 * t - 0.3 * t * t                  # phi(x); phi'' = -0.6
    phi = np.broadcast_to(f(x), ncell[::-1])
    rhs = alpha * a * phi - beta * b0 * (-0.6)
    o.set_scalars(alpha, beta); o.set_acoef(a)
    st = o.solve_abec(rhs, 1e-12, 0.0, dict(xlo=np.full((ncell[2], ncell[1]), f(0.0)), xhi=np.full((ncell[2], ncell[1]), f(hi[0]))))
    assert st.status == 0
    assert np.abs(o.get_phi() - phi).max() <= 1e-10 * np.abs(phi).max()


def test_abec_solve_matches_sparse_solve():
    """alpha a + beta A (A = the assembled homogeneous operator) solved directly vs the MLMG restatement; variable a and b."""
    rng = np.random.default_rng(6)
    ncell, hi = (16, 16, 8), (1.0, 1.0, 0.5)
    bclo, bchi = (NEUMANN, PERIODIC, NEUMANN), (DIRICHLET, PERIODIC, NEUMANN)
    dx = [hi[d] / ncell[d] for d in range(3)]
    alpha, beta = 1.0, 0.01
    o = MacOracle(ncell, (0, 0, 0), hi, bclo, bchi)
    b = random_beta(rng, ncell, bclo)
    o.set_beta(*b)
    a = rng.uniform(0.5, 2.0, ncell[::-1])
    rhs = rng.standard_normal(ncell[::-1])
    o.set_scalars(alpha, beta); o.set_acoef(a)
    st = o.solve_abec(rhs, 1e-12, 0.0)
    M = sp.diags(alpha * a.ravel()) + beta * assemble(ncell, dx, bclo, bchi, b)
    ref = spla.spsolve(M.tocsc(), rhs.ravel()).reshape(rhs.shape)
    assert st.status == 0 and st.n_vcycles <= 12
    assert np.linalg.norm(o.get_phi() - ref) <= 1e-9 * np.linalg.norm(ref)
