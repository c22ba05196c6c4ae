"""Pins for the CPU oracle (the reference's own tests hold no golden vectors for the
nodal projection, so the oracle is pinned by identities that must hold for the published algorithm).

 (1) adotx == assembled Q1 FEM stiffness (independent scipy assembly)       -> coefficient tripwire
 (2) divergence and gradient are negative adjoints under the dot-mask weights
 (3) the MG solve agrees with a sparse direct solve of the assembled operator
 (4) rhs = 0 -> phi = 0, u unchanged;  A symmetric, constants in its nullspace
"""
import numpy as np
import pytest
import scipy.sparse.linalg as spla

from oracle import fem_assembly as fem
from oracle.oracle import Oracle, PERIODIC, DIRICHLET, NEUMANN, LEX, TILED, chop_boxes

RNG = np.random.default_rng(1234)


def _neumann_scale(ncell, bclo, bchi):
    """2^(number of Neumann faces a node lies on), full nodal array [k,j,i]."""
    nx, ny, nz = ncell
    s = np.ones((nz + 1, ny + 1, nx + 1))
    for d, ax in ((0, 2), (1, 1), (2, 0)):
        sl = [slice(None)] * 3
        if bclo[d] == NEUMANN:
            sl[ax] = 0
            s[tuple(sl)] *= 2
        if bchi[d] == NEUMANN:
            sl[ax] = ncell[d]
            s[tuple(sl)] *= 2
    return s


def _dirichlet_mask(ncell, bclo, bchi):
    nx, ny, nz = ncell
    m = np.zeros((nz + 1, ny + 1, nx + 1), dtype=bool)
    for d, ax in ((0, 2), (1, 1), (2, 0)):
        sl = [slice(None)] * 3
        if bclo[d] == DIRICHLET:
            sl[ax] = 0
            m[tuple(sl)] = True
        if bchi[d] == DIRICHLET:
            sl[ax] = ncell[d]
            m[tuple(sl)] = True
    return m


BC_CASES = [
    ((PERIODIC,) * 3, (PERIODIC,) * 3),
    ((NEUMANN,) * 3, (NEUMANN,) * 3),
    ((PERIODIC, PERIODIC, NEUMANN), (PERIODIC, PERIODIC, NEUMANN)),
    ((NEUMANN, NEUMANN, PERIODIC), (DIRICHLET, NEUMANN, PERIODIC)),
    ((DIRICHLET, PERIODIC, NEUMANN), (DIRICHLET, PERIODIC, DIRICHLET)),
]


@pytest.mark.parametrize("bclo,bchi", BC_CASES)
@pytest.mark.parametrize("boxes", [None, 4])
def test_adotx_equals_fem_assembly(bclo, bchi, boxes):
    ncell = (8, 12, 4)
    plo, phi_ = (0.0, 0.0, 0.0), (2.4, 6.0, 2.8)       # dx = (0.3, 0.5, 0.7): anisotropic
    dx = [(phi_[d] - plo[d]) / ncell[d] for d in range(3)]
    sigma = RNG.uniform(0.5, 2.0, (ncell[2], ncell[1], ncell[0]))
    periodic = tuple(b == PERIODIC for b in bclo)
    A, ushape = fem.assemble(ncell, dx, sigma, periodic)
    xu = RNG.standard_normal(ushape)
    x = fem.expand_periodic(xu, ncell, periodic)
    dm = _dirichlet_mask(ncell, bclo, bchi)
    x[dm] = 0.0
    xu = fem.restrict_unique(x, ncell, periodic)
    y_fem = fem.expand_periodic((A @ xu.ravel()).reshape(ushape), ncell, periodic)
    y_fem = y_fem * _neumann_scale(ncell, bclo, bchi)
    y_fem[dm] = 0.0
    o = Oracle(ncell, plo, phi_, bclo, bchi, boxes=None if boxes is None else chop_boxes(ncell, boxes))
    o.set_sigma(sigma)
    o.prepare_sigma()
    y = o.apply(0, x)
    assert np.abs(y - y_fem).max() <= 1e-13 * np.abs(y_fem).max()


@pytest.mark.parametrize("bclo,bchi", BC_CASES[:3])
def test_div_is_negative_adjoint_of_grad(bclo, bchi):
    ncell = (8, 4, 12)
    plo, phi_ = (0.0, 0.0, 0.0), (0.8, 0.6, 2.4)
    o = Oracle(ncell, plo, phi_, bclo, bchi)
    u = RNG.standard_normal((3,) + o.cshape())
    phi = RNG.standard_normal(o.nshape())
    # make phi consistent on periodic images
    periodic = tuple(b == PERIODIC for b in bclo)
    phi = fem.expand_periodic(fem.restrict_unique(phi, ncell, periodic), ncell, periodic)
    o.set_sigma(np.ones(o.cshape()))
    o.prepare_sigma()
    o.set_vel(u)
    rhs = o.divu()
    lhs = o.dot(0, rhs, phi)                 # dot-mask weighted: owners once, 1/2 per Neumann face
    _, g = o.mknewu(phi, np.zeros_like(u))
    rhs2 = -(u * g).sum()
    assert abs(lhs - rhs2) <= 1e-12 * max(abs(lhs), 1.0)


def test_symmetry_and_nullspace():
    ncell = (8, 8, 8)
    bclo = bchi = (PERIODIC, NEUMANN, NEUMANN)
    o = Oracle(ncell, bclo=bclo, bchi=bchi)
    o.set_sigma(RNG.uniform(0.5, 2.0, o.cshape()))
    o.prepare_sigma()
    periodic = (True, False, False)
    mk = lambda: fem.expand_periodic(fem.restrict_unique(RNG.standard_normal(o.nshape()), ncell, periodic), ncell, periodic)
    x, y = mk(), mk()
    a = o.dot(0, o.apply(0, x), y)
    b = o.dot(0, x, o.apply(0, y))
    assert abs(a - b) <= 1e-12 * abs(a)
    one = np.ones(o.nshape())
    assert np.abs(o.apply(0, one)).max() <= 1e-10


@pytest.mark.parametrize("bclo,bchi", [BC_CASES[0], BC_CASES[2], BC_CASES[3]])
def test_projection_matches_sparse_direct_solve(bclo, bchi):
    ncell = (16, 16, 16)
    plo, phi_ = (0.0, 0.0, 0.0), (1.0, 1.0, 1.0)
    dx = [1.0 / 16] * 3
    sigma = RNG.uniform(0.5, 2.0, (16, 16, 16))
    u = RNG.standard_normal((3, 16, 16, 16))
    periodic = tuple(b == PERIODIC for b in bclo)
    o = Oracle(ncell, plo, phi_, bclo, bchi)
    o.set_sigma(sigma)
    o.set_vel(u)
    st = o.project(1e-12, 0.0)
    assert st.status == 0 and st.n_vcycles <= 12
    phi = o.get_phi()
    # independent: assembled FEM system, rhs from the oracle's divu (un-doubled at Neumann faces)
    A, ushape = fem.assemble(ncell, dx, sigma, periodic)
    rhs = o.get_rhs() / _neumann_scale(ncell, bclo, bchi)
    dm = _dirichlet_mask(ncell, bclo, bchi)
    free = ~fem.restrict_unique(dm, ncell, periodic).ravel()
    b = fem.restrict_unique(rhs, ncell, periodic).ravel()
    Af = A[free][:, free].tocsc()
    singular = not dm.any()
    if singular:
        # pin one unknown; the system is consistent because the oracle removed the mean of rhs
        keep = np.arange(Af.shape[0] - 1)
        xf = np.zeros(Af.shape[0])
        xf[keep] = spla.spsolve(Af[keep][:, keep].tocsc(), b[free][keep])
    else:
        xf = spla.spsolve(Af, b[free])
    xu = np.zeros(A.shape[0])
    xu[free] = xf
    phi_d = fem.expand_periodic(xu.reshape(ushape), ncell, periodic)
    if singular:
        phi = phi - phi.mean()
        phi_d = phi_d - phi_d.mean()
    assert np.linalg.norm(phi - phi_d) <= 1e-8 * np.linalg.norm(phi_d)


def test_zero_rhs_is_noop():
    ncell = (16, 16, 16)
    o = Oracle(ncell)
    z, y, x = np.meshgrid(*(3 * [(np.arange(16) + 0.5) / 16]), indexing="ij")
    # discretely divergence-free: u depends on y,z only; v on x,z only; w on x,y only
    u = np.stack([np.sin(2 * np.pi * y) * np.cos(2 * np.pi * z), np.cos(2 * np.pi * x) + z * 0, np.sin(2 * np.pi * (x + y))])
    o.set_vel(u)
    o.set_sigma(np.ones(o.cshape()))
    st = o.project()
    assert st.n_vcycles == 0
    assert np.abs(o.get_vel() - u).max() <= 1e-13
    assert np.abs(o.get_phi()).max() == 0.0


def test_box_layout_and_ordering_change_iterates_not_solution():
    """F5a: lexicographic GS iterates depend on the box layout; the converged projection does not."""
    ncell = (32, 32, 32)
    sigma = RNG.uniform(0.5, 2.0, (32, 32, 32))
    u = RNG.standard_normal((3, 32, 32, 32))
    outs = []
    for kw in (dict(), dict(boxes=chop_boxes(ncell, 16)), dict(smoother=TILED, tile=(16, 8, 1 << 30)),
               dict(boxes=chop_boxes(ncell, (32, 32, 16)), smoother=TILED, tile=(32, 32, 1 << 30), agg_level=2)):
        o = Oracle(ncell, **kw)
        o.set_sigma(sigma)
        o.set_vel(u)
        st = o.project(1e-13, 0.0)
        assert st.status == 0
        outs.append((st.n_vcycles, o.get_vel(), o.get_gradphi()))
    n0, u0, g0 = outs[0]
    for n, uu, gg in outs[1:]:
        assert abs(n - n0) <= 2
        assert np.linalg.norm(uu - u0) <= 1e-10 * np.linalg.norm(u0)
        assert np.linalg.norm(gg - g0) <= 1e-10 * np.linalg.norm(g0)
