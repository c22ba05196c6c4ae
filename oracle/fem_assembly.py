"""ORACLE — TEST INFRASTRUCTURE ONLY.

Independent pin for the restated 27-point operator: assemble the Galerkin Q1 (trilinear) finite
element stiffness with piecewise-constant sigma,  A = - sum_c sigma_c K_c / (dx dy dz),
K_c = K1x (x) M1y (x) M1z + M1x (x) K1y (x) M1z + M1x (x) M1y (x) K1z  (SURVEY.md H1), with scipy.sparse.
Nothing here shares code with oracle/stencil.hpp; agreement of the two is the tripwire for a
mis-remembered coefficient of amrex::mlndlap_adotx_aa.
"""
import numpy as np
import scipy.sparse as sp

PERIODIC, DIRICHLET, NEUMANN = 0, 1, 2


def _k1(h):
    return np.array([[1.0, -1.0], [-1.0, 1.0]]) / h


def _m1(h):
    return np.array([[2.0, 1.0], [1.0, 2.0]]) * h / 6.0


def element_stiffness(dx):
    kx, ky, kz = _k1(dx[0]), _k1(dx[1]), _k1(dx[2])
    mx, my, mz = _m1(dx[0]), _m1(dx[1]), _m1(dx[2])
    # local node order: a (x) fastest, then b (y), then c (z)  ->  kron(z, kron(y, x))
    return np.kron(mz, np.kron(my, kx)) + np.kron(mz, np.kron(ky, mx)) + np.kron(kz, np.kron(my, mx))


def assemble(ncell, dx, sigma, periodic=(True, True, True)):
    """Return (A, nshape): A acts on the unique nodal unknowns, flattened [k, j, i] C-order.
    Periodic dims have n unique nodes, others n+1.  Natural (Neumann) BCs elsewhere."""
    nx, ny, nz = ncell
    dims = [n if p else n + 1 for n, p in zip(ncell, periodic)]   # x, y, z
    Ke = element_stiffness(dx) / (dx[0] * dx[1] * dx[2])

    def nid(i, j, k):
        return (i % dims[0] if periodic[0] else i) + dims[0] * ((j % dims[1] if periodic[1] else j)
                + dims[1] * (k % dims[2] if periodic[2] else k))

    k, j, i = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    k, j, i = k.ravel(), j.ravel(), i.ravel()
    sig = np.asarray(sigma).reshape(nz, ny, nx).ravel()
    loc = [(a, b, c) for c in (0, 1) for b in (0, 1) for a in (0, 1)]
    ids = np.stack([nid(i + a, j + b, k + c) for (a, b, c) in loc], axis=1)    # (ncells, 8)
    rows = np.repeat(ids, 8, axis=1).ravel()
    cols = np.tile(ids, (1, 8)).ravel()
    vals = (-sig[:, None] * Ke.ravel()[None, :]).ravel()
    N = dims[0] * dims[1] * dims[2]
    A = sp.coo_matrix((vals, (rows, cols)), shape=(N, N)).tocsr()
    return A, (dims[2], dims[1], dims[0])


def expand_periodic(xu, ncell, periodic):
    """Unique-node array [k,j,i] -> full (n+1)^3 nodal array with periodic images duplicated."""
    x = xu
    for ax, d in ((2, 0), (1, 1), (0, 2)):
        if periodic[d]:
            x = np.concatenate([x, np.take(x, [0], axis=ax)], axis=ax)
    return x


def restrict_unique(xf, ncell, periodic):
    x = xf
    for ax, d in ((2, 0), (1, 1), (0, 2)):
        if periodic[d]:
            x = np.take(x, np.arange(ncell[d]), axis=ax)
    return x
