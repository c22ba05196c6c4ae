"""Kernel logic vs the oracle, on CPU.

The product sources (incflo_b200/csrc) are compiled against a host emulation of the CUDA execution model
(tests/emu, TEST-ONLY) and called through the same C ABI as on the GPU.  This checks index maps, tile /
halo / pipeline schedules and the MG driver before GPU time is spent; the same comparisons run on the
B200 in tests/test_gpu_parity.py.  Bit-level equality is not expected (different summation order), so
every comparison carries an explicit tolerance relative to the field's size.
"""
import numpy as np
import pytest

import nphelp
from oracle.oracle import Oracle, TILED

BIG = 1 << 30
P, D, N = 0, 1, 2
BCS = [((P, P, P), (P, P, P)), ((N, N, N), (N, N, N)), ((P, N, D), (P, N, N)), ((D, P, N), (N, P, D))]


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def make(backend, ncell, bclo, bchi, sigma, prob_hi=(1.0, 1.0, 1.0), **opts):
    drv = nphelp.Driver(ncell, (0, 0, 0), prob_hi, bclo, bchi, **backend, **opts)
    drv.set_sigma(sigma)
    orc = Oracle(ncell, (0, 0, 0), prob_hi, bclo, bchi, smoother=TILED, tile=(63, 31, BIG), tile_z_target=-1)
    orc.set_sigma(sigma)
    orc.prepare_sigma()
    return drv, orc


def nodal(rng, shape, bclo, bchi):
    return nphelp.zero_dirichlet(nphelp.periodic_consistent(rng.standard_normal(shape), bclo), bclo, bchi)


@pytest.mark.parametrize("bclo,bchi", BCS)
@pytest.mark.parametrize("const_sigma", [False, True])
def test_apply_restrict_interp_sigma(backend, bclo, bchi, const_sigma):
    rng = np.random.default_rng(7)
    ncell = (16, 8, 12)
    sigma = np.full((12, 8, 16), 0.7) if const_sigma else rng.uniform(0.5, 2.0, (12, 8, 16))
    drv, orc = make(backend, ncell, bclo, bchi, sigma, prob_hi=(1.6, 0.4, 2.4))
    assert drv.P.num_levels == orc.num_levels
    for l in range(orc.num_levels):
        assert drv.P.level_ncell(l) == orc.level_ncell(l)
        assert relerr(drv.get_sigma(l), orc.get_sigma(l)) < 1e-14
        x = nodal(rng, orc.nshape(l), bclo, bchi)
        assert relerr(drv.apply(l, x), orc.apply(l, x)) < 1e-13
    for l in range(1, orc.num_levels):
        f = nodal(rng, orc.nshape(l - 1), bclo, bchi)
        c = nodal(rng, orc.nshape(l), bclo, bchi)
        assert relerr(drv.restrict(l, f), orc.restrict(l, f)) < 1e-14
        assert relerr(drv.interp_add(l - 1, f, c), orc.interp_add(l - 1, f, c)) < 1e-13


@pytest.mark.parametrize("bclo,bchi", BCS)
@pytest.mark.parametrize("ncell", [(136, 36, 24), (64, 16, 40)])
def test_weighted_interpolation_multi_tile(backend, bclo, bchi, ncell):
    """k_interp_add_w2 (sigma-weighted prolongation marching in z: 31 x 7 coarse cells per CTA, values of the neighbouring
    threads through shared memory, aligned pair stores) on levels with several tiles in x and y and several z-chunks, against
    the oracle's per-node definition (mlndlap_interpadd_aa) on every level pair."""
    rng = np.random.default_rng(33)
    sigma = rng.uniform(0.5, 2.0, ncell[::-1])
    drv, orc = make(backend, ncell, bclo, bchi, sigma, prob_hi=(1.7, 0.45, 0.6))
    for l in range(1, orc.num_levels):
        f = nodal(rng, orc.nshape(l - 1), bclo, bchi)
        c = nodal(rng, orc.nshape(l), bclo, bchi)
        assert relerr(drv.interp_add(l - 1, f, c), orc.interp_add(l - 1, f, c)) < 1e-13, l


@pytest.mark.parametrize("bclo,bchi", BCS)
@pytest.mark.parametrize("ncell", [(16, 8, 12), (136, 68, 8), (8, 8, 72)])
@pytest.mark.parametrize("const_sigma", [False, True])
def test_fused_residual_restriction(backend, bclo, bchi, ncell, const_sigma):
    """k_resid_restrict + k_restrict_faces (computeResOfCorrection + restriction in one pass, the fine residual stored on the
    boundary shell only) against oracle.apply + oracle.restrict, on every level pair; 136 x 68: three overlapped 62 x 30
    tiles in x and y (coarse nodes next to tile seams), 72 planes: several z-chunks."""
    rng = np.random.default_rng(21)
    sigma = np.full(ncell[::-1], 0.9) if const_sigma else rng.uniform(0.5, 2.0, ncell[::-1])
    drv, orc = make(backend, ncell, bclo, bchi, sigma, prob_hi=(1.7, 0.85, 0.9))
    for l in range(orc.num_levels - 1):
        cor = nodal(rng, orc.nshape(l), bclo, bchi)
        res = nodal(rng, orc.nshape(l), bclo, bchi)
        ref = orc.restrict(l + 1, nphelp.zero_dirichlet(res - orc.apply(l, cor), bclo, bchi))
        got = drv.resid_restrict(l, cor, res)
        assert relerr(got, ref) < 1e-13, l


@pytest.mark.parametrize("bclo,bchi", BCS)
@pytest.mark.parametrize("ncell", [(16, 16, 16), (70, 36, 8)])
@pytest.mark.parametrize("const_sigma", [False, True])
def test_smoother_matches_oracle_tiled_order(backend, bclo, bchi, ncell, const_sigma):
    """Same ordering (overlapped 63x31 tiles, k-sequential, 4 colours, 2 sweeps) -> same iterates."""
    rng = np.random.default_rng(11)
    sigma = np.full(ncell[::-1], 1.3) if const_sigma else rng.uniform(0.5, 2.0, ncell[::-1])
    drv, orc = make(backend, ncell, bclo, bchi, sigma, prob_hi=(1.0, 0.7, 0.4))
    x = nodal(rng, orc.nshape(0), bclo, bchi)
    r = nodal(rng, orc.nshape(0), bclo, bchi)
    assert relerr(drv.smooth(0, x, r, 2), orc.smooth(0, x, r, 2)) < 1e-13


@pytest.mark.parametrize("const_sigma", [False, True])
def test_smoother_z_chunks(backend, const_sigma):
    rng = np.random.default_rng(12)
    ncell = (8, 8, 96)                      # 97 planes -> several z-chunks
    bclo = bchi = (P, P, N)
    sigma = np.full(ncell[::-1], 0.8) if const_sigma else rng.uniform(0.5, 2.0, ncell[::-1])
    drv, orc = make(backend, ncell, bclo, bchi, sigma)
    x = nodal(rng, orc.nshape(0), bclo, bchi)
    r = nodal(rng, orc.nshape(0), bclo, bchi)
    assert relerr(drv.smooth(0, x, r, 1), orc.smooth(0, x, r, 1)) < 1e-13


@pytest.mark.parametrize("bclo,bchi", BCS)
def test_divu_mknewu_dot(backend, bclo, bchi):
    rng = np.random.default_rng(5)
    ncell = (8, 12, 16)
    sigma = rng.uniform(0.5, 2.0, ncell[::-1])
    drv, orc = make(backend, ncell, bclo, bchi, sigma, prob_hi=(0.8, 0.6, 3.2))
    u = rng.standard_normal((3,) + ncell[::-1])
    orc.set_vel(u)
    assert relerr(drv.divu(u), orc.divu()) < 1e-13
    phi = nodal(rng, orc.nshape(0), bclo, bchi)
    u2, g2 = drv.mknewu(phi, u)
    uo, go = orc.mknewu(phi, u)
    assert relerr(u2, uo) < 1e-13 and relerr(g2, go) < 1e-13
    y = nodal(rng, orc.nshape(0), bclo, bchi)
    assert abs(drv.dot(0, phi, y) - orc.dot(0, phi, y)) < 1e-11 * np.abs(phi).sum()


@pytest.mark.parametrize("bclo,bchi", BCS)
def test_bottom_solver_and_vcycle(backend, bclo, bchi):
    rng = np.random.default_rng(3)
    ncell = (16, 16, 32)
    sigma = rng.uniform(0.5, 2.0, ncell[::-1])
    drv, orc = make(backend, ncell, bclo, bchi, sigma)
    lb = orc.num_levels - 1
    b = nodal(rng, orc.nshape(lb), bclo, bchi)
    xg, itg = drv.bottom_solve(b)
    xo, ito = orc.bottom_solve(b)
    assert itg == ito
    assert relerr(xg, xo) < 1e-9
    r0 = nodal(rng, orc.nshape(0), bclo, bchi)
    assert relerr(drv.vcycle(r0), orc.vcycle(r0)) < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("variable_sigma", [False, True])
def test_smoother_and_operator_are_deterministic_on_gpu(variable_sigma):
    """Race proxy (compute-sanitizer is closed on this pool: profiles/r2_racecheck_closed.log): the smoother orders its colour
    steps with __syncwarp / named barriers and the one-pass pipeline overlaps two sweeps in one shared-memory ring; a missing
    barrier shows up as run-to-run differences.  20 repeats of two smooths and one operator application on a level with
    partial tiles (70 x 36 x 40 cells: several z-chunks, masked threads) must be bitwise identical, and equal the
    sequential host emulation of the same kernels to round-off (tests above)."""
    import ctypes as C
    import torch
    from incflo_b200 import capi
    torch.cuda.set_device(0)
    rng = np.random.default_rng(4)
    ncell = (70, 36, 40)
    drv = nphelp.Driver(ncell, (0, 0, 0), (1.0, 0.5, 0.6), (2, 0, 0), (2, 0, 0), lib=capi.load(), like="cuda", device="cuda:0")
    drv.set_sigma(rng.uniform(0.5, 2.0, ncell[::-1]) if variable_sigma else np.full(ncell[::-1], 1.25))
    x0, r = rng.standard_normal((41, 37, 71)), rng.standard_normal((41, 37, 71))
    ref_s = ref_a = None
    for rep in range(20):
        s = drv.smooth(0, x0, r, 2)
        a = drv.apply(0, x0)
        if ref_s is None:
            ref_s, ref_a = s, a
        assert np.array_equal(s, ref_s) and np.array_equal(a, ref_a), rep
