"""np_apply_projection (incflo::ApplyProjection with its pointwise passes fused into the solver)
against the numpy restatement of /root/reference/src/projection/incflo_apply_nodal_projection.cpp:57-224 around the
oracle's projector.  All three forms of the reference: plain, incremental, and the small-dt form (:65, :107-114, :155-162)."""
import numpy as np
import pytest

from incflo_b200 import problems
from incflo_b200.apply_projection import LevelData, apply_projection
from incflo_b200.projector import Fab, Geometry, LPInfo, NodalProjector
from oracle.apply_projection import apply_projection_ref
from oracle.oracle import TILED

BIG = 1 << 30
RTOL, ATOL = 1e-11, 1e-14


def l2(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(np.asarray(b)), 1e-300)


def _fields(name, ncell, kw, seed):
    prob = problems.PROBLEMS[name](ncell, **kw)
    nx, ny, nz = ncell
    rng = np.random.default_rng(seed)
    x = (np.arange(nx) + 0.5) / nx
    y = (np.arange(ny) + 0.5) / ny
    z = (np.arange(nz) + 0.5) / nz
    X, Y, Z = x[None, None, :], y[None, :, None], z[:, None, None]
    per_z = prob["bclo"][2] == 0
    fz = np.sin(2 * np.pi * Z) if per_z else np.cos(np.pi * Z)
    gp = np.stack([0.3 * np.cos(2 * np.pi * X) * np.sin(2 * np.pi * Y) * fz,
                   -0.2 * np.sin(4 * np.pi * X) * np.cos(2 * np.pi * Y) * fz,
                   0.1 * np.sin(2 * np.pi * X) * np.sin(2 * np.pi * Y) * (np.cos(2 * np.pi * Z) if per_z else np.sin(np.pi * Z))])
    gp = np.ascontiguousarray(np.broadcast_to(gp, (3, nz, ny, nx)))
    p = rng.standard_normal((nz + 1, ny + 1, nx + 1))
    vel_old = 0.5 * prob["vel"] + 0.01 * rng.standard_normal(prob["vel"].shape)
    scaling = 0.37
    if name == "rayleigh_taylor":
        rho = 1.0 / prob["sigma"]                       # problems.rayleigh_taylor returns sigma = dt/rho with dt = 1
    else:
        rho = np.full((nz, ny, nx), 1.25)
    return prob, prob["vel"], vel_old, rho, gp, p, scaling


CASES = [
    ("taylor_green", (32, 32, 32), {}, dict(incremental=False, time=0.0, dt=0.1, prev_dt=0.1)),
    ("taylor_green", (32, 32, 32), {}, dict(incremental=True, time=0.3, dt=0.1, prev_dt=0.1)),
    ("rayleigh_taylor", (16, 16, 32), {}, dict(incremental=False, time=0.5, dt=0.005, prev_dt=0.1)),     # small-dt form
    ("rayleigh_taylor", (16, 16, 32), {}, dict(incremental=False, time=0.0, dt=0.1, prev_dt=0.1)),
]


@pytest.mark.parametrize("name,ncell,kw,mode", CASES)
def test_apply_projection_matches_reference_restatement(backend, name, ncell, kw, mode):
    prob, vel, vel_old, rho, gp, p, scaling = _fields(name, ncell, kw, seed=3)
    like, dev = backend["like"], backend.get("device", "cuda")
    geom = Geometry(ncell, prob["prob_lo"], prob["prob_hi"], [b == 0 for b in prob["bclo"]])
    fv = Fab.from_array(vel, 2, False, like, dev)          # nghost_state() = 2 (MOL)
    # garbage in the ghost cells: ApplyProjection zeroes them (setBndry(0), :142)
    fv.data += 7.0
    vv = fv.valid()
    vv -= 7.0
    fo = Fab.from_array(vel_old, 2, False, like, dev)
    fr = Fab.from_array(rho, 1, False, like, dev)
    fg = Fab.from_array(gp, 0, False, like, dev)
    fp = Fab.from_array(p, 0, True, like, dev)
    P = NodalProjector(fv, Fab.zeros(ncell, 1, 0, False, like, dev), geom, LPInfo(), lib=backend["lib"])
    P.setDomainBC(prob["bclo"], prob["bchi"])
    st = apply_projection(P, LevelData(fv, fo, fg, fp), fr, mode["time"], scaling, mode["incremental"], mode["dt"],
                          mode["prev_dt"], RTOL, ATOL)
    u_ref, gp_ref, p_ref, so = apply_projection_ref(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"], vel, vel_old,
                                                    rho, gp, p, mode["time"], scaling, mode["incremental"], mode["dt"],
                                                    mode["prev_dt"], RTOL, ATOL, smoother=TILED, tile=(63, 31, BIG),
                                                    tile_z_target=-1)
    assert st.n_vcycles == so.n_vcycles
    assert l2(fv.numpy_valid(), u_ref) <= 1e-10
    assert l2(fg.numpy_valid(), gp_ref) <= 1e-10
    dp, dp_ref = fp.numpy_valid()[0] - (p if mode["incremental"] else 0.0), p_ref - (p if mode["incremental"] else 0.0)
    assert l2(dp - dp.mean(), dp_ref - dp_ref.mean()) <= 1e-10          # phi is defined up to a constant (singular system)
    P.close()
