"""TEST INFRASTRUCTURE ONLY - numpy restatement of incflo::ApplyProjection for one level
(/root/reference/src/projection/incflo_apply_nodal_projection.cpp:57-224), around the oracle's projector.
Arrays are valid-region numpy arrays: vel, vel_old, gp (3, nz, ny, nx); rho (nz, ny, nx); p (nz+1, ny+1, nx+1)."""
import numpy as np

from .oracle import Oracle


def apply_projection_ref(ncell, prob_lo, prob_hi, bclo, bchi, vel, vel_old, rho, gp, p, time, scaling_factor, incremental,
                         dt, prev_dt, rtol, atol, **oracle_kw):
    small = bool(time > 0.0 and dt < 0.1 * prev_dt)                       # :65
    u = np.array(vel, dtype=np.float64, copy=True)
    if not incremental:                                                   # :79-105  add (grad p / rho) back to u*
        u += gp * (scaling_factor / rho)[None]
    if small or incremental:                                              # :107-114
        u -= vel_old
    sigma = scaling_factor / rho                                          # :116-131
    o = Oracle(ncell, prob_lo, prob_hi, bclo, bchi, **oracle_kw)          # :148-152 (ghost cells of vel are zero: setBndry(0), :142)
    o.set_sigma(sigma)
    o.set_vel(u)
    st = o.project(rtol, atol)                                            # :153
    u = o.get_vel()
    if small or incremental:                                              # :155-162
        u = u + vel_old
    phi, gphi = o.get_phi(), o.get_gradphi()                              # :165-166
    o.close()
    if incremental:                                                       # :168-203
        gp_new, p_new = gp + gphi, p + phi
    else:
        gp_new, p_new = gphi.copy(), phi.copy()
    return u, gp_new, p_new, st
