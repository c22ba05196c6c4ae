"""Host-side mirror of incflo::ApplyProjection (/root/reference/src/projection/incflo_apply_nodal_projection.cpp:57-224)
over the fused C-ABI entry np_apply_projection.

    ld = LevelData(velocity, velocity_o, gp, p)                  # the members of incflo's LevelData the function touches
    apply_projection(proj, ld, density, time, scaling_factor, incremental, dt, prev_dt, rtol, atol)

`proj` is a NodalProjector built for the level's box layout (its `sigma` argument is not used by this path).  Names and
argument meaning follow the reference; the small-dt switch is the reference's (:65)."""
import ctypes as C
from dataclasses import dataclass

from . import capi
from .projector import Fab, NodalProjector


@dataclass
class LevelData:
    velocity: Fab       # cell, 3 comp, nghost_state() ghosts: u* in, u^{n+1} out
    velocity_o: Fab     # cell, 3 comp: u^n (read only when incremental or small dt); may be None otherwise
    gp: Fab             # cell, 3 comp: grad(p)
    p: Fab              # nodal, 1 comp: pressure


def apply_projection(proj: NodalProjector, ld: LevelData, density: Fab, time: float, scaling_factor: float,
                     incremental: bool, dt: float, prev_dt: float, rtol: float, atol: float, zero_vel_ghosts: bool = True,
                     stream=None):
    small = bool(time > 0.0 and dt < 0.1 * prev_dt)                      # proj_for_small_dt (:65)
    args = capi.np_apply_args(float(scaling_factor), int(bool(incremental)), int(small), int(bool(zero_vel_ghosts)))
    st = capi.np_stats()
    need_old = incremental or small
    vo = ld.velocity_o.c() if (need_old and ld.velocity_o is not None) else None
    v, r, g, p = ld.velocity.c(), density.c(), ld.gp.c(), ld.p.c()
    rc = proj.L.np_apply_projection(proj.h, C.byref(v), C.byref(vo) if vo is not None else None, C.byref(r), C.byref(g),
                                    C.byref(p), C.byref(args), rtol, atol, C.c_void_p(stream) if stream else None, C.byref(st))
    proj.stats = st
    capi.check(proj.L, rc)                                               # reference behaviour: amrex::Abort on MLMG failure
    return st
