"""Times incflo::ApplyProjection on one B200 at n^3 two ways: (a) the reference's structure - the caller's pointwise
passes (here torch element-wise kernels standing in for amrex::ParallelFor) around np_project - and (b) the fused
np_apply_projection.  `python tools/time_apply.py [n] [var]`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from incflo_b200 import capi, problems                                      # noqa: E402
from incflo_b200.apply_projection import LevelData, apply_projection       # noqa: E402
from incflo_b200.projector import Fab, Geometry, LPInfo, NodalProjector   # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    var = len(sys.argv) > 2 and sys.argv[2] == "var"
    dev = "cuda:0"
    torch.cuda.set_device(0)
    L = capi.load()
    ncell = (n, n, n)
    u0 = problems.double_shear_torch(ncell, (1.0, 1.0, 1.0), 0, n, dev, ngrow=2)
    vel = Fab(u0.clone(), 2)
    rho = Fab(torch.ones((1, n + 2, n + 2, n + 2), dtype=torch.float64, device=dev), 1)
    if var:
        z = ((torch.arange(n + 2, dtype=torch.float64, device=dev) - 0.5) / n)[:, None, None]
        rho.data[0] = 1.0 + 1.5 * (1.0 + torch.tanh((z - 0.5) / 0.01)).expand(n + 2, n + 2, n + 2)
    gp = Fab(torch.zeros((3, n, n, n), dtype=torch.float64, device=dev), 0)
    gp.data.normal_().mul_(1e-3)
    p = Fab(torch.zeros((1, n + 1, n + 1, n + 1), dtype=torch.float64, device=dev), 0, True)
    sig = Fab(torch.zeros((1, n, n, n), dtype=torch.float64, device=dev), 0)
    gp0 = gp.data.clone()
    P = NodalProjector(vel, sig, Geometry(ncell, (0, 0, 0), (1, 1, 1), [True] * 3), LPInfo(), lib=L)
    P.setDomainBC((0, 0, 0), (0, 0, 0))
    dt = 1e-3
    g = 2

    def unfused():
        rv = rho.data[:, 1:-1, 1:-1, 1:-1]
        vv = vel.data[:, g:-g, g:-g, g:-g]
        vv += gp.data * (dt / rv)                                   # :79-105
        torch.div(dt, rv, out=sig.data)                             # :116-131
        inner = vv.clone()                                          # setBndry(0), :142
        vel.data.zero_(); vel.data[:, g:-g, g:-g, g:-g] = inner
        st = P.project(1e-11, 1e-14)                                # :148-153
        gp.data.copy_(P.getGradPhi().valid())                       # :168-203
        p.data.copy_(P.getPhi().valid())
        return st

    def fused():
        return apply_projection(P, LevelData(vel, None, gp, p), rho, 0.0, dt, False, dt, dt, 1e-11, 1e-14)

    for name, fn in (("caller passes + np_project", unfused), ("np_apply_projection (fused)", fused)):
        ts = []
        for it in range(4):
            vel.data.copy_(u0); gp.data.copy_(gp0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record()
            st = fn()
            e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        print(f"{name:32s} {min(ts[1:]):8.2f} ms   ({st.n_vcycles} V-cycles, solve {st.ms_solve:.2f} ms, |u| {float(vel.valid().norm()):.10e})")


if __name__ == "__main__":
    main()
