"""Device timing of the MAC projection (include/macproj.h) on one GPU: V-cycles, ms per V-cycle, launches.
   python tools/time_mac.py [--ncell 512] [--workload double_shear|rayleigh_taylor] [--reps 3]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from incflo_b200 import capi, problems  # noqa: E402
from incflo_b200.mac_projector import MacProjector  # noqa: E402
from incflo_b200.projector import Fab, Geometry, LPInfo  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--ncell", type=int, default=512)
ap.add_argument("--workload", default="double_shear")
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--verbose", type=int, default=0)
a = ap.parse_args()
torch.cuda.set_device(0)
n = (a.ncell,) * 3
hi = (1.0, 1.0, 1.0)
umac, beta, bclo, bchi = problems.mac_problem_torch(a.workload, n, hi, "cuda:0")
u0 = [u.clone() for u in umac]
P = MacProjector(Geometry(n, (0, 0, 0), hi, [b == 0 for b in bclo]), lib=capi.load())
P.setDomainBC(bclo, bchi)
P.initProjector(LPInfo().setMaxCoarseningLevel(100), [Fab(b, 0) for b in beta], verbose=a.verbose)
for r in range(a.reps):
    for u, s in zip(umac, u0):
        u.copy_(s)
    P.setUMAC([Fab(u, 0) for u in umac])
    st = P.project(1e-11, 1e-14)
    dof = n[0] * n[1] * n[2]
    print(f"{a.workload} {a.ncell}^3: vcycles {st.n_vcycles} levels {st.n_levels} bottom_iters {st.bottom_iters_total} "
          f"setup {st.ms_setup:.2f} solve {st.ms_solve:.2f} finish {st.ms_finish:.2f} ms; {st.ms_solve / max(st.n_vcycles, 1):.3f} ms/V-cycle "
          f"({dof * st.n_vcycles / (st.ms_solve * 1e-3):.3e} DOF/s per V-cycle) launches {st.kernel_launches} const_beta {st.sigma_is_const} "
          f"rel res {st.res_norm_final / max(st.rhs_norm0, 1e-300):.2e}", flush=True)
