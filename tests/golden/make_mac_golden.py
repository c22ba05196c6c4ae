"""Generates tests/golden/mac_*.npz from the MAC ORACLE (oracle/mac_oracle.cpp; the reference itself cannot run here:
AMReX is absent).  Inputs are regenerated from incflo_b200.problems.MAC_PROBLEMS; each file holds the V-cycle count, the
bottom-iteration count, the residual history and strided samples of the projected face velocities and of phi.
Run:  python tests/golden/make_mac_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from incflo_b200 import problems          # noqa: E402
from oracle.mac import MacOracle          # noqa: E402

MAC_CASES = {
    "mac_double_shear_32x32x16": ("double_shear", (32, 32, 16), dict(prob_hi=(1.0, 1.0, 0.5))),
    "mac_rayleigh_taylor_16x16x32": ("rayleigh_taylor", (16, 16, 32), {}),
}


def run_mac(name, ncell, kw):
    prob = problems.MAC_PROBLEMS[name](ncell, **kw)
    o = MacOracle(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"])
    o.set_beta(*prob["beta"]); o.set_umac(*prob["umac"])
    st = o.project(1e-11, 1e-14)
    u = o.get_umac(); p = o.get_phi(); p = p - p.mean()
    return dict(vc=st.n_vcycles, bot=st.bottom_iters_total, nlev=st.n_levels, rhs_norm0=st.rhs_norm0,
                hist=np.array([st.res_history[i] for i in range(st.n_vcycles)]),
                u_s=u[0][::5, ::3, ::7].copy(), v_s=u[1][::5, ::3, ::7].copy(), w_s=u[2][::5, ::3, ::7].copy(),
                p_s=p[::5, ::3, ::7].copy())


if __name__ == "__main__":
    for tag, (name, ncell, kw) in MAC_CASES.items():
        d = run_mac(name, ncell, kw)
        np.savez_compressed(os.path.join(HERE, tag + ".npz"), **d)
        print(tag, "vc", d["vc"], "bottom iterations", d["bot"])
