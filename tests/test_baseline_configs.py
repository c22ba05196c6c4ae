"""BASELINE.json configs[0], [1], [3] at (or near) their stated sizes: the CUDA path through the C ABI against the oracle.

  configs[0]  Taylor-Green vortices 64^3            (test_no_eb/benchmark.taylor_green_vortices, prob_init_fluid.cpp:150-160)
  configs[1]  double shear layer 256^3, sigma = 1   (test/benchmark.double_shear_layer:11-17, prob_init_fluid.cpp:470-475)
  configs[3]  Rayleigh-Taylor, variable sigma       (test_no_eb/benchmark.rayleigh_taylor:12-26, prob_init_fluid.cpp:305-321)
              at 128 x 128 x 256 (the file's 1:1:2 aspect; the oracle finishes it in seconds); 512^3 behind NP_TEST_512=1.

Two comparisons each (SURVEY.md 8(c) "parity definition"):
  * oracle with the SAME smoother ordering (TILED): same V-cycle and bottom-iteration counts, residual history,
    u / grad(phi) / phi-minus-mean within 1e-10 relative L2 at rtol 1e-11;
  * oracle with the REFERENCE's ordering (lexicographic Gauss-Seidel, AMReX-default 32^3 boxes) at a tolerance both
    reach (rtol 1e-13; 3e-12 at 256^3 and 1e-11 at 512^3, where the fp64 round-off floor of the residual is 1.2e-12 / ~5e-12):
    fields within 1e-10, V-cycle counts within +-1 (both counts are printed and, with NP_PARITY_LOG set, logged).
The oracle is the checker here, never the thing measured.  Nothing reads /root/reference at run time."""
import json
import os

import numpy as np
import pytest

from incflo_b200 import capi, problems
from incflo_b200.projector import Fab, Geometry, LPInfo, NodalProjector

pytestmark = pytest.mark.gpu

BIG = 1 << 30
TOL = 1e-10          # north_star: u and phi within 1e-10 relative L2 in fp64


def _l2(a, b):
    d, n = float(np.linalg.norm(a - b)), float(np.linalg.norm(b))
    return d / n if n > 0 else d             # both zero (a trivial projection): 0


def _gpu_project(prob, ncell, rtol, atol):
    import torch
    torch.cuda.set_device(0)
    dev = "cuda:0"
    vel = Fab.from_array(prob["vel"], 1, False, "torch", dev)
    sig = Fab.from_array(prob["sigma"], 0, False, "torch", dev)
    P = NodalProjector(vel, sig, Geometry(ncell, prob["prob_lo"], prob["prob_hi"], [b == 0 for b in prob["bclo"]]), LPInfo(),
                       lib=capi.load())
    P.setDomainBC(prob["bclo"], prob["bchi"])
    st = P.project(rtol, atol)
    out = dict(st=st, u=vel.numpy_valid(), gp=P.getGradPhi().numpy_valid(), phi=P.getPhi().numpy_valid()[0],
               hist=[st.res_history[i] for i in range(st.n_vcycles)])
    P.close()
    return out


def _oracle_project(prob, ncell, rtol, atol, **kw):
    from oracle.oracle import Oracle
    o = Oracle(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"], nthreads=os.cpu_count() or 1, **kw)
    o.set_sigma(prob["sigma"]); o.set_vel(prob["vel"])
    st = o.project(rtol, atol)
    out = dict(st=st, u=o.get_vel(), gp=o.get_gradphi(), phi=o.get_phi(), hist=[st.res_history[i] for i in range(st.n_vcycles)])
    o.close()
    return out


def _compare(g, o, what):
    eu, eg = _l2(g["u"], o["u"]), _l2(g["gp"], o["gp"])
    ep = _l2(g["phi"] - g["phi"].mean(), o["phi"] - o["phi"].mean())
    rec = dict(case=what, vcycles_gpu=g["st"].n_vcycles, vcycles_oracle=o["st"].n_vcycles,
               bottom_iters_gpu=g["st"].bottom_iters_total, bottom_iters_oracle=o["st"].bottom_iters_total,
               rel_l2_u=eu, rel_l2_gradphi=eg, rel_l2_phi_minus_mean=ep)
    print("PARITY " + json.dumps(rec))
    out = os.environ.get("NP_PARITY_LOG")
    if out:
        with open(out, "a") as f:
            f.write(json.dumps(rec) + "\n")
    return eu, eg, ep


# The reference's double shear layer (u = tanh(30 (1/4 - |y - 1/2|)), v = 0.05 sin 2 pi x, w = 0) is discretely divergence-free,
# so configs[1] as shipped is a TRIVIAL projection (rhs = 0, zero V-cycles: checked as such); the z-wave of configs[2]
# (w = 0.05 sin 2 pi x sin 2 pi z) makes it a real solve at the same size.
CONFIGS = [
    ("configs0_taylor_green_64", "taylor_green", (64, 64, 64), {}),
    ("configs1_double_shear_256_as_shipped", "double_shear", (256, 256, 256), {}),
    ("configs1_double_shear_256_zwave", "double_shear", (256, 256, 256), dict(zwave=True)),
    ("configs3_rayleigh_taylor_128x128x256", "rayleigh_taylor", (128, 128, 256), {}),
]


@pytest.mark.parametrize("tag,name,ncell,kw", CONFIGS)
def test_config_same_ordering(tag, name, ncell, kw):
    from oracle.oracle import TILED
    prob = problems.PROBLEMS[name](ncell, **kw)
    g = _gpu_project(prob, ncell, 1e-11, 1e-14)
    o = _oracle_project(prob, ncell, 1e-11, 1e-14, smoother=TILED, tile=(63, 31, BIG), tile_z_target=-1)
    eu, eg, ep = _compare(g, o, tag + " vs oracle TILED rtol 1e-11")
    assert g["st"].status == 0 and o["st"].status == 0
    assert g["st"].n_vcycles == o["st"].n_vcycles and g["st"].n_levels == o["st"].n_levels
    assert g["st"].bottom_iters_total == o["st"].bottom_iters_total
    for a, b in zip(g["hist"], o["hist"]):
        # the histories agree until round-off (different summation orders) is a visible part of the residual
        assert abs(a - b) <= 1e-4 * b + 1e-12 * o["st"].rhs_norm0
    assert eu <= TOL and eg <= TOL and ep <= TOL
    assert g["st"].sigma_is_const == (0 if name == "rayleigh_taylor" else 1)


@pytest.mark.parametrize("tag,name,ncell,kw", CONFIGS)
def test_config_vs_reference_ordering(tag, name, ncell, kw):
    from oracle.oracle import LEX, chop_boxes
    prob = problems.PROBLEMS[name](ncell, **kw)
    # measured on the B200: at 256^3 the relative residual stalls at 1.2e-12 (fp64 round-off of A phi grows like 1/h^2)
    rtol = 3e-12 if ncell[0] * ncell[1] * ncell[2] >= 256 ** 3 else 1e-13
    g = _gpu_project(prob, ncell, rtol, 0.0)
    o = _oracle_project(prob, ncell, rtol, 0.0, smoother=LEX, boxes=chop_boxes(ncell, 32))
    eu, eg, ep = _compare(g, o, tag + f" vs oracle LEX / 32^3 boxes rtol {rtol:g}")
    assert g["st"].status == 0 and o["st"].status == 0
    # different hierarchies (box-limited coarsening: 5 levels in 32^3 boxes, 7-9 in one box) and orderings: counts within +-1
    assert abs(g["st"].n_vcycles - o["st"].n_vcycles) <= 1
    assert eu <= TOL and eg <= TOL and ep <= TOL


@pytest.mark.skipif(os.environ.get("NP_TEST_512") != "1", reason="512^3 against the lexicographic oracle: ~1 min and ~14 GB of host "
                    "memory; run once per round with NP_TEST_512=1 (numbers kept under profiles/)")
@pytest.mark.parametrize("name", ["double_shear", "rayleigh_taylor"])
def test_512_vs_reference_ordering(name):
    """configs[2] (N=1) and configs[3] at the full 512^3 against the reference ordering."""
    from oracle.oracle import LEX, chop_boxes
    ncell = (512, 512, 512)
    kw = dict(zwave=True) if name == "double_shear" else dict(prob_hi=(0.5, 0.5, 0.5))
    prob = problems.PROBLEMS[name](ncell, **kw)
    g = _gpu_project(prob, ncell, 1e-11, 0.0)
    o = _oracle_project(prob, ncell, 1e-11, 0.0, smoother=LEX, boxes=chop_boxes(ncell, 32))
    eu, eg, ep = _compare(g, o, f"{name} 512^3 vs oracle LEX / 32^3 boxes rtol 1e-11")
    assert g["st"].status == 0 and o["st"].status == 0
    assert abs(g["st"].n_vcycles - o["st"].n_vcycles) <= 1
    assert eu <= TOL and eg <= TOL and ep <= TOL
