"""Times the level-`l` smoother / operator kernels through the C-ABI hooks (np_smooth, np_apply) on one GPU.
An A/B tool for kernel work: `python tools/time_kernels.py [ncell] [mglev] [var]` prints ms per Fsmooth launch
(the hook runs fill_boundary + one k_gs2 launch per call; copies in/out are cancelled by differencing)."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from incflo_b200 import capi                              # noqa: E402
from incflo_b200.projector import Fab, Geometry, LPInfo, NodalProjector   # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    lev = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    var = len(sys.argv) > 3 and sys.argv[3] == "var"
    dev = "cuda"
    L = capi.load(os.environ.get("NP_LIB"))       # NP_LIB: an experimental build to compare against
    ncell = (n, n, n)
    vel = Fab.zeros(ncell, 3, 1, False, "torch", dev)
    sig = Fab.zeros(ncell, 1, 0, False, "torch", dev)
    sig.data.fill_(1.0)
    if var:
        sig.data.add_(torch.rand_like(sig.data))
    P = NodalProjector(vel, sig, Geometry(ncell, (0, 0, 0), (1, 1, 1), [True] * 3), LPInfo(), lib=L)
    P.setDomainBC((0, 0, 0), (0, 0, 0))
    capi.check(L, L.np_set_sigma(P.h, C.byref(sig.c())))
    nl = P.level_ncell(lev)
    x = Fab.zeros(nl, 1, 1, True, "torch", dev)
    r = Fab.zeros(nl, 1, 1, True, "torch", dev)
    data = os.environ.get("NP_DATA", "random")
    if data == "random":
        r.data.normal_()
        r.data.sub_(r.data.mean())
    elif data == "tiny":
        r.data.normal_()
        r.data.mul_(1e-300)
    elif data == "smooth":
        g = torch.linspace(0, 6.283185307179586, r.data.shape[-1], device=dev, dtype=torch.float64)
        r.data.copy_(torch.sin(g)[None, None, None, :] * torch.cos(g)[None, None, :, None] * torch.sin(2 * g)[None, :, None, None])

    def run(nc):
        x.data.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        capi.check(L, L.np_smooth(P.h, lev, C.byref(x.c()), C.byref(r.c()), nc))
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    if os.environ.get("NP_TIME_APPLY"):          # the operator kernel under `ncu --metrics gpu__time_duration.sum -k regex:k_stencil`
        y = Fab.zeros(nl, 1, 1, True, "torch", dev)
        for _ in range(4):
            capi.check(L, L.np_apply(P.h, lev, C.byref(x.c()), C.byref(y.c())))
        torch.cuda.synchronize()
        return
    run(2)
    t = {nc: min(run(nc) for _ in range(3)) for nc in (0, 1, 2, 3, 5, 21)}
    print("  hook call ms by ncalls:", {k: round(v, 4) for k, v in t.items()},
          " first: %.4f  second: %.4f  third: %.4f" % (t[1] - t[0], t[2] - t[1], t[3] - t[2]))
    a, b = t[1], t[21]
    print(f"level {lev} of {n}^3 ({'variable' if var else 'constant'} sigma): {(b - a) / 20:.4f} ms per Fsmooth "
          f"(fill_boundary + k_gs2), x norm {float(x.data.norm()):.6e}")


if __name__ == "__main__":
    main()
