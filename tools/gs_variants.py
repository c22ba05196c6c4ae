"""A/B of the smoother kernels (env NP_GS_VARIANT / NP_GSV_VARIANT, read by np_create) on one GPU:
`python tools/gs_variants.py [ncell] [const|var] [variants...]` times one Fsmooth (fill_boundary + the smoother launch) per variant
through the np_smooth hook and compares every variant's result with the first one's (max abs difference relative to max |x|)."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from incflo_b200 import capi                              # noqa: E402
from incflo_b200.projector import Fab, Geometry, LPInfo, NodalProjector   # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    var_sigma = len(sys.argv) > 2 and sys.argv[2] == "var"
    variants = [int(v) for v in sys.argv[3:]] or [0, 1, 2, 3, 4]
    dev = "cuda"
    L = capi.load(os.environ.get("NP_LIB"))       # NP_LIB: an experimental build
    ncell = (n, n, n)
    torch.manual_seed(0)
    r = Fab.zeros(ncell, 1, 1, True, "torch", dev)
    r.data.normal_()
    r.data.sub_(r.data.mean())
    x0 = Fab.zeros(ncell, 1, 1, True, "torch", dev)
    x0.data.normal_()
    ref = None
    for v in variants:
        os.environ["NP_GSV_VARIANT" if var_sigma else "NP_GS_VARIANT"] = str(v)
        L.np_cache_clear()
        vel = Fab.zeros(ncell, 3, 1, False, "torch", dev)
        sig = Fab.zeros(ncell, 1, 0, False, "torch", dev)
        sig.data.fill_(1.0)
        if var_sigma:
            g = torch.Generator(device=dev); g.manual_seed(1)
            sig.data.add_(torch.rand(sig.data.shape, generator=g, device=dev, dtype=torch.float64))
        P = NodalProjector(vel, sig, Geometry(ncell, (0, 0, 0), (1, 1, 1), [True] * 3), LPInfo(), lib=L)
        P.setDomainBC((0, 0, 0), (0, 0, 0))
        capi.check(L, L.np_set_sigma(P.h, C.byref(sig.c())))
        x = Fab.zeros(ncell, 1, 1, True, "torch", dev)

        def run(nc):
            x.data.copy_(x0.data)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            capi.check(L, L.np_smooth(P.h, 0, C.byref(x.c()), C.byref(r.c()), nc))
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1)

        run(2)
        t = {nc: min(run(nc) for _ in range(3)) for nc in (1, 11)}
        run(2)
        res = x.valid().clone()
        if ref is None:
            ref = res
        diff = float((res - ref).abs().max() / ref.abs().max())
        print(f"{'var' if var_sigma else 'const'} variant {v}: {(t[11] - t[1]) / 10:.4f} ms per Fsmooth (fill + smoother), "
              f"max rel diff vs first variant {diff:.3e}", flush=True)
        P.close()
        del P, vel, sig, x
    L.np_cache_clear()


if __name__ == "__main__":
    main()
