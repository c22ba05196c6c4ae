"""How much of a projection's device time is idle between kernels?  CUPTI timeline (torch.profiler) of one np_project
at n^3: prints the span of the kernel timeline, the sum of the kernel durations and the gaps by size class.
`python tools/gap_check.py [n]` - a measurement aid, not part of the product."""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from incflo_b200 import capi, problems                                      # noqa: E402
from incflo_b200.projector import Fab, Geometry, LPInfo, NodalProjector   # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = "cuda:0"
torch.cuda.set_device(0)
u0 = problems.double_shear_torch((n, n, n), (1.0, 1.0, 1.0), 0, n, dev, ngrow=2)
vel = Fab(u0.clone(), 2)
sig = Fab(torch.ones((1, n, n, n), dtype=torch.float64, device=dev), 0)
P = NodalProjector(vel, sig, Geometry((n, n, n), (0, 0, 0), (1, 1, 1), [True] * 3), LPInfo(), lib=capi.load())
P.setDomainBC((0, 0, 0), (0, 0, 0))
P.project(1e-11, 1e-14)
vel.data.copy_(u0)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    st = P.project(1e-11, 1e-14)
    torch.cuda.synchronize()
ev = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "Memcpy" not in e.name and "Memset" not in e.name),
            key=lambda e: e.time_range.start)
span = ev[-1].time_range.end - ev[0].time_range.start
busy = sum(e.time_range.end - e.time_range.start for e in ev)
gaps = [b.time_range.start - a.time_range.end for a, b in zip(ev, ev[1:])]
print(f"{len(ev)} kernels, span {span / 1e3:.2f} ms, kernel time {busy / 1e3:.2f} ms, idle {100 * (span - busy) / span:.1f} %  "
      f"({st.n_vcycles} V-cycles, ms_total {st.ms_total:.2f})")
for lo, hi in ((0, 2), (2, 5), (5, 10), (10, 50), (50, 1e9)):
    g = [x for x in gaps if lo <= x < hi]
    print(f"  gaps {lo:>3}-{hi:<6g} us: {len(g):5d}  total {sum(g) / 1e3:7.3f} ms")
