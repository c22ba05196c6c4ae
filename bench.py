#!/usr/bin/env python
"""bench.py — nodal-projection throughput on B200 (BASELINE.json metric), one JSON line on rank 0.

A "step" is one whole nodal projection (np_project through the C ABI: sigma hierarchy, divergence,
MLMG solve to rtol 1e-11, velocity correction, grad(phi), phi) of one synthetic velocity field.

Workload (config.workload): BASELINE.json configs[2] — 512^3 cells PER GPU, z-slab sharded, periodic,
constant density, double-shear-layer field (+ z-wave so the slabs couple).  At N=1 this is the
"512^3 single-level nodal projection on 1 B200" the north_star target is quoted on.  --ncell 256 gives
configs[1].  Inputs (GBs) are far larger than the 126 MB L2, so no explicit L2 flush is needed.

  value   DOF/s per V-cycle = (unique fine-level nodal unknowns, all ranks) / (timed region / V-cycles
          run in it).  The timed region is K complete np_project calls with inputs resident in HBM
          (plus the device-to-device restore of u*, which the in-place projection needs each step).
  e2e     same metric through the same C-ABI call with HOST (pinned) fabs: u*, sigma host->device and
          u, phi, grad(phi) device->host inside the timed region, every step.
  roofline  the dominant kernel (level-0 smoother, 2 Gauss-Seidel sweeps per launch), timed live with
          CUDA events around each of its launches inside the timed region (np_options.profile_kernels).
  cpu_baseline  the oracle (CPU restatement of the reference algorithm, lexicographic GS, AMReX-default
          32^3 boxes, all host cores) on a bounded sample; N=1 only.
  --impl reference   times that CPU restatement alone (the reference's own build is unavailable: AMReX
          is not in /root/reference and there is no MPI).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "nodal_projection_dof_per_s_per_vcycle"
UNIT = "DOF/s per V-cycle"
RTOL, ATOL = 1e-11, 1e-14


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def read_traffic(const_sigma, key_const="k_gs2_const", key_var="k_gs2_var"):
    """ncu-measured DRAM bytes per launch of the dominant kernel, if a capture was summarised."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get(key_const if const_sigma else key_var)
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.time(), ln.strip()))

    def mark(self):
        """Start of the timed region: samples taken before this (warm-up, also under load) are only a fallback."""
        self.t_mark = time.time()

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        t_mark = getattr(self, "t_mark", 0.0)
        timed = [ln for (t, ln) in self.lines if t >= t_mark]
        window = "timed region"
        if not timed:                         # region shorter than the sampling period: use the warm-up samples
            timed, window = [ln for (_, ln) in self.lines][-5:], "warm-up steps just before the timed region"
        sm, smmax, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in timed:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smmax.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smmax), "power_w_max": max(pw),
                "samples": len(sm), "window": window, "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------
# CPU arm: the oracle (restatement of the reference's algorithm) on the host cores
# ---------------------------------------------------------------------------------------------------
def cpu_projection(n, reps, warmup):
    """Project the double-shear(+z-wave) field at n^3 with the oracle.  Reference-faithful settings:
    lexicographic GS box by box, AMReX's default max_grid_size 32, box-parallel over all host cores."""
    from incflo_b200 import problems
    from oracle.oracle import Oracle, LEX, chop_boxes
    cores = os.cpu_count() or 1
    ncell = (n, n, n)
    prob = problems.double_shear(ncell, zwave=True)
    times, vc = [], 0
    for it in range(warmup + reps):
        o = Oracle(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"],
                   boxes=chop_boxes(ncell, 32), smoother=LEX, nthreads=cores)
        o.set_sigma(prob["sigma"])
        o.set_vel(prob["vel"])
        t0 = time.perf_counter()
        st = o.project(RTOL, ATOL)
        dt = time.perf_counter() - t0
        o.close()
        if st.status != 0:
            raise RuntimeError("oracle did not converge")
        if it >= warmup:
            times.append(dt); vc += st.n_vcycles
    total = sum(times)
    return dict(value=(n ** 3) * vc / total, cores=cores, vcycles=vc // reps, seconds=total, ms_per_step=1e3 * total / reps,
                sample=f"double_shear+zwave {n}^3 periodic, one level, 32^3 boxes (AMReX default max_grid_size), "
                       f"lexicographic GS, rtol {RTOL:g}; {reps} projection(s), {total:.1f} s")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.cpu_ncell
    if n is None:       # the workload itself when the whole run stays within a few minutes, else a bounded sample of it
        n = 512 if (args.steps + min(args.warmup, 1)) * 26.0 <= 240.0 else 256
    r = cpu_projection(n, args.steps, min(args.warmup, 1))
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, args.gpus, sample=r["sample"]),
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference incflo+AMReX MPI build unavailable on this box (AMReX absent from /root/reference, no MPI); "
                    "this is oracle/, the CPU restatement of the same algorithm"}
    print(json.dumps(line), flush=True)


def workload_config(args, G, sample=None):
    n = args.ncell
    if args.workload == "rayleigh_taylor":
        wl = (f"BASELINE configs[3]: synthetic {n}^3 cells per GPU, {n}x{n}x{n * G} domain in {G} z-slab(s), periodic x/y, "
              f"slip walls (Neumann) in z, variable density ratio 4 (sigma = dt/rho), gravity kick + perturbation, rtol {RTOL:g} atol {ATOL:g}")
    else:
        wl = (f"BASELINE configs[2]: synthetic {n}^3 cells per GPU, {n}x{n}x{n * G} domain in {G} z-slab(s), "
              f"periodic, constant density (sigma=1), double_shear_layer + z-wave, rtol {RTOL:g} atol {ATOL:g}")
    c = {"workload": wl,
         "ncell_per_gpu": [n, n, n], "boxes": f"one {n}^3 box per GPU", "vel_ngrow": args.vel_ngrow,
         "l2": "inputs (>= 1 GB per array at 512^3) exceed the 126 MB L2; no explicit flush"}
    if sample:
        c["cpu_sample"] = sample
    return c


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
def _bind_to_gpu_numa_node(torch, local):
    """What `mpirun --bind-to numa` / a job script does for an MPI rank: run this process on the cores of the NUMA node its GPU
    hangs off, so that the pinned host fabs of the e2e leg (first touched by this process) live in that node's memory and
    the host<->device copies do not cross the socket interconnect.  Returns a description, or None when nothing was changed."""
    try:
        pr = torch.cuda.get_device_properties(local)
        bus = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            txt = f.read().strip()
        cpus = set()
        for part in txt.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0)
        mine = cpus & allowed
        if not mine or mine == allowed:
            return {"node": node, "bound": False, "cpus_allowed": len(allowed)}
        os.sched_setaffinity(0, mine)
        return {"node": node, "bound": True, "cpus": len(mine)}
    except Exception as e:                      # no sysfs / no such attribute: leave the placement to the launcher
        return {"bound": False, "why": str(e)[:80]}


class _Ctx:
    """Per-process state of the GPU arm: device, library, communicator."""

    def __init__(self):
        import ctypes as C
        import torch
        import torch.distributed as dist
        from incflo_b200 import capi
        self.C, self.torch, self.dist, self.capi = C, torch, dist, capi
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; incflo_b200 has no CPU fallback (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = f"cuda:{self.local}"
        self.numa = _bind_to_gpu_numa_node(torch, self.local) if not os.environ.get("NP_BENCH_NO_NUMA") else None
        self.L = capi.load()
        self.comm = None
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device(self.dev))
            idbuf = torch.zeros(128, dtype=torch.uint8)
            if self.rank == 0:
                raw = (C.c_ubyte * 128)()
                capi.check(self.L, self.L.np_nccl_unique_id(raw))
                idbuf = torch.tensor(list(raw), dtype=torch.uint8)
            idd = idbuf.to(self.dev)
            dist.broadcast(idd, 0)
            raw = (C.c_ubyte * 128)(*idd.cpu().tolist())
            cptr = C.c_void_p()
            capi.check(self.L, self.L.np_nccl_comm_create(C.byref(cptr), raw, self.rank, self.world))
            self.comm = cptr.value
        self.stream = torch.cuda.current_stream().cuda_stream

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def close(self):
        if self.world > 1:
            if self.comm:
                self.L.np_nccl_comm_destroy(self.C.c_void_p(self.comm))
            self.dist.destroy_process_group()


def _nodal_problem(ctx, workload, nloc, G, g, world=None, rank=None):
    """u* (pristine, with g ghost layers), sigma fab, BCs, geometry pieces of one rank's n^3 (or nloc) slab."""
    import numpy as np
    from incflo_b200 import problems
    from incflo_b200.projector import Fab
    torch, dev = ctx.torch, ctx.dev
    rank = ctx.rank if rank is None else rank
    nx, ny, nz = nloc
    ncell = (nx, ny, nz * G)
    prob_hi = (1.0, float(ny) / nx, float(nz * G) / nx)          # isotropic cells, x extent 1
    koff = rank * nz
    bclo = bchi = (0, 0, 0)
    if workload == "rayleigh_taylor":
        pr = problems.rayleigh_taylor(ncell, prob_hi=prob_hi, koff=koff, nz_local=nz)
        vel0 = torch.zeros((3, nz + 2 * g, ny + 2 * g, nx + 2 * g), dtype=torch.float64, device=dev)
        vel0[:, g:g + nz, g:g + ny, g:g + nx] = torch.from_numpy(pr["vel"]).to(dev)
        sig = Fab(torch.from_numpy(np.ascontiguousarray(pr["sigma"][None])).to(dev), 0)
        bclo, bchi = pr["bclo"], pr["bchi"]
    else:
        vel0 = problems.double_shear_torch(ncell, prob_hi, koff, nz, dev, ngrow=g)
        sig = Fab(torch.ones((1, nz, ny, nx), dtype=torch.float64, device=dev), 0)
    return dict(ncell=ncell, prob_hi=prob_hi, bclo=bclo, bchi=bchi, vel0=vel0, sig=sig)


def _make_projector(ctx, args, pb, nloc, G, vel, rank=None, nranks=None, comm="ctx", **extra):
    from incflo_b200.projector import Geometry, LPInfo, NodalProjector
    rank = ctx.rank if rank is None else rank
    nranks = ctx.world if nranks is None else nranks
    nx, ny, nz = nloc
    geom = Geometry(pb["ncell"], (0, 0, 0), pb["prob_hi"], [b == 0 for b in pb["bclo"]])
    boxes = [((0, 0, r * nz), (nx - 1, ny - 1, (r + 1) * nz - 1)) for r in range(G)]
    kw = dict(profile_kernels=1, smooth_zchunk=args.smooth_zchunk, verbose=args.verbose)
    if args.agg_min_local_nz > 0:
        kw["agg_min_local_nz"] = args.agg_min_local_nz
    if args.no_overlap:
        kw["overlap_halo"] = 0
    if args.overlap:
        kw["overlap_halo"] = 1
    if args.no_p2p:
        kw["p2p_halo"] = 0
    if args.no_graph:
        kw["use_graph"] = 0
    kw.update(extra)
    P = NodalProjector(vel, pb["sig"], geom, LPInfo(), boxes=boxes, rank=rank, nranks=nranks,
                       nccl_comm=ctx.comm if comm == "ctx" else comm, lib=ctx.L, **kw)
    P.setDomainBC(pb["bclo"], pb["bchi"])
    return P


def _measure_nodal(ctx, args, workload, n, steps, warmup, e2e_steps, ctor=False, clocks=None):
    """One workload: device-resident timed loop (value), host-fab loop (e2e), optional constructor cost.  All ranks call it."""
    import time as _t
    from incflo_b200.projector import Fab
    torch, C, capi, L, dev = ctx.torch, ctx.C, ctx.capi, ctx.L, ctx.dev
    G, g = ctx.world, args.vel_ngrow
    nloc = (n, n, n)
    pb = _nodal_problem(ctx, workload, nloc, G, g)
    vel0, sig = pb["vel0"], pb["sig"]
    vel = Fab(vel0.clone(), g)
    L.np_cache_clear()
    torch.cuda.synchronize()
    t0 = _t.perf_counter()
    P = _make_projector(ctx, args, pb, nloc, G, vel)
    torch.cuda.synchronize()
    ms_create_cold = 1e3 * (_t.perf_counter() - t0)
    dof = n * n * n * G

    def step():
        vel.data.copy_(vel0)
        return P.project(RTOL, ATOL, stream=ctx.stream)

    for _ in range(warmup):
        step()
    ctx.barrier()
    if clocks:
        clocks.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    vcycles = launches = replays = captures = 0
    ms_smooth = n_smooth = ms_solve = 0.0
    hist = None
    for _ in range(steps):
        st = step()
        vcycles += st.n_vcycles; launches += st.kernel_launches; replays += st.graph_replays; captures += st.graph_captures
        ms_smooth += st.ms_smooth0; n_smooth += st.n_smooth0; ms_solve += st.ms_solve
        hist = [st.res_history[i] / max(st.rhs_norm0, 1e-300) for i in range(st.n_vcycles)]
    e1.record()
    ctx.barrier()
    ms = ctx.max_over_ranks(e0.elapsed_time(e1))
    clk = clocks.stop() if clocks else None
    out = dict(dof=dof, ms=ms, vcycles=vcycles, launches=int(launches), ms_smooth=ms_smooth, n_smooth=n_smooth, ms_solve=ms_solve,
               hist=hist, const_sigma=bool(st.sigma_is_const), levels=st.n_levels, steps=steps, warmup=warmup, clocks=clk,
               graph=dict(replays=int(replays), captures=int(captures)),
               ms_create_cold=ms_create_cold)

    # ---- e2e: host (pinned) fabs through the same C-ABI call; H2D + D2H inside the timed region ----
    if e2e_steps > 0:
        hv0 = vel0.cpu()                                            # pristine host copy (pageable)
        hv = Fab(torch.empty_like(hv0).pin_memory(), g)
        hs = Fab(sig.data.cpu().pin_memory(), 0)
        hphi = Fab(torch.zeros((1, n + 1, n + 1, n + 1), dtype=torch.float64).pin_memory(), 0, nodal=True)
        hgp = Fab(torch.zeros((3, n, n, n), dtype=torch.float64).pin_memory(), 0)
        P.vel, P.sigma = hv, hs
        h2d = hv.data.numel() * 8 + hs.data.numel() * 8
        d2h = hv.data.numel() * 8 + hphi.data.numel() * 8 + hgp.data.numel() * 8
        tot_ms, tot_vc = 0.0, 0
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for it in range(1 + e2e_steps):
            hv.data.copy_(hv0)                                      # untimed host-side reset of the in/out buffer
            ctx.barrier()
            a.record()
            st2 = P.project(RTOL, ATOL, stream=ctx.stream)
            capi.check(L, L.np_copy_phi(P.h, 0, C.byref(hphi.c())))
            capi.check(L, L.np_copy_gradphi(P.h, 0, C.byref(hgp.c())))
            b.record()
            ctx.barrier()
            if it >= 1:
                tot_ms += ctx.max_over_ranks(a.elapsed_time(b)); tot_vc += st2.n_vcycles
        out["e2e"] = {"value": dof * tot_vc / (tot_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                      "steps": e2e_steps, "ms_per_step": tot_ms / e2e_steps}
        if ctor:
            # What incflo::ApplyProjection does every time step (reference :148-166): construct, setDomainBC, project, read phi
            # and grad(phi), destroy.  Wall clock around the whole sequence (cudaMalloc / IPC / all-gather are host-synchronous),
            # max over ranks; first with the handle cache cleared before every construction, then with it (the default).
            def full_call():
                Pn = _make_projector(ctx, args, pb, nloc, G, hv)
                Pn.sigma = hs
                stn = Pn.project(RTOL, ATOL, stream=ctx.stream)
                capi.check(L, L.np_copy_phi(Pn.h, 0, C.byref(hphi.c())))
                capi.check(L, L.np_copy_gradphi(Pn.h, 0, C.byref(hgp.c())))
                Pn.close()
                return stn.n_vcycles
            P.close()
            res = {}
            for tag, clear in (("cold", True), ("cached", False)):
                tt, vc = 0.0, 0
                for it in range(1 + e2e_steps):
                    hv.data.copy_(hv0)
                    if clear:
                        L.np_cache_clear()
                    ctx.barrier()
                    t0 = _t.perf_counter()
                    v = full_call()
                    torch.cuda.synchronize()
                    dt = ctx.max_over_ranks(1e3 * (_t.perf_counter() - t0))
                    if it >= 1:
                        tt += dt; vc += v
                res[tag] = {"value": dof * vc / (tt * 1e-3), "ms_per_step": tt / e2e_steps}
            # construction alone, device-synchronised wall clock
            L.np_cache_clear()
            tc = []
            for it in range(3):
                torch.cuda.synchronize(); ctx.barrier()
                t0 = _t.perf_counter()
                Pn = _make_projector(ctx, args, pb, nloc, G, hv)
                torch.cuda.synchronize()
                t1 = _t.perf_counter()
                Pn.close()
                torch.cuda.synchronize()
                tc.append((ctx.max_over_ranks(1e3 * (t1 - t0)), ctx.max_over_ranks(1e3 * (_t.perf_counter() - t1))))
            out["e2e_with_ctor"] = {"unit": UNIT, "value": res["cached"]["value"], "ms_per_step": res["cached"]["ms_per_step"],
                                    "value_cache_off": res["cold"]["value"], "ms_per_step_cache_off": res["cold"]["ms_per_step"],
                                    "ms_create_first": tc[0][0], "ms_create_cached": min(tc[1][0], tc[2][0]), "ms_destroy_cached": min(tc[1][1], tc[2][1]),
                                    "note": "np_create + np_set_bc + np_project + np_copy_phi/gradphi + np_destroy per step with host fabs "
                                            "(what incflo::ApplyProjection does every time step); wall clock, max over ranks; "
                                            "cache_off = np_cache_clear before every construction"}
            L.np_cache_clear()
            P = None
    if P is not None:
        P.close()
    L.np_cache_clear()
    del vel, vel0, sig, pb
    torch.cuda.empty_cache()
    return out


def _roofline(out, n, G, peak, peak_src):
    const_sigma = out["const_sigma"]
    nodes = (n + 1) ** 3
    per_node = 2 * (24 if const_sigma else 32)                  # 2 GS sweeps x (read cor, rhs[, sigma]; write cor)
    alg_bytes = per_node * nodes
    avg_ms = out["ms_smooth"] / max(out["n_smooth"], 1)
    achieved = alg_bytes / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
    compulsory = (24 if const_sigma else 32) * nodes            # what the fused 2-sweep pass must move
    bpd = 342 if const_sigma else 432
    roof = {"bound": "hbm", "kernel": f"k_gs2<{'const' if const_sigma else 'variable'} sigma> (fused 2-sweep Gauss-Seidel), level 0",
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": read_traffic(const_sigma) if n == 512 else None,
            "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes, "avg_launch_ms": avg_ms,
            "launches_timed": int(out["n_smooth"]), "share_of_step": out["ms_smooth"] / out["ms"],
            "fused_pass_bytes_per_launch": compulsory,
            "frac_fused_pass": compulsory / (avg_ms * 1e-3) / 1e9 / peak if avg_ms > 0 else 0.0,
            "vcycle_algorithmic": {"bytes_per_dof": bpd,
                                   "achieved_gbs": bpd * out["dof"] / G * out["vcycles"] / (out["ms_solve"] * 1e-3) / 1e9,
                                   "note": "SURVEY 8(d) accounting over the solve phase, per GPU"}}
    roof["vcycle_algorithmic"]["frac"] = roof["vcycle_algorithmic"]["achieved_gbs"] / peak
    roof["vcycle_algorithmic"]["frac_of_8TBs_datasheet"] = roof["vcycle_algorithmic"]["achieved_gbs"] / 8000.0
    return roof


def _multi_gpu_parity(ctx, args):
    """Hardware check of the sharded path before anything is timed (N > 1): a 128 x 128 x (64 N) problem of both workloads is
    projected on N z-slabs (peer-memory or NCCL halos, replicated coarse levels) and the same global problem on rank 0
    alone; u, grad(phi) and phi-minus-mean must agree to 1e-10 relative L2 (north_star tolerance; both solves run to a
    tolerance near the fp64 floor so that the comparison is not limited by where each stops)."""
    from incflo_b200.projector import Fab
    torch, dist, dev, G = ctx.torch, ctx.dist, ctx.dev, ctx.world
    nloc, g, rtol = (128, 128, 64), 1, 1e-12
    rec = {"problem": f"128x128x{64 * G} cells in {G} z-slabs vs the same problem on one GPU, rtol {rtol:g}", "tolerance": 1e-10,
           "halo": "nccl" if args.no_p2p else "p2p (CUDA IPC peer memory; falls back to NCCL if a rank cannot map its neighbours)"}
    worst = 0.0
    for workload in ("double_shear", "rayleigh_taylor"):
        pb = _nodal_problem(ctx, workload, nloc, G, g)
        vel = Fab(pb["vel0"].clone(), g)
        P = _make_projector(ctx, args, pb, nloc, G, vel, profile_kernels=0)
        st = P.project(rtol, 0.0, stream=ctx.stream, raise_on_fail=False)
        phi = P.getPhi().valid()[0].contiguous().clone()            # local nodes k = 0 .. nz (the top plane duplicates the next rank's first)
        gp = P.getGradPhi().valid().contiguous().clone()
        u = vel.valid().contiguous().clone()
        P.close()
        parts = {}
        for name, t in (("u", u), ("gp", gp), ("phi", phi[:-1].contiguous())):
            bufs = [torch.empty_like(t) for _ in range(G)] if ctx.rank == 0 else None
            dist.gather(t, bufs, dst=0)
            if ctx.rank == 0:
                parts[name] = torch.cat(bufs, dim=-3)
        hist_n = [st.res_history[i] / max(st.rhs_norm0, 1e-300) for i in range(st.n_vcycles)]
        w = {"vcycles_sharded": st.n_vcycles, "status_sharded": st.status}
        if ctx.rank == 0:
            pb1 = _nodal_problem(ctx, workload, (nloc[0], nloc[1], nloc[2] * G), 1, g, rank=0)
            vel1 = Fab(pb1["vel0"].clone(), g)
            P1 = _make_projector(ctx, args, pb1, (nloc[0], nloc[1], nloc[2] * G), 1, vel1, rank=0, nranks=1, comm=None, profile_kernels=0)
            st1 = P1.project(rtol, 0.0, stream=ctx.stream, raise_on_fail=False)
            phi1 = P1.getPhi().valid()[0][:-1]
            gp1 = P1.getGradPhi().valid()
            u1 = vel1.valid()

            def l2(a, b):
                return float((a - b).norm() / b.norm())
            e = {"u": l2(parts["u"], u1), "gradphi": l2(parts["gp"], gp1),
                 "phi_minus_mean": l2(parts["phi"] - parts["phi"].mean(), phi1 - phi1.mean())}
            hist_1 = [st1.res_history[i] / max(st1.rhs_norm0, 1e-300) for i in range(st1.n_vcycles)]
            w.update(vcycles_one_gpu=st1.n_vcycles, status_one_gpu=st1.status, rel_l2=e,
                     rel_res_history_sharded=hist_n, rel_res_history_one_gpu=hist_1)
            worst = max(worst, max(e.values()))
            if st.status != 0 or st1.status != 0:
                worst = float("inf")
            P1.close()
        rec[workload] = w
        ctx.L.np_cache_clear()
    rec["max_rel_l2"] = worst
    ok = torch.tensor([1.0 if (ctx.rank != 0 or worst <= 1e-10) else 0.0], device=dev)
    dist.broadcast(ok, 0)
    rec["pass"] = bool(ok.item() > 0)
    return rec


def run_gpu(args):
    ctx = _Ctx()
    rank, world = ctx.rank, ctx.world
    n, G = args.ncell, world
    parity = None
    if world > 1 and not args.no_parity:
        parity = _multi_gpu_parity(ctx, args)
        if not parity["pass"]:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "error": "multi-GPU parity check failed", "multi_gpu_parity": parity}), flush=True)
            ctx.close()
            raise SystemExit(3)
    clocks = ClockSampler(ctx.local) if rank == 0 else None
    if clocks:
        clocks.start()                        # nvidia-smi needs a moment to come up: start it before the warm-up
    out = _measure_nodal(ctx, args, args.workload, n, args.steps, args.warmup, 0 if args.no_e2e else args.e2e_steps,
                         ctor=not args.no_e2e, clocks=clocks)
    # the other BASELINE configs as sub-records of the same line (N = 1 only): configs[3] Rayleigh-Taylor 512^3 (variable sigma)
    # and configs[1] 256^3; `value` / `config` stay configs[2]
    extra = {}
    if world == 1 and not args.no_workloads and args.workload == "double_shear" and n == 512:
        for tag, wl, nn, stp, wu in (("rayleigh_taylor_512", "rayleigh_taylor", 512, 2, 3), ("double_shear_256", "double_shear", 256, 3, 3)):
            extra[tag] = (nn, _measure_nodal(ctx, args, wl, nn, stp, wu, 0))
    if rank == 0:
        peak, peak_src = read_peaks()
        ms, vcycles = out["ms"], out["vcycles"]
        line = {"metric": METRIC, "value": out["dof"] * vcycles / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, G),
                "vcycles_per_step": vcycles / args.steps, "mg_levels": out["levels"], "ms_per_vcycle": ms / vcycles,
                "ms_solve_per_vcycle": out["ms_solve"] / vcycles, "rel_res_history": out["hist"],
                "clocks": out["clocks"], "e2e": out.get("e2e"), "gpu_launches": out["launches"],
                "launches_per_vcycle": out["launches"] / vcycles, "host_numa_binding": ctx.numa,
                "cuda_graph": dict(out["graph"], note="MLMG iterations of the timed region replayed as one CUDA-graph launch each (np_options.use_graph, "
                                                      "one rank only); gpu_launches counts the kernels inside the graphs too"),
                "roofline": _roofline(out, n, G, peak, peak_src)}
        if "e2e_with_ctor" in out:
            line["e2e_with_ctor"] = out["e2e_with_ctor"]
        if parity:
            line["multi_gpu_parity"] = parity
        if extra:
            line["workloads"] = {}
            for tag, (nn, o) in extra.items():
                r = _roofline(o, nn, 1, peak, peak_src)
                line["workloads"][tag] = {
                    "value": o["dof"] * o["vcycles"] / (o["ms"] * 1e-3), "unit": UNIT, "steps": o["steps"], "warmup": o["warmup"],
                    "vcycles": o["vcycles"] / o["steps"], "ms_per_step": o["ms"] / o["steps"], "ms_per_vcycle": o["ms"] / o["vcycles"],
                    "ms_solve_per_vcycle": o["ms_solve"] / o["vcycles"], "mg_levels": o["levels"], "rel_res_history": o["hist"],
                    "launches_per_vcycle": o["launches"] / o["vcycles"], "cuda_graph": o["graph"],
                    "roofline": {k: r[k] for k in ("kernel", "frac", "frac_fused_pass", "avg_launch_ms", "share_of_step", "vcycle_algorithmic")}}
        if world == 1 and not args.no_cpu:
            r = cpu_projection(args.cpu_ncell or 512, 1, 0)      # the workload itself: one 512^3 projection, ~25 s on 16 cores
            line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                                    "sample": r["sample"], "vcycles": r["vcycles"]}
        print(json.dumps(line), flush=True)
    ctx.close()


# ---------------------------------------------------------------------------------------------------
# --path mac: the MAC projection measured to the same contract.  One GPU this round.
# Algorithmic bytes per cell and V-cycle, SURVEY 8(d) rules (each array once per kernel, reference schedule 2+2 smooths x
# 2 colours): per level cor = 0 (8) + 8 colour sweeps x (cor r/w, rhs[, 3 face coefficients]) + residual of the correction
# + restriction (9) + interpolation (17), x 8/7 for the coarse levels, + level 0's sol += cor (24), residual, norm (8).
# ---------------------------------------------------------------------------------------------------
MAC_METRIC = "mac_projection_dof_per_s_per_vcycle"
MAC_BYTES = {True: round((8 + 8 * 24 + 24 + 9 + 17) * 8 / 7) + 24 + 24 + 8, False: round((8 + 8 * 48 + 48 + 9 + 17) * 8 / 7) + 24 + 48 + 8}


def mac_config(args, sample=None, G=1):
    n = args.ncell
    if args.workload == "rayleigh_taylor":
        wl = (f"MAC projection, synthetic {n}^3 cells, periodic x/y, walls (Neumann) in z, beta = 1/rho on faces with the "
              f"Rayleigh-Taylor density (ratio 4), gravity kick + perturbation on the face velocities, rtol {RTOL:g} atol {ATOL:g}")
    else:
        wl = (f"MAC projection, synthetic {n}^3 cells, periodic, constant density (beta = 1), double_shear_layer + z-wave + "
              f"perturbation sampled on faces, rtol {RTOL:g} atol {ATOL:g}")
    if G > 1:
        wl += f"; weak scaling: {n}x{n}x{n * G} domain in {G} z-slabs, one per GPU (halo planes, norms and coarse-level gather over NCCL)"
    c = {"workload": wl, "ncell_per_gpu": [n, n, n], "boxes": f"one {n}^3 box per GPU",
         "l2": "inputs (>= 1 GB per array at 512^3) exceed the 126 MB L2; no explicit flush"}
    if sample:
        c["cpu_sample"] = sample
    return c


def cpu_mac_projection(n, workload, reps):
    """The MAC oracle (CPU restatement of MacProjector / MLMG / MLABecLaplacian; red-black GS threaded over planes)."""
    from incflo_b200 import problems
    from oracle.mac import MacOracle
    cores = os.cpu_count() or 1
    ncell = (n, n, n)
    hi = (1.0, 1.0, 1.0)
    prob = problems.MAC_PROBLEMS[workload](ncell, prob_hi=hi)
    total, vc = 0.0, 0
    for _ in range(reps):
        o = MacOracle(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"], nthreads=cores)
        o.set_beta(*prob["beta"]); o.set_umac(*prob["umac"])
        t0 = time.perf_counter()
        st = o.project(RTOL, ATOL)
        total += time.perf_counter() - t0
        vc += st.n_vcycles
        if st.status != 0:
            raise RuntimeError("MAC oracle did not converge")
    return dict(value=(n ** 3) * vc / total, cores=cores, vcycles=vc // reps, ms_per_step=1e3 * total / reps,
                sample=f"MAC {workload} {n}^3, one box, red-black GS threaded over z-planes on all host cores, rtol {RTOL:g}; "
                       f"{reps} projection(s), {total:.1f} s")


def run_reference_mac(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    r = cpu_mac_projection(args.cpu_ncell, args.workload, args.steps)
    line = {"impl": "reference", "metric": MAC_METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": 0, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": mac_config(args, r["sample"]),
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
            "note": "reference incflo+AMReX build unavailable on this box; this is oracle/mac_oracle.cpp"}
    print(json.dumps(line), flush=True)


def run_gpu_mac(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    from incflo_b200 import capi, problems
    from incflo_b200.mac_projector import MacProjector
    from incflo_b200.projector import Fab, Geometry, LPInfo
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; incflo_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    L = capi.load()
    comm = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(dev))
        idbuf = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            raw = (C.c_ubyte * 128)()
            capi.check(L, L.np_nccl_unique_id(raw))
            idbuf = torch.tensor(list(raw), dtype=torch.uint8)
        idd = idbuf.to(dev)
        dist.broadcast(idd, 0)
        raw = (C.c_ubyte * 128)(*idd.cpu().tolist())
        cptr = C.c_void_p()
        capi.check(L, L.np_nccl_comm_create(C.byref(cptr), raw, rank, world))
        comm = cptr.value
    n, G = args.ncell, world
    ncell = (n, n, n * G)                        # weak scaling: one n^3 slab per GPU, stacked in z
    hi = (1.0, 1.0, float(G))
    umac0, beta, bclo, bchi = problems.mac_problem_torch(args.workload, ncell, hi, dev, koff=rank * n, nz_local=n)
    umac = [Fab(u.clone(), 0) for u in umac0]
    P = MacProjector(Geometry(ncell, (0, 0, 0), hi, [b == 0 for b in bclo]), lib=L, rank=rank, nranks=world, nccl_comm=comm)
    P.setDomainBC(bclo, bchi)
    P.initProjector(LPInfo().setMaxCoarseningLevel(100), [Fab(b, 0) for b in beta], profile_kernels=1, verbose=args.verbose,
                    use_graph=0 if args.no_graph else 1, p2p_halo=0 if args.no_p2p else 1)
    P.setUMAC(umac)
    stream = torch.cuda.current_stream().cuda_stream
    dof = n ** 3 * G

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        for f, u0 in zip(umac, umac0):
            f.data.copy_(u0)
        return P.project(RTOL, ATOL, stream=stream)

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    for _ in range(args.warmup):
        step()
    barrier()
    clocks.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    vcycles = launches = n_smooth = 0
    ms_smooth = ms_solve = 0.0
    for _ in range(args.steps):
        st = step()
        vcycles += st.n_vcycles; launches += st.kernel_launches
        ms_smooth += st.ms_smooth0; n_smooth += st.n_smooth0; ms_solve += st.ms_solve
        hist = [st.res_history[i] / max(st.rhs_norm0, 1e-300) for i in range(st.n_vcycles)]
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop() if rank == 0 else None
    tmax = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())
    const_beta = bool(st.sigma_is_const)

    # e2e: host (pinned) face fabs through the same call; umac H2D + D2H every step, beta H2D through updateBeta, phi D2H
    e2e = None
    if not args.no_e2e:
        hu0 = [u.cpu() for u in umac0]
        hu = [Fab(torch.empty_like(u).pin_memory(), 0) for u in hu0]
        hb = [Fab(b.cpu().pin_memory(), 0) for b in beta]
        hphi = Fab(torch.zeros((1, n, n, n), dtype=torch.float64).pin_memory(), 0)
        P.setUMAC(hu)
        h2d = sum(f.data.numel() * 8 for f in hu) + sum(f.data.numel() * 8 for f in hb)
        d2h = sum(f.data.numel() * 8 for f in hu) + hphi.data.numel() * 8
        tot_ms, tot_vc = 0.0, 0
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for it in range(1 + args.e2e_steps):
            for f, u0 in zip(hu, hu0):
                f.data.copy_(u0)
            barrier()
            a.record()
            P.updateBeta(hb)
            st2 = P.project(RTOL, ATOL, stream=stream)
            P.copyPhi(hphi)
            b.record()
            barrier()
            if it >= 1:
                tm = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device=dev)
                if world > 1:
                    dist.all_reduce(tm, op=dist.ReduceOp.MAX)
                tot_ms += float(tm.item()); tot_vc += st2.n_vcycles
        e2e = {"value": dof * tot_vc / (tot_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "steps": args.e2e_steps, "ms_per_step": tot_ms / args.e2e_steps}

    if rank != 0:
        P.close()
        if comm:
            L.np_nccl_comm_destroy(C.c_void_p(comm))
        dist.destroy_process_group()
        return
    dof_gpu = n ** 3                              # roofline figures are per GPU
    peak, peak_src = read_peaks()
    per_cell = 2 * (24 if const_beta else 48)            # one smooth = red + black sweep: cor r/w, rhs[, bx, by, bz] each
    avg_ms = ms_smooth / max(n_smooth, 1)
    achieved = per_cell * dof_gpu / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
    # slabs made of whole 64 x 16 tiles (every benchmark size) run the one-pass smoother on every rank count; the timed region
    # of a smooth then also holds the two-plane exchange of the iterate that precedes the launch on a sharded level
    whole_tiles = n % 64 == 0 and n % 16 == 0
    kb = 'const' if const_beta else 'variable'
    if whole_tiles:
        kname = f"k_mac_rb3<{kb} beta> (red + black sweep of one smooth in one pass), level 0" + ("" if world == 1 else " + the two-plane halo exchange before it")
    else:
        kname = f"k_mac_gsrb<{kb} beta> x 2 (one launch per colour" + (")" if world == 1 else ", a halo-plane exchange before each)") + ", level 0"
    roof = {"bound": "hbm", "kernel": kname,
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": read_traffic(const_beta, "k_mac_rb3_const", "k_mac_rb3_var"), "peak_source": peak_src,
            "fused_pass_bytes_per_launch": (24 if const_beta else 48) * dof_gpu,
            "frac_fused_pass": (24 if const_beta else 48) * dof_gpu / (avg_ms * 1e-3) / 1e9 / peak if avg_ms > 0 else 0.0,
            "note": "achieved / frac use the SURVEY 8(d) accounting (every colour sweep of the reference schedule is a pass); the kernel "
                    "does both sweeps in one pass, so frac_fused_pass (its compulsory bytes / time / peak) is the figure that says how far "
                    "the kernel itself is from the HBM roofline",
            "algorithmic_bytes_per_launch": per_cell * dof_gpu, "avg_launch_ms": avg_ms, "launches_timed": int(n_smooth),
            "share_of_step": ms_smooth / ms,
            "vcycle_algorithmic": {"bytes_per_dof": MAC_BYTES[const_beta],
                                   "achieved_gbs": MAC_BYTES[const_beta] * dof_gpu * vcycles / (ms_solve * 1e-3) / 1e9}}
    roof["vcycle_algorithmic"]["frac"] = roof["vcycle_algorithmic"]["achieved_gbs"] / peak
    line = {"metric": MAC_METRIC, "value": dof * vcycles / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": mac_config(args, G=world), "vcycles_per_step": vcycles / args.steps,
            "mg_levels": st.n_levels, "ms_per_vcycle": ms / vcycles, "ms_solve_per_vcycle": ms_solve / vcycles,
            "rel_res_history": hist, "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof}
    if world == 1 and not args.no_cpu:
        r = cpu_mac_projection(args.cpu_ncell, args.workload, 1)
        line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"],
                                "vcycles": r["vcycles"]}
    print(json.dumps(line), flush=True)
    P.close()
    if world > 1:
        if comm:
            L.np_nccl_comm_destroy(C.c_void_p(comm))
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------
# --path diffusion: the implicit scalar diffusion solve, (rho - dt div(eta grad)) s = rho s_old, one
# GPU.  Workload: Rayleigh-Taylor density (ratio 4), eta = mu0 (1 + tanh profile) on faces (variable) or uniform, periodic x / y,
# walls in z (Neumann), tracer = a blob + the RT perturbation.  Bytes per cell and V-cycle, SURVEY 8(d) rules on the reference
# schedule (2 + 2 smooths x 2 colours; each colour sweep reads cor, rhs, a[, 3 face arrays] and writes cor).
# ---------------------------------------------------------------------------------------------------
DIFF_METRIC = "scalar_diffusion_dof_per_s_per_vcycle"
DIFF_BYTES = {True: round((8 + 8 * 32 + 32 + 9 + 17) * 8 / 7) + 24 + 32 + 8, False: round((8 + 8 * 56 + 56 + 9 + 17) * 8 / 7) + 24 + 56 + 8}


def _diffusion_problem(n, dev, variable_eta, torch):
    from incflo_b200 import problems
    return problems.diffusion_problem_torch(n, dev, variable_eta)


def run_gpu_diffusion(args):
    import ctypes as C
    import torch
    from incflo_b200 import capi
    from incflo_b200.diffusion import DiffusionScalarOp
    from incflo_b200.projector import Fab, Geometry
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        raise SystemExit("bench.py --path diffusion: one GPU this round")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; incflo_b200 has no CPU fallback")
    torch.cuda.set_device(0)
    dev, n, dt = "cuda:0", args.ncell, 0.01
    variable = args.workload == "rayleigh_taylor"
    rho, tr0, faces = _diffusion_problem(n, dev, variable, torch)
    bclo = bchi = (0, 0, 2)
    op = DiffusionScalarOp(Geometry((n, n, n), (0, 0, 0), (1, 1, 1), [True, True, False]), bclo, bchi, lib=capi.load(), like="torch", device=dev,
                           profile_kernels=1, use_graph=0 if args.no_graph else 1)
    frho, feta = Fab(rho, 0), [[Fab(f, 0) for f in faces]]
    ftr = Fab(tr0.clone(), 0)
    stream = torch.cuda.current_stream().cuda_stream
    dof = n ** 3

    def step():
        ftr.data.copy_(tr0)
        return op.diffuse_scalar(ftr, frho, feta, dt, RTOL, ATOL, stream=stream)[0]

    clocks = ClockSampler(0); clocks.start()
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize(); clocks.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    vcycles = launches = n_smooth = 0
    ms_smooth = ms_solve = 0.0
    for _ in range(args.steps):
        st = step()
        vcycles += st.n_vcycles; launches += st.kernel_launches; ms_solve += st.ms_solve
        ms_smooth += st.ms_smooth0; n_smooth += st.n_smooth0
        hist = [st.res_history[i] / max(st.rhs_norm0, 1e-300) for i in range(st.n_vcycles)]
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop()
    const_b = bool(st.sigma_is_const)
    # e2e: host (pinned) fabs: tracer up and down, rho and the three face arrays up, every step
    hrho, htr0 = Fab(rho.cpu().pin_memory(), 0), tr0.cpu()
    htr = Fab(torch.empty_like(htr0).pin_memory(), 0)
    heta = [[Fab(f.cpu().pin_memory(), 0) for f in faces]]
    h2d = htr.data.numel() * 8 * 2 + hrho.data.numel() * 8 * 2 + sum(f.data.numel() * 8 for f in heta[0])
    d2h = htr.data.numel() * 8
    tot_ms, tot_vc = 0.0, 0
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(1 + args.e2e_steps):
        htr.data.copy_(htr0)
        torch.cuda.synchronize()
        a.record()
        st2 = op.diffuse_scalar(htr, hrho, heta, dt, RTOL, ATOL, stream=stream)[0]
        b.record(); torch.cuda.synchronize()
        if it >= 1:
            tot_ms += a.elapsed_time(b); tot_vc += st2.n_vcycles
    e2e = {"value": dof * tot_vc / (tot_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "steps": args.e2e_steps, "ms_per_step": tot_ms / args.e2e_steps}
    peak, peak_src = read_peaks()
    per_cell = 2 * (32 if const_b else 56)
    avg_ms = ms_smooth / max(n_smooth, 1)
    achieved = per_cell * dof / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
    one_pass = n % 64 == 0 and n >= 512                # MacSolver::fused_ok with alpha * a: whole 64 x 16 tiles and a level of >= 512 cells in x
    kname = (f"k_mac_rb3<{'const' if const_b else 'variable'} b, alpha a> (red + black sweep of one smooth in one pass), level 0" if one_pass else
             f"k_mac_gsrb<{'const' if const_b else 'variable'} b> x 2 (red + black sweep of one smooth, one launch per colour, alpha a term), level 0")
    fused_bytes = (32 if const_b else 56) * dof
    roof = {"bound": "hbm", "kernel": kname,
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
            "fused_pass_bytes_per_launch": fused_bytes,
            "frac_fused_pass": (fused_bytes / (avg_ms * 1e-3) / 1e9 / peak if avg_ms > 0 else 0.0) if one_pass else None,
            "note": "achieved / frac use the SURVEY 8(d) accounting (every colour sweep of the reference schedule is a pass); where the one-pass "
                    "kernel runs, frac_fused_pass (its compulsory bytes / time / peak) says how far the kernel itself is from the HBM roofline",
            "algorithmic_bytes_per_launch": per_cell * dof, "avg_launch_ms": avg_ms, "launches_timed": int(n_smooth),
            "share_of_step": ms_smooth / ms,
            "vcycle_algorithmic": {"bytes_per_dof": DIFF_BYTES[const_b], "achieved_gbs": DIFF_BYTES[const_b] * dof * vcycles / (ms_solve * 1e-3) / 1e9}}
    roof["vcycle_algorithmic"]["frac"] = roof["vcycle_algorithmic"]["achieved_gbs"] / peak
    wl = (f"scalar diffusion (rho - dt div(eta grad)) s = rho s_old, {n}^3 cells, periodic x/y, walls (Neumann) in z, Rayleigh-Taylor density "
          f"(ratio 4), eta {'varying in z' if variable else 'uniform'}, dt {dt}, rtol {RTOL:g} atol {ATOL:g}")
    line = {"metric": DIFF_METRIC, "value": dof * vcycles / (ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl, "ncell_per_gpu": [n, n, n], "boxes": f"one {n}^3 box", "l2": "inputs exceed the 126 MB L2; no explicit flush"},
            "vcycles_per_step": vcycles / args.steps, "mg_levels": st.n_levels, "ms_per_vcycle": ms / vcycles, "ms_solve_per_vcycle": ms_solve / vcycles,
            "rel_res_history": hist, "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof}
    if not args.no_cpu:
        line["cpu_baseline"] = cpu_diffusion(args.cpu_ncell, variable, dt)
    print(json.dumps(line), flush=True)
    op.close()


def cpu_diffusion(n, variable, dt):
    """The MAC oracle's general MLABecLaplacian solve (oracle/mac_oracle.cpp: solve_abec) on the host cores, same problem family."""
    import torch
    from oracle.mac import MacOracle
    rho, tr0, faces = _diffusion_problem(n, "cpu", variable, torch)
    rho, tr0 = rho[0].numpy(), tr0[0].numpy()
    cores = os.cpu_count() or 1
    o = MacOracle((n, n, n), (0, 0, 0), (1, 1, 1), (0, 0, 2), (0, 0, 2), nthreads=cores)
    o.set_scalars(1.0, dt); o.set_acoef(rho); o.set_beta(*[f[0].numpy() for f in faces]); o.set_phi(tr0)
    t0 = time.perf_counter()
    st = o.solve_abec(rho * tr0, RTOL, ATOL)
    dt_s = time.perf_counter() - t0
    if st.status != 0:
        raise RuntimeError("diffusion oracle did not converge")
    return {"value": n ** 3 * st.n_vcycles / dt_s, "unit": UNIT, "cores": cores, "kind": "port", "vcycles": st.n_vcycles,
            "sample": f"scalar diffusion {n}^3, one box, red-black GS threaded over z-planes on all host cores, rtol {RTOL:g}; 1 solve, {dt_s:.1f} s"}


def run_reference_diffusion(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    r = cpu_diffusion(args.cpu_ncell, args.workload == "rayleigh_taylor", 0.01)
    line = {"impl": "reference", "metric": DIFF_METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": 1, "warmup": 0,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "scalar diffusion, oracle port", "cpu_sample": r["sample"]}, "cpu_baseline": r,
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
            "note": "reference incflo+AMReX build unavailable on this box; this is oracle/mac_oracle.cpp (solve_abec)"}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--path", default="nodal", choices=["nodal", "mac", "diffusion"],
                    help="nodal: the nodal projection (BASELINE metric, default); mac: the MAC projection (SURVEY 8(f) row 1); "
                         "diffusion: the implicit scalar diffusion solve (8(f) row 3)")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--ncell", type=int, default=512, help="cells per side per GPU (512: configs[2]; 256: configs[1])")
    ap.add_argument("--vel-ngrow", type=int, default=2, help="ghost layers of the velocity fab (incflo nghost_state: 2 MOL / 3 Godunov)")
    ap.add_argument("--cpu-ncell", type=int, default=None,
                    help="size of the CPU run (oracle): default 512 for the nodal path (the workload itself, ~25 s per projection on 16 cores; "
                         "--impl reference falls back to a 256^3 sample when steps x 26 s would exceed four minutes), 256 for --path mac / diffusion")
    ap.add_argument("--workload", default="double_shear", choices=["double_shear", "rayleigh_taylor"],
                    help="double_shear: BASELINE configs[2] (constant density, default); rayleigh_taylor: configs[3] (variable sigma)")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-workloads", action="store_true", help="skip the Rayleigh-Taylor 512^3 / double shear 256^3 sub-records (N = 1)")
    ap.add_argument("--no-parity", action="store_true", help="skip the multi-GPU parity check that precedes the timed region (N > 1)")
    ap.add_argument("--smooth-zchunk", type=int, default=0, help="experiment knob: fixed smoother z-chunk on every level (0: the level rule)")
    ap.add_argument("--agg-min-local-nz", type=int, default=0, help="experiment knob: replicate MG levels whose slab is thinner than this (0: library default)")
    ap.add_argument("--no-p2p", action="store_true", help="experiment knob: halo exchange through ncclSend/ncclRecv instead of peer memory")
    ap.add_argument("--no-graph", action="store_true", help="experiment knob: np_options.use_graph = 0 (launch the MLMG iteration kernel by kernel)")
    ap.add_argument("--overlap", action="store_true", help="experiment knob: overlap_halo=1 (end z-chunks first, their planes exchanged beside the interior chunks)")
    ap.add_argument("--no-overlap", action="store_true", help="experiment knob: halo exchange on the solver stream (overlap_halo=0)")
    ap.add_argument("--verbose", type=int, default=0, help="nodal_proj.verbose (4: per-launch smoother times on stderr)")
    args = ap.parse_args()
    if args.cpu_ncell is None and args.path in ("mac", "diffusion"):
        args.cpu_ncell = 256
    if args.path == "mac":
        run_reference_mac(args) if args.impl == "reference" else run_gpu_mac(args)
    elif args.path == "diffusion":
        run_reference_diffusion(args) if args.impl == "reference" else run_gpu_diffusion(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
