"""N > 1 host logic on CPU: two ranks (processes) drive the emulation build of the product sources, with the
halo exchange / allreduce / allgather callbacks carried by torch.distributed (gloo).  Checks the z-slab
partition, the slab halo exchange, owner / duplicate handling on the shared plane, the replicated coarse
levels and the reductions against the single-process oracle run on the same two boxes.
"""
import ctypes as C
import os
import socket

import numpy as np
import pytest

import nphelp
from incflo_b200 import problems

BIG = 1 << 30


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, ncell, kw, agg_min, overlap, outdir, mbox=None):
    import torch
    import torch.distributed as dist
    from incflo_b200.projector import Fab, Geometry, LPInfo, NodalProjector

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    L = nphelp.emu_lib()

    def view(ptr, n):
        return torch.from_numpy(np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(n,)))

    SR = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p,
                     C.c_longlong, C.c_int, C.c_int)
    AR = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int)
    AG = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_longlong)

    def sendrecv(slo, nslo, rlo, nrlo, shi, nshi, rhi, nrhi, lo, hi):
        reqs = []
        if lo >= 0:   # "down" messages carry tag 1, "up" messages tag 2
            reqs.append(dist.isend(view(slo, nslo).clone(), lo, tag=1))
            reqs.append(dist.irecv(view(rlo, nrlo), lo, tag=2))
        if hi >= 0:
            reqs.append(dist.isend(view(shi, nshi).clone(), hi, tag=2))
            reqs.append(dist.irecv(view(rhi, nrhi), hi, tag=1))
        for r in reqs:
            r.wait()
        return 0

    def allreduce(buf, n, op):
        t = view(buf, n)
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == 1 else dist.ReduceOp.SUM)
        return 0

    def allgather(send, recv, cnt):
        out = view(recv, cnt * world)
        dist.all_gather_into_tensor(out, view(send, cnt).clone())
        return 0

    cbs = (SR(sendrecv), AR(allreduce), AG(allgather))
    L.np_emu_set_comm.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.np_emu_set_comm(C.cast(cbs[0], C.c_void_p), C.cast(cbs[1], C.c_void_p), C.cast(cbs[2], C.c_void_p))

    nzl = ncell[2] // world
    prob = problems.PROBLEMS[name](ncell, **kw)
    sl = slice(rank * nzl, (rank + 1) * nzl)
    vel = Fab.from_array(prob["vel"][:, sl], 1)
    sig = Fab.from_array(prob["sigma"][sl], 0)
    geom = Geometry(ncell, prob["prob_lo"], prob["prob_hi"], [b == 0 for b in prob["bclo"]])
    boxes = [((0, 0, r * nzl), (ncell[0] - 1, ncell[1] - 1, (r + 1) * nzl - 1)) for r in range(world)]
    owner = None
    if mbox:        # several boxes per rank: every slab is chopped like amr.max_grid_size would
        boxes, owner, mine = [], [], []
        for r in range(world):
            for k0 in range(r * nzl, (r + 1) * nzl, mbox[2]):
                for j0 in range(0, ncell[1], mbox[1]):
                    for i0 in range(0, ncell[0], mbox[0]):
                        boxes.append(((i0, j0, k0), (i0 + mbox[0] - 1, j0 + mbox[1] - 1, k0 + mbox[2] - 1)))
                        owner.append(r)
        mine = [b for b, r in zip(boxes, owner) if r == rank]
        cut = lambda a, lo, hi: a[..., lo[2]:hi[2] + 1, lo[1]:hi[1] + 1, lo[0]:hi[0] + 1]
        vel = [Fab.from_array(cut(prob["vel"], lo, hi), 2) for lo, hi in mine]
        sig = [Fab.from_array(cut(prob["sigma"], lo, hi), 0) for lo, hi in mine]
    P = NodalProjector(vel, sig, geom, LPInfo(), boxes=boxes, owner=owner, rank=rank, nranks=world, nccl_comm=1, lib=L,
                       agg_min_local_nz=agg_min, overlap_halo=overlap)
    P.setDomainBC(prob["bclo"], prob["bchi"])
    st = P.project(1e-11, 1e-14)
    if mbox:        # put the boxes back together as this rank's slab
        u = np.zeros((3, nzl, ncell[1], ncell[0])); gp = np.zeros_like(u); phi = np.zeros((nzl + 1, ncell[1] + 1, ncell[0] + 1))
        for (lo, hi), fv, fp, fg in zip(mine, vel, P.getPhi(), P.getGradPhi()):
            k0, k1 = lo[2] - rank * nzl, hi[2] - rank * nzl
            u[:, k0:k1 + 1, lo[1]:hi[1] + 1, lo[0]:hi[0] + 1] = fv.numpy_valid()
            gp[:, k0:k1 + 1, lo[1]:hi[1] + 1, lo[0]:hi[0] + 1] = fg.numpy_valid()
            phi[k0:k1 + 2, lo[1]:hi[1] + 2, lo[0]:hi[0] + 2] = fp.numpy_valid()[0]
    else:
        u, phi, gp = vel.numpy_valid(), P.getPhi().numpy_valid()[0], P.getGradPhi().numpy_valid()
    np.savez(os.path.join(outdir, f"r{rank}.npz"), u=u, phi=phi,
             gp=gp, vc=st.n_vcycles, nlev=st.n_levels, bot=st.bottom_iters_total,
             hist=np.array([st.res_history[i] for i in range(st.n_vcycles)]),
             levels=np.array([P.level_ncell(l) for l in range(P.num_levels)]))
    P.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name,ncell,kw,agg_min,agg_level,overlap,mbox", [
    ("double_shear", (16, 16, 64), dict(zwave=True, prob_hi=(1.0, 1.0, 4.0)), 16, 2, 0, None),
    ("double_shear", (16, 16, 64), dict(zwave=True, prob_hi=(1.0, 1.0, 4.0)), 16, 2, -1, None),   # end chunks first + exchange beside the interior
    ("rayleigh_taylor", (16, 16, 32), {}, 4, 3, -1, None),
    ("rayleigh_taylor", (16, 16, 32), {}, 4, 3, 0, (8, 8, 8)),     # 8 boxes per rank, gathered into the rank's slab
])
def test_two_ranks_match_oracle(tmp_path, name, ncell, kw, agg_min, agg_level, overlap, mbox):
    import torch.multiprocessing as mp
    from oracle.oracle import Oracle, TILED

    nphelp.emu_lib()           # build once, before the ranks start
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), name, ncell, kw, agg_min, overlap, str(tmp_path), mbox), nprocs=world, join=True)
    r = [np.load(tmp_path / f"r{k}.npz") for k in range(world)]
    nzl = ncell[2] // world

    prob = problems.PROBLEMS[name](ncell, **kw)
    boxes = [[0, 0, k * nzl, ncell[0] - 1, ncell[1] - 1, (k + 1) * nzl - 1] for k in range(world)]
    o = Oracle(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"], boxes=boxes, smoother=TILED,
               tile=(63, 31, BIG), tile_z_target=-1, agg_level=agg_level)
    o.set_sigma(prob["sigma"])
    o.set_vel(prob["vel"])
    so = o.project(1e-11, 1e-14)

    def l2(a, b):
        return np.linalg.norm(a - b) / np.linalg.norm(b)

    assert int(r[0]["vc"]) == int(r[1]["vc"]) == so.n_vcycles
    assert int(r[0]["nlev"]) == so.n_levels
    # level agg_level is the first whole-domain (replicated) level on the ranks
    assert r[0]["levels"][agg_level][2] == ncell[2] >> agg_level and r[0]["levels"][agg_level - 1][2] == nzl >> (agg_level - 1)
    for i in range(so.n_vcycles):
        assert abs(r[0]["hist"][i] - so.res_history[i]) <= 1e-6 * so.res_history[i] + 1e-14 * so.rhs_norm0
    u = np.concatenate([r[k]["u"] for k in range(world)], axis=1)
    gp = np.concatenate([r[k]["gp"] for k in range(world)], axis=1)
    phi = np.concatenate([r[0]["phi"][:-1], r[1]["phi"]], axis=0)      # the shared plane is duplicated
    assert np.abs(r[0]["phi"][-1] - r[1]["phi"][0]).max() <= 1e-14 * np.abs(phi).max()
    assert l2(u, o.get_vel()) < 1e-11 and l2(gp, o.get_gradphi()) < 1e-10
    po = o.get_phi()
    assert l2(phi - phi.mean(), po - po.mean()) < 1e-10


# ---------------------------------------------------------------------------------------------------------------------
# MAC projection on two ranks (z-slabs): halo planes before each colour, sum / max over the ranks, the gather of the
# restricted residual and of the coarse face coefficients onto the replicated levels - against the one-box MAC oracle
# (a red-black sweep does not depend on the partition, so counts and history must be the single-box ones).
# ---------------------------------------------------------------------------------------------------------------------
def _mac_worker(rank, world, port, name, ncell, kw, agg_min, outdir):
    import torch
    import torch.distributed as dist
    from incflo_b200.mac_projector import MacProjector, face_fab_from_array
    from incflo_b200.projector import Fab, Geometry, LPInfo

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    L = nphelp.emu_lib()

    def view(ptr, n):
        return torch.from_numpy(np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(n,)))

    SR = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p,
                     C.c_longlong, C.c_int, C.c_int)
    AR = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int)
    AG = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_longlong)

    def sendrecv(slo, nslo, rlo, nrlo, shi, nshi, rhi, nrhi, lo, hi):
        reqs = []
        if lo >= 0:
            reqs.append(dist.isend(view(slo, nslo).clone(), lo, tag=1))
            reqs.append(dist.irecv(view(rlo, nrlo), lo, tag=2))
        if hi >= 0:
            reqs.append(dist.isend(view(shi, nshi).clone(), hi, tag=2))
            reqs.append(dist.irecv(view(rhi, nrhi), hi, tag=1))
        for r in reqs:
            r.wait()
        return 0

    def allreduce(buf, n, op):
        dist.all_reduce(view(buf, n), op=dist.ReduceOp.MAX if op == 1 else dist.ReduceOp.SUM)
        return 0

    def allgather(send, recv, cnt):
        dist.all_gather_into_tensor(view(recv, cnt * world), view(send, cnt).clone())
        return 0

    cbs = (SR(sendrecv), AR(allreduce), AG(allgather))
    L.np_emu_set_comm.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.np_emu_set_comm(C.cast(cbs[0], C.c_void_p), C.cast(cbs[1], C.c_void_p), C.cast(cbs[2], C.c_void_p))

    nzl = ncell[2] // world
    prob = problems.MAC_PROBLEMS[name](ncell, **kw)
    k0, k1 = rank * nzl, (rank + 1) * nzl
    cut = lambda a, d: a[k0:k1 + (1 if d == 2 else 0)]
    geom = Geometry(ncell, prob["prob_lo"], prob["prob_hi"], [b == 0 for b in prob["bclo"]])
    P = MacProjector(geom, lib=L, rank=rank, nranks=world, nccl_comm=1)
    P.setDomainBC(prob["bclo"], prob["bchi"])
    P.initProjector(LPInfo().setMaxCoarseningLevel(100), [face_fab_from_array(cut(prob["beta"][d], d)) for d in range(3)],
                    agg_min_local_nz=agg_min, fused_rb_min_n=64)
    fu = [face_fab_from_array(cut(prob["umac"][d], d), 1) for d in range(3)]
    P.setUMAC(fu)
    fphi = Fab.zeros((ncell[0], ncell[1], nzl), 1, 1)
    st = P.project(fphi, 1e-11, 1e-14)
    np.savez(os.path.join(outdir, f"m{rank}.npz"), u=fu[0].numpy_valid()[0], v=fu[1].numpy_valid()[0], w=fu[2].numpy_valid()[0],
             phi=fphi.numpy_valid()[0], vc=st.n_vcycles, nlev=st.n_levels, bot=st.bottom_iters_total, const=st.sigma_is_const,
             hist=np.array([st.res_history[i] for i in range(st.n_vcycles)]),
             levels=np.array([P.level_ncell(l) for l in range(P.num_levels)]))
    P.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name,ncell,kw,agg_min,agg_level", [
    ("double_shear", (16, 16, 64), dict(prob_hi=(1.0, 1.0, 4.0)), 16, 2),     # uniform beta, periodic z: every slab face is remote
    ("rayleigh_taylor", (16, 16, 32), {}, 4, 3),                             # variable beta (coefficient gather), walls in z,
                                                                             # two sharded coarse levels before the gather
    ("double_shear", (64, 16, 32), dict(prob_hi=(4.0, 1.0, 2.0)), 8, 2),     # whole 64 x 16 tiles, uniform beta: the one-pass
                                                                             # smoother on two ghost planes per slab side
    ("double_shear_walls", (64, 16, 32), dict(prob_hi=(4.0, 1.0, 2.0)), 8, 2),   # the same between walls: physical and shared z faces
    ("rayleigh_taylor", (64, 16, 32), dict(prob_hi=(2.0, 0.5, 1.0)), 8, 2),      # variable beta: ghost coefficient planes
])
def test_mac_two_ranks_match_oracle(tmp_path, name, ncell, kw, agg_min, agg_level):
    import torch.multiprocessing as mp
    from oracle.mac import MacOracle

    nphelp.emu_lib()
    world = 2
    mp.spawn(_mac_worker, args=(world, _free_port(), name, ncell, kw, agg_min, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(tmp_path / f"m{k}.npz") for k in range(world)]
    nzl = ncell[2] // world
    prob = problems.MAC_PROBLEMS[name](ncell, **kw)
    o = MacOracle(ncell, prob["prob_lo"], prob["prob_hi"], prob["bclo"], prob["bchi"])
    o.set_beta(*prob["beta"]); o.set_umac(*prob["umac"])
    so = o.project(1e-11, 1e-14)

    def l2(a, b):
        return np.linalg.norm(a - b) / np.linalg.norm(b)

    assert int(r[0]["vc"]) == int(r[1]["vc"]) == so.n_vcycles and int(r[0]["nlev"]) == so.n_levels
    assert int(r[0]["bot"]) == so.bottom_iters_total
    assert r[0]["levels"][agg_level][2] == ncell[2] >> agg_level and r[0]["levels"][agg_level - 1][2] == nzl >> (agg_level - 1)
    for i in range(so.n_vcycles):
        assert abs(r[0]["hist"][i] - so.res_history[i]) <= 1e-6 * so.res_history[i] + 1e-14 * so.rhs_norm0
    uo = o.get_umac()
    assert l2(np.concatenate([r[k]["u"] for k in range(world)], axis=0), uo[0]) < 1e-12
    assert l2(np.concatenate([r[k]["v"] for k in range(world)], axis=0), uo[1]) < 1e-12
    w = np.concatenate([r[0]["w"][:-1], r[1]["w"]], axis=0)               # the shared face plane is held by both ranks
    assert np.abs(r[0]["w"][-1] - r[1]["w"][0]).max() <= 1e-13 * np.abs(w).max()
    assert l2(w, uo[2]) < 1e-12
    phi = np.concatenate([r[k]["phi"] for k in range(world)], axis=0)
    po = o.get_phi()
    assert l2(phi - phi.mean(), po - po.mean()) < 1e-10
