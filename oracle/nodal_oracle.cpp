// ORACLE — TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED UPSTREAM.
//
// CPU restatement of the reference's nodal-projection path:
//     incflo::ApplyProjection  ->  amrex::NodalProjector::project  ->  amrex::MLMG::solve
//     on amrex::MLNodeLaplacian (cell-centred sigma, 3D, no EB, one AMR level).
// Reference call sites: /root/reference/src/projection/incflo_apply_nodal_projection.cpp:148-153
// (ctor, setDomainBC, project) and :165-166 (getPhi, getGradPhi).  The arithmetic itself lives in
// AMReX (>= 20.11, development branch; CMakeLists.txt:94, README.md:14-19), which is not in
// /root/reference and not on this machine, and the reference's own tests hold no golden vectors
// for this boundary.  Hence "parity unpinned": this file follows the published
// AMReX algorithm (function-by-function notes below) and is pinned by self-made identities
// (tests/test_oracle_pins.py: Q1 FEM assembly, D = -G^T, sparse direct solve).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  It is never linked into libnodalproj.so.
//
// Two smoother orderings are provided:
//   LEX   — what AMReX's CPU build does: lexicographic Gauss-Seidel, box by box, 2 sweeps per
//           Fsmooth with ghosts frozen, then nodalSync (owner overrides duplicates).
//   TILED — the ordering the sm_100a kernel uses (DESIGN.md "smoother semantics"), restated on
//           the CPU so that GPU-vs-oracle iterates can be compared to round-off.
#include "stencil.hpp"
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <memory>
#include <thread>
#include <vector>

namespace orc {

enum BC { Periodic = 0, Dirichlet = 1, Neumann = 2 };
enum Smoother { LEX = 0, TILED = 1 };
enum BottomSolver { BICGCG = 0, BICGSTAB = 1, CG = 2, SMOOTHER = 3 };

struct IBox { int lo[3], hi[3]; };   // inclusive index range

static int g_nthreads = 1;

static void parallel_for(int n, const std::function<void(int)>& f)
{
    int nt = std::min(g_nthreads, n);
    if (nt <= 1) { for (int i = 0; i < n; ++i) f(i); return; }
    std::atomic<int> next(0);
    std::vector<std::thread> th;
    th.reserve(nt);
    for (int t = 0; t < nt; ++t)
        th.emplace_back([&]() { for (;;) { int i = next.fetch_add(1); if (i >= n) break; f(i); } });
    for (auto& t : th) t.join();
}

// One box's storage: Fortran order, x fastest, component slowest (the FArrayBox layout).
struct Fab {
    IBox r{};            // allocated index range (valid + ghosts)
    int nc = 1;
    std::ptrdiff_t sy = 0, sz = 0, sn = 0;
    std::vector<double> d;
    void define(const IBox& alloc, int ncomp) {
        r = alloc; nc = ncomp;
        sy = r.hi[0]-r.lo[0]+1; sz = sy*(r.hi[1]-r.lo[1]+1); sn = sz*(r.hi[2]-r.lo[2]+1);
        d.assign((size_t)sn*nc, 0.0);
    }
    void define_layout(const IBox& alloc) {     // strides only, no storage (metadata building)
        r = alloc; nc = 1;
        sy = r.hi[0]-r.lo[0]+1; sz = sy*(r.hi[1]-r.lo[1]+1); sn = sz*(r.hi[2]-r.lo[2]+1);
    }
    std::ptrdiff_t off(int i, int j, int k) const { return (i-r.lo[0]) + sy*(j-r.lo[1]) + sz*(k-r.lo[2]); }
    double& operator()(int i, int j, int k, int n = 0) { return d[off(i,j,k)+sn*n]; }
    double operator()(int i, int j, int k, int n = 0) const { return d[off(i,j,k)+sn*n]; }
    double* ptr(int i, int j, int k, int n = 0) { return d.data()+off(i,j,k)+sn*n; }
    const double* ptr(int i, int j, int k, int n = 0) const { return d.data()+off(i,j,k)+sn*n; }
};

struct CopyItem { int db; std::ptrdiff_t doff; int sb; std::ptrdiff_t soff; };

// A "MultiFab": one Fab per box.  nodal=true -> valid range is [lo, hi+1].
struct MF {
    bool nodal = false; int ng = 0, nc = 1;
    std::vector<Fab> f;
    void setVal(double v) { parallel_for((int)f.size(), [&](int b){ std::fill(f[b].d.begin(), f[b].d.end(), v); }); }
};

static IBox valid_of(const IBox& cb, bool nodal) {
    IBox v = cb; if (nodal) for (int d = 0; d < 3; ++d) v.hi[d] += 1; return v;
}
static IBox grow(const IBox& b, int g) { IBox r = b; for (int d = 0; d < 3; ++d) { r.lo[d]-=g; r.hi[d]+=g; } return r; }

struct Level {
    int n[3];                 // domain cells
    double dxinv[3];
    int bclo[3], bchi[3];
    std::vector<IBox> boxes;  // cell boxes
    MF sigma;                 // cell, 1 comp, 1 ghost
    // communication metadata (AMReX caches the same in FabArrayBase::FB / OverrideSync)
    std::vector<CopyItem> nd_ghost, nd_sync;     // nodal 1-ghost fill; non-owner <- owner
    std::vector<CopyItem> cc_ghost;              // cell 1-ghost fill (periodic + neighbours)
    std::vector<std::vector<unsigned char>> owner;   // per box, over the VALID nodal range
    bool lo_owns = false;     // TILED: a shared node belongs to the box whose LOW face it is on
    bool periodic(int d) const { return bclo[d] == Periodic; }
    bool is_dirichlet(int i, int j, int k) const {
        return (bclo[0]==Dirichlet && i==0) || (bchi[0]==Dirichlet && i==n[0])
            || (bclo[1]==Dirichlet && j==0) || (bchi[1]==Dirichlet && j==n[1])
            || (bclo[2]==Dirichlet && k==0) || (bchi[2]==Dirichlet && k==n[2]);
    }
    bool has_dirichlet() const { for (int d=0; d<3; ++d) if (bclo[d]==Dirichlet||bchi[d]==Dirichlet) return true; return false; }
    // dot-mask weight excluding ownership: 1/2 per Neumann face the node lies on
    // (restates amrex::mlndlap_set_dot_mask)
    double neumann_weight(int i, int j, int k) const {
        double w = 1.0; const int idx[3] = {i,j,k};
        for (int d = 0; d < 3; ++d) {
            if (bclo[d]==Neumann && idx[d]==0) w *= 0.5;
            if (bchi[d]==Neumann && idx[d]==n[d]) w *= 0.5;
        }
        return w;
    }
    MF make(bool nodal, int ng, int nc) const {
        MF m; m.nodal = nodal; m.ng = ng; m.nc = nc; m.f.resize(boxes.size());
        for (size_t b = 0; b < boxes.size(); ++b) m.f[b].define(grow(valid_of(boxes[b], nodal), ng), nc);
        return m;
    }
    MF make_layout(bool nodal, int ng) const {
        MF m; m.nodal = nodal; m.ng = ng; m.nc = 1; m.f.resize(boxes.size());
        for (size_t b = 0; b < boxes.size(); ++b) m.f[b].define_layout(grow(valid_of(boxes[b], nodal), ng));
        return m;
    }
};

static void owned_range(const Level& L, const IBox& cb, IBox& o);

// ---------------------------------------------------------------------------------------------
// Communication metadata.  Owner rule restates FabArray::OwnerMask: the lowest box index owns a
// shared node; within one box a periodic image on the high side is not the owner.
// ---------------------------------------------------------------------------------------------
static void build_comm(Level& L)
{
    const int nb = (int)L.boxes.size();
    // ---- nodal ----
    int dn[3]; for (int d=0; d<3; ++d) dn[d] = L.periodic(d) ? L.n[d] : L.n[d]+1;
    auto canon = [&](int idx, int d, bool& ok) {
        if (L.periodic(d)) { int m = idx % L.n[d]; if (m < 0) m += L.n[d]; ok = true; return m; }
        ok = (idx >= 0 && idx <= L.n[d]); return idx;
    };
    std::vector<int> obox((size_t)dn[0]*dn[1]*dn[2], -1);
    std::vector<std::ptrdiff_t> ooff(obox.size(), 0);
    MF proto = L.make_layout(true, 1);
    for (int b = 0; b < nb; ++b) {
        IBox v = valid_of(L.boxes[b], true);
        if (L.lo_owns) owned_range(L, L.boxes[b], v);
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i) {
            bool o0,o1,o2; int ci=canon(i,0,o0), cj=canon(j,1,o1), ck=canon(k,2,o2);
            size_t g = (size_t)ci + (size_t)dn[0]*((size_t)cj + (size_t)dn[1]*ck);
            if (obox[g] < 0) { obox[g] = b; ooff[g] = proto.f[b].off(i,j,k); }
        }
    }
    L.owner.assign(nb, {});
    L.nd_ghost.clear(); L.nd_sync.clear();
    for (int b = 0; b < nb; ++b) {
        IBox v = valid_of(L.boxes[b], true), gb = grow(v,1);
        const Fab& fb = proto.f[b];
        L.owner[b].assign((size_t)(v.hi[0]-v.lo[0]+1)*(v.hi[1]-v.lo[1]+1)*(v.hi[2]-v.lo[2]+1), 1);
        size_t vi = 0;
        for (int k=gb.lo[2]; k<=gb.hi[2]; ++k) for (int j=gb.lo[1]; j<=gb.hi[1]; ++j) for (int i=gb.lo[0]; i<=gb.hi[0]; ++i) {
            bool isvalid = i>=v.lo[0]&&i<=v.hi[0]&&j>=v.lo[1]&&j<=v.hi[1]&&k>=v.lo[2]&&k<=v.hi[2];
            bool o0,o1,o2; int ci=canon(i,0,o0), cj=canon(j,1,o1), ck=canon(k,2,o2);
            if (!(o0&&o1&&o2)) continue;          // outside a non-periodic domain: BC's business
            size_t g = (size_t)ci + (size_t)dn[0]*((size_t)cj + (size_t)dn[1]*ck);
            std::ptrdiff_t off = fb.off(i,j,k);
            if (isvalid) {
                size_t idx = (size_t)(i-v.lo[0]) + (size_t)(v.hi[0]-v.lo[0]+1)*((size_t)(j-v.lo[1]) + (size_t)(v.hi[1]-v.lo[1]+1)*(k-v.lo[2]));
                (void)vi;
                if (!(obox[g]==b && ooff[g]==off)) { L.owner[b][idx] = 0; L.nd_sync.push_back({b,off,obox[g],ooff[g]}); }
            } else if (obox[g] >= 0) {
                L.nd_ghost.push_back({b,off,obox[g],ooff[g]});
            }
        }
    }
    // ---- cells (1 ghost) ----
    std::vector<int> cbox((size_t)L.n[0]*L.n[1]*L.n[2], -1);
    std::vector<std::ptrdiff_t> coff(cbox.size(), 0);
    MF cproto = L.make_layout(false, 1);
    for (int b = 0; b < nb; ++b) {
        const IBox& v = L.boxes[b];
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i) {
            size_t g = (size_t)i + (size_t)L.n[0]*((size_t)j + (size_t)L.n[1]*k);
            cbox[g] = b; coff[g] = cproto.f[b].off(i,j,k);
        }
    }
    L.cc_ghost.clear();
    for (int b = 0; b < nb; ++b) {
        const IBox& v = L.boxes[b]; IBox gb = grow(v,1);
        for (int k=gb.lo[2]; k<=gb.hi[2]; ++k) for (int j=gb.lo[1]; j<=gb.hi[1]; ++j) for (int i=gb.lo[0]; i<=gb.hi[0]; ++i) {
            bool isvalid = i>=v.lo[0]&&i<=v.hi[0]&&j>=v.lo[1]&&j<=v.hi[1]&&k>=v.lo[2]&&k<=v.hi[2];
            if (isvalid) continue;
            int c[3] = {i,j,k}; bool ok = true;
            for (int d=0; d<3; ++d) {
                if (L.periodic(d)) { c[d] %= L.n[d]; if (c[d]<0) c[d]+=L.n[d]; }
                else if (c[d]<0 || c[d]>=L.n[d]) ok = false;
            }
            if (!ok) continue;
            size_t g = (size_t)c[0] + (size_t)L.n[0]*((size_t)c[1] + (size_t)L.n[1]*c[2]);
            if (cbox[g] >= 0) L.cc_ghost.push_back({b, cproto.f[b].off(i,j,k), cbox[g], coff[g]});
        }
    }
}

static void run_copy(MF& m, const std::vector<CopyItem>& items)
{
    // Sources are owner/valid entries, destinations are ghosts or non-owner duplicates: disjoint sets.
    const int nc = m.nc;
    const size_t N = items.size();
    int chunks = std::max(1, std::min(g_nthreads*4, (int)(N/4096)+1));
    parallel_for(chunks, [&](int c) {
        size_t b0 = N*c/chunks, b1 = N*(c+1)/chunks;
        for (size_t q = b0; q < b1; ++q) {
            const CopyItem& it = items[q];
            for (int n = 0; n < nc; ++n)
                m.f[it.db].d[it.doff + m.f[it.db].sn*n] = m.f[it.sb].d[it.soff + m.f[it.sb].sn*n];
        }
    });
}

// FabArray::FillBoundary(periodicity) on a nodal 1-ghost MF / cell 1-ghost MF.
// The cell lists were built for ng = 1 fabs; cell MFs with more ghosts use fill_cc_ghost_generic.
static void fill_boundary_nodal(const Level& L, MF& m) { run_copy(m, L.nd_ghost); }
// amrex::OverrideSync with the owner mask (MLNodeLinOp::nodalSync)
static void nodal_sync(const Level& L, MF& m) { run_copy(m, L.nd_sync); }

// MLNodeLinOp::applyBC = FillBoundary + mlndlap_applybc (Neumann/inflow: ghost = mirror image).
static void apply_bc(const Level& L, MF& m, bool skip_fillboundary = false)
{
    if (!skip_fillboundary) fill_boundary_nodal(L, m);
    parallel_for((int)m.f.size(), [&](int b) {
        Fab& f = m.f[b];
        IBox v = valid_of(L.boxes[b], true), g = grow(v,1);
        for (int d = 0; d < 3; ++d) {
            const int d1 = (d+1)%3, d2 = (d+2)%3;
            for (int side = 0; side < 2; ++side) {
                int bc = side ? L.bchi[d] : L.bclo[d];
                if (bc != Neumann) continue;
                int face = side ? L.n[d] : 0;
                if ((side ? v.hi[d] : v.lo[d]) != face) continue;
                int gi = side ? face+1 : -1, si = side ? face-1 : 1;
                for (int a2 = g.lo[d2]; a2 <= g.hi[d2]; ++a2) for (int a1 = g.lo[d1]; a1 <= g.hi[d1]; ++a1) {
                    int ig[3], is[3];
                    ig[d]=gi; ig[d1]=a1; ig[d2]=a2; is[d]=si; is[d1]=a1; is[d2]=a2;
                    f(ig[0],ig[1],ig[2]) = f(is[0],is[1],is[2]);
                }
            }
        }
    });
}

// MLNodeLaplacian::FillBoundaryCoeff: FillBoundary, then mlndlap_fillbc_cc (copy of the adjacent
// interior cell outside every non-periodic domain face).
static void fill_sigma_bc(const Level& L, MF& s)
{
    run_copy(s, L.cc_ghost);
    parallel_for((int)s.f.size(), [&](int b) {
        Fab& f = s.f[b];
        const IBox& v = L.boxes[b]; IBox g = grow(v,1);
        for (int d = 0; d < 3; ++d) {
            if (L.periodic(d)) continue;
            const int d1 = (d+1)%3, d2 = (d+2)%3;
            for (int side = 0; side < 2; ++side) {
                int edge = side ? L.n[d]-1 : 0;
                if ((side ? v.hi[d] : v.lo[d]) != edge) continue;
                int gi = side ? edge+1 : -1;
                for (int a2 = g.lo[d2]; a2 <= g.hi[d2]; ++a2) for (int a1 = g.lo[d1]; a1 <= g.hi[d1]; ++a1) {
                    int ig[3], is[3];
                    ig[d]=gi; ig[d1]=a1; ig[d2]=a2; is[d]=edge; is[d1]=a1; is[d2]=a2;
                    f(ig[0],ig[1],ig[2]) = f(is[0],is[1],is[2]);
                }
            }
        }
    });
}

// ---------------------------------------------------------------------------------------------
// Operator pieces (box loops around stencil.hpp)
// ---------------------------------------------------------------------------------------------
// MLNodeLaplacian::Fapply (mlndlap_adotx_aa); caller has applied BCs to x.
static void fapply(const Level& L, MF& y, const MF& x)
{
    StencilCoef c(L.dxinv);
    parallel_for((int)L.boxes.size(), [&](int b) {
        IBox v = valid_of(L.boxes[b], true);
        const Fab& xf = x.f[b]; const Fab& sf = L.sigma.f[b]; Fab& yf = y.f[b];
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i) {
            yf(i,j,k) = L.is_dirichlet(i,j,k) ? 0.0
                      : adotx_node(c, xf.ptr(i,j,k), xf.sy, xf.sz, sf.ptr(i,j,k), sf.sy, sf.sz);
        }
    });
}

// MLNodeLaplacian::Fsmooth, CPU branch: mlndlap_gauss_seidel_aa, lexicographic, 2 sweeps per box.
static void fsmooth_lex(const Level& L, MF& x, const MF& rhs, int nsweeps)
{
    StencilCoef c(L.dxinv);
    parallel_for((int)L.boxes.size(), [&](int b) {
        IBox v = valid_of(L.boxes[b], true);
        Fab& xf = x.f[b]; const Fab& sf = L.sigma.f[b]; const Fab& rf = rhs.f[b];
        for (int ns = 0; ns < nsweeps; ++ns)
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i) {
            if (L.is_dirichlet(i,j,k)) { xf(i,j,k) = 0.0; continue; }
            Sig8 g(sf.ptr(i,j,k), sf.sy, sf.sz);
            double s0 = diag_node(c, g);
            double* xp = xf.ptr(i,j,k);
            double Ax = xp[0]*s0 + offdiag_node(c, g, xp, xf.sy, xf.sz);
            xp[0] += (rf(i,j,k) - Ax) / s0;
        }
    });
}

// TILED ordering (the GPU kernel's).  Owned nodes of each box are cut into tiles of
// (tx,ty,tz) starting at the box's low corner; every tile is smoothed independently with all
// values outside the tile frozen at their pre-smooth state.  Inside a tile: for each sweep, for
// each plane k ascending, for colour c = 2*(j&1)+(i&1) (tile-local indices) ascending, update all
// nodes of that colour.  A node is owned by a box unless it lies on the box's high face and that
// face is not a non-periodic domain boundary.
struct TileSpec { int t[3]; int variant = 0; int overlap = 1; int z_target = 0; int owned_only = 0; };
// Experimental orderings (variant > 0) used only to study smoothing rates: nodes are
// visited in ascending key order, nodes with equal keys are updated simultaneously (Jacobi).
static long order_key(int variant, int ns, int i, int j, int k, int ni, int nj, int nk)
{
    const long B = 1L<<12;
    switch (variant) {
    case 1: { long c = 2*(j&1)+(i&1); return (ns&1) ? ((nk-1-k)*4 + (3-c)) : (k*4 + c); }   // symmetric
    case 2: { long c = 4*(k&1)+2*(j&1)+(i&1); return c; }                                   // 8-colour
    case 3: return ((long)k*B + j)*2 + (i&1);                                               // row-seq, rb in i
    case 4: return ((long)k*B + j)*B + i;                                                   // lexicographic
    case 5: { long c = (i+j+k)&1; return c; }                                               // red-black hybrid
    case 6: return 0;                                                                       // Jacobi
    case 7: { long c = 2*(j&1)+(i&1); return (long)c*B + k; }                               // 4-colour columns, k-seq inside
    default: return 0;
    }
}
static void owned_range(const Level& L, const IBox& cb, IBox& o)
{
    o = valid_of(cb, true);
    for (int d = 0; d < 3; ++d) {
        bool domain_hi = (cb.hi[d] == L.n[d]-1);
        if (!(domain_hi && !L.periodic(d))) o.hi[d] -= 1;
    }
}
static void fsmooth_tiled(const Level& L, MF& x, const MF& rhs, int nsweeps, const TileSpec& ts_in)
{
    StencilCoef c(L.dxinv);
    MF out = x;   // frozen input stays in x; results go to out
    for (size_t b = 0; b < L.boxes.size(); ++b) {
        // All VALID nodes are relaxed, duplicates included (they are overwritten by their owner in
        // nodal_sync afterwards) — the same one-node overlap between neighbouring boxes / periodic
        // images that AMReX's box-by-box Gauss-Seidel has.
        IBox o = valid_of(L.boxes[b], true);
        if (ts_in.owned_only) owned_range(L, L.boxes[b], o);   // duplicates stay frozen; nodal_sync refreshes them
        TileSpec ts = ts_in;
        if (ts.z_target != 0) {      // the GPU driver's rule (incflo_b200/csrc/np_solver.cu gs_zchunk / pick_zchunk)
            const int nn = o.hi[2]-o.lo[2]+1;
            int target = ts.z_target, zc_small = 0;
            if (target < 0) {
                const int ncols = ((o.hi[0]-o.lo[0]+1 + ts.t[0]-1)/ts.t[0]) * ((o.hi[1]-o.lo[1]+1 + ts.t[1]-1)/ts.t[1]);
                // small levels: as many chunks as fit one wave of 296 CTAs, at least 4 planes each
                const int nch1 = std::min(296 / ncols, std::max(1, nn / 4));
                if (nch1 >= 1 && (nn + nch1 - 1) / nch1 <= 32) zc_small = (nn + nch1 - 1) / nch1;
                // large levels: enough z-chunks to fill the GPU, between 8 and 32 planes each
                const int nch_want = (592 + ncols - 1) / ncols;
                target = std::min(32, std::max(8, nn / nch_want));
            }
            const int nch = std::max(1, (nn + target/2) / target);
            ts.t[2] = zc_small > 0 ? zc_small : (nn + nch - 1) / nch;
            if (ts.z_target < 0) ts.t[2] += ts.t[2] & 1;      // even: chunk interfaces on coarse-grid planes
        }
        int nt[3]; for (int d=0; d<3; ++d) nt[d] = (o.hi[d]-o.lo[d]+1 + ts.t[d]-1)/ts.t[d];
        const Fab& xin = x.f[b]; const Fab& sf = L.sigma.f[b]; const Fab& rf = rhs.f[b]; Fab& xo = out.f[b];
        parallel_for(nt[0]*nt[1]*nt[2], [&](int t) {
            int t0 = t % nt[0], t1 = (t/nt[0]) % nt[1], t2 = t/(nt[0]*nt[1]);
            IBox tb;
            tb.lo[0]=o.lo[0]+t0*ts.t[0]; tb.lo[1]=o.lo[1]+t1*ts.t[1]; tb.lo[2]=o.lo[2]+t2*ts.t[2];
            IBox wb = tb;    // nodes this tile WRITES: [lo, lo+t-1]
            for (int d=0; d<3; ++d) wb.hi[d] = std::min(tb.lo[d]+ts.t[d]-1, o.hi[d]);
            // nodes this tile RELAXES: one more on the high side (shared with the next tile, which owns it)
            for (int d=0; d<3; ++d) tb.hi[d] = std::min(tb.lo[d]+ts.t[d]-1+ts.overlap, o.hi[d]);
            Fab w; w.define(grow(tb,1), 1);
            for (int k=w.r.lo[2]; k<=w.r.hi[2]; ++k) for (int j=w.r.lo[1]; j<=w.r.hi[1]; ++j) for (int i=w.r.lo[0]; i<=w.r.hi[0]; ++i)
                w(i,j,k) = xin(i,j,k);
            if (ts.variant > 0) {
                struct Nd { long key; int i,j,k; };
                const int ni=tb.hi[0]-tb.lo[0]+1, nj=tb.hi[1]-tb.lo[1]+1, nk=tb.hi[2]-tb.lo[2]+1;
                for (int ns = 0; ns < nsweeps; ++ns) {
                    std::vector<Nd> nds; nds.reserve((size_t)ni*nj*nk);
                    for (int k=tb.lo[2]; k<=tb.hi[2]; ++k) for (int j=tb.lo[1]; j<=tb.hi[1]; ++j) for (int i=tb.lo[0]; i<=tb.hi[0]; ++i)
                        nds.push_back({order_key(ts.variant, ns, i-tb.lo[0], j-tb.lo[1], k-tb.lo[2], ni,nj,nk), i,j,k});
                    std::stable_sort(nds.begin(), nds.end(), [](const Nd& a, const Nd& b){ return a.key < b.key; });
                    std::vector<double> tmp;
                    for (size_t a = 0; a < nds.size();) {
                        size_t e = a; while (e < nds.size() && nds[e].key == nds[a].key) ++e;
                        tmp.resize(e-a);
                        for (size_t q = a; q < e; ++q) {
                            int i=nds[q].i, j=nds[q].j, k=nds[q].k;
                            if (L.is_dirichlet(i,j,k)) { tmp[q-a] = 0.0; continue; }
                            Sig8 g(sf.ptr(i,j,k), sf.sy, sf.sz);
                            tmp[q-a] = (rf(i,j,k) - offdiag_node(c, g, w.ptr(i,j,k), w.sy, w.sz)) / diag_node(c, g);
                        }
                        for (size_t q = a; q < e; ++q) w(nds[q].i,nds[q].j,nds[q].k) = tmp[q-a];
                        a = e;
                    }
                }
            } else
            for (int ns = 0; ns < nsweeps; ++ns)
            for (int k=tb.lo[2]; k<=tb.hi[2]; ++k)
            for (int col = 0; col < 4; ++col)
            for (int j=tb.lo[1]; j<=tb.hi[1]; ++j) {
                if (((j-tb.lo[1])&1) != (col>>1)) continue;      // colour parity is tile-local
                for (int i=tb.lo[0]; i<=tb.hi[0]; ++i) {
                    if (((i-tb.lo[0])&1) != (col&1)) continue;
                    if (L.is_dirichlet(i,j,k)) { w(i,j,k) = 0.0; continue; }
                    Sig8 g(sf.ptr(i,j,k), sf.sy, sf.sz);
                    double s0 = diag_node(c, g);
                    double* xp = w.ptr(i,j,k);
                    double off = offdiag_node(c, g, xp, w.sy, w.sz);
                    xp[0] = (rf(i,j,k) - off) / s0;
                }
            }
            for (int k=wb.lo[2]; k<=wb.hi[2]; ++k) for (int j=wb.lo[1]; j<=wb.hi[1]; ++j) for (int i=wb.lo[0]; i<=wb.hi[0]; ++i)
                xo(i,j,k) = w(i,j,k);
        });
    }
    x = std::move(out);
}

// MLNodeLaplacian::restriction (mlndlap_restriction); caller applied BCs to the fine MF.
static void restriction(const Level& Lf, const Level& Lc, MF& crse, const MF& fine)
{
    parallel_for((int)Lc.boxes.size(), [&](int b) {
        IBox v = valid_of(Lc.boxes[b], true);
        const Fab& ff = fine.f[b]; Fab& cf = crse.f[b];
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i)
            cf(i,j,k) = Lf.is_dirichlet(2*i,2*j,2*k) ? 0.0 : restrict_node(ff.ptr(2*i,2*j,2*k), ff.sy, ff.sz);
    });
}

// sigma-weighted prolongation (mlndlap_interpadd_aa with aa_interp_line_* / aa_interp_face_*)
struct InterpCtx {
    const Fab* c; const Fab* s;
    double quad(const int m[3], int d, int side) const {   // 4 fine cells on `side` of fine node m along d
        const int d1=(d+1)%3, d2=(d+2)%3; double w = 0;
        for (int b = -1; b <= 0; ++b) for (int a = -1; a <= 0; ++a) {
            int q[3]; q[d] = m[d] + (side ? 0 : -1); q[d1] = m[d1]+a; q[d2] = m[d2]+b;
            w += (*s)(q[0],q[1],q[2]);
        }
        return w;
    }
    double crse(const int m[3]) const { return (*c)(m[0]/2, m[1]/2, m[2]/2); }   // m even, >= 0
    double line(const int m[3], int d) const {
        int lo[3]={m[0],m[1],m[2]}, hi[3]={m[0],m[1],m[2]}; lo[d]-=1; hi[d]+=1;
        double w1 = quad(m,d,0), w2 = quad(m,d,1);
        return (w1*crse(lo) + w2*crse(hi))/(w1+w2);
    }
    double face(const int m[3], int da, int db) const {
        double wa0=quad(m,da,0), wa1=quad(m,da,1), wb0=quad(m,db,0), wb1=quad(m,db,1);
        int a0[3]={m[0],m[1],m[2]}, a1[3]={m[0],m[1],m[2]}, b0[3]={m[0],m[1],m[2]}, b1[3]={m[0],m[1],m[2]};
        a0[da]-=1; a1[da]+=1; b0[db]-=1; b1[db]+=1;
        return (wa0*line(a0,db) + wa1*line(a1,db) + wb0*line(b0,da) + wb1*line(b1,da))/(wa0+wa1+wb0+wb1);
    }
    double centre(const int m[3]) const {
        double num = 0, den = 0;
        for (int d = 0; d < 3; ++d) {
            const int d1=(d+1)%3, d2=(d+2)%3;
            int lo[3]={m[0],m[1],m[2]}, hi[3]={m[0],m[1],m[2]}; lo[d]-=1; hi[d]+=1;
            double w0 = quad(m,d,0), w1 = quad(m,d,1);
            num += w0*face(lo,d1,d2) + w1*face(hi,d1,d2); den += w0+w1;
        }
        return num/den;
    }
};

static void interp_add(const Level& Lf, MF& fine, const MF& crse, bool sigma_weighted)
{
    parallel_for((int)Lf.boxes.size(), [&](int b) {
        IBox v = valid_of(Lf.boxes[b], true);
        Fab& ff = fine.f[b]; const Fab& cf = crse.f[b];
        InterpCtx ctx{&cf, &Lf.sigma.f[b]};
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i) {
            if (Lf.is_dirichlet(i,j,k)) continue;
            const int m[3] = {i,j,k};
            const bool ox = i&1, oy = j&1, oz = k&1;   // box lows are even at every level that has a coarser one
            const int nodd = ox+oy+oz;
            double val;
            if (nodd == 0) val = cf(i/2,j/2,k/2);
            else if (!sigma_weighted) {
                double sum = 0;
                for (int c2=0; c2<=oz; ++c2) for (int c1=0; c1<=oy; ++c1) for (int c0=0; c0<=ox; ++c0)
                    sum += cf((i-ox)/2+c0, (j-oy)/2+c1, (k-oz)/2+c2);
                val = sum / (double)(1<<nodd);
            }
            else if (nodd == 1) val = ctx.line(m, ox?0:(oy?1:2));
            else if (nodd == 2) val = ox&&oy ? ctx.face(m,0,1) : (ox&&oz ? ctx.face(m,0,2) : ctx.face(m,1,2));
            else val = ctx.centre(m);
            ff(i,j,k) += val;
        }
    });
}

// MLNodeLaplacian::averageDownCoeffs (arithmetic; harmonic averaging is off in NodalProjector)
static void avgdown_sigma(const Level& Lf, Level& Lc)
{
    parallel_for((int)Lc.boxes.size(), [&](int b) {
        const IBox& v = Lc.boxes[b]; const Fab& ff = Lf.sigma.f[b]; Fab& cf = Lc.sigma.f[b];
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i) {
            double s = 0;
            for (int c=0;c<2;++c) for (int bb=0;bb<2;++bb) for (int a=0;a<2;++a) s += ff(2*i+a,2*j+bb,2*k+c);
            cf(i,j,k) = 0.125*s;
        }
    });
    fill_sigma_bc(Lc, Lc.sigma);
}

// MLNodeLaplacian::compRHS -> mlndlap_divu + mlndlap_impose_neumann_bc.  vel has >= 1 filled ghost.
static void divu(const Level& L, MF& rhs, const MF& vel)
{
    parallel_for((int)L.boxes.size(), [&](int b) {
        IBox v = valid_of(L.boxes[b], true);
        const Fab& u = vel.f[b]; Fab& r = rhs.f[b];
        const double fac[3] = {0.25*L.dxinv[0], 0.25*L.dxinv[1], 0.25*L.dxinv[2]};
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i) {
            if (L.is_dirichlet(i,j,k)) { r(i,j,k) = 0.0; continue; }
            const int idx[3] = {i,j,k};
            // zero[d][side]: cells on that side of the node along d lie outside a Neumann face
            double zero[3][2];
            for (int d=0; d<3; ++d) {
                zero[d][0] = (L.bclo[d]==Neumann && idx[d]==0) ? 0.0 : 1.0;
                zero[d][1] = (L.bchi[d]==Neumann && idx[d]==L.n[d]) ? 0.0 : 1.0;
            }
            double acc = 0;
            for (int d = 0; d < 3; ++d) {
                const int d1=(d+1)%3, d2=(d+2)%3;
                double s = 0;
                for (int b2=-1; b2<=0; ++b2) for (int b1=-1; b1<=0; ++b1) {
                    int hi[3], lo[3];
                    hi[d]=idx[d]; lo[d]=idx[d]-1; hi[d1]=lo[d1]=idx[d1]+b1; hi[d2]=lo[d2]=idx[d2]+b2;
                    double z = zero[d1][b1+1]*zero[d2][b2+1];   // tangential cells outside the wall are not seen
                    s += z*(u(hi[0],hi[1],hi[2],d) - u(lo[0],lo[1],lo[2],d));
                }
                acc += fac[d]*s;
            }
            // mlndlap_impose_neumann_bc: x2 per Neumann face the node lies on
            for (int d=0; d<3; ++d) {
                if (L.bclo[d]==Neumann && idx[d]==0) acc *= 2.0;
                if (L.bchi[d]==Neumann && idx[d]==L.n[d]) acc *= 2.0;
            }
            r(i,j,k) = acc;
        }
    });
}

// MLNodeLaplacian::getFluxes -> mlndlap_mknewu on a zeroed flux, then NodalProjector::project's
// "vel += flux; flux = -flux/sigma".  Done in one pass: u -= sigma*grad, gphi = grad.
static void mknewu(const Level& L, MF& vel, MF& gphi, const MF& phi)
{
    parallel_for((int)L.boxes.size(), [&](int b) {
        const IBox& v = L.boxes[b];
        const Fab& p = phi.f[b]; const Fab& sf = L.sigma.f[b]; Fab& u = vel.f[b]; Fab& g = gphi.f[b];
        const double fac[3] = {0.25*L.dxinv[0], 0.25*L.dxinv[1], 0.25*L.dxinv[2]};
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i) {
            const double* q = p.ptr(i,j,k); const std::ptrdiff_t y = p.sy, z = p.sz;
            double gx = fac[0]*((q[1]-q[0]) + (q[1+y]-q[y]) + (q[1+z]-q[z]) + (q[1+y+z]-q[y+z]));
            double gy = fac[1]*((q[y]-q[0]) + (q[1+y]-q[1]) + (q[y+z]-q[z]) + (q[1+y+z]-q[1+z]));
            double gz = fac[2]*((q[z]-q[0]) + (q[1+z]-q[1]) + (q[y+z]-q[y]) + (q[1+y+z]-q[1+y]));
            double s = sf(i,j,k);
            u(i,j,k,0) -= s*gx; u(i,j,k,1) -= s*gy; u(i,j,k,2) -= s*gz;
            g(i,j,k,0) = gx; g(i,j,k,1) = gy; g(i,j,k,2) = gz;
        }
    });
}

// ---------------------------------------------------------------------------------------------
// Reductions (MLNodeLinOp::xdoty with the dot mask; MultiFab::norminf over valid nodes)
// ---------------------------------------------------------------------------------------------
static double dot_masked(const Level& L, const MF& x, const MF& y)
{
    std::vector<double> part(L.boxes.size(), 0.0);
    parallel_for((int)L.boxes.size(), [&](int b) {
        IBox v = valid_of(L.boxes[b], true); size_t q = 0; double s = 0;
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i, ++q)
            if (L.owner[b][q]) s += L.neumann_weight(i,j,k)*x.f[b](i,j,k)*y.f[b](i,j,k);
        part[b] = s;
    });
    double s = 0; for (double p : part) s += p; return s;
}
static double norm_inf(const Level& L, const MF& x)
{
    std::vector<double> part(L.boxes.size(), 0.0);
    parallel_for((int)L.boxes.size(), [&](int b) {
        IBox v = valid_of(L.boxes[b], true); double s = 0;
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i)
            s = std::max(s, std::fabs(x.f[b](i,j,k)));
        part[b] = s;
    });
    double s = 0; for (double p : part) s = std::max(s,p); return s;
}
// MLNodeLinOp::getSolvabilityOffset + fixSolvabilityByOffset
static double solvability_offset(const Level& L, const MF& r)
{
    std::vector<double> p1(L.boxes.size(), 0.0), p2(L.boxes.size(), 0.0);
    parallel_for((int)L.boxes.size(), [&](int b) {
        IBox v = valid_of(L.boxes[b], true); size_t q = 0; double s1 = 0, s2 = 0;
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i, ++q)
            if (L.owner[b][q]) { double w = L.neumann_weight(i,j,k); s1 += w*r.f[b](i,j,k); s2 += w; }
        p1[b] = s1; p2[b] = s2;
    });
    double s1 = 0, s2 = 0; for (size_t b=0;b<p1.size();++b) { s1+=p1[b]; s2+=p2[b]; }
    return s1/s2;
}
static void make_solvable(const Level& L, MF& r)
{
    double off = solvability_offset(L, r);
    parallel_for((int)L.boxes.size(), [&](int b) {
        IBox v = valid_of(L.boxes[b], true);
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i)
            r.f[b](i,j,k) -= off;
    });
}

// valid-region helpers
template <class F> static void for_valid(const Level& L, MF& a, F fn) {
    parallel_for((int)L.boxes.size(), [&](int b) {
        IBox v = valid_of(L.boxes[b], a.nodal);
        for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i) fn(b,i,j,k);
    });
}

// ---------------------------------------------------------------------------------------------
// The solver
// ---------------------------------------------------------------------------------------------
struct Options {
    int max_coarsening_level = 30;
    int maxiter = 100, bottom_maxiter = 100;
    int nu1 = 2, nu2 = 2, nuf = 8, nub = 0;
    int smooth_sweeps = 2;            // GS sweeps inside one Fsmooth (AMReX CPU path)
    int bottom_solver = BICGCG;
    double bottom_rtol = 1e-4, bottom_atol = -1.0;
    int verbose = 0;
    int smoother = LEX;
    int tile[3] = {32, 32, 1<<30};
    int interp_sigma_weighted = 1;
    int bottom_normalize = 1;         // MLCGSolver calls Lp.normalize (divide by the diagonal)
    int agg_level = -1;               // TILED only: from this MG level on, one box covers the domain
    int tile_variant = 0;             // experimental orderings, see order_key()
    int tile_z_target = 0;            // >0: z tile = pick_zchunk(nodes_z, target), the GPU driver's rule
    int tile_overlap = 1;             // tiles share their high-face nodes with the next tile (like boxes do)
    int tile_owned_only = 0;          // 1: only owned nodes are relaxed (periodic / shared high-face duplicates stay frozen)
};

struct Stats {
    int n_vcycles = 0, bottom_iters_total = 0, status = 0, n_levels = 0;
    double rhs_norm0 = 0, res_norm0 = 0, res_norm_final = 0;
    double res_history[128] = {0};
    double t_setup = 0, t_solve = 0, t_total = 0;
};

struct Solver {
    Options opt;
    std::vector<Level> lev;
    std::vector<MF> res, cor, rescor;     // per MG level (MLMG::res/cor/rescor)
    MF sol, rhs, vel, gphi;               // level 0
    bool singular = false;
    int bottom_iters = 0;
    bool bicg_failed_switch = false;

    void define(const int ncell[3], const double plo[3], const double phi_[3], const int bclo[3], const int bchi[3],
                int nboxes, const int* boxes, const Options& o)
    {
        opt = o;
        Level L0;
        L0.lo_owns = (opt.smoother == TILED);
        for (int d=0; d<3; ++d) { L0.n[d]=ncell[d]; L0.dxinv[d]=ncell[d]/(phi_[d]-plo[d]); L0.bclo[d]=bclo[d]; L0.bchi[d]=bchi[d]; }
        L0.boxes.resize(nboxes);
        for (int b=0; b<nboxes; ++b) for (int d=0; d<3; ++d) { L0.boxes[b].lo[d]=boxes[6*b+d]; L0.boxes[b].hi[d]=boxes[6*b+3+d]; }
        lev.clear(); lev.push_back(L0);
        // MLLinOp::defineGrids: coarsen by 2 while every box (and the domain) stays coarsenable with
        // mg_box_min_width = 2, up to max_coarsening_level.
        while ((int)lev.size() <= opt.max_coarsening_level) {
            const Level& F = lev.back();
            std::vector<IBox> bx = F.boxes;
            int nn[3] = {F.n[0],F.n[1],F.n[2]};
            if (opt.agg_level >= 0 && (int)lev.size() == opt.agg_level) {
                bx.assign(1, IBox{{0,0,0},{nn[0]-1,nn[1]-1,nn[2]-1}});
            }
            bool ok = true;
            for (int d=0; d<3; ++d) if (nn[d] % 2 || nn[d] < 4) ok = false;
            for (auto& q : bx) for (int d=0; d<3; ++d) {
                int len = q.hi[d]-q.lo[d]+1;
                if (len % 2 || len < 4 || (q.lo[d] % 2)) ok = false;
            }
            if (!ok) break;
            Level C = F; C.boxes = bx;
            C.sigma = MF(); C.nd_ghost.clear(); C.nd_sync.clear(); C.cc_ghost.clear(); C.owner.clear();
            for (int d=0; d<3; ++d) { C.n[d] = nn[d]/2; C.dxinv[d] = F.dxinv[d]*0.5; }
            for (auto& q : C.boxes) for (int d=0; d<3; ++d) { q.lo[d] /= 2; q.hi[d] = (q.hi[d]+1)/2 - 1; }
            lev.push_back(C);
        }
        for (auto& L : lev) { build_comm(L); L.sigma = L.make(false, 1, 1); }
        const int nl = (int)lev.size();
        res.resize(nl); cor.resize(nl); rescor.resize(nl);
        for (int l=0; l<nl; ++l) { res[l]=lev[l].make(true,1,1); cor[l]=lev[l].make(true,1,1); rescor[l]=lev[l].make(true,1,1); }
        sol = lev[0].make(true,1,1); rhs = lev[0].make(true,1,1);
        vel = lev[0].make(false,1,3); gphi = lev[0].make(false,1,3);
        singular = !lev[0].has_dirichlet();   // domain is covered by construction
    }

    // Level l's boxes may differ from level l-1's when agg_level == l (TILED only): data crossing
    // that boundary goes through global arrays.
    bool agg_boundary(int lc) const { return opt.agg_level >= 0 && lc == opt.agg_level; }

    void smooth(int l, MF& x, const MF& b, bool skip_fillboundary) {
        const Level& L = lev[l];
        apply_bc(L, x, skip_fillboundary);
        if (opt.smoother == LEX) fsmooth_lex(L, x, b, opt.smooth_sweeps);
        else { TileSpec ts{{opt.tile[0],opt.tile[1],opt.tile[2]}, opt.tile_variant, opt.tile_overlap, opt.tile_z_target, opt.tile_owned_only}; fsmooth_tiled(L, x, b, opt.smooth_sweeps, ts); }
        nodal_sync(L, x);
    }
    // MLNodeLinOp::apply = applyBC + Fapply
    void apply(int l, MF& y, MF& x) { apply_bc(lev[l], x); fapply(lev[l], y, x); }
    // y = b - A x
    void residual(int l, MF& y, MF& x, const MF& b) {
        apply(l, y, x);
        for_valid(lev[l], y, [&](int q,int i,int j,int k){ y.f[q](i,j,k) = b.f[q](i,j,k) - y.f[q](i,j,k); });
    }

    // ---- global <-> boxes (owner values) ----
    void scatter_nodal(int l, MF& m, const double* g) const {
        const Level& L = lev[l]; const size_t NX=L.n[0]+1, NY=L.n[1]+1;
        for (size_t b=0; b<L.boxes.size(); ++b) { IBox v = valid_of(L.boxes[b], true);
            for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i)
                m.f[b](i,j,k) = g[i + NX*(j + NY*(size_t)k)]; }
    }
    void gather_nodal(int l, const MF& m, double* g) const {
        const Level& L = lev[l]; const size_t NX=L.n[0]+1, NY=L.n[1]+1;
        for (size_t b=0; b<L.boxes.size(); ++b) { IBox v = valid_of(L.boxes[b], true); size_t q = 0;
            for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i, ++q)
                if (L.owner[b][q]) g[i + NX*(j + NY*(size_t)k)] = m.f[b](i,j,k); }
        // periodic high faces have no owner of their own: copy the low image
        for (int d=0; d<3; ++d) if (L.periodic(d)) {
            for (size_t k=0; k<=(size_t)L.n[2]; ++k) for (size_t j=0; j<=(size_t)L.n[1]; ++j) for (size_t i=0; i<=(size_t)L.n[0]; ++i) {
                size_t idx[3]={i,j,k}; if ((int)idx[d] != L.n[d]) continue;
                size_t s[3]={i,j,k}; s[d]=0;
                g[i+NX*(j+NY*k)] = g[s[0]+NX*(s[1]+NY*s[2])];
            }
        }
    }
    void scatter_cell(int l, MF& m, const double* g, int nc) const {
        const Level& L = lev[l]; const size_t NX=L.n[0], NY=L.n[1], NZ=L.n[2];
        for (size_t b=0; b<L.boxes.size(); ++b) { const IBox& v = L.boxes[b];
            for (int n=0;n<nc;++n) for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i)
                m.f[b](i,j,k,n) = g[i + NX*(j + NY*(k + NZ*(size_t)n))]; }
    }
    void gather_cell(int l, const MF& m, double* g, int nc) const {
        const Level& L = lev[l]; const size_t NX=L.n[0], NY=L.n[1], NZ=L.n[2];
        for (size_t b=0; b<L.boxes.size(); ++b) { const IBox& v = L.boxes[b];
            for (int n=0;n<nc;++n) for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i)
                g[i + NX*(j + NY*(k + NZ*(size_t)n))] = m.f[b](i,j,k,n); }
    }
    size_t nnodes(int l) const { return (size_t)(lev[l].n[0]+1)*(lev[l].n[1]+1)*(lev[l].n[2]+1); }
    size_t ncells(int l) const { return (size_t)lev[l].n[0]*lev[l].n[1]*lev[l].n[2]; }

    void restrict_to(int lc, MF& crse, MF& fine) {      // fine is level lc-1
        apply_bc(lev[lc-1], fine);
        if (!agg_boundary(lc)) { restriction(lev[lc-1], lev[lc], crse, fine); return; }
        // different box arrays: restrict box-wise on the fine layout, move through a global array
        Level tmpL = lev[lc]; tmpL.boxes = lev[lc-1].boxes;
        for (auto& q : tmpL.boxes) for (int d=0; d<3; ++d) { q.lo[d] /= 2; q.hi[d] = (q.hi[d]+1)/2 - 1; }
        MF t = tmpL.make(true,1,1);
        restriction(lev[lc-1], tmpL, t, fine);
        std::vector<double> g(nnodes(lc));
        Solver* self = this; (void)self;
        { // gather with tmpL's owners
            const size_t NX=tmpL.n[0]+1, NY=tmpL.n[1]+1;
            for (size_t b=0; b<tmpL.boxes.size(); ++b) { IBox v = valid_of(tmpL.boxes[b], true);
                for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i)
                    g[i + NX*(j + NY*(size_t)k)] = t.f[b](i,j,k); }
        }
        scatter_nodal(lc, crse, g.data());
    }
    void interp_to(int lf, MF& fine, MF& crse) {        // crse is level lf+1
        if (!agg_boundary(lf+1)) { interp_add(lev[lf], fine, crse, opt.interp_sigma_weighted); return; }
        std::vector<double> g(nnodes(lf+1));
        gather_nodal(lf+1, crse, g.data());
        Level tmpL = lev[lf+1]; tmpL.boxes = lev[lf].boxes;
        for (auto& q : tmpL.boxes) for (int d=0; d<3; ++d) { q.lo[d] /= 2; q.hi[d] = (q.hi[d]+1)/2 - 1; }
        MF t = tmpL.make(true,1,1);
        { const size_t NX=tmpL.n[0]+1, NY=tmpL.n[1]+1;
          for (size_t b=0; b<tmpL.boxes.size(); ++b) { IBox v = valid_of(tmpL.boxes[b], true);
            for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i)
                t.f[b](i,j,k) = g[i + NX*(j + NY*(size_t)k)]; } }
        interp_add(lev[lf], fine, t, opt.interp_sigma_weighted);
    }

    void set_sigma_levels() {
        fill_sigma_bc(lev[0], lev[0].sigma);
        for (size_t l = 1; l < lev.size(); ++l) {
            if (!agg_boundary((int)l)) { avgdown_sigma(lev[l-1], lev[l]); continue; }
            Level tmpL = lev[l]; tmpL.boxes = lev[l-1].boxes;
            for (auto& q : tmpL.boxes) for (int d=0; d<3; ++d) { q.lo[d] /= 2; q.hi[d] = (q.hi[d]+1)/2 - 1; }
            build_comm(tmpL); tmpL.sigma = tmpL.make(false,1,1);
            avgdown_sigma(lev[l-1], tmpL);
            std::vector<double> g(ncells((int)l));
            { const size_t NX=tmpL.n[0], NY=tmpL.n[1];
              for (size_t b=0; b<tmpL.boxes.size(); ++b) { const IBox& v = tmpL.boxes[b];
                for (int k=v.lo[2]; k<=v.hi[2]; ++k) for (int j=v.lo[1]; j<=v.hi[1]; ++j) for (int i=v.lo[0]; i<=v.hi[0]; ++i)
                    g[i + NX*(j + NY*(size_t)k)] = tmpL.sigma.f[b](i,j,k); } }
            scatter_cell((int)l, lev[l].sigma, g.data(), 1);
            fill_sigma_bc(lev[l], lev[l].sigma);
        }
    }

    // ---- MLCGSolver ----
    void normalize(int l, MF& x) {      // mlndlap_normalize_aa
        if (!opt.bottom_normalize) return;
        const Level& L = lev[l]; StencilCoef c(L.dxinv);
        for_valid(L, x, [&](int b,int i,int j,int k){
            if (L.is_dirichlet(i,j,k)) { x.f[b](i,j,k) = 0.0; return; }
            Sig8 g(L.sigma.f[b].ptr(i,j,k), L.sigma.f[b].sy, L.sigma.f[b].sz);
            x.f[b](i,j,k) /= diag_node(c, g);
        });
    }
    int solve_bicgstab(int l, MF& solx, const MF& b, double eps_rel, double eps_abs, int& niters) {
        const Level& L = lev[l];
        MF ph=L.make(true,1,1), sh=L.make(true,1,1), sorig=L.make(true,1,1), p=L.make(true,1,1), r=L.make(true,1,1),
           s=L.make(true,1,1), rh=L.make(true,1,1), v=L.make(true,1,1), t=L.make(true,1,1);
        residual(l, r, solx, b);            // correctionResidual
        normalize(l, r);
        sorig = solx; rh = r; solx.setVal(0.0);
        double rnorm = norm_inf(L, r); const double rnorm0 = rnorm;
        niters = 0;
        if (rnorm0 == 0 || rnorm0 < eps_abs) return 0;
        int ret = 0, iter = 1; double rho_1 = 0, alpha = 0, omega = 0;
        for (; iter <= opt.bottom_maxiter; ++iter) {
            const double rho = dot_masked(L, rh, r);
            if (rho == 0) { ret = 1; break; }
            if (iter == 1) p = r;
            else {
                const double beta = (rho/rho_1)*(alpha/omega);
                for_valid(L, p, [&](int q,int i,int j,int k){ double pv = p.f[q](i,j,k) - omega*v.f[q](i,j,k); p.f[q](i,j,k) = r.f[q](i,j,k) + beta*pv; });
            }
            ph = p;
            apply(l, v, ph); normalize(l, v);
            double rhTv = dot_masked(L, rh, v);
            if (rhTv != 0) alpha = rho/rhTv; else { ret = 2; break; }
            for_valid(L, solx, [&](int q,int i,int j,int k){ solx.f[q](i,j,k) += alpha*ph.f[q](i,j,k); s.f[q](i,j,k) = r.f[q](i,j,k) - alpha*v.f[q](i,j,k); });
            rnorm = norm_inf(L, s);
            if (rnorm < eps_rel*rnorm0 || rnorm < eps_abs) break;
            sh = s;
            apply(l, t, sh); normalize(l, t);
            double tt = dot_masked(L, t, t), tsv = dot_masked(L, t, s);
            if (tt != 0) omega = tsv/tt; else { ret = 3; break; }
            for_valid(L, solx, [&](int q,int i,int j,int k){ solx.f[q](i,j,k) += omega*sh.f[q](i,j,k); r.f[q](i,j,k) = s.f[q](i,j,k) - omega*t.f[q](i,j,k); });
            rnorm = norm_inf(L, r);
            if (rnorm < eps_rel*rnorm0 || rnorm < eps_abs) break;
            if (omega == 0) { ret = 4; break; }
            rho_1 = rho;
        }
        niters = std::min(iter, opt.bottom_maxiter);
        if (ret == 0 && rnorm > eps_rel*rnorm0 && rnorm > eps_abs) ret = 8;
        if ((ret == 0 || ret == 8) && rnorm < rnorm0) {
            for_valid(L, solx, [&](int q,int i,int j,int k){ solx.f[q](i,j,k) += sorig.f[q](i,j,k); });
        } else { solx = sorig; }
        return ret;
    }
    int solve_cg(int l, MF& solx, const MF& b, double eps_rel, double eps_abs, int& niters) {
        const Level& L = lev[l];
        MF p=L.make(true,1,1), sorig=L.make(true,1,1), r=L.make(true,1,1), q=L.make(true,1,1);
        residual(l, r, solx, b);
        sorig = solx; solx.setVal(0.0);
        double rnorm = norm_inf(L, r); const double rnorm0 = rnorm;
        niters = 0;
        if (rnorm0 == 0 || rnorm0 < eps_abs) return 0;
        int ret = 0, iter = 1; double rho_1 = 0;
        for (; iter <= opt.bottom_maxiter; ++iter) {
            double rho = dot_masked(L, r, r);
            if (rho == 0) { ret = 1; break; }
            if (iter == 1) p = r;
            else { double beta = rho/rho_1; for_valid(L, p, [&](int bb,int i,int j,int k){ p.f[bb](i,j,k) = r.f[bb](i,j,k) + beta*p.f[bb](i,j,k); }); }
            apply(l, q, p);
            double pw = dot_masked(L, p, q), alpha;
            if (pw != 0) alpha = rho/pw; else { ret = 1; break; }
            for_valid(L, solx, [&](int bb,int i,int j,int k){ solx.f[bb](i,j,k) += alpha*p.f[bb](i,j,k); r.f[bb](i,j,k) -= alpha*q.f[bb](i,j,k); });
            rnorm = norm_inf(L, r);
            if (rnorm < eps_rel*rnorm0 || rnorm < eps_abs) break;
            rho_1 = rho;
        }
        niters = std::min(iter, opt.bottom_maxiter);
        if (ret == 0 && rnorm > eps_rel*rnorm0 && rnorm > eps_abs) ret = 8;
        if ((ret == 0 || ret == 8) && rnorm < rnorm0) {
            for_valid(L, solx, [&](int bb,int i,int j,int k){ solx.f[bb](i,j,k) += sorig.f[bb](i,j,k); });
        } else { solx = sorig; }
        return ret;
    }

    // ---- MLMG ----
    void bottom_solve() {
        const int l = (int)lev.size()-1;
        MF& x = cor[l]; MF& b = res[l];
        x.setVal(0.0);
        if (opt.bottom_solver == SMOOTHER) {
            bool skip = true;
            for (int i = 0; i < opt.nuf; ++i) { smooth(l, x, b, skip); skip = false; }
            return;
        }
        MF bb = b;
        if (singular) make_solvable(lev[l], bb);
        int type = (opt.bottom_solver == CG || (opt.bottom_solver == BICGCG && bicg_failed_switch)) ? CG : BICGSTAB;
        int nit = 0;
        auto run = [&](int ty) { return ty == CG ? solve_cg(l, x, bb, opt.bottom_rtol, opt.bottom_atol, nit)
                                                 : solve_bicgstab(l, x, bb, opt.bottom_rtol, opt.bottom_atol, nit); };
        int ret = run(type);
        bottom_iters += nit;
        if (ret != 0 && ret != 9) {
            x.setVal(0.0);
            if (opt.bottom_solver == BICGCG) {
                int other = (type == CG) ? BICGSTAB : CG;
                ret = run(other); bottom_iters += nit;
                if (ret == 0) bicg_failed_switch = (other == CG);
                else x.setVal(0.0);
            }
        }
        const int n = (ret == 0) ? opt.nub : opt.nuf;
        for (int i = 0; i < n; ++i) smooth(l, x, b, false);
    }

    void vcycle() {
        const int nl = (int)lev.size();
        for (int l = 0; l < nl-1; ++l) {
            cor[l].setVal(0.0);
            bool skip = true;
            for (int i = 0; i < opt.nu1; ++i) { smooth(l, cor[l], res[l], skip); skip = false; }
            residual(l, rescor[l], cor[l], res[l]);            // computeResOfCorrection
            restrict_to(l+1, res[l+1], rescor[l]);
        }
        bottom_solve();
        for (int l = nl-2; l >= 0; --l) {
            interp_to(l, cor[l], cor[l+1]);                    // addInterpCorrection
            for (int i = 0; i < opt.nu2; ++i) smooth(l, cor[l], res[l], false);
        }
    }

    // MLMG::solve with sol = 0 initial guess
    void solve(double rtol, double atol, Stats& st) {
        const Level& L0 = lev[0];
        sol.setVal(0.0);
        bottom_iters = 0; bicg_failed_switch = false;
        if (singular) make_solvable(L0, rhs);
        residual(0, res[0], sol, rhs);
        double resnorm0 = norm_inf(L0, res[0]), rhsnorm0 = norm_inf(L0, rhs);
        double max_norm = (rhsnorm0 >= resnorm0) ? rhsnorm0 : resnorm0;
        const double res_target = std::max(atol, std::max(rtol, 1e-16)*max_norm);
        st.rhs_norm0 = rhsnorm0; st.res_norm0 = resnorm0; st.n_vcycles = 0; st.status = 0; st.res_norm_final = resnorm0;
        st.n_levels = (int)lev.size();
        if (resnorm0 <= res_target) return;
        bool converged = false;
        for (int iter = 0; iter < opt.maxiter; ++iter) {
            if (singular) make_solvable(L0, res[0]);
            vcycle();
            for_valid(L0, sol, [&](int b,int i,int j,int k){ sol.f[b](i,j,k) += cor[0].f[b](i,j,k); });
            residual(0, res[0], sol, rhs);
            double rn = norm_inf(L0, res[0]);
            if (iter < 128) st.res_history[iter] = rn;
            st.n_vcycles = iter+1; st.res_norm_final = rn;
            if (opt.verbose >= 2) std::printf("MLMG: Iteration %3d Fine resid/bnorm = %.6e\n", iter+1, rn/max_norm);
            if (rn <= res_target) { converged = true; break; }
            if (rn > 1e20*max_norm) { st.status = 2; break; }
        }
        if (!converged && st.status == 0) st.status = 1;
        st.bottom_iters_total = bottom_iters;
    }

    // NodalProjector::project
    void project(double rtol, double atol, Stats& st) {
        auto t0 = std::chrono::steady_clock::now();
        set_sigma_levels();
        // vel.FillBoundary: 1 ghost layer from neighbours / periodic images; physical ghosts stay as given
        run_copy(vel, lev[0].cc_ghost);
        divu(lev[0], rhs, vel);
        auto t1 = std::chrono::steady_clock::now();
        solve(rtol, atol, st);
        auto t2 = std::chrono::steady_clock::now();
        mknewu(lev[0], vel, gphi, sol);
        auto t3 = std::chrono::steady_clock::now();
        st.t_setup = std::chrono::duration<double>(t1-t0).count();
        st.t_solve = std::chrono::duration<double>(t2-t1).count();
        st.t_total = std::chrono::duration<double>(t3-t0).count();
    }
};

} // namespace orc

// ---------------------------------------------------------------------------------------------
// C entry points for ctypes (tests / bench cpu_baseline only)
// Global arrays: x fastest.  Nodal: (n0+1)(n1+1)(n2+1).  Cell: n0*n1*n2 [*ncomp, component slowest].
// ---------------------------------------------------------------------------------------------
using namespace orc;
extern "C" {

struct orc_options {
    int max_coarsening_level, maxiter, bottom_maxiter, nu1, nu2, nuf, nub, smooth_sweeps, bottom_solver;
    double bottom_rtol, bottom_atol;
    int verbose, smoother, tile[3], interp_sigma_weighted, bottom_normalize, agg_level, nthreads, tile_variant, tile_overlap, tile_z_target, tile_owned_only;
};
struct orc_stats {
    int n_vcycles, bottom_iters_total, status, n_levels;
    double rhs_norm0, res_norm0, res_norm_final;
    double res_history[128];
    double t_setup, t_solve, t_total;
};

void orc_default_options(orc_options* o) {
    Options d;
    o->max_coarsening_level=d.max_coarsening_level; o->maxiter=d.maxiter; o->bottom_maxiter=d.bottom_maxiter;
    o->nu1=d.nu1; o->nu2=d.nu2; o->nuf=d.nuf; o->nub=d.nub; o->smooth_sweeps=d.smooth_sweeps; o->bottom_solver=d.bottom_solver;
    o->bottom_rtol=d.bottom_rtol; o->bottom_atol=d.bottom_atol; o->verbose=0; o->smoother=LEX;
    o->tile[0]=d.tile[0]; o->tile[1]=d.tile[1]; o->tile[2]=d.tile[2];
    o->interp_sigma_weighted=1; o->bottom_normalize=1; o->agg_level=-1; o->nthreads=1; o->tile_variant=0; o->tile_overlap=1; o->tile_z_target=0; o->tile_owned_only=0;
}

void* orc_create(const int* ncell, const double* plo, const double* phi, const int* bclo, const int* bchi,
                 int nboxes, const int* boxes, const orc_options* o)
{
    Options op;
    op.max_coarsening_level=o->max_coarsening_level; op.maxiter=o->maxiter; op.bottom_maxiter=o->bottom_maxiter;
    op.nu1=o->nu1; op.nu2=o->nu2; op.nuf=o->nuf; op.nub=o->nub; op.smooth_sweeps=o->smooth_sweeps; op.bottom_solver=o->bottom_solver;
    op.bottom_rtol=o->bottom_rtol; op.bottom_atol=o->bottom_atol; op.verbose=o->verbose; op.smoother=o->smoother;
    op.tile[0]=o->tile[0]; op.tile[1]=o->tile[1]; op.tile[2]=o->tile[2];
    op.interp_sigma_weighted=o->interp_sigma_weighted; op.bottom_normalize=o->bottom_normalize; op.agg_level=o->agg_level; op.tile_variant=o->tile_variant; op.tile_overlap=o->tile_overlap; op.tile_z_target=o->tile_z_target; op.tile_owned_only=o->tile_owned_only;
    g_nthreads = std::max(1, o->nthreads);
    auto* s = new Solver();
    s->define(ncell, plo, phi, bclo, bchi, nboxes, boxes, op);
    return s;
}
void orc_destroy(void* h) { delete (Solver*)h; }
int orc_num_levels(void* h) { return (int)((Solver*)h)->lev.size(); }
void orc_level_ncell(void* h, int l, int* n) { auto* s=(Solver*)h; for (int d=0;d<3;++d) n[d]=s->lev[l].n[d]; }

void orc_set_vel(void* h, const double* u) {
    auto* s=(Solver*)h; s->vel.setVal(0.0); s->scatter_cell(0, s->vel, u, 3);
}
// Write one physical ghost cell (outside a non-periodic face) of the velocity: inflow data.
void orc_set_vel_ghost(void* h, int i, int j, int k, int n, double v) {
    auto* s=(Solver*)h; const Level& L=s->lev[0];
    for (size_t b=0;b<L.boxes.size();++b) { const IBox g = grow(L.boxes[b],1);
        if (i>=g.lo[0]&&i<=g.hi[0]&&j>=g.lo[1]&&j<=g.hi[1]&&k>=g.lo[2]&&k<=g.hi[2]) s->vel.f[b](i,j,k,n)=v; }
}
void orc_set_sigma(void* h, const double* sg) { auto* s=(Solver*)h; s->scatter_cell(0, s->lev[0].sigma, sg, 1); }
void orc_prepare_sigma(void* h) { ((Solver*)h)->set_sigma_levels(); }
void orc_get_sigma(void* h, int l, double* sg) { auto* s=(Solver*)h; s->gather_cell(l, s->lev[l].sigma, sg, 1); }
void orc_project(void* h, double rtol, double atol, orc_stats* out) {
    auto* s=(Solver*)h; Stats st; s->project(rtol, atol, st);
    out->n_vcycles=st.n_vcycles; out->bottom_iters_total=st.bottom_iters_total; out->status=st.status; out->n_levels=st.n_levels;
    out->rhs_norm0=st.rhs_norm0; out->res_norm0=st.res_norm0; out->res_norm_final=st.res_norm_final;
    std::memcpy(out->res_history, st.res_history, sizeof(st.res_history));
    out->t_setup=st.t_setup; out->t_solve=st.t_solve; out->t_total=st.t_total;
}
void orc_get_vel(void* h, double* u) { auto* s=(Solver*)h; s->gather_cell(0, s->vel, u, 3); }
void orc_get_gradphi(void* h, double* g) { auto* s=(Solver*)h; s->gather_cell(0, s->gphi, g, 3); }
void orc_get_phi(void* h, double* p) { auto* s=(Solver*)h; s->gather_nodal(0, s->sol, p); }
void orc_get_rhs(void* h, double* p) { auto* s=(Solver*)h; s->gather_nodal(0, s->rhs, p); }

// ---- kernel-level hooks (operate on MG level l with global nodal arrays) ----
void orc_divu(void* h, double* rhs_out) {       // uses the velocity set by orc_set_vel (+ghosts)
    auto* s=(Solver*)h; run_copy(s->vel, s->lev[0].cc_ghost); divu(s->lev[0], s->rhs, s->vel); s->gather_nodal(0, s->rhs, rhs_out);
}
void orc_apply(void* h, int l, const double* x, double* y) {
    auto* s=(Solver*)h; MF X=s->lev[l].make(true,1,1), Y=s->lev[l].make(true,1,1);
    s->scatter_nodal(l, X, x); s->apply(l, Y, X); s->gather_nodal(l, Y, y);
}
void orc_smooth(void* h, int l, double* x, const double* rhs, int ncalls) {
    auto* s=(Solver*)h; MF X=s->lev[l].make(true,1,1), B=s->lev[l].make(true,1,1);
    s->scatter_nodal(l, X, x); s->scatter_nodal(l, B, rhs);
    for (int i=0;i<ncalls;++i) s->smooth(l, X, B, false);
    s->gather_nodal(l, X, x);
}
void orc_restrict(void* h, int lc, const double* fine, double* crse) {
    auto* s=(Solver*)h; MF F=s->lev[lc-1].make(true,1,1), C=s->lev[lc].make(true,1,1);
    s->scatter_nodal(lc-1, F, fine); s->restrict_to(lc, C, F); s->gather_nodal(lc, C, crse);
}
void orc_interp_add(void* h, int lf, double* fine, const double* crse) {
    auto* s=(Solver*)h; MF F=s->lev[lf].make(true,1,1), C=s->lev[lf+1].make(true,1,1);
    s->scatter_nodal(lf, F, fine); s->scatter_nodal(lf+1, C, crse); s->interp_to(lf, F, C); s->gather_nodal(lf, F, fine);
}
void orc_mknewu(void* h, const double* phi, double* u_inout, double* gphi_out) {
    auto* s=(Solver*)h; s->scatter_nodal(0, s->sol, phi); s->scatter_cell(0, s->vel, u_inout, 3);
    mknewu(s->lev[0], s->vel, s->gphi, s->sol); s->gather_cell(0, s->vel, u_inout, 3); s->gather_cell(0, s->gphi, gphi_out, 3);
}
double orc_dot(void* h, int l, const double* x, const double* y) {
    auto* s=(Solver*)h; MF X=s->lev[l].make(true,1,1), Y=s->lev[l].make(true,1,1);
    s->scatter_nodal(l, X, x); s->scatter_nodal(l, Y, y); return dot_masked(s->lev[l], X, Y);
}
// bottom solve alone: x = bottom_solve(b) on the coarsest level
void orc_bottom_solve(void* h, const double* b, double* x, int* iters) {
    auto* s=(Solver*)h; int l=(int)s->lev.size()-1;
    s->scatter_nodal(l, s->res[l], b); s->bottom_iters = 0; s->bottom_solve(); s->gather_nodal(l, s->cor[l], x); *iters = s->bottom_iters;
}
// one V-cycle on a given level-0 residual: cor = V(res)
void orc_vcycle(void* h, const double* res0, double* cor0) {
    auto* s=(Solver*)h; s->scatter_nodal(0, s->res[0], res0);
    if (s->singular) make_solvable(s->lev[0], s->res[0]);
    s->vcycle(); s->gather_nodal(0, s->cor[0], cor0);
}

} // extern "C"
