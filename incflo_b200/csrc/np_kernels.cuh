// np_kernels.cuh — the fp64 kernels of the nodal projection, written for sm_100a.
//
// Kernel families:
//   K1 k_divu            nodal divergence of the cell velocity     (amrex::mlndlap_divu + impose_neumann_bc)
//   K2 k_stencil         y = A x  /  y = (b - off) - A x marching in z, constant or variable sigma (mlndlap_adotx_aa)
//   K3 k_gs2             Gauss-Seidel smoother, two sweeps fused in one pass (mlndlap_gauss_seidel_aa)
//   K4 k_restrict        full-weighting restriction                (mlndlap_restriction)
//   K5 k_interp_add      prolongation + add                        (mlndlap_interpadd_aa)
//   K6 k_bottom_*        BiCGStab / CG on the coarsest level in one CTA (MLCGSolver)
//   K7 k_mknewu          u -= sigma grad(phi), gradphi = grad(phi) (mlndlap_mknewu + NodalProjector::project)
//   K8 k_avgdown_sigma   8->1 arithmetic coarsening of sigma, k_fill_sigma ghost fill
// plus boundary fill (FillBoundary + nodalSync + mlndlap_applybc in one pass), reductions and copies.
// The reference has no file for any of these (the arithmetic is in AMReX); the call site they serve is
// /root/reference/src/projection/incflo_apply_nodal_projection.cpp:153.
#pragma once
#include "np_types.h"
#include <math.h>

// ------------------------------------------------------------------------------------------------
// Stencil on generic accessors.  X(di,dj,dk): nodal value at offset from the node.  S(a,b,c): sigma of the
// cell at offset a,b,c in {-1,0} from CELL (i,j,k).
// ------------------------------------------------------------------------------------------------
struct Sig8 {
    double mmm, pmm, mpm, ppm, mmp, pmp, mpp, ppp;   // x,y,z side: m = -1, p = 0
};
template <class SA> __device__ __forceinline__ Sig8 np_load_sig(const SA& s) {
    Sig8 g;
    g.mmm = s(-1,-1,-1); g.pmm = s(0,-1,-1); g.mpm = s(-1,0,-1); g.ppm = s(0,0,-1);
    g.mmp = s(-1,-1,0);  g.pmp = s(0,-1,0);  g.mpp = s(-1,0,0);  g.ppp = s(0,0,0);
    return g;
}
__device__ __forceinline__ double np_diag(const NpCoef& c, const Sig8& g) {
    return -4.0 * c.fxyz * (((g.mmm + g.pmm) + (g.mpm + g.ppm)) + ((g.mmp + g.pmp) + (g.mpp + g.ppp)));
}
template <class XA> __device__ __forceinline__ double np_offdiag(const NpCoef& c, const Sig8& g, const XA& x) {
    double corners = x(-1,-1,-1)*g.mmm + x(1,-1,-1)*g.pmm + x(-1,1,-1)*g.mpm + x(1,1,-1)*g.ppm
                   + x(-1,-1, 1)*g.mmp + x(1,-1, 1)*g.pmp + x(-1,1, 1)*g.mpp + x(1,1, 1)*g.ppp;
    double eyz = x(0,-1,-1)*(g.mmm+g.pmm) + x(0,1,-1)*(g.mpm+g.ppm) + x(0,-1,1)*(g.mmp+g.pmp) + x(0,1,1)*(g.mpp+g.ppp);
    double exz = x(-1,0,-1)*(g.mmm+g.mpm) + x(1,0,-1)*(g.pmm+g.ppm) + x(-1,0,1)*(g.mmp+g.mpp) + x(1,0,1)*(g.pmp+g.ppp);
    double exy = x(-1,-1,0)*(g.mmm+g.mmp) + x(1,-1,0)*(g.pmm+g.pmp) + x(-1,1,0)*(g.mpm+g.mpp) + x(1,1,0)*(g.ppm+g.ppp);
    double fx = x(-1,0,0)*((g.mmm+g.mpm)+(g.mmp+g.mpp)) + x(1,0,0)*((g.pmm+g.ppm)+(g.pmp+g.ppp));
    double fy = x(0,-1,0)*((g.mmm+g.pmm)+(g.mmp+g.pmp)) + x(0,1,0)*((g.mpm+g.ppm)+(g.mpp+g.ppp));
    double fz = x(0,0,-1)*((g.mmm+g.pmm)+(g.mpm+g.ppm)) + x(0,0,1)*((g.mmp+g.pmp)+(g.mpp+g.ppp));
    return c.fxyz*corners + c.fmx2y2z*eyz + c.f2xmy2z*exz + c.f2x2ymz*exy
         + c.f4xm2ym2z*fx + c.fm2x4ym2z*fy + c.fm2xm2y4z*fz;
}
// constant sigma: the 26 neighbours collapse into 7 weight classes (weights already multiplied by sigma)
template <class XA> __device__ __forceinline__ double np_offdiag_const(const NpCoef& c, double sg, const XA& x) {
    double corners = ((x(-1,-1,-1) + x(1,-1,-1)) + (x(-1,1,-1) + x(1,1,-1))) + ((x(-1,-1,1) + x(1,-1,1)) + (x(-1,1,1) + x(1,1,1)));
    double eyz = (x(0,-1,-1) + x(0,1,-1)) + (x(0,-1,1) + x(0,1,1));
    double exz = (x(-1,0,-1) + x(1,0,-1)) + (x(-1,0,1) + x(1,0,1));
    double exy = (x(-1,-1,0) + x(1,-1,0)) + (x(-1,1,0) + x(1,1,0));
    double fx = x(-1,0,0) + x(1,0,0), fy = x(0,-1,0) + x(0,1,0), fz = x(0,0,-1) + x(0,0,1);
    return sg * (c.fxyz*corners + 2.0*(c.fmx2y2z*eyz + c.f2xmy2z*exz + c.f2x2ymz*exy)
               + 4.0*(c.f4xm2ym2z*fx + c.fm2x4ym2z*fy + c.fm2xm2y4z*fz));
}
__device__ __forceinline__ double np_diag_const(const NpCoef& c, double sg) { return -32.0 * c.fxyz * sg; }

struct GlobX {   // nodal array in HBM, pointer at the node
    const double* p; long long sy, sz;
    __device__ __forceinline__ double operator()(int di, int dj, int dk) const { return p[di + sy*dj + sz*dk]; }
};
struct GlobS {   // cell array in HBM, pointer at cell (i,j,k)
    const double* p; long long sy, sz;
    __device__ __forceinline__ double operator()(int a, int b, int c) const { return p[a + sy*b + sz*c]; }
};

// ------------------------------------------------------------------------------------------------
// Block reductions (warp shuffles, then one smem pass; every thread of the block must call them)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double np_warp_sum(double v) {
    for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
    return v;
}
__device__ __forceinline__ double np_warp_max(double v) {
    for (int m = 16; m > 0; m >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, m));
    return v;
}
// result valid in every thread; fixed summation order -> deterministic
// running max of |v| that does not lose a NaN: fmax() drops NaN operands, so a NaN entry becomes +Inf, which every later
// fmax keeps and the host turns into NP_ERR_DIVERGED (AMReX's MLMG fails on such input)
__device__ __forceinline__ double np_absmax(double m, double v) {
    const double a = fabs(v);
    return (a == a) ? fmax(m, a) : INFINITY;
}
template <bool IS_MAX> __device__ __forceinline__ double np_block_reduce(double v) {
    __shared__ double red[33];
    const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
    const int nth = blockDim.x * blockDim.y * blockDim.z;
    v = IS_MAX ? np_warp_max(v) : np_warp_sum(v);
    __syncthreads();                       // protect red[] from a previous call
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    if (tid == 0) {
        double a = red[0];
        for (int w = 1; w < (nth + 31) / 32; ++w) a = IS_MAX ? fmax(a, red[w]) : a + red[w];
        red[32] = a;
    }
    __syncthreads();
    return red[32];
}

// ------------------------------------------------------------------------------------------------
// Boundary fill: FillBoundary(periodic) + nodalSync (high duplicates <- owner) + mlndlap_applybc
// (Neumann mirror) in ONE pass over the shell.  Sources are owned nodes, destinations are not, so the
// pass is race-free.  zmode 0: z handled like x,y.  zmode 1: rank-interior z faces are filled by the
// halo exchange; only x,y (on planes lo_k..hi_k) and domain-end z mirrors are done here.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_fill_boundary(LevelDev L, double* x)
{
    const int e0 = L.n[0] + 3, e1 = L.n[1] + 3, e2 = L.n[2] + 3;     // extended extents
    // three slab groups of 3 planes each: i in {-1,n,n+1}, j in {...}, k in {...}
    const long long g0 = 3LL * e1 * e2, g1 = 3LL * e0 * e2, g2 = 3LL * e0 * e1;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < g0 + g1 + g2;
         t += (long long)gridDim.x * blockDim.x) {
        int i, j, k;
        if (t < g0)           { long long q = t;           int s = (int)(q % 3); q /= 3; j = (int)(q % e1) - 1; k = (int)(q / e1) - 1; i = s == 0 ? -1 : L.n[0] + s - 1; }
        else if (t < g0 + g1) { long long q = t - g0;      int s = (int)(q % 3); q /= 3; i = (int)(q % e0) - 1; k = (int)(q / e0) - 1; j = s == 0 ? -1 : L.n[1] + s - 1; }
        else                  { long long q = t - g0 - g1; int s = (int)(q % 3); q /= 3; i = (int)(q % e0) - 1; j = (int)(q / e0) - 1; k = s == 0 ? -1 : L.n[2] + s - 1; }
        const int si = np_map_index(i, L.n[0], L.map_lo[0], L.map_hi[0]);
        const int sj = np_map_index(j, L.n[1], L.map_lo[1], L.map_hi[1]);
        const int sk = np_map_index(k, L.n[2], L.map_lo[2], L.map_hi[2]);
        if (si == i && sj == j && sk == k) continue;
        // a source must be a valid node; (mapped) indices that stay outside belong to Dirichlet /
        // rank-interior faces and are left alone
        if (si < 0 || si > L.n[0] || sj < 0 || sj > L.n[1] || sk < 0 || sk > L.n[2]) continue;
        x[np_noff(L, i, j, k)] = x[np_noff(L, si, sj, sk)];
    }
}

// The x / y part of k_fill_boundary for nk planes starting at k0 only: what the planes handed to a neighbour rank need
// before they leave (the halo exchange sends whole planes, ghost columns included).
__global__ void __launch_bounds__(256) k_fill_xy_planes(LevelDev L, double* x, int k0, int nk)
{
    const int e0 = L.n[0] + 3, e1 = L.n[1] + 3;
    const long long per = 3LL * e1 + 3LL * e0;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < per * nk; t += (long long)gridDim.x * blockDim.x) {
        const int k = k0 + (int)(t / per);
        long long q = t % per;
        int i, j;
        if (q < 3LL * e1) { const int s = (int)(q % 3); j = (int)(q / 3) - 1; i = s == 0 ? -1 : L.n[0] + s - 1; }
        else { q -= 3LL * e1; const int s = (int)(q % 3); i = (int)(q / 3) - 1; j = s == 0 ? -1 : L.n[1] + s - 1; }
        const int si = np_map_index(i, L.n[0], L.map_lo[0], L.map_hi[0]);
        const int sj = np_map_index(j, L.n[1], L.map_lo[1], L.map_hi[1]);
        if (si == i && sj == j) continue;
        if (si < 0 || si > L.n[0] || sj < 0 || sj > L.n[1]) continue;
        x[np_noff(L, i, j, k)] = x[np_noff(L, si, sj, k)];
    }
}

// ------------------------------------------------------------------------------------------------
// Halo exchange over peer memory (NVLink 5 / NVSwitch): this rank STORES its boundary planes straight into the ghost
// planes of the neighbours' arrays (opened through CUDA IPC) and the ranks pace each other with four flags per rank
// that the neighbours write: ready_from_{lo,hi} ("I no longer read the ghost planes you filled last time") and
// done_from_{lo,hi} ("your ghost planes of this exchange have landed").  One launch per exchange replaces a grouped
// ncclSend/ncclRecv (about 25 us of latency each on the small MG levels).  Every rank runs the same sequence of
// exchanges, so a plain epoch counter identifies them.
// ------------------------------------------------------------------------------------------------
enum { NP_FLAG_READY_LO = 0, NP_FLAG_READY_HI = 1, NP_FLAG_DONE_LO = 2, NP_FLAG_DONE_HI = 3 };
struct HaloP2P {
    const double* send_lo; double* dst_lo; long long n_lo;     // my planes for the lower neighbour -> its array
    const double* send_hi; double* dst_hi; long long n_hi;
    unsigned long long* mine;                                  // my four flags (written by the neighbours)
    unsigned long long* lo_flags; unsigned long long* hi_flags;   // the neighbours' flag arrays (null: no neighbour)
    unsigned long long* epoch_word;                            // my count of finished exchanges (device memory, so that a CUDA
                                                               // graph can replay the launch: nothing changes in its arguments);
                                                               // epoch_word[1]: error word (first exchange a wait gave up on)
    unsigned int* counter;                                     // CTAs of this launch that finished their stores
};
__device__ __forceinline__ void np_flag_store(unsigned long long* p, unsigned long long v)
{
#ifdef NP_EMU
    *p = v;
#else
    asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
#endif
}
__device__ __forceinline__ unsigned long long np_flag_load(const unsigned long long* p)
{
#ifdef NP_EMU
    return *p;
#else
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
#endif
}
// Wait until *p >= epoch.  Every rank runs the same sequence of exchanges, so the wait is short; if a neighbour never gets there
// (it returned early on an error of its own, or died), the spin gives up after NP_SPIN_BUDGET clocks (about ten seconds),
// records the exchange in this rank's error word (checked by the host after the solve: NP_ERR_NCCL) and every later wait of
// this rank returns at once, so that the stream drains instead of hanging the GPU.
#define NP_SPIN_BUDGET 20000000000ll
__device__ __forceinline__ void np_wait_flag(const unsigned long long* p, unsigned long long epoch, unsigned long long* errw)
{
    if (*(volatile unsigned long long*)errw != 0ull) return;
#ifdef NP_EMU
    (void)p; (void)epoch;
#else
    const long long t0 = clock64();
    unsigned n = 0;
    while (np_flag_load(p) < epoch) {
        if ((++n & 4095u) == 0u && clock64() - t0 > NP_SPIN_BUDGET) { *(volatile unsigned long long*)errw = epoch; __threadfence_system(); return; }
    }
#endif
}
__global__ void __launch_bounds__(256) k_halo_p2p(HaloP2P a)
{
    // the number of this exchange: every rank runs the same sequence of exchanges, so the counts agree.  The word is advanced
    // by the last CTA to finish, i.e. after every CTA of this launch has read it.
    unsigned long long epoch = 0;
    if (threadIdx.x == 0) epoch = *(volatile unsigned long long*)a.epoch_word + 1ull;
    // 1) this kernel runs after everything that read my ghost planes: tell the neighbours they may overwrite them
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (a.lo_flags) np_flag_store(a.lo_flags + NP_FLAG_READY_HI, epoch);
        if (a.hi_flags) np_flag_store(a.hi_flags + NP_FLAG_READY_LO, epoch);
    }
    // 2) wait until the neighbours have said the same about theirs
    if (threadIdx.x == 0) {
        if (a.lo_flags) np_wait_flag(a.mine + NP_FLAG_READY_LO, epoch, a.epoch_word + 1);
        if (a.hi_flags) np_wait_flag(a.mine + NP_FLAG_READY_HI, epoch, a.epoch_word + 1);
    }
    __syncthreads();
    // 3) my boundary planes -> their ghost planes (16-byte stores; plane sizes are even)
    const long long t0 = blockIdx.x * (long long)blockDim.x + threadIdx.x, nt = (long long)gridDim.x * blockDim.x;
    if (a.lo_flags) {
        const double2* s2 = reinterpret_cast<const double2*>(a.send_lo); double2* d2 = reinterpret_cast<double2*>(a.dst_lo);
        for (long long q = t0; q < a.n_lo / 2; q += nt) d2[q] = s2[q];
    }
    if (a.hi_flags) {
        const double2* s2 = reinterpret_cast<const double2*>(a.send_hi); double2* d2 = reinterpret_cast<double2*>(a.dst_hi);
        for (long long q = t0; q < a.n_hi / 2; q += nt) d2[q] = s2[q];
    }
    // 4) the last CTA to finish publishes "done" and 5) waits for the neighbours' planes to have landed here
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int prev = atomicAdd(a.counter, 1u);
        if (prev == gridDim.x - 1) {
            *a.counter = 0u;
            *(volatile unsigned long long*)a.epoch_word = epoch;
            __threadfence_system();
            if (a.lo_flags) np_flag_store(a.lo_flags + NP_FLAG_DONE_HI, epoch);
            if (a.hi_flags) np_flag_store(a.hi_flags + NP_FLAG_DONE_LO, epoch);
            if (a.lo_flags) np_wait_flag(a.mine + NP_FLAG_DONE_LO, epoch, a.epoch_word + 1);
            if (a.hi_flags) np_wait_flag(a.mine + NP_FLAG_DONE_HI, epoch, a.epoch_word + 1);
        }
    }
}

// sigma ghost layer: periodic wrap, else copy of the adjacent interior cell (MLNodeLaplacian::FillBoundaryCoeff)
// fill_z = 0 leaves the z ghost layers to the halo exchange (rank-interior faces).
__global__ void __launch_bounds__(256) k_fill_sigma(LevelDev L, double* s, int per0, int per1, int per2, int zlo_local, int zhi_local)
{
    const int e0 = L.n[0] + 2, e1 = L.n[1] + 2, e2 = L.n[2] + 2;
    const long long g0 = 2LL * e1 * e2, g1 = 2LL * e0 * e2, g2 = 2LL * e0 * e1;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < g0 + g1 + g2;
         t += (long long)gridDim.x * blockDim.x) {
        int i, j, k;
        if (t < g0)           { long long q = t;           int sd = (int)(q % 2); q /= 2; j = (int)(q % e1) - 1; k = (int)(q / e1) - 1; i = sd ? L.n[0] : -1; }
        else if (t < g0 + g1) { long long q = t - g0;      int sd = (int)(q % 2); q /= 2; i = (int)(q % e0) - 1; k = (int)(q / e0) - 1; j = sd ? L.n[1] : -1; }
        else                  { long long q = t - g0 - g1; int sd = (int)(q % 2); q /= 2; i = (int)(q % e0) - 1; j = (int)(q / e0) - 1; k = sd ? L.n[2] : -1; }
        int si = i, sj = j, sk = k;
        if (i < 0) si = per0 ? i + L.n[0] : 0; else if (i >= L.n[0]) si = per0 ? i - L.n[0] : L.n[0] - 1;
        if (j < 0) sj = per1 ? j + L.n[1] : 0; else if (j >= L.n[1]) sj = per1 ? j - L.n[1] : L.n[1] - 1;
        if (k < 0)          { if (!zlo_local) continue; sk = per2 ? k + L.n[2] : 0; }
        else if (k >= L.n[2]) { if (!zhi_local) continue; sk = per2 ? k - L.n[2] : L.n[2] - 1; }
        s[np_coff(L, i, j, k)] = s[np_coff(L, si, sj, sk)];
    }
}

// final stage after a reducing kernel: scal[0] = max(part_max), scal[2] = sum(part_sum) (fixed order),
// and with want_off scal[6] = scal[2] / wtot, the solvability offset (MLNodeLinOp::fixSolvabilityByOffset)
__global__ void __launch_bounds__(1024) k_reduce_norm_off(const double* __restrict__ part_max, const double* __restrict__ part_sum,
                                                          long long n, double* scal, double wtot, int want_off)
{
    double m = 0.0, sm = 0.0;
    for (long long t = threadIdx.x; t < n; t += blockDim.x) { m = fmax(m, part_max[t]); if (want_off) sm += part_sum[t]; }
    m = np_block_reduce<true>(m);
    sm = np_block_reduce<false>(sm);
    if (threadIdx.x == 0) { scal[0] = m; if (want_off) { scal[2] = sm; scal[6] = sm / wtot; } }
}
__global__ void k_set_offset(double* scal, double wtot) { if (threadIdx.x == 0 && blockIdx.x == 0) scal[6] = scal[2] / wtot; }

// second stage of a reduction: one block folds `n` partials into out[slot] (max or sum), fixed order
template <bool IS_MAX> __global__ void __launch_bounds__(256) k_reduce_final(const double* __restrict__ part, long long n, double* out, int slot)
{
    double v = 0.0;
    for (long long t = threadIdx.x; t < n; t += blockDim.x) v = IS_MAX ? fmax(v, part[t]) : v + part[t];
    v = np_block_reduce<IS_MAX>(v);
    if (threadIdx.x == 0) out[slot] = v;
}

// inf-norm partials of a nodal array over valid nodes
__global__ void __launch_bounds__(256) k_norminf(LevelDev L, const double* __restrict__ x, double* __restrict__ block_max)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    double v = 0.0;
    if (i <= L.n[0] && j <= L.n[1]) v = fabs(x[np_noff(L, i, j, k)]);
    double m = np_block_reduce<true>(v);
    if (threadIdx.x == 0 && threadIdx.y == 0)
        block_max[blockIdx.x + gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z)] = m;
}

// dot-mask weighted sums over OWNED nodes: part0 = sum w*x*y (y may be null -> sum w*x), part1 = sum w
__global__ void __launch_bounds__(256) k_wdot(LevelDev L, const double* __restrict__ x, const double* __restrict__ y,
                                              double* __restrict__ part0, double* __restrict__ part1)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    double a = 0.0, w = 0.0;
    if (i <= L.owned_hi[0] && j <= L.owned_hi[1] && k <= L.owned_hi[2]) {
        w = np_neumann_weight(L, i, j, k);
        const long long o = np_noff(L, i, j, k);
        a = w * x[o] * (y ? y[o] : 1.0);
    }
    double sa = np_block_reduce<false>(a);
    double sw = np_block_reduce<false>(w);
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        const long long bidx = blockIdx.x + gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z);
        part0[bidx] = sa; part1[bidx] = sw;
    }
}

// x[valid] -= scal[0]/scal[1]   (MLNodeLinOp::fixSolvabilityByOffset; scalars stay on the device)
__global__ void __launch_bounds__(256) k_sub_offset(LevelDev L, double* __restrict__ x, const double* __restrict__ scal)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i <= L.n[0] && j <= L.n[1]) x[np_noff(L, i, j, k)] -= scal[0] / scal[1];
}

// y[valid] += x[valid]
__global__ void __launch_bounds__(256) k_add(LevelDev L, double* __restrict__ y, const double* __restrict__ x)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i <= L.n[0] && j <= L.n[1]) { const long long o = np_noff(L, i, j, k); y[o] += x[o]; }
}

// ------------------------------------------------------------------------------------------------
// K3: the smoother.  DESIGN.md "smoother semantics":
//   * the valid nodes of the box are cut into tiles that WRITE (63 x 31 x zc) nodes and RELAX one more
//     node on each high side (shared with the next tile, which owns it) — the same one-node overlap
//     AMReX's boxes have — with everything outside the relaxed region frozen at its pre-smooth value;
//   * inside a tile: sweep 1 then sweep 2 (the two sweeps of one Fsmooth), each visiting planes k
//     ascending and, inside a plane, the four colours c = 2*(j&1)+(i&1) ascending (Gauss-Seidel).
// One CTA owns one tile column chunk and streams it through shared memory ONCE: warps 0-3 run sweep 1
// on plane t while warps 4-7 run sweep 2 on plane t-2 (which only needs sweep-1 values of planes
// <= t-1), so both sweeps cost one pass over HBM.  A ring of six planes lives in shared memory, filled
// by cp.async one plane ahead; every plane holds the frozen halo and the latest in-tile values.
// Each thread owns a 2 x 8 (x,y) block of nodes in registers.  Planes are stored with even and odd
// columns in separate arrays so that the own pair, the left and the right halo column are all
// conflict-free 64-bit shared loads.
// Constant sigma: x_new = B - (a_fx*(W+E) + a_fy*(S+N) + a_exy*(4 in-plane diagonals)), where
// B = rhs/diag - (9-point stencil of x[k-1]+x[k+1]) is formed once per plane before the colour loop
// and all weights are pre-divided by the diagonal on the host (no division in the kernel).
// ------------------------------------------------------------------------------------------------
#define GS_TX 64                 // relaxed nodes per tile in x
#define GS_TY 32                 // relaxed nodes per tile in y
#define GS_WX 63                 // written nodes per tile in x
#define GS_WY 31
#define GS_HP 34                 // pitch of a half-row (even or odd columns)
#define GS_HALF (34 * 34 + 4)    // doubles per half-plane; +4 puts the odd array 16 banks away from the even one
#define GS_PLANE (2 * GS_HALF)
#define GS_RING 6
#define GS_SMEM_BYTES (GS_RING * GS_PLANE * 8)

struct GsCoef { double a_fx, a_fy, a_exy, a_fz, a_exz, a_eyz, a_c, dinv; };
struct GsArgs {
    GsCoef c;
    double* xout; const double* xin; const double* rhs; const double* sig;
    const double* boff;      // optional device scalar subtracted from rhs (lazy makeSolvable), or null
    const double* addto;     // optional: xout = addto + result (xout may alias addto), or null
    int zc, zero_in;         // z-chunk; 1: xin is known to be zero and is not read
    int zmap0, zmapS;        // z-chunk of a CTA = zmap0 + blockIdx.z * zmapS (a launch may cover only the end chunks)
};

__device__ __forceinline__ void np_cp_async8(double* smem_dst, const double* gsrc, bool valid)
{
#ifdef NP_EMU
    *smem_dst = valid ? *gsrc : 0.0;
#else
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    const int sz = valid ? 8 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" :: "r"(d), "l"(gsrc), "r"(sz) : "memory");
#endif
}
__device__ __forceinline__ void np_cp_async8_all(double* smem_dst, const double* gsrc)
{
#ifdef NP_EMU
    *smem_dst = *gsrc;
#else
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" :: "r"(d), "l"(gsrc) : "memory");
#endif
}
__device__ __forceinline__ void np_prefetch_l1(const void* p)
{
#ifndef NP_EMU
    asm volatile("prefetch.global.L1 [%0];\n" :: "l"(p));
#endif
}
// barrier among the 128 threads of one warp group (named barrier 1 + grp); the emulation's barrier is CTA-wide
__device__ __forceinline__ void np_group_barrier(int grp)
{
#ifdef NP_EMU
    __syncthreads();
#else
    if (grp == 0) asm volatile("bar.sync 1, 128;\n" ::: "memory");
    else          asm volatile("bar.sync 2, 128;\n" ::: "memory");
#endif
}
__device__ __forceinline__ void np_cp_async_commit()
{
#ifndef NP_EMU
    asm volatile("cp.async.commit_group;\n" ::: "memory");
#endif
}
__device__ __forceinline__ void np_cp_async_wait_all()
{
#ifndef NP_EMU
    asm volatile("cp.async.wait_all;\n" ::: "memory");
#endif
}

// Cooperative plane loader shared by k_gs2 and k_stencil: 198 of the 256 threads each own one of the 66
// columns (tile halo included) of three interleaved row sets, so a plane costs 12 cp.async per thread with
// one add per address.  Source offsets are relative to the plane, destinations are split-layout slots.
struct GsLoader {
    int src0, dst0, nrow_ok;     // first element; rows (counted from this thread's first) that exist
    __device__ __forceinline__ void init(const LevelDev& L, int tid, int i0, int j0, bool zero_in) {
        const int c = tid % 66, r0 = tid / 66;           // column slot g = c+1, first row r0 (0..2); tid >= 198 idle
        const int g = c + 1, i = i0 + g - 2;
        // row index `row` (0..33) is node j = j0 + row - 1; it exists when j <= n1 + 1
        const int rows_exist = min(34, L.n[1] + 3 - j0);            // rows 0 .. rows_exist-1 exist
        nrow_ok = (i <= L.n[0] + 1 && !zero_in) ? rows_exist : 0;
        src0 = (int)(L.nsy * (long long)(j0 + r0) + (i + 1));      // row r0 -> j+1 = j0 + r0
        dst0 = ((g & 1) ? GS_HALF : 0) + r0 * GS_HP + (g >> 1);
        row_first = (tid < 198) ? r0 : 34;              // idle threads: no row passes the range test below
    }
    int row_first;
    // ALLVALID: every row and column of the tile plane exists (full tile, input is read): no zero-fill, no range tests
    template <bool ALLVALID>
    __device__ __forceinline__ void load(double* dst_plane, const double* src_plane, long long nsy) const {
        if (row_first >= 34) return;                         // idle threads: skip (a branch; warp 7 never enters)
        const double* sp = src_plane + src0;
        double* dp = dst_plane + dst0;
        const long long step = 3 * nsy;
#pragma unroll
        for (int it = 0; it < 12; ++it) {
            if (it < 11 || row_first == 0) {                 // rows row_first + 3*it < 34
                if (ALLVALID) np_cp_async8_all(dp + 3 * GS_HP * it, sp);
                else {
                    const bool valid = row_first + 3 * it < nrow_ok;
                    np_cp_async8(dp + 3 * GS_HP * it, valid ? sp : src_plane, valid);
                }
            }
            sp += step;
        }
    }
};

// The four colour steps of one plane for a thread's 2 x 8 block (constant sigma).  MASKED: the tile has
// nodes outside the domain or on Dirichlet faces, which keep their value / are set to zero.
template <bool MASKED>
__device__ __forceinline__ void gs_colours_cs(bool on, double (&x)[2][8], const double (&B)[2][8], double* Pc, const GsCoef& C,
                                              int ofs, unsigned act, unsigned dmask, int grp)
{
    // ofs = row0 * GS_HP + lane: (row r, own even column) is Pc[ofs + r*GS_HP + 1], own odd column adds GS_HALF,
    // left halo column Pc[GS_HALF + ofs + r*GS_HP], right halo column Pc[ofs + r*GS_HP + 2]
    double t0 = 0.0, t1 = 0.0, d0 = 0.0, d1 = 0.0;
    const double nfx = -C.a_fx, nfy = -C.a_fy, nxy = -C.a_exy;
    // ---- colour 0: a = 0, r even ----
    if (on) {
        double xl[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) xl[q] = Pc[GS_HALF + ofs + (q - 1) * GS_HP];              // rows -1..7
        t0 = Pc[ofs - GS_HP + 1]; t1 = Pc[GS_HALF + ofs - GS_HP + 1];
#pragma unroll
        for (int r = 0; r < 8; r += 2) {
            const double up0 = r == 0 ? t0 : x[0][r - 1], up1 = r == 0 ? t1 : x[1][r - 1];
            const double v = fma(nxy, (xl[r] + xl[r + 2]) + (up1 + x[1][r + 1]), fma(nfy, up0 + x[0][r + 1], fma(nfx, xl[r + 1] + x[1][r], B[0][r])));
            if (!MASKED) x[0][r] = v; else if ((act >> r) & 1u) x[0][r] = ((dmask >> r) & 1u) ? 0.0 : v;
            Pc[ofs + r * GS_HP + 1] = x[0][r];
        }
    }
    __syncwarp();            // the next colour's new in-plane neighbours sit in the same row, i.e. the same warp
    // ---- colour 1: a = 1, r even ----
    if (on) {
        double xr[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) xr[q] = Pc[ofs + (q - 1) * GS_HP + 2];                    // rows -1..7
#pragma unroll
        for (int r = 0; r < 8; r += 2) {
            const double up0 = r == 0 ? t0 : x[0][r - 1], up1 = r == 0 ? t1 : x[1][r - 1];
            const double v = fma(nxy, (up0 + x[0][r + 1]) + (xr[r] + xr[r + 2]), fma(nfy, up1 + x[1][r + 1], fma(nfx, x[0][r] + xr[r + 1], B[1][r])));
            if (!MASKED) x[1][r] = v; else if ((act >> (8 + r)) & 1u) x[1][r] = ((dmask >> (8 + r)) & 1u) ? 0.0 : v;
            Pc[GS_HALF + ofs + r * GS_HP + 1] = x[1][r];
        }
    }
    np_group_barrier(grp);   // only this sweep's warps exchange values inside a plane
    // ---- colour 2: a = 0, r odd ----
    if (on) {
        double xl[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) xl[q] = Pc[GS_HALF + ofs + q * GS_HP];                    // rows 0..8
        d0 = Pc[ofs + 8 * GS_HP + 1]; d1 = Pc[GS_HALF + ofs + 8 * GS_HP + 1];
#pragma unroll
        for (int r = 1; r < 8; r += 2) {
            const double dn0 = r == 7 ? d0 : x[0][r + 1], dn1 = r == 7 ? d1 : x[1][r + 1];
            const double v = fma(nxy, (xl[r - 1] + xl[r + 1]) + (x[1][r - 1] + dn1), fma(nfy, x[0][r - 1] + dn0, fma(nfx, xl[r] + x[1][r], B[0][r])));
            if (!MASKED) x[0][r] = v; else if ((act >> r) & 1u) x[0][r] = ((dmask >> r) & 1u) ? 0.0 : v;
            Pc[ofs + r * GS_HP + 1] = x[0][r];
        }
    }
    __syncwarp();            // the next colour's new in-plane neighbours sit in the same row, i.e. the same warp
    // ---- colour 3: a = 1, r odd ----
    if (on) {
        double xr[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) xr[q] = Pc[ofs + q * GS_HP + 2];                          // rows 0..8
#pragma unroll
        for (int r = 1; r < 8; r += 2) {
            const double dn0 = r == 7 ? d0 : x[0][r + 1], dn1 = r == 7 ? d1 : x[1][r + 1];
            const double v = fma(nxy, (x[0][r - 1] + dn0) + (xr[r - 1] + xr[r + 1]), fma(nfy, x[1][r - 1] + dn1, fma(nfx, x[0][r] + xr[r], B[1][r])));
            if (!MASKED) x[1][r] = v; else if ((act >> (8 + r)) & 1u) x[1][r] = ((dmask >> (8 + r)) & 1u) ? 0.0 : v;
            Pc[GS_HALF + ofs + r * GS_HP + 1] = x[1][r];
        }
    }
}

// 1/d for a normal, non-zero d to within an ulp or two: hardware seed (20 bits) + two Newton steps.  A full IEEE
// division costs several times as much in fp64 and sits on the critical path of every Gauss-Seidel update.
__device__ __forceinline__ double np_rcp(double d)
{
#ifdef NP_EMU
    return 1.0 / d;
#else
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    double e = fma(-d, r, 1.0);
    r = fma(r, e, r);
    e = fma(-d, r, 1.0);
    return fma(r, e, r);
#endif
}
// Variable sigma: the same four colour steps; V[c][q] = sigma_below + sigma_above of the cell column (c, q) around the
// thread's block (c = 0..2 <-> cells gi-1..gi+1, q = 0..8 <-> cells gj-1..gj+7), so node (a, r) sees the four columns
// V[a][r], V[a+1][r], V[a][r+1], V[a+1][r+1].  B holds rhs minus the out-of-plane part; x_new = (B - in-plane) / diag.
__device__ __forceinline__ double gs_upd_var(const NpCoef& c, double Bv, double v00, double v10, double v01, double v11,
                                             double xW, double xE, double xS, double xN, double xSW, double xSE, double xNW, double xNE)
{
    const double inpl = c.f4xm2ym2z * (xW * (v00 + v01) + xE * (v10 + v11)) + c.fm2x4ym2z * (xS * (v00 + v10) + xN * (v01 + v11))
                      + c.f2x2ymz * ((xSW * v00 + xSE * v10) + (xNW * v01 + xNE * v11));
    const double diag = -4.0 * c.fxyz * ((v00 + v10) + (v01 + v11));
    return (Bv - inpl) * np_rcp(diag);
}
__device__ __forceinline__ void gs_colours_var(bool on, double (&x)[2][8], const double (&B)[2][8], const double (&V)[3][9], double* Pc,
                                               const NpCoef& c, int ofs, unsigned act, unsigned dmask, int grp)
{
    double t0 = 0.0, t1 = 0.0, d0 = 0.0, d1 = 0.0;
    if (on) {   // colour 0: a = 0, r even
        double xl[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) xl[q] = Pc[GS_HALF + ofs + (q - 1) * GS_HP];              // rows -1..7
        t0 = Pc[ofs - GS_HP + 1]; t1 = Pc[GS_HALF + ofs - GS_HP + 1];
#pragma unroll
        for (int r = 0; r < 8; r += 2) {
            const double up0 = r == 0 ? t0 : x[0][r - 1], up1 = r == 0 ? t1 : x[1][r - 1];
            const double v = gs_upd_var(c, B[0][r], V[0][r], V[1][r], V[0][r + 1], V[1][r + 1],
                                        xl[r + 1], x[1][r], up0, x[0][r + 1], xl[r], up1, xl[r + 2], x[1][r + 1]);
            if ((act >> r) & 1u) x[0][r] = ((dmask >> r) & 1u) ? 0.0 : v;
            Pc[ofs + r * GS_HP + 1] = x[0][r];
        }
    }
    __syncwarp();            // the next colour's new in-plane neighbours sit in the same row, i.e. the same warp
    if (on) {   // colour 1: a = 1, r even
        double xr[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) xr[q] = Pc[ofs + (q - 1) * GS_HP + 2];                    // rows -1..7
#pragma unroll
        for (int r = 0; r < 8; r += 2) {
            const double up0 = r == 0 ? t0 : x[0][r - 1], up1 = r == 0 ? t1 : x[1][r - 1];
            const double v = gs_upd_var(c, B[1][r], V[1][r], V[2][r], V[1][r + 1], V[2][r + 1],
                                        x[0][r], xr[r + 1], up1, x[1][r + 1], up0, xr[r], x[0][r + 1], xr[r + 2]);
            if ((act >> (8 + r)) & 1u) x[1][r] = ((dmask >> (8 + r)) & 1u) ? 0.0 : v;
            Pc[GS_HALF + ofs + r * GS_HP + 1] = x[1][r];
        }
    }
    np_group_barrier(grp);
    if (on) {   // colour 2: a = 0, r odd
        double xl[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) xl[q] = Pc[GS_HALF + ofs + q * GS_HP];                    // rows 0..8
        d0 = Pc[ofs + 8 * GS_HP + 1]; d1 = Pc[GS_HALF + ofs + 8 * GS_HP + 1];
#pragma unroll
        for (int r = 1; r < 8; r += 2) {
            const double dn0 = r == 7 ? d0 : x[0][r + 1], dn1 = r == 7 ? d1 : x[1][r + 1];
            const double v = gs_upd_var(c, B[0][r], V[0][r], V[1][r], V[0][r + 1], V[1][r + 1],
                                        xl[r], x[1][r], x[0][r - 1], dn0, xl[r - 1], x[1][r - 1], xl[r + 1], dn1);
            if ((act >> r) & 1u) x[0][r] = ((dmask >> r) & 1u) ? 0.0 : v;
            Pc[ofs + r * GS_HP + 1] = x[0][r];
        }
    }
    __syncwarp();            // the next colour's new in-plane neighbours sit in the same row, i.e. the same warp
    if (on) {   // colour 3: a = 1, r odd
        double xr[9];
#pragma unroll
        for (int q = 0; q < 9; ++q) xr[q] = Pc[ofs + q * GS_HP + 2];                          // rows 0..8
#pragma unroll
        for (int r = 1; r < 8; r += 2) {
            const double dn0 = r == 7 ? d0 : x[0][r + 1], dn1 = r == 7 ? d1 : x[1][r + 1];
            const double v = gs_upd_var(c, B[1][r], V[1][r], V[2][r], V[1][r + 1], V[2][r + 1],
                                        x[0][r], xr[r], x[1][r - 1], dn1, x[0][r - 1], xr[r - 1], dn0, xr[r + 1]);
            if ((act >> (8 + r)) & 1u) x[1][r] = ((dmask >> (8 + r)) & 1u) ? 0.0 : v;
            Pc[GS_HALF + ofs + r * GS_HP + 1] = x[1][r];
        }
    }
}

// out-of-plane part of one neighbouring plane P (x[k-1] or x[k+1]) through the cell layer S between it and plane k
__device__ __forceinline__ void gs_zpart_var(const NpCoef& c, double (&B)[2][8], const double (&S)[3][9], const double* P, int ofs)
{
    double xa[4], xb[4], xc[4];     // node rows r-1, r, r+1; columns: left halo, own even, own odd, right halo
    {
        const int ro = ofs - GS_HP;
        xa[0] = P[GS_HALF + ro]; xa[1] = P[ro + 1]; xa[2] = P[GS_HALF + ro + 1]; xa[3] = P[ro + 2];
        xb[0] = P[GS_HALF + ofs]; xb[1] = P[ofs + 1]; xb[2] = P[GS_HALF + ofs + 1]; xb[3] = P[ofs + 2];
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int ro = ofs + (r + 1) * GS_HP;
        xc[0] = P[GS_HALF + ro]; xc[1] = P[ro + 1]; xc[2] = P[GS_HALF + ro + 1]; xc[3] = P[ro + 2];
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const double l00 = S[a][r], l10 = S[a + 1][r], l01 = S[a][r + 1], l11 = S[a + 1][r + 1];
            B[a][r] -= c.fm2xm2y4z * (xb[a + 1] * ((l00 + l10) + (l01 + l11)))
                     + c.f2xmy2z * (xb[a] * (l00 + l01) + xb[a + 2] * (l10 + l11))
                     + c.fmx2y2z * (xa[a + 1] * (l00 + l10) + xc[a + 1] * (l01 + l11))
                     + c.fxyz * ((xa[a] * l00 + xa[a + 2] * l10) + (xc[a] * l01 + xc[a + 2] * l11));
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) { xa[q] = xb[q]; xb[q] = xc[q]; }
    }
}

// Body of k_gs2.  FULL: the whole 64 x 32 tile lies inside the domain and no face is Dirichlet, so every node is
// relaxed and the masks fold away at compile time (only the 63 x 31 write window remains).
template <bool CS, bool FULL>
__device__ __forceinline__ void gs2_run(const LevelDev& L, const GsArgs& A, double* ring)
{
    const GsCoef& C = A.c;
    double* __restrict__ xout = A.xout; const double* __restrict__ xin = A.xin;
    const double* __restrict__ rhs = A.rhs; const double* __restrict__ sig = A.sig;
    const int zc = A.zc;
    const double roff = A.boff ? A.boff[0] : 0.0;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int grp = warp >> 2;                           // 0: first sweep (plane t), 1: second sweep (plane t-2)
    const int wy = warp & 3;                             // tile rows 8*wy .. 8*wy+7
    const int i0 = blockIdx.x * GS_WX, j0 = blockIdx.y * GS_WY, kb = (A.zmap0 + (int)blockIdx.z * A.zmapS) * zc;
    const int nk  = min(kb + zc, L.n[2]) - kb + 1;       // relaxed planes p = 0..nk-1 (global k = kb+p)
    const int nkw = min(kb + zc - 1, L.n[2]) - kb + 1;   // written planes
    const int gi = i0 + 2 * lane, gj = j0 + 8 * wy;      // first node of this thread's 2 x 8 block
    const int row0 = 8 * wy + 1;                         // smem row of block row 0
    const int ofs = row0 * GS_HP + lane;

    // per-thread node masks, bit a*8+r: relaxed (valid node), Dirichlet in x/y, written by this tile
    unsigned act = 0, dir = 0, wrt = 0;
    if (FULL) {
        act = 0xffffu;
        wrt = 0xffffu & ~((lane == 31 ? 0xff00u : 0u) | (wy == 3 ? 0x8080u : 0u));
    } else {
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int i = gi + a, j = gj + r;
                const bool v = (i <= L.n[0] && j <= L.n[1]);
                const bool dd = v && ((L.face[0] == NP_FACE_DIRICHLET && i == 0) || (L.face[3] == NP_FACE_DIRICHLET && i == L.n[0])
                                   || (L.face[1] == NP_FACE_DIRICHLET && j == 0) || (L.face[4] == NP_FACE_DIRICHLET && j == L.n[1]));
                const bool w = v && (2 * lane + a < GS_WX) && (8 * wy + r < GS_WY);
                act |= (unsigned)v << (a * 8 + r); dir |= (unsigned)dd << (a * 8 + r); wrt |= (unsigned)w << (a * 8 + r);
            }
    }
    constexpr bool full = FULL;

    GsLoader ld; ld.init(L, tid, i0, j0, A.zero_in != 0);
    auto slot = [&](int p) -> double* { return ring + ((p + 1) % GS_RING) * GS_PLANE; };     // p >= -1
    auto load_plane = [&](int p) {                                                           // planes -1 .. nk
        if (FULL && !A.zero_in) ld.template load<true>(slot(p), xin + L.nsz * (long long)(kb + p + 1), L.nsy);
        else                    ld.template load<false>(slot(p), xin + L.nsz * (long long)(kb + p + 1), L.nsy);
    };

    load_plane(-1); load_plane(0); load_plane(1);
    np_cp_async_commit(); np_cp_async_wait_all();
    __syncthreads();

    double x[2][8], B[2][8];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int r = 0; r < 8; ++r) { x[a][r] = 0.0; B[a][r] = 0.0; }

    auto load_rhs = [&](int pn) {                        // B <- raw rhs of plane pn (roff where the node is not relaxed)
        const double* rp = rhs + np_noff(L, gi, gj, kb + pn);
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            B[0][r] = (full || ((act >> r) & 1u)) ? rp[L.nsy * r] : roff;
            B[1][r] = (full || ((act >> (8 + r)) & 1u)) ? rp[L.nsy * r + 1] : roff;
        }
    };
    if (CS && grp == 0) load_rhs(0);

    // variable sigma: the 3 x 9 cells (gi-1+c, gj-1+q) of a cell layer around this thread's block.  Addresses cost more than
    // the loads if they are formed per element (ncu: a third of the kernel's instructions were IMAD / LEA): the thread keeps
    // one pointer to its first cell in layer -1; a full tile then uses row pointers and immediate column offsets, any other
    // tile 9 + 3 precomputed element offsets with the indices clamped into the array (clamped cells only feed nodes
    // outside the domain, which are never updated).
    const double* sig_t = nullptr;
    int sro[9], sco[3];
    if (!CS) {
        if (FULL) sig_t = sig + (long long)gi + L.csy * (long long)gj;          // cell (gi-1, gj-1) of layer -1
        else {
            sig_t = sig;
#pragma unroll
            for (int q = 0; q < 9; ++q) sro[q] = (int)(L.csy * (long long)(min(gj - 1 + q, L.n[1]) + 1));
#pragma unroll
            for (int c = 0; c < 3; ++c) sco[c] = min(gi - 1 + c, L.n[0]) + 1;
        }
    }
    auto load_sig = [&](double (&S)[3][9], int kc) {                            // cell layer kc (k index), -1 <= kc <= n2
        const double* sb = sig_t + L.csz * (long long)(kc + 1);
        if (FULL) {
#pragma unroll
            for (int q = 0; q < 9; ++q) {
#pragma unroll
                for (int c = 0; c < 3; ++c) S[c][q] = sb[c];
                sb += L.csy;
            }
        } else {
#pragma unroll
            for (int q = 0; q < 9; ++q)
#pragma unroll
                for (int c = 0; c < 3; ++c) S[c][q] = sb[sro[q] + sco[c]];
        }
    };

    for (int t = 0; t <= nk + 1; ++t) {
        if (t + 2 <= nk) load_plane(t + 2);
        np_cp_async_commit();
        const int p = t - 2 * grp;
        const bool on = (p >= 0 && p <= nk - 1) && (FULL || act != 0u);     // a thread whose block lies outside the level only keeps the barriers
        const int k = kb + p;
        double* Pc = on ? slot(p) : ring;
        const bool zdir = !FULL && on && ((L.face[2] == NP_FACE_DIRICHLET && k == 0) || (L.face[5] == NP_FACE_DIRICHLET && k == L.n[2]));
        const unsigned dmask = zdir ? act : dir;
        const long long gbase = on ? np_noff(L, gi, gj, k) : 0;
        if (A.addto && on && grp == 1 && p < nkw) {     // the add-to values are needed at the end of the step: pull them into L1 now
#pragma unroll
            for (int r = 0; r < 8; ++r) np_prefetch_l1(A.addto + gbase + L.nsy * r);
        }

        if (CS) {
            if (on) {
                // ---- B = rhs/diag - (out-of-plane part): 9-point stencil of S = x[k-1] + x[k+1] ----
                // B arrives holding the raw right-hand side of this plane (loaded at the end of the previous step)
                // x still holds this thread's block of plane p-1 as the previous step left it (first plane: fetch it)
                const double* Pm = slot(p - 1); const double* Pp = slot(p + 1);
                if (p == 0) {
#pragma unroll
                    for (int r = 0; r < 8; ++r) { x[0][r] = Pm[ofs + r * GS_HP + 1]; x[1][r] = Pm[GS_HALF + ofs + r * GS_HP + 1]; }
                }
#pragma unroll
                for (int rr = -1; rr <= 8; ++rr) {
                    const int ro = ofs + rr * GS_HP;
                    const bool own = (rr >= 0 && rr <= 7);
                    const double sL = Pm[GS_HALF + ro] + Pp[GS_HALF + ro];
                    const double s0 = (own ? x[0][rr & 7] : Pm[ro + 1]) + Pp[ro + 1];
                    const double s1 = (own ? x[1][rr & 7] : Pm[GS_HALF + ro + 1]) + Pp[GS_HALF + ro + 1];
                    const double sR = Pm[ro + 2] + Pp[ro + 2];
                    const double h0 = sL + s1, h1 = s0 + sR;
                    const double q0 = C.a_eyz * s0 + C.a_c * h0, q1 = C.a_eyz * s1 + C.a_c * h1;
                    if (rr >= 0 && rr <= 7) {
                        // row rr's own contribution; B[.][rr] already holds -q of row rr-1 (set below)
                        B[0][rr] -= C.a_fz * s0 + C.a_exz * h0;
                        B[1][rr] -= C.a_fz * s1 + C.a_exz * h1;
                    }
                    if (rr - 1 >= 0 && rr - 1 <= 7) { B[0][rr - 1] -= q0; B[1][rr - 1] -= q1; }
                    if (rr + 1 >= 0 && rr + 1 <= 7) {      // first touch of row rr+1: rhs/diag comes in here
                        B[0][rr + 1] = fma(B[0][rr + 1] - roff, C.dinv, -q0); B[1][rr + 1] = fma(B[1][rr + 1] - roff, C.dinv, -q1);
                    }
                }
                // ---- own values of plane p ----
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    x[0][r] = Pc[ofs + r * GS_HP + 1];
                    x[1][r] = Pc[GS_HALF + ofs + r * GS_HP + 1];
                }
            }
            gs_colours_cs<!FULL>(on, x, B, Pc, C, ofs, act, dmask, grp);
            // B is free now: fetch the raw right-hand side of this group's next plane so the loads fly across the barrier
            if (p + 1 >= 0 && p + 1 <= nk - 1) load_rhs(p + 1);
        } else {
            // ---- variable sigma: B = rhs - (out-of-plane part through the two cell layers); V = column sums of sigma ----
            double V[3][9];
            if (on && p + 1 <= nk - 1) {
                // this thread's next plane: pull its right-hand side and the sigma layer above it into L1 while this plane is
                // being relaxed (one CTA per SM leaves a large L1), so the loads at the top of the next step are hits
#pragma unroll
                for (int r = 0; r < 8; ++r) np_prefetch_l1(rhs + gbase + L.nsz + L.nsy * r);
                const double* sn = sig_t + L.csz * (long long)(k + 2);
                if (FULL) {
#pragma unroll
                    for (int q = 0; q < 9; ++q) { np_prefetch_l1(sn + 1); sn += L.csy; }
                } else {
#pragma unroll
                    for (int q = 0; q < 9; ++q) np_prefetch_l1(sn + sro[q] + sco[1]);
                }
            }
            if (on) {
                const double* rp = rhs + gbase;
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    B[0][r] = ((act >> r) & 1u) ? rp[L.nsy * r] - roff : 0.0;
                    B[1][r] = ((act >> (8 + r)) & 1u) ? rp[L.nsy * r + 1] - roff : 0.0;
                }
                // cells (gi-1+c, gj-1+q) of the layers below (k-1) and above (k) plane k; indices clamped into the array
                // (clamped cells only feed nodes outside the domain, which are never updated)
                double S[3][9];
                load_sig(S, k - 1);                                             // cells below plane k
                gs_zpart_var(L.c, B, S, slot(p - 1), ofs);
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int q = 0; q < 9; ++q) V[c][q] = S[c][q];
                load_sig(S, k);                                                 // cells above
                gs_zpart_var(L.c, B, S, slot(p + 1), ofs);
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int q = 0; q < 9; ++q) V[c][q] += S[c][q];
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    x[0][r] = Pc[ofs + r * GS_HP + 1];
                    x[1][r] = Pc[GS_HALF + ofs + r * GS_HP + 1];
                }
            } else {
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int q = 0; q < 9; ++q) V[c][q] = 1.0;
            }
            gs_colours_var(on, x, B, V, Pc, L.c, ofs, act, dmask, grp);
        }
        np_cp_async_wait_all();
        __syncthreads();
        if (on && grp == 1 && p < nkw) {
            double* op = xout + gbase;
            if (A.addto) {
                const double* ap = A.addto + gbase;
                double av[2][8];
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    av[0][r] = ((wrt >> r) & 1u) ? ap[L.nsy * r] : 0.0;
                    av[1][r] = ((wrt >> (8 + r)) & 1u) ? ap[L.nsy * r + 1] : 0.0;
                }
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    if ((wrt >> r) & 1u) op[L.nsy * r] = av[0][r] + x[0][r];
                    if ((wrt >> (8 + r)) & 1u) op[L.nsy * r + 1] = av[1][r] + x[1][r];
                }
            } else {
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    if ((wrt >> r) & 1u) op[L.nsy * r] = x[0][r];
                    if ((wrt >> (8 + r)) & 1u) op[L.nsy * r + 1] = x[1][r];
                }
            }
        }
    }
}

template <bool CS>
__global__ void __launch_bounds__(256, CS ? 2 : 1) k_gs2(LevelDev L, GsArgs A)
{
    NP_DYN_SMEM(double, ring);
    const bool any_dir = L.face[0] == NP_FACE_DIRICHLET || L.face[1] == NP_FACE_DIRICHLET || L.face[2] == NP_FACE_DIRICHLET
                      || L.face[3] == NP_FACE_DIRICHLET || L.face[4] == NP_FACE_DIRICHLET || L.face[5] == NP_FACE_DIRICHLET;
    const bool full = ((int)blockIdx.x * GS_WX + GS_TX - 1 <= L.n[0]) && ((int)blockIdx.y * GS_WY + GS_TY - 1 <= L.n[1]) && !any_dir;
    if (full) gs2_run<CS, true>(L, A, ring);
    else      gs2_run<CS, false>(L, A, ring);
}

#include "np_gs3.cuh"
#include "np_gs_small.cuh"

// ------------------------------------------------------------------------------------------------
// K2, constant sigma: y = A x  (MODE 0)  or  y = (b - *boff) - A x  (MODE 1), marching in z.
// One CTA owns a 64 x 32 column chunk; each plane of x is staged ONCE in shared memory (cp.async, two
// planes ahead, split even/odd layout as in k_gs2); each thread owns a 2 x 4 block of nodes and keeps,
// per node, two running values in registers (the pending sum V[k-1] + U[k] and the last V):
//     y[k] = (V[k-1] + U[k]) + V[k+1],   U = same-plane weights (incl. diagonal), V = adjacent-plane weights.
// REDUCE: per-CTA partials of max|y| (all valid nodes) and sum(w*y) (owned nodes, dot-mask weights) —
// the residual norm and the solvability offset of the next V-cycle come out of the same pass.
// ------------------------------------------------------------------------------------------------
#define ST_RING 3
#define ST_SMEM_BYTES (ST_RING * GS_PLANE * 8)
struct StCoef { double d, w_fx, w_fy, w_exy, w_fz, w_exz, w_eyz, w_c; };

// CS = false (variable sigma): the same march; the plane arriving at step q contributes, through the cell layers below
// and above it, D[q] to y[q-1], M[q] to y[q] and UP[q] to y[q+1]:  y[k] = UP[k-1] + M[k] + D[k+1]  (for constant sigma
// D = UP = V).  Sigma comes through L1 (each cell is used by up to 8 nodes of neighbouring threads).
// x/y factor of the dot-mask weight of node (i, j): 0 when another rank / the periodic image owns it, 1/2 per Neumann face
__device__ __forceinline__ double np_wgt_xy(const LevelDev& L, int i, int j)
{
    if (i > L.owned_hi[0] || j > L.owned_hi[1]) return 0.0;
    double w = 1.0;
    if (L.face[0] == NP_FACE_NEUMANN && i == 0) w *= 0.5;
    if (L.face[3] == NP_FACE_NEUMANN && i == L.n[0]) w *= 0.5;
    if (L.face[1] == NP_FACE_NEUMANN && j == 0) w *= 0.5;
    if (L.face[4] == NP_FACE_NEUMANN && j == L.n[1]) w *= 0.5;
    return w;
}
// FULL: the tile lies inside the domain, no face is Dirichlet and every dot-mask weight is 1: the masks fold away.
// RR (MLMG::computeResOfCorrection + MLNodeLaplacian::restriction in one pass, "the residual is never stored"): the tile
// strides are RR_SX x RR_SY = 62 x 30 so that neighbouring tiles overlap by two fine nodes; a tile then holds the whole 3 x 3
// fine neighbourhood of its 31 x 15 coarse nodes (fine 2I = i0+2 .. i0+62).  The z part of the full weighting (1/4, 1/2, 1/4)
// is accumulated in registers while the planes go by; at every odd fine plane the accumulated tile goes through shared
// memory for the x / y part and one coarse plane is stored.  Coarse nodes ON the faces of the box need ghost residuals
// (periodic images, Neumann mirrors, the neighbour rank's planes): the kernel stores the fine residual on the boundary shell
// only (nodes within one of a face) and k_restrict_faces finishes those coarse nodes after the usual ghost fill.
#define RR_SX 62
#define RR_SY 30
#define RR_TP 66                                        // pitch of the accumulated tile (even: 16-byte row pairs)
struct RrArgs { LevelDev Lc; double* crse; int kcz; };   // coarse level descriptor / array, coarse planes per z-chunk
template <int MODE, bool REDUCE, bool CS, bool FULL, bool RR = false>
__device__ __forceinline__ void st_run(const LevelDev& L, const StCoef& C, double* __restrict__ y, const double* __restrict__ x,
                                       const double* __restrict__ b, const double* __restrict__ boff,
                                       double* __restrict__ part_max, double* __restrict__ part_sum, int zc,
                                       const double* __restrict__ sig, double* ring, const RrArgs* rr = nullptr)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int i0 = blockIdx.x * (RR ? RR_SX : GS_TX), j0 = blockIdx.y * (RR ? RR_SY : GS_TY);
    // RR: this CTA produces the coarse planes Ka .. Kb (1 <= K <= nc2 - 1) from the fine planes 2 Ka - 1 .. 2 Kb + 1; the first
    // and the last chunk also run the fine planes 0 and n2 (their residual is part of the stored shell)
    const int Ka = RR ? 1 + (int)blockIdx.z * rr->kcz : 0, Kb = RR ? min(Ka + rr->kcz - 1, rr->Lc.n[2] - 1) : 0;
    const int kb = RR ? (Ka == 1 ? 0 : 2 * Ka - 1) : (int)blockIdx.z * zc;
    const int nkl = RR ? (Kb == rr->Lc.n[2] - 1 ? L.n[2] : 2 * Kb + 1) - kb + 1
                       : min(kb + zc - 1, L.n[2]) - kb + 1;       // output planes kb .. kb+nkl-1
    const int gi = i0 + 2 * lane, gj = j0 + 4 * warp;
    const int row0 = 4 * warp + 1;
    const double off = (MODE == 1 && boff) ? boff[0] : 0.0;

    unsigned act = FULL ? 0xffu : 0u, dir = 0;
    bool wone = true;                                    // REDUCE: all eight dot-mask weights of this thread are 1
    if (!FULL) {
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int i = gi + a, j = gj + r;
                const bool v = (i <= L.n[0] && j <= L.n[1]);
                const bool dd = v && ((L.face[0] == NP_FACE_DIRICHLET && i == 0) || (L.face[3] == NP_FACE_DIRICHLET && i == L.n[0])
                                   || (L.face[1] == NP_FACE_DIRICHLET && j == 0) || (L.face[4] == NP_FACE_DIRICHLET && j == L.n[1]));
                act |= (unsigned)v << (a * 4 + r); dir |= (unsigned)dd << (a * 4 + r);
                if (REDUCE) wone = wone && (np_wgt_xy(L, i, j) == 1.0);
            }
    }

    GsLoader ld; ld.init(L, tid, i0, j0, false);
    auto load_plane = [&](int k) { ld.template load<FULL>(ring + ((k + 1) % ST_RING) * GS_PLANE, x + L.nsz * (long long)(k + 1), L.nsy); };   // k in [-1, n2+1]

    // per node: A1 = UP[q-2] + M[q-1] (y[q-1] still waiting for D[q]) and A2 = UP[q-1]
    double A1[2][4], A2[2][4];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int r = 0; r < 4; ++r) { A1[a][r] = 0.0; A2[a][r] = 0.0; }
    double tmax = 0.0, tsum = 0.0;
    double acc[2][4];                                    // RR: z-weighted residual of the coarse plane in the making
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int r = 0; r < 4; ++r) acc[a][r] = 0.0;
    double* T = ring + ST_RING * GS_PLANE;               // RR: the accumulated tile, 32 rows x RR_TP

    load_plane(kb - 1); np_cp_async_commit();
    load_plane(kb);     np_cp_async_commit();
    for (int s = 0; s <= nkl + 1; ++s) {
        const int q = kb - 1 + s;                        // plane arriving this step
        // planes q (needed now) and q+1 were issued; keep at most the newest group in flight
#ifndef NP_EMU
        if (s <= nkl) asm volatile("cp.async.wait_group 1;\n" ::: "memory"); else asm volatile("cp.async.wait_group 0;\n" ::: "memory");
#endif
        __syncthreads();                                 // plane q visible; everyone is done reading plane q-1
        if (s + 2 <= nkl + 1) load_plane(q + 2);         // goes into the slot of plane q-1
        np_cp_async_commit();
        const double* P = ring + ((q + 1) % ST_RING) * GS_PLANE;
        // right-hand side of the plane written this step: issued before the shared-memory work, used after it
        double bv[2][4];
        const long long gbase = np_noff(L, gi, gj, q - 1);
        if (MODE == 1 && s >= 2) {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                bv[0][r] = (FULL || ((act >> r) & 1u)) ? b[gbase + L.nsy * r] : 0.0;
                bv[1][r] = (FULL || ((act >> (4 + r)) & 1u)) ? b[gbase + L.nsy * r + 1] : 0.0;
            }
        }
        double Un[2][4], Vn[2][4], Dn[2][4];             // M[q], UP[q], D[q] of the node block
        if (CS) {
            double c0[6], c1[6], h0[6], h1[6];
#pragma unroll
            for (int rr = 0; rr < 6; ++rr) {
                const int ro = (row0 - 1 + rr) * GS_HP + lane;
                const double xl = P[GS_HALF + ro], x0 = P[ro + 1], x1 = P[GS_HALF + ro + 1], xr = P[ro + 2];
                c0[rr] = x0; c1[rr] = x1; h0[rr] = xl + x1; h1[rr] = x0 + xr;
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const double cc0 = c0[r] + c0[r + 2], hh0 = h0[r] + h0[r + 2], cc1 = c1[r] + c1[r + 2], hh1 = h1[r] + h1[r + 2];
                Un[0][r] = C.d * c0[r + 1] + C.w_fx * h0[r + 1] + C.w_fy * cc0 + C.w_exy * hh0;
                Un[1][r] = C.d * c1[r + 1] + C.w_fx * h1[r + 1] + C.w_fy * cc1 + C.w_exy * hh1;
                Vn[0][r] = C.w_fz * c0[r + 1] + C.w_exz * h0[r + 1] + C.w_eyz * cc0 + C.w_c * hh0;
                Vn[1][r] = C.w_fz * c1[r + 1] + C.w_exz * h1[r + 1] + C.w_eyz * cc1 + C.w_c * hh1;
                Dn[0][r] = Vn[0][r]; Dn[1][r] = Vn[1][r];
            }
        } else {
            double xw[6][4];                             // node rows -1..4; columns: left halo, own even, own odd, right halo
#pragma unroll
            for (int rr = 0; rr < 6; ++rr) {
                const int ro = (row0 - 1 + rr) * GS_HP + lane;
                xw[rr][0] = P[GS_HALF + ro]; xw[rr][1] = P[ro + 1]; xw[rr][2] = P[GS_HALF + ro + 1]; xw[rr][3] = P[ro + 2];
            }
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int r = 0; r < 4; ++r) { Un[a][r] = 0.0; }
#pragma unroll
            for (int lay = 0; lay < 2; ++lay) {          // 0: cells below plane q (k index q-1), 1: cells above (k index q)
                const int kc = min(max(q - 1 + lay, -1), L.n[2]);       // clamped layers only feed planes that are not written
                const double* sb = sig + L.csz * (long long)(kc + 1);
                double S[3][5];
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int qq = 0; qq < 5; ++qq) {
                        const int ci = min(gi - 1 + c, L.n[0]), cj = min(gj - 1 + qq, L.n[1]);
                        S[c][qq] = sb[(ci + 1) + L.csy * (long long)(cj + 1)];
                    }
#pragma unroll
                for (int a = 0; a < 2; ++a)
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const double l00 = S[a][r], l10 = S[a + 1][r], l01 = S[a][r + 1], l11 = S[a + 1][r + 1];
                        const double p0 = xw[r + 1][a + 1] * ((l00 + l10) + (l01 + l11));
                        const double px = xw[r + 1][a] * (l00 + l01) + xw[r + 1][a + 2] * (l10 + l11);
                        const double py = xw[r][a + 1] * (l00 + l10) + xw[r + 2][a + 1] * (l01 + l11);
                        const double pxy = (xw[r][a] * l00 + xw[r][a + 2] * l10) + (xw[r + 2][a] * l01 + xw[r + 2][a + 2] * l11);
                        const double cross = L.c.fm2xm2y4z * p0 + L.c.f2xmy2z * px + L.c.fmx2y2z * py + L.c.fxyz * pxy;
                        const double same = -4.0 * L.c.fxyz * p0 + L.c.f4xm2ym2z * px + L.c.fm2x4ym2z * py + L.c.f2x2ymz * pxy;
                        Un[a][r] += same;
                        if (lay == 0) Dn[a][r] = cross; else Vn[a][r] = cross;
                    }
            }
        }
        if (s >= 2) {                                    // output plane k = q-1
            const int k = q - 1;
            const bool zdir = !FULL && ((L.face[2] == NP_FACE_DIRICHLET && k == 0) || (L.face[5] == NP_FACE_DIRICHLET && k == L.n[2]));
            double wz = 1.0;
            if (REDUCE) {
                if (k > L.owned_hi[2]) wz = 0.0;
                if ((L.face[2] == NP_FACE_NEUMANN && k == 0) || (L.face[5] == NP_FACE_NEUMANN && k == L.n[2])) wz *= 0.5;
            }
            double vv[2][4];                             // RR: the residual of this plane (0 outside the level / on Dirichlet faces)
            const bool kshell = RR && (k <= 1 || k >= L.n[2] - 1);
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    vv[a][r] = 0.0;
                    if (FULL || ((act >> (a * 4 + r)) & 1u)) {
                        const long long o = gbase + a + L.nsy * r;
                        double v = 0.0;
                        if (FULL || !(zdir || ((dir >> (a * 4 + r)) & 1u))) {
                            const double ax = A1[a][r] + Dn[a][r];
                            v = (MODE == 1) ? (bv[a][r] - off) - ax : ax;
                        }
                        if (!RR) y[o] = v;
                        else {
                            vv[a][r] = v;
                            const int i = gi + a, j = gj + r;
                            if (kshell || i <= 1 || i >= L.n[0] - 1 || j <= 1 || j >= L.n[1] - 1) y[o] = v;      // boundary shell only
                        }
                        if (REDUCE) {
                            tmax = np_absmax(tmax, v);
                            tsum += ((FULL || wone) ? wz : np_wgt_xy(L, gi + a, gj + r) * wz) * v;      // boundary threads redo the weight
                        }
                    }
                }
            if (RR) {
                const bool odd = (k & 1) != 0;
                const double wk = odd ? 0.25 : 0.5;
#pragma unroll
                for (int a = 0; a < 2; ++a)
#pragma unroll
                    for (int r = 0; r < 4; ++r) acc[a][r] += wk * vv[a][r];
                if (odd) {
                    const int K = (k - 1) >> 1;          // fine planes 2K-1, 2K, 2K+1 are in: coarse plane K
                    if (K >= Ka && K <= Kb) {            // (uniform over the CTA)
#pragma unroll
                        for (int r = 0; r < 4; ++r)
                            *reinterpret_cast<double2*>(T + (4 * warp + r) * RR_TP + 2 * lane) = make_double2(acc[0][r], acc[1][r]);
                        __syncthreads();                 // the next write of T is two steps (two loop barriers) away
                        const LevelDev& Lc = rr->Lc;
                        for (int t = tid; t < 31 * 15; t += 256) {
                            const int m = t % 31, rw = t / 31;
                            const int I = (i0 >> 1) + 1 + m, J = (j0 >> 1) + 1 + rw;
                            if (I <= Lc.n[0] - 1 && J <= Lc.n[1] - 1) {
                                const double* tp = T + (2 * rw + 2) * RR_TP + (2 * m + 2);
                                rr->crse[np_noff(Lc, I, J, K)] = 0.25 * tp[0] + 0.125 * ((tp[-1] + tp[1]) + (tp[-RR_TP] + tp[RR_TP]))
                                    + 0.0625 * ((tp[-RR_TP - 1] + tp[-RR_TP + 1]) + (tp[RR_TP - 1] + tp[RR_TP + 1]));
                            }
                        }
                    }
#pragma unroll
                    for (int a = 0; a < 2; ++a)
#pragma unroll
                        for (int r = 0; r < 4; ++r) acc[a][r] = 0.25 * vv[a][r];
                }
            }
        }
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int r = 0; r < 4; ++r) { A1[a][r] = A2[a][r] + Un[a][r]; A2[a][r] = Vn[a][r]; }
    }
    if (REDUCE) {
        const double m = np_block_reduce<true>(tmax);
        const double sm = np_block_reduce<false>(tsum);
        if (tid == 0) {
            const long long bidx = blockIdx.x + gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z);
            part_max[bidx] = m; part_sum[bidx] = sm;
        }
    }
}


template <int MODE, bool REDUCE, bool CS>
__global__ void __launch_bounds__(256, CS ? 2 : 1) k_stencil(LevelDev L, StCoef C, double* __restrict__ y, const double* __restrict__ x,
                                                          const double* __restrict__ b, const double* __restrict__ boff,
                                                          double* __restrict__ part_max, double* __restrict__ part_sum, int zc,
                                                          const double* __restrict__ sig)
{
    NP_DYN_SMEM(double, ring);
    const int i0 = blockIdx.x * GS_TX, j0 = blockIdx.y * GS_TY, i1 = i0 + GS_TX - 1, j1 = j0 + GS_TY - 1;
    bool full = i1 <= L.n[0] && j1 <= L.n[1];
#pragma unroll
    for (int f = 0; f < 6; ++f) full = full && L.face[f] != NP_FACE_DIRICHLET;
    if (REDUCE)
        full = full && i1 <= L.owned_hi[0] && j1 <= L.owned_hi[1]
            && !(L.face[0] == NP_FACE_NEUMANN && i0 == 0) && !(L.face[3] == NP_FACE_NEUMANN && i1 == L.n[0])
            && !(L.face[1] == NP_FACE_NEUMANN && j0 == 0) && !(L.face[4] == NP_FACE_NEUMANN && j1 == L.n[1]);
    if (full) st_run<MODE, REDUCE, CS, true>(L, C, y, x, b, boff, part_max, part_sum, zc, sig, ring);
    else      st_run<MODE, REDUCE, CS, false>(L, C, y, x, b, boff, part_max, part_sum, zc, sig, ring);
}

#define RR_SMEM_BYTES (ST_SMEM_BYTES + 32 * RR_TP * 8)
// K2 + K4 fused: crse(interior) = R(b - *boff - A x); y receives the residual on the boundary shell only (see st_run, RR)
template <bool CS>
__global__ void __launch_bounds__(256, CS ? 2 : 1) k_resid_restrict(LevelDev L, StCoef C, double* __restrict__ y, const double* __restrict__ x,
                                                                    const double* __restrict__ b, const double* __restrict__ boff,
                                                                    const double* __restrict__ sig, RrArgs rr)
{
    NP_DYN_SMEM(double, ring);
    const int i0 = blockIdx.x * RR_SX, j0 = blockIdx.y * RR_SY;
    bool full = i0 + GS_TX - 1 <= L.n[0] && j0 + GS_TY - 1 <= L.n[1];
#pragma unroll
    for (int f = 0; f < 6; ++f) full = full && L.face[f] != NP_FACE_DIRICHLET;
    if (full) st_run<1, false, CS, true, true>(L, C, y, x, b, boff, nullptr, nullptr, 0, sig, ring, &rr);
    else      st_run<1, false, CS, false, true>(L, C, y, x, b, boff, nullptr, nullptr, 0, sig, ring, &rr);
}
// the coarse nodes on the six faces of the box (I = 0, nc0, J = 0, nc1, K = 0, nc2) from the fine residual stored on the
// boundary shell, whose ghosts k_fill_boundary / the halo exchange have filled: plain mlndlap_restriction there
__global__ void __launch_bounds__(256) k_restrict_faces(LevelDev Lc, LevelDev Lf, double* __restrict__ crse, const double* __restrict__ fine)
{
    const int e0 = Lc.n[0] + 1, e1 = Lc.n[1] + 1, e2 = Lc.n[2] + 1;
    const long long g0 = 2LL * e1 * e2, g1 = 2LL * e0 * e2, g2 = 2LL * e0 * e1;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < g0 + g1 + g2; t += (long long)gridDim.x * blockDim.x) {
        int i, j, k;
        if (t < g0)           { long long q = t;           const int sd = (int)(q % 2); q /= 2; j = (int)(q % e1); k = (int)(q / e1); i = sd ? Lc.n[0] : 0; }
        else if (t < g0 + g1) { long long q = t - g0;      const int sd = (int)(q % 2); q /= 2; i = (int)(q % e0); k = (int)(q / e0); j = sd ? Lc.n[1] : 0; }
        else                  { long long q = t - g0 - g1; const int sd = (int)(q % 2); q /= 2; i = (int)(q % e0); j = (int)(q / e0); k = sd ? Lc.n[2] : 0; }
        double v = 0.0;
        if (!np_is_dirichlet(Lf, 2*i, 2*j, 2*k)) {
            GlobX f{fine + np_noff(Lf, 2*i, 2*j, 2*k), Lf.nsy, Lf.nsz};
            const double faces = (f(-1,0,0) + f(1,0,0)) + (f(0,-1,0) + f(0,1,0)) + (f(0,0,-1) + f(0,0,1));
            const double edges = ((f(-1,-1,0) + f(1,-1,0)) + (f(-1,1,0) + f(1,1,0)))
                               + ((f(-1,0,-1) + f(1,0,-1)) + (f(-1,0,1) + f(1,0,1)))
                               + ((f(0,-1,-1) + f(0,1,-1)) + (f(0,-1,1) + f(0,1,1)));
            const double corners = ((f(-1,-1,-1) + f(1,-1,-1)) + (f(-1,1,-1) + f(1,1,-1)))
                                 + ((f(-1,-1,1) + f(1,-1,1)) + (f(-1,1,1) + f(1,1,1)));
            v = 0.125 * f(0,0,0) + 0.0625 * faces + 0.03125 * edges + 0.015625 * corners;
        }
        crse[np_noff(Lc, i, j, k)] = v;
    }
}

// ------------------------------------------------------------------------------------------------
// K4: restriction, one thread per coarse valid node.  Lf: fine level (for the Dirichlet test), Lc: coarse.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_restrict(LevelDev Lc, LevelDev Lf, double* __restrict__ crse, const double* __restrict__ fine)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i > Lc.n[0] || j > Lc.n[1]) return;
    double v = 0.0;
    if (!np_is_dirichlet(Lf, 2*i, 2*j, 2*k)) {
        GlobX f{fine + np_noff(Lf, 2*i, 2*j, 2*k), Lf.nsy, Lf.nsz};
        double faces = (f(-1,0,0) + f(1,0,0)) + (f(0,-1,0) + f(0,1,0)) + (f(0,0,-1) + f(0,0,1));
        double edges = ((f(-1,-1,0) + f(1,-1,0)) + (f(-1,1,0) + f(1,1,0)))
                     + ((f(-1,0,-1) + f(1,0,-1)) + (f(-1,0,1) + f(1,0,1)))
                     + ((f(0,-1,-1) + f(0,1,-1)) + (f(0,-1,1) + f(0,1,1)));
        double corners = ((f(-1,-1,-1) + f(1,-1,-1)) + (f(-1,1,-1) + f(1,1,-1)))
                       + ((f(-1,-1,1) + f(1,-1,1)) + (f(-1,1,1) + f(1,1,1)));
        v = 0.125 * f(0,0,0) + 0.0625 * faces + 0.03125 * edges + 0.015625 * corners;
    }
    crse[np_noff(Lc, i, j, k)] = v;
}

// ------------------------------------------------------------------------------------------------
// K5: fine += P(crse).  WEIGHTED: sigma-weighted lines/faces/centres (amrex aa_interp_line_* etc.);
// otherwise plain trilinear (identical when sigma is uniform).  koff2: coarse plane index of fine
// plane 0 is koffc (non-zero only when the coarse level is a replicated copy of the whole domain).
// ------------------------------------------------------------------------------------------------
struct InterpCtx {
    const double* c; long long csy_n, csz_n;      // coarse nodal array, pointer at coarse node (0,0,0) incl. k offset
    const double* s; long long ssy, ssz;          // fine sigma, pointer at fine cell (0,0,0)
    __device__ __forceinline__ double crse(int i, int j, int k) const { return c[(i >> 1) + csy_n * (j >> 1) + csz_n * (k >> 1)]; }
    __device__ __forceinline__ double sg(int i, int j, int k) const { return s[i + ssy * j + ssz * k]; }
    // 4 fine cells on the low (side 0) / high (side 1) side of fine node (i,j,k) along d
    __device__ __forceinline__ double quad(int i, int j, int k, int d, int side) const {
        if (d == 0) { int q = i - 1 + side; return (sg(q,j-1,k-1) + sg(q,j,k-1)) + (sg(q,j-1,k) + sg(q,j,k)); }
        if (d == 1) { int q = j - 1 + side; return (sg(i-1,q,k-1) + sg(i,q,k-1)) + (sg(i-1,q,k) + sg(i,q,k)); }
        int q = k - 1 + side; return (sg(i-1,j-1,q) + sg(i,j-1,q)) + (sg(i-1,j,q) + sg(i,j,q));
    }
    __device__ __forceinline__ double line(int i, int j, int k, int d) const {
        const double w1 = quad(i,j,k,d,0), w2 = quad(i,j,k,d,1);
        const int di = d == 0, dj = d == 1, dk = d == 2;
        return (w1 * crse(i-di, j-dj, k-dk) + w2 * crse(i+di, j+dj, k+dk)) * np_rcp(w1 + w2);
    }
    __device__ __forceinline__ double face(int i, int j, int k, int da, int db) const {
        const double wa0 = quad(i,j,k,da,0), wa1 = quad(i,j,k,da,1), wb0 = quad(i,j,k,db,0), wb1 = quad(i,j,k,db,1);
        const int ai = da == 0, aj = da == 1, ak = da == 2, bi = db == 0, bj = db == 1, bk = db == 2;
        return (wa0 * line(i-ai, j-aj, k-ak, db) + wa1 * line(i+ai, j+aj, k+ak, db)
              + wb0 * line(i-bi, j-bj, k-bk, da) + wb1 * line(i+bi, j+bj, k+bk, da)) * np_rcp((wa0 + wa1) + (wb0 + wb1));
    }
    __device__ __forceinline__ double centre(int i, int j, int k) const {
        const double x0 = quad(i,j,k,0,0), x1 = quad(i,j,k,0,1), y0 = quad(i,j,k,1,0), y1 = quad(i,j,k,1,1),
                     z0 = quad(i,j,k,2,0), z1 = quad(i,j,k,2,1);
        const double num = (x0 * face(i-1,j,k,1,2) + x1 * face(i+1,j,k,1,2))
                         + (y0 * face(i,j-1,k,2,0) + y1 * face(i,j+1,k,2,0))
                         + (z0 * face(i,j,k-1,0,1) + z1 * face(i,j,k+1,0,1));
        return num * np_rcp(((x0 + x1) + (y0 + y1)) + (z0 + z1));
    }
};

// K5, sigma-weighted (mlndlap_interpadd_aa): one thread per COARSE node (I,J,K) owns the 2x2x2 fine nodes
// (2I+a, 2J+b, 2K+c) and evaluates the coarse cell's 12 edge (line) values once, its 6 face values from those and the
// centre from the faces — the nested definition of the reference, without recomputing any of it per fine node.
__device__ __forceinline__ double np_interp_face(const InterpCtx& ctx, int i, int j, int k, int da, int db,
                                                 double la0, double la1, double lb0, double lb1)
{   // la0/la1: db-lines at the node's low/high neighbour along da;  lb0/lb1: da-lines at its neighbours along db
    const double wa0 = ctx.quad(i,j,k,da,0), wa1 = ctx.quad(i,j,k,da,1), wb0 = ctx.quad(i,j,k,db,0), wb1 = ctx.quad(i,j,k,db,1);
    return (wa0 * la0 + wa1 * la1 + wb0 * lb0 + wb1 * lb1) * np_rcp((wa0 + wa1) + (wb0 + wb1));
}
__global__ void __launch_bounds__(128) k_interp_add_w(LevelDev Lf, LevelDev Lc, double* __restrict__ fine, const double* __restrict__ crse,
                                                      const double* __restrict__ sig, int koffc)
{
    const int I = blockIdx.x * blockDim.x + threadIdx.x;
    const int J = blockIdx.y * blockDim.y + threadIdx.y;
    const int K = blockIdx.z;
    if (2 * I > Lf.n[0] || 2 * J > Lf.n[1]) return;
    const int i = 2 * I, j = 2 * J, k = 2 * K;
    const bool ax = (i + 1 <= Lf.n[0]), ay = (j + 1 <= Lf.n[1]), az = (k + 1 <= Lf.n[2]);
    InterpCtx ctx{crse + np_noff(Lc, 0, 0, koffc), Lc.nsy, Lc.nsz, sig + np_coff(Lf, 0, 0, 0), Lf.csy, Lf.csz};
    double v[2][2][2];                     // [c][b][a]
    v[0][0][0] = ctx.crse(i, j, k);
    // ---- the 12 edges of the coarse cell ----
    double lx[2][2], ly[2][2], lz[2][2];   // lx[b][c] at (i+1, j+2b, k+2c); ly[a][c] at (i+2a, j+1, k+2c); lz[a][b] at (i+2a, j+2b, k+1)
#pragma unroll
    for (int p = 0; p < 2; ++p)
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            lx[p][q] = (ax && (p == 0 || ay) && (q == 0 || az)) ? ctx.line(i + 1, j + 2 * p, k + 2 * q, 0) : 0.0;
            ly[p][q] = (ay && (p == 0 || ax) && (q == 0 || az)) ? ctx.line(i + 2 * p, j + 1, k + 2 * q, 1) : 0.0;
            lz[p][q] = (az && (p == 0 || ax) && (q == 0 || ay)) ? ctx.line(i + 2 * p, j + 2 * q, k + 1, 2) : 0.0;
        }
    v[0][0][1] = lx[0][0]; v[0][1][0] = ly[0][0]; v[1][0][0] = lz[0][0];
    // ---- its 6 faces (same argument order as the per-node definition: xy -> (0,1), xz -> (2,0), yz -> (1,2)) ----
    double fxy[2], fxz[2], fyz[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        fxy[q] = (ax && ay && (q == 0 || az)) ? np_interp_face(ctx, i + 1, j + 1, k + 2 * q, 0, 1, ly[0][q], ly[1][q], lx[0][q], lx[1][q]) : 0.0;
        fxz[q] = (ax && az && (q == 0 || ay)) ? np_interp_face(ctx, i + 1, j + 2 * q, k + 1, 2, 0, lx[q][0], lx[q][1], lz[0][q], lz[1][q]) : 0.0;
        fyz[q] = (ay && az && (q == 0 || ax)) ? np_interp_face(ctx, i + 2 * q, j + 1, k + 1, 1, 2, lz[q][0], lz[q][1], ly[q][0], ly[q][1]) : 0.0;
    }
    v[0][1][1] = fxy[0]; v[1][0][1] = fxz[0]; v[1][1][0] = fyz[0];
    // ---- centre ----
    v[1][1][1] = 0.0;
    if (ax && ay && az) {
        const int ic = i + 1, jc = j + 1, kc = k + 1;
        const double x0 = ctx.quad(ic,jc,kc,0,0), x1 = ctx.quad(ic,jc,kc,0,1), y0 = ctx.quad(ic,jc,kc,1,0), y1 = ctx.quad(ic,jc,kc,1,1),
                     z0 = ctx.quad(ic,jc,kc,2,0), z1 = ctx.quad(ic,jc,kc,2,1);
        const double num = (x0 * fyz[0] + x1 * fyz[1]) + (y0 * fxz[0] + y1 * fxz[1]) + (z0 * fxy[0] + z1 * fxy[1]);
        v[1][1][1] = num * np_rcp(((x0 + x1) + (y0 + y1)) + (z0 + z1));
    }
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int b = 0; b < 2; ++b)
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                if ((a && !ax) || (b && !ay) || (c && !az)) continue;
                if (np_is_dirichlet(Lf, i + a, j + b, k + c)) continue;
                fine[np_noff(Lf, i + a, j + b, k + c)] += v[c][b][a];
            }
}

#include "np_interp_w.cuh"

// K5, plain trilinear (constant sigma, or interp_sigma_weighted = 0): one thread per COARSE node (I,J,K) updates the
// 2x2x2 fine nodes (2I-1..2I, 2J-1..2J, 2K-1..2K) from the coarse nodes (I-1..I, J-1..J, K-1..K); fine pairs in x are
// 16-byte aligned in the padded layout (node i at offset i+1), so each row is one 128-bit read-modify-write.
__global__ void __launch_bounds__(256) k_interp_add_tri(LevelDev Lf, LevelDev Lc, double* __restrict__ fine, const double* __restrict__ crse, int koffc)
{
    const int I = blockIdx.x * blockDim.x + threadIdx.x;
    const int J = blockIdx.y * blockDim.y + threadIdx.y;
    const int K = blockIdx.z;
    if (I > Lc.n[0] || J > Lc.n[1]) return;
    const double* c = crse + np_noff(Lc, I, J, K + koffc);
    const long long cy = Lc.nsy, cz = Lc.nsz;
    double px[2][2][2];                    // [dz][dy][a]: interpolated in x
#pragma unroll
    for (int dz = 0; dz < 2; ++dz)
#pragma unroll
        for (int dy = 0; dy < 2; ++dy) {
            const double lo = c[-1 + cy * (dy - 1) + cz * (dz - 1)], hi = c[cy * (dy - 1) + cz * (dz - 1)];
            px[dz][dy][0] = 0.5 * (lo + hi); px[dz][dy][1] = hi;
        }
    const bool xdir_lo = (Lf.face[0] == NP_FACE_DIRICHLET), xdir_hi = (Lf.face[3] == NP_FACE_DIRICHLET);
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
        const int fk = 2 * K - 1 + cc;
        if (fk < 0) continue;
        if ((Lf.face[2] == NP_FACE_DIRICHLET && fk == 0) || (Lf.face[5] == NP_FACE_DIRICHLET && fk == Lf.n[2])) continue;
#pragma unroll
        for (int bb = 0; bb < 2; ++bb) {
            const int fj = 2 * J - 1 + bb;
            if (fj < 0) continue;
            if ((Lf.face[1] == NP_FACE_DIRICHLET && fj == 0) || (Lf.face[4] == NP_FACE_DIRICHLET && fj == Lf.n[1])) continue;
            double v[2];
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                const double y0 = bb ? px[0][1][a] : 0.5 * (px[0][0][a] + px[0][1][a]);
                const double y1 = bb ? px[1][1][a] : 0.5 * (px[1][0][a] + px[1][1][a]);
                v[a] = cc ? y1 : 0.5 * (y0 + y1);
            }
            double* fp = fine + np_noff(Lf, 2 * I - 1, fj, fk);       // offset 2I: 16-byte aligned
            const bool w0 = (I > 0), w1 = !((xdir_lo && I == 0) || (xdir_hi && 2 * I == Lf.n[0]));
            if (w0 && w1) {
                double2 f2 = *reinterpret_cast<double2*>(fp);
                f2.x += v[0]; f2.y += v[1];
                *reinterpret_cast<double2*>(fp) = f2;
            } else {
                if (w0) fp[0] += v[0];
                if (w1) fp[1] += v[1];
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K8: sigma coarsening (arithmetic mean of the 8 children), interior cells only
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_avgdown_sigma(LevelDev Lc, LevelDev Lf, double* __restrict__ sc, const double* __restrict__ sf)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= Lc.n[0] || j >= Lc.n[1]) return;
    const double* p = sf + np_coff(Lf, 2*i, 2*j, 2*k);
    const long long y = Lf.csy, z = Lf.csz;
    sc[np_coff(Lc, i, j, k)] = 0.125 * (((p[0] + p[1]) + (p[y] + p[y+1])) + ((p[z] + p[z+1]) + (p[z+y] + p[z+y+1])));
}

__global__ void __launch_bounds__(256) k_reduce_minmax(const double* __restrict__ bmin, const double* __restrict__ bmax, long long n, double* out)
{
    double mx = -1e300, mn = 1e300;
    for (long long t = threadIdx.x; t < n; t += blockDim.x) { mx = fmax(mx, bmax[t]); mn = fmin(mn, bmin[t]); }
    mx = np_block_reduce<true>(mx);
    mn = -np_block_reduce<true>(-mn);
    if (threadIdx.x == 0) { out[0] = mn; out[1] = mx; }
}

// ------------------------------------------------------------------------------------------------
// Velocity ghost fill (1 layer, periodic x/y[/z]) on the caller's fab — what compRHS's
// vel.FillBoundary does.  Physical ghosts (inflow data / zeros) are left untouched.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_fill_vel(LevelDev L, FabView u, int per0, int per1, int per2)
{
    const int e0 = L.n[0] + 2, e1 = L.n[1] + 2, e2 = L.n[2] + 2;
    const long long g0 = 2LL * e1 * e2, g1 = 2LL * e0 * e2, g2 = 2LL * e0 * e1;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < g0 + g1 + g2;
         t += (long long)gridDim.x * blockDim.x) {
        int i, j, k;
        if (t < g0)           { long long q = t;           int sd = (int)(q % 2); q /= 2; j = (int)(q % e1) - 1; k = (int)(q / e1) - 1; i = sd ? L.n[0] : -1; }
        else if (t < g0 + g1) { long long q = t - g0;      int sd = (int)(q % 2); q /= 2; i = (int)(q % e0) - 1; k = (int)(q / e0) - 1; j = sd ? L.n[1] : -1; }
        else                  { long long q = t - g0 - g1; int sd = (int)(q % 2); q /= 2; i = (int)(q % e0) - 1; j = (int)(q / e0) - 1; k = sd ? L.n[2] : -1; }
        int si = i, sj = j, sk = k;
        if (i < 0 || i >= L.n[0]) { if (!per0) continue; si = i < 0 ? i + L.n[0] : i - L.n[0]; }
        if (j < 0 || j >= L.n[1]) { if (!per1) continue; sj = j < 0 ? j + L.n[1] : j - L.n[1]; }
        if (k < 0 || k >= L.n[2]) { if (!per2) continue; sk = k < 0 ? k + L.n[2] : k - L.n[2]; }
        for (int n = 0; n < 3; ++n) u.p[np_foff(u, i, j, k, n)] = u.p[np_foff(u, si, sj, sk, n)];
    }
}

// ------------------------------------------------------------------------------------------------
// K1: rhs = div(u) at valid nodes, with the Neumann conventions of mlndlap_divu / impose_neumann_bc
// ------------------------------------------------------------------------------------------------
// Each block marches through a contiguous run of node planes: the cell layer between two node planes is read once and
// its three in-layer sums (x-differences, y-differences, z-values over the 4 cells around the node column) serve the
// node plane below and the one above:  rhs(k) = 1/4 [ dxinv_x (zzl Dx(k-1) + zzh Dx(k)) + dxinv_y (zzl Dy(k-1) + zzh Dy(k))
//                                                    + dxinv_z (Sz(k) - Sz(k-1)) ]  (x2 per Neumann face, mlndlap_impose_neumann_bc).
template <bool WSUM>
__global__ void __launch_bounds__(256) k_divu(LevelDev L, double* __restrict__ rhs, FabView u, double* __restrict__ part_sum)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const bool inside = (i <= L.n[0] && j <= L.n[1]);
    double tsum = 0.0;
    const int per = (L.n[2] + 1 + gridDim.z - 1) / gridDim.z;            // node planes per block
    const int k0 = blockIdx.z * per, k1 = min(k0 + per - 1, L.n[2]);
    if (inside && k0 <= k1) {
        // weight of the cells on the low / high side of the node (0 outside a Neumann wall)
        const double zxl = (L.face[0] == NP_FACE_NEUMANN && i == 0) ? 0.0 : 1.0, zxh = (L.face[3] == NP_FACE_NEUMANN && i == L.n[0]) ? 0.0 : 1.0;
        const double zyl = (L.face[1] == NP_FACE_NEUMANN && j == 0) ? 0.0 : 1.0, zyh = (L.face[4] == NP_FACE_NEUMANN && j == L.n[1]) ? 0.0 : 1.0;
        double fxy = 1.0;                                                  // x2 per Neumann face in x / y the node lies on
        if (zxl == 0.0) fxy *= 2.0;
        if (zxh == 0.0) fxy *= 2.0;
        if (zyl == 0.0) fxy *= 2.0;
        if (zyh == 0.0) fxy *= 2.0;
        const bool dxy = (L.face[0] == NP_FACE_DIRICHLET && i == 0) || (L.face[3] == NP_FACE_DIRICHLET && i == L.n[0])
                      || (L.face[1] == NP_FACE_DIRICHLET && j == 0) || (L.face[4] == NP_FACE_DIRICHLET && j == L.n[1]);
        const long long y = u.sy;
        // in-layer sums of cell layer c (cells (i-1..i, j-1..j, c))
        auto layer = [&](int c, double& dx, double& dy, double& sz) {
            const double* ux = u.p + np_foff(u, i, j, c, 0);
            const double* uy = ux + u.sn; const double* uz = ux + 2 * u.sn;
            dx = zyl * (ux[-y] - ux[-1 - y]) + zyh * (ux[0] - ux[-1]);
            dy = zxl * (uy[-1] - uy[-1 - y]) + zxh * (uy[0] - uy[-y]);
            sz = (zxl * zyl * uz[-1 - y] + zxh * zyl * uz[-y]) + (zxl * zyh * uz[-1] + zxh * zyh * uz[0]);
        };
        double dxm, dym, szm;
        layer(k0 - 1, dxm, dym, szm);
        for (int k = k0; k <= k1; ++k) {
            double dxc, dyc, szc;
            layer(k, dxc, dyc, szc);
            const double zzl = (L.face[2] == NP_FACE_NEUMANN && k == 0) ? 0.0 : 1.0, zzh = (L.face[5] == NP_FACE_NEUMANN && k == L.n[2]) ? 0.0 : 1.0;
            double acc = 0.25 * (L.dxinv[0] * (zzl * dxm + zzh * dxc) + L.dxinv[1] * (zzl * dym + zzh * dyc) + L.dxinv[2] * (szc - szm)) * fxy;
            if (zzl == 0.0) acc *= 2.0;
            if (zzh == 0.0) acc *= 2.0;
            if (dxy || (L.face[2] == NP_FACE_DIRICHLET && k == 0) || (L.face[5] == NP_FACE_DIRICHLET && k == L.n[2])) acc = 0.0;
            rhs[np_noff(L, i, j, k)] = acc;
            if (WSUM && i <= L.owned_hi[0] && j <= L.owned_hi[1] && k <= L.owned_hi[2]) tsum += np_neumann_weight(L, i, j, k) * acc;
            dxm = dxc; dym = dyc; szm = szc;
        }
    }
    if (WSUM) {
        const double sm = np_block_reduce<false>(tsum);
        if (threadIdx.x == 0 && threadIdx.y == 0)
            part_sum[blockIdx.x + gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z)] = sm;
    }
}
// scal[slot_off] = sum(part) / wtot (single rank), scal[slot_sum] = sum(part): the offset that makes rhs solvable
__global__ void __launch_bounds__(1024) k_reduce_sum_off(const double* __restrict__ part, long long n, double* scal, int slot_sum, int slot_off, double wtot)
{
    double sm = 0.0;
    for (long long t = threadIdx.x; t < n; t += blockDim.x) sm += part[t];
    sm = np_block_reduce<false>(sm);
    if (threadIdx.x == 0) { scal[slot_sum] = sm; scal[slot_off] = sm / wtot; }
}
__global__ void k_div_scalar(double* scal, int slot_sum, int slot_off, double wtot) { if (threadIdx.x == 0 && blockIdx.x == 0) scal[slot_off] = scal[slot_sum] / wtot; }

// min / max of the caller's sigma fab (valid cells), plane-strided so that few partials remain
__global__ void __launch_bounds__(256) k_minmax_cell(int n0, int n1, int n2, FabView src, double* __restrict__ bmin, double* __restrict__ bmax)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    double mx = -1e300, mn = 1e300;
    if (i < n0 && j < n1)
        for (int k = blockIdx.z; k < n2; k += gridDim.z) { const double v = src.p[np_foff(src, i, j, k, 0)]; mx = fmax(mx, v); mn = fmin(mn, v); }
    mx = np_block_reduce<true>(mx);
    mn = -np_block_reduce<true>(-mn);
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        const long long bidx = blockIdx.x + gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z);
        bmin[bidx] = mn; bmax[bidx] = mx;
    }
}
__global__ void __launch_bounds__(256) k_copy_sigma_plain(LevelDev L, double* __restrict__ s, FabView src)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i < L.n[0] && j < L.n[1]) s[np_coff(L, i, j, k)] = src.p[np_foff(src, i, j, k, 0)];
}

// ------------------------------------------------------------------------------------------------
// K7: u -= sigma * grad(phi), gradphi = grad(phi); one thread per cell, all three outputs in one pass
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_mknewu(LevelDev L, FabView u, FabView g, const double* __restrict__ phi, const double* __restrict__ sig)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= L.n[0] || j >= L.n[1]) return;
    const double* q = phi + np_noff(L, i, j, k);
    const long long y = L.nsy, z = L.nsz;
    const double gx = 0.25 * L.dxinv[0] * (((q[1] - q[0]) + (q[1+y] - q[y])) + ((q[1+z] - q[z]) + (q[1+y+z] - q[y+z])));
    const double gy = 0.25 * L.dxinv[1] * (((q[y] - q[0]) + (q[1+y] - q[1])) + ((q[y+z] - q[z]) + (q[1+y+z] - q[1+z])));
    const double gz = 0.25 * L.dxinv[2] * (((q[z] - q[0]) + (q[1+z] - q[1])) + ((q[y+z] - q[y]) + (q[1+y+z] - q[1+y])));
    const double s = L.const_sigma ? L.sigma0 : sig[np_coff(L, i, j, k)];
    double* up = u.p + np_foff(u, i, j, k, 0);
    up[0] -= s * gx; up[u.sn] -= s * gy; up[2 * u.sn] -= s * gz;
    double* gp = g.p + np_foff(g, i, j, k, 0);
    gp[0] = gx; gp[g.sn] = gy; gp[2 * g.sn] = gz;
}

// ------------------------------------------------------------------------------------------------
// K9: the caller's pointwise passes around the projection (incflo::ApplyProjection,
// /root/reference/src/projection/incflo_apply_nodal_projection.cpp:79-131 before, :156-203 after), folded into two
// kernels so that u, rho, gp are streamed once on each side and the caller's sigma MultiFab never exists.
//   pre  (:79-131): u += gp * scaling/rho unless incremental; u -= u_old for the incremental / small-dt forms;
//                   sigma = scaling/rho goes straight into the level-0 coefficient array; per-block min / max of sigma.
//   post (:153-203): u -= sigma grad(phi) (+ u_old); gp = grad(phi) or gp += grad(phi); p = phi or p += phi.
// ------------------------------------------------------------------------------------------------
struct ApPre { double scaling; int add_gp, sub_old; };
__global__ void __launch_bounds__(256) k_apply_pre(LevelDev L, FabView u, FabView uo, FabView rho, FabView gp, double* __restrict__ sig,
                                                   ApPre a, double* __restrict__ bmin, double* __restrict__ bmax)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    double mx = -1e300, mn = 1e300;
    if (i < L.n[0] && j < L.n[1]) {
        const double s = a.scaling / rho.p[np_foff(rho, i, j, k, 0)];
        sig[np_coff(L, i, j, k)] = s;
        mx = s; mn = s;
        if (a.add_gp || a.sub_old) {
            double* up = u.p + np_foff(u, i, j, k, 0);
#pragma unroll
            for (int n = 0; n < 3; ++n) {
                double v = up[u.sn * n];
                if (a.add_gp) v += gp.p[np_foff(gp, i, j, k, n)] * s;
                if (a.sub_old) v -= uo.p[np_foff(uo, i, j, k, n)];
                up[u.sn * n] = v;
            }
        }
    }
    mx = np_block_reduce<true>(mx);
    mn = -np_block_reduce<true>(-mn);
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        const long long bidx = blockIdx.x + gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z);
        bmin[bidx] = mn; bmax[bidx] = mx;
    }
}
// MultiFab::setBndry(0) on a 3-component cell fab: every ghost cell of every layer (:142)
__global__ void __launch_bounds__(256) k_zero_cell_ghosts(int n0, int n1, int n2, FabView u)
{
    const int g = u.ng;
    const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x) - g;
    const int j = (int)(blockIdx.y * blockDim.y + threadIdx.y) - g;
    const int k = (int)blockIdx.z - g;
    if (i >= n0 + g || j >= n1 + g) return;
    if (i >= 0 && i < n0 && j >= 0 && j < n1 && k >= 0 && k < n2) return;
    double* up = u.p + np_foff(u, i, j, k, 0);
    up[0] = 0.0; up[u.sn] = 0.0; up[2 * u.sn] = 0.0;
}
struct ApPost { int add_old, accumulate; };
__global__ void __launch_bounds__(256) k_apply_post(LevelDev L, FabView u, FabView uo, FabView g, const double* __restrict__ phi,
                                                    const double* __restrict__ sig, ApPost a)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= L.n[0] || j >= L.n[1]) return;
    const double* q = phi + np_noff(L, i, j, k);
    const long long y = L.nsy, z = L.nsz;
    double gr[3];
    gr[0] = 0.25 * L.dxinv[0] * (((q[1] - q[0]) + (q[1+y] - q[y])) + ((q[1+z] - q[z]) + (q[1+y+z] - q[y+z])));
    gr[1] = 0.25 * L.dxinv[1] * (((q[y] - q[0]) + (q[1+y] - q[1])) + ((q[y+z] - q[z]) + (q[1+y+z] - q[1+z])));
    gr[2] = 0.25 * L.dxinv[2] * (((q[z] - q[0]) + (q[1+z] - q[1])) + ((q[y+z] - q[y]) + (q[1+y+z] - q[1+y])));
    const double s = L.const_sigma ? L.sigma0 : sig[np_coff(L, i, j, k)];
    double* up = u.p + np_foff(u, i, j, k, 0);
    double* gp = g.p + np_foff(g, i, j, k, 0);
#pragma unroll
    for (int n = 0; n < 3; ++n) {
        double v = up[u.sn * n] - s * gr[n];
        if (a.add_old) v += uo.p[np_foff(uo, i, j, k, n)];
        up[u.sn * n] = v;
        gp[g.sn * n] = a.accumulate ? gp[g.sn * n] + gr[n] : gr[n];
    }
}
// p = phi or p += phi on the valid nodes of the caller's nodal fab (:187-200)
__global__ void __launch_bounds__(256) k_apply_post_p(LevelDev L, const double* __restrict__ phi, FabView p, int accumulate)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i > L.n[0] || j > L.n[1]) return;
    double* pp = p.p + np_foff(p, i, j, k, 0);
    const double v = phi[np_noff(L, i, j, k)];
    *pp = accumulate ? *pp + v : v;
}

// ------------------------------------------------------------------------------------------------
// Copies between the padded internal nodal layout and a caller fab (1 comp).  ng_copy ghost layers of the
// fab are transferred too (clamped to 1).
// ------------------------------------------------------------------------------------------------
template <bool TO_FAB>
__global__ void __launch_bounds__(256) k_copy_nodal(LevelDev L, double* __restrict__ x, FabView f, int ngc)
{
    const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x) - ngc;
    const int j = (int)(blockIdx.y * blockDim.y + threadIdx.y) - ngc;
    const int k = (int)blockIdx.z - ngc;
    if (i > L.n[0] + ngc || j > L.n[1] + ngc) return;
    if (TO_FAB) f.p[np_foff(f, i, j, k, 0)] = x[np_noff(L, i, j, k)];
    else        x[np_noff(L, i, j, k)] = f.p[np_foff(f, i, j, k, 0)];
}
// cell fab -> cell fab (ncomp components, interior + ngc ghosts), used for host-staged / copy-out paths
__global__ void __launch_bounds__(256) k_copy_cell(int n0, int n1, FabView dst, FabView src, int ncomp, int ngc)
{
    const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x) - ngc;
    const int j = (int)(blockIdx.y * blockDim.y + threadIdx.y) - ngc;
    const int k = (int)blockIdx.z - ngc;
    if (i >= n0 + ngc || j >= n1 + ngc) return;
    for (int n = 0; n < ncomp; ++n) dst.p[np_foff(dst, i, j, k, n)] = src.p[np_foff(src, i, j, k, n)];
}
// Several boxes per rank (np_project with nlocal > 1): one box of the caller's MultiFab <-> this rank's slab.
// Indices are slab-local.  `dst` / `src` are views whose index origin (first valid cell / node) sits at dorg / sorg.
// The region [rlo, rhi] is copied; with guard != 0 (gather of vel, whose ghost layer is part of the region) a cell is
// taken only if it lies inside the box's valid range [vlo, vhi] or outside the problem domain [dlo, dhi] - the physical
// ghost cells ApplyProjection prepared (zero / inflow); ghost cells over other boxes belong to those boxes.
struct BoxCopy { int rlo[3], rhi[3], dorg[3], sorg[3], vlo[3], vhi[3], dlo[3], dhi[3]; int guard; };
__global__ void __launch_bounds__(256) k_box_copy(FabView dst, FabView src, BoxCopy a, int ncomp)
{
    const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x) + a.rlo[0];
    const int j = (int)(blockIdx.y * blockDim.y + threadIdx.y) + a.rlo[1];
    const int k = (int)blockIdx.z + a.rlo[2];
    if (i > a.rhi[0] || j > a.rhi[1]) return;
    if (a.guard) {
        const bool in_box = i >= a.vlo[0] && i <= a.vhi[0] && j >= a.vlo[1] && j <= a.vhi[1] && k >= a.vlo[2] && k <= a.vhi[2];
        const bool outside = i < a.dlo[0] || i > a.dhi[0] || j < a.dlo[1] || j > a.dhi[1] || k < a.dlo[2] || k > a.dhi[2];
        if (!in_box && !outside) return;
    }
    for (int n = 0; n < ncomp; ++n)
        dst.p[np_foff(dst, i - a.dorg[0], j - a.dorg[1], k - a.dorg[2], n)] = src.p[np_foff(src, i - a.sorg[0], j - a.sorg[1], k - a.sorg[2], n)];
}
__global__ void __launch_bounds__(256) k_copy_sigma_out(LevelDev L, const double* __restrict__ s, FabView dst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= L.n[0] || j >= L.n[1]) return;
    dst.p[np_foff(dst, i, j, k, 0)] = L.const_sigma ? L.sigma0 : s[np_coff(L, i, j, k)];
}

// ------------------------------------------------------------------------------------------------
// K6: bottom solver — the whole BiCGStab / CG iteration of amrex::MLCGSolver in ONE CTA: the coarsest
// level is tiny (3^3 .. 33^3 nodes), so the ~7 reductions and 2 operator applications per iteration are
// block barriers instead of kernel launches / allreduces.  Vectors live in HBM (L2-resident).
// ------------------------------------------------------------------------------------------------
#define NP_BOTTOM_SMEM_MAX (160 * 1024)      // bytes of dynamic shared memory k_bottom_solve may use for its eight vectors
struct BottomArgs {
    LevelDev L;
    double* x; const double* b;             // solution (out), right-hand side (in; not modified)
    double *r, *rh, *p, *v, *s, *t, *bb;    // work arrays (full nodal arrays with ghosts)
    const double* sig;
    int maxiter; double eps_rel, eps_abs;
    int singular, normalize, type;          // type: 0 BiCGStab, 1 CG
    int in_smem;                            // 1: the vectors live in dynamic shared memory (8 * L.ntot doubles)
    int* out;                               // out[0] = ret code, out[1] = iterations
};

struct BottomCtx {
    const BottomArgs& A;
    int tid, nth;
    long long nvalid;
    __device__ BottomCtx(const BottomArgs& a) : A(a) {
        tid = threadIdx.x; nth = blockDim.x;
        nvalid = (long long)(a.L.n[0] + 1) * (a.L.n[1] + 1) * (a.L.n[2] + 1);
    }
    __device__ __forceinline__ void ijk(long long q, int& i, int& j, int& k) const {      // (the level has < 2^31 nodes: 32-bit divisions)
        const int e0 = A.L.n[0] + 1, e1 = A.L.n[1] + 1, q32 = (int)q;
        const int t = q32 / e0;
        i = q32 - t * e0; k = t / e1; j = t - k * e1;
    }
    __device__ void fill(double* x) const {
        const LevelDev& L = A.L;
        const int e0 = L.n[0] + 3, e1 = L.n[1] + 3, e2 = L.n[2] + 3;
        for (int q = tid; q < e0 * e1 * e2; q += nth) {
            const int t = q / e0, i = q - t * e0 - 1, k = t / e1 - 1, j = t - (k + 1) * e1 - 1;
            const int si = np_map_index(i, L.n[0], L.map_lo[0], L.map_hi[0]);
            const int sj = np_map_index(j, L.n[1], L.map_lo[1], L.map_hi[1]);
            const int sk = np_map_index(k, L.n[2], L.map_lo[2], L.map_hi[2]);
            if (si == i && sj == j && sk == k) continue;
            if (si < 0 || si > L.n[0] || sj < 0 || sj > L.n[1] || sk < 0 || sk > L.n[2]) continue;
            x[np_noff(L, i, j, k)] = x[np_noff(L, si, sj, sk)];
        }
        __syncthreads();
    }
    // y = A x (x gets its ghosts filled), optionally divided by the diagonal (MLCGSolver's Lp.normalize)
    __device__ void apply(double* y, double* x, bool normalize) const {
        const LevelDev& L = A.L;
        fill(x);
        for (long long q = tid; q < nvalid; q += nth) {
            int i, j, k; ijk(q, i, j, k);
            const long long o = np_noff(L, i, j, k);
            double val = 0.0;
            if (!np_is_dirichlet(L, i, j, k)) {
                GlobX xa{x + o, L.nsy, L.nsz};
                double d, off;
                if (L.const_sigma) { d = np_diag_const(L.c, L.sigma0); off = np_offdiag_const(L.c, L.sigma0, xa); }
                else {
                    GlobS sa{A.sig + np_coff(L, i, j, k), L.csy, L.csz};
                    Sig8 g = np_load_sig(sa);
                    d = np_diag(L.c, g); off = np_offdiag(L.c, g, xa);
                }
                val = xa(0,0,0) * d + off;
                if (normalize) val /= d;
            }
            y[o] = val;
        }
        __syncthreads();
    }
    __device__ void normalize_vec(double* x) const {
        const LevelDev& L = A.L;
        for (long long q = tid; q < nvalid; q += nth) {
            int i, j, k; ijk(q, i, j, k);
            const long long o = np_noff(L, i, j, k);
            if (np_is_dirichlet(L, i, j, k)) { x[o] = 0.0; continue; }
            double d;
            if (L.const_sigma) d = np_diag_const(L.c, L.sigma0);
            else { GlobS sa{A.sig + np_coff(L, i, j, k), L.csy, L.csz}; Sig8 g = np_load_sig(sa); d = np_diag(L.c, g); }
            x[o] /= d;
        }
        __syncthreads();
    }
    __device__ double dot(const double* x, const double* y) const {
        const LevelDev& L = A.L;
        double a = 0.0;
        for (long long q = tid; q < nvalid; q += nth) {
            int i, j, k; ijk(q, i, j, k);
            if (i > L.owned_hi[0] || j > L.owned_hi[1] || k > L.owned_hi[2]) continue;
            const long long o = np_noff(L, i, j, k);
            a += np_neumann_weight(L, i, j, k) * x[o] * y[o];
        }
        return np_block_reduce<false>(a);
    }
    __device__ double norminf(const double* x) const {
        const LevelDev& L = A.L;
        double a = 0.0;
        for (long long q = tid; q < nvalid; q += nth) { int i, j, k; ijk(q, i, j, k); a = fmax(a, fabs(x[np_noff(L, i, j, k)])); }
        return np_block_reduce<true>(a);
    }
    __device__ double wsum(const double* x, double& wtot) const {
        const LevelDev& L = A.L;
        double a = 0.0, w = 0.0;
        for (long long q = tid; q < nvalid; q += nth) {
            int i, j, k; ijk(q, i, j, k);
            if (i > L.owned_hi[0] || j > L.owned_hi[1] || k > L.owned_hi[2]) continue;
            const double ww = np_neumann_weight(L, i, j, k);
            a += ww * x[np_noff(L, i, j, k)]; w += ww;
        }
        a = np_block_reduce<false>(a); wtot = np_block_reduce<false>(w);
        return a;
    }
};

__global__ void __launch_bounds__(1024) k_bottom_solve(BottomArgs A)
{
    BottomCtx C(A);
    const LevelDev& L = A.L;
    const int tid = C.tid, nth = C.nth;
    const long long ntot = L.ntot;
    // the eight vectors of the iteration: in shared memory when the level fits (the usual 3^3 .. 9^3-node bottom level: every
    // phase of an iteration is then a barrier away from the next instead of an L2 round trip - 67 -> ~10 us per bottom solve
    // at 3^3 nodes), else the caller's work arrays in HBM
    NP_DYN_SMEM(double, bsm);
    double *X = A.x, *BB = A.bb, *R = A.r, *RH = A.rh, *P = A.p, *V = A.v, *S = A.s, *T = A.t;
    if (A.in_smem) { X = bsm; BB = X + ntot; R = BB + ntot; RH = R + ntot; P = RH + ntot; V = P + ntot; S = V + ntot; T = S + ntot;
                     for (long long q = tid; q < ntot; q += nth) { R[q] = 0.0; RH[q] = 0.0; V[q] = 0.0; T[q] = 0.0; } }
    // bb = b (made solvable when the operator is singular: MLMG::bottomSolve -> makeSolvable)
    for (long long q = tid; q < ntot; q += nth) { BB[q] = A.b[q]; X[q] = 0.0; P[q] = 0.0; S[q] = 0.0; }
    __syncthreads();
    if (A.singular) {
        double wtot; double sum = C.wsum(BB, wtot);
        const double off = sum / wtot;
        for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); BB[np_noff(L, i, j, k)] -= off; }
        __syncthreads();
    }
    int ret = 0, iter = 1;
    double rnorm, rnorm0;
    if (A.type == 0) {
        // ---------------- BiCGStab (MLCGSolver::solve_bicgstab), initial guess 0 -> r = b ----------------
        for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); const long long o = np_noff(L, i, j, k); R[o] = BB[o]; }
        __syncthreads();
        if (A.normalize) C.normalize_vec(R);
        for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); const long long o = np_noff(L, i, j, k); RH[o] = R[o]; }
        __syncthreads();
        rnorm = C.norminf(R); rnorm0 = rnorm;
        if (rnorm0 == 0.0 || rnorm0 < A.eps_abs) { if (A.in_smem) for (long long q = tid; q < ntot; q += nth) A.x[q] = 0.0;
            if (tid == 0) { A.out[0] = 0; A.out[1] = 0; } return; }
        double rho_1 = 0.0, alpha = 0.0, omega = 0.0;
        for (; iter <= A.maxiter; ++iter) {
            const double rho = C.dot(RH, R);
            if (rho == 0.0) { ret = 1; break; }
            if (iter == 1) {
                for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); const long long o = np_noff(L, i, j, k); P[o] = R[o]; }
            } else {
                const double beta = (rho / rho_1) * (alpha / omega);
                for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); const long long o = np_noff(L, i, j, k);
                    const double pv = P[o] - omega * V[o]; P[o] = R[o] + beta * pv; }
            }
            __syncthreads();
            C.apply(V, P, A.normalize);
            const double rhTv = C.dot(RH, V);
            if (rhTv != 0.0) alpha = rho / rhTv; else { ret = 2; break; }
            for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); const long long o = np_noff(L, i, j, k);
                X[o] += alpha * P[o]; S[o] = R[o] - alpha * V[o]; }
            __syncthreads();
            rnorm = C.norminf(S);
            if (rnorm < A.eps_rel * rnorm0 || rnorm < A.eps_abs) break;
            C.apply(T, S, A.normalize);
            const double tt = C.dot(T, T), ts = C.dot(T, S);
            if (tt != 0.0) omega = ts / tt; else { ret = 3; break; }
            for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); const long long o = np_noff(L, i, j, k);
                X[o] += omega * S[o]; R[o] = S[o] - omega * T[o]; }
            __syncthreads();
            rnorm = C.norminf(R);
            if (rnorm < A.eps_rel * rnorm0 || rnorm < A.eps_abs) break;
            if (omega == 0.0) { ret = 4; break; }
            rho_1 = rho;
        }
    } else {
        // ---------------- CG (MLCGSolver::solve_cg) ----------------
        for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); const long long o = np_noff(L, i, j, k); R[o] = BB[o]; }
        __syncthreads();
        rnorm = C.norminf(R); rnorm0 = rnorm;
        if (rnorm0 == 0.0 || rnorm0 < A.eps_abs) { if (A.in_smem) for (long long q = tid; q < ntot; q += nth) A.x[q] = 0.0;
            if (tid == 0) { A.out[0] = 0; A.out[1] = 0; } return; }
        double rho_1 = 0.0;
        for (; iter <= A.maxiter; ++iter) {
            const double rho = C.dot(R, R);
            if (rho == 0.0) { ret = 1; break; }
            if (iter == 1) {
                for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); const long long o = np_noff(L, i, j, k); P[o] = R[o]; }
            } else {
                const double beta = rho / rho_1;
                for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); const long long o = np_noff(L, i, j, k); P[o] = R[o] + beta * P[o]; }
            }
            __syncthreads();
            C.apply(V, P, false);
            const double pw = C.dot(P, V);
            double alpha;
            if (pw != 0.0) alpha = rho / pw; else { ret = 1; break; }
            for (long long q = tid; q < C.nvalid; q += nth) { int i, j, k; C.ijk(q, i, j, k); const long long o = np_noff(L, i, j, k);
                X[o] += alpha * P[o]; R[o] -= alpha * V[o]; }
            __syncthreads();
            rnorm = C.norminf(R);
            if (rnorm < A.eps_rel * rnorm0 || rnorm < A.eps_abs) break;
            rho_1 = rho;
        }
    }
    if (ret == 0 && rnorm > A.eps_rel * rnorm0 && rnorm > A.eps_abs) ret = 8;
    if (!((ret == 0 || ret == 8) && rnorm < rnorm0)) {
        for (long long q = tid; q < ntot; q += nth) X[q] = 0.0;     // sol = sorig (= 0)
    }
    if (A.in_smem) { __syncthreads(); for (long long q = tid; q < ntot; q += nth) A.x[q] = X[q]; }
    if (tid == 0) { A.out[0] = ret; A.out[1] = iter < A.maxiter ? iter : A.maxiter; }
}

// ------------------------------------------------------------------------------------------------
// Multi-GPU agglomeration: every rank contributed its local coarse slab (layout Ll, ntot_l / ctot_l
// elements per rank) to an all-gather; these kernels assemble the replicated whole-domain level Lg.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_assemble_nodal(LevelDev Lg, LevelDev Ll, double* __restrict__ dst, const double* __restrict__ recv, int nranks)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int K = blockIdx.z;
    if (i > Lg.n[0] || j > Lg.n[1]) return;
    int r = K / Ll.n[2]; if (r > nranks - 1) r = nranks - 1;
    const int kl = K - r * Ll.n[2];
    dst[np_noff(Lg, i, j, K)] = recv[(long long)r * Ll.ntot + np_noff(Ll, i, j, kl)];
}
__global__ void __launch_bounds__(256) k_assemble_cell(LevelDev Lg, LevelDev Ll, double* __restrict__ dst, const double* __restrict__ recv)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int K = blockIdx.z;
    if (i >= Lg.n[0] || j >= Lg.n[1]) return;
    const int r = K / Ll.n[2];
    const int kl = K - r * Ll.n[2];
    dst[np_coff(Lg, i, j, K)] = recv[(long long)r * Ll.ctot + np_coff(Ll, i, j, kl)];
}
__global__ void k_negate_one(double* p) { if (threadIdx.x == 0 && blockIdx.x == 0) p[0] = -p[0]; }
