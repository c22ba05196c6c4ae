// mac_kernels.cuh — fp64 sm_100a kernels of the MAC projection: the cell-centred
// variable-coefficient 7-point operator of amrex::MLABecLaplacian (alpha = 0, beta = 1, face-centred b = 1/rho) as
// amrex::MacProjector drives it from /root/reference/src/convection/incflo_compute_MAC_projected_velocities.cpp:36-131.
//
// HBM layout: a cell array of a level has NO ghost layer: (i,j,k) -> i + sx*(j + n1*k), sx = n0 rounded up to even, so an
// aligned pair of cells (i even) is one 16-byte access.  Boundary conditions (MLCellLinOp::applyBC) are evaluated where a
// stencil reaches across a domain face - periodic wrap, Neumann copy, Dirichlet polynomial ghost - instead of a ghost-fill
// launch before every colour.  Face coefficients: bx (i in 0..n0) pitch fxs = n0+1 rounded up to even, rows j, planes k;
// by pitch sx with n1+1 rows per plane; bz pitch sx, n2+1 planes.  A uniform b (constant density) is not read at all.
#pragma once
#include "np_kernels.cuh"

enum { MAC_PER = 0, MAC_DIR = 1, MAC_NEU = 2, MAC_REMOTE = 3 };   // REMOTE: a z face shared with the neighbouring rank's slab:
                                                                  // the neighbour value is in the ghost plane the halo exchange filled

struct MacLevel {
    int n[3];
    long long sx, sxy, ntot;      // cell pitch, plane pitch, allocation
    long long fxs;                // row pitch of bx
    double dxinv[3], dh[3];       // 1/dx, 1/dx^2
    int bclo[3], bchi[3];
    double dc0[3], dc1[3];        // Dirichlet ghost = dc0*phi0 + dc1*phi1 (mllinop_apply_bc, maxorder 3: -2, 1/3)
    double flo[3], fhi[3];        // coefficient of the first interior cell in its ghost (smoother's boundary factors)
    int const_beta; double beta0;
    int koff;                     // global z index of local plane 0 (colour parity of a slab); 0 on one rank / replicated levels
    int kcoff;                    // plane of the next coarser level's array that holds the parent of local plane 0
    // MLABecLaplacian's alpha * a term (the scalar / velocity-component diffusion solves): cell array
    // in the layout of the level's cell arrays, or null (alpha = 0: the MAC projection).  dh[] then carries beta / dx^2.
    double ascal; const double* acoef;
};
__device__ __forceinline__ double mac_aterm(const MacLevel& L, long long c) { return L.acoef ? L.ascal * L.acoef[c] : 0.0; }
struct MacBeta { const double *bx, *by, *bz; };

__host__ __device__ __forceinline__ long long mac_c(const MacLevel& L, int i, int j, int k) { return i + L.sx * (j + (long long)L.n[1] * k); }
__host__ __device__ __forceinline__ long long mac_fx(const MacLevel& L, int i, int j, int k) { return i + L.fxs * (j + (long long)L.n[1] * k); }
__host__ __device__ __forceinline__ long long mac_fy(const MacLevel& L, int i, int j, int k) { return i + L.sx * (j + (long long)(L.n[1] + 1) * k); }
__host__ __device__ __forceinline__ long long mac_fz(const MacLevel& L, int i, int j, int k) { return i + L.sx * (j + (long long)L.n[1] * k); }

// value of the neighbour of cell c (value xc) across its low / high side in direction d; `st` = element stride of d,
// `idx` = the cell's index along d.  Inside the box a plain load; on a domain face the boundary condition.
// `xw` is the array periodic images are read from: x itself, or a snapshot taken before the launch when the level has an
// odd number of cells in a periodic direction (then the image has the cell's own colour and is being updated).
__device__ __forceinline__ double mac_nb_lo(const MacLevel& L, const double* __restrict__ x, long long c, double xc, int d, int idx, long long st,
                                            const double* __restrict__ xw)
{
    if (idx > 0 || L.bclo[d] == MAC_REMOTE) return x[c - st];
    if (L.bclo[d] == MAC_PER) return xw[c + st * (L.n[d] - 1)];
    if (L.bclo[d] == MAC_NEU) return xc;
    return L.dc0[d] * xc + L.dc1[d] * x[c + st];
}
__device__ __forceinline__ double mac_nb_hi(const MacLevel& L, const double* __restrict__ x, long long c, double xc, int d, int idx, long long st,
                                            const double* __restrict__ xw)
{
    if (idx < L.n[d] - 1 || L.bchi[d] == MAC_REMOTE) return x[c + st];
    if (L.bchi[d] == MAC_PER) return xw[c - st * (L.n[d] - 1)];
    if (L.bchi[d] == MAC_NEU) return xc;
    return L.dc0[d] * xc + L.dc1[d] * x[c - st];
}
struct MacB6 { double xm, xp, ym, yp, zm, zp; };
template <bool CB> __device__ __forceinline__ MacB6 mac_load_b(const MacLevel& L, const MacBeta& B, int i, int j, int k)
{
    MacB6 b;
    if (CB) { b.xm = b.xp = b.ym = b.yp = b.zm = b.zp = L.beta0; return b; }
    const long long fx = mac_fx(L, i, j, k), fy = mac_fy(L, i, j, k), fz = mac_fz(L, i, j, k);
    b.xm = B.bx[fx]; b.xp = B.bx[fx + 1];
    b.ym = B.by[fy]; b.yp = B.by[fy + L.sx];
    b.zm = B.bz[fz]; b.zp = B.bz[fz + L.sxy];
    return b;
}

// ------------------------------------------------------------------------------------------------
// MLABecLaplacian::Fsmooth (abec_gsrb), one colour: cells with (i+j+k+colour) even.  A thread owns an aligned pair of
// cells; exactly one of them has the colour.  It reads the pair as one 16-byte load (centre + one x neighbour), the
// other five neighbours as scalars, and stores the pair as a whole (no half-written sectors).
//   zero_in : phi is taken as 0 and not read (first pre-smooth after cor = 0: no memset, no read)
//   add_to  : sol += phi for the pair after the update (last post-smooth of level 0 on the black pass: MLMG's sol += cor)
//   boff    : device scalar subtracted from rhs on the fly (lazy makeSolvable of the level-0 residual), or null
// ------------------------------------------------------------------------------------------------
template <bool CB>
__global__ void __launch_bounds__(256) k_mac_gsrb(MacLevel L, MacBeta B, double* __restrict__ phi, const double* __restrict__ rhs, int colour,
                                                  int zero_in, double* __restrict__ add_to, const double* __restrict__ boff,
                                                  const double* __restrict__ xw)
{
    const int i0 = 2 * (int)(blockIdx.x * blockDim.x + threadIdx.x);
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i0 >= L.n[0] || j >= L.n[1]) return;
    const int a = (j + k + L.koff + colour) & 1;
    const int i = i0 + a;
    const long long c0 = mac_c(L, i0, j, k);
    const bool pair = i0 + 1 < L.n[0];
    double p0 = 0.0, p1 = 0.0;
    if (!zero_in) {
        if (pair) { const double2 v = *reinterpret_cast<const double2*>(phi + c0); p0 = v.x; p1 = v.y; }
        else p0 = phi[c0];
    }
    if (i < L.n[0]) {
        const long long c = c0 + a;
        const double xc = a ? p1 : p0;
        double rho = 0.0;
        const MacB6 b = mac_load_b<CB>(L, B, i, j, k);
        if (!zero_in) {
            const double xm = a ? p0 : mac_nb_lo(L, phi, c, xc, 0, i, 1, xw);
            const double xp = (!a && pair) ? p1 : mac_nb_hi(L, phi, c, xc, 0, i, 1, xw);
            const double ym = mac_nb_lo(L, phi, c, xc, 1, j, L.sx, xw), yp = mac_nb_hi(L, phi, c, xc, 1, j, L.sx, xw);
            const double zm = mac_nb_lo(L, phi, c, xc, 2, k, L.sxy, xw), zp = mac_nb_hi(L, phi, c, xc, 2, k, L.sxy, xw);
            rho = L.dh[0] * (b.xm * xm + b.xp * xp) + L.dh[1] * (b.ym * ym + b.yp * yp) + L.dh[2] * (b.zm * zm + b.zp * zp);
        }
        const double cf0 = i == 0 ? L.flo[0] : 0.0, cf3 = i == L.n[0] - 1 ? L.fhi[0] : 0.0;
        const double cf1 = j == 0 ? L.flo[1] : 0.0, cf4 = j == L.n[1] - 1 ? L.fhi[1] : 0.0;
        const double cf2 = k == 0 ? L.flo[2] : 0.0, cf5 = k == L.n[2] - 1 ? L.fhi[2] : 0.0;
        const double gamma = mac_aterm(L, c) + L.dh[0] * (b.xm + b.xp) + L.dh[1] * (b.ym + b.yp) + L.dh[2] * (b.zm + b.zp);
        const double g_m_d = gamma - (L.dh[0] * (b.xm * cf0 + b.xp * cf3) + L.dh[1] * (b.ym * cf1 + b.yp * cf4) + L.dh[2] * (b.zm * cf2 + b.zp * cf5));
        double r = rhs[c];
        if (boff) r -= *boff;
        const double res = r - (gamma * xc - rho);
        const double xn = xc + 1.15 / g_m_d * res;
        if (a) p1 = xn; else p0 = xn;
    }
    if (pair) {
        *reinterpret_cast<double2*>(phi + c0) = make_double2(p0, p1);
        if (add_to) { double2 s = *reinterpret_cast<double2*>(add_to + c0); s.x += p0; s.y += p1; *reinterpret_cast<double2*>(add_to + c0) = s; }
    } else {
        phi[c0] = p0;
        if (add_to) add_to[c0] += p0;
    }
}

// ------------------------------------------------------------------------------------------------
// One whole MLCellLinOp::smooth (red sweep, then black sweep) in ONE pass over the level.  Out of place: reads `in`,
// writes `out`.  A block owns a tile of RB_TX x RB_TY cells and a chunk of zc planes and marches through the chunk with a
// ring of four planes in shared memory, each holding the tile plus a halo of 2:
//     step p:  load plane p (old values)  |  red cells of plane p-1 (tile + 1) updated in the ring  |  black cells of
//              plane p-2 (tile) computed from the new reds around them and stored, with their red partners, as 16-byte pairs
// Red cells one cell outside the tile / chunk are recomputed by the neighbouring blocks as well (that is what makes the
// blocks independent); a red-black sweep does not depend on the order inside a colour, so the result is the reference's.
// Periodic images are loaded through wrapped indices and treated like any halo cell; Neumann / Dirichlet ghosts are
// evaluated from the ring where a stencil reaches across the face.  Needs an even number of cells in periodic directions.
//   zero_in : `in` is taken as 0 (first pre-smooth)          crse : in += crse(parent) on load (addInterpCorrection fused)
//   add_to  : add_to += out on store (MLMG's sol += cor)     boff : scalar subtracted from rhs (lazy makeSolvable)
// ------------------------------------------------------------------------------------------------
#define RB_TX 64
#define RB_TY 16
#define RB_PX (RB_TX + 4)
#define RB_PY (RB_TY + 4)
#define RB_PS (RB_PX * RB_PY)
struct RbArgs {
    double* out; const double* in; const double* rhs; const double* boff; double* add_to; const double* crse;
    int zero_in, zc;
};
__device__ __forceinline__ int rb_wrap(int g, int n, int periodic) { return (g >= 0 && g < n) ? g : (periodic ? (g % n + n) % n : -1); }

template <bool CB>
__device__ __forceinline__ double rb_update(const MacLevel& L, const MacBeta& B, const double* ring, int q, int lx, int ly,
                                            int gx, int gy, int gz, const double* __restrict__ rhs, double off)
{
    const double* P = ring + (q & 3) * RB_PS;
    const int o = lx + RB_PX * ly;
    const double xc = P[o];
    const bool px = L.bclo[0] == MAC_PER, py = L.bclo[1] == MAC_PER, pz = L.bclo[2] == MAC_PER;
    const double xm = (gx > 0 || px) ? P[o - 1] : (L.bclo[0] == MAC_NEU ? xc : L.dc0[0] * xc + L.dc1[0] * P[o + 1]);
    const double xp = (gx < L.n[0] - 1 || px) ? P[o + 1] : (L.bchi[0] == MAC_NEU ? xc : L.dc0[0] * xc + L.dc1[0] * P[o - 1]);
    const double ym = (gy > 0 || py) ? P[o - RB_PX] : (L.bclo[1] == MAC_NEU ? xc : L.dc0[1] * xc + L.dc1[1] * P[o + RB_PX]);
    const double yp = (gy < L.n[1] - 1 || py) ? P[o + RB_PX] : (L.bchi[1] == MAC_NEU ? xc : L.dc0[1] * xc + L.dc1[1] * P[o - RB_PX]);
    const double* Pm = ring + ((q - 1) & 3) * RB_PS; const double* Pp = ring + ((q + 1) & 3) * RB_PS;
    const double zm = (gz > 0 || pz) ? Pm[o] : (L.bclo[2] == MAC_NEU ? xc : L.dc0[2] * xc + L.dc1[2] * Pp[o]);
    const double zp = (gz < L.n[2] - 1 || pz) ? Pp[o] : (L.bchi[2] == MAC_NEU ? xc : L.dc0[2] * xc + L.dc1[2] * Pm[o]);
    const MacB6 b = mac_load_b<CB>(L, B, gx, gy, gz);
    const double rho = L.dh[0] * (b.xm * xm + b.xp * xp) + L.dh[1] * (b.ym * ym + b.yp * yp) + L.dh[2] * (b.zm * zm + b.zp * zp);
    const double cf0 = gx == 0 ? L.flo[0] : 0.0, cf3 = gx == L.n[0] - 1 ? L.fhi[0] : 0.0;
    const double cf1 = gy == 0 ? L.flo[1] : 0.0, cf4 = gy == L.n[1] - 1 ? L.fhi[1] : 0.0;
    const double cf2 = gz == 0 ? L.flo[2] : 0.0, cf5 = gz == L.n[2] - 1 ? L.fhi[2] : 0.0;
    const double gamma = L.dh[0] * (b.xm + b.xp) + L.dh[1] * (b.ym + b.yp) + L.dh[2] * (b.zm + b.zp);
    const double g_m_d = gamma - (L.dh[0] * (b.xm * cf0 + b.xp * cf3) + L.dh[1] * (b.ym * cf1 + b.yp * cf4) + L.dh[2] * (b.zm * cf2 + b.zp * cf5));
    const double r = rhs[mac_c(L, gx, gy, gz)] - off;
    const double res = r - (gamma * xc - rho);
    return xc + 1.15 / g_m_d * res;
}

template <bool CB>
__global__ void __launch_bounds__(256) k_mac_rb2(MacLevel L, MacLevel Lc, MacBeta B, RbArgs A)
{
    __shared__ double ring[4 * RB_PS];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * RB_TX, y0 = blockIdx.y * RB_TY, z0 = blockIdx.z * A.zc, z1 = min(z0 + A.zc, L.n[2]);
    const int px = L.bclo[0] == MAC_PER, py = L.bclo[1] == MAC_PER, pz = L.bclo[2] == MAC_PER;
    const double off = A.boff ? *A.boff : 0.0;
    for (int p = z0 - 2; p <= z1 + 1; ++p) {
        // ---- load plane p (old values; periodic images through wrapped indices; cells outside the domain are never read) ----
        {
            const int gz = rb_wrap(p, L.n[2], pz);
            double* P = ring + (p & 3) * RB_PS;
            if (gz >= 0) {
                for (int q = tid; q < RB_PS; q += 256) {
                    const int lx = q % RB_PX, ly = q / RB_PX;
                    const int gx = rb_wrap(x0 - 2 + lx, L.n[0], px), gy = rb_wrap(y0 - 2 + ly, L.n[1], py);
                    double v = 0.0;
                    if (gx >= 0 && gy >= 0) {
                        if (!A.zero_in) v = A.in[mac_c(L, gx, gy, gz)];
                        if (A.crse) v += A.crse[mac_c(Lc, gx >> 1, gy >> 1, gz >> 1)];
                    }
                    P[q] = v;
                }
            }
        }
        __syncthreads();
        // ---- red cells of plane p-1 over tile + 1 ----
        {
            const int qz = p - 1;
            const int gz = rb_wrap(qz, L.n[2], pz);
            if (gz >= 0 && qz >= z0 - 1 && qz <= z1) {
                double* P = ring + (qz & 3) * RB_PS;
                const int npair = (RB_TX + 2) / 2;                      // 33 pair slots per row, rows 1 .. RB_TY+2
                for (int q = tid; q < npair * (RB_TY + 2); q += 256) {
                    const int ly = 1 + q / npair, pr = q % npair;
                    const int gy = rb_wrap(y0 - 2 + ly, L.n[1], py);
                    if (gy < 0) continue;
                    int lx = 1 + 2 * pr;
                    int gx = rb_wrap(x0 - 2 + lx, L.n[0], px);
                    // the red cell of the pair (lx, lx+1): parity of the wrapped global index (n is even in periodic directions)
                    const int gxa = x0 - 2 + lx;                         // unwrapped parity is the same
                    if ((gxa + (y0 - 2 + ly) + qz) & 1) { ++lx; gx = rb_wrap(x0 - 2 + lx, L.n[0], px); }
                    if (lx > RB_TX + 2 || gx < 0) continue;
                    P[lx + RB_PX * ly] = rb_update<CB>(L, B, ring, qz, lx, ly, gx, gy, gz, A.rhs, off);
                }
            }
        }
        __syncthreads();
        // ---- black cells of plane p-2 over the tile, stored with their red partners ----
        {
            const int r = p - 2;
            if (r >= z0 && r < z1) {
                const double* P = ring + (r & 3) * RB_PS;
                for (int q = tid; q < (RB_TX / 2) * RB_TY; q += 256) {
                    const int ly = 2 + q / (RB_TX / 2), lx0 = 2 + 2 * (q % (RB_TX / 2));
                    const int gx0 = x0 - 2 + lx0, gy = y0 - 2 + ly;
                    if (gx0 >= L.n[0] || gy >= L.n[1]) continue;
                    const int a = ((gx0 + gy + r) & 1) ? 0 : 1;          // which cell of the pair is black (odd parity)
                    double v0 = P[lx0 + RB_PX * ly], v1 = P[lx0 + 1 + RB_PX * ly];
                    if (gx0 + a < L.n[0]) {
                        const double nb = rb_update<CB>(L, B, ring, r, lx0 + a, ly, gx0 + a, gy, r, A.rhs, off);
                        if (a) v1 = nb; else v0 = nb;
                    }
                    const long long c = mac_c(L, gx0, gy, r);
                    if (gx0 + 1 < L.n[0]) {
                        *reinterpret_cast<double2*>(A.out + c) = make_double2(v0, v1);
                        if (A.add_to) { double2 t = *reinterpret_cast<double2*>(A.add_to + c); t.x += v0; t.y += v1; *reinterpret_cast<double2*>(A.add_to + c) = t; }
                    } else {
                        A.out[c] = v0;
                        if (A.add_to) A.add_to[c] += v0;
                    }
                }
            }
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------
// k_mac_rb3: the same one-pass smooth as k_mac_rb2 for levels whose x / y extents are multiples of the tile (every large
// level of a 2^n grid), with all index work hoisted out of the plane loop.  A thread owns two aligned cell pairs of the tile
// (rows ty and ty+8), at most two cells of the halo-2 frame to load and at most one cell of the tile+1 ring to relax; their
// shared-memory offsets, wrapped in-plane global offsets and boundary flags are computed once.  Per plane a thread issues
// two 16-byte loads + <= 2 scalar loads, <= 3 red and 2 black updates (a branch-free body when no domain face is near)
// and two 16-byte stores.
// ------------------------------------------------------------------------------------------------
enum { RBF_XLO = 1, RBF_XHI = 2, RBF_YLO = 4, RBF_YHI = 8, RBF_ZLO = 16, RBF_ZHI = 32 };
struct RbCell { int so; int go; int gfx; int flags; int dxm; };   // smem offset, in-plane cell offset, in-plane bx offset, RBF_*, offset of the x-lo neighbour
// k_mac_rb3's ring planes keep the even and the odd columns of a row in separate halves, so the cells of one colour in a row
// (and their x neighbours, which all have the other parity) are contiguous: no shared-memory bank conflicts in the sweeps
#define RB3_SO(lx, ly) (((lx) >> 1) + ((lx) & 1) * (RB_PX / 2) + RB_PX * (ly))
#define RB3_DXM(lx) (((lx) & 1) ? -(RB_PX / 2) : (RB_PX / 2 - 1))

// HA: the operator has MLABecLaplacian's alpha * a term (the diffusion solves): one more cell value in the diagonal
template <bool CB, bool HA>
__device__ __forceinline__ double rb3_update(const MacLevel& L, const MacBeta& B, const double* __restrict__ P, const double* __restrict__ Pm,
                                             const double* __restrict__ Pp, const RbCell& cl, int zflags, long long zc, long long zfx, long long zfy,
                                             double rhs_c, double off)
{
    const int o = cl.so;
    const double xc = P[o];
    double at = 0.0;
    if (HA) at = L.ascal * L.acoef[zc + cl.go];
    double bxm, bxp, bym, byp, bzm, bzp;
    if (CB) { bxm = bxp = bym = byp = bzm = bzp = L.beta0; }
    else {
        const double* px = B.bx + zfx + cl.gfx; bxm = px[0]; bxp = px[1];
        const double* py = B.by + zfy + cl.go;  bym = py[0]; byp = py[L.sx];
        const double* pz = B.bz + zc + cl.go;   bzm = pz[0]; bzp = pz[L.sxy];
    }
    const double gamma = HA ? at + L.dh[0] * (bxm + bxp) + L.dh[1] * (bym + byp) + L.dh[2] * (bzm + bzp)
                            : L.dh[0] * (bxm + bxp) + L.dh[1] * (bym + byp) + L.dh[2] * (bzm + bzp);
    const double r = rhs_c - off;
    const int fl = cl.flags | zflags;
    double xm = P[o + cl.dxm], xp = P[o + cl.dxm + 1], ym = P[o - RB_PX], yp = P[o + RB_PX], zm = Pm[o], zp = Pp[o];
    const double xm0 = xm;
    double g_m_d = gamma;
    if (fl) {       // a Neumann / Dirichlet face next to the cell: ghost values from the boundary condition
        double delta = 0.0;
        if (fl & RBF_XLO) { xm = L.bclo[0] == MAC_NEU ? xc : L.dc0[0] * xc + L.dc1[0] * xp; delta += L.dh[0] * bxm * L.flo[0]; }
        if (fl & RBF_XHI) { xp = L.bchi[0] == MAC_NEU ? xc : L.dc0[0] * xc + L.dc1[0] * xm0; delta += L.dh[0] * bxp * L.fhi[0]; }
        if (fl & RBF_YLO) { ym = L.bclo[1] == MAC_NEU ? xc : L.dc0[1] * xc + L.dc1[1] * yp; delta += L.dh[1] * bym * L.flo[1]; }
        if (fl & RBF_YHI) { yp = L.bchi[1] == MAC_NEU ? xc : L.dc0[1] * xc + L.dc1[1] * P[o - RB_PX]; delta += L.dh[1] * byp * L.fhi[1]; }
        if (fl & RBF_ZLO) { zm = L.bclo[2] == MAC_NEU ? xc : L.dc0[2] * xc + L.dc1[2] * zp; delta += L.dh[2] * bzm * L.flo[2]; }
        if (fl & RBF_ZHI) { zp = L.bchi[2] == MAC_NEU ? xc : L.dc0[2] * xc + L.dc1[2] * Pm[o]; delta += L.dh[2] * bzp * L.fhi[2]; }
        g_m_d = gamma - delta;
    }
    const double rho = L.dh[0] * (bxm * xm + bxp * xp) + L.dh[1] * (bym * ym + byp * yp) + L.dh[2] * (bzm * zm + bzp * zp);
    const double res = r - (gamma * xc - rho);
    return xc + 1.15 / g_m_d * res;
}

__device__ __forceinline__ int rb3_flags(const MacLevel& L, int gx, int gy)
{
    int f = 0;
    if (L.bclo[0] != MAC_PER) { if (gx == 0) f |= RBF_XLO; if (gx == L.n[0] - 1) f |= RBF_XHI; }
    if (L.bclo[1] != MAC_PER) { if (gy == 0) f |= RBF_YLO; if (gy == L.n[1] - 1) f |= RBF_YHI; }
    return f;
}

// plane index of the ring's plane p in the arrays of k_mac_rb3: inside the slab p itself; across a periodic face the wrapped
// plane; across a face shared with the neighbouring rank one of the two ghost planes the halo exchange filled (p = -2, -1, n2,
// n2+1: plain addresses below / above the slab); RB_NOPLANE outside a physical face
#define RB_NOPLANE (-(1 << 30))
__device__ __forceinline__ int rb3_zmap(const MacLevel& L, int p)
{
    const int n = L.n[2];
    if (p >= 0 && p < n) return p;
    const int bc = p < 0 ? L.bclo[2] : L.bchi[2];
    if (bc == MAC_REMOTE) return (p >= -2 && p < n + 2) ? p : RB_NOPLANE;
    if (bc == MAC_PER) return (p % n + n) % n;
    return RB_NOPLANE;
}

template <bool CB, bool HA = false>
__global__ void __launch_bounds__(256) k_mac_rb3(MacLevel L, MacLevel Lc, MacBeta B, RbArgs A)
{
    __shared__ __align__(16) double ring[4 * RB_PS];
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int x0 = blockIdx.x * RB_TX, y0 = blockIdx.y * RB_TY, z0 = blockIdx.z * A.zc, z1 = min(z0 + A.zc, L.n[2]);
    const int px = L.bclo[0] == MAC_PER, py = L.bclo[1] == MAC_PER;
    const bool zlo_phys = L.bclo[2] == MAC_NEU || L.bclo[2] == MAC_DIR, zhi_phys = L.bchi[2] == MAC_NEU || L.bchi[2] == MAC_DIR;
    const double off = A.boff ? *A.boff : 0.0;
    const int sx = (int)L.sx, fxs = (int)L.fxs, sxc = (int)Lc.sx;
    // ---- the thread's fixed cells ----
    RbCell own0, own1;          // first cell of the thread's two pairs (the second cell is +1 in every offset)
    int of0[2], of1[2];         // boundary flags of the two cells of each pair
    int par0, par1, pco0, pco1; // parity of (gx0 + gy); in-plane offset of the pair's coarse parent
    {
        const int lx0 = 2 + 2 * tx, gx0 = x0 + 2 * tx;
        const int ly0 = 2 + ty, gy0 = y0 + ty, ly1 = ly0 + 8, gy1 = gy0 + 8;
        par0 = (gx0 + gy0 + L.koff) & 1; par1 = (gx0 + gy1 + L.koff) & 1;
        pco0 = (gx0 >> 1) + sxc * (gy0 >> 1); pco1 = (gx0 >> 1) + sxc * (gy1 >> 1);
        own0.so = RB3_SO(lx0, ly0); own0.dxm = RB3_DXM(lx0); own0.go = gx0 + sx * gy0; own0.gfx = gx0 + fxs * gy0; own0.flags = 0;
        own1.so = RB3_SO(lx0, ly1); own1.dxm = RB3_DXM(lx0); own1.go = gx0 + sx * gy1; own1.gfx = gx0 + fxs * gy1; own1.flags = 0;
        of0[0] = rb3_flags(L, gx0, gy0); of0[1] = rb3_flags(L, gx0 + 1, gy0);
        of1[0] = rb3_flags(L, gx0, gy1); of1[1] = rb3_flags(L, gx0 + 1, gy1);
    }
    auto cell_of = [&](int m, int a) {
        RbCell c = m ? own1 : own0;
        c.so += a * (RB_PX / 2); c.go += a; c.gfx += a; c.dxm = a ? -(RB_PX / 2) : (RB_PX / 2 - 1);
        c.flags = m ? (a ? of1[1] : of1[0]) : (a ? of0[1] : of0[0]);
        return c;
    };
    // halo-2 frame: rows 0,1 and 18,19 (68 wide), columns 0,1,66,67 of rows 2..17: 336 cells, threads take tid and tid + 256
    int fso[2], fgo[2], fpc[2];
    for (int e = 0; e < 2; ++e) {
        const int f = tid + 256 * e;
        fso[e] = -1; fgo[e] = -1; fpc[e] = 0;
        if (f < 336) {
            int lx, ly;
            if (f < 272) { const int rr = f / RB_PX; lx = f - rr * RB_PX; ly = rr < 2 ? rr : RB_PY - 4 + rr; }
            else { const int g = f - 272; ly = 2 + (g >> 2); const int cidx = g & 3; lx = cidx < 2 ? cidx : RB_PX - 4 + cidx; }
            const int gx = rb_wrap(x0 - 2 + lx, L.n[0], px), gy = rb_wrap(y0 - 2 + ly, L.n[1], py);
            fso[e] = RB3_SO(lx, ly);
            if (gx >= 0 && gy >= 0) { fgo[e] = gx + sx * gy; fpc[e] = (gx >> 1) + sxc * (gy >> 1); }
        }
    }
    // tile+1 ring: rows 1 and 18 (lx 1..66), columns 1 and 66 of rows 2..17: 164 cells, one per thread
    RbCell rc; rc.so = -1; rc.go = -1; rc.gfx = 0; rc.flags = 0; rc.dxm = 0;
    int rpar = 0;
    if (tid < 164) {
        int lx, ly;
        if (tid < 132) { const int rr = tid / 66; lx = 1 + tid - rr * 66; ly = rr ? RB_PY - 2 : 1; }
        else { const int g = tid - 132; ly = 2 + (g >> 1); lx = (g & 1) ? RB_PX - 2 : 1; }
        const int gxu = x0 - 2 + lx, gyu = y0 - 2 + ly;
        const int gx = rb_wrap(gxu, L.n[0], px), gy = rb_wrap(gyu, L.n[1], py);
        rpar = (gxu + gyu + L.koff) & 1;
        if (gx >= 0 && gy >= 0) { rc.so = RB3_SO(lx, ly); rc.dxm = RB3_DXM(lx); rc.go = gx + sx * gy; rc.gfx = gx + fxs * gy; rc.flags = rb3_flags(L, gx, gy); }
    }
    const long long fxz = L.fxs * (long long)L.n[1], fyz = L.sx * (long long)(L.n[1] + 1);

    // plane loads run one step ahead: the values of plane p+1 travel in registers while plane p is relaxed
    double2 pre_own[2]; double pre_fr[2];
    auto fetch_plane = [&](int p) {
        const int gz = rb3_zmap(L, p);
        pre_own[0] = pre_own[1] = make_double2(0.0, 0.0); pre_fr[0] = pre_fr[1] = 0.0;
        if (gz == RB_NOPLANE || p > z1 + 1) return;
        const long long zc = L.sxy * (long long)gz, zcc = Lc.sxy * (long long)(gz >> 1);
#pragma unroll
        for (int m = 0; m < 2; ++m) {
            double2 v = make_double2(0.0, 0.0);
            if (!A.zero_in) v = *reinterpret_cast<const double2*>(A.in + zc + (m ? own1.go : own0.go));
            if (A.crse) { const double cv = A.crse[zcc + (m ? pco1 : pco0)]; v.x += cv; v.y += cv; }
            pre_own[m] = v;
        }
        for (int e = 0; e < 2; ++e) if (fso[e] >= 0 && fgo[e] >= 0) {
            double v = 0.0;
            if (!A.zero_in) v = A.in[zc + fgo[e]];
            if (A.crse) v += A.crse[zcc + fpc[e]];
            pre_fr[e] = v;
        }
    };
    // so do the right-hand sides of the cells relaxed in the next step: reds of plane p (own pairs + ring), blacks of plane p-1
    double rr0 = 0.0, rr1 = 0.0, rrr = 0.0, rk0 = 0.0, rk1 = 0.0;
    auto fetch_rhs = [&](int p) {           // for the step that will load plane p: red plane p-1, black plane p-2
        if (!CB) return;                    // variable coefficients: the registers are needed elsewhere (measured slower)
        const int q = p - 1, r = p - 2;
        const int gzq = rb3_zmap(L, q);
        if (gzq != RB_NOPLANE && q >= z0 - 1 && q <= z1) {
            const long long zc = L.sxy * (long long)gzq;
            rr0 = A.rhs[zc + own0.go + ((par0 + q) & 1)]; rr1 = A.rhs[zc + own1.go + ((par1 + q) & 1)];
            if (rc.so >= 0 && (((rpar + q) & 1) == 0)) rrr = A.rhs[zc + rc.go];
        }
        if (r >= z0 && r < z1) {
            const long long zc = L.sxy * (long long)r;
            rk0 = A.rhs[zc + own0.go + 1 - ((par0 + r) & 1)]; rk1 = A.rhs[zc + own1.go + 1 - ((par1 + r) & 1)];
        }
    };
    fetch_plane(z0 - 2);
    for (int p = z0 - 2; p <= z1 + 1; ++p) {
        const double cr0 = rr0, cr1 = rr1, crr = rrr, ck0 = rk0, ck1 = rk1;      // this step's right-hand sides
        // ---- plane p: registers -> ring; start the loads of plane p+1 ----
        {
            double* P = ring + (p & 3) * RB_PS;
            P[own0.so] = pre_own[0].x; P[own0.so + RB_PX / 2] = pre_own[0].y;
            P[own1.so] = pre_own[1].x; P[own1.so + RB_PX / 2] = pre_own[1].y;
            for (int e = 0; e < 2; ++e) if (fso[e] >= 0) P[fso[e]] = pre_fr[e];
            fetch_plane(p + 1);
            fetch_rhs(p + 1);
            if (HA) {       // the a values the next step's red sweep (plane p) and, one step later, its black sweep read: into L1 now
                const int gza = rb3_zmap(L, p);
                if (gza != RB_NOPLANE && p <= z1) { const long long za = L.sxy * (long long)gza; np_prefetch_l1(L.acoef + za + own0.go); np_prefetch_l1(L.acoef + za + own1.go); }
            }
            if (!CB) {      // variable coefficients: pull the lines the next step's red sweep (plane p) reads into L1 now; the black
                            // sweep of that plane, one step later, finds them there
                const int gzn = rb3_zmap(L, p);
                if (gzn != RB_NOPLANE && p <= z1) {
                    const long long zc = L.sxy * (long long)gzn, zfx = fxz * gzn, zfy = fyz * gzn;
                    np_prefetch_l1(B.bx + zfx + own0.gfx); np_prefetch_l1(B.bx + zfx + own1.gfx);
                    np_prefetch_l1(B.by + zfy + own0.go); np_prefetch_l1(B.by + zfy + own0.go + L.sx);
                    np_prefetch_l1(B.by + zfy + own1.go); np_prefetch_l1(B.by + zfy + own1.go + L.sx);
                    np_prefetch_l1(B.bz + zc + L.sxy + own0.go); np_prefetch_l1(B.bz + zc + L.sxy + own1.go);
                    np_prefetch_l1(A.rhs + zc + own0.go); np_prefetch_l1(A.rhs + zc + own1.go);
                }
            }
        }
        __syncthreads();
        // ---- red cells of plane p-1: own pairs + ring ----
        {
            const int q = p - 1;
            const int gz = rb3_zmap(L, q);
            if (gz != RB_NOPLANE && q >= z0 - 1 && q <= z1) {
                double* P = ring + (q & 3) * RB_PS;
                const double* Pm = ring + ((q - 1) & 3) * RB_PS; const double* Pp = ring + ((q + 1) & 3) * RB_PS;
                const int zfl = ((zlo_phys && gz == 0) ? RBF_ZLO : 0) | ((zhi_phys && gz == L.n[2] - 1) ? RBF_ZHI : 0);
                const long long zc = L.sxy * (long long)gz, zfx = fxz * gz, zfy = fyz * gz;
                // reds only read blacks (and their own value), so the in-place writes of other threads cannot be seen here
                const RbCell c0 = cell_of(0, (par0 + q) & 1), c1 = cell_of(1, (par1 + q) & 1);
                const double n0 = rb3_update<CB, HA>(L, B, P, Pm, Pp, c0, zfl, zc, zfx, zfy, CB ? cr0 : A.rhs[zc + c0.go], off);
                const double n1 = rb3_update<CB, HA>(L, B, P, Pm, Pp, c1, zfl, zc, zfx, zfy, CB ? cr1 : A.rhs[zc + c1.go], off);
                P[c0.so] = n0; P[c1.so] = n1;
                if (rc.so >= 0 && (((rpar + q) & 1) == 0)) P[rc.so] = rb3_update<CB, HA>(L, B, P, Pm, Pp, rc, zfl, zc, zfx, zfy, CB ? crr : A.rhs[zc + rc.go], off);
            }
        }
        __syncthreads();
        // ---- black cells of plane p-2, stored with their red partners ----
        {
            const int r = p - 2;
            if (r >= z0 && r < z1) {
                const double* P = ring + (r & 3) * RB_PS;
                const double* Pm = ring + ((r - 1) & 3) * RB_PS; const double* Pp = ring + ((r + 1) & 3) * RB_PS;
                const int zfl = ((zlo_phys && r == 0) ? RBF_ZLO : 0) | ((zhi_phys && r == L.n[2] - 1) ? RBF_ZHI : 0);
                const long long zc = L.sxy * (long long)r, zfx = fxz * r, zfy = fyz * r;
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    const int a = 1 - (((m ? par1 : par0) + r) & 1);
                    const int so = m ? own1.so : own0.so;
                    double2 v = make_double2(P[so], P[so + RB_PX / 2]);
                    const RbCell cb = cell_of(m, a);
                    const double nb = rb3_update<CB, HA>(L, B, P, Pm, Pp, cb, zfl, zc, zfx, zfy, CB ? (m ? ck1 : ck0) : A.rhs[zc + cb.go], off);
                    if (a) v.y = nb; else v.x = nb;
                    const long long c = zc + (m ? own1.go : own0.go);
                    *reinterpret_cast<double2*>(A.out + c) = v;
                    if (A.add_to) { double2 t = *reinterpret_cast<double2*>(A.add_to + c); t.x += v.x; t.y += v.y; *reinterpret_cast<double2*>(A.add_to + c) = t; }
                }
            }
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------
// MLABecLaplacian::Fapply (mlabeclap_adotx) and the residuals built on it:
//   MODE 0: y = A x            MODE 1: y = (b - off) - A x  (MLMG::computeResidual / computeResOfCorrection)
// REDUCE: per-block partials of max|y| and sum(y) (inf-norm for the convergence test, mean for makeSolvable).
// ------------------------------------------------------------------------------------------------
#define MAC_KZ 8     // planes one block of k_mac_adotx marches through
template <int MODE, bool REDUCE, bool CB>
__global__ void __launch_bounds__(256) k_mac_adotx(MacLevel L, MacBeta B, double* __restrict__ y, const double* __restrict__ x,
                                                   const double* __restrict__ bb, const double* __restrict__ boff,
                                                   double* __restrict__ part_max, double* __restrict__ part_sum)
{
    // a block marches through MAC_KZ planes: x(k-1), x(k), x(k+1) of a column roll through registers, so every plane is
    // loaded once per block (not three times), and the block-wide reduction is paid once per MAC_KZ cells of a thread.  A
    // thread owns an aligned pair of cells: every access away from a domain face is a 16-byte one.
    const int i0 = 2 * (int)(blockIdx.x * blockDim.x + threadIdx.x);
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k0 = blockIdx.z * MAC_KZ, k1 = min(k0 + MAC_KZ, L.n[2]);
    double vmax = 0.0, vsum = 0.0;
    const bool on = i0 < L.n[0] && j < L.n[1];
    if (on) {
        const bool pair = i0 + 1 < L.n[0];
        const double off = (MODE == 1 && boff) ? *boff : 0.0;
        long long c = mac_c(L, i0, j, k0);
        auto ld2 = [&](const double* p) { return pair ? *reinterpret_cast<const double2*>(p) : make_double2(p[0], 0.0); };
        // both cells' neighbours across the low / high side of direction d (1: y, 2: z): one 16-byte load inside, the boundary
        // expression cell by cell on a domain face
        auto nb2 = [&](const double2& xc, long long cc, int d, int idx, long long st, bool hi) {
            const bool inside = hi ? (idx < L.n[d] - 1 || L.bchi[d] == MAC_REMOTE) : (idx > 0 || L.bclo[d] == MAC_REMOTE);
            if (inside) return ld2(x + (hi ? cc + st : cc - st));
            double2 r;
            r.x = hi ? mac_nb_hi(L, x, cc, xc.x, d, idx, st, x) : mac_nb_lo(L, x, cc, xc.x, d, idx, st, x);
            r.y = pair ? (hi ? mac_nb_hi(L, x, cc + 1, xc.y, d, idx, st, x) : mac_nb_lo(L, x, cc + 1, xc.y, d, idx, st, x)) : 0.0;
            return r;
        };
        double2 xc = ld2(x + c);
        double2 zm = nb2(xc, c, 2, k0, L.sxy, false);
        for (int k = k0; k < k1; ++k, c += L.sxy) {
            const double2 zp = nb2(xc, c, 2, k, L.sxy, true);
            const double2 ym = nb2(xc, c, 1, j, L.sx, false), yp = nb2(xc, c, 1, j, L.sx, true);
            const double xm0 = mac_nb_lo(L, x, c, xc.x, 0, i0, 1, x);
            const double xp1 = pair ? mac_nb_hi(L, x, c + 1, xc.y, 0, i0 + 1, 1, x) : 0.0;
            const double xp0 = pair ? xc.y : mac_nb_hi(L, x, c, xc.x, 0, i0, 1, x);
            double2 bv = make_double2(0.0, 0.0);
            if (MODE == 1) bv = ld2(bb + c);
            double2 out;
            MacB6 b0, b1;
            if (CB || !pair) { b0 = mac_load_b<CB>(L, B, i0, j, k); b1 = b0; }
            else {      // the coefficients of both cells with 16-byte loads: three x faces, two rows of y faces, two planes of z faces
                const long long fx = mac_fx(L, i0, j, k), fy = mac_fy(L, i0, j, k), fz = mac_fz(L, i0, j, k);
                const double2 bx01 = *reinterpret_cast<const double2*>(B.bx + fx); const double bx2 = B.bx[fx + 2];
                const double2 bym = *reinterpret_cast<const double2*>(B.by + fy), byp = *reinterpret_cast<const double2*>(B.by + fy + L.sx);
                const double2 bzm = *reinterpret_cast<const double2*>(B.bz + fz), bzp = *reinterpret_cast<const double2*>(B.bz + fz + L.sxy);
                b0.xm = bx01.x; b0.xp = bx01.y; b0.ym = bym.x; b0.yp = byp.x; b0.zm = bzm.x; b0.zp = bzp.x;
                b1.xm = bx01.y; b1.xp = bx2;    b1.ym = bym.y; b1.yp = byp.y; b1.zm = bzm.y; b1.zp = bzp.y;
            }
            {
                const MacB6& b = b0;
                const double ax = mac_aterm(L, c) * xc.x - L.dh[0] * (b.xp * (xp0 - xc.x) - b.xm * (xc.x - xm0)) - L.dh[1] * (b.yp * (yp.x - xc.x) - b.ym * (xc.x - ym.x))
                                - L.dh[2] * (b.zp * (zp.x - xc.x) - b.zm * (xc.x - zm.x));
                out.x = MODE == 0 ? ax : (bv.x - off) - ax;
            }
            out.y = 0.0;
            if (pair) {
                const MacB6& b = b1;
                const double ax = mac_aterm(L, c + 1) * xc.y - L.dh[0] * (b.xp * (xp1 - xc.y) - b.xm * (xc.y - xc.x)) - L.dh[1] * (b.yp * (yp.y - xc.y) - b.ym * (xc.y - ym.y))
                                - L.dh[2] * (b.zp * (zp.y - xc.y) - b.zm * (xc.y - zm.y));
                out.y = MODE == 0 ? ax : (bv.y - off) - ax;
                *reinterpret_cast<double2*>(y + c) = out;
            } else y[c] = out.x;
            if (REDUCE) { vmax = np_absmax(np_absmax(vmax, out.x), out.y); vsum += out.x + out.y; }
            // roll (inside the chunk zp is a plain interior value; the boundary expression only occurs on the last plane)
            zm = xc; xc = zp;
        }
    }
    if (REDUCE) {
        const double m = np_block_reduce<true>(vmax);
        const double sm = np_block_reduce<false>(vsum);
        if (threadIdx.x == 0 && threadIdx.y == 0) {
            const long long q = blockIdx.x + gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z);
            part_max[q] = m; part_sum[q] = sm;
        }
    }
}
// MLMG::computeResOfCorrection + MLCellLinOp::restriction in one pass: crse = average_down(b - off - A x).  The residual of
// the correction is only ever restricted, so it is never stored: a thread owns one coarse cell = 2x2x2 fine cells.
template <bool CB>
__global__ void __launch_bounds__(256) k_mac_resid_restrict(MacLevel Lc, MacLevel L, MacBeta B, double* __restrict__ crse, const double* __restrict__ x,
                                                            const double* __restrict__ bb, const double* __restrict__ boff)
{
    const int ic = blockIdx.x * blockDim.x + threadIdx.x;
    const int jc = blockIdx.y * blockDim.y + threadIdx.y;
    const int kc = blockIdx.z;
    if (ic >= Lc.n[0] || jc >= Lc.n[1]) return;
    const double off = boff ? *boff : 0.0;
    double acc = 0.0;
#pragma unroll
    for (int dk = 0; dk < 2; ++dk)
#pragma unroll
        for (int dj = 0; dj < 2; ++dj) {
            const int i = 2 * ic, j = 2 * jc + dj, k = 2 * kc + dk;
            const long long c = mac_c(L, i, j, k);
            const double2 xv = *reinterpret_cast<const double2*>(x + c);
            const double2 bv = *reinterpret_cast<const double2*>(bb + c);
            MacB6 bp[2];
            if (CB) { bp[0] = mac_load_b<CB>(L, B, i, j, k); bp[1] = bp[0]; }
            else {      // the coefficients of the pair with 16-byte loads
                const long long fx = mac_fx(L, i, j, k), fy = mac_fy(L, i, j, k), fz = mac_fz(L, i, j, k);
                const double2 bx01 = *reinterpret_cast<const double2*>(B.bx + fx); const double bx2 = B.bx[fx + 2];
                const double2 bym = *reinterpret_cast<const double2*>(B.by + fy), byp = *reinterpret_cast<const double2*>(B.by + fy + L.sx);
                const double2 bzm = *reinterpret_cast<const double2*>(B.bz + fz), bzp = *reinterpret_cast<const double2*>(B.bz + fz + L.sxy);
                bp[0].xm = bx01.x; bp[0].xp = bx01.y; bp[0].ym = bym.x; bp[0].yp = byp.x; bp[0].zm = bzm.x; bp[0].zp = bzp.x;
                bp[1].xm = bx01.y; bp[1].xp = bx2;    bp[1].ym = bym.y; bp[1].yp = byp.y; bp[1].zm = bzm.y; bp[1].zp = bzp.y;
            }
#pragma unroll
            for (int di = 0; di < 2; ++di) {
                const double xc = di ? xv.y : xv.x;
                const long long cc = c + di;
                const MacB6 b = bp[di];
                const double xm = di ? xv.x : mac_nb_lo(L, x, cc, xc, 0, i, 1, x);
                const double xp = di ? mac_nb_hi(L, x, cc, xc, 0, i + 1, 1, x) : xv.y;
                const double ym = mac_nb_lo(L, x, cc, xc, 1, j, L.sx, x), yp = mac_nb_hi(L, x, cc, xc, 1, j, L.sx, x);
                const double zm = mac_nb_lo(L, x, cc, xc, 2, k, L.sxy, x), zp = mac_nb_hi(L, x, cc, xc, 2, k, L.sxy, x);
                const double ax = mac_aterm(L, cc) * xc - L.dh[0] * (b.xp * (xp - xc) - b.xm * (xc - xm)) - L.dh[1] * (b.yp * (yp - xc) - b.ym * (xc - ym))
                                - L.dh[2] * (b.zp * (zp - xc) - b.zm * (xc - zm));
                acc += ((di ? bv.y : bv.x) - off) - ax;
            }
        }
    crse[mac_c(Lc, ic, jc, kc)] = 0.125 * acc;
}
// one block folds the partials: scal[0] = max, scal[2] = sum, scal[6] = sum / ncells (the solvability offset)
__global__ void __launch_bounds__(1024) k_mac_reduce(const double* __restrict__ part_max, const double* __restrict__ part_sum, long long n,
                                                     double* scal, double ncells)
{
    double m = 0.0, s = 0.0;
    for (long long t = threadIdx.x; t < n; t += blockDim.x) { if (part_max) m = fmax(m, part_max[t]); if (part_sum) s += part_sum[t]; }
    m = np_block_reduce<true>(m);
    s = np_block_reduce<false>(s);
    if (threadIdx.x == 0) { if (part_max) scal[0] = m; if (part_sum) { scal[2] = s; scal[6] = s / ncells; } }
}

// x -= *off (makeSolvable of the right-hand side, once per projection) with the inf-norm partials of the result
__global__ void __launch_bounds__(256) k_mac_sub_norm(MacLevel L, double* __restrict__ x, const double* __restrict__ off, double* __restrict__ part_max)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    double v = 0.0;
    if (i < L.n[0] && j < L.n[1]) { const long long c = mac_c(L, i, j, k); v = x[c] - *off; x[c] = v; }
    const double m = np_block_reduce<true>(fabs(v));
    if (threadIdx.x == 0 && threadIdx.y == 0) part_max[blockIdx.x + gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z)] = m;
}
// y += x over the cells of a level (MLMG's sol += cor when no post-smooth can carry it)
__global__ void __launch_bounds__(256) k_mac_axpy_cell(MacLevel L, double* __restrict__ y, const double* __restrict__ x)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i < L.n[0] && j < L.n[1]) { const long long c = mac_c(L, i, j, k); y[c] += x[c]; }
}

// MLCellLinOp::restriction (amrex::average_down): mean of the 8 children; the fine offset (lazy makeSolvable) is not needed
// here because only residuals of corrections are restricted.
__global__ void __launch_bounds__(256) k_mac_restrict(MacLevel Lc, MacLevel Lf, double* __restrict__ crse, const double* __restrict__ fine)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= Lc.n[0] || j >= Lc.n[1]) return;
    const long long f = mac_c(Lf, 2 * i, 2 * j, 2 * k);
    const double2 a = *reinterpret_cast<const double2*>(fine + f), b = *reinterpret_cast<const double2*>(fine + f + Lf.sx);
    const double2 c = *reinterpret_cast<const double2*>(fine + f + Lf.sxy), d = *reinterpret_cast<const double2*>(fine + f + Lf.sxy + Lf.sx);
    crse[mac_c(Lc, i, j, k)] = 0.125 * (a.x + a.y + b.x + b.y + c.x + c.y + d.x + d.y);
}
// MLCellLinOp::interpolation (mlmg_lin_cc_interp_r2): fine += crse(parent), one aligned fine pair per thread
__global__ void __launch_bounds__(256) k_mac_interp_add(MacLevel Lf, MacLevel Lc, double* __restrict__ fine, const double* __restrict__ crse)
{
    const int ic = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (ic >= Lc.n[0] || j >= Lf.n[1]) return;
    const double v = crse[mac_c(Lc, ic, j >> 1, (k >> 1) + Lf.kcoff)];
    double2* p = reinterpret_cast<double2*>(fine + mac_c(Lf, 2 * ic, j, k));
    double2 f = *p; f.x += v; f.y += v; *p = f;
}
// amrex::average_down_faces: mean of the 4 fine faces on a coarse face, direction d
__global__ void __launch_bounds__(256) k_mac_avgdown_face(MacLevel Lc, MacLevel Lf, double* __restrict__ bc, const double* __restrict__ bf, int d)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    const int e0 = Lc.n[0] + (d == 0), e1 = Lc.n[1] + (d == 1);
    if (i >= e0 || j >= e1) return;
    if (d == 0) {
        const long long f = mac_fx(Lf, 2 * i, 2 * j, 2 * k), r = Lf.fxs, p = Lf.fxs * Lf.n[1];
        bc[mac_fx(Lc, i, j, k)] = 0.25 * (bf[f] + bf[f + r] + bf[f + p] + bf[f + p + r]);
    } else if (d == 1) {
        const long long f = mac_fy(Lf, 2 * i, 2 * j, 2 * k), p = Lf.sx * (Lf.n[1] + 1);
        bc[mac_fy(Lc, i, j, k)] = 0.25 * (bf[f] + bf[f + 1] + bf[f + p] + bf[f + p + 1]);
    } else {
        const long long f = mac_fz(Lf, 2 * i, 2 * j, 2 * k);
        bc[mac_fz(Lc, i, j, k)] = 0.25 * (bf[f] + bf[f + 1] + bf[f + Lf.sx] + bf[f + Lf.sx + 1]);
    }
}
// MacProjector::updateBeta: the caller's face fab of direction d -> internal layout, with min / max partials so that a
// uniform coefficient (constant density) is detected without another pass
__global__ void __launch_bounds__(256) k_mac_copy_beta(MacLevel L, double* __restrict__ dst, FabView src, int d, double* __restrict__ pmin, double* __restrict__ pmax)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    const int e0 = L.n[0] + (d == 0), e1 = L.n[1] + (d == 1);
    double lo = 1e300, hi = -1e300;
    if (i < e0 && j < e1) {
        const double v = src.p[np_foff(src, i, j, k, 0)];
        dst[d == 0 ? mac_fx(L, i, j, k) : d == 1 ? mac_fy(L, i, j, k) : mac_fz(L, i, j, k)] = v;
        lo = hi = v;
    }
    const double mx = np_block_reduce<true>(hi), mn = -np_block_reduce<true>(-lo);
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        const long long q = blockIdx.x + gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z);
        pmin[q] = mn; pmax[q] = mx;
    }
}
__global__ void __launch_bounds__(1024) k_mac_reduce_minmax(const double* __restrict__ pmin, const double* __restrict__ pmax, long long n, double* out, int first)
{
    double lo = 1e300, hi = -1e300;
    for (long long t = threadIdx.x; t < n; t += blockDim.x) { lo = fmin(lo, pmin[t]); hi = fmax(hi, pmax[t]); }
    hi = np_block_reduce<true>(hi); lo = -np_block_reduce<true>(-lo);
    if (threadIdx.x == 0) { out[0] = first ? lo : fmin(out[0], lo); out[1] = first ? hi : fmax(out[1], hi); }
}

// MacProjector::project: rhs = -div(umac) (amrex::computeDivergence) with the partial sums of its mean
__global__ void __launch_bounds__(256) k_mac_divu(MacLevel L, double* __restrict__ rhs, FabView ux, FabView uy, FabView uz,
                                                  double* __restrict__ part_max, double* __restrict__ part_sum)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    double out = 0.0;
    if (i < L.n[0] && j < L.n[1]) {
        const double d = L.dxinv[0] * (ux.p[np_foff(ux, i + 1, j, k, 0)] - ux.p[np_foff(ux, i, j, k, 0)])
                       + L.dxinv[1] * (uy.p[np_foff(uy, i, j + 1, k, 0)] - uy.p[np_foff(uy, i, j, k, 0)])
                       + L.dxinv[2] * (uz.p[np_foff(uz, i, j, k + 1, 0)] - uz.p[np_foff(uz, i, j, k, 0)]);
        out = -d;
        rhs[mac_c(L, i, j, k)] = out;
    }
    const double m = np_block_reduce<true>(fabs(out));
    const double s = np_block_reduce<false>(out);
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        const long long q = blockIdx.x + gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z);
        part_max[q] = m; part_sum[q] = s;
    }
}
// MLMG::getFluxes (mlabeclap_flux_x/y/z at face centres) and MacProjector::project's umac += flux.  One thread per cell does
// its three low faces and, on the last cell of a direction, the high boundary face.
//   ADD = true : u_d -= b_d * dphi/dx_d           ADD = false: flux_d = -b_d * dphi/dx_d   (getLinOp().getFluxes, :63)
template <bool ADD, bool CB>
__global__ void __launch_bounds__(256) k_mac_fluxes(MacLevel L, MacBeta B, const double* __restrict__ phi, FabView ux, FabView uy, FabView uz)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= L.n[0] || j >= L.n[1]) return;
    const long long c = mac_c(L, i, j, k);
    const double xc = phi[c];
    const MacB6 b = mac_load_b<CB>(L, B, i, j, k);
    const double fxm = -b.xm * L.dxinv[0] * (xc - mac_nb_lo(L, phi, c, xc, 0, i, 1, phi));
    const double fym = -b.ym * L.dxinv[1] * (xc - mac_nb_lo(L, phi, c, xc, 1, j, L.sx, phi));
    const double fzm = -b.zm * L.dxinv[2] * (xc - mac_nb_lo(L, phi, c, xc, 2, k, L.sxy, phi));
    double* p;
    p = ux.p + np_foff(ux, i, j, k, 0); *p = ADD ? *p + fxm : fxm;
    p = uy.p + np_foff(uy, i, j, k, 0); *p = ADD ? *p + fym : fym;
    p = uz.p + np_foff(uz, i, j, k, 0); *p = ADD ? *p + fzm : fzm;
    if (i == L.n[0] - 1) { const double f = -b.xp * L.dxinv[0] * (mac_nb_hi(L, phi, c, xc, 0, i, 1, phi) - xc); p = ux.p + np_foff(ux, i + 1, j, k, 0); *p = ADD ? *p + f : f; }
    if (j == L.n[1] - 1) { const double f = -b.yp * L.dxinv[1] * (mac_nb_hi(L, phi, c, xc, 1, j, L.sx, phi) - xc); p = uy.p + np_foff(uy, i, j + 1, k, 0); *p = ADD ? *p + f : f; }
    if (k == L.n[2] - 1) { const double f = -b.zp * L.dxinv[2] * (mac_nb_hi(L, phi, c, xc, 2, k, L.sxy, phi) - xc); p = uz.p + np_foff(uz, i, j, k + 1, 0); *p = ADD ? *p + f : f; }
}
// cell fab (any ghost count) <-> internal cell array
template <bool TO_FAB>
__global__ void __launch_bounds__(256) k_mac_copy_cell(MacLevel L, double* __restrict__ x, FabView f)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= L.n[0] || j >= L.n[1]) return;
    if (TO_FAB) f.p[np_foff(f, i, j, k, 0)] = x[mac_c(L, i, j, k)];
    else x[mac_c(L, i, j, k)] = f.p[np_foff(f, i, j, k, 0)];
}
// face array of direction d, internal layout -> fab (test hook mp_get_beta)
__global__ void __launch_bounds__(256) k_mac_beta_out(MacLevel L, const double* __restrict__ src, FabView dst, int d)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= L.n[0] + (d == 0) || j >= L.n[1] + (d == 1)) return;
    dst.p[np_foff(dst, i, j, k, 0)] = L.const_beta ? L.beta0 : src[d == 0 ? mac_fx(L, i, j, k) : d == 1 ? mac_fy(L, i, j, k) : mac_fz(L, i, j, k)];
}

// ------------------------------------------------------------------------------------------------
// Bottom solver: the whole BiCGStab / CG iteration of amrex::MLCGSolver on the coarsest level in ONE CTA (the level has
// 2^3 .. a few 10^4 cells): reductions are block barriers, not launches.  x = 0 on entry, so r0 = b.
// out[0] = MLCGSolver return code (0 ok, 1-4 breakdown, 8 not converged), out[1] = iterations.
// ------------------------------------------------------------------------------------------------
// MLMG with inhomogeneous Dirichlet data (MLCellLinOp::setLevelBC: the boundary values sit in the ghost cells of phi): the
// ghost of the first interior cell is  dc0 phi0 + dc1 phi1 + cbc value  (maxorder 3: cbc = 8/3).  MLMG evaluates the first
// residual with it and iterates on the homogeneous correction equation; moving  dh b_face cbc value  to the right-hand side
// once is the same thing.  One launch per Dirichlet face; phi is the caller's fab (>= 1 ghost cell).
__global__ void __launch_bounds__(256) k_mac_dirichlet_rhs(MacLevel L, MacBeta B, double* __restrict__ rhs, FabView phi, int d, int side, double cbc)
{
    const int a = (d + 1) % 3, b = (d + 2) % 3, t0 = a < b ? a : b, t1 = a < b ? b : a;
    const int p = blockIdx.x * blockDim.x + threadIdx.x, q = blockIdx.y * blockDim.y + threadIdx.y;
    if (p >= L.n[t0] || q >= L.n[t1]) return;
    int idx[3]; idx[d] = side ? L.n[d] - 1 : 0; idx[t0] = p; idx[t1] = q;
    int gh[3] = {idx[0], idx[1], idx[2]}; gh[d] += side ? 1 : -1;                 // the ghost cell that carries the wall value
    int fi[3] = {idx[0], idx[1], idx[2]}; if (side) fi[d] += 1;                  // the face
    double bf = L.beta0;
    if (!L.const_beta) bf = d == 0 ? B.bx[mac_fx(L, fi[0], fi[1], fi[2])] : d == 1 ? B.by[mac_fy(L, fi[0], fi[1], fi[2])] : B.bz[mac_fz(L, fi[0], fi[1], fi[2])];
    rhs[mac_c(L, idx[0], idx[1], idx[2])] += L.dh[d] * bf * cbc * phi.p[np_foff(phi, gh[0], gh[1], gh[2], 0)];
}
// rhs = rho * tracer (DiffusionScalarOp::diffuse_scalar, /root/reference/src/diffusion/DiffusionScalarOp.cpp:207-217)
__global__ void __launch_bounds__(256) k_mac_rho_times(MacLevel L, double* __restrict__ rhs, FabView rho, FabView tra)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z;
    if (i >= L.n[0] || j >= L.n[1]) return;
    rhs[mac_c(L, i, j, k)] = rho.p[np_foff(rho, i, j, k, 0)] * tra.p[np_foff(tra, i, j, k, 0)];
}

struct MacBottomArgs {
    MacLevel L; MacBeta B;
    double* x; const double* b;
    double *r, *rh, *p, *v, *s, *t;
    int maxiter, use_cg, singular;
    double eps_rel, eps_abs;
    int* out;
};
template <bool CB> __device__ __forceinline__ double mac_diag(const MacLevel& L, const MacBeta& B, int i, int j, int k)
{
    const MacB6 b = mac_load_b<CB>(L, B, i, j, k);
    return mac_aterm(L, mac_c(L, i, j, k)) + L.dh[0] * (b.xm + b.xp) + L.dh[1] * (b.ym + b.yp) + L.dh[2] * (b.zm + b.zp);
}
template <bool CB> __device__ __forceinline__ double mac_ax(const MacLevel& L, const MacBeta& B, const double* x, int i, int j, int k)
{
    const long long c = mac_c(L, i, j, k);
    const double xc = x[c];
    const MacB6 b = mac_load_b<CB>(L, B, i, j, k);
    const double xm = mac_nb_lo(L, x, c, xc, 0, i, 1, x), xp = mac_nb_hi(L, x, c, xc, 0, i, 1, x);
    const double ym = mac_nb_lo(L, x, c, xc, 1, j, L.sx, x), yp = mac_nb_hi(L, x, c, xc, 1, j, L.sx, x);
    const double zm = mac_nb_lo(L, x, c, xc, 2, k, L.sxy, x), zp = mac_nb_hi(L, x, c, xc, 2, k, L.sxy, x);
    return mac_aterm(L, c) * xc - L.dh[0] * (b.xp * (xp - xc) - b.xm * (xc - xm)) - L.dh[1] * (b.yp * (yp - xc) - b.ym * (xc - ym)) - L.dh[2] * (b.zp * (zp - xc) - b.zm * (xc - zm));
}
#define MAC_FOR_CELLS(q, i, j, k, c) \
    for (long long q = threadIdx.x; q < ncell; q += blockDim.x) { \
        const int i = (int)(q % A.L.n[0]), j = (int)((q / A.L.n[0]) % A.L.n[1]), k = (int)(q / ((long long)A.L.n[0] * A.L.n[1])); \
        const long long c = mac_c(A.L, i, j, k);
#define MAC_END }

template <bool CB>
__global__ void __launch_bounds__(1024) k_mac_bottom(MacBottomArgs A)
{
    const long long ncell = (long long)A.L.n[0] * A.L.n[1] * A.L.n[2];
    // makeSolvable(b): subtract the mean (singular problems); r = b (x = 0), normalised for BiCGStab
    double mean = 0.0;
    if (A.singular) {
        double s = 0.0;
        MAC_FOR_CELLS(q, i, j, k, c) s += A.b[c]; MAC_END
        mean = np_block_reduce<false>(s) / (double)ncell;
    }
    double m = 0.0;
    MAC_FOR_CELLS(q, i, j, k, c)
        double v = A.b[c] - mean;
        if (!A.use_cg) v /= mac_diag<CB>(A.L, A.B, i, j, k);
        A.r[c] = v; A.rh[c] = v; A.x[c] = 0.0; m = fmax(m, fabs(v));
    MAC_END
    double rnorm = np_block_reduce<true>(m);
    const double rnorm0 = rnorm;
    int ret = 0, iter = 1;
    if (rnorm0 == 0.0 || rnorm0 < A.eps_abs) { if (threadIdx.x == 0) { A.out[0] = 0; A.out[1] = 0; } return; }
    double rho_1 = 0.0, alpha = 0.0, omega = 0.0;
    if (!A.use_cg) {
        for (; iter <= A.maxiter; ++iter) {
            double s = 0.0;
            MAC_FOR_CELLS(q, i, j, k, c) s += A.rh[c] * A.r[c]; MAC_END
            const double rho = np_block_reduce<false>(s);
            if (rho == 0.0) { ret = 1; break; }
            if (iter == 1) { MAC_FOR_CELLS(q, i, j, k, c) A.p[c] = A.r[c]; MAC_END }
            else { const double beta = (rho / rho_1) * (alpha / omega); MAC_FOR_CELLS(q, i, j, k, c) A.p[c] = A.r[c] + beta * (A.p[c] - omega * A.v[c]); MAC_END }
            __syncthreads();
            s = 0.0;
            MAC_FOR_CELLS(q, i, j, k, c) const double v = mac_ax<CB>(A.L, A.B, A.p, i, j, k) / mac_diag<CB>(A.L, A.B, i, j, k); A.v[c] = v; s += A.rh[c] * v; MAC_END
            const double rhTv = np_block_reduce<false>(s);
            if (rhTv != 0.0) alpha = rho / rhTv; else { ret = 2; break; }
            m = 0.0;
            MAC_FOR_CELLS(q, i, j, k, c) A.x[c] += alpha * A.p[c]; const double v = A.r[c] - alpha * A.v[c]; A.s[c] = v; m = fmax(m, fabs(v)); MAC_END
            rnorm = np_block_reduce<true>(m);
            if (rnorm < A.eps_rel * rnorm0 || rnorm < A.eps_abs) break;
            double tt = 0.0, ts = 0.0;          // the barrier inside the reduction above made s visible
            MAC_FOR_CELLS(q, i, j, k, c) const double v = mac_ax<CB>(A.L, A.B, A.s, i, j, k) / mac_diag<CB>(A.L, A.B, i, j, k); A.t[c] = v; tt += v * v; ts += v * A.s[c]; MAC_END
            tt = np_block_reduce<false>(tt); ts = np_block_reduce<false>(ts);
            if (tt != 0.0) omega = ts / tt; else { ret = 3; break; }
            m = 0.0;
            MAC_FOR_CELLS(q, i, j, k, c) A.x[c] += omega * A.s[c]; const double v = A.s[c] - omega * A.t[c]; A.r[c] = v; m = fmax(m, fabs(v)); MAC_END
            rnorm = np_block_reduce<true>(m);
            if (rnorm < A.eps_rel * rnorm0 || rnorm < A.eps_abs) break;
            if (omega == 0.0) { ret = 4; break; }
            rho_1 = rho;
        }
    } else {
        for (; iter <= A.maxiter; ++iter) {
            double s = 0.0;
            MAC_FOR_CELLS(q, i, j, k, c) s += A.r[c] * A.r[c]; MAC_END
            const double rho = np_block_reduce<false>(s);
            if (rho == 0.0) { ret = 1; break; }
            if (iter == 1) { MAC_FOR_CELLS(q, i, j, k, c) A.p[c] = A.r[c]; MAC_END }
            else { const double beta = rho / rho_1; MAC_FOR_CELLS(q, i, j, k, c) A.p[c] = A.r[c] + beta * A.p[c]; MAC_END }
            __syncthreads();
            s = 0.0;
            MAC_FOR_CELLS(q, i, j, k, c) const double v = mac_ax<CB>(A.L, A.B, A.p, i, j, k); A.v[c] = v; s += A.p[c] * v; MAC_END
            const double pw = np_block_reduce<false>(s);
            if (pw != 0.0) alpha = rho / pw; else { ret = 1; break; }
            m = 0.0;
            MAC_FOR_CELLS(q, i, j, k, c) A.x[c] += alpha * A.p[c]; const double v = A.r[c] - alpha * A.v[c]; A.r[c] = v; m = fmax(m, fabs(v)); MAC_END
            rnorm = np_block_reduce<true>(m);
            if (rnorm < A.eps_rel * rnorm0 || rnorm < A.eps_abs) break;
            rho_1 = rho;
        }
    }
    if (ret == 0 && rnorm > A.eps_rel * rnorm0 && rnorm > A.eps_abs) ret = 8;
    if (!((ret == 0 || ret == 8) && rnorm < rnorm0)) { MAC_FOR_CELLS(q, i, j, k, c) A.x[c] = 0.0; MAC_END }
    if (threadIdx.x == 0) { A.out[0] = ret; A.out[1] = iter < A.maxiter ? iter : A.maxiter; }
}
