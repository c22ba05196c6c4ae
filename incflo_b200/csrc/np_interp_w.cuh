// np_interp_w.cuh — K5, sigma-weighted prolongation fine += P(crse) (amrex mlndlap_interpadd_aa: aa_interp_line_x/y/z,
// aa_interp_face_xy/xz/yz and the cell-centre formula), marching in z.
//
// k_interp_add_w (np_kernels.cuh) gives one thread a whole coarse cell: 12 edges, 6 faces, the centre, 216 sigma loads through
// L1 and eight scattered 8-byte read-modify-writes - 1.67 ms per 512^3 launch, 21 % of HBM bandwidth.  Here a thread (I, J)
// owns the LOW corner of its coarse cell - per coarse plane K the seven fine nodes
//     even plane 2K+2:  x-line (2I+1, 2J)   y-line (2I, 2J+1)    xy-face (2I+1, 2J+1)        (+ the coarse node itself)
//     odd  plane 2K+1:  z-line (2I, 2J)     xz-face (2I+1, 2J)   yz-face (2I, 2J+1)   centre (2I+1, 2J+1)
// and marches in K, so every edge, face and centre value is computed exactly once.  The 27 fine cells around those nodes
// (2I-1..2I+1, 2J-1..2J+1, 2K..2K+2) live in registers, 18 new loads per step (the top layer becomes the bottom one).  The
// values a node needs from the neighbouring threads I+1 and J+1 (their lines for a face, their faces for the centre) go
// through five small shared-memory arrays; the lines and the xy-face of the plane below are carried in registers.  These
// are exactly the values of fine column 2I+2, so a thread stores the 16-byte aligned pairs (2I+1, 2I+2) of its rows: four
// aligned 128-bit read-modify-writes per step, fully coalesced.  A CTA is 32 x 8 threads of which column 31 and row 7 only
// provide values to their neighbours (they are written by the next tile): 31 x 7 coarse cells per CTA and plane.
// The operation order inside every formula is that of k_interp_add_w (np_kernels.cuh: InterpCtx), so results are bitwise equal.
#pragma once

#define IW_TX 31
#define IW_TY 7

// sum of the four cells on side SIDE along D of the node whose lower cells are block cells (X0, Y0, Z0)
template <int X0, int Y0, int Z0, int D, int SIDE>
__device__ __forceinline__ double iw_quad(const double (&s)[3][3][3])
{
    if (D == 0) { constexpr int q = X0 + SIDE; return (s[Z0][Y0][q] + s[Z0][Y0 + 1][q]) + (s[Z0 + 1][Y0][q] + s[Z0 + 1][Y0 + 1][q]); }
    if (D == 1) { constexpr int q = Y0 + SIDE; return (s[Z0][q][X0] + s[Z0][q][X0 + 1]) + (s[Z0 + 1][q][X0] + s[Z0 + 1][q][X0 + 1]); }
    constexpr int q = Z0 + SIDE; return (s[q][Y0][X0] + s[q][Y0][X0 + 1]) + (s[q][Y0 + 1][X0] + s[q][Y0 + 1][X0 + 1]);
}
template <int X0, int Y0, int Z0, int D>
__device__ __forceinline__ double iw_line(const double (&s)[3][3][3], double clo, double chi)
{
    const double w1 = iw_quad<X0, Y0, Z0, D, 0>(s), w2 = iw_quad<X0, Y0, Z0, D, 1>(s);
    return (w1 * clo + w2 * chi) * np_rcp(w1 + w2);
}
// la0 / la1: DB-lines at the node's low / high neighbour along DA;  lb0 / lb1: DA-lines at its neighbours along DB
template <int X0, int Y0, int Z0, int DA, int DB>
__device__ __forceinline__ double iw_face(const double (&s)[3][3][3], double la0, double la1, double lb0, double lb1)
{
    const double wa0 = iw_quad<X0, Y0, Z0, DA, 0>(s), wa1 = iw_quad<X0, Y0, Z0, DA, 1>(s);
    const double wb0 = iw_quad<X0, Y0, Z0, DB, 0>(s), wb1 = iw_quad<X0, Y0, Z0, DB, 1>(s);
    return (wa0 * la0 + wa1 * la1 + wb0 * lb0 + wb1 * lb1) * np_rcp((wa0 + wa1) + (wb0 + wb1));
}

__device__ __forceinline__ void iw_rmw2(double* p, double v0, double v1)
{
    double2 f = *reinterpret_cast<double2*>(p);
    f.x += v0; f.y += v1;
    *reinterpret_cast<double2*>(p) = f;
}

__global__ void __launch_bounds__(256, 2) k_interp_add_w2(LevelDev Lf, LevelDev Lc, double* __restrict__ fine, const double* __restrict__ crse,
                                                          const double* __restrict__ sig, int koffc, int kcz)
{
    __shared__ double sh_lx[8][33], sh_ly[8][33], sh_lz[8][33], sh_fxz[8][33], sh_fyz[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int I = blockIdx.x * IW_TX + tx, J = blockIdx.y * IW_TY + ty;
    const int nc0 = Lf.n[0] >> 1, nc1 = Lf.n[1] >> 1, nc2 = Lf.n[2] >> 1;
    const int K0 = blockIdx.z * kcz, K1 = min(K0 + kcz, nc2);         // coarse cells K0 .. K1-1: fine planes 2K0+1 .. 2K1
    const bool ex = (I <= nc0 && J <= nc1);                            // the coarse node (I, J) exists
    const bool hx = ex && I < nc0 && tx < IW_TX;                       // computes the nodes of fine column 2I+1 ...
    const bool hy = ex && J < nc1 && ty < IW_TY;                       // ... and of fine row 2J+1
    const bool wr = hx && ty < IW_TY;                                  // stores (column 31 / row 7 only provide)
    const int i = 2 * I, j = 2 * J;
    // Dirichlet nodes keep their value: their increment is replaced by zero
    const bool dx0 = (Lf.face[0] == NP_FACE_DIRICHLET && I == 0), dx2 = (Lf.face[3] == NP_FACE_DIRICHLET && I + 1 == nc0);
    const bool dy0 = (Lf.face[1] == NP_FACE_DIRICHLET && J == 0) || (Lf.face[4] == NP_FACE_DIRICHLET && J == nc1);     // fine row 2J

    const double* cp = crse + np_noff(Lc, ex ? I : 0, ex ? J : 0, koffc);           // coarse node (I, J, 0)
    const int cxs = min(i - 1 + 2, Lf.n[0]) - (i - 1);                              // offset of block column X = 2 (clamped into the ghost layer)
    const int cys = min(j - 1 + 2, Lf.n[1]) - (j - 1);
    const double* sp = sig + np_coff(Lf, ex ? i - 1 : 0, ex ? j - 1 : 0, 0);        // fine cell (2I-1, 2J-1, 0)
    double* fp = fine + np_noff(Lf, ex ? i + 1 : 0, ex ? j : 0, 0);                  // fine node (2I+1, 2J, 0): 16-byte aligned

    double s[3][3][3];
#pragma unroll
    for (int z = 0; z < 3; ++z)
#pragma unroll
        for (int y = 0; y < 3; ++y)
#pragma unroll
            for (int x = 0; x < 3; ++x) s[z][y][x] = 1.0;
    double c_prev = 0.0, lx_prev = 0.0, ly_prev = 0.0, fxy_prev = 0.0;      // plane 2K of the previous step: coarse node, lines, xy-face

    for (int K = K0 - 1; K < K1; ++K) {
        const bool odd = (K >= K0);                                   // the first step only builds the even plane 2K0 (the carry)
        // ---- sigma: layers 2K+1, 2K+2 arrive, layer 2K (the previous top) moves down ----
#pragma unroll
        for (int y = 0; y < 3; ++y)
#pragma unroll
            for (int x = 0; x < 3; ++x) s[0][y][x] = s[2][y][x];
        if (ex) {
#pragma unroll
            for (int z = 1; z < 3; ++z) {
                const double* sl = sp + Lf.csz * (long long)(2 * K + z);
#pragma unroll
                for (int y = 0; y < 3; ++y) {
                    const double* sr = sl + Lf.csy * (long long)(y == 2 ? cys : y);
                    s[z][y][0] = sr[0]; s[z][y][1] = sr[1]; s[z][y][2] = sr[cxs];
                }
            }
        }
        // next step's lines into L1 (ncu: 12 of 17 stall cycles per issue were long-scoreboard waits on the loads at the top of
        // a step and on the read-modify-write at its end; two CTAs of 120 registers per SM have nothing else to hide them behind)
        if (ex && K + 1 < K1) {
#pragma unroll
            for (int z = 1; z < 3; ++z)
#pragma unroll
                for (int y = 0; y < 3; ++y) np_prefetch_l1(sp + Lf.csz * (long long)(2 * K + 2 + z) + Lf.csy * (long long)(y == 2 ? cys : y));
            np_prefetch_l1(cp + Lc.nsz * (long long)(K + 2));
            if (wr) {
                const double* fn = fp + Lf.nsz * (long long)(2 * K + 3);
                np_prefetch_l1(fn); np_prefetch_l1(fn + Lf.nsy); np_prefetch_l1(fn + Lf.nsz); np_prefetch_l1(fn + Lf.nsz + Lf.nsy);
            }
        }
        // ---- phase A: the coarse node and the three lines ----
        const long long ck = Lc.nsz * (long long)(K + 1);
        double c00 = 0.0, c10 = 0.0, lxE = 0.0, lyE = 0.0, lz = 0.0;
        if (ex) {
            c00 = cp[ck];
            if (hx) { c10 = cp[ck + 1]; lxE = iw_line<1, 0, 1, 0>(s, c00, c10); }
            if (hy) lyE = iw_line<0, 1, 1, 1>(s, c00, cp[ck + Lc.nsy]);
            if (odd) lz = iw_line<0, 0, 0, 2>(s, c_prev, c00);
        }
        sh_lx[ty][tx] = lxE; sh_ly[ty][tx] = lyE; sh_lz[ty][tx] = lz;
        __syncthreads();
        // ---- phase B: the three faces ----
        double fxyE = 0.0, fxz = 0.0, fyz = 0.0, lyE_hi = 0.0, lz_hi = 0.0;
        if (hx) { lyE_hi = sh_ly[ty][tx + 1]; lz_hi = sh_lz[ty][tx + 1]; }
        if (hx && hy) fxyE = iw_face<1, 1, 1, 0, 1>(s, lyE, lyE_hi, lxE, sh_lx[ty + 1][tx]);
        if (odd) {
            if (hx) fxz = iw_face<1, 0, 0, 2, 0>(s, lx_prev, lxE, lz, lz_hi);
            if (hy) fyz = iw_face<0, 1, 0, 1, 2>(s, lz, sh_lz[ty + 1][tx], ly_prev, lyE);
        }
        sh_fxz[ty][tx] = fxz; sh_fyz[ty][tx] = fyz;
        __syncthreads();
        // ---- phase C: the centre, then the stores ----
        double cen = 0.0, fyz_hi = 0.0;
        if (odd && hx && hy) {
            fyz_hi = sh_fyz[ty][tx + 1];
            const double x0 = iw_quad<1, 1, 0, 0, 0>(s), x1 = iw_quad<1, 1, 0, 0, 1>(s), y0 = iw_quad<1, 1, 0, 1, 0>(s), y1 = iw_quad<1, 1, 0, 1, 1>(s),
                         z0 = iw_quad<1, 1, 0, 2, 0>(s), z1 = iw_quad<1, 1, 0, 2, 1>(s);
            const double num = (x0 * fyz + x1 * fyz_hi) + (y0 * fxz + y1 * sh_fxz[ty + 1][tx]) + (z0 * fxy_prev + z1 * fxyE);
            cen = num * np_rcp(((x0 + x1) + (y0 + y1)) + (z0 + z1));
        }
        if (wr) {
            const int ke = 2 * K + 2, ko = 2 * K + 1;
            const bool dze = (Lf.face[2] == NP_FACE_DIRICHLET && ke == 0) || (Lf.face[5] == NP_FACE_DIRICHLET && ke == Lf.n[2]);
            // even plane 2K+2 (written by every step but the carry-building one, unless that one builds plane 0)
            if (odd || K0 == 0) {
                double* pe = fp + Lf.nsz * (long long)ke;                // node (2I+1, 2J, ke); pe[-1] is column 0 when I == 0
                const bool z0r = dze || dy0;                           // whole row 2J is Dirichlet
                iw_rmw2(pe, z0r ? 0.0 : lxE, (z0r || dx2) ? 0.0 : c10);
                if (I == 0) pe[-1] += (z0r || dx0) ? 0.0 : c00;
                if (hy) {
                    iw_rmw2(pe + Lf.nsy, dze ? 0.0 : fxyE, (dze || dx2) ? 0.0 : lyE_hi);
                    if (I == 0) pe[Lf.nsy - 1] += (dze || dx0) ? 0.0 : lyE;
                }
            }
            if (odd) {                                                 // odd plane 2K+1: never on a z face
                double* po = fp + Lf.nsz * (long long)ko;
                iw_rmw2(po, dy0 ? 0.0 : fxz, (dy0 || dx2) ? 0.0 : lz_hi);
                if (I == 0) po[-1] += (dy0 || dx0) ? 0.0 : lz;
                if (hy) {
                    iw_rmw2(po + Lf.nsy, cen, dx2 ? 0.0 : fyz_hi);
                    if (I == 0) po[Lf.nsy - 1] += dx0 ? 0.0 : fyz;
                }
            }
        }
        c_prev = c00; lx_prev = lxE; ly_prev = lyE; fxy_prev = fxyE;
    }
}
