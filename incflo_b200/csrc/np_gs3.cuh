// np_gs3.cuh — K3, the Gauss-Seidel smoother (amrex::mlndlap_gauss_seidel_aa, two sweeps per Fsmooth) with the thread
// block shape as a template parameter.
//
// Same algorithm, ordering and pipeline as k_gs2 (np_kernels.cuh): tiles write 63 x 31 x zc nodes and
// relax 64 x 32 x (zc+1); sweep 1 then sweep 2, planes ascending, four colours c = 2*(j&1)+(i&1) inside a plane; one pass
// over HBM: sweep group 0 runs sweep 1 on plane t while sweep group 1 runs sweep 2 on plane t-2; a ring of six planes in
// shared memory filled one plane ahead by cp.async.  What is a parameter here is how a plane is cut among threads: a
// thread owns a BW x BH block of nodes, a sweep group has (64/BW) x (32/BH) threads, a CTA two groups.
//   k_gs3<2, 4, variable sigma>: 512 threads of <= 128 registers instead of k_gs2<var>'s 256 threads of 244 registers:
//       16 warps per SM instead of 8 for a kernel that ncu shows latency-bound (issue slots 46 % busy, fp64 pipe 44 %).
// In-plane halo values are kept in registers and re-read only where another thread's colour step has changed them.
// Shared-memory layout of a plane (66 x 34 with the tile halo): row-major, every row split into BW sub-arrays by
// (column-1) mod BW, so that the 64/BW threads of a row read consecutive doubles whichever of their columns (or halo
// columns) they touch: every access is a conflict-free 64-bit access.  Sub-array a holds position
// pos = (g-1) div BW + 1 of tile column g (g = 0, the left halo: sub-array BW-1, pos 0; g = 65, the right halo:
// sub-array 0, pos NLX+1); the sub-arrays are packed to 66 doubles per row.
#pragma once

template <int BW, int BH> struct G3 {
    static constexpr int NLX = GS_TX / BW;            // threads along x in a sweep group
    static constexpr int NTY = GS_TY / BH;            // thread rows in a sweep group
    static constexpr int GT = NLX * NTY;              // threads per sweep group
    static constexpr int CTA = 2 * GT;
    static constexpr int NW = CTA / 32;
    static constexpr int RP = GS_TX + 2;              // doubles per row
    static constexpr int ROWS = GS_TY + 2;
    static constexpr int PLANE = RP * ROWS;
    static constexpr int SMEM_BYTES = GS_RING * PLANE * 8;
    // first element of sub-array a minus its first position
    __host__ __device__ static constexpr int base(int a) {
        return a == 0 ? -1 : (a == BW - 1 ? (NLX + 1) + (a - 1) * NLX : (NLX + 1) + (a - 1) * NLX - 1);
    }
    __host__ __device__ static constexpr int own(int a) { return base(a) + 1; }     // + lx: own column a of x-lane lx
    static constexpr int LEFT = (BW == 1) ? 0 : ((NLX + 1) + (BW - 2) * NLX);      // base(BW-1): + lx: left halo column
    static constexpr int RIGHT = 1;                                                 // base(0) + 2: + lx: right halo column
    __device__ static __forceinline__ int slot_of_column(int g) {                  // loader: tile column g = 0..65
        return g == 0 ? LEFT : base((g - 1) % BW) + (g - 1) / BW + 1;
    }
};

// barrier among the threads of one sweep group (named barrier 1 + grp); the emulation's barrier is CTA-wide
template <int NT> __device__ __forceinline__ void np_group_barrier_n(int grp)
{
#ifdef NP_EMU
    __syncthreads();
#else
    if (grp == 0) asm volatile("bar.sync 1, %0;\n" :: "n"(NT) : "memory");
    else          asm volatile("bar.sync 2, %0;\n" :: "n"(NT) : "memory");
#endif
}

// Plane loader: warp w copies rows w, w+NW, ... (lanes -> tile columns lane and lane+32), and the two remaining columns
// (64, 65) of all 34 rows are copied by three warp-wide instructions (warps 0-2, lane -> row 16 w + lane/2).
template <int BW, int BH> struct G3Loader {
    typedef G3<BW, BH> G;
    int src0, dst0, dst1;        // row `warp`, columns lane / lane+32
    int srct, dstt;              // tail element (columns 64 / 65)
    int rows_ok; bool c0_ok, c1_ok, t_ok; int trow;
    int warp;
    __device__ __forceinline__ void init(const LevelDev& L, int tid, int i0, int j0, bool zero_in) {
        const int lane = tid & 31; warp = tid >> 5;
        const int rows_exist = min(G::ROWS, L.n[1] + 3 - j0);               // row index rho <-> node j = j0 - 1 + rho
        rows_ok = zero_in ? 0 : rows_exist;
        // tile column g <-> node i = i0 - 1 + g, array offset i + 1 = i0 + g; it exists when i <= n0 + 1
        c0_ok = (i0 - 1 + lane <= L.n[0] + 1); c1_ok = (i0 - 1 + lane + 32 <= L.n[0] + 1);
        src0 = (int)(L.nsy * (long long)(j0 + warp) + (i0 + lane));
        dst0 = warp * G::RP + G::slot_of_column(lane); dst1 = warp * G::RP + G::slot_of_column(lane + 32);
        trow = 16 * warp + (lane >> 1);
        const int gt = 64 + (lane & 1);
        t_ok = (warp < 3) && trow < rows_ok && (i0 - 1 + gt <= L.n[0] + 1);
        srct = (int)(L.nsy * (long long)(j0 + trow) + (i0 + gt));
        dstt = trow * G::RP + G::slot_of_column(gt);
        if (!(warp < 3 && trow < G::ROWS)) trow = -1;
    }
    template <bool ALLVALID>
    __device__ __forceinline__ void load(double* dst_plane, const double* src_plane, long long nsy) const {
        const double* sp = src_plane + src0;
        double* dp = dst_plane;
        const long long step = (long long)G::NW * nsy;
        constexpr int NIT = (G::ROWS + G::NW - 1) / G::NW;
#pragma unroll
        for (int it = 0; it < NIT; ++it) {
            const int row = warp + it * G::NW;
            if (it < NIT - 1 || row < G::ROWS) {
                if (ALLVALID) {
                    np_cp_async8_all(dp + dst0 + it * G::NW * G::RP, sp);
                    np_cp_async8_all(dp + dst1 + it * G::NW * G::RP, sp + 32);
                } else {
                    const bool rv = row < rows_ok;
                    np_cp_async8(dp + dst0 + it * G::NW * G::RP, (rv && c0_ok) ? sp : src_plane, rv && c0_ok);
                    np_cp_async8(dp + dst1 + it * G::NW * G::RP, (rv && c1_ok) ? sp + 32 : src_plane, rv && c1_ok);
                }
            }
            sp += step;
        }
        if (trow >= 0) {
            if (ALLVALID) np_cp_async8_all(dp + dstt, src_plane + srct);
            else          np_cp_async8(dp + dstt, t_ok ? src_plane + srct : src_plane, t_ok);
        }
    }
};

__host__ __device__ constexpr int g3_ci(int r, int n) { return r < 0 ? 0 : (r >= n ? 0 : r); }     // index kept legal in dead branches

// value of node (a, r) of the thread's neighbourhood, a in [-1, BW], r in [-1, BH]; all indices are compile-time constants
// after unrolling, so this is a register pick
template <int BW, int BH>
__device__ __forceinline__ double g3_nb(const double (&x)[BW][BH], const double (&Lh)[BH + 2], const double (&Rh)[BH + 2],
                                        const double (&Th)[BW], const double (&Bh)[BW], int a, int r)
{
    if (a < 0) return Lh[r + 1];
    if (a >= BW) return Rh[r + 1];
    if (r < 0) return Th[a];
    if (r >= BH) return Bh[a];
    return x[a][r];
}

// One colour (CX, CY) = parities of (a, r): every node of that colour in the thread's block, neighbours from registers.
//   constant sigma: x_new = B - (a_fx (W + E) + a_fy (S + N) + a_exy (SW + NW + SE + NE)),   B = rhs/diag - out-of-plane part
//   variable sigma: x_new = (B - in-plane part) / diag with V[a][r] = sigma_below + sigma_above of cell column (a-1, r-1)
//                   relative to the block, so node (a, r) sees the columns V[a][r], V[a+1][r], V[a][r+1], V[a+1][r+1]
template <int BW, int BH, bool CS, bool MASKED, int CX, int CY>
__device__ __forceinline__ void g3_relax(double (&x)[BW][BH], const double (&B)[BW][BH], const double (&V)[BW + 1][BH + 1],
                                         const double (&Lh)[BH + 2], const double (&Rh)[BH + 2], const double (&Th)[BW], const double (&Bh)[BW],
                                         double* Pc, int ofs, const GsCoef& C, const NpCoef& c, unsigned act, unsigned dmask)
{
    typedef G3<BW, BH> G;
#pragma unroll
    for (int r = CY; r < BH; r += 2)
#pragma unroll
        for (int a = CX; a < BW; a += 2) {
            const double W = g3_nb<BW, BH>(x, Lh, Rh, Th, Bh, a - 1, r), E = g3_nb<BW, BH>(x, Lh, Rh, Th, Bh, a + 1, r);
            const double S = g3_nb<BW, BH>(x, Lh, Rh, Th, Bh, a, r - 1), N = g3_nb<BW, BH>(x, Lh, Rh, Th, Bh, a, r + 1);
            const double SW = g3_nb<BW, BH>(x, Lh, Rh, Th, Bh, a - 1, r - 1), NW = g3_nb<BW, BH>(x, Lh, Rh, Th, Bh, a - 1, r + 1);
            const double SE = g3_nb<BW, BH>(x, Lh, Rh, Th, Bh, a + 1, r - 1), NE = g3_nb<BW, BH>(x, Lh, Rh, Th, Bh, a + 1, r + 1);
            double v;
            if (CS) v = fma(-C.a_exy, (SW + NW) + (SE + NE), fma(-C.a_fy, S + N, fma(-C.a_fx, W + E, B[a][r])));
            else    v = gs_upd_var(c, B[a][r], V[a][r], V[a + 1][r], V[a][r + 1], V[a + 1][r + 1], W, E, S, N, SW, SE, NW, NE);
            const int bit = a * BH + r;
            if (!MASKED) x[a][r] = v; else if ((act >> bit) & 1u) x[a][r] = ((dmask >> bit) & 1u) ? 0.0 : v;
            Pc[ofs + r * G::RP + G::own(a)] = x[a][r];
        }
}

// The four colour steps of one plane.  Halo registers: Lh / Rh = left / right halo column (rows -1..BH), Th / Bh = row -1 /
// row BH over the own columns.  They are loaded once (old values) and re-read only where another thread has relaxed them:
//   before colour 1: even rows of Rh (colour 0 of the right neighbour, same warp: __syncwarp);
//   before colour 2: even rows of Lh (colour 1), Bh and the corner Rh[BH] (colours 0 / 1 of the thread row below: group barrier);
//   before colour 3: odd rows of Rh (colour 2 of the right neighbour: __syncwarp).
// Row BH is not read before the group barrier: the thread row below may be writing it.  A warp covers whole rows of the
// tile (NLX divides 32), so "same rows" means "same warp".
template <int BW, int BH, bool CS, bool MASKED>
__device__ __forceinline__ void g3_colours(bool on, double (&x)[BW][BH], const double (&B)[BW][BH], const double (&V)[BW + 1][BH + 1],
                                           double* Pc, const GsCoef& C, const NpCoef& c, int ofs, unsigned act, unsigned dmask, int grp)
{
    typedef G3<BW, BH> G;
    constexpr int RP = G::RP;
    double Lh[BH + 2], Rh[BH + 2], Th[BW], Bh[BW];
#pragma unroll
    for (int q = 0; q < BH + 2; ++q) { Lh[q] = 0.0; Rh[q] = 0.0; }
#pragma unroll
    for (int a = 0; a < BW; ++a) { Th[a] = 0.0; Bh[a] = 0.0; }
    if (on) {
#pragma unroll
        for (int r = -1; r < BH; ++r) { Lh[r + 1] = Pc[ofs + r * RP + G::LEFT]; Rh[r + 1] = Pc[ofs + r * RP + G::RIGHT]; }
#pragma unroll
        for (int a = 0; a < BW; ++a) Th[a] = Pc[ofs - RP + G::own(a)];
        g3_relax<BW, BH, CS, MASKED, 0, 0>(x, B, V, Lh, Rh, Th, Bh, Pc, ofs, C, c, act, dmask);
    }
    __syncwarp();
    if (on) {
#pragma unroll
        for (int r = 0; r < BH; r += 2) Rh[r + 1] = Pc[ofs + r * RP + G::RIGHT];
        g3_relax<BW, BH, CS, MASKED, 1, 0>(x, B, V, Lh, Rh, Th, Bh, Pc, ofs, C, c, act, dmask);
    }
    np_group_barrier_n<G::GT>(grp);   // only this sweep's warps exchange values inside a plane
    if (on) {
#pragma unroll
        for (int r = 0; r <= BH; r += 2) Lh[r + 1] = Pc[ofs + r * RP + G::LEFT];
#pragma unroll
        for (int a = 0; a < BW; ++a) Bh[a] = Pc[ofs + BH * RP + G::own(a)];
        Rh[BH + 1] = Pc[ofs + BH * RP + G::RIGHT];
        g3_relax<BW, BH, CS, MASKED, 0, 1>(x, B, V, Lh, Rh, Th, Bh, Pc, ofs, C, c, act, dmask);
    }
    __syncwarp();
    if (on) {
#pragma unroll
        for (int r = 1; r < BH; r += 2) Rh[r + 1] = Pc[ofs + r * RP + G::RIGHT];
        g3_relax<BW, BH, CS, MASKED, 1, 1>(x, B, V, Lh, Rh, Th, Bh, Pc, ofs, C, c, act, dmask);
    }
}

// Constant sigma: B = rhs/diag - (9-point stencil of S = x[k-1] + x[k+1]).  B arrives holding the raw right-hand side; x
// holds the thread's block of plane p-1 as the previous step left it.
template <int BW, int BH>
__device__ __forceinline__ void g3_zpart_cs(const double (&x)[BW][BH], double (&B)[BW][BH], const double* Pm, const double* Pp,
                                            int ofs, const GsCoef& C, double roff)
{
    typedef G3<BW, BH> G;
#pragma unroll
    for (int rr = -1; rr <= BH; ++rr) {
        const int ro = ofs + rr * G::RP;
        const bool own = (rr >= 0 && rr < BH);
        double s[BW + 2];
        s[0] = Pm[ro + G::LEFT] + Pp[ro + G::LEFT];
#pragma unroll
        for (int a = 0; a < BW; ++a) s[a + 1] = (own ? x[a][g3_ci(rr, BH)] : Pm[ro + G::own(a)]) + Pp[ro + G::own(a)];
        s[BW + 1] = Pm[ro + G::RIGHT] + Pp[ro + G::RIGHT];
#pragma unroll
        for (int a = 0; a < BW; ++a) {
            const double h = s[a] + s[a + 2];
            const double q = C.a_eyz * s[a + 1] + C.a_c * h;
            if (own) B[a][g3_ci(rr, BH)] -= C.a_fz * s[a + 1] + C.a_exz * h;       // row rr's own contribution
            if (rr - 1 >= 0 && rr - 1 < BH) B[a][g3_ci(rr - 1, BH)] -= q;
            if (rr + 1 >= 0 && rr + 1 < BH)      // first touch of row rr+1: rhs/diag comes in here
                B[a][g3_ci(rr + 1, BH)] = fma(B[a][g3_ci(rr + 1, BH)] - roff, C.dinv, -q);
        }
    }
}

// Variable sigma: out-of-plane part of one neighbouring plane P (x[k-1] or x[k+1]) through the cell layer S between it and
// plane k; S[c][q] = sigma of cell (gi-1+c, gj-1+q) of that layer.
template <int BW, int BH>
__device__ __forceinline__ void g3_row(double (&w)[BW + 2], const double* P, int ro)
{
    typedef G3<BW, BH> G;
    w[0] = P[ro + G::LEFT];
#pragma unroll
    for (int a = 0; a < BW; ++a) w[a + 1] = P[ro + G::own(a)];
    w[BW + 1] = P[ro + G::RIGHT];
}
template <int BW, int BH>
__device__ __forceinline__ void g3_zpart_var(const NpCoef& c, double (&B)[BW][BH], const double (&S)[BW + 1][BH + 1], const double* P, int ofs)
{
    typedef G3<BW, BH> G;
    double xa[BW + 2], xb[BW + 2], xc[BW + 2];     // node rows r-1, r, r+1; columns -1 .. BW
    g3_row<BW, BH>(xa, P, ofs - G::RP); g3_row<BW, BH>(xb, P, ofs);
#pragma unroll
    for (int r = 0; r < BH; ++r) {
        g3_row<BW, BH>(xc, P, ofs + (r + 1) * G::RP);
#pragma unroll
        for (int a = 0; a < BW; ++a) {
            const double l00 = S[a][r], l10 = S[a + 1][r], l01 = S[a][r + 1], l11 = S[a + 1][r + 1];
            B[a][r] -= c.fm2xm2y4z * (xb[a + 1] * ((l00 + l10) + (l01 + l11)))
                     + c.f2xmy2z * (xb[a] * (l00 + l01) + xb[a + 2] * (l10 + l11))
                     + c.fmx2y2z * (xa[a + 1] * (l00 + l10) + xc[a + 1] * (l01 + l11))
                     + c.fxyz * ((xa[a] * l00 + xa[a + 2] * l10) + (xc[a] * l01 + xc[a + 2] * l11));
        }
#pragma unroll
        for (int q = 0; q < BW + 2; ++q) { xa[q] = xb[q]; xb[q] = xc[q]; }
    }
}

// Body of k_gs3.  FULL: the whole 64 x 32 tile lies inside the domain and no face is Dirichlet, so every node is relaxed and
// the masks fold away at compile time (only the 63 x 31 write window remains).
template <int BW, int BH, bool CS, bool FULL>
__device__ __forceinline__ void gs3_run(const LevelDev& L, const GsArgs& A, double* ring)
{
    typedef G3<BW, BH> G;
    constexpr int RP = G::RP;
    const GsCoef& C = A.c;
    double* __restrict__ xout = A.xout; const double* __restrict__ xin = A.xin;
    const double* __restrict__ rhs = A.rhs; const double* __restrict__ sig = A.sig;
    const int zc = A.zc;
    const double roff = A.boff ? A.boff[0] : 0.0;
    const int tid = threadIdx.x;
    const int grp = tid / G::GT;                         // 0: first sweep (plane t), 1: second sweep (plane t-2)
    const int tg = tid - grp * G::GT;
    const int i0 = blockIdx.x * GS_WX, j0 = blockIdx.y * GS_WY, kb = (A.zmap0 + (int)blockIdx.z * A.zmapS) * zc;
    // Which row block a warp owns.  On a tile that is not full only the first act_ty row blocks have nodes; warps go to the four
    // schedulers of the SM round-robin, so the second sweep group's row blocks are rotated to start on the scheduler after the
    // first group's last busy one (a small level is a chain of plane steps bound by the issue slots of its busiest scheduler).
    // A thin tile (at most 32 live columns, BW = 2: sixteen lanes hold them) packs two row blocks into a warp, so half as many
    // warps are busy and each still owns whole rows (the __syncwarp between colours 0 -> 1 and 2 -> 3 stays sufficient).
    int lx = tg % G::NLX;
    int ty = tg / G::NLX;
    bool dead = false;                                   // a thread without a row block (thin tiles only)
    if (!FULL) {
        const int act_ty = (min(GS_TY, L.n[1] + 1 - j0) + BH - 1) / BH;
        if (G::NLX == 32) {
            constexpr int NWG = G::GT / 32;              // warps per sweep group (= NTY)
            const bool thin = min(GS_TX, L.n[0] + 1 - i0) <= 32;
            const int busy = thin ? (act_ty + 1) / 2 : act_ty;
            int w = tg >> 5;
            const int lane = tg & 31;
            if (grp == 1 && busy < NWG) w = (w - busy % NWG + NWG) % NWG;
            if (thin) { lx = lane & 15; ty = 2 * w + (lane >> 4); if (ty >= G::NTY) { dead = true; ty = 0; } }
            else ty = w;
        } else if (grp == 1 && act_ty < G::NTY) ty = (ty - act_ty % G::NTY + G::NTY) % G::NTY;
    }
    const int nk  = min(kb + zc, L.n[2]) - kb + 1;       // relaxed planes p = 0..nk-1 (global k = kb+p)
    const int nkw = min(kb + zc - 1, L.n[2]) - kb + 1;   // written planes
    const int gi = i0 + BW * lx, gj = j0 + BH * ty;      // first node of this thread's block
    const int ofs = (BH * ty + 1) * RP + lx;

    // per-thread node masks, bit a*BH+r: relaxed (valid node), Dirichlet in x/y, written by this tile
    unsigned act = 0, dir = 0, wrt = 0;
#pragma unroll
    for (int a = 0; a < BW; ++a)
#pragma unroll
        for (int r = 0; r < BH; ++r) {
            const int i = gi + a, j = gj + r;
            const bool v = FULL || (!dead && i <= L.n[0] && j <= L.n[1]);
            const bool dd = !FULL && v && ((L.face[0] == NP_FACE_DIRICHLET && i == 0) || (L.face[3] == NP_FACE_DIRICHLET && i == L.n[0])
                               || (L.face[1] == NP_FACE_DIRICHLET && j == 0) || (L.face[4] == NP_FACE_DIRICHLET && j == L.n[1]));
            const bool w = v && (BW * lx + a < GS_WX) && (BH * ty + r < GS_WY);
            act |= (unsigned)v << (a * BH + r); dir |= (unsigned)dd << (a * BH + r); wrt |= (unsigned)w << (a * BH + r);
        }

    G3Loader<BW, BH> ld; ld.init(L, tid, i0, j0, A.zero_in != 0);
    auto slot = [&](int p) -> double* { return ring + ((p + 1) % GS_RING) * G::PLANE; };     // p >= -1
    auto load_plane = [&](int p) {                                                           // planes -1 .. nk
        if (FULL && !A.zero_in) ld.template load<true>(slot(p), xin + L.nsz * (long long)(kb + p + 1), L.nsy);
        else                    ld.template load<false>(slot(p), xin + L.nsz * (long long)(kb + p + 1), L.nsy);
    };

    load_plane(-1); load_plane(0); load_plane(1);
    np_cp_async_commit(); np_cp_async_wait_all();
    __syncthreads();

    double x[BW][BH], B[BW][BH];
#pragma unroll
    for (int a = 0; a < BW; ++a)
#pragma unroll
        for (int r = 0; r < BH; ++r) { x[a][r] = 0.0; B[a][r] = 0.0; }

    auto load_rhs = [&](int pn) {                        // B <- raw rhs of plane pn (roff where the node is not relaxed)
        const double* rp = rhs + np_noff(L, gi, gj, kb + pn);
#pragma unroll
        for (int r = 0; r < BH; ++r)
#pragma unroll
            for (int a = 0; a < BW; ++a)
                B[a][r] = (FULL || ((act >> (a * BH + r)) & 1u)) ? rp[L.nsy * r + a] : roff;
    };
    // sigma of the cells (gi-1+c, gj-1+q) of cell layer kc (k index); indices clamped into the array (clamped cells only
    // feed nodes outside the domain, which are never updated)
    auto load_sig = [&](double (&S)[BW + 1][BH + 1], int kc) {
        const double* sb = sig + L.csz * (long long)(kc + 1);
#pragma unroll
        for (int c = 0; c <= BW; ++c)
#pragma unroll
            for (int q = 0; q <= BH; ++q) {
                const int ci = FULL ? gi - 1 + c : min(gi - 1 + c, L.n[0]), cj = FULL ? gj - 1 + q : min(gj - 1 + q, L.n[1]);
                S[c][q] = sb[(ci + 1) + L.csy * (long long)(cj + 1)];
            }
    };
    if (CS && grp == 0) load_rhs(0);

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
            for (int r = 0; r < BH; ++r) np_prefetch_l1(A.addto + gbase + L.nsy * r);
        }
        double V[BW + 1][BH + 1];
        if (CS) {
#pragma unroll
            for (int c = 0; c <= BW; ++c)
#pragma unroll
                for (int q = 0; q <= BH; ++q) V[c][q] = 0.0;
            if (on) {
                const double* Pm = slot(p - 1); const double* Pp = slot(p + 1);
                if (p == 0) {           // first plane of the chunk: the frozen plane below comes from the ring
#pragma unroll
                    for (int r = 0; r < BH; ++r)
#pragma unroll
                        for (int a = 0; a < BW; ++a) x[a][r] = Pm[ofs + r * RP + G::own(a)];
                }
                g3_zpart_cs<BW, BH>(x, B, Pm, Pp, ofs, C, roff);
#pragma unroll
                for (int r = 0; r < BH; ++r)
#pragma unroll
                    for (int a = 0; a < BW; ++a) x[a][r] = Pc[ofs + r * RP + G::own(a)];
            }
        } else {
            if (on && p + 1 <= nk - 1) {
                // this thread's next plane: pull its right-hand side and the sigma layer above it into L1 while this plane is
                // being relaxed, so the loads at the top of the next step are hits
#pragma unroll
                for (int r = 0; r < BH; ++r) np_prefetch_l1(rhs + gbase + L.nsz + L.nsy * r);
                const double* sn = sig + L.csz * (long long)(k + 2) + (min(gi, L.n[0]) + 1);
#pragma unroll
                for (int q = 0; q <= BH; ++q) np_prefetch_l1(sn + L.csy * (long long)(min(gj - 1 + q, L.n[1]) + 1));
            }
            if (on) {
                const double* rp = rhs + gbase;
#pragma unroll
                for (int r = 0; r < BH; ++r)
#pragma unroll
                    for (int a = 0; a < BW; ++a)
                        B[a][r] = (FULL || ((act >> (a * BH + r)) & 1u)) ? rp[L.nsy * r + a] - roff : 0.0;
                double S[BW + 1][BH + 1];
                load_sig(S, k - 1);                                   // cells below plane k
                g3_zpart_var<BW, BH>(L.c, B, S, slot(p - 1), ofs);
#pragma unroll
                for (int c = 0; c <= BW; ++c)
#pragma unroll
                    for (int q = 0; q <= BH; ++q) V[c][q] = S[c][q];
                load_sig(S, k);                                       // cells above
                g3_zpart_var<BW, BH>(L.c, B, S, slot(p + 1), ofs);
#pragma unroll
                for (int c = 0; c <= BW; ++c)
#pragma unroll
                    for (int q = 0; q <= BH; ++q) V[c][q] += S[c][q];
#pragma unroll
                for (int r = 0; r < BH; ++r)
#pragma unroll
                    for (int a = 0; a < BW; ++a) x[a][r] = Pc[ofs + r * RP + G::own(a)];
            } else {
#pragma unroll
                for (int c = 0; c <= BW; ++c)
#pragma unroll
                    for (int q = 0; q <= BH; ++q) V[c][q] = 1.0;
            }
        }
        g3_colours<BW, BH, CS, !FULL>(on, x, B, V, Pc, C, L.c, ofs, act, dmask, grp);
        // constant sigma: B is free now: fetch the raw right-hand side of this group's next plane so the loads fly across the barrier
        if (CS && p + 1 >= 0 && p + 1 <= nk - 1) load_rhs(p + 1);
        np_cp_async_wait_all();
        __syncthreads();
        if (on && grp == 1 && p < nkw) {
            double* op = xout + gbase;
            if (A.addto) {
                const double* ap = A.addto + gbase;
                double av[BW][BH];
#pragma unroll
                for (int r = 0; r < BH; ++r)
#pragma unroll
                    for (int a = 0; a < BW; ++a) av[a][r] = ((wrt >> (a * BH + r)) & 1u) ? ap[L.nsy * r + a] : 0.0;
#pragma unroll
                for (int r = 0; r < BH; ++r)
#pragma unroll
                    for (int a = 0; a < BW; ++a) if ((wrt >> (a * BH + r)) & 1u) op[L.nsy * r + a] = av[a][r] + x[a][r];
            }
            else {
#pragma unroll
                for (int r = 0; r < BH; ++r)
#pragma unroll
                    for (int a = 0; a < BW; ++a) if ((wrt >> (a * BH + r)) & 1u) op[L.nsy * r + a] = x[a][r];
            }
        }
    }
}

template <int BW, int BH, bool CS, int MINB>
__global__ void __launch_bounds__(2 * (GS_TX / BW) * (GS_TY / BH), MINB) k_gs3(LevelDev L, GsArgs A)
{
    NP_DYN_SMEM(double, ring);
    const bool any_dir = L.face[0] == NP_FACE_DIRICHLET || L.face[1] == NP_FACE_DIRICHLET || L.face[2] == NP_FACE_DIRICHLET
                      || L.face[3] == NP_FACE_DIRICHLET || L.face[4] == NP_FACE_DIRICHLET || L.face[5] == NP_FACE_DIRICHLET;
    const bool full = ((int)blockIdx.x * GS_WX + GS_TX - 1 <= L.n[0]) && ((int)blockIdx.y * GS_WY + GS_TY - 1 <= L.n[1]) && !any_dir;
    if (full) gs3_run<BW, BH, CS, true>(L, A, ring);
    else      gs3_run<BW, BH, CS, false>(L, A, ring);
}
