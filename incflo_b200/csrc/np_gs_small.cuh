// np_gs_small.cuh — K3 for the small multigrid levels (amrex::mlndlap_gauss_seidel_aa, two sweeps per Fsmooth).
//
// Same algorithm, ordering, tiles and pipeline as k_gs2 (np_kernels.cuh): tiles write 63 x 31 x zc nodes
// and relax 64 x 32 x (zc+1); sweep 1 then sweep 2, planes ascending, four colours c = 2*(j&1)+(i&1) inside a plane; sweep
// group 0 runs sweep 1 on plane t while group 1 runs sweep 2 on plane t-2; six-plane shared-memory ring filled by cp.async.
// What differs is how a plane is cut among threads.  On a level that cannot fill the GPU (at most one CTA per SM, seven
// planes per CTA) the time of a launch is the length of the dependent instruction stream of one plane step times nine
// steps; k_gs2's 2 x 8 blocks in registers make that stream about 700 instructions per thread with 8 warps per SM to
// hide it behind (2 us per step, 22-27 us per launch whatever the level size).  Here a thread owns a 2 x 2 block - one node
// of every colour - and a CTA has 2 x 512 threads, so a colour step is one update per thread and 32 warps share the SM.
// Values are exchanged through shared memory only (no register blocking): more shared-memory traffic per update, which does
// not matter where nothing else runs on the SM.  A small level's chunk has at most 12 planes (zc <= 9), so the whole chunk
// is brought into shared memory by one bulk cp.async phase before the march (no per-step loader, no per-step wait), and
// the right-hand side of a thread's four nodes is fetched one step ahead into registers.
#pragma once

#define GSS_MAX_ZC 9                     // (zc + 3) planes of GS_PLANE doubles must fit 227 KB of shared memory
#define GSS_SMEM_BYTES(zc) ((size_t)((zc) + 3) * GS_PLANE * 8)

#define GSS_GT 512                       // threads per sweep group: 32 x 16 blocks of 2 x 2 nodes
#define GSS_CTA (2 * GSS_GT)

__device__ __forceinline__ void np_group_barrier512(int grp)
{
#ifdef NP_EMU
    __syncthreads();
#else
    if (grp == 0) asm volatile("bar.sync 1, 512;\n" ::: "memory");
    else          asm volatile("bar.sync 2, 512;\n" ::: "memory");
#endif
}

// element (li, lj) of a ring plane, li = -1..64, lj = -1..32 relative to the tile origin (split even/odd column layout of k_gs2)
__device__ __forceinline__ int gss_at(int li, int lj) { return ((li & 1) ? GS_HALF : 0) + (lj + 1) * GS_HP + ((li + 2) >> 1); }

template <bool CS>
__global__ void __launch_bounds__(GSS_CTA, 1) k_gs_small(LevelDev L, GsArgs A)
{
    NP_DYN_SMEM(double, ring);
    const GsCoef& C = A.c;
    const NpCoef& c = L.c;
    double* __restrict__ xout = A.xout; const double* __restrict__ xin = A.xin;
    const double* __restrict__ rhs = A.rhs; const double* __restrict__ sig = A.sig;
    const int zc = A.zc;
    const double roff = A.boff ? A.boff[0] : 0.0;
    const int tid = threadIdx.x, grp = tid / GSS_GT, tg = tid % GSS_GT;
    const int bx = tg & 31, by = tg >> 5;                    // this thread's block: tile nodes (2bx + a, 2by + b)
    const int i0 = blockIdx.x * GS_WX, j0 = blockIdx.y * GS_WY, kb = (A.zmap0 + (int)blockIdx.z * A.zmapS) * zc;
    const int nk  = min(kb + zc, L.n[2]) - kb + 1;           // relaxed planes p = 0..nk-1 (global k = kb+p)
    const int nkw = min(kb + zc - 1, L.n[2]) - kb + 1;       // written planes
    const int li0 = 2 * bx, lj0 = 2 * by, gi = i0 + li0, gj = j0 + lj0;

    // node masks, bit q = 2b + a (= the colour of the node): relaxed, Dirichlet in x / y, written by this tile
    unsigned act = 0, dir = 0, wrt = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int a = q & 1, b = q >> 1, i = gi + a, j = gj + b;
        const bool v = (i <= L.n[0] && j <= L.n[1]);
        const bool dd = v && ((L.face[0] == NP_FACE_DIRICHLET && i == 0) || (L.face[3] == NP_FACE_DIRICHLET && i == L.n[0])
                           || (L.face[1] == NP_FACE_DIRICHLET && j == 0) || (L.face[4] == NP_FACE_DIRICHLET && j == L.n[1]));
        const bool w = v && (li0 + a < GS_WX) && (lj0 + b < GS_WY);
        act |= (unsigned)v << q; dir |= (unsigned)dd << q; wrt |= (unsigned)w << q;
    }

    // bulk load: planes -1 .. nk of the tile (66 x 34 elements each, halo included) over all threads of the CTA; elements
    // beyond the level's ghost layer, and everything when the input is known to be zero, are zero-filled
    auto slot = [&](int p) -> double* { return ring + (p + 1) * GS_PLANE; };               // p >= -1
    {
        int dsto[3]; long long srco[3]; bool val[3];
#pragma unroll
        for (int m = 0; m < 3; ++m) {
            const int e = tid + m * GSS_CTA;
            const int row = e / 66, col = e - 66 * row;       // lj = row - 1, li = col - 1
            const int i = i0 + col - 1, j = j0 + row - 1;
            val[m] = e < 66 * 34 && !A.zero_in && i <= L.n[0] + 1 && j <= L.n[1] + 1;
            dsto[m] = e < 66 * 34 ? gss_at(col - 1, row - 1) : -1;
            srco[m] = val[m] ? L.nsy * (long long)(j + 1) + (i + 1) : 0;
        }
        for (int p = -1; p <= nk; ++p) {
            const double* sp = xin + L.nsz * (long long)(kb + p + 1);
            double* dp = slot(p);
#pragma unroll
            for (int m = 0; m < 3; ++m) if (dsto[m] >= 0) np_cp_async8(dp + dsto[m], sp + srco[m], val[m]);
        }
        np_cp_async_commit();
    }
    // right-hand side of this thread's nodes, one plane ahead
    double rh[4] = {0.0, 0.0, 0.0, 0.0};
    auto load_rhs = [&](int pn) {
        const long long o = np_noff(L, gi, gj, kb + pn);
#pragma unroll
        for (int q = 0; q < 4; ++q) if ((act >> q) & 1u) rh[q] = rhs[o + L.nsy * (q >> 1) + (q & 1)];
    };
    if (act != 0u) load_rhs(0);
    np_cp_async_wait_all();
    __syncthreads();

    double xn[4] = {0.0, 0.0, 0.0, 0.0};
    for (int t = 0; t <= nk + 1; ++t) {
        const int p = t - 2 * grp;
        const bool on = (p >= 0 && p <= nk - 1) && act != 0u;
        const int k = kb + p;
        double* Pc = on ? slot(p) : ring;
        const bool zdir = on && ((L.face[2] == NP_FACE_DIRICHLET && k == 0) || (L.face[5] == NP_FACE_DIRICHLET && k == L.n[2]));
        const unsigned dmask = zdir ? act : dir;
        const long long gbase = on ? np_noff(L, gi, gj, k) : 0;

        // ---- B = right-hand side minus the out-of-plane part (planes p-1 and p+1 do not change during this step) ----
        double B[4] = {0.0, 0.0, 0.0, 0.0};
        double V[4][4];                                       // variable sigma: the four cell-column sums around node q
        if (on) {
            const double* Pm = slot(p - 1); const double* Pp = slot(p + 1);
            if (CS) {
                double S[4][4];                               // x[k-1] + x[k+1] on (li0-1 .. li0+2) x (lj0-1 .. lj0+2)
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int q = 0; q < 4; ++q) { const int o = gss_at(li0 - 1 + q, lj0 - 1 + r); S[r][q] = Pm[o] + Pp[o]; }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int a = q & 1, b = q >> 1;
                    if ((act >> q) & 1u) {
                        const double rv = rh[q] - roff;
                        const double oop = C.a_fz * S[b + 1][a + 1] + C.a_exz * (S[b + 1][a] + S[b + 1][a + 2]) + C.a_eyz * (S[b][a + 1] + S[b + 2][a + 1])
                                         + C.a_c * ((S[b][a] + S[b][a + 2]) + (S[b + 2][a] + S[b + 2][a + 2]));
                        B[q] = fma(rv, C.dinv, -oop);
                    }
                }
            } else {
                // cells (gi-1 .. gi+1, gj-1 .. gj+1) of the layers below (k-1) and above (k) plane k
                double Sl[2][3][3];
#pragma unroll
                for (int z = 0; z < 2; ++z)
#pragma unroll
                    for (int r = 0; r < 3; ++r)
#pragma unroll
                        for (int q = 0; q < 3; ++q) {
                            const int ci = min(gi - 1 + q, L.n[0]), cj = min(gj - 1 + r, L.n[1]);      // clamped cells only feed nodes outside the level
                            Sl[z][r][q] = sig[np_coff(L, ci, cj, k - 1 + z)];
                        }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int a = q & 1, b = q >> 1;
                    V[q][0] = Sl[0][b][a] + Sl[1][b][a];         V[q][1] = Sl[0][b][a + 1] + Sl[1][b][a + 1];
                    V[q][2] = Sl[0][b + 1][a] + Sl[1][b + 1][a]; V[q][3] = Sl[0][b + 1][a + 1] + Sl[1][b + 1][a + 1];
                    if ((act >> q) & 1u) {
                        double bv = rh[q] - roff;
#pragma unroll
                        for (int z = 0; z < 2; ++z) {
                            const double* P = z ? Pp : Pm;
                            const double l00 = Sl[z][b][a], l10 = Sl[z][b][a + 1], l01 = Sl[z][b + 1][a], l11 = Sl[z][b + 1][a + 1];
                            const int li = li0 + a, lj = lj0 + b;
                            bv -= c.fm2xm2y4z * (P[gss_at(li, lj)] * ((l00 + l10) + (l01 + l11)))
                                + c.f2xmy2z * (P[gss_at(li - 1, lj)] * (l00 + l01) + P[gss_at(li + 1, lj)] * (l10 + l11))
                                + c.fmx2y2z * (P[gss_at(li, lj - 1)] * (l00 + l10) + P[gss_at(li, lj + 1)] * (l01 + l11))
                                + c.fxyz * ((P[gss_at(li - 1, lj - 1)] * l00 + P[gss_at(li + 1, lj - 1)] * l10)
                                          + (P[gss_at(li - 1, lj + 1)] * l01 + P[gss_at(li + 1, lj + 1)] * l11));
                        }
                        B[q] = bv;
                    }
                }
            }
        }

        if (act != 0u && p + 1 >= 0 && p + 1 <= nk - 1) load_rhs(p + 1);      // rh is free now: the next plane's values fly across the colour steps
        // ---- the four colour steps: colour q is node q of every block ----
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            if (on && ((act >> q) & 1u)) {
                const int li = li0 + (q & 1), lj = lj0 + (q >> 1);
                const double xW = Pc[gss_at(li - 1, lj)], xE = Pc[gss_at(li + 1, lj)], xS = Pc[gss_at(li, lj - 1)], xN = Pc[gss_at(li, lj + 1)];
                const double xSW = Pc[gss_at(li - 1, lj - 1)], xSE = Pc[gss_at(li + 1, lj - 1)], xNW = Pc[gss_at(li - 1, lj + 1)], xNE = Pc[gss_at(li + 1, lj + 1)];
                double v;
                if (CS) v = fma(-C.a_exy, (xSW + xSE) + (xNW + xNE), fma(-C.a_fy, xS + xN, fma(-C.a_fx, xW + xE, B[q])));
                else    v = gs_upd_var(c, B[q], V[q][0], V[q][1], V[q][2], V[q][3], xW, xE, xS, xN, xSW, xSE, xNW, xNE);
                if ((dmask >> q) & 1u) v = 0.0;
                xn[q] = v;
                Pc[gss_at(li, lj)] = v;
            }
            // colours 0 -> 1 and 2 -> 3 only exchange values inside a row of blocks, i.e. inside a warp
            if (q == 0 || q == 2) __syncwarp();
            else if (q == 1) np_group_barrier512(grp);
        }
        __syncthreads();
        if (on && grp == 1 && p < nkw) {
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if ((wrt >> q) & 1u) {
                    const long long o = gbase + L.nsy * (q >> 1) + (q & 1);
                    xout[o] = A.addto ? A.addto[o] + xn[q] : xn[q];
                }
        }
    }
}
