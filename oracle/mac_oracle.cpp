// ORACLE — TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED UPSTREAM.
//
// CPU restatement of the reference's MAC-projection path:
//     incflo::compute_MAC_projected_velocities -> amrex::MacProjector::project -> amrex::MLMG::solve
//     on amrex::MLABecLaplacian (alpha = 0, beta = 1, face-centred b = 1/rho, 3D, no EB, one AMR level, one box).
// Reference call sites: /root/reference/src/convection/incflo_compute_MAC_projected_velocities.cpp:36-46
// (initProjector / setDomainBC / updateBeta), :115 (setUMAC), :127-131 (project); options src/incflo.H:330-344.
// The arithmetic lives in AMReX (>= 20.11), which is neither in /root/reference nor on this machine, and the
// reference's tests hold no golden vectors at this boundary: "parity unpinned".  This file follows the published
// AMReX algorithm function by function (names in the comments) and is pinned by self-made identities
// (tests/test_mac_oracle_pins.py: assembled 7-point finite-volume matrix incl. boundary rows, sparse direct solve,
// discrete divergence of the result).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
//
// Layout: cell arrays carry one ghost layer, x fastest: (i,j,k) -> (i+1) + sx*((j+1) + sy*(k+1)), sx = nx+2.
// Face arrays: bx (nx+1, ny, nz), by (nx, ny+1, nz), bz (nx, ny, nz+1), x fastest, no ghosts.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

namespace mor {

enum { PERIODIC = 0, DIRICHLET = 1, NEUMANN = 2 };
enum { BICGCG = 0, BICGSTAB = 1, CG = 2, SMOOTHER = 3 };
static int g_nthreads = 1;

static void pfor(int n, long long work_per_item, const std::function<void(int)>& f)
{
    const int nt = std::min(g_nthreads, n);
    if (nt <= 1 || work_per_item * n < 200000) { for (int i = 0; i < n; ++i) f(i); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < nt; ++t) th.emplace_back([&, t]() { for (int i = t; i < n; i += nt) f(i); });
    for (auto& x : th) x.join();
}

struct Level {
    int n[3]; double dxinv[3];
    long long sx, sy, ntot;
    std::vector<double> bx, by, bz;
    std::vector<double> a;      // cell coefficient of the alpha term (MLABecLaplacian::setACoeffs), empty: none
    long long c(int i, int j, int k) const { return (i + 1) + sx * ((j + 1) + sy * (long long)(k + 1)); }
    long long fx(int i, int j, int k) const { return i + (long long)(n[0] + 1) * (j + (long long)n[1] * k); }
    long long fy(int i, int j, int k) const { return i + (long long)n[0] * (j + (long long)(n[1] + 1) * k); }
    long long fz(int i, int j, int k) const { return i + (long long)n[0] * (j + (long long)n[1] * k); }
    std::vector<double> make() const { return std::vector<double>(ntot, 0.0); }
};

struct Options {
    int max_coarsening_level = 100, maxiter = 200, bottom_maxiter = 200, nu1 = 2, nu2 = 2, nuf = 8, nub = 0;
    int bottom_solver = BICGCG; double bottom_rtol = 1e-4, bottom_atol = -1.0; int verbose = 0, maxorder = 3;
};
struct Stats {
    int n_vcycles = 0, bottom_iters_total = 0, status = 0, n_levels = 0;
    double rhs_norm0 = 0, res_norm0 = 0, res_norm_final = 0, res_history[128] = {0};
    double t_setup = 0, t_solve = 0, t_total = 0;
};

typedef std::vector<double> V;

struct Solver {
    Options opt; std::vector<Level> lev; int bclo[3], bchi[3]; bool singular = true;
    double alpha = 0.0, bscal = 1.0;      // MLABecLaplacian::setScalars(alpha, beta): (alpha a - beta div b grad) phi = rhs
    std::vector<V> res, cor, rescor;
    V sol, rhs, umac[3];
    int bottom_iters = 0; bool bicg_failed_switch = false;

    void define(const int* ncell, const double* plo, const double* phi, const int* lo, const int* hi, const Options& o)
    {
        opt = o;
        for (int d = 0; d < 3; ++d) { bclo[d] = lo[d]; bchi[d] = hi[d]; if (lo[d] == DIRICHLET || hi[d] == DIRICHLET) singular = false; }
        int n[3] = {ncell[0], ncell[1], ncell[2]}; double dxi[3];
        for (int d = 0; d < 3; ++d) dxi[d] = ncell[d] / (phi[d] - plo[d]);
        for (;;) {      // MLLinOp::defineGrids: coarsen by 2 while the box stays coarsenable with min width 2
            Level L; for (int d = 0; d < 3; ++d) { L.n[d] = n[d]; L.dxinv[d] = dxi[d]; }
            L.sx = n[0] + 2; L.sy = n[1] + 2; L.ntot = L.sx * L.sy * (n[2] + 2);
            L.bx.assign((size_t)(n[0] + 1) * n[1] * n[2], 1.0); L.by.assign((size_t)n[0] * (n[1] + 1) * n[2], 1.0); L.bz.assign((size_t)n[0] * n[1] * (n[2] + 1), 1.0);
            lev.push_back(L);
            if ((int)lev.size() > opt.max_coarsening_level) break;
            bool ok = true; for (int d = 0; d < 3; ++d) if (n[d] % 2 || n[d] < 4) ok = false;
            if (!ok) break;
            for (int d = 0; d < 3; ++d) { n[d] /= 2; dxi[d] *= 0.5; }
        }
        for (auto& L : lev) { res.push_back(L.make()); cor.push_back(L.make()); rescor.push_back(L.make()); }
        sol = lev[0].make(); rhs = lev[0].make();
        const Level& L0 = lev[0];
        umac[0].assign(L0.bx.size(), 0.0); umac[1].assign(L0.by.size(), 0.0); umac[2].assign(L0.bz.size(), 0.0);
    }

    // MLCellLinOp::applyBC -> FillBoundary (periodic images) + mllinop_apply_bc_{x,y,z}, homogeneous mode: Neumann ghost =
    // first interior cell; Dirichlet ghost = polynomial through the wall value (0) at the face and the first
    // min(blen+1, maxorder)-1 interior cells, evaluated at the ghost centre: maxorder 3 -> -2 phi0 + phi1/3, 2 -> -phi0.
    void apply_bc(int l, V& x) const
    {
        const Level& L = lev[l];
        for (int d = 0; d < 3; ++d) {
            const int a = (d + 1) % 3, b = (d + 2) % 3;
            const int nd = L.n[d];
            const int nx_ = std::min(nd + 1, opt.maxorder);
            const double c0 = nx_ >= 3 ? -2.0 : -1.0, c1 = nx_ >= 3 ? 1.0 / 3.0 : 0.0;
            pfor(L.n[b], L.n[a], [&](int q) {
                for (int p = 0; p < L.n[a]; ++p) {
                    int idx[3];
                    auto at = [&](int v) { idx[d] = v; idx[a] = p; idx[b] = q; return L.c(idx[0], idx[1], idx[2]); };
                    if (bclo[d] == PERIODIC) { x[at(-1)] = x[at(nd - 1)]; x[at(nd)] = x[at(0)]; continue; }
                    x[at(-1)] = bclo[d] == NEUMANN ? x[at(0)] : c0 * x[at(0)] + c1 * x[at(1)];
                    x[at(nd)] = bchi[d] == NEUMANN ? x[at(nd - 1)] : c0 * x[at(nd - 1)] + c1 * x[at(nd - 2)];
                }
            });
        }
    }
    // coefficient of the first interior cell in its ghost cell (MLCellLinOp's "undrrelxr" factors, mllinop_comp_interp_coef0):
    // used by the smoother to take the ghost's dependence on the cell out of the diagonal.  0 on periodic faces.
    double bc_fac(int d, int side, int nd) const
    {
        const int t = side ? bchi[d] : bclo[d];
        if (t == PERIODIC) return 0.0;
        if (t == NEUMANN) return 1.0;
        return std::min(nd + 1, opt.maxorder) >= 3 ? -2.0 : -1.0;
    }

    // MLABecLaplacian::Fapply -> mlabeclap_adotx with alpha = 0, beta = 1: y = -div(b grad x); x's ghosts are filled.
    void adotx(int l, V& y, const V& x) const
    {
        const Level& L = lev[l];
        const double dhx = bscal * L.dxinv[0] * L.dxinv[0], dhy = bscal * L.dxinv[1] * L.dxinv[1], dhz = bscal * L.dxinv[2] * L.dxinv[2];
        const bool ha = alpha != 0.0 && !L.a.empty();
        pfor(L.n[2], (long long)L.n[0] * L.n[1], [&](int k) {
            for (int j = 0; j < L.n[1]; ++j) for (int i = 0; i < L.n[0]; ++i) {
                const long long c = L.c(i, j, k);
                y[c] = (ha ? alpha * L.a[c] * x[c] : 0.0) - dhx * (L.bx[L.fx(i + 1, j, k)] * (x[c + 1] - x[c]) - L.bx[L.fx(i, j, k)] * (x[c] - x[c - 1]))
                       - dhy * (L.by[L.fy(i, j + 1, k)] * (x[c + L.sx] - x[c]) - L.by[L.fy(i, j, k)] * (x[c] - x[c - L.sx]))
                       - dhz * (L.bz[L.fz(i, j, k + 1)] * (x[c + L.sx * L.sy] - x[c]) - L.bz[L.fz(i, j, k)] * (x[c] - x[c - L.sx * L.sy]));
            }
        });
    }
    void apply(int l, V& y, V& x) const { apply_bc(l, x); adotx(l, y, x); }
    void residual(int l, V& r, V& x, const V& b) const
    {
        apply(l, r, x);
        const Level& L = lev[l];
        for (int k = 0; k < L.n[2]; ++k) for (int j = 0; j < L.n[1]; ++j) for (int i = 0; i < L.n[0]; ++i) { const long long c = L.c(i, j, k); r[c] = b[c] - r[c]; }
    }
    // MLABecLaplacian::Fsmooth -> abec_gsrb: cells with (i+j+k+redblack) even, over-relaxation omega = 1.15, the boundary
    // factors taken out of the diagonal (gamma - delta).
    void gsrb(int l, V& phi, const V& b, int redblack) const
    {
        const Level& L = lev[l];
        const double dhx = bscal * L.dxinv[0] * L.dxinv[0], dhy = bscal * L.dxinv[1] * L.dxinv[1], dhz = bscal * L.dxinv[2] * L.dxinv[2];
        const bool ha = alpha != 0.0 && !L.a.empty();
        const double omega = 1.15;
        const double f0 = bc_fac(0, 0, L.n[0]), f3 = bc_fac(0, 1, L.n[0]), f1 = bc_fac(1, 0, L.n[1]), f4 = bc_fac(1, 1, L.n[1]),
                     f2 = bc_fac(2, 0, L.n[2]), f5 = bc_fac(2, 1, L.n[2]);
        const long long sz = L.sx * L.sy;
        pfor(L.n[2], (long long)L.n[0] * L.n[1] / 2, [&](int k) {
            for (int j = 0; j < L.n[1]; ++j) for (int i = (j + k + redblack) % 2; i < L.n[0]; i += 2) {
                const long long c = L.c(i, j, k);
                const double bxm = L.bx[L.fx(i, j, k)], bxp = L.bx[L.fx(i + 1, j, k)], bym = L.by[L.fy(i, j, k)], byp = L.by[L.fy(i, j + 1, k)],
                             bzm = L.bz[L.fz(i, j, k)], bzp = L.bz[L.fz(i, j, k + 1)];
                const double cf0 = i == 0 ? f0 : 0.0, cf3 = i == L.n[0] - 1 ? f3 : 0.0, cf1 = j == 0 ? f1 : 0.0, cf4 = j == L.n[1] - 1 ? f4 : 0.0,
                             cf2 = k == 0 ? f2 : 0.0, cf5 = k == L.n[2] - 1 ? f5 : 0.0;
                const double gamma = (ha ? alpha * L.a[c] : 0.0) + dhx * (bxm + bxp) + dhy * (bym + byp) + dhz * (bzm + bzp);
                const double g_m_d = gamma - (dhx * (bxm * cf0 + bxp * cf3) + dhy * (bym * cf1 + byp * cf4) + dhz * (bzm * cf2 + bzp * cf5));
                const double rho = dhx * (bxm * phi[c - 1] + bxp * phi[c + 1]) + dhy * (bym * phi[c - L.sx] + byp * phi[c + L.sx])
                                 + dhz * (bzm * phi[c - sz] + bzp * phi[c + sz]);
                const double r = b[c] - (gamma * phi[c] - rho);
                phi[c] += omega / g_m_d * r;
            }
        });
    }
    // MLCellLinOp::smooth: for each colour applyBC then Fsmooth
    void smooth(int l, V& x, const V& b, bool skip_fillboundary) const
    {
        for (int rb = 0; rb < 2; ++rb) { if (!(skip_fillboundary && rb == 0)) apply_bc(l, x); gsrb(l, x, b, rb); }
    }
    // MLCellLinOp::restriction -> amrex::average_down: mean of the 8 children
    void restriction(int lc, V& crse, const V& fine) const
    {
        const Level &Lc = lev[lc], &Lf = lev[lc - 1];
        pfor(Lc.n[2], (long long)Lc.n[0] * Lc.n[1], [&](int k) {
            for (int j = 0; j < Lc.n[1]; ++j) for (int i = 0; i < Lc.n[0]; ++i) {
                double s = 0; for (int c = 0; c < 2; ++c) for (int b = 0; b < 2; ++b) for (int a = 0; a < 2; ++a) s += fine[Lf.c(2 * i + a, 2 * j + b, 2 * k + c)];
                crse[Lc.c(i, j, k)] = 0.125 * s;
            }
        });
    }
    // MLCellLinOp::interpolation -> mlmg_lin_cc_interp_r2: piecewise constant, fine += crse(parent)
    void interp_add(int lf, V& fine, const V& crse) const
    {
        const Level &Lf = lev[lf], &Lc = lev[lf + 1];
        pfor(Lf.n[2], (long long)Lf.n[0] * Lf.n[1], [&](int k) {
            for (int j = 0; j < Lf.n[1]; ++j) for (int i = 0; i < Lf.n[0]; ++i) fine[Lf.c(i, j, k)] += crse[Lc.c(i / 2, j / 2, k / 2)];
        });
    }
    // MLABecLaplacian::averageDownCoeffs -> amrex::average_down_faces: mean of the 4 fine faces on a coarse face
    void average_down_beta()
    {
        for (size_t l = 1; l < lev.size(); ++l) {
            Level& C = lev[l]; const Level& F = lev[l - 1];
            if (!F.a.empty()) { C.a = C.make(); restriction((int)l, C.a, F.a); }      // MLABecLaplacian::averageDownCoeffs: a by average_down
            for (int k = 0; k < C.n[2]; ++k) for (int j = 0; j < C.n[1]; ++j) for (int i = 0; i <= C.n[0]; ++i) {
                double s = 0; for (int c = 0; c < 2; ++c) for (int b = 0; b < 2; ++b) s += F.bx[F.fx(2 * i, 2 * j + b, 2 * k + c)];
                C.bx[C.fx(i, j, k)] = 0.25 * s;
            }
            for (int k = 0; k < C.n[2]; ++k) for (int j = 0; j <= C.n[1]; ++j) for (int i = 0; i < C.n[0]; ++i) {
                double s = 0; for (int c = 0; c < 2; ++c) for (int a = 0; a < 2; ++a) s += F.by[F.fy(2 * i + a, 2 * j, 2 * k + c)];
                C.by[C.fy(i, j, k)] = 0.25 * s;
            }
            for (int k = 0; k <= C.n[2]; ++k) for (int j = 0; j < C.n[1]; ++j) for (int i = 0; i < C.n[0]; ++i) {
                double s = 0; for (int b = 0; b < 2; ++b) for (int a = 0; a < 2; ++a) s += F.bz[F.fz(2 * i + a, 2 * j + b, 2 * k)];
                C.bz[C.fz(i, j, k)] = 0.25 * s;
            }
        }
    }
    // MLABecLaplacian::normalize -> mlabeclap_normalize: divide by alpha*a + dhx*(bX(i)+bX(i+1)) + ... (MLCGSolver)
    void normalize(int l, V& x) const
    {
        const Level& L = lev[l];
        const double dhx = bscal * L.dxinv[0] * L.dxinv[0], dhy = bscal * L.dxinv[1] * L.dxinv[1], dhz = bscal * L.dxinv[2] * L.dxinv[2];
        const bool ha = alpha != 0.0 && !L.a.empty();
        for (int k = 0; k < L.n[2]; ++k) for (int j = 0; j < L.n[1]; ++j) for (int i = 0; i < L.n[0]; ++i)
            x[L.c(i, j, k)] /= (ha ? alpha * L.a[L.c(i, j, k)] : 0.0) + dhx * (L.bx[L.fx(i, j, k)] + L.bx[L.fx(i + 1, j, k)]) + dhy * (L.by[L.fy(i, j, k)] + L.by[L.fy(i, j + 1, k)])
                             + dhz * (L.bz[L.fz(i, j, k)] + L.bz[L.fz(i, j, k + 1)]);
    }
    template <class F> void for_valid(const Level& L, F f) const
    {
        for (int k = 0; k < L.n[2]; ++k) for (int j = 0; j < L.n[1]; ++j) for (int i = 0; i < L.n[0]; ++i) f(L.c(i, j, k));
    }
    double dot(const Level& L, const V& x, const V& y) const { double s = 0; for_valid(L, [&](long long c) { s += x[c] * y[c]; }); return s; }
    double norm_inf(const Level& L, const V& x) const { double s = 0; for_valid(L, [&](long long c) { s = std::max(s, std::fabs(x[c])); }); return s; }
    // MLMG::makeSolvable (cell-centred): subtract the mean
    void make_solvable(const Level& L, V& r) const
    {
        double s = 0; for_valid(L, [&](long long c) { s += r[c]; });
        const double off = s / ((double)L.n[0] * L.n[1] * L.n[2]);
        for_valid(L, [&](long long c) { r[c] -= off; });
    }

    // ---- MLCGSolver ----
    int solve_bicgstab(int l, V& solx, const V& b, double eps_rel, double eps_abs, int& niters)
    {
        const Level& L = lev[l];
        V ph = L.make(), sh = L.make(), p = L.make(), r = L.make(), s = L.make(), rh, v = L.make(), t = L.make();
        residual(l, r, solx, b); normalize(l, r);
        V sorig = solx; rh = r; std::fill(solx.begin(), solx.end(), 0.0);
        double rnorm = norm_inf(L, r); const double rnorm0 = rnorm;
        niters = 0;
        if (rnorm0 == 0 || rnorm0 < eps_abs) return 0;
        int ret = 0, iter = 1; double rho_1 = 0, alpha = 0, omega = 0;
        for (; iter <= opt.bottom_maxiter; ++iter) {
            const double rho = dot(L, rh, r);
            if (rho == 0) { ret = 1; break; }
            if (iter == 1) p = r;
            else { const double beta = (rho / rho_1) * (alpha / omega); for_valid(L, [&](long long c) { p[c] = r[c] + beta * (p[c] - omega * v[c]); }); }
            ph = p;
            apply(l, v, ph); normalize(l, v);
            const double rhTv = dot(L, rh, v);
            if (rhTv != 0) alpha = rho / rhTv; else { ret = 2; break; }
            for_valid(L, [&](long long c) { solx[c] += alpha * ph[c]; s[c] = r[c] - alpha * v[c]; });
            rnorm = norm_inf(L, s);
            if (rnorm < eps_rel * rnorm0 || rnorm < eps_abs) break;
            sh = s;
            apply(l, t, sh); normalize(l, t);
            const double tt = dot(L, t, t), ts = dot(L, t, s);
            if (tt != 0) omega = ts / tt; else { ret = 3; break; }
            for_valid(L, [&](long long c) { solx[c] += omega * sh[c]; r[c] = s[c] - omega * t[c]; });
            rnorm = norm_inf(L, r);
            if (rnorm < eps_rel * rnorm0 || rnorm < eps_abs) break;
            if (omega == 0) { ret = 4; break; }
            rho_1 = rho;
        }
        niters = std::min(iter, opt.bottom_maxiter);
        if (ret == 0 && rnorm > eps_rel * rnorm0 && rnorm > eps_abs) ret = 8;
        if ((ret == 0 || ret == 8) && rnorm < rnorm0) for_valid(L, [&](long long c) { solx[c] += sorig[c]; });
        else solx = sorig;
        return ret;
    }
    int solve_cg(int l, V& solx, const V& b, double eps_rel, double eps_abs, int& niters)
    {
        const Level& L = lev[l];
        V p = L.make(), r = L.make(), q = L.make();
        residual(l, r, solx, b);
        V sorig = solx; std::fill(solx.begin(), solx.end(), 0.0);
        double rnorm = norm_inf(L, r); const double rnorm0 = rnorm;
        niters = 0;
        if (rnorm0 == 0 || rnorm0 < eps_abs) return 0;
        int ret = 0, iter = 1; double rho_1 = 0;
        for (; iter <= opt.bottom_maxiter; ++iter) {
            const double rho = dot(L, r, r);
            if (rho == 0) { ret = 1; break; }
            if (iter == 1) p = r; else { const double beta = rho / rho_1; for_valid(L, [&](long long c) { p[c] = r[c] + beta * p[c]; }); }
            apply(l, q, p);
            const double pw = dot(L, p, q); double alpha;
            if (pw != 0) alpha = rho / pw; else { ret = 1; break; }
            for_valid(L, [&](long long c) { solx[c] += alpha * p[c]; r[c] -= alpha * q[c]; });
            rnorm = norm_inf(L, r);
            if (rnorm < eps_rel * rnorm0 || rnorm < eps_abs) break;
            rho_1 = rho;
        }
        niters = std::min(iter, opt.bottom_maxiter);
        if (ret == 0 && rnorm > eps_rel * rnorm0 && rnorm > eps_abs) ret = 8;
        if ((ret == 0 || ret == 8) && rnorm < rnorm0) for_valid(L, [&](long long c) { solx[c] += sorig[c]; });
        else solx = sorig;
        return ret;
    }

    // ---- MLMG ----
    void bottom_solve()
    {
        const int l = (int)lev.size() - 1;
        V& x = cor[l]; V& b = res[l];
        std::fill(x.begin(), x.end(), 0.0);
        if (opt.bottom_solver == SMOOTHER) { bool skip = true; for (int i = 0; i < opt.nuf; ++i) { smooth(l, x, b, skip); skip = false; } return; }
        V bb = b;
        if (singular) make_solvable(lev[l], bb);
        const int type = (opt.bottom_solver == CG || (opt.bottom_solver == BICGCG && bicg_failed_switch)) ? CG : BICGSTAB;
        int nit = 0;
        auto run = [&](int ty) { return ty == CG ? solve_cg(l, x, bb, opt.bottom_rtol, opt.bottom_atol, nit) : solve_bicgstab(l, x, bb, opt.bottom_rtol, opt.bottom_atol, nit); };
        int ret = run(type);
        bottom_iters += nit;
        if (ret != 0 && ret != 9) {
            std::fill(x.begin(), x.end(), 0.0);
            if (opt.bottom_solver == BICGCG) {
                const int other = type == CG ? BICGSTAB : CG;
                ret = run(other); bottom_iters += nit;
                if (ret == 0) bicg_failed_switch = (other == CG); else std::fill(x.begin(), x.end(), 0.0);
            }
        }
        const int n = ret == 0 ? opt.nub : opt.nuf;
        for (int i = 0; i < n; ++i) smooth(l, x, b, false);
    }
    void vcycle()
    {
        const int nl = (int)lev.size();
        for (int l = 0; l < nl - 1; ++l) {
            std::fill(cor[l].begin(), cor[l].end(), 0.0);
            bool skip = true;
            for (int i = 0; i < opt.nu1; ++i) { smooth(l, cor[l], res[l], skip); skip = false; }
            residual(l, rescor[l], cor[l], res[l]);            // computeResOfCorrection
            restriction(l + 1, res[l + 1], rescor[l]);
        }
        bottom_solve();
        for (int l = nl - 2; l >= 0; --l) {
            interp_add(l, cor[l], cor[l + 1]);                 // addInterpCorrection
            for (int i = 0; i < opt.nu2; ++i) smooth(l, cor[l], res[l], false);
        }
    }
    // MLMG::solve; sol holds the initial guess (MacProjector::project: 0, or mac_phi)
    void solve(double rtol, double atol, Stats& st)
    {
        const Level& L0 = lev[0];
        bottom_iters = 0; bicg_failed_switch = false;
        if (singular) make_solvable(L0, rhs);
        residual(0, res[0], sol, rhs);
        const double resnorm0 = norm_inf(L0, res[0]), rhsnorm0 = norm_inf(L0, rhs);
        const double max_norm = rhsnorm0 >= resnorm0 ? rhsnorm0 : resnorm0;
        const double res_target = std::max(atol, std::max(rtol, 1e-16) * max_norm);
        st.rhs_norm0 = rhsnorm0; st.res_norm0 = resnorm0; st.n_vcycles = 0; st.status = 0; st.res_norm_final = resnorm0;
        st.n_levels = (int)lev.size();
        if (resnorm0 <= res_target) return;
        bool converged = false;
        for (int iter = 0; iter < opt.maxiter; ++iter) {
            if (singular) make_solvable(L0, res[0]);
            vcycle();
            for_valid(L0, [&](long long c) { sol[c] += cor[0][c]; });
            residual(0, res[0], sol, rhs);
            const double rn = norm_inf(L0, res[0]);
            if (iter < 128) st.res_history[iter] = rn;
            st.n_vcycles = iter + 1; st.res_norm_final = rn;
            if (opt.verbose >= 2) std::printf("MLMG: Iteration %3d Fine resid/bnorm = %.6e\n", iter + 1, rn / max_norm);
            if (rn <= res_target) { converged = true; break; }
            if (rn > 1e20 * max_norm) { st.status = 2; break; }
        }
        if (!converged && st.status == 0) st.status = 1;
        st.bottom_iters_total = bottom_iters;
    }
    // amrex::computeDivergence of the face velocities
    void divergence(V& d) const
    {
        const Level& L = lev[0];
        for (int k = 0; k < L.n[2]; ++k) for (int j = 0; j < L.n[1]; ++j) for (int i = 0; i < L.n[0]; ++i)
            d[L.c(i, j, k)] = L.dxinv[0] * (umac[0][L.fx(i + 1, j, k)] - umac[0][L.fx(i, j, k)]) + L.dxinv[1] * (umac[1][L.fy(i, j + 1, k)] - umac[1][L.fy(i, j, k)])
                            + L.dxinv[2] * (umac[2][L.fz(i, j, k + 1)] - umac[2][L.fz(i, j, k)]);
    }
    // MacProjector::project: rhs = -div(umac); phi = 0 (or the caller's mac_phi); MLMG::solve; getFluxes (mlabeclap_flux:
    // -b dphi/dn on every face, ghosts from applyBC in solution mode = homogeneous here); umac += flux.
    void project(double rtol, double atol, Stats& st, bool have_guess)
    {
        auto t0 = std::chrono::steady_clock::now();
        average_down_beta();
        const Level& L = lev[0];
        divergence(rhs);
        for_valid(L, [&](long long c) { rhs[c] = -rhs[c]; });
        if (!have_guess) std::fill(sol.begin(), sol.end(), 0.0);
        auto t1 = std::chrono::steady_clock::now();
        solve(rtol, atol, st);
        auto t2 = std::chrono::steady_clock::now();
        apply_bc(0, sol);
        for (int k = 0; k < L.n[2]; ++k) for (int j = 0; j < L.n[1]; ++j) for (int i = 0; i <= L.n[0]; ++i)
            umac[0][L.fx(i, j, k)] -= L.bx[L.fx(i, j, k)] * L.dxinv[0] * (sol[L.c(i, j, k)] - sol[L.c(i - 1, j, k)]);
        for (int k = 0; k < L.n[2]; ++k) for (int j = 0; j <= L.n[1]; ++j) for (int i = 0; i < L.n[0]; ++i)
            umac[1][L.fy(i, j, k)] -= L.by[L.fy(i, j, k)] * L.dxinv[1] * (sol[L.c(i, j, k)] - sol[L.c(i, j - 1, k)]);
        for (int k = 0; k <= L.n[2]; ++k) for (int j = 0; j < L.n[1]; ++j) for (int i = 0; i < L.n[0]; ++i)
            umac[2][L.fz(i, j, k)] -= L.bz[L.fz(i, j, k)] * L.dxinv[2] * (sol[L.c(i, j, k)] - sol[L.c(i, j, k - 1)]);
        auto t3 = std::chrono::steady_clock::now();
        st.t_setup = std::chrono::duration<double>(t1 - t0).count();
        st.t_solve = std::chrono::duration<double>(t2 - t1).count();
        st.t_total = std::chrono::duration<double>(t3 - t0).count();
    }
    // A general MLABecLaplacian solve (the scalar diffusion solves of the reference, SURVEY.md 8(f) row 3:
    // /root/reference/src/diffusion/DiffusionScalarOp.cpp:189-250 - alpha = 1, beta = dt, a = rho, b = eta on faces,
    // setLevelBC(phi) with the Dirichlet values in phi's ghost cells).  `sol` holds the initial guess, `rhs` the right-hand
    // side; bcv[f] (f = x-lo, y-lo, z-lo, x-hi, y-hi, z-hi) the Dirichlet values on the faces of side f or null (0).
    // MLMG evaluates the first residual with the inhomogeneous ghost  c0 phi0 + c1 phi1 + c_bc value  and iterates on the
    // homogeneous correction equation; moving the boundary term to the right-hand side once is the same thing.
    void solve_abec(double rtol, double atol, Stats& st, const double* const bcv[6])
    {
        singular = true;
        for (int d = 0; d < 3; ++d) if (bclo[d] == DIRICHLET || bchi[d] == DIRICHLET) singular = false;
        if (alpha != 0.0 && !lev[0].a.empty()) singular = false;
        average_down_beta();
        const Level& L = lev[0];
        for (int d = 0; d < 3; ++d) for (int side = 0; side < 2; ++side) {
            const double* v = bcv[d + 3 * side];
            if (!v || (side ? bchi[d] : bclo[d]) != DIRICHLET) continue;
            const int a = (d + 1) % 3, b = (d + 2) % 3;
            const double cbc = std::min(L.n[d] + 1, opt.maxorder) >= 3 ? 8.0 / 3.0 : 2.0;      // weight of the wall value in the ghost
            const double dh = bscal * L.dxinv[d] * L.dxinv[d];
            // face arrays of side values are indexed [q][p] with p the lower of the two transverse directions (x fastest)
            const int t0 = std::min(a, b), t1 = std::max(a, b);
            for (int q = 0; q < L.n[t1]; ++q) for (int p = 0; p < L.n[t0]; ++p) {
                int idx[3]; idx[d] = side ? L.n[d] - 1 : 0; idx[t0] = p; idx[t1] = q;
                int fidx[3] = {idx[0], idx[1], idx[2]}; if (side) fidx[d] += 1;
                const double bf = d == 0 ? L.bx[L.fx(fidx[0], fidx[1], fidx[2])] : d == 1 ? L.by[L.fy(fidx[0], fidx[1], fidx[2])] : L.bz[L.fz(fidx[0], fidx[1], fidx[2])];
                rhs[L.c(idx[0], idx[1], idx[2])] += dh * bf * cbc * v[p + (long long)L.n[t0] * q];
            }
        }
        solve(rtol, atol, st);
    }
    void scatter(int l, V& dst, const double* src) const
    {
        const Level& L = lev[l]; long long q = 0;
        for (int k = 0; k < L.n[2]; ++k) for (int j = 0; j < L.n[1]; ++j) for (int i = 0; i < L.n[0]; ++i) dst[L.c(i, j, k)] = src[q++];
    }
    void gather(int l, const V& src, double* dst) const
    {
        const Level& L = lev[l]; long long q = 0;
        for (int k = 0; k < L.n[2]; ++k) for (int j = 0; j < L.n[1]; ++j) for (int i = 0; i < L.n[0]; ++i) dst[q++] = src[L.c(i, j, k)];
    }
};

} // namespace mor

// ---------------------------------------------------------------------------------------------
// C entry points for ctypes (tests / bench cpu_baseline only).  Global arrays, x fastest, no ghosts.
// ---------------------------------------------------------------------------------------------
extern "C" {
using namespace mor;
struct mor_options { int max_coarsening_level, maxiter, bottom_maxiter, nu1, nu2, nuf, nub, bottom_solver, verbose, maxorder, nthreads; double bottom_rtol, bottom_atol; };
struct mor_stats { int n_vcycles, bottom_iters_total, status, n_levels; double rhs_norm0, res_norm0, res_norm_final; double res_history[128]; double t_setup, t_solve, t_total; };

void mor_default_options(mor_options* o)
{
    Options d; o->max_coarsening_level = d.max_coarsening_level; o->maxiter = d.maxiter; o->bottom_maxiter = d.bottom_maxiter; o->nu1 = d.nu1; o->nu2 = d.nu2;
    o->nuf = d.nuf; o->nub = d.nub; o->bottom_solver = d.bottom_solver; o->verbose = 0; o->maxorder = d.maxorder; o->nthreads = 1; o->bottom_rtol = d.bottom_rtol; o->bottom_atol = d.bottom_atol;
}
void* mor_create(const int* ncell, const double* plo, const double* phi, const int* bclo, const int* bchi, const mor_options* o)
{
    Options op; op.max_coarsening_level = o->max_coarsening_level; op.maxiter = o->maxiter; op.bottom_maxiter = o->bottom_maxiter; op.nu1 = o->nu1; op.nu2 = o->nu2;
    op.nuf = o->nuf; op.nub = o->nub; op.bottom_solver = o->bottom_solver; op.verbose = o->verbose; op.maxorder = o->maxorder; op.bottom_rtol = o->bottom_rtol; op.bottom_atol = o->bottom_atol;
    g_nthreads = std::max(1, o->nthreads);
    auto* s = new Solver(); s->define(ncell, plo, phi, bclo, bchi, op); return s;
}
void mor_destroy(void* h) { delete (Solver*)h; }
int mor_num_levels(void* h) { return (int)((Solver*)h)->lev.size(); }
void mor_level_ncell(void* h, int l, int* n) { for (int d = 0; d < 3; ++d) n[d] = ((Solver*)h)->lev[l].n[d]; }
void mor_set_beta(void* h, const double* bx, const double* by, const double* bz)
{
    auto* s = (Solver*)h; Level& L = s->lev[0];
    std::copy(bx, bx + L.bx.size(), L.bx.begin()); std::copy(by, by + L.by.size(), L.by.begin()); std::copy(bz, bz + L.bz.size(), L.bz.begin());
    s->average_down_beta();
}
void mor_get_beta(void* h, int l, double* bx, double* by, double* bz)
{
    const Level& L = ((Solver*)h)->lev[l];
    std::copy(L.bx.begin(), L.bx.end(), bx); std::copy(L.by.begin(), L.by.end(), by); std::copy(L.bz.begin(), L.bz.end(), bz);
}
void mor_set_umac(void* h, const double* u, const double* v, const double* w)
{
    auto* s = (Solver*)h; const double* src[3] = {u, v, w};
    for (int d = 0; d < 3; ++d) std::copy(src[d], src[d] + s->umac[d].size(), s->umac[d].begin());
}
void mor_get_umac(void* h, double* u, double* v, double* w)
{
    auto* s = (Solver*)h; double* dst[3] = {u, v, w};
    for (int d = 0; d < 3; ++d) std::copy(s->umac[d].begin(), s->umac[d].end(), dst[d]);
}
void mor_set_phi(void* h, const double* p) { auto* s = (Solver*)h; s->scatter(0, s->sol, p); }
void mor_get_phi(void* h, double* p) { auto* s = (Solver*)h; s->gather(0, s->sol, p); }
void mor_project(void* h, double rtol, double atol, int have_guess, mor_stats* out)
{
    auto* s = (Solver*)h; Stats st; s->project(rtol, atol, st, have_guess != 0);
    out->n_vcycles = st.n_vcycles; out->bottom_iters_total = st.bottom_iters_total; out->status = st.status; out->n_levels = st.n_levels;
    out->rhs_norm0 = st.rhs_norm0; out->res_norm0 = st.res_norm0; out->res_norm_final = st.res_norm_final;
    std::memcpy(out->res_history, st.res_history, sizeof(st.res_history));
    out->t_setup = st.t_setup; out->t_solve = st.t_solve; out->t_total = st.t_total;
}
// ---- general ABecLaplacian solve (scalar diffusion; CPU groundwork only, no CUDA counterpart yet) ----
void mor_set_scalars(void* h, double alpha, double beta) { auto* s = (Solver*)h; s->alpha = alpha; s->bscal = beta; }
void mor_set_acoef(void* h, const double* a) { auto* s = (Solver*)h; s->lev[0].a = s->lev[0].make(); s->scatter(0, s->lev[0].a, a); }
void mor_set_rhs(void* h, const double* r) { auto* s = (Solver*)h; s->scatter(0, s->rhs, r); }
void mor_solve_abec(void* h, double rtol, double atol, const double* xlo, const double* ylo, const double* zlo, const double* xhi,
                    const double* yhi, const double* zhi, mor_stats* out)
{
    auto* s = (Solver*)h; Stats st; const double* bcv[6] = {xlo, ylo, zlo, xhi, yhi, zhi};
    s->solve_abec(rtol, atol, st, bcv);
    out->n_vcycles = st.n_vcycles; out->bottom_iters_total = st.bottom_iters_total; out->status = st.status; out->n_levels = st.n_levels;
    out->rhs_norm0 = st.rhs_norm0; out->res_norm0 = st.res_norm0; out->res_norm_final = st.res_norm_final;
    std::memcpy(out->res_history, st.res_history, sizeof(st.res_history));
}
// kernel-level hooks
void mor_apply(void* h, int l, const double* x, double* y)
{
    auto* s = (Solver*)h; V X = s->lev[l].make(), Y = s->lev[l].make(); s->scatter(l, X, x); s->apply(l, Y, X); s->gather(l, Y, y);
}
void mor_smooth(void* h, int l, double* x, const double* b, int ncalls)
{
    auto* s = (Solver*)h; V X = s->lev[l].make(), B = s->lev[l].make(); s->scatter(l, X, x); s->scatter(l, B, b);
    for (int i = 0; i < ncalls; ++i) s->smooth(l, X, B, false);
    s->gather(l, X, x);
}
void mor_restrict(void* h, int lc, const double* fine, double* crse)
{
    auto* s = (Solver*)h; V F = s->lev[lc - 1].make(), Cc = s->lev[lc].make(); s->scatter(lc - 1, F, fine); s->restriction(lc, Cc, F); s->gather(lc, Cc, crse);
}
void mor_interp_add(void* h, int lf, double* fine, const double* crse)
{
    auto* s = (Solver*)h; V F = s->lev[lf].make(), Cc = s->lev[lf + 1].make(); s->scatter(lf, F, fine); s->scatter(lf + 1, Cc, crse); s->interp_add(lf, F, Cc); s->gather(lf, F, fine);
}
void mor_divergence(void* h, double* d) { auto* s = (Solver*)h; V D = s->lev[0].make(); s->divergence(D); s->gather(0, D, d); }
int mor_bottom_solve(void* h, const double* b, double* x)
{
    auto* s = (Solver*)h; const int l = (int)s->lev.size() - 1;
    s->scatter(l, s->res[l], b); s->bottom_iters = 0; s->bottom_solve(); s->gather(l, s->cor[l], x); return s->bottom_iters;
}
} // extern "C"
