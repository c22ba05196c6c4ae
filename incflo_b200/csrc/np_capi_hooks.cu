// np_capi_hooks.cu — kernel-level test hooks of include/nodalproj.h (one kernel family each), so the
// parity tests can compare every kernel with the oracle through the C ABI.
#include "np_solver.h"
#include "np_kernels.cuh"



namespace {
struct Tmp {                       // device scratch in the internal nodal layout of level l
    double* p = nullptr;
    Tmp(NpSolver& S, int l) { void* q = nullptr; np_rt::dev_malloc(&q, (size_t)S.lev[l].d.ntot * sizeof(double)); p = (double*)q;
                              if (p) np_rt::memset_async(p, 0, (size_t)S.lev[l].d.ntot * sizeof(double), S.stream); }
    ~Tmp() { if (p) np_rt::dev_free(p); }
    Tmp(const Tmp&) = delete;
    Tmp& operator=(const Tmp&) = delete;
};
// valid nodes of a caller fab -> internal array
int fab_in(NpSolver& S, int l, const np_fab* f, double* dst)
{
    if (!f || !f->p || f->ngrow < 0) return NP_ERR_INVALID;
    const LevelDev& d = S.lev[l].d;
    FabView v = S.view_nodal(*f, l);
    double* dev; bool staged;
    if (S.import_fab(*f, v.sn, &dev, &staged)) return NP_ERR_CUDA;
    v.p = dev;
    NP_LAUNCH(false, (k_copy_nodal<false>), grid_nodes(d), NP_BLK, 0, S.stream, d, dst, v, 0);
    np_rt::stream_sync(S.stream);
    if (staged) np_rt::dev_free(dev);
    return NP_OK;
}
int fab_out(NpSolver& S, int l, const double* src, np_fab* f)
{
    if (!f || !f->p || f->ngrow < 0) return NP_ERR_INVALID;
    const LevelDev& d = S.lev[l].d;
    FabView v = S.view_nodal(*f, l);
    double* dev; bool staged;
    if (S.import_fab(*f, v.sn, &dev, &staged)) return NP_ERR_CUDA;
    v.p = dev;
    NP_LAUNCH(false, (k_copy_nodal<true>), grid_nodes(d), NP_BLK, 0, S.stream, d, const_cast<double*>(src), v, 0);
    S.export_fab(*f, v.sn, dev, staged);
    np_rt::stream_sync(S.stream);
    return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
}
bool bad_level(np_handle* h, int l) { return !h || l < 0 || l >= (int)h->S.lev.size(); }
} // namespace

extern "C" {

int np_set_sigma(np_handle* h, const np_fab* sigma)
{
    if (!h || !sigma || !sigma->p) return NP_ERR_INVALID;
    NpSolver& S = h->S;
    FabView sv = S.view_cell(*sigma, 0, 1);
    double* dev; bool staged;
    if (S.import_fab(*sigma, sv.sn, &dev, &staged)) return NP_ERR_CUDA;
    sv.p = dev;
    int rc = S.set_sigma(sv);
    np_rt::stream_sync(S.stream);
    if (staged) np_rt::dev_free(dev);
    return rc;
}

int np_get_sigma(np_handle* h, int l, np_fab* dst)
{
    if (bad_level(h, l) || !dst || !dst->p) return NP_ERR_INVALID;
    NpSolver& S = h->S;
    if (!S.have_sigma) return NP_ERR_STATE;
    const LevelDev& d = S.lev[l].d;
    FabView v = S.view_cell(*dst, l, 1);
    double* dev; bool staged;
    if (S.import_fab(*dst, v.sn, &dev, &staged)) return NP_ERR_CUDA;
    v.p = dev;
    NP_LAUNCH(false, k_copy_sigma_out, grid_cells(d), NP_BLK, 0, S.stream, d, (const double*)S.lev[l].sigma, v);
    S.export_fab(*dst, v.sn, dev, staged);
    np_rt::stream_sync(S.stream);
    return NP_OK;
}

int np_apply(np_handle* h, int l, const np_fab* x, np_fab* y)
{
    if (bad_level(h, l)) return NP_ERR_INVALID;
    NpSolver& S = h->S; if (!S.have_sigma) return NP_ERR_STATE;
    Tmp tx(S, l), ty(S, l);
    int rc = fab_in(S, l, x, tx.p); if (rc) return rc;
    S.apply(l, ty.p, tx.p);
    return fab_out(S, l, ty.p, y);
}

int np_smooth(np_handle* h, int l, np_fab* x, const np_fab* rhs, int ncalls)
{
    if (bad_level(h, l)) return NP_ERR_INVALID;
    NpSolver& S = h->S; if (!S.have_sigma) return NP_ERR_STATE;
    NpLevel& L = S.lev[l];
    int rc = fab_in(S, l, x, L.cor); if (rc) return rc;
    rc = fab_in(S, l, rhs, L.res); if (rc) return rc;
    L.cor_ghost_clean = false;
    S.lvl0_off = nullptr;
    for (int i = 0; i < ncalls; ++i) S.smooth(l, false, nullptr);
    return fab_out(S, l, L.cor, x);
}

int np_restrict(np_handle* h, int lc, const np_fab* fine, np_fab* crse)
{
    if (bad_level(h, lc) || lc < 1) return NP_ERR_INVALID;
    NpSolver& S = h->S;
    Tmp tf(S, lc - 1), tc(S, lc);
    int rc = fab_in(S, lc - 1, fine, tf.p); if (rc) return rc;
    S.restrict_to(lc, tc.p, tf.p);
    return fab_out(S, lc, tc.p, crse);
}

// K2 + K4 fused (k_resid_restrict + k_restrict_faces): crse = R(res - A cor) on level l -> l + 1
int np_resid_restrict(np_handle* h, int l, const np_fab* cor, const np_fab* res, np_fab* crse)
{
    if (bad_level(h, l) || bad_level(h, l + 1)) return NP_ERR_INVALID;
    NpSolver& S = h->S; if (!S.have_sigma) return NP_ERR_STATE;
    if (S.lev[l + 1].replicated != S.lev[l].replicated) return NP_ERR_INVALID;
    NpLevel& L = S.lev[l];
    int rc = fab_in(S, l, cor, L.cor); if (rc) return rc;
    rc = fab_in(S, l, res, L.res); if (rc) return rc;
    L.cor_ghost_clean = false;
    S.resid_restrict(l, nullptr);
    return fab_out(S, l + 1, S.lev[l + 1].res, crse);
}

int np_interp_add(np_handle* h, int lf, np_fab* fine, const np_fab* crse)
{
    if (bad_level(h, lf) || bad_level(h, lf + 1)) return NP_ERR_INVALID;
    NpSolver& S = h->S; if (!S.have_sigma) return NP_ERR_STATE;
    Tmp tf(S, lf), tc(S, lf + 1);
    int rc = fab_in(S, lf, fine, tf.p); if (rc) return rc;
    rc = fab_in(S, lf + 1, crse, tc.p); if (rc) return rc;
    S.interp_add(lf, tf.p, tc.p);
    return fab_out(S, lf, tf.p, fine);
}

int np_divu(np_handle* h, const np_fab* vel, np_fab* rhs)
{
    if (!h || !vel || !vel->p || vel->ngrow < 1 || vel->ncomp < 3) return NP_ERR_INVALID;
    NpSolver& S = h->S;
    FabView uv = S.view_cell(*vel, 0, 3);
    double* dev; bool staged;
    if (S.import_fab(*vel, uv.sn * 3, &dev, &staged)) return NP_ERR_CUDA;
    uv.p = dev;
    // same two launches project() uses
    np_fill_cell3(S, uv);
    NP_LAUNCH(true, (k_divu<false>), grid_nodes(S.lev[0].d), NP_BLK, 0, S.stream, S.lev[0].d, S.rhs, uv, S.part1);
    np_rt::stream_sync(S.stream);
    if (staged) np_rt::dev_free(dev);
    return fab_out(S, 0, S.rhs, rhs);
}

int np_mknewu(np_handle* h, const np_fab* phi, np_fab* vel, np_fab* gradphi)
{
    if (!h || !vel || !vel->p || vel->ncomp < 3 || !gradphi || !gradphi->p || gradphi->ncomp < 3) return NP_ERR_INVALID;
    NpSolver& S = h->S; if (!S.have_sigma) return NP_ERR_STATE;
    const LevelDev& d = S.lev[0].d;
    int rc = fab_in(S, 0, phi, S.sol); if (rc) return rc;
    FabView uv = S.view_cell(*vel, 0, 3), gv = S.view_cell(*gradphi, 0, 3);
    double *du, *dg; bool su, sg;
    if (S.import_fab(*vel, uv.sn * 3, &du, &su) || S.import_fab(*gradphi, gv.sn * 3, &dg, &sg)) return NP_ERR_CUDA;
    uv.p = du; gv.p = dg;
    NP_LAUNCH(false, k_mknewu, grid_cells(d), NP_BLK, 0, S.stream, d, uv, gv, (const double*)S.sol, (const double*)S.lev[0].sigma);
    S.export_fab(*vel, uv.sn * 3, du, su);
    S.export_fab(*gradphi, gv.sn * 3, dg, sg);
    np_rt::stream_sync(S.stream);
    return np_rt::last_error() ? NP_ERR_CUDA : NP_OK;
}

int np_bottom_solve(np_handle* h, const np_fab* b, np_fab* x, int* iters)
{
    if (!h) return NP_ERR_INVALID;
    NpSolver& S = h->S; if (!S.have_sigma) return NP_ERR_STATE;
    const int l = (int)S.lev.size() - 1;
    int rc = fab_in(S, l, b, S.lev[l].res); if (rc) return rc;
    S.bicg_failed_switch = false;
    int it = S.bottom_solve();
    if (iters) *iters = it;
    return fab_out(S, l, S.lev[l].cor, x);
}

int np_vcycle(np_handle* h, const np_fab* res0, np_fab* cor0)
{
    if (!h) return NP_ERR_INVALID;
    NpSolver& S = h->S; if (!S.have_sigma) return NP_ERR_STATE;
    int rc = fab_in(S, 0, res0, S.lev[0].res); if (rc) return rc;
    S.bicg_failed_switch = false;
    S.lvl0_off = nullptr;
    if (S.singular) S.make_solvable(0, S.lev[0].res);
    S.vcycle(false);
    return fab_out(S, 0, S.lev[0].cor, cor0);
}

int np_dot(np_handle* h, int l, const np_fab* x, const np_fab* y, double* out)
{
    if (bad_level(h, l) || !out) return NP_ERR_INVALID;
    NpSolver& S = h->S;
    const LevelDev& d = S.lev[l].d;
    Tmp tx(S, l), ty(S, l);
    int rc = fab_in(S, l, x, tx.p); if (rc) return rc;
    rc = fab_in(S, l, y, ty.p); if (rc) return rc;
    const long long nb = (long long)grid_nodes(d).x * grid_nodes(d).y * grid_nodes(d).z;
    const double* px = tx.p; const double* py = ty.p;
    NP_LAUNCH(true, k_wdot, grid_nodes(d), NP_BLK, 0, S.stream, d, px, py, S.part0, S.part1);
    NP_LAUNCH(true, (k_reduce_final<false>), dim3(1), dim3(256), 0, S.stream, (const double*)S.part0, nb, S.scal, 8);
    if (S.comm.nranks > 1 && !S.lev[l].replicated) S.comm.allreduce(S.comm.ctx, S.scal + 8, 1, 0, (void*)S.stream);
    np_rt::copy_async(S.h_scal + 8, S.scal + 8, sizeof(double), S.stream);
    np_rt::stream_sync(S.stream);
    *out = S.h_scal[8];
    return NP_OK;
}

} // extern "C"
