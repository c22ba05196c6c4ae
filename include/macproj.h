/* macproj.h — C ABI of the B200-native MAC projection (part of libnodalproj.so).  SURVEY.md 8(f) row 1.
 *
 * The second Poisson solve of every incflo time step: amrex::MacProjector on the cell-centred operator
 * -div(b grad phi) with face-centred b = 1/rho (amrex::MLABecLaplacian, alpha = 0, beta = 1), which makes the
 * face (MAC) velocities discretely divergence-free.  Entry points and the reference interface each one replaces
 * (paths relative to the reference tree; the arithmetic lives in AMReX >= 20.11, which the reference does not vendor):
 *
 *   mp_create       <- MacProjector ctor + initProjector(lp_info, inv_rho)
 *                      src/convection/incflo_compute_MAC_projected_velocities.cpp:39-43, "mac_proj.*" options
 *                      src/incflo.H:316-344 (maxiter 200, bottom_maxiter 200, max_coarsening_level 100, rtol 1e-11, atol 1e-14)
 *   mp_set_bc       <- MacProjector::setDomainBC                  ...MAC_projected_velocities.cpp:44
 *                      (BC mapping: incflo::get_projection_bc, src/projection/incflo_apply_nodal_projection.cpp:6-36)
 *   mp_update_beta  <- initProjector's beta / MacProjector::updateBeta(inv_rho)         ...:43, :46
 *   mp_project      <- MacProjector::setUMAC(mac_vec) + project(rtol, atol)             ...:115, :131
 *                      or project(mac_phi, rtol, atol) (initial guess in, solution out)  ...:127
 *   mp_get_fluxes   <- macproj->getLinOp().getFluxes(m_fluxes, mac_phi, FaceCenter)     ...:63
 *   mp_destroy      <- ~MacProjector
 *
 * Fabs use AMReX's FArrayBox layout (include/nodalproj.h).  A face fab of direction d is nodal in d: valid index range
 * [lo, hi+1] in d, [lo, hi] in the other two.  Device or host pointers (host fabs are staged).  One AMR level.  One rank: any number
 * of boxes tiling the domain (the *_boxes entry points).  Several ranks (one GPU each): one box per rank, equal z-slabs
 * ordered by rank with an even number of planes; halo planes, norms and the gather onto the replicated coarse levels go
 * through NCCL (nccl_comm: an initialised ncclComm_t, e.g. from np_nccl_comm_create).
 * There is no CPU compute path: without a CUDA device mp_create fails with NP_ERR_NO_DEVICE.
 */
#ifndef MACPROJ_H
#define MACPROJ_H
#include "nodalproj.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mp_handle mp_handle;

typedef struct {
    int max_coarsening_level;   /* mac_proj.mg_max_coarsening_level        default 100 */
    int maxiter;                /* mac_proj.maxiter                        default 200 */
    int bottom_maxiter;         /* mac_proj.bottom_maxiter                 default 200 */
    int num_pre_smooth;         /* nu1                                     default 2 */
    int num_post_smooth;        /* nu2                                     default 2 */
    int num_final_smooth;       /* nuf (bottom fallback)                   default 8 */
    int bottom_solver;          /* np_bottom, mac_proj.bottom_solver       default bicgcg */
    int verbose;                /* mac_proj.verbose                        default 0 */
    int bottom_verbose;         /* mac_proj.bottom_verbose                 default 0 */
    int maxorder;               /* MLLinOp::setMaxOrder: Dirichlet ghost polynomial order, default 3 */
    double bottom_rtol;         /* mac_proj.bottom_rtol                    default 1e-4 */
    double bottom_atol;         /* mac_proj.bottom_atol                    default -1 */
    int const_beta_fastpath;    /* skip every coefficient read when b is uniform (constant density); default 1 */
    int fused_rb_min_n;         /* levels with at least this many cells in x run the red and the black sweep of a smooth in
                                   ONE pass (k_mac_rb3 / k_mac_rb2); smaller levels use one launch per colour; default 256 (measured:
                                   a 128^3 level is faster with two launches), <= 0: never.  With MLABecLaplacian's alpha * a term
                                   (mp_set_acoeffs) the one-pass kernel runs from twice this size up, on whole-tile levels */
    int fused_rb_const;         /* 1 (default): also when b is uniform; 0: one launch per colour there */
    int agg_min_local_nz;       /* multi-GPU: replicate MG levels whose slab would be thinner than this; default 32 */
    int profile_kernels;        /* bracket every level-0 smooth with CUDA events (np_stats.ms_smooth0 / n_smooth0); default 0 */
    int use_graph;              /* 1 (default): one MLMG iteration is captured as a CUDA graph and replayed, as np_options.use_graph;
                                   0: launch kernel by kernel */
    int p2p_halo;               /* multi-GPU: 1 (default): halo planes are stored straight into the neighbours' arrays over NVLink (CUDA IPC
                                   peer memory + flags, k_halo_p2p), as np_options.p2p_halo; falls back to grouped ncclSend/ncclRecv when
                                   peer memory cannot be opened on every rank; 0: NCCL */
} mp_options;

void mp_default_options(mp_options* o);

int mp_create(mp_handle** h, const np_geom* geom, int nboxes_global, const np_box* all_boxes, const int* owner_rank,
              int my_rank, int nranks, void* nccl_comm_or_null, const mp_options* opt);
int mp_set_bc(mp_handle* h, const int lo[3], const int hi[3]);
/* beta[d]: face fab of direction d, 1 comp (inv_rho_x/y/z).  Copied into the solver's hierarchy (coarse levels =
 * amrex::average_down_faces). */
int mp_update_beta(mp_handle* h, const np_fab beta[3]);
/* umac[d]: face fab of direction d, 1 comp, modified in place: umac -= b grad(phi) with div(b grad phi) = div(umac).
 * mac_phi: NULL (phi starts from 0 and stays inside the handle) or a cell fab, 1 comp, any ghost count: initial guess
 * in, solution out (valid cells). */
int mp_project(mp_handle* h, np_fab umac[3], np_fab* mac_phi_or_null, double rtol, double atol, void* cuda_stream, np_stats* out);
/* flux[d] = -b_d dphi/dx_d at face centres from a cell fab phi (homogeneous domain BCs as in the solve). */
int mp_get_fluxes(mp_handle* h, np_fab flux[3], const np_fab* phi);
int mp_copy_phi(mp_handle* h, np_fab* dst);          /* the last solution -> cell fab (valid cells) */

/* ---- the same MLMG / MLABecLaplacian machinery as the reference's implicit diffusion solves use it,
 *      (alpha a - beta div(b grad)) phi = rhs,   DiffusionScalarOp::diffuse_scalar and ::diffuse_vel_components
 *      (src/diffusion/DiffusionScalarOp.cpp:139-260, :262-380: alpha = 1, beta = dt, a = rho, b = eta averaged to faces).
 *   mp_set_scalars  <- MLABecLaplacian::setScalars(1.0, dt)                 DiffusionScalarOp.cpp:179, :315
 *   mp_set_acoeffs  <- setACoeffs(lev, density)  (cell fab, 1 comp; coarse levels by average_down)   :181, :317
 *   mp_update_beta  <- setBCoeffs(lev, b)        (face fabs)                                         :197, :324
 *   mp_set_bc       <- setDomainBC (constructor, :52-53: Dirichlet at inflow / no-slip faces, Neumann else)
 *   mp_solve        <- setLevelBC(lev, &phi) + MLMG::solve(phi, rhs, rtol, atol)                      :227, :250
 * phi: cell fab, 1 comp: initial guess in the valid cells and, next to Dirichlet faces, the wall values in the first ghost
 * layer (what setLevelBC reads; needs ngrow >= 1 then); solution out.  rhs: cell fab.  rho_or_null: when given, the
 * right-hand side is rho * rhs (diffuse_scalar passes the tracer and the density, :207-217) and no rhs fab has to be
 * formed.  One rank, one box this round; alpha != 0 runs the one-launch-per-colour smoother. */
int mp_set_scalars(mp_handle* h, double alpha, double beta);
int mp_set_acoeffs(mp_handle* h, const np_fab* a);
int mp_solve(mp_handle* h, np_fab* phi, const np_fab* rhs, const np_fab* rho_or_null, double rtol, double atol, void* cuda_stream, np_stats* out);

/* Several boxes (amr.max_grid_size chops the level; the boxes given to mp_create must tile the domain, all on this rank):
 * arrays of nlocal fabs per direction, in all_boxes order.  The boxes are gathered into domain-shaped device fabs, the
 * one-box path runs on those, and the results are scattered back (a face shared by two boxes is written to both):
 * bit-identical to the one-box calls, one extra streaming pass over every fab on either side. */
int mp_update_beta_boxes(mp_handle* h, int nlocal, const np_fab* bx, const np_fab* by, const np_fab* bz);
int mp_project_boxes(mp_handle* h, int nlocal, np_fab* ux, np_fab* uy, np_fab* uz, np_fab* mac_phi_or_null, double rtol, double atol,
                     void* cuda_stream, np_stats* out);
int mp_get_fluxes_boxes(mp_handle* h, int nlocal, np_fab* fx, np_fab* fy, np_fab* fz, const np_fab* phi);
void mp_destroy(mp_handle* h);
int mp_num_levels(mp_handle* h);
int mp_level_ncell(mp_handle* h, int mglev, int ncell[3]);

/* ---- kernel-level test hooks: cell fabs are 1 comp, 0 ghosts, at MG level mglev; need mp_update_beta first ---- */
int mp_get_beta(mp_handle* h, int mglev, np_fab beta[3]);                               /* averageDownCoeffs */
int mp_apply(mp_handle* h, int mglev, const np_fab* x, np_fab* y);                      /* Fapply */
int mp_smooth(mp_handle* h, int mglev, np_fab* x, const np_fab* rhs, int ncalls);       /* ncalls x (red, black) */
int mp_restrict(mp_handle* h, int cmglev, const np_fab* fine, np_fab* crse);
int mp_interp_add(mp_handle* h, int fmglev, np_fab* fine, const np_fab* crse);
int mp_bottom_solve(mp_handle* h, const np_fab* b, np_fab* x, int* iters);

#ifdef __cplusplus
}
#endif
#endif /* MACPROJ_H */
