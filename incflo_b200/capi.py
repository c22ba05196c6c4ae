"""ctypes binding of libnodalproj.so (include/nodalproj.h) — the C ABI a reference-side shim would bind.

There is no CPU compute path in the product: `load()` raises if the CUDA library has not been built,
and np_create returns NP_ERR_NO_DEVICE without a GPU.  (The CPU tests load a *test-only* emulation
build of the same sources by passing an explicit path; nothing in this package does that.)
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnodalproj.so")

NP_BC_PERIODIC, NP_BC_DIRICHLET, NP_BC_NEUMANN = 0, 1, 2
NP_BOTTOM_BICGCG, NP_BOTTOM_BICGSTAB, NP_BOTTOM_CG, NP_BOTTOM_SMOOTHER = 0, 1, 2, 3
NP_OK, NP_ERR_INVALID, NP_ERR_NO_DEVICE, NP_ERR_CUDA, NP_ERR_NCCL, NP_ERR_NOT_CONVERGED, NP_ERR_DIVERGED, NP_ERR_STATE, NP_ERR_TOO_LARGE = range(9)


class np_box(C.Structure):
    _fields_ = [("lo", C.c_int * 3), ("hi", C.c_int * 3)]


class np_fab(C.Structure):
    _fields_ = [("p", C.c_void_p), ("ngrow", C.c_int), ("ncomp", C.c_int)]


class np_geom(C.Structure):
    _fields_ = [("ncell", C.c_int * 3), ("prob_lo", C.c_double * 3), ("prob_hi", C.c_double * 3),
                ("periodic", C.c_int * 3)]


class np_options(C.Structure):
    _fields_ = [("max_coarsening_level", C.c_int), ("maxiter", C.c_int), ("bottom_maxiter", C.c_int),
                ("num_pre_smooth", C.c_int), ("num_post_smooth", C.c_int), ("num_final_smooth", C.c_int),
                ("bottom_solver", C.c_int), ("verbose", C.c_int), ("bottom_verbose", C.c_int),
                ("bottom_rtol", C.c_double), ("bottom_atol", C.c_double),
                ("smooth_zchunk", C.c_int), ("interp_sigma_weighted", C.c_int), ("bottom_normalize", C.c_int),
                ("agg_min_local_nz", C.c_int), ("use_graph", C.c_int), ("const_sigma_fastpath", C.c_int),
                ("profile_kernels", C.c_int), ("overlap_halo", C.c_int), ("p2p_halo", C.c_int)]


class np_stats(C.Structure):
    _fields_ = [("n_vcycles", C.c_int), ("bottom_iters_total", C.c_int), ("status", C.c_int), ("n_levels", C.c_int),
                ("rhs_norm0", C.c_double), ("res_norm0", C.c_double), ("res_norm_final", C.c_double),
                ("res_history", C.c_double * 128),
                ("ms_setup", C.c_double), ("ms_solve", C.c_double), ("ms_finish", C.c_double), ("ms_total", C.c_double),
                ("kernel_launches", C.c_longlong), ("sigma_is_const", C.c_int),
                ("n_smooth0", C.c_int), ("ms_smooth0", C.c_double), ("graph_replays", C.c_int), ("graph_captures", C.c_int)]


class np_apply_args(C.Structure):
    _fields_ = [("scaling_factor", C.c_double), ("incremental", C.c_int), ("proj_for_small_dt", C.c_int),
                ("zero_vel_ghosts", C.c_int)]


class mp_options(C.Structure):
    _fields_ = [("max_coarsening_level", C.c_int), ("maxiter", C.c_int), ("bottom_maxiter", C.c_int),
                ("num_pre_smooth", C.c_int), ("num_post_smooth", C.c_int), ("num_final_smooth", C.c_int),
                ("bottom_solver", C.c_int), ("verbose", C.c_int), ("bottom_verbose", C.c_int), ("maxorder", C.c_int),
                ("bottom_rtol", C.c_double), ("bottom_atol", C.c_double), ("const_beta_fastpath", C.c_int),
                ("fused_rb_min_n", C.c_int), ("fused_rb_const", C.c_int), ("agg_min_local_nz", C.c_int),
                ("profile_kernels", C.c_int), ("use_graph", C.c_int), ("p2p_halo", C.c_int)]


# every symbol include/macproj.h declares
MAC_SYMBOLS = ["mp_default_options", "mp_create", "mp_set_bc", "mp_update_beta", "mp_project", "mp_get_fluxes", "mp_copy_phi",
               "mp_destroy", "mp_num_levels", "mp_level_ncell", "mp_get_beta", "mp_apply", "mp_smooth", "mp_restrict",
               "mp_interp_add", "mp_bottom_solve", "mp_update_beta_boxes", "mp_project_boxes", "mp_get_fluxes_boxes",
               "mp_set_scalars", "mp_set_acoeffs", "mp_solve"]

# every symbol include/nodalproj.h declares
SYMBOLS = ["np_default_options", "np_create", "np_set_bc", "np_project", "np_apply_projection", "np_get_phi", "np_get_gradphi",
           "np_copy_phi", "np_copy_gradphi", "np_destroy", "np_strerror", "np_num_levels", "np_level_ncell",
           "np_last_stats", "np_set_sigma", "np_get_sigma", "np_apply", "np_smooth", "np_restrict",
           "np_interp_add", "np_divu", "np_mknewu", "np_bottom_solve", "np_vcycle", "np_dot",
           "np_nccl_unique_id", "np_nccl_comm_create", "np_nccl_comm_destroy", "np_is_device_pointer", "np_apply_noalloc", "np_cache_clear", "np_resid_restrict"]

_libs = {}


class NodalProjError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{msg} (code {code})")
        self.code = code


def load(path=None):
    """Load the C-ABI library.  Default: the in-tree CUDA build; raises if it is missing."""
    path = path or LIB_PATH
    if path in _libs:
        return _libs[path]
    if not os.path.exists(path):
        raise ImportError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(incflo_b200 has no CPU fallback)")
    L = C.CDLL(path, mode=C.RTLD_LOCAL)
    L.np_strerror.restype = C.c_char_p
    L.np_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(np_geom), C.c_int, C.POINTER(np_box), C.POINTER(C.c_int),
                            C.c_int, C.c_int, C.c_void_p, C.POINTER(np_options)]
    L.np_set_bc.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.np_project.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab), C.c_double, C.c_double,
                             C.c_void_p, C.POINTER(np_stats)]
    L.np_apply_projection.argtypes = [C.c_void_p, C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(np_fab),
                                      C.POINTER(np_fab), C.POINTER(np_apply_args), C.c_double, C.c_double, C.c_void_p,
                                      C.POINTER(np_stats)]
    for name in ("np_get_phi", "np_get_gradphi", "np_copy_phi", "np_copy_gradphi"):
        getattr(L, name).argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab)]
    L.np_destroy.argtypes = [C.c_void_p]
    L.np_destroy.restype = None
    L.np_num_levels.argtypes = [C.c_void_p]
    L.np_level_ncell.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    L.np_last_stats.argtypes = [C.c_void_p, C.POINTER(np_stats)]
    L.np_set_sigma.argtypes = [C.c_void_p, C.POINTER(np_fab)]
    L.np_get_sigma.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab)]
    L.np_apply.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab)]
    L.np_apply_noalloc.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab)]
    L.np_smooth.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab), C.c_int]
    L.np_restrict.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab)]
    L.np_interp_add.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab)]
    L.np_resid_restrict.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(np_fab)]
    L.np_divu.argtypes = [C.c_void_p, C.POINTER(np_fab), C.POINTER(np_fab)]
    L.np_mknewu.argtypes = [C.c_void_p, C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(np_fab)]
    L.np_bottom_solve.argtypes = [C.c_void_p, C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(C.c_int)]
    L.np_vcycle.argtypes = [C.c_void_p, C.POINTER(np_fab), C.POINTER(np_fab)]
    L.np_dot.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(C.c_double)]
    L.np_nccl_unique_id.argtypes = [C.c_void_p]
    L.np_nccl_comm_create.argtypes = [C.POINTER(C.c_void_p), C.c_void_p, C.c_int, C.c_int]
    L.np_nccl_comm_destroy.argtypes = [C.c_void_p]
    L.np_is_device_pointer.argtypes = [C.c_void_p]
    # ---- MAC projection (include/macproj.h) ----
    L.mp_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(np_geom), C.c_int, C.POINTER(np_box), C.POINTER(C.c_int),
                            C.c_int, C.c_int, C.c_void_p, C.POINTER(mp_options)]
    L.mp_set_bc.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.mp_update_beta.argtypes = [C.c_void_p, C.POINTER(np_fab)]
    L.mp_project.argtypes = [C.c_void_p, C.POINTER(np_fab), C.POINTER(np_fab), C.c_double, C.c_double, C.c_void_p,
                             C.POINTER(np_stats)]
    L.mp_set_scalars.argtypes = [C.c_void_p, C.c_double, C.c_double]
    L.mp_set_acoeffs.argtypes = [C.c_void_p, C.POINTER(np_fab)]
    L.mp_solve.argtypes = [C.c_void_p, C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(np_fab), C.c_double, C.c_double, C.c_void_p,
                           C.POINTER(np_stats)]
    L.mp_get_fluxes.argtypes = [C.c_void_p, C.POINTER(np_fab), C.POINTER(np_fab)]
    L.mp_copy_phi.argtypes = [C.c_void_p, C.POINTER(np_fab)]
    L.mp_update_beta_boxes.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(np_fab)]
    L.mp_project_boxes.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(np_fab),
                                   C.c_double, C.c_double, C.c_void_p, C.POINTER(np_stats)]
    L.mp_get_fluxes_boxes.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(np_fab)]
    L.mp_destroy.argtypes = [C.c_void_p]
    L.mp_destroy.restype = None
    L.mp_num_levels.argtypes = [C.c_void_p]
    L.mp_level_ncell.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    L.mp_get_beta.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab)]
    L.mp_apply.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab)]
    L.mp_smooth.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab), C.c_int]
    L.mp_restrict.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab)]
    L.mp_interp_add.argtypes = [C.c_void_p, C.c_int, C.POINTER(np_fab), C.POINTER(np_fab)]
    L.mp_bottom_solve.argtypes = [C.c_void_p, C.POINTER(np_fab), C.POINTER(np_fab), C.POINTER(C.c_int)]
    _libs[path] = L
    return L


def check(L, rc):
    if rc != 0:
        raise NodalProjError(rc, L.np_strerror(rc).decode())


def mac_default_options(L, **kw):
    o = mp_options()
    L.mp_default_options(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise KeyError(k)
        setattr(o, k, v)
    return o


def default_options(L, **kw):
    o = np_options()
    L.np_default_options(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise KeyError(k)
        setattr(o, k, v)
    return o
