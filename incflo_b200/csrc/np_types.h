// np_types.h — device-visible descriptors of one multigrid level on one rank.
//
// HBM layout (DESIGN.md "data layout"): every nodal array of a level holds the valid nodes
// [0..n] of this rank's box plus ONE ghost layer, x fastest:
//     element (i,j,k), -1 <= i <= n0+1  ->  base[(i+1) + nsy*(j+1) + nsz*(k+1)]
// with nsy = n0+3 rounded up to even (16-byte rows, what a TMA tensor map needs).  The cell-centred
// sigma of a level has one ghost layer:  (i,j,k), -1 <= i <= n0  ->  base[(i+1) + csy*(j+1) + csz*(k+1)].
// This is the FArrayBox layout of the reference (nodal / cell fab with 1 ghost) except for the padded
// row pitch, so phi and grad(phi) are handed back through an unpadded copy / directly.
#pragma once
#include "np_rt.h"

enum { NP_FACE_NONE = 0, NP_FACE_DIRICHLET = 1, NP_FACE_NEUMANN = 2 };
// how ghost/duplicate nodes of a dimension are produced
enum { NP_MAP_NONE = 0, NP_MAP_PERIODIC = 1, NP_MAP_MIRROR = 2 };

struct NpCoef {   // 27-point weights, SURVEY.md Appendix A (amrex::mlndlap_adotx_aa)
    double fxyz, fmx2y2z, f2xmy2z, f2x2ymz, f4xm2ym2z, fm2x4ym2z, fm2xm2y4z;
};

struct LevelDev {
    int n[3];                 // local cells; valid nodes are [0..n[d]]
    long long nsy, nsz, ntot; // nodal pitches / allocation (elements)
    long long csy, csz, ctot; // cell pitches / allocation (elements)
    double dxinv[3];
    NpCoef c;
    int face[6];              // NP_FACE_* of the local box faces (x-lo,y-lo,z-lo,x-hi,y-hi,z-hi); NONE for
                              // periodic and rank-interior faces
    int map_lo[3], map_hi[3]; // NP_MAP_* : how index -1 (lo) and indices n, n+1 (hi) are filled locally
    int owned_hi[3];          // last index this rank owns per dim (n, or n-1 when the high face is a duplicate)
    int const_sigma;          // 1: sigma is uniform, value in sigma0 (constant-coefficient kernels)
    double sigma0;
};

__host__ __device__ __forceinline__ long long np_noff(const LevelDev& L, int i, int j, int k) {
    return (long long)(i + 1) + L.nsy * (long long)(j + 1) + L.nsz * (long long)(k + 1);
}
__host__ __device__ __forceinline__ long long np_coff(const LevelDev& L, int i, int j, int k) {
    return (long long)(i + 1) + L.csy * (long long)(j + 1) + L.csz * (long long)(k + 1);
}
__host__ __device__ __forceinline__ bool np_is_dirichlet(const LevelDev& L, int i, int j, int k) {
    return (L.face[0] == NP_FACE_DIRICHLET && i == 0) || (L.face[3] == NP_FACE_DIRICHLET && i == L.n[0])
        || (L.face[1] == NP_FACE_DIRICHLET && j == 0) || (L.face[4] == NP_FACE_DIRICHLET && j == L.n[1])
        || (L.face[2] == NP_FACE_DIRICHLET && k == 0) || (L.face[5] == NP_FACE_DIRICHLET && k == L.n[2]);
}
// dot-mask weight of an owned node: 1/2 per Neumann face it lies on (amrex::mlndlap_set_dot_mask)
__host__ __device__ __forceinline__ double np_neumann_weight(const LevelDev& L, int i, int j, int k) {
    double w = 1.0;
    if (L.face[0] == NP_FACE_NEUMANN && i == 0) w *= 0.5;
    if (L.face[3] == NP_FACE_NEUMANN && i == L.n[0]) w *= 0.5;
    if (L.face[1] == NP_FACE_NEUMANN && j == 0) w *= 0.5;
    if (L.face[4] == NP_FACE_NEUMANN && j == L.n[1]) w *= 0.5;
    if (L.face[2] == NP_FACE_NEUMANN && k == 0) w *= 0.5;
    if (L.face[5] == NP_FACE_NEUMANN && k == L.n[2]) w *= 0.5;
    return w;
}
// source index of a ghost / duplicate index along one dimension (identity when nothing to do)
__host__ __device__ __forceinline__ int np_map_index(int idx, int n, int map_lo, int map_hi) {
    if (idx < 0) {
        if (map_lo == NP_MAP_PERIODIC) return idx + n;
        if (map_lo == NP_MAP_MIRROR) return -idx;
        return idx;
    }
    if (idx >= n) {
        if (map_hi == NP_MAP_PERIODIC) return idx - n;
        if (map_hi == NP_MAP_MIRROR && idx > n) return 2 * n - idx;
    }
    return idx;
}

// A caller-side fab (MultiFab layout, include/nodalproj.h): element (i,j,k,comp) of a box whose valid
// range starts at local index 0.
struct FabView {
    double* p;
    int ng;
    long long sy, sz, sn;
};
__host__ __device__ __forceinline__ long long np_foff(const FabView& f, int i, int j, int k, int comp) {
    return (long long)(i + f.ng) + f.sy * (long long)(j + f.ng) + f.sz * (long long)(k + f.ng) + f.sn * comp;
}
