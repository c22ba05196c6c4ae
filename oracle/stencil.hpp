// ORACLE — TEST INFRASTRUCTURE ONLY.  Parity unpinned upstream (see oracle/README.md).
//
// Point formulas of the nodal projection, restated for the CPU oracle.
//
// The arithmetic of this path lives in AMReX (>= 20.11, branch development), which is NOT
// vendored in /root/reference; the only local anchor is the call site
//   /root/reference/src/projection/incflo_apply_nodal_projection.cpp:148-153,165-166.
// Upstream names each function restates (no local file:line exists):
//   adotx_node      <- amrex::mlndlap_adotx_aa        (AMReX_MLNodeLap_3D_K.H)
//   diag_node       <- s0 in amrex::mlndlap_gauss_seidel_aa / mlndlap_normalize_aa
//   divu_node       <- amrex::mlndlap_divu
//   grad_cell       <- amrex::mlndlap_mknewu
//   restrict_node   <- amrex::mlndlap_restriction
//   interp_node     <- amrex::mlndlap_interpadd_aa (sigma-weighted) / trilinear
// The formulas are SURVEY.md Appendix A; tests/test_oracle_pins.py checks them against an
// independent Q1 finite-element assembly (oracle/fem_assembly.py).
#pragma once
#include <cstddef>

namespace orc {

struct StencilCoef {
    double fxyz, fmx2y2z, f2xmy2z, f2x2ymz, f4xm2ym2z, fm2x4ym2z, fm2xm2y4z;
    explicit StencilCoef(const double dxinv[3]) {
        const double fx = dxinv[0]*dxinv[0]/36.0, fy = dxinv[1]*dxinv[1]/36.0, fz = dxinv[2]*dxinv[2]/36.0;
        fxyz = fx+fy+fz;
        fmx2y2z = -fx+2*fy+2*fz;  f2xmy2z = 2*fx-fy+2*fz;  f2x2ymz = 2*fx+2*fy-fz;
        f4xm2ym2z = 4*fx-2*fy-2*fz;  fm2x4ym2z = -2*fx+4*fy-2*fz;  fm2xm2y4z = -2*fx-2*fy+4*fz;
    }
};

// x points at node (i,j,k) with strides (1,xy,xz); s points at CELL (i,j,k) with strides (1,sy,sz).
// The 8 cells touching the node are s[a + b*sy + c*sz], a,b,c in {-1,0}.
struct Sig8 {
    double mmm, pmm, mpm, ppm, mmp, pmp, mpp, ppp;   // (x,y,z) side: m = -1, p = 0
    Sig8(const double* s, std::ptrdiff_t sy, std::ptrdiff_t sz)
        : mmm(s[-1-sy-sz]), pmm(s[-sy-sz]), mpm(s[-1-sz]), ppm(s[-sz]),
          mmp(s[-1-sy]),    pmp(s[-sy]),    mpp(s[-1]),    ppp(s[0]) {}
    double sum() const { return mmm+pmm+mpm+ppm+mmp+pmp+mpp+ppp; }
};

inline double diag_node(const StencilCoef& c, const Sig8& g) { return -4.0*c.fxyz*g.sum(); }

// Off-diagonal part of (A x)(i,j,k)
inline double offdiag_node(const StencilCoef& c, const Sig8& g, const double* x,
                           std::ptrdiff_t xy, std::ptrdiff_t xz)
{
    const double* xm = x - xz;   // plane k-1
    const double* xp = x + xz;   // plane k+1
    double corners = xm[-1-xy]*g.mmm + xm[1-xy]*g.pmm + xm[-1+xy]*g.mpm + xm[1+xy]*g.ppm
                   + xp[-1-xy]*g.mmp + xp[1-xy]*g.pmp + xp[-1+xy]*g.mpp + xp[1+xy]*g.ppp;
    double eyz = xm[-xy]*(g.mmm+g.pmm) + xm[xy]*(g.mpm+g.ppm)
               + xp[-xy]*(g.mmp+g.pmp) + xp[xy]*(g.mpp+g.ppp);
    double exz = xm[-1]*(g.mmm+g.mpm) + xm[1]*(g.pmm+g.ppm)
               + xp[-1]*(g.mmp+g.mpp) + xp[1]*(g.pmp+g.ppp);
    double exy = x[-1-xy]*(g.mmm+g.mmp) + x[1-xy]*(g.pmm+g.pmp)
               + x[-1+xy]*(g.mpm+g.mpp) + x[1+xy]*(g.ppm+g.ppp);
    double fx = x[-1]*(g.mmm+g.mpm+g.mmp+g.mpp) + x[1]*(g.pmm+g.ppm+g.pmp+g.ppp);
    double fy = x[-xy]*(g.mmm+g.pmm+g.mmp+g.pmp) + x[xy]*(g.mpm+g.ppm+g.mpp+g.ppp);
    double fz = xm[0]*(g.mmm+g.pmm+g.mpm+g.ppm) + xp[0]*(g.mmp+g.pmp+g.mpp+g.ppp);
    return c.fxyz*corners + c.fmx2y2z*eyz + c.f2xmy2z*exz + c.f2x2ymz*exy
         + c.f4xm2ym2z*fx + c.fm2x4ym2z*fy + c.fm2xm2y4z*fz;
}

inline double adotx_node(const StencilCoef& c, const double* x, std::ptrdiff_t xy, std::ptrdiff_t xz,
                         const double* s, std::ptrdiff_t sy, std::ptrdiff_t sz)
{
    Sig8 g(s, sy, sz);
    return x[0]*diag_node(c, g) + offdiag_node(c, g, x, xy, xz);
}

// Full-weighting restriction at coarse node <- fine node f (strides 1,fy,fz): 1/8,1/16,1/32,1/64.
inline double restrict_node(const double* f, std::ptrdiff_t fy, std::ptrdiff_t fz)
{
    double faces = f[-1]+f[1]+f[-fy]+f[fy]+f[-fz]+f[fz];
    double edges = f[-1-fy]+f[1-fy]+f[-1+fy]+f[1+fy]
                 + f[-1-fz]+f[1-fz]+f[-1+fz]+f[1+fz]
                 + f[-fy-fz]+f[fy-fz]+f[-fy+fz]+f[fy+fz];
    double corners = f[-1-fy-fz]+f[1-fy-fz]+f[-1+fy-fz]+f[1+fy-fz]
                   + f[-1-fy+fz]+f[1-fy+fz]+f[-1+fy+fz]+f[1+fy+fz];
    return 0.125*f[0] + 0.0625*faces + 0.03125*edges + 0.015625*corners;
}

// ---- sigma-weighted prolongation helpers (restating aa_interp_line_* / aa_interp_face_*) ----
// All take the FINE node (i,j,k) through `s` = pointer to fine CELL (i,j,k) and `c` = pointer to the
// coarse node whose fine image is the LOW neighbour along the interpolated direction(s).
struct SigQuad {   // 4-cell sums on each side of a fine node
    double xlo, xhi, ylo, yhi, zlo, zhi;
    explicit SigQuad(const Sig8& g)
        : xlo(g.mmm+g.mpm+g.mmp+g.mpp), xhi(g.pmm+g.ppm+g.pmp+g.ppp),
          ylo(g.mmm+g.pmm+g.mmp+g.pmp), yhi(g.mpm+g.ppm+g.mpp+g.ppp),
          zlo(g.mmm+g.pmm+g.mpm+g.ppm), zhi(g.mmp+g.pmp+g.mpp+g.ppp) {}
};

} // namespace orc
