// TEST INFRASTRUCTURE ONLY.  Drives include/NodalProjectorB200.H exactly as incflo::ApplyProjection drives
// amrex::NodalProjector (/root/reference/src/projection/incflo_apply_nodal_projection.cpp:148-166), against the mock
// AMReX types of tests/mock_amrex and whichever build of the C ABI it is linked with.  Problem: Taylor-Green + the
// perturbation of incflo_b200/problems.py:taylor_green, N^3 cells, periodic, sigma = 1.  Writes u, phi, gradphi (valid
// regions, C order [comp][k][j][i]) to the file given as argv[2].
#include "NodalProjectorB200.H"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <vector>
using namespace amrex;

int main(int argc, char** argv)
{
    const int N = argc > 1 ? std::atoi(argv[1]) : 16;
    const char* out = argc > 2 ? argv[2] : "shim_out.bin";
    if (argc > 3) ParmParse::table()["nodal_proj.maxiter"] = argv[3];
    const int MG = argc > 4 ? std::atoi(argv[4]) : N;       // amr.max_grid_size: N/MG boxes per direction
    Box dom({0,0,0}, {N-1,N-1,N-1});
    BoxArray ba; DistributionMapping dm;
    for (int k = 0; k < N; k += MG) for (int j = 0; j < N; j += MG) for (int i = 0; i < N; i += MG) {
        ba.b.push_back(Box({i,j,k}, {std::min(i+MG,N)-1, std::min(j+MG,N)-1, std::min(k+MG,N)-1})); dm.p.push_back(0);
    }
    Geometry geom; geom.dom = dom;
    for (int d = 0; d < 3; ++d) { geom.plo[d] = 0.0; geom.phi[d] = 1.0; geom.per[d] = 1; }
    MultiFab vel(ba, dm, 3, 2), sigma(ba, dm, 1, 0);
    const double twopi = 2.0 * 3.1415926535897932, dx = 1.0 / N;
    for (MFIter mfi(vel); mfi.isValid(); ++mfi) {
        FArrayBox& u = vel[mfi]; FArrayBox& s = sigma[mfi];
        const Box vb = mfi.validbox();
        for (int k = vb.lo[2]; k <= vb.hi[2]; ++k) for (int j = vb.lo[1]; j <= vb.hi[1]; ++j) for (int i = vb.lo[0]; i <= vb.hi[0]; ++i) {
            const double x = (i + 0.5) * dx, y = (j + 0.5) * dx, z = (k + 0.5) * dx;
            const double p = 0.1 * std::sin(twopi * x) * std::sin(twopi * y) * std::sin(twopi * z);
            u(i,j,k,0) = std::sin(twopi * x) * std::cos(twopi * y) + p;
            u(i,j,k,1) = -std::cos(twopi * x) * std::sin(twopi * y) + p;
            u(i,j,k,2) = p;
            s(i,j,k) = 1.0;
        }
    }
    LPInfo info; info.setMaxCoarseningLevel(30);
    std::unique_ptr<b200::NodalProjector> nodal_projector;
    Vector<MultiFab*> v{&vel}; Vector<const MultiFab*> sg{&sigma};
    nodal_projector.reset(new b200::NodalProjector(v, sg, {geom}, info));
    std::array<LinOpBCType,3> bc{LinOpBCType::Periodic, LinOpBCType::Periodic, LinOpBCType::Periodic};
    nodal_projector->setDomainBC(bc, bc);
    nodal_projector->project(1e-11, 1e-14);
    auto phi = nodal_projector->getPhi();
    auto gradphi = nodal_projector->getGradPhi();
    FILE* f = std::fopen(out, "wb");
    if (!f) return 2;
    int hdr[2] = {N, nodal_projector->stats().n_vcycles};
    std::fwrite(hdr, sizeof(int), 2, f);
    // assemble the global arrays from the boxes (shared nodes: every box must agree)
    std::vector<double> U(3L*N*N*N), G(3L*N*N*N), P((long)(N+1)*(N+1)*(N+1));
    std::vector<char> seen(P.size(), 0);
    int mismatch = 0;
    for (MFIter mfi(vel); mfi.isValid(); ++mfi) {
        const Box vb = mfi.validbox();
        FArrayBox& u = vel[mfi];
        FArrayBox& p = const_cast<FArrayBox&>((*phi[0])[mfi]);
        FArrayBox& g = const_cast<FArrayBox&>((*gradphi[0])[mfi]);
        for (int n = 0; n < 3; ++n) for (int k = vb.lo[2]; k <= vb.hi[2]; ++k) for (int j = vb.lo[1]; j <= vb.hi[1]; ++j) for (int i = vb.lo[0]; i <= vb.hi[0]; ++i) {
            U[i + (long)N*(j + (long)N*(k + (long)N*n))] = u(i,j,k,n);
            G[i + (long)N*(j + (long)N*(k + (long)N*n))] = g(i,j,k,n);
        }
        for (int k = vb.lo[2]; k <= vb.hi[2]+1; ++k) for (int j = vb.lo[1]; j <= vb.hi[1]+1; ++j) for (int i = vb.lo[0]; i <= vb.hi[0]+1; ++i) {
            const long q = i + (long)(N+1)*(j + (long)(N+1)*k);
            if (seen[q] && P[q] != p(i,j,k)) ++mismatch;
            P[q] = p(i,j,k); seen[q] = 1;
        }
    }
    if (mismatch) { std::fprintf(stderr, "shim: %d shared nodes differ between boxes\n", mismatch); return 4; }
    std::fwrite(U.data(), 8, U.size(), f);
    std::fwrite(P.data(), 8, P.size(), f);
    std::fwrite(G.data(), 8, G.size(), f);
    std::fclose(f);
    std::printf("shim: N=%d vcycles=%d\n", N, hdr[1]);
    return 0;
}
