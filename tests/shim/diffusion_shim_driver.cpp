// TEST INFRASTRUCTURE ONLY.  Drives include/DiffusionSolverB200.H the way DiffusionScalarOp::diffuse_scalar drives
// MLABecLaplacian + MLMG (/root/reference/src/diffusion/DiffusionScalarOp.cpp:139-260), against the mock AMReX types of
// tests/mock_amrex and whichever build of the C ABI it is linked with.  Problem (restated in tests/test_shim.py): N^3 cells,
// mass inflow (Dirichlet, tracer = 1 in the ghost layer) at x-lo, Neumann at x-hi and in y, periodic z.  Writes the tracer to argv[2].
#include "DiffusionSolverB200.H"
#include <cmath>
#include <cstdio>
#include <memory>
#include <vector>
using namespace amrex;

int main(int argc, char** argv)
{
    const int N = argc > 1 ? std::atoi(argv[1]) : 16;
    const char* out = argc > 2 ? argv[2] : "diff_shim_out.bin";
    const double twopi = 2.0 * 3.1415926535897932, dx = 1.0 / N, dt = 0.01;
    Box dom({0,0,0}, {N-1,N-1,N-1});
    BoxArray ba; DistributionMapping dm;
    ba.b.push_back(dom); dm.p.push_back(0);
    Geometry geom; geom.dom = dom;
    for (int d = 0; d < 3; ++d) { geom.plo[d] = 0.0; geom.phi[d] = 1.0; geom.per[d] = d == 2; }
    MultiFab tracer(ba, dm, 1, 1), density(ba, dm, 1, 0), rhs(ba, dm, 1, 0);
    std::unique_ptr<MultiFab> eta[3];
    for (MFIter mfi(tracer); mfi.isValid(); ++mfi) {
        FArrayBox& t = tracer[mfi]; FArrayBox& r = density[mfi]; FArrayBox& q = rhs[mfi];
        for (int k = -1; k <= N; ++k) for (int j = -1; j <= N; ++j) for (int i = -1; i <= N; ++i) {
            const double x = (i + 0.5) * dx, y = (j + 0.5) * dx, z = (k + 0.5) * dx;
            t(i,j,k) = i < 0 ? 1.0 : x * y * (1.0 - z);                       // inflow value in the x-lo ghost layer
        }
        for (int k = 0; k < N; ++k) for (int j = 0; j < N; ++j) for (int i = 0; i < N; ++i) {
            const double x = (i + 0.5) * dx, y = (j + 0.5) * dx;
            r(i,j,k) = 1.0 + 0.5 * std::sin(twopi * x) * std::cos(twopi * y);
            q(i,j,k) = r(i,j,k) * t(i,j,k);                                  // rhs = rho * tracer (:207-217)
        }
    }
    for (int d = 0; d < 3; ++d) {
        eta[d].reset(new MultiFab(convert(ba, IntVect::TheDimensionVector(d)), dm, 1, 0));
        for (MFIter mfi(*eta[d]); mfi.isValid(); ++mfi) {
            FArrayBox& e = (*eta[d])[mfi]; const Box fb = mfi.validbox();
            for (int k = fb.lo[2]; k <= fb.hi[2]; ++k) for (int j = fb.lo[1]; j <= fb.hi[1]; ++j) for (int i = fb.lo[0]; i <= fb.hi[0]; ++i) {
                const double z = (k + (d == 2 ? 0.0 : 0.5)) * dx;
                e(i,j,k) = 0.02 + 0.01 * std::cos(twopi * z);
            }
        }
    }
    LPInfo info; info.setMaxCoarseningLevel(100);
    b200::ABecSolver op({geom}, {ba}, {dm}, info);
    op.setMaxOrder(2);
    op.setDomainBC({LinOpBCType::Dirichlet, LinOpBCType::Neumann, LinOpBCType::Periodic},
                   {LinOpBCType::Neumann, LinOpBCType::Neumann, LinOpBCType::Periodic});
    op.setScalars(1.0, dt);
    op.setACoeffs(0, density);
    op.setBCoeffs(0, {eta[0].get(), eta[1].get(), eta[2].get()});
    op.setLevelBC(0, &tracer);
    op.setMaxIter(100); op.setBottomMaxIter(100); op.setBottomSolver("bicgcg");
    op.solve({&tracer}, {&rhs}, 1e-11, 1e-14);
    FILE* f = std::fopen(out, "wb");
    if (!f) return 2;
    int hdr[2] = {N, op.stats().n_vcycles};
    std::fwrite(hdr, sizeof(int), 2, f);
    std::vector<double> T((size_t)N * N * N);
    for (MFIter mfi(tracer); mfi.isValid(); ++mfi) {
        FArrayBox& t = tracer[mfi];
        for (int k = 0; k < N; ++k) for (int j = 0; j < N; ++j) for (int i = 0; i < N; ++i) T[i + (size_t)N * (j + (size_t)N * k)] = t(i,j,k);
    }
    std::fwrite(T.data(), 8, T.size(), f);
    std::fclose(f);
    std::printf("diffusion shim: N=%d vcycles=%d\n", N, hdr[1]);
    return 0;
}
