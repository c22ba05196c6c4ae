// TEST INFRASTRUCTURE ONLY.  Drives include/MacProjectorB200.H the way incflo::compute_MAC_projected_velocities drives
// amrex::MacProjector (/root/reference/src/convection/incflo_compute_MAC_projected_velocities.cpp:36-131), against the mock
// AMReX types of tests/mock_amrex and whichever build of the C ABI it is linked with.  Problem: incflo_b200/problems.py
// mac_double_shear on N^3 cells (periodic, beta = 1).  Writes umac (3 face arrays, C order [k][j][i]) and mac_phi to argv[2].
#include "MacProjectorB200.H"
#include <algorithm>
#include <cmath>
#include <vector>
#include <cstdio>
#include <memory>
using namespace amrex;

int main(int argc, char** argv)
{
    const int N = argc > 1 ? std::atoi(argv[1]) : 16;
    const char* out = argc > 2 ? argv[2] : "mac_shim_out.bin";
    const int MG = argc > 3 ? std::atoi(argv[3]) : N;       // amr.max_grid_size
    Box dom({0,0,0}, {N-1,N-1,N-1});
    BoxArray ba; DistributionMapping dm;
    for (int k = 0; k < N; k += MG) for (int j = 0; j < N; j += MG) for (int i = 0; i < N; i += MG) {
        ba.b.push_back(Box({i,j,k}, {std::min(i+MG,N)-1, std::min(j+MG,N)-1, std::min(k+MG,N)-1})); dm.p.push_back(0);
    }
    Geometry geom; geom.dom = dom;
    for (int d = 0; d < 3; ++d) { geom.plo[d] = 0.0; geom.phi[d] = 1.0; geom.per[d] = 1; }
    std::unique_ptr<MultiFab> umac[3], rho_inv[3];
    const double twopi = 2.0 * 3.1415926535897932, dx = 1.0 / N;
    for (int d = 0; d < 3; ++d) {
        const BoxArray fba = convert(ba, IntVect::TheDimensionVector(d));
        umac[d].reset(new MultiFab(fba, dm, 1, 0)); rho_inv[d].reset(new MultiFab(fba, dm, 1, 0));
        for (MFIter mfi(*umac[d]); mfi.isValid(); ++mfi) {
            FArrayBox& u = (*umac[d])[mfi]; FArrayBox& b = (*rho_inv[d])[mfi];
            const Box fb = mfi.validbox();
            for (int k = fb.lo[2]; k <= fb.hi[2]; ++k) for (int j = fb.lo[1]; j <= fb.hi[1]; ++j) for (int i = fb.lo[0]; i <= fb.hi[0]; ++i) {
                const double x = (i + (d == 0 ? 0.0 : 0.5)) * dx, y = (j + (d == 1 ? 0.0 : 0.5)) * dx, z = (k + (d == 2 ? 0.0 : 0.5)) * dx;
                const double p = 0.1 * std::sin(twopi * x) * std::sin(twopi * y) * std::sin(twopi * z);
                const double f = d == 0 ? std::tanh(30.0 * (0.25 - std::fabs(y - 0.5))) : d == 1 ? 0.05 * std::sin(twopi * x)
                                        : 0.05 * std::sin(twopi * x) * std::sin(twopi * z);
                u(i,j,k) = f + p; b(i,j,k) = 1.0;
            }
        }
    }
    MultiFab mac_phi(ba, dm, 1, 1);
    std::unique_ptr<b200::MacProjector> macproj;
    macproj.reset(new b200::MacProjector({geom}));
    Vector<Array<MultiFab const*,3>> inv_rho(1);
    Vector<Array<MultiFab*,3>> mac_vec(1);
    for (int d = 0; d < 3; ++d) { inv_rho[0][d] = rho_inv[d].get(); mac_vec[0][d] = umac[d].get(); }
    std::array<LinOpBCType,3> bc{LinOpBCType::Periodic, LinOpBCType::Periodic, LinOpBCType::Periodic};
    if (macproj->needInitialization()) {
        LPInfo lp_info; lp_info.setMaxCoarseningLevel(100);
        macproj->initProjector(lp_info, inv_rho);
        macproj->setDomainBC(bc, bc);
    } else {
        macproj->updateBeta(inv_rho);
    }
    macproj->setUMAC(mac_vec);
    Vector<MultiFab*> phi{&mac_phi};
    macproj->project(phi, 1e-11, 1e-14);
    FILE* f = std::fopen(out, "wb");
    if (!f) return 2;
    int hdr[2] = {N, macproj->stats().n_vcycles};
    std::fwrite(hdr, sizeof(int), 2, f);
    // assemble the global arrays (a face shared by two boxes must agree in both)
    int mismatch = 0;
    for (int d = 0; d < 3; ++d) {
        const int e[3] = {N + (d == 0), N + (d == 1), N + (d == 2)};
        std::vector<double> U((size_t)e[0] * e[1] * e[2]); std::vector<char> seen(U.size(), 0);
        for (MFIter mfi(*umac[d]); mfi.isValid(); ++mfi) {
            FArrayBox& u = (*umac[d])[mfi]; const Box fb = mfi.validbox();
            for (int k = fb.lo[2]; k <= fb.hi[2]; ++k) for (int j = fb.lo[1]; j <= fb.hi[1]; ++j) for (int i = fb.lo[0]; i <= fb.hi[0]; ++i) {
                const size_t q = i + (size_t)e[0] * (j + (size_t)e[1] * k);
                if (seen[q] && U[q] != u(i,j,k)) ++mismatch;
                U[q] = u(i,j,k); seen[q] = 1;
            }
        }
        std::fwrite(U.data(), 8, U.size(), f);
    }
    if (mismatch) { std::fprintf(stderr, "mac shim: %d shared faces differ between boxes\n", mismatch); return 4; }
    std::vector<double> Pg((size_t)N * N * N);
    for (MFIter mfi(mac_phi); mfi.isValid(); ++mfi) {
        FArrayBox& p = mac_phi[mfi]; const Box vb = mfi.validbox();
        for (int k = vb.lo[2]; k <= vb.hi[2]; ++k) for (int j = vb.lo[1]; j <= vb.hi[1]; ++j) for (int i = vb.lo[0]; i <= vb.hi[0]; ++i)
            Pg[i + (size_t)N * (j + (size_t)N * k)] = p(i,j,k);
    }
    std::fwrite(Pg.data(), 8, Pg.size(), f);
    std::fclose(f);
    std::printf("mac shim: N=%d vcycles=%d\n", N, hdr[1]);
    return 0;
}
