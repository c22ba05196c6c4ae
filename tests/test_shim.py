"""The C++ shim (include/NodalProjectorB200.H) driven the way incflo::ApplyProjection drives amrex::NodalProjector,
compiled against the mock AMReX types (AMReX itself is unavailable here) and linked with
  * the emulation build of the C ABI on CPU, or
  * libnodalproj.so on the GPU box (host fabs -> staged inside np_project),
then compared with the oracle."""
import os
import subprocess

import numpy as np
import pytest

import nphelp
from incflo_b200 import capi, problems
from oracle.oracle import Oracle, TILED

ROOT = nphelp.ROOT


def _build(libpath, exe, src="shim_driver.cpp"):
    libdir, libname = os.path.dirname(libpath), os.path.basename(libpath)[3:-3]
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "tests", "mock_amrex"),
                           os.path.join(ROOT, "tests", "shim", src), "-o", exe, "-L" + libdir, "-l" + libname,
                           "-Wl,-rpath," + libdir])


def _run_mac_and_check(libpath, tmp_path, n, max_grid=None):
    """include/MacProjectorB200.H driven like incflo::compute_MAC_projected_velocities, against the MAC oracle."""
    from oracle.mac import MacOracle, face_shapes
    exe, out = str(tmp_path / "mac_shim_driver"), str(tmp_path / "mac_out.bin")
    _build(libpath, exe, "mac_shim_driver.cpp")
    subprocess.check_call([exe, str(n), out] + ([str(max_grid)] if max_grid else []))
    raw = np.fromfile(out, dtype=np.uint8)
    hdr, body = raw[:8].view(np.int32), raw[8:].view(np.float64)
    shapes = face_shapes((n, n, n))
    u, pos = [], 0
    for s in shapes:
        cnt = int(np.prod(s)); u.append(body[pos:pos + cnt].reshape(s)); pos += cnt
    phi = body[pos:].reshape(n, n, n)
    prob = problems.mac_double_shear((n, n, n))
    o = MacOracle((n, n, n))
    o.set_beta(*prob["beta"]); o.set_umac(*prob["umac"])
    so = o.project(1e-11, 1e-14)
    l2 = lambda a, b: np.linalg.norm(a - b) / np.linalg.norm(b)
    assert hdr[0] == n and hdr[1] == so.n_vcycles
    for a, b in zip(u, o.get_umac()):
        assert l2(a, b) < 1e-11
    po = o.get_phi()
    assert l2(phi - phi.mean(), po - po.mean()) < 1e-10


@pytest.mark.parametrize("max_grid", [None, 8])
def test_mac_shim_on_emulation(tmp_path, max_grid):
    nphelp.emu_lib()
    _run_mac_and_check(nphelp.EMU_LIB, tmp_path, 16, max_grid)


@pytest.mark.gpu
def test_mac_shim_on_gpu(tmp_path):
    _run_mac_and_check(capi.LIB_PATH, tmp_path, 32, 16)


def _run_diffusion_and_check(libpath, tmp_path, n):
    """include/DiffusionSolverB200.H driven like DiffusionScalarOp::diffuse_scalar, against the oracle's solve_abec."""
    from oracle.mac import MacOracle
    exe, out = str(tmp_path / "diffusion_shim_driver"), str(tmp_path / "diff_out.bin")
    _build(libpath, exe, "diffusion_shim_driver.cpp")
    subprocess.check_call([exe, str(n), out])
    raw = np.fromfile(out, dtype=np.uint8)
    hdr, tr = raw[:8].view(np.int32), raw[8:].view(np.float64).reshape(n, n, n)
    g = (np.arange(n) + 0.5) / n
    x, y, z = g[None, None, :], g[None, :, None], g[:, None, None]
    t0 = np.broadcast_to(x * y * (1.0 - z), (n, n, n)).copy()
    rho = np.broadcast_to(1.0 + 0.5 * np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y), (n, n, n)).copy()
    zf = np.arange(n + 1) / n
    eta = [np.broadcast_to((0.02 + 0.01 * np.cos(2 * np.pi * z)), (n, n, n + 1)).copy(),
           np.broadcast_to((0.02 + 0.01 * np.cos(2 * np.pi * z)), (n, n + 1, n)).copy(),
           np.broadcast_to((0.02 + 0.01 * np.cos(2 * np.pi * zf))[:, None, None], (n + 1, n, n)).copy()]
    o = MacOracle((n, n, n), (0, 0, 0), (1, 1, 1), (1, 2, 0), (2, 2, 0), maxorder=2)
    o.set_scalars(1.0, 0.01); o.set_acoef(rho); o.set_beta(*eta); o.set_phi(t0)
    so = o.solve_abec(rho * t0, 1e-11, 1e-14, dict(xlo=np.ones((n, n))))
    assert so.status == 0 and hdr[0] == n and hdr[1] == so.n_vcycles
    assert np.linalg.norm(tr - o.get_phi()) / np.linalg.norm(o.get_phi()) < 1e-10


def test_diffusion_shim_on_emulation(tmp_path):
    nphelp.emu_lib()
    _run_diffusion_and_check(nphelp.EMU_LIB, tmp_path, 16)


@pytest.mark.gpu
def test_diffusion_shim_on_gpu(tmp_path):
    _run_diffusion_and_check(capi.LIB_PATH, tmp_path, 32)


def _run_and_check(libpath, tmp_path, n, max_grid=None):
    exe = str(tmp_path / "shim_driver")
    out = str(tmp_path / "out.bin")
    _build(libpath, exe)
    subprocess.check_call([exe, str(n), out] + (["100", str(max_grid)] if max_grid else []))
    raw = np.fromfile(out, dtype=np.uint8)
    hdr = raw[:8].view(np.int32)
    body = raw[8:].view(np.float64)
    nc, nn = n ** 3, (n + 1) ** 3
    u = body[:3 * nc].reshape(3, n, n, n)
    phi = body[3 * nc:3 * nc + nn].reshape(n + 1, n + 1, n + 1)
    gp = body[3 * nc + nn:].reshape(3, n, n, n)
    prob = problems.taylor_green((n, n, n))
    o = Oracle((n, n, n), smoother=TILED, tile=(63, 31, 1 << 30), tile_z_target=-1)
    o.set_sigma(prob["sigma"]); o.set_vel(prob["vel"])
    so = o.project(1e-11, 1e-14)
    l2 = lambda a, b: np.linalg.norm(a - b) / np.linalg.norm(b)
    assert hdr[0] == n and hdr[1] == so.n_vcycles
    po = o.get_phi()
    assert l2(u, o.get_vel()) < 1e-11 and l2(gp, o.get_gradphi()) < 1e-10 and l2(phi - phi.mean(), po - po.mean()) < 1e-10


def test_shim_on_emulation(tmp_path):
    nphelp.emu_lib()
    _run_and_check(nphelp.EMU_LIB, tmp_path, 16)


def test_shim_on_emulation_multibox(tmp_path):
    """amr.max_grid_size = 8 on 16^3: eight boxes on the rank, walked with MFIter."""
    nphelp.emu_lib()
    _run_and_check(nphelp.EMU_LIB, tmp_path, 16, 8)


@pytest.mark.gpu
def test_shim_on_gpu(tmp_path):
    _run_and_check(capi.LIB_PATH, tmp_path, 32)


@pytest.mark.gpu
def test_shim_on_gpu_multibox(tmp_path):
    _run_and_check(capi.LIB_PATH, tmp_path, 32, 16)


def test_shim_abort_on_mlmg_failure(tmp_path):
    nphelp.emu_lib()
    exe = str(tmp_path / "shim_driver")
    _build(nphelp.EMU_LIB, exe)
    p = subprocess.run([exe, "16", str(tmp_path / "o.bin"), "1"], capture_output=True, text=True)
    assert p.returncode == 3 and "MLMG failed to converge" in p.stderr
