"""Synthetic inputs for the nodal projection.

The formulas are the reference's initial conditions, evaluated at cell centres with numpy:
  taylor_green      /root/reference/src/prob/prob_init_fluid.cpp:150-160
  double_shear      /root/reference/src/prob/prob_init_fluid.cpp:470-475   (probtype 21)
  rayleigh_taylor   /root/reference/src/prob/prob_init_fluid.cpp:295-321   (density; sigma = dt/rho)
plus the perturbations SURVEY.md 8d adds so that the projection is non-trivial.
Arrays are C-order [k, j, i]; velocity is [3, k, j, i].
"""
import numpy as np

PERIODIC, DIRICHLET, NEUMANN = 0, 1, 2
TWOPI = 2.0 * 3.1415926535897932


def cell_centres(ncell, prob_lo, prob_hi, koff=0, nz_local=None):
    """Cell-centre coordinate arrays (z, y, x) broadcastable to [k, j, i].  koff/nz_local select a z-slab."""
    nx, ny, nz = ncell
    dx = [(prob_hi[d] - prob_lo[d]) / ncell[d] for d in range(3)]
    nzl = nz if nz_local is None else nz_local
    x = prob_lo[0] + (np.arange(nx) + 0.5) * dx[0]
    y = prob_lo[1] + (np.arange(ny) + 0.5) * dx[1]
    z = prob_lo[2] + (np.arange(koff, koff + nzl) + 0.5) * dx[2]
    return z[:, None, None], y[None, :, None], x[None, None, :]


def _perturb(x, y, z, Lz=1.0):
    return 0.1 * np.sin(TWOPI * x) * np.sin(TWOPI * y) * np.sin(TWOPI * z / Lz)


def taylor_green(ncell, prob_lo=(0, 0, 0), prob_hi=(1, 1, 1), **kw):
    """Config 1: TGV + perturbation, sigma = 1, periodic."""
    z, y, x = cell_centres(ncell, prob_lo, prob_hi, **kw)
    p = _perturb(x, y, z)
    u = np.sin(TWOPI * x) * np.cos(TWOPI * y) + p
    v = -np.cos(TWOPI * x) * np.sin(TWOPI * y) + p
    w = 0.0 * x * y + p
    shape = np.broadcast_shapes(z.shape, y.shape, x.shape)
    vel = np.stack([np.broadcast_to(a, shape) for a in (u, v, w)]).astype(np.float64)
    return dict(vel=np.ascontiguousarray(vel), sigma=np.ones(shape), bclo=(0, 0, 0), bchi=(0, 0, 0),
                prob_lo=tuple(prob_lo), prob_hi=tuple(prob_hi))


def double_shear(ncell, prob_lo=(0, 0, 0), prob_hi=(1, 1, 1), zwave=False, **kw):
    """Configs 2/3: double shear layer (probtype 21), sigma = 1, periodic.
    zwave adds w = 0.05 sin(2 pi x) sin(2 pi z / Lz) (config 3, so z is non-trivial across slabs)."""
    z, y, x = cell_centres(ncell, prob_lo, prob_hi, **kw)
    shape = np.broadcast_shapes(z.shape, y.shape, x.shape)
    u = np.tanh(30.0 * (0.25 - np.abs(y - 0.5)))
    v = 0.05 * np.sin(TWOPI * x)
    Lz = prob_hi[2] - prob_lo[2]
    w = 0.05 * np.sin(TWOPI * x) * np.sin(TWOPI * z / Lz) if zwave else 0.0 * x
    vel = np.stack([np.broadcast_to(a, shape) for a in (u, v, w)]).astype(np.float64)
    return dict(vel=np.ascontiguousarray(vel), sigma=np.ones(shape), bclo=(0, 0, 0), bchi=(0, 0, 0),
                prob_lo=tuple(prob_lo), prob_hi=tuple(prob_hi))


def rayleigh_taylor(ncell, prob_lo=(0, 0, 0), prob_hi=(0.5, 0.5, 1.0), dt=1.0, **kw):
    """Config 4: variable density (ratio 4), periodic x/y, slip walls in z (Neumann); sigma = dt/rho.
    u* = gravity kick (0, 0, -0.1 dt) + perturbation."""
    z, y, x = cell_centres(ncell, prob_lo, prob_hi, **kw)
    shape = np.broadcast_shapes(z.shape, y.shape, x.shape)
    rho_1, rho_2, width = 0.5, 2.0, 0.005
    splitx, splity = 0.5 * (prob_lo[0] + prob_hi[0]), 0.5 * (prob_lo[1] + prob_hi[1])
    L_x = prob_hi[0] - prob_lo[0]
    r2d = np.minimum(np.hypot(x - splitx, y - splity), 0.5 * L_x)
    pertheight = 0.5 - 0.01 * np.cos(2.0 * np.pi * r2d / L_x)
    rho = rho_1 + ((rho_2 - rho_1) / 2.0) * (1.0 + np.tanh((z - pertheight) / width))
    rho = np.broadcast_to(rho, shape)
    Lx, Ly, Lz = [prob_hi[d] - prob_lo[d] for d in range(3)]
    p = 0.1 * np.sin(TWOPI * x / Lx) * np.sin(TWOPI * y / Ly) * np.sin(TWOPI * z / Lz)
    vel = np.stack([np.broadcast_to(p, shape), np.broadcast_to(p, shape),
                    np.broadcast_to(-0.1 * dt + p, shape)]).astype(np.float64)
    return dict(vel=np.ascontiguousarray(vel), sigma=np.ascontiguousarray(dt / rho), bclo=(0, 0, NEUMANN),
                bchi=(0, 0, NEUMANN), prob_lo=tuple(prob_lo), prob_hi=tuple(prob_hi))


def random_fields(ncell, seed=0):
    """Kernel-level parity inputs: phi ~ N(0,1) nodal, sigma ~ U[0.5, 2]."""
    rng = np.random.default_rng(seed)
    nx, ny, nz = ncell
    phi = rng.standard_normal((nz + 1, ny + 1, nx + 1))
    sigma = rng.uniform(0.5, 2.0, (nz, ny, nx))
    vel = rng.standard_normal((3, nz, ny, nx))
    return dict(phi=phi, sigma=sigma, vel=vel)


PROBLEMS = {"taylor_green": taylor_green, "double_shear": double_shear, "rayleigh_taylor": rayleigh_taylor}


def double_shear_torch(ncell, prob_hi, koff, nz_local, device, ngrow=1, zwave=True):
    """Config 3 field (double_shear(zwave=True)) built directly on `device` for one z-slab, as the
    [3, nz+2g, ny+2g, nx+2g] storage of a velocity fab with zero ghosts.  Same formulas as double_shear."""
    import torch
    nx, ny, nz = ncell
    g = ngrow
    dx = [prob_hi[d] / ncell[d] for d in range(3)]
    f64 = dict(dtype=torch.float64, device=device)
    x = ((torch.arange(nx, **f64) + 0.5) * dx[0])[None, None, :]
    y = ((torch.arange(ny, **f64) + 0.5) * dx[1])[None, :, None]
    z = ((torch.arange(koff, koff + nz_local, **f64) + 0.5) * dx[2])[:, None, None]
    vel = torch.zeros((3, nz_local + 2 * g, ny + 2 * g, nx + 2 * g), **f64)
    inner = (slice(g, g + nz_local), slice(g, g + ny), slice(g, g + nx))
    vel[0][inner] = torch.tanh(30.0 * (0.25 - torch.abs(y - 0.5))).expand(nz_local, ny, nx)
    vel[1][inner] = (0.05 * torch.sin(TWOPI * x)).expand(nz_local, ny, nx)
    if zwave:
        vel[2][inner] = 0.05 * torch.sin(TWOPI * x) * torch.sin(TWOPI * z / prob_hi[2])
    return vel


# ---- MAC projection inputs: the same flows sampled at face centres, beta = 1/rho on faces ----
def _face_coords(ncell, prob_lo, prob_hi, d):
    """(z, y, x) coordinate arrays of the faces of direction d, broadcastable to the face array [k, j, i]."""
    n = list(ncell)
    dx = [(prob_hi[q] - prob_lo[q]) / ncell[q] for q in range(3)]
    c = []
    for q in range(3):
        if q == d:
            c.append(prob_lo[q] + np.arange(n[q] + 1) * dx[q])
        else:
            c.append(prob_lo[q] + (np.arange(n[q]) + 0.5) * dx[q])
    return c[2][:, None, None], c[1][None, :, None], c[0][None, None, :]


def mac_double_shear(ncell, prob_lo=(0, 0, 0), prob_hi=(1, 1, 1)):
    """Double shear layer (/root/reference/src/prob/prob_init_fluid.cpp:470-475) + the z wave and perturbation of the
    nodal configs at face centres; constant density (beta = 1), periodic."""
    L = [prob_hi[d] - prob_lo[d] for d in range(3)]
    umac, beta = [], []
    for d in range(3):
        z, y, x = _face_coords(ncell, prob_lo, prob_hi, d)
        shape = np.broadcast_shapes(z.shape, y.shape, x.shape)
        p = 0.1 * np.sin(TWOPI * x / L[0]) * np.sin(TWOPI * y / L[1]) * np.sin(TWOPI * z / L[2])
        f = (np.tanh(30.0 * (0.25 - np.abs(y / L[1] - 0.5))), 0.05 * np.sin(TWOPI * x / L[0]),
             0.05 * np.sin(TWOPI * x / L[0]) * np.sin(TWOPI * z / L[2]))[d]
        umac.append(np.ascontiguousarray(np.broadcast_to(f + p, shape), dtype=np.float64))
        beta.append(np.ones(shape))
    return dict(umac=umac, beta=beta, bclo=(0, 0, 0), bchi=(0, 0, 0), prob_lo=tuple(prob_lo), prob_hi=tuple(prob_hi))


def mac_rayleigh_taylor(ncell, prob_lo=(0, 0, 0), prob_hi=(0.5, 0.5, 1.0)):
    """Rayleigh-Taylor density (/root/reference/src/prob/prob_init_fluid.cpp:295-321) on faces: beta = 1/rho, ratio 4;
    periodic x/y, walls in z (Neumann, no flow through the wall faces); gravity kick + perturbation."""
    L = [prob_hi[d] - prob_lo[d] for d in range(3)]
    rho_1, rho_2, width = 0.5, 2.0, 0.005
    sx, sy = 0.5 * (prob_lo[0] + prob_hi[0]), 0.5 * (prob_lo[1] + prob_hi[1])
    umac, beta = [], []
    for d in range(3):
        z, y, x = _face_coords(ncell, prob_lo, prob_hi, d)
        shape = np.broadcast_shapes(z.shape, y.shape, x.shape)
        xw = prob_lo[0] + np.mod(x - prob_lo[0], L[0])       # the periodic image face carries the same density
        yw = prob_lo[1] + np.mod(y - prob_lo[1], L[1])
        r2d = np.minimum(np.hypot(xw - sx, yw - sy), 0.5 * L[0])
        h = 0.5 - 0.01 * np.cos(2.0 * np.pi * r2d / L[0])
        rho = rho_1 + ((rho_2 - rho_1) / 2.0) * (1.0 + np.tanh((z - h) / width))
        p = 0.1 * np.sin(TWOPI * x / L[0]) * np.sin(TWOPI * y / L[1]) * np.sin(TWOPI * z / L[2])
        f = np.broadcast_to((p, p, -0.1 + p)[d], shape).astype(np.float64).copy()
        if d == 2:
            f[0] = 0.0
            f[-1] = 0.0
        umac.append(np.ascontiguousarray(f))
        beta.append(np.ascontiguousarray(np.broadcast_to(1.0 / rho, shape), dtype=np.float64))
    return dict(umac=umac, beta=beta, bclo=(0, 0, NEUMANN), bchi=(0, 0, NEUMANN), prob_lo=tuple(prob_lo),
                prob_hi=tuple(prob_hi))


def mac_double_shear_walls(ncell, prob_lo=(0, 0, 0), prob_hi=(1, 1, 1)):
    """mac_double_shear between slip walls in z (Neumann, no flow through the wall faces): uniform beta with physical z faces."""
    d = mac_double_shear(ncell, prob_lo, prob_hi)
    d["umac"] = [np.array(u) for u in d["umac"]]
    d["umac"][2][0] = 0.0
    d["umac"][2][-1] = 0.0
    d["bclo"], d["bchi"] = (0, 0, NEUMANN), (0, 0, NEUMANN)
    return d


MAC_PROBLEMS = {"double_shear": mac_double_shear, "rayleigh_taylor": mac_rayleigh_taylor, "double_shear_walls": mac_double_shear_walls}


def mac_problem_torch(name, ncell, prob_hi, device, ngrow=0, koff=0, nz_local=None):
    """The MAC_PROBLEMS fields built directly on `device` as face-fab storage [1, nz(+1)+2g, ny(+1)+2g, nx(+1)+2g]
    (zero ghosts): (umac[3], beta[3], bclo, bchi).  Same formulas as mac_double_shear / mac_rayleigh_taylor.
    koff / nz_local select a z-slab of the domain `ncell` (multi-GPU): planes koff .. koff + nz_local."""
    import torch
    f64 = dict(dtype=torch.float64, device=device)
    g = ngrow
    L = [float(prob_hi[d]) for d in range(3)]
    dx = [L[d] / ncell[d] for d in range(3)]
    umac, beta = [], []
    for d in range(3):
        nzl = ncell[2] if nz_local is None else nz_local
        n = [ncell[0], ncell[1], nzl]
        n[d] += 1
        c = [(torch.arange(n[q], **f64) + (koff if q == 2 else 0) + (0.0 if q == d else 0.5)) * dx[q] for q in range(3)]
        x, y, z = c[0][None, None, :], c[1][None, :, None], c[2][:, None, None]
        p = 0.1 * torch.sin(TWOPI * x / L[0]) * torch.sin(TWOPI * y / L[1]) * torch.sin(TWOPI * z / L[2])
        if name == "double_shear":
            f = (torch.tanh(30.0 * (0.25 - torch.abs(y / L[1] - 0.5))), 0.05 * torch.sin(TWOPI * x / L[0]),
                 0.05 * torch.sin(TWOPI * x / L[0]) * torch.sin(TWOPI * z / L[2]))[d] + p
            b = torch.ones((n[2], n[1], n[0]), **f64)
        else:
            xw, yw = torch.remainder(x, L[0]), torch.remainder(y, L[1])
            r2d = torch.clamp(torch.hypot(xw - 0.5 * L[0], yw - 0.5 * L[1]).expand(1, n[1], n[0]), max=0.5 * L[0])
            h = 0.5 - 0.01 * torch.cos(2.0 * np.pi * r2d / L[0])
            rho = 0.5 + 0.75 * (1.0 + torch.tanh((z - h) / 0.005))
            b = 1.0 / rho
            f = ((p, p, -0.1 + p)[d]).expand(n[2], n[1], n[0]).clone()
            if d == 2:
                if koff == 0:
                    f[0] = 0.0
                if koff + nzl == ncell[2]:
                    f[-1] = 0.0
        uf = torch.zeros((1, n[2] + 2 * g, n[1] + 2 * g, n[0] + 2 * g), **f64)
        bf = torch.zeros_like(uf)
        inner = (0, slice(g, g + n[2]), slice(g, g + n[1]), slice(g, g + n[0]))
        uf[inner] = f.expand(n[2], n[1], n[0])
        bf[inner] = b.expand(n[2], n[1], n[0])
        umac.append(uf)
        beta.append(bf)
        del f, b, p
    bc = ((0, 0, 0), (0, 0, 0)) if name == "double_shear" else ((0, 0, NEUMANN), (0, 0, NEUMANN))
    return umac, beta, bc[0], bc[1]


def diffusion_problem_torch(n, dev, variable_eta):
    """Scalar diffusion workload of bench.py --path diffusion and tests/test_fullsize.py: Rayleigh-Taylor density (ratio 4, the
    interface of /root/reference/src/prob/prob_init_fluid.cpp's init_rayleigh_taylor), a Gaussian blob plus a sine mode as the
    tracer, eta uniform or varying in z.  Returns (rho [1,n,n,n], tracer [1,n,n,n], [bx, by, bz] face arrays [1,...])."""
    import torch
    f64 = dict(dtype=torch.float64, device=dev)
    g = (torch.arange(n, **f64) + 0.5) / n
    x, y, z = g[None, None, :], g[None, :, None], g[:, None, None]
    r2d = torch.clamp(torch.hypot(x - 0.5, y - 0.5), max=0.5)
    h = 0.5 - 0.01 * torch.cos(2 * 3.141592653589793 * r2d)
    rho = (0.5 + 0.75 * (1.0 + torch.tanh((z - h) / 0.005))).expand(n, n, n).contiguous()
    tracer = (torch.exp(-((x - 0.5) ** 2 + (y - 0.5) ** 2 + (z - 0.4) ** 2) / 0.02)
              + 0.1 * torch.sin(6.283185307179586 * x) * torch.sin(6.283185307179586 * y) * torch.sin(6.283185307179586 * z)).contiguous()
    mu0 = 1.0e-3
    faces = []
    for d in range(3):
        shp = [n, n, n]; shp[2 - d] += 1
        if variable_eta:
            zz = (torch.arange(shp[0], **f64) + (0.0 if d == 2 else 0.5)) / n
            faces.append((mu0 * (1.0 + 0.5 * torch.tanh((zz - 0.5) / 0.05)))[:, None, None].expand(*shp).contiguous()[None])
        else:
            faces.append(torch.full([1] + shp, mu0, **f64))
    return rho[None].contiguous(), tracer[None].contiguous(), faces
