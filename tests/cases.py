"""Seeded input states for the parity tests, shaped after BASELINE.json's configs
 at sizes the CPU oracle finishes in seconds."""
from __future__ import annotations

import numpy as np

from oracle_lib import (ALL_FIELDS, BConds, DIRICHLET, NEUMANN, PERIODIC, SYMMETRY, State)


def random_state(nx, ny, seed=0, rho_ratio=1000.0, lx=1.0, ly=1.0):
    """Every field random; densities jump between gas and liquid so both branches of
    hybrid_interp (src/FS.hpp:132-150) are taken."""
    rng = np.random.default_rng(seed)
    st = State(nx, ny, lx / nx, ly / ny, rho_gas=1.0, rho_liquid=rho_ratio, visc_gas=1e-3,
               visc_liquid=1e-2)
    for n in ALL_FIELDS:
        st.a[n][:] = rng.uniform(-1, 1, st.a[n].shape)
    for n in ("rho_u", "rho_v", "rho_u_old", "rho_v_old"):
        blob = rng.uniform(size=st.a[n].shape) < 0.5
        st.a[n][:] = np.where(blob, 1.0, rho_ratio) * (1 + 1e-4 * rng.uniform(size=st.a[n].shape))
    st.a["visc"][:] = np.abs(st.a["visc"]) * 1e-2 + 1e-3
    return st


def ellipse_vf(nx, ny, dx, dy, sharp=1.5):
    """Smoothed indicator of the ellipse (2(x-1/2))^2 + (y-1/2)^2 <= 0.25^2
    (test/StationaryDrop.cpp:46-52), ghost inclusive."""
    x = (np.arange(-1, nx + 1) + 0.5) * dx
    y = (np.arange(-1, ny + 1) + 0.5) * dy
    X, Y = np.meshgrid(x, y, indexing="ij")
    d = np.sqrt((2 * (X - 0.5)) ** 2 + (Y - 0.5) ** 2) - 0.25
    return np.clip(0.5 - d / (sharp * dx), 0.0, 1.0)


def drop_vf_quadrature(nx, ny, dx, dy, npt=16):
    """the initial volume fraction of test/StationaryDrop.cpp:47-52,209-212: npt-point Gauss quadrature per
    direction (src/Quadrature.hpp:37-66) of the indicator of the ellipse (2 (x - 1/2))^2 + (y - 1/2)^2 <= 1/16 over
    every cell incl. ghosts, divided by the cell area, then apply_neumann_bconds; ghost inclusive"""
    xi, w = np.polynomial.legendre.leggauss(npt)
    x = np.arange(-1, nx + 2) * dx
    y = np.arange(-1, ny + 2) * dy
    X = (x[1:] - x[:-1])[:, None] / 2.0 * xi[None, :] + (x[1:] + x[:-1])[:, None] / 2.0   # (nx + 2, npt)
    Y = (y[1:] - y[:-1])[:, None] / 2.0 * xi[None, :] + (y[1:] + y[:-1])[:, None] / 2.0
    inside = ((2.0 * (X[:, None, :, None] - 0.5)) ** 2 + (Y[None, :, None, :] - 0.5) ** 2 <= 0.25 ** 2)
    integral = np.einsum("ijab,a,b->ij", inside.astype(np.float64), w, w)
    vf = ((x[1:] - x[:-1])[:, None] / 2.0 * (y[1:] - y[:-1])[None, :] / 2.0 * integral) / (dx * dy)
    vf[0, :], vf[-1, :] = vf[1, :], vf[-2, :]
    vf[:, 0], vf[:, -1] = vf[:, 1], vf[:, -2]
    return vf


def cavity_bc():
    """scaling/LidDrivenFlow.cpp:33-41: no-slip walls, lid U = 1 on top."""
    return BConds(kind=(DIRICHLET,) * 4, u_val=(0.0, 0.0, 0.0, 1.0), v_val=(0.0,) * 4)


def channel_bc():
    """test/LaminarChannel.cpp:42-50: inlet U = 1, Neumann outlet, no-slip walls."""
    return BConds(kind=(DIRICHLET, NEUMANN, DIRICHLET, DIRICHLET), u_val=(1.0, 0.0, 0.0, 0.0),
                  v_val=(0.0,) * 4)


def couette_bc():
    """test/Couette.cpp:45-50: Neumann left / right, no-slip bottom, lid U = 1 on top."""
    return BConds(kind=(NEUMANN, NEUMANN, DIRICHLET, DIRICHLET), u_val=(0.0, 0.0, 0.0, 1.0), v_val=(0.0,) * 4)


def couette_state(oracle, seed=None):
    """test/Couette.cpp:21-41: 210 x 21 cells on 10 x 1, rho = 0.9, mu = 0.1, zero initial flow
    (seed: a small random velocity instead, so that the pressure solves have work to do)"""
    nx, ny = 210, 21
    st = State(nx, ny, 10.0 / nx, 1.0 / ny, rho_gas=0.9, rho_liquid=0.9, visc_gas=0.1, visc_liquid=0.1)
    oracle.calc_rho_visc_const(st)
    if seed is not None:
        rng = np.random.default_rng(seed)
        st.a["U"][:] = 0.1 * rng.uniform(-1, 1, st.a["U"].shape)
        st.a["V"][:] = 0.1 * rng.uniform(-1, 1, st.a["V"].shape)
    return st


def couette_l1_errors(U, nx, ny):
    """test/Couette.cpp:226-240: trapezoid L1 error of U against U_top y at x = NX/4, NX/2, 3NX/4"""
    ym = (np.arange(-1, ny + 1) + 0.5) / ny
    out = []
    for i in (nx // 4, nx // 2, 3 * nx // 4):
        d = np.abs(U[i + 1, :] - ym)
        out.append(float(np.sum((ym[1:] - ym[:-1]) * 0.5 * (d[1:] + d[:-1]))))
    return out


CHANNEL215_TOTAL_FLOW = 0.5  # (Y_MAX - Y_MIN) * U * RHO, test/PeriodicChannel.cpp:39, test/CustomDirichletChannel.cpp:39


def channel215_state(oracle, seed=None, backflow=False):
    """test/PeriodicChannel.cpp:21-39,147-149 = test/CustomDirichletChannel.cpp:21-40,146-148: 215 x 43 cells on
    5 x 1, rho = 0.5, mu = 1e-3, U = 1 on the interior faces, V = 0 (seed: a small random V on top;
    backflow: the last face rows carry U in [-0.5, 1] so that a clipped Neumann outlet has something to clip)"""
    nx, ny = 215, 43
    st = State(nx, ny, 5.0 / nx, 1.0 / ny, rho_gas=0.5, rho_liquid=0.5, visc_gas=1e-3, visc_liquid=1e-3)
    oracle.calc_rho_visc_const(st)
    st.a["U"][1:nx + 2, 1:ny + 1] = 1.0
    if seed is not None:
        rng = np.random.default_rng(seed)
        st.a["V"][1:nx + 1, 1:ny + 2] = 0.05 * rng.uniform(-1, 1, (nx, ny + 1))
        if backflow:
            st.a["U"][nx - 1:nx + 2, 1:ny + 1] = rng.uniform(-0.5, 1.0, (3, ny))
    return st


def periodic_channel_bc():
    """test/PeriodicChannel.cpp:46-52: periodic left / right, no-slip walls"""
    return BConds(kind=(PERIODIC, PERIODIC, DIRICHLET, DIRICHLET), u_val=(0.0,) * 4, v_val=(0.0,) * 4)


def custom_channel_bc(ny=43, clipped=False):
    """test/CustomDirichletChannel.cpp:46-56 (function inlet U(y) = dpdx / (2 mu) (y^2 - y), Neumann outlet,
    no-slip walls); clipped=True is scaling/Channel.cpp:48-59"""
    visc, rho = 1e-3, 0.5
    dpdx = -12.0 * visc * CHANNEL215_TOTAL_FLOW / rho
    ym = (np.arange(-1, ny + 1) + 0.5) / ny
    prof = dpdx / (2.0 * visc) * (ym * ym - ym)
    return BConds(kind=(DIRICHLET, NEUMANN, DIRICHLET, DIRICHLET), u_val=(0.0,) * 4, v_val=(0.0,) * 4,
                  u_prof=(prof, None, None, None), clipped=(False, clipped, False, False))


def channel215_cases():
    """name -> (bconds, ghost-column correction mode (include/fsgpu.h FSGPU_FLOW_*), total flow, backflow)"""
    return {"periodic215x43": (periodic_channel_bc(), 3, CHANNEL215_TOTAL_FLOW, False),   # test/PeriodicChannel.cpp
            "customdirichlet215x43": (custom_channel_bc(), 2, 0.0, False),             # test/CustomDirichletChannel.cpp
            "clippedchannel215x43": (custom_channel_bc(clipped=True), 2, 0.0, True)}   # scaling/Channel.cpp


def profile_bcs(nx, ny, seed=5):
    """two FlowBConds with function-pointer Dirichlet sides (src/BoundaryConditions.hpp:15-17) given
    as per-point tables: every side and component, and a mix with constants / Neumann"""
    rng = np.random.default_rng(seed)
    lr_u, lr_v, bt_u, bt_v = ny + 2, ny + 3, nx + 3, nx + 2
    full = BConds(kind=(DIRICHLET,) * 4, u_val=(0.0,) * 4, v_val=(0.0,) * 4,
                  u_prof=(rng.uniform(-1, 1, lr_u), rng.uniform(-1, 1, lr_u), rng.uniform(-1, 1, bt_u),
                          rng.uniform(-1, 1, bt_u)),
                  v_prof=(rng.uniform(-1, 1, lr_v), rng.uniform(-1, 1, lr_v), rng.uniform(-1, 1, bt_v),
                          rng.uniform(-1, 1, bt_v)))
    mixed = BConds(kind=(DIRICHLET, NEUMANN, DIRICHLET, DIRICHLET), u_val=(0.0, 0.0, 0.3, 1.0),
                   v_val=(0.1, 0.0, 0.0, -0.2),
                   u_prof=(rng.uniform(0, 1, lr_u), None, None, rng.uniform(0, 1, bt_u)),
                   v_prof=(None, None, rng.uniform(-1, 1, bt_v), None))
    return {"full": full, "mixed": mixed}


def periodic_bc():
    return BConds(kind=(PERIODIC,) * 4)


def cavity_state(oracle, nx, ny, drop=False, lx=1.0, ly=1.0, rho=1.0, mu=1e-3, seed=None):
    st = State(nx, ny, lx / nx, ly / ny, rho_gas=rho, rho_liquid=1000.0 if drop else rho,
               visc_gas=mu, visc_liquid=mu)
    if drop:
        vf = ellipse_vf(nx, ny, st.dx, st.dy)
        oracle.apply_neumann_bconds(vf)
        oracle.calc_rho_visc_vof(st, vf)
        st.vf = vf
    else:
        oracle.calc_rho_visc_const(st)
    st.a["rho_u_old"][:] = st.a["rho_u"]
    st.a["rho_v_old"][:] = st.a["rho_v"]
    if seed is not None:
        rng = np.random.default_rng(seed)
        st.a["U"][:] = 0.1 * rng.uniform(-1, 1, st.a["U"].shape)
        st.a["V"][:] = 0.1 * rng.uniform(-1, 1, st.a["V"].shape)
    return st


def tgv_state(oracle, n=128):
    """test/TaylorGreenVortex.cpp:21-48,109-113 (note: the test fills V over U's index range)."""
    L = 2 * np.pi
    st = State(n, n, L / n, L / n, rho_gas=0.9, rho_liquid=0.9, visc_gas=0.1, visc_liquid=0.1)
    oracle.calc_rho_visc_const(st)
    x = np.arange(-1, n + 2) * st.dx          # face coordinates
    xm = (np.arange(-1, n + 1) + 0.5) * st.dx  # centre coordinates
    st.a["U"][:] = np.sin(x)[:, None] * np.cos(xm)[None, :]
    V = np.zeros((n + 2, n + 3))
    # V(i, j) = -cos(xm(i)) sin(y(j)) for (i, j) in U's interior range [0, N] x [0, N)
    Vfull = -np.cos(xm)[:, None] * np.sin(x)[None, :]
    V[1:n + 2, 1:n + 1] = Vfull[1:n + 2, 1:n + 1]
    st.a["V"][:] = V
    return st


def twophase_inputs(st, seed=0, random_vf=False):
    """vf, curv, interface_length (cell fields) for calc_pressure_jump (src/FS.hpp:439-466): curvature
    scattered around 1/R = 4 and a positive interface length in mixed cells only (plus, for the
    random variant, isolated zero-length cells next to non-zero ones so both branches of the
    length-weighted average are taken)."""
    rng = np.random.default_rng(seed)
    shape = (st.nx + 2, st.ny + 2)
    if random_vf:
        vf = rng.uniform(0, 1, shape)
        ilen = np.where(rng.uniform(size=shape) < 0.5, 0.0, rng.uniform(0.1, 1.0, shape) * st.dx)
    else:
        vf = getattr(st, "vf", None)
        if vf is None:
            vf = ellipse_vf(st.nx, st.ny, st.dx, st.dy)
        ilen = np.where((vf > 1e-8) & (vf < 1 - 1e-8), rng.uniform(0.5, 1.4, shape) * st.dx, 0.0)
    curv = 4.0 * (1 + 0.05 * rng.uniform(-1, 1, shape))
    return np.ascontiguousarray(vf), curv, ilen


def ib_inputs(st, seed=0):
    """ib_corr_u_stag / ib_corr_v_stag (src/IB.hpp:62-107 computes lambda = (d - dist)/(dist d^2) near
    the wall): zero in the free flow, ~1/d^2 near the wall, inf inside the solid"""
    rng = np.random.default_rng(seed)
    out = []
    for n, d in (("U", st.dx), ("V", st.dy)):
        sh = st.a[n].shape
        r = rng.uniform(size=sh)
        lam = rng.uniform(0.1, 3.0, sh) / (d * d)
        out.append(np.where(r < 0.5, 0.0, np.where(r < 0.9, lam, np.inf)))
    return out


def to_gpu(F, st, bc=None, device=0):
    """oracle_lib.State -> fluidsolver_b200.FS (upload every field)."""
    fs = F.FS(st.nx, st.ny, st.dx, st.dy, st.rho_gas, st.rho_liquid, st.visc_gas, st.visc_liquid,
              st.sigma, device=device)
    for n in ALL_FIELDS:
        fs.upload(n.upper(), st.a[n])
    if bc is not None:
        for s in range(4):
            fs.set_bconds(s, bc.kind[s], bc.u_val[s], bc.v_val[s], bc.u_prof[s], bc.v_prof[s],
                          bc.clipped[s])
    return fs


def rel_inf(a, b):
    """||a - b||_inf / ||b||_inf (norm-relative, SURVEY.md section 7 'FMA contraction')"""
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def rel_l2(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))


def vel_err(U, V, Uref, Vref):
    """relative L2 error of the velocity FIELD (U, V) -- north_star's "velocity ... fields after N
    steps within 1e-10 relative L2"; one norm over both components, so a component that is
    ~1e-4 of the other (V in a channel) is not judged against its own tiny norm."""
    num = np.linalg.norm((U - Uref).ravel()) ** 2 + np.linalg.norm((V - Vref).ravel()) ** 2
    den = np.linalg.norm(Uref.ravel()) ** 2 + np.linalg.norm(Vref.ravel()) ** 2
    return float(np.sqrt(num / max(den, 1e-300)))


# ---- file formats of src/IO.hpp and src/VTKWriter.hpp, restated with numpy (checked against files
# ---- written by the unmodified reference: tests/golden/io_random24x10.npz) ------------------------
NPY_STATE_FILES = {"visc": "VISC", "p": "P", "p_jump_u_stag": "P_JUMP_U", "p_jump_v_stag": "P_JUMP_V",
                   "rho_u_stag": "RHO_U", "rho_v_stag": "RHO_V", "U": "U", "V": "V", "rho_u_stag_old": "RHO_U_OLD",
                   "rho_v_stag_old": "RHO_V_OLD", "U_old": "U_OLD", "V_old": "V_OLD"}


def grid_coords(st, x_min=0.0, y_min=0.0):
    """fs.x, fs.y of init_grid (src/FS.hpp:55-73) including ghost entries"""
    return (x_min + np.arange(-1, st.nx + 2, dtype=np.float64) * st.dx,
            y_min + np.arange(-1, st.ny + 2, dtype=np.float64) * st.dy)


def npy_bytes(a):
    """src/IO.hpp:25-95: version 1.0 header padded to a multiple of 64 bytes, then the C-order payload"""
    a = np.ascontiguousarray(a, dtype="<f8")
    h = "{'descr': '<f8', 'fortran_order': False, 'shape': (" + "".join(f"{n}, " for n in a.shape) + "), }"
    while (6 + 2 + 2 + len(h) + 1) % 64:
        h += " "
    h += "\n"
    return b"\x93NUMPY\x01\x00" + len(h).to_bytes(2, "little") + h.encode() + a.tobytes()


def vtk_bytes(t, x, y, scalars, vectors):
    """src/VTKWriter.hpp:39-152. x, y incl. ghosts; scalars [(name, cell array incl. ghosts)], vectors
    [(name, ax, ay)]; interior cells, j outer / i inner, big-endian doubles"""
    nx, ny = x.size - 3, y.size - 3
    be = lambda a: np.ascontiguousarray(a, dtype=">f8").tobytes()  # noqa: E731
    out = (f"# vtk DataFile Version 2.0\nState of FluidSolver at time t={t:g}\nBINARY\nDATASET STRUCTURED_GRID\n"
           f"DIMENSIONS {nx + 1} {ny + 1} 1\nPOINTS {(nx + 1) * (ny + 1)} double\n").encode()
    pts = np.zeros((ny + 1, nx + 1, 3))
    pts[:, :, 0] = x[1:nx + 2][None, :]
    pts[:, :, 1] = y[1:ny + 2][:, None]
    out += be(pts) + f"\n\nCELL_DATA {nx * ny}\n".encode()
    for name, a in scalars:
        out += f"SCALARS {name} double 1\nLOOKUP_TABLE default\n".encode() + be(a[1:-1, 1:-1].T) + b"\n\n"
    for name, ax, ay in vectors:
        v = np.zeros((ny, nx, 3))
        v[:, :, 0], v[:, :, 1] = ax[1:-1, 1:-1].T, ay[1:-1, 1:-1].T
        out += f"VECTORS {name} double\n".encode() + be(v) + b"\n\n"
    return out


# ---- the convergence study of scaling/Channel.cpp (non-periodic variant) -----------------------------
def channel_study_case(oracle, n):
    """scaling/Channel.cpp:19-59,81-83,152-154: NY = 2^n + 1, NX = 5 NY on 5 x 1, rho = 0.5, mu = 1e-3,
    U = 1, function inlet + clipped Neumann outlet + no-slip walls; -> (state, bconds)"""
    ny = (1 << n) + 1
    nx = 5 * ny
    st = State(nx, ny, 5.0 / nx, 1.0 / ny, rho_gas=0.5, rho_liquid=0.5, visc_gas=1e-3, visc_liquid=1e-3)
    oracle.calc_rho_visc_const(st)
    st.a["U"][1:nx + 2, 1:ny + 1] = 1.0
    bc = custom_channel_bc(ny, clipped=True)
    oracle.apply_velocity_bconds(st, bc)
    return st, bc


def channel_study_errors(p, U, V, nx, ny, dx, visc=1e-3):
    """scaling/Channel.cpp:253-300 -> (pressure_error, dpdx_error, U_error, V_error)"""
    P = p[1:, 1:]    # P[i, j] = p(i, j), i >= 0
    pe = np.abs(P[:nx, :ny] - P[:nx, :1]).sum() / (nx * ny)
    jm = ny // 2
    ref_dpdx = (P[nx // 2 + 1, jm] - P[nx // 2, jm]) / dx
    dpdx = (P[1:nx, jm] - P[0:nx - 1, jm]) / dx
    de = np.abs(ref_dpdx - dpdx).sum() / (nx - 1)
    i = nx // 2
    ym = (np.arange(ny) + 0.5) / ny
    g = (P[i, :ny] - p[i, 1:ny + 1]) / dx    # (p(i, j) - p(i - 1, j)) / dx
    diff = np.abs(U[i + 1, 1:ny + 1] - g / (2 * visc) * (ym * ym - ym))
    ue = float(np.sum((ym[1:] - ym[:-1]) * 0.5 * (diff[1:] + diff[:-1])))
    ve = np.abs(V[1:nx + 1, 1:ny + 2]).sum() / (nx * (ny + 1))
    return float(pe), float(de), ue, float(ve)


def run_channel_study(adjust_dt, step, t_end=20.0):
    """the time loop of scaling/Channel.cpp:169-172,235: adjust_dt() -> dt, step(dt); -> number of steps"""
    t, n = 0.0, 0
    while t < t_end:
        dt = min(adjust_dt(), t_end - t)
        step(dt)
        t += dt
        n += 1
    return n


def sample_points(st, n=400, seed=33):
    """query points for eval_flow_field_at / bilinear_interpolate: uniform over a box 1.5 cells larger than the
    domain (so every clamped branch is taken), plus points exactly on cell centres and on the first / last one"""
    rng = np.random.default_rng(seed)
    lx, ly = st.nx * st.dx, st.ny * st.dy
    px = rng.uniform(-1.5 * st.dx, lx + 1.5 * st.dx, n)
    py = rng.uniform(-1.5 * st.dy, ly + 1.5 * st.dy, n)
    xm = (np.arange(st.nx) + 0.5) * st.dx
    ym = (np.arange(st.ny) + 0.5) * st.dy
    ex = np.array([xm[0], xm[-1], xm[3], xm[0], xm[-1], 0.5 * lx])
    ey = np.array([ym[0], ym[-1], ym[5], ym[-1], ym[0], ym[2]])
    return np.concatenate([px, ex]), np.concatenate([py, ey])


def conserved_known_answer():
    """the reference's own test of calc_conserved_quantities (test/Operators.cpp:217-283): rho = 12 x + x y^2,
    rho U = rho sin(x + y), rho V = rho cos(x + y) on [2, 4] x [-1, 3], 200 x 300 cells; -> (state, expected (mass,
    momentum_x, momentum_y) by Gauss-Legendre quadrature, tolerance 10 min(dx, dy)^2)"""
    nx, ny, dx, dy = 200, 300, 2.0 / 200, 4.0 / 300
    st = State(nx, ny, dx, dy)
    x = 2.0 + np.arange(-1, nx + 2) * dx
    y = -1.0 + np.arange(-1, ny + 2) * dy
    xm, ym = (x[:-1] + x[1:]) / 2, (y[:-1] + y[1:]) / 2
    rho = lambda X, Y: 12.0 * X + X * Y * Y  # noqa: E731
    XU, YU = np.meshgrid(x, ym, indexing="ij")
    XV, YV = np.meshgrid(xm, y, indexing="ij")
    st.a["rho_u"][:], st.a["U"][:] = rho(XU, YU), np.sin(XU + YU)
    st.a["rho_v"][:], st.a["V"][:] = rho(XV, YV), np.cos(XV + YV)
    gx, gw = np.polynomial.legendre.leggauss(96)
    QX, QY = np.meshgrid(3.0 + gx, 1.0 + 2.0 * gx, indexing="ij")   # [2, 4] x [-1, 3]
    W = np.outer(gw, gw) * 2.0
    expected = tuple(float((W * f).sum()) for f in (rho(QX, QY), rho(QX, QY) * np.sin(QX + QY), rho(QX, QY) * np.cos(QX + QY)))
    return st, expected, 10.0 * min(dx, dy) ** 2


def twophase_quadrature_case(oracle, n=128):
    """BASELINE configs[3] with the reference's own inputs: volume fraction by Gauss quadrature of the ellipse
    indicator (test/StationaryDrop.cpp:47-52,209-212), rho 1 : 1000 and mu from calc_rho / calc_visc(vf, fs),
    sigma = 1/20, curvature 1 / R of the circle R = 0.25 and interface length dx on the interface cells;
    lid-driven boundary conditions; -> (state, bconds, (vf, curv, interface_length))"""
    st = cavity_state(oracle, n, n, drop=True, seed=1)
    vf = drop_vf_quadrature(n, n, st.dx, st.dy)
    st.vf = vf
    oracle.calc_rho_visc_vof(st, vf)
    for k in ("rho_u", "rho_v"):
        st.a[k + "_old"][:] = st.a[k]
    st.sigma = 0.05
    bc = cavity_bc()
    oracle.apply_velocity_bconds(st, bc)
    band = ((vf > 1e-8) & (vf < 1.0 - 1e-8)).astype(np.float64)
    return st, bc, (vf, band * 4.0, band * st.dx)


def channel215_measures(periodic, p, U, V, nx=215, ny=43, dx=5.0 / 215, visc=1e-3):
    """the end-of-run measures of test/PeriodicChannel.cpp:264-337 (periodic) and test/CustomDirichletChannel.cpp:
    252-331: (max deviation of p along y, max deviation of dp/dx from its reference, worst trapezoid L1 error of the
    U profile over all stations x_i, max |V| on the interior faces)"""
    P = p[1:, 1:]
    ym = (np.arange(-1, ny + 1) + 0.5) / ny
    dev_p = float(np.abs(P[:nx, :ny] - P[:nx, :1]).max())
    dpdx = (P[1:nx, ny // 2] - P[0:nx - 1, ny // 2]) / dx
    dpdx_ref = (P[nx // 2 + 1, ny // 2] - P[nx // 2, ny // 2]) / dx if periodic else -12.0 * visc
    dev_dpdx = float(np.abs(dpdx - dpdx_ref).max())
    l1 = 0.0
    for i in range(nx + 1):
        g = (p[i + 1, :] - p[i, :]) / dx if periodic else dpdx_ref
        d = np.abs(U[i + 1, :] - g / (2 * visc) * (ym * ym - ym))
        l1 = max(l1, float(np.sum((ym[1:] - ym[:-1]) * 0.5 * (d[1:] + d[:-1]))))
    return dev_p, dev_dpdx, l1, float(np.abs(V[1:nx + 1, 1:ny + 2]).max())


def ib_channel_case(oracle, n=64):
    """scaling/LuchiniIBChannel.cpp:17-75 at n x n cells: 5 x 5 domain, channel of height 1 between immersed walls
    (solid for y <= 2 and y >= 3), rho = 10, mu = 1e-2, Poiseuille inlet inside the channel, Neumann outlet;
    -> (state, bconds); the wall correction fields come from the reference (tests/golden/run_ibchannel64.npz)"""
    visc, rho, p0, length, height, off = 1e-2, 10.0, 0.2, 5.0, 1.0, 2.0
    st = State(n, n, 5.0 / n, 5.0 / n, rho_gas=rho, rho_liquid=rho, visc_gas=visc, visc_liquid=visc)
    oracle.calc_rho_visc_const(st)
    ym = (np.arange(-1, n + 1) + 0.5) * st.dy
    yo = ym - off
    prof = np.where((ym <= off) | (ym >= off + height), 0.0, p0 * height / (visc * length) * yo * (1.0 - yo / height))
    # half the developed profile plus a seeded perturbation: the pressure solves have a right-hand side well above
    # rounding (on the developed x-independent flow the divergence vanishes and PCG would iterate on noise)
    rng = np.random.default_rng(17)
    st.a["U"][:] = 0.5 * prof[None, :] + 0.02 * rng.uniform(-1, 1, st.a["U"].shape)
    st.a["V"][:] = 0.02 * rng.uniform(-1, 1, st.a["V"].shape)
    bc = BConds(kind=(DIRICHLET, NEUMANN, DIRICHLET, DIRICHLET), u_val=(0.0,) * 4, v_val=(0.0,) * 4,
                u_prof=(prof, None, None, None))
    oracle.apply_velocity_bconds(st, bc)
    return st, bc
