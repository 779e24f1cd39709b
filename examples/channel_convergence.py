#!/usr/bin/env python
"""The grid-convergence study of scaling/Channel.cpp (non-periodic variant) on the B200 path.

For NY = 2^n + 1, NX = 5 NY (n = 3 .. 7 as in the reference; --max-n 10 adds the three sizes the
reference leaves commented out, scaling/Channel.cpp:353-355): Poiseuille inlet profile given as a
function of y, clipped Neumann outlet, no-slip walls, the dy-weighted outflow correction after the
boundary conditions (scaling/Channel.cpp:208-213), NUM_SUBITER = 2, tol 1e-6, 50 iterations,
CFL 0.9, T_END = 20; then the four error measures of :253-300, printed in the reference's CSV format
(:303-314).  Every step runs through the C ABI of libfsgpu.so (fsgpu_step); only the final p, U, V
come back to the host for the error measures.

    python examples/channel_convergence.py [--csv] [--max-n 7] [--t-end 20] [--vtk DIR]
"""
import argparse
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

import fluidsolver_b200 as F  # noqa: E402

VISC, RHO, U_INIT, TOTAL_FLOW = 1e-3, 0.5, 1.0, 0.5


def errors(p, U, V, nx, ny, dx):
    """scaling/Channel.cpp:253-300"""
    P = p[1:, 1:]
    pe = np.abs(P[:nx, :ny] - P[:nx, :1]).sum() / (nx * ny)
    jm = ny // 2
    ref = (P[nx // 2 + 1, jm] - P[nx // 2, jm]) / dx
    de = np.abs(ref - (P[1:nx, jm] - P[0:nx - 1, jm]) / dx).sum() / (nx - 1)
    i = nx // 2
    ym = (np.arange(ny) + 0.5) / ny
    g = (P[i, :ny] - p[i, 1:ny + 1]) / dx
    diff = np.abs(U[i + 1, 1:ny + 1] - g / (2 * VISC) * (ym * ym - ym))
    ue = np.sum((ym[1:] - ym[:-1]) * 0.5 * (diff[1:] + diff[:-1]))
    ve = np.abs(V[1:nx + 1, 1:ny + 2]).sum() / (nx * (ny + 1))
    return pe, de, ue, ve


def run(n, t_end, vtk_dir=None):
    ny = (1 << n) + 1
    nx = 5 * ny
    dx, dy = 5.0 / nx, 1.0 / ny
    fs = F.FS(nx, ny, dx, dy, RHO, RHO, VISC, VISC)
    fs.calc_rho_visc_const()
    U = np.zeros(F.field_shape("U", nx, ny))
    U[1:nx + 2, 1:ny + 1] = U_INIT
    fs.upload("U", U)
    ym = (np.arange(-1, ny + 1) + 0.5) / ny
    dpdx = -12.0 * VISC * TOTAL_FLOW / RHO
    fs.set_bconds(F.LEFT, F.DIRICHLET, 0.0, 0.0, u_prof=dpdx / (2.0 * VISC) * (ym * ym - ym))
    fs.set_bconds(F.RIGHT, F.NEUMANN, clipped=True)
    fs.set_bconds(F.BOTTOM, F.DIRICHLET, 0.0, 0.0)
    fs.set_bconds(F.TOP, F.DIRICHLET, 0.0, 0.0)
    fs.apply_velocity_bconds()
    fs.set_flow_control(F.FLOW_OUTFLOW_FIX_DY)
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=50)
    ps.set_pressure_operator()
    x, y = fs.grid_coords()
    fs.sync()
    t0 = time.perf_counter()
    t, steps, iters, frame = 0.0, 0, 0, 0
    while t < t_end:
        dt = min(fs.adjust_dt(0.9, 0.1), t_end - t)
        _, it = fs.step(ps, dt, 2, F.STEP_OUTFLOW_FIX)
        t += dt
        steps += 1
        iters += it
        if vtk_dir and t >= frame:  # DT_WRITE = 1
            fs.calc_divergence()
            fs.interpolate_cell_centres()
            fs.write_vtk(os.path.join(vtk_dir, f"n{n}_state_{frame:06d}.vtk"), t, x, y,
                         [("pressure", "P"), ("divergence", "DIV")], [("velocity", "UI", "VI")])
            frame += 1
    fs.sync()
    secs = time.perf_counter() - t0
    out = errors(fs.download("P"), fs.download("U"), fs.download("V"), nx, ny, dx)
    ps.close()
    fs.close()
    return (nx, ny, dx, dy) + out + (secs, steps, iters)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--csv", action="store_true")
    ap.add_argument("--min-n", type=int, default=3)
    ap.add_argument("--max-n", type=int, default=7)
    ap.add_argument("--t-end", type=float, default=20.0)
    ap.add_argument("--vtk", default=None, help="directory for one VTK frame per time unit")
    a = ap.parse_args()
    if a.vtk:
        os.makedirs(a.vtk, exist_ok=True)
    if a.csv:
        print("NX,NY,dx,dy,pressure_error,dpdx_error,U_error,V_error,runtime_s")
    for n in range(a.min_n, a.max_n + 1):
        nx, ny, dx, dy, pe, de, ue, ve, secs, steps, iters = run(n, a.t_end, a.vtk)
        if a.csv:
            print(f"{nx},{ny},{dx:.6e},{dy:.6e},{pe:.6e},{de:.6e},{ue:.6e},{ve:.6e},{secs:.6e}", flush=True)
        else:
            print(f"NX = {nx}, NY = {ny}: pressure_error {pe:.6e}, dpdx_error {de:.6e}, U_error {ue:.6e}, "
                  f"V_error {ve:.6e}; {steps} steps, {iters} PCG iterations, {secs:.3f} s "
                  f"({nx * ny * steps / secs / 1e6:.2f} MCell-updates/s)", flush=True)


if __name__ == "__main__":
    main()
