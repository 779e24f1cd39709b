"""Timestep throughput of the BASELINE.json configurations bench.py does not time (bench.py times
configs[2], the 8192^2 cavity): configs[1] channel 4096 x 1024 (outflow correction, 5 sub-iterations),
configs[3] variable-density drop 4096^2 (operator rebuilt every sub-iteration, 3 sub-iterations) and
the same drop with consistent mass transport + capillary block switched on.  One JSON line each:
ms/step over `steps` timed steps after `warmup` (CUDA events on the library's stream), MCell-updates/s,
PCG iterations per solve.

usage: python scripts/bench_configs.py [steps] [warmup] [scale]    (scale 2 halves every grid dimension)
       python scripts/bench_configs.py --cpu [scale]   the UNMODIFIED reference (oracle/_ref/libfsref.so, OpenMP + HYPRE,
                                                       all host cores) on the channel and the drop at the same sizes:
                                                       1 warm-up + 2 timed timesteps each, one JSON line each
"""
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CPU = len(sys.argv) > 1 and sys.argv[1] in ("--cpu", "--cpu-child")
if not CPU:
    import fluidsolver_b200 as F  # noqa: E402
    import torch  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 and not CPU else 20
warmup = int(sys.argv[2]) if len(sys.argv) > 2 and not CPU else 5
scale = int(sys.argv[2]) if CPU and len(sys.argv) > 2 else (int(sys.argv[3]) if len(sys.argv) > 3 and not CPU else 1)


def ellipse_vf(nx, ny, dx, dy, npt=16, rows=64):
    """the initial volume fraction of test/StationaryDrop.cpp:47-52,209-212: 16-point Gauss quadrature per direction
    (src/Quadrature.hpp:37-66) of the indicator of the ellipse (2 (x - 1/2))^2 + (y - 1/2)^2 <= 1/16 over every cell,
    divided by the cell area, then apply_neumann_bconds (the restatement tests/cases.py pins against the reference,
    evaluated in blocks of rows)"""
    xi, w = np.polynomial.legendre.leggauss(npt)
    x = np.arange(-1, nx + 2) * dx
    y = np.arange(-1, ny + 2) * dy
    X = (x[1:] - x[:-1])[:, None] / 2.0 * xi[None, :] + (x[1:] + x[:-1])[:, None] / 2.0
    Y = (y[1:] - y[:-1])[:, None] / 2.0 * xi[None, :] + (y[1:] + y[:-1])[:, None] / 2.0
    vf = np.zeros((nx + 2, ny + 2))
    ylo, yhi = (Y.min(axis=1) - 0.5), (Y.max(axis=1) - 0.5)
    for i0 in range(0, nx + 2, rows):
        xs = X[i0:i0 + rows]
        # columns the ellipse can touch for these rows; everything else is 0 (outside) -- the ellipse is convex,
        # so cells fully inside are found by the same quadrature at no special cost
        cols = np.nonzero((np.minimum(np.abs(ylo), np.abs(yhi)) <= 0.25) | (ylo * yhi <= 0))[0]
        if cols.size == 0 or np.abs(2.0 * (xs - 0.5)).min() > 0.25:
            continue
        inside = ((2.0 * (xs[:, None, :, None] - 0.5)) ** 2 + (Y[cols][None, :, None, :] - 0.5) ** 2 <= 0.25 ** 2)
        vf[i0:i0 + rows, cols] = np.einsum("ijab,a,b->ij", inside.astype(np.float64), w, w)
    vf *= (x[1:] - x[:-1])[:, None] / 2.0 * (y[1:] - y[:-1])[None, :] / 2.0 / (dx * dy)
    vf[0, :], vf[-1, :] = vf[1, :], vf[-2, :]
    vf[:, 0], vf[:, -1] = vf[:, 1], vf[:, -2]
    return vf


def timed(name, fs, ps, nsub, flags, cfl=0.9, dt_max=0.1):
    for _ in range(warmup):
        fs.step(ps, fs.adjust_dt(cfl, dt_max), nsub, flags)
    st = torch.cuda.ExternalStream(fs.stream())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fs.sync()
    e0.record(st)
    its, res = 0, 0.0
    for _ in range(steps):
        res, it = fs.step(ps, fs.adjust_dt(cfl, dt_max), nsub, flags)
        its += it
    e1.record(st)
    fs.sync()
    ms = e0.elapsed_time(e1) / steps
    um, vm, dm = fs.monitor()
    assert res < 1e-6, f"{name}: the last solve did not converge (residual {res})"
    print(json.dumps({"config": name, "nx": fs.nx, "ny": fs.ny, "steps": steps, "warmup": warmup,
                      "ms_per_step": round(ms, 3), "mcell_updates_per_s": round(fs.nx * fs.ny / ms / 1e3, 1),
                      "pcg_iters_per_solve": round(its / (steps * nsub), 2), "last_residual": res,
                      "levels": ps.num_levels, "max_U": um, "max_V": vm, "max_div": dm}), flush=True)


def channel():
    nx, ny = 4096 // scale, 1024 // scale
    fs = F.FS(nx, ny, 100.0 / nx, 1.0 / ny, 0.5, 0.5, 1e-3, 1e-3)
    fs.calc_rho_visc_const()
    fs.fill("U", 1.0)
    fs.set_bconds(F.LEFT, F.DIRICHLET, 1.0, 0.0)
    fs.set_bconds(F.RIGHT, F.NEUMANN)
    fs.set_bconds(F.BOTTOM, F.DIRICHLET, 0.0, 0.0)
    fs.set_bconds(F.TOP, F.DIRICHLET, 0.0, 0.0)
    fs.apply_velocity_bconds()
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=50, max_levels=40)
    ps.set_pressure_operator()
    timed("channel", fs, ps, 5, F.STEP_OUTFLOW_FIX)
    ps.close()
    fs.close()


def drop(twophase):
    n = 4096 // scale
    fs = F.FS(n, n, 1.0 / n, 1.0 / n, 1.0, 1000.0, 1e-3, 1e-3, sigma=0.05 if twophase else 0.0)
    vf = ellipse_vf(n, n, 1.0 / n, 1.0 / n)
    fs.upload("VF", vf)
    fs.calc_rho_visc_vof()
    fs.save_old_density()
    for s, u in enumerate((0.0, 0.0, 0.0, 1.0)):
        fs.set_bconds(s, F.DIRICHLET, u, 0.0)
    fs.apply_velocity_bconds()
    flags = F.STEP_REBUILD_OPERATOR
    if twophase:  # curvature 1 / R of a circle of the drop's height, interface cells flagged
        band = ((vf > 1e-8) & (vf < 1 - 1e-8)).astype(np.float64)
        fs.upload("CURV", band / 0.25)
        fs.upload("INTERFACE_LENGTH", band)
        flags |= F.STEP_DENSITY | F.STEP_CAPILLARY
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=50, max_levels=40)
    ps.set_pressure_operator()
    timed("drop_twophase" if twophase else "drop", fs, ps, 3, flags)
    ps.close()
    fs.close()


def cpu_child():
    """the reference's own path on the host cores (the protocol of bench.py's CPU legs: threads bound, all cores)"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from cases import cavity_bc, cavity_state, channel_bc
    from oracle_lib import Oracle, Ref
    o = Oracle()
    cores = os.cpu_count() or 1
    Ref.lib().fsref_set_threads(cores)
    for name in ("channel", "drop"):
        if name == "channel":
            nx, ny, nsub, kw = 4096 // scale, 1024 // scale, 5, dict(outflow_fix=1)
            st, bc = cavity_state(o, nx, ny, lx=100.0, ly=1.0, rho=0.5), channel_bc()
            st.a["U"][:] = 1.0
        else:
            nx = ny = 4096 // scale
            nsub, kw = 3, dict(rebuild=1)
            st, bc = cavity_state(o, nx, ny), cavity_bc()
            st.rho_liquid = 1000.0
            vf = ellipse_vf(nx, ny, st.dx, st.dy)
            o.calc_rho_visc_vof(st, vf)
            st.a["rho_u_old"][:] = st.a["rho_u"]
            st.a["rho_v_old"][:] = st.a["rho_v"]
        o.apply_velocity_bconds(st, bc)
        r = Ref(nx, ny)
        r.load(st)
        r.set_bconds(bc)
        r.solver_create(1e-6, 50)
        r.solver_set_operator()
        r.step(r.adjust_dt(0.9, 0.1), nsub, **kw)
        t0 = time.perf_counter()
        its, k = 0, 2
        for _ in range(k):
            it, res = r.step(r.adjust_dt(0.9, 0.1), nsub, **kw)
            its += it
        ms = (time.perf_counter() - t0) / k * 1e3
        print(json.dumps({"config": name, "impl": "reference", "nx": nx, "ny": ny, "steps": k, "warmup": 1,
                          "ms_per_step": round(ms, 1), "mcell_updates_per_s": round(nx * ny / ms / 1e3, 3),
                          "pcg_iters_per_solve": round(its / (k * nsub), 2), "last_residual": res,
                          "cores": cores, "solver": "HYPRE Struct PCG + PFMG, PS_PFMG_MAX_LEVELS 16"}), flush=True)
        del r


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--cpu":
        env = dict(os.environ, OMP_PROC_BIND="close", OMP_PLACES="cores")
        env.pop("OMP_NUM_THREADS", None)
        sys.exit(subprocess.call([sys.executable, os.path.abspath(__file__), "--cpu-child", str(scale)], env=env))
    if len(sys.argv) > 1 and sys.argv[1] == "--cpu-child":
        cpu_child()
        sys.exit(0)
    channel()
    drop(False)
    drop(True)
