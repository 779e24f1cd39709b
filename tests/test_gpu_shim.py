"""The C++ host layer (include/shim/FluidSolverGPU.hpp): examples/lid_driven_cavity.cpp is the
reference's scaling/LidDrivenFlow.cpp loop written against the shim.  Its result must equal the
same call sequence driven through the ctypes binding (same library, same kernels) bit for bit, and
that sequence is pinned to the oracle by test_gpu_solver.py."""
import os
import re
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "examples", "lid_driven_cavity")


def test_cpp_driver_matches_binding(F):
    if not os.path.exists(EXE):
        from fluidsolver_b200 import build
        build.build_examples()
    nsteps, n = 3, 256
    r = subprocess.run([EXE, str(nsteps)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"checksum N (\d+) steps (\d+) U2 (\S+) V2 (\S+) P2 (\S+)", r.stdout)
    assert m and int(m.group(1)) == n and int(m.group(2)) == nsteps
    cu, cv, cp = (float(m.group(k)) for k in (3, 4, 5))
    its_cpp = [int(x) for x in re.findall(r"p_iter (\d+)", r.stdout)]

    fs = F.FS(n, n, 1.0 / n, 1.0 / n, 1.0, 1.0, 1e-3, 1e-3)
    fs.calc_rho_visc_const()
    for s, u in enumerate((0.0, 0.0, 0.0, 1.0)):
        fs.set_bconds(s, F.DIRICHLET, u, 0.0)
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=50)
    ps.set_pressure_operator()
    fs.apply_velocity_bconds()
    its = []
    for _ in range(nsteps):
        dt = fs.adjust_dt(0.9, 0.1)
        fs.save_old_state()
        tot = 0
        for _ in range(5):
            fs.calc_mid_time(); fs.calc_dmomdt(); fs.update_velocity(dt); fs.apply_velocity_bconds()
            fs.calc_divergence(); ps.set_pressure_rhs(dt)
            _, it = ps.solve()
            tot += it
            fs.shift_pressure_to_zero(); fs.correct_velocity(dt)
        its.append(tot)
    U, V, P = fs.download("U"), fs.download("V"), fs.download("P")
    su = float((U[1:n + 2, 1:n + 1] ** 2).sum())
    sv = float((V[1:n + 1, 1:n + 2] ** 2).sum())
    sp = float(((P[1:n + 1, 1:n + 1] - P[1, 1]) ** 2).sum())
    ps.close(); fs.close()
    assert its_cpp == its
    assert cu == pytest.approx(su, rel=1e-12) and cv == pytest.approx(sv, rel=1e-12)
    assert cp == pytest.approx(sp, rel=1e-10)


def test_cpp_two_phase_driver_matches_binding(F):
    """examples/two_phase_drop.cpp (the test/StationaryDrop.cpp sub-iteration on the shim, frozen VOF)
    vs the same un-fused call sequence through the ctypes binding"""
    exe = os.path.join(ROOT, "examples", "two_phase_drop")
    if not os.path.exists(exe):
        from fluidsolver_b200 import build
        build.build_examples()
    nsteps, n = 2, 128
    r = subprocess.run([exe, str(nsteps)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"checksum N (\d+) steps (\d+) U2 (\S+) V2 (\S+) RHOI (\S+)", r.stdout)
    assert m and int(m.group(1)) == n and int(m.group(2)) == nsteps
    cu, cv, cr = (float(m.group(k)) for k in (3, 4, 5))
    its_cpp = [int(x) for x in re.findall(r"p_iter (\d+)", r.stdout)]
    mass_cpp = [float(x) for x in re.findall(r"mass (\S+)", r.stdout)]

    dx = 1.0 / n
    xm = (np.arange(-1, n + 1) * dx + np.arange(0, n + 2) * dx) / 2  # (x(i) + x(i+1)) / 2 as init_grid
    X, Y = np.meshgrid(xm, xm, indexing="ij")
    d = np.sqrt(4.0 * (X - 0.5) * (X - 0.5) + (Y - 0.5) * (Y - 0.5)) - 0.25
    vf = np.clip(0.5 - d / (1.5 * dx), 0.0, 1.0)
    curv = np.full_like(vf, 4.0)
    ilen = np.where((vf > 1e-8) & (vf < 1.0 - 1e-8), dx, 0.0)
    fs = F.FS(n, n, dx, dx, 1.0, 1000.0, 1e-3, 1e-3, sigma=0.05)
    fs.upload("VF", vf)
    fs.calc_rho_visc_vof()
    fs.save_old_density()
    for s, u in enumerate((0.0, 0.0, 0.0, 1.0)):
        fs.set_bconds(s, F.DIRICHLET, u, 0.0)
    ps = F.LinearSolver(fs, tol=1e-10, max_iter=200)
    fs.apply_velocity_bconds()
    its, mass = [], []
    for step in range(nsteps):
        dt = fs.adjust_dt(0.9, 0.1)
        fs.save_old_state()
        tot = 0
        for sub in range(3):
            fs.calc_mid_time(); fs.calc_drhodt(); fs.update_density(dt)
            fs.apply_neumann_bconds("RHO_U"); fs.apply_neumann_bconds("RHO_V")
            fs.calc_dmomdt(); fs.update_velocity(dt); fs.apply_velocity_bconds()
            fs.calc_divergence()
            if step == 0 and sub == 0:
                fs.upload("CURV", curv); fs.upload("INTERFACE_LENGTH", ilen)
                fs.calc_pressure_jump()
            fs.capillary_divergence(dt)
            ps.set_pressure_operator(); ps.set_pressure_rhs(dt)
            _, it = ps.solve()
            tot += it
            fs.shift_pressure_to_zero(); fs.correct_velocity(dt)
        its.append(tot)
        fs.interpolate_cell_centres()
        mass.append(fs.calc_conserved_quantities()[0])
    U, V, R = fs.download("U"), fs.download("V"), fs.download("RHOI")
    su = float((U[1:n + 2, 1:n + 1] ** 2).sum())
    sv = float((V[1:n + 1, 1:n + 2] ** 2).sum())
    sr = float(R[1:n + 1, 1:n + 1].sum())
    ps.close(); fs.close()
    assert its_cpp == its
    assert mass_cpp == pytest.approx(mass, rel=1e-14)
    assert cu == pytest.approx(su, rel=1e-12) and cv == pytest.approx(sv, rel=1e-12)
    assert cr == pytest.approx(sr, rel=1e-12)
