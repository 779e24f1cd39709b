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
