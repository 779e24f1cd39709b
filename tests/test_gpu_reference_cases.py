"""The reference's own integration tests, run on the GPU path with the reference's pass criteria
(SURVEY.md section 4), and BASELINE.json's configs at full size through size-independent properties.

* test/TaylorGreenVortex.cpp  (configs[0]): 128^2, periodic, T = 5 -> L1 error of U and V against the
  analytic decaying vortex <= 3 h^2 (:193-240)
* test/LaminarChannel.cpp     (configs[1]): 750 x 15, T = 60 -> pressure constant across the channel
  and dp/dx constant for x > 60 (1e-4), U profile vs dpdx/(2 mu)(y^2 - y), trapezoid L1 <= 7.5e-3
  (:270-340), outflow/inflow mass balance; and the channel scaled to 4096 x 1024 on one B200, held on the
  analytic Poiseuille solution
* configs[2], configs[3] at full size (8192^2 cavity, 4096^2 rho-ratio-1000 drop): the projection
  property -- after a step the discrete divergence is at the solver tolerance -- iteration counts in
  the reference's range, finite fields, conserved mass.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(fs, ps, F, t_end, cfl, dt_max, nsub, flags=0, max_steps=10**6):
    t, steps, its = 0.0, 0, 0
    while t < t_end and steps < max_steps:
        dt = min(fs.adjust_dt(cfl, dt_max), t_end - t)
        res, it = fs.step(ps, dt, nsub, flags)
        assert np.isfinite(res)
        t += dt
        steps += 1
        its += it
    return t, steps, its


def test_taylor_green_vortex_analytic(F):
    """test/TaylorGreenVortex.cpp:20-48,109-113,127-181,193-240"""
    n, visc, rho, t_end = 128, 1e-1, 0.9, 5.0
    L = 2.0 * np.pi
    dx = L / n
    fs = F.FS(n, n, dx, dx, rho, rho, visc, visc)
    fs.calc_rho_visc_const()
    for s in range(4):
        fs.set_bconds(s, F.PERIODIC)
    x = np.arange(-1, n + 2) * dx
    xm = (x[:-1] + x[1:]) / 2
    U = np.zeros((n + 3, n + 2))
    V = np.zeros((n + 2, n + 3))
    # the test initialises both components over U's interior index range (:109-113)
    U[1:n + 2, 1:n + 1] = (np.sin(x)[:, None] * np.cos(xm)[None, :])[1:n + 2, 1:n + 1]
    V[1:n + 2, 1:n + 1] = (-np.cos(xm)[:, None] * np.sin(x)[None, :])[1:n + 2, 1:n + 1]
    fs.upload("U", U)
    fs.upload("V", V)
    fs.apply_velocity_bconds()
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=500)
    ps.set_pressure_operator()
    t, steps, its = _run(fs, ps, F, t_end, 0.5, 1e-2, 2)
    assert t == pytest.approx(t_end) and steps >= 500
    U, V = fs.download("U"), fs.download("V")
    f = np.exp(-2.0 * visc / rho * t_end)
    ua = np.sin(x[1:n + 2])[:, None] * np.cos(xm[1:n + 1])[None, :] * f
    va = -np.cos(xm[1:n + 1])[:, None] * np.sin(x[1:n + 2])[None, :] * f
    l1u = np.abs(U[1:n + 2, 1:n + 1] - ua).sum() * dx * dx
    l1v = np.abs(V[1:n + 1, 1:n + 2] - va).sum() * dx * dx
    tol = 3.0 * dx * dx
    print(f"\n[TGV 128^2, T=5] {steps} steps, {its / (2 * steps):.1f} its/solve, L1 error U {l1u:.3e} V {l1v:.3e} "
          f"(tolerance 3 h^2 = {tol:.3e})")
    assert np.isfinite(U).all() and np.isfinite(V).all()
    assert l1u <= tol and l1v <= tol
    ps.close()
    fs.close()


def _channel(F, nx, ny, t_end, max_levels=0):
    """test/LaminarChannel.cpp:22-51: 100 x 1 channel, inlet U = 1, Neumann outlet with the outflow
    fix, no-slip walls, rho = 0.5, mu = 1e-3"""
    visc, rho = 1e-3, 0.5
    dx, dy = 100.0 / nx, 1.0 / ny
    fs = F.FS(nx, ny, dx, dy, rho, rho, visc, visc)
    fs.calc_rho_visc_const()
    fs.set_bconds(F.LEFT, F.DIRICHLET, 1.0, 0.0)
    fs.set_bconds(F.RIGHT, F.NEUMANN)
    fs.set_bconds(F.BOTTOM, F.DIRICHLET, 0.0, 0.0)
    fs.set_bconds(F.TOP, F.DIRICHLET, 0.0, 0.0)
    fs.fill("U", 1.0)   # :151-153: U = U_IN on the interior, then the boundary conditions
    fs.apply_velocity_bconds()
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=50, max_levels=max_levels)
    ps.set_pressure_operator()
    t, steps, its = _run(fs, ps, F, t_end, 0.9, 1e-1, 5, F.STEP_OUTFLOW_FIX)
    U, p, RU = fs.download("U"), fs.download("P"), fs.download("RHO_U")
    ps.close()
    fs.close()
    return U, p, RU, dx, dy, steps, its


def _check_poiseuille(U, p, dx, dy, nx, ny, tol_p=1e-4, tol_u=7.5e-3, visc=1e-3):
    """the three checks of test/LaminarChannel.cpp:270-340; arrays are ghost inclusive (index + 1)"""
    x = np.arange(-1, nx + 2) * dx
    ym = (np.arange(-1, ny + 1) + 0.5) * dy
    i60 = int(np.argmax(x > 60.0))  # as the test: the position in the ghost-inclusive x array, used as i
    P = lambda i, j: p[i + 1, j + 1]            # noqa: E731
    # pressure constant along y for x > 60
    pi = p[i60 + 1:nx + 1, 1:ny + 1]
    dev_p = np.abs(pi - pi[:, :1]).max()
    # dp/dx constant after x = 60 (mid-height)
    jm = ny // 2
    col = p[:, jm + 1]
    dpdx = (col[i60 + 2:nx + 2] - col[i60 + 1:nx + 1]) / dx   # i = i60+1 .. NX  (p.extent(0) = NX)
    dpdx = dpdx[:nx - (i60 + 1)]
    ref = (P(i60 + 1, jm) - P(i60, jm)) / dx
    dev_dpdx = np.abs(dpdx - ref).max()
    # U profile against dpdx / (2 mu) (y^2 - y), trapezoid L1 over ym incl. ghosts
    worst = 0.0
    for xt in (60.0, 70.0, 80.0, 90.0):
        i = int(xt / 100.0 * (nx + 1))
        d = (p[i + 1, :] - p[i, :]) / dx
        ua = d / (2 * visc) * (ym * ym - ym)
        diff = np.abs(U[i + 1, :] - ua)
        worst = max(worst, float(np.sum((ym[1:] - ym[:-1]) * 0.5 * (diff[1:] + diff[:-1]))))
    return dev_p, dev_dpdx, worst, ref


def test_laminar_channel_poiseuille(F):
    """the reference's own case: 750 x 15 to T = 60"""
    nx, ny = 750, 15
    U, p, RU, dx, dy, steps, its = _channel(F, nx, ny, 60.0)
    dev_p, dev_dpdx, l1, dpdx = _check_poiseuille(U, p, dx, dy, nx, ny)
    inflow, outflow = (RU[0] * U[0]).sum(), (RU[nx + 2] * U[nx + 2]).sum()
    print(f"\n[channel 750x15, T=60] {steps} steps, {its / (5 * steps):.1f} its/solve, dp/dx {dpdx:.5f}, "
          f"|p - p(y0)| {dev_p:.2e}, |dpdx - ref| {dev_dpdx:.2e}, U profile L1 {l1:.3e}, "
          f"mass error {outflow - inflow:.2e}")
    assert np.isfinite(U).all() and np.isfinite(p).all()
    assert dev_p <= 1e-4 and dev_dpdx <= 1e-4 and l1 <= 7.5e-3
    assert abs(outflow - inflow) <= 1e-8 * abs(inflow)   # the outflow fix (:55-66, monitor criterion 1e-8)
    # fully developed flow: dp/dx = -12 mu U / H^2 = -0.012; at NY = 15 the scheme's wall closure gives -0.0106
    assert -0.0125 <= dpdx <= -0.0100


def test_channel_4096x1024_poiseuille(F):
    """BASELINE.json configs[1]: the channel scaled to 4096 x 1024 on one B200, verified against the
    analytic Poiseuille profile.  At dy = 1/1024 the viscous limit of adjust_dt (4 mu / (dy^2 rho),
    src/FS.hpp:116-121) gives dt = 1.07e-4, so developing the flow from a uniform inlet (T = 60 in the
    reference's 750 x 15 test) would take 560 000 steps; instead the analytic solution
    U(y) = 6 y (1 - y), the fixed point of the scheme up to the O(dy^2) wall closure, is prescribed at
    the inlet (function-pointer Dirichlet variant, src/BoundaryConditions.hpp:15-17) and as initial
    state, and must be held: after a few steps dp/dx = -12 mu U / H^2 and U stays on the parabola."""
    nx, ny, visc, rho = 4096, 1024, 1e-3, 0.5
    dx, dy = 100.0 / nx, 1.0 / ny
    ym = (np.arange(-1, ny + 1) + 0.5) * dy
    prof = 6.0 * ym * (1.0 - ym)
    fs = F.FS(nx, ny, dx, dy, rho, rho, visc, visc)
    fs.calc_rho_visc_const()
    fs.set_bconds(F.LEFT, F.DIRICHLET, 0.0, 0.0, u_prof=prof)
    fs.set_bconds(F.RIGHT, F.NEUMANN)
    fs.set_bconds(F.BOTTOM, F.DIRICHLET, 0.0, 0.0)
    fs.set_bconds(F.TOP, F.DIRICHLET, 0.0, 0.0)
    fs.upload("U", np.broadcast_to(prof, (nx + 3, ny + 2)).copy())
    fs.apply_velocity_bconds()
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=50, max_levels=40)
    ps.set_pressure_operator()
    nsteps, its = 4, 0
    for _ in range(nsteps):
        dt = fs.adjust_dt(0.9, 1e-1)
        assert dt == pytest.approx(0.9 / (4 * visc / (dy * dy * rho)), rel=1e-12)
        res, it = fs.step(ps, dt, 5, F.STEP_OUTFLOW_FIX)
        assert np.isfinite(res)
        its += it
    U, p, RU = fs.download("U"), fs.download("P"), fs.download("RHO_U")
    ps.close()
    fs.close()
    inflow, outflow = (RU[0] * U[0]).sum(), (RU[nx + 2] * U[nx + 2]).sum()
    i0, i1 = int(20.0 / dx), int(80.0 / dx)
    pm = p[i0:i1, 1:ny + 1]
    dpdx = (pm[1:, ny // 2] - pm[:-1, ny // 2]) / dx
    err_u = np.abs(U[i0:i1, 1:ny + 1] - prof[None, 1:ny + 1]).max()
    print(f"\n[channel 4096x1024 on the Poiseuille solution] {nsteps} steps, {its / (5 * nsteps):.1f} its/solve, "
          f"dp/dx in [{dpdx.min():.6f}, {dpdx.max():.6f}] (analytic {-12 * visc:.6f}), max|U - U_analytic| {err_u:.2e}, "
          f"|p - p(y0)| {np.abs(pm - pm[:, :1]).max():.2e}, mass error {outflow - inflow:.2e}")
    assert np.isfinite(U).all() and np.isfinite(p).all()
    assert abs(outflow - inflow) <= 1e-8 * abs(inflow)
    assert np.abs(dpdx + 12 * visc).max() <= 0.02 * 12 * visc
    assert np.abs(pm - pm[:, :1]).max() <= 1e-4
    assert err_u <= 1e-4
    assert its / (5 * nsteps) <= 50


def _projection_ratio(fs, ps, dt, n, rebuild=False):
    """one UN-fused sub-iteration; returns the spread (std) of the interior divergence before and
    after the projection.  The reference removes the mean of the rhs over ALL cells of the ghost
    ring-inclusive system (src/LinearSystem.hpp:141-147), so the projected interior divergence is a
    constant plus the solver residual: the spread is what the solve must kill."""
    fs.momentum_update(dt)
    fs.apply_velocity_bconds()
    fs.calc_divergence()
    before = fs.download("DIV")[1:n + 1, 1:n + 1].std()
    if rebuild:
        ps.set_pressure_operator()
    ps.set_pressure_rhs(dt)
    res, it = ps.solve()
    fs.shift_pressure_to_zero()
    fs.correct_velocity(dt)
    fs.calc_divergence()
    after = fs.download("DIV")[1:n + 1, 1:n + 1].std()
    return before, after, res, it


def test_cavity_full_size_projection(F):
    """BASELINE.json configs[2] at full size (8192^2): two timesteps of the bench workload, then one
    un-fused sub-iteration whose projection must remove the interior divergence to the solver tolerance"""
    n = 8192
    fs = F.FS(n, n, 1.0 / n, 1.0 / n, 1.0, 1.0, 1e-3, 1e-3)
    fs.calc_rho_visc_const()
    for s, u in enumerate((0.0, 0.0, 0.0, 1.0)):
        fs.set_bconds(s, F.DIRICHLET, u, 0.0)
    fs.apply_velocity_bconds()
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=50, max_levels=40)
    ps.set_pressure_operator()
    um, vm, d0 = fs.monitor()
    assert um == 0.0 and vm == 0.0 and fs.abs_max("U", include_ghost=True) == 2.0   # only the lid ghost row (2 U_top - 0) is non-zero yet
    for _ in range(2):
        dt = fs.adjust_dt(0.9, 0.1)
        res, it = fs.step(ps, dt, 5)
        assert res < 1e-6 and 5 <= it <= 5 * 20
    fs.save_old_state()
    before, after, res, it1 = _projection_ratio(fs, ps, dt, n)
    mass, mx, my = fs.calc_conserved_quantities()
    print(f"\n[cavity {n}^2] its/solve {it / 5:.1f}, dt {dt:.3e}, interior divergence spread {before:.3e} -> "
          f"{after:.3e} ({it1} its, residual {res:.1e}), momentum_x {mx:.3e}")
    assert res < 1e-6 and np.isfinite(after)
    assert after <= 1e-3 * before
    ai, ag = fs.abs_max("U"), fs.abs_max("U", include_ghost=True)   # lid ghost row = 2 U_top - U(., NY-1)
    assert 0.0 < ai <= ag and ag >= 1.0
    assert mass == pytest.approx(1.0, rel=1e-12) and mx > 0.0
    ps.close()
    fs.close()


def test_drop_4096_variable_density(F):
    """BASELINE.json configs[3]: frozen VOF drop, density ratio 1000, 4096^2, operator rebuilt every
    sub-iteration (NUM_SUBITER 3); the variable-coefficient multigrid-PCG must converge in the
    reference's iteration range (8-19) and project the velocity."""
    n = 4096
    dx = 1.0 / n
    fs = F.FS(n, n, dx, dx, 1.0, 1000.0, 1e-3, 1e-3)
    xm = (np.arange(-1, n + 1) + 0.5) * dx
    X, Y = np.meshgrid(xm, xm, indexing="ij")
    d = np.sqrt((2 * (X - 0.5)) ** 2 + (Y - 0.5) ** 2) - 0.25
    vf = np.clip(0.5 - d / (1.5 * dx), 0.0, 1.0)
    del X, Y, d
    fs.upload("VF", vf)
    fs.calc_rho_visc_vof()
    fs.save_old_density()
    for s, u in enumerate((0.0, 0.0, 0.0, 1.0)):
        fs.set_bconds(s, F.DIRICHLET, u, 0.0)
    fs.apply_velocity_bconds()
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=50, max_levels=40)
    ps.set_pressure_operator()
    mass0 = fs.calc_conserved_quantities()[0]
    for _ in range(2):
        dt = fs.adjust_dt(0.9, 0.1)
        res, it = fs.step(ps, dt, 3, F.STEP_REBUILD_OPERATOR)
        assert res < 1e-6, res
    fs.save_old_state()
    before, after, res1, it1 = _projection_ratio(fs, ps, dt, n, rebuild=True)
    mass1 = fs.calc_conserved_quantities()[0]
    print(f"\n[drop 4096^2, rho 1:1000] its/solve {it / 3:.1f}, interior divergence spread {before:.3e} -> "
          f"{after:.3e} ({it1} its, residual {res1:.1e}), mass {mass1:.9f}")
    assert 3 <= it / 3 <= 25
    assert res1 < 1e-6 and after <= 1e-3 * before
    assert mass1 == mass0          # frozen VOF: densities untouched by the step
    # the ellipse has semi-axes 0.125 and 0.25: mass = 1 + 999 * area up to the smoothed interface
    assert mass0 == pytest.approx(1.0 + 999.0 * np.pi * 0.125 * 0.25, rel=2e-3)
    ps.close()
    fs.close()


def test_capped_hierarchy_vs_unmodified_reference_2048(F, oracle):
    """GPU against the UNMODIFIED reference (oracle/_ref/libfsref.so: its headers + vendored HYPRE) at a size where
    the reference's 16-level cap is active (src/LinearSolver_StructHypre.hpp:221-229: from 2048^2 upwards the
    hierarchy stops at a 32 x 32 coarsest grid with 64 sweeps).  Lid-driven cavity 2048^2 with the constants of
    scaling/LidDrivenFlow.cpp:16-42, developed for two timesteps, then one more step on both sides from the same
    state: max_levels = 16, PCG tol 1e-6.  Both sides stop at a 1e-6 residual, so the fields agree to the solver
    tolerance (not to rounding), the iteration counts per timestep are equal and so is dt."""
    from cases import cavity_bc, cavity_state, rel_l2, to_gpu, vel_err
    from oracle_lib import Ref, have_ref
    if not have_ref():
        pytest.skip("oracle/_ref/libfsref.so not built (it is built where /root/reference exists and travels with the repo)")
    n, nsub, tol = 2048, 5, 1e-6
    st, bc = cavity_state(oracle, n, n), cavity_bc()
    oracle.apply_velocity_bconds(st, bc)
    r = Ref(n, n)
    r.load(st)
    r.set_bconds(bc)
    r.solver_create(tol, 50)
    r.solver_set_operator()
    for _ in range(2):  # develop the flow on the reference, then hand the SAME state to both sides
        r.step(r.adjust_dt(0.9, 0.1), nsub)
    r.store(st)
    fs = to_gpu(F, st, bc)
    ps = F.LinearSolver(fs, tol=tol, max_iter=50, max_levels=16)
    ps.set_pressure_operator()
    assert ps.num_levels == 16
    dt_r = r.adjust_dt(0.9, 0.1)
    dt_g = fs.adjust_dt(0.9, 0.1)
    assert dt_g == pytest.approx(dt_r, rel=1e-12)
    it_r, res_r = r.step(dt_r, nsub)
    res_g, it_g = fs.step(ps, dt_g, nsub)
    r.store(st)
    U, V, P = fs.download("U"), fs.download("V"), fs.download("P")
    ev = vel_err(U, V, st.U, st.V)
    ep = rel_l2(P - P.mean(), st.p - st.p.mean())
    print(f"\n[2048^2 vs libfsref] PCG iterations gpu {it_g} reference {it_r}, residuals {res_g:.2e} {res_r:.2e}, "
          f"velocity rel L2 {ev:.2e}, pressure rel L2 {ep:.2e}")
    assert res_g < tol and res_r < tol
    assert it_g == it_r
    assert ev <= 1e-6 and ep <= 1e-4
    ps.close()
    fs.close()
