"""GPU parity of everything added after round 1's GPU budget was spent (SURVEY.md section 8f rows 2 and 3 and the
reference's own unit / integration tests).  The file sorts last on purpose: a surprise here cannot hide the
hardware-verified GPU tests behind `pytest -x`.

* test/Couette.cpp: analytic criterion at T = 10 (every pressure rhs exactly zero -> HYPRE's zero-rhs exit,
  pcg.c:473-488; here the PCG_ZERO_RHS path) and seeded timesteps against the reference golden
* function-pointer Dirichlet sides, PSDirichlet solves, the misc golden (sigma CFL, monitor maxima, shift, rho / mu)
* test/PeriodicChannel.cpp, test/CustomDirichletChannel.cpp, scaling/Channel.cpp: flux corrections, timesteps against
  reference goldens, slab variants, the convergence study, the drivers' own pass criteria at T = 60 (Python and the
  C++ shim example)
* file output from the device buffers: .npy / VTK bytes against files written by the reference, other sizes, slabs
* integrate / L1_norm / min / max / NaN scan / gradients / point sampling: reference goldens, the reference's
  known-answer tests (test/Operators.cpp, test/Utility.cpp), slab variants
* two-phase timesteps on the drop initialised by quadrature as the reference's drivers do
"""
import os

import numpy as np
import pytest

from cases import couette_bc, couette_l1_errors, couette_state, rel_l2, to_gpu, vel_err

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_couette_analytic(F, oracle):
    st, bc = couette_state(oracle), couette_bc()
    oracle.apply_velocity_bconds(st, bc)
    fs = to_gpu(F, st, bc)
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=500)
    ps.set_pressure_operator()
    t, steps, its = 0.0, 0, 0
    while t < 10.0:
        dt = min(fs.adjust_dt(0.9, 0.1), 10.0 - t)
        res, it = fs.step(ps, dt, 2)
        assert np.isfinite(res)
        its += it
        t += dt
        steps += 1
    U, RU = fs.download("U"), fs.download("RHO_U")
    errs = couette_l1_errors(U, st.nx, st.ny)
    inflow, outflow = (RU[0] * U[0]).sum(), (RU[st.nx + 2] * U[st.nx + 2]).sum()
    print(f"\n[Couette 210x21, T=10] {steps} steps, {its} PCG iterations, L1 errors {errs}, "
          f"mass error {outflow - inflow:.2e}")
    # (the oracle takes the zero-rhs exit in every solve; the GPU does too as long as FMA contraction keeps
    # the x-differences of equal fluxes exactly zero -- the count is printed, not asserted)
    assert steps > 2000
    assert max(errs) <= 1e-5
    assert abs(outflow - inflow) <= 1e-8
    assert fs.abs_max("V") <= 1e-10
    ps.close()
    fs.close()


def test_couette_bconds_timesteps(F, oracle):
    g = np.load(os.path.join(GOLD, "run_couette210x21.npz"))
    st, bc = couette_state(oracle, seed=5), couette_bc()
    oracle.apply_velocity_bconds(st, bc)
    fs = to_gpu(F, st, bc)
    so = oracle.solver(st.nx, st.ny, tol=1e-13, max_iter=500, snap_zero_diag=1)
    so.set_operator(oracle.fill_pressure_operator(st))
    ps = F.LinearSolver(fs, tol=1e-13, max_iter=500)
    ps.set_pressure_operator()
    for dt_ref in g["dts"]:
        assert fs.adjust_dt(0.9, 0.1) == pytest.approx(float(dt_ref), rel=1e-9)
        _, res_o = oracle.step(st, bc, so, float(dt_ref), 2)
        res_g, _ = fs.step(ps, float(dt_ref), 2)
        # HYPRE's PCG stalls at 1e-5 .. 1e-9 on this system (see tests/test_oracle.py), and where
        # exactly is decided at rounding level; the same algorithm must end in the same range
        assert res_g < max(1e-4, 100 * res_o), (res_g, res_o)
    U, V, p = fs.download("U"), fs.download("V"), fs.download("P")
    for ru, rv, rp, who in ((st.U, st.V, st.a["p"], "oracle"), (g["U"], g["V"], g["p"], "reference")):
        ev, ep = vel_err(U, V, ru, rv), rel_l2(p - p.mean(), rp - rp.mean())
        print(f"\n[Couette BCs, seeded, vs {who}] velocity rel L2 {ev:.3e}, pressure rel L2 {ep:.3e}")
        assert ev <= 1e-6 and ep <= 1e-4
    ps.close()
    fs.close()


def test_dirichlet_profiles_vs_reference_golden(F):
    """function-pointer Dirichlet sides (src/BoundaryConditions.hpp:15-88; the inlet of
    test/CustomDirichletChannel.cpp) against the committed output of the unmodified reference"""
    from cases import profile_bcs, random_state
    g = np.load(os.path.join(GOLD, "bc_profiles48x20.npz"))
    st = random_state(48, 20, seed=3)
    for key, bc in profile_bcs(48, 20).items():
        fs = to_gpu(F, st, bc)
        fs.apply_velocity_bconds()
        assert np.array_equal(fs.download("U"), g[f"{key}_U"]), key
        assert np.array_equal(fs.download("V"), g[f"{key}_V"]), key
        fs.close()


@pytest.mark.parametrize("psd", [1, 2, 3, 4])
def test_solve_with_dirichlet_side_vs_reference_golden(F, oracle, psd):
    """PSDirichlet LEFT / RIGHT / BOTTOM / TOP (src/LinearSystem.hpp:10,32-155) on the rho-ratio-1000
    drop with all-Neumann velocity BCs (test/StationaryDrop.cpp:67-72,223-224): the solution HYPRE finds.  The
    reference keeps the coupling of the neighbouring row TO the identity row (a non-symmetric matrix, which PCG is not
    built for); the product eliminates that column as well (same solution, symmetric operator, csrc/solver_setup.cu
    k_build_l0) and must not need more iterations than HYPRE: measured 21 against 25 / 23 / 21 / 21"""
    from cases import cavity_state
    from oracle_lib import BConds, NEUMANN
    g = np.load(os.path.join(GOLD, "solve_psd128.npz"))
    st = cavity_state(oracle, 128, 128, drop=True, seed=2)
    bc = BConds(kind=(NEUMANN,) * 4)
    oracle.apply_velocity_bconds(st, bc)
    fs = to_gpu(F, st, bc)
    ps = F.LinearSolver(fs, tol=1e-10, max_iter=300, psd=psd)
    ps.set_pressure_operator()
    fs.calc_divergence()
    ps.set_pressure_rhs(1e-2)
    res, it = ps.solve()
    dp = fs.download("DELTA_P")
    print(f"\n[PSDirichlet {psd}] its gpu {it} reference {int(g[f'it{psd}'])}, residual {res:.2e}, "
          f"rel L2 {rel_l2(dp, g[f'dp{psd}']):.2e}")
    assert res < 1e-10 and 0.7 * int(g[f"it{psd}"]) <= it <= int(g[f"it{psd}"]) + 1
    assert rel_l2(dp, g[f"dp{psd}"]) <= 1e-8
    ps.close()
    fs.close()


def test_misc_vs_reference_golden(F, oracle):
    """adjust_dt with the surface-tension limit (src/FS.hpp:108-114), the monitor maxima (interior entries,
    src/Utility.hpp:44-88), shift_pressure_to_zero and calc_rho / calc_visc(vf, fs) against the committed
    output of the unmodified reference (tests/golden/misc_random48x20.npz)"""
    from cases import random_state, rel_inf, twophase_inputs
    g = np.load(os.path.join(GOLD, "misc_random48x20.npz"))
    st = random_state(48, 20, seed=9)
    st.sigma = 0.05
    vf = twophase_inputs(st, seed=2, random_vf=True)[0]
    oracle.apply_neumann_bconds(vf)
    fs = to_gpu(F, st)
    assert fs.adjust_dt(0.9, 1e9) == pytest.approx(float(g["adjust_dt_sigma"]), rel=1e-13)
    um, vm, dm = fs.monitor()
    assert um == g["diag"][0] and vm == g["diag"][1] and dm == pytest.approx(g["diag"][2], rel=1e-12)
    assert fs.abs_max("U") == g["diag"][0] and fs.abs_max("DIV") == pytest.approx(g["diag"][2], rel=1e-12)
    dp = np.random.default_rng(4).uniform(-1, 1, (st.nx + 2, st.ny + 2)) + 0.3
    fs.upload("DELTA_P", dp)
    fs.shift_pressure_to_zero()
    assert rel_inf(fs.download("DELTA_P"), g["dp_shifted"]) <= 1e-12
    fs.upload("VF", vf)
    fs.calc_rho_visc_vof()
    for n, k in (("RHO_U", "rho_u"), ("RHO_V", "rho_v"), ("VISC", "visc")):
        assert rel_inf(fs.download(n), g[k]) <= 1e-12, n
    fs.close()


CHANNEL215 = ["periodic215x43", "customdirichlet215x43", "clippedchannel215x43"]


def _channel215(oracle, name):
    from cases import channel215_cases, channel215_state
    bc, mode, total, backflow = channel215_cases()[name]
    st = channel215_state(oracle, seed=8, backflow=backflow)
    oracle.apply_velocity_bconds(st, bc)
    return st, bc, mode, total


# ---- file output from the device buffers ------------------------------
def _io_case():
    from cases import grid_coords, random_state
    st = random_state(24, 10, seed=2410)
    return (st,) + grid_coords(st)


def _vtk_inputs(fs):
    fs.calc_divergence()
    fs.interpolate_cell_centres()
    return [("pressure", "P"), ("divergence", "DIV")], [("velocity", "UI", "VI")]


@pytest.mark.parametrize("chunk", [None, "1024"])
def test_npy_and_vtk_files_vs_reference_files(F, tmp_path, monkeypatch, chunk):
    """fsgpu_to_npy_state and fsgpu_write_vtk byte for byte against the files the unmodified reference
    writes for the same state (tests/golden/io_random24x10.npz); chunk = 1024 bytes forces the
    double-buffered staging through many chunks"""
    if chunk:
        monkeypatch.setenv("FSGPU_OUT_CHUNK_BYTES", chunk)
    g = np.load(os.path.join(GOLD, "io_random24x10.npz"))
    st, x, y = _io_case()
    fs = to_gpu(F, st)
    d = tmp_path / "npy"
    fs.to_npy_state(str(d), x, y)
    names = sorted(k[4:] for k in g.files if k.startswith("npy_"))
    assert sorted(f[:-4] for f in os.listdir(d)) == names
    for k in names:
        assert (d / (k + ".npy")).read_bytes() == g["npy_" + k].tobytes(), k
    sc, ve = _vtk_inputs(fs)
    f = tmp_path / "state_000000.vtk"
    f.write_bytes(b"\0" * 40000)  # an older, longer file
    fs.write_vtk(str(f), 0.125, x, y, sc, ve)
    assert f.read_bytes() == g["vtk"].tobytes()
    fs.close()


@pytest.mark.parametrize("nx,ny", [(130, 67), (64, 200)])
def test_vtk_and_npy_other_sizes(F, oracle, tmp_path, monkeypatch, nx, ny):
    """sizes that are no multiple of the transpose tile, against the numpy restatement of the two
    formats (tests/cases.py, itself checked against the reference's files)"""
    from cases import grid_coords, npy_bytes, random_state, twophase_inputs, vtk_bytes
    monkeypatch.setenv("FSGPU_OUT_CHUNK_BYTES", "65536")
    st = random_state(nx, ny, seed=nx + ny)
    x, y = grid_coords(st, x_min=-0.25, y_min=1.5)
    vf = twophase_inputs(st, seed=1, random_vf=True)[0]
    fs = to_gpu(F, st)
    fs.upload("VF", vf)
    sc, ve = _vtk_inputs(fs)
    f = tmp_path / "s.vtk"
    fs.write_vtk(str(f), 1e-5, x, y, sc + [("vof", "VF")], ve)
    div = oracle.calc_divergence(st)
    ui, vi = oracle.interpolate_UV(st)
    ref = vtk_bytes(1e-5, x, y, [("pressure", st.a["p"]), ("divergence", fs.download("DIV")), ("vof", vf)],
                    [("velocity", fs.download("UI"), fs.download("VI"))])
    assert f.read_bytes() == ref
    assert np.abs(fs.download("DIV") - div).max() <= 1e-12 * np.abs(div).max()
    assert np.abs(fs.download("UI") - ui).max() <= 1e-14 and np.abs(fs.download("VI") - vi).max() <= 1e-14
    for name in ("U", "V", "P"):
        fs.to_npy(name, str(tmp_path / (name + ".npy")))
        assert (tmp_path / (name + ".npy")).read_bytes() == npy_bytes(st.a[name if name != "P" else "p"])
    fs.close()


def test_output_files_from_slabs(F, oracle, tmp_path):
    """3 slab contexts write their rows (npy) / columns (VTK) of the same files: identical bytes"""
    from test_gpu_slab import run_ranks, run_single
    from cases import cavity_bc, grid_coords, random_state
    st = random_state(50, 23, seed=7)
    x, y = grid_coords(st)

    def body(tag):
        def run(fs, r):
            sc, ve = _vtk_inputs(fs)
            fs.to_npy_state(str(tmp_path / tag), x, y)
            fs.write_vtk(str(tmp_path / (tag + ".vtk")), 2.0, x, y, sc, ve)
        return run

    run_ranks(F, st, cavity_bc(), 3, body("slabs"), fields=("U",))
    run_single(F, st, cavity_bc(), body("one"), fields=("U",))
    assert (tmp_path / "slabs.vtk").read_bytes() == (tmp_path / "one.vtk").read_bytes()
    files = sorted(os.listdir(tmp_path / "one"))
    assert len(files) == 16 and files == sorted(os.listdir(tmp_path / "slabs"))
    for f in files:
        assert (tmp_path / "slabs" / f).read_bytes() == (tmp_path / "one" / f).read_bytes(), f


# ---- integrate / L1_norm / calc_grad_of_centered_points (src/Operators.hpp:51-76,164-201) ----------
def test_integrate_l1_grad_vs_reference_golden(F, oracle):
    from cases import random_state, rel_inf
    g = np.load(os.path.join(GOLD, "operators_random48x20.npz"))
    st = random_state(48, 20, seed=11)
    f = np.random.default_rng(12).uniform(-1, 1, (st.nx + 2, st.ny + 2)) + 0.2
    fs = to_gpu(F, st)
    fs.upload("VF", f)
    sums = [fn("VF", gh) for gh in (False, True) for fn in (fs.integrate, fs.l1_norm)]
    assert np.allclose(sums, g["sums"], rtol=1e-12, atol=0)
    # a U-like and a V-like field: interior = the faces the reference's for_each_i visits
    for name, a in (("U", st.U), ("V", st.V)):
        assert fs.integrate(name) == pytest.approx(oracle.integrate(a, st.dx, st.dy), rel=1e-12, abs=1e-13)
        assert fs.l1_norm(name, True) == pytest.approx(oracle.integrate(a, st.dx, st.dy, True, True), rel=1e-12)
    fs.calc_grad_of_centered_points("VF", "GRAD_X", "GRAD_Y")
    fs.calc_grad_of_centered_points("GRAD_X", "GRAD_XX", "CURV")   # dfdxx, dfdxy
    fs.calc_grad_of_centered_points("GRAD_Y", "RHOI", "GRAD_YY")   # dfdyx, dfdyy
    for name, k in (("GRAD_X", "gx"), ("GRAD_Y", "gy"), ("GRAD_XX", "gxx"), ("CURV", "gxy"), ("RHOI", "gyx"),
                    ("GRAD_YY", "gyy")):
        assert rel_inf(fs.download(name), g[k]) <= 1e-12, k
    with pytest.raises(F.FsgpuError):
        fs.calc_grad_of_centered_points("VF", "VF", "GRAD_Y")
    with pytest.raises(F.FsgpuError):
        fs.calc_grad_of_centered_points("U", "GRAD_X", "GRAD_Y")
    fs.close()


def test_gradient_centered_points_known_answer(F):
    """the reference's own test/Operators.cpp:86-215 on the GPU (quadratic f, derivatives exact within 2e-10)"""
    nx, ny, dx, dy = 200, 300, 2.0 / 200, 4.0 / 300
    x = 2.0 + (np.arange(-1, nx + 1) + 0.5) * dx - 0.5
    y = -1.0 + (np.arange(-1, ny + 1) + 0.5) * dy - 0.5
    X, Y = np.meshgrid(x, y, indexing="ij")
    fs = F.FS(nx, ny, dx, dy)
    fs.upload("VF", X * X + X * Y + Y * Y)
    fs.calc_grad_of_centered_points("VF", "GRAD_X", "GRAD_Y")
    fs.calc_grad_of_centered_points("GRAD_X", "GRAD_XX", "CURV")
    fs.calc_grad_of_centered_points("GRAD_Y", "RHOI", "GRAD_YY")
    eps = 2e-10
    assert np.abs(fs.download("GRAD_X") - (2 * X + Y)).max() <= eps
    assert np.abs(fs.download("GRAD_Y") - (X + 2 * Y)).max() <= eps
    for name, want in (("GRAD_XX", 2.0), ("GRAD_YY", 2.0), ("CURV", 1.0), ("RHOI", 1.0)):
        assert np.abs(fs.download(name) - want).max() <= eps, name
    fs.close()


def test_integrate_grad_slabs(F, oracle):
    """the same operators on 3 slab contexts: sums combined over the ranks, gradient rows next to a slab
    boundary read the neighbour's rows"""
    from test_gpu_slab import compare, run_ranks, run_single
    from cases import cavity_bc, random_state
    st = random_state(50, 23, seed=7)
    st.vf = np.random.default_rng(3).uniform(0, 1, (st.nx + 2, st.ny + 2))

    def body(fs, r):
        out = [fs.integrate("VF"), fs.integrate("VF", True), fs.l1_norm("U"), fs.l1_norm("V", True)]
        fs.calc_grad_of_centered_points("VF", "GRAD_X", "GRAD_Y")
        fs.calc_grad_of_centered_points("GRAD_X", "GRAD_XX", "CURV")
        return out

    names = ("GRAD_X", "GRAD_Y", "GRAD_XX", "CURV")
    a, ra = run_ranks(F, st, cavity_bc(), 3, body, fields=names)
    b, rb = run_single(F, st, cavity_bc(), body, fields=names)
    assert all(x == ra[0] for x in ra)
    assert ra[0] == pytest.approx(rb, rel=1e-12)
    assert rb[1] == pytest.approx(oracle.integrate(st.vf, st.dx, st.dy, True), rel=1e-12)
    compare(a, b, names, tol=1e-13)


def test_point_sampling_vs_reference_golden(F, oracle):
    """eval_flow_field_at / bilinear_interpolate (src/Operators.hpp:85-161) at seeded points, incl. points outside the
    cell-centre box, against the unmodified reference; then the reference's own known-answer test
    (test/Operators.cpp:24-84, bilinear fields reproduced within 100 eps)"""
    from cases import grid_coords, random_state, rel_inf, sample_points
    g = np.load(os.path.join(GOLD, "sampling_random48x20.npz"))
    st = random_state(48, 20, seed=21)
    px, py = sample_points(st)
    x, y = grid_coords(st)
    fs = to_gpu(F, st)
    fs.interpolate_cell_centres()
    u, v = fs.eval_flow_field_at(x, y, px, py)
    assert rel_inf(u, g["u"]) <= 1e-12 and rel_inf(v, g["v"]) <= 1e-12
    assert rel_inf(fs.bilinear_interpolate("P", x, y, px, py), g["p"]) <= 1e-12
    with pytest.raises(F.FsgpuError):
        fs.bilinear_interpolate("U", x, y, px, py)
    fs.close()

    nx, ny, dx, dy = 200, 300, 2.0 / 200, 4.0 / 300
    x = 2.0 + np.arange(-1, nx + 2) * dx
    y = -1.0 + np.arange(-1, ny + 2) * dy
    xm, ym = (x[:-1] + x[1:]) / 2, (y[:-1] + y[1:]) / 2
    fs = F.FS(nx, ny, dx, dy)
    fs.upload("U", 12.0 * x[:, None] + ym[None, :])
    fs.upload("V", xm[:, None] * y[None, :] - 27.31415)
    fs.interpolate_cell_centres()
    k = np.arange(50)
    PX, PY = (a.ravel() for a in np.meshgrid((xm[nx] - xm[1]) / 49.0 * k + xm[1], (ym[ny] - ym[1]) / 49.0 * k + ym[1],
                                             indexing="ij"))
    u, v = fs.eval_flow_field_at(x, y, PX, PY)
    eps = 400.0 * np.finfo(float).eps   # the reference asks 100 eps of its own (differently contracted) arithmetic
    assert np.abs(u - (12.0 * PX + PY)).max() <= eps and np.abs(v - (PX * PY - 27.31415)).max() <= eps
    fs.close()


def test_conserved_quantities_known_answer(F):
    """the reference's own test of calc_conserved_quantities (test/Operators.cpp:217-283) on the GPU"""
    from cases import conserved_known_answer
    st, expected, tol = conserved_known_answer()
    fs = to_gpu(F, st)
    got = fs.calc_conserved_quantities()
    assert np.abs(np.array(got) - np.array(expected)).max() <= tol
    fs.close()


def test_twophase_quadrature_timesteps(F, oracle):
    """BASELINE configs[3] with the reference drivers' own initialisation (volume fraction by Gauss quadrature of
    the ellipse indicator: a sharp 1 : 1000 density jump), density update + capillary block + operator rebuild
    through fsgpu_step, against the committed output of the unmodified reference"""
    from cases import rel_inf, twophase_quadrature_case
    g = np.load(os.path.join(GOLD, "run_twophase_q128.npz"))
    st, bc, cap = twophase_quadrature_case(oracle)
    fs = to_gpu(F, st, bc)
    for n, a in zip(("VF", "CURV", "INTERFACE_LENGTH"), cap):
        fs.upload(n, a)
    # calc_rho / calc_visc(vf, fs) on the device reproduce the uploaded (oracle-made) densities
    rho_u, visc = fs.download("RHO_U"), fs.download("VISC")
    fs.calc_rho_visc_vof()
    assert rel_inf(fs.download("RHO_U"), rho_u) <= 1e-12 and rel_inf(fs.download("VISC"), visc) <= 1e-12
    ps = F.LinearSolver(fs, tol=1e-13, max_iter=500)
    ps.set_pressure_operator()
    flags = F.STEP_REBUILD_OPERATOR | F.STEP_DENSITY | F.STEP_CAPILLARY
    for k, dt_ref in enumerate(g["dts"]):
        assert fs.adjust_dt(0.9, 0.1) == pytest.approx(float(dt_ref), rel=1e-9)
        res, it = fs.step(ps, float(dt_ref), 3, flags)
        assert res < 1e-12 and abs(it - int(g["its"][k])) <= max(3, 0.35 * int(g["its"][k]))
    U, V, p = fs.download("U"), fs.download("V"), fs.download("P")
    ev = vel_err(U, V, g["U"], g["V"])
    ep = rel_l2(p - p.mean(), g["p"] - g["p"].mean())
    print(f"\n[twophase_q128] velocity rel L2 {ev:.3e}, pressure rel L2 {ep:.3e}")
    assert ev <= 1e-10 and ep <= 1e-8
    assert rel_inf(fs.download("RHO_U"), g["rho_u"]) <= 1e-12 and rel_inf(fs.download("P_JUMP_U"), g["pj_u"]) <= 1e-12
    ps.close()
    fs.close()


def test_min_max_known_answer_and_slabs(F):
    """min / max / abs_max with and without ghosts: the reference's own test (test/Utility.cpp:89-136: ghosts -100,
    interior rows i even ? i : -i on 10 x 10, here with the path's single ghost layer), random fields against numpy,
    and 3 slab contexts against one"""
    from test_gpu_slab import run_ranks, run_single
    from cases import cavity_bc, random_state
    fs = F.FS(10, 10, 0.1, 0.1)
    a = np.full((12, 12), -100.0)
    a[1:-1, 1:-1] = np.array([i if i % 2 == 0 else -i for i in range(10)], dtype=np.float64)[:, None]
    fs.upload("P", a)
    assert (fs.min("P"), fs.min("P", True)) == (-9.0, -100.0)
    assert (fs.max("P"), fs.max("P", True)) == (8.0, 8.0)
    assert (fs.abs_max("P"), fs.abs_max("P", True)) == (9.0, 100.0)
    fs.close()
    st = random_state(50, 23, seed=9)

    def body(fs, r):
        return [f(n, gh) for n in ("U", "V", "P") for gh in (False, True) for f in (fs.min, fs.max)]

    _, ra = run_ranks(F, st, cavity_bc(), 3, body, fields=("U",))
    _, rb = run_single(F, st, cavity_bc(), body, fields=("U",))
    assert all(x == ra[0] for x in ra) and ra[0] == rb
    want = [f(arr if gh else arr[1:-1, 1:-1]) for arr in (st.U, st.V, st.a["p"]) for gh in (False, True)
            for f in (np.min, np.max)]
    assert rb == want


def test_count_nonfinite(F):
    """has_nan / has_inf / has_nan_or_inf (src/Container.hpp:191-204) as one device pass over the ghost-inclusive array"""
    from cases import random_state
    st = random_state(50, 23, seed=4)
    fs = to_gpu(F, st)
    assert fs.count_nonfinite("U") == (0, 0) and fs.count_nonfinite("DELTA_P") == (0, 0)
    a = st.U.copy()
    a[0, 0] = np.nan            # a ghost corner counts: the reference scans the whole array
    a[17, 5] = np.nan
    a[-1, -1] = np.inf
    a[3, 3] = -np.inf
    a[4, 4] = np.inf
    fs.upload("U", a)
    assert fs.count_nonfinite("U") == (2, 3)
    assert fs.count_nonfinite("V") == (0, 0)
    fs.close()


def test_point_sampling_slabs(F, oracle):
    """eval_flow_field_at / bilinear_interpolate on 3 slab contexts: the slab owning a point's row evaluates it,
    the results are summed over the ranks; identical to one context"""
    from test_gpu_slab import run_ranks, run_single
    from cases import cavity_bc, grid_coords, random_state, sample_points
    st = random_state(50, 23, seed=5)
    px, py = sample_points(st)
    x, y = grid_coords(st)

    def body(fs, r):
        fs.interpolate_cell_centres()
        u, v = fs.eval_flow_field_at(x, y, px, py)
        return [u, v, fs.bilinear_interpolate("P", x, y, px, py)]

    _, ra = run_ranks(F, st, cavity_bc(), 3, body, fields=("U",))
    _, rb = run_single(F, st, cavity_bc(), body, fields=("U",))
    for q in range(3):
        assert all(np.array_equal(r[q], ra[0][q]) for r in ra)
        assert np.array_equal(ra[0][q], rb[q])


# ---- the long / tolerance-sensitive runs last (pytest -x stops at the first failure) ---------------------------
@pytest.mark.parametrize("name", CHANNEL215)
def test_channel215_drivers_vs_reference_golden(F, oracle, name):
    """test/PeriodicChannel.cpp (total-flow forcing of both ghost columns), test/CustomDirichletChannel.cpp
    (function inlet, dy-weighted outflow correction) and scaling/Channel.cpp (clipped Neumann outlet with
    backflow): fsgpu_flow_control alone, then timesteps through fsgpu_step with
    fsgpu_set_flow_control, against the committed output of the unmodified reference"""
    from cases import rel_inf
    g = np.load(os.path.join(GOLD, f"run_{name}.npz"))
    st, bc, mode, total = _channel215(oracle, name)
    fs = to_gpu(F, st, bc)
    io = fs.flow_control(mode, total)
    assert np.allclose(io, g["flow_io"], rtol=1e-13, atol=0)
    assert rel_inf(fs.download("U"), g["flow_U"]) <= 1e-12
    if mode == F.FLOW_FORCE_TOTAL:
        assert np.allclose(fs.flow_control(mode, total), total, rtol=1e-12)
    fs.upload("U", st.U)
    fs.set_flow_control(mode, total)
    ps = F.LinearSolver(fs, tol=1e-13, max_iter=500)
    ps.set_pressure_operator()
    worst = 0.0
    for k, dt_ref in enumerate(g["dts"]):
        assert fs.adjust_dt(0.9, 0.1) == pytest.approx(float(dt_ref), rel=1e-9)
        res, it = fs.step(ps, float(dt_ref), 2, F.STEP_OUTFLOW_FIX)
        worst = max(worst, res)
        assert res < 1e-6
        if res < 1e-12:
            assert abs(it - int(g["its"][k])) <= max(3, 0.35 * int(g["its"][k]))
    U, V, p = fs.download("U"), fs.download("V"), fs.download("P")
    ev = vel_err(U, V, g["U"], g["V"])
    ep = rel_l2(p - p.mean(), g["p"] - g["p"].mean())
    print(f"\n[{name}] velocity rel L2 {ev:.3e}, pressure rel L2 {ep:.3e}, worst residual {worst:.2e}")
    # the reference reaches 1e-13 in every solve of these runs; should the PCG here leave one at its
    # breakdown floor instead (as HYPRE itself does on the TGV / channel systems, tests/test_gpu_solver.py),
    # the fields agree to that floor
    floor = worst >= 1e-12
    assert ev <= (1e-7 if floor else 1e-10) and ep <= (1e-5 if floor else 1e-8)
    ps.close()
    fs.close()


@pytest.mark.parametrize("name", CHANNEL215[:2])
def test_channel215_drivers_slabs(F, oracle, name):
    """the same drivers on 3 slab contexts (in-process transport) vs one context: the inlet sum lives
    on the first slab, the outlet sum on the last, the periodic wrap-around between the two"""
    from test_gpu_slab import run_ranks, run_single
    st, bc, mode, total = _channel215(oracle, name)

    def body(fs, r):
        fs.set_flow_control(mode, total)
        ps = F.LinearSolver(fs, tol=1e-13, max_iter=500)
        ps.set_pressure_operator()
        hist = [fs.flow_control(mode, total)]
        for _ in range(2):
            dt = fs.adjust_dt(0.9, 0.1)
            fs.step(ps, dt, 2, F.STEP_OUTFLOW_FIX)
            hist.append(dt)
        ps.close()
        return hist

    a, ra = run_ranks(F, st, bc, 3, body, fields=("U", "V", "P"))
    b, rb = run_single(F, st, bc, body, fields=("U", "V", "P"))
    assert all(x == ra[0] for x in ra)
    assert np.allclose(ra[0][0], rb[0], rtol=1e-13) and ra[0][1:] == pytest.approx(rb[1:], rel=1e-9)
    # channel systems: same bound as tests/test_gpu_slab.py::test_slab_timesteps (PCG breakdown floor)
    assert vel_err(a["U"], a["V"], b["U"], b["V"]) <= 1e-7


def test_channel_convergence_study(F, oracle):
    """scaling/Channel.cpp (non-periodic) on the GPU at the driver's own settings, T_END = 20, n = 3, 4, 5:
    the error measures of :253-300 against the unmodified reference's (tests/golden/channel_study.npz) and
    the second-order trend of the profile error.  The reference solves to tol 1e-6, so two implementations
    agree to about that in the fields."""
    from cases import channel_study_case, channel_study_errors, run_channel_study
    g = np.load(os.path.join(GOLD, "channel_study.npz"))
    ue = []
    for k, n in enumerate(g["n"]):
        st, bc = channel_study_case(oracle, int(n))
        fs = to_gpu(F, st, bc)
        fs.set_flow_control(F.FLOW_OUTFLOW_FIX_DY)
        ps = F.LinearSolver(fs, tol=1e-6, max_iter=50)
        ps.set_pressure_operator()
        steps = run_channel_study(lambda: fs.adjust_dt(0.9, 0.1), lambda dt: fs.step(ps, dt, 2, F.STEP_OUTFLOW_FIX))
        e = np.array(channel_study_errors(fs.download("P"), fs.download("U"), fs.download("V"), st.nx, st.ny, st.dx))
        print(f"\n[channel study n={n}] steps {steps} (reference {int(g['errors'][k][4])}), errors {e}, "
              f"reference {g['errors'][k][:4]}")
        assert abs(steps - int(g["errors"][k][4])) <= 1
        # the profile error is a discretisation error (2 %); the pressure / dpdx / V measures are at the level the
        # 1e-6 solves leave behind and move with the iteration at which a solve stops (20 %)
        assert np.all(np.abs(e / g["errors"][k][:4] - 1.0) <= np.array([0.2, 0.2, 0.02, 0.2])), (n, e)
        if k == 0:
            assert np.abs(fs.download("U") - g["U_n3"]).max() <= 1e-5
        ue.append(e[2])
        ps.close()
        fs.close()
    assert ue[0] / ue[1] > 3.0 and ue[1] / ue[2] > 3.0


def test_custom_dirichlet_channel_reference_criteria(F, oracle):
    """the reference's own test/CustomDirichletChannel.cpp on the GPU at the driver's settings (tol 1e-6, 50
    iterations, NUM_SUBITER 2, CFL 0.9, T_END = 60), judged by the driver's criteria (:240-331): outflow = inflow
    within 1e-8 once t > 10, pressure constant along y and dp/dx = -12 mu within 1e-4, U profile within 2e-3 (L1) of
    the Poiseuille profile at every station, |V| <= 3e-4"""
    from cases import channel215_cases, channel215_measures, channel215_state, run_channel_study
    bc, mode, total, _ = channel215_cases()["customdirichlet215x43"]
    st = channel215_state(oracle)
    oracle.apply_velocity_bconds(st, bc)
    fs = to_gpu(F, st, bc)
    fs.set_flow_control(mode, total)
    ps = F.LinearSolver(fs, tol=1e-6, max_iter=50)
    ps.set_pressure_operator()
    t, worst = [0.0], [0.0]

    def step(dt):
        fs.step(ps, dt, 2, F.STEP_OUTFLOW_FIX)
        t[0] += dt
        if t[0] > 10.0:   # rho U dy summed over the two ghost columns (CustomDirichletChannel.cpp:61-72)
            flux = 0.5 * fs.download("U")[[0, -1], :].sum(axis=1) * st.dy
            worst[0] = max(worst[0], abs(flux[1] - flux[0]))

    steps = run_channel_study(lambda: fs.adjust_dt(0.9, 0.1), step, t_end=60.0)
    dev_p, dev_dpdx, l1, vmax = channel215_measures(False, fs.download("P"), fs.download("U"), fs.download("V"))
    print(f"\n[customdirichlet215x43, {steps} steps] p dev {dev_p:.2e}, dpdx dev {dev_dpdx:.2e}, U L1 {l1:.3e}, "
          f"max|V| {vmax:.3e}, worst flux mismatch {worst[0]:.2e}")
    assert dev_p <= 1e-4 and dev_dpdx <= 1e-4 and l1 <= 2e-3 and vmax <= 3e-4 and worst[0] <= 1e-8
    ps.close()
    fs.close()


def test_cpp_periodic_channel_driver(tmp_path):
    """examples/periodic_channel.cpp: the reference's test/PeriodicChannel.cpp on the C++ shim, run to its own
    T_END = 60 with the reference's own pass criteria (:264-337) as the exit code; the VTK frames it writes
    are legacy-VTK files of the expected size"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "examples", "periodic_channel")
    if not os.path.exists(exe):
        from fluidsolver_b200 import build
        build.build_examples()
    out = tmp_path / "vtk"
    out.mkdir()
    r = subprocess.run([exe, "60", str(out)], capture_output=True, text=True, timeout=900)
    print("\n" + r.stdout[-600:])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "all tests of test/PeriodicChannel.cpp passed" in r.stdout
    frames = sorted(os.listdir(out))
    assert len(frames) >= 60 and frames[0] == "state_000000.vtk"
    nx, ny = 215, 43
    head = (out / frames[-1]).read_bytes()[:200]
    assert head.startswith(b"# vtk DataFile Version 2.0\nState of FluidSolver at time t=")
    size = os.path.getsize(out / frames[-1])
    assert size > (nx + 1) * (ny + 1) * 24 + nx * ny * (8 + 8 + 24)
