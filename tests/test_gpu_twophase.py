"""GPU parity for the rows either side of the single-phase path: consistent
mass transport, the capillary pressure jump, immersed-boundary velocity updates, diagnostics.
Every kernel is called through the C ABI (libfsgpu.so) and compared with the CPU oracle (pinned to
the reference by tests/test_oracle.py) on the same seeded inputs, and with the committed outputs
of the unmodified reference (tests/golden/ops8f_*.npz, run_twophase128.npz).

Tolerances: stencil outputs 1e-12 norm-relative (north_star); N-step fields 1e-10 relative L2.
"""
import os

import numpy as np
import pytest

from cases import (cavity_bc, cavity_state, ib_inputs, random_state, rel_inf, rel_l2, to_gpu,
                   twophase_inputs, vel_err)

pytestmark = pytest.mark.gpu
TOL = 1e-12
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SIZES = [(48, 20), (64, 64), (20, 48), (130, 67)]


def _case(oracle, name):
    if name == "random48x20":
        return random_state(48, 20, seed=48020), True
    return cavity_state(oracle, 64, 64, drop=True, seed=3), False


@pytest.mark.parametrize("nx,ny", SIZES)
def test_drhodt_update_density(F, oracle, nx, ny):
    """calc_drhodt, update_density (src/FS.hpp:318-420) and the fused density_step"""
    st = random_state(nx, ny, seed=7 * nx + ny)
    dt = 1e-3
    fs = to_gpu(F, st)
    dru, drv = oracle.calc_drhodt(st)
    fs.calc_drhodt()
    gU, gV = fs.download("DRHO_U"), fs.download("DRHO_V")
    assert rel_inf(gU, dru) <= TOL and rel_inf(gV, drv) <= TOL
    assert np.abs(gU[0]).max() == 0 and np.abs(gU[-1]).max() == 0 and np.abs(gV[:, 0]).max() == 0
    oracle.update_density(st, dru, drv, dt)
    fs.update_density(dt)
    assert rel_inf(fs.download("RHO_U"), st.rho_u) <= TOL
    assert rel_inf(fs.download("RHO_V"), st.rho_v) <= TOL
    # one-pass variant incl. apply_neumann_bconds on both densities (test/StationaryDrop.cpp:274-293)
    oracle.apply_neumann_bconds(st.rho_u)
    oracle.apply_neumann_bconds(st.rho_v)
    fs.density_step(dt)
    assert rel_inf(fs.download("RHO_U"), st.rho_u) <= TOL
    assert rel_inf(fs.download("RHO_V"), st.rho_v) <= TOL
    fs.close()


@pytest.mark.parametrize("name", ["random48x20", "drop64"])
def test_8f_operators_vs_reference_golden(F, oracle, name):
    """the same checks as tests/test_oracle.py::test_8f_operators_vs_golden, GPU side: outputs of
    the unmodified reference"""
    g = np.load(os.path.join(GOLD, f"ops8f_{name}.npz"))
    st, random_vf = _case(oracle, name)
    dt = 1e-2
    vf, curv, ilen = twophase_inputs(st, seed=7, random_vf=random_vf)
    ibu, ibv = ib_inputs(st, seed=11)
    fs = to_gpu(F, st)
    fs.set_props(st.rho_gas, st.rho_liquid, st.visc_gas, st.visc_liquid, st.sigma)
    fs.calc_drhodt()
    fs.update_density(dt)
    assert rel_inf(fs.download("RHO_U"), g["rho_u"]) <= TOL
    assert rel_inf(fs.download("RHO_V"), g["rho_v"]) <= TOL
    fs.upload("RHO_U", st.rho_u)
    fs.upload("RHO_V", st.rho_v)
    for n, a in (("VF", vf), ("CURV", curv), ("INTERFACE_LENGTH", ilen), ("IB_CORR_U", ibu),
                 ("IB_CORR_V", ibv)):
        fs.upload(n, a)
    fs.calc_pressure_jump()
    assert rel_inf(fs.download("P_JUMP_U"), g["pj_u"]) <= TOL
    assert rel_inf(fs.download("P_JUMP_V"), g["pj_v"]) <= TOL
    fs.upload("P_JUMP_U", st.p_jump_u)
    fs.upload("P_JUMP_V", st.p_jump_v)
    fs.interpolate_cell_centres()
    assert rel_inf(fs.download("UI"), g["Ui"]) <= TOL and rel_inf(fs.download("VI"), g["Vi"]) <= TOL
    rhoi = (st.rho_u[:-1, :] + st.rho_u[1:, :] + st.rho_v[:, :-1] + st.rho_v[:, 1:]) / 4.0
    assert rel_inf(fs.download("RHOI"), rhoi) <= TOL  # interpolate_UV_staggered_field, Operators.hpp:22-29
    # the reference sums serially, the GPU in a fixed tree: 1e-12 relative to the sum of magnitudes
    got, ref = np.array(fs.calc_conserved_quantities()), g["conserved"]
    scale = np.abs(st.rho_u).max() * max(1.0, np.abs(st.U).max()) * st.nx * st.ny * st.dx * st.dy
    assert np.abs(got - ref).max() <= TOL * scale
    fs.correct_velocity_ib_implicit_euler(dt)
    assert rel_inf(fs.download("U"), g["ibe_U"]) <= TOL and rel_inf(fs.download("V"), g["ibe_V"]) <= TOL
    fs.upload("U", st.U)
    fs.upload("V", st.V)
    fs.calc_dmomdt()
    fs.update_velocity_ib_semi_analytical(dt)
    assert rel_inf(fs.download("U"), g["ibs_U"]) <= TOL and rel_inf(fs.download("V"), g["ibs_V"]) <= TOL
    fs.close()


@pytest.mark.parametrize("nx,ny", SIZES)
def test_capillary_divergence_and_monitor(F, oracle, nx, ny):
    """test/StationaryDrop.cpp:331-353 and the monitor row (calc_divergence + 3 abs_max)"""
    st = random_state(nx, ny, seed=3 * nx + ny)
    st.sigma = 0.07
    dt = 2e-3
    vf, curv, ilen = twophase_inputs(st, seed=5, random_vf=True)
    fs = to_gpu(F, st)
    for n, a in (("VF", vf), ("CURV", curv), ("INTERFACE_LENGTH", ilen)):
        fs.upload(n, a)
    div = oracle.calc_divergence(st)
    um, vm, dm = fs.monitor()
    assert um == oracle.abs_max(st.U) and vm == oracle.abs_max(st.V)
    assert dm == pytest.approx(oracle.abs_max(div), rel=1e-12)
    assert rel_inf(fs.download("DIV"), div) <= TOL
    oracle.capillary_divergence(st, vf, curv, ilen, dt, div)
    fs.capillary_divergence(dt)
    assert rel_inf(fs.download("DIV"), div) <= TOL
    assert rel_inf(fs.download("P_JUMP_U"), st.p_jump_u) <= TOL
    assert rel_inf(fs.download("P_JUMP_V"), st.p_jump_v) <= TOL
    fs.close()


def _twophase_case(oracle):
    st, bc = cavity_state(oracle, 128, 128, drop=True, seed=1), cavity_bc()
    st.sigma = 0.05
    cap = twophase_inputs(st, seed=7)
    oracle.apply_velocity_bconds(st, bc)
    return st, bc, cap


def _twophase_gpu(F, st, bc, cap, **kw):
    fs = to_gpu(F, st, bc, **kw)
    for n, a in zip(("VF", "CURV", "INTERFACE_LENGTH"), cap):
        fs.upload(n, a)
    return fs


def test_twophase_timesteps(F, oracle):
    """two-phase sub-iteration through fsgpu_step (density update + capillary block + operator
    rebuild) vs the oracle AND vs the committed output of the unmodified reference"""
    g = np.load(os.path.join(GOLD, "run_twophase128.npz"))
    st, bc, cap = _twophase_case(oracle)
    fs = _twophase_gpu(F, st, bc, cap)
    flags = F.STEP_REBUILD_OPERATOR | F.STEP_DENSITY | F.STEP_CAPILLARY
    so = oracle.solver(st.nx, st.ny, tol=1e-13, max_iter=500, snap_zero_diag=1)
    so.set_operator(oracle.fill_pressure_operator(st))
    ps = F.LinearSolver(fs, tol=1e-13, max_iter=500)
    ps.set_pressure_operator()
    for dt_ref in g["dts"]:
        dt = fs.adjust_dt(0.9, 0.1)
        assert dt == pytest.approx(float(dt_ref), rel=1e-9)
        it_o, _ = oracle.step(st, bc, so, float(dt_ref), 3, rebuild=1, density=1, capillary=cap)
        res, it_g = fs.step(ps, float(dt_ref), 3, flags)
        assert res < 1e-12 and abs(it_g - it_o) <= max(3, 0.35 * it_o)
    U, V, p = fs.download("U"), fs.download("V"), fs.download("P")
    for ref_u, ref_v, ref_p, ref_ru, ref_pj, who in (
            (st.U, st.V, st.a["p"], st.rho_u, st.p_jump_u, "oracle"),
            (g["U"], g["V"], g["p"], g["rho_u"], g["pj_u"], "reference")):
        ev = vel_err(U, V, ref_u, ref_v)
        ep = rel_l2(p - p.mean(), ref_p - ref_p.mean())
        print(f"\n[twophase128 vs {who}] velocity rel L2 {ev:.3e}, pressure rel L2 {ep:.3e}")
        assert ev <= 1e-10 and ep <= 1e-8
        assert rel_inf(fs.download("RHO_U"), ref_ru) <= TOL
        assert rel_inf(fs.download("P_JUMP_U"), ref_pj) <= TOL
    ps.close()
    fs.close()


@pytest.mark.parametrize("R", [2, 3])
def test_twophase_timesteps_slabs(F, oracle, R):
    """the same step on R slab contexts (in-process transport) vs one context"""
    from test_gpu_slab import compare, run_ranks, run_single
    st, bc, cap = _twophase_case(oracle)
    flags = F.STEP_REBUILD_OPERATOR | F.STEP_DENSITY | F.STEP_CAPILLARY

    def body(fs, r):
        for n, a in zip(("VF", "CURV", "INTERFACE_LENGTH"), cap):
            fs.upload(n, a)
        ps = F.LinearSolver(fs, tol=1e-13, max_iter=500)
        ps.set_pressure_operator()
        hist = []
        for _ in range(2):
            dt = fs.adjust_dt(0.9, 0.1)
            res, it = fs.step(ps, dt, 3, flags)
            hist.append(dt)
        fs.interpolate_cell_centres()
        hist.append(fs.calc_conserved_quantities())
        hist.append(fs.monitor())
        ps.close()
        return hist

    names = ("U", "V", "P", "RHO_U", "RHO_V", "P_JUMP_U", "P_JUMP_V", "UI", "VI", "RHOI", "DIV")
    a, ra = run_ranks(F, st, bc, R, body, fields=names)
    b, rb = run_single(F, st, bc, body, fields=names)
    assert all(x == ra[0] for x in ra)
    assert ra[0][0] == pytest.approx(rb[0], rel=1e-12)
    assert np.allclose(ra[0][2], rb[2], rtol=1e-11) and np.allclose(ra[0][3][:2], rb[3][:2], rtol=1e-9)
    assert vel_err(a["U"], a["V"], b["U"], b["V"]) <= 1e-10
    compare(a, b, ("RHO_U", "RHO_V", "P_JUMP_U", "P_JUMP_V", "UI", "VI", "RHOI"), tol=1e-9)
