"""GPU parity, stencil operators: every kernel called through the C ABI (libfsgpu.so) against
the CPU oracle (oracle/fs_oracle.c) on the same seeded inputs.

Tolerance: north_star asks for "stencil outputs within 1e-12 relative"; SURVEY.md section 7 shows
this must be norm-relative (||gpu - ref||_inf / ||ref||_inf), never pointwise.  TOL is that 1e-12.
"""
import numpy as np
import pytest

from cases import (cavity_bc, cavity_state, channel_bc, random_state, rel_inf, tgv_state, to_gpu)
from oracle_lib import (ALL_FIELDS, BConds, DIRICHLET, NEUMANN, PERIODIC, SYMMETRY, PSD_NONE, PSD_LEFT,
                        PSD_RIGHT, PSD_BOTTOM, PSD_TOP)

pytestmark = pytest.mark.gpu
TOL = 1e-12
SIZES = [(64, 64), (48, 20), (20, 48), (130, 67), (37, 301)]


def _down(fs, names):
    return {n: fs.download(n.upper()) for n in names}


@pytest.mark.parametrize("nx,ny", SIZES)
def test_mid_time_dmomdt_update(F, oracle, nx, ny):
    st = random_state(nx, ny, seed=nx * 1000 + ny)
    fs = to_gpu(F, st)
    # calc_mid_time
    oracle.calc_mid_time(st)
    fs.calc_mid_time()
    assert rel_inf(fs.download("U"), st.U) <= TOL
    assert rel_inf(fs.download("V"), st.V) <= TOL
    # calc_dmomdt
    dU, dV = oracle.calc_dmomdt(st)
    fs.calc_dmomdt()
    gU, gV = fs.download("DMOM_U"), fs.download("DMOM_V")
    assert rel_inf(gU, dU) <= TOL and rel_inf(gV, dV) <= TOL
    # ghosts of the outputs are zero, as after fill(dmomUdt, 0) (FS.hpp:173-174)
    assert gU[0].max() == 0 and gU[-1].max() == 0 and np.abs(gU[:, 0]).max() == 0
    # update_velocity
    dt = 1e-3
    oracle.update_velocity(st, dU, dV, dt)
    fs.update_velocity(dt)
    assert rel_inf(fs.download("U"), st.U) <= TOL
    assert rel_inf(fs.download("V"), st.V) <= TOL
    fs.close()


@pytest.mark.parametrize("nx,ny", SIZES)
@pytest.mark.parametrize("ratio", [1.0, 1000.0])
def test_fused_momentum_update(F, oracle, nx, ny, ratio):
    """fsgpu_momentum_update == calc_mid_time + calc_dmomdt + update_velocity (interior faces;
    ghosts are whatever apply_velocity_bconds writes next, so BCs are applied on both sides)."""
    st = random_state(nx, ny, seed=7 + nx, rho_ratio=ratio)
    bc = BConds(kind=(DIRICHLET, NEUMANN, SYMMETRY, DIRICHLET), u_val=(1.0, 0, 0, 0.5), v_val=(0.1, 0, 0, 0))
    fs = to_gpu(F, st, bc)
    dt = 2e-3
    oracle.calc_mid_time(st)
    dU, dV = oracle.calc_dmomdt(st)
    oracle.update_velocity(st, dU, dV, dt)
    oracle.apply_velocity_bconds(st, bc)
    fs.momentum_update(dt)
    fs.apply_velocity_bconds()
    assert rel_inf(fs.download("U"), st.U) <= TOL
    assert rel_inf(fs.download("V"), st.V) <= TOL
    fs.close()


KINDS = [DIRICHLET, NEUMANN, PERIODIC, SYMMETRY]


@pytest.mark.parametrize("kinds", [(k,) * 4 for k in KINDS] + [(DIRICHLET, NEUMANN, DIRICHLET, DIRICHLET),
                                                                (NEUMANN, DIRICHLET, SYMMETRY, NEUMANN),
                                                                (PERIODIC, PERIODIC, DIRICHLET, SYMMETRY)])
@pytest.mark.parametrize("clipped", [0, 1])
def test_apply_velocity_bconds(F, oracle, kinds, clipped):
    nx, ny = 37, 22
    st = random_state(nx, ny, seed=3)
    bc = BConds(kind=kinds, u_val=(0.3, -0.2, 0.1, 1.0), v_val=(0.05, 0.0, -0.4, 0.2), clipped=(clipped,) * 4)
    fs = to_gpu(F, st, bc)
    oracle.apply_velocity_bconds(st, bc)
    fs.apply_velocity_bconds()
    assert np.array_equal(fs.download("U"), st.U)  # pure copies / sign flips / 2*c - x: bit exact
    assert np.array_equal(fs.download("V"), st.V)
    fs.close()


def test_dirichlet_profiles(F, oracle):
    """function-pointer Dirichlet (BoundaryConditions.hpp:15-17) as per-point profiles"""
    nx, ny = 40, 24
    st = random_state(nx, ny, seed=11)
    rng = np.random.default_rng(5)
    up = (rng.uniform(size=ny + 2), None, rng.uniform(size=nx + 3), None)
    vp = (rng.uniform(size=ny + 3), None, None, rng.uniform(size=nx + 2))
    bc = BConds(kind=(DIRICHLET,) * 4, u_val=(0, 0.5, 0, 1.0), v_val=(0, 0.1, 0.2, 0), u_prof=up, v_prof=vp)
    fs = to_gpu(F, st, bc)
    oracle.apply_velocity_bconds(st, bc)
    fs.apply_velocity_bconds()
    assert np.array_equal(fs.download("U"), st.U)
    assert np.array_equal(fs.download("V"), st.V)
    fs.close()


def test_outflow_fix(F, oracle):
    nx, ny = 60, 30
    st = random_state(nx, ny, seed=2)
    fs = to_gpu(F, st)
    e_ref = oracle.outflow_fix(st)
    e_gpu = fs.outflow_fix()
    assert abs(e_gpu - e_ref) <= 1e-12 * max(1.0, abs(e_ref))
    assert rel_inf(fs.download("U"), st.U) <= TOL
    fs.close()


@pytest.mark.parametrize("nx,ny", SIZES)
@pytest.mark.parametrize("psd", [PSD_NONE, PSD_LEFT, PSD_RIGHT, PSD_BOTTOM, PSD_TOP])
def test_divergence_rhs(F, oracle, nx, ny, psd):
    st = random_state(nx, ny, seed=5)
    fs = to_gpu(F, st)
    div = oracle.calc_divergence(st)
    fs.calc_divergence()
    assert rel_inf(fs.download("DIV"), div) <= TOL
    dt = 1e-2
    rhs = oracle.fill_pressure_rhs(st, div, dt, psd)  # Layout::F: [j][i]
    ps = F.LinearSolver(fs, psd=psd)
    ps.set_pressure_rhs(dt)
    assert rel_inf(fs.download("RHS"), rhs.T) <= TOL
    ps.close()
    fs.close()


@pytest.mark.parametrize("nx,ny", SIZES)
@pytest.mark.parametrize("psd", [PSD_NONE, PSD_LEFT, PSD_RIGHT, PSD_BOTTOM, PSD_TOP])
def test_fused_divergence_rhs(F, oracle, nx, ny, psd):
    """k_rhs<true> (fsgpu_ps_set_rhs_from_velocity): calc_divergence + fill_pressure_rhs in ONE pass over U, V --
    the kernel fsgpu_substep and bench.py run -- against the two reference calls (src/Operators.hpp:32-41,
    src/LinearSystem.hpp:125-155), all five PSDirichlet sides"""
    st = random_state(nx, ny, seed=15)
    fs = to_gpu(F, st)
    dt = 7e-3
    div = oracle.calc_divergence(st)
    rhs = oracle.fill_pressure_rhs(st, div, dt, psd)
    ps = F.LinearSolver(fs, psd=psd)
    ps.set_pressure_rhs_from_velocity(dt)
    assert rel_inf(fs.download("DIV"), div) <= TOL
    assert rel_inf(fs.download("RHS"), rhs.T) <= TOL
    ps.close()
    fs.close()


@pytest.mark.parametrize("nx,ny", SIZES)
@pytest.mark.parametrize("ratio", [1.0, 1000.0])
def test_fused_projection(F, oracle, nx, ny, ratio):
    """k_project<true> (fsgpu_project): shift_pressure_to_zero + p += delta_p + the velocity correction in ONE pass
    -- the kernel fsgpu_substep and bench.py run -- against the three driver lines (src/Operators.hpp:78-82,
    test/LaminarChannel.cpp:222-233) on a delta_p with a large mean"""
    st = random_state(nx, ny, seed=19 + nx, rho_ratio=ratio)
    rng = np.random.default_rng(nx + ny)
    dp = rng.uniform(-1, 1, (nx + 2, ny + 2)) + 3.0
    fs = to_gpu(F, st)
    fs.upload("DELTA_P", dp)
    dt = 5e-3
    oracle.shift_pressure_to_zero(st, dp)
    oracle.correct(st, dp, dt)
    fs.project(dt)
    assert rel_inf(fs.download("DELTA_P"), dp) <= TOL
    for n in ("U", "V", "P"):
        assert rel_inf(fs.download(n), st.a[n.lower() if n == "P" else n]) <= TOL, n
    # twice in a row: the delta_p / rhs buffer swap must leave a usable pair behind
    dp2 = rng.uniform(-1, 1, (nx + 2, ny + 2))
    fs.upload("DELTA_P", dp2)
    oracle.shift_pressure_to_zero(st, dp2)
    oracle.correct(st, dp2, dt)
    fs.project(dt)
    assert rel_inf(fs.download("DELTA_P"), dp2) <= TOL
    for n in ("U", "V", "P"):
        assert rel_inf(fs.download(n), st.a[n.lower() if n == "P" else n]) <= TOL, n
    fs.close()


@pytest.mark.parametrize("ratio", [1.0, 1000.0])
@pytest.mark.parametrize("psd", [PSD_NONE, PSD_LEFT, PSD_RIGHT, PSD_BOTTOM, PSD_TOP])
def test_substep_kernel_chain(F, oracle, ratio, psd):
    """every output of the kernel chain fsgpu_substep launches (and bench.py times), in its order, on the random
    density-jump state: k_momentum<1> -> BCs -> k_rhs<true> -> [solve] -> k_project<true>"""
    nx, ny = 66, 50
    st = random_state(nx, ny, seed=23, rho_ratio=ratio)
    bc = BConds(kind=(DIRICHLET, NEUMANN, DIRICHLET, SYMMETRY), u_val=(1.0, 0, 0, 0.5), v_val=(0.1, 0, 0, 0))
    oracle.apply_velocity_bconds(st, bc)
    fs = to_gpu(F, st, bc)
    dt = 2e-3
    oracle.calc_mid_time(st)
    dU, dV = oracle.calc_dmomdt(st)
    oracle.update_velocity(st, dU, dV, dt)
    oracle.apply_velocity_bconds(st, bc)
    fs.momentum_update(dt)
    fs.apply_velocity_bconds()
    assert rel_inf(fs.download("U"), st.U) <= TOL and rel_inf(fs.download("V"), st.V) <= TOL
    div = oracle.calc_divergence(st)
    rhs = oracle.fill_pressure_rhs(st, div, dt, psd)
    ps = F.LinearSolver(fs, psd=psd)
    ps.set_pressure_rhs_from_velocity(dt)
    assert rel_inf(fs.download("DIV"), div) <= TOL
    assert rel_inf(fs.download("RHS"), rhs.T) <= TOL
    dp = np.random.default_rng(3).uniform(-1, 1, (nx + 2, ny + 2))
    fs.upload("DELTA_P", dp)
    oracle.shift_pressure_to_zero(st, dp)
    oracle.correct(st, dp, dt)
    fs.project(dt)
    for n in ("U", "V", "P"):
        assert rel_inf(fs.download(n), st.a[n.lower() if n == "P" else n]) <= TOL, n
    ps.close()
    fs.close()


def test_shift_and_correct(F, oracle):
    nx, ny = 50, 34
    st = random_state(nx, ny, seed=9)
    rng = np.random.default_rng(1)
    dp = rng.uniform(-1, 1, (nx + 2, ny + 2))
    fs = to_gpu(F, st)
    fs.upload("DELTA_P", dp)
    oracle.shift_pressure_to_zero(st, dp)
    fs.shift_pressure_to_zero()
    assert rel_inf(fs.download("DELTA_P"), dp) <= TOL
    dt = 5e-3
    oracle.correct(st, dp, dt)
    fs.correct_velocity(dt)
    for n in ("U", "V", "P"):
        assert rel_inf(fs.download(n), st.a[n.lower() if n == "P" else n]) <= TOL, n
    fs.close()


def test_adjust_dt_absmax(F, oracle):
    nx, ny = 45, 61
    st = random_state(nx, ny, seed=4)
    fs = to_gpu(F, st)
    assert fs.adjust_dt(0.5, 1e9) == pytest.approx(oracle.adjust_dt(st, 0.5, 1e9), rel=1e-14)
    assert fs.adjust_dt(0.5, 1e-9) == 1e-9
    for n in ("U", "V", "P"):   # abs_max<INCLUDE_GHOST = false> (interior) and abs_max<true>, src/Utility.hpp:44-88
        a = st.a[n.lower() if n == "P" else n]
        assert fs.abs_max(n) == oracle.abs_max(a) == np.abs(a[1:-1, 1:-1]).max()
        assert fs.abs_max(n, include_ghost=True) == oracle.abs_max(a, include_ghost=True) == np.abs(a).max()
    fs.close()


def test_rho_visc_vof(F, oracle):
    nx, ny = 64, 48
    st = cavity_state(oracle, nx, ny, drop=True)
    fs = F.FS(nx, ny, st.dx, st.dy, st.rho_gas, st.rho_liquid, 1e-3, 5e-3)
    st.visc_liquid = 5e-3
    oracle.calc_rho_visc_vof(st, st.vf)
    fs.upload("VF", st.vf)
    fs.calc_rho_visc_vof()
    for n in ("RHO_U", "RHO_V", "VISC"):
        assert rel_inf(fs.download(n), st.a[n.lower()]) <= 1e-15, n
    fs.save_old_density()
    assert np.array_equal(fs.download("RHO_U_OLD"), fs.download("RHO_U"))
    fs.close()


def test_upload_download_roundtrip_and_pads(F):
    """ragged sizes: every field class survives the padded device layout bit for bit"""
    for nx, ny in [(2, 2), (3, 17), (31, 5), (129, 255)]:
        fs = F.FS(nx, ny, 0.1, 0.2)
        rng = np.random.default_rng(nx)
        for n in F.FIELDS:
            a = rng.uniform(-1, 1, F.field_shape(n, nx, ny))
            fs.upload(n, a)
            assert np.array_equal(fs.download(n), a), (n, nx, ny)
        fs.close()


@pytest.mark.parametrize("y", [1.0 / 8192, 1.0 / 96, 1.0 / 3, 100.0 / 215, 0.1, 1e-3, 7.123456789e-5, 3.0, 2.0 - 2.0 ** -52])
def test_division_by_uniform_divisor_is_exact(F, y):
    """calc_divergence / fill_pressure_rhs / the projection divide by dx, dy, dt as the reference does
    (src/Operators.hpp:36-39, src/LinearSystem.hpp:134, scaling/LidDrivenFlow.cpp:155-159); the kernels form
    those quotients from a reciprocal with one FMA correction (csrc/common.cuh DivBy), which must be the
    correctly rounded quotient: compared bit for bit with the FP64 division on 2^24 random numerators.
    The last divisor has an all-ones significand -- the one case the shortcut does not cover, where the
    kernels fall back to the division itself."""
    fast, bad = F.selftest_division(y, 1 << 24, seed=12345)
    assert bad == 0
    assert fast == (0 if y == 2.0 - 2.0 ** -52 else 1)
