"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libfsref.so = reference
headers + vendored HYPRE 3.0.0, built by oracle/Makefile).  Run in the build container, where
/root/reference exists:   python tests/golden/make_golden.py

Inputs are the seeded states of tests/cases.py (re-created by the tests), so only the reference's
OUTPUTS are stored: each hot-path operator's result, one pressure solve, and the fields after a few
whole timesteps at tight solver tolerance.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from cases import cavity_bc, cavity_state, channel_bc, periodic_bc, random_state, tgv_state  # noqa: E402
from oracle_lib import Oracle, PSD_NONE, PSD_RIGHT, Ref  # noqa: E402

o = Oracle()  # only used to build the INPUT states (constant / VOF densities, initial BCs)


def operators(name, st, bc, dt=1e-2):
    r = Ref(st.nx, st.ny)
    out = {}
    r.load(st)
    out["adjust_dt"] = np.array(r.adjust_dt(0.5, 1e9))
    r.calc_mid_time()
    s2 = st.copy()
    r.store(s2)
    out["mid_U"], out["mid_V"] = s2.U.copy(), s2.V.copy()
    dU, dV = r.calc_dmomdt()
    out["dmomU"], out["dmomV"] = dU, dV
    r.update_velocity(dU, dV, dt)
    r.set_bconds(bc)
    r.apply_velocity_bconds()
    r.store(s2)
    out["upd_U"], out["upd_V"] = s2.U.copy(), s2.V.copy()
    div = r.calc_divergence()
    out["div"] = div
    for psd in (PSD_NONE, PSD_RIGHT):
        out[f"rhs_psd{psd}"] = r.fill_pressure_rhs(div, dt, psd)
    out["op_psd0"] = r.fill_pressure_operator(PSD_NONE)
    dru, drv = r.calc_drhodt()
    out["drho_u"], out["drho_v"] = dru, drv
    np.savez_compressed(os.path.join(HERE, f"ops_{name}.npz"), **out)
    print("ops", name, {k: v.shape for k, v in out.items()})


def solve_and_steps(name, st, bc, nsub, nsteps, cfl, dtmax, outflow_fix=0, rebuild=0):
    """reference PCG+PFMG: one solve at tol 1e-6 and 1e-12, then nsteps timesteps at 1e-13"""
    out = {}
    o.apply_velocity_bconds(st, bc)
    r = Ref(st.nx, st.ny)
    r.load(st)
    r.set_bconds(bc)
    div = r.calc_divergence()
    for tol in (1e-6, 1e-12):
        r.solver_create(tol, 200)
        r.solver_set_operator()
        dp, res, it = r.solver_solve(div, 1e-2)
        out[f"dp_tol{tol:g}"], out[f"res_tol{tol:g}"], out[f"it_tol{tol:g}"] = dp, np.array(res), np.array(it)
    r.solver_create(1e-13, 500)
    r.solver_set_operator()
    dts, its = [], []
    for _ in range(nsteps):
        dt = r.adjust_dt(cfl, dtmax)
        it, res = r.step(dt, nsub, outflow_fix=outflow_fix, rebuild=rebuild)
        dts.append(dt)
        its.append(it)
    s2 = st.copy()
    r.store(s2)
    out["dts"], out["its"] = np.array(dts), np.array(its)
    out["U"], out["V"], out["p"] = s2.U, s2.V, s2.a["p"]
    np.savez_compressed(os.path.join(HERE, f"run_{name}.npz"), **out)
    print("run", name, "its", its, "dt", dts[:2])


if __name__ == "__main__":
    from oracle_lib import BConds, DIRICHLET, NEUMANN, SYMMETRY
    mixed = BConds(kind=(DIRICHLET, NEUMANN, SYMMETRY, DIRICHLET), u_val=(1.0, 0, 0, 0.5), v_val=(0.1, 0, 0, 0))
    operators("random48x20", random_state(48, 20, seed=48020), mixed)
    operators("random64", random_state(64, 64, seed=64064), mixed)
    operators("tgv32", tgv_state(o, 32), periodic_bc())
    solve_and_steps("tgv32", tgv_state(o, 32), periodic_bc(), 2, 3, 0.5, 1e-2)
    solve_and_steps("cavity64", cavity_state(o, 64, 64, seed=1), cavity_bc(), 5, 3, 0.9, 0.1)
    st = cavity_state(o, 256, 64, lx=100.0, ly=1.0, rho=0.5)
    st.a["U"][:] = 1.0
    solve_and_steps("channel256x64", st, channel_bc(), 5, 3, 0.9, 0.1, outflow_fix=1)
    # variable density: the reference's PCG+PFMG only converges when the 16-level hierarchy does not
    # reach 1-2 point grids (64^2 and 96^2 end in NaN residuals in the reference itself), so 128^2
    solve_and_steps("drop128", cavity_state(o, 128, 128, drop=True, seed=1), cavity_bc(), 3, 2, 0.9, 0.1, rebuild=1)
