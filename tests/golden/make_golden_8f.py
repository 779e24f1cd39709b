"""Golden vectors for the rows either side of the single-phase path, from the
UNMODIFIED reference (oracle/_ref/libfsref.so):   python tests/golden/make_golden_8f.py

ops8f_<case>.npz : update_density, calc_pressure_jump, interpolate_U/V, calc_conserved_quantities,
                   correct_velocity_ib_implicit_euler, update_velocity_ib_semi_analytical
run_twophase128.npz : two timesteps of the two-phase sub-iteration (density update + capillary block
                   + operator rebuild, test/StationaryDrop.cpp:270-384) on the frozen-VOF drop
Inputs are the seeded states of tests/cases.py, re-created by the tests.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from cases import (sample_points, CHANNEL215_TOTAL_FLOW, channel_study_case, channel_study_errors, run_channel_study, cavity_bc, cavity_state, channel215_state, couette_bc,  # noqa: E402
                   couette_state, custom_channel_bc, ib_inputs, periodic_channel_bc, profile_bcs, random_state,
                   twophase_inputs)
from oracle_lib import Oracle, Ref  # noqa: E402

o = Oracle()  # input construction only


def ops8f(name, st, random_vf, dt=1e-2):
    r = Ref(st.nx, st.ny)
    out = {}
    s2 = st.copy()
    r.load(st)
    dru, drv = r.calc_drhodt()
    r.update_density(dru, drv, dt)
    r.store(s2)
    out["rho_u"], out["rho_v"] = s2.rho_u.copy(), s2.rho_v.copy()
    r.load(st)
    vf, curv, ilen = twophase_inputs(st, seed=7, random_vf=random_vf)
    r.calc_pressure_jump(vf, curv, ilen)
    r.store(s2)
    out["pj_u"], out["pj_v"] = s2.p_jump_u.copy(), s2.p_jump_v.copy()
    r.load(st)
    out["Ui"], out["Vi"] = r.interpolate_UV()
    out["conserved"] = np.array(r.conserved())
    ibu, ibv = ib_inputs(st, seed=11)
    r.ib_implicit_euler(ibu, ibv, dt)
    r.store(s2)
    out["ibe_U"], out["ibe_V"] = s2.U.copy(), s2.V.copy()
    r.load(st)
    dU, dV = r.calc_dmomdt()
    r.ib_semi_analytical(dU, dV, dt, ibu, ibv)
    r.store(s2)
    out["ibs_U"], out["ibs_V"] = s2.U.copy(), s2.V.copy()
    np.savez_compressed(os.path.join(HERE, f"ops8f_{name}.npz"), **out)
    print("ops8f", name, {k: v.shape for k, v in out.items()})


def twophase_state(n=128):
    st = cavity_state(o, n, n, drop=True, seed=1)
    st.sigma = 0.05
    return st


def run_twophase(name, nsub=3, nsteps=2):
    st, bc = twophase_state(), cavity_bc()
    cap = twophase_inputs(st, seed=7)
    o.apply_velocity_bconds(st, bc)
    r = Ref(st.nx, st.ny)
    r.load(st)
    r.set_bconds(bc)
    r.solver_create(1e-13, 500)
    r.solver_set_operator()
    dts, its = [], []
    for _ in range(nsteps):
        dt = r.adjust_dt(0.9, 0.1)
        it, res = r.step(dt, nsub, rebuild=1, density=1, capillary=cap)
        dts.append(dt)
        its.append(it)
    s2 = st.copy()
    r.store(s2)
    out = {"dts": np.array(dts), "its": np.array(its), "U": s2.U, "V": s2.V, "p": s2.a["p"],
           "rho_u": s2.rho_u, "rho_v": s2.rho_v, "pj_u": s2.p_jump_u, "pj_v": s2.p_jump_v}
    np.savez_compressed(os.path.join(HERE, f"run_{name}.npz"), **out)
    print("run", name, "its", its, "dt", dts, "max|U|", np.abs(s2.U).max())


def run_couette(name, nsub=2, nsteps=3):
    """test/Couette.cpp boundary conditions (Neumann left/right) on a seeded velocity field: three
    timesteps at tight tolerance"""
    st, bc = couette_state(o, seed=5), couette_bc()
    o.apply_velocity_bconds(st, bc)
    r = Ref(st.nx, st.ny)
    r.load(st)
    r.set_bconds(bc)
    r.solver_create(1e-13, 500)
    r.solver_set_operator()
    dts, its = [], []
    for _ in range(nsteps):
        dt = r.adjust_dt(0.9, 0.1)
        it, res = r.step(dt, nsub)
        dts.append(dt)
        its.append(it)
    s2 = st.copy()
    r.store(s2)
    np.savez_compressed(os.path.join(HERE, f"run_{name}.npz"), dts=np.array(dts), its=np.array(its), U=s2.U,
                        V=s2.V, p=s2.a["p"])
    print("run", name, "its", its, "dt", dts)


CHANNEL215 = {  # name: (bconds, flow-control mode of oracle/fs_oracle.h, total flow)
    "periodic215x43": (periodic_channel_bc, 3, CHANNEL215_TOTAL_FLOW),  # test/PeriodicChannel.cpp
    "customdirichlet215x43": (custom_channel_bc, 2, 0.0),               # test/CustomDirichletChannel.cpp
    "clippedchannel215x43": (lambda: custom_channel_bc(clipped=True), 2, 0.0),  # scaling/Channel.cpp
}


def run_channel215(name, nsub=2, nsteps=3):
    """the 215 x 43 channel drivers with their ghost-column corrections:
    the correction alone on the initial state, then timesteps at tight tolerance.  The backflow case
    stops after two: in its third step HYPRE's PCG leaves the 1e-13 request at 1.2e-7 (one of its
    breakdown exits, krylov/pcg.c:598-652,863-899, on the singular all-Neumann system), which no
    second implementation reproduces bit for bit."""
    mk_bc, mode, total = CHANNEL215[name]
    if name.startswith("clipped"):
        nsteps = 2
    st, bc = channel215_state(o, seed=8, backflow=name.startswith("clipped")), mk_bc()
    o.apply_velocity_bconds(st, bc)
    r = Ref(st.nx, st.ny)
    r.load(st)
    r.set_bconds(bc)
    out = {"flow_io": np.array(r.flow_control(mode, total))}
    s2 = st.copy()
    r.store(s2)
    out["flow_U"] = s2.U.copy()
    r.load(st)
    r.solver_create(1e-13, 500)
    r.solver_set_operator()
    dts, its = [], []
    for _ in range(nsteps):
        dt = r.adjust_dt(0.9, 0.1)
        it, res = r.step(dt, nsub, outflow_fix=mode, total_flow=total)
        dts.append(dt)
        its.append(it)
    r.store(s2)
    out.update(dts=np.array(dts), its=np.array(its), U=s2.U, V=s2.V, p=s2.a["p"])
    np.savez_compressed(os.path.join(HERE, f"run_{name}.npz"), **out)
    print("run", name, "its", its, "dt", dts, "io", out["flow_io"])


def bc_profiles(name="bc_profiles48x20"):
    """apply_velocity_bconds with function-pointer Dirichlet sides (src/BoundaryConditions.hpp:15-88)"""
    st = random_state(48, 20, seed=3)
    out = {}
    for key, bc in profile_bcs(48, 20).items():
        r = Ref(st.nx, st.ny)
        r.load(st)
        r.set_bconds(bc)
        r.apply_velocity_bconds()
        s2 = st.copy()
        r.store(s2)
        out[f"{key}_U"], out[f"{key}_V"] = s2.U, s2.V
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
    print(name, list(out))


def misc(name="misc_random48x20"):
    """calc_rho(vf, fs) / calc_visc(vf, fs) (src/FS.hpp:592-650), adjust_dt with surface tension
    (src/FS.hpp:108-114) and the monitor maxima (src/Utility.hpp:80-88)"""
    st = random_state(48, 20, seed=9)
    st.sigma = 0.05
    vf = twophase_inputs(st, seed=2, random_vf=True)[0]
    o.apply_neumann_bconds(vf)
    r = Ref(st.nx, st.ny)
    r.load(st)
    out = {"adjust_dt_sigma": np.array(r.adjust_dt(0.9, 1e9)), "diag": np.array(r.diagnostics())}
    dp = np.random.default_rng(4).uniform(-1, 1, (st.nx + 2, st.ny + 2)) + 0.3
    r.shift_pressure_to_zero(dp)  # src/Operators.hpp:52-62,78-82 (in place)
    out["dp_shifted"] = dp
    r.calc_rho_visc_vof(vf)
    s2 = st.copy()
    r.store(s2)
    out["rho_u"], out["rho_v"], out["visc"] = s2.rho_u, s2.rho_v, s2.visc
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
    print(name, out["adjust_dt_sigma"], out["diag"])


def io_files(name="io_random24x10"):
    """the files the unmodified reference writes for a seeded state: to_npy(directory, fs) (src/IO.hpp:231-269)
    and the drivers' VTK DataWriter with pressure, divergence and cell-centre velocity
    (src/VTKWriter.hpp, test/PeriodicChannel.cpp:131-134); every file as a byte array"""
    import tempfile
    st = random_state(24, 10, seed=2410)
    r = Ref(st.nx, st.ny)
    r.load(st)
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        d = os.path.join(tmp, "npy")  # to_npy removes and re-creates its directory
        r.to_npy(d)
        for f in sorted(os.listdir(d)):
            out["npy_" + f[:-4]] = np.fromfile(os.path.join(d, f), dtype=np.uint8)
        v = os.path.join(tmp, "vtk")
        os.makedirs(v)
        r.write_vtk(v, 0.125)
        (f,) = os.listdir(v)
        out["vtk"] = np.fromfile(os.path.join(v, f), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
    print(name, {k: v.size for k, v in out.items()})


def channel_study(name="channel_study", ns=(3, 4, 5)):
    """scaling/Channel.cpp run by the unmodified reference at its own settings (tol 1e-6, 50 iterations,
    NUM_SUBITER 2, CFL 0.9, T_END 20) for NY = 2^n + 1: the four error measures of :253-300"""
    out = {"n": np.array(ns)}
    rows = []
    for n in ns:
        st, bc = channel_study_case(o, n)
        r = Ref(st.nx, st.ny)
        r.load(st)
        r.set_bconds(bc)
        r.solver_create(1e-6, 50)
        r.solver_set_operator()
        steps = run_channel_study(lambda: r.adjust_dt(0.9, 0.1), lambda dt: r.step(dt, 2, outflow_fix=2))
        s2 = st.copy()
        r.store(s2)
        e = channel_study_errors(s2.a["p"], s2.U, s2.V, st.nx, st.ny, st.dx)
        rows.append(e + (steps,))
        print("study n", n, st.nx, st.ny, "steps", steps, "errors", e)
        if n == ns[0]:
            out["U_n%d" % n], out["p_n%d" % n] = s2.U, s2.a["p"]
    out["errors"] = np.array(rows)
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)


def operators_misc(name="operators_random48x20"):
    """integrate / L1_norm (src/Operators.hpp:51-76) with and without ghosts, and calc_grad_of_centered_points
    (:164-201) applied twice as src/Curvature.hpp:337-339 does, on a seeded cell field"""
    st = random_state(48, 20, seed=11)
    f = np.random.default_rng(12).uniform(-1, 1, (st.nx + 2, st.ny + 2)) + 0.2
    r = Ref(st.nx, st.ny)
    r.load(st)
    out = {"sums": np.array([r.integrate(f, g, l1) for g in (False, True) for l1 in (False, True)])}
    out["gx"], out["gy"] = r.grad_centered(f)
    out["gxx"], out["gxy"] = r.grad_centered(out["gx"])
    out["gyx"], out["gyy"] = r.grad_centered(out["gy"])
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
    print(name, out["sums"])


def sampling(name="sampling_random48x20"):
    """eval_flow_field_at on (Ui, Vi) and bilinear_interpolate on p (src/Operators.hpp:85-161) at seeded points, some of
    them outside the cell-centre box (clamped branches) and some exactly on cell centres"""
    st = random_state(48, 20, seed=21)
    px, py = sample_points(st)
    r = Ref(st.nx, st.ny)
    r.load(st)
    u, v = r.sample(0, px, py)
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), u=u, v=v, p=r.sample(1, px, py))
    print(name, u[:3], v[:3])


def vf_drop(name="vf_drop64"):
    """the initial volume fraction of test/StationaryDrop.cpp:47-52,209-212 (Gauss quadrature of the ellipse indicator)
    from the unmodified reference on 64 x 64 and 48 x 20 cells of the unit-cell-size grids the tests use"""
    out = {}
    for nx, ny in ((64, 64), (48, 20)):
        st = random_state(nx, ny, seed=1)   # only the grid matters: dx = 1 / nx, dy = 1 / ny
        r = Ref(nx, ny)
        r.load(st)
        out[f"vf{nx}x{ny}"] = r.vf_drop()
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
    print(name, {k: float(v[1:-1, 1:-1].mean()) for k, v in out.items()})


def run_twophase_quadrature(name="run_twophase_q128", nsub=3, nsteps=2):
    """two timesteps of the two-phase sub-iteration on the drop initialised as the reference's drivers do (Gauss
    quadrature of the ellipse indicator: a sharp 1 : 1000 density jump)"""
    from cases import twophase_quadrature_case
    st, bc, cap = twophase_quadrature_case(o)
    r = Ref(st.nx, st.ny)
    r.load(st)
    r.set_bconds(bc)
    r.solver_create(1e-13, 500)
    r.solver_set_operator()
    dts, its = [], []
    for _ in range(nsteps):
        dt = r.adjust_dt(0.9, 0.1)
        it, res = r.step(dt, nsub, rebuild=1, density=1, capillary=cap)
        dts.append(dt)
        its.append(it)
    s2 = st.copy()
    r.store(s2)
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), dts=np.array(dts), its=np.array(its), U=s2.U, V=s2.V,
                        p=s2.a["p"], rho_u=s2.rho_u, pj_u=s2.p_jump_u)
    print("run", name, "its", its, "dt", dts)


def run_ibchannel(name="run_ibchannel64", nsub=5, nsteps=3):
    """scaling/LuchiniIBChannel.cpp: the wall correction fields of calc_ib_correction_shape and three timesteps with
    each of the two immersed-boundary velocity updates (:212-221), with the driver's own constants: CFL 0.25, five
    sub-iterations, pressure tolerance 1e-6 and at most 50 iterations (:28-52).
    The pressure operator is singular (Dirichlet walls and inlet, flux-corrected outlet) with non-dyadic coefficients
    (rho = 10, dx = 5/64): the diagonal of PFMG's one-point coarsest grid is cancellation noise, HYPRE's
    zero-diagonal test (pfmg_setup.c:361) does not fire and its PCG breaks down below about 1e-7 (10 iterations and
    6e-8 at tol 1e-10; 561 iterations in the first step at 1e-13). At the driver's 1e-6 the solves converge, and that
    is what this fixture pins: fields to the solver tolerance, iteration counts loosely."""
    from cases import ib_channel_case
    out = {}
    for mode in (1, 2):
        st, bc = ib_channel_case(o)
        r = Ref(st.nx, st.ny)
        r.load(st)
        r.set_bconds(bc)
        ibu, ibv = r.ib_channel_correction(2.0, 3.0)
        out["ib_u"], out["ib_v"] = ibu, ibv
        r.solver_create(1e-6, 50)
        r.solver_set_operator()
        dts, its = [], []
        for _ in range(nsteps):
            dt = r.adjust_dt(0.25, 0.1)
            it, res = r.step(dt, nsub, outflow_fix=1, ib=(mode, ibu, ibv))
            dts.append(dt)
            its.append(it)
        s2 = st.copy()
        r.store(s2)
        out.update({f"dts{mode}": np.array(dts), f"its{mode}": np.array(its), f"U{mode}": s2.U, f"V{mode}": s2.V,
                    f"p{mode}": s2.a["p"]})
        print("run", name, "mode", mode, "its", its, "dt", dts, "solid points", int(np.isinf(ibu).sum()))
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)


def monitor_log(name="monitor_log.txt"):
    """src/Monitor.hpp and should_save (src/IO.hpp:97-108) of the unmodified reference on the rows / step sequence of
    tests/test_output_cpu.py"""
    from test_output_cpu import _reference_monitor_and_saves
    log, seq = _reference_monitor_and_saves(os.path.join(HERE, name))
    with open(os.path.join(HERE, name), "ab") as f:
        f.write(b"#saves " + " ".join(map(str, seq)).encode() + b"\n")
    print(log.decode(), seq)


def solve_psd(name="solve_psd128"):
    """one PCG+PFMG solve per PSDirichlet side (src/LinearSystem.hpp:10,32-155; the two-phase drivers use
    PSDirichlet::RIGHT with all-Neumann velocity BCs, test/StationaryDrop.cpp:67-72,223-224) on the
    rho-ratio-1000 drop"""
    from oracle_lib import BConds, NEUMANN
    st = cavity_state(o, 128, 128, drop=True, seed=2)
    bc = BConds(kind=(NEUMANN,) * 4)
    o.apply_velocity_bconds(st, bc)
    out = {}
    for psd in (1, 2, 3, 4):
        r = Ref(st.nx, st.ny)
        r.load(st)
        r.set_bconds(bc)
        div = r.calc_divergence()
        r.solver_create(1e-10, 300, psd)
        r.solver_set_operator()
        dp, res, it = r.solver_solve(div, 1e-2)
        out[f"dp{psd}"], out[f"res{psd}"], out[f"it{psd}"] = dp, np.array(res), np.array(it)
        print("psd", psd, "its", it, "res", res)
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)


if __name__ == "__main__":
    if "channel215" in sys.argv:
        for nm in CHANNEL215:
            run_channel215(nm)
        sys.exit(0)
    if "twophase_q" in sys.argv:
        run_twophase_quadrature()
        sys.exit(0)
    if "ib" in sys.argv:
        run_ibchannel()
        sys.exit(0)
    if "vf" in sys.argv:
        vf_drop()
        sys.exit(0)
    if "sampling" in sys.argv:
        sampling()
        sys.exit(0)
    if "monitor" in sys.argv:
        monitor_log()
        sys.exit(0)
    if "operators" in sys.argv:
        operators_misc()
        sys.exit(0)
    if "study" in sys.argv:
        channel_study()
        sys.exit(0)
    if "io" in sys.argv:
        io_files()
        sys.exit(0)
    if "psd" in sys.argv:
        solve_psd()
        sys.exit(0)
    if "misc" in sys.argv:
        misc()
        sys.exit(0)
    if "profiles" in sys.argv:
        bc_profiles()
        sys.exit(0)
    if "couette" in sys.argv:
        run_couette("couette210x21")
        sys.exit(0)
    ops8f("random48x20", random_state(48, 20, seed=48020), True)
    ops8f("drop64", cavity_state(o, 64, 64, drop=True, seed=3), False)
    run_twophase("twophase128")
    run_couette("couette210x21")
    bc_profiles()
    misc()
    solve_psd()
    for nm in CHANNEL215:
        run_channel215(nm)
    io_files()
    channel_study()
    operators_misc()
    monitor_log()
    sampling()
    vf_drop()
    run_twophase_quadrature()
