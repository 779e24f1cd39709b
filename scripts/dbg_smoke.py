"""debug: cavity step at several sizes, GPU vs oracle iteration counts"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import fluidsolver_b200 as F
from cases import cavity_bc, cavity_state, rel_l2, to_gpu
from oracle_lib import Oracle
o = Oracle()
for n, tol, mi in ((64, 1e-12, 200), (64, 1e-13, 500), (96, 1e-12, 200), (48, 1e-12, 200), (80, 1e-12, 200)):
    st, bc = cavity_state(o, n, n), cavity_bc()
    o.apply_velocity_bconds(st, bc)
    fs = to_gpu(F, st, bc)
    so = o.solver(n, n, tol=tol, max_iter=mi)
    so.set_operator(o.fill_pressure_operator(st))
    ps = F.LinearSolver(fs, tol=tol, max_iter=mi)
    ps.set_pressure_operator()
    dt = fs.adjust_dt(0.9, 0.1)
    fs.save_old_state(); o_dt = o.adjust_dt(st, 0.9, 0.1)
    for k in range(5):
        res, it = fs.substep(ps, dt)
        print(f"n={n} tol={tol} substep {k}: gpu iters {it} res {res:.3e} levels {ps.num_levels}", flush=True)
    it_o, res_o = o.step(st, bc, so, o_dt, 5)
    print(f"   oracle total iters {it_o} res {res_o:.3e}; U err {rel_l2(fs.download('U'), st.U):.2e}")
    ps.close(); fs.close()
