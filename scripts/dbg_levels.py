"""debug: GPU vs oracle after k PCG iterations as a function of the multigrid depth"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import fluidsolver_b200 as F
from cases import cavity_bc, cavity_state, rel_l2, to_gpu
from oracle_lib import Oracle
o = Oracle()
for n in (int(x) for x in (sys.argv[1:] or ["64", "48"])):
    st, bc = cavity_state(o, n, n, seed=3), cavity_bc()
    o.apply_velocity_bconds(st, bc)
    op = o.fill_pressure_operator(st)
    div = o.calc_divergence(st) if hasattr(o, "calc_divergence") else None
    for ml in range(1, 17):
        fs = to_gpu(F, st, bc)
        fs.calc_divergence()
        ps = F.LinearSolver(fs, tol=1e-30, max_iter=3, max_levels=ml)
        ps.set_pressure_operator()
        ps.set_pressure_rhs(1e-2)
        rhs = fs.download("RHS")
        res, it = ps.solve()
        dp = fs.download("DELTA_P")
        so = o.solver(n, n, tol=1e-30, max_iter=3, max_levels=ml)
        so.set_operator(op)
        sol, res_o, it_o = so.solve(rhs)
        print(f"n={n} max_levels={ml} nl={ps.num_levels}: gpu it {it} res {res:.3e} | oracle it {it_o} res {res_o:.3e} | dp err {rel_l2(dp, sol):.2e}", flush=True)
        ps.close(); fs.close()
