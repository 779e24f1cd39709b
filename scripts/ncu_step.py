"""One warm 8192^2 cavity timestep + one more sub-iteration for ncu (scripts/gpu_ncu2.sh): with
FSGPU_NCU_CLASSES set, only launches of the chosen kernel classes fall inside cudaProfilerStart/Stop."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fluidsolver_b200 as F  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
fs = F.FS(n, n, 1.0 / n, 1.0 / n, 1.0, 1.0, 1e-3, 1e-3)
fs.calc_rho_visc_const()
for s, u in enumerate((0.0, 0.0, 0.0, 1.0)):
    fs.set_bconds(s, F.DIRICHLET, u, 0.0)
fs.apply_velocity_bconds()
ps = F.LinearSolver(fs, tol=1e-6, max_iter=50, max_levels=40)
ps.set_pressure_operator()
res, it = fs.step(ps, fs.adjust_dt(0.9, 0.1), 1)
fs.sync()
print("ok", res, it, F.launch_count())
