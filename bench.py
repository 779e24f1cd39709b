#!/usr/bin/env python
"""bench.py -- timestep MCell-updates/s (FP64, incl. every Poisson solve) of the stepping path.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (config.workload): lid-driven cavity, scaling/LidDrivenFlow.cpp:16-42 constants (unit
square, lid U = 1, rho = 1, mu = 1e-3, NUM_SUBITER 5, PCG tol 1e-6 / 50 its, CFL 0.9, dt_max 0.1),
8192^2 cells on one B200 -- BASELINE.json's configs[2], the configuration its target is quoted on.
One "step" = adjust_dt + save_old_state + 5 projection sub-iterations (momentum update, BCs,
divergence/rhs, multigrid-PCG solve, pressure correction).  Synthetic data: zero initial flow.

Prints ONE JSON line (rank 0).  `value` is timed with CUDA events on the library's stream with
all inputs resident in HBM; `e2e` is the same step driven through the C ABI with HOST buffers
(pinned), uploads and downloads inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NSUB, TOL, MAX_ITER, CFL, DT_MAX = 5, 1e-6, 50, 0.9, 0.1
CPU_SAMPLE_N = 2048  # grid of the bounded CPU sample (reference instantiation in ref_harness_sizes.inc)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=0, help="grid size override (default 8192; 16384 for N > 1)")
    ap.add_argument("--max-levels", type=int, default=-1, help="multigrid depth (-1: library default)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--prof", default="", help="kernel class to time for the roofline (default: auto)")
    return ap.parse_args()


# ---- clocks: sampled DURING the timed region ----------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---- the CPU leg: the UNMODIFIED reference (oracle/_ref/libfsref.so), else the C oracle --------
def cpu_leg(steps, warmup, n=CPU_SAMPLE_N):
    """Times the reference's own OpenMP + HYPRE path on this box's host cores for the same
    workload at n^2 (a bounded sample: the 8192^2 step takes the CPU about a minute)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np  # noqa: F401
    from cases import cavity_bc, cavity_state
    from oracle_lib import Oracle, Ref, have_ref
    o = Oracle() if os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")) else None
    cores = os.cpu_count() or 1
    if o is None:
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])
        o = Oracle()
    st, bc = cavity_state(o, n, n), cavity_bc()
    o.apply_velocity_bconds(st, bc)
    its = 0
    if have_ref():
        kind = "reference"
        Ref.lib().fsref_set_threads(cores)
        r = Ref(n, n)
        r.load(st)
        r.set_bconds(bc)
        r.solver_create(TOL, MAX_ITER)
        r.solver_set_operator()
        for _ in range(warmup):
            r.step(r.adjust_dt(CFL, DT_MAX), NSUB)
        t0 = time.perf_counter()
        for _ in range(steps):
            it, _ = r.step(r.adjust_dt(CFL, DT_MAX), NSUB)
            its += it
        el = time.perf_counter() - t0
    else:
        kind, cores = "port", 1
        so = o.solver(n, n, tol=TOL, max_iter=MAX_ITER)
        so.set_operator(o.fill_pressure_operator(st))
        for _ in range(warmup):
            o.step(st, bc, so, o.adjust_dt(st, CFL, DT_MAX), NSUB)
        t0 = time.perf_counter()
        for _ in range(steps):
            it, _ = o.step(st, bc, so, o.adjust_dt(st, CFL, DT_MAX), NSUB)
            its += it
        el = time.perf_counter() - t0
    return {"value": n * n * steps / el / 1e6, "unit": "MCell-updates/s", "cores": cores, "kind": kind,
            "sample": f"lid-driven cavity {n}^2, {steps} timesteps x {NSUB} sub-iterations after {warmup} warm-up, "
                      f"{its / max(steps * NSUB, 1):.1f} PCG its/solve, OpenMP threads = {cores}",
            "ms_per_step": el / steps * 1e3}
