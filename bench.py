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
# every kernel is loaded when the library is, not at its first launch (a first launch under lazy loading waits for
# the device to drain, which a rank whose kernel is waiting for a neighbour's flag does not do quickly)
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
# CPU-baseline protocol: all host cores, threads pinned (OMP_PROC_BIND=close,
# OMP_PLACES=cores; without binding, 8 spinning threads on the build container's 8 vCPUs ran the reference 4x
# slower).  The binding must be in the environment before any OpenMP runtime is loaded AND it also pins the
# process's initial thread to the first core -- under torchrun that put the launch threads of all N ranks on
# core 0 (round 1's 8-GPU collapse).  So the CPU legs run in a child process of their own with that
# environment (cpu_leg -> _cpu_leg_child), and the GPU ranks never see it.
CPU_CHILD_ENV = {"OMP_PROC_BIND": "close", "OMP_PLACES": "cores"}

NSUB, TOL, MAX_ITER, CFL, DT_MAX = 5, 1e-6, 50, 0.9, 0.1
# The CPU legs run the SAME configuration as the GPU arm: the 8192^2 cavity (the reference
# harness instantiates it, oracle/ref_harness_sizes.inc).  Bounded in the number of timesteps, not in the grid:
# CPU_WARMUP warm-up + CPU_STEPS timed timesteps (a timestep takes the host tens of seconds).  FSBENCH_CPU_N
# overrides the grid for quick local runs only.
CPU_SAMPLE_N = int(os.environ.get("FSBENCH_CPU_N", "8192"))
CPU_STEPS, CPU_WARMUP = int(os.environ.get("FSBENCH_CPU_STEPS", "2")), 1


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--grid", dest="n", type=int, default=0, help="grid size override (default 8192; 16384 for N > 1)")
    ap.add_argument("--max-levels", type=int, default=40,
                    help="multigrid depth cap (default 40 = uncapped: coarsen down to one point; 16 = the "
                         "reference's PS_PFMG_MAX_LEVELS, src/LinearSolver_StructHypre.hpp:221-229)")
    ap.add_argument("--no-ref-depth", action="store_true", help="skip the secondary 16-level measurement")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-anchor", action="store_true", help="N > 1: skip the 1-GPU timing of the same grid")
    ap.add_argument("--halo", default="", choices=["", "nccl", "p2p"],
                    help="N > 1: sets FSGPU_HALO (default: the library's, i.e. peer-mapped mailboxes + fused push; nccl = NCCL send/recv)")
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
    """The CPU leg in a child process (its OpenMP binding must not leak into the GPU ranks)."""
    env = dict(os.environ)
    for k, v in CPU_CHILD_ENV.items():
        env.setdefault(k, v)
    for k in ("OMP_NUM_THREADS", "RANK", "WORLD_SIZE", "LOCAL_RANK"):  # torchrun sets OMP_NUM_THREADS=1
        env.pop(k, None)
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--cpu-child", str(steps), str(warmup), str(n)],
                       env=env, stdout=subprocess.PIPE, text=True, check=True)
    return json.loads(r.stdout.strip().splitlines()[-1])


def _cpu_leg_child(steps, warmup, n):
    """Times the reference's own OpenMP + HYPRE path on this box's host cores for the same
    workload and grid as the GPU arm (bounded in the number of timesteps)."""
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


def single_gpu_anchor(F, torch, n, dev, max_levels, steps=3):
    """The whole n^2 cavity on ONE GPU with the settings of the timed run: the 1-GPU point of the strong-scaling
    curve on the same problem (16384^2 needs ~66 GB of the 180 GB)."""
    try:
        fs = F.FS(n, n, 1.0 / n, 1.0 / n, 1.0, 1.0, 1e-3, 1e-3, device=dev)
        fs.calc_rho_visc_const()
        for s, u in enumerate((0.0, 0.0, 0.0, 1.0)):
            fs.set_bconds(s, F.DIRICHLET, u, 0.0)
        fs.apply_velocity_bconds()
        ps = F.LinearSolver(fs, tol=TOL, max_iter=MAX_ITER, max_levels=max_levels)
        ps.set_pressure_operator()
        stream = torch.cuda.ExternalStream(fs.stream(), device=torch.device("cuda", dev))
        fs.step(ps, fs.adjust_dt(CFL, DT_MAX), NSUB)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fs.sync()
        e0.record(stream)
        its = 0
        for _ in range(steps):
            its += fs.step(ps, fs.adjust_dt(CFL, DT_MAX), NSUB)[1]
        e1.record(stream)
        fs.sync()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        umax = fs.abs_max("U")
        ps.close()
        fs.close()
        return {"n_gpus": 1, "grid": n, "value": n * n * steps / (ms * 1e-3) / 1e6, "unit": "MCell-updates/s",
                "ms_per_step": ms / steps, "steps": steps, "warmup": 1,
                "pcg_iters_per_solve": its / (steps * NSUB), "max_abs_U_after_4_steps": umax}
    except Exception as e:  # e.g. a grid that does not fit one GPU
        return {"n_gpus": 1, "grid": n, "unavailable": str(e)[:200]}


# ---- algorithmic bytes per cell of each kernel class (DESIGN.md "Kernels and rooflines") -------
# doubles moved once per launch by the algorithm as implemented (8 B each), per fine-level cell
ALGO_BYTES = {
    "matvec_dot": 40.0,        # read p, d, cx, cy; write s
    "xr_update": 48.0,         # read x, p, r, s; write x, r
    "p_update": 24.0,          # read z, p; write p
    "vcycle_down_l0": 36.0,    # read r, d, cx, cy; write b_coarse (0.5)   (dinv, pc recomputed in registers)
    "vcycle_up_l0": 44.0,      # read r, d, cx, cy, x_coarse (0.5); write z (dinv, pc recomputed in registers)
    "momentum": 112.0,         # 12 fields read, U and V written
    "div_rhs": 32.0,           # read U, V; write div, rhs (the mean is subtracted by the solver's first readers)
    "shift_correct": 80.0,     # read dp, p, U, V, rho_u, rho_v; write dp', p, U, V (one pass; the sum of dp rides on PCG)
}


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    n_gpus = a.gpus
    if a.impl == "reference":
        if rank != 0:
            return 0
        # same grid and constants as the GPU arm; the timed steps are capped (each costs the host tens of
        # seconds) and the counts actually run are the ones reported
        ks, kw = max(1, min(a.steps, CPU_STEPS)), min(a.warmup, CPU_WARMUP)
        cb = cpu_leg(ks, kw)
        line = {"impl": "reference", "metric": "timestep MCell-updates/s (FP64, incl. Poisson solve)",
                "value": cb["value"], "unit": "MCell-updates/s", "n_gpus": n_gpus, "steps": ks,
                "warmup": kw, "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"lid-driven cavity {CPU_SAMPLE_N}^2 (scaling/LidDrivenFlow.cpp:16-42), NUM_SUBITER {NSUB}, "
                                       f"PCG tol {TOL} max_iter {MAX_ITER}, CFL {CFL}",
                           "solver": "the reference's own: HYPRE Struct PCG + PFMG, PS_PFMG_MAX_LEVELS 16",
                           "steps_requested": a.steps, "warmup_requested": a.warmup},
                "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": cb["value"], "unit": "MCell-updates/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line), flush=True)
        return 0

    import numpy as np
    import torch
    import fluidsolver_b200 as F

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    dist = None
    comm_id = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        # the library's own NCCL communicator: rank 0 draws the id, torch.distributed carries it
        box = [F.comm_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        comm_id = box[0]
    if a.halo:
        os.environ["FSGPU_HALO"] = a.halo
    # N = 1: BASELINE configs[2] (8192^2, the configuration the metric is quoted on);
    # N > 1: configs[4], the 16384^2 strong-scaling case, slab-decomposed over the N GPUs
    n = a.n or (8192 if world == 1 else 16384)

    # N > 1: the same-problem anchor of the scaling curve -- rank 0 alone runs the WHOLE n^2 grid on its one GPU
    # (same solver settings, 1 warm-up + 3 timed steps) before the slabs are created; the others wait.
    anchor = None
    if world > 1 and not a.no_anchor:
        if rank == 0:
            anchor = single_gpu_anchor(F, torch, n, local, max(a.max_levels, 0))
        dist.barrier()

    fs = F.FS(n, n, 1.0 / n, 1.0 / n, 1.0, 1.0, 1e-3, 1e-3, device=local, rank=rank, nranks=world,
              comm_id=comm_id)
    fs.calc_rho_visc_const()
    for s, u in enumerate((0.0, 0.0, 0.0, 1.0)):
        fs.set_bconds(s, F.DIRICHLET, u, 0.0)
    fs.apply_velocity_bconds()
    # Multigrid depth.  The reference caps PFMG at 16 semicoarsened levels (a knob of its HYPRE
    # wrapper); from 2048^2 upwards that leaves a coarsest grid far too large for its few Jacobi
    # sweeps and the PCG iteration count climbs (13 at 8192^2, 23 at 16384^2 instead of 7).  The
    # GPU solver coarsens down to a single point by default here -- same stopping tolerance, same
    # operator, same smoother -- and the capped configuration is timed next to it
    # (config.reference_depth_16).
    ps = F.LinearSolver(fs, tol=TOL, max_iter=MAX_ITER, max_levels=max(a.max_levels, 0))
    ps.set_pressure_operator()
    stream = torch.cuda.ExternalStream(fs.stream(), device=torch.device("cuda", local))

    def one_step():
        dt = fs.adjust_dt(CFL, DT_MAX)
        return fs.step(ps, dt, NSUB)

    for _ in range(a.warmup):
        one_step()
    prof_cls = a.prof or "vcycle_up_l0"
    F.prof_enable((prof_cls,))
    F.prof_read()
    clocks = ClockSampler(local)
    if rank == 0:  # one poller per job: N pollers contend for the driver lock beside the kernels they time
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fs.sync()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    l0 = F.launch_count()
    e0.record(stream)
    its, last_res = 0, 0.0
    for _ in range(a.steps):
        last_res, it = one_step()
        its += it
    e1.record(stream)
    fs.sync()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    if dist:  # device time, MAX over ranks
        t = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    launches = F.launch_count() - l0
    clk = clocks.stop()
    prof = F.prof_read()
    F.prof_enable(())
    if not last_res < TOL:  # a step whose solves stop short of the tolerance is not the workload
        raise SystemExit(f"bench.py: the last pressure solve did not converge (residual {last_res} >= {TOL})")
    value = n * n * a.steps / (ms * 1e-3) / 1e6
    k_bar = its / (a.steps * NSUB)

    # roofline of the dominant kernel, timed live with CUDA events on the launching stream
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    a_lo, a_hi = F.slab_range(n, rank, world)
    cells = (a_hi - a_lo) * (n + 2)  # level-0 unknowns this rank's kernels sweep per launch
    kms, kcnt = prof.get(prof_cls, (0.0, 0))
    achieved = ALGO_BYTES[prof_cls] * cells / (kms / max(kcnt, 1) * 1e-3) / 1e9 if kcnt else None
    # DRAM bytes per launch of that kernel from the committed ncu --set full capture (same grid only)
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if tr["grid"] == n and world == 1 and prof_cls in tr["dram_bytes_per_launch"]:
            e = tr["dram_bytes_per_launch"][prof_cls]
            traffic = e["read"] + e["write"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": prof_cls, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak if achieved else None, "traffic": traffic,
                "traffic_source": "profiles/ncu_traffic.json (dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full)" if traffic else None,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                "algorithmic_bytes_per_launch": ALGO_BYTES[prof_cls] * cells,
                "avg_launch_us": kms / max(kcnt, 1) * 1e3, "launches_timed": kcnt,
                "launches_note": "launches that did work: the copies PCG enqueues one iteration past convergence "
                                 "(they return at their first instruction) are excluded from the average",
                "share_of_step": kms / ms if ms else None}
    # whole-step roofline: SURVEY.md section 8d  B_step(k) = NSUB (232 + 160 + 264 k) + 120 bytes per cell
    b_step = NSUB * (232 + 160 + 264 * k_bar) + 120
    step_gbs = b_step * n * n * a.steps / (ms * 1e-3) / 1e9
    roofline["step_algorithmic_GBs"] = step_gbs
    roofline["step_frac"] = step_gbs / (peak * world)  # N GPUs: N x the single-GPU peak

    # e2e: the same step through the C ABI with HOST buffers (pinned), copies inside the timed region
    e2e = None
    if not a.no_e2e:
        names = ("U", "V", "P")
        host, r0s = {}, {}
        for nm in names:  # each rank pins only its own rows (+ halo rows) of the global arrays
            r0, r1 = fs.local_rows(nm)
            t = torch.empty((r1 - r0, F.field_shape(nm, n, n)[1]), dtype=torch.float64, pin_memory=True)
            host[nm], r0s[nm] = t.numpy(), r0
            fs.download_rows(nm, host[nm], r0)
        ke = max(1, min(a.steps, 3))
        nbytes = sum(h.nbytes for h in host.values())
        fs.sync()
        if dist:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(ke):
            for nm in names:
                fs.upload_rows(nm, host[nm], r0s[nm])
            one_step()
            for nm in names:
                fs.download_rows(nm, host[nm], r0s[nm])
        fs.sync()
        if dist:
            dist.barrier()
        el = time.perf_counter() - t0
        if dist:
            t = torch.tensor([el, float(nbytes)], dtype=torch.float64, device=f"cuda:{local}")
            tm = t.clone()
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            el, nbytes = float(tm[0].item()), int(t[1].item())
        e2e = {"value": n * n * ke / el / 1e6, "unit": "MCell-updates/s", "h2d_bytes_per_step": nbytes,
               "d2h_bytes_per_step": nbytes, "steps": ke, "ms_per_step": el / ke * 1e3,
               "api": "fsgpu_upload(U,V,P) -> fsgpu_adjust_dt -> fsgpu_step -> fsgpu_download(U,V,P), pinned host "
                      "buffers" + (" (each rank its own slab rows; bytes summed over ranks)" if dist else "")}
    umax = fs.abs_max("U")
    direct = fs.halo_is_direct()
    levels = (ps.num_levels, ps.num_distributed_levels)
    ps.close()
    # the same steps with the reference's 16-level cap (secondary number, device-resident timing)
    ref_depth = None
    if not a.no_ref_depth and levels[0] > 16:
        ps16 = F.LinearSolver(fs, tol=TOL, max_iter=MAX_ITER, max_levels=16)
        ps16.set_pressure_operator()
        k16 = max(1, min(a.steps, 3))
        fs.step(ps16, fs.adjust_dt(CFL, DT_MAX), NSUB)  # warm-up
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fs.sync()
        if dist:
            dist.barrier()
        f0.record(stream)
        it16 = 0
        for _ in range(k16):
            it16 += fs.step(ps16, fs.adjust_dt(CFL, DT_MAX), NSUB)[1]
        f1.record(stream)
        fs.sync()
        torch.cuda.synchronize()
        ms16 = f0.elapsed_time(f1)
        if dist:
            t = torch.tensor([ms16], dtype=torch.float64, device=f"cuda:{local}")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms16 = float(t.item())
        ref_depth = {"value": n * n * k16 / (ms16 * 1e-3) / 1e6, "unit": "MCell-updates/s", "mg_levels": 16,
                     "pcg_iters_per_solve": it16 / (k16 * NSUB), "steps": k16}
        ps16.close()
    fs.close()
    if dist:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0

    cpu = None
    if rank == 0 and n_gpus == 1 and not a.no_cpu:
        cpu = cpu_leg(CPU_STEPS, CPU_WARMUP)
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}

    line = {"metric": "timestep MCell-updates/s (FP64, incl. Poisson solve)", "value": value,
            "unit": "MCell-updates/s", "n_gpus": n_gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"lid-driven cavity {n}^2 (scaling/LidDrivenFlow.cpp:16-42), NUM_SUBITER {NSUB}, "
                                   f"PCG tol {TOL} max_iter {MAX_ITER}, CFL {CFL}",
                       "pcg_iters_per_solve": k_bar, "pcg_last_residual": last_res, "mg_levels": levels[0],
                       "mg_depth": "uncapped (coarsest grid = 2 points)" if a.max_levels >= 40 else f"capped at {a.max_levels}",
                       "reference_depth_16": ref_depth, "same_problem_1gpu": anchor,
                       "decomposition": None if world == 1 else
                       f"{world} slabs of i-rows, " + ("peer-mapped transport (boundary rows pushed by the producing kernels on "
                                                     "the multigrid levels, mailbox kernels elsewhere, in-kernel scalar "
                                                     "all-reduce), NCCL for large reductions" if direct else
                                                     "NCCL halo exchange + all-reduce") + ", multigrid levels "
                       f"0..{levels[1] - 1} distributed, {levels[1]}..{levels[0] - 1} replicated", "l2": "inputs larger than L2 "
                       f"(each field {cells * 8 / 1e6:.0f} MB vs 126 MB L2)", "max_abs_U": umax},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches * world, "clocks": clk}
    print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    if len(sys.argv) >= 5 and sys.argv[1] == "--cpu-child":
        print(json.dumps(_cpu_leg_child(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]))), flush=True)
        sys.exit(0)
    sys.exit(main())
