"""Per-kernel-class device time of the bench workload (CUDA events on the library's stream).
usage: python scripts/prof_classes.py [n] [steps]   -> table: class, launches, total ms, ms/step, share
Under torchrun (N ranks, slab decomposition; FSGPU_HALO as for bench.py) every class interval includes the halo
exchanges / all-reduces enqueued inside its scope; ranks 0 and 1 print their tables."""
import os
import sys

os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")  # see bench.py

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fluidsolver_b200 as F  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
ml = int(os.environ.get("MAX_LEVELS", "40"))
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
cid = None
if world > 1:
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    box = [F.comm_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    cid = box[0]
fs = F.FS(n, n, 1.0 / n, 1.0 / n, 1.0, 1.0, 1e-3, 1e-3, device=local, rank=rank, nranks=world, comm_id=cid)
fs.calc_rho_visc_const()
for s, u in enumerate((0.0, 0.0, 0.0, 1.0)):
    fs.set_bconds(s, F.DIRICHLET, u, 0.0)
fs.apply_velocity_bconds()
ps = F.LinearSolver(fs, tol=1e-6, max_iter=50, max_levels=ml)
ps.set_pressure_operator()
for _ in range(2):
    fs.step(ps, fs.adjust_dt(0.9, 0.1), 5)
# one class at a time: the event pairs of different classes would otherwise perturb each other
import torch  # noqa: E402
tot = {}
for cls in F.PROF_CLASSES:
    F.prof_enable((cls,))
    F.prof_read()
    its = 0
    for _ in range(steps):
        its += fs.step(ps, fs.adjust_dt(0.9, 0.1), 5)[1]
    r = F.prof_read()
    if cls in r:
        tot[cls] = r[cls]
F.prof_enable(())
st = torch.cuda.ExternalStream(fs.stream(), device=torch.device("cuda", local))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
fs.sync()
e0.record(st)
for _ in range(steps):
    fs.step(ps, fs.adjust_dt(0.9, 0.1), 5)
e1.record(st)
fs.sync()
ms = e0.elapsed_time(e1) / steps
if world > 1:
    dist.barrier()
    import time
    time.sleep(0.3 * rank)  # tables one after the other
if rank > 1:
    sys.exit(0)
print(f"rank {rank} of {world} FSGPU_HALO={os.environ.get('FSGPU_HALO', 'default')} levels distributed {ps.num_distributed_levels}")
print(f"n {n} steps {steps} levels {ps.num_levels} its/solve {its / (steps * 5):.1f} step {ms:.2f} ms")
s = 0.0
for cls, (t, c) in tot.items():
    print(f"{cls:22s} {c // steps:6d} launches/step {t / steps:9.3f} ms/step {t / steps / ms * 100:6.1f} %  {t / c * 1e3:9.1f} us/launch")
    s += t / steps
print(f"{'sum of classes':22s} {s:9.3f} ms/step ({s / ms * 100:.1f} % of the step)")
