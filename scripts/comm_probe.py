"""Per-operation cost of the slab transport, for diagnosing multi-GPU scaling (run under torchrun):

  NCCL_DEBUG=INFO python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29555 scripts/comm_probe.py [grid] 2> nccl.log      (grep -E 'via P2P|via SHM|NET/' nccl.log)
  FSGPU_HALO=p2p ... the same with the mailbox halo exchange

Prints, on rank 0: us per halo exchange (2 rows of one field, stream-ordered, no host sync in the loop),
us per scalar all-reduce round trip (kernel + all-reduce + host read), and one timestep's per-class times.
Round 1 measured 710 ms per 16384^2 step on 8 GPUs with 1.7 % of it inside the dominant kernel
(profiles/bench_r01h_n8_16384.json): ~20 ms per PCG iteration for ~31 stream-ordered NCCL calls."""
import os
import sys

os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")  # see bench.py
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import fluidsolver_b200 as F  # noqa: E402

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
cid = None
if world > 1:
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


def timed(fn, reps):
    fn()
    fs.sync()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    fs.sync()
    return (time.perf_counter() - t0) / reps * 1e6


halo = timed(lambda: fs.apply_neumann_bconds("P"), 300)   # 2 edge kernels + one 2-row halo exchange
red = timed(lambda: fs.abs_max("P"), 300)                 # reduction kernel + all-reduce + host read
ps = F.LinearSolver(fs, tol=1e-6, max_iter=50, max_levels=int(os.environ.get("MAX_LEVELS", "40")))
ps.set_pressure_operator()
for _ in range(2):
    fs.step(ps, fs.adjust_dt(0.9, 0.1), 5)
fs.sync()
t0 = time.perf_counter()
its = fs.step(ps, fs.adjust_dt(0.9, 0.1), 5)[1]
fs.sync()
step_ms = (time.perf_counter() - t0) * 1e3
if rank == 0:
    print(f"world {world} grid {n}^2 FSGPU_HALO={os.environ.get('FSGPU_HALO', 'default')} "
          f"levels {ps.num_levels} (distributed {ps.num_distributed_levels})")
    print(f"  halo exchange (apply_neumann_bconds: 2 kernels + exchange): {halo:8.1f} us")
    print(f"  scalar all-reduce round trip (abs_max):                    {red:8.1f} us")
    print(f"  one timestep: {step_ms:.1f} ms, {its} PCG iterations -> {step_ms / max(its, 1):.2f} ms per iteration")
ps.close()
fs.close()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
