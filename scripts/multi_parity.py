"""Slab decomposition against the single-GPU run of the same problem, on real GPUs (under torchrun):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29650 \
      scripts/multi_parity.py [grid] [steps]

Every rank steps its slab of the lid-driven cavity (the library's default transport, or FSGPU_HALO / FSGPU_PUSH as
set) and, on its own GPU, the whole grid in a single context; it then compares the rows it owns.  Prints the
relative L2 difference of (U, V) and p over all ranks, the iteration counts and dt of both runs.  The 1-GPU path is
pinned to the oracle / the unmodified reference by tests/; this ties the N-GPU path to it at a size and rank count the
1-GPU test box cannot hold."""
import os
import sys

os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")  # see bench.py

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import fluidsolver_b200 as F  # noqa: E402

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
tol = float(os.environ.get("PARITY_TOL", "1e-12"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
box = [F.comm_id() if rank == 0 else None]
dist.broadcast_object_list(box, src=0)


def run(fs):
    fs.calc_rho_visc_const()
    for s, u in enumerate((0.0, 0.0, 0.0, 1.0)):
        fs.set_bconds(s, F.DIRICHLET, u, 0.0)
    fs.apply_velocity_bconds()
    ps = F.LinearSolver(fs, tol=tol, max_iter=200, max_levels=40)
    ps.set_pressure_operator()
    hist = []
    for _ in range(steps):
        dt = fs.adjust_dt(0.9, 0.1)
        res, it = fs.step(ps, dt, 5)
        hist.append((dt, it, res))
    info = (ps.num_levels, ps.num_distributed_levels, fs.halo_is_direct())
    out = {}
    for nm in ("U", "V", "P"):
        r0, r1 = fs.local_rows(nm)
        a = np.empty((r1 - r0, F.field_shape(nm, n, n)[1]))
        fs.download_rows(nm, a, r0)
        out[nm] = (r0, a)
    ps.close()
    return hist, info, out


slab = F.FS(n, n, 1.0 / n, 1.0 / n, 1.0, 1.0, 1e-3, 1e-3, device=local, rank=rank, nranks=world, comm_id=box[0])
h_s, info, o_s = run(slab)
a_lo, a_hi = F.slab_range(n, rank, world)
slab.close()
one = F.FS(n, n, 1.0 / n, 1.0 / n, 1.0, 1.0, 1e-3, 1e-3, device=local)
h_1, _, o_1 = run(one)
one.close()
num = np.zeros(4)
for k, nm in enumerate(("U", "V", "P")):
    r0, a = o_s[nm]
    q0, b = o_1[nm]
    b = b[r0 - q0:r0 - q0 + a.shape[0]]
    own = slice(a_lo - r0, a_lo - r0 + (a_hi - a_lo))  # the rows this rank owns (halo rows excluded)
    a, b = a[own], b[own]
    if nm == "P":
        num[2] += float(((a - b) ** 2).sum()); num[3] += float((b ** 2).sum())
    else:
        num[0] += float(((a - b) ** 2).sum()); num[1] += float((b ** 2).sum())
t = torch.tensor(num, dtype=torch.float64, device=f"cuda:{local}")
dist.all_reduce(t)
num = t.cpu().numpy()
if rank == 0:
    print(f"world {world} grid {n}^2 steps {steps} tol {tol}: levels {info[0]} (distributed {info[1]}), peer-mapped transport {info[2]}, "
          f"FSGPU_HALO={os.environ.get('FSGPU_HALO', 'default')} FSGPU_PUSH={os.environ.get('FSGPU_PUSH', 'default')}")
    print("  slab run   (dt, PCG iterations, residual) per step:", [(f"{d:.6e}", i, f"{r:.1e}") for d, i, r in h_s])
    print("  1-GPU run  (dt, PCG iterations, residual) per step:", [(f"{d:.6e}", i, f"{r:.1e}") for d, i, r in h_1])
    print(f"  rel L2 difference over all ranks: velocity {np.sqrt(num[0] / max(num[1], 1e-300)):.3e}, "
          f"pressure (un-shifted) {np.sqrt(num[2] / max(num[3], 1e-300)):.3e}")
dist.barrier()
dist.destroy_process_group()
