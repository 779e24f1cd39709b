"""Worker of tests/test_slab_cpu.py: one rank of a world_size-N gloo group (CPU only).

Exercises the HOST side of the slab decomposition exactly as bench.py / a driver would under
torchrun, minus the device: rank bootstrap (communicator id drawn by rank 0 and broadcast), the
library's partition rules (fsgpu_slab_range / fsgpu_slab_coarse_range through the C ABI), the
halo-exchange protocol of fluidsolver_b200/csrc/comm.cu restated with gloo send/recv on numpy rows
(send my first/last h owned rows, receive into the h rows above/below), row-sharded host buffers
(slab_rows) and the all-reduce of a partial sum.  Exit code 0 = all checks passed on this rank.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fluidsolver_b200 as F  # noqa: E402

H = 2  # FS_RO: halo rows of the device layout


def halo_exchange(local, lo, hi, row0, rank, world):
    """comm_halo's protocol: rows are addressed by GLOBAL index, local[k] is global row row0 + k"""
    reqs = []
    up, dn = rank - 1, rank + 1
    bufs = []
    if up >= 0:
        reqs.append(dist.isend(torch.from_numpy(local[lo - row0:lo - row0 + H].copy()), up))
        b = torch.empty((H, local.shape[1]), dtype=torch.float64)
        reqs.append(dist.irecv(b, up))
        bufs.append((lo - H - row0, b))
    if dn < world:
        reqs.append(dist.isend(torch.from_numpy(local[hi - H - row0:hi - row0].copy()), dn))
        b = torch.empty((H, local.shape[1]), dtype=torch.float64)
        reqs.append(dist.irecv(b, dn))
        bufs.append((hi - row0, b))
    for r in reqs:
        r.wait()
    for k, b in bufs:
        local[k:k + H] = b.numpy()


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    nx, ny = 61, 23
    # 1. bootstrap: the 128-byte communicator id travels from rank 0 (random bytes stand in for
    #    ncclGetUniqueId, which needs no GPU either but opens sockets)
    box = [os.urandom(F.COMM_ID_BYTES) if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    ids = [None] * world
    dist.all_gather_object(ids, box[0])
    assert all(i == ids[0] for i in ids) and len(ids[0]) == F.COMM_ID_BYTES
    # 2. partition: every rank computes every rank's range and they tile [0, nx + 2)
    mine = F.slab_range(nx, rank, world)
    allr = [None] * world
    dist.all_gather_object(allr, mine)
    assert allr == [F.slab_range(nx, r, world) for r in range(world)]
    assert allr[0][0] == 0 and allr[-1][1] == nx + 2
    assert all(allr[r][1] == allr[r + 1][0] for r in range(world - 1))
    lo, hi = mine
    # 3. row-sharded host buffers + halo exchange + stencil on owned rows == global stencil
    rng = np.random.default_rng(7)
    G = rng.uniform(-1, 1, (nx + 2, ny + 2))       # the same global field on every rank
    r0, r1 = F.slab_rows("P", nx, ny, rank, world)
    assert r0 == max(lo - H, 0) and r1 == min(hi + H, nx + 2)
    local = np.zeros((hi - lo + 2 * H, ny + 2))    # rows lo - H .. hi + H (pads zero at the walls)
    row0 = lo - H
    local[lo - row0:hi - row0] = G[lo:hi]          # owned rows only; halos come from the exchange
    halo_exchange(local, lo, hi, row0, rank, world)
    Gp = np.zeros((nx + 2 + 2 * H, ny + 2)); Gp[H:-H] = G
    assert np.array_equal(local, Gp[lo:hi + 2 * H]), "halo rows differ from the neighbour's rows"
    # 5-point stencil reaching two rows out (restriction of a residual), owned rows only
    def wide(a, i):
        return a[i - 2] - 4 * a[i - 1] + 6 * a[i] - 4 * a[i + 1] + a[i + 2]
    own = np.stack([wide(local, i - row0) for i in range(lo, hi)])
    ref = np.stack([wide(Gp, i + H) for i in range(lo, hi)])
    assert np.array_equal(own, ref)
    # 4. coarse-level ranges follow the fine partition and tile each coarse level
    n0, lo_l, hi_l, box_lo = nx + 2, lo, hi, -1
    for level in range(5):
        clo_box = -(-box_lo // 2)                 # ceil: PFMG keeps the even global indices
        foff = 2 * clo_box - box_lo
        n0c = (n0 - foff + 1) // 2
        c = F.slab_coarse_range(lo_l, hi_l, foff)
        allc = [None] * world
        dist.all_gather_object(allc, c)
        assert allc[0][0] == 0 and allc[-1][1] == n0c, (level, allc, n0c)
        assert all(allc[r][1] == allc[r + 1][0] for r in range(world - 1))
        for ci in range(c[0], c[1]):               # every coarse row sits on a fine row I own
            assert lo_l <= 2 * ci + foff < hi_l
        n0, lo_l, hi_l, box_lo = n0c, c[0], c[1], clo_box
    # 5. a global sum from per-slab partial sums (the PCG dot products, rhs mean, sum delta_p)
    part = torch.tensor([G[lo:hi].sum()], dtype=torch.float64)
    dist.all_reduce(part, op=dist.ReduceOp.SUM)
    assert abs(part.item() - G.sum()) <= 1e-12 * np.abs(G).sum()
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank}/{world} ok")


if __name__ == "__main__":
    main()
