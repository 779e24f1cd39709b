// fluidsolver_b200/csrc/p2p.cuh -- device side of the peer-mapped mailboxes (comm.cu) that kernels of
// other translation units use: the scalar all-reduce executed INSIDE the reduction kernel that produced
// the local sum (solver_pcg.cu, solver_cycle.cu), so a PCG dot product needs no collective launch.
//
// Mailbox layout (one device allocation per rank, mapped by every rank: cudaIpc between processes,
// plain pointers inside one process):
//   byte 0   : halo flag[0]  last epoch whose rows from rank - 1 are complete    (own 128-byte line)
//   byte 128 : halo flag[1]  ... from rank + 1
//   byte 192 : arrive counter of the local halo kernel
//   byte 200 : device copy of the error flag (kernels test this one: no PCIe read per exchange)
//   byte 208 : epoch of the scalar all-reduce (u64, advanced on the device: host-launched and in-kernel
//              reductions share it, and every rank runs the same sequence of them)
//   byte 256 : 4 halo slots of `cap` doubles: (side 0 | side 1) x (epoch parity 0 | 1)
//   red_off  : flag[P2P_MAX_RANKS] (u64), then values[2][P2P_MAX_RANKS][P2P_RED_MAX]
#pragma once

#include <stddef.h>

constexpr int P2P_MAX_RANKS = 16, P2P_RED_MAX = 16;

struct P2PRedDev {
  char* peer[P2P_MAX_RANKS];  // every rank's mailbox as mapped HERE (peer[rank] is mine)
  int rank, nranks;
  size_t red_off;
  int* err;                   // mapped pinned flag the host polls
  long long timeout;          // clock64 ticks a wait may last before it gives up
};

// ---- halo rows pushed by the kernel that produces them ("fused push", comm.cu / solver_cycle.cu) ----------
// A producer kernel of a distributed multigrid level stores every point of its h boundary rows twice: into its
// own array and, over NVLink, into the halo rows of the neighbouring rank's array (the same global row
// indices there).  Each CTA fences its remote stores and takes a ticket; the last one raises the epoch flag in
// the neighbours' mailboxes.  The kernels that read those halo rows wait for their own flags in their prologue.
// No exchange kernel, no unpacking: what is left of an exchange is the flag's flight time.
// Safe without double buffering: between two writes of the same halo rows lies a PCG all-reduce, which the
// writer can only pass after the reader's V-cycle -- every kernel that read the old rows -- has finished.
#define P2P_PUSH_LEVELS 40
struct HaloPush {
  int on, h;                      // on == 0: plain kernel
  long long i0, i1;               // rows this rank owns of the produced array (global indices)
  double* dst[2];                 // the array's virtual base in rank - 1's / rank + 1's address space (null: none)
  unsigned long long* raise[2];   // that neighbour's flag for rows arriving from me
  unsigned long long epoch;
  unsigned int* ticket;           // arrival counter (this rank's memory; wraps to 0)
};
struct HaloWait {
  int on;
  const unsigned long long* flag[2];  // my flags: rows from rank - 1 / rank + 1 (null: no such neighbour)
  unsigned long long epoch;
  int* err;                       // mapped pinned error flag
  char* mine;                     // my mailbox (device copy of the error flag)
  long long timeout;
};

#ifdef __CUDACC__
// consumer prologue (all threads of every CTA): the halo rows of this epoch have landed
__device__ __forceinline__ void halo_wait(const HaloWait& w) {
  if (!w.on) return;
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    const long long t0 = clock64();
    for (int sd = 0; sd < 2; ++sd) {
      if (!w.flag[sd]) continue;
      while (*reinterpret_cast<const volatile unsigned long long*>(w.flag[sd]) < w.epoch) {
        if (clock64() - t0 > w.timeout || *reinterpret_cast<volatile int*>(w.mine + 200)) {
          w.err[1] = -(int)w.epoch;
          *w.err   = 1;
          *reinterpret_cast<volatile int*>(w.mine + 200) = 1;
          break;
        }
      }
    }
    __threadfence_system();
  }
  __syncthreads();
}
// producer epilogue (all threads of every CTA, after their last store)
__device__ __forceinline__ void halo_push_done(const HaloPush& p) {
  if (!p.on) return;
  __threadfence_system();  // my remote stores are visible before my ticket is
  __syncthreads();
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    const unsigned nb = gridDim.x * gridDim.y;
    if (atomicInc(p.ticket, nb - 1) == nb - 1) {
      __threadfence_system();
      if (p.dst[0]) *reinterpret_cast<volatile unsigned long long*>(p.raise[0]) = p.epoch;
      if (p.dst[1]) *reinterpret_cast<volatile unsigned long long*>(p.raise[1]) = p.epoch;
    }
  }
}
// the neighbour's copy of row `row`, or null when the row is not one of the h rows next to that neighbour
__device__ __forceinline__ double* halo_push_row(const HaloPush& p, long long row) {
  if (!p.on) return nullptr;
  if (p.dst[0] && row < p.i0 + p.h) return p.dst[0];
  if (p.dst[1] && row >= p.i1 - p.h) return p.dst[1];
  return nullptr;
}

__device__ __forceinline__ volatile unsigned long long* p2p_red_flag(char* base, size_t off, int src) {
  return reinterpret_cast<volatile unsigned long long*>(base + off) + src;
}
__device__ __forceinline__ double* p2p_red_vals(char* base, size_t off, int par, int src) {
  return reinterpret_cast<double*>(base + off + P2P_MAX_RANKS * sizeof(unsigned long long)) +
         ((size_t)par * P2P_MAX_RANKS + src) * P2P_RED_MAX;
}
__device__ __forceinline__ unsigned long long* p2p_red_epoch(char* mine) {
  return reinterpret_cast<unsigned long long*>(mine + 208);
}
__device__ __forceinline__ volatile int* p2p_dev_err(char* mine) {
  return reinterpret_cast<volatile int*>(mine + 200);
}

// Sum of one double over all ranks, executed by ONE thread (the thread that holds the local sum: thread
// 0 of the last block of a reduction kernel).  All-to-all stores into the peers' value slots, a flag
// per (destination, source), then a fold in rank order -- the same order on every rank, so the result
// is bitwise identical everywhere (the PCG control flow relies on that).  Two value slots per source
// (epoch parity) suffice: a rank can only be one reduction ahead of a peer.
// (not inlined: it runs once per launch in one thread, and inlined it costs the hot kernels registers -- k_matvec_dot
// went from 32 to 40 and lost two resident CTAs per SM)
static __device__ __noinline__ double p2p_allreduce_sum1(const P2PRedDev& a, double v) {
  char* mine = a.peer[a.rank];
  if (*p2p_dev_err(mine)) return v;  // an earlier wait timed out: fail fast, the host reports it
  const unsigned long long e = ++*p2p_red_epoch(mine);
  const int par = (int)(e & 1ull);
  for (int r = 0; r < a.nranks; ++r) p2p_red_vals(a.peer[r], a.red_off, par, a.rank)[0] = v;
  __threadfence_system();
  for (int r = 0; r < a.nranks; ++r) *p2p_red_flag(a.peer[r], a.red_off, a.rank) = e;
  const long long t0 = clock64();
  double acc = 0.0;
  for (int r = 0; r < a.nranks; ++r) {
    while (*p2p_red_flag(mine, a.red_off, r) < e) {
      if (clock64() - t0 > a.timeout) {
        *a.err = 1;
        *p2p_dev_err(mine) = 1;
        __threadfence_system();
        return v;
      }
    }
    __threadfence_system();
    const double x = __ldcg(p2p_red_vals(mine, a.red_off, par, r));
    acc = r == 0 ? x : acc + x;
  }
  return acc;
}
#endif
