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

#ifdef __CUDACC__
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
