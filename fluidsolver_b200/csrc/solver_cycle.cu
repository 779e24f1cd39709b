// fluidsolver_b200/csrc/solver_cycle.cu -- the multigrid V-cycle used as CG preconditioner.
//
// Follows hypre_PFMGSolve with max_iter 1 and a zero initial guess
// (Submodules/hypre/src/struct_ls/pfmg_solve.c:129-454) level by level, but each level is TWO
// fused kernels instead of HYPRE's five passes (relax, matvec, axpy, restrict / interpolate, relax):
//
//   k_down<CDIR,ACTIVE>  pre-relaxation from a zero guess (x = w b / d, struct_ls/jacobi.c:142)
//                        + residual r = b - A x + restriction b_c = P^T r, without ever storing
//                        x or r: x is recomputed from b where the post-relaxation needs it
//   k_up<CDIR,ACTIVE>    x = (pre-relaxed x) + P x_c   (interpolation, pfmg_solve.c:353-355)
//                        + post-relaxation x += w (b - A x) / d
//   k_coarse             all sweeps on the coarsest grid inside one CTA (shared memory)
//
// Pads of every array are zero, so stencils read past the grid edge without branches.
#include "solver.h"

#include <stdlib.h>

#include <algorithm>
#include <utility>

// resident CTAs per SM the marching kernels are compiled for (register cap = 65536 / (128 * MINB))
#ifndef FS_MARCH_MINB
#define FS_MARCH_MINB 8
#endif
// k_up_march operand policy: 0 load dinv and pc; 1 recompute both (two divisions per point);
// 2 load dinv, recompute pc (half a division per point on an i-coarsened level)
#ifndef FS_UP_MODE
#define FS_UP_MODE 1
#endif
#ifndef FS_UP_PREFETCH
#define FS_UP_PREFETCH 0
#endif
// marching kernels: rows ahead whose operands are pulled into L2 with prefetch.global.L2 (0 = off).
// The kernels are latency bound (ncu: long-scoreboard stalls, DRAM 61 % busy, 30 warps/SM at 64
// registers); an L2 prefetch adds memory-level parallelism without costing a register.  A/B variant
// for scripts/build_variants.py (-DFS_PF_ROWS=4), not yet timed.
#ifndef FS_PF_ROWS
#define FS_PF_ROWS 0
#endif
// k_up_march: operand rows fetched FS_UP_ASYNC rows ahead with cp.async (LDGSTS) into a thread-private
// shared-memory ring instead of into registers (0 = off).  Every thread copies and later reads only
// its own words, so no barrier is needed -- cp.async.wait_group alone orders the ring; the depth
// is bought with shared memory (7 KB per row ahead and CTA) instead of registers.  A/B variant for
// scripts/build_variants.py (-DFS_UP_ASYNC=4), NOT yet run on hardware.
#ifndef FS_UP_ASYNC
#define FS_UP_ASYNC 0
#endif
// the same ring for the five operand words per row of k_down_march0 / k_down_march1 on ACTIVE levels
#ifndef FS_DOWN_ASYNC
#define FS_DOWN_ASYNC 0
#endif
#if FS_DOWN_ASYNC > 6
#error "FS_DOWN_ASYNC: at most 6 rows ahead"
#endif
#if FS_UP_ASYNC > 0 && (FS_UP_MODE != 1 || FS_UP_PREFETCH != 0 || FS_UP_ASYNC > 6)
#error "FS_UP_ASYNC needs FS_UP_MODE == 1, FS_UP_PREFETCH == 0 and a depth of at most 6 rows (48 KB static shared memory)"
#endif

namespace {

constexpr int BX = 64, BY = 4;

// one lane in eight touches the line (a warp row is 256 B = 2-3 lines of 128 B)
__device__ __forceinline__ void pf_l2(const double* p, int lane) {
#if FS_PF_ROWS > 0
  if ((lane & 7) == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#else
  (void)p; (void)lane;
#endif
}

// 8-byte asynchronous global -> shared copy (LDGSTS) and the two group operations of the variants above
__device__ __forceinline__ void cp_async8(double* dst, const double* src) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// 1 / v with the convention of the stored dinv / pc arrays (0 where v == 0: pads, halo-less walls,
// a singular coarsest point).  The active-level kernels recompute dinv = 1/d and the interpolation
// denominator pc = 1/(d + t_m + t_p) from d, cx, cy instead of loading them: two FP64 divisions
// per point buy 16 B/point less HBM traffic, and the values are bit-identical to the stored ones
// (same operands, same association as k_interp / k_rap, correctly rounded division).
__device__ __forceinline__ double rcp0(double v) { return v != 0.0 ? 1.0 / v : 0.0; }

template <int CDIR>
struct Coef {
  const double *cx, *cy;
  int64_t P;
  __device__ __forceinline__ double am(int64_t q) const { return CDIR == 0 ? __ldg(cx + q) : __ldg(cy + q); }
  __device__ __forceinline__ double ap(int64_t q) const { return CDIR == 0 ? __ldg(cx + q + P) : __ldg(cy + q + 1); }
};

// ---- marching restriction kernels (same strip / chunk scheme as k_up_march below) -------------
// k_down_march0: coarsening along i.  Per fine row the warp forms X = w b / d, then the residual
// R of the row above from three rows of X held in registers (left/right by shuffle), and on every
// second row the coarse rhs  b_c = R + pp(row-1) R(row-1) + pm(row+1) R(row+1).
template <bool ACTIVE>
__global__ void __launch_bounds__(128, FS_MARCH_MINB)
k_down_march0(const LevelDev f, const double* __restrict__ fb, const LevelDev c,
              double* __restrict__ cb, double w, int mrc, const PcgScal* scal) {
  if (scal->done) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n1 = (int)f.n1, c_lo = (int)c.i0, c_hi = (int)c.i1, P = (int)f.pitch, Pc = (int)c.pitch,
            foff = (int)c.foff;
  const int strips = (n1 + 119) / 120;
  const int total  = strips * ((c_hi - c_lo + mrc - 1) / mrc);
  for (int t = blockIdx.x; t < total; t += gridDim.x) {
    const int ch = t / strips, sp = t - ch * strips;
    int lj = sp * 120 + warp * 30 - 1 + lane;
    const bool store_lane = lane >= 1 && lane <= 30 && lj < n1;
    if (lj > n1) lj = n1;
    const int c0 = c_lo + ch * mrc, c1 = min(c0 + mrc, c_hi);
    const int f0 = 2 * c0 + foff, fL = 2 * (c1 - 1) + foff;
    double X2 = 0.0, X1 = 0.0, b1 = 0.0, cx2 = 0.0, cx1 = 0.0, cy1 = 0.0, cy1p = 0.0, Ra = 0.0, Rb = 0.0;
    double pc1 = 0.0, pc2 = 0.0, pc3 = 0.0;  // ACTIVE: pc of rows li - 1, li - 2, li - 3
    int li = f0 - 2;
    int o  = (li + FS_RO) * P + lj + FS_CO;
#if FS_DOWN_ASYNC > 0
    __shared__ double ring[FS_DOWN_ASYNC][5][128];
    const int tx = threadIdx.x;
    auto issue_row = [&](int o_, int sl) {
      cp_async8(&ring[sl][0][tx], fb + o_);
      cp_async8(&ring[sl][1][tx], f.cx + o_);
      cp_async8(&ring[sl][2][tx], f.d + o_);
      cp_async8(&ring[sl][3][tx], f.cy + o_);
      cp_async8(&ring[sl][4][tx], f.cy + o_ + 1);
    };
    int slot = 0;
    if (ACTIVE) {
#pragma unroll
      for (int dd = 0; dd < FS_DOWN_ASYNC; ++dd) {
        if (li + dd <= fL + 2) issue_row(o + dd * P, dd);
        cp_async_commit();
      }
    }
#endif
#pragma unroll 2
    for (; li <= fL + 2; ++li, o += P) {
#if FS_PF_ROWS > 0
      if (li + FS_PF_ROWS <= fL + 2) {
        const int op = o + FS_PF_ROWS * P;
        pf_l2(fb + op, lane); pf_l2(f.cx + op, lane);
        if (ACTIVE) { pf_l2(f.d + op, lane); pf_l2(f.cy + op, lane); }
      }
#endif
#if FS_DOWN_ASYNC > 0
      double b0, cx0, X0 = 0.0, cy0 = 0.0, cy0p = 0.0, pc0 = 0.0;
      if (ACTIVE) {
        cp_async_wait<FS_DOWN_ASYNC - 1>();  // row li has landed in its slot
        b0   = ring[slot][0][tx];
        cx0  = ring[slot][1][tx];
        const double d0 = ring[slot][2][tx];
        cy0  = ring[slot][3][tx];
        cy0p = ring[slot][4][tx];
        X0   = w * b0 * rcp0(d0);
        pc0  = rcp0(d0 + cy0 + cy0p);
      } else {
        b0  = __ldg(fb + o);
        cx0 = __ldg(f.cx + o);
      }
#else
      const double b0 = __ldg(fb + o), cx0 = __ldg(f.cx + o);
      double X0 = 0.0, cy0 = 0.0, cy0p = 0.0, pc0 = 0.0;
      if (ACTIVE) {
        const double d0 = __ldg(f.d + o);
        cy0  = __ldg(f.cy + o);
        cy0p = __ldg(f.cy + o + 1);
        X0   = w * b0 * rcp0(d0);
        pc0  = rcp0(d0 + cy0 + cy0p);
      }
#endif
      // residual of row li - 1
      double Rn;
      if (ACTIVE) {
        const double xl = __shfl_up_sync(0xffffffffu, X1, 1), xr = __shfl_down_sync(0xffffffffu, X1, 1);
        double v = w * b1;
        v += cx1 * X2;
        v += cx0 * X0;
        v += cy1 * xl;
        v += cy1p * xr;
        Rn = b1 - v;
      } else {
        Rn = b1;
      }
      // coarse point on row li - 2 (a C row when li - f0 is even)
      if (li >= f0 + 2 && ((li - f0) & 1) == 0) {
        const double pcm = ACTIVE ? pc3 : __ldg(f.pc + o - 3 * P);
        const double pcp = ACTIVE ? pc1 : __ldg(f.pc + o - P);
        double v = Rb;
        v += (-cx2 * pcm) * Ra;
        v += (-cx1 * pcp) * Rn;
        if (store_lane) cb[(((li - 2 - foff) >> 1) + FS_RO) * Pc + lj + FS_CO] = v;
      }
      X2 = X1; X1 = X0; b1 = b0; cx2 = cx1; cx1 = cx0; cy1 = cy0; cy1p = cy0p; Ra = Rb; Rb = Rn;
      pc3 = pc2; pc2 = pc1; pc1 = pc0;
#if FS_DOWN_ASYNC > 0
      if (ACTIVE) {  // refill the slot just consumed
        if (li + FS_DOWN_ASYNC <= fL + 2) issue_row(o + FS_DOWN_ASYNC * P, slot);
        cp_async_commit();
        slot = slot + 1 == FS_DOWN_ASYNC ? 0 : slot + 1;
      }
#endif
    }
  }
}

// k_down_march1: coarsening along j.  Lanes 1..30 hold a valid residual (X of the neighbours by
// shuffle), lanes 2..29 that sit on a C column emit  b_c = R + pp(j-1) R(j-1) + pm(j+1) R(j+1).
template <bool ACTIVE>
__global__ void __launch_bounds__(128, FS_MARCH_MINB)
k_down_march1(const LevelDev f, const double* __restrict__ fb, const LevelDev c,
              double* __restrict__ cb, double w, int mr, const PcgScal* scal) {
  if (scal->done) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int f_lo = (int)f.i0, f_hi = (int)f.i1, n1 = (int)f.n1, P = (int)f.pitch, Pc = (int)c.pitch,
            foff = (int)c.foff;
  const int strips = (n1 + 111) / 112;
  const int total  = strips * ((f_hi - f_lo + mr - 1) / mr);
  for (int t = blockIdx.x; t < total; t += gridDim.x) {
    const int ch = t / strips, sp = t - ch * strips;
    int lj = sp * 112 + warp * 28 - 2 + lane;
    const int pos = lj - foff;
    const bool store_lane = lane >= 2 && lane <= 29 && lj < n1 && (pos & 1) == 0;
    const int oc_col = (pos >> 1) + FS_CO;
    if (lj > n1) lj = n1;
    const int r0 = f_lo + ch * mr, r1 = min(r0 + mr, f_hi);
    double X2 = 0.0, X1 = 0.0, b1 = 0.0, cx1 = 0.0, cy1 = 0.0, cy1p = 0.0, d1 = 0.0;
    int li = r0 - 1;
    int o  = (li + FS_RO) * P + lj + FS_CO;
#if FS_DOWN_ASYNC > 0
    __shared__ double ring[FS_DOWN_ASYNC][5][128];
    const int tx = threadIdx.x;
    auto issue_row = [&](int o_, int sl) {
      cp_async8(&ring[sl][0][tx], fb + o_);
      cp_async8(&ring[sl][1][tx], f.cy + o_);
      cp_async8(&ring[sl][2][tx], f.cy + o_ + 1);
      cp_async8(&ring[sl][3][tx], f.d + o_);
      cp_async8(&ring[sl][4][tx], f.cx + o_);
    };
    int slot = 0;
    if (ACTIVE) {
#pragma unroll
      for (int dd = 0; dd < FS_DOWN_ASYNC; ++dd) {
        if (li + dd <= r1) issue_row(o + dd * P, dd);
        cp_async_commit();
      }
    }
#endif
#pragma unroll 2
    for (; li <= r1; ++li, o += P) {
#if FS_PF_ROWS > 0
      if (li + FS_PF_ROWS <= r1) {
        const int op = o + FS_PF_ROWS * P;
        pf_l2(fb + op, lane); pf_l2(f.cy + op, lane);
        if (ACTIVE) { pf_l2(f.d + op, lane); pf_l2(f.cx + op, lane); }
      }
#endif
#if FS_DOWN_ASYNC > 0
      double b0, cy0, cy0p, X0 = 0.0, cx0 = 0.0, d0 = 0.0;
      if (ACTIVE) {
        cp_async_wait<FS_DOWN_ASYNC - 1>();  // row li has landed in its slot
        b0   = ring[slot][0][tx];
        cy0  = ring[slot][1][tx];
        cy0p = ring[slot][2][tx];
        d0   = ring[slot][3][tx];
        cx0  = ring[slot][4][tx];
        X0   = w * b0 * rcp0(d0);
      } else {
        b0   = __ldg(fb + o);
        cy0  = __ldg(f.cy + o);
        cy0p = __ldg(f.cy + o + 1);
      }
#else
      const double b0 = __ldg(fb + o);
      const double cy0 = __ldg(f.cy + o), cy0p = __ldg(f.cy + o + 1);
      double X0 = 0.0, cx0 = 0.0, d0 = 0.0;
      if (ACTIVE) {
        d0  = __ldg(f.d + o);
        X0  = w * b0 * rcp0(d0);
        cx0 = __ldg(f.cx + o);
      }
#endif
      if (li > r0) {  // row li - 1
        double Rn;
        if (ACTIVE) {
          const double xl = __shfl_up_sync(0xffffffffu, X1, 1), xr = __shfl_down_sync(0xffffffffu, X1, 1);
          double v = w * b1;
          v += cx1 * X2;
          v += cx0 * X0;
          v += cy1 * xl;
          v += cy1p * xr;
          Rn = b1 - v;
        } else {
          Rn = b1;
        }
        const double pcv = ACTIVE ? rcp0(d1 + cx1 + cx0) : __ldg(f.pc + o - P);
        const double pcl = __shfl_up_sync(0xffffffffu, pcv, 1), pcr = __shfl_down_sync(0xffffffffu, pcv, 1);
        const double Rl = __shfl_up_sync(0xffffffffu, Rn, 1), Rr = __shfl_down_sync(0xffffffffu, Rn, 1);
        double v = Rn;
        v += (-cy1 * pcl) * Rl;
        v += (-cy1p * pcr) * Rr;
        if (store_lane) cb[(li - 1 + FS_RO) * Pc + oc_col] = v;
      }
      X2 = X1; X1 = X0; b1 = b0; cx1 = cx0; cy1 = cy0; cy1p = cy0p; d1 = d0;
#if FS_DOWN_ASYNC > 0
      if (ACTIVE) {  // refill the slot just consumed
        if (li + FS_DOWN_ASYNC <= r1) issue_row(o + FS_DOWN_ASYNC * P, slot);
        cp_async_commit();
        slot = slot + 1 == FS_DOWN_ASYNC ? 0 : slot + 1;
      }
#endif
    }
  }
}

// value at fine point (li, lj) after [zero-guess relaxation] + interpolation of the coarse x
// (bq = rhs at the point, only used when ACTIVE)
template <int CDIR, bool ACTIVE>
__device__ __forceinline__ double x_interp(const LevelDev& f, double bq, const LevelDev& c,
                                           const double* __restrict__ cx_, int64_t li, int64_t lj,
                                           int64_t q, double w) {
  const Coef<CDIR> C{f.cx, f.cy, f.pitch};
  double x = ACTIVE ? w * bq * __ldg(f.dinv + q) : 0.0;
  const int64_t pos = (CDIR == 0 ? li : lj) - c.foff;  // even <=> C-point
  const int64_t csd = CDIR == 0 ? c.pitch : 1;
  if ((pos & 1) == 0) {
    const int64_t cp = pos >> 1;  // arithmetic shift: pos = -2 -> -1 (pad)
    x += __ldg(cx_ + (CDIR == 0 ? c.idx(cp, lj) : c.idx(li, cp)));
  } else {
    const int64_t cm = (pos - 1) >> 1;
    const int64_t kc = CDIR == 0 ? c.idx(cm, lj) : c.idx(li, cm);
    const double pc  = __ldg(f.pc + q);
    double v = (-C.am(q) * pc) * __ldg(cx_ + kc);
    v += (-C.ap(q) * pc) * __ldg(cx_ + kc + csd);
    x += v;
  }
  return x;
}

// inactive level: interpolation only (no relaxation on this level); persistent tile loop, so the
// copies enqueued past PCG convergence cost a few us instead of an empty 1e5-CTA grid
template <int CDIR>
__global__ void __launch_bounds__(BX* BY)
k_up_interp(const LevelDev f, double* __restrict__ fx, const LevelDev c,
            const double* __restrict__ cxv, const PcgScal* scal) {
  if (scal->done) return;
  for_each_point(f.i0, f.i1, f.n1, [&](int64_t li, int64_t lj) {
    const int64_t k = f.idx(li, lj);
    fx[k] = x_interp<CDIR, false>(f, 0.0, c, cxv, li, lj, k, 0.0);
  });
}

// active level, "marching" formulation.  A warp owns a strip of 32 consecutive columns (30 of
// them produce output, lanes 0 and 31 are the left/right halo) and walks down MR rows.  Per row
// it forms  x0 = w b / d + P x_c  ONCE per point, keeps x0 of three consecutive rows in
// registers, takes the left/right neighbours by warp shuffle, and emits
// x = x0 + w (b - A x0) / d  for the row above.  No shared memory, no block barrier, 32-bit
// offsets advanced by one pitch per row; every array is read once (plus the 2/32 column and
// 2/MR row overlap).  CTAs loop over (strip, chunk) tasks, so the optional dot product <b, x>
// (= <r, z> of PCG on level 0) ends in ~1e3 partials.
// DOT: 0 none, 1 PCG iteration (convergence test + beta), 2 first V-cycle (bi).
// FS_MWARPS: warps (= adjacent 30-column strips) per CTA of k_up_march; A/B variant (-DFS_MWARPS=8 keeps the
// resident warps per SM: the CTA count per SM scales down with it)
#ifndef FS_MWARPS
#define FS_MWARPS 4
#endif
constexpr int MOUT = 30, MWARPS = FS_MWARPS;

template <int CDIR, int DOT>
__global__ void __launch_bounds__(32 * MWARPS, FS_MARCH_MINB * 4 / MWARPS)
k_up_march(const LevelDev f, const double* __restrict__ fb, double* __restrict__ fx, const LevelDev c,
           const double* __restrict__ xc, double w, int mr, PcgScal* scal, double* part,
           unsigned int* ticket) {
  if (scal->done) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int f_lo = (int)f.i0, f_hi = (int)f.i1, n1 = (int)f.n1, P = (int)f.pitch, Pc = (int)c.pitch,
            foff = (int)c.foff;
  const int strips = (n1 + MOUT * MWARPS - 1) / (MOUT * MWARPS);
  const int total  = strips * ((f_hi - f_lo + mr - 1) / mr);
  double acc[1] = {0.0};
#if FS_UP_ASYNC > 0
  // [row slot][operand][thread]: a warp's lanes touch 32 consecutive words of a bank row
  __shared__ double ring[FS_UP_ASYNC][7][32 * MWARPS];
  auto cpa = [](double* dst, const double* src) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(src) : "memory");
  };
#endif
  for (int t = blockIdx.x; t < total; t += gridDim.x) {
    const int ch = t / strips, sp = t - ch * strips;
    int lj = sp * (MOUT * MWARPS) + warp * MOUT - 1 + lane;  // lane 0 sits one column left
    const bool store_lane = lane >= 1 && lane <= MOUT && lj < n1;
    if (lj > n1) lj = n1;  // the zero pad column right of the grid
    const int r0 = f_lo + ch * mr, r1 = min(r0 + mr, f_hi);
    double x0p = 0.0, x0c = 0.0, bc = 0.0, dic = 0.0, dc = 0.0, cyc = 0.0, cy1c = 0.0, cxc = 0.0;
    int li = r0 - 1;
    int o  = (li + FS_RO) * P + lj + FS_CO;
    double cxn = __ldg(f.cx + o);
    // operands of one row, loaded one row AHEAD of their use (software pipelining: the loads of
    // row li + 1 are in flight while row li is computed, independent of what the compiler hoists)
    struct Row { double b, d, di, cy, cy1, cxnn, xm, xp, pc; };
    auto load_row = [&](int li_, int o_) {
      Row q;
      q.cxnn = __ldg(f.cx + o_ + P);
      q.b    = __ldg(fb + o_);
      q.cy   = __ldg(f.cy + o_);
      q.cy1  = __ldg(f.cy + o_ + 1);
#if FS_UP_MODE == 1
      q.d = __ldg(f.d + o_);  q.di = 0.0;  q.pc = 0.0;
#elif FS_UP_MODE == 2
      q.d = __ldg(f.d + o_);  q.di = __ldg(f.dinv + o_);  q.pc = 0.0;
#else
      q.d = __ldg(f.d + o_);  q.di = __ldg(f.dinv + o_);  q.pc = 0.0;
#endif
      if (CDIR == 0) {
        const int pos = li_ - foff, par = pos & 1;
        const int oc  = (((pos - par) >> 1) + FS_RO) * Pc + lj + FS_CO;
        q.xm = __ldg(xc + oc);
        q.xp = par ? __ldg(xc + oc + Pc) : 0.0;
#if FS_UP_MODE == 0
        if (par) q.pc = __ldg(f.pc + o_);
#endif
      } else {
        const int pos = lj - foff, par = pos & 1;
        const int oc  = (li_ + FS_RO) * Pc + ((pos - par) >> 1) + FS_CO;
        q.xm = __ldg(xc + oc);
        q.xp = __ldg(xc + oc + 1);
#if FS_UP_MODE == 0
        q.pc = __ldg(f.pc + o_);
#endif
      }
      return q;
    };
#if FS_UP_PREFETCH
    Row nx = load_row(li, o);
#endif
#if FS_UP_ASYNC > 0
    // the seven operand words of row li_ (the same addresses load_row reads) into ring slot sl
    auto issue_row = [&](int li_, int o_, int sl) {
      const int tx = threadIdx.x;
      cpa(&ring[sl][0][tx], f.cx + o_ + P);
      cpa(&ring[sl][1][tx], fb + o_);
      cpa(&ring[sl][2][tx], f.cy + o_);
      cpa(&ring[sl][3][tx], f.cy + o_ + 1);
      cpa(&ring[sl][4][tx], f.d + o_);
      if (CDIR == 0) {
        const int pos = li_ - foff, par = pos & 1;
        const int oc  = (((pos - par) >> 1) + FS_RO) * Pc + lj + FS_CO;
        cpa(&ring[sl][5][tx], xc + oc);
        if (par) cpa(&ring[sl][6][tx], xc + oc + Pc);
      } else {
        const int pos = lj - foff, par = pos & 1;
        const int oc  = (li_ + FS_RO) * Pc + ((pos - par) >> 1) + FS_CO;
        cpa(&ring[sl][5][tx], xc + oc);
        cpa(&ring[sl][6][tx], xc + oc + 1);
      }
    };
    auto read_row = [&](int li_, int sl) {
      const int tx = threadIdx.x;
      Row q;
      q.cxnn = ring[sl][0][tx];
      q.b    = ring[sl][1][tx];
      q.cy   = ring[sl][2][tx];
      q.cy1  = ring[sl][3][tx];
      q.d    = ring[sl][4][tx];
      q.di = 0.0;  q.pc = 0.0;
      q.xm = ring[sl][5][tx];
      q.xp = (CDIR == 0 && (((li_ - foff) & 1) == 0)) ? 0.0 : ring[sl][6][tx];
      return q;
    };
#pragma unroll
    for (int dd = 0; dd < FS_UP_ASYNC; ++dd) {  // one group per row, empty past the chunk's last row
      if (li + dd <= r1) issue_row(li + dd, o + dd * P, dd);
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
    int slot = 0;
#endif
#pragma unroll 2
    for (; li <= r1; ++li, o += P) {
#if FS_PF_ROWS > 0
      if (li + FS_PF_ROWS <= r1) {  // cx is read one row further down than the others
        const int op = o + FS_PF_ROWS * P;
        pf_l2(fb + op, lane); pf_l2(f.d + op, lane); pf_l2(f.cy + op, lane); pf_l2(f.cx + op + P, lane);
      }
#endif
#if FS_UP_ASYNC > 0
      // all but the newest FS_UP_ASYNC - 1 groups have landed: row li is in its slot
      asm volatile("cp.async.wait_group %0;" ::"n"(FS_UP_ASYNC - 1) : "memory");
      const Row q = read_row(li, slot);
#elif FS_UP_PREFETCH
      const Row q = nx;
      if (li < r1) nx = load_row(li + 1, o + P);
#else
      const Row q = load_row(li, o);
#endif
      const double cxnn = q.cxnn, bn = q.b, dn = q.d, cyn = q.cy, cy1n = q.cy1;
#if FS_UP_MODE == 1
      const double din = rcp0(dn);
#else
      const double din = q.di;
#endif
      double xi;
      if (CDIR == 0) {
        const int par = (li - foff) & 1;
        if (par == 0) {  // C row (warp uniform)
          xi = q.xm;
        } else {
#if FS_UP_MODE == 0
          const double pcv = q.pc;
#else
          const double pcv = rcp0(dn + cyn + cy1n);
#endif
          double v = (-cxn * pcv) * q.xm;
          v += (-cxnn * pcv) * q.xp;
          xi = v;
        }
      } else {
        const int par = (lj - foff) & 1;
#if FS_UP_MODE == 0
        const double pcv = q.pc;
#else
        const double pcv = rcp0(dn + cxn + cxnn);
#endif
        double v = (-cyn * pcv) * q.xm;
        v += (-cy1n * pcv) * q.xp;
        xi = par ? v : q.xm;
      }
      const double x0n = w * bn * din + xi;
      if (li > r0) {  // emit row li - 1
        const double dd = dc;
        const double xl = __shfl_up_sync(0xffffffffu, x0c, 1), xr = __shfl_down_sync(0xffffffffu, x0c, 1);
        double v = dd * x0c;
        v += cxc * x0p;
        v += cxn * x0n;
        v += cyc * xl;
        v += cy1c * xr;
        const double xo = x0c + w * (bc - v) * dic;
        if (store_lane) {
          fx[o - P] = xo;
          if (DOT) acc[0] += bc * xo;
        }
      }
      x0p = x0c; x0c = x0n; bc = bn; dic = din; dc = dn; cyc = cyn; cy1c = cy1n; cxc = cxn; cxn = cxnn;
#if FS_UP_ASYNC > 0
      // refill the slot just consumed (its words are in registers; the clobber keeps the reads above)
      if (li + FS_UP_ASYNC <= r1) issue_row(li + FS_UP_ASYNC, o + FS_UP_ASYNC * P, slot);
      asm volatile("cp.async.commit_group;" ::: "memory");
      slot = slot + 1 == FS_UP_ASYNC ? 0 : slot + 1;
#endif
    }
  }
  if (DOT) {
    double tot[1];
    if (grid_reduce<ReduceSum, 1>(acc, part, ticket, tot)) {
      if (threadIdx.x == 0) pcg_local_rz<DOT == 2>(scal, tot[0]);
    }
  }
}

// <r, z> as its own pass (only for one-level hierarchies, where no k_up runs)
template <bool INIT>
__global__ void __launch_bounds__(BX* BY)
k_dot_rz(const LevelDev L, const double* __restrict__ r, const double* __restrict__ z, PcgScal* scal,
         double* part, unsigned int* ticket) {
  if (scal->done) return;
  double v[1] = {0.0};
  for_each_point(L.i0, L.i1, L.n1, [&](int64_t li, int64_t lj) {
    const int64_t k = L.idx(li, lj);
    v[0] += r[k] * z[k];
  });
  double tot[1];
  if (grid_reduce<ReduceSum, 1>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0) pcg_local_rz<INIT>(scal, tot[0]);
  }
}

// ---- coarsest level: `sweeps` Jacobi sweeps with weight 1 from a zero guess
// (pfmg_setup.c:400-412 sets the sweep count, pfmg_relax.c / jacobi.c the iteration).
// One CTA; x ping-pongs between two shared-memory tiles with a one-point zero ring.
// STAGE: the five coefficient arrays and b are staged in shared memory as well (7 tiles), so a
// sweep touches no global memory at all; otherwise they are re-read through L1/L2 every sweep.
template <bool STAGE>
__global__ void __launch_bounds__(1024)
k_coarse_smem(const LevelDev L, const double* __restrict__ b, double* __restrict__ x, int sweeps,
              int active, const PcgScal* scal) {
  if (scal->done) return;
  extern __shared__ double sm[];
  const int64_t n0 = L.n0, n1 = L.n1, w1 = n1 + 2, tot = (n0 + 2) * w1, n = n0 * n1;
  double* xa = sm;
  double* xb = sm + tot;
  double *sd = sm + 2 * tot, *sdi = sm + 3 * tot, *scx = sm + 4 * tot, *scy = sm + 5 * tot, *sb = sm + 6 * tot;
  const bool on = active && !scal->coarse_zero_diag;
  for (int64_t t = threadIdx.x; t < (STAGE ? 7 : 2) * tot; t += blockDim.x) sm[t] = 0.0;
  __syncthreads();
  if (on) {
    for (int64_t t = threadIdx.x; t < n; t += blockDim.x) {
      const int64_t li = t / n1, lj = t - li * n1, k = L.idx(li, lj), m = (li + 1) * w1 + lj + 1;
      xa[m] = b[k] * L.dinv[k];
      if (STAGE) {  // couplings past the edge stay 0, as the pads of the global arrays
        sd[m] = L.d[k]; sdi[m] = L.dinv[k]; scx[m] = L.cx[k]; scy[m] = L.cy[k]; sb[m] = b[k];
      }
    }
    __syncthreads();
    for (int it = 1; it < sweeps; ++it) {
      for (int64_t t = threadIdx.x; t < n; t += blockDim.x) {
        const int64_t li = t / n1, lj = t - li * n1;
        const int64_t m = (li + 1) * w1 + lj + 1;
        if (STAGE) {
          double v = sd[m] * xa[m];
          v += scx[m] * xa[m - w1];
          v += scx[m + w1] * xa[m + w1];
          v += scy[m] * xa[m - 1];
          v += scy[m + 1] * xa[m + 1];
          xb[m] = xa[m] + (sb[m] - v) * sdi[m];
        } else {
          const int64_t k = L.idx(li, lj), P = L.pitch;
          double v = L.d[k] * xa[m];
          v += L.cx[k] * xa[m - w1];
          v += L.cx[k + P] * xa[m + w1];
          v += L.cy[k] * xa[m - 1];
          v += L.cy[k + 1] * xa[m + 1];
          xb[m] = xa[m] + (b[k] - v) * L.dinv[k];
        }
      }
      __syncthreads();
      double* t2 = xa; xa = xb; xb = t2;
    }
  }
  for (int64_t t = threadIdx.x; t < n; t += blockDim.x) {
    const int64_t li = t / n1, lj = t - li * n1;
    x[L.idx(li, lj)] = xa[(li + 1) * w1 + lj + 1];
  }
}

// fallback for a coarsest grid too large for shared memory, and for one-level hierarchies:
// xo = (ZERO ? 0 : xi) + w (b - A xi) / d
template <bool ZERO>
__global__ void __launch_bounds__(BX* BY)
k_jacobi(const LevelDev L, const double* __restrict__ b, const double* __restrict__ xi,
         double* __restrict__ xo, double w, int on, const PcgScal* scal) {
  if (scal->done) return;
  const int64_t lj = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t li = L.i0 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (li >= L.i1 || lj >= L.n1) return;
  const int64_t k = L.idx(li, lj), P = L.pitch;
  if (!on || scal->coarse_zero_diag * (on == 2)) {
    xo[k] = 0.0;
    return;
  }
  if (ZERO) {
    xo[k] = w * b[k] * L.dinv[k];
    return;
  }
  double v = L.d[k] * xi[k];
  v += L.cx[k] * xi[k - P];
  v += L.cx[k + P] * xi[k + P];
  v += L.cy[k] * xi[k - 1];
  v += L.cy[k + 1] * xi[k + 1];
  xo[k] = xi[k] + w * (b[k] - v) * L.dinv[k];
}

// ---- the tail of the hierarchy in ONE kernel ------------------------------------------------------
// From the first level whose padded grid fits a 48 KB shared-memory tile (<= FS_TAIL_TILE points)
// down to the coarsest, a V-cycle leg is a handful of points per thread: as separate launches
// those ~2 x 15 levels cost 5-10 us each (launch + drain), 6 % of a timestep at 8192^2.  Here one
// CTA walks them all -- restriction down, the coarsest sweeps, interpolation + post-relaxation up --
// with a block barrier between levels.  b and x of the tail levels stay in their global arrays
// (L1/L2 resident; plain loads, never ld.global.nc, because this kernel writes them); the
// pre-relaxed iterate x0 of the level being post-relaxed lives in shared memory.  Same formulas
// and association as k_down_march* / k_up_march / k_coarse_smem.
#define FS_TAIL_MAX 24
#define FS_TAIL_TILE 6144
struct TailArgs {
  int n;       // number of tail levels; v[0] is the first tail level, v[n - 1] the coarsest
  int sweeps;  // coarsest-level Jacobi sweeps
  LevelDev v[FS_TAIL_MAX];
  int cdir[FS_TAIL_MAX], active[FS_TAIL_MAX];
  double w[FS_TAIL_MAX];
};

// residual of the zero-guess pre-relaxation at q: b - A (w b / d)   (inactive level: b)
__device__ __forceinline__ double tail_resid(const LevelDev& f, const double* b, bool active, double w,
                                             int64_t q) {
  const double bq = b[q];
  if (!active) return bq;
  const int64_t P = f.pitch;
  double v = w * bq;
  v += __ldg(f.cx + q) * (w * b[q - P] * __ldg(f.dinv + q - P));
  v += __ldg(f.cx + q + P) * (w * b[q + P] * __ldg(f.dinv + q + P));
  v += __ldg(f.cy + q) * (w * b[q - 1] * __ldg(f.dinv + q - 1));
  v += __ldg(f.cy + q + 1) * (w * b[q + 1] * __ldg(f.dinv + q + 1));
  return bq - v;
}

__global__ void __launch_bounds__(1024)
k_tail_vcycle(const TailArgs a, const PcgScal* scal) {
  if (scal->done) return;
  extern __shared__ double sm[];
  const int tid = threadIdx.x, nth = blockDim.x;
  // ---- down: b of level k -> b of level k + 1
  for (int k = 0; k + 1 < a.n; ++k) {
    const LevelDev& f = a.v[k];
    const LevelDev& c = a.v[k + 1];
    const int64_t P = f.pitch, nc = c.n0 * c.n1;
    const bool act = a.active[k] != 0;
    const double w = a.w[k];
    const int cd = a.cdir[k];
    for (int64_t t = tid; t < nc; t += nth) {
      const int64_t ci = t / c.n1, cj = t - ci * c.n1;
      const int64_t li = cd == 0 ? 2 * ci + c.foff : ci, lj = cd == 0 ? cj : 2 * cj + c.foff;
      const int64_t q = f.idx(li, lj), st = cd == 0 ? P : 1;
      const double am = cd == 0 ? __ldg(f.cx + q) : __ldg(f.cy + q);
      const double ap = cd == 0 ? __ldg(f.cx + q + P) : __ldg(f.cy + q + 1);
      double v = tail_resid(f, f.b, act, w, q);
      v += (-am * __ldg(f.pc + q - st)) * tail_resid(f, f.b, act, w, q - st);
      v += (-ap * __ldg(f.pc + q + st)) * tail_resid(f, f.b, act, w, q + st);
      c.b[c.idx(ci, cj)] = v;
    }
    __syncthreads();
  }
  // ---- coarsest level: `sweeps` Jacobi sweeps (weight 1) from a zero guess, as k_coarse_smem
  {
    const LevelDev& L = a.v[a.n - 1];
    const int64_t n0 = L.n0, n1 = L.n1, w1 = n1 + 2, tot = (n0 + 2) * w1, n = n0 * n1, P = L.pitch;
    double* xa = sm;
    double* xb = sm + tot;
    const bool on = a.active[a.n - 1] && !scal->coarse_zero_diag;
    for (int64_t t = tid; t < 2 * tot; t += nth) sm[t] = 0.0;
    __syncthreads();
    if (on) {
      for (int64_t t = tid; t < n; t += nth) {
        const int64_t li = t / n1, lj = t - li * n1, q = L.idx(li, lj);
        xa[(li + 1) * w1 + lj + 1] = L.b[q] * __ldg(L.dinv + q);
      }
      __syncthreads();
      for (int it = 1; it < a.sweeps; ++it) {
        for (int64_t t = tid; t < n; t += nth) {
          const int64_t li = t / n1, lj = t - li * n1, q = L.idx(li, lj), m = (li + 1) * w1 + lj + 1;
          double v = __ldg(L.d + q) * xa[m];
          v += __ldg(L.cx + q) * xa[m - w1];
          v += __ldg(L.cx + q + P) * xa[m + w1];
          v += __ldg(L.cy + q) * xa[m - 1];
          v += __ldg(L.cy + q + 1) * xa[m + 1];
          xb[m] = xa[m] + (L.b[q] - v) * __ldg(L.dinv + q);
        }
        __syncthreads();
        double* t2 = xa; xa = xb; xb = t2;
      }
    }
    for (int64_t t = tid; t < n; t += nth) {
      const int64_t li = t / n1, lj = t - li * n1;
      L.x[L.idx(li, lj)] = xa[(li + 1) * w1 + lj + 1];
    }
    __syncthreads();
  }
  // ---- up: x of level k + 1 -> x of level k (interpolation + post-relaxation)
  for (int k = a.n - 2; k >= 0; --k) {
    const LevelDev& f = a.v[k];
    const LevelDev& c = a.v[k + 1];
    const int64_t P = f.pitch, n1 = f.n1, w1 = n1 + 2, tot = (f.n0 + 2) * w1, n = f.n0 * n1;
    const bool act = a.active[k] != 0;
    const double w = a.w[k];
    const int cd = a.cdir[k];
    if (act) {
      for (int64_t t = tid; t < tot; t += nth) sm[t] = 0.0;
      __syncthreads();
    }
    for (int64_t t = tid; t < n; t += nth) {
      const int64_t li = t / n1, lj = t - li * n1, q = f.idx(li, lj);
      double x = act ? w * f.b[q] * __ldg(f.dinv + q) : 0.0;
      const int64_t pos = (cd == 0 ? li : lj) - c.foff;  // even <=> C-point
      if ((pos & 1) == 0) {
        const int64_t cp = pos >> 1;
        x += c.x[cd == 0 ? c.idx(cp, lj) : c.idx(li, cp)];
      } else {
        const int64_t cm = (pos - 1) >> 1;
        const int64_t kc = cd == 0 ? c.idx(cm, lj) : c.idx(li, cm);
        const double pc  = __ldg(f.pc + q);
        const double am = cd == 0 ? __ldg(f.cx + q) : __ldg(f.cy + q);
        const double ap = cd == 0 ? __ldg(f.cx + q + P) : __ldg(f.cy + q + 1);
        double v = (-am * pc) * c.x[kc];
        v += (-ap * pc) * c.x[kc + (cd == 0 ? c.pitch : 1)];
        x += v;
      }
      if (act) sm[(li + 1) * w1 + lj + 1] = x;
      else f.x[q] = x;
    }
    __syncthreads();
    if (act) {
      for (int64_t t = tid; t < n; t += nth) {
        const int64_t li = t / n1, lj = t - li * n1, q = f.idx(li, lj), m = (li + 1) * w1 + lj + 1;
        const double x0 = sm[m];
        double v = __ldg(f.d + q) * x0;
        v += __ldg(f.cx + q) * sm[m - w1];
        v += __ldg(f.cx + q + P) * sm[m + w1];
        v += __ldg(f.cy + q) * sm[m - 1];
        v += __ldg(f.cy + q + 1) * sm[m + 1];
        f.x[q] = x0 + w * (f.b[q] - v) * __ldg(f.dinv + q);
      }
      __syncthreads();
    }
  }
}

dim3 lvl_grid(const LevelDev& v) { return dim3((unsigned)cdiv(v.n1, BX), (unsigned)cdiv(v.i1 - v.i0, BY)); }

}  // namespace

static int coarse_solve(fsgpu_ps* s, int l, const double* b, double* x) {
  cudaStream_t st = s->c->stream;
  const Level& L  = s->L[l];
  const size_t tile = (size_t)((L.v.n0 + 2) * (L.v.n1 + 2)) * sizeof(double);
  const size_t cap  = 200 * 1024;
  if (2 * tile <= cap) {
    // per device, and cheap: set on every call (several contexts / devices may share the process)
    if (7 * tile <= cap) {
      FS_CUDA(cudaFuncSetAttribute(k_coarse_smem<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cap));
      k_coarse_smem<true><<<1, 1024, 7 * tile, st>>>(L.v, b, x, s->coarse_sweeps, L.active, s->scal);
    } else {
      FS_CUDA(cudaFuncSetAttribute(k_coarse_smem<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cap));
      k_coarse_smem<false><<<1, 1024, 2 * tile, st>>>(L.v, b, x, s->coarse_sweeps, L.active, s->scal);
    }
    FS_LAUNCHED();
    return 0;
  }
  // global-memory ping-pong; `on == 2` lets the kernel honour the zero-diagonal flag
  double* cur = x;
  double* oth = s->ctmp;
  if ((s->coarse_sweeps & 1) == 0) std::swap(cur, oth);  // so that the last sweep lands in x
  k_jacobi<true><<<lvl_grid(L.v), dim3(BX, BY), 0, st>>>(L.v, b, nullptr, cur, 1.0, L.active ? 2 : 0, s->scal);
  FS_LAUNCHED();
  for (int it = 1; it < s->coarse_sweeps; ++it) {
    k_jacobi<false><<<lvl_grid(L.v), dim3(BX, BY), 0, st>>>(L.v, b, cur, oth, 1.0, L.active ? 2 : 0, s->scal);
    FS_LAUNCHED();
    std::swap(cur, oth);
  }
  return 0;
}

// rows per marching task: 64 when that still gives every SM several tasks, else smaller chunks
// (more halo-row overlap, but small levels are latency bound, not bandwidth bound)
static int pick_rows(int64_t rows, int64_t strips) {
  int mr = 64;
  while (mr > 4 && strips * cdiv(rows, mr) < (int64_t)FS_NUM_SMS * 16) mr >>= 1;
  // tiny levels are pure latency: a task's rows are a serial chain of dependent loads, so hand
  // every warp as few rows as possible while the whole level still fits one wave of CTAs
  while (mr > 1 && strips * cdiv(rows, mr) < (int64_t)FS_NUM_SMS * 2) mr >>= 1;
  return mr;
}
static unsigned march_blocks(int64_t rows, int64_t strips, int mr) {
  const int64_t total = strips * cdiv(rows, mr);
  const int64_t cap = (int64_t)FS_NUM_SMS * 8;
  return (unsigned)(total < cap ? (total > 0 ? total : 1) : cap);
}

// [r0, r1): the COARSE rows to produce (r0 < 0: all rows this rank owns / the whole replicated level).  A
// j-coarsened level keeps its rows, so there the range is the fine rows swept as well.
static int launch_down(fsgpu_ps* s, int l, const double* fb, int64_t r0 = -1, int64_t r1 = -1) {
  cudaStream_t st = s->c->stream;
  Level* L        = s->L;
  LevelDev f = L[l].v;
  LevelDev c = L[l + 1].v;
  if (!L[l].replicated) {  // a distributed level writes the coarse rows this rank owns
    c.i0 = L[l + 1].own0;
    c.i1 = L[l + 1].own1;
  }
  if (r0 >= 0) {
    if (r1 <= r0) return 0;
    if (L[l].cdir == 0) { c.i0 = r0; c.i1 = r1; }
    else { f.i0 = r0; f.i1 = r1; }
  }
  if (L[l].cdir == 0) {
    const int64_t strips = cdiv(f.n1, 120);
    const int mrc = pick_rows(c.i1 - c.i0, strips);
    const unsigned nb = march_blocks(c.i1 - c.i0, strips, mrc);
    if (L[l].active) k_down_march0<true><<<nb, 128, 0, st>>>(f, fb, c, c.b, L[l].weight, mrc, s->scal);
    else k_down_march0<false><<<nb, 128, 0, st>>>(f, fb, c, c.b, L[l].weight, mrc, s->scal);
  } else {
    const int64_t strips = cdiv(f.n1, 112);
    const int mr = pick_rows(f.i1 - f.i0, strips);
    const unsigned nb = march_blocks(f.i1 - f.i0, strips, mr);
    if (L[l].active) k_down_march1<true><<<nb, 128, 0, st>>>(f, fb, c, c.b, L[l].weight, mr, s->scal);
    else k_down_march1<false><<<nb, 128, 0, st>>>(f, fb, c, c.b, L[l].weight, mr, s->scal);
  }
  FS_LAUNCHED();
  return 0;
}

// [r0, r1): the rows of level l to produce (r0 < 0: all rows this rank sweeps)
template <int CDIR>
static int launch_up(fsgpu_ps* s, int l, const double* fb, double* fx, int dot_mode, int64_t r0 = -1,
                     int64_t r1 = -1) {
  cudaStream_t st = s->c->stream;
  Level* L        = s->L;
  fsgpu_ctx* c    = s->c;
  LevelDev f      = L[l].v;
  if (r0 >= 0) {
    if (r1 <= r0) return 0;
    f.i0 = r0; f.i1 = r1;
  }
  if (!L[l].active) {
    k_up_interp<CDIR><<<persist_blocks(f.i1 - f.i0, f.n1), dim3(BX, BY), 0, st>>>(f, fx, L[l + 1].v, L[l + 1].v.x, s->scal);
    FS_LAUNCHED();
    return 0;
  }
  const int64_t strips = cdiv(f.n1, MOUT * MWARPS);
  const int mr      = pick_rows(f.i1 - f.i0, strips);
  const unsigned nb = march_blocks(f.i1 - f.i0, strips, mr);
  const int bl = 32 * MWARPS;
#define UP_ARGS f, fb, fx, L[l + 1].v, L[l + 1].v.x, L[l].weight, mr, s->scal, c->part, c->ticket
  if (dot_mode == 1) k_up_march<CDIR, 1><<<nb, bl, 0, st>>>(UP_ARGS);
  else if (dot_mode == 2) k_up_march<CDIR, 2><<<nb, bl, 0, st>>>(UP_ARGS);
  else k_up_march<CDIR, 0><<<nb, bl, 0, st>>>(UP_ARGS);
#undef UP_ARGS
  FS_LAUNCHED();
  return 0;
}

// first level of the one-kernel tail (k_tail_vcycle), or nl if the hierarchy has none: at least two
// levels, every one of them whole on this rank (single rank, or replicated levels) and small
// enough for the shared-memory tile.  FSGPU_NO_TAIL=1 keeps the per-level launches (A/B runs).
static int tail_start(const fsgpu_ps* s) {
  static const bool off = getenv("FSGPU_NO_TAIL") != nullptr;
  if (off) return s->nl;
  int lt = s->nl;
  for (int l = s->nl - 1; l >= 1; --l) {
    const LevelDev& v = s->L[l].v;
    if ((v.n0 + 2) * (v.n1 + 2) > FS_TAIL_TILE || v.i0 != 0 || v.i1 != v.n0) break;
    if (s->c->nranks > 1 && l < s->ld) break;
    lt = l;
  }
  if (s->nl - lt < 2 || s->nl - lt > FS_TAIL_MAX) return s->nl;
  return lt;
}
static int launch_tail(fsgpu_ps* s, int lt) {
  TailArgs a;
  a.n = s->nl - lt;
  a.sweeps = s->coarse_sweeps;
  size_t tile = 0;
  for (int k = 0; k < a.n; ++k) {
    const Level& L = s->L[lt + k];
    a.v[k] = L.v; a.cdir[k] = L.cdir; a.active[k] = L.active; a.w[k] = L.weight;
    if (k + 1 < a.n) tile = std::max(tile, (size_t)((L.v.n0 + 2) * (L.v.n1 + 2)));
  }
  const LevelDev& C = s->L[s->nl - 1].v;
  const size_t smem = std::max(tile, (size_t)(2 * (C.n0 + 2) * (C.n1 + 2))) * sizeof(double);
  if (smem > 48 * 1024)
    FS_CUDA(cudaFuncSetAttribute(k_tail_vcycle, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_tail_vcycle<<<1, 1024, smem, s->c->stream>>>(a, s->scal);
  FS_LAUNCHED();
  return 0;
}

int ps_vcycle(fsgpu_ps* s, const double* r, double* z, int dot_mode) {
  cudaStream_t st = s->c->stream;
  Level* L        = s->L;
  const int nl    = s->nl;
  if (fs_ensure_part(s->c, (size_t)FS_NUM_SMS * 8 * 4)) return 1;
  const bool dist = s->c->nranks > 1;
  if (nl == 1) {  // relax from zero, relax again (pfmg_solve.c with a single level)
    k_jacobi<true><<<lvl_grid(L[0].v), dim3(BX, BY), 0, st>>>(L[0].v, r, nullptr, s->ctmp, L[0].weight, 1, s->scal);
    FS_LAUNCHED();
    if (ps_halo(s, 0, &s->ctmp, 1, 1)) return 1;
    k_jacobi<false><<<lvl_grid(L[0].v), dim3(BX, BY), 0, st>>>(L[0].v, r, s->ctmp, z, L[0].weight, 1, s->scal);
    FS_LAUNCHED();
    const unsigned pg = persist_blocks(L[0].v.i1 - L[0].v.i0, L[0].v.n1);
    if (dot_mode == 1) k_dot_rz<false><<<pg, dim3(BX, BY), 0, st>>>(L[0].v, r, z, s->scal, s->c->part, s->c->ticket);
    if (dot_mode == 2) k_dot_rz<true><<<pg, dim3(BX, BY), 0, st>>>(L[0].v, r, z, s->scal, s->c->part, s->c->ticket);
    if (dot_mode) FS_LAUNCHED();
    if (dot_mode && ps_finish_rz(s, dot_mode == 2)) return 1;
    return 0;
  }
  // down: level l rhs -> level l+1 rhs.  Slab decomposition: a distributed coarse level gets
  // FS_RO halo rows of its rhs from the neighbours; the first replicated level is gathered
  // (every rank writes its rows into a zeroed full array, then a sum all-reduce).
  const int lt = tail_start(s);  // levels [lt, nl) run inside k_tail_vcycle (lt == nl: none)
  const int lbot = lt < nl ? lt : nl - 1;
  for (int l = 0; l <= lbot - 1; ++l) {
    const double* fb = l == 0 ? r : L[l].v.b;
    ProfScope ps_(l == 0 ? PROF_DOWN0 : PROF_DOWN, st);
    const bool gather = dist && l + 1 == s->ld;
    const size_t cn = (size_t)(L[l + 1].v.n0 * L[l + 1].v.pitch);
    double* crow0 = L[l + 1].v.b + FS_RO * L[l + 1].v.pitch;  // row 0 of the replicated level
    if (gather) FS_CUDA(cudaMemsetAsync(crow0, 0, cn * sizeof(double), st));
    if (dist && l + 1 < s->ld) {
      // the coarse rhs needs FS_RO halo rows: boundary rows first, their exchange on the transport's
      // stream while the interior rows are restricted
      const SlabSplit sp = ps_split(s, l + 1, FS_RO);
      bool async = false;
      if (sp.on) {
        if (launch_down(s, l, fb, sp.bnd.a0, sp.bnd.a1) || launch_down(s, l, fb, sp.bnd.b0, sp.bnd.b1)) return 1;
        if (comm_halo_async(s->c, &L[l + 1].v.b, 1, L[l + 1].v.pitch, L[l + 1].own0, L[l + 1].own1, FS_RO, &async))
          return 1;
        if (launch_down(s, l, fb, sp.in.a0, sp.in.a1)) return 1;
      } else if (launch_down(s, l, fb)) {
        return 1;
      }
      if (async) { if (comm_halo_wait(s->c)) return 1; }
      else if (ps_halo(s, l + 1, &L[l + 1].v.b, 1, FS_RO)) return 1;
      continue;
    }
    if (launch_down(s, l, fb)) return 1;
    if (gather && comm_allreduce(s->c, crow0, (int64_t)cn, COMM_SUM)) return 1;
  }
  {
    ProfScope ps_(PROF_COARSE, st);
    if (lt < nl) { if (launch_tail(s, lt)) return 1; }
    else if (coarse_solve(s, nl - 1, L[nl - 1].v.b, L[nl - 1].v.x)) return 1;
  }
  // up: level l+1 x -> level l x; the level-0 kernel also carries PCG's <r, z>.  A distributed
  // level's x needs one halo row before the next finer level interpolates from it.
  for (int l = lbot - 1; l >= 0; --l) {
    const double* fb = l == 0 ? r : L[l].v.b;
    double* fx       = l == 0 ? z : L[l].v.x;
    ProfScope ps_(l == 0 ? PROF_UP0 : PROF_UP, st);
    const int dm = l == 0 ? dot_mode : 0;
    auto up = [&](int64_t r0, int64_t r1) {
      return L[l].cdir == 0 ? launch_up<0>(s, l, fb, fx, dm, r0, r1) : launch_up<1>(s, l, fb, fx, dm, r0, r1);
    };
    if (dist && l >= 1 && l < s->ld) {  // x of a distributed level: one halo row for the next finer level
      const SlabSplit sp = ps_split(s, l, 1);
      bool async = false;
      if (sp.on) {
        if (up(sp.bnd.a0, sp.bnd.a1) || up(sp.bnd.b0, sp.bnd.b1)) return 1;
        if (comm_halo_async(s->c, &fx, 1, L[l].v.pitch, L[l].own0, L[l].own1, 1, &async)) return 1;
        if (up(sp.in.a0, sp.in.a1)) return 1;
      } else if (up(-1, -1)) {
        return 1;
      }
      if (async) { if (comm_halo_wait(s->c)) return 1; }
      else if (ps_halo(s, l, &fx, 1, 1)) return 1;
    } else if (up(-1, -1)) {
      return 1;
    }
  }
  if (dot_mode && ps_finish_rz(s, dot_mode == 2)) return 1;
  return 0;
}
