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

namespace {

constexpr int BX = 64, BY = 4;

// 1 / v with the convention of the stored dinv / pc arrays (0 where v == 0: pads, halo-less walls,
// a singular coarsest point).  The active-level kernels recompute dinv = 1/d and the interpolation
// denominator pc = 1/(d + t_m + t_p) from d, cx, cy instead of loading them: two FP64 divisions
// per point buy 16 B/point less HBM traffic, and the values are bit-identical to the stored ones
// (same operands, same association as k_interp / k_rap, correctly rounded division).
__device__ __forceinline__ double rcp0(double v) { return v != 0.0 ? 1.0 / v : 0.0; }

// The per-point arithmetic of the marching kernels, with the multiply-add contraction written out:
// a sum of two products can be fused either way round, and the register-pipelined and the TMA-fed
// version of a kernel must round alike to stay bit-identical (tests/test_gpu_solver.py).
// residual of the zero-guess Jacobi iterate X = w b / d:  b - (w b + sum of the four couplings * X)
__device__ __forceinline__ double mg_resid(double w, double b, double ca, double Xa, double cb, double Xb,
                                           double cl, double Xl, double cr, double Xr) {
  double v = __dmul_rn(w, b);
  v = fma(ca, Xa, v);
  v = fma(cb, Xb, v);
  v = fma(cl, Xl, v);
  v = fma(cr, Xr, v);
  return b - v;
}
// restriction  R + (-cm pm) Rm + (-cp pp) Rp
__device__ __forceinline__ double mg_restrict(double R, double cm, double pm, double Rm, double cp, double pp,
                                              double Rp) {
  double v = fma(__dmul_rn(-cm, pm), Rm, R);
  return fma(__dmul_rn(-cp, pp), Rp, v);
}
// interpolation at an F point  (-cm pc) xm + (-cp pc) xp
__device__ __forceinline__ double mg_interp(double cm, double cp, double pc, double xm, double xp) {
  const double v = __dmul_rn(__dmul_rn(-cm, pc), xm);
  return fma(__dmul_rn(-cp, pc), xp, v);
}
// pre-relaxed iterate + interpolated correction  w b / d + xi
__device__ __forceinline__ double mg_x0(double w, double b, double dinv, double xi) {
  return fma(__dmul_rn(w, b), dinv, xi);
}
// post-relaxation  x0 + w (b - A x0) / d
__device__ __forceinline__ double mg_post(double w, double b, double d, double dinv, double x0, double ca,
                                          double xa, double cb, double xb, double cl, double xl, double cr,
                                          double xr) {
  double v = __dmul_rn(d, x0);
  v = fma(ca, xa, v);
  v = fma(cb, xb, v);
  v = fma(cl, xl, v);
  v = fma(cr, xr, v);
  return fma(__dmul_rn(w, b - v), dinv, x0);
}

// The rhs of the pressure system with its mean still in it (fill_pressure_rhs subtracts the mean over all
// (NX+2)(NY+2) entries, src/LinearSystem.hpp:150-152): fsgpu_substep leaves the subtraction to the first
// readers of the rhs -- the level-0 kernels of the first V-cycle here and the first x / r update -- instead of
// a pass of its own over the array.  sum == null: nothing to subtract.  Pads stay zero: callers mask.
struct BShift {
  const double* sum;  // device: the sum of all entries (on every rank)
  double size;        // number of entries
  __device__ __forceinline__ double mean() const { return sum ? __ldg(sum) / size : 0.0; }
};

template <int CDIR>
struct Coef {
  const double *cx, *cy;
  int64_t P;
  __device__ __forceinline__ double am(int64_t q) const { return CDIR == 0 ? __ldg(cx + q) : __ldg(cy + q); }
  __device__ __forceinline__ double ap(int64_t q) const { return CDIR == 0 ? __ldg(cx + q + P) : __ldg(cy + q + 1); }
};

// ---- work partition of the marching kernels --------------------------------------------------------
// A level is cut into tasks of `mr` rows x one column strip, numbered block-row by block-row, and dealt
// round-robin to the resident CTAs: the CTAs that run together then sweep neighbouring strips of the same
// few row blocks (a compact set of DRAM pages / TLB entries; handing every CTA one long private run of
// rows instead measured 8-30 % SLOWER at 8192^2).  Balance comes from the host's choice of mr
// (pick_rows): the task count lands just under a multiple of the grid size, so no round of the deal is
// left partly filled (a fixed mr = 64 gave 8832 tasks on 1184 resident CTAs: 7.46 rounds paid as 8).
struct MarchCursor {
  int t, total, strips, mr, lo, hi;
  __device__ MarchCursor(int strips_, int lo_, int hi_, int mr_)
      : t(blockIdx.x), strips(strips_), mr(mr_), lo(lo_), hi(hi_) {
    total = strips * ((hi - lo + mr - 1) / mr);
  }
  __device__ bool next(int& sp, int& r0, int& r1) {
    if (t >= total) return false;
    const int ch = t / strips;
    sp = t - ch * strips;
    r0 = lo + ch * mr;
    r1 = min(r0 + mr, hi);
    t += gridDim.x;
    return true;
  }
};

// ---- marching restriction kernels (same strip / task scheme as k_up_march below) --------------
// k_down_march0: coarsening along i.  Per fine row the warp forms X = w b / d, then the residual
// R of the row above from three rows of X held in registers (left/right by shuffle), and on every
// second row the coarse rhs  b_c = R + pp(row-1) R(row-1) + pm(row+1) R(row+1).
template <bool ACTIVE, bool SHIFT, bool PUSH>
__global__ void __launch_bounds__(128, FS_MARCH_MINB)
k_down_march0(const LevelDev f, const double* __restrict__ fb, const LevelDev c,
              double* __restrict__ cb, double w, int mr, const PcgScal* scal, const BShift bs, const HaloWait wt,
              const HaloPush ps) {
  if (scal->done) return;
  halo_wait(wt);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n1 = (int)f.n1, P = (int)f.pitch, Pc = (int)c.pitch, foff = (int)c.foff;
  const double bm = SHIFT ? bs.mean() : 0.0;
  MarchCursor cur((n1 + 119) / 120, (int)c.i0, (int)c.i1, mr);  // rows of a task: coarse rows
  int sp, c0, c1;
  while (cur.next(sp, c0, c1)) {
    int lj = sp * 120 + warp * 30 - 1 + lane;
    const bool store_lane = lane >= 1 && lane <= 30 && lj < n1;
    const bool colin = lj >= 0 && lj < n1;
    if (lj > n1) lj = n1;
    const int f0 = 2 * c0 + foff, fL = 2 * (c1 - 1) + foff;
    double X2 = 0.0, X1 = 0.0, b1 = 0.0, cx2 = 0.0, cx1 = 0.0, cy1 = 0.0, cy1p = 0.0, Ra = 0.0, Rb = 0.0;
    double pc1 = 0.0, pc2 = 0.0, pc3 = 0.0;  // ACTIVE: pc of rows li - 1, li - 2, li - 3
    int li = f0 - 2;
    int o  = (li + FS_RO) * P + lj + FS_CO;
#pragma unroll 2
    for (; li <= fL + 2; ++li, o += P) {
      double b0 = __ldg(fb + o);
      const double cx0 = __ldg(f.cx + o);
      if (SHIFT && colin && li >= 0 && li < (int)f.n0) b0 -= bm;
      double X0 = 0.0, cy0 = 0.0, cy0p = 0.0, pc0 = 0.0;
      if (ACTIVE) {
        const double d0 = __ldg(f.d + o);
        cy0  = __ldg(f.cy + o);
        cy0p = __ldg(f.cy + o + 1);
        X0   = __dmul_rn(__dmul_rn(w, b0), rcp0(d0));
        pc0  = rcp0(d0 + cy0 + cy0p);
      }
      // residual of row li - 1
      double Rn;
      if (ACTIVE) {
        const double xl = __shfl_up_sync(0xffffffffu, X1, 1), xr = __shfl_down_sync(0xffffffffu, X1, 1);
        Rn = mg_resid(w, b1, cx1, X2, cx0, X0, cy1, xl, cy1p, xr);
      } else {
        Rn = b1;
      }
      // coarse point on row li - 2 (a C row when li - f0 is even)
      if (li >= f0 + 2 && ((li - f0) & 1) == 0) {
        const double pcm = ACTIVE ? pc3 : __ldg(f.pc + o - 3 * P);
        const double pcp = ACTIVE ? pc1 : __ldg(f.pc + o - P);
        const double v = mg_restrict(Rb, cx2, pcm, Ra, cx1, pcp, Rn);
        const int crow = (li - 2 - foff) >> 1;
        if (store_lane) cb[(crow + FS_RO) * Pc + lj + FS_CO] = v;
        if (PUSH) {  // a boundary row of the slab: also into the neighbour's halo
          double* pr = halo_push_row(ps, crow);
          if (pr && store_lane) pr[(int64_t)(crow + FS_RO) * Pc + lj + FS_CO] = v;
        }
      }
      X2 = X1; X1 = X0; b1 = b0; cx2 = cx1; cx1 = cx0; cy1 = cy0; cy1p = cy0p; Ra = Rb; Rb = Rn;
      pc3 = pc2; pc2 = pc1; pc1 = pc0;
    }
  }
  if (PUSH) halo_push_done(ps);
}

// k_down_march1: coarsening along j.  Lanes 1..30 hold a valid residual (X of the neighbours by
// shuffle), lanes 2..29 that sit on a C column emit  b_c = R + pp(j-1) R(j-1) + pm(j+1) R(j+1).
template <bool ACTIVE, bool SHIFT, bool PUSH>
__global__ void __launch_bounds__(128, FS_MARCH_MINB)
k_down_march1(const LevelDev f, const double* __restrict__ fb, const LevelDev c,
              double* __restrict__ cb, double w, int mr, const PcgScal* scal, const BShift bs, const HaloWait wt,
              const HaloPush ps) {
  if (scal->done) return;
  halo_wait(wt);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n1 = (int)f.n1, P = (int)f.pitch, Pc = (int)c.pitch, foff = (int)c.foff;
  const double bm = SHIFT ? bs.mean() : 0.0;
  MarchCursor cur((n1 + 111) / 112, (int)f.i0, (int)f.i1, mr);
  int sp, r0, r1;
  while (cur.next(sp, r0, r1)) {
    int lj = sp * 112 + warp * 28 - 2 + lane;
    const int pos = lj - foff;
    const bool store_lane = lane >= 2 && lane <= 29 && lj < n1 && (pos & 1) == 0;
    const bool colin = lj >= 0 && lj < n1;
    const int oc_col = (pos >> 1) + FS_CO;
    if (lj > n1) lj = n1;
    double X2 = 0.0, X1 = 0.0, b1 = 0.0, cx1 = 0.0, cy1 = 0.0, cy1p = 0.0, d1 = 0.0;
    int li = r0 - 1;
    int o  = (li + FS_RO) * P + lj + FS_CO;
#pragma unroll 2
    for (; li <= r1; ++li, o += P) {
      double b0 = __ldg(fb + o);
      if (SHIFT && colin && li >= 0 && li < (int)f.n0) b0 -= bm;
      const double cy0 = __ldg(f.cy + o), cy0p = __ldg(f.cy + o + 1);
      double X0 = 0.0, cx0 = 0.0, d0 = 0.0;
      if (ACTIVE) {
        d0  = __ldg(f.d + o);
        X0  = __dmul_rn(__dmul_rn(w, b0), rcp0(d0));
        cx0 = __ldg(f.cx + o);
      }
      if (li > r0) {  // row li - 1
        double Rn;
        if (ACTIVE) {
          const double xl = __shfl_up_sync(0xffffffffu, X1, 1), xr = __shfl_down_sync(0xffffffffu, X1, 1);
          Rn = mg_resid(w, b1, cx1, X2, cx0, X0, cy1, xl, cy1p, xr);
        } else {
          Rn = b1;
        }
        const double pcv = ACTIVE ? rcp0(d1 + cx1 + cx0) : __ldg(f.pc + o - P);
        const double pcl = __shfl_up_sync(0xffffffffu, pcv, 1), pcr = __shfl_down_sync(0xffffffffu, pcv, 1);
        const double Rl = __shfl_up_sync(0xffffffffu, Rn, 1), Rr = __shfl_down_sync(0xffffffffu, Rn, 1);
        const double v = mg_restrict(Rn, cy1, pcl, Rl, cy1p, pcr, Rr);
        if (store_lane) cb[(li - 1 + FS_RO) * Pc + oc_col] = v;
        if (PUSH) {
          double* pr = halo_push_row(ps, li - 1);
          if (pr && store_lane) pr[(int64_t)(li - 1 + FS_RO) * Pc + oc_col] = v;
        }
      }
      X2 = X1; X1 = X0; b1 = b0; cx1 = cx0; cy1 = cy0; cy1p = cy0p; d1 = d0;
    }
  }
  if (PUSH) halo_push_done(ps);
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
            const double* __restrict__ cxv, const PcgScal* scal, const HaloWait wt, const HaloPush ps) {
  if (scal->done) return;
  halo_wait(wt);
  for_each_point(f.i0, f.i1, f.n1, [&](int64_t li, int64_t lj) {
    const int64_t k = f.idx(li, lj);
    const double v = x_interp<CDIR, false>(f, 0.0, c, cxv, li, lj, k, 0.0);
    fx[k] = v;
    double* pr = halo_push_row(ps, li);
    if (pr) pr[k] = v;
  });
  halo_push_done(ps);
}

// active level, "marching" formulation.  A warp owns a strip of 32 consecutive columns (30 of
// them produce output, lanes 0 and 31 are the left/right halo) and walks down MR rows.  Per row
// it forms  x0 = w b / d + P x_c  ONCE per point, keeps x0 of three consecutive rows in
// registers, takes the left/right neighbours by warp shuffle, and emits
// x = x0 + w (b - A x0) / d  for the row above.  No shared memory, no block barrier, 32-bit
// offsets advanced by one pitch per row; every array is read once (plus the 2/32 column and
// 2/MR row overlap).  dinv and pc are recomputed from d, cx, cy (rcp0 above).  CTAs loop over
// (strip, chunk) tasks, so the optional dot product <b, x> (= <r, z> of PCG on level 0) ends in
// ~1e3 partials.
// DOT: 0 none, 1 PCG iteration (convergence test + beta), 2 first V-cycle (bi).
// FS_UP_WARPS warps (= adjacent 30-column strips) per CTA; the CTAs per SM scale down with it (32 warps per SM)
#ifndef FS_UP_WARPS
#define FS_UP_WARPS 8
#endif
constexpr int MOUT = 30, MWARPS = FS_UP_WARPS, UP_MINB = 32 / FS_UP_WARPS;

template <int CDIR, int DOT, bool PUSH>
__global__ void __launch_bounds__(32 * MWARPS, UP_MINB)
k_up_march(const LevelDev f, const double* __restrict__ fb, double* __restrict__ fx, const LevelDev c,
           const double* __restrict__ xc, double w, int mr, PcgScal* scal, double* part, unsigned int* ticket,
           const BShift bs, const HaloWait wt, const HaloPush ps) {
  if (scal->done) return;
  halo_wait(wt);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n1 = (int)f.n1, P = (int)f.pitch, Pc = (int)c.pitch, foff = (int)c.foff;
  const bool shift = DOT == 2 && bs.sum != nullptr;  // only the first V-cycle of a solve can see an unshifted rhs
  const double bm  = shift ? bs.mean() : 0.0;
  MarchCursor cur((n1 + MOUT * MWARPS - 1) / (MOUT * MWARPS), (int)f.i0, (int)f.i1, mr);
  double acc[1] = {0.0};
  int sp, r0, r1;
  while (cur.next(sp, r0, r1)) {
    int lj = sp * (MOUT * MWARPS) + warp * MOUT - 1 + lane;  // lane 0 sits one column left
    const bool store_lane = lane >= 1 && lane <= MOUT && lj < n1;
    const bool colin = lj >= 0 && lj < n1;
    if (lj > n1) lj = n1;  // the zero pad column right of the grid
    double x0p = 0.0, x0c = 0.0, bc = 0.0, dic = 0.0, dc = 0.0, cyc = 0.0, cy1c = 0.0, cxc = 0.0;
    int li = r0 - 1;
    int o  = (li + FS_RO) * P + lj + FS_CO;
    double cxn = __ldg(f.cx + o);
    // operands of one row
    struct Row { double b, d, cy, cy1, cxnn, xm, xp; };
    auto load_row = [&](int li_, int o_) {
      Row q;
      q.cxnn = __ldg(f.cx + o_ + P);
      q.b    = __ldg(fb + o_);
      q.cy   = __ldg(f.cy + o_);
      q.cy1  = __ldg(f.cy + o_ + 1);
      q.d    = __ldg(f.d + o_);
      if (CDIR == 0) {
        const int pos = li_ - foff, par = pos & 1;
        const int oc  = (((pos - par) >> 1) + FS_RO) * Pc + lj + FS_CO;
        q.xm = __ldg(xc + oc);
        q.xp = par ? __ldg(xc + oc + Pc) : 0.0;
      } else {
        const int pos = lj - foff, par = pos & 1;
        const int oc  = (li_ + FS_RO) * Pc + ((pos - par) >> 1) + FS_CO;
        q.xm = __ldg(xc + oc);
        q.xp = __ldg(xc + oc + 1);
      }
      return q;
    };
#pragma unroll 2
    for (; li <= r1; ++li, o += P) {
      const Row q = load_row(li, o);
      const double cxnn = q.cxnn, dn = q.d, cyn = q.cy, cy1n = q.cy1;
      const double bn = (DOT == 2 && shift && colin && li >= 0 && li < (int)f.n0) ? q.b - bm : q.b;
      const double din = rcp0(dn);
      double xi;
      if (CDIR == 0) {
        const int par = (li - foff) & 1;
        if (par == 0) {  // C row (warp uniform)
          xi = q.xm;
        } else {
          xi = mg_interp(cxn, cxnn, rcp0(dn + cyn + cy1n), q.xm, q.xp);
        }
      } else {
        const int par = (lj - foff) & 1;
        const double v = mg_interp(cyn, cy1n, rcp0(dn + cxn + cxnn), q.xm, q.xp);
        xi = par ? v : q.xm;
      }
      const double x0n = mg_x0(w, bn, din, xi);
      if (li > r0) {  // emit row li - 1
        const double xl = __shfl_up_sync(0xffffffffu, x0c, 1), xr = __shfl_down_sync(0xffffffffu, x0c, 1);
        const double xo = mg_post(w, bc, dc, dic, x0c, cxc, x0p, cxn, x0n, cyc, xl, cy1c, xr);
        if (store_lane) {
          fx[o - P] = xo;
          if (DOT) acc[0] += bc * xo;
        }
        if (PUSH) {
          double* pr = halo_push_row(ps, li - 1);
          if (pr && store_lane) pr[o - P] = xo;
        }
      }
      x0p = x0c; x0c = x0n; bc = bn; dic = din; dc = dn; cyc = cyn; cy1c = cy1n; cxc = cxn; cxn = cxnn;
    }
  }
  if (PUSH) halo_push_done(ps);
  if (DOT) {
    double tot[1];
    if (grid_reduce<ReduceSum, 1>(acc, part, ticket, tot)) {
      if (threadIdx.x == 0) pcg_local_rz<DOT == 2>(scal, tot[0]);
    }
  }
}

// ---- the same kernel fed by the TMA engine ---------------------------------------------------------
// k_up_march keeps its loads in registers, so the bytes a warp has in flight are bounded by its
// register budget (ncu: long-scoreboard stalls, DRAM 61 % busy at 30 resident warps per SM).  Here the
// operand rows of a CTA's 120-column strip are streamed into a shared-memory ring by bulk asynchronous
// copies (cp.async.bulk -> UBLKCP: one instruction per 1 KB row segment, issued by ONE thread, no
// register staging), FS_RING_STAGES rows deep, behind one full / one empty mbarrier per stage.  The four
// warps consume a row as soon as its bytes have landed and keep the x0 window in registers exactly
// as above (same arithmetic, same association: results are bit-identical).  Segments start at an even
// physical column (16-byte aligned source) two columns left of the strip's first output column.
#ifndef FS_RING_STAGES
#define FS_RING_STAGES (FS_UP_WARPS >= 8 ? 3 : 4)
#endif
// doubles per operand row segment (the strip + its halo columns, a multiple of 16 bytes) and per coarse
// segment of a j-coarsened level (half the strip + slack); segments per stage
constexpr int RING_SEG = 32 * MWARPS, RING_CSEG = RING_SEG / 2 + 8, RING_SLOTS = 6;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* b, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* b, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* b, unsigned parity) {
  unsigned ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(double* dst, const double* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// the barriers of the ring (one instance per CTA): `full` is armed by the issuing thread with the bytes of
// the row, `empty` collects one arrival per consumer warp.
// Measured alternatives (8192^2, level 0, profiles/r02_ring_variants.txt): 3, 4 and 6 stages run alike;
// sharing the copies out over the four warps (lane 0 of each announcing and issuing its own segment) was
// 16 % slower than one issuing thread; b and d by LDG beside a four-array ring 48 % slower; the restriction
// kernels gained nothing from the ring (507 vs 495 us) and keep their register pipeline.
template <int NS>
struct RingBars {
  unsigned long long full[NS], empty[NS];
  __device__ void init() {
    if (threadIdx.x == 0) {
      for (int q = 0; q < NS; ++q) { mbar_init(&full[q], 1); mbar_init(&empty[q], MWARPS); }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
  }
};

template <int CDIR, int DOT, bool PUSH>
__global__ void __launch_bounds__(32 * MWARPS, UP_MINB)
k_up_ring(const LevelDev f, const double* __restrict__ fb, double* __restrict__ fx, const LevelDev c,
          const double* __restrict__ xc, double w, int mr, PcgScal* scal, double* part, unsigned int* ticket,
          const BShift bs, const HaloWait wt, const HaloPush ps) {
  if (scal->done) return;
  halo_wait(wt);
  constexpr int S = FS_RING_STAGES;
  const bool shift = DOT == 2 && bs.sum != nullptr;
  const double bm  = shift ? bs.mean() : 0.0;
  __shared__ __align__(128) double ring[S][RING_SLOTS][RING_SEG];
  __shared__ __align__(8) RingBars<S> bar;
  unsigned long long* const full = bar.full;
  unsigned long long* const empty = bar.empty;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n1 = (int)f.n1, foff = (int)c.foff;
  const int64_t P = f.pitch, Pc = c.pitch;
  const int strips = (n1 + MOUT * MWARPS - 1) / (MOUT * MWARPS);
  bar.init();
  // producer (thread 0): the CTA's rows form one sequence across its tasks.  Segments of a row:
  // 0 cx (of the row below), 1 b, 2 cy, 3 d, 4 / 5 coarse x (5: odd rows of CDIR 0 only)
  MarchCursor pcur(strips, (int)f.i0, (int)f.i1, mr);
  int psp = 0, pr0 = 0, pr1 = 0, pq = 0;
  bool pmore = pcur.next(psp, pr0, pr1);
  unsigned pn = 0;
  auto issue = [&]() {
    if (!pmore) return;
    const int li = pr0 - 1 + pq;
    const int st = (int)(pn % S);
    mbar_wait(&empty[st], ((pn / S) & 1u) ^ 1u);  // the consumers are done with the previous tenant
    const int col0 = psp * (MOUT * MWARPS) + 2;   // physical column (even): logical column of lane 0 minus 1
    const int64_t o = (int64_t)(li + FS_RO) * P + col0;
    constexpr unsigned SB = RING_SEG * sizeof(double);
    if (CDIR == 0) {
      const int pos = li - foff, par = pos & 1;
      const int64_t oc = (int64_t)(((pos - par) >> 1) + FS_RO) * Pc + col0;
      mbar_expect_tx(&full[st], (par ? 6u : 5u) * SB);
      bulk_g2s(ring[st][4], xc + oc, SB, &full[st]);
      if (par) bulk_g2s(ring[st][5], xc + oc + Pc, SB, &full[st]);
    } else {
      constexpr unsigned CB = RING_CSEG * sizeof(double);
      const int cj0 = ((psp * (MOUT * MWARPS) - 2 - foff) >> 1) & ~1;  // even, <= the leftmost coarse column used
      mbar_expect_tx(&full[st], 4u * SB + CB);
      bulk_g2s(ring[st][4], xc + (int64_t)(li + FS_RO) * Pc + cj0 + FS_CO, CB, &full[st]);
    }
    bulk_g2s(ring[st][0], f.cx + o + P, SB, &full[st]);
    bulk_g2s(ring[st][1], fb + o, SB, &full[st]);
    bulk_g2s(ring[st][2], f.cy + o, SB, &full[st]);
    bulk_g2s(ring[st][3], f.d + o, SB, &full[st]);
    ++pn;
    if (++pq == pr1 - pr0 + 2) { pq = 0; pmore = pcur.next(psp, pr0, pr1); }
  };
  if (threadIdx.x == 0)
    for (int q = 0; q < S - 1; ++q) issue();
  unsigned cn = 0;
  double acc[1] = {0.0};
  MarchCursor cur(strips, (int)f.i0, (int)f.i1, mr);
  int sp, r0, r1;
  while (cur.next(sp, r0, r1)) {
    const int ljr = sp * (MOUT * MWARPS) + warp * MOUT - 1 + lane;  // lane 0 sits one column left
    const bool store_lane = lane >= 1 && lane <= MOUT && ljr < n1;
    const bool colok = ljr <= n1;       // lanes right of the zero pad column compute with zeros
    const int lj = colok ? ljr : n1;
    const int ti = warp * MOUT + lane + 1;  // my column inside the segments
    double x0p = 0.0, x0c = 0.0, bc = 0.0, dic = 0.0, dc = 0.0, cyc = 0.0, cy1c = 0.0, cxc = 0.0;
    int li = r0 - 1;
    int64_t o = (int64_t)(li + FS_RO) * P + lj + FS_CO;
    double cxn = __ldg(f.cx + o);
    int tc = 0;
    if (CDIR == 1) {
      const int cj0 = ((sp * (MOUT * MWARPS) - 2 - foff) >> 1) & ~1;
      const int pos = ljr - foff;
      tc = ((pos - (pos & 1)) >> 1) - cj0;
    }
    for (; li <= r1; ++li, o += P) {
      if (threadIdx.x == 0) issue();
      const int st = (int)(cn % S);
      mbar_wait(&full[st], (cn / S) & 1u);
      double cxnn = ring[st][0][ti], bn = ring[st][1][ti], cyn = ring[st][2][ti], cy1n = ring[st][2][ti + 1],
             dn = ring[st][3][ti], xm, xp = 0.0;
      const int par = CDIR == 0 ? ((li - foff) & 1) : ((ljr - foff) & 1);
      if (CDIR == 0) {
        xm = ring[st][4][ti];
        if (par) xp = ring[st][5][ti];
      } else {
        xm = ring[st][4][tc];
        xp = ring[st][4][tc + 1];
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[st]);
      ++cn;
      if (!colok) { cxnn = 0.0; bn = 0.0; cyn = 0.0; cy1n = 0.0; dn = 0.0; xm = 0.0; xp = 0.0; }
      if (DOT == 2 && shift && ljr >= 0 && ljr < n1 && li >= 0 && li < (int)f.n0) bn -= bm;
      const double din = rcp0(dn);
      double xi;
      if (CDIR == 0) {
        if (par == 0) {  // C row (warp uniform)
          xi = xm;
        } else {
          xi = mg_interp(cxn, cxnn, rcp0(dn + cyn + cy1n), xm, xp);
        }
      } else {
        const double v = mg_interp(cyn, cy1n, rcp0(dn + cxn + cxnn), xm, xp);
        xi = par ? v : xm;
      }
      const double x0n = mg_x0(w, bn, din, xi);
      if (li > r0) {  // emit row li - 1
        const double xl = __shfl_up_sync(0xffffffffu, x0c, 1), xr = __shfl_down_sync(0xffffffffu, x0c, 1);
        const double xo = mg_post(w, bc, dc, dic, x0c, cxc, x0p, cxn, x0n, cyc, xl, cy1c, xr);
        if (store_lane) {
          fx[o - P] = xo;
          if (DOT) acc[0] += bc * xo;
        }
        if (PUSH) {
          double* pr = halo_push_row(ps, li - 1);
          if (pr && store_lane) pr[o - P] = xo;
        }
      }
      x0p = x0c; x0c = x0n; bc = bn; dic = din; dc = dn; cyc = cyn; cy1c = cy1n; cxc = cxn; cxn = cxnn;
    }
  }
  if (PUSH) halo_push_done(ps);
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

// Rows per marching task.  Bandwidth-bound levels: the height in [40, 96] whose task count fills the
// rounds of the round-robin deal best (tasks / (rounds x grid)), discounted by the two rows a task re-reads
// around its range.  Levels with fewer tasks than a few rounds: smaller tasks (they are latency bound; a
// task's rows are a serial chain of loads), down to one row while the level still fits one wave.
static int pick_rows(int64_t rows, int64_t strips, int minb = FS_MARCH_MINB) {
  const int64_t G = (int64_t)FS_NUM_SMS * minb;
  if (strips * cdiv(rows, 64) >= 2 * G) {
    int best = 64;
    double best_score = 0.0;
    for (int mr = 96; mr >= 40; --mr) {
      const int64_t total = strips * cdiv(rows, mr);
      const double score = (double)total / (double)(cdiv(total, G) * G) * mr / (mr + 2.0);
      if (score > best_score + 1e-9) { best_score = score; best = mr; }
    }
    return best;
  }
  int mr = 64;
  while (mr > 4 && strips * cdiv(rows, mr) < 2 * G) mr >>= 1;
  while (mr > 1 && strips * cdiv(rows, mr) < G / 4) mr >>= 1;
  return mr;
}
static unsigned march_blocks(int64_t rows, int64_t strips, int mr, int minb = FS_MARCH_MINB) {
  const int64_t total = strips * cdiv(rows, mr);
  const int64_t cap = (int64_t)FS_NUM_SMS * minb;
  return (unsigned)(total < cap ? (total > 0 ? total : 1) : cap);
}

// levels large enough to be bandwidth bound take the TMA-fed kernels (fsgpu_ps::ring_min_points, read from
// FSGPU_RING / FSGPU_RING_MIN_POINTS when the solver is created: threshold on the points a launch sweeps)
static bool ring_level(const fsgpu_ps* s, int64_t rows, int64_t n1) {
  return s->ring_min_points > 0 && rows * n1 >= s->ring_min_points;
}
// FS_MARCH_MINB CTAs with a 16-24 KB ring each only fit with the shared-memory carve-out at its maximum
static bool ring_carveout(const void* kernel) {
  cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  return true;
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
  const BShift bs{l == 0 ? s->b_shift : nullptr, s->b_shift_size};  // only level 0 reads the rhs itself
  // fused push (p2p.cuh): this kernel reads the halo rows of b_l (pushed by the neighbours' previous down kernel)
  // and stores the boundary rows of b_{l+1} into the neighbours' halos as well
  HaloWait wt{};
  HaloPush ps{};
  if (s->push_ready && r0 < 0) {
    if (l >= 1 && l < s->ld) {
      wt.on = 1; wt.epoch = s->push_epoch;
      comm_push_flags(s->c, l, 0, nullptr, &wt);
    }
    if (l + 1 < s->ld) {
      ps.on = 1; ps.h = FS_RO; ps.i0 = L[l + 1].own0; ps.i1 = L[l + 1].own1; ps.epoch = s->push_epoch;
      ps.dst[0] = s->nb_b[l + 1][0]; ps.dst[1] = s->nb_b[l + 1][1];
      comm_push_flags(s->c, l + 1, 0, &ps, nullptr);
    }
  }
#define DOWN_ARGS(MR) f, fb, c, c.b, L[l].weight, MR, s->scal, bs, wt, ps
#define DOWN_LAUNCH(K, MR)                                                                              \
  do {                                                                                                  \
    const bool a_ = L[l].active != 0, h_ = bs.sum != nullptr, p_ = ps.on != 0;                          \
    if (a_ && h_ && p_) K<true, true, true><<<nb, 128, 0, st>>>(DOWN_ARGS(MR));                         \
    else if (a_ && h_) K<true, true, false><<<nb, 128, 0, st>>>(DOWN_ARGS(MR));                         \
    else if (a_ && p_) K<true, false, true><<<nb, 128, 0, st>>>(DOWN_ARGS(MR));                         \
    else if (a_) K<true, false, false><<<nb, 128, 0, st>>>(DOWN_ARGS(MR));                              \
    else if (h_ && p_) K<false, true, true><<<nb, 128, 0, st>>>(DOWN_ARGS(MR));                         \
    else if (h_) K<false, true, false><<<nb, 128, 0, st>>>(DOWN_ARGS(MR));                              \
    else if (p_) K<false, false, true><<<nb, 128, 0, st>>>(DOWN_ARGS(MR));                              \
    else K<false, false, false><<<nb, 128, 0, st>>>(DOWN_ARGS(MR));                                     \
  } while (0)
  if (L[l].cdir == 0) {
    const int64_t strips = cdiv(f.n1, 120);
    const int mrc = pick_rows(c.i1 - c.i0, strips);
    const unsigned nb = march_blocks(c.i1 - c.i0, strips, mrc);
    DOWN_LAUNCH(k_down_march0, mrc);
  } else {
    const int64_t strips = cdiv(f.n1, 112);
    const int mr = pick_rows(f.i1 - f.i0, strips);
    const unsigned nb = march_blocks(f.i1 - f.i0, strips, mr);
    DOWN_LAUNCH(k_down_march1, mr);
  }
#undef DOWN_LAUNCH
#undef DOWN_ARGS
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
  // fused push: this kernel reads the halo row of x_{l+1} (pushed by the neighbours' previous up kernel) and
  // stores the boundary rows of x_l into the neighbours' halos as well (levels >= 1: level 0's z needs no halo)
  HaloWait wt{};
  HaloPush ps{};
  if (s->push_ready && r0 < 0) {
    if (l + 1 < s->ld) {
      wt.on = 1; wt.epoch = s->push_epoch;
      comm_push_flags(c, l + 1, 1, nullptr, &wt);
    }
    if (l >= 1 && l < s->ld) {
      ps.on = 1; ps.h = 1; ps.i0 = L[l].own0; ps.i1 = L[l].own1; ps.epoch = s->push_epoch;
      ps.dst[0] = s->nb_x[l][0]; ps.dst[1] = s->nb_x[l][1];
      comm_push_flags(c, l, 1, &ps, nullptr);
    }
  }
  if (!L[l].active) {
    k_up_interp<CDIR><<<persist_blocks(f.i1 - f.i0, f.n1), dim3(BX, BY), 0, st>>>(f, fx, L[l + 1].v, L[l + 1].v.x, s->scal, wt, ps);
    FS_LAUNCHED();
    return 0;
  }
  const int64_t strips = cdiv(f.n1, MOUT * MWARPS);
  const int mr      = pick_rows(f.i1 - f.i0, strips, UP_MINB);
  const unsigned nb = march_blocks(f.i1 - f.i0, strips, mr, UP_MINB);
  const int bl = 32 * MWARPS;
  const BShift bs{l == 0 ? s->b_shift : nullptr, s->b_shift_size};
#define UP_ARGS f, fb, fx, L[l + 1].v, L[l + 1].v.x, L[l].weight, mr, s->scal, c->part, c->ticket, bs, wt, ps
  if (ring_level(s, f.i1 - f.i0, f.n1)) {  // bandwidth-bound levels: operands through the TMA ring
    static const bool carve = ring_carveout((const void*)k_up_ring<CDIR, 0, false>) &&
                              ring_carveout((const void*)k_up_ring<CDIR, 0, true>) &&
                              ring_carveout((const void*)k_up_ring<CDIR, 1, false>) &&
                              ring_carveout((const void*)k_up_ring<CDIR, 2, false>);
    (void)carve;
    if (dot_mode == 1) k_up_ring<CDIR, 1, false><<<nb, bl, 0, st>>>(UP_ARGS);
    else if (dot_mode == 2) k_up_ring<CDIR, 2, false><<<nb, bl, 0, st>>>(UP_ARGS);
    else if (ps.on) k_up_ring<CDIR, 0, true><<<nb, bl, 0, st>>>(UP_ARGS);
    else k_up_ring<CDIR, 0, false><<<nb, bl, 0, st>>>(UP_ARGS);
  } else if (dot_mode == 1) k_up_march<CDIR, 1, false><<<nb, bl, 0, st>>>(UP_ARGS);
  else if (dot_mode == 2) k_up_march<CDIR, 2, false><<<nb, bl, 0, st>>>(UP_ARGS);
  else if (ps.on) k_up_march<CDIR, 0, true><<<nb, bl, 0, st>>>(UP_ARGS);
  else k_up_march<CDIR, 0, false><<<nb, bl, 0, st>>>(UP_ARGS);
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
  if (s->push_ready) s->push_epoch = comm_push_next_epoch(s->c);  // one per V-cycle, the same count on every rank
  for (int l = 0; l <= lbot - 1; ++l) {
    const double* fb = l == 0 ? r : L[l].v.b;
    ProfScope ps_(l == 0 ? PROF_DOWN0 : PROF_DOWN, st);
    const bool gather = dist && l + 1 == s->ld;
    const size_t cn = (size_t)(L[l + 1].v.n0 * L[l + 1].v.pitch);
    double* crow0 = L[l + 1].v.b + FS_RO * L[l + 1].v.pitch;  // row 0 of the replicated level
    if (gather) FS_CUDA(cudaMemsetAsync(crow0, 0, cn * sizeof(double), st));
    if (dist && l + 1 < s->ld && s->push_ready) {  // the kernel pushes its boundary rows itself (p2p.cuh)
      if (launch_down(s, l, fb)) return 1;
      continue;
    }
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
    if (dist && l >= 1 && l < s->ld && s->push_ready) {
      if (up(-1, -1)) return 1;
    } else if (dist && l >= 1 && l < s->ld) {  // x of a distributed level: one halo row for the next finer level
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
