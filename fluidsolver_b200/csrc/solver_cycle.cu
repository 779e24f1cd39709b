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

#include <utility>

namespace {

constexpr int BX = 64, BY = 4;

template <int CDIR>
struct Coef {
  const double *cx, *cy;
  int64_t P;
  __device__ __forceinline__ double am(int64_t q) const { return CDIR == 0 ? __ldg(cx + q) : __ldg(cy + q); }
  __device__ __forceinline__ double ap(int64_t q) const { return CDIR == 0 ? __ldg(cx + q + P) : __ldg(cy + q + 1); }
};

// residual after the zero-guess Jacobi sweep at fine point q (b is the level's rhs)
template <bool ACTIVE>
__device__ __forceinline__ double resid0(const LevelDev& f, const double* __restrict__ b, int64_t q,
                                         double w) {
  const double bq = __ldg(b + q);
  if (!ACTIVE) return bq;
  const int64_t P = f.pitch;
  auto X = [&](int64_t r) { return w * __ldg(b + r) * __ldg(f.dinv + r); };
  double v = w * bq;  // d * x(q)
  v += __ldg(f.cx + q) * X(q - P);
  v += __ldg(f.cx + q + P) * X(q + P);
  v += __ldg(f.cy + q) * X(q - 1);
  v += __ldg(f.cy + q + 1) * X(q + 1);
  return bq - v;
}

template <int CDIR, bool ACTIVE>
__global__ void __launch_bounds__(BX* BY)
k_down(const LevelDev f, const double* __restrict__ fb, const LevelDev c, double* __restrict__ cb,
       double w, const PcgScal* scal) {
  if (scal->done) return;
  const int64_t cj = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t ci = (int64_t)blockIdx.y * BY + threadIdx.y;
  if (ci >= c.n0 || cj >= c.n1) return;
  const int64_t fi = CDIR == 0 ? 2 * ci + c.foff : ci;
  const int64_t fj = CDIR == 1 ? 2 * cj + c.foff : cj;
  const int64_t k = f.idx(fi, fj), sd = CDIR == 0 ? f.pitch : 1;
  const Coef<CDIR> C{f.cx, f.cy, f.pitch};
  const double pp_m = -C.ap(k - sd) * __ldg(f.pc + k - sd);  // F-point below, weight to this C-point
  const double pm_p = -C.am(k + sd) * __ldg(f.pc + k + sd);  // F-point above, weight to this C-point
  double v = resid0<ACTIVE>(f, fb, k, w);
  v += pp_m * resid0<ACTIVE>(f, fb, k - sd, w);
  v += pm_p * resid0<ACTIVE>(f, fb, k + sd, w);
  cb[c.idx(ci, cj)] = v;
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

// inactive level: interpolation only (no relaxation on this level)
template <int CDIR>
__global__ void __launch_bounds__(BX* BY)
k_up_interp(const LevelDev f, double* __restrict__ fx, const LevelDev c,
            const double* __restrict__ cxv, const PcgScal* scal) {
  if (scal->done) return;
  const int64_t lj = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t li = (int64_t)blockIdx.y * BY + threadIdx.y;
  if (li >= f.n0 || lj >= f.n1) return;
  const int64_t k = f.idx(li, lj);
  fx[k] = x_interp<CDIR, false>(f, 0.0, c, cxv, li, lj, k, 0.0);
}

// active level: x0 = w b / d + P x_c is formed ONCE per point in a shared-memory tile
// (32 rows x 64 columns + a one-point ring), then x = x0 + w (b - A x0) / d reads its four
// neighbours from the tile.  A CTA walks tiles with a grid stride (persistent), so the optional
// dot product <b, x> (= <r, z> of PCG on level 0) ends in ~600 partials.
// DOT: 0 none, 1 PCG iteration (convergence test + beta), 2 first V-cycle (bi).
constexpr int UT_I = 32, UT_J = 64, UT_Y = 4, UT_SW = UT_J + 2, UT_PER = UT_I / UT_Y;

template <int CDIR, int DOT>
__global__ void __launch_bounds__(UT_J* UT_Y, 3)
k_up_tiled(const LevelDev f, const double* __restrict__ fb, double* __restrict__ fx, const LevelDev c,
           const double* __restrict__ cxv, double w, PcgScal* scal, double* part,
           unsigned int* ticket) {
  if (scal->done) return;
  __shared__ double sx[(UT_I + 2) * UT_SW];
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * UT_J + tx;
  const int64_t tiles_j = (f.n1 + UT_J - 1) / UT_J, nt = tiles_j * ((f.n0 + UT_I - 1) / UT_I);
  const int64_t P = f.pitch;
  double acc[1] = {0.0};
  for (int64_t t = blockIdx.x; t < nt; t += gridDim.x) {
    const int64_t ti = t / tiles_j, tj = t - ti * tiles_j;
    const int64_t i0 = ti * UT_I, j0 = tj * UT_J;
    const int64_t lj = j0 + tx;
    double bv[UT_PER], x0v[UT_PER];
    // phase 1a: this thread's own points
#pragma unroll
    for (int m = 0; m < UT_PER; ++m) {
      const int r = ty + UT_Y * m;
      const int64_t li = i0 + r;
      double b = 0.0, x0 = 0.0;
      if (li < f.n0 && lj < f.n1) {
        const int64_t k = f.idx(li, lj);
        b  = __ldg(fb + k);
        x0 = x_interp<CDIR, true>(f, b, c, cxv, li, lj, k, w);
      }
      bv[m] = b;
      x0v[m] = x0;
      sx[(r + 1) * UT_SW + tx + 1] = x0;
    }
    // phase 1b: the ring (2 rows of UT_SW + 2 columns of UT_I = 196 points), one per thread
    if (tid < 2 * UT_SW + 2 * UT_I) {
      int r, cc;
      if (tid < UT_SW) { r = -1; cc = tid - 1; }
      else if (tid < 2 * UT_SW) { r = UT_I; cc = tid - UT_SW - 1; }
      else if (tid < 2 * UT_SW + UT_I) { r = tid - 2 * UT_SW; cc = -1; }
      else { r = tid - 2 * UT_SW - UT_I; cc = UT_J; }
      const int64_t li = i0 + r, l2 = j0 + cc;
      double x0 = 0.0;
      if (li >= 0 && li < f.n0 && l2 >= 0 && l2 < f.n1) {
        const int64_t k = f.idx(li, l2);
        x0 = x_interp<CDIR, true>(f, __ldg(fb + k), c, cxv, li, l2, k, w);
      }
      sx[(r + 1) * UT_SW + cc + 1] = x0;
    }
    __syncthreads();
    // phase 2: post-relaxation
#pragma unroll
    for (int m = 0; m < UT_PER; ++m) {
      const int r = ty + UT_Y * m;
      const int64_t li = i0 + r;
      if (li < f.n0 && lj < f.n1) {
        const int64_t k = f.idx(li, lj);
        const int sm = (r + 1) * UT_SW + tx + 1;
        double v = __ldg(f.d + k) * x0v[m];
        v += __ldg(f.cx + k) * sx[sm - UT_SW];
        v += __ldg(f.cx + k + P) * sx[sm + UT_SW];
        v += __ldg(f.cy + k) * sx[sm - 1];
        v += __ldg(f.cy + k + 1) * sx[sm + 1];
        const double xo = x0v[m] + w * (bv[m] - v) * __ldg(f.dinv + k);
        fx[k] = xo;
        if (DOT) acc[0] += bv[m] * xo;
      }
    }
    __syncthreads();  // the tile is overwritten by the next iteration
  }
  if (DOT) {
    double tot[1];
    if (grid_reduce<ReduceSum, 1>(acc, part, ticket, tot)) {
      if (tid == 0) pcg_after_rz<DOT == 2>(scal, tot[0]);
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
  for_each_point(L.n0, L.n1, [&](int64_t li, int64_t lj) {
    const int64_t k = L.idx(li, lj);
    v[0] += r[k] * z[k];
  });
  double tot[1];
  if (grid_reduce<ReduceSum, 1>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0) pcg_after_rz<INIT>(scal, tot[0]);
  }
}

// ---- coarsest level: `sweeps` Jacobi sweeps with weight 1 from a zero guess
// (pfmg_setup.c:400-412 sets the sweep count, pfmg_relax.c / jacobi.c the iteration).
// One CTA; x ping-pongs between two shared-memory tiles with a one-point zero ring.
__global__ void __launch_bounds__(1024)
k_coarse_smem(const LevelDev L, const double* __restrict__ b, double* __restrict__ x, int sweeps,
              int active, const PcgScal* scal) {
  if (scal->done) return;
  extern __shared__ double sm[];
  const int64_t n0 = L.n0, n1 = L.n1, w1 = n1 + 2, tot = (n0 + 2) * w1, n = n0 * n1;
  double* xa = sm;
  double* xb = sm + tot;
  const bool on = active && !scal->coarse_zero_diag;
  for (int64_t t = threadIdx.x; t < 2 * tot; t += blockDim.x) sm[t] = 0.0;
  __syncthreads();
  if (on) {
    for (int64_t t = threadIdx.x; t < n; t += blockDim.x) {
      const int64_t li = t / n1, lj = t - li * n1, k = L.idx(li, lj);
      xa[(li + 1) * w1 + lj + 1] = b[k] * L.dinv[k];
    }
    __syncthreads();
    for (int it = 1; it < sweeps; ++it) {
      for (int64_t t = threadIdx.x; t < n; t += blockDim.x) {
        const int64_t li = t / n1, lj = t - li * n1, k = L.idx(li, lj), P = L.pitch;
        const int64_t m = (li + 1) * w1 + lj + 1;
        double v = L.d[k] * xa[m];
        v += L.cx[k] * xa[m - w1];
        v += L.cx[k + P] * xa[m + w1];
        v += L.cy[k] * xa[m - 1];
        v += L.cy[k + 1] * xa[m + 1];
        xb[m] = xa[m] + (b[k] - v) * L.dinv[k];
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
  const int64_t li = (int64_t)blockIdx.y * BY + threadIdx.y;
  if (li >= L.n0 || lj >= L.n1) return;
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

dim3 lvl_grid(const LevelDev& v) { return dim3((unsigned)cdiv(v.n1, BX), (unsigned)cdiv(v.n0, BY)); }

}  // namespace

#define DISPATCH(KER, cdir, act, grid, ...)                                                    \
  do {                                                                                         \
    if ((cdir) == 0) {                                                                         \
      if (act) KER<0, true><<<grid, dim3(BX, BY), 0, st>>>(__VA_ARGS__);                       \
      else KER<0, false><<<grid, dim3(BX, BY), 0, st>>>(__VA_ARGS__);                          \
    } else {                                                                                   \
      if (act) KER<1, true><<<grid, dim3(BX, BY), 0, st>>>(__VA_ARGS__);                       \
      else KER<1, false><<<grid, dim3(BX, BY), 0, st>>>(__VA_ARGS__);                          \
    }                                                                                          \
    FS_LAUNCHED();                                                                             \
  } while (0)

static int coarse_solve(fsgpu_ps* s, int l, const double* b, double* x) {
  cudaStream_t st = s->c->stream;
  const Level& L  = s->L[l];
  const size_t smem = (size_t)(2 * (L.v.n0 + 2) * (L.v.n1 + 2)) * sizeof(double);
  if (smem <= 200 * 1024) {
    static bool attr_set = false;
    if (!attr_set) {
      FS_CUDA(cudaFuncSetAttribute(k_coarse_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      attr_set = true;
    }
    k_coarse_smem<<<1, 1024, smem, st>>>(L.v, b, x, s->coarse_sweeps, L.active, s->scal);
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

static unsigned up_blocks(const LevelDev& v) {
  const int64_t nt = cdiv(v.n1, UT_J) * cdiv(v.n0, UT_I);
  const int64_t cap = (int64_t)FS_NUM_SMS * 3;
  return (unsigned)(nt < cap ? nt : cap);
}

template <int CDIR>
static int launch_up(fsgpu_ps* s, int l, const double* fb, double* fx, int dot_mode) {
  cudaStream_t st = s->c->stream;
  Level* L        = s->L;
  fsgpu_ctx* c    = s->c;
  if (!L[l].active) {
    k_up_interp<CDIR><<<lvl_grid(L[l].v), dim3(BX, BY), 0, st>>>(L[l].v, fx, L[l + 1].v, L[l + 1].v.x, s->scal);
    FS_LAUNCHED();
    return 0;
  }
  const unsigned nb = up_blocks(L[l].v);
  const dim3 bl(UT_J, UT_Y);
#define UP_ARGS L[l].v, fb, fx, L[l + 1].v, L[l + 1].v.x, L[l].weight, s->scal, c->part, c->ticket
  if (dot_mode == 1) k_up_tiled<CDIR, 1><<<nb, bl, 0, st>>>(UP_ARGS);
  else if (dot_mode == 2) k_up_tiled<CDIR, 2><<<nb, bl, 0, st>>>(UP_ARGS);
  else k_up_tiled<CDIR, 0><<<nb, bl, 0, st>>>(UP_ARGS);
#undef UP_ARGS
  FS_LAUNCHED();
  return 0;
}

int ps_vcycle(fsgpu_ps* s, const double* r, double* z, int dot_mode) {
  cudaStream_t st = s->c->stream;
  Level* L        = s->L;
  const int nl    = s->nl;
  if (fs_ensure_part(s->c, (size_t)FS_NUM_SMS * 8 * 4)) return 1;
  if (nl == 1) {  // relax from zero, relax again (pfmg_solve.c with a single level)
    k_jacobi<true><<<lvl_grid(L[0].v), dim3(BX, BY), 0, st>>>(L[0].v, r, nullptr, s->ctmp, L[0].weight, 1, s->scal);
    FS_LAUNCHED();
    k_jacobi<false><<<lvl_grid(L[0].v), dim3(BX, BY), 0, st>>>(L[0].v, r, s->ctmp, z, L[0].weight, 1, s->scal);
    FS_LAUNCHED();
    const unsigned pg = persist_blocks(L[0].v.n0, L[0].v.n1);
    if (dot_mode == 1) k_dot_rz<false><<<pg, dim3(BX, BY), 0, st>>>(L[0].v, r, z, s->scal, s->c->part, s->c->ticket);
    if (dot_mode == 2) k_dot_rz<true><<<pg, dim3(BX, BY), 0, st>>>(L[0].v, r, z, s->scal, s->c->part, s->c->ticket);
    if (dot_mode) FS_LAUNCHED();
    return 0;
  }
  // down: level l rhs -> level l+1 rhs
  for (int l = 0; l <= nl - 2; ++l) {
    const double* fb = l == 0 ? r : L[l].v.b;
    ProfScope ps_(l == 0 ? PROF_DOWN0 : PROF_DOWN, st);
    DISPATCH(k_down, L[l].cdir, L[l].active, lvl_grid(L[l + 1].v), L[l].v, fb, L[l + 1].v,
             L[l + 1].v.b, L[l].weight, s->scal);
  }
  {
    ProfScope ps_(PROF_COARSE, st);
    if (coarse_solve(s, nl - 1, L[nl - 1].v.b, L[nl - 1].v.x)) return 1;
  }
  // up: level l+1 x -> level l x; the level-0 kernel also carries PCG's <r, z>
  for (int l = nl - 2; l >= 0; --l) {
    const double* fb = l == 0 ? r : L[l].v.b;
    double* fx       = l == 0 ? z : L[l].v.x;
    ProfScope ps_(l == 0 ? PROF_UP0 : PROF_UP, st);
    const int dm = l == 0 ? dot_mode : 0;
    if (L[l].cdir == 0) { if (launch_up<0>(s, l, fb, fx, dm)) return 1; }
    else { if (launch_up<1>(s, l, fb, fx, dm)) return 1; }
  }
  return 0;
}
