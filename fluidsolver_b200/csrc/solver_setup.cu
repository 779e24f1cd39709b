// fluidsolver_b200/csrc/solver_setup.cu -- pressure operator + multigrid hierarchy setup.
//
//   k_build_l0   LinearSystem::fill_pressure_operator      src/LinearSystem.hpp:32-122
//                + the per-direction stencil-strength sums of hypre_PFMGComputeDxyz
//                  (Submodules/hypre/src/struct_ls/pfmg_coarsen.c:374-503)
//   schedule()   hypre_PFMGCoarsen                          pfmg_coarsen.c:508-637 (host, O(levels))
//   k_interp     interpolation denominators                 pfmg_setup_interp.c:327-341
//   k_rap<CDIR>  non-Galerkin 5-point coarse operator       pfmg_setup_rap5.c:339-379
#include "solver.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

namespace {

constexpr int BX = 64, BY = 4;

__global__ void __launch_bounds__(BX* BY)
k_build_l0(const double* __restrict__ ru, const double* __restrict__ rv, LevelDev L, const Geom g,
           int psd, double* part, unsigned int* ticket, double* out) {
  double v[4] = {0.0, 0.0, 0.0, 0.0};
  for_each_point(g.a_lo, g.a_hi, g.ny + 2, [&](int64_t a, int64_t b) {
    const int64_t k = g.idx(a, b), P = g.pitch;
    const double vol = g.dx * g.dy, dx2 = g.dx * g.dx, dy2 = g.dy * g.dy;
    double W = 0.0, E = 0.0, S = 0.0, N = 0.0, dxp, dyp;
    // x part, LinearSystem.hpp:41-57
    if (a == 0) {
      dxp = -vol * -1.0 / (dx2 * ru[k + P]);
      E   = -vol * 1.0 / (dx2 * ru[k + P]);
    } else if (a == g.nx + 1) {
      dxp = -vol * -1.0 / (dx2 * ru[k]);
      W   = -vol * 1.0 / (dx2 * ru[k]);
    } else {
      dxp = -vol * (-1.0 / (dx2 * ru[k]) + -1.0 / (dx2 * ru[k + P]));
      W   = -vol * 1.0 / (dx2 * ru[k]);
      E   = -vol * 1.0 / (dx2 * ru[k + P]);
    }
    // y part, LinearSystem.hpp:60-76
    if (b == 0) {
      dyp = -vol * -1.0 / (dy2 * rv[k + 1]);
      N   = -vol * 1.0 / (dy2 * rv[k + 1]);
    } else if (b == g.ny + 1) {
      dyp = -vol * -1.0 / (dy2 * rv[k]);
      S   = -vol * 1.0 / (dy2 * rv[k]);
    } else {
      dyp = -vol * (-1.0 / (dy2 * rv[k]) + -1.0 / (dy2 * rv[k + 1]));
      S   = -vol * 1.0 / (dy2 * rv[k]);
      N   = -vol * 1.0 / (dy2 * rv[k + 1]);
    }
    double d = 0.0;
    d += dxp;
    d += dyp;
    // one-sided Dirichlet (LinearSystem.hpp:79-121): identity row.  The reference leaves the
    // neighbour's coupling to that row in place (a non-symmetric matrix whose extra column only
    // ever multiplies the zero solution of the identity row); here the column is eliminated too,
    // which keeps the operator symmetric for CG and has the same solution.
    const bool me = (psd == FSGPU_PSD_LEFT && a == 0) || (psd == FSGPU_PSD_RIGHT && a == g.nx + 1) ||
                    (psd == FSGPU_PSD_BOTTOM && b == 0) || (psd == FSGPU_PSD_TOP && b == g.ny + 1);
    if (me) {
      d = 1.0;
      W = E = S = N = 0.0;
    } else {
      if (psd == FSGPU_PSD_LEFT && a == 1) W = 0.0;
      if (psd == FSGPU_PSD_RIGHT && a == g.nx) E = 0.0;
      if (psd == FSGPU_PSD_BOTTOM && b == 1) S = 0.0;
      if (psd == FSGPU_PSD_TOP && b == g.ny) N = 0.0;
    }
    L.d[k]    = d;
    L.dinv[k] = 1.0 / d;
    L.cx[k]   = W;
    L.cy[k]   = S;
    // pfmg_coarsen.c:374-503 (cxyz / sqcxyz of pfmg_cxyz_vc.c): sign from the diagonal
    const double sign = d < 0.0 ? 1.0 : -1.0;
    const double w0 = sign * (W + E), w1 = sign * (S + N);
    v[0] += w0; v[1] += w0 * w0; v[2] += w1; v[3] += w1 * w1;
  });
  double tot[4];
  if (grid_reduce<ReduceSum, 4>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0)
      for (int q = 0; q < 4; ++q) out[q] = tot[q];
  }
}

template <int CDIR>
struct Coef {  // accessors "along cdir" (a_m, a_p) and "transverse" (t_m, t_p) in symmetric storage
  const double *cx, *cy;
  int64_t P;
  __device__ __forceinline__ double am(int64_t q) const { return CDIR == 0 ? __ldg(cx + q) : __ldg(cy + q); }
  __device__ __forceinline__ double ap(int64_t q) const { return CDIR == 0 ? __ldg(cx + q + P) : __ldg(cy + q + 1); }
  __device__ __forceinline__ double tm(int64_t q) const { return CDIR == 0 ? __ldg(cy + q) : __ldg(cx + q); }
  __device__ __forceinline__ double tp(int64_t q) const { return CDIR == 0 ? __ldg(cy + q + 1) : __ldg(cx + q + P); }
};

// pc = 1 / (d + t_m + t_p) (0 where that sum is 0), pfmg_setup_interp.c:327-341
template <int CDIR>
__global__ void __launch_bounds__(BX* BY) k_interp(LevelDev f) {
  const int64_t lj = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t li = f.i0 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (li >= f.i1 || lj >= f.n1) return;
  const int64_t k = f.idx(li, lj);
  const Coef<CDIR> C{f.cx, f.cy, f.pitch};
  const double center = f.d[k] + C.tm(k) + C.tp(k);
  f.pc[k] = center != 0.0 ? 1.0 / center : 0.0;
}

// coarse operator, pfmg_setup_rap5.c:339-379 (thread per coarse point)
template <int CDIR>
__global__ void __launch_bounds__(BX* BY) k_rap(LevelDev f, LevelDev c) {
  const int64_t cj = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t ci = c.i0 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (ci >= c.i1 || cj >= c.n1) return;
  const int64_t fi = CDIR == 0 ? 2 * ci + c.foff : ci;
  const int64_t fj = CDIR == 1 ? 2 * cj + c.foff : cj;
  const int64_t k = f.idx(fi, fj), kc = c.idx(ci, cj);
  const int64_t sd = CDIR == 0 ? f.pitch : 1;
  const Coef<CDIR> C{f.cx, f.cy, f.pitch};
  const double pcm = __ldg(f.pc + k - sd), pcp = __ldg(f.pc + k + sd);
  const double pm_m = -C.am(k - sd) * pcm;  // F-point below: weight from ITS minus neighbour
  const double pp_m = -C.ap(k - sd) * pcm;  // F-point below: weight from this C-point
  const double pm_p = -C.am(k + sd) * pcp;  // F-point above: weight from this C-point
  const double am = C.am(k), ap = C.ap(k), tm = C.tm(k), tp = C.tp(k);
  double west = tm + 0.5 * C.tm(k - sd) + 0.5 * C.tm(k + sd);
  double east = tp + 0.5 * C.tp(k - sd) + 0.5 * C.tp(k + sd);
  if (tm == 0.0) west = 0.0;
  if (tp == 0.0) east = 0.0;
  const double cm = am * pm_m;
  double dc = f.d[k] + tm + tp + am * pp_m + ap * pm_p - west - east;
  // A singular (all-Neumann) operator has zero row sums on every level, so on a grid coarsened
  // down to one point the diagonal is 0 in exact arithmetic.  HYPRE tests `== 0.0`
  // (hypre_StructMatrixZeroDiagonal, pfmg_setup.c:349-361) on the ROUNDED value: exact with dyadic
  // coefficients, ~1e-16 otherwise -- which it then divides by in the coarsest relaxation (TGV, IB
  // channel: PCG stalls at 1e-8 or breaks down).  NOT in HYPRE (same guard as the oracle's
  // snap_zero_diag = 1): pure cancellation noise becomes 0, so the test cannot hinge on the last bit
  // of this symmetric-storage, reciprocal-weight setup, which is not bit-identical to HYPRE's and
  // cannot inherit its exact zeros (measured: without the guard the 64^2 cavity breaks down after
  // 8 iterations where HYPRE's arithmetic converges in 12).
  const double scale = fabs(f.d[k]) + fabs(tm) + fabs(tp) + fabs(am * pp_m) + fabs(ap * pm_p) +
                       fabs(west) + fabs(east);
  if (fabs(dc) <= 1e-13 * scale) dc = 0.0;
  (CDIR == 0 ? c.cx : c.cy)[kc] = cm;
  (CDIR == 0 ? c.cy : c.cx)[kc] = west;
  c.d[kc]    = dc;
  c.dinv[kc] = dc != 0.0 ? 1.0 / dc : 0.0;
}

// zero diagonal on the coarsest grid => no relaxation there, pfmg_setup.c:349-357 (its own pass
// so that it sees the whole gathered level when the hierarchy is decomposed)
__global__ void __launch_bounds__(BX* BY) k_zero_diag(LevelDev c, PcgScal* scal) {
  const int64_t cj = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t ci = c.i0 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (ci >= c.i1 || cj >= c.n1) return;
  if (c.d[c.idx(ci, cj)] == 0.0) scal->coarse_zero_diag = 1;
}

int64_t ceil_div2(int64_t a) { return a >= 0 ? (a + 1) / 2 : -((-a) / 2); }
int64_t floor_div2(int64_t a) { return a >= 0 ? a / 2 : -((-a + 1) / 2); }
int hlog2(int64_t p) {
  int e = 0;
  if (p <= 0) return 0;
  while (p > 1) { e++; p /= 2; }
  return e;
}
int64_t round_up(int64_t a, int64_t m) { return (a + m - 1) / m * m; }

}  // namespace

// hypre_PFMGComputeDxyz tail + hypre_PFMGCoarsen on the host (pfmg_coarsen.c:437-637)
static void schedule(fsgpu_ps* s, const double sums[4]) {
  const Geom& g = s->c->g;
  const int64_t n0 = g.nx + 2, n1 = g.ny + 2;
  int max_levels = 1 + (hlog2(n0) + 2) + (hlog2(n1) + 2);  // hypre_PFMGComputeMaxLevels
  const int want = s->o.max_levels > 0 ? s->o.max_levels : 16;
  if (want < max_levels) max_levels = want;
  if (max_levels > FS_MAX_LEVELS) max_levels = FS_MAX_LEVELS;

  double cxyz[2] = {sums[0], sums[2]}, sq[2] = {sums[1], sums[3]}, dxyz[2];
  const double gs = (double)(n0 * n1);
  double mean[2] = {cxyz[0] / gs, cxyz[1] / gs}, dev[2] = {sq[0] / gs, sq[1] / gs};
  double cmax = cxyz[0] > cxyz[1] ? cxyz[0] : cxyz[1];
  if (cmax < 0.0) cmax = 0.0;
  if (cmax == 0.0) { cxyz[0] = cxyz[1] = 1.0; cmax = 1.0; }
  for (int d = 0; d < 2; ++d) {
    const double max_aniso = 1.79769313486231570e+308 / 1000;
    if (cxyz[d] > cmax / max_aniso) dxyz[d] = sqrt(1.0 / (cxyz[d] / cmax));
    else dxyz[d] = sqrt(max_aniso);
  }
  int dxyz_flag = 0;
  for (int d = 0; d < 2; ++d) {
    dev[d] -= mean[d] * mean[d];
    if (dev[d] > 0.1 * (mean[d] * mean[d])) { dxyz_flag = 1; break; }
  }

  Level* L = s->L;
  int coarsen[2] = {1, 1};
  int64_t lo[2] = {-1, -1}, hi[2] = {g.nx, g.ny};
  L[0].v.n0 = n0; L[0].v.n1 = n1; L[0].lo[0] = L[0].lo[1] = -1; L[0].v.foff = 0;
  int l;
  for (l = 0; l < max_levels; ++l) {
    double min_dxyz = 1 + dxyz[0] + dxyz[1], alpha = 0.0, beta = 0.0;
    int cdir = -1;
    for (int d = 0; d < 2; ++d) {
      if (hi[d] > lo[d] && dxyz[d] < min_dxyz) { min_dxyz = dxyz[d]; cdir = d; }
      alpha += 1.0 / (dxyz[d] * dxyz[d]);
    }
    L[l].weight = 1.0;
    if (cdir != -1) {
      if (dxyz_flag) L[l].weight = 2.0 / 3.0;
      else {
        for (int d = 0; d < 2; ++d)
          if (d != cdir) beta += 1.0 / (dxyz[d] * dxyz[d]);
        L[l].weight = 2.0 / (3.0 - beta / alpha);
      }
      if (l == max_levels - 1) cdir = -1;
      /* Hierarchies deeper than the reference's cap of 16 (src/LinearSolver_StructHypre.hpp:221) stop at a
       * two-point coarsest grid instead of a single point.  The one-point level of an all-Neumann system is
       * the zero operator; on the anisotropic channel grid (4096 x 1024 on 100 x 1) carrying it made the
       * V-cycle indefinite and PCG broke down after ~4 iterations at a 1e-2 residual, while one level
       * fewer converges like the capped hierarchy (and the cavity's iteration counts do not change). */
      if (cdir != -1 && want > 16) {
        int64_t nc[2] = {hi[0] - lo[0] + 1, hi[1] - lo[1] + 1};
        nc[cdir] = floor_div2(hi[cdir]) - ceil_div2(lo[cdir]) + 1;
        if (nc[0] * nc[1] <= 1) cdir = -1;
      }
    }
    if (cdir == -1) { L[l].active = 1; L[l].cdir = -1; break; }
    L[l].cdir = cdir;
    if (coarsen[cdir] != 0) { L[l].active = 1; coarsen[0] = coarsen[1] = 0; }
    else L[l].active = 0;
    coarsen[cdir] = 1;
    dxyz[cdir] *= 2;
    const int64_t lof = lo[cdir];
    lo[cdir] = ceil_div2(lo[cdir]);
    hi[cdir] = floor_div2(hi[cdir]);
    L[l + 1].v.n0 = hi[0] - lo[0] + 1; L[l + 1].v.n1 = hi[1] - lo[1] + 1;
    L[l + 1].lo[0] = lo[0]; L[l + 1].lo[1] = lo[1];
    L[l + 1].v.foff = 2 * lo[cdir] - lof;
  }
  s->nl = l + 1;
  const int skip = s->o.skip_relax < 0 ? 1 : s->o.skip_relax;
  if (!skip)
    for (l = 0; l < s->nl; ++l) L[l].active = 1;
  const Level& c = L[s->nl - 1];
  const int64_t cmaxsize = c.v.n0 > c.v.n1 ? c.v.n0 : c.v.n1;
  s->coarse_sweeps = 4 * s->nl < cmaxsize ? 4 * s->nl : (int)cmaxsize;  // pfmg_setup.c:400-412
}

// Slab decomposition of the hierarchy.  Every level's rows are split by
// following the fine partition down: a coarse row belongs to the rank that owns the fine row it
// sits on.  Levels stay distributed while every rank keeps >= 2 FS_RO rows and the level has
// more than `repl_points` points; from level `ld` on the (small) levels are gathered and every
// rank runs the rest of the V-cycle redundantly on the whole level -- no exchanges there, and
// no rank is ever left with fewer rows than a halo is wide.
static int partition(fsgpu_ps* s) {
  const int R = s->c->nranks, me = s->c->rank;
  Level* L = s->L;
  if (R <= 1) {
    for (int l = 0; l < s->nl; ++l) {
      L[l].own0 = 0; L[l].own1 = L[l].v.n0; L[l].replicated = false;
      L[l].v.i0 = 0; L[l].v.i1 = L[l].v.n0;
    }
    s->ld = s->nl;
    return 0;
  }
  // 2^18: at 16384^2 on 8 GPUs 103.6 -> 101.8 ms per step against 2^16, 2^20 no better (profiles/r02_multigpu.txt)
  int64_t repl_points = 1 << 18;
  if (const char* e = getenv("FSGPU_REPL_POINTS")) repl_points = atoll(e);
  std::vector<int64_t> lo(R), hi(R);
  for (int r = 0; r < R; ++r)
    if (fsgpu_slab_range(s->c->g.nx, r, R, &lo[r], &hi[r])) return 1;
  int ld = -1;
  for (int l = 0; l < s->nl; ++l) {
    int64_t min_rows = hi[0] - lo[0];
    for (int r = 1; r < R; ++r) min_rows = hi[r] - lo[r] < min_rows ? hi[r] - lo[r] : min_rows;
    const bool keep = min_rows >= 2 * FS_RO && L[l].v.n0 * L[l].v.n1 > repl_points;
    if (ld < 0 && l >= 1 && !keep) ld = l;
    L[l].own0 = lo[me]; L[l].own1 = hi[me];
    if (l + 1 < s->nl && L[l].cdir == 0) {
      const int64_t foff = L[l + 1].v.foff;
      for (int r = 0; r < R; ++r) fsgpu_slab_coarse_range(lo[r], hi[r], foff, &lo[r], &hi[r]);
    }
  }
  if (ld < 0) ld = s->nl;
  if (s->nl >= 2 && ld > s->nl - 1) ld = s->nl - 1;  // the coarsest level is always replicated
  s->ld = ld;
  for (int l = 0; l < s->nl; ++l) {
    L[l].replicated = l >= ld;
    L[l].v.i0 = L[l].replicated ? 0 : L[l].own0;
    L[l].v.i1 = L[l].replicated ? L[l].v.n0 : L[l].own1;
  }
  return 0;
}

// Fused push (p2p.cuh): map the neighbours' slabs and learn where their b and x arrays of every distributed
// level live, so that the marching kernels can store boundary rows straight into the neighbours' halo rows.
// Collective; every (re)allocation of the levels repeats it.  Leaves push_ready false when the mailbox
// transport is off or a rank cannot map a neighbour (the exchanges then run as separate kernels).
static int map_neighbour_levels(fsgpu_ps* s) {
  fsgpu_ctx* c = s->c;
  s->push_ready = false;
  if (s->nb_token) { comm_unmap_neighbours(c, s->nb_token); s->nb_token = nullptr; }
  if (c->nranks <= 1 || !comm_push_enabled(c) || s->ld > P2P_PUSH_LEVELS) return 0;
  void *up = nullptr, *dn = nullptr;
  if (comm_map_neighbours(c, s->slab, &up, &dn, &s->nb_token)) return 1;
  if (!s->nb_token) return 0;
  // offsets (in doubles, signed) of the virtual bases of b and x relative to each rank's slab
  struct Rec { long long off_b[FS_MAX_LEVELS], off_x[FS_MAX_LEVELS]; };
  Rec mine{};
  for (int l = 1; l < s->ld; ++l) {
    mine.off_b[l] = (long long)(s->L[l].v.b - s->slab);
    mine.off_x[l] = (long long)(s->L[l].v.x - s->slab);
  }
  std::vector<Rec> all((size_t)c->nranks);
  if (comm_allgather_host(c, &mine, sizeof(Rec), all.data())) return 1;
  void* nbp[2] = {up, dn};
  const int peer[2] = {c->rank - 1, c->rank + 1 < c->nranks ? c->rank + 1 : -1};
  for (int l = 0; l < FS_MAX_LEVELS; ++l)
    for (int sd = 0; sd < 2; ++sd) {
      const bool on = l >= 1 && l < s->ld && peer[sd] >= 0 && nbp[sd];
      s->nb_b[l][sd] = on ? static_cast<double*>(nbp[sd]) + all[(size_t)peer[sd]].off_b[l] : nullptr;
      s->nb_x[l][sd] = on ? static_cast<double*>(nbp[sd]) + all[(size_t)peer[sd]].off_x[l] : nullptr;
    }
  s->push_ready = true;
  return 0;
}

// (re)allocate the level arrays for the current schedule; returns 0 on success
static int alloc_levels(fsgpu_ps* s) {
  const Geom& g = s->c->g;
  Level* L = s->L;
  // level 0 shares the flow-field geometry
  L[0].v.pitch = g.pitch; L[0].rows = g.rows;
  size_t total = 0;
  const size_t n_l0 = (size_t)g.count();
  total += n_l0 * (5 + 3);  // d, dinv, cx, cy, pc + p, s, z
  for (int l = 1; l < s->nl; ++l) {
    L[l].v.pitch = round_up(L[l].v.n1 + FS_CO + 2, 16);
    L[l].rows    = (L[l].v.i1 - L[l].v.i0) + 2 * FS_RO;  // owned (or all) rows + pads / halos
    total += (size_t)(L[l].rows * L[l].v.pitch) * 7;  // + x, b
  }
  const Level& c = L[s->nl - 1];
  total += (size_t)(c.rows * c.v.pitch);  // coarsest ping-pong
  total += 512;  // slack: the 1 KB bulk-copy segments of a level's last strip may run past the last row
  if (total > s->slab_doubles) {
    if (s->slab) cudaFree(s->slab);
    s->slab = nullptr;
    s->slab_doubles = 0;
    FS_CUDA(cudaMalloc(&s->slab, total * sizeof(double)));
    s->slab_doubles = total;
  }
  FS_CUDA(cudaMemsetAsync(s->slab, 0, s->slab_doubles * sizeof(double), s->c->stream));
  double* q = s->slab;
  int cur = 0;  // arrays are handed out as VIRTUAL bases of level `cur`
  auto take = [&](size_t n) { double* r = q - L[cur].voff(); q += n; return r; };
  L[0].v.d = take(n_l0); L[0].v.dinv = take(n_l0); L[0].v.cx = take(n_l0); L[0].v.cy = take(n_l0);
  L[0].v.pc = take(n_l0);
  s->p = take(n_l0); s->s = take(n_l0); s->z = take(n_l0);
  L[0].v.x = nullptr; L[0].v.b = nullptr;  // bound per V-cycle call
  for (int l = 1; l < s->nl; ++l) {
    const size_t n = (size_t)(L[l].rows * L[l].v.pitch);
    cur = l;
    L[l].v.d = take(n); L[l].v.dinv = take(n); L[l].v.cx = take(n); L[l].v.cy = take(n);
    L[l].v.pc = take(n); L[l].v.x = take(n); L[l].v.b = take(n);
  }
  cur = s->nl - 1;
  s->ctmp = take((size_t)(c.rows * c.v.pitch));
  if (comm_alloc_fence(s->c)) return 1;
  return map_neighbour_levels(s);
}

static dim3 lvl_grid(const LevelDev& v) {
  return dim3((unsigned)cdiv(v.n1, BX), (unsigned)cdiv(v.i1 - v.i0, BY));
}

// exchange FS_RO halo rows of a distributed level's operator arrays
static int halo_coeffs(fsgpu_ps* s, int l) {
  LevelDev& v = s->L[l].v;
  double* ptrs[4] = {v.d, v.dinv, v.cx, v.cy};
  return ps_halo(s, l, ptrs, 4, FS_RO);
}

int ps_setup(fsgpu_ps* s) {
  fsgpu_ctx* c = s->c;
  const Geom& g = c->g;
  const bool dist = c->nranks > 1;
  ProfScope ps_(PROF_SETUP, c->stream);
  // sizes of level 0 are fixed; the first setup needs level-0 arrays before the schedule is
  // known, so allocate for a provisional single level, then re-run once the schedule exists.
  if (!s->levels_allocated) {
    s->nl = 1;
    s->L[0].v.n0 = g.nx + 2; s->L[0].v.n1 = g.ny + 2;
    s->L[0].cdir = -1;
    if (partition(s)) return 1;
    if (alloc_levels(s)) return 1;
  }
  const unsigned gr = persist_blocks(g.a_hi - g.a_lo, g.ny + 2);
  if (fs_ensure_part(c, (size_t)gr * 4)) return 1;
  k_build_l0<<<gr, dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_RHO_U], c->f[FSGPU_RHO_V], s->L[0].v, g,
                                                 s->o.psd, c->part, c->ticket, c->red_out + 4);
  FS_LAUNCHED();
  if (comm_allreduce(c, c->red_out + 4, 4, COMM_SUM)) return 1;
  FS_CUDA(cudaMemcpyAsync(c->h_red + 4, c->red_out + 4, 4 * sizeof(double), cudaMemcpyDeviceToHost,
                          c->stream));
  FS_CUDA(cudaStreamSynchronize(c->stream));
  // remember the previous schedule to see whether the arrays can be reused
  int old_nl = s->levels_allocated ? s->nl : 0;
  int old_cdir[FS_MAX_LEVELS];
  for (int l = 0; l < old_nl; ++l) old_cdir[l] = s->L[l].cdir;
  schedule(s, c->h_red + 4);
  if (partition(s)) return 1;
  bool same = (old_nl == s->nl);
  for (int l = 0; same && l < s->nl; ++l) same = (old_cdir[l] == s->L[l].cdir);
  if (!same) {
    if (alloc_levels(s)) return 1;
    s->levels_allocated = true;
    // the slab was cleared: rebuild level 0
    k_build_l0<<<gr, dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_RHO_U], c->f[FSGPU_RHO_V], s->L[0].v,
                                                   g, s->o.psd, c->part, c->ticket, c->red_out + 4);
    FS_LAUNCHED();
  }
  if (halo_coeffs(s, 0)) return 1;
  FS_CUDA(cudaMemsetAsync(&s->scal->coarse_zero_diag, 0, sizeof(int), c->stream));
  for (int l = 0; l + 1 < s->nl; ++l) {
    const LevelDev& f = s->L[l].v;
    LevelDev cc = s->L[l + 1].v;
    const bool gather = dist && l + 1 == s->ld;  // first replicated level: built in slabs, then summed
    if (!s->L[l].replicated) { cc.i0 = s->L[l + 1].own0; cc.i1 = s->L[l + 1].own1; }
    if (s->L[l].cdir == 0) k_interp<0><<<lvl_grid(f), dim3(BX, BY), 0, c->stream>>>(f);
    else k_interp<1><<<lvl_grid(f), dim3(BX, BY), 0, c->stream>>>(f);
    FS_LAUNCHED();
    if (!s->L[l].replicated) {
      double* pc = f.pc;
      if (ps_halo(s, l, &pc, 1, FS_RO)) return 1;
    }
    const size_t cn = (size_t)(cc.n0 * cc.pitch);
    double* arrs[4] = {cc.d, cc.dinv, cc.cx, cc.cy};
    if (gather)
      for (double* a : arrs) FS_CUDA(cudaMemsetAsync(a + FS_RO * cc.pitch, 0, cn * sizeof(double), c->stream));
    if (s->L[l].cdir == 0) k_rap<0><<<lvl_grid(cc), dim3(BX, BY), 0, c->stream>>>(f, cc);
    else k_rap<1><<<lvl_grid(cc), dim3(BX, BY), 0, c->stream>>>(f, cc);
    FS_LAUNCHED();
    if (gather) {
      for (double* a : arrs)
        if (comm_allreduce(c, a + FS_RO * cc.pitch, (int64_t)cn, COMM_SUM)) return 1;
    } else if (dist && l + 1 < s->ld) {
      if (halo_coeffs(s, l + 1)) return 1;
    }
  }
  if (s->nl >= 2) {
    const LevelDev& cl = s->L[s->nl - 1].v;
    k_zero_diag<<<lvl_grid(cl), dim3(BX, BY), 0, c->stream>>>(cl, s->scal);
    FS_LAUNCHED();
  }
  s->is_setup = true;
  return 0;
}
