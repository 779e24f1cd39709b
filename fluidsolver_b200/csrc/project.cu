// fluidsolver_b200/csrc/project.cu -- divergence / rhs / projection / reductions.
//
//   k_divergence   calc_divergence               src/Operators.hpp:32-41 (all cells incl. ghosts)
//   k_rhs          fill_pressure_rhs             src/LinearSystem.hpp:125-155
//   k_sum          integrate<true>               src/Operators.hpp:52-62
//   k_project      shift_pressure_to_zero + p += dp + velocity correction
//                                                src/Operators.hpp:78-82, scaling/LidDrivenFlow.cpp:153-160,
//                                                test/LaminarChannel.cpp:222-233
//   k_cfl          adjust_dt                     src/FS.hpp:100-128
//   k_absmax       abs_max                       src/Utility.hpp:80-88
//   k_rho_visc_vof calc_rho(vf,fs)/calc_visc(vf,fs)  src/FS.hpp:592-650
#include "ctx.h"

#include <math.h>

#include <algorithm>

namespace {

constexpr int BX = 64, BY = 4;

// uniform divisors of a launch (DivBy: the reference's divisions, correctly rounded, without the division)
struct Divs { DivBy dx, dy, dt; };

__device__ __forceinline__ double div_at(const double* __restrict__ U, const double* __restrict__ V,
                                         int64_t k, int64_t P, const Divs& q) {
  return q.dx(__ldg(U + k + P) - __ldg(U + k)) + q.dy(__ldg(V + k + 1) - __ldg(V + k));
}

__global__ void __launch_bounds__(BX* BY)
k_divergence(const double* __restrict__ U, const double* __restrict__ V, double* __restrict__ div,
             const Geom g, const Divs q) {
  const int64_t b = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t a = g.a_lo + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (a >= g.a_hi || b > g.ny + 1) return;
  const int64_t k = g.idx(a, b);
  div[k] = div_at(U, V, k, g.pitch, q);
}

// rhs = -vol * div / dt on all cells; zero on the Dirichlet side; block partial sums of rhs.
// FROM_UV: div is computed from U, V here (and stored) instead of being read.
template <bool FROM_UV>
__global__ void __launch_bounds__(BX* BY)
k_rhs(const double* __restrict__ U, const double* __restrict__ V, double* __restrict__ div,
      double* __restrict__ rhs, const Geom g, const Divs q, int psd, double* part, unsigned int* ticket,
      double* out) {
  double v[1] = {0.0};
  const double vol = g.dx * g.dy;
  for_each_point(g.a_lo, g.a_hi, g.ny + 2, [&](int64_t a, int64_t b) {
    const int64_t k = g.idx(a, b);
    double d;
    if (FROM_UV) {  // the reference's operation order (Operators.hpp:36-39): true divisions by dx, dy
      d      = div_at(U, V, k, g.pitch, q);
      div[k] = d;
    } else {
      d = div[k];
    }
    double r = q.dt(__dmul_rn(-vol, d));  // -vol * div / dt, LinearSystem.hpp:134
    if ((psd == FSGPU_PSD_LEFT && a == 0) || (psd == FSGPU_PSD_RIGHT && a == g.nx + 1) ||
        (psd == FSGPU_PSD_BOTTOM && b == 0) || (psd == FSGPU_PSD_TOP && b == g.ny + 1))
      r = 0.0;
    rhs[k] = r;
    v[0] += r;
  });
  double tot[1];
  if (grid_reduce<ReduceSum, 1>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0) out[0] = tot[0];
  }
}

// f -= (*sum) * s0 * s1 on all cells (the association of integrate(), src/Operators.hpp:61: integral * dx * dy;
// the rhs mean passes s0 = 1 / size, s1 = 1)
__global__ void __launch_bounds__(BX* BY)
k_sub_scaled(double* __restrict__ f, const double* __restrict__ sum, double s0, double s1, bool divide, const Geom g) {
  const int64_t b = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t a = g.a_lo + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (a >= g.a_hi || b > g.ny + 1) return;
  const double c = divide ? __ldg(sum) / s0 : __ldg(sum) * s0 * s1;
  f[g.idx(a, b)] -= c;
}

__global__ void __launch_bounds__(BX* BY)
k_sum(const double* __restrict__ f, const Geom g, double* part, unsigned int* ticket, double* out) {
  double v[1] = {0.0};
  for_each_point(g.a_lo, g.a_hi, g.ny + 2, [&](int64_t a, int64_t b) { v[0] += f[g.idx(a, b)]; });
  double tot[1];
  if (grid_reduce<ReduceSum, 1>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0) out[0] = tot[0];
  }
}

// The projection of one sub-iteration in ONE pass: shift_pressure_to_zero (src/Operators.hpp:78-82) + p += delta_p +
// the velocity correction (scaling/LidDrivenFlow.cpp:153-160, test/LaminarChannel.cpp:222-233), with the
// reference's operation order ((delta / dx) * dt) / rho.
// SHIFT: delta_p is used as dpi - c with c = (*sum) * dx * dy, *sum = the sum of ALL entries of dpi, and the
// shifted field is stored to dpo, a DIFFERENT buffer (a neighbour's un-shifted value is still needed by the threads
// next to it, so the shift cannot be stored in place inside the same pass); the caller swaps the two buffers.
#ifndef FS_PROJECT_MINB
#define FS_PROJECT_MINB 8
#endif
template <bool SHIFT>
__global__ void __launch_bounds__(BX* BY, FS_PROJECT_MINB)
k_project(const double* __restrict__ dpi, double* __restrict__ dpo, double* __restrict__ p, double* __restrict__ U,
          double* __restrict__ V, const double* __restrict__ ru, const double* __restrict__ rv,
          const double* __restrict__ sum, const Geom g, const Divs q, double dt) {
  const double c = SHIFT ? __ldg(sum) * g.dx * g.dy : 0.0;
  const int64_t P = g.pitch;
  // persistent tile loop (as the reduction kernels): four rows per thread in flight, 8 CTAs per SM
  for_each_point(g.a_lo, g.a_hi, g.ny + 2, [&](int64_t a, int64_t b) {
    const int64_t k = g.idx(a, b);
    const double d0 = SHIFT ? __ldg(dpi + k) - c : __ldg(dpi + k);
    // U(i,j) for i in [0,NX], j in [0,NY): a in [1,NX+1], b in [1,NY]
    if (a >= 1 && b >= 1 && b <= g.ny) {
      const double dm = SHIFT ? __ldg(dpi + k - P) - c : __ldg(dpi + k - P);
      U[k] -= __dmul_rn(q.dx(d0 - dm), dt) / __ldg(ru + k);
    }
    // V(i,j) for i in [0,NX), j in [0,NY]: a in [1,NX], b in [1,NY+1]
    if (a >= 1 && a <= g.nx && b >= 1) {
      const double dm = SHIFT ? __ldg(dpi + k - 1) - c : __ldg(dpi + k - 1);
      V[k] -= __dmul_rn(q.dy(d0 - dm), dt) / __ldg(rv + k);
    }
    p[k] += d0;
    if (SHIFT) dpo[k] = d0;
  });
}

// in-place shift as its own pass (public fsgpu_shift_pressure_to_zero): k_sum -> k_sub_scaled.

__global__ void __launch_bounds__(BX* BY)
k_cfl(const double* __restrict__ U, const double* __restrict__ V, const double* __restrict__ ru,
      const double* __restrict__ rv, const double* __restrict__ mu, const Geom g, double* part,
      unsigned int* ticket, double* out) {
  double v[4] = {0.0, 0.0, 0.0, 0.0};  // the reference's maxima start at 0 (FS.hpp:102-106)
  const int64_t r0 = g.a_lo > 1 ? g.a_lo : 1, r1 = g.a_hi < g.nx + 1 ? g.a_hi : g.nx + 1;
  for_each_point(r0, r1, g.ny, [&](int64_t a, int64_t j) {  // a = i + 1 over the owned interior rows
    const int64_t k = g.idx(a, j + 1), P = g.pitch;
    v[0] = fmax(v[0], (U[k] + U[k + P]) / 2 / g.dx);
    v[1] = fmax(v[1], (V[k] + V[k + 1]) / 2 / g.dy);
    const double rho = (ru[k] + ru[k + P] + rv[k] + rv[k + 1]) / 4.0;
    v[2] = fmax(v[2], 4.0 * mu[k] / (g.dx * g.dx * rho));
    v[3] = fmax(v[3], 4.0 * mu[k] / (g.dy * g.dy * rho));
  });
  double tot[4];
  if (grid_reduce<ReduceMax, 4>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0)
      for (int q = 0; q < 4; ++q) out[q] = fmax(tot[q], 0.0);
  }
}

// max |f| (OP 0), sum f (OP 1), sum |f| (OP 2), max f (OP 3) or max -f (OP 4) over rows [r0, r1) x columns
// [b0, b0 + nb) (ghost-inclusive indices): abs_max, integrate, L1_norm, max, min (src/Utility.hpp:80-115,
// src/Operators.hpp:51-76)
template <int OP>
__global__ void __launch_bounds__(BX* BY)
k_reduce_field(const double* __restrict__ f, int64_t r0, int64_t r1, int64_t b0, int64_t nb, const Geom g,
               double scale, double* part, unsigned int* ticket, double* out) {
  double v[1] = {OP >= 3 ? ReduceMax::id() : 0.0};
  for_each_point(r0, r1, nb, [&](int64_t a, int64_t j) {
    const double x = f[g.idx(a, b0 + j)];
    if (OP == 0) v[0] = fmax(v[0], fabs(x));
    else if (OP == 3) v[0] = fmax(v[0], x);
    else if (OP == 4) v[0] = fmax(v[0], -x);
    else v[0] += OP == 2 ? fabs(x) : x;
  });
  double tot[1];
  if (OP == 0 || OP >= 3) {
    if (grid_reduce<ReduceMax, 1>(v, part, ticket, tot)) {
      if (threadIdx.x == 0 && threadIdx.y == 0) out[0] = OP == 0 ? fmax(tot[0], 0.0) : tot[0];
    }
  } else {
    if (grid_reduce<ReduceSum, 1>(v, part, ticket, tot)) {
      if (threadIdx.x == 0 && threadIdx.y == 0) out[0] = tot[0] * scale;
    }
  }
}

// has_nan / has_inf (src/Container.hpp:191-204) over the whole ghost-inclusive extent: counts of both in one pass
__global__ void __launch_bounds__(BX* BY)
k_count_nonfinite(const double* __restrict__ f, int64_t r0, int64_t r1, int64_t nb, const Geom g, double* part,
                  unsigned int* ticket, double* out) {
  double v[2] = {0.0, 0.0};
  for_each_point(r0, r1, nb, [&](int64_t a, int64_t j) {
    const double x = f[g.idx(a, j)];
    if (isnan(x)) v[0] += 1.0;
    else if (isinf(x)) v[1] += 1.0;
  });
  double tot[2];
  if (grid_reduce<ReduceSum, 2>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0) { out[0] = tot[0]; out[1] = tot[1]; }
  }
}

// fill the (m+2) x (n+2) ghost-inclusive extent of a field
// (rows [r0, r1): the owned rows and the halo rows inside the field's extent)
__global__ void k_fill(double* __restrict__ f, int64_t r0, int64_t r1, int64_t n2, double v, const Geom g) {
  const int64_t b = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t a = r0 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (a >= r1 || b >= n2) return;
  f[g.idx(a, b)] = v;
}

// FS.hpp:592-650: face densities are the mean of the two adjacent cell densities, viscosity is
// the harmonic mix with the VF_LOW / VF_HIGH clamps of src/IR.hpp:16-17; interior only, the
// caller applies apply_neumann_bconds to each.
__global__ void __launch_bounds__(BX* BY)
k_rho_visc_vof(const double* __restrict__ vf, double* __restrict__ ru, double* __restrict__ rv,
               double* __restrict__ mu, const Geom g, double rg, double rl, double mg, double ml) {
  const int64_t j = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t i = (g.a_lo > 1 ? g.a_lo : 1) - 1 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (i > g.nx || i + 1 >= g.a_hi || j > g.ny) return;
  const int64_t k = g.idx(i + 1, j + 1), P = g.pitch;
  const double f0 = vf[k];
  const double r0 = f0 * rl + (1.0 - f0) * rg;
  if (j < g.ny) {
    const double fm = vf[k - P];
    ru[k]           = ((fm * rl + (1.0 - fm) * rg) + r0) / 2.0;
  }
  if (i < g.nx) {
    const double fm = vf[k - 1];
    rv[k]           = ((fm * rl + (1.0 - fm) * rg) + r0) / 2.0;
  }
  if (i < g.nx && j < g.ny) {
    const double VF_LOW = 1e-8, VF_HIGH = 1.0 - 1e-8;
    double m;
    if (f0 < VF_LOW) m = mg;
    else if (f0 > VF_HIGH) m = ml;
    else m = (ml * mg) / (ml * (1.0 - f0) + mg * f0);
    mu[k] = m;
  }
}

}  // namespace

int fs_ensure_part(fsgpu_ctx* c, size_t doubles) {
  if (doubles <= c->part_cap) return 0;
  if (c->part) cudaFree(c->part);
  c->part = nullptr;
  FS_CUDA(cudaMalloc(&c->part, doubles * sizeof(double)));
  c->part_cap = doubles;
  return comm_alloc_fence(c);
}

static Divs make_divs(const Geom& g, double dt) { return Divs{make_div_by(g.dx), make_div_by(g.dy), make_div_by(dt)}; }
static dim3 cell_grid(const Geom& g) {  // the owned cell rows
  return dim3((unsigned)cdiv(g.ny + 2, BX), (unsigned)cdiv(g.a_hi - g.a_lo, BY));
}
static dim3 ext_grid(int64_t m2, int64_t n2) {
  return dim3((unsigned)cdiv(n2, BX), (unsigned)cdiv(m2, BY));
}

int fs_launch_divergence(fsgpu_ctx* c) {
  k_divergence<<<cell_grid(c->g), dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_U], c->f[FSGPU_V],
                                                                c->f[FSGPU_DIV], c->g, make_divs(c->g, 1.0));
  FS_LAUNCHED();
  return 0;
}

// defer_mean: leave the mean in (its sum, all-reduced, stays in red_out[0]); the solver's first readers of the
// rhs subtract it (fsgpu_substep; solver_cycle.cu BShift).  *deferred tells the caller whether that applies.
int fs_launch_rhs(fsgpu_ctx* c, double dt, int psd, bool div_from_uv, bool defer_mean, bool* deferred) {
  if (deferred) *deferred = false;
  const dim3 gr = cell_grid(c->g);
  const unsigned pg = persist_blocks(c->g.a_hi - c->g.a_lo, c->g.ny + 2);
  if (fs_ensure_part(c, (size_t)pg * 4)) return 1;
  if (div_from_uv)
    k_rhs<true><<<pg, dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_U], c->f[FSGPU_V], c->f[FSGPU_DIV],
                                                    c->f[FSGPU_RHS], c->g, make_divs(c->g, dt), psd, c->part, c->ticket,
                                                    c->red_out);
  else
    k_rhs<false><<<pg, dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_U], c->f[FSGPU_V], c->f[FSGPU_DIV],
                                                     c->f[FSGPU_RHS], c->g, make_divs(c->g, dt), psd, c->part, c->ticket,
                                                     c->red_out);
  FS_LAUNCHED();
  if (psd == FSGPU_PSD_NONE) {  // subtract the arithmetic mean over all (NX+2)(NY+2) entries: sum / size (LinearSystem.hpp:150-152)
    if (comm_allreduce(c, c->red_out, 1, COMM_SUM)) return 1;
    if (defer_mean) {
      if (deferred) *deferred = true;
      return 0;
    }
    k_sub_scaled<<<gr, dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_RHS], c->red_out,
                                                     (double)((c->g.nx + 2) * (c->g.ny + 2)), 1.0, true, c->g);
    FS_LAUNCHED();
  }
  return 0;
}

// sum of all entries of delta_p -> red_out[1] (all ranks)
static int launch_sum_dp(fsgpu_ctx* c) {
  const unsigned gr = persist_blocks(c->g.a_hi - c->g.a_lo, c->g.ny + 2);
  if (fs_ensure_part(c, (size_t)gr * 4)) return 1;
  k_sum<<<gr, dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_DELTA_P], c->g, c->part, c->ticket,
                                            c->red_out + 1);
  FS_LAUNCHED();
  return comm_allreduce(c, c->red_out + 1, 1, COMM_SUM);
}

int fs_launch_shift_pressure(fsgpu_ctx* c) {
  if (launch_sum_dp(c)) return 1;
  k_sub_scaled<<<cell_grid(c->g), dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_DELTA_P], c->red_out + 1, c->g.dx,
                                                                c->g.dy, false, c->g);
  FS_LAUNCHED();
  return 0;
}

// shift: shift_pressure_to_zero, p += delta_p and the velocity correction in one pass (k_project<true>);
// sum_ready: red_out[1] already holds the sum of all entries of delta_p on every rank (the solver leaves it
// there, solver_pcg.cu), otherwise k_sum runs first.  The shifted delta_p lands in the buffer of FSGPU_RHS
// (dead after the solve) and the two field pointers are swapped.
int fs_launch_correct(fsgpu_ctx* c, double dt, bool shift, bool sum_ready) {
  if (shift && !sum_ready && launch_sum_dp(c)) return 1;
  {  // the correction of face a reads delta_p of row a - 1
    const int fl[1] = {FSGPU_DELTA_P};
    if (fs_exchange_fields(c, fl, 1, 1)) return 1;
  }
  if (shift) {
    k_project<true><<<persist_blocks(c->g.a_hi - c->g.a_lo, c->g.ny + 2), dim3(BX, BY), 0, c->stream>>>(
        c->f[FSGPU_DELTA_P], c->f[FSGPU_RHS], c->f[FSGPU_P], c->f[FSGPU_U], c->f[FSGPU_V], c->f[FSGPU_RHO_U],
        c->f[FSGPU_RHO_V], c->red_out + 1, c->g, make_divs(c->g, dt), dt);
    std::swap(c->f[FSGPU_DELTA_P], c->f[FSGPU_RHS]);
  } else {
    k_project<false><<<persist_blocks(c->g.a_hi - c->g.a_lo, c->g.ny + 2), dim3(BX, BY), 0, c->stream>>>(
        c->f[FSGPU_DELTA_P], nullptr, c->f[FSGPU_P], c->f[FSGPU_U], c->f[FSGPU_V], c->f[FSGPU_RHO_U],
        c->f[FSGPU_RHO_V], c->red_out + 1, c->g, make_divs(c->g, dt), dt);
  }
  FS_LAUNCHED();
  const int fl[3] = {FSGPU_U, FSGPU_V, FSGPU_P};
  return fs_exchange_fields(c, fl, 3, 1);
}

int fs_launch_cfl(fsgpu_ctx* c) {
  const unsigned gr = persist_blocks(c->g.a_hi - c->g.a_lo, c->g.ny);
  if (fs_ensure_part(c, (size_t)gr * 4)) return 1;
  k_cfl<<<gr, dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_U], c->f[FSGPU_V], c->f[FSGPU_RHO_U],
                                            c->f[FSGPU_RHO_V], c->f[FSGPU_VISC], c->g, c->part,
                                            c->ticket, c->red_out);
  FS_LAUNCHED();
  return comm_allreduce(c, c->red_out, 4, COMM_MAX);
}

// abs_max<INCLUDE_GHOST>(field), src/Utility.hpp:44-88: the reference's default reduces over the
// INTERIOR of the field (for_each_i), abs_max<true> over the whole ghost-inclusive extent
// op 0: abs_max -> red_out[0] (max over ranks); op 1 / 2: integrate / L1_norm (* dx * dy, sum over ranks);
// op 3: max; op 4: -min
int fs_launch_reduce_field(fsgpu_ctx* c, int field, bool include_ghost, int op) {
  const Geom& g = c->g;
  int64_t m2, n2;
  field_extent(g, field, &m2, &n2);
  const int64_t lo = include_ghost ? 0 : 1;
  const int64_t own_hi = g.last() ? m2 : g.a_hi;  // a U-like field's extra row belongs to the last rank
  const int64_t r0 = std::max<int64_t>(lo, g.a_lo), r1 = std::min<int64_t>(m2 - lo, own_hi);
  const int64_t nb = n2 - 2 * lo;
  const unsigned nbl = persist_blocks(r1 > r0 ? r1 - r0 : 0, nb);
  if (fs_ensure_part(c, (size_t)nbl * 4)) return 1;
  const dim3 blk(BX, BY);
  // each rank scales its partial sum; the reference scales the total (<= 1 ulp apart per rank)
  const double sc = g.dx * g.dy;
  if (op == 0)
    k_reduce_field<0><<<nbl, blk, 0, c->stream>>>(c->f[field], r0, r1, lo, nb, g, 1.0, c->part, c->ticket, c->red_out);
  else if (op == 1)
    k_reduce_field<1><<<nbl, blk, 0, c->stream>>>(c->f[field], r0, r1, lo, nb, g, sc, c->part, c->ticket, c->red_out);
  else if (op == 2)
    k_reduce_field<2><<<nbl, blk, 0, c->stream>>>(c->f[field], r0, r1, lo, nb, g, sc, c->part, c->ticket, c->red_out);
  else if (op == 3)
    k_reduce_field<3><<<nbl, blk, 0, c->stream>>>(c->f[field], r0, r1, lo, nb, g, 1.0, c->part, c->ticket, c->red_out);
  else  // the minimum travels as max(-f); the caller negates
    k_reduce_field<4><<<nbl, blk, 0, c->stream>>>(c->f[field], r0, r1, lo, nb, g, 1.0, c->part, c->ticket, c->red_out);
  FS_LAUNCHED();
  return comm_allreduce(c, c->red_out, 1, (op == 1 || op == 2) ? COMM_SUM : COMM_MAX);
}
int fs_launch_count_nonfinite(fsgpu_ctx* c, int field) {  // -> red_out[0] NaN count, red_out[1] Inf count
  const Geom& g = c->g;
  int64_t m2, n2;
  field_extent(g, field, &m2, &n2);
  const int64_t r0 = g.a_lo, r1 = g.last() ? m2 : g.a_hi;
  const unsigned nbl = persist_blocks(r1 - r0, n2);
  if (fs_ensure_part(c, (size_t)nbl * 4)) return 1;
  k_count_nonfinite<<<nbl, dim3(BX, BY), 0, c->stream>>>(c->f[field], r0, r1, n2, g, c->part, c->ticket, c->red_out);
  FS_LAUNCHED();
  return comm_allreduce(c, c->red_out, 2, COMM_SUM);
}
int fs_launch_absmax(fsgpu_ctx* c, int field, bool include_ghost) {
  return fs_launch_reduce_field(c, field, include_ghost, 0);
}

int fs_launch_fill(fsgpu_ctx* c, int field, double v) {
  int64_t m, n;
  field_extent(c->g, field, &m, &n);
  const int64_t r0 = c->g.a_lo - FS_RO > 0 ? c->g.a_lo - FS_RO : 0;
  const int64_t r1 = c->g.a_hi + FS_RO < m ? c->g.a_hi + FS_RO : m;
  k_fill<<<ext_grid(r1 - r0, n), dim3(BX, BY), 0, c->stream>>>(c->f[field], r0, r1, n, v, c->g);
  FS_LAUNCHED();
  return 0;
}

int fs_launch_rho_visc_vof(fsgpu_ctx* c) {
  k_rho_visc_vof<<<dim3((unsigned)cdiv(c->g.ny + 1, BX), (unsigned)cdiv(c->g.a_hi - c->g.a_lo, BY)),
                   dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_VF], c->f[FSGPU_RHO_U], c->f[FSGPU_RHO_V],
                                                 c->f[FSGPU_VISC], c->g, c->rho_gas, c->rho_liquid,
                                                 c->visc_gas, c->visc_liquid);
  FS_LAUNCHED();
  if (fs_launch_neumann(c, c->f[FSGPU_RHO_U], c->g.nx + 1, c->g.ny)) return 1;
  if (fs_launch_neumann(c, c->f[FSGPU_RHO_V], c->g.nx, c->g.ny + 1)) return 1;
  if (fs_launch_neumann(c, c->f[FSGPU_VISC], c->g.nx, c->g.ny)) return 1;
  const int fl[3] = {FSGPU_RHO_U, FSGPU_RHO_V, FSGPU_VISC};
  return fs_exchange_fields(c, fl, 3, FS_RO);
}
