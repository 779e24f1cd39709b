// fluidsolver_b200/csrc/twophase.cu -- the rows either side of the single-phase path:
// consistent mass transport, the capillary pressure jump, immersed-boundary velocity updates and
// the per-step diagnostics.  Same layout, block shape and slab rules as stencil.cu / project.cu.
//
//   k_drhodt<MID,UPD>     calc_drhodt (+ update_density fused)    src/FS.hpp:318-420
//   k_update_density      update_density                          src/FS.hpp:409-420
//   k_pressure_jump       calc_pressure_jump                      src/FS.hpp:439-466
//   k_capillary_div       div += dt div(delta_pj / rho)           test/StationaryDrop.cpp:334-353
//   k_ib_implicit_euler   correct_velocity_ib_implicit_euler      src/IB.hpp:110-126
//   k_ib_semi_analytical  update_velocity_ib_semi_analytical      src/IB.hpp:129-186
//   k_interpolate         interpolate_U / _V / _UV_staggered_field src/Operators.hpp:10-29
//   k_conserved           calc_conserved_quantities               src/FS.hpp:653-676
//   k_monitor             calc_divergence + 3 x abs_max           scaling/LidDrivenFlow.cpp:166-170
#include "ctx.h"

#include <math.h>

#include <algorithm>
#include <vector>

namespace {

constexpr int BX = 64, BY = 4;

// density half of hybrid_interp (src/FS.hpp:132-150) as calc_drhodt uses it (velocity slots 0)
__device__ __forceinline__ double rho_hybrid(double eps, double rm, double rp, double tm, double tp) {
  if (!(fabs(rp - rm) > eps)) return (rp + rm) / 2.0;
  return (tp + tm >= 0.0) ? rm : rp;
}

struct RhoArgs {
  const double *U, *V, *Uo, *Vo;  // curr (or, MID, to-be-averaged) and old velocities
  const double *ruo, *rvo;        // old face densities
  double *outU, *outV;            // drho/dt (UPD = false) or the new face densities (UPD = true)
  double dx, dy, dt, eps;
};

template <bool MID>
struct Vel2 {
  const double *c, *o;
  __device__ __forceinline__ double operator()(int64_t k) const {
    if (MID) return 0.5 * (__ldg(c + k) + __ldg(o + k));
    return __ldg(c + k);
  }
};

// mass fluxes of FS.hpp:340-403 at linear index k = (i, j)
template <bool MID>
__device__ __forceinline__ double mflux_FXU(const RhoArgs& a, const Vel2<MID>& U, int64_t k, int64_t P) {
  const double um = U(k), up = U(k + P);
  return -rho_hybrid(a.eps, __ldg(a.ruo + k), __ldg(a.ruo + k + P), um, up) * ((um + up) / 2.0);
}
template <bool MID>
__device__ __forceinline__ double mflux_FYU(const RhoArgs& a, const Vel2<MID>& V, int64_t k, int64_t P) {
  const double v0 = V(k - P), v1 = V(k);
  return -rho_hybrid(a.eps, __ldg(a.ruo + k - 1), __ldg(a.ruo + k), v0, v1) * ((v0 + v1) / 2);
}
template <bool MID>
__device__ __forceinline__ double mflux_FXV(const RhoArgs& a, const Vel2<MID>& U, int64_t k, int64_t P) {
  const double u0 = U(k - 1), u1 = U(k);
  return -rho_hybrid(a.eps, __ldg(a.rvo + k - P), __ldg(a.rvo + k), u0, u1) * ((u0 + u1) / 2.0);
}
template <bool MID>
__device__ __forceinline__ double mflux_FYV(const RhoArgs& a, const Vel2<MID>& V, int64_t k) {
  const double vm = V(k), vp = V(k + 1);
  return -rho_hybrid(a.eps, __ldg(a.rvo + k), __ldg(a.rvo + k + 1), vm, vp) * ((vm + vp) / 2.0);
}

// One thread per (i, j), i in [0, NX], j in [0, NY]: the U face and the V face with that index.
// MID: calc_mid_time fused in; UPD: update_density fused in (rho = rho_old + dt * drho/dt).
template <bool MID, bool UPD>
__global__ void __launch_bounds__(BX* BY)
k_drhodt(const RhoArgs a, const Geom g) {
  const int64_t j = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t i = (g.a_lo > 1 ? g.a_lo : 1) - 1 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (i > g.nx || i + 1 >= g.a_hi || j > g.ny) return;
  const int64_t P = g.pitch, k = g.idx(i + 1, j + 1);
  const Vel2<MID> U{a.U, a.Uo}, V{a.V, a.Vo};
  if (j < g.ny) {  // FS.hpp:368-371
    const double d = (mflux_FXU<MID>(a, U, k, P) - mflux_FXU<MID>(a, U, k - P, P)) / a.dx +
                     (mflux_FYU<MID>(a, V, k + 1, P) - mflux_FYU<MID>(a, V, k, P)) / a.dy;
    a.outU[k] = UPD ? __ldg(a.ruo + k) + a.dt * d : d;
  }
  if (i < g.nx) {  // FS.hpp:405-408
    const double d = (mflux_FXV<MID>(a, U, k + P, P) - mflux_FXV<MID>(a, U, k, P)) / a.dx +
                     (mflux_FYV<MID>(a, V, k) - mflux_FYV<MID>(a, V, k - 1)) / a.dy;
    a.outV[k] = UPD ? __ldg(a.rvo + k) + a.dt * d : d;
  }
}

__global__ void __launch_bounds__(BX* BY)
k_update_density(const double* __restrict__ ruo, const double* __restrict__ rvo,
                 const double* __restrict__ du, const double* __restrict__ dv,
                 double* __restrict__ ru, double* __restrict__ rv, double dt, const Geom g) {
  const int64_t j = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t i = (g.a_lo > 1 ? g.a_lo : 1) - 1 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (i > g.nx || i + 1 >= g.a_hi || j > g.ny) return;
  const int64_t k = g.idx(i + 1, j + 1);
  if (j < g.ny) ru[k] = ruo[k] + dt * du[k];
  if (i < g.nx) rv[k] = rvo[k] + dt * dv[k];
}

// FS.hpp:439-466 on the interior faces; the caller has zeroed both outputs (fill(..., 0.0))
__global__ void __launch_bounds__(BX* BY)
k_pressure_jump(const double* __restrict__ vf, const double* __restrict__ curv,
                const double* __restrict__ ilen, double* __restrict__ pju, double* __restrict__ pjv,
                double sigma, const Geom g) {
  const int64_t j = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t i = (g.a_lo > 1 ? g.a_lo : 1) - 1 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (i > g.nx || i + 1 >= g.a_hi || j > g.ny) return;
  const int64_t k = g.idx(i + 1, j + 1), P = g.pitch;
  const double pl = __ldg(ilen + k), c0 = __ldg(curv + k), f0 = __ldg(vf + k);
  if (j < g.ny) {
    const double ml = __ldg(ilen + k - P);
    const double ci = (pl + ml) > 0.0 ? (c0 * pl + __ldg(curv + k - P) * ml) / (pl + ml) : 0.0;
    pju[k] = sigma * ci * (f0 - __ldg(vf + k - P)) / g.dx;
  }
  if (i < g.nx) {
    const double ml = __ldg(ilen + k - 1);
    const double ci = (pl + ml) > 0.0 ? (c0 * pl + __ldg(curv + k - 1) * ml) / (pl + ml) : 0.0;
    pjv[k] = sigma * ci * (f0 - __ldg(vf + k - 1)) / g.dy;
  }
}

// test/StationaryDrop.cpp:334-353 on all cells incl. ghosts.  delta_pj = pj - pj_old is formed on
// the fly for the four faces of a cell (the reference stores it first; the arithmetic is the same).
__global__ void __launch_bounds__(BX* BY)
k_capillary_div(const double* __restrict__ pju, const double* __restrict__ pjv,
                const double* __restrict__ ou, const double* __restrict__ ov,
                const double* __restrict__ ru, const double* __restrict__ rv,
                double* __restrict__ div, double dt, const Geom g) {
  const int64_t b = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t a = g.a_lo + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (a >= g.a_hi || b > g.ny + 1) return;
  const int64_t k = g.idx(a, b), P = g.pitch;
  const double ue = (__ldg(pju + k + P) - __ldg(ou + k + P)) / __ldg(ru + k + P);
  const double uw = (__ldg(pju + k) - __ldg(ou + k)) / __ldg(ru + k);
  const double vn = (__ldg(pjv + k + 1) - __ldg(ov + k + 1)) / __ldg(rv + k + 1);
  const double vs = (__ldg(pjv + k) - __ldg(ov + k)) / __ldg(rv + k);
  div[k] += dt * ((ue - uw) / g.dx + (vn - vs) / g.dy);
}

// src/IB.hpp:110-126
__global__ void __launch_bounds__(BX* BY)
k_ib_implicit_euler(double* __restrict__ U, double* __restrict__ V, const double* __restrict__ mu,
                    const double* __restrict__ ru, const double* __restrict__ rv,
                    const double* __restrict__ ibu, const double* __restrict__ ibv, double dt,
                    const Geom g) {
  const int64_t j = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t i = (g.a_lo > 1 ? g.a_lo : 1) - 1 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (i > g.nx || i + 1 >= g.a_hi || j > g.ny) return;
  const int64_t k = g.idx(i + 1, j + 1), P = g.pitch;
  const double m0 = __ldg(mu + k);
  if (j < g.ny) {
    const double visc = (m0 + __ldg(mu + k - P)) / 2.0;
    U[k] /= 1.0 + dt * visc / __ldg(ru + k) * __ldg(ibu + k);
  }
  if (i < g.nx) {
    const double visc = (m0 + __ldg(mu + k - 1)) / 2.0;
    V[k] /= 1.0 + dt * visc / __ldg(rv + k) * __ldg(ibv + k);
  }
}

// src/IB.hpp:129-186: U = (B rho_old U_old + dt F) / (A rho), B = l dt / (exp(l dt) - 1), A = l dt + B
__device__ __forceinline__ double ib_semi(double lambda, double dt, double ro, double uo, double f,
                                          double r) {
  if (isinf(lambda)) return 0.0;
  const double B = fabs(lambda) < 1e-6 ? 1.0 : (lambda * dt) / (exp(lambda * dt) - 1.0);
  const double A = lambda * dt + B;
  return (B * ro * uo + dt * f) / (A * r);
}
struct IbArgs {
  double *U, *V;
  const double *Uo, *Vo, *ruo, *rvo, *ru, *rv, *mu, *dU, *dV, *ibu, *ibv;
  double dt;
};
__global__ void __launch_bounds__(BX* BY)
k_ib_semi_analytical(const IbArgs a, const Geom g) {
  const int64_t j = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t i = (g.a_lo > 1 ? g.a_lo : 1) - 1 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (i > g.nx || i + 1 >= g.a_hi || j > g.ny) return;
  const int64_t k = g.idx(i + 1, j + 1), P = g.pitch;
  const double m0 = __ldg(a.mu + k);
  if (j < g.ny) {
    const double visc = (m0 + __ldg(a.mu + k - P)) / 2.0, rho = __ldg(a.ru + k);
    a.U[k] = ib_semi(visc / rho * __ldg(a.ibu + k), a.dt, __ldg(a.ruo + k), __ldg(a.Uo + k),
                     __ldg(a.dU + k), rho);
  }
  if (i < g.nx) {
    const double visc = (m0 + __ldg(a.mu + k - 1)) / 2.0, rho = __ldg(a.rv + k);
    a.V[k] = ib_semi(visc / rho * __ldg(a.ibv + k), a.dt, __ldg(a.rvo + k), __ldg(a.Vo + k),
                     __ldg(a.dV + k), rho);
  }
}

// src/Operators.hpp:10-29 on all cells incl. ghosts
__global__ void __launch_bounds__(BX* BY)
k_interpolate(const double* __restrict__ U, const double* __restrict__ V,
              const double* __restrict__ ru, const double* __restrict__ rv, double* __restrict__ Ui,
              double* __restrict__ Vi, double* __restrict__ rhoi, const Geom g) {
  const int64_t b = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t a = g.a_lo + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (a >= g.a_hi || b > g.ny + 1) return;
  const int64_t k = g.idx(a, b), P = g.pitch;
  Ui[k]   = (__ldg(U + k) + __ldg(U + k + P)) / 2;
  Vi[k]   = (__ldg(V + k) + __ldg(V + k + 1)) / 2;
  rhoi[k] = (__ldg(ru + k) + __ldg(ru + k + P) + __ldg(rv + k) + __ldg(rv + k + 1)) / 4.0;
}

// src/FS.hpp:653-676 over the interior cells -> out[0..2] = mass, momentum_x, momentum_y
__global__ void __launch_bounds__(BX* BY)
k_conserved(const double* __restrict__ U, const double* __restrict__ V, const double* __restrict__ ru,
            const double* __restrict__ rv, const Geom g, double* part, unsigned int* ticket,
            double* out) {
  double v[3] = {0.0, 0.0, 0.0};
  const int64_t r0 = g.a_lo > 1 ? g.a_lo : 1, r1 = g.a_hi < g.nx + 1 ? g.a_hi : g.nx + 1;
  for_each_point(r0, r1, g.ny, [&](int64_t a, int64_t j) {
    const int64_t k = g.idx(a, j + 1), P = g.pitch;
    const double r0u = ru[k], r1u = ru[k + P], r0v = rv[k], r1v = rv[k + 1];
    v[0] += (r0u + r1u + r0v + r1v) / 4.0 * g.dx * g.dy;
    v[1] += (r0u * U[k] + r1u * U[k + P]) / 2.0 * g.dx * g.dy;
    v[2] += (r0v * V[k] + r1v * V[k + 1]) / 2.0 * g.dx * g.dy;
  });
  double tot[3];
  if (grid_reduce<ReduceSum, 3>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0)
      for (int q = 0; q < 3; ++q) out[q] = tot[q];
  }
}

// the monitor row of every driver (scaling/LidDrivenFlow.cpp:166-170): div = calc_divergence (stored,
// all cells incl. ghosts), then abs_max(curr.U), abs_max(curr.V), abs_max(div) -- abs_max's default
// reduces over the INTERIOR entries of each field (src/Utility.hpp:44-88):
// U: a in [1, NX+1], b in [1, NY];  V: a in [1, NX], b in [1, NY+1];  div: a in [1, NX], b in [1, NY]
__global__ void __launch_bounds__(BX* BY)
k_monitor(const double* __restrict__ U, const double* __restrict__ V, double* __restrict__ div,
          const Geom g, double* part, unsigned int* ticket, double* out) {
  double v[3] = {0.0, 0.0, 0.0};
  for_each_point(g.a_lo, g.a_hi, g.ny + 2, [&](int64_t a, int64_t b) {  // the owned cell rows
    const int64_t k = g.idx(a, b), P = g.pitch;
    const double u = U[k], w = V[k];
    const double d = (U[k + P] - u) / g.dx + (V[k + 1] - w) / g.dy;
    div[k] = d;
    const bool ia = a >= 1 && a <= g.nx, ib = b >= 1 && b <= g.ny;
    if (a >= 1 && ib) v[0] = fmax(v[0], fabs(u));                 // rows 1 .. NX+1 (a_hi <= NX+2)
    if (ia && b >= 1) v[1] = fmax(v[1], fabs(w));                 // columns 1 .. NY+1
    if (ia && ib) v[2] = fmax(v[2], fabs(d));
  });
  double tot[3];
  if (grid_reduce<ReduceMax, 3>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0)
      for (int q = 0; q < 3; ++q) out[q] = fmax(tot[q], 0.0);
  }
}

// calc_grad_of_centered_points, src/Operators.hpp:164-201 (NGHOST = 1): on every cell incl. ghosts, per
// direction the central difference where both neighbours exist and the second-order one-sided
// difference on the outermost ring
__global__ void __launch_bounds__(BX* BY)
k_grad_centered(const double* __restrict__ f, double* __restrict__ gx, double* __restrict__ gy, const Geom g) {
  const int64_t b = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t a = g.a_lo + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (a >= g.a_hi || b > g.ny + 1) return;
  const int64_t k = g.idx(a, b), P = g.pitch;
  const double f0 = f[k], tdx = 2.0 * g.dx, tdy = 2.0 * g.dy;
  double x, y;
  if (a == 0) x = (-3.0 * f0 + 4.0 * f[k + P] - f[k + 2 * P]) / tdx;
  else if (a == g.nx + 1) x = (3.0 * f0 - 4.0 * f[k - P] + f[k - 2 * P]) / tdx;
  else x = (f[k + P] - f[k - P]) / tdx;
  if (b == 0) y = (-3.0 * f0 + 4.0 * f[k + 1] - f[k + 2]) / tdy;
  else if (b == g.ny + 1) y = (3.0 * f0 - 4.0 * f[k - 1] + f[k - 2]) / tdy;
  else y = (f[k + 1] - f[k - 1]) / tdy;
  gx[k] = x;
  gy[k] = y;
}

// bilinear sampling of cell fields at arbitrary points, src/Operators.hpp:85-161.  xm, ym: cell-centre
// coordinates incl. their ghost entry (index + 1).  MODE 0 = eval_flow_field_at (index pairs clamped into
// [0, N-1], two fields), MODE 1 = bilinear_interpolate (clamped by position only, one field).
struct SampleArgs {
  const double *fa, *fb, *xm, *ym, *px, *py;
  double *oa, *ob;
  int64_t npts;
};
template <int MODE>
__device__ __forceinline__ void sample_indices(double pos, const double* __restrict__ grid, int64_t N, double delta,
                                               int64_t& prev, int64_t& next) {
  const int64_t p = (int64_t)floor((pos - grid[1]) / delta);
  const int64_t q = (int64_t)floor((pos - grid[1]) / delta + 1.0);
  if (pos <= grid[1] || (MODE == 0 && p < 0)) { prev = next = 0; return; }
  if (pos >= grid[N] || (MODE == 0 && q >= N)) { prev = next = N - 1; return; }
  prev = p;
  next = q;
}
template <int MODE>
__global__ void __launch_bounds__(256)
k_sample(const SampleArgs s, const Geom g) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= s.npts) return;
  const double dx = s.xm[2] - s.xm[1], dy = s.ym[2] - s.ym[1];
  const double x = s.px[k], y = s.py[k];
  int64_t ip, in, jp, jn;
  sample_indices<MODE>(x, s.xm, g.nx, dx, ip, in);
  sample_indices<MODE>(y, s.ym, g.ny, dy, jp, jn);
  if (ip + 1 < g.a_lo || ip + 1 >= g.a_hi) {  // another slab owns the point's row: it contributes, the sum combines
    s.oa[k] = 0.0;
    if (MODE == 0) s.ob[k] = 0.0;
    return;
  }
  const int64_t k00 = g.idx(ip + 1, jp + 1), k10 = g.idx(in + 1, jp + 1), k01 = g.idx(ip + 1, jn + 1),
                k11 = g.idx(in + 1, jn + 1);
  const double tx = x - s.xm[ip + 1], ty = y - s.ym[jp + 1];
  {
    const double a = (s.fa[k10] - s.fa[k00]) / dx * tx + s.fa[k00];
    const double b = (s.fa[k11] - s.fa[k01]) / dx * tx + s.fa[k01];
    s.oa[k] = (b - a) / dy * ty + a;
  }
  if (MODE == 0) {
    const double a = (s.fb[k10] - s.fb[k00]) / dx * tx + s.fb[k00];
    const double b = (s.fb[k11] - s.fb[k01]) / dx * tx + s.fb[k01];
    s.ob[k] = (b - a) / dy * ty + a;
  }
}

}  // namespace

static dim3 face_grid(const Geom& g) {
  return dim3((unsigned)cdiv(g.ny + 1, BX), (unsigned)cdiv(g.a_hi - g.a_lo, BY));
}
static dim3 cell_grid(const Geom& g) {
  return dim3((unsigned)cdiv(g.ny + 2, BX), (unsigned)cdiv(g.a_hi - g.a_lo, BY));
}

// extension fields live outside the core slab and are allocated (zeroed) on first use
int fs_ensure_field(fsgpu_ctx* c, int field) {
  if (c->f[field]) return 0;
  const size_t n = (size_t)c->g.count();
  double* base = nullptr;
  FS_CUDA(cudaMalloc(&base, n * sizeof(double)));
  FS_CUDA(cudaMemsetAsync(base, 0, n * sizeof(double), c->stream));
  c->ext_alloc[field] = base;
  c->f[field] = base - c->g.voff();
  return comm_alloc_fence(c);
}
static int ensure_fields(fsgpu_ctx* c, std::initializer_list<int> fl) {
  for (int f : fl)
    if (fs_ensure_field(c, f)) return 1;
  return 0;
}

static RhoArgs rho_args(fsgpu_ctx* c, double dt) {
  RhoArgs a;
  a.U = c->f[FSGPU_U]; a.V = c->f[FSGPU_V]; a.Uo = c->f[FSGPU_U_OLD]; a.Vo = c->f[FSGPU_V_OLD];
  a.ruo = c->f[FSGPU_RHO_U_OLD]; a.rvo = c->f[FSGPU_RHO_V_OLD];
  a.outU = nullptr; a.outV = nullptr;
  a.dx = c->g.dx; a.dy = c->g.dy; a.dt = dt;
  a.eps = 1e-3 * fmin(c->rho_gas, c->rho_liquid);  // calc_rho_eps, FS.hpp:153-155
  return a;
}

int fs_launch_drhodt(fsgpu_ctx* c) {
  if (ensure_fields(c, {FSGPU_DRHO_U, FSGPU_DRHO_V})) return 1;
  // ghosts of the outputs are 0 (fill(drho_*, 0), FS.hpp:331-332): never written after allocation
  RhoArgs a = rho_args(c, 0.0);
  a.outU = c->f[FSGPU_DRHO_U]; a.outV = c->f[FSGPU_DRHO_V];
  k_drhodt<false, false><<<face_grid(c->g), dim3(BX, BY), 0, c->stream>>>(a, c->g);
  FS_LAUNCHED();
  return 0;
}

int fs_launch_update_density(fsgpu_ctx* c, double dt) {
  if (ensure_fields(c, {FSGPU_DRHO_U, FSGPU_DRHO_V})) return 1;
  k_update_density<<<face_grid(c->g), dim3(BX, BY), 0, c->stream>>>(
      c->f[FSGPU_RHO_U_OLD], c->f[FSGPU_RHO_V_OLD], c->f[FSGPU_DRHO_U], c->f[FSGPU_DRHO_V],
      c->f[FSGPU_RHO_U], c->f[FSGPU_RHO_V], dt, c->g);
  FS_LAUNCHED();
  const int fl[2] = {FSGPU_RHO_U, FSGPU_RHO_V};
  return fs_exchange_fields(c, fl, 2, FS_RO);
}

// calc_drhodt + update_density + apply_neumann_bconds(rho_u / rho_v) (test/StationaryDrop.cpp:274-293)
// in one field pass; mid: velocities are still un-averaged (calc_mid_time fused in)
int fs_launch_density_step(fsgpu_ctx* c, double dt, bool mid) {
  RhoArgs a = rho_args(c, dt);
  a.outU = c->f[FSGPU_RHO_U]; a.outV = c->f[FSGPU_RHO_V];
  if (mid) k_drhodt<true, true><<<face_grid(c->g), dim3(BX, BY), 0, c->stream>>>(a, c->g);
  else k_drhodt<false, true><<<face_grid(c->g), dim3(BX, BY), 0, c->stream>>>(a, c->g);
  FS_LAUNCHED();
  if (fs_launch_neumann(c, c->f[FSGPU_RHO_U], c->g.nx + 1, c->g.ny)) return 1;
  if (fs_launch_neumann(c, c->f[FSGPU_RHO_V], c->g.nx, c->g.ny + 1)) return 1;
  const int fl[2] = {FSGPU_RHO_U, FSGPU_RHO_V};
  return fs_exchange_fields(c, fl, 2, FS_RO);
}

int fs_launch_pressure_jump(fsgpu_ctx* c) {
  if (ensure_fields(c, {FSGPU_CURV, FSGPU_INTERFACE_LENGTH})) return 1;
  const size_t bytes = (size_t)c->g.count() * sizeof(double);
  FS_CUDA(cudaMemsetAsync(alloc_base(c->g, c->f[FSGPU_P_JUMP_U]), 0, bytes, c->stream));
  FS_CUDA(cudaMemsetAsync(alloc_base(c->g, c->f[FSGPU_P_JUMP_V]), 0, bytes, c->stream));
  k_pressure_jump<<<face_grid(c->g), dim3(BX, BY), 0, c->stream>>>(
      c->f[FSGPU_VF], c->f[FSGPU_CURV], c->f[FSGPU_INTERFACE_LENGTH], c->f[FSGPU_P_JUMP_U],
      c->f[FSGPU_P_JUMP_V], c->sigma, c->g);
  FS_LAUNCHED();
  const int fl[2] = {FSGPU_P_JUMP_U, FSGPU_P_JUMP_V};
  return fs_exchange_fields(c, fl, 2, 1);
}

int fs_launch_capillary_divergence(fsgpu_ctx* c, double dt) {
  if (ensure_fields(c, {FSGPU_PJ_OLD_U, FSGPU_PJ_OLD_V})) return 1;
  const size_t bytes = (size_t)c->g.count() * sizeof(double);
  FS_CUDA(cudaMemcpyAsync(alloc_base(c->g, c->f[FSGPU_PJ_OLD_U]), alloc_base(c->g, c->f[FSGPU_P_JUMP_U]),
                          bytes, cudaMemcpyDeviceToDevice, c->stream));
  FS_CUDA(cudaMemcpyAsync(alloc_base(c->g, c->f[FSGPU_PJ_OLD_V]), alloc_base(c->g, c->f[FSGPU_P_JUMP_V]),
                          bytes, cudaMemcpyDeviceToDevice, c->stream));
  if (fs_launch_pressure_jump(c)) return 1;
  if (fs_launch_neumann(c, c->f[FSGPU_P_JUMP_U], c->g.nx + 1, c->g.ny)) return 1;
  if (fs_launch_neumann(c, c->f[FSGPU_P_JUMP_V], c->g.nx, c->g.ny + 1)) return 1;
  const int fl[2] = {FSGPU_P_JUMP_U, FSGPU_P_JUMP_V};
  if (fs_exchange_fields(c, fl, 2, FS_RO)) return 1;
  k_capillary_div<<<cell_grid(c->g), dim3(BX, BY), 0, c->stream>>>(
      c->f[FSGPU_P_JUMP_U], c->f[FSGPU_P_JUMP_V], c->f[FSGPU_PJ_OLD_U], c->f[FSGPU_PJ_OLD_V],
      c->f[FSGPU_RHO_U], c->f[FSGPU_RHO_V], c->f[FSGPU_DIV], dt, c->g);
  FS_LAUNCHED();
  return 0;
}

int fs_launch_ib_implicit_euler(fsgpu_ctx* c, double dt) {
  if (ensure_fields(c, {FSGPU_IB_CORR_U, FSGPU_IB_CORR_V})) return 1;
  k_ib_implicit_euler<<<face_grid(c->g), dim3(BX, BY), 0, c->stream>>>(
      c->f[FSGPU_U], c->f[FSGPU_V], c->f[FSGPU_VISC], c->f[FSGPU_RHO_U], c->f[FSGPU_RHO_V],
      c->f[FSGPU_IB_CORR_U], c->f[FSGPU_IB_CORR_V], dt, c->g);
  FS_LAUNCHED();
  return 0;
}

int fs_launch_ib_semi_analytical(fsgpu_ctx* c, double dt) {
  if (ensure_fields(c, {FSGPU_IB_CORR_U, FSGPU_IB_CORR_V})) return 1;
  IbArgs a;
  a.U = c->f[FSGPU_U]; a.V = c->f[FSGPU_V]; a.Uo = c->f[FSGPU_U_OLD]; a.Vo = c->f[FSGPU_V_OLD];
  a.ruo = c->f[FSGPU_RHO_U_OLD]; a.rvo = c->f[FSGPU_RHO_V_OLD];
  a.ru = c->f[FSGPU_RHO_U]; a.rv = c->f[FSGPU_RHO_V]; a.mu = c->f[FSGPU_VISC];
  a.dU = c->f[FSGPU_DMOM_U]; a.dV = c->f[FSGPU_DMOM_V];
  a.ibu = c->f[FSGPU_IB_CORR_U]; a.ibv = c->f[FSGPU_IB_CORR_V];
  a.dt = dt;
  k_ib_semi_analytical<<<face_grid(c->g), dim3(BX, BY), 0, c->stream>>>(a, c->g);
  FS_LAUNCHED();
  return 0;
}

int fs_launch_interpolate(fsgpu_ctx* c) {
  if (ensure_fields(c, {FSGPU_UI, FSGPU_VI, FSGPU_RHOI})) return 1;
  k_interpolate<<<cell_grid(c->g), dim3(BX, BY), 0, c->stream>>>(
      c->f[FSGPU_U], c->f[FSGPU_V], c->f[FSGPU_RHO_U], c->f[FSGPU_RHO_V], c->f[FSGPU_UI],
      c->f[FSGPU_VI], c->f[FSGPU_RHOI], c->g);
  FS_LAUNCHED();
  return 0;
}

int fs_launch_conserved(fsgpu_ctx* c) {  // -> red_out[4..6]
  const unsigned gr = persist_blocks(c->g.a_hi - c->g.a_lo, c->g.ny);
  if (fs_ensure_part(c, (size_t)gr * 4)) return 1;
  k_conserved<<<gr, dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_U], c->f[FSGPU_V], c->f[FSGPU_RHO_U],
                                                  c->f[FSGPU_RHO_V], c->g, c->part, c->ticket,
                                                  c->red_out + 4);
  FS_LAUNCHED();
  return comm_allreduce(c, c->red_out + 4, 3, COMM_SUM);
}

int fs_launch_grad_centered(fsgpu_ctx* c, int f, int dfdx, int dfdy) {
  if (fs_exchange_fields(c, &f, 1, FS_RO)) return 1;  // neighbours a +- 1 (a +- 2 on the first / last row)
  k_grad_centered<<<cell_grid(c->g), dim3(BX, BY), 0, c->stream>>>(c->f[f], c->f[dfdx], c->f[dfdy], c->g);
  FS_LAUNCHED();
  const int fl[2] = {dfdx, dfdy};
  return fs_exchange_fields(c, fl, 2, FS_RO);
}

// host arrays in, host arrays out: xm, ym are fs.xm / fs.ym incl. ghosts (NX + 2 and NY + 2 values)
int fs_sample_points(fsgpu_ctx* c, int mode, int fa, int fb, const double* xm, const double* ym, int64_t npts,
                     const double* px, const double* py, double* oa, double* ob) {
  const int64_t nx = c->g.nx, ny = c->g.ny;
  if (npts <= 0) return 0;
  std::vector<double> h((size_t)(nx + ny + 4));
  std::copy(xm, xm + nx + 2, h.begin());
  std::copy(ym, ym + ny + 2, h.begin() + (nx + 2));
  {  // row ip + 2 may be the first halo row of this slab
    const int fl[2] = {fa, fb};
    if (fs_exchange_fields(c, fl, mode == 0 ? 2 : 1, 1)) return 1;
  }
  double* d = nullptr;
  const size_t nc = (size_t)(nx + ny + 4), np = (size_t)npts;
  FS_CUDA(cudaMalloc(&d, (nc + 4 * np) * sizeof(double)));
  int rc = cudaMemcpyAsync(d, h.data(), nc * sizeof(double), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
           cudaMemcpyAsync(d + nc, px, np * sizeof(double), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
           cudaMemcpyAsync(d + nc + np, py, np * sizeof(double), cudaMemcpyHostToDevice, c->stream) != cudaSuccess;
  if (!rc) {
    SampleArgs s{c->f[fa], mode == 0 ? c->f[fb] : nullptr, d, d + nx + 2, d + nc, d + nc + np, d + nc + 2 * np,
                 d + nc + 3 * np, npts};
    const unsigned nb = (unsigned)((npts + 255) / 256);
    if (mode == 0) k_sample<0><<<nb, 256, 0, c->stream>>>(s, c->g);
    else k_sample<1><<<nb, 256, 0, c->stream>>>(s, c->g);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    rc = cudaPeekAtLastError() != cudaSuccess ||
         comm_allreduce(c, d + nc + 2 * np, (int64_t)(mode == 0 ? 2 * np : np), COMM_SUM) != 0 ||
         cudaMemcpyAsync(oa, d + nc + 2 * np, np * sizeof(double), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
         (mode == 0 && cudaMemcpyAsync(ob, d + nc + 3 * np, np * sizeof(double), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess);
  }
  rc = rc || cudaStreamSynchronize(c->stream) != cudaSuccess;
  cudaFree(d);
  if (rc) fs_set_error("point sampling: %s", cudaGetErrorString(cudaGetLastError()));
  return rc ? 1 : 0;
}

int fs_launch_monitor(fsgpu_ctx* c) {  // -> red_out[4..6]
  const unsigned gr = persist_blocks(c->g.a_hi - c->g.a_lo, c->g.ny + 2);
  if (fs_ensure_part(c, (size_t)gr * 4)) return 1;
  k_monitor<<<gr, dim3(BX, BY), 0, c->stream>>>(c->f[FSGPU_U], c->f[FSGPU_V], c->f[FSGPU_DIV], c->g,
                                                c->part, c->ticket, c->red_out + 4);
  FS_LAUNCHED();
  return comm_allreduce(c, c->red_out + 4, 3, COMM_MAX);
}
