// fluidsolver_b200/csrc/stencil.cu -- staggered-mesh stencil kernels of the timestep path.
//
// Reference operations restated as CUDA kernels (no code shared with the reference; each kernel
// cites what it replaces):
//   k_momentum<...>      calc_mid_time + calc_dmomdt + update_velocity   Operators.hpp:44-49,
//                        FS.hpp:158-315, FS.hpp:423-436 (fusion pattern: bench/update-velo-fuse.cpp)
// (boundary conditions and the channel flow corrections: bconds.cu; divergence, rhs, projection and the
// reductions: project.cu)
#include "ctx.h"

#include <math.h>

#include <algorithm>
#include <utility>

namespace {

constexpr int BX = 64, BY = 4;  // thread block: 64 columns (j, contiguous) x 4 rows (i)

// src/FS.hpp:132-150
__device__ __forceinline__ void hybrid_interp(double eps, double rm, double rp, double vm, double vp,
                                              double tm, double tp, double& rho, double& vel) {
  if (!(fabs(rp - rm) > eps)) {
    rho = (rp + rm) / 2.0;
    vel = (vp + vm) / 2.0;
  } else if (tp + tm >= 0.0) {
    rho = rm;
    vel = vm;
  } else {
    rho = rp;
    vel = vp;
  }
}

struct MomArgs {
  const double *U, *V, *Uo, *Vo;    // curr (or mid-time) and old velocities
  const double *ruo, *rvo;          // old densities
  const double *ru, *rv;            // curr densities
  const double *mu, *p, *pju, *pjv;
  double *outU, *outV;              // dmom (MODE 0) or new velocity (MODE 1/2)
  double dx, dy, dt, eps;
  double rdx, rdy;  // 1/dx, 1/dy: the fused pass multiplies (<= 1 ulp per quotient, tolerance 1e-12)
};

// MID: velocities are averaged with the old ones on the fly (calc_mid_time fused in).
template <bool MID>
struct Vel {
  const double *c, *o;
  __device__ __forceinline__ double operator()(int64_t k) const {
    if (MID) return 0.5 * (__ldg(c + k) + __ldg(o + k));
    return __ldg(c + k);
  }
};

// The four fluxes of src/FS.hpp:180-254.  k is the linear index of (i, j); P the pitch.
template <bool MID>
__device__ __forceinline__ double flux_FXU(const MomArgs& a, const Vel<MID>& U, int64_t k, int64_t P) {
  const double um = U(k), up = U(k + P);
  double rh, uh;
  hybrid_interp(a.eps, __ldg(a.ruo + k), __ldg(a.ruo + k + P), um, up, um, up, rh, uh);
  const double Ui   = (up + um) / 2;
  const double dUdx = MID ? (up - um) * a.rdx : (up - um) / a.dx;
  return -rh * uh * Ui + 2.0 * __ldg(a.mu + k) * dUdx - __ldg(a.p + k);
}
template <bool MID>
__device__ __forceinline__ double flux_FYV(const MomArgs& a, const Vel<MID>& V, int64_t k, int64_t P) {
  const double vm = V(k), vp = V(k + 1);
  double rh, vh;
  hybrid_interp(a.eps, __ldg(a.rvo + k), __ldg(a.rvo + k + 1), vm, vp, vm, vp, rh, vh);
  const double Vi   = (vm + vp) / 2;
  const double dVdy = MID ? (vp - vm) * a.rdy : (vp - vm) / a.dy;
  return -rh * vh * Vi + 2.0 * __ldg(a.mu + k) * dVdy - __ldg(a.p + k);
}
// corner fluxes FYU(i,j) and FXV(i,j) share the corner viscosity and the shear rate
template <bool MID>
__device__ __forceinline__ void flux_corner(const MomArgs& a, const Vel<MID>& U, const Vel<MID>& V,
                                            int64_t k, int64_t P, double& fyu, double& fxv) {
  const double u0 = U(k - 1), u1 = U(k);   // U(i,j-1), U(i,j)
  const double v0 = V(k - P), v1 = V(k);   // V(i-1,j), V(i,j)
  const double vc = (__ldg(a.mu + k) + __ldg(a.mu + k - P) + __ldg(a.mu + k - 1) +
                     __ldg(a.mu + k - P - 1)) / 4.0;
  const double dUdy = MID ? (u1 - u0) * a.rdy : (u1 - u0) / a.dy;
  const double dVdx = MID ? (v1 - v0) * a.rdx : (v1 - v0) / a.dx;
  const double sh   = vc * (dUdy + dVdx);
  double rh, h;
  hybrid_interp(a.eps, __ldg(a.ruo + k - 1), __ldg(a.ruo + k), u0, u1, v0, v1, rh, h);
  fyu = -rh * h * ((v0 + v1) / 2) + sh;
  hybrid_interp(a.eps, __ldg(a.rvo + k - P), __ldg(a.rvo + k), v0, v1, u0, u1, rh, h);
  fxv = -rh * ((u1 + u0) / 2) * h + sh;
}

// MODE 0: outU/outV = dmomUdt/dmomVdt (calc_dmomdt);  MODE 1: mid-time + dmomdt + update fused.
template <int MODE>
__global__ void __launch_bounds__(BX* BY)
k_momentum(const MomArgs a, const Geom g) {
  const int64_t j = (int64_t)blockIdx.x * BX + threadIdx.x;  // 0..NY
  // 0..NX, restricted to the face rows a = i + 1 this slab owns
  const int64_t i = (g.a_lo > 1 ? g.a_lo : 1) - 1 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (i > g.nx || i + 1 >= g.a_hi || j > g.ny) return;
  const int64_t P = g.pitch, k = g.idx(i + 1, j + 1);
  constexpr bool MID = (MODE == 1);
  const Vel<MID> U{a.U, a.Uo}, V{a.V, a.Vo};
  double fyu00, fxv00;                      // corner (i, j)
  flux_corner<MID>(a, U, V, k, P, fyu00, fxv00);
  if (j < g.ny) {                           // U face (i, j), i in [0, NX], j in [0, NY)
    double fyu01, t;                        // corner (i, j+1)
    flux_corner<MID>(a, U, V, k + 1, P, fyu01, t);
    const double fxu0 = flux_FXU<MID>(a, U, k, P), fxum = flux_FXU<MID>(a, U, k - P, P);
    const double dm = (MID ? (fxu0 - fxum) * a.rdx + (fyu01 - fyu00) * a.rdy
                           : (fxu0 - fxum) / a.dx + (fyu01 - fyu00) / a.dy) + __ldg(a.pju + k);
    if (MODE == 0) a.outU[k] = dm;
    else a.outU[k] = (__ldg(a.ruo + k) * __ldg(a.Uo + k) + a.dt * dm) / __ldg(a.ru + k);
  }
  if (i < g.nx) {                           // V face (i, j), i in [0, NX), j in [0, NY]
    double fxv10, t;                        // corner (i+1, j)
    flux_corner<MID>(a, U, V, k + P, P, t, fxv10);
    const double fyv0 = flux_FYV<MID>(a, V, k, P), fyvm = flux_FYV<MID>(a, V, k - 1, P);
    const double dm = (MID ? (fxv10 - fxv00) * a.rdx + (fyv0 - fyvm) * a.rdy
                           : (fxv10 - fxv00) / a.dx + (fyv0 - fyvm) / a.dy) + __ldg(a.pjv + k);
    if (MODE == 0) a.outV[k] = dm;
    else a.outV[k] = (__ldg(a.rvo + k) * __ldg(a.Vo + k) + a.dt * dm) / __ldg(a.rv + k);
  }
}

// cur = 0.5 (cur + old) on every entry incl. ghosts (pads are 0 and stay 0). Operators.hpp:44-49
__global__ void k_mid_time(double* __restrict__ cu, const double* __restrict__ ou,
                           double* __restrict__ cv, const double* __restrict__ ov, int64_t n) {
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n;
       k += (int64_t)gridDim.x * blockDim.x) {
    cu[k] = 0.5 * (cu[k] + ou[k]);
    cv[k] = 0.5 * (cv[k] + ov[k]);
  }
}

// FS.hpp:423-436 (un-fused form)
__global__ void k_update_velocity(const MomArgs a, const double* __restrict__ dU,
                                  const double* __restrict__ dV, double* __restrict__ U,
                                  double* __restrict__ V, const Geom g) {
  const int64_t j = (int64_t)blockIdx.x * BX + threadIdx.x;
  const int64_t i = (g.a_lo > 1 ? g.a_lo : 1) - 1 + (int64_t)blockIdx.y * BY + threadIdx.y;
  if (i > g.nx || i + 1 >= g.a_hi || j > g.ny) return;
  const int64_t k = g.idx(i + 1, j + 1);
  if (j < g.ny) U[k] = (a.ruo[k] * a.Uo[k] + a.dt * dU[k]) / a.ru[k];
  if (i < g.nx) V[k] = (a.rvo[k] * a.Vo[k] + a.dt * dV[k]) / a.rv[k];
}

}  // namespace

static MomArgs mom_args(fsgpu_ctx* c, double dt) {
  MomArgs a;
  a.U = c->f[FSGPU_U]; a.V = c->f[FSGPU_V]; a.Uo = c->f[FSGPU_U_OLD]; a.Vo = c->f[FSGPU_V_OLD];
  a.ruo = c->f[FSGPU_RHO_U_OLD]; a.rvo = c->f[FSGPU_RHO_V_OLD];
  a.ru = c->f[FSGPU_RHO_U]; a.rv = c->f[FSGPU_RHO_V];
  a.mu = c->f[FSGPU_VISC]; a.p = c->f[FSGPU_P];
  a.pju = c->f[FSGPU_P_JUMP_U]; a.pjv = c->f[FSGPU_P_JUMP_V];
  a.outU = c->f[FSGPU_DMOM_U]; a.outV = c->f[FSGPU_DMOM_V];
  a.dx = c->g.dx; a.dy = c->g.dy; a.dt = dt;
  a.rdx = 1.0 / c->g.dx; a.rdy = 1.0 / c->g.dy;
  a.eps = 1e-3 * fmin(c->rho_gas, c->rho_liquid);  // calc_rho_eps, FS.hpp:153-155
  return a;
}
static dim3 face_grid(const Geom& g) {
  return dim3((unsigned)cdiv(g.ny + 1, BX), (unsigned)cdiv(g.a_hi - g.a_lo, BY));
}

int fs_launch_mid_time(fsgpu_ctx* c) {
  const int64_t n = c->g.count();
  // whole local allocation (the reference averages every entry incl. ghosts; halo rows are
  // averaged redundantly, which keeps them equal to the neighbour's rows)
  const Geom& g = c->g;
  k_mid_time<<<(unsigned)std::min<int64_t>(cdiv(n, 256), 148 * 16), 256, 0, c->stream>>>(
      alloc_base(g, c->f[FSGPU_U]), alloc_base(g, c->f[FSGPU_U_OLD]), alloc_base(g, c->f[FSGPU_V]),
      alloc_base(g, c->f[FSGPU_V_OLD]), n);
  FS_LAUNCHED();
  return 0;
}

int fs_launch_dmomdt(fsgpu_ctx* c) {
  // ghosts of the outputs are 0 (fill(dmomUdt, 0), FS.hpp:173-174): only interior is written and
  // the rest of the array is never touched after allocation (zero-initialised).
  k_momentum<0><<<face_grid(c->g), dim3(BX, BY), 0, c->stream>>>(mom_args(c, 0.0), c->g);
  FS_LAUNCHED();
  return 0;
}

int fs_launch_update_velocity(fsgpu_ctx* c, double dt) {
  k_update_velocity<<<face_grid(c->g), dim3(BX, BY), 0, c->stream>>>(
      mom_args(c, dt), c->f[FSGPU_DMOM_U], c->f[FSGPU_DMOM_V], c->f[FSGPU_U], c->f[FSGPU_V], c->g);
  FS_LAUNCHED();
  return 0;
}

int fs_launch_momentum_update(fsgpu_ctx* c, double dt) {
  MomArgs a = mom_args(c, dt);
  a.outU = c->U_alt;
  a.outV = c->V_alt;
  k_momentum<1><<<face_grid(c->g), dim3(BX, BY), 0, c->stream>>>(a, c->g);
  FS_LAUNCHED();
  // ping-pong: the new interior faces live in the alternate buffers; every ghost entry is
  // rewritten by apply_velocity_bconds, which the caller runs next (as every reference driver).
  std::swap(c->f[FSGPU_U], c->U_alt);
  std::swap(c->f[FSGPU_V], c->V_alt);
  return 0;
}
