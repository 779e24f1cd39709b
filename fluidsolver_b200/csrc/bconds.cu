// fluidsolver_b200/csrc/bconds.cu -- velocity boundary conditions on the ghost ring.
//
// Replaces apply_velocity_bconds (src/BoundaryConditions.hpp:215-224) with its four kinds
// Dirichlet (:14-90), Neumann (:93-139), Periodic (:142-167), Symmetry (:170-206), the scalar
// apply_neumann_bconds (:227-252) and the channel drivers' outflow fix
// (test/LaminarChannel.cpp:55-66,201-205).
//
// The reference applies left, right, bottom, top in that order; bottom/top run over the ghost
// inclusive i-range and read the values left/right have just written at i = -1 / NX(+1), so the
// work is two launches: k_bc_lr (left + right, which never touch each other's rows) and then
// k_bc_bt (bottom + top).
#include "ctx.h"

namespace {

struct BcArgs {
  int kind[4];
  double u[4], v[4];
  int clipped[4];
  const double* up[4];
  const double* vp[4];
  double *U, *V;
};

// left (s = 0) and right (s = 1): thread b = j + 1 in [0, NY + 2]
__global__ void k_bc_lr(const BcArgs a, const Geom g) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int s     = blockIdx.y;  // 0 left, 1 right
  if (b > g.ny + 2) return;
  // the left rows live on the first slab, the right rows on the last one.  x-periodic sides of a
  // decomposed grid are filled by the wrap-around exchange in fs_launch_bconds instead.
  if (s == 0 ? !g.first() : !g.last()) return;
  if (a.kind[s] == FSGPU_BC_PERIODIC && !(g.first() && g.last())) return;
  const int64_t P = g.pitch;
  const bool in_u = b <= g.ny + 1;  // U rows have NY+2 entries, V rows NY+3
  // U: ghost row ag, boundary face row af, first interior-side row a1; V: ghost row vg, inner vi
  const int64_t ag = s == 0 ? 0 : g.nx + 2, af = s == 0 ? 1 : g.nx + 1, a1 = s == 0 ? 2 : g.nx;
  const int64_t vg = s == 0 ? 0 : g.nx + 1, vi = s == 0 ? 1 : g.nx;
  double* U = a.U;
  double* V = a.V;
  switch (a.kind[s]) {
    case FSGPU_BC_DIRICHLET: {
      if (in_u) {
        const double ub = a.up[s] ? a.up[s][b] : a.u[s];
        U[g.idx(ag, b)] = ub;
        U[g.idx(af, b)] = ub;
      }
      const double vb = a.vp[s] ? a.vp[s][b] : a.v[s];
      V[g.idx(vg, b)] = 2.0 * vb - V[g.idx(vi, b)];
    } break;
    case FSGPU_BC_NEUMANN: {
      if (in_u) {
        double x = U[g.idx(af, b)];
        if (a.clipped[s]) x = s == 0 ? fmin(x, 0.0) : fmax(x, 0.0);
        U[g.idx(ag, b)] = x;
      }
      V[g.idx(vg, b)] = V[g.idx(vi, b)];
    } break;
    case FSGPU_BC_PERIODIC: {
      // left: U(-1,j) = U(NX-1,j), V(-1,j) = V(NX-1,j); right: U(NX+1,j) = U(1,j), V(NX,j) = V(0,j)
      if (in_u) U[g.idx(ag, b)] = U[g.idx(s == 0 ? g.nx : 2, b)];
      V[g.idx(vg, b)] = V[g.idx(s == 0 ? g.nx : 1, b)];
    } break;
    default: {  // symmetry
      if (in_u) {
        U[g.idx(ag, b)] = -U[g.idx(a1, b)];
        U[g.idx(af, b)] = 0.0;
      }
      V[g.idx(vg, b)] = V[g.idx(vi, b)];
    } break;
  }
  (void)P;
}

// bottom (s = 2) and top (s = 3): thread a = i + 1 in [0, NX + 2]
__global__ void k_bc_bt(const BcArgs a, const Geom g) {
  const int64_t r = g.a_lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int s     = 2 + blockIdx.y;
  if (r >= g.u_hi()) return;
  const bool in_v = r <= g.nx + 1;  // V has NX+2 rows, U NX+3
  const int64_t ug = s == 2 ? 0 : g.ny + 1, ui = s == 2 ? 1 : g.ny;                      // U cols
  const int64_t bg = s == 2 ? 0 : g.ny + 2, bf = s == 2 ? 1 : g.ny + 1, b1 = s == 2 ? 2 : g.ny;  // V cols
  double* U = a.U;
  double* V = a.V;
  switch (a.kind[s]) {
    case FSGPU_BC_DIRICHLET: {
      const double ub = a.up[s] ? a.up[s][r] : a.u[s];
      U[g.idx(r, ug)] = 2.0 * ub - U[g.idx(r, ui)];
      if (in_v) {
        const double vb = a.vp[s] ? a.vp[s][r] : a.v[s];
        V[g.idx(r, bg)] = vb;
        V[g.idx(r, bf)] = vb;
      }
    } break;
    case FSGPU_BC_NEUMANN: {
      U[g.idx(r, ug)] = U[g.idx(r, ui)];
      if (in_v) {
        double x = V[g.idx(r, bf)];
        if (a.clipped[s]) x = s == 2 ? fmin(x, 0.0) : fmax(x, 0.0);
        V[g.idx(r, bg)] = x;
      }
    } break;
    case FSGPU_BC_PERIODIC: {
      // bottom: U(i,-1) = U(i,NY-1), V(i,-1) = V(i,NY-1); top: U(i,NY) = U(i,0), V(i,NY+1) = V(i,1)
      U[g.idx(r, ug)] = U[g.idx(r, s == 2 ? g.ny : 1)];
      if (in_v) V[g.idx(r, bg)] = V[g.idx(r, s == 2 ? g.ny : 2)];
    } break;
    default: {  // symmetry
      U[g.idx(r, ug)] = U[g.idx(r, ui)];
      if (in_v) {
        V[g.idx(r, bg)] = -V[g.idx(r, b1)];
        V[g.idx(r, bf)] = 0.0;
      }
    } break;
  }
}

// The channel drivers' corrections of the two ghost columns after apply_velocity_bconds:
//   mode 1  test/LaminarChannel.cpp:55-66 (sums over j = -1..NY) and :201-205
//   mode 2  test/CustomDirichletChannel.cpp:61-72,197-202 = scaling/Channel.cpp:65-76,208-213: the sums carry
//           a factor dy, the correction formula is the same
//   mode 3  test/PeriodicChannel.cpp:56-67,193-203 = scaling/Channel.cpp:196-206: both columns are shifted so
//           that each carries total_flow
// Two one-block kernels so that a decomposed grid can combine the inlet sum (first slab) and the
// outlet sum (last slab) in between: io[0] = inflow, io[1] = outflow (w = 1 or dy per term).
__global__ void k_outflow_sums(const double* __restrict__ U, const double* __restrict__ ru, const Geom g,
                               double w, double* io) {
  __shared__ double sm[32];
  double in = 0.0, out = 0.0;
  for (int64_t b = threadIdx.x; b <= g.ny + 1; b += blockDim.x) {
    if (g.first()) in += ru[g.idx(0, b)] * U[g.idx(0, b)] * w;
    if (g.last()) out += ru[g.idx(g.nx + 2, b)] * U[g.idx(g.nx + 2, b)] * w;
  }
  in  = block_reduce<ReduceSum>(in, sm);
  out = block_reduce<ReduceSum>(out, sm);
  if (threadIdx.x == 0) {
    io[0] = in;
    io[1] = out;
  }
}
__global__ void k_outflow_apply(double* __restrict__ U, const double* __restrict__ ru, const Geom g, int mode,
                                double total_flow, const double* io, double* mass_error_out) {
  const double me = io[1] - io[0];
  if (threadIdx.x == 0 && mass_error_out) *mass_error_out = me;
  const double n2 = (double)(g.ny + 2);
  if (mode == FSGPU_FLOW_FORCE_TOTAL) {
    const double ie = total_flow - io[0], oe = total_flow - io[1];
    for (int64_t b = threadIdx.x; b <= g.ny + 1; b += blockDim.x) {
      if (g.first()) {
        const int64_t k = g.idx(0, b);
        U[k] += ie / (ru[k] * g.dy * n2);
      }
      if (g.last()) {
        const int64_t k = g.idx(g.nx + 2, b);
        U[k] += oe / (ru[k] * g.dy * n2);
      }
    }
    return;
  }
  if (!g.last()) return;
  for (int64_t b = threadIdx.x; b <= g.ny + 1; b += blockDim.x) {
    const int64_t k = g.idx(g.nx + 2, b);
    U[k] -= me / (ru[k] * n2);
  }
}

// BoundaryConditions.hpp:227-252 on an (m+2) x (n+2) field: left/right first, then bottom/top
__global__ void k_neumann_lr(double* f, int64_t m, int64_t n, const Geom g) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b > n + 1) return;
  if (g.first()) f[g.idx(0, b)] = f[g.idx(1, b)];
  if (g.last()) f[g.idx(m + 1, b)] = f[g.idx(m, b)];
}
__global__ void k_neumann_bt(double* f, int64_t m, int64_t n, const Geom g) {
  const int64_t a = g.a_lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (a > m + 1 || a >= g.u_hi()) return;
  f[g.idx(a, 0)]     = f[g.idx(a, 1)];
  f[g.idx(a, n + 1)] = f[g.idx(a, n)];
}

}  // namespace

int fs_launch_bconds(fsgpu_ctx* c) {
  BcArgs a;
  for (int s = 0; s < 4; ++s) {
    a.kind[s] = c->bc[s].kind; a.u[s] = c->bc[s].u; a.v[s] = c->bc[s].v;
    a.clipped[s] = c->bc[s].clipped; a.up[s] = c->bc[s].u_prof; a.vp[s] = c->bc[s].v_prof;
  }
  a.U = c->f[FSGPU_U];
  a.V = c->f[FSGPU_V];
  k_bc_lr<<<dim3((unsigned)cdiv(c->g.ny + 3, 128), 2), 128, 0, c->stream>>>(a, c->g);
  FS_LAUNCHED();
  const Geom& g = c->g;
  if (c->nranks > 1 && (c->bc[FSGPU_LEFT].kind == FSGPU_BC_PERIODIC || c->bc[FSGPU_RIGHT].kind == FSGPU_BC_PERIODIC)) {
    // x-periodic wrap-around between the first and the last slab (BoundaryConditions.hpp:142-167):
    // left ghosts U(-1,.) = U(NX-1,.), V(-1,.) = V(NX-1,.); right ghosts U(NX+1,.) = U(1,.), V(NX,.) = V(0,.)
    if (c->bc[FSGPU_LEFT].kind != c->bc[FSGPU_RIGHT].kind) {
      fs_set_error("periodic left/right sides must come as a pair");
      return 1;
    }
    const double* snd[2];
    double* rcv[2];
    int peer = -1;
    if (g.first()) {
      peer = c->nranks - 1;
      snd[0] = a.U + g.idx(2, 0) - FS_CO;  rcv[0] = a.U + g.idx(0, 0) - FS_CO;
      snd[1] = a.V + g.idx(1, 0) - FS_CO;  rcv[1] = a.V + g.idx(0, 0) - FS_CO;
    } else if (g.last()) {
      peer = 0;
      snd[0] = a.U + g.idx(g.nx, 0) - FS_CO;  rcv[0] = a.U + g.idx(g.nx + 2, 0) - FS_CO;
      snd[1] = a.V + g.idx(g.nx, 0) - FS_CO;  rcv[1] = a.V + g.idx(g.nx + 1, 0) - FS_CO;
    }
    if (comm_sendrecv(c, peer, snd, rcv, 2, g.pitch)) return 1;
  }
  k_bc_bt<<<dim3((unsigned)cdiv(g.u_hi() - g.a_lo, 128), 2), 128, 0, c->stream>>>(a, g);
  FS_LAUNCHED();
  const int fl[2] = {FSGPU_U, FSGPU_V};
  return fs_exchange_fields(c, fl, 2, 1);
}

// io (device, 2 doubles) keeps inflow and outflow as they were before the correction
int fs_launch_flow_control(fsgpu_ctx* c, int mode, double total_flow, double* d_mass_error) {
  double* io = c->red_out + 8;
  const double w = mode == FSGPU_FLOW_OUTFLOW_FIX ? 1.0 : c->g.dy;
  k_outflow_sums<<<1, 1024, 0, c->stream>>>(c->f[FSGPU_U], c->f[FSGPU_RHO_U], c->g, w, io);
  FS_LAUNCHED();
  if (comm_allreduce(c, io, 2, COMM_SUM)) return 1;  // inflow lives on the first slab, outflow on the last
  k_outflow_apply<<<1, 1024, 0, c->stream>>>(c->f[FSGPU_U], c->f[FSGPU_RHO_U], c->g, mode, total_flow, io,
                                             d_mass_error);
  FS_LAUNCHED();
  return 0;
}

int fs_launch_neumann(fsgpu_ctx* c, double* f, int64_t m, int64_t n) {
  k_neumann_lr<<<(unsigned)cdiv(n + 2, 128), 128, 0, c->stream>>>(f, m, n, c->g);
  FS_LAUNCHED();
  k_neumann_bt<<<(unsigned)cdiv(c->g.u_hi() - c->g.a_lo, 128), 128, 0, c->stream>>>(f, m, n, c->g);
  FS_LAUNCHED();
  return 0;
}
