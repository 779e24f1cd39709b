/* oracle/fs_oracle.c -- TEST INFRASTRUCTURE ONLY (checker, never product; see fs_oracle.h).
 *
 * Plain-C, scalar, single-threaded restatement of the stencil operators on the reference's
 * timestep path.  Each function cites the reference lines it follows.  Arithmetic is written
 * in the same operation order as the reference; the library is built with
 * -ffp-contract=off, so it is the un-fused (no FMA) evaluation of those expressions.
 *
 * Parity pin: tests/test_oracle_vs_ref.py checks every function here against the UNMODIFIED
 * reference compiled into oracle/_ref/libfsref.so, and tests/golden/ holds outputs of that
 * library for the GPU box, where /root/reference does not exist.
 */
#include "fs_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define NX (fs->nx)
#define NY (fs->ny)
/* accessors in the reference's storage (src/Container.hpp:105-111) */
#define IU(i, j) ((j) + 1 + ((i) + 1) * (NY + 2))  /* U-like: (NX+3) x (NY+2) */
#define IV(i, j) ((j) + 1 + ((i) + 1) * (NY + 3))  /* V-like: (NX+2) x (NY+3) */
#define IC(i, j) ((j) + 1 + ((i) + 1) * (NY + 2))  /* cell:   (NX+2) x (NY+2) */
/* Layout::F cell index (i fastest), src/Container.hpp:109 */
#define IF(i, j) ((i) + 1 + ((j) + 1) * (NX + 2))

static inline double dmin(double a, double b) { return a < b ? a : b; }
static inline double dmax(double a, double b) { return a > b ? a : b; }
static inline double sqr(double a) { return a * a; }

/* src/FS.hpp:100-128 (note: no absolute value on the convective terms, as in the reference) */
double fso_adjust_dt(const fso_fs* fs, double cfl_max, double dt_max) {
  double cx = 0.0, cy = 0.0, vx = 0.0, vy = 0.0, st = 0.0;
  if (fs->sigma > 0.0) {
    st = 1.0 / sqrt(((fs->rho_gas + fs->rho_liquid) * pow(fs->dx * fs->dy, 3.0 / 2.0)) /
                    (4.0 * M_PI * fs->sigma));
  }
  for (int64_t i = 0; i < NX; ++i) {
    for (int64_t j = 0; j < NY; ++j) {
      cx = dmax(cx, (fs->U[IU(i, j)] + fs->U[IU(i + 1, j)]) / 2 / fs->dx);
      cy = dmax(cy, (fs->V[IV(i, j)] + fs->V[IV(i, j + 1)]) / 2 / fs->dy);
      const double rho = (fs->rho_u[IU(i, j)] + fs->rho_u[IU(i + 1, j)] + fs->rho_v[IV(i, j)] +
                          fs->rho_v[IV(i, j + 1)]) /
                         4.0;
      vx = dmax(vx, 4.0 * fs->visc[IC(i, j)] / (sqr(fs->dx) * rho));
      vy = dmax(vy, 4.0 * fs->visc[IC(i, j)] / (sqr(fs->dy) * rho));
    }
  }
  const double m = dmax(dmax(dmax(cx, cy), dmax(vx, vy)), st);
  return dmin(cfl_max / m, dt_max);
}

/* src/FS.hpp:76-97 */
void fso_save_old_state(fso_fs* fs) {
  const size_t nu = (size_t)((NX + 3) * (NY + 2)), nv = (size_t)((NX + 2) * (NY + 3));
  memcpy(fs->rho_u_old, fs->rho_u, nu * sizeof(double));
  memcpy(fs->rho_v_old, fs->rho_v, nv * sizeof(double));
  memcpy(fs->U_old, fs->U, nu * sizeof(double));
  memcpy(fs->V_old, fs->V, nv * sizeof(double));
}

/* src/Operators.hpp:44-49: all entries incl. ghosts */
void fso_calc_mid_time(double* cur, const double* old, int64_t n) {
  for (int64_t k = 0; k < n; ++k) cur[k] = 0.5 * (cur[k] + old[k]);
}

/* src/FS.hpp:132-150 */
static inline void hybrid_interp(double eps, double rm, double rp, double vm, double vp, double tm,
                                 double tp, double* rho, double* vel) {
  const int upwind = fabs(rp - rm) > eps;
  if (!upwind) {
    *rho = (rp + rm) / 2.0;
    *vel = (vp + vm) / 2.0;
  } else if (tp + tm >= 0.0) {
    *rho = rm;
    *vel = vm;
  } else {
    *rho = rp;
    *vel = vp;
  }
}

/* the four flux lambdas of src/FS.hpp:180-254 */
static double flux_FXU(const fso_fs* fs, double eps, int64_t i, int64_t j) {
  double rh, uh;
  hybrid_interp(eps, fs->rho_u_old[IU(i, j)], fs->rho_u_old[IU(i + 1, j)], fs->U[IU(i, j)],
                fs->U[IU(i + 1, j)], fs->U[IU(i, j)], fs->U[IU(i + 1, j)], &rh, &uh);
  const double Ui   = (fs->U[IU(i + 1, j)] + fs->U[IU(i, j)]) / 2;
  const double dUdx = (fs->U[IU(i + 1, j)] - fs->U[IU(i, j)]) / fs->dx;
  return -rh * uh * Ui + 2.0 * fs->visc[IC(i, j)] * dUdx - fs->p[IC(i, j)];
}
static double flux_FYU(const fso_fs* fs, double eps, int64_t i, int64_t j) {
  double rh, uh;
  hybrid_interp(eps, fs->rho_u_old[IU(i, j - 1)], fs->rho_u_old[IU(i, j)], fs->U[IU(i, j - 1)],
                fs->U[IU(i, j)], fs->V[IV(i - 1, j)], fs->V[IV(i, j)], &rh, &uh);
  const double Vi = (fs->V[IV(i - 1, j)] + fs->V[IV(i, j)]) / 2;
  const double vc = (fs->visc[IC(i, j)] + fs->visc[IC(i - 1, j)] + fs->visc[IC(i, j - 1)] +
                     fs->visc[IC(i - 1, j - 1)]) /
                    4.0;
  const double dUdy = (fs->U[IU(i, j)] - fs->U[IU(i, j - 1)]) / fs->dy;
  const double dVdx = (fs->V[IV(i, j)] - fs->V[IV(i - 1, j)]) / fs->dx;
  return -rh * uh * Vi + vc * (dUdy + dVdx);
}
static double flux_FXV(const fso_fs* fs, double eps, int64_t i, int64_t j) {
  double rh, vh;
  hybrid_interp(eps, fs->rho_v_old[IV(i - 1, j)], fs->rho_v_old[IV(i, j)], fs->V[IV(i - 1, j)],
                fs->V[IV(i, j)], fs->U[IU(i, j - 1)], fs->U[IU(i, j)], &rh, &vh);
  const double Ui = (fs->U[IU(i, j)] + fs->U[IU(i, j - 1)]) / 2;
  const double vc = (fs->visc[IC(i, j)] + fs->visc[IC(i - 1, j)] + fs->visc[IC(i, j - 1)] +
                     fs->visc[IC(i - 1, j - 1)]) /
                    4.0;
  const double dUdy = (fs->U[IU(i, j)] - fs->U[IU(i, j - 1)]) / fs->dy;
  const double dVdx = (fs->V[IV(i, j)] - fs->V[IV(i - 1, j)]) / fs->dx;
  return -rh * Ui * vh + vc * (dUdy + dVdx);
}
static double flux_FYV(const fso_fs* fs, double eps, int64_t i, int64_t j) {
  double rh, vh;
  hybrid_interp(eps, fs->rho_v_old[IV(i, j)], fs->rho_v_old[IV(i, j + 1)], fs->V[IV(i, j)],
                fs->V[IV(i, j + 1)], fs->V[IV(i, j)], fs->V[IV(i, j + 1)], &rh, &vh);
  const double Vi   = (fs->V[IV(i, j)] + fs->V[IV(i, j + 1)]) / 2;
  const double dVdy = (fs->V[IV(i, j + 1)] - fs->V[IV(i, j)]) / fs->dy;
  return -rh * vh * Vi + 2.0 * fs->visc[IC(i, j)] * dVdy - fs->p[IC(i, j)];
}

/* src/FS.hpp:158-315, default build: flux scratch fields, then their divergence */
void fso_calc_dmomdt(const fso_fs* fs, double* dmomU, double* dmomV) {
  const double eps = 1e-3 * dmin(fs->rho_gas, fs->rho_liquid); /* FS.hpp:153-155 */
  const size_t nu = (size_t)((NX + 3) * (NY + 2)), nv = (size_t)((NX + 2) * (NY + 3));
  /* FXU, FYV: Field2D<NX,NY,1>; FYU, FXV: Field2D<NX+1,NY+1,0> (FS.hpp:163-169) */
  double* FXU = malloc(sizeof(double) * (size_t)((NX + 2) * (NY + 2)));
  double* FYV = malloc(sizeof(double) * (size_t)((NX + 2) * (NY + 2)));
  double* FYU = malloc(sizeof(double) * (size_t)((NX + 1) * (NY + 1)));
  double* FXV = malloc(sizeof(double) * (size_t)((NX + 1) * (NY + 1)));
#define ICR(i, j) ((j) + (i) * (NY + 1))
  memset(dmomU, 0, nu * sizeof(double));
  memset(dmomV, 0, nv * sizeof(double));
  for (int64_t i = -1; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j) FXU[IC(i, j)] = flux_FXU(fs, eps, i, j);
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY + 1; ++j) FYU[ICR(i, j)] = flux_FYU(fs, eps, i, j);
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j)
      dmomU[IU(i, j)] = (FXU[IC(i, j)] - FXU[IC(i - 1, j)]) / fs->dx +
                        (FYU[ICR(i, j + 1)] - FYU[ICR(i, j)]) / fs->dy + fs->p_jump_u[IU(i, j)];
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY + 1; ++j) FXV[ICR(i, j)] = flux_FXV(fs, eps, i, j);
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = -1; j < NY + 1; ++j) FYV[IC(i, j)] = flux_FYV(fs, eps, i, j);
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY + 1; ++j)
      dmomV[IV(i, j)] = (FXV[ICR(i + 1, j)] - FXV[ICR(i, j)]) / fs->dx +
                        (FYV[IC(i, j)] - FYV[IC(i, j - 1)]) / fs->dy + fs->p_jump_v[IV(i, j)];
#undef ICR
  free(FXU);
  free(FYV);
  free(FYU);
  free(FXV);
}

/* src/FS.hpp:423-436 */
void fso_update_velocity(fso_fs* fs, const double* dmomU, const double* dmomV, double dt) {
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j)
      fs->U[IU(i, j)] = (fs->rho_u_old[IU(i, j)] * fs->U_old[IU(i, j)] + dt * dmomU[IU(i, j)]) /
                        fs->rho_u[IU(i, j)];
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY + 1; ++j)
      fs->V[IV(i, j)] = (fs->rho_v_old[IV(i, j)] * fs->V_old[IV(i, j)] + dt * dmomV[IV(i, j)]) /
                        fs->rho_v[IV(i, j)];
}

/* ---- boundary conditions: src/BoundaryConditions.hpp:14-224 --------------------------------
 * Sides are applied in the order left, right, bottom, top (BoundaryConditions.hpp:220-223), each
 * over the full ghost-inclusive range of the 1-D coordinate it iterates (for_each_a). */
static void bc_left(fso_fs* fs, const fso_bconds* bc) {
  const int s = FSO_LEFT;
  double *U = fs->U, *V = fs->V;
  switch (bc->kind[s]) {
    case FSO_BC_DIRICHLET: /* :19-33 */
      for (int64_t j = -1; j < NY + 1; ++j) {
        const double ub = bc->u_prof[s] ? bc->u_prof[s][j + 1] : bc->u_val[s];
        U[IU(-1, j)] = ub;
        U[IU(0, j)]  = ub;
      }
      for (int64_t j = -1; j < NY + 2; ++j) {
        const double vb = bc->v_prof[s] ? bc->v_prof[s][j + 1] : bc->v_val[s];
        V[IV(-1, j)] = 2.0 * vb - V[IV(0, j)];
      }
      break;
    case FSO_BC_NEUMANN: /* :93-102 */
      for (int64_t j = -1; j < NY + 1; ++j)
        U[IU(-1, j)] = bc->clipped[s] ? dmin(U[IU(0, j)], 0.0) : U[IU(0, j)];
      for (int64_t j = -1; j < NY + 2; ++j) V[IV(-1, j)] = V[IV(0, j)];
      break;
    case FSO_BC_PERIODIC: /* :143-147 */
      for (int64_t j = -1; j < NY + 1; ++j) U[IU(-1, j)] = U[IU(NX - 1, j)];
      for (int64_t j = -1; j < NY + 2; ++j) V[IV(-1, j)] = V[IV(NX - 1, j)];
      break;
    case FSO_BC_SYMMETRY: /* :170-177 */
      for (int64_t j = -1; j < NY + 1; ++j) {
        U[IU(-1, j)] = -U[IU(1, j)];
        U[IU(0, j)]  = 0.0;
      }
      for (int64_t j = -1; j < NY + 2; ++j) V[IV(-1, j)] = V[IV(0, j)];
      break;
  }
}
static void bc_right(fso_fs* fs, const fso_bconds* bc) {
  const int s = FSO_RIGHT;
  double *U = fs->U, *V = fs->V;
  switch (bc->kind[s]) {
    case FSO_BC_DIRICHLET: /* :36-51 */
      for (int64_t j = -1; j < NY + 1; ++j) {
        const double ub = bc->u_prof[s] ? bc->u_prof[s][j + 1] : bc->u_val[s];
        U[IU(NX, j)]     = ub;
        U[IU(NX + 1, j)] = ub;
      }
      for (int64_t j = -1; j < NY + 2; ++j) {
        const double vb = bc->v_prof[s] ? bc->v_prof[s][j + 1] : bc->v_val[s];
        V[IV(NX, j)] = 2.0 * vb - V[IV(NX - 1, j)];
      }
      break;
    case FSO_BC_NEUMANN: /* :105-114 */
      for (int64_t j = -1; j < NY + 1; ++j)
        U[IU(NX + 1, j)] = bc->clipped[s] ? dmax(U[IU(NX, j)], 0.0) : U[IU(NX, j)];
      for (int64_t j = -1; j < NY + 2; ++j) V[IV(NX, j)] = V[IV(NX - 1, j)];
      break;
    case FSO_BC_PERIODIC: /* :150-153 */
      for (int64_t j = -1; j < NY + 1; ++j) U[IU(NX + 1, j)] = U[IU(1, j)];
      for (int64_t j = -1; j < NY + 2; ++j) V[IV(NX, j)] = V[IV(0, j)];
      break;
    case FSO_BC_SYMMETRY: /* :180-186 */
      for (int64_t j = -1; j < NY + 1; ++j) {
        U[IU(NX + 1, j)] = -U[IU(NX - 1, j)];
        U[IU(NX, j)]     = 0.0;
      }
      for (int64_t j = -1; j < NY + 2; ++j) V[IV(NX, j)] = V[IV(NX - 1, j)];
      break;
  }
}
static void bc_bottom(fso_fs* fs, const fso_bconds* bc) {
  const int s = FSO_BOTTOM;
  double *U = fs->U, *V = fs->V;
  switch (bc->kind[s]) {
    case FSO_BC_DIRICHLET: /* :54-69 */
      for (int64_t i = -1; i < NX + 2; ++i) {
        const double ub = bc->u_prof[s] ? bc->u_prof[s][i + 1] : bc->u_val[s];
        U[IU(i, -1)] = 2.0 * ub - U[IU(i, 0)];
      }
      for (int64_t i = -1; i < NX + 1; ++i) {
        const double vb = bc->v_prof[s] ? bc->v_prof[s][i + 1] : bc->v_val[s];
        V[IV(i, -1)] = vb;
        V[IV(i, 0)]  = vb;
      }
      break;
    case FSO_BC_NEUMANN: /* :117-125 */
      for (int64_t i = -1; i < NX + 2; ++i) U[IU(i, -1)] = U[IU(i, 0)];
      for (int64_t i = -1; i < NX + 1; ++i)
        V[IV(i, -1)] = bc->clipped[s] ? dmin(V[IV(i, 0)], 0.0) : V[IV(i, 0)];
      break;
    case FSO_BC_PERIODIC: /* :156-159 */
      for (int64_t i = -1; i < NX + 2; ++i) U[IU(i, -1)] = U[IU(i, NY - 1)];
      for (int64_t i = -1; i < NX + 1; ++i) V[IV(i, -1)] = V[IV(i, NY - 1)];
      break;
    case FSO_BC_SYMMETRY: /* :189-195 */
      for (int64_t i = -1; i < NX + 2; ++i) U[IU(i, -1)] = U[IU(i, 0)];
      for (int64_t i = -1; i < NX + 1; ++i) {
        V[IV(i, -1)] = -V[IV(i, 1)];
        V[IV(i, 0)]  = 0.0;
      }
      break;
  }
}
static void bc_top(fso_fs* fs, const fso_bconds* bc) {
  const int s = FSO_TOP;
  double *U = fs->U, *V = fs->V;
  switch (bc->kind[s]) {
    case FSO_BC_DIRICHLET: /* :72-87 */
      for (int64_t i = -1; i < NX + 2; ++i) {
        const double ub = bc->u_prof[s] ? bc->u_prof[s][i + 1] : bc->u_val[s];
        U[IU(i, NY)] = 2.0 * ub - U[IU(i, NY - 1)];
      }
      for (int64_t i = -1; i < NX + 1; ++i) {
        const double vb = bc->v_prof[s] ? bc->v_prof[s][i + 1] : bc->v_val[s];
        V[IV(i, NY)]     = vb;
        V[IV(i, NY + 1)] = vb;
      }
      break;
    case FSO_BC_NEUMANN: /* :128-137 */
      for (int64_t i = -1; i < NX + 2; ++i) U[IU(i, NY)] = U[IU(i, NY - 1)];
      for (int64_t i = -1; i < NX + 1; ++i)
        V[IV(i, NY + 1)] = bc->clipped[s] ? dmax(V[IV(i, NY)], 0.0) : V[IV(i, NY)];
      break;
    case FSO_BC_PERIODIC: /* :162-165 */
      for (int64_t i = -1; i < NX + 2; ++i) U[IU(i, NY)] = U[IU(i, 0)];
      for (int64_t i = -1; i < NX + 1; ++i) V[IV(i, NY + 1)] = V[IV(i, 1)];
      break;
    case FSO_BC_SYMMETRY: /* :198-205 */
      for (int64_t i = -1; i < NX + 2; ++i) U[IU(i, NY)] = U[IU(i, NY - 1)];
      for (int64_t i = -1; i < NX + 1; ++i) {
        V[IV(i, NY + 1)] = -V[IV(i, NY - 1)];
        V[IV(i, NY)]     = 0.0;
      }
      break;
  }
}
void fso_apply_velocity_bconds(fso_fs* fs, const fso_bconds* bc) {
  bc_left(fs, bc);
  bc_right(fs, bc);
  bc_bottom(fs, bc);
  bc_top(fs, bc);
}

/* test/LaminarChannel.cpp:55-66 (sums over j = -1..NY in order) and :201-205 */
double fso_outflow_fix(fso_fs* fs) {
  double inflow = 0, outflow = 0;
  for (int64_t j = -1; j < NY + 1; ++j) {
    inflow += fs->rho_u[IU(-1, j)] * fs->U[IU(-1, j)];
    outflow += fs->rho_u[IU(NX + 1, j)] * fs->U[IU(NX + 1, j)];
  }
  const double mass_error = outflow - inflow;
  for (int64_t j = -1; j < NY + 1; ++j)
    fs->U[IU(NX + 1, j)] -= mass_error / (fs->rho_u[IU(NX + 1, j)] * (double)(NY + 2));
  return mass_error;
}

/* the channel drivers' corrections of the two ghost columns after apply_velocity_bconds:
 * mode 1 test/LaminarChannel.cpp:55-66,201-205 (plain sums);
 * mode 2 test/CustomDirichletChannel.cpp:61-72,197-202 = scaling/Channel.cpp:65-76,208-213 (sums weighted
 *        with dy, the same correction formula -> only a fraction dy of the error is removed);
 * mode 3 test/PeriodicChannel.cpp:56-67,193-203 = scaling/Channel.cpp:196-206: both columns are shifted
 *        so that each carries total_flow. */
double fso_flow_control(fso_fs* fs, int mode, double total_flow, double* io) {
  if (mode == 1) {
    double in = 0, out = 0;
    for (int64_t j = -1; j < NY + 1; ++j) {
      in += fs->rho_u[IU(-1, j)] * fs->U[IU(-1, j)];
      out += fs->rho_u[IU(NX + 1, j)] * fs->U[IU(NX + 1, j)];
    }
    if (io) { io[0] = in; io[1] = out; }
    return fso_outflow_fix(fs);
  }
  double inflow = 0, outflow = 0;
  for (int64_t j = -1; j < NY + 1; ++j) {
    inflow += fs->rho_u[IU(-1, j)] * fs->U[IU(-1, j)] * fs->dy;
    outflow += fs->rho_u[IU(NX + 1, j)] * fs->U[IU(NX + 1, j)] * fs->dy;
  }
  if (io) { io[0] = inflow; io[1] = outflow; }
  const double mass_error = outflow - inflow;
  if (mode == 2) {
    for (int64_t j = -1; j < NY + 1; ++j)
      fs->U[IU(NX + 1, j)] -= mass_error / (fs->rho_u[IU(NX + 1, j)] * (double)(NY + 2));
  } else {
    const double ie = total_flow - inflow, oe = total_flow - outflow;
    for (int64_t j = -1; j < NY + 1; ++j) {
      fs->U[IU(-1, j)] += ie / (fs->rho_u[IU(-1, j)] * fs->dy * (double)(NY + 2));
      fs->U[IU(NX + 1, j)] += oe / (fs->rho_u[IU(NX + 1, j)] * fs->dy * (double)(NY + 2));
    }
  }
  return mass_error;
}

/* src/Operators.hpp:32-41: all cells incl. ghosts */
void fso_calc_divergence(const fso_fs* fs, double* div) {
  for (int64_t i = -1; i < NX + 1; ++i)
    for (int64_t j = -1; j < NY + 1; ++j)
      div[IC(i, j)] = (fs->U[IU(i + 1, j)] - fs->U[IU(i, j)]) / fs->dx +
                      (fs->V[IV(i, j + 1)] - fs->V[IV(i, j)]) / fs->dy;
}

/* src/LinearSystem.hpp:32-122.  op[IF(i,j)*5 + {C,W,E,S,N}], Layout::F. */
void fso_fill_pressure_operator(const fso_fs* fs, int psd, double* op) {
  const double vol = fs->dx * fs->dy;
  const double* ru = fs->rho_u;
  const double* rv = fs->rho_v;
  for (int64_t i = -1; i < NX + 1; ++i) {
    for (int64_t j = -1; j < NY + 1; ++j) {
      double* s = op + 5 * IF(i, j);
      s[0] = s[1] = s[2] = s[3] = s[4] = 0.0;
      if (i == -1) { /* :41-45 */
        s[0] += -vol * -1.0 / (sqr(fs->dx) * ru[IU(i + 1, j)]);
        s[1] += 0.0;
        s[2] += -vol * 1.0 / (sqr(fs->dx) * ru[IU(i + 1, j)]);
      } else if (i == NX) { /* :46-50 */
        s[0] += -vol * -1.0 / (sqr(fs->dx) * ru[IU(i, j)]);
        s[1] += -vol * 1.0 / (sqr(fs->dx) * ru[IU(i, j)]);
        s[2] += 0.0;
      } else { /* :51-57 */
        s[0] += -vol * (-1.0 / (sqr(fs->dx) * ru[IU(i, j)]) + -1.0 / (sqr(fs->dx) * ru[IU(i + 1, j)]));
        s[1] += -vol * 1.0 / (sqr(fs->dx) * ru[IU(i, j)]);
        s[2] += -vol * 1.0 / (sqr(fs->dx) * ru[IU(i + 1, j)]);
      }
      if (j == -1) { /* :60-64 */
        s[0] += -vol * -1.0 / (sqr(fs->dy) * rv[IV(i, j + 1)]);
        s[3] += 0.0;
        s[4] += -vol * 1.0 / (sqr(fs->dy) * rv[IV(i, j + 1)]);
      } else if (j == NY) { /* :65-69 */
        s[0] += -vol * -1.0 / (sqr(fs->dy) * rv[IV(i, j)]);
        s[3] += -vol * 1.0 / (sqr(fs->dy) * rv[IV(i, j)]);
        s[4] += 0.0;
      } else { /* :70-76 */
        s[0] += -vol * (-1.0 / (sqr(fs->dy) * rv[IV(i, j)]) + -1.0 / (sqr(fs->dy) * rv[IV(i, j + 1)]));
        s[3] += -vol * 1.0 / (sqr(fs->dy) * rv[IV(i, j)]);
        s[4] += -vol * 1.0 / (sqr(fs->dy) * rv[IV(i, j + 1)]);
      }
    }
  }
  /* identity row on one side, :79-121 */
#define IDROW(i, j)                  \
  do {                               \
    double* s = op + 5 * IF(i, j);   \
    s[0] = 1.0;                      \
    s[1] = s[2] = s[3] = s[4] = 0.0; \
  } while (0)
  if (psd == FSO_PSD_LEFT)
    for (int64_t j = -1; j < NY + 1; ++j) IDROW(-1, j);
  if (psd == FSO_PSD_RIGHT)
    for (int64_t j = -1; j < NY + 1; ++j) IDROW(NX, j);
  if (psd == FSO_PSD_BOTTOM)
    for (int64_t i = -1; i < NX + 1; ++i) IDROW(i, -1);
  if (psd == FSO_PSD_TOP)
    for (int64_t i = -1; i < NX + 1; ++i) IDROW(i, NY);
#undef IDROW
}

/* src/LinearSystem.hpp:125-155.  rhs in Layout::F; the mean is std::reduce over the array in
 * storage order (serial left fold), :147-149. */
void fso_fill_pressure_rhs(const fso_fs* fs, const double* div, double dt, int psd, double* rhs) {
  const double vol = fs->dx * fs->dy;
  for (int64_t i = -1; i < NX + 1; ++i)
    for (int64_t j = -1; j < NY + 1; ++j) rhs[IF(i, j)] = -vol * div[IC(i, j)] / dt;
  if (psd == FSO_PSD_LEFT)
    for (int64_t j = -1; j < NY + 1; ++j) rhs[IF(-1, j)] = 0.0;
  else if (psd == FSO_PSD_RIGHT)
    for (int64_t j = -1; j < NY + 1; ++j) rhs[IF(NX, j)] = 0.0;
  else if (psd == FSO_PSD_BOTTOM)
    for (int64_t i = -1; i < NX + 1; ++i) rhs[IF(i, -1)] = 0.0;
  else if (psd == FSO_PSD_TOP)
    for (int64_t i = -1; i < NX + 1; ++i) rhs[IF(i, NY)] = 0.0;
  else {
    const int64_t n = (NX + 2) * (NY + 2);
    double sum = 0.0;
    for (int64_t k = 0; k < n; ++k) sum += rhs[k];
    const double mean = sum / (double)n;
    for (int64_t k = 0; k < n; ++k) rhs[k] -= mean;
  }
}

/* src/Operators.hpp:78-82 with integrate<true> :52-62 (row-major serial sum, then * dx * dy) */
void fso_shift_pressure_to_zero(const fso_fs* fs, double* dp) {
  const int64_t n = (NX + 2) * (NY + 2);
  double integral = 0.0;
  for (int64_t k = 0; k < n; ++k) integral += dp[k];
  const double c = integral * fs->dx * fs->dy;
  for (int64_t k = 0; k < n; ++k) dp[k] -= c;
}

/* p += dp on all cells; U,V corrected on interior faces with the variable-density form
 * test/LaminarChannel.cpp:222-233 (== scaling/LidDrivenFlow.cpp:153-160 for constant rho) */
void fso_correct(fso_fs* fs, const double* dp, double dt) {
  const int64_t n = (NX + 2) * (NY + 2);
  for (int64_t k = 0; k < n; ++k) fs->p[k] += dp[k];
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j)
      fs->U[IU(i, j)] -= (dp[IC(i, j)] - dp[IC(i - 1, j)]) / fs->dx * dt / fs->rho_u[IU(i, j)];
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY + 1; ++j)
      fs->V[IV(i, j)] -= (dp[IC(i, j)] - dp[IC(i, j - 1)]) / fs->dy * dt / fs->rho_v[IV(i, j)];
}

/* src/Utility.hpp:80-88 (over the whole array incl. ghosts) */
/* src/Utility.hpp:44-88: abs_max<INCLUDE_GHOST = false>(field) of an (m+2) x (n+2) field reduces over the
 * INTERIOR m x n entries (for_each_i) unless include_ghost (for_each_a) */
double fso_abs_max(const double* a, int64_t m, int64_t n, int include_ghost) {
  double r = 0.0;
  const int64_t lo = include_ghost ? -1 : 0;
  for (int64_t i = lo; i < m - lo; ++i)
    for (int64_t j = lo; j < n - lo; ++j) r = dmax(r, fabs(a[(j + 1) + (i + 1) * (n + 2)]));
  return r;
}

/* integrate<INCLUDE_GHOST> / L1_norm<INCLUDE_GHOST>, src/Operators.hpp:51-76, of an (m+2) x (n+2) field:
 * serial sum in row-major order, then * dx * dy */
double fso_integrate(const double* a, int64_t m, int64_t n, double dx, double dy, int include_ghost, int l1) {
  double r = 0.0;
  const int64_t lo = include_ghost ? -1 : 0;
  for (int64_t i = lo; i < m - lo; ++i)
    for (int64_t j = lo; j < n - lo; ++j) {
      const double v = a[(j + 1) + (i + 1) * (n + 2)];
      r += l1 ? fabs(v) : v;
    }
  return r * dx * dy;
}

/* calc_grad_of_centered_points, src/Operators.hpp:164-201, NGHOST = 1: central differences where both
 * neighbours exist, second-order one-sided differences on the outermost (ghost) ring */
void fso_grad_centered(const double* f, int64_t m, int64_t n, double dx, double dy, double* dfdx, double* dfdy) {
#define F(i, j) f[(j) + 1 + ((i) + 1) * (n + 2)]
#define GX(i, j) dfdx[(j) + 1 + ((i) + 1) * (n + 2)]
#define GY(i, j) dfdy[(j) + 1 + ((i) + 1) * (n + 2)]
  for (int64_t i = 0; i < m; ++i)
    for (int64_t j = 0; j < n; ++j) {
      GX(i, j) = (F(i + 1, j) - F(i - 1, j)) / (2.0 * dx);
      GY(i, j) = (F(i, j + 1) - F(i, j - 1)) / (2.0 * dy);
    }
  for (int64_t i = -1; i < m + 1; ++i) {
    if (i > -1 && i < m) {
      GX(i, -1) = (F(i + 1, -1) - F(i - 1, -1)) / (2.0 * dx);
      GX(i, n)  = (F(i + 1, n) - F(i - 1, n)) / (2.0 * dx);
    }
    GY(i, -1) = (-3.0 * F(i, -1) + 4.0 * F(i, 0) - F(i, 1)) / (2.0 * dy);
    GY(i, n)  = (3.0 * F(i, n) - 4.0 * F(i, n - 1) + F(i, n - 2)) / (2.0 * dy);
  }
  for (int64_t j = -1; j < n + 1; ++j) {
    GX(-1, j) = (-3.0 * F(-1, j) + 4.0 * F(0, j) - F(1, j)) / (2.0 * dx);
    GX(m, j)  = (3.0 * F(m, j) - 4.0 * F(m - 1, j) + F(m - 2, j)) / (2.0 * dx);
    if (j > -1 && j < n) {
      GY(-1, j) = (F(-1, j + 1) - F(-1, j - 1)) / (2.0 * dy);
      GY(m, j)  = (F(m, j + 1) - F(m, j - 1)) / (2.0 * dy);
    }
  }
#undef F
#undef GX
#undef GY
}

/* bilinear sampling of cell-centred fields, src/Operators.hpp:85-161.  xm, ym: the cell-centre coordinates incl.
 * their ghost entries (m + 2 and n + 2 values).  mode 0 = eval_flow_field_at (:118-161, both index pairs clamped
 * into [0, N-1]); mode 1 = bilinear_interpolate (:85-115, clamped by position only). */
static void sample_indices(int mode, double pos, const double* grid /* grid[k + 1] = grid(k) */, int64_t N, double delta,
                           int64_t* prev, int64_t* next) {
  const int64_t p = (int64_t)floor((pos - grid[1]) / delta);
  const int64_t q = (int64_t)floor((pos - grid[1]) / delta + 1.0);
  if (mode == 0) {
    if (pos <= grid[1] || p < 0) { *prev = *next = 0; return; }
    if (pos >= grid[N] || q >= N) { *prev = *next = N - 1; return; }
  } else {
    if (pos <= grid[1]) { *prev = *next = 0; return; }
    if (pos >= grid[N]) { *prev = *next = N - 1; return; }
  }
  *prev = p;
  *next = q;
}
static double sample_one(const double* f, int64_t n, const double* xm, const double* ym, double dx, double dy,
                         int64_t ip, int64_t in, int64_t jp, int64_t jn, double x, double y) {
#define F(i, j) f[(j) + 1 + ((i) + 1) * (n + 2)]
  const double a = (F(in, jp) - F(ip, jp)) / dx * (x - xm[ip + 1]) + F(ip, jp);
  const double b = (F(in, jn) - F(ip, jn)) / dx * (x - xm[ip + 1]) + F(ip, jn);
#undef F
  return (b - a) / dy * (y - ym[jp + 1]) + a;
}
void fso_sample(int mode, const double* fa, const double* fb, int64_t m, int64_t n, const double* xm, const double* ym,
                int64_t npts, const double* px, const double* py, double* oa, double* ob) {
  const double dx = xm[2] - xm[1], dy = ym[2] - ym[1];  /* xm(1) - xm(0) */
  for (int64_t k = 0; k < npts; ++k) {
    int64_t ip, in, jp, jn;
    sample_indices(mode, px[k], xm, m, dx, &ip, &in);
    sample_indices(mode, py[k], ym, n, dy, &jp, &jn);
    oa[k] = sample_one(fa, n, xm, ym, dx, dy, ip, in, jp, jn, px[k], py[k]);
    if (fb && ob) ob[k] = sample_one(fb, n, xm, ym, dx, dy, ip, in, jp, jn, px[k], py[k]);
  }
}

/* src/BoundaryConditions.hpp:227-252 for an (m+2) x (n+2) field */
void fso_apply_neumann_bconds(double* f, int64_t m, int64_t n) {
#define F(i, j) f[(j) + 1 + ((i) + 1) * (n + 2)]
  for (int64_t j = -1; j < n + 1; ++j) {
    F(-1, j) = F(0, j);
    F(m, j)  = F(m - 1, j);
  }
  for (int64_t i = -1; i < m + 1; ++i) {
    F(i, -1) = F(i, 0);
    F(i, n)  = F(i, n - 1);
  }
#undef F
}

/* src/FS.hpp:569-589 */
void fso_calc_rho_visc_const(fso_fs* fs) {
  const int64_t nu = (NX + 3) * (NY + 2), nv = (NX + 2) * (NY + 3), nc = (NX + 2) * (NY + 2);
  for (int64_t k = 0; k < nu; ++k) fs->rho_u[k] = fs->rho_u_old[k] = fs->rho_gas;
  for (int64_t k = 0; k < nv; ++k) fs->rho_v[k] = fs->rho_v_old[k] = fs->rho_gas;
  for (int64_t k = 0; k < nc; ++k) fs->visc[k] = fs->visc_gas;
}

/* src/FS.hpp:592-650 (harmonic viscosity; VF_LOW/VF_HIGH from src/IR.hpp:16-17) */
void fso_calc_rho_visc_vof(fso_fs* fs, const double* vf) {
  const double VF_LOW = 1e-8, VF_HIGH = 1.0 - 1e-8;
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j) {
      const double rm = vf[IC(i - 1, j)] * fs->rho_liquid + (1.0 - vf[IC(i - 1, j)]) * fs->rho_gas;
      const double rp = vf[IC(i, j)] * fs->rho_liquid + (1.0 - vf[IC(i, j)]) * fs->rho_gas;
      fs->rho_u[IU(i, j)] = (rm + rp) / 2.0;
    }
  fso_apply_neumann_bconds(fs->rho_u, NX + 1, NY);
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY + 1; ++j) {
      const double rm = vf[IC(i, j - 1)] * fs->rho_liquid + (1.0 - vf[IC(i, j - 1)]) * fs->rho_gas;
      const double rp = vf[IC(i, j)] * fs->rho_liquid + (1.0 - vf[IC(i, j)]) * fs->rho_gas;
      fs->rho_v[IV(i, j)] = (rm + rp) / 2.0;
    }
  fso_apply_neumann_bconds(fs->rho_v, NX, NY + 1);
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY; ++j) {
      const double f = vf[IC(i, j)];
      if (f < VF_LOW) fs->visc[IC(i, j)] = fs->visc_gas;
      else if (f > VF_HIGH) fs->visc[IC(i, j)] = fs->visc_liquid;
      else
        fs->visc[IC(i, j)] =
            (fs->visc_liquid * fs->visc_gas) / (fs->visc_liquid * (1.0 - f) + fs->visc_gas * f);
    }
  fso_apply_neumann_bconds(fs->visc, NX, NY);
}

/* src/FS.hpp:439-466 */
void fso_calc_pressure_jump(fso_fs* fs, const double* vf, const double* curv, const double* ilen) {
  memset(fs->p_jump_u, 0, sizeof(double) * (size_t)((NX + 3) * (NY + 2)));
  memset(fs->p_jump_v, 0, sizeof(double) * (size_t)((NX + 2) * (NY + 3)));
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j) {
      const double ml = ilen[IC(i - 1, j)], pl = ilen[IC(i, j)];
      const double c =
          (pl + ml) > 0.0 ? (curv[IC(i, j)] * pl + curv[IC(i - 1, j)] * ml) / (pl + ml) : 0.0;
      fs->p_jump_u[IU(i, j)] = fs->sigma * c * (vf[IC(i, j)] - vf[IC(i - 1, j)]) / fs->dx;
    }
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY + 1; ++j) {
      const double ml = ilen[IC(i, j - 1)], pl = ilen[IC(i, j)];
      const double c =
          (pl + ml) > 0.0 ? (curv[IC(i, j)] * pl + curv[IC(i, j - 1)] * ml) / (pl + ml) : 0.0;
      fs->p_jump_v[IV(i, j)] = fs->sigma * c * (vf[IC(i, j)] - vf[IC(i, j - 1)]) / fs->dy;
    }
}

/* src/Operators.hpp:10-19: all cells incl. ghosts */
void fso_interpolate_UV(const fso_fs* fs, double* Ui, double* Vi) {
  for (int64_t i = -1; i < NX + 1; ++i)
    for (int64_t j = -1; j < NY + 1; ++j) {
      Ui[IC(i, j)] = (fs->U[IU(i, j)] + fs->U[IU(i + 1, j)]) / 2;
      Vi[IC(i, j)] = (fs->V[IV(i, j)] + fs->V[IV(i, j + 1)]) / 2;
    }
}

/* src/FS.hpp:653-676 */
void fso_conserved(const fso_fs* fs, double* mass, double* mom_x, double* mom_y) {
  double m = 0, mx = 0, my = 0;
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY; ++j) {
      m += (fs->rho_u[IU(i, j)] + fs->rho_u[IU(i + 1, j)] + fs->rho_v[IV(i, j)] +
            fs->rho_v[IV(i, j + 1)]) /
           4.0 * fs->dx * fs->dy;
      mx += (fs->rho_u[IU(i, j)] * fs->U[IU(i, j)] + fs->rho_u[IU(i + 1, j)] * fs->U[IU(i + 1, j)]) /
            2.0 * fs->dx * fs->dy;
      my += (fs->rho_v[IV(i, j)] * fs->V[IV(i, j)] + fs->rho_v[IV(i, j + 1)] * fs->V[IV(i, j + 1)]) /
            2.0 * fs->dx * fs->dy;
    }
  *mass = m;
  *mom_x = mx;
  *mom_y = my;
}

/* src/FS.hpp:318-406: mass fluxes with the same hybrid density selection (velocity slots 0) */
static inline double rho_hybrid(double eps, double rm, double rp, double tm, double tp) {
  double r, v;
  hybrid_interp(eps, rm, rp, 0.0, 0.0, tm, tp, &r, &v);
  return r;
}
void fso_calc_drhodt(const fso_fs* fs, double* drho_u, double* drho_v) {
  const double eps = 1e-3 * dmin(fs->rho_gas, fs->rho_liquid);
  const size_t nu = (size_t)((NX + 3) * (NY + 2)), nv = (size_t)((NX + 2) * (NY + 3));
  double* FXU = malloc(sizeof(double) * (size_t)((NX + 2) * (NY + 2)));
  double* FYV = malloc(sizeof(double) * (size_t)((NX + 2) * (NY + 2)));
  double* FYU = malloc(sizeof(double) * (size_t)((NX + 1) * (NY + 1)));
  double* FXV = malloc(sizeof(double) * (size_t)((NX + 1) * (NY + 1)));
#define ICR(i, j) ((j) + (i) * (NY + 1))
  memset(drho_u, 0, nu * sizeof(double));
  memset(drho_v, 0, nv * sizeof(double));
  const double *U = fs->U, *V = fs->V, *ru = fs->rho_u_old, *rv = fs->rho_v_old;
  for (int64_t i = -1; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j) { /* :340-351 */
      const double r = rho_hybrid(eps, ru[IU(i, j)], ru[IU(i + 1, j)], U[IU(i, j)], U[IU(i + 1, j)]);
      FXU[IC(i, j)] = -r * ((U[IU(i, j)] + U[IU(i + 1, j)]) / 2.0);
    }
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY + 1; ++j) { /* :355-366 */
      const double r = rho_hybrid(eps, ru[IU(i, j - 1)], ru[IU(i, j)], V[IV(i - 1, j)], V[IV(i, j)]);
      FYU[ICR(i, j)] = -r * ((V[IV(i - 1, j)] + V[IV(i, j)]) / 2);
    }
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j)
      drho_u[IU(i, j)] = (FXU[IC(i, j)] - FXU[IC(i - 1, j)]) / fs->dx +
                         (FYU[ICR(i, j + 1)] - FYU[ICR(i, j)]) / fs->dy;
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY + 1; ++j) { /* :377-388 */
      const double r = rho_hybrid(eps, rv[IV(i - 1, j)], rv[IV(i, j)], U[IU(i, j - 1)], U[IU(i, j)]);
      FXV[ICR(i, j)] = -r * ((U[IU(i, j - 1)] + U[IU(i, j)]) / 2.0);
    }
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = -1; j < NY + 1; ++j) { /* :392-403 */
      const double r = rho_hybrid(eps, rv[IV(i, j)], rv[IV(i, j + 1)], V[IV(i, j)], V[IV(i, j + 1)]);
      FYV[IC(i, j)] = -r * ((V[IV(i, j)] + V[IV(i, j + 1)]) / 2.0);
    }
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY + 1; ++j)
      drho_v[IV(i, j)] = (FXV[ICR(i + 1, j)] - FXV[ICR(i, j)]) / fs->dx +
                         (FYV[IC(i, j)] - FYV[IC(i, j - 1)]) / fs->dy;
#undef ICR
  free(FXU);
  free(FYV);
  free(FYU);
  free(FXV);
}

/* src/FS.hpp:409-420 */
void fso_update_density(fso_fs* fs, const double* drho_u, const double* drho_v, double dt) {
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j)
      fs->rho_u[IU(i, j)] = fs->rho_u_old[IU(i, j)] + dt * drho_u[IU(i, j)];
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY + 1; ++j)
      fs->rho_v[IV(i, j)] = fs->rho_v_old[IV(i, j)] + dt * drho_v[IV(i, j)];
}

/* src/IB.hpp:110-126 */
void fso_correct_velocity_ib_implicit_euler(fso_fs* fs, const double* ib_u, const double* ib_v, double dt) {
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j) {
      const double visc = (fs->visc[IC(i, j)] + fs->visc[IC(i - 1, j)]) / 2.0;
      const double rho  = fs->rho_u[IU(i, j)];
      fs->U[IU(i, j)] /= 1.0 + dt * visc / rho * ib_u[IU(i, j)];
    }
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY + 1; ++j) {
      const double visc = (fs->visc[IC(i, j)] + fs->visc[IC(i, j - 1)]) / 2.0;
      const double rho  = fs->rho_v[IV(i, j)];
      fs->V[IV(i, j)] /= 1.0 + dt * visc / rho * ib_v[IV(i, j)];
    }
}

/* src/IB.hpp:129-186: U = (B rho_old U_old + dt F) / (A rho), B = l dt / (exp(l dt) - 1), A = l dt + B */
static inline double ib_semi(double lambda, double dt, double ro, double uo, double f, double r) {
  if (isinf(lambda)) return 0.0;
  const double B = fabs(lambda) < 1e-6 ? 1.0 : (lambda * dt) / (exp(lambda * dt) - 1.0);
  const double A = lambda * dt + B;
  return (B * ro * uo + dt * f) / (A * r);
}
void fso_update_velocity_ib_semi_analytical(fso_fs* fs, const double* dmomU, const double* dmomV,
                                            double dt, const double* ib_u, const double* ib_v) {
  for (int64_t i = 0; i < NX + 1; ++i)
    for (int64_t j = 0; j < NY; ++j) {
      const double visc = (fs->visc[IC(i, j)] + fs->visc[IC(i - 1, j)]) / 2.0;
      const double rho  = fs->rho_u[IU(i, j)];
      fs->U[IU(i, j)] = ib_semi(visc / rho * ib_u[IU(i, j)], dt, fs->rho_u_old[IU(i, j)],
                                fs->U_old[IU(i, j)], dmomU[IU(i, j)], rho);
    }
  for (int64_t i = 0; i < NX; ++i)
    for (int64_t j = 0; j < NY + 1; ++j) {
      const double visc = (fs->visc[IC(i, j)] + fs->visc[IC(i, j - 1)]) / 2.0;
      const double rho  = fs->rho_v[IV(i, j)];
      fs->V[IV(i, j)] = ib_semi(visc / rho * ib_v[IV(i, j)], dt, fs->rho_v_old[IV(i, j)],
                                fs->V_old[IV(i, j)], dmomV[IV(i, j)], rho);
    }
}

/* test/StationaryDrop.cpp:331-353 */
void fso_capillary_divergence(fso_fs* fs, const double* vf, const double* curv, const double* ilen,
                              double dt, double* div) {
  const size_t nu = (size_t)((NX + 3) * (NY + 2)), nv = (size_t)((NX + 2) * (NY + 3));
  double* du = malloc(sizeof(double) * nu);
  double* dv = malloc(sizeof(double) * nv);
  memcpy(du, fs->p_jump_u, sizeof(double) * nu); /* :332-333 copy(fs.p_jump_*, delta_pj_*) */
  memcpy(dv, fs->p_jump_v, sizeof(double) * nv);
  fso_calc_pressure_jump(fs, vf, curv, ilen);
  fso_apply_neumann_bconds(fs->p_jump_u, NX + 1, NY);
  fso_apply_neumann_bconds(fs->p_jump_v, NX, NY + 1);
  for (size_t k = 0; k < nu; ++k) du[k] = fs->p_jump_u[k] - du[k];
  for (size_t k = 0; k < nv; ++k) dv[k] = fs->p_jump_v[k] - dv[k];
  for (int64_t i = -1; i < NX + 1; ++i)
    for (int64_t j = -1; j < NY + 1; ++j)
      div[IC(i, j)] += dt * ((du[IU(i + 1, j)] / fs->rho_u[IU(i + 1, j)] -
                              du[IU(i, j)] / fs->rho_u[IU(i, j)]) /
                                 fs->dx +
                             (dv[IV(i, j + 1)] / fs->rho_v[IV(i, j + 1)] -
                              dv[IV(i, j)] / fs->rho_v[IV(i, j)]) /
                                 fs->dy);
  free(du);
  free(dv);
}

/* One timestep's sub-iteration loop, scaling/LidDrivenFlow.cpp:127-161
 * (+ test/LaminarChannel.cpp:201-205 outflow fix, + operator rebuild of two-phase drivers). */
int fso_step(fso_fs* fs, const fso_bconds* bc, fso_solver* s, const fso_step_opts* o, double dt,
             double* res) {
  const int64_t nu = (NX + 3) * (NY + 2), nv = (NX + 2) * (NY + 3), nc = (NX + 2) * (NY + 2);
  double* dmomU = malloc(sizeof(double) * (size_t)nu);
  double* dmomV = malloc(sizeof(double) * (size_t)nv);
  double* div   = malloc(sizeof(double) * (size_t)nc);
  double* rhs   = malloc(sizeof(double) * (size_t)nc);
  double* dp    = malloc(sizeof(double) * (size_t)nc);
  double* op    = o->rebuild_operator_each_subiter ? malloc(sizeof(double) * 5 * (size_t)nc) : NULL;
  int iters = 0;
  fso_save_old_state(fs);
  for (int sub = 0; sub < o->num_subiter; ++sub) {
    fso_calc_mid_time(fs->U, fs->U_old, nu);
    fso_calc_mid_time(fs->V, fs->V_old, nv);
    if (o->density_update) { /* test/StationaryDrop.cpp:274-293; dmomU/V double as drho scratch */
      fso_calc_drhodt(fs, dmomU, dmomV);
      fso_update_density(fs, dmomU, dmomV, dt);
      fso_apply_neumann_bconds(fs->rho_u, NX + 1, NY);
      fso_apply_neumann_bconds(fs->rho_v, NX, NY + 1);
    }
    fso_calc_dmomdt(fs, dmomU, dmomV);
    if (o->ib_mode == 2) {
      fso_update_velocity_ib_semi_analytical(fs, dmomU, dmomV, dt, o->ib_u, o->ib_v);
    } else {
      fso_update_velocity(fs, dmomU, dmomV, dt);
      if (o->ib_mode == 1) fso_correct_velocity_ib_implicit_euler(fs, o->ib_u, o->ib_v, dt);
    }
    fso_apply_velocity_bconds(fs, bc);
    if (o->outflow_fix) fso_flow_control(fs, o->outflow_fix, o->total_flow, NULL);
    fso_calc_divergence(fs, div);
    if (o->capillary) fso_capillary_divergence(fs, o->vf, o->curv, o->ilen, dt, div);
    if (op) {
      fso_fill_pressure_operator(fs, o->psd, op);
      fso_solver_set_operator(s, op);
    }
    fso_fill_pressure_rhs(fs, div, dt, o->psd, rhs);
    iters += fso_solver_solve(s, rhs, dp, res);
    fso_shift_pressure_to_zero(fs, dp);
    fso_correct(fs, dp, dt);
  }
  free(dmomU);
  free(dmomV);
  free(div);
  free(rhs);
  free(dp);
  free(op);
  return iters;
}
