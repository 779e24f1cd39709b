/* oracle/fs_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's (gidonbauer/FluidSolver) per-timestep
 * stepping path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this; the product (fluidsolver_b200/) never does.
 *
 * All arrays use the reference's own storage (src/Container.hpp:105-111, Layout::C,
 * NGHOST = 1): a field with interior extent M x N is (M+2) x (N+2) doubles, j fastest,
 *      idx(i, j) = (j + 1) + (i + 1) * (N + 2),      i in [-1, M], j in [-1, N].
 *   U, rho_u, p_jump_u, dmomU : M = NX+1, N = NY       -> (NX+3) x (NY+2)
 *   V, rho_v, p_jump_v, dmomV : M = NX,   N = NY+1     -> (NX+2) x (NY+3)
 *   visc, p, div, delta_p, vf : M = NX,   N = NY       -> (NX+2) x (NY+2)
 */
#ifndef FS_ORACLE_H_
#define FS_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* boundary-condition kinds, src/BoundaryConditions.hpp:14-200 */
enum { FSO_BC_DIRICHLET = 0, FSO_BC_NEUMANN = 1, FSO_BC_PERIODIC = 2, FSO_BC_SYMMETRY = 3 };
/* sides, order of application in apply_velocity_bconds (BoundaryConditions.hpp:220-223) */
enum { FSO_LEFT = 0, FSO_RIGHT = 1, FSO_BOTTOM = 2, FSO_TOP = 3 };
/* PSDirichlet, src/LinearSystem.hpp:10 */
enum { FSO_PSD_NONE = 0, FSO_PSD_LEFT = 1, FSO_PSD_RIGHT = 2, FSO_PSD_BOTTOM = 3, FSO_PSD_TOP = 4 };

typedef struct {
  int kind[4];
  double u_val[4], v_val[4];
  /* optional per-point Dirichlet values (function-pointer variant evaluated by the caller):
   * left/right: u_prof has NY+2 entries (j=-1..NY), v_prof NY+3 (j=-1..NY+1);
   * bottom/top: u_prof has NX+3 entries (i=-1..NX+1), v_prof NX+2 (i=-1..NX). NULL = constant. */
  const double* u_prof[4];
  const double* v_prof[4];
  int clipped[4];
} fso_bconds;

typedef struct {
  int64_t nx, ny;
  double dx, dy;
  double rho_gas, rho_liquid, visc_gas, visc_liquid, sigma;
  double *visc, *p, *p_jump_u, *p_jump_v;
  double *rho_u_old, *rho_v_old, *U_old, *V_old;
  double *rho_u, *rho_v, *U, *V;
} fso_fs;

/* src/FS.hpp:100-128 */
double fso_adjust_dt(const fso_fs* fs, double cfl_max, double dt_max);
/* src/FS.hpp:76-97 */
void fso_save_old_state(fso_fs* fs);
/* src/Operators.hpp:44-49; n = total array length incl. ghosts */
void fso_calc_mid_time(double* cur, const double* old, int64_t n);
/* src/FS.hpp:158-315 (default, un-fused build) */
void fso_calc_dmomdt(const fso_fs* fs, double* dmomU, double* dmomV);
/* src/FS.hpp:423-436 */
void fso_update_velocity(fso_fs* fs, const double* dmomU, const double* dmomV, double dt);
/* src/BoundaryConditions.hpp:215-224 */
void fso_apply_velocity_bconds(fso_fs* fs, const fso_bconds* bc);
/* test/LaminarChannel.cpp:55-66,201-205 ; returns mass_error before the fix */
double fso_outflow_fix(fso_fs* fs);
/* mode 1/2/3 as fso_step_opts.outflow_fix; io[0] = inflow, io[1] = outflow before the correction
 * (dy-weighted for modes 2 and 3); returns outflow - inflow */
double fso_flow_control(fso_fs* fs, int mode, double total_flow, double* io);
/* src/Operators.hpp:32-41 */
void fso_calc_divergence(const fso_fs* fs, double* div);
/* src/LinearSystem.hpp:32-122 ; op is [(nx+2)*(ny+2)][5] in Layout::F (i fastest),
 * entries {C, W(i-1), E(i+1), S(j-1), N(j+1)} */
void fso_fill_pressure_operator(const fso_fs* fs, int psd, double* op);
/* src/LinearSystem.hpp:125-155 ; rhs in Layout::F (i fastest) */
void fso_fill_pressure_rhs(const fso_fs* fs, const double* div, double dt, int psd, double* rhs);
/* src/Operators.hpp:52-62,78-82 */
void fso_shift_pressure_to_zero(const fso_fs* fs, double* dp);
/* driver-inlined, scaling/LidDrivenFlow.cpp:153-160 / test/LaminarChannel.cpp:224-233 */
void fso_correct(fso_fs* fs, const double* dp, double dt);
/* src/Utility.hpp:44-88: abs_max<INCLUDE_GHOST = false> of an (m+2) x (n+2) field: interior entries only
 * unless include_ghost */
double fso_abs_max(const double* a, int64_t m, int64_t n, int include_ghost);
/* integrate / L1_norm, src/Operators.hpp:51-76; calc_grad_of_centered_points, :164-201 */
double fso_integrate(const double* a, int64_t m, int64_t n, double dx, double dy, int include_ghost, int l1);
/* eval_flow_field_at (mode 0, fa = Ui, fb = Vi) / bilinear_interpolate (mode 1, fb = ob = NULL), src/Operators.hpp:85-161 */
void fso_sample(int mode, const double* fa, const double* fb, int64_t m, int64_t n, const double* xm, const double* ym,
                int64_t npts, const double* px, const double* py, double* oa, double* ob);
void fso_grad_centered(const double* f, int64_t m, int64_t n, double dx, double dy, double* dfdx, double* dfdy);
/* src/FS.hpp:569-650 with src/BoundaryConditions.hpp:227-252 */
void fso_calc_rho_visc_const(fso_fs* fs);
void fso_calc_rho_visc_vof(fso_fs* fs, const double* vf);
void fso_apply_neumann_bconds(double* f, int64_t m, int64_t n);
/* src/FS.hpp:439-466 */
void fso_calc_pressure_jump(fso_fs* fs, const double* vf, const double* curv, const double* ilen);
/* src/FS.hpp:318-420 */
void fso_calc_drhodt(const fso_fs* fs, double* drho_u, double* drho_v);
void fso_update_density(fso_fs* fs, const double* drho_u, const double* drho_v, double dt);
/* src/Operators.hpp:10-19, src/FS.hpp:653-676 */
void fso_interpolate_UV(const fso_fs* fs, double* Ui, double* Vi);
void fso_conserved(const fso_fs* fs, double* mass, double* mom_x, double* mom_y);
/* src/IB.hpp:110-126 and :129-186 (immersed-boundary variants of the velocity update) */
void fso_correct_velocity_ib_implicit_euler(fso_fs* fs, const double* ib_u, const double* ib_v, double dt);
void fso_update_velocity_ib_semi_analytical(fso_fs* fs, const double* dmomU, const double* dmomV,
                                            double dt, const double* ib_u, const double* ib_v);
/* the capillary block of the two-phase drivers, test/StationaryDrop.cpp:331-353: keeps the old
 * pressure jump, recomputes it (calc_pressure_jump + apply_neumann_bconds), and adds
 * dt * div(delta_pj / rho) to div on all cells */
void fso_capillary_divergence(fso_fs* fs, const double* vf, const double* curv, const double* ilen,
                              double dt, double* div);

/* ---- pressure solver: restatement of HYPRE 3.0.0 Struct PCG + PFMG as configured by
 * src/LinearSolver_StructHypre.hpp:196-290 (see oracle/pfmg_oracle.c) ------------------------ */
typedef struct fso_solver fso_solver;
typedef struct {
  double tol;
  int max_iter;
  int max_levels;   /* 16 = reference (PS_PFMG_MAX_LEVELS) */
  int relax_type;   /* 1 = weighted Jacobi (reference default), 2 = symmetric red/black GS */
  int skip_relax;   /* 1 = reference default */
  int snap_zero_diag; /* 0 = HYPRE's arithmetic as is. 1 = coarse diagonals that are pure cancellation noise become 0
                         (what k_rap of the CUDA path does, see setup_rap in pfmg_oracle.c): for singular operators
                         whose coefficients are not dyadic, where HYPRE's `== 0.0` test hinges on the last bit */
} fso_solver_opts;
fso_solver* fso_solver_create(int64_t nx, int64_t ny, const fso_solver_opts* o);
void fso_solver_destroy(fso_solver* s);
/* op/rhs/sol in Layout::F as produced by fso_fill_pressure_*; */
void fso_solver_set_operator(fso_solver* s, const double* op);
/* returns iterations; *res = sqrt(<Cr,r>/<Cb,b>); sol is Layout::C (the reference's delta_p) */
int fso_solver_solve(fso_solver* s, const double* rhs, double* sol, double* res);
int fso_solver_num_levels(const fso_solver* s);

/* ---- whole sub-iteration / timestep, scaling/LidDrivenFlow.cpp:121-174 -------------------- */
typedef struct {
  int num_subiter;
  int outflow_fix;   /* 0 none; 1 test/LaminarChannel.cpp:201-205; 2 the dy-weighted variant of
                        test/CustomDirichletChannel.cpp:61-72,197-202 (scaling/Channel.cpp:208-213);
                        3 the total-flow forcing of test/PeriodicChannel.cpp:56-67,193-203 */
  int psd;
  int rebuild_operator_each_subiter; /* two-phase drivers, test/StationaryDrop.cpp:357 */
  int density_update; /* consistent mass transport, test/StationaryDrop.cpp:274-293 */
  int capillary;      /* pressure-jump block, test/StationaryDrop.cpp:331-353 (needs vf, curv, ilen) */
  const double *vf, *curv, *ilen; /* cell fields (NX+2)x(NY+2); frozen during the step */
  double total_flow;  /* outflow_fix == 3: the flux both ghost columns are forced to */
  int ib_mode;        /* immersed-boundary drivers, scaling/LuchiniIBChannel.cpp:212-221: 0 none; 1 update_velocity +
                         correct_velocity_ib_implicit_euler; 2 update_velocity_ib_semi_analytical (src/IB.hpp:110-186) */
  const double *ib_u, *ib_v; /* ib_corr_u_stag (U-like), ib_corr_v_stag (V-like); Inf marks solid points */
} fso_step_opts;
/* returns total PCG iterations of the step; *res = last residual */
int fso_step(fso_fs* fs, const fso_bconds* bc, fso_solver* s, const fso_step_opts* o, double dt,
             double* res);

#ifdef __cplusplus
}
#endif
#endif
