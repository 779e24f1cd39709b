/* include/fsgpu.h -- C ABI of libfsgpu.so, the B200 (sm_100a) implementation of the per-timestep
 * stepping path of gidonbauer/FluidSolver.
 *
 * The reference has no FFI layer: its "API" is C++ function templates over an `FS` object
 * (src/FS.hpp, src/Operators.hpp, src/BoundaryConditions.hpp, src/LinearSolver_StructHypre.hpp).
 * Every entry point below replaces exactly one of those calls and cites it; the header-only C++
 * shim under include/shim/ re-creates the reference's template signatures on top of this ABI, so
 * a reference driver (test/TaylorGreenVortex.cpp, scaling/LidDrivenFlow.cpp, ...) switches by
 * changing its include path.
 *
 * Conventions
 *  - plain C, pointers and sizes only; every function returns 0 on success, non-zero on error
 *    (fsgpu_last_error() gives the message).  There is NO CPU fallback: without a CUDA device
 *    fsgpu_create fails.
 *  - host arrays are in the reference's storage (src/Container.hpp:105-111, Layout::C, NGHOST=1):
 *    ghost inclusive, j fastest.  U-like fields (NX+3)x(NY+2), V-like (NX+2)x(NY+3), cell fields
 *    (NX+2)x(NY+2).  The device keeps its own padded layout (DESIGN.md "Data layout").
 *  - one host thread per context (as the reference: single caller thread, parallelism inside).
 *    All device work of a context is ordered on one CUDA stream; calls that return a scalar
 *    synchronise, the others are asynchronous.
 */
#ifndef FSGPU_H_
#define FSGPU_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fsgpu_ctx fsgpu_ctx;
typedef struct fsgpu_ps fsgpu_ps;

/* ---- field ids: the members of the reference's `FS` / `State` (src/FS.hpp:15-52) plus the
 * scratch fields the drivers own (scaling/LidDrivenFlow.cpp:63-70) ------------------------------ */
enum {
  FSGPU_VISC = 0,      /* fs.visc            cell   */
  FSGPU_P,             /* fs.p               cell   */
  FSGPU_P_JUMP_U,      /* fs.p_jump_u_stag   U-like */
  FSGPU_P_JUMP_V,      /* fs.p_jump_v_stag   V-like */
  FSGPU_RHO_U_OLD,     /* fs.old.rho_u_stag  U-like */
  FSGPU_RHO_V_OLD,     /* fs.old.rho_v_stag  V-like */
  FSGPU_U_OLD,         /* fs.old.U           U-like */
  FSGPU_V_OLD,         /* fs.old.V           V-like */
  FSGPU_RHO_U,         /* fs.curr.rho_u_stag U-like */
  FSGPU_RHO_V,         /* fs.curr.rho_v_stag V-like */
  FSGPU_U,             /* fs.curr.U          U-like */
  FSGPU_V,             /* fs.curr.V          V-like */
  FSGPU_DMOM_U,        /* drhoUdt            U-like */
  FSGPU_DMOM_V,        /* drhoVdt            V-like */
  FSGPU_DIV,           /* div                cell   */
  FSGPU_DELTA_P,       /* delta_p            cell   */
  FSGPU_VF,            /* volume fraction    cell   */
  FSGPU_RHS,           /* pressure rhs, (NX+2)x(NY+2) in Layout::C (the reference keeps it in
                          Layout::F, src/LinearSystem.hpp:18-24; transpose to compare) */
  /* ---- extension fields: device memory is allocated on first use ---- */
  FSGPU_DRHO_U,        /* drho_u_stagdt      U-like  test/StationaryDrop.cpp:107 */
  FSGPU_DRHO_V,        /* drho_v_stagdt      V-like */
  FSGPU_CURV,          /* vof.curv           cell    src/VOF.hpp (input of calc_pressure_jump) */
  FSGPU_INTERFACE_LENGTH, /* vof.interface_length cell */
  FSGPU_PJ_OLD_U,      /* previous p_jump_u_stag (the drivers' delta_pj_u_stag scratch) U-like */
  FSGPU_PJ_OLD_V,      /* previous p_jump_v_stag V-like */
  FSGPU_UI,            /* Ui   cell-centred U     cell   src/Operators.hpp:10-13 */
  FSGPU_VI,            /* Vi   cell-centred V     cell   src/Operators.hpp:16-19 */
  FSGPU_RHOI,          /* rhoi cell-centred rho   cell   src/Operators.hpp:22-29 */
  FSGPU_IB_CORR_U,     /* ib_corr_u_stag     U-like  src/IB.hpp:62-107 (input) */
  FSGPU_IB_CORR_V,     /* ib_corr_v_stag     V-like */
  FSGPU_GRAD_X,        /* dfdx, dfdy of calc_grad_of_centered_points and two spare cell fields for its */
  FSGPU_GRAD_Y,        /* second derivatives (src/Curvature.hpp:337-339)   cell */
  FSGPU_GRAD_XX,
  FSGPU_GRAD_YY,
  FSGPU_NUM_FIELDS
};
#define FSGPU_NUM_CORE_FIELDS (FSGPU_RHS + 1)

/* src/BoundaryConditions.hpp:14-200 */
enum { FSGPU_BC_DIRICHLET = 0, FSGPU_BC_NEUMANN = 1, FSGPU_BC_PERIODIC = 2, FSGPU_BC_SYMMETRY = 3 };
/* order of application in apply_velocity_bconds, src/BoundaryConditions.hpp:220-223 */
enum { FSGPU_LEFT = 0, FSGPU_RIGHT = 1, FSGPU_BOTTOM = 2, FSGPU_TOP = 3 };
/* PSDirichlet, src/LinearSystem.hpp:10 */
enum { FSGPU_PSD_NONE = 0, FSGPU_PSD_LEFT = 1, FSGPU_PSD_RIGHT = 2, FSGPU_PSD_BOTTOM = 3, FSGPU_PSD_TOP = 4 };
/* smoother of the multigrid preconditioner: weighted Jacobi, what the reference gets from HYPRE's defaults
 * (Submodules/hypre/src/struct_ls/pfmg.c:32; src/LinearSolver_StructHypre.hpp never calls SetRelaxType).
 * It is the only smoother: iterate-level parity with the reference depends on it. */
enum { FSGPU_RELAX_JACOBI = 1 };

const char* fsgpu_last_error(void);
/* number of CUDA kernels launched by this library since load (bench.py's gpu_launches) */
int64_t fsgpu_launch_count(void);
int fsgpu_device_count(int* n);
/* per-kernel-class device timing (CUDA events on the context's stream) for bench.py's roofline:
 * bit k of class_mask enables class k: 0 momentum, 1 bconds, 2 div+rhs, 3 shift+correct,
 * 4 matvec+dot, 5 x/r update, 6 <r,z>, 7 p update, 8 V-cycle down level 0, 9 V-cycle up level 0,
 * 10 down levels>=1, 11 up levels>=1, 12 coarsest solve, 13 operator setup.  fsgpu_prof_read
 * synchronises, returns total milliseconds and launch-group counts per class, and clears. */
int fsgpu_prof_enable(unsigned int class_mask);
int fsgpu_prof_read(double* ms_total, int64_t* count, int nclass);

/* ---- context = the reference's `FS<Float,NX,NY,1>` object (src/FS.hpp:28-52) on one device ---- */
/* replaces: `FS<Float,NX,NY,NGHOST> fs{...}; init_grid(...)` (src/FS.hpp:55-73) */
int fsgpu_create(int64_t nx, int64_t ny, double dx, double dy, int device, fsgpu_ctx** out);
int fsgpu_destroy(fsgpu_ctx* c);

/* ---- slab decomposition over the GPUs of one box (the reference itself is
 * single-node OpenMP, so there is no reference call behind these) -------------------------------
 * The global (NX+2) x (NY+2) ghost-inclusive grid is cut into `nranks` slabs of whole i-rows;
 * rank r owns the rows a = i + 1 in [a_lo, a_hi) returned by fsgpu_slab_range (U-like fields: the
 * last rank also owns the extra row a = NX + 2).  Every other call of this header keeps its
 * meaning and is COLLECTIVE: all ranks call it in the same order with the same arguments; scalars
 * returned (dt, residual, iterations, maxima) are identical on every rank.  Host arrays passed to
 * fsgpu_upload / fsgpu_download stay GLOBAL arrays in the reference layout: upload reads this
 * rank's rows (+ halo rows) from it, download writes this rank's rows into it.
 * Transports: NCCL = one process per GPU, `id` from fsgpu_comm_id on rank 0 and broadcast by the
 * launcher (torch.distributed, MPI, a file ...); LOCAL = all ranks are contexts of ONE process,
 * each driven by its own host thread, `id` any 128-byte key shared by the group (test transport,
 * lets the decomposed path run on a single GPU). */
enum { FSGPU_TRANSPORT_NCCL = 0, FSGPU_TRANSPORT_LOCAL = 1 };
#define FSGPU_COMM_ID_BYTES 128
int fsgpu_comm_id(void* id128); /* ncclGetUniqueId */
int fsgpu_create_slab(int64_t nx, int64_t ny, double dx, double dy, int device, int rank, int nranks,
                      const void* id128, int transport, fsgpu_ctx** out);
/* With FSGPU_TRANSPORT_NCCL (one process per GPU) halo rows and scalar all-reduces travel through peer-mapped
 * memory by default (cudaIpc between the processes): mailboxes with one kernel per exchange, the PCG dot
 * products all-reduced inside their reduction kernels, and the multigrid kernels of the distributed levels
 * storing their boundary rows straight into the neighbours' halo rows (FSGPU_PUSH=0: exchange kernels
 * there too).  Every rank falls back to NCCL send/recv + all-reduce if any rank cannot map a peer, or
 * with FSGPU_HALO=nccl.  The in-process transport uses the mailboxes with FSGPU_HALO=p2p only.
 * *on = 1 if the peer-mapped path is active. */
int fsgpu_halo_is_direct(fsgpu_ctx* c, int* on);
/* host-only partition rule (no device needed): rows [*a_lo, *a_hi) of rank `rank` */
int fsgpu_slab_range(int64_t nx, int rank, int nranks, int64_t* a_lo, int64_t* a_hi);
/* host-only: the rows [*c_lo, *c_hi) of the next coarser multigrid level that belong to the owner of
 * the fine rows [lo, hi) when the level is coarsened along i with coarse point c sitting on fine row
 * 2 c + foff (a coarse row goes with the fine row it sits on; coarsening along j keeps the rows) */
int fsgpu_slab_coarse_range(int64_t lo, int64_t hi, int64_t foff, int64_t* c_lo, int64_t* c_hi);
/* fs.rho_gas, fs.rho_liquid, fs.visc_gas, fs.visc_liquid, fs.sigma (src/FS.hpp:29-33) */
int fsgpu_set_props(fsgpu_ctx* c, double rho_gas, double rho_liquid, double visc_gas,
                    double visc_liquid, double sigma);
/* Field2D::get_data() access (src/Container.hpp:150-155): whole field incl. ghosts */
int fsgpu_upload(fsgpu_ctx* c, int field, const double* host);
int fsgpu_download(fsgpu_ctx* c, int field, double* host);
int fsgpu_fill(fsgpu_ctx* c, int field, double value); /* fill(), src/Container.hpp:172-180 */
int fsgpu_sync(fsgpu_ctx* c);
/* raw device pointer + layout of a field, for zero-copy interop (torch tensors, other kernels):
 * element (i, j) of the reference lives at ptr[(i + 1 + row0) * pitch + (j + 1 + col0)] */
int fsgpu_device_field(fsgpu_ctx* c, int field, double** ptr, int64_t* pitch, int64_t* row0,
                       int64_t* col0);
int fsgpu_stream(fsgpu_ctx* c, void** cuda_stream);

/* FlowBConds (src/BoundaryConditions.hpp:207-212): one call per side.  u_prof / v_prof are the
 * function-pointer Dirichlet variant (src/BoundaryConditions.hpp:15-17) evaluated by the caller
 * at the ghost-inclusive coordinates of that side (NULL = constant u_val / v_val):
 * left/right: u_prof NY+2 values (j=-1..NY), v_prof NY+3 (j=-1..NY+1);
 * bottom/top: u_prof NX+3 values (i=-1..NX+1), v_prof NX+2 (i=-1..NX). */
int fsgpu_set_bconds(fsgpu_ctx* c, int side, int kind, double u_val, double v_val,
                     const double* u_prof, const double* v_prof, int clipped);

/* adjust_dt, src/FS.hpp:100-128 */
int fsgpu_adjust_dt(fsgpu_ctx* c, double cfl_max, double dt_max, double* dt);
/* save_old_state / save_old_velocity / save_old_density, src/FS.hpp:76-97 */
int fsgpu_save_old_state(fsgpu_ctx* c);
int fsgpu_save_old_velocity(fsgpu_ctx* c);
int fsgpu_save_old_density(fsgpu_ctx* c);
/* calc_mid_time(curr.U, old.U); calc_mid_time(curr.V, old.V), src/Operators.hpp:44-49 */
int fsgpu_calc_mid_time(fsgpu_ctx* c);
/* calc_dmomdt(fs, drhoUdt, drhoVdt), src/FS.hpp:158-315 -> FSGPU_DMOM_U / FSGPU_DMOM_V */
int fsgpu_calc_dmomdt(fsgpu_ctx* c);
/* update_velocity(drhoUdt, drhoVdt, dt, fs), src/FS.hpp:423-436 */
int fsgpu_update_velocity(fsgpu_ctx* c, double dt);
/* the three calls above in one kernel (the pattern of bench/update-velo-fuse.cpp:57-164):
 * mid-time average, fluxes and velocity update in one pass over the fields */
int fsgpu_momentum_update(fsgpu_ctx* c, double dt);
/* apply_velocity_bconds(fs, bconds, t), src/BoundaryConditions.hpp:215-224 */
int fsgpu_apply_velocity_bconds(fsgpu_ctx* c);
/* driver-level outflow correction of the channel cases, test/LaminarChannel.cpp:55-66,201-205;
 * mass_error (may be NULL) receives outflow - inflow before the fix */
int fsgpu_outflow_fix(fsgpu_ctx* c, double* mass_error);
/* the other channel drivers' variants of that correction, applied to the ghost columns U(-1, .) and
 * U(NX+1, .) right after apply_velocity_bconds; inflow / outflow (may be NULL) receive the two
 * fluxes as they were before the correction (dy-weighted for the last two modes) */
enum {
  FSGPU_FLOW_OUTFLOW_FIX = 1,    /* test/LaminarChannel.cpp:55-66,201-205 (= fsgpu_outflow_fix) */
  FSGPU_FLOW_OUTFLOW_FIX_DY = 2, /* test/CustomDirichletChannel.cpp:61-72,197-202, scaling/Channel.cpp:208-213 */
  FSGPU_FLOW_FORCE_TOTAL = 3     /* test/PeriodicChannel.cpp:56-67,193-203, scaling/Channel.cpp:196-206:
                                    both columns shifted so that each carries total_flow */
};
int fsgpu_flow_control(fsgpu_ctx* c, int mode, double total_flow, double* inflow, double* outflow);
/* which of the three FSGPU_STEP_OUTFLOW_FIX applies inside fsgpu_substep / fsgpu_step
 * (default FSGPU_FLOW_OUTFLOW_FIX) */
int fsgpu_set_flow_control(fsgpu_ctx* c, int mode, double total_flow);
/* calc_divergence(curr.U, curr.V, dx, dy, div), src/Operators.hpp:32-41 -> FSGPU_DIV */
int fsgpu_calc_divergence(fsgpu_ctx* c);
/* shift_pressure_to_zero(dx, dy, delta_p), src/Operators.hpp:78-82 (on FSGPU_DELTA_P) */
int fsgpu_shift_pressure_to_zero(fsgpu_ctx* c);
/* the driver-inlined projection, scaling/LidDrivenFlow.cpp:153-160 and, for variable density,
 * test/LaminarChannel.cpp:222-233: p += delta_p; U,V -= grad(delta_p)*dt/rho on interior faces */
int fsgpu_correct_velocity(fsgpu_ctx* c, double dt);
/* the three driver lines shift_pressure_to_zero + p += delta_p + velocity correction in one pass (what
 * fsgpu_substep runs); FSGPU_DELTA_P holds the shifted field afterwards, FSGPU_RHS is scratch */
int fsgpu_project(fsgpu_ctx* c, double dt);
/* abs_max(field) = abs_max<INCLUDE_GHOST = false>, src/Utility.hpp:44-88: the maximum over the
 * field's INTERIOR entries (what every driver's monitor row uses); fsgpu_abs_max_all is
 * abs_max<true>(field), the whole ghost-inclusive extent */
int fsgpu_abs_max(fsgpu_ctx* c, int field, double* out);
int fsgpu_abs_max_all(fsgpu_ctx* c, int field, double* out);
/* has_nan(field) / has_inf(field) / has_nan_or_inf(field), src/Container.hpp:191-204 (the NaN scans the drivers run
 * on delta_p and the momentum increments): number of NaN and of +-Inf entries of the whole ghost-inclusive array */
int fsgpu_count_nonfinite(fsgpu_ctx* c, int field, int64_t* n_nan, int64_t* n_inf);
/* max<INCLUDE_GHOST>(field) and min<INCLUDE_GHOST>(field), src/Utility.hpp:91-115 */
int fsgpu_field_max(fsgpu_ctx* c, int field, int include_ghost, double* out);
int fsgpu_field_min(fsgpu_ctx* c, int field, int include_ghost, double* out);
/* integrate<INCLUDE_GHOST>(dx, dy, field) and L1_norm<INCLUDE_GHOST>(dx, dy, field), src/Operators.hpp:51-76:
 * (sum of the entries, or of their magnitudes) * dx * dy over the interior or the whole extent */
int fsgpu_integrate(fsgpu_ctx* c, int field, int include_ghost, double* out);
int fsgpu_l1_norm(fsgpu_ctx* c, int field, int include_ghost, double* out);
/* calc_grad_of_centered_points(f, dx, dy, dfdx, dfdy), src/Operators.hpp:164-201: three distinct cell
 * fields (e.g. FSGPU_VF -> FSGPU_GRAD_X, FSGPU_GRAD_Y); every cell incl. ghosts is written */
int fsgpu_calc_grad_of_centered_points(fsgpu_ctx* c, int f, int dfdx, int dfdy);
/* eval_flow_field_at(xm, ym, Ui, Vi, x, y), src/Operators.hpp:118-161, at n points (px[k], py[k]) on FSGPU_UI /
 * FSGPU_VI as they are (fsgpu_interpolate_cell_centres fills them), and bilinear_interpolate(xm, ym, field, x, y),
 * :85-115, on any cell field.  xm, ym: fs.xm / fs.ym incl. ghosts (NX + 2 and NY + 2 doubles, fs.xm.get_data());
 * all arrays are host arrays.  With slab contexts the call is collective with the same points on every rank: the
 * slab owning a point's row evaluates it and the results are summed over the ranks. */
int fsgpu_eval_flow_field_at(fsgpu_ctx* c, const double* xm, const double* ym, int64_t n, const double* px,
                             const double* py, double* u, double* v);
int fsgpu_bilinear_interpolate(fsgpu_ctx* c, int field, const double* xm, const double* ym, int64_t n,
                               const double* px, const double* py, double* out);
/* calc_rho(fs)/calc_visc(fs), src/FS.hpp:569-589 and calc_rho(vf,fs)/calc_visc(vf,fs),
 * src/FS.hpp:592-650 (uses FSGPU_VF) */
int fsgpu_calc_rho_visc_const(fsgpu_ctx* c);
int fsgpu_calc_rho_visc_vof(fsgpu_ctx* c);
/* apply_neumann_bconds(field), src/BoundaryConditions.hpp:227-252 */
int fsgpu_apply_neumann_bconds(fsgpu_ctx* c, int field);

/* ---- rows either side of the single-phase path ---------------------------------- */
/* calc_drhodt(fs, drho_u_stagdt, drho_v_stagdt), src/FS.hpp:318-406 -> FSGPU_DRHO_U / FSGPU_DRHO_V */
int fsgpu_calc_drhodt(fsgpu_ctx* c);
/* update_density(drho_u_stagdt, drho_v_stagdt, dt, fs), src/FS.hpp:409-420 */
int fsgpu_update_density(fsgpu_ctx* c, double dt);
/* the density block of the two-phase drivers in one pass (test/StationaryDrop.cpp:274-293):
 * calc_drhodt + update_density + apply_neumann_bconds(curr.rho_u_stag / rho_v_stag) */
int fsgpu_density_step(fsgpu_ctx* c, double dt);
/* calc_pressure_jump(vf, curv, interface_length, fs), src/FS.hpp:439-466; reads FSGPU_VF,
 * FSGPU_CURV, FSGPU_INTERFACE_LENGTH and fs.sigma, writes FSGPU_P_JUMP_U / _V (ghosts 0) */
int fsgpu_calc_pressure_jump(fsgpu_ctx* c);
/* the capillary block of the two-phase drivers (test/StationaryDrop.cpp:331-353): keep the old
 * jump (FSGPU_PJ_OLD_*), calc_pressure_jump, apply_neumann_bconds on both jumps, and
 * div += dt * div((pj - pj_old) / rho) on all cells of FSGPU_DIV */
int fsgpu_capillary_divergence(fsgpu_ctx* c, double dt);
/* correct_velocity_ib_implicit_euler(ib_corr_u, ib_corr_v, dt, fs), src/IB.hpp:110-126 */
int fsgpu_correct_velocity_ib_implicit_euler(fsgpu_ctx* c, double dt);
/* update_velocity_ib_semi_analytical(drhoUdt, drhoVdt, dt, ib_corr_u, ib_corr_v, fs),
 * src/IB.hpp:129-186; reads FSGPU_DMOM_U / _V and FSGPU_IB_CORR_U / _V */
int fsgpu_update_velocity_ib_semi_analytical(fsgpu_ctx* c, double dt);
/* interpolate_U, interpolate_V, interpolate_UV_staggered_field(rho_u, rho_v), src/Operators.hpp:10-29
 * -> FSGPU_UI, FSGPU_VI, FSGPU_RHOI (all cells incl. ghosts), one pass */
int fsgpu_interpolate_cell_centres(fsgpu_ctx* c);
/* calc_conserved_quantities(fs, mass, momentum_x, momentum_y), src/FS.hpp:653-676 */
int fsgpu_calc_conserved_quantities(fsgpu_ctx* c, double* mass, double* momentum_x, double* momentum_y);
/* the monitor row of every driver (scaling/LidDrivenFlow.cpp:166-170) in one pass:
 * calc_divergence -> FSGPU_DIV (all cells), then abs_max(curr.U), abs_max(curr.V), abs_max(div)
 * (interior entries of each field, as abs_max's default) */
int fsgpu_monitor(fsgpu_ctx* c, double* u_max, double* v_max, double* div_max);

/* ---- pressure solver = `LinearSolver_StructHypre<Float,NX,NY,1>` -------------------------------
 * constructor (src/LinearSolver_StructHypre.hpp:43-62): tol, max_iter, PSDirichlet.  The solver
 * is always PCG preconditioned by one PFMG-style V-cycle (the reference's default
 * HypreSolver::PCG + HyprePrecond::PFMG, :31-32).
 * One DELIBERATE deviation from HYPRE's arithmetic: a coarse-grid diagonal that is pure cancellation noise
 * (|d| <= 1e-13 * sum of the magnitudes of its terms) is stored as exactly 0, so that the zero-diagonal test of
 * pfmg_setup.c:349-361 fires on the one-point coarsest grid of a singular (all-Neumann) operator whatever the
 * rounding.  HYPRE tests the rounded value: exact with dyadic coefficients, ~1e-16 otherwise, which it then
 * divides by (TGV: PCG stalls near 1e-8; IB channel: breakdown).  The setup here (symmetric 3-array storage,
 * reciprocal interpolation weights) is not bit-identical to HYPRE's, so it cannot inherit HYPRE's exact zeros
 * either.  The oracle has the same guard as a switch (oracle/fs_oracle.h snap_zero_diag), pinned against the
 * unmodified reference with and without it. */
typedef struct {
  double tol;
  int max_iter;
  int psd;        /* FSGPU_PSD_* */
  int relax;      /* FSGPU_RELAX_*; 0 = default (Jacobi, the reference's configuration) */
  int max_levels; /* 0 = default 16 = PS_PFMG_MAX_LEVELS (src/LinearSolver_StructHypre.hpp:221); values above 16
                   * coarsen further but stop at a two-point coarsest grid */
  int skip_relax; /* -1 = default (1, HYPRE's pfmg.c:38); 0/1 */
} fsgpu_ps_opts;
int fsgpu_ps_create(fsgpu_ctx* c, const fsgpu_ps_opts* o, fsgpu_ps** out);
int fsgpu_ps_destroy(fsgpu_ps* s);
/* set_pressure_operator(fs), src/LinearSolver_StructHypre.hpp:91-106 (+ LinearSystem.hpp:32-122
 * and the PFMG setup of Submodules/hypre/src/struct_ls/pfmg_setup.c) */
int fsgpu_ps_set_operator(fsgpu_ps* s);
/* set_pressure_rhs(fs, div, dt), src/LinearSolver_StructHypre.hpp:109-116 (+ LinearSystem.hpp:125-155);
 * reads FSGPU_DIV, writes FSGPU_RHS */
int fsgpu_ps_set_rhs(fsgpu_ps* s, double dt);
/* calc_divergence + set_pressure_rhs in one pass over curr.U, curr.V (what fsgpu_substep runs): writes
 * FSGPU_DIV and FSGPU_RHS; same arithmetic as the two separate calls */
int fsgpu_ps_set_rhs_from_velocity(fsgpu_ps* s, double dt);
/* solve(delta_p, &residual, &num_iter), src/LinearSolver_StructHypre.hpp:119-192 -> FSGPU_DELTA_P.
 * Non-convergence is not an error (the reference only warns, :177-183): returns 0 with
 * *iters == max_iter and *residual >= tol. */
int fsgpu_ps_solve(fsgpu_ps* s, double* residual, int* iters);
int fsgpu_ps_num_levels(fsgpu_ps* s, int* n);
/* slab decomposition: number of leading multigrid levels that are distributed over the ranks; the
 * remaining (coarse) levels are gathered and solved redundantly on every rank.  Equal to the number
 * of levels for a single rank. */
int fsgpu_ps_num_distributed_levels(fsgpu_ps* s, int* n);

/* ---- self-test --------------------------------------------------------------------------------
 * The kernels that divide by dx, dy, dt (calc_divergence, fill_pressure_rhs, the projection) form the
 * quotient from a host-made reciprocal with one FMA correction instead of an FP64 division
 * (csrc/common.cuh: DivBy); it must equal the division bit for bit.  This compares the two on n
 * pseudo-random numerators (random sign and significand, magnitudes 2^-200 ... 2^200, zeros) for the
 * divisor y: *mismatches counts the differing quotients, *fast is 0 when y is one of the divisors the
 * shortcut does not cover (significand all ones), for which the kernels divide. */
int fsgpu_selftest_division(double y, int64_t n, uint64_t seed, int* fast, int64_t* mismatches);

/* ---- file output from the device buffers --------------------------
 * Files are byte-identical to the reference's.  x and y are the reference's fs.x / fs.y arrays
 * including their ghost entries (NX + 3 and NY + 3 doubles, fs.x.get_data()).  With slab contexts
 * the calls are collective: every rank writes its own rows (npy) or columns (VTK) of the same
 * file at their offsets, rank 0 adds headers and coordinates; the file is complete once every
 * rank has returned. */
/* to_npy(filename, Field2D), src/IO.hpp:204-228: .npy v1.0, '<f8', C order, shape incl. ghosts */
int fsgpu_to_npy(fsgpu_ctx* c, int field, const char* filename);
/* to_npy(directory, fs), src/IO.hpp:231-269: x, xm, y, ym (when x and y are given) and the twelve
 * state fields under the reference's file names.  The directory is created if missing; unlike the
 * reference (init_output_directory, src/IO.hpp:158-175) it is NOT emptied first. */
int fsgpu_to_npy_state(fsgpu_ctx* c, const char* directory, const double* x, const double* y);
/* the host half of the two calls above (src/IO.hpp:25-95 + payload), exported so that the byte
 * format can be checked without a GPU; n = 0 writes a 1-D array of m entries */
int fsgpu_npy_write_host(const char* filename, const double* data, int64_t m, int64_t n);
/* VTKWriter::write, src/VTKWriter.hpp:39-152: legacy VTK, BINARY (big-endian), structured grid on
 * the interior nodes, interior cell data in j-major order.  scalar_fields[n_scalars] and
 * vector_fields[2 * n_vectors] (x then y component) are cell fields (FSGPU_P, FSGPU_DIV, FSGPU_UI,
 * FSGPU_VI, FSGPU_VF ...); the transposition and byte swap run on the device. */
int fsgpu_write_vtk(fsgpu_ctx* c, const char* filename, double t, const double* x, const double* y,
                    int n_scalars, const int* scalar_fields, const char* const* scalar_names, int n_vectors,
                    const int* vector_fields, const char* const* vector_names);

/* ---- fused fast path used by our own drivers: one projection sub-iteration
 * (scaling/LidDrivenFlow.cpp:127-161) and one whole timestep (:121-163) ----------------------- */
enum {
  FSGPU_STEP_OUTFLOW_FIX = 1,       /* test/LaminarChannel.cpp:201-205, or the variant chosen with
                                       fsgpu_set_flow_control */
  FSGPU_STEP_REBUILD_OPERATOR = 2,  /* set_pressure_operator every sub-iteration (two-phase
                                       drivers, test/StationaryDrop.cpp:357) */
  FSGPU_STEP_DENSITY = 4,           /* consistent mass transport before the momentum update
                                       (test/StationaryDrop.cpp:274-293) */
  FSGPU_STEP_CAPILLARY = 8          /* capillary block after the divergence
                                       (test/StationaryDrop.cpp:331-353); FSGPU_VF, FSGPU_CURV and
                                       FSGPU_INTERFACE_LENGTH are inputs, frozen during the step */
};
int fsgpu_substep(fsgpu_ctx* c, fsgpu_ps* s, double dt, int flags, double* residual, int* iters);
/* save_old_state + num_subiter sub-iterations; *iters = total PCG iterations of the step */
int fsgpu_step(fsgpu_ctx* c, fsgpu_ps* s, double dt, int num_subiter, int flags, double* residual,
               int* iters);

#ifdef __cplusplus
}
#endif
#endif /* FSGPU_H_ */
