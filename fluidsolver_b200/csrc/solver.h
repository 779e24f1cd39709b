// fluidsolver_b200/csrc/solver.h -- GPU-resident multigrid-preconditioned CG (internal).
//
// Replaces LinearSolver_StructHypre (src/LinearSolver_StructHypre.hpp) = HYPRE Struct PCG
// (Submodules/hypre/src/krylov/pcg.c) preconditioned by one PFMG V-cycle
// (Submodules/hypre/src/struct_ls/pfmg*.c).  The multigrid hierarchy follows PFMG's construction
// (semicoarsening direction chosen per level, operator-dependent 1-D interpolation, non-Galerkin
// 5-point coarse operators, skipped relaxation on the second level of an x/y pair), so with the
// Jacobi smoother the iterates equal the reference's up to rounding.
//
// Storage: the operator is symmetric, so a level keeps 3 coefficient arrays instead of 5:
//   d [k]  centre;  cx[k] coupling (li-1,lj)<->(li,lj);  cy[k] coupling (li,lj-1)<->(li,lj)
// plus dinv = 1/d and pc = 1/(d + transverse couplings), the interpolation denominator of
// pfmg_setup_interp.c:327-341.  All arrays share the padded layout of common.cuh.
#pragma once

#include "ctx.h"

#define FS_MAX_LEVELS 40

struct LevelDev {
  int64_t n0, n1;     // points along i (d = 0) and j (d = 1)
  int64_t pitch;      // doubles per row
  int64_t foff;       // fine index along cdir of THIS level's point 0 in its parent (0 or 1)
  double *d, *dinv, *cx, *cy, *pc;
  double *x, *b;
  __host__ __device__ inline int64_t idx(int64_t li, int64_t lj) const {
    return (li + FS_RO) * pitch + (lj + FS_CO);
  }
};

struct Level {
  LevelDev v{};
  int64_t lo[2] = {0, 0};  // global index of point (0, 0) (the fine box starts at -1)
  int64_t rows   = 0;      // allocated rows
  int cdir       = -1;     // direction coarsened to reach the next level; -1 on the coarsest
  int active     = 1;      // relax on this level
  double weight  = 1.0;    // Jacobi weight
};

// scalars of the PCG recurrence, resident on the device (mirrored to pinned host memory)
struct PcgScal {
  double gamma, sdotp, alpha, beta, bi, iprod, eps;
  int iter, done, first, coarse_zero_diag;
};
// done codes
enum { PCG_RUNNING = 0, PCG_CONVERGED = 1, PCG_BREAKDOWN = 2, PCG_ZERO_RHS = 3 };

#ifdef __CUDACC__
// the scalar part of "gamma' = <r, C r>" (executed by one thread of the last reduction block).
// INIT: bi = <C b, b> before the loop (pcg.c:430-488); otherwise the convergence test of
// pcg.c:816 and beta = gamma'/gamma.
template <bool INIT>
__device__ __forceinline__ void pcg_after_rz(PcgScal* scal, double g) {
  if (INIT) {
    scal->bi    = g;
    scal->gamma = g;
    scal->iprod = 0.0;
    scal->beta  = 0.0;
    if (!(g > 0.0)) scal->done = PCG_ZERO_RHS;  // pcg.c:473-488
  } else {
    scal->iprod = g;
    if (g / scal->bi < scal->eps) scal->done = PCG_CONVERGED;  // pcg.c:816
    else if (g <= 0.0) scal->done = PCG_BREAKDOWN;
    else {
      scal->beta  = g / scal->gamma;
      scal->gamma = g;
    }
  }
}
#endif

struct fsgpu_ps {
  fsgpu_ctx* c = nullptr;
  fsgpu_ps_opts o{};
  int nl = 0;
  Level L[FS_MAX_LEVELS];
  int coarse_sweeps = 0;
  double* slab = nullptr;  // every level's arrays + PCG vectors
  size_t slab_doubles = 0;
  double *p = nullptr, *s = nullptr, *z = nullptr;  // PCG vectors (level-0 layout)
  double* ctmp = nullptr;                            // coarsest-level ping-pong (fallback path)
  PcgScal* scal = nullptr;   // device
  PcgScal* h_scal = nullptr; // pinned, FS_LAG + 1 slots
  cudaEvent_t ev[4] = {};
  bool is_setup = false;
  bool levels_allocated = false;
  double last_dt = 0.0;
};

// solver_setup.cu
int ps_setup(fsgpu_ps* s);
// solver_cycle.cu: z = C r, one V-cycle from a zero guess (r, z in level-0 layout).
// dot_mode 0: nothing more; 1: also gamma' = <r, z> with the convergence test and beta;
// 2: also bi = <r, z> (first V-cycle).  The dot product rides on the last kernel of the cycle.
int ps_vcycle(fsgpu_ps* s, const double* r, double* z, int dot_mode);
// solver_pcg.cu
int ps_solve(fsgpu_ps* s, double* residual, int* iters);
