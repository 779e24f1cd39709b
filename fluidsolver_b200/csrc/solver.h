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
#include "p2p.cuh"

#define FS_MAX_LEVELS 40

// Slab decomposition: rows are addressed by their GLOBAL level index li; the array pointers are
// virtual bases (allocation base - first stored row * pitch) exactly like the flow fields
// (common.cuh), and kernels sweep the rows [i0, i1) this rank owns ([0, n0) for one rank and on
// the replicated coarse levels).
struct LevelDev {
  int64_t n0, n1;     // points along i (d = 0) and j (d = 1) of the GLOBAL level grid
  int64_t i0, i1;     // rows swept by a kernel launched on this level
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
  int64_t own0 = 0, own1 = 0;  // rows [own0, own1) belong to this rank under the slab partition
  bool replicated = false;     // level >= ld: every rank holds (and solves) the whole level
  int64_t voff() const { return replicated ? 0 : own0 * v.pitch; }
  int cdir       = -1;     // direction coarsened to reach the next level; -1 on the coarsest
  int active     = 1;      // relax on this level
  double weight  = 1.0;    // Jacobi weight
};

// scalars of the PCG recurrence, resident on the device (mirrored to pinned host memory)
struct PcgScal {
  double gamma, sdotp, alpha, beta, bi, iprod, eps;
  int iter, done, first, coarse_zero_diag;
  // slab decomposition: nranks > 1 makes the reduction kernels park their LOCAL sum in red[]
  // instead of finishing the recurrence; after the all-reduce k_pcg_after_* finishes it.
  // With the mailbox transport (p2p != null) the thread holding the local sum all-reduces it over
  // peer memory on the spot (p2p_allreduce_sum1) and finishes the recurrence itself: a dot product
  // of the distributed PCG is then ONE kernel, as on a single GPU.
  int nranks;
  double red[4];
  const P2PRedDev* p2p;
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
// the scalar part of "<s, p>" (k_matvec_dot): alpha = gamma / <s,p> with HYPRE's breakdown exits
__device__ __forceinline__ void pcg_after_sp(PcgScal* scal, double sp) {
  scal->iter += 1;
  scal->sdotp = sp;
  if (sp == 0.0) {  // pcg.c:760-775
    if (scal->iter == 1) scal->iprod = scal->gamma;
    scal->done = PCG_BREAKDOWN;
  } else {
    const double alpha = scal->gamma / sp;
    scal->alpha = alpha;
    if (alpha <= 0.0) {  // pcg.c:777-790
      if (scal->iter == 1) scal->iprod = scal->gamma;
      scal->done = PCG_BREAKDOWN;
    }
  }
}
// called by the last block of a reduction kernel with this rank's sum
template <bool INIT>
__device__ __forceinline__ void pcg_local_rz(PcgScal* scal, double g) {
  if (scal->nranks > 1) {
    if (scal->p2p) pcg_after_rz<INIT>(scal, p2p_allreduce_sum1(*scal->p2p, g));
    else scal->red[0] = g;
  } else {
    pcg_after_rz<INIT>(scal, g);
  }
}
__device__ __forceinline__ void pcg_local_sp(PcgScal* scal, double sp) {
  if (scal->nranks > 1) {
    if (scal->p2p) pcg_after_sp(scal, p2p_allreduce_sum1(*scal->p2p, sp));
    else scal->red[0] = sp;
  } else {
    pcg_after_sp(scal, sp);
  }
}
#endif

struct fsgpu_ps {
  fsgpu_ctx* c = nullptr;
  fsgpu_ps_opts o{};
  int nl = 0;
  int ld = 0;  // levels [0, ld) are distributed over the ranks, [ld, nl) replicated (ld == nl: 1 rank)
  Level L[FS_MAX_LEVELS];
  int coarse_sweeps = 0;
  double* slab = nullptr;  // every level's arrays + PCG vectors
  size_t slab_doubles = 0;
  double *p = nullptr, *s = nullptr, *z = nullptr;  // PCG vectors (level-0 layout)
  double* ctmp = nullptr;                            // coarsest-level ping-pong (fallback path)
  PcgScal* scal = nullptr;   // device
  PcgScal* h_scal = nullptr; // pinned, FS_LAG + 1 slots
  cudaEvent_t ev[4] = {};
  // launches sweeping at least this many points use the TMA-fed marching kernels (0: never).  Environment, read
  // by fsgpu_ps_create: FSGPU_RING=0 turns them off, FSGPU_RING_MIN_POINTS moves the threshold.
  int64_t ring_min_points = 1 << 20;
  // fsgpu_substep: the rhs still carries its mean; b_shift (device: the sum of all its entries, on every rank)
  // is non-null from the start of the solve until the first x / r update has absorbed it (solver_cycle.cu BShift)
  bool rhs_mean_pending = false;
  const double* b_shift = nullptr;
  double b_shift_size = 1.0;
  // fused push (p2p.cuh): the neighbours' level arrays as mapped here (virtual bases of b and x of every
  // distributed level, [level][0: rank - 1, 1: rank + 1]), set up by alloc_levels when the mailbox transport is on
  bool push_ready = false;
  void* nb_token = nullptr;
  double* nb_b[FS_MAX_LEVELS][2] = {};
  double* nb_x[FS_MAX_LEVELS][2] = {};
  unsigned long long push_epoch = 0;
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
// slab decomposition: all-reduce the parked local sum and finish the recurrence (no-ops for 1 rank)
int ps_finish_rz(fsgpu_ps* s, bool init);
int ps_finish_sp(fsgpu_ps* s);
// halo rows of level-l arrays given by virtual base pointers
int ps_halo(fsgpu_ps* s, int l, double* const* ptrs, int n, int h);
int ps_solve(fsgpu_ps* s, double* residual, int* iters);
// boundary-first split of level l's owned rows for an exchange of h halo rows (overlap)
struct SlabSplit { RowSplit bnd, in; bool on; };
SlabSplit ps_split(fsgpu_ps* s, int l, int h);
