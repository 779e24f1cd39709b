// fluidsolver_b200/csrc/ctx.h -- host-side context of libfsgpu.so (internal).
#pragma once

#include "../../include/fsgpu.h"
#include "common.cuh"

struct BCondSide {
  int kind    = FSGPU_BC_NEUMANN;
  double u    = 0.0, v = 0.0;
  int clipped = 0;
  double* u_prof = nullptr;  // device, optional
  double* v_prof = nullptr;
};

struct Comm;  // comm.cu: NCCL (one process per GPU) or the in-process test transport

struct fsgpu_ctx {
  Geom g{};
  int device = 0;
  int rank = 0, nranks = 1;   // slab decomposition along i
  Comm* comm = nullptr;       // null for a single rank
  cudaStream_t stream = nullptr;
  double rho_gas = 1, rho_liquid = 1, visc_gas = 1e-3, visc_liquid = 1e-3, sigma = 0;
  int flow_mode = FSGPU_FLOW_OUTFLOW_FIX;  // what FSGPU_STEP_OUTFLOW_FIX applies (fsgpu_set_flow_control)
  double flow_total = 0.0;
  double* slab = nullptr;              // one allocation for all fields
  double* f[FSGPU_NUM_FIELDS] = {};    // VIRTUAL field base pointers (allocation base - g.voff())
  double* ext_alloc[FSGPU_NUM_FIELDS] = {};  // extension fields: own allocations, made on first use
  double *U_alt = nullptr, *V_alt = nullptr;  // ping-pong targets of the fused momentum update
  BCondSide bc[4];
  // reduction scratch
  double* part = nullptr;       // block partials
  size_t part_cap = 0;          // doubles
  unsigned int* ticket = nullptr;
  double* red_out = nullptr;    // device result slots (8 doubles)
  double* h_red = nullptr;      // pinned mirror
  struct OutStage* out = nullptr;  // output.cu: pinned / device staging of the file writers (made on first use)
};

// argument checks of the C-ABI entry points (api.cu, output.cu)
#define FS_CHECK_CTX(c)                           \
  do {                                            \
    if (!(c)) {                                   \
      fs_set_error("%s: null context", __func__); \
      return 1;                                   \
    }                                             \
    FS_CUDA(cudaSetDevice((c)->device));          \
  } while (0)
#define FS_CHECK_FIELD(f)                                  \
  do {                                                     \
    if ((f) < 0 || (f) >= FSGPU_NUM_FIELDS) {              \
      fs_set_error("%s: bad field id %d", __func__, (f));  \
      return 1;                                            \
    }                                                      \
    if (fs_ensure_field(c, (f))) return 1;                 \
  } while (0)

// field classes: extents of the ghost-inclusive reference array
static inline void field_extent(const Geom& g, int field, int64_t* m, int64_t* n) {
  switch (field) {
    case FSGPU_P_JUMP_U: case FSGPU_RHO_U_OLD: case FSGPU_U_OLD: case FSGPU_RHO_U: case FSGPU_U:
    case FSGPU_DMOM_U: case FSGPU_DRHO_U: case FSGPU_PJ_OLD_U: case FSGPU_IB_CORR_U:
      *m = g.nx + 3; *n = g.ny + 2; break;
    case FSGPU_P_JUMP_V: case FSGPU_RHO_V_OLD: case FSGPU_V_OLD: case FSGPU_RHO_V: case FSGPU_V:
    case FSGPU_DMOM_V: case FSGPU_DRHO_V: case FSGPU_PJ_OLD_V: case FSGPU_IB_CORR_V:
      *m = g.nx + 2; *n = g.ny + 3; break;
    default:
      *m = g.nx + 2; *n = g.ny + 2; break;
  }
}

// allocation base of an array whose virtual base is v (for whole-array memset / memcpy)
static inline double* alloc_base(const Geom& g, double* v) { return v + g.voff(); }

// ---- comm.cu: halo exchange and reductions between the slabs ------------------------------------
// All of these are no-ops for a single rank.  With NCCL everything is enqueued on c->stream.
enum { COMM_SUM = 0, COMM_MAX = 1 };
int comm_create(fsgpu_ctx* c, int rank, int nranks, const void* id128, int transport);
void comm_destroy(fsgpu_ctx* c);
int comm_is_direct(fsgpu_ctx* c);
int comm_failed(fsgpu_ctx* c);  // a direct (FSGPU_HALO=p2p) halo exchange timed out
// arrays given by VIRTUAL base pointers sharing `pitch`; this rank owns rows [i0, i1) (global
// indices).  Fills the h rows above and below from the neighbour ranks' owned rows.
int comm_halo(fsgpu_ctx* c, double* const* ptrs, int n, int64_t pitch, int64_t i0, int64_t i1, int h);
// the same exchange on the transport's own stream, ordered after what the compute stream holds NOW; the
// compute stream goes on (interior rows) and meets it again at comm_halo_wait.  *async_out = false: not
// available for this exchange (no mailbox, too large, too few rows) -- use comm_halo instead.
int comm_halo_async(fsgpu_ctx* c, double* const* ptrs, int n, int64_t pitch, int64_t i0, int64_t i1, int h,
                    bool* async_out);
int comm_halo_wait(fsgpu_ctx* c);
int comm_alloc_fence(fsgpu_ctx* c);  // after an allocation made once the ranks talk (in-process ranks only; see comm.cu)
int64_t comm_overlap_min_points(fsgpu_ctx* c);  // 0: no overlap; else the smallest slab (points) worth splitting
struct P2PRedDev;
const P2PRedDev* comm_red_dev(fsgpu_ctx* c);  // device struct of the in-kernel scalar all-reduce (null: none)
int comm_allreduce(fsgpu_ctx* c, double* dev, int64_t n, int op);
// fused push (p2p.cuh): collectives used when the solver (re)allocates its levels
struct HaloPush;
struct HaloWait;
int comm_allgather_host(fsgpu_ctx* c, const void* mine, size_t bytes, void* all);
int comm_map_neighbours(fsgpu_ctx* c, void* base, void** up, void** dn, void** token);
void comm_unmap_neighbours(fsgpu_ctx* c, void* token);
int comm_push_enabled(fsgpu_ctx* c);
unsigned long long comm_push_next_epoch(fsgpu_ctx* c);  // epochs are per context: its mailbox flags outlive a solver
void comm_push_flags(fsgpu_ctx* c, int level, int kind, HaloPush* p, HaloWait* w);
// n doubles to / from one peer rank (the x-periodic wrap-around); peer < 0: not a participant
int comm_sendrecv(fsgpu_ctx* c, int peer, const double* const* send, double* const* recv, int narr, int64_t n);
// halo rows of flow fields (ids), h rows each
int fs_exchange_fields(fsgpu_ctx* c, const int* fields, int n, int h);

// ---- output.cu -------------------------------------------------------------------------------
void fs_output_release(fsgpu_ctx* c);

// ---- stencil.cu -------------------------------------------------------------------------------
int fs_launch_mid_time(fsgpu_ctx* c);
int fs_launch_dmomdt(fsgpu_ctx* c);
int fs_launch_update_velocity(fsgpu_ctx* c, double dt);
int fs_launch_momentum_update(fsgpu_ctx* c, double dt);
int fs_launch_bconds(fsgpu_ctx* c);
int fs_launch_flow_control(fsgpu_ctx* c, int mode, double total_flow, double* d_mass_error);
int fs_launch_divergence(fsgpu_ctx* c);
// rhs = -vol*div/dt (+ Dirichlet row zeroing); if psd == NONE the mean over all cells is
// subtracted (two passes).  div_from_uv: compute div on the fly from U,V and store it too.
int fs_launch_rhs(fsgpu_ctx* c, double dt, int psd, bool div_from_uv, bool defer_mean = false, bool* deferred = nullptr);
int fs_launch_shift_pressure(fsgpu_ctx* c);
int fs_launch_correct(fsgpu_ctx* c, double dt, bool shift, bool sum_ready);
int fs_launch_cfl(fsgpu_ctx* c);               // 4 maxima -> red_out[0..3]
int fs_launch_absmax(fsgpu_ctx* c, int field, bool include_ghost);  // -> red_out[0]
int fs_launch_count_nonfinite(fsgpu_ctx* c, int field);  // -> red_out[0] NaNs, red_out[1] Infs
int fs_launch_reduce_field(fsgpu_ctx* c, int field, bool include_ghost, int op);  // 0 max|f|, 1 sum, 2 sum|f| (* dx dy)
int fs_launch_fill(fsgpu_ctx* c, int field, double v);
int fs_launch_rho_visc_vof(fsgpu_ctx* c);
int fs_launch_neumann(fsgpu_ctx* c, double* f, int64_t m, int64_t n);
int fs_ensure_part(fsgpu_ctx* c, size_t doubles);

// ---- twophase.cu ------------------------------------------------------------
int fs_ensure_field(fsgpu_ctx* c, int field);  // allocate an extension field on first use
int fs_launch_drhodt(fsgpu_ctx* c);
int fs_launch_update_density(fsgpu_ctx* c, double dt);
int fs_launch_density_step(fsgpu_ctx* c, double dt, bool mid);
int fs_launch_pressure_jump(fsgpu_ctx* c);
int fs_launch_capillary_divergence(fsgpu_ctx* c, double dt);
int fs_launch_ib_implicit_euler(fsgpu_ctx* c, double dt);
int fs_launch_ib_semi_analytical(fsgpu_ctx* c, double dt);
int fs_launch_interpolate(fsgpu_ctx* c);
int fs_launch_conserved(fsgpu_ctx* c);  // -> red_out[4..6]
int fs_launch_monitor(fsgpu_ctx* c);    // -> red_out[4..6]
int fs_launch_grad_centered(fsgpu_ctx* c, int f, int dfdx, int dfdy);
int fs_sample_points(fsgpu_ctx* c, int mode, int fa, int fb, const double* xm, const double* ym, int64_t npts,
                     const double* px, const double* py, double* oa, double* ob);
