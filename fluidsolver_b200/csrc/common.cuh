// fluidsolver_b200/csrc/common.cuh -- shared definitions of libfsgpu.so (sm_100a only).
//
// Device data layout (DESIGN.md "Data layout in HBM").  Every level-0 array -- flow fields and
// solver vectors alike -- is one padded 2-D array of doubles with the SAME geometry, so a kernel
// addresses U, V, rho, p, rhs, delta_p ... with one linear index:
//
//     element (a, b), a = i + 1 >= 0, b = j + 1 >= 0   (ghost-inclusive, zero based; i, j are the
//     reference's indices, src/Container.hpp:105-111)  lives at  (a + FS_RO) * pitch + (b + FS_CO)
//
// FS_RO = 2 zero rows above/below and FS_CO = 4 / >= 2 zero columns left/right stay zero for the
// lifetime of the context: stencils may read one or two points past the edge without a branch
// (coefficients there are zero), and b = 0 sits on a 32-byte boundary for vector access.
// `pitch` is a multiple of 16 doubles (128 B lines).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#define FS_RO 2
#define FS_CO 4

//
// Slab decomposition.  A context owns the ghost-inclusive rows
// a in [a_lo, a_hi) of the GLOBAL grid (one rank: [0, nx + 2)); U-like fields have one more row
// (a = nx + 2), owned by the last rank.  Kernels keep addressing rows by their GLOBAL index: the
// field pointers handed to them are "virtual" bases (allocation base - a_lo * pitch), so
// idx(a, b) lands inside the local allocation for a in [a_lo - FS_RO, a_hi + 1 + FS_RO).  The
// FS_RO rows above/below the owned range are zero pads at a physical boundary and halo rows
// (copies of the neighbour rank's rows) at an interior slab boundary; either way a stencil reads
// them without a branch.
struct Geom {
  int64_t nx, ny;      // interior cells of the GLOBAL grid
  int64_t pitch;       // doubles per row
  int64_t rows;        // allocated rows = (a_hi - a_lo) + 1 + 2 * FS_RO
  int64_t a_lo, a_hi;  // owned ghost-inclusive cell rows [a_lo, a_hi)
  double dx, dy;
  __host__ __device__ inline int64_t idx(int64_t a, int64_t b) const {
    return (a + FS_RO) * pitch + (b + FS_CO);
  }
  __host__ __device__ inline int64_t count() const { return rows * pitch; }
  // offset of the virtual base relative to the allocation base (subtract it from the latter)
  __host__ __device__ inline int64_t voff() const { return a_lo * pitch; }
  __host__ __device__ inline bool first() const { return a_lo == 0; }
  __host__ __device__ inline bool last() const { return a_hi == nx + 2; }
  // owned rows of a U-like field end one later on the last rank
  __host__ __device__ inline int64_t u_hi() const { return a_hi + (last() ? 1 : 0); }
};

// ---- x / y for a launch-uniform divisor, correctly rounded ---------------------------------------
// The reference divides by dx, dy, dt (src/Operators.hpp:36-39, src/LinearSystem.hpp:134,
// scaling/LidDrivenFlow.cpp:155-159) and the kernels must round as it does.  The compiler's FP64 division
// costs ~35 instructions per quotient and drops to a subroutine for zero / tiny numerators -- on the
// cavity's mostly quiescent field that made k_rhs instruction bound (ncu: 340 instructions per point,
// 25 % DRAM).  With r = RN(1 / y) formed once on the host,
//     q = RN(x r);   q' = RN(q + RN(x - q y) r)          (the inner remainder is exact under FMA)
// is the correctly rounded quotient (Markstein 1990) unless the significand of y is all ones, which
// make_div_by tests for (fast = 0 -> the ordinary division); huge / non-finite q takes it as well.
// Results in the denormal range may differ from the division's in their last (denormal) bit.
struct DivBy {
  double y, r;
  int fast;
#ifdef __CUDACC__
  __device__ __forceinline__ double operator()(double x) const {
    const double q = __dmul_rn(x, r);
    if (!fast || !(fabs(q) < 1e300)) return x / y;
    return fma(fma(-q, y, x), r, q);
  }
#endif
};
static inline DivBy make_div_by(double y) {
  DivBy d;
  d.y = y;
  d.r = 1.0 / y;
  unsigned long long bits;
  __builtin_memcpy(&bits, &y, sizeof(bits));
  const unsigned long long man = bits & 0xfffffffffffffull, ex = (bits >> 52) & 0x7ff;
  const double ar = d.r < 0 ? -d.r : d.r;
  d.fast = (man != 0xfffffffffffffull && ex != 0 && ex != 0x7ff && ar > 1e-290 && ar < 1e290) ? 1 : 0;
  return d;
}

// ---- deterministic grid-wide reductions -----------------------------------------------------
// Each block reduces with warp shuffles + one shared-memory stage and stores its partial at
// part[blockIdx]; the LAST block to finish (threadfence + atomic ticket) folds all partials in a
// fixed order, so results do not depend on block scheduling.
struct ReduceSum {
  __device__ static inline double op(double a, double b) { return a + b; }
  __device__ static inline double id() { return 0.0; }
};
struct ReduceMax {
  __device__ static inline double op(double a, double b) { return a > b ? a : b; }
  __device__ static inline double id() { return -1.0e308; }
};

template <class R>
__device__ inline double warp_reduce(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = R::op(v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}

// all threads of the block must call; result valid in thread 0
template <class R>
__device__ inline double block_reduce(double v, double* sm /* >= 32 doubles */) {
  const int tid  = threadIdx.x + threadIdx.y * blockDim.x;
  const int nthr = blockDim.x * blockDim.y;
  const int lane = tid & 31, w = tid >> 5;
  v = warp_reduce<R>(v);
  __syncthreads();
  if (lane == 0) sm[w] = v;
  __syncthreads();
  if (w == 0) {
    v = (lane < (nthr + 31) / 32) ? sm[lane] : R::id();
    v = warp_reduce<R>(v);
  }
  return v;
}

// NV values per thread reduced at once.  Returns true in ALL threads of the last block, with the
// grid totals in out[0..NV) (valid in thread 0 of that block).  part must hold NV * nblocks doubles.
template <class R, int NV>
__device__ inline bool grid_reduce(const double (&v)[NV], double* part, unsigned int* ticket,
                                   double (&out)[NV]) {
  __shared__ double sm[32];
  __shared__ bool last;
  const int tid     = threadIdx.x + threadIdx.y * blockDim.x;
  const int nthr    = blockDim.x * blockDim.y;
  const unsigned nb = gridDim.x * gridDim.y;
  const unsigned bid = blockIdx.x + blockIdx.y * gridDim.x;
#pragma unroll
  for (int q = 0; q < NV; ++q) {
    const double r = block_reduce<R>(v[q], sm);
    if (tid == 0) part[(size_t)q * nb + bid] = r;
  }
  __threadfence();
  if (tid == 0) {
    const unsigned t = atomicInc(ticket, nb - 1);  // wraps to 0 after the last block
    last = (t == nb - 1);
  }
  __syncthreads();
  if (!last) return false;
  __threadfence();
#pragma unroll
  for (int q = 0; q < NV; ++q) {
    double a = R::id();
    for (unsigned k = tid; k < nb; k += nthr) a = R::op(a, __ldcg(&part[(size_t)q * nb + k]));
    a = block_reduce<R>(a, sm);
    out[q] = a;
  }
  return true;
}

// ---- persistent tile loop for kernels that end in a grid-wide reduction ------------------------
// A (64 x 4)-thread block walks 16-row x 64-column tiles with a grid stride, so the grid is a few
// CTAs per SM (persist_blocks) instead of one CTA per 256 points: the reduction then costs
// ~1e3 partials + tickets instead of ~3e5 serialised atomics on one address (measured: 0.8 ms per
// reduction at 8192^2 before this change).
// rows [r0, r1) x columns [0, n1)
template <class F>
__device__ __forceinline__ void for_each_point(int64_t r0, int64_t r1, int64_t n1, F&& f) {
  // tile arithmetic in 32 bits (a 64-bit division per tile and thread was a quarter of the instructions of
  // the element-wise kernels); only the element index the callback forms is 64-bit
  const int n0 = (int)(r1 - r0);
  const unsigned tiles_j = (unsigned)((n1 + 63) / 64), nt = tiles_j * (unsigned)((n0 + 15) / 16);
  for (unsigned t = blockIdx.x; t < nt; t += gridDim.x) {
    const unsigned ti = t / tiles_j, tj = t - ti * tiles_j;
    const int lj = (int)(tj * 64 + threadIdx.x);
    if (lj < (int)n1) {
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        const int i = (int)(ti * 16 + threadIdx.y) + 4 * m;
        if (i < n0) f(r0 + i, (int64_t)lj);
      }
    }
  }
}
// two row ranges [a0, a1) and [b0, b1) (either may be empty): the boundary rows at both ends of a slab,
// swept by ONE launch ahead of the interior so that their halo exchange overlaps the interior launch
struct RowSplit { int64_t a0, a1, b0, b1; };
template <class F>
__device__ __forceinline__ void for_each_point2(const RowSplit& R, int64_t n1, F&& f) {
  if (R.a1 > R.a0) for_each_point(R.a0, R.a1, n1, f);
  if (R.b1 > R.b0) for_each_point(R.b0, R.b1, n1, f);
}
#define FS_NUM_SMS 148
static inline unsigned persist_blocks(int64_t n0, int64_t n1, int per_sm = 8) {
  const int64_t nt = ((n1 + 63) / 64) * ((n0 + 15) / 16);
  const int64_t cap = (int64_t)FS_NUM_SMS * per_sm;
  return (unsigned)(nt < cap ? (nt > 0 ? nt : 1) : cap);
}

// ---- error handling / launch accounting (api.cu) ---------------------------------------------
void fs_set_error(const char* fmt, ...);
#include <atomic>
extern std::atomic<int64_t> g_launches;  // contexts may be driven by different host threads
#define FS_CUDA(x)                                                                       \
  do {                                                                                   \
    cudaError_t e_ = (x);                                                                \
    if (e_ != cudaSuccess) {                                                             \
      fs_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #x, cudaGetErrorString(e_));   \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)
#define FS_LAUNCHED()                                                                    \
  do {                                                                                   \
    g_launches.fetch_add(1, std::memory_order_relaxed);                                                                     \
    cudaError_t e_ = cudaPeekAtLastError();                                              \
    if (e_ != cudaSuccess) {                                                             \
      fs_set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---- optional per-kernel-class timing with CUDA events on the launching stream (api.cu) -------
// bench.py enables ONE class at a time (fsgpu_prof_enable) to get the average launch duration of
// the dominant kernel inside the timed region; disabled (the default) it costs one branch.
enum {
  PROF_MOMENTUM = 0, PROF_BC, PROF_RHS, PROF_CORRECT, PROF_MATVEC, PROF_XR, PROF_DOT, PROF_PUPD,
  PROF_DOWN0, PROF_UP0, PROF_DOWN, PROF_UP, PROF_COARSE, PROF_SETUP, PROF_OTHER, PROF_NCLASS
};
extern unsigned int g_prof_mask;
extern thread_local int g_prof_tag;  // stamped on every record (PCG: the iteration number)
size_t fs_prof_mark(void);
// drop records [from, end) with tag > last_real_tag, and those of class cls_empty_at_last with tag == last_real_tag
void fs_prof_drop_after(size_t from, int last_real_tag, int cls_empty_at_last);
void fs_prof_begin(int cls, cudaStream_t st);
void fs_prof_end(int cls, cudaStream_t st);
// FSGPU_NCU_CLASSES=<bit mask of PROF_*> (read once): launches inside scopes of those classes are bracketed by
// cudaProfilerStart / Stop, so `ncu --profile-from-start off` captures exactly them (e.g. only the
// level-0 kernels of the V-cycle) instead of the thousands of small coarse-level launches of a step
extern unsigned int g_ncu_mask;
void fs_ncu_range(bool begin);
struct ProfScope {
  int cls;
  cudaStream_t st;
  bool on, ncu;
  ProfScope(int c, cudaStream_t s) : cls(c), st(s), on((g_prof_mask >> c) & 1u), ncu((g_ncu_mask >> c) & 1u) {
    if (on) fs_prof_begin(cls, st);
    if (ncu) fs_ncu_range(true);
  }
  ~ProfScope() {
    if (ncu) fs_ncu_range(false);
    if (on) fs_prof_end(cls, st);
  }
};
