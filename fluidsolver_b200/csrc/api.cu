// fluidsolver_b200/csrc/api.cu -- the C ABI of include/fsgpu.h.
#include "solver.h"

#include <cuda_profiler_api.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <new>
#include <set>
#include <vector>

std::atomic<int64_t> g_launches{0};
unsigned int g_prof_mask = 0;
unsigned int g_ncu_mask = getenv("FSGPU_NCU_CLASSES") ? (unsigned int)strtoul(getenv("FSGPU_NCU_CLASSES"), nullptr, 0) : 0u;
void fs_ncu_range(bool begin) {
  if (begin) cudaProfilerStart();
  else cudaProfilerStop();
}
thread_local int g_prof_tag = 0;

namespace {
// Kernels are loaded with the library, not at their first launch: under CUDA's default lazy loading a first launch
// waits for the device to drain, and a device whose running kernel waits for a neighbour rank's flag (mailbox
// transport, fused push) drains late -- or, with in-process ranks on one device, never.
// Only takes effect if CUDA has not started yet and the user has not chosen a mode.
struct EagerModuleLoading {
  EagerModuleLoading() { setenv("CUDA_MODULE_LOADING", "EAGER", 0); }
} g_eager_module_loading;
struct ProfRec { int cls; cudaEvent_t a, b; int tag; };
// per host thread: the in-process transport drives one context per thread, and a thread reads back its own records
thread_local std::vector<ProfRec> g_prof_recs;
thread_local std::vector<cudaEvent_t> g_prof_pool;
thread_local cudaEvent_t g_prof_open[PROF_NCLASS];
cudaEvent_t prof_event() {
  cudaEvent_t e;
  if (!g_prof_pool.empty()) { e = g_prof_pool.back(); g_prof_pool.pop_back(); }
  else cudaEventCreate(&e);
  return e;
}
}  // namespace
void fs_prof_begin(int cls, cudaStream_t st) {
  g_prof_open[cls] = prof_event();
  cudaEventRecord(g_prof_open[cls], st);
}
void fs_prof_end(int cls, cudaStream_t st) {
  cudaEvent_t e = prof_event();
  cudaEventRecord(e, st);
  g_prof_recs.push_back({cls, g_prof_open[cls], e, g_prof_tag});
}
// PCG launches one iteration past convergence (its kernels return at their first instruction);
// the solver tags records with the iteration they belong to and, once it knows how many
// iterations were real, drops the records of the empty ones so that averages are over launches
// that did work.
size_t fs_prof_mark(void) { return g_prof_recs.size(); }
void fs_prof_drop_after(size_t from, int last_real_tag, int cls_empty_at_last) {
  size_t w = from;
  for (size_t k = from; k < g_prof_recs.size(); ++k) {
    if (g_prof_recs[k].tag > last_real_tag ||
        (g_prof_recs[k].tag == last_real_tag && g_prof_recs[k].cls == cls_empty_at_last)) {
      g_prof_pool.push_back(g_prof_recs[k].a);
      g_prof_pool.push_back(g_prof_recs[k].b);
    } else {
      g_prof_recs[w++] = g_prof_recs[k];
    }
  }
  g_prof_recs.resize(w);
}
static thread_local char g_err[512] = "";

void fs_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}


// DivBy (common.cuh) against the hardware-exact division on n pseudo-random numerators
__global__ void k_selftest_division(const DivBy q, int64_t n, unsigned long long seed, unsigned long long* bad) {
  unsigned long long local = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    unsigned long long z = seed + 0x9e3779b97f4a7c15ull * (unsigned long long)(i + 1);  // splitmix64
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    z ^= z >> 31;
    // random sign and significand, exponent in [2^-200, 2^200) (every 64th sample exactly 0)
    const unsigned long long ex = 1023 - 200 + ((z >> 52) % 400);
    double x = __longlong_as_double((long long)((z & 0x800fffffffffffffull) | (ex << 52)));
    if ((i & 63) == 0) x = 0.0;
    const double a = q(x), b = x / q.y;
    if (__double_as_longlong(a) != __double_as_longlong(b) && fabs(b) >= 2.3e-308) ++local;
  }
  if (local) atomicAdd(bad, local);
}

extern "C" {

const char* fsgpu_last_error(void) { return g_err; }

int fsgpu_selftest_division(double y, int64_t n, uint64_t seed, int* fast, int64_t* mismatches) {
  const DivBy q = make_div_by(y);
  unsigned long long* d = nullptr;
  FS_CUDA(cudaMalloc(&d, sizeof(*d)));
  FS_CUDA(cudaMemset(d, 0, sizeof(*d)));
  k_selftest_division<<<148 * 8, 256>>>(q, n, (unsigned long long)seed, d);
  unsigned long long h = 0;
  FS_CUDA(cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost));
  cudaFree(d);
  if (fast) *fast = q.fast;
  if (mismatches) *mismatches = (int64_t)h;
  return 0;
}
int64_t fsgpu_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int fsgpu_prof_enable(unsigned int class_mask) {
  g_prof_mask = class_mask;
  return 0;
}
// sums the recorded intervals per class (synchronises the device) and clears the records
int fsgpu_prof_read(double* ms_total, int64_t* count, int nclass) {
  FS_CUDA(cudaDeviceSynchronize());
  for (int q = 0; q < nclass; ++q) { ms_total[q] = 0.0; count[q] = 0; }
  for (const ProfRec& r : g_prof_recs) {
    float ms = 0.f;
    FS_CUDA(cudaEventElapsedTime(&ms, r.a, r.b));
    if (r.cls < nclass) { ms_total[r.cls] += ms; count[r.cls] += 1; }
    g_prof_pool.push_back(r.a);
    g_prof_pool.push_back(r.b);
  }
  g_prof_recs.clear();
  return 0;
}

int fsgpu_device_count(int* n) {
  FS_CUDA(cudaGetDeviceCount(n));
  return 0;
}

int fsgpu_slab_range(int64_t nx, int rank, int nranks, int64_t* a_lo, int64_t* a_hi) {
  if (nranks < 1 || rank < 0 || rank >= nranks || nx < 2) {
    fs_set_error("fsgpu_slab_range: bad arguments (rank %d of %d)", rank, nranks);
    return 1;
  }
  const int64_t n = nx + 2;
  if (n / nranks < 2 * FS_RO) {
    fs_set_error("fsgpu_slab_range: %lld rows over %d ranks leaves fewer than %d rows per slab",
                 (long long)n, nranks, 2 * FS_RO);
    return 1;
  }
  if (a_lo) *a_lo = n * rank / nranks;
  if (a_hi) *a_hi = n * (rank + 1) / nranks;
  return 0;
}

int fsgpu_slab_coarse_range(int64_t lo, int64_t hi, int64_t foff, int64_t* c_lo, int64_t* c_hi) {
  auto ceil_div2 = [](int64_t a) { return a >= 0 ? (a + 1) / 2 : -((-a) / 2); };
  if (c_lo) *c_lo = ceil_div2(lo - foff);
  if (c_hi) *c_hi = ceil_div2(hi - foff);
  return 0;
}

int fsgpu_create_slab(int64_t nx, int64_t ny, double dx, double dy, int device, int rank, int nranks,
                      const void* id128, int transport, fsgpu_ctx** out) {
  if (!out || nx < 2 || ny < 2 || !(dx > 0) || !(dy > 0)) {
    fs_set_error("fsgpu_create: bad arguments (nx=%lld ny=%lld)", (long long)nx, (long long)ny);
    return 1;
  }
  int64_t a_lo, a_hi;
  if (fsgpu_slab_range(nx, rank, nranks, &a_lo, &a_hi)) return 1;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    fs_set_error("fsgpu_create: no CUDA device (this library has no CPU fallback)");
    return 1;
  }
  FS_CUDA(cudaSetDevice(device));
  fsgpu_ctx* c = new (std::nothrow) fsgpu_ctx();
  if (!c) { fs_set_error("out of host memory"); return 1; }
  c->device = device;
  c->g.nx = nx; c->g.ny = ny; c->g.dx = dx; c->g.dy = dy;
  c->g.a_lo = a_lo; c->g.a_hi = a_hi;
  c->g.pitch = (ny + 3 + FS_CO + 2 + 15) / 16 * 16;
  c->g.rows  = (a_hi - a_lo) + 1 + 2 * FS_RO;
  const size_t n = (size_t)c->g.count();
  const int nf = FSGPU_NUM_CORE_FIELDS + 2;
  FS_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  // + slack: bulk-copy row segments (solver_cycle.cu) of the last strip may run past the last row's pitch
  FS_CUDA(cudaMalloc(&c->slab, (n * nf + 512) * sizeof(double)));
  FS_CUDA(cudaMemsetAsync(c->slab, 0, (n * nf + 512) * sizeof(double), c->stream));
  for (int f = 0; f < FSGPU_NUM_CORE_FIELDS; ++f) c->f[f] = c->slab + n * f - c->g.voff();
  c->U_alt = c->slab + n * FSGPU_NUM_CORE_FIELDS - c->g.voff();
  c->V_alt = c->slab + n * (FSGPU_NUM_CORE_FIELDS + 1) - c->g.voff();
  // block partials of every grid-wide reduction: the persistent grids are <= 148 x 8 CTAs x 4 values, so
  // the hot path never allocates (fs_ensure_part only grows it for a larger request)
  if (fs_ensure_part(c, (size_t)FS_NUM_SMS * 8 * 4 * 2)) { fsgpu_destroy(c); return 1; }
  FS_CUDA(cudaMalloc(&c->ticket, sizeof(unsigned int)));
  FS_CUDA(cudaMemsetAsync(c->ticket, 0, sizeof(unsigned int), c->stream));
  FS_CUDA(cudaMalloc(&c->red_out, 16 * sizeof(double)));
  FS_CUDA(cudaMemsetAsync(c->red_out, 0, 16 * sizeof(double), c->stream));
  FS_CUDA(cudaMallocHost(&c->h_red, 16 * sizeof(double)));
  FS_CUDA(cudaStreamSynchronize(c->stream));
  if (comm_create(c, rank, nranks, id128, transport)) {
    fsgpu_destroy(c);
    return 1;
  }
  *out = c;
  return 0;
}

int fsgpu_create(int64_t nx, int64_t ny, double dx, double dy, int device, fsgpu_ctx** out) {
  return fsgpu_create_slab(nx, ny, dx, dy, device, 0, 1, nullptr, FSGPU_TRANSPORT_NCCL, out);
}

static std::set<fsgpu_ps*> g_live_solvers;  // so that destroy order (solver / context) is free
static std::mutex g_live_mu;                // contexts may be driven by different host threads
static int ps_destroy_impl(fsgpu_ps* s);

int fsgpu_destroy(fsgpu_ctx* c) {
  if (!c) return 0;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  std::vector<fsgpu_ps*> mine;
  {
    std::lock_guard<std::mutex> lk(g_live_mu);
    for (fsgpu_ps* s : g_live_solvers)
      if (s->c == c) mine.push_back(s);
  }
  for (fsgpu_ps* s : mine) ps_destroy_impl(s);
  comm_destroy(c);
  for (int s = 0; s < 4; ++s) {
    if (c->bc[s].u_prof) cudaFree(c->bc[s].u_prof);
    if (c->bc[s].v_prof) cudaFree(c->bc[s].v_prof);
  }
  cudaFree(c->slab);
  for (int f = 0; f < FSGPU_NUM_FIELDS; ++f)
    if (c->ext_alloc[f]) cudaFree(c->ext_alloc[f]);
  cudaFree(c->part);
  cudaFree(c->ticket);
  fs_output_release(c);
  cudaFree(c->red_out);
  cudaFreeHost(c->h_red);
  cudaStreamDestroy(c->stream);
  delete c;
  return 0;
}

int fsgpu_halo_is_direct(fsgpu_ctx* c, int* on) {
  FS_CHECK_CTX(c);
  if (on) *on = comm_is_direct(c);
  return 0;
}

int fsgpu_set_props(fsgpu_ctx* c, double rho_gas, double rho_liquid, double visc_gas,
                    double visc_liquid, double sigma) {
  FS_CHECK_CTX(c);
  c->rho_gas = rho_gas; c->rho_liquid = rho_liquid;
  c->visc_gas = visc_gas; c->visc_liquid = visc_liquid; c->sigma = sigma;
  return 0;
}

int fsgpu_upload(fsgpu_ctx* c, int field, const double* host) {
  FS_CHECK_CTX(c);
  FS_CHECK_FIELD(field);
  int64_t m, n;
  field_extent(c->g, field, &m, &n);
  // this rank's rows and its halo rows, straight from the global array
  const int64_t r0 = c->g.a_lo - FS_RO > 0 ? c->g.a_lo - FS_RO : 0;
  const int64_t r1 = c->g.a_hi + FS_RO < m ? c->g.a_hi + FS_RO : m;
  FS_CUDA(cudaMemcpy2DAsync(c->f[field] + c->g.idx(r0, 0), c->g.pitch * sizeof(double),
                            host + r0 * n, n * sizeof(double), n * sizeof(double), r1 - r0,
                            cudaMemcpyHostToDevice, c->stream));
  // pageable host memory: the copy may still be staged when the call returns
  FS_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

int fsgpu_download(fsgpu_ctx* c, int field, double* host) {
  FS_CHECK_CTX(c);
  FS_CHECK_FIELD(field);
  int64_t m, n;
  field_extent(c->g, field, &m, &n);
  // the rows this rank owns (a U-like field's extra row belongs to the last rank)
  const int64_t r0 = c->g.a_lo, r1 = c->g.last() ? m : c->g.a_hi;
  FS_CUDA(cudaMemcpy2DAsync(host + r0 * n, n * sizeof(double), c->f[field] + c->g.idx(r0, 0),
                            c->g.pitch * sizeof(double), n * sizeof(double), r1 - r0,
                            cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

int fsgpu_fill(fsgpu_ctx* c, int field, double value) {
  FS_CHECK_CTX(c);
  FS_CHECK_FIELD(field);
  return fs_launch_fill(c, field, value);
}

int fsgpu_sync(fsgpu_ctx* c) {
  FS_CHECK_CTX(c);
  FS_CUDA(cudaStreamSynchronize(c->stream));
  return comm_failed(c);
}

int fsgpu_device_field(fsgpu_ctx* c, int field, double** ptr, int64_t* pitch, int64_t* row0,
                       int64_t* col0) {
  FS_CHECK_CTX(c);
  FS_CHECK_FIELD(field);
  if (ptr) *ptr = c->f[field];
  if (pitch) *pitch = c->g.pitch;
  if (row0) *row0 = FS_RO;
  if (col0) *col0 = FS_CO;
  return 0;
}

int fsgpu_stream(fsgpu_ctx* c, void** cuda_stream) {
  FS_CHECK_CTX(c);
  *cuda_stream = (void*)c->stream;
  return 0;
}

int fsgpu_set_bconds(fsgpu_ctx* c, int side, int kind, double u_val, double v_val,
                     const double* u_prof, const double* v_prof, int clipped) {
  FS_CHECK_CTX(c);
  if (side < 0 || side > 3 || kind < 0 || kind > 3) {
    fs_set_error("fsgpu_set_bconds: bad side/kind %d/%d", side, kind);
    return 1;
  }
  BCondSide& b = c->bc[side];
  b.kind = kind; b.u = u_val; b.v = v_val; b.clipped = clipped;
  const bool lr = side == FSGPU_LEFT || side == FSGPU_RIGHT;
  const size_t nu = (size_t)(lr ? c->g.ny + 2 : c->g.nx + 3);
  const size_t nv = (size_t)(lr ? c->g.ny + 3 : c->g.nx + 2);
  FS_CUDA(cudaStreamSynchronize(c->stream));
  if (b.u_prof) { cudaFree(b.u_prof); b.u_prof = nullptr; }
  if (b.v_prof) { cudaFree(b.v_prof); b.v_prof = nullptr; }
  if (u_prof) {
    FS_CUDA(cudaMalloc(&b.u_prof, nu * sizeof(double)));
    FS_CUDA(cudaMemcpy(b.u_prof, u_prof, nu * sizeof(double), cudaMemcpyHostToDevice));
  }
  if (v_prof) {
    FS_CUDA(cudaMalloc(&b.v_prof, nv * sizeof(double)));
    FS_CUDA(cudaMemcpy(b.v_prof, v_prof, nv * sizeof(double), cudaMemcpyHostToDevice));
  }
  return 0;
}

static int read_red(fsgpu_ctx* c, int off, int n) {
  FS_CUDA(cudaMemcpyAsync(c->h_red + off, c->red_out + off, n * sizeof(double),
                          cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(cudaStreamSynchronize(c->stream));
  return comm_failed(c);
}

int fsgpu_adjust_dt(fsgpu_ctx* c, double cfl_max, double dt_max, double* dt) {
  FS_CHECK_CTX(c);
  if (fs_launch_cfl(c)) return 1;
  if (read_red(c, 0, 4)) return 1;
  double st = 0.0;
  if (c->sigma > 0.0) {  // FS.hpp:108-114
    st = 1.0 / sqrt(((c->rho_gas + c->rho_liquid) * pow(c->g.dx * c->g.dy, 3.0 / 2.0)) /
                    (4.0 * M_PI * c->sigma));
  }
  double m = st;
  for (int q = 0; q < 4; ++q) m = fmax(m, c->h_red[q]);
  *dt = fmin(cfl_max / m, dt_max);
  return 0;
}

static int copy_field(fsgpu_ctx* c, int dst, int src) {
  FS_CUDA(cudaMemcpyAsync(alloc_base(c->g, c->f[dst]), alloc_base(c->g, c->f[src]),
                          (size_t)c->g.count() * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  return 0;
}
int fsgpu_save_old_velocity(fsgpu_ctx* c) {
  FS_CHECK_CTX(c);
  return copy_field(c, FSGPU_U_OLD, FSGPU_U) || copy_field(c, FSGPU_V_OLD, FSGPU_V);
}
int fsgpu_save_old_density(fsgpu_ctx* c) {
  FS_CHECK_CTX(c);
  return copy_field(c, FSGPU_RHO_U_OLD, FSGPU_RHO_U) || copy_field(c, FSGPU_RHO_V_OLD, FSGPU_RHO_V);
}
int fsgpu_save_old_state(fsgpu_ctx* c) {
  return fsgpu_save_old_density(c) || fsgpu_save_old_velocity(c);
}

int fsgpu_calc_mid_time(fsgpu_ctx* c) { FS_CHECK_CTX(c); return fs_launch_mid_time(c); }
int fsgpu_calc_dmomdt(fsgpu_ctx* c) { FS_CHECK_CTX(c); return fs_launch_dmomdt(c); }
int fsgpu_update_velocity(fsgpu_ctx* c, double dt) { FS_CHECK_CTX(c); return fs_launch_update_velocity(c, dt); }
int fsgpu_momentum_update(fsgpu_ctx* c, double dt) { FS_CHECK_CTX(c); return fs_launch_momentum_update(c, dt); }
int fsgpu_apply_velocity_bconds(fsgpu_ctx* c) { FS_CHECK_CTX(c); return fs_launch_bconds(c); }
int fsgpu_calc_divergence(fsgpu_ctx* c) { FS_CHECK_CTX(c); return fs_launch_divergence(c); }
int fsgpu_shift_pressure_to_zero(fsgpu_ctx* c) { FS_CHECK_CTX(c); return fs_launch_shift_pressure(c); }
int fsgpu_correct_velocity(fsgpu_ctx* c, double dt) { FS_CHECK_CTX(c); return fs_launch_correct(c, dt, false, false); }
int fsgpu_project(fsgpu_ctx* c, double dt) { FS_CHECK_CTX(c); return fs_launch_correct(c, dt, true, false); }

static bool flow_mode_ok(int mode) {
  if (mode >= FSGPU_FLOW_OUTFLOW_FIX && mode <= FSGPU_FLOW_FORCE_TOTAL) return true;
  fs_set_error("flow control: unknown mode %d", mode);
  return false;
}
int fsgpu_flow_control(fsgpu_ctx* c, int mode, double total_flow, double* inflow, double* outflow) {
  FS_CHECK_CTX(c);
  if (!flow_mode_ok(mode)) return 1;
  if (fs_launch_flow_control(c, mode, total_flow, c->red_out + 2)) return 1;
  if (inflow || outflow) {
    if (read_red(c, 8, 2)) return 1;
    if (inflow) *inflow = c->h_red[8];
    if (outflow) *outflow = c->h_red[9];
  }
  return 0;
}
int fsgpu_outflow_fix(fsgpu_ctx* c, double* mass_error) {
  double in = 0.0, out = 0.0;
  if (fsgpu_flow_control(c, FSGPU_FLOW_OUTFLOW_FIX, 0.0, mass_error ? &in : nullptr, mass_error ? &out : nullptr))
    return 1;
  if (mass_error) *mass_error = out - in;
  return 0;
}
int fsgpu_set_flow_control(fsgpu_ctx* c, int mode, double total_flow) {
  FS_CHECK_CTX(c);
  if (!flow_mode_ok(mode)) return 1;
  c->flow_mode  = mode;
  c->flow_total = total_flow;
  return 0;
}

static int abs_max_impl(fsgpu_ctx* c, int field, bool include_ghost, double* out) {
  FS_CHECK_CTX(c);
  FS_CHECK_FIELD(field);
  if (fs_launch_absmax(c, field, include_ghost)) return 1;
  if (read_red(c, 0, 1)) return 1;
  *out = c->h_red[0];
  return 0;
}
static int reduce_impl(fsgpu_ctx* c, int field, int include_ghost, int op, double* out) {
  FS_CHECK_CTX(c);
  FS_CHECK_FIELD(field);
  if (!out) { fs_set_error("null result pointer"); return 1; }
  if (fs_launch_reduce_field(c, field, include_ghost != 0, op)) return 1;
  if (read_red(c, 0, 1)) return 1;
  *out = c->h_red[0];
  return 0;
}
int fsgpu_count_nonfinite(fsgpu_ctx* c, int field, int64_t* n_nan, int64_t* n_inf) {
  FS_CHECK_CTX(c);
  FS_CHECK_FIELD(field);
  if (fs_launch_count_nonfinite(c, field)) return 1;
  if (read_red(c, 0, 2)) return 1;
  if (n_nan) *n_nan = (int64_t)c->h_red[0];
  if (n_inf) *n_inf = (int64_t)c->h_red[1];
  return 0;
}
int fsgpu_field_max(fsgpu_ctx* c, int field, int include_ghost, double* out) {
  return reduce_impl(c, field, include_ghost, 3, out);
}
int fsgpu_field_min(fsgpu_ctx* c, int field, int include_ghost, double* out) {
  if (reduce_impl(c, field, include_ghost, 4, out)) return 1;
  *out = -*out;
  return 0;
}
int fsgpu_integrate(fsgpu_ctx* c, int field, int include_ghost, double* out) {
  return reduce_impl(c, field, include_ghost, 1, out);
}
int fsgpu_l1_norm(fsgpu_ctx* c, int field, int include_ghost, double* out) {
  return reduce_impl(c, field, include_ghost, 2, out);
}
int fsgpu_calc_grad_of_centered_points(fsgpu_ctx* c, int f, int dfdx, int dfdy) {
  FS_CHECK_CTX(c);
  const int ids[3] = {f, dfdx, dfdy};
  for (int q = 0; q < 3; ++q) {
    FS_CHECK_FIELD(ids[q]);
    int64_t m, n;
    field_extent(c->g, ids[q], &m, &n);
    if (m != c->g.nx + 2 || n != c->g.ny + 2) {
      fs_set_error("fsgpu_calc_grad_of_centered_points: field %d is not a cell field", ids[q]);
      return 1;
    }
  }
  if (f == dfdx || f == dfdy || dfdx == dfdy) {
    fs_set_error("fsgpu_calc_grad_of_centered_points: the three fields must differ");
    return 1;
  }
  if (c->g.nx < 2 || c->g.ny < 2) { fs_set_error("fsgpu_calc_grad_of_centered_points: grid too small"); return 1; }
  return fs_launch_grad_centered(c, f, dfdx, dfdy);
}
static int sample_impl(fsgpu_ctx* c, int mode, int fa, int fb, const double* x, const double* y, int64_t n,
                       const double* px, const double* py, double* oa, double* ob) {
  FS_CHECK_CTX(c);
  FS_CHECK_FIELD(fa);
  FS_CHECK_FIELD(fb);
  int64_t m2, n2;
  field_extent(c->g, fa, &m2, &n2);
  if (m2 != c->g.nx + 2 || n2 != c->g.ny + 2) { fs_set_error("point sampling: field %d is not a cell field", fa); return 1; }
  if (n < 0 || !x || !y || (n > 0 && (!px || !py || !oa || (mode == 0 && !ob)))) {
    fs_set_error("point sampling: bad arguments");
    return 1;
  }
  if (c->g.nx < 2 || c->g.ny < 2) { fs_set_error("point sampling: grid too small"); return 1; }
  return fs_sample_points(c, mode, fa, fb, x, y, n, px, py, oa, ob);
}
int fsgpu_eval_flow_field_at(fsgpu_ctx* c, const double* x, const double* y, int64_t n, const double* px,
                             const double* py, double* u, double* v) {
  return sample_impl(c, 0, FSGPU_UI, FSGPU_VI, x, y, n, px, py, u, v);
}
int fsgpu_bilinear_interpolate(fsgpu_ctx* c, int field, const double* x, const double* y, int64_t n,
                               const double* px, const double* py, double* out) {
  return sample_impl(c, 1, field, field, x, y, n, px, py, out, nullptr);
}
int fsgpu_abs_max(fsgpu_ctx* c, int field, double* out) { return abs_max_impl(c, field, false, out); }
int fsgpu_abs_max_all(fsgpu_ctx* c, int field, double* out) { return abs_max_impl(c, field, true, out); }

int fsgpu_calc_rho_visc_const(fsgpu_ctx* c) {
  FS_CHECK_CTX(c);
  // calc_rho(fs) fills curr and old densities, calc_visc(fs) the viscosity (FS.hpp:569-589)
  return fs_launch_fill(c, FSGPU_RHO_U, c->rho_gas) || fs_launch_fill(c, FSGPU_RHO_V, c->rho_gas) ||
         fs_launch_fill(c, FSGPU_RHO_U_OLD, c->rho_gas) || fs_launch_fill(c, FSGPU_RHO_V_OLD, c->rho_gas) ||
         fs_launch_fill(c, FSGPU_VISC, c->visc_gas);
}
int fsgpu_calc_rho_visc_vof(fsgpu_ctx* c) { FS_CHECK_CTX(c); return fs_launch_rho_visc_vof(c); }
int fsgpu_apply_neumann_bconds(fsgpu_ctx* c, int field) {
  FS_CHECK_CTX(c);
  FS_CHECK_FIELD(field);
  int64_t m, n;
  field_extent(c->g, field, &m, &n);
  if (fs_launch_neumann(c, c->f[field], m - 2, n - 2)) return 1;
  return fs_exchange_fields(c, &field, 1, FS_RO);
}

// ---- rows either side of the single-phase path (twophase.cu) ------------------------------------
int fsgpu_calc_drhodt(fsgpu_ctx* c) { FS_CHECK_CTX(c); return fs_launch_drhodt(c); }
int fsgpu_update_density(fsgpu_ctx* c, double dt) { FS_CHECK_CTX(c); return fs_launch_update_density(c, dt); }
int fsgpu_density_step(fsgpu_ctx* c, double dt) { FS_CHECK_CTX(c); return fs_launch_density_step(c, dt, false); }
int fsgpu_calc_pressure_jump(fsgpu_ctx* c) { FS_CHECK_CTX(c); return fs_launch_pressure_jump(c); }
int fsgpu_capillary_divergence(fsgpu_ctx* c, double dt) { FS_CHECK_CTX(c); return fs_launch_capillary_divergence(c, dt); }
int fsgpu_correct_velocity_ib_implicit_euler(fsgpu_ctx* c, double dt) {
  FS_CHECK_CTX(c);
  return fs_launch_ib_implicit_euler(c, dt);
}
int fsgpu_update_velocity_ib_semi_analytical(fsgpu_ctx* c, double dt) {
  FS_CHECK_CTX(c);
  return fs_launch_ib_semi_analytical(c, dt);
}
int fsgpu_interpolate_cell_centres(fsgpu_ctx* c) { FS_CHECK_CTX(c); return fs_launch_interpolate(c); }
int fsgpu_calc_conserved_quantities(fsgpu_ctx* c, double* mass, double* momentum_x, double* momentum_y) {
  FS_CHECK_CTX(c);
  if (fs_launch_conserved(c)) return 1;
  if (read_red(c, 4, 3)) return 1;
  if (mass) *mass = c->h_red[4];
  if (momentum_x) *momentum_x = c->h_red[5];
  if (momentum_y) *momentum_y = c->h_red[6];
  return 0;
}
int fsgpu_monitor(fsgpu_ctx* c, double* u_max, double* v_max, double* div_max) {
  FS_CHECK_CTX(c);
  if (fs_launch_monitor(c)) return 1;
  if (read_red(c, 4, 3)) return 1;
  if (u_max) *u_max = c->h_red[4];
  if (v_max) *v_max = c->h_red[5];
  if (div_max) *div_max = c->h_red[6];
  return 0;
}

// ---- solver -----------------------------------------------------------------------------------
int fsgpu_ps_create(fsgpu_ctx* c, const fsgpu_ps_opts* o, fsgpu_ps** out) {
  FS_CHECK_CTX(c);
  if (!o || !out) { fs_set_error("fsgpu_ps_create: null argument"); return 1; }
  if (o->relax != 0 && o->relax != FSGPU_RELAX_JACOBI) {
    fs_set_error("fsgpu_ps_create: smoother %d not available in this build", o->relax);
    return 1;
  }
  fsgpu_ps* s = new (std::nothrow) fsgpu_ps();
  if (!s) { fs_set_error("out of host memory"); return 1; }
  s->c = c;
  s->o = *o;
  if (s->o.relax == 0) s->o.relax = FSGPU_RELAX_JACOBI;
  if (const char* e = getenv("FSGPU_RING_MIN_POINTS")) s->ring_min_points = atoll(e) > 0 ? atoll(e) : 1;
  if (const char* e = getenv("FSGPU_RING")) { if (atoi(e) == 0) s->ring_min_points = 0; }
  {  // registered first: every failure below is undone by the one teardown routine
    std::lock_guard<std::mutex> lk(g_live_mu);
    g_live_solvers.insert(s);
  }
  auto alloc = [&]() {
    FS_CUDA(cudaMalloc(&s->scal, sizeof(PcgScal)));
    FS_CUDA(cudaMemset(s->scal, 0, sizeof(PcgScal)));
    FS_CUDA(cudaMallocHost(&s->h_scal, 4 * sizeof(PcgScal)));
    for (int i = 0; i < 4; ++i) FS_CUDA(cudaEventCreateWithFlags(&s->ev[i], cudaEventDisableTiming));
    return comm_alloc_fence(c);
  };
  if (alloc()) {
    ps_destroy_impl(s);
    return 1;
  }
  *out = s;
  return 0;
}

int fsgpu_ps_destroy(fsgpu_ps* s) {
  // a solver already torn down with its context (fsgpu_destroy) is a no-op here
  {
    std::lock_guard<std::mutex> lk(g_live_mu);
    if (!s || !g_live_solvers.count(s)) return 0;
  }
  return ps_destroy_impl(s);
}

static int ps_destroy_impl(fsgpu_ps* s) {
  {
    std::lock_guard<std::mutex> lk(g_live_mu);
    if (!g_live_solvers.erase(s)) return 0;
  }
  cudaSetDevice(s->c->device);
  cudaStreamSynchronize(s->c->stream);
  if (s->nb_token) comm_unmap_neighbours(s->c, s->nb_token);
  cudaFree(s->slab);
  cudaFree(s->scal);
  if (s->h_scal) cudaFreeHost(s->h_scal);
  for (int i = 0; i < 4; ++i)
    if (s->ev[i]) cudaEventDestroy(s->ev[i]);
  delete s;
  return 0;
}

int fsgpu_ps_set_operator(fsgpu_ps* s) {
  if (!s) { fs_set_error("null solver"); return 1; }
  FS_CHECK_CTX(s->c);
  return ps_setup(s);
}

int fsgpu_ps_set_rhs(fsgpu_ps* s, double dt) {
  if (!s) { fs_set_error("null solver"); return 1; }
  FS_CHECK_CTX(s->c);
  return fs_launch_rhs(s->c, dt, s->o.psd, false);
}

int fsgpu_ps_set_rhs_from_velocity(fsgpu_ps* s, double dt) {
  if (!s) { fs_set_error("null solver"); return 1; }
  FS_CHECK_CTX(s->c);
  return fs_launch_rhs(s->c, dt, s->o.psd, true);
}

int fsgpu_ps_solve(fsgpu_ps* s, double* residual, int* iters) {
  if (!s) { fs_set_error("null solver"); return 1; }
  FS_CHECK_CTX(s->c);
  if (!s->is_setup) { fs_set_error("fsgpu_ps_solve: set_pressure_operator has not been called"); return 1; }
  return ps_solve(s, residual, iters);
}

int fsgpu_ps_num_levels(fsgpu_ps* s, int* n) {
  if (!s) { fs_set_error("null solver"); return 1; }
  *n = s->nl;
  return 0;
}

int fsgpu_ps_num_distributed_levels(fsgpu_ps* s, int* n) {
  if (!s) { fs_set_error("null solver"); return 1; }
  *n = s->ld;
  return 0;
}

// ---- fused sub-iteration / step, scaling/LidDrivenFlow.cpp:121-163 ----------------------------
int fsgpu_substep(fsgpu_ctx* c, fsgpu_ps* s, double dt, int flags, double* residual, int* iters) {
  FS_CHECK_CTX(c);
  if (!s || s->c != c) { fs_set_error("fsgpu_substep: solver does not belong to this context"); return 1; }
  if ((flags & FSGPU_STEP_DENSITY) && fs_launch_density_step(c, dt, true)) return 1;
  {
    ProfScope ps_(PROF_MOMENTUM, c->stream);
    if (fs_launch_momentum_update(c, dt)) return 1;  // mid-time + dmomdt + update
  }
  {
    ProfScope ps_(PROF_BC, c->stream);
    if (fs_launch_bconds(c)) return 1;
  }
  if ((flags & FSGPU_STEP_OUTFLOW_FIX) && fs_launch_flow_control(c, c->flow_mode, c->flow_total, c->red_out + 2))
    return 1;
  if ((flags & FSGPU_STEP_REBUILD_OPERATOR) || !s->is_setup) {
    if (ps_setup(s)) return 1;
  }
  if (flags & FSGPU_STEP_CAPILLARY) {  // the divergence is corrected before it becomes the rhs
    if (fs_launch_divergence(c) || fs_launch_capillary_divergence(c, dt)) return 1;
  }
  {
    ProfScope ps_(PROF_RHS, c->stream);  // divergence + rhs (+ mean removal)
    bool deferred = false;
    // (a one-level hierarchy has no marching kernels to absorb the mean: subtract it here)
    if (fs_launch_rhs(c, dt, s->o.psd, !(flags & FSGPU_STEP_CAPILLARY), s->nl > 1, &deferred)) return 1;
    s->rhs_mean_pending = deferred;
  }
  if (ps_solve(s, residual, iters)) return 1;
  {
    ProfScope ps_(PROF_CORRECT, c->stream);
    if (fs_launch_correct(c, dt, true, true)) return 1;  // shift + p += dp + velocity correction, one pass
  }
  return 0;
}

int fsgpu_step(fsgpu_ctx* c, fsgpu_ps* s, double dt, int num_subiter, int flags, double* residual,
               int* iters) {
  if (fsgpu_save_old_state(c)) return 1;
  int total = 0;
  for (int k = 0; k < num_subiter; ++k) {
    int it = 0;
    if (fsgpu_substep(c, s, dt, flags, residual, &it)) return 1;
    total += it;
  }
  if (iters) *iters = total;
  return 0;
}

}  // extern "C"
