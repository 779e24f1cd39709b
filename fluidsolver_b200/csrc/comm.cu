// fluidsolver_b200/csrc/comm.cu -- communication between the slabs of a decomposed grid.
//
// 1-D slabs along i; every halo is a block of whole, contiguous rows.
// Two transports behind one interface:
//
//   FSGPU_TRANSPORT_NCCL   one process per GPU (torchrun / mpirun).  Halo rows travel as grouped
//                          ncclSend/ncclRecv pairs over NVLink, scalars as ncclAllReduce; all of
//                          it is enqueued on the context's stream, so a PCG iteration still needs
//                          no host synchronisation.  libnccl.so.2 is dlopen'ed on first use (in a
//                          torch process this resolves to the copy torch already loaded), so the
//                          library has no link-time dependency on NCCL.
//                          FSGPU_HALO=p2p (opt-in): halo rows and scalar reductions bypass NCCL and
//                          go through peer-mapped mailboxes, one kernel per exchange (see below).
//   FSGPU_TRANSPORT_LOCAL  all ranks are contexts of ONE process, each driven by its own host
//                          thread (any mix of devices, typically all on one GPU).  Copies are
//                          cudaMemcpyAsync between the contexts, ordered by host barriers.  It is
//                          synchronous and slow by design: it exists so that the decomposed path
//                          (partitioning, halo widths, replicated coarse levels, reductions) can be
//                          tested bit for bit on a single-GPU box.
#include "ctx.h"
#include "p2p.cuh"

#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <nccl.h>
#include <string.h>

#include <chrono>
#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

namespace {

// ---- NCCL through dlopen ------------------------------------------------------------------------
struct NcclApi {
  void* h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*)                                              = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int)                       = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t)                                                 = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t)                                                 = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t,
                            cudaStream_t)                                                 = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t)      = nullptr;
  ncclResult_t (*GroupStart)()                                                            = nullptr;
  ncclResult_t (*GroupEnd)()                                                              = nullptr;
  const char* (*GetErrorString)(ncclResult_t)                                             = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

int nccl_load() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.h) return 0;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    fs_set_error("cannot load libnccl.so.2: %s", dlerror());
    return 1;
  }
#define SYM(field, name)                                              \
  do {                                                                \
    *(void**)(&g_nccl.field) = dlsym(h, name);                        \
    if (!g_nccl.field) {                                              \
      fs_set_error("libnccl.so.2 lacks %s", name);                    \
      return 1;                                                       \
    }                                                                 \
  } while (0)
  SYM(GetUniqueId, "ncclGetUniqueId");
  SYM(CommInitRank, "ncclCommInitRank");
  SYM(CommDestroy, "ncclCommDestroy");
  SYM(AllReduce, "ncclAllReduce");
  SYM(AllGather, "ncclAllGather");
  SYM(Send, "ncclSend");
  SYM(Recv, "ncclRecv");
  SYM(GroupStart, "ncclGroupStart");
  SYM(GroupEnd, "ncclGroupEnd");
  SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  g_nccl.h = h;
  return 0;
}

#define FS_NCCL(x)                                                                          \
  do {                                                                                      \
    ncclResult_t r_ = (x);                                                                  \
    if (r_ != ncclSuccess) {                                                                \
      fs_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #x, g_nccl.GetErrorString(r_));   \
      return 1;                                                                             \
    }                                                                                       \
  } while (0)

// ---- the in-process transport -------------------------------------------------------------------
struct LocalGroup {
  int n = 0;
  std::mutex mu;
  std::condition_variable cv;
  int waiting = 0;
  long generation = 0;
  bool broken = false;
  std::vector<fsgpu_ctx*> ctx;
  std::vector<std::vector<const double*>> pub;  // pointers published by each rank for the current op
  std::vector<std::vector<double>> host;        // reduction staging
  std::vector<char*> box;                       // FSGPU_HALO=p2p: every rank's mailbox
  int members = 0;

  // generation barrier with a timeout, so a failed rank cannot hang the others for ever
  int barrier() {
    std::unique_lock<std::mutex> lk(mu);
    if (broken) return 1;
    const long gen = generation;
    if (++waiting == n) {
      waiting = 0;
      ++generation;
      cv.notify_all();
      return 0;
    }
    const bool ok = cv.wait_for(lk, std::chrono::seconds(120), [&] { return generation != gen || broken; });
    // a barrier that completed is a success even if a faster rank has left the group since
    // (fsgpu_destroy marks the group broken): only an incomplete generation is a failure
    if (generation != gen) return 0;
    if (!ok || broken) {
      broken = true;
      cv.notify_all();
      return 1;
    }
    return 0;
  }
};
// ---- direct halo exchange through peer-mapped mailboxes (FSGPU_HALO=p2p) ---------------------------
// Every rank owns one device allocation, its MAILBOX, that the other ranks map (layout: p2p.cuh).
// One exchange = ONE kernel per rank on the context's stream: (A) store my boundary rows straight
// into the neighbours' slots over NVLink, fence, the last CTA to arrive raises the neighbours' flags
// to the epoch number; (B) wait for my own flags to reach the epoch, copy my slots into my halo rows.
// No host round trip, no NCCL launch.  Two slots per side are enough: a rank can only start epoch
// e + 2 after finishing e + 1, which needed the neighbour's flag e + 1, which the neighbour raises
// after its epoch-e kernel (including the unpacking of slot e & 1) has completed in stream order.
// The waits are bounded (a peer that never arrives sets *err instead of hanging the GPU): FSGPU_P2P_TIMEOUT_S
// seconds, default 20 -- long enough for any host-side skew between the ranks of one job (file output, a slower
// setup), short enough that a dead peer surfaces as an error at the next call that returns a scalar.
struct HaloP2PArgs {
  double* ptr[4];
  int n, h;
  int64_t pitch, i0, i1;
  char *mine, *up, *dn;  // mailbox bases: mine, rank - 1's, rank + 1's (null at the ends of the chain)
  size_t cap;            // doubles per slot
  unsigned long long e;  // epoch, counts exchanges from 1
  int* err;
  long long timeout;     // clock64 ticks
};
__device__ __forceinline__ double* p2p_slot(char* base, int side, int par, size_t cap) {
  return reinterpret_cast<double*>(base + 256) + (size_t)(side * 2 + par) * cap;
}
__device__ __forceinline__ volatile unsigned long long* p2p_flag(char* base, int side) {
  return reinterpret_cast<volatile unsigned long long*>(base + 128 * side);
}
constexpr int P2P_CTAS = 8, P2P_THREADS = 512;

__global__ void __launch_bounds__(P2P_THREADS)
k_halo_p2p(const HaloP2PArgs a) {
  if (*reinterpret_cast<volatile int*>(a.mine + 200)) return;  // an earlier exchange timed out: fail fast, the host reports it
  const size_t cnt2 = (size_t)a.h * (size_t)a.pitch / 2;  // double2 per array and direction
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
  const int par = (int)(a.e & 1ull);
  // (A) my first h rows go to the lower halo of rank - 1 (its side 1), my last h rows to the upper
  // halo of rank + 1 (its side 0)
  for (int k = 0; k < a.n; ++k) {
    if (a.up) {
      const double2* src = reinterpret_cast<const double2*>(a.ptr[k] + (a.i0 + FS_RO) * a.pitch);
      double2* dst = reinterpret_cast<double2*>(p2p_slot(a.up, 1, par, a.cap)) + (size_t)k * cnt2;
      for (size_t t = tid; t < cnt2; t += nth) dst[t] = src[t];
    }
    if (a.dn) {
      const double2* src = reinterpret_cast<const double2*>(a.ptr[k] + (a.i1 - a.h + FS_RO) * a.pitch);
      double2* dst = reinterpret_cast<double2*>(p2p_slot(a.dn, 0, par, a.cap)) + (size_t)k * cnt2;
      for (size_t t = tid; t < cnt2; t += nth) dst[t] = src[t];
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* arrive = reinterpret_cast<unsigned int*>(a.mine + 192);
    if (atomicAdd(arrive, 1u) == gridDim.x - 1) {  // every CTA's rows are out
      *arrive = 0;
      __threadfence_system();
      if (a.up) *p2p_flag(a.up, 1) = a.e;
      if (a.dn) *p2p_flag(a.dn, 0) = a.e;
    }
    // (B) wait for the neighbours' rows of this epoch
    const long long t0 = clock64();
    bool ok = true;
    while (ok && ((a.up && *p2p_flag(a.mine, 0) < a.e) || (a.dn && *p2p_flag(a.mine, 1) < a.e))) {
      if (clock64() - t0 > a.timeout) ok = false;
    }
    if (!ok) {
      if (blockIdx.x == 0) {  // mapped pinned memory: the host sees it at its next scalar read
        a.err[1] = (int)a.e;
        a.err[2] = a.up ? (int)*p2p_flag(a.mine, 0) : -1;
        a.err[3] = a.dn ? (int)*p2p_flag(a.mine, 1) : -1;
      }
      *a.err = 1;
      *reinterpret_cast<volatile int*>(a.mine + 200) = 1;
    }
    __threadfence_system();
  }
  __syncthreads();
  for (int k = 0; k < a.n; ++k) {
    if (a.up) {
      const double2* src = reinterpret_cast<const double2*>(p2p_slot(a.mine, 0, par, a.cap)) + (size_t)k * cnt2;
      double2* dst = reinterpret_cast<double2*>(a.ptr[k] + (a.i0 - a.h + FS_RO) * a.pitch);
      for (size_t t = tid; t < cnt2; t += nth) dst[t] = __ldcg(src + t);
    }
    if (a.dn) {
      const double2* src = reinterpret_cast<const double2*>(p2p_slot(a.mine, 1, par, a.cap)) + (size_t)k * cnt2;
      double2* dst = reinterpret_cast<double2*>(a.ptr[k] + (a.i1 + FS_RO) * a.pitch);
      for (size_t t = tid; t < cnt2; t += nth) dst[t] = __ldcg(src + t);
    }
  }
}

// Scalar all-reduce (n <= P2P_RED_MAX doubles) through the same mailboxes: every rank stores its
// values into slot [parity][its rank] of EVERY rank's reduce area and raises that rank's flag for it;
// then it waits for all flags of this epoch and folds the slots in rank order -- the same order on
// every rank, so the result is bitwise identical everywhere (what the PCG control flow relies on).
// The epoch lives in the mailbox (p2p.cuh): the PCG dot products run the same protocol from inside
// their reduction kernels (p2p_allreduce_sum1) and advance the same counter.
struct RedP2PArgs {
  P2PRedDev d;
  double* dev;
  int n, op;
};
__global__ void __launch_bounds__(P2P_MAX_RANKS * P2P_RED_MAX)
k_allreduce_p2p(const RedP2PArgs a) {
  const int t = threadIdx.x;
  char* mine = a.d.peer[a.d.rank];
  if (*p2p_dev_err(mine)) return;
  __shared__ unsigned long long se;
  if (t == 0) se = ++*p2p_red_epoch(mine);
  __syncthreads();
  const unsigned long long e = se;
  const int par = (int)(e & 1ull);
  if (t < a.d.nranks * a.n) {
    const int pr = t / a.n, q = t - pr * a.n;
    p2p_red_vals(a.d.peer[pr], a.d.red_off, par, a.d.rank)[q] = a.dev[q];
  }
  __threadfence_system();
  __syncthreads();
  if (t < a.d.nranks) {
    __threadfence_system();
    *p2p_red_flag(a.d.peer[t], a.d.red_off, a.d.rank) = e;
    const long long t0 = clock64();
    bool ok = true;
    while (ok && *p2p_red_flag(mine, a.d.red_off, t) < e)
      if (clock64() - t0 > a.d.timeout) ok = false;
    if (!ok) {
      *a.d.err = 1;
      *p2p_dev_err(mine) = 1;
    }
    __threadfence_system();
  }
  __syncthreads();
  if (t < a.n) {
    double acc = __ldcg(p2p_red_vals(mine, a.d.red_off, par, 0) + t);
    for (int r = 1; r < a.d.nranks; ++r) {
      const double v = __ldcg(p2p_red_vals(mine, a.d.red_off, par, r) + t);
      acc = a.op == COMM_MAX ? (acc > v ? acc : v) : acc + v;
    }
    a.dev[t] = acc;
  }
}

struct P2PBox {
  bool on = false;
  char* base = nullptr;               // my mailbox
  size_t cap = 0;                     // doubles per slot
  unsigned long long epoch = 0;
  char* nb[2] = {nullptr, nullptr};   // rank - 1's and rank + 1's mailbox, mapped here
  std::vector<char*> peer;            // every rank's mailbox (peer[rank] == base)
  std::vector<char> ipc;              // mapped with cudaIpcOpenMemHandle (to be closed)
  size_t red_off = 0;                 // byte offset of the scalar all-reduce area
  size_t push_off = 0;                // byte offset of the fused-push flags: [level][kind b | x][side] u64
  unsigned int* push_ticket = nullptr;  // arrival counter of the producer kernels
  bool push = false;                  // fused push allowed (FSGPU_PUSH, read when the context is created)
  unsigned long long push_epoch = 0;  // one per V-cycle of any solver of this context
  int *h_err = nullptr, *d_err = nullptr;  // mapped pinned flag the kernel raises on a timeout
  long long timeout = 0;              // clock64 ticks a device-side wait may last
  P2PRedDev red{};                    // what the reduction kernels need (host copy)
  P2PRedDev* d_red = nullptr;         // ... and its device copy (PcgScal::p2p)
  // overlap: exchanges enqueued with comm_halo_async run on their own stream, fenced by two events
  cudaStream_t xstream = nullptr;
  cudaEvent_t ev_ready = nullptr, ev_done = nullptr;
  bool pending = false;
  bool overlap = false;               // FSGPU_OVERLAP=1 (read when the context is created)
  int64_t overlap_min_points = 262144;  // FSGPU_OVERLAP_MIN_POINTS: smaller slabs of a level exchange in stream order
};
std::map<std::string, std::shared_ptr<LocalGroup>> g_groups;
std::mutex g_groups_mu;

}  // namespace

struct Comm {
  int transport = 0;
  ncclComm_t nccl = nullptr;
  std::shared_ptr<LocalGroup> grp;
  std::string key;
  P2PBox p2p;
};

#define LOCAL_BARRIER(cm)                                                    \
  do {                                                                       \
    if ((cm)->grp->barrier()) {                                              \
      fs_set_error("local transport: a rank failed or timed out in a barrier"); \
      return 1;                                                              \
    }                                                                        \
  } while (0)

// Mailbox setup; collective.  Returns 0 also when the direct path cannot be used (it then stays
// off on EVERY rank: the decision is all-reduced) -- the NCCL / staged exchange keeps working.
static int p2p_setup(fsgpu_ctx* c, Comm* cm) {
  P2PBox& b = cm->p2p;
  // One process per GPU: the mailboxes (with the fused push) are the default transport -- 99.2 ms per 16384^2 step on
  // 8 GPUs against 102.6 with NCCL send/recv (profiles/r02_multigpu.txt); FSGPU_HALO=nccl keeps everything on NCCL.
  // The in-process test transport takes them only on request (FSGPU_HALO=p2p).
  const char* env = getenv("FSGPU_HALO");
  const bool want = env ? strcmp(env, "p2p") == 0 : cm->transport == FSGPU_TRANSPORT_NCCL;
  if (!want) return 0;
  b.cap = (size_t)((8 * c->g.pitch + 15) / 16 * 16);  // the widest exchange: 4 arrays x FS_RO rows
  b.red_off = 256 + 4 * b.cap * sizeof(double);
  b.push_off = b.red_off + P2P_MAX_RANKS * sizeof(unsigned long long) +
               2 * (size_t)P2P_MAX_RANKS * P2P_RED_MAX * sizeof(double);
  const size_t bytes = b.push_off + (size_t)P2P_PUSH_LEVELS * 4 * sizeof(unsigned long long);
  int ok = c->nranks <= P2P_MAX_RANKS ? 1 : 0;
  b.peer.assign((size_t)c->nranks, nullptr);
  b.ipc.assign((size_t)c->nranks, 0);
  if (cudaMalloc(&b.base, bytes) != cudaSuccess) { b.base = nullptr; ok = 0; }
  if (ok && cudaMemsetAsync(b.base, 0, bytes, c->stream) != cudaSuccess) ok = 0;
  if (ok && cudaHostAlloc(&b.h_err, 4 * sizeof(int), cudaHostAllocMapped) != cudaSuccess) { b.h_err = nullptr; ok = 0; }
  if (ok) {
    for (int q = 0; q < 4; ++q) b.h_err[q] = 0;
    if (cudaHostGetDevicePointer(&b.d_err, b.h_err, 0) != cudaSuccess) ok = 0;
  }
  const int peer[2] = {c->rank - 1, c->rank + 1 < c->nranks ? c->rank + 1 : -1};
  if (cm->transport == FSGPU_TRANSPORT_NCCL) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof(mine));
    if (ok && cudaIpcGetMemHandle(&mine, b.base) != cudaSuccess) ok = 0;
    char* d_all = nullptr;  // every rank takes part in the collectives below, whatever its `ok`
    FS_CUDA(cudaMalloc(&d_all, 64 * (size_t)c->nranks + sizeof(int)));
    FS_CUDA(cudaMemcpyAsync(d_all + 64 * c->rank, &mine, 64, cudaMemcpyHostToDevice, c->stream));
    FS_NCCL(g_nccl.AllGather(d_all + 64 * c->rank, d_all, 64, ncclChar, cm->nccl, c->stream));
    std::vector<cudaIpcMemHandle_t> all((size_t)c->nranks);
    FS_CUDA(cudaMemcpyAsync(all.data(), d_all, 64 * (size_t)c->nranks, cudaMemcpyDeviceToHost, c->stream));
    FS_CUDA(cudaStreamSynchronize(c->stream));
    for (int r = 0; r < c->nranks && ok; ++r) {  // every rank: the scalar all-reduce is all-to-all
      if (r == c->rank) { b.peer[(size_t)r] = b.base; continue; }
      void* q = nullptr;
      if (cudaIpcOpenMemHandle(&q, all[(size_t)r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ok = 0;
      else { b.peer[(size_t)r] = (char*)q; b.ipc[(size_t)r] = 1; }
    }
    for (int sd = 0; sd < 2; ++sd)
      if (peer[sd] >= 0) b.nb[sd] = b.peer[(size_t)peer[sd]];
    cudaGetLastError();  // a failed attempt must not poison later launch checks
    int* d_ok = reinterpret_cast<int*>(d_all + 64 * (size_t)c->nranks);
    FS_CUDA(cudaMemcpyAsync(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice, c->stream));
    FS_NCCL(g_nccl.AllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, cm->nccl, c->stream));
    FS_CUDA(cudaMemcpyAsync(&ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    FS_CUDA(cudaStreamSynchronize(c->stream));
    cudaFree(d_all);
  } else {
    LocalGroup& g = *cm->grp;
    {
      std::lock_guard<std::mutex> lk(g.mu);
      if (g.box.size() != (size_t)c->nranks) g.box.assign((size_t)c->nranks, nullptr);
      g.box[(size_t)c->rank] = ok ? b.base : nullptr;
    }
    FS_CUDA(cudaStreamSynchronize(c->stream));
    LOCAL_BARRIER(cm);
    for (int r = 0; r < c->nranks; ++r)
      if (!g.box[(size_t)r]) ok = 0;  // same verdict on every rank
    for (int r = 0; r < c->nranks; ++r) b.peer[(size_t)r] = g.box[(size_t)r];
    for (int sd = 0; sd < 2; ++sd)
      if (peer[sd] >= 0) b.nb[sd] = g.box[(size_t)peer[sd]];
    LOCAL_BARRIER(cm);
  }
  if (ok) {
    double secs = 20.0;
    if (const char* e = getenv("FSGPU_P2P_TIMEOUT_S")) secs = atof(e) > 0.0 ? atof(e) : secs;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, c->device);
    b.timeout = (long long)(secs * 1e3 * (khz > 0 ? khz : 2000000));
    for (int r = 0; r < P2P_MAX_RANKS; ++r) b.red.peer[r] = r < c->nranks ? b.peer[(size_t)r] : nullptr;
    b.red.rank = c->rank; b.red.nranks = c->nranks; b.red.red_off = b.red_off;
    b.red.err = b.d_err; b.red.timeout = b.timeout;
    FS_CUDA(cudaMalloc(&b.d_red, sizeof(P2PRedDev)));
    FS_CUDA(cudaMemcpy(b.d_red, &b.red, sizeof(P2PRedDev), cudaMemcpyHostToDevice));
    // boundary-first overlap is opt-in (FSGPU_OVERLAP=1): at 16384^2 on 8 GPUs the two extra launches and the
    // cross-stream events per exchange cost more than the ~10 us they hide (115.0 vs 111.6 ms per step,
    // profiles/r02_multigpu.txt)
    b.overlap = getenv("FSGPU_OVERLAP") != nullptr && atoi(getenv("FSGPU_OVERLAP")) != 0;
    // One process per GPU: on unless FSGPU_PUSH=0.  In-process ranks share ONE device: a consumer kernel whose
    // CTAs fill the GPU while they wait would keep the neighbour rank's producer from ever being scheduled, so
    // there it is opt-in (FSGPU_PUSH=1; the tests' grids are a handful of CTAs).
    {
      const char* e = getenv("FSGPU_PUSH");
      b.push = cm->transport == FSGPU_TRANSPORT_NCCL ? !(e && atoi(e) == 0) : (e && atoi(e) != 0);
    }
    FS_CUDA(cudaMalloc(&b.push_ticket, sizeof(unsigned int)));
    FS_CUDA(cudaMemset(b.push_ticket, 0, sizeof(unsigned int)));
    if (const char* e = getenv("FSGPU_OVERLAP_MIN_POINTS")) b.overlap_min_points = atoll(e);
    FS_CUDA(cudaStreamCreateWithFlags(&b.xstream, cudaStreamNonBlocking));
    FS_CUDA(cudaEventCreateWithFlags(&b.ev_ready, cudaEventDisableTiming));
    FS_CUDA(cudaEventCreateWithFlags(&b.ev_done, cudaEventDisableTiming));
  }
  b.on = ok != 0;
  if (!b.on && c->rank == 0)
    fprintf(stderr, "[fsgpu] FSGPU_HALO=p2p: peer mailboxes unavailable, halo exchange stays on the default transport\n");
  return 0;
}

static void p2p_teardown(Comm* cm) {
  P2PBox& b = cm->p2p;
  for (size_t r = 0; r < b.peer.size(); ++r)
    if (b.ipc[r] && b.peer[r]) cudaIpcCloseMemHandle(b.peer[r]);
  if (b.xstream) { cudaStreamSynchronize(b.xstream); cudaStreamDestroy(b.xstream); }
  if (b.ev_ready) cudaEventDestroy(b.ev_ready);
  if (b.ev_done) cudaEventDestroy(b.ev_done);
  if (b.d_red) cudaFree(b.d_red);
  if (b.push_ticket) cudaFree(b.push_ticket);
  if (b.base) cudaFree(b.base);
  if (b.h_err) cudaFreeHost(b.h_err);
  b = P2PBox();
}

int comm_create(fsgpu_ctx* c, int rank, int nranks, const void* id128, int transport) {
  c->rank = rank;
  c->nranks = nranks;
  if (nranks <= 1) return 0;
  if (!id128) { fs_set_error("comm_create: null communicator id"); return 1; }
  Comm* cm = new Comm();
  cm->transport = transport;
  if (transport == FSGPU_TRANSPORT_NCCL) {
    if (nccl_load()) { delete cm; return 1; }
    ncclUniqueId id;
    static_assert(sizeof(ncclUniqueId) == FSGPU_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    memcpy(&id, id128, sizeof(id));
    ncclResult_t r = g_nccl.CommInitRank(&cm->nccl, nranks, id, rank);
    if (r != ncclSuccess) {
      fs_set_error("ncclCommInitRank(rank %d of %d) -> %s", rank, nranks, g_nccl.GetErrorString(r));
      delete cm;
      return 1;
    }
  } else if (transport == FSGPU_TRANSPORT_LOCAL) {
    cm->key.assign((const char*)id128, FSGPU_COMM_ID_BYTES);
    std::lock_guard<std::mutex> lk(g_groups_mu);
    auto& g = g_groups[cm->key];
    if (!g) {
      g = std::make_shared<LocalGroup>();
      g->n = nranks;
      g->ctx.assign(nranks, nullptr);
      g->pub.resize(nranks);
      g->host.resize(nranks);
    }
    if (g->n != nranks || rank < 0 || rank >= nranks || g->ctx[rank]) {
      fs_set_error("local transport: inconsistent group (rank %d of %d)", rank, nranks);
      delete cm;
      return 1;
    }
    g->ctx[rank] = c;
    g->members++;
    cm->grp = g;
  } else {
    fs_set_error("comm_create: unknown transport %d", transport);
    delete cm;
    return 1;
  }
  c->comm = cm;
  if (p2p_setup(c, cm)) {
    comm_destroy(c);
    return 1;
  }
  return 0;
}

// 1 when halo rows and scalar reductions go through the peer-mapped mailboxes (FSGPU_HALO=p2p took)
int comm_is_direct(fsgpu_ctx* c) { return c->comm && c->comm->p2p.on ? 1 : 0; }

// non-zero once a direct halo exchange has timed out waiting for a neighbour
int comm_failed(fsgpu_ctx* c) {
  Comm* cm = c->comm;
  if (!cm || !cm->p2p.on || !*(volatile int*)cm->p2p.h_err) return 0;
  const volatile int* e = cm->p2p.h_err;
  if (e[1])
    fs_set_error("mailbox transport: rank %d timed out in halo exchange %d (flags seen: from rank-1 %d, from rank+1 %d)",
                 c->rank, e[1], e[2], e[3]);
  else
    fs_set_error("mailbox transport: rank %d timed out waiting for a peer in a scalar all-reduce", c->rank);
  return 1;
}

void comm_destroy(fsgpu_ctx* c) {
  Comm* cm = c->comm;
  if (!cm) return;
  cudaStreamSynchronize(c->stream);
  p2p_teardown(cm);
  if (cm->transport == FSGPU_TRANSPORT_NCCL) {
    if (cm->nccl) g_nccl.CommDestroy(cm->nccl);
  } else {
    std::lock_guard<std::mutex> lk(g_groups_mu);
    {
      std::lock_guard<std::mutex> lk2(cm->grp->mu);
      cm->grp->broken = true;  // a group is dead as soon as one member leaves
      cm->grp->cv.notify_all();
    }
    if (--cm->grp->members == 0) g_groups.erase(cm->key);
  }
  delete cm;
  c->comm = nullptr;
}

// ---- fused push: neighbours' arrays mapped here ---------------------------------------------------
// collective: every rank contributes `bytes` bytes of host memory and receives all ranks' contributions
int comm_allgather_host(fsgpu_ctx* c, const void* mine, size_t bytes, void* all) {
  Comm* cm = c->comm;
  if (!cm || c->nranks <= 1) { memcpy(all, mine, bytes); return 0; }
  if (cm->transport == FSGPU_TRANSPORT_NCCL) {
    char* d = nullptr;
    FS_CUDA(cudaMalloc(&d, bytes * (size_t)c->nranks));
    FS_CUDA(cudaMemcpyAsync(d + bytes * c->rank, mine, bytes, cudaMemcpyHostToDevice, c->stream));
    FS_NCCL(g_nccl.AllGather(d + bytes * c->rank, d, bytes, ncclChar, cm->nccl, c->stream));
    FS_CUDA(cudaMemcpyAsync(all, d, bytes * (size_t)c->nranks, cudaMemcpyDeviceToHost, c->stream));
    FS_CUDA(cudaStreamSynchronize(c->stream));
    cudaFree(d);
    return 0;
  }
  LocalGroup& g = *cm->grp;
  g.host[c->rank].assign((bytes + 7) / 8, 0.0);
  memcpy(g.host[c->rank].data(), mine, bytes);
  FS_CUDA(cudaStreamSynchronize(c->stream));
  LOCAL_BARRIER(cm);
  for (int r = 0; r < c->nranks; ++r) memcpy((char*)all + bytes * r, g.host[r].data(), bytes);
  LOCAL_BARRIER(cm);
  return 0;
}

// collective: `base` is this rank's instance of an allocation every rank has (the solver's slab).  Returns the
// neighbours' instances as mapped in THIS rank's address space (cudaIpc between processes, the plain pointers
// inside one process; null at the ends of the chain).  *token: what comm_unmap_neighbours needs.
struct NeighbourMap { void* ptr[2]; bool ipc[2]; };
int comm_map_neighbours(fsgpu_ctx* c, void* base, void** up, void** dn, void** token) {
  *up = *dn = nullptr;
  *token = nullptr;
  Comm* cm = c->comm;
  if (!cm || c->nranks <= 1 || !cm->p2p.on) return 0;
  NeighbourMap* m = new NeighbourMap();
  m->ptr[0] = m->ptr[1] = nullptr;
  m->ipc[0] = m->ipc[1] = false;
  const int peer[2] = {c->rank - 1, c->rank + 1 < c->nranks ? c->rank + 1 : -1};
  if (cm->transport == FSGPU_TRANSPORT_NCCL) {
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof(mine));
    int ok = cudaIpcGetMemHandle(&mine, base) == cudaSuccess ? 1 : 0;
    std::vector<char> all((size_t)c->nranks * (sizeof(mine) + sizeof(int)));
    char rec[sizeof(mine) + sizeof(int)];
    memcpy(rec, &mine, sizeof(mine));
    memcpy(rec + sizeof(mine), &ok, sizeof(int));
    if (comm_allgather_host(c, rec, sizeof(rec), all.data())) { delete m; return 1; }
    for (int r = 0; r < c->nranks; ++r) {
      int okr;
      memcpy(&okr, all.data() + (size_t)r * sizeof(rec) + sizeof(mine), sizeof(int));
      ok = ok && okr;
    }
    for (int sd = 0; sd < 2 && ok; ++sd) {
      if (peer[sd] < 0) continue;
      cudaIpcMemHandle_t h;
      memcpy(&h, all.data() + (size_t)peer[sd] * sizeof(rec), sizeof(h));
      void* q = nullptr;
      if (cudaIpcOpenMemHandle(&q, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ok = 0;
      else { m->ptr[sd] = q; m->ipc[sd] = true; }
    }
    cudaGetLastError();
    // all ranks or none (a rank that cannot map must not be pushed to either)
    int allok = ok;
    std::vector<int> oks((size_t)c->nranks);
    if (comm_allgather_host(c, &ok, sizeof(int), oks.data())) { delete m; return 1; }
    for (int r = 0; r < c->nranks; ++r) allok = allok && oks[(size_t)r];
    if (!allok) {
      for (int sd = 0; sd < 2; ++sd)
        if (m->ipc[sd] && m->ptr[sd]) cudaIpcCloseMemHandle(m->ptr[sd]);
      delete m;
      return 0;
    }
  } else {
    std::vector<void*> all((size_t)c->nranks);
    if (comm_allgather_host(c, &base, sizeof(void*), all.data())) { delete m; return 1; }
    for (int sd = 0; sd < 2; ++sd)
      if (peer[sd] >= 0) m->ptr[sd] = all[(size_t)peer[sd]];
  }
  *up = m->ptr[0];
  *dn = m->ptr[1];
  *token = m;
  return 0;
}
void comm_unmap_neighbours(fsgpu_ctx* c, void* token) {
  (void)c;
  NeighbourMap* m = static_cast<NeighbourMap*>(token);
  if (!m) return;
  for (int sd = 0; sd < 2; ++sd)
    if (m->ipc[sd] && m->ptr[sd]) cudaIpcCloseMemHandle(m->ptr[sd]);
  delete m;
}
// 1: producer kernels may push halo rows into mapped neighbour arrays (mailbox transport on, FSGPU_PUSH)
int comm_push_enabled(fsgpu_ctx* c) { return c->comm && c->comm->p2p.on && c->comm->p2p.push ? 1 : 0; }
unsigned long long comm_push_next_epoch(fsgpu_ctx* c) { return ++c->comm->p2p.push_epoch; }
// flags of the exchange (level, kind): what a producer raises in its neighbours' mailboxes, what a consumer waits for
void comm_push_flags(fsgpu_ctx* c, int level, int kind, HaloPush* p, HaloWait* w) {
  P2PBox& b = c->comm->p2p;
  auto flag = [&](char* base, int side) {
    return reinterpret_cast<unsigned long long*>(base + b.push_off) + ((size_t)level * 2 + kind) * 2 + side;
  };
  // rows I send up land in rank - 1's LOWER halo: its flag for side 1; rows I send down: rank + 1's side 0
  if (p) {
    p->raise[0] = b.nb[0] ? flag(b.nb[0], 1) : nullptr;
    p->raise[1] = b.nb[1] ? flag(b.nb[1], 0) : nullptr;
    p->ticket   = b.push_ticket;
  }
  if (w) {
    w->flag[0] = b.nb[0] ? flag(b.base, 0) : nullptr;
    w->flag[1] = b.nb[1] ? flag(b.base, 1) : nullptr;
    w->err = b.d_err;
    w->mine = b.base;
    w->timeout = b.timeout;
  }
}

// true when this exchange fits the mailbox slots (the same decision on every rank)
static bool halo_fits_mailbox(const Comm* cm, int n, int64_t pitch, int h) {
  return cm->p2p.on && n <= 4 && (size_t)n * (size_t)h * (size_t)pitch <= cm->p2p.cap;
}
static int halo_mailbox_launch(fsgpu_ctx* c, double* const* ptrs, int n, int64_t pitch, int64_t i0, int64_t i1,
                               int h, cudaStream_t st) {
  Comm* cm = c->comm;
  HaloP2PArgs a;
  for (int k = 0; k < 4; ++k) a.ptr[k] = k < n ? ptrs[k] : nullptr;
  a.n = n; a.h = h; a.pitch = pitch; a.i0 = i0; a.i1 = i1;
  a.mine = cm->p2p.base;
  a.up = c->rank > 0 ? cm->p2p.nb[0] : nullptr;
  a.dn = c->rank + 1 < c->nranks ? cm->p2p.nb[1] : nullptr;
  a.cap = cm->p2p.cap;
  a.e = ++cm->p2p.epoch;
  a.err = cm->p2p.d_err;
  a.timeout = cm->p2p.timeout;
  // one CTA per 32 KB sent to a neighbour (coarse levels: a single CTA, whose arrival counter is trivially complete)
  const size_t bytes = (size_t)n * (size_t)h * (size_t)pitch * sizeof(double);
  const unsigned ctas = (unsigned)(bytes / 32768 < 1 ? 1 : bytes / 32768 > P2P_CTAS ? P2P_CTAS : bytes / 32768);
  k_halo_p2p<<<ctas, P2P_THREADS, 0, st>>>(a);
  FS_LAUNCHED();
  return 0;
}

// Overlap: the caller has just enqueued the kernel(s) that wrote the h boundary rows
// at both ends of [i0, i1) and will enqueue the interior part next.  The exchange then runs on the
// transport's own stream, behind an event on the compute stream, while the interior kernel computes;
// comm_halo_wait makes the compute stream wait for it before the first consumer of the halo rows.
// Returns 0 with *async_out = false when the exchange is not one the mailbox can carry: the caller then
// uses the blocking comm_halo after its interior kernel.
int comm_halo_async(fsgpu_ctx* c, double* const* ptrs, int n, int64_t pitch, int64_t i0, int64_t i1, int h,
                    bool* async_out) {
  Comm* cm = c->comm;
  *async_out = false;
  if (!cm || c->nranks <= 1 || n <= 0) return 0;
  if (!halo_fits_mailbox(cm, n, pitch, h) || i1 - i0 < 2 * h) return 0;
  if (comm_halo_wait(c)) return 1;  // one exchange in flight at a time: the slot protocol is per epoch
  P2PBox& b = cm->p2p;
  FS_CUDA(cudaEventRecord(b.ev_ready, c->stream));
  FS_CUDA(cudaStreamWaitEvent(b.xstream, b.ev_ready, 0));
  if (halo_mailbox_launch(c, ptrs, n, pitch, i0, i1, h, b.xstream)) return 1;
  FS_CUDA(cudaEventRecord(b.ev_done, b.xstream));
  b.pending = true;
  *async_out = true;
  return 0;
}
int comm_halo_wait(fsgpu_ctx* c) {
  Comm* cm = c->comm;
  if (!cm || !cm->p2p.pending) return 0;
  FS_CUDA(cudaStreamWaitEvent(c->stream, cm->p2p.ev_done, 0));
  cm->p2p.pending = false;
  return 0;
}
// In-process ranks on ONE device (the LOCAL test transport) with the mailboxes on: a call that synchronises
// the device implicitly (cudaFree, cudaMalloc, pinned allocations) on one host thread would wait for a peer
// rank's exchange kernel that is itself waiting for a kernel this thread has not enqueued yet.  Every
// allocation site reachable after comm_create therefore ends in this fence: drain my stream, meet the
// other ranks.  Nothing of the kind exists with one process per GPU (a device only waits for its own
// work, and that work only for kernels the peers have already enqueued), where this is a no-op.
int comm_alloc_fence(fsgpu_ctx* c) {
  Comm* cm = c->comm;
  if (!cm || cm->transport != FSGPU_TRANSPORT_LOCAL || !cm->p2p.on) return 0;
  FS_CUDA(cudaStreamSynchronize(c->stream));
  LOCAL_BARRIER(cm);
  return 0;
}

// > 0: slabs of a level with at least that many points split their producer kernels (boundary rows first)
// and exchange on the transport's stream; 0: no overlap (no mailbox, or FSGPU_OVERLAP not set)
int64_t comm_overlap_min_points(fsgpu_ctx* c) {
  if (!c->comm || !c->comm->p2p.on || !c->comm->p2p.overlap) return 0;
  return c->comm->p2p.overlap_min_points > 0 ? c->comm->p2p.overlap_min_points : 1;
}
const P2PRedDev* comm_red_dev(fsgpu_ctx* c) { return c->comm && c->comm->p2p.on ? c->comm->p2p.d_red : nullptr; }

int comm_halo(fsgpu_ctx* c, double* const* ptrs, int n, int64_t pitch, int64_t i0, int64_t i1, int h) {
  Comm* cm = c->comm;
  if (!cm || c->nranks <= 1 || n <= 0) return 0;
  if (i1 - i0 < h) {
    fs_set_error("comm_halo: rank %d owns %lld rows, fewer than the halo width %d", c->rank,
                 (long long)(i1 - i0), h);
    return 1;
  }
  if (comm_halo_wait(c)) return 1;
  const size_t cnt = (size_t)h * (size_t)pitch;
  auto row = [&](double* p, int64_t i) { return p + (i + FS_RO) * pitch; };
  const bool up = c->rank > 0, dn = c->rank + 1 < c->nranks;
  if (halo_fits_mailbox(cm, n, pitch, h)) return halo_mailbox_launch(c, ptrs, n, pitch, i0, i1, h, c->stream);
  if (cm->transport == FSGPU_TRANSPORT_NCCL) {
    FS_NCCL(g_nccl.GroupStart());
    for (int k = 0; k < n; ++k) {
      if (up) {
        FS_NCCL(g_nccl.Send(row(ptrs[k], i0), cnt, ncclDouble, c->rank - 1, cm->nccl, c->stream));
        FS_NCCL(g_nccl.Recv(row(ptrs[k], i0 - h), cnt, ncclDouble, c->rank - 1, cm->nccl, c->stream));
      }
      if (dn) {
        FS_NCCL(g_nccl.Send(row(ptrs[k], i1 - h), cnt, ncclDouble, c->rank + 1, cm->nccl, c->stream));
        FS_NCCL(g_nccl.Recv(row(ptrs[k], i1), cnt, ncclDouble, c->rank + 1, cm->nccl, c->stream));
      }
    }
    FS_NCCL(g_nccl.GroupEnd());
    return 0;
  }
  LocalGroup& g = *cm->grp;
  g.pub[c->rank].assign(ptrs, ptrs + n);
  FS_CUDA(cudaStreamSynchronize(c->stream));
  LOCAL_BARRIER(cm);
  for (int k = 0; k < n; ++k) {
    // the neighbour's virtual base addresses the same global rows
    if (up) {
      const double* src = g.pub[c->rank - 1][k];
      FS_CUDA(cudaMemcpyAsync(row(ptrs[k], i0 - h), src + (i0 - h + FS_RO) * pitch, cnt * sizeof(double),
                              cudaMemcpyDefault, c->stream));
    }
    if (dn) {
      const double* src = g.pub[c->rank + 1][k];
      FS_CUDA(cudaMemcpyAsync(row(ptrs[k], i1), src + (i1 + FS_RO) * pitch, cnt * sizeof(double),
                              cudaMemcpyDefault, c->stream));
    }
  }
  FS_CUDA(cudaStreamSynchronize(c->stream));
  LOCAL_BARRIER(cm);
  return 0;
}

int comm_allreduce(fsgpu_ctx* c, double* dev, int64_t n, int op) {
  Comm* cm = c->comm;
  if (!cm || c->nranks <= 1 || n <= 0) return 0;
  if (comm_halo_wait(c)) return 1;
  if (cm->p2p.on && n <= P2P_RED_MAX) {  // scalars: one kernel, rank-ordered fold
    RedP2PArgs a;
    a.d = cm->p2p.red;
    a.dev = dev; a.n = (int)n; a.op = op;
    k_allreduce_p2p<<<1, P2P_MAX_RANKS * P2P_RED_MAX, 0, c->stream>>>(a);
    FS_LAUNCHED();
    return 0;
  }
  if (cm->transport == FSGPU_TRANSPORT_NCCL) {
    FS_NCCL(g_nccl.AllReduce(dev, dev, (size_t)n, ncclDouble, op == COMM_MAX ? ncclMax : ncclSum,
                             cm->nccl, c->stream));
    return 0;
  }
  LocalGroup& g = *cm->grp;
  std::vector<double>& mine = g.host[c->rank];
  mine.resize((size_t)n);
  FS_CUDA(cudaMemcpyAsync(mine.data(), dev, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  FS_CUDA(cudaStreamSynchronize(c->stream));
  LOCAL_BARRIER(cm);
  std::vector<double> acc(g.host[0]);
  for (int r = 1; r < c->nranks; ++r) {
    const std::vector<double>& o = g.host[r];
    if ((int64_t)o.size() != n) { fs_set_error("local allreduce: size mismatch between ranks"); return 1; }
    for (int64_t k = 0; k < n; ++k) acc[k] = op == COMM_MAX ? (acc[k] > o[k] ? acc[k] : o[k]) : acc[k] + o[k];
  }
  FS_CUDA(cudaMemcpyAsync(dev, acc.data(), (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  FS_CUDA(cudaStreamSynchronize(c->stream));
  LOCAL_BARRIER(cm);  // nobody may overwrite its staging buffer before everyone has summed
  return 0;
}

int comm_sendrecv(fsgpu_ctx* c, int peer, const double* const* send, double* const* recv, int narr,
                  int64_t n) {
  Comm* cm = c->comm;
  if (!cm || c->nranks <= 1) return 0;
  if (cm->transport == FSGPU_TRANSPORT_NCCL) {
    if (peer < 0) return 0;
    FS_NCCL(g_nccl.GroupStart());
    for (int k = 0; k < narr; ++k) {
      FS_NCCL(g_nccl.Send(send[k], (size_t)n, ncclDouble, peer, cm->nccl, c->stream));
      FS_NCCL(g_nccl.Recv(recv[k], (size_t)n, ncclDouble, peer, cm->nccl, c->stream));
    }
    FS_NCCL(g_nccl.GroupEnd());
    return 0;
  }
  LocalGroup& g = *cm->grp;
  if (peer >= 0) g.pub[c->rank].assign(send, send + narr);
  FS_CUDA(cudaStreamSynchronize(c->stream));
  LOCAL_BARRIER(cm);
  if (peer >= 0) {
    for (int k = 0; k < narr; ++k)
      FS_CUDA(cudaMemcpyAsync(recv[k], g.pub[peer][k], (size_t)n * sizeof(double), cudaMemcpyDefault, c->stream));
    FS_CUDA(cudaStreamSynchronize(c->stream));
  }
  LOCAL_BARRIER(cm);
  return 0;
}

int fs_exchange_fields(fsgpu_ctx* c, const int* fields, int n, int h) {
  if (c->nranks <= 1) return 0;
  double* ptrs[FSGPU_NUM_FIELDS];
  for (int k = 0; k < n; ++k) ptrs[k] = c->f[fields[k]];
  return comm_halo(c, ptrs, n, c->g.pitch, c->g.a_lo, c->g.a_hi, h);
}

extern "C" int fsgpu_comm_id(void* id128) {
  if (!id128) { fs_set_error("fsgpu_comm_id: null buffer"); return 1; }
  if (nccl_load()) return 1;
  ncclUniqueId id;
  FS_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return 0;
}
