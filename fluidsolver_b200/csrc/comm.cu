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
//   FSGPU_TRANSPORT_LOCAL  all ranks are contexts of ONE process, each driven by its own host
//                          thread (any mix of devices, typically all on one GPU).  Copies are
//                          cudaMemcpyAsync between the contexts, ordered by host barriers.  It is
//                          synchronous and slow by design: it exists so that the decomposed path
//                          (partitioning, halo widths, replicated coarse levels, reductions) can be
//                          tested bit for bit on a single-GPU box.
#include "ctx.h"

#include <dlfcn.h>
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
std::map<std::string, std::shared_ptr<LocalGroup>> g_groups;
std::mutex g_groups_mu;

}  // namespace

struct Comm {
  int transport = 0;
  ncclComm_t nccl = nullptr;
  std::shared_ptr<LocalGroup> grp;
  std::string key;
};

#define LOCAL_BARRIER(cm)                                                    \
  do {                                                                       \
    if ((cm)->grp->barrier()) {                                              \
      fs_set_error("local transport: a rank failed or timed out in a barrier"); \
      return 1;                                                              \
    }                                                                        \
  } while (0)

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
  return 0;
}

void comm_destroy(fsgpu_ctx* c) {
  Comm* cm = c->comm;
  if (!cm) return;
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

int comm_halo(fsgpu_ctx* c, double* const* ptrs, int n, int64_t pitch, int64_t i0, int64_t i1, int h) {
  Comm* cm = c->comm;
  if (!cm || c->nranks <= 1 || n <= 0) return 0;
  if (i1 - i0 < h) {
    fs_set_error("comm_halo: rank %d owns %lld rows, fewer than the halo width %d", c->rank,
                 (long long)(i1 - i0), h);
    return 1;
  }
  const size_t cnt = (size_t)h * (size_t)pitch;
  auto row = [&](double* p, int64_t i) { return p + (i + FS_RO) * pitch; };
  const bool up = c->rank > 0, dn = c->rank + 1 < c->nranks;
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
