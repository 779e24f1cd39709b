// fluidsolver_b200/csrc/output.cu -- file output straight from the device buffers.
//
//   fsgpu_npy_write_host   detail::write_npy_header + payload          src/IO.hpp:25-95
//   fsgpu_to_npy           to_npy(filename, Field2D)                   src/IO.hpp:204-228
//   fsgpu_to_npy_state     to_npy(directory, fs)                       src/IO.hpp:231-269
//   fsgpu_write_vtk        VTKWriter::write (legacy VTK, BINARY)       src/VTKWriter.hpp:39-152
//
// The files are byte-identical to the reference's.  Where the reference walks a field element by
// element through an ofstream (VTK: transposed to j-major order and byte-swapped to big-endian), the
// bytes of the file are produced on the device (tile transpose + byte swap in one pass), copied to
// pinned memory in chunks and written with pwrite; the copy of chunk k + 1 overlaps the write of
// chunk k.  With several slab contexts every rank writes its own rows / columns of the SAME file at
// their offsets (no gather); rank 0 adds the text parts.
#include "ctx.h"

#include <errno.h>
#include <fcntl.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <string>
#include <vector>

struct OutStage {
  double* d[2]      = {nullptr, nullptr};  // device: packed file bytes of one chunk
  double* h[2]      = {nullptr, nullptr};  // pinned mirror
  cudaEvent_t ev[2] = {nullptr, nullptr};
  size_t cap        = 0;                   // doubles per buffer
};

namespace {

constexpr int TT = 32;  // transpose tile

__device__ __forceinline__ unsigned long long big_endian(double v) {
  const unsigned long long u = (unsigned long long)__double_as_longlong(v);
  const unsigned lo = (unsigned)u, hi = (unsigned)(u >> 32);
  return ((unsigned long long)__byte_perm(lo, 0, 0x0123) << 32) | (unsigned long long)__byte_perm(hi, 0, 0x0123);
}

// Cell data of the VTK file: interior cells, j outer / i inner (src/VTKWriter.hpp:66-71,81-88), while
// the device layout has j contiguous: one shared-memory tile transpose, both sides coalesced.
// dst[((j - j0) * ni + (i - i0)) * E + e], E = 1: f0;  E = 3: (f0, f1, 0).
template <int E>
__global__ void __launch_bounds__(TT * 8)
k_vtk_cells(unsigned long long* __restrict__ dst, const double* __restrict__ f0, const double* __restrict__ f1,
            const Geom g, int64_t i0, int64_t ni, int64_t j0, int64_t nj) {
  __shared__ double t0[TT][TT + 1];
  __shared__ double t1[E == 3 ? TT : 1][TT + 1];
  const int64_t ib = i0 + (int64_t)blockIdx.x * TT, jb = j0 + (int64_t)blockIdx.y * TT;
  for (int r = threadIdx.y; r < TT; r += 8) {  // threadIdx.x along j (contiguous in memory)
    const int64_t i = ib + r, j = jb + threadIdx.x;
    if (i < i0 + ni && j < j0 + nj) {
      const int64_t k = g.idx(i + 1, j + 1);
      t0[r][threadIdx.x] = f0[k];
      if (E == 3) t1[r][threadIdx.x] = f1[k];
    }
  }
  __syncthreads();
  for (int r = threadIdx.y; r < TT; r += 8) {  // threadIdx.x along i (contiguous in the file)
    const int64_t j = jb + r, i = ib + threadIdx.x;
    if (i < i0 + ni && j < j0 + nj) {
      const int64_t o = ((j - j0) * ni + (i - i0)) * E;
      dst[o] = big_endian(t0[threadIdx.x][r]);
      if (E == 3) {
        dst[o + 1] = big_endian(t1[threadIdx.x][r]);
        dst[o + 2] = 0ull;
      }
    }
  }
}

// POINTS of the structured grid (src/VTKWriter.hpp:50-58): (x(i), y(j), 0) for j outer, i inner;
// x, y are the reference's Field1D arrays including their ghost entry (index + 1)
__global__ void k_vtk_points(unsigned long long* __restrict__ dst, const double* __restrict__ x,
                             const double* __restrict__ y, int64_t i0, int64_t ni, int64_t j0, int64_t nj) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= ni * nj) return;
  const int64_t i = i0 + q % ni, j = j0 + q / ni;
  dst[3 * q]     = big_endian(x[i + 1]);
  dst[3 * q + 1] = big_endian(y[j + 1]);
  dst[3 * q + 2] = 0ull;
}

int pwrite_all(int fd, const void* buf, size_t n, int64_t off, const char* what) {
  const char* p = (const char*)buf;
  while (n > 0) {
    const ssize_t w = pwrite(fd, p, n, (off_t)off);
    if (w < 0) {
      if (errno == EINTR) continue;
      fs_set_error("could not write %s: %s", what, strerror(errno));
      return 1;
    }
    p += w; off += w; n -= (size_t)w;
  }
  return 0;
}

// every rank opens (creating if needed) and sizes the file; none truncates what another wrote
int open_sized(const char* filename, int64_t total, int* fd) {
  *fd = open(filename, O_WRONLY | O_CREAT, 0644);
  if (*fd < 0) {
    fs_set_error("could not open file `%s`: %s", filename, strerror(errno));
    return 1;
  }
  if (ftruncate(*fd, (off_t)total) != 0) {
    fs_set_error("could not size file `%s`: %s", filename, strerror(errno));
    close(*fd);
    return 1;
  }
  return 0;
}

OutStage* stage(fsgpu_ctx* c, size_t min_doubles) {
  size_t want = (size_t)4 << 20;  // 32 MiB per buffer
  if (const char* e = getenv("FSGPU_OUT_CHUNK_BYTES")) want = std::max<size_t>(64, (size_t)atoll(e) / 8);
  want = std::max(want, min_doubles);
  if (c->out && c->out->cap >= want) return c->out;
  fs_output_release(c);
  OutStage* s = new OutStage();
  c->out = s;
  for (int b = 0; b < 2; ++b) {
    if (cudaMalloc(&s->d[b], want * sizeof(double)) != cudaSuccess ||
        cudaMallocHost(&s->h[b], want * sizeof(double)) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev[b], cudaEventDisableTiming) != cudaSuccess) {
      fs_set_error("output staging: %s", cudaGetErrorString(cudaGetLastError()));
      fs_output_release(c);
      return nullptr;
    }
  }
  s->cap = want;
  return s;
}

// One array section of a file: `row_b` bytes per row for the whole grid, starting at `base`; this
// rank contributes bytes [col_b, col_b + my_b) of rows [r0, r1).  land(d, h, ra, rb) enqueues on
// c->stream whatever puts rows [ra, rb) of this rank's part, packed, into the pinned buffer h (d is
// the device scratch of the same size).
template <class Land>
int stream_rows(fsgpu_ctx* c, int fd, int64_t base, int64_t row_b, int64_t col_b, int64_t my_b, int64_t r0,
                int64_t r1, const char* what, Land land) {
  if (my_b <= 0 || r1 <= r0) return 0;
  OutStage* s = stage(c, (size_t)(my_b / 8));
  if (!s) return 1;
  const int64_t rows_per = std::max<int64_t>(1, (int64_t)(s->cap * 8) / my_b);
  const int64_t nchunk   = (r1 - r0 + rows_per - 1) / rows_per;
  auto issue = [&](int64_t k) -> int {
    const int b = (int)(k & 1);
    const int64_t ra = r0 + k * rows_per, rb = std::min(r1, ra + rows_per);
    if (land(s->d[b], s->h[b], ra, rb)) return 1;
    FS_CUDA(cudaEventRecord(s->ev[b], c->stream));
    return 0;
  };
  if (issue(0)) return 1;
  for (int64_t k = 0; k < nchunk; ++k) {
    if (k + 1 < nchunk && issue(k + 1)) return 1;  // its buffer was written out one turn ago
    const int b = (int)(k & 1);
    const int64_t ra = r0 + k * rows_per, rb = std::min(r1, ra + rows_per);
    FS_CUDA(cudaEventSynchronize(s->ev[b]));
    const char* h = (const char*)s->h[b];
    if (my_b == row_b) {
      if (pwrite_all(fd, h, (size_t)((rb - ra) * my_b), base + ra * row_b, what)) return 1;
    } else {
      for (int64_t r = ra; r < rb; ++r)
        if (pwrite_all(fd, h + (r - ra) * my_b, (size_t)my_b, base + r * row_b + col_b, what)) return 1;
    }
  }
  return 0;
}

// src/IO.hpp:25-95: magic, version 1.0, little-endian header length, the dict padded with spaces so
// that the preamble is a multiple of 64 bytes, newline
std::string npy_preamble(int64_t m, int64_t n) {
  std::string h = "{'descr': '<f8', 'fortran_order': False, 'shape': (";
  h += std::to_string(m) + ", ";
  if (n > 0) h += std::to_string(n) + ", ";
  h += "), }";
  while ((6 + 2 + 2 + h.size() + 1) % 64 != 0) h.push_back(' ');
  h.push_back('\n');
  std::string out("\x93NUMPY", 6);
  out.push_back('\x01');
  out.push_back('\x00');
  const uint16_t len = (uint16_t)h.size();
  out.append((const char*)&len, 2);
  return out + h;
}

int mkdirs(const std::string& dir) {
  std::string cur;
  for (size_t k = 0; k <= dir.size(); ++k) {
    if (k == dir.size() || dir[k] == '/') {
      if (!cur.empty() && mkdir(cur.c_str(), 0755) != 0 && errno != EEXIST) {
        fs_set_error("could not create directory `%s`: %s", cur.c_str(), strerror(errno));
        return 1;
      }
    }
    if (k < dir.size()) cur.push_back(dir[k]);
  }
  return 0;
}

}  // namespace

void fs_output_release(fsgpu_ctx* c) {
  OutStage* s = c->out;
  if (!s) return;
  for (int b = 0; b < 2; ++b) {
    if (s->d[b]) cudaFree(s->d[b]);
    if (s->h[b]) cudaFreeHost(s->h[b]);
    if (s->ev[b]) cudaEventDestroy(s->ev[b]);
  }
  delete s;
  c->out = nullptr;
}

extern "C" {

int fsgpu_npy_write_host(const char* filename, const double* data, int64_t m, int64_t n) {
  if (!filename || !data || m < 0 || n < 0) {
    fs_set_error("fsgpu_npy_write_host: bad arguments");
    return 1;
  }
  const std::string pre = npy_preamble(m, n);
  const int64_t bytes   = m * (n > 0 ? n : 1) * 8;
  int fd;
  if (open_sized(filename, (int64_t)pre.size() + bytes, &fd)) return 1;
  const int rc = pwrite_all(fd, pre.data(), pre.size(), 0, filename) ||
                 pwrite_all(fd, data, (size_t)bytes, (int64_t)pre.size(), filename);
  close(fd);
  return rc;
}

int fsgpu_to_npy(fsgpu_ctx* c, int field, const char* filename) {
  FS_CHECK_CTX(c);
  FS_CHECK_FIELD(field);
  if (!filename) { fs_set_error("fsgpu_to_npy: null file name"); return 1; }
  int64_t m, n;
  field_extent(c->g, field, &m, &n);
  const std::string pre = npy_preamble(m, n);
  int fd;
  if (open_sized(filename, (int64_t)pre.size() + m * n * 8, &fd)) return 1;
  int rc = 0;
  if (c->g.first()) rc = pwrite_all(fd, pre.data(), pre.size(), 0, filename);
  // the rows this rank owns (a U-like field's extra row belongs to the last rank), as fsgpu_download
  const int64_t r0 = c->g.a_lo, r1 = c->g.last() ? m : c->g.a_hi;
  const double* f = c->f[field];
  const Geom g = c->g;
  cudaStream_t st = c->stream;
  rc = rc || stream_rows(c, fd, (int64_t)pre.size(), n * 8, 0, n * 8, r0, r1, filename,
                         [&](double*, double* h, int64_t ra, int64_t rb) -> int {
                           FS_CUDA(cudaMemcpy2DAsync(h, n * sizeof(double), f + g.idx(ra, 0),
                                                     g.pitch * sizeof(double), n * sizeof(double), rb - ra,
                                                     cudaMemcpyDeviceToHost, st));
                           return 0;
                         });
  close(fd);
  return rc;
}

int fsgpu_to_npy_state(fsgpu_ctx* c, const char* directory, const double* x, const double* y) {
  FS_CHECK_CTX(c);
  if (!directory) { fs_set_error("fsgpu_to_npy_state: null directory"); return 1; }
  const std::string dir(directory);
  if (mkdirs(dir)) return 1;
  const int64_t nx = c->g.nx, ny = c->g.ny;
  if (c->g.first() && x && y) {  // the four Field1D of init_grid, src/FS.hpp:55-73
    std::vector<double> xm((size_t)nx + 2), ym((size_t)ny + 2);
    for (int64_t i = 0; i < nx + 2; ++i) xm[i] = (x[i] + x[i + 1]) / 2;
    for (int64_t j = 0; j < ny + 2; ++j) ym[j] = (y[j] + y[j + 1]) / 2;
    if (fsgpu_npy_write_host((dir + "/x.npy").c_str(), x, nx + 3, 0) ||
        fsgpu_npy_write_host((dir + "/xm.npy").c_str(), xm.data(), nx + 2, 0) ||
        fsgpu_npy_write_host((dir + "/y.npy").c_str(), y, ny + 3, 0) ||
        fsgpu_npy_write_host((dir + "/ym.npy").c_str(), ym.data(), ny + 2, 0))
      return 1;
  }
  static const struct { const char* name; int field; } files[] = {
      {"visc", FSGPU_VISC}, {"p", FSGPU_P}, {"p_jump_u_stag", FSGPU_P_JUMP_U}, {"p_jump_v_stag", FSGPU_P_JUMP_V},
      {"rho_u_stag", FSGPU_RHO_U}, {"rho_v_stag", FSGPU_RHO_V}, {"U", FSGPU_U}, {"V", FSGPU_V},
      {"rho_u_stag_old", FSGPU_RHO_U_OLD}, {"rho_v_stag_old", FSGPU_RHO_V_OLD}, {"U_old", FSGPU_U_OLD},
      {"V_old", FSGPU_V_OLD}};
  for (const auto& f : files)
    if (fsgpu_to_npy(c, f.field, (dir + "/" + f.name + ".npy").c_str())) return 1;
  return 0;
}

int fsgpu_write_vtk(fsgpu_ctx* c, const char* filename, double t, const double* x, const double* y,
                    int n_scalars, const int* scalar_fields, const char* const* scalar_names, int n_vectors,
                    const int* vector_fields, const char* const* vector_names) {
  FS_CHECK_CTX(c);
  if (!filename || !x || !y || n_scalars < 0 || n_vectors < 0 || (n_scalars && (!scalar_fields || !scalar_names)) ||
      (n_vectors && (!vector_fields || !vector_names))) {
    fs_set_error("fsgpu_write_vtk: bad arguments");
    return 1;
  }
  const Geom g = c->g;
  const int64_t nx = g.nx, ny = g.ny;
  for (int q = 0; q < n_scalars + 2 * n_vectors; ++q) {
    const int field = q < n_scalars ? scalar_fields[q] : vector_fields[q - n_scalars];
    FS_CHECK_FIELD(field);
    int64_t m, n;
    field_extent(g, field, &m, &n);
    if (m != nx + 2 || n != ny + 2) {
      fs_set_error("fsgpu_write_vtk: field %d is not a cell field", field);
      return 1;
    }
  }
  // ---- layout of the file (identical on every rank) ----
  char num[64];
  snprintf(num, sizeof num, "%g", t);  // what `out << t` prints for a double
  std::string head = "# vtk DataFile Version 2.0\nState of FluidSolver at time t=" + std::string(num) +
                     "\nBINARY\nDATASET STRUCTURED_GRID\nDIMENSIONS " + std::to_string(nx + 1) + " " +
                     std::to_string(ny + 1) + " 1\nPOINTS " + std::to_string((nx + 1) * (ny + 1)) + " double\n";
  struct Text { int64_t off; std::string s; };
  std::vector<Text> text;
  std::vector<int64_t> base;  // points, then every scalar, then every vector
  int64_t off = 0;
  text.push_back({off, head});
  off += (int64_t)head.size();
  base.push_back(off);
  off += (nx + 1) * (ny + 1) * 24;
  text.push_back({off, "\n\nCELL_DATA " + std::to_string(nx * ny) + "\n"});
  off += (int64_t)text.back().s.size();
  for (int q = 0; q < n_scalars; ++q) {
    text.push_back({off, "SCALARS " + std::string(scalar_names[q]) + " double 1\nLOOKUP_TABLE default\n"});
    off += (int64_t)text.back().s.size();
    base.push_back(off);
    off += nx * ny * 8;
    text.push_back({off, "\n\n"});
    off += 2;
  }
  for (int q = 0; q < n_vectors; ++q) {
    text.push_back({off, "VECTORS " + std::string(vector_names[q]) + " double\n"});
    off += (int64_t)text.back().s.size();
    base.push_back(off);
    off += nx * ny * 24;
    text.push_back({off, "\n\n"});
    off += 2;
  }
  int fd;
  if (open_sized(filename, off, &fd)) return 1;
  int rc = 0;
  if (g.first())
    for (const Text& tx : text) rc = rc || pwrite_all(fd, tx.s.data(), tx.s.size(), tx.off, filename);

  // ---- this rank's columns: its interior cells, and the nodes on their left (+ the last one) ----
  const int64_t ci0 = std::max<int64_t>(g.a_lo, 1) - 1, ci1 = std::min<int64_t>(g.a_hi, nx + 1) - 1;
  const int64_t pi0 = ci0, pi1 = ci1 + (g.last() ? 1 : 0);
  cudaStream_t st = c->stream;
  double* dxy = nullptr;  // x (nx + 3) then y (ny + 3)
  if (!rc && cudaMalloc(&dxy, (size_t)(nx + ny + 6) * sizeof(double)) != cudaSuccess) {
    fs_set_error("fsgpu_write_vtk: %s", cudaGetErrorString(cudaGetLastError()));
    rc = 1;
  }
  if (!rc) {
    rc = cudaMemcpyAsync(dxy, x, (size_t)(nx + 3) * sizeof(double), cudaMemcpyHostToDevice, st) != cudaSuccess ||
         cudaMemcpyAsync(dxy + nx + 3, y, (size_t)(ny + 3) * sizeof(double), cudaMemcpyHostToDevice, st) != cudaSuccess;
    if (rc) fs_set_error("fsgpu_write_vtk: coordinate upload failed");
  }
  if (!rc) {
    const int64_t ni = pi1 - pi0;
    rc = stream_rows(c, fd, base[0], (nx + 1) * 24, pi0 * 24, ni * 24, 0, ny + 1, filename,
                     [&](double* d, double* h, int64_t ja, int64_t jb) -> int {
                       const int64_t cnt = ni * (jb - ja);
                       k_vtk_points<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(
                           (unsigned long long*)d, dxy, dxy + nx + 3, pi0, ni, ja, jb - ja);
                       FS_LAUNCHED();
                       FS_CUDA(cudaMemcpyAsync(h, d, (size_t)cnt * 24, cudaMemcpyDeviceToHost, st));
                       return 0;
                     });
  }
  const int64_t ni = ci1 - ci0;
  for (int q = 0; q < n_scalars + n_vectors && !rc; ++q) {
    const bool vec   = q >= n_scalars;
    const int64_t eb = vec ? 24 : 8;
    const double* f0 = c->f[vec ? vector_fields[2 * (q - n_scalars)] : scalar_fields[q]];
    const double* f1 = vec ? c->f[vector_fields[2 * (q - n_scalars) + 1]] : nullptr;
    rc = stream_rows(c, fd, base[1 + q], nx * eb, ci0 * eb, ni * eb, 0, ny, filename,
                     [&](double* d, double* h, int64_t ja, int64_t jb) -> int {
                       const dim3 grid((unsigned)((ni + TT - 1) / TT), (unsigned)((jb - ja + TT - 1) / TT));
                       if (vec)
                         k_vtk_cells<3><<<grid, dim3(TT, 8), 0, st>>>((unsigned long long*)d, f0, f1, g, ci0, ni, ja,
                                                                      jb - ja);
                       else
                         k_vtk_cells<1><<<grid, dim3(TT, 8), 0, st>>>((unsigned long long*)d, f0, f0, g, ci0, ni, ja,
                                                                      jb - ja);
                       FS_LAUNCHED();
                       FS_CUDA(cudaMemcpyAsync(h, d, (size_t)(ni * (jb - ja) * eb), cudaMemcpyDeviceToHost, st));
                       return 0;
                     });
  }
  if (dxy) {
    cudaStreamSynchronize(st);
    cudaFree(dxy);
  }
  close(fd);
  return rc;
}

}  // extern "C"
