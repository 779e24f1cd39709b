// fluidsolver_b200/csrc/solver_pcg.cu -- preconditioned conjugate gradients on the device.
//
// hypre_PCGSolve (Submodules/hypre/src/krylov/pcg.c:328-1027) with the settings of
// src/LinearSolver_StructHypre.hpp:332-338 (two_norm 0, rel_change 0, a_tol 0, zero guess):
//   p = C b; bi = <p, b>;  r = b;  gamma = bi
//   loop: s = A p; alpha = gamma / <s,p>; x += alpha p; r -= alpha s; s = C r; gamma' = <r,s>;
//         stop if gamma'/bi < tol^2; beta = gamma'/gamma; p = s + beta p
// (HYPRE evaluates C r a second time before the loop; with x0 = 0 that equals C b bit for bit,
// so it is computed once here.)
//
// All recurrence scalars live in device memory (PcgScal); dot products are deterministic
// two-stage reductions whose last block updates the scalars, so an iteration needs no host
// round trip.  The host runs FS_LAG iterations ahead and reads the `done` flag through pinned
// memory; kernels of iterations enqueued past convergence return immediately.
#include "solver.h"

#include <math.h>
#include <stdlib.h>

#define FS_LAG 1

namespace {

constexpr int BX = 64, BY = 4;

// s = A p and <s, p>;  last block: alpha = gamma / <s,p> with HYPRE's breakdown exits
__global__ void __launch_bounds__(BX* BY, 8)
k_matvec_dot(const LevelDev L, const double* __restrict__ p, double* __restrict__ s, PcgScal* scal,
             double* part, unsigned int* ticket) {
  if (scal->done) return;
  double v[1] = {0.0};
  const int64_t P = L.pitch;
  for_each_point(L.i0, L.i1, L.n1, [&](int64_t li, int64_t lj) {
    const int64_t k = L.idx(li, lj);
    const double pk = __ldg(p + k);
    double a = __ldg(L.d + k) * pk;
    a += __ldg(L.cx + k) * __ldg(p + k - P);
    a += __ldg(L.cx + k + P) * __ldg(p + k + P);
    a += __ldg(L.cy + k) * __ldg(p + k - 1);
    a += __ldg(L.cy + k + 1) * __ldg(p + k + 1);
    s[k] = a;
    v[0] += a * pk;
  });
  double tot[1];
  if (grid_reduce<ReduceSum, 1>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0) pcg_local_sp(scal, tot[0]);
  }
}

// x += alpha p; r -= alpha s.  Persistent grid (as the reduction kernels): the copy of this kernel
// that is enqueued past convergence then costs ~3 us instead of the 140 us a 264k-CTA empty grid takes.
// The sum of all entries of the new x rides along (-> *sum_out, this rank's rows): after the last iteration
// it is the integral shift_pressure_to_zero needs (src/Operators.hpp:78-82), so the projection needs no
// separate summation pass over delta_p.
// Slab overlap: the launch is split in two (boundary rows first, RowSplit; ACC adds the second launch's sum to the
// first's), so that the exchange of r's halo rows runs beside the interior launch.
// bsum != null (first iteration after fsgpu_substep): r still is the rhs with its mean in it, which goes here.
template <bool ACC>
__global__ void __launch_bounds__(BX* BY)
k_xr_update(const LevelDev L, const RowSplit rr, double* __restrict__ x, double* __restrict__ r,
            const double* __restrict__ p, const double* __restrict__ s, const PcgScal* scal, double* part,
            unsigned int* ticket, double* sum_out, const double* bsum, double bsize) {
  if (scal->done) return;
  const double alpha = scal->alpha;
  const double bm = bsum ? __ldg(bsum) / bsize : 0.0;
  double v[1] = {0.0};
  for_each_point2(rr, L.n1, [&](int64_t li, int64_t lj) {
    const int64_t k = L.idx(li, lj);
    const double xn = x[k] + alpha * p[k];
    x[k] = xn;
    v[0] += xn;
    const double rk = bsum ? r[k] - bm : r[k];
    r[k] = rk - alpha * s[k];
  });
  double tot[1];
  if (grid_reduce<ReduceSum, 1>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0) sum_out[0] = ACC ? sum_out[0] + tot[0] : tot[0];
  }
}

// gamma' = <r, z>.  INIT: this is bi = <C b, b>; otherwise the convergence test + beta.
template <bool INIT>
__global__ void __launch_bounds__(BX* BY)
k_dot_rz(const LevelDev L, const double* __restrict__ r, const double* __restrict__ z, PcgScal* scal,
         double* part, unsigned int* ticket) {
  if (scal->done) return;
  double v[1] = {0.0};
  for_each_point(L.i0, L.i1, L.n1, [&](int64_t li, int64_t lj) {
    const int64_t k = L.idx(li, lj);
    v[0] += r[k] * z[k];
  });
  double tot[1];
  if (grid_reduce<ReduceSum, 1>(v, part, ticket, tot)) {
    if (threadIdx.x == 0 && threadIdx.y == 0) pcg_local_rz<INIT>(scal, tot[0]);
  }
}

// p = z + beta p   (FIRST: p = z); persistent grid for the same reason
template <bool FIRST>
__global__ void __launch_bounds__(BX* BY)
k_p_update(const LevelDev L, const RowSplit rr, double* __restrict__ p, const double* __restrict__ z,
           const PcgScal* scal) {
  if (scal->done) return;
  const double beta = FIRST ? 0.0 : scal->beta;
  for_each_point2(rr, L.n1, [&](int64_t li, int64_t lj) {
    const int64_t k = L.idx(li, lj);
    if (FIRST) p[k] = z[k];
    else p[k] = z[k] + beta * p[k];
  });
}

// (the reduction kernels return at once after convergence and leave red[] stale: skip then too)
__global__ void k_after_sp(PcgScal* scal) {
  if (!scal->done) pcg_after_sp(scal, scal->red[0]);
}
template <bool INIT>
__global__ void k_after_rz(PcgScal* scal) {
  if (!scal->done) pcg_after_rz<INIT>(scal, scal->red[0]);
}

__global__ void k_scal_init(PcgScal* scal, double eps, int nranks, const P2PRedDev* p2p, double* sum_x) {
  sum_x[0] = 0.0;  // x = 0
  scal->nranks = nranks;
  scal->p2p = p2p;
  scal->gamma = scal->sdotp = scal->alpha = scal->beta = scal->bi = scal->iprod = 0.0;
  scal->eps   = eps;
  scal->iter  = 0;
  scal->done  = PCG_RUNNING;
  scal->first = 1;
}

}  // namespace

int ps_finish_sp(fsgpu_ps* s) {
  if (s->c->nranks <= 1 || comm_red_dev(s->c)) return 0;  // mailbox transport: done inside the reduction kernel
  if (comm_allreduce(s->c, s->scal->red, 1, COMM_SUM)) return 1;
  k_after_sp<<<1, 1, 0, s->c->stream>>>(s->scal);
  FS_LAUNCHED();
  return 0;
}
int ps_finish_rz(fsgpu_ps* s, bool init) {
  if (s->c->nranks <= 1 || comm_red_dev(s->c)) return 0;
  if (comm_allreduce(s->c, s->scal->red, 1, COMM_SUM)) return 1;
  if (init) k_after_rz<true><<<1, 1, 0, s->c->stream>>>(s->scal);
  else k_after_rz<false><<<1, 1, 0, s->c->stream>>>(s->scal);
  FS_LAUNCHED();
  return 0;
}
int ps_halo(fsgpu_ps* s, int l, double* const* ptrs, int n, int h) {
  if (s->c->nranks <= 1) return 0;
  const Level& L = s->L[l];
  return comm_halo(s->c, ptrs, n, L.v.pitch, L.own0, L.own1, h);
}

// Boundary-first split of a level's owned rows [i0, i1) for an exchange of h halo rows: `bnd` = the h rows
// next to each neighbouring slab, `in` = the rest.  on == false: too small to be worth it (or no mailbox).
SlabSplit ps_split(fsgpu_ps* s, int l, int h) {
  fsgpu_ctx* c = s->c;
  const int64_t min_pts = comm_overlap_min_points(c);
  const Level& L = s->L[l];
  const int64_t i0 = L.own0, i1 = L.own1;
  SlabSplit sp;
  const bool up = c->rank > 0, dn = c->rank + 1 < c->nranks;
  const int64_t lo = i0 + (up ? h : 0), hi = i1 - (dn ? h : 0);
  sp.on  = min_pts > 0 && !L.replicated && i1 - i0 >= 4 * h && (i1 - i0) * L.v.n1 >= min_pts;
  sp.bnd = RowSplit{i0, lo, hi, i1};
  sp.in  = RowSplit{lo, hi, 0, 0};
  return sp;
}

int ps_solve(fsgpu_ps* s, double* residual, int* iters) {
  fsgpu_ctx* c    = s->c;
  cudaStream_t st = c->stream;
  const LevelDev& L0 = s->L[0].v;
  const dim3 bl(BX, BY);
  const unsigned pg = persist_blocks(L0.i1 - L0.i0, L0.n1);
  if (fs_ensure_part(c, (size_t)pg * 4)) return 1;
  double* x = c->f[FSGPU_DELTA_P];
  double* r = c->f[FSGPU_RHS];  // r = b - A 0 = b: the rhs buffer becomes the residual
  const size_t prof0 = fs_prof_mark();
  g_prof_tag = 0;
  k_scal_init<<<1, 1, 0, st>>>(s->scal, s->o.tol * s->o.tol, c->nranks, comm_red_dev(c), c->red_out + 1);
  FS_LAUNCHED();
  FS_CUDA(cudaMemsetAsync(alloc_base(c->g, x), 0, (size_t)c->g.count() * sizeof(double), st));
  // slab decomposition, per iteration: r needs FS_RO halo rows before the V-cycle (restriction
  // reaches two rows out), p one halo row before the matvec
  if (ps_halo(s, 0, &r, 1, FS_RO)) return 1;
  // red_out[12]: a copy of the rhs sum that later reductions of the solve leave alone
  if (s->rhs_mean_pending) {
    FS_CUDA(cudaMemcpyAsync(c->red_out + 12, c->red_out, sizeof(double), cudaMemcpyDeviceToDevice, st));
    s->b_shift      = c->red_out + 12;
    s->b_shift_size = (double)((c->g.nx + 2) * (c->g.ny + 2));
  }
  s->rhs_mean_pending = false;
  if (ps_vcycle(s, r, s->z, 2)) return 1;  // z = C b and bi = <z, b>
  const RowSplit all{L0.i0, L0.i1, 0, 0};
  const SlabSplit sp1 = ps_split(s, 0, 1), sp2 = ps_split(s, 0, FS_RO);  // p: one halo row, r: FS_RO
  const unsigned pg1 = persist_blocks(2, L0.n1), pg2 = persist_blocks(2 * FS_RO, L0.n1);
  k_p_update<true><<<pg, bl, 0, st>>>(L0, all, s->p, s->z, s->scal);
  FS_LAUNCHED();
  if (ps_halo(s, 0, &s->p, 1, 1)) return 1;
  const int nslot = FS_LAG + 1;
  int launched = 0;
  bool stop = false;
  for (int it = 1; it <= s->o.max_iter && !stop; ++it) {
    g_prof_tag = it;
    {
      ProfScope ps_(PROF_MATVEC, st);
      k_matvec_dot<<<pg, bl, 0, st>>>(L0, s->p, s->s, s->scal, c->part, c->ticket);
      FS_LAUNCHED();
      if (ps_finish_sp(s)) return 1;
    }
    {
      ProfScope ps_(PROF_XR, st);
      bool async = false;
      if (sp2.on) {  // boundary rows, their exchange on the transport's stream, interior rows beside it
        k_xr_update<false><<<pg2, bl, 0, st>>>(L0, sp2.bnd, x, r, s->p, s->s, s->scal, c->part, c->ticket, c->red_out + 1,
                                               s->b_shift, s->b_shift_size);
        FS_LAUNCHED();
        if (comm_halo_async(c, &r, 1, L0.pitch, s->L[0].own0, s->L[0].own1, FS_RO, &async)) return 1;
        k_xr_update<true><<<pg, bl, 0, st>>>(L0, sp2.in, x, r, s->p, s->s, s->scal, c->part, c->ticket, c->red_out + 1,
                                             s->b_shift, s->b_shift_size);
        FS_LAUNCHED();
      } else {
        k_xr_update<false><<<pg, bl, 0, st>>>(L0, all, x, r, s->p, s->s, s->scal, c->part, c->ticket, c->red_out + 1,
                                              s->b_shift, s->b_shift_size);
        FS_LAUNCHED();
      }
      s->b_shift = nullptr;  // r is the residual from here on
      if (async) { if (comm_halo_wait(c)) return 1; }
      else if (ps_halo(s, 0, &r, 1, FS_RO)) return 1;
    }
    if (ps_vcycle(s, r, s->z, 1)) return 1;  // z = C r and gamma' = <r, z>
    {
      ProfScope ps_(PROF_PUPD, st);
      bool async = false;
      if (sp1.on) {
        k_p_update<false><<<pg1, bl, 0, st>>>(L0, sp1.bnd, s->p, s->z, s->scal);
        FS_LAUNCHED();
        if (comm_halo_async(c, &s->p, 1, L0.pitch, s->L[0].own0, s->L[0].own1, 1, &async)) return 1;
        k_p_update<false><<<pg, bl, 0, st>>>(L0, sp1.in, s->p, s->z, s->scal);
        FS_LAUNCHED();
      } else {
        k_p_update<false><<<pg, bl, 0, st>>>(L0, all, s->p, s->z, s->scal);
        FS_LAUNCHED();
      }
      if (async) { if (comm_halo_wait(c)) return 1; }
      else if (ps_halo(s, 0, &s->p, 1, 1)) return 1;
    }
    const int slot = it % nslot;
    FS_CUDA(cudaMemcpyAsync(&s->h_scal[slot], s->scal, sizeof(PcgScal), cudaMemcpyDeviceToHost, st));
    FS_CUDA(cudaEventRecord(s->ev[slot], st));
    launched = it;
    if (it > FS_LAG) {  // look at the state FS_LAG iterations back while this one runs
      const int ps = (it - FS_LAG) % nslot;
      FS_CUDA(cudaEventSynchronize(s->ev[ps]));
      if (s->h_scal[ps].done) stop = true;
    }
  }
  (void)launched;
  s->b_shift = nullptr;
  // the sum of all entries of the solution, on every rank (fs_launch_correct's sum_ready)
  if (comm_allreduce(c, c->red_out + 1, 1, COMM_SUM)) return 1;
  FS_CUDA(cudaMemcpyAsync(&s->h_scal[0], s->scal, sizeof(PcgScal), cudaMemcpyDeviceToHost, st));
  FS_CUDA(cudaStreamSynchronize(st));
  const PcgScal& h = s->h_scal[0];
  g_prof_tag = 0;
  if (comm_failed(c)) return 1;
  // (the p update of the iteration that met the convergence test is empty as well)
  if (g_prof_mask) fs_prof_drop_after(prof0, h.done == PCG_ZERO_RHS ? 0 : h.iter, h.done ? PROF_PUPD : -1);
  if (h.done == PCG_ZERO_RHS) {
    if (residual) *residual = 0.0;
    if (iters) *iters = 0;
    return 0;
  }
  if (residual) *residual = sqrt(h.iprod / h.bi);  // pcg.c:1016
  if (iters) *iters = h.iter;
  return 0;
}
