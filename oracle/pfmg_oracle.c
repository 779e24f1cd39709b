/* oracle/pfmg_oracle.c -- TEST INFRASTRUCTURE ONLY (checker, never product; see fs_oracle.h).
 *
 * Plain-C restatement of the pressure solve the reference performs through HYPRE 3.0.0
 * (vendored under Submodules/hypre): Struct PCG preconditioned by ONE PFMG V-cycle, configured
 * as in src/LinearSolver_StructHypre.hpp:217-238 (MaxIter 1, MaxLevels 16, RAPType 1, Tol 0,
 * zero guess, 1 pre + 1 post relaxation, HYPRE defaults relax_type 1 = weighted Jacobi and
 * skip_relax 1 from struct_ls/pfmg.c:32,38).
 *
 * What is restated, with the HYPRE file each part follows:
 *   PCG loop and stopping rule <Cr,r>/<Cb,b> < tol^2 ............ krylov/pcg.c:328-1027
 *   anisotropy measure dxyz + coefficient-variation flag ......... struct_ls/pfmg_coarsen.c:374-503
 *   level schedule (cdir, active, Jacobi weights) ................ struct_ls/pfmg_coarsen.c:508-637
 *   operator-dependent 1-D interpolation ......................... struct_ls/pfmg_setup_interp.c:327-341
 *   non-Galerkin 5-point coarse operator ......................... struct_ls/pfmg_setup_rap5.c:339-379
 *   V-cycle with inactive levels, R = P^T ........................ struct_ls/pfmg_solve.c:129-454
 *   weighted Jacobi .............................................. struct_ls/jacobi.c:142-360
 *   coarsest-level sweeps min(4*num_levels, cmaxsize), weight 1 .. struct_ls/pfmg_setup.c:400-412
 *
 * relax_type 2 (symmetric red/black Gauss-Seidel, struct_ls/red_black_gs.c) and a larger
 * max_levels are the variant the GPU solver uses; they are here so the GPU kernels can be
 * compared against a CPU evaluation of the very same cycle.
 *
 * Index convention: d = 0 is the reference's i (x), d = 1 is j (y).  Level arrays are row-major
 * [i][j] (j fastest).  Coarse points are the EVEN global indices (hypre_PFMGSetCIndex = 0,
 * stride 2); the fine box starts at global index -1 in both directions.
 */
#include "fs_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define MAXL 64

typedef struct {
  int64_t n[2];
  int64_t lo[2];
  double* a[5]; /* 0 centre, 1 d0-minus (W), 2 d0-plus (E), 3 d1-minus (S), 4 d1-plus (N) */
  double *pm, *pp; /* interpolation weights towards the minus / plus neighbour in cdir */
  double *x, *b, *r;
  int cdir;   /* direction coarsened to reach the next level; -1 on the coarsest */
  int active; /* relax on this level */
  double weight;
} level;

struct fso_solver {
  int64_t nx, ny;
  fso_solver_opts o;
  int nl;
  level L[MAXL];
  int coarse_sweeps;
  double *p, *s, *r, *z; /* PCG vectors on level 0 */
  int is_setup;
};

static int64_t ceil_div2(int64_t a) { return a >= 0 ? (a + 1) / 2 : -((-a) / 2); }
static int64_t floor_div2(int64_t a) { return a >= 0 ? a / 2 : -((-a + 1) / 2); }
static double* dalloc(int64_t n) { return (double*)calloc((size_t)(n > 0 ? n : 1), sizeof(double)); }
static int hlog2(int64_t p) { /* utilities: hypre_Log2 = floor(log2(p)), <=0 -> 0 */
  int e = 0;
  if (p <= 0) return 0;
  while (p > 1) { e++; p /= 2; }
  return e;
}

static void level_alloc(level* l) {
  const int64_t n = l->n[0] * l->n[1];
  for (int e = 0; e < 5; ++e) l->a[e] = dalloc(n);
  l->pm = dalloc(n);
  l->pp = dalloc(n);
  l->x  = dalloc(n);
  l->b  = dalloc(n);
  l->r  = dalloc(n);
}
static void level_free(level* l) {
  for (int e = 0; e < 5; ++e) free(l->a[e]);
  free(l->pm); free(l->pp); free(l->x); free(l->b); free(l->r);
  memset(l, 0, sizeof(*l));
}

fso_solver* fso_solver_create(int64_t nx, int64_t ny, const fso_solver_opts* o) {
  fso_solver* s = (fso_solver*)calloc(1, sizeof(*s));
  s->nx = nx; s->ny = ny; s->o = *o;
  const int64_t n = (nx + 2) * (ny + 2);
  s->p = dalloc(n); s->s = dalloc(n); s->r = dalloc(n); s->z = dalloc(n);
  return s;
}
static void free_levels(fso_solver* s) {
  for (int l = 0; l < s->nl; ++l) level_free(&s->L[l]);
  s->nl = 0;
}
void fso_solver_destroy(fso_solver* s) {
  if (!s) return;
  free_levels(s);
  free(s->p); free(s->s); free(s->r); free(s->z);
  free(s);
}
int fso_solver_num_levels(const fso_solver* s) { return s->nl; }

#define AT(l, i, j) ((i) * (l)->n[1] + (j))

/* y = A x on a level (5-point, zero outside the grid) */
static void matvec(const level* l, const double* x, double* y) {
  const int64_t n0 = l->n[0], n1 = l->n[1];
  for (int64_t i = 0; i < n0; ++i)
    for (int64_t j = 0; j < n1; ++j) {
      const int64_t k = AT(l, i, j);
      double v = l->a[0][k] * x[k];
      if (i > 0) v += l->a[1][k] * x[k - n1];
      if (i < n0 - 1) v += l->a[2][k] * x[k + n1];
      if (j > 0) v += l->a[3][k] * x[k - 1];
      if (j < n1 - 1) v += l->a[4][k] * x[k + 1];
      y[k] = v;
    }
}
static double dot(const double* a, const double* b, int64_t n) {
  double s = 0.0;
  for (int64_t k = 0; k < n; ++k) s += a[k] * b[k];
  return s;
}

/* struct_ls/pfmg_setup_interp.c:327-341 */
static void setup_interp(level* l) {
  const int d = l->cdir;
  const int em = d == 0 ? 1 : 3, ep = d == 0 ? 2 : 4;      /* entries along cdir */
  const int tm = d == 0 ? 3 : 1, tp = d == 0 ? 4 : 2;      /* transverse entries */
  const int64_t n = l->n[0] * l->n[1];
  for (int64_t k = 0; k < n; ++k) {
    const double center = l->a[0][k] + l->a[tm][k] + l->a[tp][k];
    if (center != 0.0) {
      l->pm[k] = -l->a[em][k] / center;
      l->pp[k] = -l->a[ep][k] / center;
    } else {
      l->pm[k] = l->pp[k] = 0.0;
    }
  }
}

/* struct_ls/pfmg_setup_rap5.c:339-379 */
static void setup_rap(const level* f, level* c, int snap) {
  const int d = f->cdir;
  const int em = d == 0 ? 1 : 3, ep = d == 0 ? 2 : 4;
  const int tm = d == 0 ? 3 : 1, tp = d == 0 ? 4 : 2;
  const int64_t sd = d == 0 ? f->n[1] : 1; /* fine stride along cdir */
  for (int64_t ic = 0; ic < c->n[0]; ++ic)
    for (int64_t jc = 0; jc < c->n[1]; ++jc) {
      /* fine local index of this coarse point: global 2*(lo_c + kc), local minus lo_f */
      const int64_t fi = d == 0 ? 2 * (c->lo[0] + ic) - f->lo[0] : ic;
      const int64_t fj = d == 1 ? 2 * (c->lo[1] + jc) - f->lo[1] : jc;
      const int64_t k = AT(f, fi, fj), kc = AT(c, ic, jc);
      const int64_t pos = d == 0 ? fi : fj, nd = f->n[d];
      const int has_m = pos > 0, has_p = pos < nd - 1;
      const double pm_m = has_m ? f->pm[k - sd] : 0.0; /* weight of F-point below from ITS minus */
      const double pp_m = has_m ? f->pp[k - sd] : 0.0; /* weight of F-point below from this C-point */
      const double pm_p = has_p ? f->pm[k + sd] : 0.0; /* weight of F-point above from this C-point */
      const double pp_p = has_p ? f->pp[k + sd] : 0.0;
      c->a[em][kc] = f->a[em][k] * pm_m;
      c->a[ep][kc] = f->a[ep][k] * pp_p;
      double west = f->a[tm][k] + 0.5 * (has_m ? f->a[tm][k - sd] : 0.0) +
                    0.5 * (has_p ? f->a[tm][k + sd] : 0.0);
      double east = f->a[tp][k] + 0.5 * (has_m ? f->a[tp][k - sd] : 0.0) +
                    0.5 * (has_p ? f->a[tp][k + sd] : 0.0);
      if (f->a[tm][k] == 0.0) west = 0.0;
      if (f->a[tp][k] == 0.0) east = 0.0;
      c->a[tm][kc] = west;
      c->a[tp][kc] = east;
      double dc = f->a[0][k] + f->a[tm][k] + f->a[tp][k] + f->a[em][k] * pp_m +
                  f->a[ep][k] * pm_p - west - east;
      /* opts.snap_zero_diag; NOT in HYPRE (same guard as k_rap, fluidsolver_b200/csrc/solver_setup.cu): a singular all-Neumann operator
       * has zero row sums on every level, so the one-point coarsest grid's diagonal is 0 in exact arithmetic and
       * pure cancellation noise otherwise (3e-16 on the rho = 10, dx = 5/64 channel of scaling/LuchiniIBChannel.cpp).
       * hypre_StructMatrixZeroDiagonal tests == 0.0; noise that fails that test is divided by in the coarsest
       * relaxation and PCG breaks down (HYPRE itself: 10 iterations, residual 6e-8, at tol 1e-10 on that system).
       * Off by default: where HYPRE's own solve stalls at the rounding floor (tgv32 at tol 1e-13) the guard changes
       * the iteration count. */
      const double scale = fabs(f->a[0][k]) + fabs(f->a[tm][k]) + fabs(f->a[tp][k]) + fabs(f->a[em][k] * pp_m) +
                           fabs(f->a[ep][k] * pm_p) + fabs(west) + fabs(east);
      if (snap && fabs(dc) <= 1e-13 * scale) dc = 0.0;
      c->a[0][kc] = dc;
    }
}

/* op: Layout::F, 5 entries per cell {C, W, E, S, N} (src/LinearSystem.hpp:19) */
void fso_solver_set_operator(fso_solver* s, const double* op) {
  free_levels(s);
  level* L = s->L;
  L[0].n[0] = s->nx + 2; L[0].n[1] = s->ny + 2;
  L[0].lo[0] = -1; L[0].lo[1] = -1;
  level_alloc(&L[0]);
  const int64_t n0 = L[0].n[0], n1 = L[0].n[1];
  for (int64_t i = 0; i < n0; ++i)
    for (int64_t j = 0; j < n1; ++j)
      for (int e = 0; e < 5; ++e) L[0].a[e][i * n1 + j] = op[5 * (i + j * n0) + e];

  /* hypre_PFMGComputeMaxLevels, pfmg_coarsen.c:17-33 */
  int max_levels = 1 + (hlog2(n0) + 2) + (hlog2(n1) + 2);
  if (s->o.max_levels > 0 && s->o.max_levels < max_levels) max_levels = s->o.max_levels;
  if (max_levels > MAXL) max_levels = MAXL;

  /* hypre_PFMGComputeDxyz, pfmg_coarsen.c:374-503 (cxyz from pfmg_cxyz_vc.c) */
  double cxyz[2] = {0, 0}, sq[2] = {0, 0}, dxyz[2];
  for (int64_t k = 0; k < n0 * n1; ++k) {
    const double sign = L[0].a[0][k] < 0.0 ? 1.0 : -1.0;
    const double w0 = sign * (L[0].a[1][k] + L[0].a[2][k]);
    const double w1 = sign * (L[0].a[3][k] + L[0].a[4][k]);
    cxyz[0] += w0; sq[0] += w0 * w0;
    cxyz[1] += w1; sq[1] += w1 * w1;
  }
  const double gs = (double)(n0 * n1);
  double mean[2] = {cxyz[0] / gs, cxyz[1] / gs}, dev[2] = {sq[0] / gs, sq[1] / gs};
  double cmax = cxyz[0] > cxyz[1] ? cxyz[0] : cxyz[1];
  if (cmax < 0.0) cmax = 0.0;
  if (cmax == 0.0) { cxyz[0] = cxyz[1] = 1.0; cmax = 1.0; }
  for (int d = 0; d < 2; ++d) {
    const double max_aniso = 1.79769313486231570e+308 / 1000;
    if (cxyz[d] > cmax / max_aniso) dxyz[d] = sqrt(1.0 / (cxyz[d] / cmax));
    else dxyz[d] = sqrt(max_aniso);
  }
  int dxyz_flag = 0;
  for (int d = 0; d < 2; ++d) {
    dev[d] -= mean[d] * mean[d];
    if (dev[d] > 0.1 * (mean[d] * mean[d])) { dxyz_flag = 1; break; }
  }

  /* hypre_PFMGCoarsen, pfmg_coarsen.c:508-637 */
  int coarsen[2] = {1, 1};
  int64_t lo[2] = {-1, -1}, hi[2] = {s->nx, s->ny};
  int l;
  for (l = 0; l < max_levels; ++l) {
    double min_dxyz = 1 + dxyz[0] + dxyz[1], alpha = 0.0, beta = 0.0;
    int cdir = -1;
    for (int d = 0; d < 2; ++d) {
      if (hi[d] > lo[d] && dxyz[d] < min_dxyz) { min_dxyz = dxyz[d]; cdir = d; }
      alpha += 1.0 / (dxyz[d] * dxyz[d]);
    }
    L[l].weight = 1.0;
    if (cdir != -1) {
      if (dxyz_flag) L[l].weight = 2.0 / 3.0;
      else {
        for (int d = 0; d < 2; ++d)
          if (d != cdir) beta += 1.0 / (dxyz[d] * dxyz[d]);
        L[l].weight = 2.0 / (3.0 - beta / alpha);
      }
      if (l == max_levels - 1) cdir = -1;
      /* Hierarchies deeper than the reference's cap of 16 (src/LinearSolver_StructHypre.hpp:221) stop at a
       * two-point coarsest grid instead of a single point.  The one-point level of an all-Neumann system is
       * the zero operator; on the anisotropic channel grid (4096 x 1024 on 100 x 1) carrying it made the
       * V-cycle indefinite and PCG broke down after ~4 iterations at a 1e-2 residual, while one level
       * fewer converges like the capped hierarchy (and the cavity's iteration counts do not change). */
      if (cdir != -1 && s->o.max_levels > 16) {
        int64_t nc[2] = {hi[0] - lo[0] + 1, hi[1] - lo[1] + 1};
        nc[cdir] = floor_div2(hi[cdir]) - ceil_div2(lo[cdir]) + 1;
        if (nc[0] * nc[1] <= 1) cdir = -1;
      }
    }
    if (cdir == -1) { L[l].active = 1; L[l].cdir = -1; break; }
    L[l].cdir = cdir;
    if (coarsen[cdir] != 0) { L[l].active = 1; coarsen[0] = coarsen[1] = 0; }
    else L[l].active = 0;
    coarsen[cdir] = 1;
    dxyz[cdir] *= 2;
    lo[cdir] = ceil_div2(lo[cdir]);
    hi[cdir] = floor_div2(hi[cdir]);
    L[l + 1].n[0] = hi[0] - lo[0] + 1; L[l + 1].n[1] = hi[1] - lo[1] + 1;
    L[l + 1].lo[0] = lo[0]; L[l + 1].lo[1] = lo[1];
  }
  s->nl = l + 1;
  if (!s->o.skip_relax)
    for (l = 0; l < s->nl; ++l) L[l].active = 1;

  for (l = 0; l < s->nl - 1; ++l) {
    level_alloc(&L[l + 1]);
    setup_interp(&L[l]);
    setup_rap(&L[l], &L[l + 1], s->o.snap_zero_diag);
  }
  /* zero diagonal on the coarsest grid => no relaxation there, pfmg_setup.c:349-357 */
  level* c = &L[s->nl - 1];
  for (int64_t k = 0; k < c->n[0] * c->n[1]; ++k)
    if (c->a[0][k] == 0.0) c->active = 0;
  const int64_t cmaxsize = c->n[0] > c->n[1] ? c->n[0] : c->n[1];
  s->coarse_sweeps = 4 * s->nl < cmaxsize ? 4 * s->nl : (int)cmaxsize; /* pfmg_setup.c:400-412 */
  s->is_setup = 1;
}

/* weighted Jacobi, struct_ls/jacobi.c:142-360 */
static void jacobi(level* l, double w, int sweeps, int zero_guess) {
  const int64_t n = l->n[0] * l->n[1];
  int it = 0;
  if (sweeps == 0) { if (zero_guess) memset(l->x, 0, sizeof(double) * (size_t)n); return; }
  if (zero_guess) {
    for (int64_t k = 0; k < n; ++k) l->x[k] = w * l->b[k] / l->a[0][k];
    it = 1;
  }
  for (; it < sweeps; ++it) {
    matvec(l, l->x, l->r);
    for (int64_t k = 0; k < n; ++k) l->r[k] = l->b[k] - l->r[k];
    for (int64_t k = 0; k < n; ++k) l->x[k] += w * l->r[k] / l->a[0][k];
  }
}

/* one colour of Gauss-Seidel: points with (gi + gj) parity == colour (global indices) */
static void gs_colour(level* l, int colour) {
  const int64_t n0 = l->n[0], n1 = l->n[1];
  for (int64_t i = 0; i < n0; ++i)
    for (int64_t j = 0; j < n1; ++j) {
      if ((((i + l->lo[0]) + (j + l->lo[1])) & 1) != colour) continue;
      const int64_t k = AT(l, i, j);
      if (l->a[0][k] == 0.0) continue;
      double v = l->b[k];
      if (i > 0) v -= l->a[1][k] * l->x[k - n1];
      if (i < n0 - 1) v -= l->a[2][k] * l->x[k + n1];
      if (j > 0) v -= l->a[3][k] * l->x[k - 1];
      if (j < n1 - 1) v -= l->a[4][k] * l->x[k + 1];
      l->x[k] = v / l->a[0][k];
    }
}
/* symmetric red/black GS: pre = red, black ; post = black, red */
static void rbgs(level* l, int sweeps, int zero_guess, int post) {
  const int64_t n = l->n[0] * l->n[1];
  if (zero_guess) memset(l->x, 0, sizeof(double) * (size_t)n);
  for (int it = 0; it < sweeps; ++it) {
    gs_colour(l, post ? 1 : 0);
    gs_colour(l, post ? 0 : 1);
  }
}
static void relax(fso_solver* s, int li, int zero_guess, int post) {
  level* l = &s->L[li];
  if (s->o.relax_type == 2) rbgs(l, 1, zero_guess, post);
  else jacobi(l, l->weight, 1, zero_guess);
}

/* b_c = P^T r  (semi_restrict / StructMatvec with transpose, pfmg_solve.c:205-207) */
static void restrict_residual(const level* f, level* c) {
  const int d = f->cdir;
  const int64_t sd = d == 0 ? f->n[1] : 1;
  for (int64_t ic = 0; ic < c->n[0]; ++ic)
    for (int64_t jc = 0; jc < c->n[1]; ++jc) {
      const int64_t fi = d == 0 ? 2 * (c->lo[0] + ic) - f->lo[0] : ic;
      const int64_t fj = d == 1 ? 2 * (c->lo[1] + jc) - f->lo[1] : jc;
      const int64_t k = AT(f, fi, fj);
      const int64_t pos = d == 0 ? fi : fj, nd = f->n[d];
      double v = f->r[k];
      if (pos > 0) v += f->pp[k - sd] * f->r[k - sd];
      if (pos < nd - 1) v += f->pm[k + sd] * f->r[k + sd];
      c->b[AT(c, ic, jc)] = v;
    }
}
/* x_f += P x_c  (pfmg_solve.c:353-355) */
static void interpolate_add(level* f, const level* c) {
  const int d = f->cdir;
  for (int64_t i = 0; i < f->n[0]; ++i)
    for (int64_t j = 0; j < f->n[1]; ++j) {
      const int64_t k = AT(f, i, j);
      const int64_t g = (d == 0 ? i + f->lo[0] : j + f->lo[1]); /* global index along cdir */
      const int64_t oi = d == 0 ? 0 : i, oj = d == 1 ? 0 : j;
      if ((g & 1) == 0) {
        const int64_t kc = g / 2 - c->lo[d];
        f->x[k] += c->x[d == 0 ? AT(c, kc, oj) : AT(c, oi, kc)];
      } else {
        const int64_t cm = floor_div2(g - 1) - c->lo[d], cp = floor_div2(g + 1) - c->lo[d];
        double v = 0.0;
        if (cm >= 0) v += f->pm[k] * c->x[d == 0 ? AT(c, cm, oj) : AT(c, oi, cm)];
        if (cp < c->n[d]) v += f->pp[k] * c->x[d == 0 ? AT(c, cp, oj) : AT(c, oi, cp)];
        f->x[k] += v;
      }
    }
}

/* z = C r : one V-cycle from a zero guess, struct_ls/pfmg_solve.c:129-454 with max_iter 1 */
static void vcycle(fso_solver* s, const double* r, double* z) {
  level* L = s->L;
  const int nl = s->nl;
  const int64_t n = L[0].n[0] * L[0].n[1];
  memcpy(L[0].b, r, sizeof(double) * (size_t)n);
  relax(s, 0, 1, 0);
  if (nl > 1) {
    matvec(&L[0], L[0].x, L[0].r);
    for (int64_t k = 0; k < n; ++k) L[0].r[k] = L[0].b[k] - L[0].r[k];
    restrict_residual(&L[0], &L[1]);
    int l;
    for (l = 1; l <= nl - 2; ++l) {
      const int64_t nn = L[l].n[0] * L[l].n[1];
      if (L[l].active) {
        relax(s, l, 1, 0);
        matvec(&L[l], L[l].x, L[l].r);
        for (int64_t k = 0; k < nn; ++k) L[l].r[k] = L[l].b[k] - L[l].r[k];
      } else {
        memset(L[l].x, 0, sizeof(double) * (size_t)nn);
        memcpy(L[l].r, L[l].b, sizeof(double) * (size_t)nn);
      }
      restrict_residual(&L[l], &L[l + 1]);
    }
    /* bottom: l == nl-1 */
    if (L[l].active) {
      if (s->o.relax_type == 2) {
        rbgs(&L[l], s->coarse_sweeps, 1, 0);
        rbgs(&L[l], s->coarse_sweeps, 0, 1);
      } else {
        jacobi(&L[l], 1.0, s->coarse_sweeps, 1);
      }
    } else {
      memset(L[l].x, 0, sizeof(double) * (size_t)(L[l].n[0] * L[l].n[1]));
    }
    for (l = nl - 2; l >= 1; --l) {
      interpolate_add(&L[l], &L[l + 1]);
      if (L[l].active) relax(s, l, 0, 1);
    }
    interpolate_add(&L[0], &L[1]);
  }
  relax(s, 0, 0, 1);
  memcpy(z, L[0].x, sizeof(double) * (size_t)n);
}

/* krylov/pcg.c:328-1027 with the settings of src/LinearSolver_StructHypre.hpp:332-338
 * (two_norm 0, rel_change 0, a_tol 0): x0 = 0. rhs is Layout::F, sol Layout::C. */
int fso_solver_solve(fso_solver* s, const double* rhs, double* sol, double* res) {
  const int64_t n0 = s->nx + 2, n1 = s->ny + 2, n = n0 * n1;
  double* b = dalloc(n);
  double* x = dalloc(n);
  for (int64_t i = 0; i < n0; ++i)
    for (int64_t j = 0; j < n1; ++j) b[i * n1 + j] = rhs[i + j * n0];
  double *p = s->p, *sv = s->s, *r = s->r;
  int it = 0;
  double i_prod = 0.0;
  vcycle(s, b, p);                         /* p = C b */
  const double bi_prod = dot(p, b, n);     /* <C b, b> */
  if (!(bi_prod > 0.0)) {                  /* pcg.c:473-488: zero rhs -> x = 0 */
    memset(sol, 0, sizeof(double) * (size_t)n);
    if (res) *res = 0.0;
    free(b); free(x);
    return 0;
  }
  const double eps = s->o.tol * s->o.tol;
  memcpy(r, b, sizeof(double) * (size_t)n); /* r = b - A*0 */
  vcycle(s, r, p);                          /* p = C r */
  double gamma = dot(r, p, n);
  while (it + 1 <= s->o.max_iter) {
    it++;
    matvec(&s->L[0], p, sv);
    const double sdotp = dot(sv, p, n);
    if (sdotp == 0.0) { if (it == 1) i_prod = gamma; break; }
    const double alpha = gamma / sdotp;
    if (alpha <= 0.0) { if (it == 1) i_prod = gamma; break; }
    const double gamma_old = gamma;
    for (int64_t k = 0; k < n; ++k) x[k] += alpha * p[k];
    for (int64_t k = 0; k < n; ++k) r[k] -= alpha * sv[k];
    vcycle(s, r, sv);                       /* s = C r */
    gamma = dot(r, sv, n);
    i_prod = gamma;
    if (i_prod / bi_prod < eps) break;      /* converged */
    if (gamma <= 0.0) break;
    const double beta = gamma / gamma_old;
    for (int64_t k = 0; k < n; ++k) p[k] = sv[k] + beta * p[k];
  }
  memcpy(sol, x, sizeof(double) * (size_t)n);
  if (res) *res = sqrt(i_prod / bi_prod);  /* pcg.c:1016 */
  free(b); free(x);
  return it;
}
