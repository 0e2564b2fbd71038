/*
 * doma_oracle.c — TEST INFRASTRUCTURE ONLY.  CPU restatement of DOMA's sweep hot path.
 *
 * This file is the parity oracle for the CUDA path in doma_b200/csrc.  It is NOT part of the
 * product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.  The product library (libdoma_b200.so) never links or calls it.
 *
 * Every function restates one member of the reference's TransportSolver3D<T> / TransportSolver2D<T>
 * in plain C over flat arrays, with the reference's expression order so that, at one thread, the
 * results are bit-identical to the reference built with its own flags (makefile:1: -O2, no FMA
 * contraction on x86-64).  Parity is PINNED: tests/test_oracle_vs_reference.py checks this oracle
 * against oracle/_ref/ref_harness (the unmodified reference compiled from /root/reference) and
 * against the golden fixtures under tests/golden/ that ref_harness generated.
 *
 * Citations are relative to the DOMA tree: "3D" = src/TransportSolver3D.cpp, "2D" =
 * src/TransportSolver2D.cpp.  The entry points mirror include/doma_b200.h one to one
 * (doma_oracle_X <-> doma_b200_X) and take the same doma_b200_problem.
 */
#include "doma_b200.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

typedef struct doma_oracle_ctx
{
  int dims, nx, ny, nz, G, M, D, n_ord, scheme, mode, threads;
  long N;
  double dt;
  double *dx, *dy, *dz;
  int *mat;
  int *mat_flags, *n_delayed, *step_type;
  double *sigma_t, *sigma_s, *nu_sigma_f, *chi_p, *src, *inv_v, *chi_d, *beta, *lambda;
  double *step_t_insert, *step_t_remove, *step_src;
  int *ord_octant;
  double *mu, *eta, *xi, *w;
  /* state (reference: per-cell members of CartesianCell3D, include/CartesianCell3D.h:84-121) */
  double *phi_tot, *phi_prev, *phi_last; /* [G][N] */
  double *dnp_cur, *dnp_last;            /* [D][N] */
  double *q, *phi_cur;                   /* [N] */
  long sweeps;
  char err[256];
} doma_oracle_ctx;

static char g_create_err[256];

static void *
dup_arr(const void * p, size_t bytes)
{
  void * r = calloc(bytes ? bytes : 1, 1);
  if (p && bytes)
    memcpy(r, p, bytes);
  return r;
}

int
doma_oracle_create(const doma_b200_problem * p, doma_oracle_ctx ** out)
{
  if (!p || !out || (p->dims != 2 && p->dims != 3) || p->nx < 1 || p->ny < 1 || p->n_groups < 1 || p->n_mat < 1)
  {
    snprintf(g_create_err, sizeof g_create_err, "doma_oracle_create: bad problem");
    return 1;
  }
  doma_oracle_ctx * c = (doma_oracle_ctx *)calloc(1, sizeof *c);
  c->dims = p->dims;
  c->nx = p->nx;
  c->ny = p->ny;
  c->nz = p->dims == 3 ? p->nz : 1;
  c->N = (long)c->nx * c->ny * c->nz;
  c->G = p->n_groups;
  c->M = p->n_mat;
  c->D = p->n_delayed_max;
  c->n_ord = p->n_ord;
  c->scheme = p->scheme;
  c->mode = p->mode;
  c->dt = p->dt;
  c->threads = 1;
  const size_t G = c->G, M = c->M, D = c->D > 0 ? c->D : 0;
  c->dx = dup_arr(p->dx, sizeof(double) * c->nx);
  c->dy = dup_arr(p->dy, sizeof(double) * c->ny);
  c->dz = dup_arr(p->dims == 3 ? p->dz : NULL, sizeof(double) * c->nz);
  if (p->dims != 3)
    c->dz[0] = 1.0;
  c->mat = dup_arr(p->mat_of_cell, sizeof(int) * c->N);
  c->mat_flags = dup_arr(p->mat_flags, sizeof(int) * M);
  c->n_delayed = dup_arr(p->n_delayed, sizeof(int) * M);
  c->step_type = dup_arr(p->step_type, sizeof(int) * M);
  if (!p->step_type)
    for (size_t m = 0; m < M; ++m)
      c->step_type[m] = DOMA_STEP_NONE;
  c->sigma_t = dup_arr(p->sigma_t, sizeof(double) * M * G);
  c->sigma_s = dup_arr(p->sigma_s, sizeof(double) * M * G * G);
  c->nu_sigma_f = dup_arr(p->nu_sigma_f, sizeof(double) * M * G);
  c->chi_p = dup_arr(p->chi_p, sizeof(double) * M * G);
  c->src = dup_arr(p->src, sizeof(double) * M * G);
  c->inv_v = dup_arr(p->inv_v, sizeof(double) * M * G);
  c->chi_d = dup_arr(p->chi_d, sizeof(double) * M * G * D);
  c->beta = dup_arr(p->beta, sizeof(double) * M * G * D);
  c->lambda = dup_arr(p->lambda, sizeof(double) * M * D);
  c->step_t_insert = dup_arr(p->step_t_insert, sizeof(double) * M);
  c->step_t_remove = dup_arr(p->step_t_remove, sizeof(double) * M);
  c->step_src = dup_arr(p->step_src, sizeof(double) * M * G);
  c->ord_octant = dup_arr(p->ord_octant, sizeof(int) * c->n_ord);
  c->mu = dup_arr(p->mu, sizeof(double) * c->n_ord);
  c->eta = dup_arr(p->eta, sizeof(double) * c->n_ord);
  c->xi = dup_arr(p->xi, sizeof(double) * c->n_ord);
  c->w = dup_arr(p->w, sizeof(double) * c->n_ord);
  for (long i = 0; i < c->N; ++i)
    if (c->mat[i] < 0 || c->mat[i] >= c->M)
    {
      snprintf(g_create_err, sizeof g_create_err, "doma_oracle_create: cell %ld has material %d outside [0,%d)", i, c->mat[i], c->M);
      free(c);
      return 2;
    }
  c->phi_tot = calloc(G * c->N, sizeof(double));
  c->phi_prev = calloc(G * c->N, sizeof(double));
  c->phi_last = calloc(G * c->N, sizeof(double));
  c->dnp_cur = calloc((D ? D : 1) * c->N, sizeof(double));
  c->dnp_last = calloc((D ? D : 1) * c->N, sizeof(double));
  c->q = calloc(c->N, sizeof(double));
  c->phi_cur = calloc(c->N, sizeof(double));
  *out = c;
  return 0;
}

int
doma_oracle_destroy(doma_oracle_ctx * c)
{
  if (!c)
    return 0;
  void * ptrs[] = {c->dx, c->dy, c->dz, c->mat, c->mat_flags, c->n_delayed, c->step_type, c->sigma_t, c->sigma_s,
                   c->nu_sigma_f, c->chi_p, c->src, c->inv_v, c->chi_d, c->beta, c->lambda, c->step_t_insert,
                   c->step_t_remove, c->step_src, c->ord_octant, c->mu, c->eta, c->xi, c->w, c->phi_tot, c->phi_prev,
                   c->phi_last, c->dnp_cur, c->dnp_last, c->q, c->phi_cur};
  for (size_t i = 0; i < sizeof ptrs / sizeof ptrs[0]; ++i)
    free(ptrs[i]);
  free(c);
  return 0;
}

const char *
doma_oracle_last_error(const doma_oracle_ctx * c)
{
  return c ? c->err : g_create_err;
}

int
doma_oracle_set_threads(doma_oracle_ctx * c, int threads)
{
  c->threads = threads > 0 ? threads : 1;
  return 0;
}

int doma_oracle_set_stream(doma_oracle_ctx * c, void * s) { (void)c; (void)s; return 0; }
int doma_oracle_synchronize(doma_oracle_ctx * c) { (void)c; return 0; }

static inline double
cell_measure(const doma_oracle_ctx * c, long i)
{
  /* CartesianCell3D::_volume = l_x*l_y*l_z (src/CartesianCell3D.cpp:19); 2D _area = l_x*l_y */
  const int ix = (int)(i % c->nx), iy = (int)((i / c->nx) % c->ny), iz = (int)(i / ((long)c->nx * c->ny));
  return c->dims == 3 ? c->dx[ix] * c->dy[iy] * c->dz[iz] : c->dx[ix] * c->dy[iy];
}

/* ------------------------------------------------------------------------------------------------
 * state management
 * ---------------------------------------------------------------------------------------------- */

/* initializeSolve (3D:499-529): phi_tot = 1.0 (eigen) / 0.0; phi_prev = 0; q = 0; phi_cur = 0.
 * NB the reference uses vector::resize(G, value), which only fills NEW elements; on the first (and,
 * in every shipped driver path, only) call the vectors are empty, so this is a plain fill. */
int
doma_oracle_reset_state(doma_oracle_ctx * c, double phi_init)
{
  for (long i = 0; i < c->G * c->N; ++i)
  {
    c->phi_tot[i] = phi_init;
    c->phi_prev[i] = 0.0;
  }
  memset(c->q, 0, sizeof(double) * c->N);
  memset(c->phi_cur, 0, sizeof(double) * c->N);
  return 0;
}

/* initZeroIC (3D:140-156) */
int
doma_oracle_init_zero_ic(doma_oracle_ctx * c)
{
  memset(c->phi_tot, 0, sizeof(double) * c->G * c->N);
  memset(c->phi_prev, 0, sizeof(double) * c->G * c->N);
  memset(c->phi_last, 0, sizeof(double) * c->G * c->N);
  if (c->D > 0)
  {
    memset(c->dnp_cur, 0, sizeof(double) * c->D * c->N);
    memset(c->dnp_last, 0, sizeof(double) * c->D * c->N);
  }
  return 0;
}

/* updateStepFluxes (3D:212-234) */
int
doma_oracle_update_step_fluxes(doma_oracle_ctx * c)
{
  memcpy(c->phi_last, c->phi_tot, sizeof(double) * c->G * c->N);
  memset(c->phi_tot, 0, sizeof(double) * c->G * c->N);
  memset(c->q, 0, sizeof(double) * c->N);
  memset(c->phi_cur, 0, sizeof(double) * c->N);
  return 0;
}

/* stepDNPs (3D:236-260) */
int
doma_oracle_step_dnps(doma_oracle_ctx * c)
{
  const int G = c->G, D = c->D;
  const long N = c->N;
#pragma omp parallel for num_threads(c->threads)
  for (long i = 0; i < N; ++i)
  {
    const int m = c->mat[i];
    const int nd = c->n_delayed[m];
    for (int d = 0; d < nd; ++d)
    {
      double c_src = 0.0;
      for (int g = 0; g < G; ++g)
        c_src += c->nu_sigma_f[m * G + g] * c->phi_tot[g * N + i] * c->beta[(m * G + g) * D + d];
      const double lam = c->lambda[m * D + d];
      const double decay = exp(-lam * c->dt);
      c->dnp_cur[d * N + i] = (c_src / lam) * (1.0 - decay) + c->dnp_last[d * N + i] * decay;
      c->dnp_last[d * N + i] = c->dnp_cur[d * N + i];
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * source build
 * ---------------------------------------------------------------------------------------------- */

/* updateMultigroupSource(g,t) (3D:339-402; 2D same lines with 0.5 instead of 0.25) */
int
doma_oracle_update_mg_source(doma_oracle_ctx * c, int32_t g, double t)
{
  const int G = c->G, D = c->D;
  const long N = c->N;
  const double f = c->dims == 3 ? 0.25 : 0.5;
  const int transient = c->mode == DOMA_MODE_TRANSIENT;
#pragma omp parallel for num_threads(c->threads)
  for (long i = 0; i < N; ++i)
  {
    const int m = c->mat[i];
    const int nd = c->n_delayed[m];
    const int has_chi = (c->mat_flags[m] & DOMA_MAT_HAS_CHI_P) != 0;
    c->phi_cur[i] = 0.0;
    double q = (c->mat_flags[m] & DOMA_MAT_HAS_SRC) ? f * c->src[m * G + g] / M_PI : 0.0;
    if (transient && c->step_type[m] != DOMA_STEP_NONE)
    {
      const double s = f * c->step_src[m * G + g] / M_PI;
      switch (c->step_type[m])
      {
        case DOMA_STEP_BOTH: q += (c->step_t_insert[m] <= t && t <= c->step_t_remove[m]) ? s : 0.0; break;
        case DOMA_STEP_INSERT: q += (c->step_t_insert[m] <= t) ? s : 0.0; break;
        case DOMA_STEP_REMOVE: q += (t <= c->step_t_remove[m]) ? s : 0.0; break;
        default: break;
      }
    }
    for (int gp = 0; gp < G; ++gp)
    {
      const double phi = c->phi_tot[gp * N + i];
      if (gp != g)
        q += f * c->sigma_s[(m * G + g) * G + gp] * phi / M_PI;
      if (has_chi && nd == 0)
        q += f * c->chi_p[m * G + g] * c->nu_sigma_f[m * G + gp] * phi / M_PI;
      else if (has_chi && nd > 0)
      {
        double g_beta = 0.0;
        for (int d = 0; d < nd; ++d)
          g_beta += c->beta[(m * G + gp) * D + d];
        q += f * (1.0 - g_beta) * c->chi_p[m * G + g] * c->nu_sigma_f[m * G + gp] * phi / M_PI;
      }
      /* the delayed and 1/(v dt) terms sit inside the g' loop -> added G times */
      if (nd > 0 && transient)
        for (int d = 0; d < nd; ++d)
          q += f * c->chi_d[(m * G + g) * D + d] * c->dnp_cur[d * N + i] * c->lambda[m * D + d] / M_PI;
      if (transient)
        q += f * c->phi_last[g * N + i] * c->inv_v[m * G + g] / c->dt / M_PI;
    }
    c->q[i] = q;
    c->phi_prev[g * N + i] = c->phi_tot[g * N + i];
    c->phi_tot[g * N + i] = 0.0;
  }
  return 0;
}

/* updateMultigroupSourceEigen(g) (3D:404-430); "+=" onto the stale q */
int
doma_oracle_update_mg_source_eigen(doma_oracle_ctx * c, int32_t g, double k)
{
  const int G = c->G;
  const long N = c->N;
  const double f = c->dims == 3 ? 0.25 : 0.5;
#pragma omp parallel for num_threads(c->threads)
  for (long i = 0; i < N; ++i)
  {
    const int m = c->mat[i];
    const int has_chi = (c->mat_flags[m] & DOMA_MAT_HAS_CHI_P) != 0;
    c->phi_cur[i] = 0.0;
    double q = c->q[i];
    for (int gp = 0; gp < G; ++gp)
    {
      const double phi = c->phi_tot[gp * N + i];
      if (gp != g)
        q += f * c->sigma_s[(m * G + g) * G + gp] * phi / M_PI;
      if (has_chi)
        q += f * c->chi_p[m * G + g] * c->nu_sigma_f[m * G + gp] * phi / M_PI / k;
    }
    c->q[i] = q;
    c->phi_prev[g * N + i] = c->phi_tot[g * N + i];
    c->phi_tot[g * N + i] = 0.0;
  }
  return 0;
}

/* updateScatteringSource(g) (3D:531-545) */
int
doma_oracle_update_scatter_source(doma_oracle_ctx * c, int32_t g)
{
  const int G = c->G;
  const long N = c->N;
  const double f = c->dims == 3 ? 0.25 : 0.5;
#pragma omp parallel for num_threads(c->threads)
  for (long i = 0; i < N; ++i)
  {
    const int m = c->mat[i];
    c->phi_tot[g * N + i] += c->phi_cur[i];
    c->q[i] = f * c->sigma_s[(m * G + g) * G + g] * c->phi_cur[i] / M_PI;
    c->phi_cur[i] = 0.0;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * reductions
 * ---------------------------------------------------------------------------------------------- */

/* computeScatteringResidual(g) (3D:547-563) */
int
doma_oracle_scatter_residual(doma_oracle_ctx * c, int32_t g, double * out)
{
  const long N = c->N;
  double diff = 0.0, total = 0.0;
#pragma omp parallel for reduction(+ : diff, total) num_threads(c->threads)
  for (long i = 0; i < N; ++i)
  {
    const double v = cell_measure(c, i);
    const double cur = c->phi_cur[i], s = c->phi_tot[g * N + i] + c->phi_cur[i];
    diff = diff + (cur * cur) * v;
    total = total + (s * s) * v;
  }
  *out = total > 1e-8 ? sqrt(diff) / sqrt(total) : 0.0;
  return 0;
}

int doma_oracle_sweep(doma_oracle_ctx * c, int32_t g);

/* sourceIteration(g) (3D:458-497): do { sweep; residual; update } while (it < max_it && tol < residual) */
int
doma_oracle_source_iteration(doma_oracle_ctx * c, int32_t g, double tol, int32_t max_it, int32_t * iters_out,
                             double * residual_out, double * residual_history)
{
  int32_t it = 0;
  double res = 0.0;
  do
  {
    doma_oracle_sweep(c, g);
    doma_oracle_scatter_residual(c, g, &res);
    doma_oracle_update_scatter_source(c, g);
    if (residual_history)
      residual_history[it] = res;
    it++;
  } while (it < max_it && tol < res);
  *iters_out = it;
  *residual_out = res;
  return 0;
}

/* computeMGFluxResidual() (3D:318-337) */
int
doma_oracle_mg_residual(doma_oracle_ctx * c, double * out)
{
  const long N = c->N;
  const int G = c->G;
  double diff = 0.0, total = 0.0;
#pragma omp parallel for reduction(+ : diff, total) num_threads(c->threads)
  for (long i = 0; i < N; ++i)
  {
    const double v = cell_measure(c, i);
    for (int g = 0; g < G; ++g)
    {
      const double d = c->phi_tot[g * N + i] - c->phi_prev[g * N + i], s = c->phi_tot[g * N + i];
      diff = diff + (d * d) * v;
      total = total + (s * s) * v;
    }
  }
  *out = total > 1e-8 ? sqrt(diff) / sqrt(total) : 0.0;
  return 0;
}

/* updateEigenvalue() (3D:432-456) */
int
doma_oracle_update_k(doma_oracle_ctx * c, double * k_inout, double * k_prev_out)
{
  const long N = c->N;
  const int G = c->G;
  const double k = *k_inout;
  double num = 0.0, den = 0.0;
#pragma omp parallel for reduction(+ : num, den) num_threads(c->threads)
  for (long i = 0; i < N; ++i)
  {
    const int m = c->mat[i];
    if (!(c->mat_flags[m] & DOMA_MAT_HAS_NU_SIGMA_F))
      continue;
    const double v = cell_measure(c, i);
    for (int g = 0; g < G; ++g)
    {
      num = num + v * c->nu_sigma_f[m * G + g] * c->phi_tot[g * N + i];
      den = den + v * c->nu_sigma_f[m * G + g] * c->phi_prev[g * N + i] / k;
    }
  }
  *k_prev_out = k;
  *k_inout = num / den;
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * sweep
 * ---------------------------------------------------------------------------------------------- */

/* One ordinate over the whole mesh in the reference's lexicographic order z -> y -> x
 * (sweepPPP..sweepMMM, 3D:655-874; 2D:635-725), direction per axis from the octant's sign pattern.
 * The reference keeps six face slots per cell per thread (src/BrickMesh3D.cpp:176-183) and reads
 * the upwind neighbour's downwind slot; every slot read was written earlier in the same traversal,
 * so carrying the x face in a scalar, the y faces in a row and the z faces in a plane holds exactly
 * the same values.  Boundary faces are vacuum: 0.0 (src/CartesianCell3D.cpp:31-38). */
static void
sweep_ordinate_3d(const doma_oracle_ctx * c, int g, int oct, double mu, double eta, double xi, double w,
                  double * swept, double * yrow, double * zplane)
{
  const int nx = c->nx, ny = c->ny, nz = c->nz, G = c->G;
  const int sx = (oct & 4) ? -1 : 1, sy = (oct & 2) ? -1 : 1, sz = (oct & 1) ? -1 : 1;
  const int twdd = c->scheme == DOMA_SCHEME_TWDD;
  const double theta = 0.9, theta_s = 0.9;
  for (int kk = 0; kk < nz; ++kk)
  {
    const int k = sz > 0 ? kk : nz - 1 - kk;
    for (int jj = 0; jj < ny; ++jj)
    {
      const int j = sy > 0 ? jj : ny - 1 - jj;
      double xf = 0.0;
      for (int ii = 0; ii < nx; ++ii)
      {
        const int i = sx > 0 ? ii : nx - 1 - ii;
        const long id = ((long)k * ny + j) * nx + i;
        const double lx = c->dx[i], ly = c->dy[j], lz = c->dz[k];
        const double sig = c->sigma_t[c->mat[id] * G + g];
        const double q = c->q[id];
        const double x = xf, y = jj == 0 ? 0.0 : yrow[i], z = kk == 0 ? 0.0 : zplane[(long)j * nx + i];
        double center, ox, oy, oz;
        if (twdd)
        {
          /* include/TWDiamondDifference3D.h:38-86 */
          const double A = ly * lz, B = lx * lz, C = lx * ly, V = lx * ly * lz;
          double a = 0.5;
          if (x > 1e-8)
          {
            a = q * V * theta_s + (eta * B * y + xi * C * z) * theta + mu * A * x;
            a /= (sig * V + 2.0 * eta * B + 2.0 * xi * C) * x;
            a = -1.0 * (a - 1.0);
          }
          double b = 0.5;
          if (y > 1e-8)
          {
            b = q * V * theta_s + (mu * A * x + xi * C * z) * theta + eta * B * y;
            b /= (sig * V + 2.0 * mu * A + 2.0 * xi * C) * y;
            b = -1.0 * (b - 1.0);
          }
          double cc = 0.5;
          if (z > 1e-8)
          {
            cc = q * V * theta_s + (mu * A * x + eta * B * y) * theta + xi * C * z;
            cc /= (sig * V + 2.0 * mu * A + 2.0 * eta * B) * z;
            cc = -1.0 * (cc - 1.0);
          }
          a = a < 0.5 ? 0.5 : a;
          a = a > 1.0 ? 1.0 : a;
          b = b < 0.5 ? 0.5 : b;
          b = b > 1.0 ? 1.0 : b;
          cc = cc < 0.5 ? 0.5 : cc;
          cc = cc > 1.0 ? 1.0 : cc;
          center = q * V + (mu * A / a * x) + (eta * B / b * y) + (xi * C / cc * z);
          center /= sig * V + (mu * A / a) + (eta * B / b) + (xi * C / cc);
          ox = (center / a) - (((1.0 - a) / a) * x);
          oy = (center / b) - (((1.0 - b) / b) * y);
          oz = (center / cc) - (((1.0 - cc) / cc) * z);
        }
        else
        {
          /* include/DiamondDifference3D.h:35-40 */
          center = q + (2.0 * x * mu / lx) + (2.0 * y * eta / ly) + (2.0 * z * xi / lz);
          center /= (2.0 * mu / lx) + (2.0 * eta / ly) + (2.0 * xi / lz) + sig;
          ox = 2.0 * center - x;
          oy = 2.0 * center - y;
          oz = 2.0 * center - z;
        }
        swept[id] += w * center;
        xf = ox;
        yrow[i] = oy;
        zplane[(long)j * nx + i] = oz;
      }
    }
  }
}

static void
sweep_ordinate_2d(const doma_oracle_ctx * c, int g, int oct, double mu, double eta, double xi, double w,
                  double * swept, double * yrow)
{
  const int nx = c->nx, ny = c->ny, G = c->G;
  const int sx = (oct & 4) ? -1 : 1, sy = (oct & 2) ? -1 : 1;
  const int twdd = c->scheme == DOMA_SCHEME_TWDD;
  const int transient = c->mode == DOMA_MODE_TRANSIENT;
  const double theta = 0.9, theta_s = 0.9;
  for (int jj = 0; jj < ny; ++jj)
  {
    const int j = sy > 0 ? jj : ny - 1 - jj;
    double xf = 0.0;
    for (int ii = 0; ii < nx; ++ii)
    {
      const int i = sx > 0 ? ii : nx - 1 - ii;
      const long id = (long)j * nx + i;
      const double lx = c->dx[i], ly = c->dy[j];
      const int m = c->mat[id];
      /* 2D passes (_mode,_dt) to the cell solve (2D:648-651) */
      double tot = c->sigma_t[m * G + g];
      if (transient)
        tot += c->inv_v[m * G + g] / c->dt;
      const double q = c->q[id];
      const double x = xf, y = jj == 0 ? 0.0 : yrow[i];
      double center, ox, oy;
      if (twdd)
      {
        /* include/TWDiamondDifference2D.h:36-71, including the bare "+ abs_xi"  */
        const double A = ly, B = lx, V = lx * ly;
        double a = 0.5;
        if (x > 1e-8)
        {
          a = q * V * theta_s + (eta * B * y) * theta + mu * A * x;
          a /= (tot * V + 2.0 * eta * B) * x;
          a = -1.0 * (a - 1.0);
        }
        double b = 0.5;
        if (y > 1e-8)
        {
          b = q * V * theta_s + (mu * A * x + xi) * theta + eta * B * y;
          b /= (tot * V + 2.0 * mu * A) * y;
          b = -1.0 * (b - 1.0);
        }
        a = a < 0.5 ? 0.5 : a;
        a = a > 1.0 ? 1.0 : a;
        b = b < 0.5 ? 0.5 : b;
        b = b > 1.0 ? 1.0 : b;
        center = q * V + (mu * A / a * x) + (eta * B / b * y);
        center /= tot * V + (mu * A / a) + (eta * B / b);
        ox = (center / a) - (((1.0 - a) / a) * x);
        oy = (center / b) - (((1.0 - b) / b) * y);
      }
      else
      {
        /* include/DiamondDifference2D.h:34-40 */
        center = q + (2.0 * x * mu / lx) + (2.0 * y * eta / ly);
        center /= (2.0 * mu / lx) + (2.0 * eta / ly) + tot;
        ox = 2.0 * center - x;
        oy = 2.0 * center - y;
      }
      swept[id] += w * center;
      xf = ox;
      yrow[i] = oy;
    }
  }
}

/* sweep(g) (3D:565-622, 2D:565-602) with parallelSweepOctant (3D:624-653): octants in enum order,
 * `omp parallel for` (static schedule) over the ordinates of ONE octant, each thread accumulating
 * into its own swept[tid][cell]; then phi_cur[c] += swept[0][c] + ... in tid order. */
int
doma_oracle_sweep(doma_oracle_ctx * c, int32_t g)
{
  const long N = c->N;
  const int T = c->threads;
  double * swept = calloc((size_t)T * N, sizeof(double));
  double * yrow = malloc(sizeof(double) * (size_t)T * c->nx);
  double * zplane = malloc(sizeof(double) * (size_t)T * c->nx * c->ny);
  if (!swept || !yrow || !zplane)
  {
    snprintf(c->err, sizeof c->err, "doma_oracle_sweep: out of memory");
    free(swept); free(yrow); free(zplane);
    return 3;
  }
  for (int oct = 0; oct < 8; ++oct)
  {
    /* this context's ordinates of octant `oct`, in the order given (the reference's order) */
    int first = -1, count = 0;
    for (int n = 0; n < c->n_ord; ++n)
      if (c->ord_octant[n] == oct)
      {
        if (first < 0)
          first = n;
        ++count;
      }
    if (count == 0)
      continue;
    int * idx = malloc(sizeof(int) * count);
    for (int n = 0, k = 0; n < c->n_ord; ++n)
      if (c->ord_octant[n] == oct)
        idx[k++] = n;
#pragma omp parallel for schedule(static) num_threads(T)
    for (int n = 0; n < count; ++n)
    {
#ifdef _OPENMP
      const int tid = omp_get_thread_num();
#else
      const int tid = 0;
#endif
      const int o = idx[n];
      const double mu = fabs(c->mu[o]), eta = fabs(c->eta[o]), xi = fabs(c->xi[o]);
      if (c->dims == 3)
        sweep_ordinate_3d(c, g, oct, mu, eta, xi, c->w[o], swept + (size_t)tid * N, yrow + (size_t)tid * c->nx,
                          zplane + (size_t)tid * c->nx * c->ny);
      else
        sweep_ordinate_2d(c, g, oct, mu, eta, xi, c->w[o], swept + (size_t)tid * N, yrow + (size_t)tid * c->nx);
    }
    free(idx);
  }
#pragma omp parallel for num_threads(T)
  for (long i = 0; i < N; ++i)
    for (int t = 0; t < T; ++t)
      c->phi_cur[i] += swept[(size_t)t * N + i];
  free(swept); free(yrow); free(zplane);
  c->sweeps++;
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * host <-> state
 * ---------------------------------------------------------------------------------------------- */
static double *
arr_ptr(doma_oracle_ctx * c, int which, int index)
{
  const long N = c->N;
  switch (which)
  {
    case DOMA_ARR_PHI_TOT: return index >= 0 && index < c->G ? c->phi_tot + index * N : NULL;
    case DOMA_ARR_PHI_PREV: return index >= 0 && index < c->G ? c->phi_prev + index * N : NULL;
    case DOMA_ARR_PHI_LAST: return index >= 0 && index < c->G ? c->phi_last + index * N : NULL;
    case DOMA_ARR_DNP_CUR: return index >= 0 && index < c->D ? c->dnp_cur + index * N : NULL;
    case DOMA_ARR_DNP_LAST: return index >= 0 && index < c->D ? c->dnp_last + index * N : NULL;
    case DOMA_ARR_Q: return c->q;
    case DOMA_ARR_PHI_CUR: return c->phi_cur;
    default: return NULL;
  }
}

int
doma_oracle_get_array(doma_oracle_ctx * c, int32_t which, int32_t index, double * host_out)
{
  const double * p = arr_ptr(c, which, index);
  if (!p)
  {
    snprintf(c->err, sizeof c->err, "doma_oracle_get_array: bad array %d[%d]", which, index);
    return 4;
  }
  memcpy(host_out, p, sizeof(double) * c->N);
  return 0;
}

int
doma_oracle_set_array(doma_oracle_ctx * c, int32_t which, int32_t index, const double * host_in)
{
  double * p = arr_ptr(c, which, index);
  if (!p)
  {
    snprintf(c->err, sizeof c->err, "doma_oracle_set_array: bad array %d[%d]", which, index);
    return 4;
  }
  memcpy(p, host_in, sizeof(double) * c->N);
  return 0;
}

int
doma_oracle_get_info(doma_oracle_ctx * c, int32_t what, int64_t * out)
{
  switch (what)
  {
    case DOMA_INFO_KERNEL_LAUNCHES: *out = 0; return 0;
    case DOMA_INFO_N_CELLS: *out = c->N; return 0;
    case DOMA_INFO_N_ORD: *out = c->n_ord; return 0;
    default: *out = 0; return 0;
  }
}
