// elementwise.cu — source build, state management and reduction kernels (FP64 CUDA cores, HBM-bound).
//
// One thread per cell, structure-of-arrays, coalesced along x.  This translation unit is compiled
// with -fmad=false and every expression keeps the reference's operation order, so the per-cell
// values are bit-identical to the reference's (g++ -O2 on x86-64 does not contract to FMA,
// SURVEY Appendix A); only the order of the two-stage reductions differs (fixed, deterministic).
// Citations: "3D" = src/TransportSolver3D.cpp of the DOMA tree (2D twin: same lines of
// src/TransportSolver2D.cpp with 0.5 instead of 0.25 and _area instead of _volume).
#include "ctx.h"

#include <math.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

namespace
{
constexpr int EW_THREADS = 256;
constexpr int RED_BLOCKS = 1184; // 148 SMs x 8 resident blocks of 256 threads

inline int
ew_blocks(int64_t n)
{
  return (int)((n + EW_THREADS - 1) / EW_THREADS);
}

struct Geo
{
  const double *dx, *dy, *dz;
  int nx, ny, dims;
};

// Walks the cells id0, id0 + stride, id0 + 2 stride, ... of a grid-stride loop and yields each cell's measure
// (CartesianCell3D::_volume = l_x*l_y*l_z, src/CartesianCell3D.cpp:19; 2D: _area = l_x*l_y) without a 64-bit
// division per cell: (i, j, k) is divided out once, then advanced by the precomputed digits of the stride with carries.
struct CellWalk
{
  const Geo & g;
  int i, j, k, si, sj, sk;
  __device__ CellWalk(const Geo & geo, int64_t id0, int64_t stride) : g(geo)
  {
    i = (int)(id0 % geo.nx);
    const int64_t r = id0 / geo.nx;
    j = (int)(r % geo.ny);
    k = (int)(r / geo.ny);
    si = (int)(stride % geo.nx);
    const int64_t rs = stride / geo.nx;
    sj = (int)(rs % geo.ny);
    sk = (int)(rs / geo.ny);
  }
  __device__ double measure() const { return g.dims == 3 ? g.dx[i] * g.dy[j] * g.dz[k] : g.dx[i] * g.dy[j]; }
  __device__ void next()
  {
    i += si;
    int c = i >= g.nx;
    i -= c ? g.nx : 0;
    j += sj + c;
    c = j >= g.ny;
    j -= c ? g.ny : 0;
    k += sk + c;
  }
};

Geo
geo_of(const doma_b200_ctx * c)
{
  return Geo{c->dx, c->dy, c->dz, c->nx, c->ny, c->dims};
}

__global__ void
fill_kernel(double * __restrict__ p, int64_t n, double v)
{
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    p[i] = v;
}

// device-to-device moves of the asynchronous host-buffer sweeps, as kernels: a cudaMemcpyAsync D2D may be put on the
// copy engine that is busy with the neighbouring group's PCIe transfer and then waits 2.5 ms for it (measured)
__global__ void
stage_in_kernel(double * __restrict__ q, const double * __restrict__ q_stage, double * __restrict__ phi_cur, int64_t lo, int64_t hi)
{
  const int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < hi)
  {
    q[i] = q_stage[i];
    phi_cur[i] = 0.0;
  }
}

__global__ void
copy_kernel(double * __restrict__ dst, const double * __restrict__ src, int64_t n)
{
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    dst[i] = src[i];
}

// stepDNPs (3D:236-260)
__global__ void
step_dnps_kernel(int64_t N, int64_t lo, int64_t hi, int G, int D, double dt, const int32_t * __restrict__ mat, const int32_t * __restrict__ n_delayed,
                 const double * __restrict__ nu_sigma_f, const double * __restrict__ beta, const double * __restrict__ lambda,
                 const double * __restrict__ phi_tot, double * __restrict__ dnp_cur, double * __restrict__ dnp_last)
{
  const int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= hi)
    return;
  const int m = mat[i];
  const int nd = n_delayed[m];
  for (int d = 0; d < nd; ++d)
  {
    double c_src = 0.0;
    for (int g = 0; g < G; ++g)
      c_src += nu_sigma_f[m * G + g] * phi_tot[g * N + i] * beta[(m * G + g) * D + d];
    const double lam = lambda[m * D + d];
    const double decay = exp(-lam * dt);
    const double v = (c_src / lam) * (1.0 - decay) + dnp_last[d * N + i] * decay;
    dnp_cur[d * N + i] = v;
    dnp_last[d * N + i] = v;
  }
}

// updateMultigroupSource(g,t) (3D:339-402).  GT = compile-time group count (1..8; 0 = run-time G): the G
// loads of phi_tot are issued up front instead of one per iteration behind an IEEE division.
template <int GT>
__global__ void
mg_source_kernel(int64_t N, int64_t lo, int64_t hi, int G, int D, int g, double t, double f, int transient, double dt,
                 const int32_t * __restrict__ mat, const int32_t * __restrict__ mat_flags,
                 const int32_t * __restrict__ n_delayed, const int32_t * __restrict__ step_type,
                 const double * __restrict__ step_t_insert, const double * __restrict__ step_t_remove,
                 const double * __restrict__ step_src, const double * __restrict__ src, const double * __restrict__ sigma_s,
                 const double * __restrict__ chi_p, const double * __restrict__ nu_sigma_f, const double * __restrict__ beta,
                 const double * __restrict__ chi_d, const double * __restrict__ lambda, const double * __restrict__ inv_v,
                 const double * __restrict__ dnp_cur, const double * __restrict__ phi_last, double * __restrict__ phi_tot,
                 double * __restrict__ phi_prev, double * __restrict__ q_out, double * __restrict__ phi_cur)
{
  const int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= hi)
    return;
  const int m = mat[i];
  const int nd = n_delayed[m];
  const bool has_chi = (mat_flags[m] & DOMA_MAT_HAS_CHI_P) != 0;
  phi_cur[i] = 0.0;
  double ph[GT > 0 ? GT : 1];
  if (GT > 0)
  {
#pragma unroll
    for (int gp = 0; gp < GT; ++gp)
      ph[gp] = phi_tot[gp * N + i];
  }
  double q = (mat_flags[m] & DOMA_MAT_HAS_SRC) ? f * src[m * G + g] / M_PI : 0.0;
  if (transient && step_type[m] != DOMA_STEP_NONE)
  {
    const double s = f * step_src[m * G + g] / M_PI;
    switch (step_type[m])
    {
      case DOMA_STEP_BOTH: q += (step_t_insert[m] <= t && t <= step_t_remove[m]) ? s : 0.0; break;
      case DOMA_STEP_INSERT: q += (step_t_insert[m] <= t) ? s : 0.0; break;
      case DOMA_STEP_REMOVE: q += (t <= step_t_remove[m]) ? s : 0.0; break;
      default: break;
    }
  }
#pragma unroll
  for (int gp = 0; gp < (GT > 0 ? GT : G); ++gp)
  {
    const double phi = GT > 0 ? ph[gp] : phi_tot[gp * N + i];
    // A term whose cross section is exactly zero adds +-0 and leaves q unchanged (for finite phi), so its
    // IEEE division is skipped: the common case for the up-scatter half of sigma_s and for chi_p = 0.
    const double ss = sigma_s[(m * G + g) * G + gp];
    if (gp != g && ss != 0.0)
      q += f * ss * phi / M_PI;
    const double chi = has_chi ? chi_p[m * G + g] : 0.0, nsf = has_chi ? nu_sigma_f[m * G + gp] : 0.0;
    if (chi == 0.0 || nsf == 0.0)
      ;
    else if (has_chi && nd == 0)
      q += f * chi_p[m * G + g] * nu_sigma_f[m * G + gp] * phi / M_PI;
    else if (has_chi && nd > 0)
    {
      double g_beta = 0.0;
      for (int d = 0; d < nd; ++d)
        g_beta += beta[(m * G + gp) * D + d];
      q += f * (1.0 - g_beta) * chi_p[m * G + g] * nu_sigma_f[m * G + gp] * phi / M_PI;
    }
    // these two terms sit inside the g' loop of the reference -> added G times
    if (nd > 0 && transient)
      for (int d = 0; d < nd; ++d)
        q += f * chi_d[(m * G + g) * D + d] * dnp_cur[d * N + i] * lambda[m * D + d] / M_PI;
    if (transient)
      q += f * phi_last[g * N + i] * inv_v[m * G + g] / dt / M_PI;
  }
  q_out[i] = q;
  phi_prev[g * N + i] = phi_tot[g * N + i];
  phi_tot[g * N + i] = 0.0;
}

// updateMultigroupSourceEigen(g) (3D:404-430): "+=" onto the stale q
template <int GT>
__global__ void
mg_source_eigen_kernel(int64_t N, int64_t lo, int64_t hi, int G, int g, double k, double f, const int32_t * __restrict__ mat,
                       const int32_t * __restrict__ mat_flags, const double * __restrict__ sigma_s,
                       const double * __restrict__ chi_p, const double * __restrict__ nu_sigma_f,
                       double * __restrict__ phi_tot, double * __restrict__ phi_prev, double * __restrict__ q_io,
                       double * __restrict__ phi_cur)
{
  const int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= hi)
    return;
  const int m = mat[i];
  const bool has_chi = (mat_flags[m] & DOMA_MAT_HAS_CHI_P) != 0;
  phi_cur[i] = 0.0;
  double ph[GT > 0 ? GT : 1];
  if (GT > 0)
  {
#pragma unroll
    for (int gp = 0; gp < GT; ++gp)
      ph[gp] = phi_tot[gp * N + i];
  }
  double q = q_io[i];
#pragma unroll
  for (int gp = 0; gp < (GT > 0 ? GT : G); ++gp)
  {
    const double phi = GT > 0 ? ph[gp] : phi_tot[gp * N + i];
    const double ss = sigma_s[(m * G + g) * G + gp]; // exact zeros skipped, see mg_source_kernel
    if (gp != g && ss != 0.0)
      q += f * ss * phi / M_PI;
    if (has_chi && chi_p[m * G + g] != 0.0 && nu_sigma_f[m * G + gp] != 0.0)
      q += f * chi_p[m * G + g] * nu_sigma_f[m * G + gp] * phi / M_PI / k;
  }
  q_io[i] = q;
  phi_prev[g * N + i] = phi_tot[g * N + i];
  phi_tot[g * N + i] = 0.0;
}

// updateScatteringSource(g) (3D:531-545)
__global__ void
scatter_source_kernel(int64_t N, int64_t lo, int64_t hi, int G, int g, double f, const int32_t * __restrict__ mat,
                      const double * __restrict__ sigma_s, double * __restrict__ phi_tot_g, double * __restrict__ q,
                      double * __restrict__ phi_cur)
{
  const int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= hi)
    return;
  const int m = mat[i];
  const double cur = phi_cur[i];
  phi_tot_g[i] += cur;
  q[i] = f * sigma_s[(m * G + g) * G + g] * cur / M_PI;
  phi_cur[i] = 0.0;
}

// total cross section seen by the cell solve for group g, expanded per cell so the sweep reads it
// without the material indirection: sigma_t, plus inv_v/dt for 2D transients (the 3D sweeps never
// pass mode/dt to the cell solve).  Run once per group at the first sweep (doma_launch_sig_eff).
__global__ void
sig_cell_kernel(int64_t N, int G, int g, int add_inv_v, double dt, const int32_t * __restrict__ mat,
                const double * __restrict__ sigma_t, const double * __restrict__ inv_v, double * __restrict__ sig_cell)
{
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N)
    return;
  const int m = mat[i];
  double tot = sigma_t[m * G + g];
  if (add_inv_v)
    tot += inv_v[m * G + g] / dt;
  sig_cell[i] = tot;
}

// ---- reductions: block partials in a fixed grid, then one block sums the partials in order ----
__device__ void
block_reduce2(double a, double b, double * __restrict__ partial, int slot)
{
  __shared__ double sa[EW_THREADS / 32], sb[EW_THREADS / 32];
  for (int o = 16; o > 0; o >>= 1)
  {
    a += __shfl_down_sync(0xffffffffu, a, o);
    b += __shfl_down_sync(0xffffffffu, b, o);
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0)
  {
    sa[w] = a;
    sb[w] = b;
  }
  __syncthreads();
  if (threadIdx.x == 0)
  {
    double ta = 0.0, tb = 0.0;
    for (int i = 0; i < EW_THREADS / 32; ++i)
    {
      ta += sa[i];
      tb += sb[i];
    }
    partial[slot] = ta;
    partial[RED_BLOCKS + slot] = tb;
  }
}

// computeScatteringResidual(g) (3D:547-563): sum V*cur^2, sum V*(tot+cur)^2
__global__ void
scatter_residual_kernel(int64_t lo, int64_t hi, Geo geo, const double * __restrict__ phi_tot_g, const double * __restrict__ phi_cur,
                        double * __restrict__ partial)
{
  double diff = 0.0, total = 0.0;
  const int64_t first = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
  CellWalk walk(geo, first < hi ? first : 0, stride);
  for (int64_t i = first; i < hi; i += stride, walk.next())
  {
    const double v = walk.measure();
    const double cur = phi_cur[i], s = phi_tot_g[i] + cur;
    diff = diff + (cur * cur) * v;
    total = total + (s * s) * v;
  }
  block_reduce2(diff, total, partial, blockIdx.x);
}

// computeScatteringResidual(g) immediately followed by updateScatteringSource(g), as sourceIteration always
// calls them (3D:475-477), in ONE pass over the cells: same per-cell expressions, same grid and reduction
// order as the two separate kernels, so residual and state are bit-identical to the unfused path.
// 16 B read + 24 B written per cell instead of 32 B + 24 B.
__global__ void
scatter_residual_update_kernel(int64_t lo, int64_t hi, int G, int g, double f, Geo geo, const int32_t * __restrict__ mat,
                               const double * __restrict__ sigma_s, double * __restrict__ phi_tot_g,
                               double * __restrict__ phi_cur, double * __restrict__ q, double * __restrict__ partial,
                               const int * __restrict__ stop)
{
  if (stop && *stop)
    return; // device-side source iteration: already converged
  double diff = 0.0, total = 0.0;
  const int64_t first = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
  CellWalk walk(geo, first < hi ? first : 0, stride);
  for (int64_t i = first; i < hi; i += stride, walk.next())
  {
    const double v = walk.measure();
    const double cur = phi_cur[i], s = phi_tot_g[i] + cur;
    diff = diff + (cur * cur) * v;
    total = total + (s * s) * v;
    phi_tot_g[i] = s;
    q[i] = f * sigma_s[(mat[i] * G + g) * G + g] * cur / M_PI;
    phi_cur[i] = 0.0;
  }
  block_reduce2(diff, total, partial, blockIdx.x);
}

// computeMGFluxResidual() (3D:318-337)
__global__ void
mg_residual_kernel(int64_t N, int64_t lo, int64_t hi, int G, Geo geo, const double * __restrict__ phi_tot, const double * __restrict__ phi_prev,
                   double * __restrict__ partial)
{
  double diff = 0.0, total = 0.0;
  const int64_t first = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
  CellWalk walk(geo, first < hi ? first : 0, stride);
  for (int64_t i = first; i < hi; i += stride, walk.next())
  {
    const double v = walk.measure();
    for (int g = 0; g < G; ++g)
    {
      const double s = phi_tot[g * N + i], d = s - phi_prev[g * N + i];
      diff = diff + (d * d) * v;
      total = total + (s * s) * v;
    }
  }
  block_reduce2(diff, total, partial, blockIdx.x);
}

// updateEigenvalue() (3D:432-456): num = sum V nuSf phi_tot, den = sum V nuSf phi_prev / k
__global__ void
k_sums_kernel(int64_t N, int64_t lo, int64_t hi, int G, double k, Geo geo, const int32_t * __restrict__ mat, const int32_t * __restrict__ mat_flags,
              const double * __restrict__ nu_sigma_f, const double * __restrict__ phi_tot,
              const double * __restrict__ phi_prev, double * __restrict__ partial)
{
  double num = 0.0, den = 0.0;
  const int64_t first = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
  CellWalk walk(geo, first < hi ? first : 0, stride);
  for (int64_t i = first; i < hi; i += stride, walk.next())
  {
    const int m = mat[i];
    if (!(mat_flags[m] & DOMA_MAT_HAS_NU_SIGMA_F))
      continue;
    const double v = walk.measure();
    for (int g = 0; g < G; ++g)
    {
      num = num + v * nu_sigma_f[m * G + g] * phi_tot[g * N + i];
      den = den + v * nu_sigma_f[m * G + g] * phi_prev[g * N + i] / k;
    }
  }
  block_reduce2(num, den, partial, blockIdx.x);
}

__global__ void
publish_pair_kernel(const double * __restrict__ out, double * __restrict__ out_host)
{
  out_host[0] = out[0];
  out_host[1] = out[1];
  __threadfence_system();
}

__global__ void
final_reduce_kernel(const double * __restrict__ partial, double * __restrict__ out, double * __restrict__ out_host)
{
  // 2 warps: warp 0 sums the first RED_BLOCKS partials, warp 1 the second; lane-strided then a
  // fixed shuffle tree -> deterministic
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double a = 0.0;
  for (int i = lane; i < RED_BLOCKS; i += 32)
    a += partial[w * RED_BLOCKS + i];
  for (int o = 16; o > 0; o >>= 1)
    a += __shfl_down_sync(0xffffffffu, a, o);
  if (lane == 0)
  {
    out[w] = a;
    // the pair also lands in pinned host memory straight from the kernel (zero-copy store): a 16-byte
    // cudaMemcpyAsync would queue on the D2H copy engine behind a 134 MB result transfer of an
    // asynchronous host-buffer sweep and cost the stream 2.5 ms (measured, scripts/e2e_probe.py)
    if (out_host)
    {
      out_host[w] = a;
      __threadfence_system();
    }
  }
}

// Closing kernel of one device-side source iteration (sourceIteration, 3D:464-481): the two sums -> the residual exactly
// as the host path forms it (3D:562: total > 1e-8 ? sqrt(diff)/sqrt(total) : 0; IEEE sqrt and division, this file is
// compiled without FMA contraction), appended to the history; the loop condition (3D:481) decides whether the kernels
// already enqueued for the next iterations still have work.
__global__ void
si_close_kernel(const double * __restrict__ partial, SiState * __restrict__ st, SiState * __restrict__ host, double tol, int max_it)
{
  __shared__ double sums[2];
  if (st->stop)
    return;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double a = 0.0;
  for (int i = lane; i < RED_BLOCKS; i += 32)
    a += partial[w * RED_BLOCKS + i];
  for (int o = 16; o > 0; o >>= 1)
    a += __shfl_down_sync(0xffffffffu, a, o);
  if (lane == 0)
    sums[w] = a;
  __syncthreads();
  if (threadIdx.x == 0)
  {
    const double res = sums[1] > 1e-8 ? sqrt(sums[0]) / sqrt(sums[1]) : 0.0;
    const int it = st->it;
    const int stop = !(it + 1 < max_it && tol < res);
    st->res = res;
    st->it = it + 1;
    st->stop = stop;
    // mirror for the host (pinned, mapped): what it reads after the batch's synchronisation
    if (it < DOMA_SI_HIST)
      host->hist[it] = res;
    host->res = res;
    host->it = it + 1;
    host->stop = stop;
    __threadfence_system();
  }
}

int
finish_reduce(doma_b200_ctx * c)
{
  // cell-sharded ranks hold partial sums of their own cells: sum the pair over the ranks first (every rank gets the
  // same bits, so every rank takes the same branch in the loops), then publish
  final_reduce_kernel<<<1, 64, 0, c->stream>>>(c->red_partial, c->red_out, c->cell_sharded ? nullptr : c->h_red_dev);
  c->launches += 2;
  DOMA_CUDA(c, cudaGetLastError());
  if (c->cell_sharded)
  {
    const int rc = doma_comm_sum_pair(c);
    if (rc)
      return rc;
    publish_pair_kernel<<<1, 1, 0, c->stream>>>(c->red_out, c->h_red_dev);
    c->launches += 1;
    DOMA_CUDA(c, cudaGetLastError());
  }
  DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}
} // namespace

int
doma_launch_reset_state(doma_b200_ctx * c, double phi_init)
{
  const int64_t GN = (int64_t)c->G * c->N;
  fill_kernel<<<ew_blocks(GN), EW_THREADS, 0, c->stream>>>(c->phi_tot, GN, phi_init);
  c->launches += 1;
  DOMA_CUDA(c, cudaGetLastError());
  DOMA_CUDA(c, cudaMemsetAsync(c->phi_prev, 0, sizeof(double) * GN, c->stream));
  DOMA_CUDA(c, cudaMemsetAsync(c->q, 0, sizeof(double) * c->N, c->stream));
  DOMA_CUDA(c, cudaMemsetAsync(c->phi_cur, 0, sizeof(double) * c->N, c->stream));
  if (c->cell_sharded)
  {
    DOMA_CUDA(c, cudaMemsetAsync(c->phi_red, 0, sizeof(double) * c->slice, c->stream));
    c->q_slice_only = false;
    c->phi_cur_reduced = true;
  }
  return 0;
}

int
doma_launch_stage_in(doma_b200_ctx * c)
{
  const int64_t lo = ew_lo(c), hi = ew_hi(c);
  stage_in_kernel<<<ew_blocks(hi - lo), EW_THREADS, 0, c->stream>>>(c->q, c->q_stage, ew_phi_cur(c), lo, hi);
  c->launches += 1;
  DOMA_CUDA(c, cudaGetLastError());
  if (c->cell_sharded)
  {
    if (!c->phi_cur_reduced) // a sweep without its reduction left partial sums in the local accumulator
      DOMA_CUDA(c, cudaMemsetAsync(c->phi_cur, 0, sizeof(double) * c->N, c->stream));
    c->q_slice_only = true;
    c->phi_cur_reduced = true;
  }
  return 0;
}

int
doma_launch_snapshot_phi_cur(doma_b200_ctx * c)
{
  // cell-sharded: the summed increment of the own cells (the host buffer receives that slice)
  const int64_t lo = ew_lo(c), hi = ew_hi(c);
  copy_kernel<<<ew_blocks(hi - lo), EW_THREADS, 0, c->stream>>>(c->phi_snap + lo, ew_phi_cur(c) + lo, hi - lo);
  c->launches += 1;
  DOMA_CUDA(c, cudaGetLastError());
  return 0;
}

int
doma_launch_init_zero_ic(doma_b200_ctx * c)
{
  const int64_t GN = (int64_t)c->G * c->N;
  DOMA_CUDA(c, cudaMemsetAsync(c->phi_tot, 0, sizeof(double) * GN, c->stream));
  DOMA_CUDA(c, cudaMemsetAsync(c->phi_prev, 0, sizeof(double) * GN, c->stream));
  if (c->phi_last)
    DOMA_CUDA(c, cudaMemsetAsync(c->phi_last, 0, sizeof(double) * GN, c->stream));
  if (c->dnp_cur)
  {
    DOMA_CUDA(c, cudaMemsetAsync(c->dnp_cur, 0, sizeof(double) * c->D * c->N, c->stream));
    DOMA_CUDA(c, cudaMemsetAsync(c->dnp_last, 0, sizeof(double) * c->D * c->N, c->stream));
  }
  return 0;
}

int
doma_launch_update_step_fluxes(doma_b200_ctx * c)
{
  const int64_t GN = (int64_t)c->G * c->N;
  if (!c->phi_last)
  {
    snprintf(c->err, DOMA_MAX_ERR, "update_step_fluxes: context was not created in transient mode");
    return 5;
  }
  DOMA_CUDA(c, cudaMemcpyAsync(c->phi_last, c->phi_tot, sizeof(double) * GN, cudaMemcpyDeviceToDevice, c->stream));
  DOMA_CUDA(c, cudaMemsetAsync(c->phi_tot, 0, sizeof(double) * GN, c->stream));
  DOMA_CUDA(c, cudaMemsetAsync(c->q, 0, sizeof(double) * c->N, c->stream));
  DOMA_CUDA(c, cudaMemsetAsync(c->phi_cur, 0, sizeof(double) * c->N, c->stream));
  if (c->cell_sharded)
  {
    DOMA_CUDA(c, cudaMemsetAsync(c->phi_red, 0, sizeof(double) * c->slice, c->stream));
    c->q_slice_only = false;
    c->phi_cur_reduced = true;
  }
  return 0;
}

int
doma_launch_step_dnps(doma_b200_ctx * c)
{
  if (c->D == 0 || !c->dnp_cur)
    return 0;
  step_dnps_kernel<<<ew_blocks(ew_hi(c) - ew_lo(c)), EW_THREADS, 0, c->stream>>>(c->N, ew_lo(c), ew_hi(c), c->G, c->D, c->dt, c->mat, c->n_delayed, c->nu_sigma_f,
                                                                  c->beta, c->lambda, c->phi_tot, c->dnp_cur, c->dnp_last);
  c->launches += 1;
  DOMA_CUDA(c, cudaGetLastError());
  return 0;
}

int
doma_launch_update_mg_source(doma_b200_ctx * c, int g, double t)
{
  const double f = c->dims == 3 ? 0.25 : 0.5;
  const int transient = c->mode == DOMA_MODE_TRANSIENT;
#define MGS(GT)                                                                                                          \
  mg_source_kernel<GT><<<ew_blocks(ew_hi(c) - ew_lo(c)), EW_THREADS, 0, c->stream>>>(                                    \
      c->N, ew_lo(c), ew_hi(c), c->G, c->D, g, t, f, transient, c->dt, c->mat, c->mat_flags, c->n_delayed, c->step_type, c->step_t_insert, \
      c->step_t_remove, c->step_src, c->src, c->sigma_s, c->chi_p, c->nu_sigma_f, c->beta, c->chi_d, c->lambda,        \
      c->inv_v, c->dnp_cur, c->phi_last, c->phi_tot, c->phi_prev, c->q, ew_phi_cur(c))
  switch (c->G)
  {
    case 1: MGS(1); break;
    case 2: MGS(2); break;
    case 3: MGS(3); break;
    case 4: MGS(4); break;
    case 5: MGS(5); break;
    case 6: MGS(6); break;
    case 7: MGS(7); break;
    case 8: MGS(8); break;
    default: MGS(0); break;
  }
#undef MGS
  c->launches += 1;
  DOMA_CUDA(c, cudaGetLastError());
  c->q_slice_only = c->cell_sharded;
  return 0;
}

int
doma_launch_update_mg_source_eigen(doma_b200_ctx * c, int g, double k)
{
  const double f = c->dims == 3 ? 0.25 : 0.5;
#define MGE(GT)                                                                                                    \
  mg_source_eigen_kernel<GT><<<ew_blocks(ew_hi(c) - ew_lo(c)), EW_THREADS, 0, c->stream>>>(                        \
      c->N, ew_lo(c), ew_hi(c), c->G, g, k, f, c->mat, c->mat_flags, c->sigma_s, c->chi_p, c->nu_sigma_f, c->phi_tot,  \
      c->phi_prev, c->q, ew_phi_cur(c))
  switch (c->G)
  {
    case 1: MGE(1); break;
    case 2: MGE(2); break;
    case 3: MGE(3); break;
    case 4: MGE(4); break;
    case 5: MGE(5); break;
    case 6: MGE(6); break;
    case 7: MGE(7); break;
    case 8: MGE(8); break;
    default: MGE(0); break;
  }
#undef MGE
  c->launches += 1;
  DOMA_CUDA(c, cudaGetLastError());
  c->q_slice_only = c->cell_sharded;
  return 0;
}

int
doma_launch_update_scatter_source(doma_b200_ctx * c, int g)
{
  const double f = c->dims == 3 ? 0.25 : 0.5;
  scatter_source_kernel<<<ew_blocks(ew_hi(c) - ew_lo(c)), EW_THREADS, 0, c->stream>>>(
      c->N, ew_lo(c), ew_hi(c), c->G, g, f, c->mat, c->sigma_s, c->phi_tot + (int64_t)g * c->N, c->q, ew_phi_cur(c));
  c->launches += 1;
  DOMA_CUDA(c, cudaGetLastError());
  c->q_slice_only = c->cell_sharded;
  return 0;
}

int
doma_launch_scatter_residual(doma_b200_ctx * c, int g)
{
  scatter_residual_kernel<<<RED_BLOCKS, EW_THREADS, 0, c->stream>>>(ew_lo(c), ew_hi(c), geo_of(c), c->phi_tot + (int64_t)g * c->N,
                                                                    ew_phi_cur(c), c->red_partial);
  DOMA_CUDA(c, cudaGetLastError());
  return finish_reduce(c);
}

int
doma_launch_scatter_residual_update(doma_b200_ctx * c, int g)
{
  const double f = c->dims == 3 ? 0.25 : 0.5;
  scatter_residual_update_kernel<<<RED_BLOCKS, EW_THREADS, 0, c->stream>>>(ew_lo(c), ew_hi(c), c->G, g, f, geo_of(c), c->mat,
                                                                           c->sigma_s, c->phi_tot + (int64_t)g * c->N,
                                                                           ew_phi_cur(c), c->q, c->red_partial, nullptr);
  DOMA_CUDA(c, cudaGetLastError());
  c->q_slice_only = c->cell_sharded;
  return finish_reduce(c);
}

int
doma_launch_si_iteration_tail(doma_b200_ctx * c, int g, double tol, int max_it)
{
  const double f = c->dims == 3 ? 0.25 : 0.5;
  scatter_residual_update_kernel<<<RED_BLOCKS, EW_THREADS, 0, c->stream>>>(0, c->N, c->G, g, f, geo_of(c), c->mat, c->sigma_s,
                                                                           c->phi_tot + (int64_t)g * c->N, c->phi_cur, c->q,
                                                                           c->red_partial, &c->si_dev->stop);
  DOMA_CUDA(c, cudaGetLastError());
  si_close_kernel<<<1, 64, 0, c->stream>>>(c->red_partial, c->si_dev, c->si_host_dev, tol, max_it);
  DOMA_CUDA(c, cudaGetLastError());
  c->launches += 2;
  return 0;
}

int
doma_launch_mg_residual(doma_b200_ctx * c)
{
  mg_residual_kernel<<<RED_BLOCKS, EW_THREADS, 0, c->stream>>>(c->N, ew_lo(c), ew_hi(c), c->G, geo_of(c), c->phi_tot, c->phi_prev,
                                                               c->red_partial);
  DOMA_CUDA(c, cudaGetLastError());
  return finish_reduce(c);
}

int
doma_launch_k_sums(doma_b200_ctx * c, double k)
{
  k_sums_kernel<<<RED_BLOCKS, EW_THREADS, 0, c->stream>>>(c->N, ew_lo(c), ew_hi(c), c->G, k, geo_of(c), c->mat, c->mat_flags,
                                                          c->nu_sigma_f, c->phi_tot, c->phi_prev, c->red_partial);
  DOMA_CUDA(c, cudaGetLastError());
  return finish_reduce(c);
}

// The total cross section the cell solve uses, per cell and per group, is static (materials, and dt for 2D transients,
// are fixed at create): expanded ONCE for all groups, so a sweep only picks its group's array.  G x N doubles of HBM
// (1 GB at 256^3 x 8 groups) instead of a 12-byte-per-cell pass in front of every sweep of a new group.
int
doma_launch_sig_eff(doma_b200_ctx * c, int g)
{
  (void)g;
  if (c->sig_cell_group == c->G)
    return 0; // built
  const int add = (c->dims == 2 && c->mode == DOMA_MODE_TRANSIENT) ? 1 : 0;
  for (int gg = 0; gg < c->G; ++gg)
  {
    sig_cell_kernel<<<ew_blocks(c->N), EW_THREADS, 0, c->stream>>>(c->N, c->G, gg, add, c->dt, c->mat, c->sigma_t, c->inv_v,
                                                                   c->sig_cell + (int64_t)gg * c->N);
    c->launches += 1;
    DOMA_CUDA(c, cudaGetLastError());
  }
  c->sig_cell_group = c->G;
  return 0;
}
