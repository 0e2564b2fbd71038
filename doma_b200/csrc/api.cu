// api.cu — the C ABI of include/doma_b200.h: context lifetime, host<->device state, reductions'
// host side, the NCCL phi all-reduce.  No torch types, plain pointers and sizes.
#include "ctx.h"

#include <dlfcn.h>
#include <math.h>
#include <nvtx3/nvToolsExt.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

namespace
{
char g_create_err[DOMA_MAX_ERR] = "";

template <typename T>
int
upload(doma_b200_ctx * c, T ** dst, const T * src, size_t n, T fill = T())
{
  *dst = nullptr;
  const size_t cnt = n ? n : 1;
  DOMA_CUDA(c, cudaMalloc((void **)dst, sizeof(T) * cnt));
  c->device_bytes += sizeof(T) * cnt;
  std::vector<T> tmp(cnt, fill);
  if (src && n)
    memcpy(tmp.data(), src, sizeof(T) * n);
  DOMA_CUDA(c, cudaMemcpy(*dst, tmp.data(), sizeof(T) * cnt, cudaMemcpyHostToDevice));
  return 0;
}

int
alloc_zero(doma_b200_ctx * c, double ** dst, size_t n)
{
  *dst = nullptr;
  const size_t cnt = n ? n : 1;
  // DOMA_PAD doubles behind every per-cell array: the cell-sharded collectives work on nranks equal slices
  DOMA_CUDA(c, cudaMalloc((void **)dst, sizeof(double) * (cnt + DOMA_PAD)));
  DOMA_CUDA(c, cudaMemset(*dst, 0, sizeof(double) * (cnt + DOMA_PAD)));
  c->device_bytes += sizeof(double) * (cnt + DOMA_PAD);
  return 0;
}

template <typename T>
void
free_dev(T *& p)
{
  if (p)
    cudaFree(p);
  p = nullptr;
}
} // namespace

// Regroup this context's ordinates into sweep units: per octant, ceil(count/32) units of near-equal size; lanes
// beyond a unit's ordinates are padded with a harmless direction and weight 0.  Also (re)allocates the face planes
// and rebuilds the tile table.  The new buffers are allocated first and the context is only changed once
// everything has succeeded, so a failure (out of memory) leaves the previous decomposition usable.
int
doma_build_units(doma_b200_ctx * c)
{
  DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  const int L = DOMA_LANES;
  std::vector<int> uo;
  std::vector<double> mu, eta, xi, w;
  std::vector<SweepUnit> units;
  // ordinates of every octant, in the context's order
  std::vector<int> by_oct[8];
  for (int n = 0; n < c->n_ord; ++n)
    by_oct[c->h_ord_octant[n]].push_back(n);
  size_t most = 0;
  for (int oct = 0; oct < 8; ++oct)
    most = std::max(most, by_oct[oct].size());
  auto push_lane = [&](int n) {
    if (n >= 0)
    {
      // parallelSweepOctant takes std::abs of each component (src/TransportSolver3D.cpp:645-649)
      mu.push_back(fabs(c->h_mu[n]));
      eta.push_back(fabs(c->h_eta[n]));
      xi.push_back(fabs(c->h_xi[n]));
      w.push_back(c->h_w[n]);
    }
    else
    {
      mu.push_back(1.0);
      eta.push_back(1.0);
      xi.push_back(1.0);
      w.push_back(0.0);
    }
  };
  // <= 16 ordinates per octant (the shipped assembly cases): two octants that differ only in the x direction share a
  // warp, 16 lanes each (sweep.cu, PAIR) - otherwise every warp would run half empty
  const bool paired = c->pairing_enabled && most > 0 && most <= DOMA_LANES / 2;
  if (paired)
  {
    for (int oct = 0; oct < 4; ++oct) // bit 2 clear: x ascending; its partner is oct | 4
    {
      const std::vector<int> &ia = by_oct[oct], &ib = by_oct[oct | 4];
      if (ia.empty() && ib.empty())
        continue;
      units.push_back(SweepUnit{oct, (int)(ia.size() + ib.size())});
      uo.push_back(oct);
      for (int l = 0; l < DOMA_LANES / 2; ++l)
        push_lane(l < (int)ia.size() ? ia[l] : -1);
      for (int l = 0; l < DOMA_LANES / 2; ++l)
        push_lane(l < (int)ib.size() ? ib[l] : -1);
    }
  }
  else
    for (int oct = 0; oct < 8; ++oct)
    {
      const std::vector<int> & idx = by_oct[oct];
      if (idx.empty())
        continue;
      const int cnt = (int)idx.size();
      const int nb = (cnt + L - 1) / L;
      int pos = 0;
      for (int b = 0; b < nb; ++b)
      {
        const int sz = cnt / nb + (b < cnt % nb ? 1 : 0);
        units.push_back(SweepUnit{oct, sz});
        uo.push_back(oct);
        for (int l = 0; l < DOMA_LANES; ++l)
          push_lane(l < sz ? idx[pos + l] : -1);
        pos += sz;
      }
    }
  // ---- new device buffers (nothing of the context is touched until all of them exist) ----
  const size_t U = units.size();
  int * d_uo = nullptr;
  double *d_mu = nullptr, *d_eta = nullptr, *d_xi = nullptr, *d_w = nullptr, *fx = nullptr, *fy = nullptr, *fz = nullptr;
  const int64_t bytes_before = c->device_bytes;
  // a face slot is 16 bytes: the LL protocol stores {low half, tag, high half, tag}; the flag protocol uses the first 8
  const size_t slot = 2;
  const size_t nfx = U * c->nz * c->ny * L * slot, nfy = U * c->nz * c->nx * L * slot, nfz = c->dims == 3 ? U * c->ny * c->nx * L * slot : 0;
  auto fail = [&](int rc) {
    free_dev(d_uo), free_dev(d_mu), free_dev(d_eta), free_dev(d_xi), free_dev(d_w), free_dev(fx), free_dev(fy), free_dev(fz);
    c->device_bytes = bytes_before;
    return rc;
  };
  int rc = 0;
  if ((rc = upload(c, &d_uo, uo.data(), uo.size())) || (rc = upload(c, &d_mu, mu.data(), mu.size())) ||
      (rc = upload(c, &d_eta, eta.data(), eta.size())) || (rc = upload(c, &d_xi, xi.data(), xi.size())) ||
      (rc = upload(c, &d_w, w.data(), w.size())))
    return fail(rc);
  if (U)
  {
    cudaError_t e = cudaMalloc(&fx, sizeof(double) * nfx);
    if (e == cudaSuccess)
      e = cudaMalloc(&fy, sizeof(double) * nfy);
    if (e == cudaSuccess && nfz)
      e = cudaMalloc(&fz, sizeof(double) * nfz);
    if (e == cudaSuccess)
      e = cudaMemset(fx, 0, sizeof(double) * nfx);
    if (e == cudaSuccess)
      e = cudaMemset(fy, 0, sizeof(double) * nfy);
    if (e == cudaSuccess && nfz)
      e = cudaMemset(fz, 0, sizeof(double) * nfz);
    if (e != cudaSuccess)
    {
      snprintf(c->err, DOMA_MAX_ERR, "build_units: face planes (%zu bytes): %s", sizeof(double) * (nfx + nfy + nfz), cudaGetErrorString(e));
      return fail(100 + (int)e);
    }
  }
  // ---- commit ----
  free_dev(c->unit_octant), free_dev(c->u_mu), free_dev(c->u_eta), free_dev(c->u_xi), free_dev(c->u_w);
  free_dev(c->face_x), free_dev(c->face_y), free_dev(c->face_z);
  c->device_bytes = bytes_before - c->face_bytes + (int64_t)(sizeof(double) * (nfx + nfy + nfz));
  c->face_bytes = (int64_t)(sizeof(double) * (nfx + nfy + nfz));
  c->unit_octant = d_uo, c->u_mu = d_mu, c->u_eta = d_eta, c->u_xi = d_xi, c->u_w = d_w;
  c->face_x = fx, c->face_y = fy, c->face_z = fz;
  c->units_paired = paired;
  free(c->h_units);
  c->n_units = (int)U;
  c->h_units = (SweepUnit *)malloc(sizeof(SweepUnit) * (U + 1));
  memcpy(c->h_units, units.data(), sizeof(SweepUnit) * U);
  c->ll_epoch = 0; // the planes are zero again: LL tags restart
  if (!c->tile_auto && c->tile_y > 0 && c->tile_z > 0)
    return doma_sweep_setup_tiles(c, c->tile_x, c->tile_y, c->dims == 3 ? c->tile_z : 1); // keep a tile the caller chose
  return doma_choose_tile(c);
}

namespace
{
int
create_impl(doma_b200_ctx * c, const doma_b200_problem * p)
{
  int n_dev = 0;
  DOMA_CUDA(c, cudaGetDeviceCount(&n_dev));
  if (n_dev < 1 || p->device < 0 || p->device >= n_dev)
  {
    snprintf(c->err, DOMA_MAX_ERR, "create: CUDA device %d not available (%d devices); there is no CPU fallback", p->device, n_dev);
    return 2;
  }
  c->device = p->device;
  DOMA_CUDA(c, cudaSetDevice(c->device));
  cudaDeviceProp prop;
  DOMA_CUDA(c, cudaGetDeviceProperties(&prop, c->device));
  c->n_sms = prop.multiProcessorCount;
  c->dims = p->dims;
  c->nx = p->nx, c->ny = p->ny, c->nz = p->dims == 3 ? p->nz : 1;
  c->N = (int64_t)c->nx * c->ny * c->nz;
  c->G = p->n_groups, c->M = p->n_mat, c->D = p->n_delayed_max > 0 ? p->n_delayed_max : 0;
  c->n_ord = p->n_ord;
  c->scheme = p->scheme, c->mode = p->mode, c->dt = p->dt;
  c->pairing_enabled = true;
  c->rows2 = -1;
  if (const char * r2 = getenv("DOMA_B200_ROWS2")) // experiments: 0 / 1 force the one-row / two-row kernel
    c->rows2 = atoi(r2);
  if (const char * mb = getenv("DOMA_B200_MAX_BLOCKS")) // experiments only (scripts/small_case_grid.py)
    c->max_blocks = atoi(mb);
  if (const char * ct = getenv("DOMA_B200_COMM_TIMING")) // measurement mode: per-collective device times, printed at destroy
    c->comm_timing = *ct && *ct != '0';
  if (c->comm_timing)
    cudaMalloc(&c->comm_scratch, 16);
  c->variant = DOMA_SWEEP_AUTO;
  {
    const char * chk = getenv("DOMA_B200_CHECK"); // documented in include/doma_b200.h (doma_b200_set_check)
    c->check = chk && *chk && *chk != '0';
  }
  DOMA_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  c->own_stream = true;
  DOMA_CUDA(c, cudaEventCreate(&c->ev0));
  DOMA_CUDA(c, cudaEventCreate(&c->ev1));

  for (int64_t i = 0; i < c->N; ++i)
    if (p->mat_of_cell[i] < 0 || p->mat_of_cell[i] >= c->M)
    {
      snprintf(c->err, DOMA_MAX_ERR, "create: cell %lld has material %d outside [0,%d)", (long long)i, p->mat_of_cell[i], c->M);
      return 3;
    }
  const size_t G = c->G, M = c->M, D = c->D;
  int rc = 0;
  std::vector<double> one(1, 1.0);
  const double * dz_src = p->dims == 3 ? p->dz : one.data();
#define UP(dst, src, n)                      \
  if ((rc = upload(c, &c->dst, src, (n)))) \
    return rc;
  UP(dx, p->dx, (size_t)c->nx);
  UP(dy, p->dy, (size_t)c->ny);
  UP(dz, dz_src, (size_t)c->nz);
  {
    std::vector<double> r;
    r.resize(c->nx);
    for (int i = 0; i < c->nx; ++i)
      r[i] = 1.0 / p->dx[i];
    UP(rdx, r.data(), r.size());
    r.resize(c->ny);
    for (int i = 0; i < c->ny; ++i)
      r[i] = 1.0 / p->dy[i];
    UP(rdy, r.data(), r.size());
    r.resize(c->nz);
    for (int i = 0; i < c->nz; ++i)
      r[i] = 1.0 / dz_src[i];
    UP(rdz, r.data(), r.size());
  }
  UP(mat, p->mat_of_cell, (size_t)c->N);
  UP(mat_flags, p->mat_flags, M);
  UP(n_delayed, p->n_delayed, M);
  if (p->step_type)
  {
    UP(step_type, p->step_type, M);
  }
  else if ((rc = upload<int32_t>(c, &c->step_type, nullptr, M, DOMA_STEP_NONE)))
    return rc;
  UP(sigma_t, p->sigma_t, M * G);
  UP(sigma_s, p->sigma_s, M * G * G);
  UP(nu_sigma_f, p->nu_sigma_f, M * G);
  UP(chi_p, p->chi_p, M * G);
  UP(src, p->src, M * G);
  UP(inv_v, p->inv_v, M * G);
  UP(chi_d, D ? p->chi_d : nullptr, M * G * D);
  UP(beta, D ? p->beta : nullptr, M * G * D);
  UP(lambda, D ? p->lambda : nullptr, M * D);
  UP(step_t_insert, p->step_t_insert, M);
  UP(step_t_remove, p->step_t_remove, M);
  UP(step_src, p->step_src, M * G);
#undef UP
  if ((rc = alloc_zero(c, &c->sig_cell, (size_t)c->N * c->G)))
    return rc;
  c->sig_cell_group = -1;
  const size_t N = (size_t)c->N;
  if ((rc = alloc_zero(c, &c->phi_tot, G * N)) || (rc = alloc_zero(c, &c->phi_prev, G * N)) || (rc = alloc_zero(c, &c->q, N)) ||
      (rc = alloc_zero(c, &c->phi_cur, N)))
    return rc;
  if (c->mode == DOMA_MODE_TRANSIENT)
  {
    if ((rc = alloc_zero(c, &c->phi_last, G * N)))
      return rc;
    if (D && ((rc = alloc_zero(c, &c->dnp_cur, D * N)) || (rc = alloc_zero(c, &c->dnp_last, D * N))))
      return rc;
  }
  // host copy of the ordinates (the sweep variant can be changed after create)
  {
    const size_t n = (size_t)(c->n_ord > 0 ? c->n_ord : 1);
    c->h_ord_octant = (int32_t *)calloc(n, sizeof(int32_t));
    c->h_mu = (double *)calloc(n, sizeof(double)), c->h_eta = (double *)calloc(n, sizeof(double));
    c->h_xi = (double *)calloc(n, sizeof(double)), c->h_w = (double *)calloc(n, sizeof(double));
    int per_oct[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < c->n_ord; ++i)
    {
      const int o = p->ord_octant[i];
      if (o < 0 || o > 7 || (c->dims == 2 && (o & 1)))
      {
        snprintf(c->err, DOMA_MAX_ERR, "create: ordinate %d has octant %d%s", i, o,
                 c->dims == 2 ? "; 2D problems sweep only the xi>0 octants (PPP, PMP, MPP, MMP)" : "");
        return 9;
      }
      c->h_ord_octant[i] = o, c->h_mu[i] = p->mu[i], c->h_eta[i] = p->eta[i], c->h_xi[i] = p->xi[i], c->h_w[i] = p->w[i];
      per_oct[o] += 1;
    }
    (void)per_oct;
    DOMA_CUDA(c, cudaMalloc(&c->work_counter, sizeof(int)));
    c->tile_auto = true;
    if ((rc = doma_build_units(c)))
      return rc;
  }
  DOMA_CUDA(c, cudaMalloc(&c->red_partial, sizeof(double) * 2 * 1184));
  DOMA_CUDA(c, cudaMalloc(&c->red_out, sizeof(double) * 2));
  DOMA_CUDA(c, cudaHostAlloc(&c->h_red, sizeof(double) * 2, cudaHostAllocMapped));
  DOMA_CUDA(c, cudaHostGetDevicePointer(&c->h_red_dev, c->h_red, 0));
  DOMA_CUDA(c, cudaHostAlloc(&c->si, sizeof(SiState), cudaHostAllocMapped));
  DOMA_CUDA(c, cudaHostGetDevicePointer(&c->si_host_dev, c->si, 0));
  memset(c->si, 0, sizeof(SiState));
  DOMA_CUDA(c, cudaMalloc(&c->si_dev, sizeof(SiState)));
  return 0;
}

// ---- NCCL through dlopen: no link-time dependency, the torch-bundled libnccl.so.2 is reused ----
struct NcclApi
{
  void * lib;
  int (*GetUniqueId)(void *);
  int (*CommInitRank)(void **, int, char[128], int); // ncclUniqueId is a 128-byte struct passed by value
  int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t);
  int (*ReduceScatter)(const void *, void *, size_t, int, int, void *, cudaStream_t);
  int (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t);
  int (*CommDestroy)(void *);
  const char * (*GetErrorString)(int);
};
NcclApi g_nccl = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};

struct NcclId
{
  char b[128];
};

int
load_nccl(char * err)
{
  if (g_nccl.lib)
    return 0;
  const char * names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char * n : names)
    if ((g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL)))
      break;
  if (!g_nccl.lib)
  {
    snprintf(err, DOMA_MAX_ERR, "comm: cannot dlopen libnccl.so.2: %s", dlerror());
    return 20;
  }
  g_nccl.GetUniqueId = (int (*)(void *))dlsym(g_nccl.lib, "ncclGetUniqueId");
  g_nccl.CommInitRank = (int (*)(void **, int, char[128], int))dlsym(g_nccl.lib, "ncclCommInitRank");
  g_nccl.AllReduce = (int (*)(const void *, void *, size_t, int, int, void *, cudaStream_t))dlsym(g_nccl.lib, "ncclAllReduce");
  g_nccl.ReduceScatter = (int (*)(const void *, void *, size_t, int, int, void *, cudaStream_t))dlsym(g_nccl.lib, "ncclReduceScatter");
  g_nccl.AllGather = (int (*)(const void *, void *, size_t, int, void *, cudaStream_t))dlsym(g_nccl.lib, "ncclAllGather");
  g_nccl.CommDestroy = (int (*)(void *))dlsym(g_nccl.lib, "ncclCommDestroy");
  g_nccl.GetErrorString = (const char * (*)(int))dlsym(g_nccl.lib, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.ReduceScatter || !g_nccl.AllGather || !g_nccl.CommDestroy)
  {
    snprintf(err, DOMA_MAX_ERR, "comm: libnccl lacks a required symbol");
    return 21;
  }
  return 0;
}
} // namespace

extern "C" {

int
doma_b200_create(const doma_b200_problem * p, doma_b200_ctx ** out)
{
  // every pointer the create path dereferences is checked here: a bad caller gets rc != 0 and a message, not a fault
  const char * missing = nullptr;
  if (!p || !out)
    missing = "problem / out pointer";
  else if (p->dims != 2 && p->dims != 3)
    missing = "dims (2 or 3)";
  else if (p->nx < 1 || p->ny < 1 || (p->dims == 3 && p->nz < 1))
    missing = "nx, ny, nz (>= 1)";
  else if (p->n_groups < 1 || p->n_mat < 1 || p->n_ord < 0 || p->n_delayed_max < 0)
    missing = "n_groups, n_mat (>= 1), n_ord, n_delayed_max (>= 0)";
  else if ((int64_t)p->n_mat * p->n_groups * p->n_groups > INT32_MAX || (int64_t)p->n_mat * p->n_groups * (p->n_delayed_max + 1) > INT32_MAX)
    missing = "n_mat * n_groups^2 and n_mat * n_groups * n_delayed_max (must fit 31 bits)";
  else if (!p->dx || !p->dy || (p->dims == 3 && !p->dz))
    missing = "dx, dy, dz";
  else if (!p->mat_of_cell || !p->mat_flags || !p->n_delayed)
    missing = "mat_of_cell, mat_flags, n_delayed";
  else if (!p->sigma_t || !p->sigma_s || !p->nu_sigma_f || !p->chi_p || !p->src || !p->inv_v)
    missing = "sigma_t, sigma_s, nu_sigma_f, chi_p, src, inv_v (zeros when a material has none)";
  else if (p->n_delayed_max > 0 && (!p->chi_d || !p->beta || !p->lambda))
    missing = "chi_d, beta, lambda (n_delayed_max > 0)";
  else if (p->step_type && (!p->step_t_insert || !p->step_t_remove || !p->step_src))
    missing = "step_t_insert, step_t_remove, step_src (step_type given)";
  else if (p->n_ord > 0 && (!p->ord_octant || !p->mu || !p->eta || !p->xi || !p->w))
    missing = "ord_octant, mu, eta, xi, w (n_ord > 0)";
  else if (p->scheme != DOMA_SCHEME_DD && p->scheme != DOMA_SCHEME_TWDD)
    missing = "scheme (DOMA_SCHEME_DD or DOMA_SCHEME_TWDD)";
  else if (p->mode != DOMA_MODE_FIXED && p->mode != DOMA_MODE_EIGEN && p->mode != DOMA_MODE_TRANSIENT)
    missing = "mode (DOMA_MODE_*)";
  if (missing)
  {
    snprintf(g_create_err, DOMA_MAX_ERR, "create: invalid problem description: %s", missing);
    if (out)
      *out = nullptr;
    return 1;
  }
  doma_b200_ctx * c = (doma_b200_ctx *)calloc(1, sizeof(doma_b200_ctx));
  const int rc = create_impl(c, p);
  if (rc)
  {
    memcpy(g_create_err, c->err, DOMA_MAX_ERR);
    doma_b200_destroy(c);
    *out = nullptr;
    return rc;
  }
  *out = c;
  return 0;
}

int
doma_b200_destroy(doma_b200_ctx * c)
{
  if (c && c->comm_timing && c->comm_calls[1])
    fprintf(stderr, "doma_b200 comm timing (rank %d of %d, ms per call x calls): all-gather(q) %.3f x %d, reduce(phi_cur) %.3f x %d, pair all-reduce %.3f x %d, accumulator clear %.3f x %d\n",
            c->rank, c->nranks, c->comm_calls[0] ? c->comm_ms[0] / c->comm_calls[0] : 0.0, c->comm_calls[0], c->comm_ms[1] / c->comm_calls[1], c->comm_calls[1],
            c->comm_calls[2] ? c->comm_ms[2] / c->comm_calls[2] : 0.0, c->comm_calls[2], c->comm_calls[3] ? c->comm_ms[3] / c->comm_calls[3] : 0.0, c->comm_calls[3]);
  if (!c)
    return 0;
  cudaSetDevice(c->device);
  doma_b200_dump_wait(c);
  if (c->stream)
    cudaStreamSynchronize(c->stream);
  if (c->nccl_comm && g_nccl.CommDestroy)
    g_nccl.CommDestroy(c->nccl_comm);
  void * ptrs[] = {c->dx, c->dy, c->dz, c->rdx, c->rdy, c->rdz, c->mat, c->mat_flags, c->n_delayed, c->step_type, c->sigma_t,
                   c->sigma_s, c->nu_sigma_f, c->chi_p, c->src, c->inv_v, c->chi_d, c->beta, c->lambda, c->step_t_insert,
                   c->step_t_remove, c->step_src, c->sig_cell, c->phi_tot, c->phi_prev, c->phi_last, c->dnp_cur, c->dnp_last,
                   c->q, c->phi_cur, c->unit_octant, c->u_mu, c->u_eta, c->u_xi, c->u_w, c->tile_order, c->face_x, c->face_y,
                   c->face_z, c->tile_done, c->work_counter, c->red_partial, c->red_out, c->phi_red, c->gather_tmp, c->si_dev, c->comm_scratch};
  for (void * p : ptrs)
    if (p)
      cudaFree(p);
  if (c->async_ready)
  {
    cudaStreamSynchronize(c->copy_in), cudaStreamSynchronize(c->copy_out);
    cudaFree(c->q_stage), cudaFree(c->phi_snap);
    cudaEventDestroy(c->ev_q_staged), cudaEventDestroy(c->ev_q_consumed), cudaEventDestroy(c->ev_snap), cudaEventDestroy(c->ev_copied);
    cudaStreamDestroy(c->copy_in), cudaStreamDestroy(c->copy_out);
  }
  if (c->h_red)
    cudaFreeHost(c->h_red);
  if (c->si)
    cudaFreeHost(c->si);
  if (c->ev0)
    cudaEventDestroy(c->ev0);
  if (c->ev1)
    cudaEventDestroy(c->ev1);
  if (c->own_stream && c->stream)
    cudaStreamDestroy(c->stream);
  free(c->h_units);
  free(c->h_ord_octant), free(c->h_mu), free(c->h_eta), free(c->h_xi), free(c->h_w);
  free(c);
  return 0;
}

const char *
doma_b200_last_error(const doma_b200_ctx * c)
{
  return c ? c->err : g_create_err;
}

int
doma_b200_set_stream(doma_b200_ctx * c, void * s)
{
  DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  if (c->async_ready)
  {
    DOMA_CUDA(c, cudaStreamSynchronize(c->copy_in));
    DOMA_CUDA(c, cudaStreamSynchronize(c->copy_out));
  }
  if (s)
  {
    if (c->own_stream)
      cudaStreamDestroy(c->stream);
    c->stream = (cudaStream_t)s;
    c->own_stream = false;
  }
  else if (!c->own_stream)
  {
    DOMA_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    c->own_stream = true;
  }
  return 0;
}

int
doma_b200_synchronize(doma_b200_ctx * c)
{
  DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  if (c->async_ready)
  {
    DOMA_CUDA(c, cudaStreamSynchronize(c->copy_in));
    DOMA_CUDA(c, cudaStreamSynchronize(c->copy_out));
    c->copy_pending = false;
  }
  return 0;
}

// every entry point: null check, device selection, one NVTX range named after the ABI call (visible in nsys / ncu)
struct NvtxRange
{
  explicit NvtxRange(const char * name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};
#define ENTER(c)                    \
  if (!(c))                         \
    return 1;                       \
  NvtxRange nvtx_range_(__func__);  \
  DOMA_CUDA(c, cudaSetDevice((c)->device))
#define CHECK_GROUP(c, g)                                                             \
  if ((g) < 0 || (g) >= (c)->G)                                                       \
  {                                                                                   \
    snprintf((c)->err, DOMA_MAX_ERR, "%s: group %d outside [0,%d)", __func__, (int)(g), (c)->G); \
    return 4;                                                                         \
  }

int doma_b200_reset_state(doma_b200_ctx * c, double phi_init) { ENTER(c); return doma_launch_reset_state(c, phi_init); }
int doma_b200_init_zero_ic(doma_b200_ctx * c) { ENTER(c); return doma_launch_init_zero_ic(c); }
int doma_b200_update_step_fluxes(doma_b200_ctx * c) { ENTER(c); return doma_launch_update_step_fluxes(c); }
int doma_b200_step_dnps(doma_b200_ctx * c) { ENTER(c); return doma_launch_step_dnps(c); }

int
doma_b200_update_mg_source(doma_b200_ctx * c, int32_t g, double t)
{
  ENTER(c);
  CHECK_GROUP(c, g);
  return doma_launch_update_mg_source(c, g, t);
}

int
doma_b200_update_mg_source_eigen(doma_b200_ctx * c, int32_t g, double k)
{
  ENTER(c);
  CHECK_GROUP(c, g);
  return doma_launch_update_mg_source_eigen(c, g, k);
}

int
doma_b200_update_scatter_source(doma_b200_ctx * c, int32_t g)
{
  ENTER(c);
  CHECK_GROUP(c, g);
  return doma_launch_update_scatter_source(c, g);
}

int
doma_b200_sweep(doma_b200_ctx * c, int32_t g)
{
  ENTER(c);
  CHECK_GROUP(c, g);
  return doma_launch_sweep(c, g);
}

// Host-buffer form of sweep(g) for end-to-end timing: q (N doubles, host) -> device, phi_cur = 0,
// sweep, (all-reduce over ranks when a communicator exists,) phi_cur -> host.  Synchronous.
int
doma_b200_sweep_from_host(doma_b200_ctx * c, int32_t g, const double * q_host, double * phi_cur_out)
{
  ENTER(c);
  CHECK_GROUP(c, g);
  if (!q_host || !phi_cur_out)
  {
    snprintf(c->err, DOMA_MAX_ERR, "sweep_from_host: null host pointer");
    return 1;
  }
  // cell-sharded ranks move their own cells only: 1/P of q up (then an all-gather over NVLink), 1/P of phi down
  const int64_t lo = ew_lo(c), hi = ew_hi(c);
  DOMA_CUDA(c, cudaMemcpyAsync(c->q + lo, q_host + lo, sizeof(double) * (hi - lo), cudaMemcpyHostToDevice, c->stream));
  c->q_slice_only = c->cell_sharded;
  DOMA_CUDA(c, cudaMemsetAsync(c->phi_cur, 0, sizeof(double) * c->N, c->stream));
  int rc = doma_launch_sweep(c, g);
  if (rc)
    return rc;
  if (c->nccl_comm && (rc = doma_b200_allreduce_phi_cur(c)))
    return rc;
  DOMA_CUDA(c, cudaMemcpyAsync(phi_cur_out + lo, ew_phi_cur(c) + lo, sizeof(double) * (hi - lo), cudaMemcpyDeviceToHost, c->stream));
  DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

// ---- asynchronous host-buffer sweeps: the copies of one group overlap the kernels of its neighbours ----
// copy_in:  H2D q_host -> q_stage                      (may run while the previous sweep is still executing)
// stream:   q_stage -> q, phi_cur = 0, sweep, all-reduce, phi_cur -> phi_snap
// copy_out: D2H phi_snap -> phi_cur_out                (runs under the next group's source build and sweep)
namespace
{
int
async_setup(doma_b200_ctx * c)
{
  if (c->async_ready)
    return 0;
  DOMA_CUDA(c, cudaMalloc(&c->q_stage, sizeof(double) * c->N));
  DOMA_CUDA(c, cudaMalloc(&c->phi_snap, sizeof(double) * c->N));
  c->device_bytes += 2 * (int64_t)sizeof(double) * c->N;
  DOMA_CUDA(c, cudaStreamCreateWithFlags(&c->copy_in, cudaStreamNonBlocking));
  DOMA_CUDA(c, cudaStreamCreateWithFlags(&c->copy_out, cudaStreamNonBlocking));
  DOMA_CUDA(c, cudaEventCreateWithFlags(&c->ev_q_staged, cudaEventDisableTiming));
  DOMA_CUDA(c, cudaEventCreateWithFlags(&c->ev_q_consumed, cudaEventDisableTiming));
  DOMA_CUDA(c, cudaEventCreateWithFlags(&c->ev_snap, cudaEventDisableTiming));
  DOMA_CUDA(c, cudaEventCreateWithFlags(&c->ev_copied, cudaEventDisableTiming));
  // "free" state: nothing staged, nothing pending
  DOMA_CUDA(c, cudaEventRecord(c->ev_q_consumed, c->stream));
  DOMA_CUDA(c, cudaEventRecord(c->ev_copied, c->copy_out));
  c->async_ready = true;
  return 0;
}
} // namespace

int
doma_b200_stage_q_from_host(doma_b200_ctx * c, const double * q_host)
{
  ENTER(c);
  if (!q_host)
  {
    snprintf(c->err, DOMA_MAX_ERR, "stage_q_from_host: null host pointer");
    return 1;
  }
  if (c->q_is_staged)
  {
    snprintf(c->err, DOMA_MAX_ERR, "stage_q_from_host: a staged source is still waiting for its sweep");
    return 1;
  }
  int rc = async_setup(c);
  if (rc)
    return rc;
  DOMA_CUDA(c, cudaStreamWaitEvent(c->copy_in, c->ev_q_consumed, 0)); // the previous staged source has left q_stage
  DOMA_CUDA(c, cudaMemcpyAsync(c->q_stage + ew_lo(c), q_host + ew_lo(c), sizeof(double) * (ew_hi(c) - ew_lo(c)), cudaMemcpyHostToDevice,
                               c->copy_in)); // cell-sharded: the own cells only; the sweep all-gathers q over NVLink
  DOMA_CUDA(c, cudaEventRecord(c->ev_q_staged, c->copy_in));
  c->q_is_staged = true;
  return 0;
}

int
doma_b200_sweep_from_host_async(doma_b200_ctx * c, int32_t g, const double * q_host, double * phi_cur_out)
{
  ENTER(c);
  CHECK_GROUP(c, g);
  if (!phi_cur_out || (!q_host && !c->q_is_staged))
  {
    snprintf(c->err, DOMA_MAX_ERR, "sweep_from_host_async: %s", !phi_cur_out ? "null output pointer" : "no source: pass q_host or stage one first");
    return 1;
  }
  int rc = 0;
  if (q_host && (rc = doma_b200_stage_q_from_host(c, q_host)))
    return rc;
  DOMA_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_q_staged, 0));
  if ((rc = doma_launch_stage_in(c)))
    return rc;
  DOMA_CUDA(c, cudaEventRecord(c->ev_q_consumed, c->stream));
  c->q_is_staged = false;
  if ((rc = doma_launch_sweep(c, g)))
    return rc;
  if (c->nccl_comm && (rc = doma_b200_allreduce_phi_cur(c)))
    return rc;
  DOMA_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copied, 0)); // the previous result has left phi_snap
  if ((rc = doma_launch_snapshot_phi_cur(c)))
    return rc;
  DOMA_CUDA(c, cudaEventRecord(c->ev_snap, c->stream));
  DOMA_CUDA(c, cudaStreamWaitEvent(c->copy_out, c->ev_snap, 0));
  DOMA_CUDA(c, cudaMemcpyAsync(phi_cur_out + ew_lo(c), c->phi_snap + ew_lo(c), sizeof(double) * (ew_hi(c) - ew_lo(c)),
                               cudaMemcpyDeviceToHost, c->copy_out)); // cell-sharded: the summed increment of the own cells
  DOMA_CUDA(c, cudaEventRecord(c->ev_copied, c->copy_out));
  c->copy_pending = true;
  return 0;
}

int
doma_b200_host_wait(doma_b200_ctx * c)
{
  ENTER(c);
  if (c->copy_pending)
  {
    DOMA_CUDA(c, cudaEventSynchronize(c->ev_copied));
    c->copy_pending = false;
  }
  return 0;
}

// sourceIteration(g) (3D:458-497) owned by the library: no host round trip between the sweep, the all-reduce
// and the residual/update pass; one 16-byte readback per iteration for the loop condition.
int
doma_b200_source_iteration(doma_b200_ctx * c, int32_t g, double tol, int32_t max_it, int32_t * iters_out, double * residual_out,
                           double * residual_history)
{
  ENTER(c);
  CHECK_GROUP(c, g);
  if (!iters_out || !residual_out)
  {
    snprintf(c->err, DOMA_MAX_ERR, "source_iteration: null output pointer");
    return 1;
  }
  int32_t it = 0;
  double res = 0.0;
  if (!c->nccl_comm && !c->check && max_it <= DOMA_SI_HIST && !getenv("DOMA_B200_HOST_SI"))
  {
    // One rank: the loop runs on the device.  Iterations are enqueued in batches; the closing kernel of each iteration
    // evaluates the loop condition and raises a stop flag that turns every kernel still in the queue into a no-op, so
    // the host synchronises once per batch instead of once per sweep (the 2D transient case runs 2.6 M sweeps of
    // 0.2 ms).  Same kernels, same values, same iteration counts as the host-side loop below.
    const int cap = std::max(max_it, 1); // a do-while: the first sweep always runs (3D:464-481)
    c->si->it = 0, c->si->stop = 0, c->si->res = 0.0;
    DOMA_CUDA(c, cudaMemsetAsync(c->si_dev, 0, 2 * sizeof(int) + sizeof(double), c->stream)); // it, stop, res
    int enqueued = 0;
    int batch = 4;
    while (!c->si->stop)
    {
      const int n = std::min(batch, cap - enqueued);
      for (int i = 0; i < n; ++i)
      {
        int rc = doma_launch_sweep(c, g, &c->si_dev->stop, enqueued + i == 0); // the first sweep always runs: it is the timed one
        if (rc)
          return rc;
        if ((rc = doma_launch_si_iteration_tail(c, g, tol, max_it)))
          return rc;
      }
      enqueued += n;
      DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
      batch = std::min(batch * 2, 32);
      if (enqueued >= cap)
        break;
    }
    it = c->si->it;
    res = c->si->res;
    if (residual_history)
      memcpy(residual_history, c->si->hist, sizeof(double) * it);
    *iters_out = it;
    *residual_out = res;
    return 0;
  }
  do
  {
    int rc = doma_launch_sweep(c, g);
    if (rc)
      return rc;
    if (c->nccl_comm && (rc = doma_b200_allreduce_phi_cur(c)))
      return rc;
    if ((rc = doma_launch_scatter_residual_update(c, g)))
      return rc;
    res = c->h_red[1] > 1e-8 ? sqrt(c->h_red[0]) / sqrt(c->h_red[1]) : 0.0; // 3D:562
    if (residual_history)
      residual_history[it] = res;
    it++;
  } while (it < max_it && tol < res);
  *iters_out = it;
  *residual_out = res;
  return 0;
}

int
doma_b200_scatter_residual(doma_b200_ctx * c, int32_t g, double * out)
{
  ENTER(c);
  CHECK_GROUP(c, g);
  const int rc = doma_launch_scatter_residual(c, g);
  if (rc)
    return rc;
  // 3D:562: total_L2 > 1e-8 ? sqrt(diff)/sqrt(total) : 0
  *out = c->h_red[1] > 1e-8 ? sqrt(c->h_red[0]) / sqrt(c->h_red[1]) : 0.0;
  return 0;
}

int
doma_b200_mg_residual(doma_b200_ctx * c, double * out)
{
  ENTER(c);
  const int rc = doma_launch_mg_residual(c);
  if (rc)
    return rc;
  *out = c->h_red[1] > 1e-8 ? sqrt(c->h_red[0]) / sqrt(c->h_red[1]) : 0.0;
  return 0;
}

int
doma_b200_update_k(doma_b200_ctx * c, double * k_inout, double * k_prev_out)
{
  ENTER(c);
  const double k = *k_inout;
  const int rc = doma_launch_k_sums(c, k);
  if (rc)
    return rc;
  *k_prev_out = k;
  *k_inout = c->h_red[0] / c->h_red[1];
  return 0;
}

static double *
array_ptr(doma_b200_ctx * c, int which, int index)
{
  const int64_t N = c->N;
  switch (which)
  {
    case DOMA_ARR_PHI_TOT: return index >= 0 && index < c->G ? c->phi_tot + index * N : nullptr;
    case DOMA_ARR_PHI_PREV: return index >= 0 && index < c->G ? c->phi_prev + index * N : nullptr;
    case DOMA_ARR_PHI_LAST: return c->phi_last && index >= 0 && index < c->G ? c->phi_last + index * N : nullptr;
    case DOMA_ARR_DNP_CUR: return c->dnp_cur && index >= 0 && index < c->D ? c->dnp_cur + index * N : nullptr;
    case DOMA_ARR_DNP_LAST: return c->dnp_last && index >= 0 && index < c->D ? c->dnp_last + index * N : nullptr;
    case DOMA_ARR_Q: return c->q;
    case DOMA_ARR_PHI_CUR: return c->phi_cur;
    default: return nullptr;
  }
}

namespace
{
// cell-sharded: the slices of one per-cell array on all ranks -> gather_tmp (a collective: every rank calls it)
int
gather_slices(doma_b200_ctx * c, const double * array)
{
  const int r = g_nccl.AllGather(array + c->cell_lo, c->gather_tmp, (size_t)c->slice, /*ncclDouble*/ 8, c->nccl_comm, c->stream);
  if (r)
  {
    snprintf(c->err, DOMA_MAX_ERR, "ncclAllGather: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    return 24;
  }
  c->launches += 1;
  return 0;
}

// the device address a whole per-cell array can be copied to the host from (cell-sharded: after a gather)
int
whole_array(doma_b200_ctx * c, int which, const double * p, const double ** out)
{
  *out = p;
  if (!c->cell_sharded || which == DOMA_ARR_Q)
    return which == DOMA_ARR_Q ? doma_comm_gather_q(c) : 0;
  if (which == DOMA_ARR_PHI_CUR)
  {
    if (!c->phi_cur_reduced)
      return 0; // between a sweep and its reduction: this rank's partial sums
    p = c->phi_red - c->cell_lo;
  }
  const int rc = gather_slices(c, p);
  *out = c->gather_tmp;
  return rc;
}
} // namespace

int
doma_b200_get_array(doma_b200_ctx * c, int32_t which, int32_t index, double * host_out)
{
  ENTER(c);
  const double * p = array_ptr(c, which, index);
  if (!p || !host_out)
  {
    snprintf(c->err, DOMA_MAX_ERR, "get_array: no array %d[%d] in this context", which, index);
    return 4;
  }
  const int rc = whole_array(c, which, p, &p);
  if (rc)
    return rc;
  DOMA_CUDA(c, cudaMemcpyAsync(host_out, p, sizeof(double) * c->N, cudaMemcpyDeviceToHost, c->stream));
  DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int
doma_b200_set_array(doma_b200_ctx * c, int32_t which, int32_t index, const double * host_in)
{
  ENTER(c);
  double * p = array_ptr(c, which, index);
  if (!p || !host_in)
  {
    snprintf(c->err, DOMA_MAX_ERR, "set_array: no array %d[%d] in this context", which, index);
    return 4;
  }
  if (c->cell_sharded && which == DOMA_ARR_PHI_CUR)
  {
    // the increment the residual / update kernels read = the summed one of the own cells; the local accumulator restarts
    DOMA_CUDA(c, cudaMemcpyAsync(c->phi_red, host_in + c->cell_lo, sizeof(double) * (c->cell_hi - c->cell_lo), cudaMemcpyHostToDevice, c->stream));
    DOMA_CUDA(c, cudaMemsetAsync(c->phi_cur, 0, sizeof(double) * c->N, c->stream));
    c->phi_cur_reduced = true;
  }
  else
    DOMA_CUDA(c, cudaMemcpyAsync(p, host_in, sizeof(double) * c->N, cudaMemcpyHostToDevice, c->stream));
  if (which == DOMA_ARR_Q)
    c->q_slice_only = false;
  DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  return 0;
}

int
doma_b200_device_ptr(doma_b200_ctx * c, int32_t which, int32_t index, void ** out)
{
  if (!c || !out)
    return 1;
  double * p = array_ptr(c, which, index);
  if (!p)
  {
    snprintf(c->err, DOMA_MAX_ERR, "device_ptr: no array %d[%d] in this context", which, index);
    return 4;
  }
  *out = p;
  return 0;
}

int
doma_b200_get_info(doma_b200_ctx * c, int32_t what, int64_t * out)
{
  if (!c || !out)
    return 1;
  switch (what)
  {
    case DOMA_INFO_KERNEL_LAUNCHES: *out = c->launches; break;
    case DOMA_INFO_SWEEP_VARIANT: *out = doma_sweep_uses_ll(c) ? DOMA_SWEEP_LL : DOMA_SWEEP_FLAGS; break;
    case DOMA_INFO_N_CELLS: *out = c->N; break;
    case DOMA_INFO_N_ORD: *out = c->n_ord; break;
    case DOMA_INFO_FACE_BYTES: *out = c->face_bytes; break;
    case DOMA_INFO_DEVICE_BYTES: *out = c->device_bytes; break;
    case DOMA_INFO_N_UNITS: *out = c->n_units; break;
    case DOMA_INFO_N_TILES: *out = c->n_tiles; break;
    case DOMA_INFO_TILE_Y: *out = c->tile_y; break;
    case DOMA_INFO_TILE_Z: *out = c->tile_z; break;
    case DOMA_INFO_CELL_LO: *out = ew_lo(c); break;
    case DOMA_INFO_CELL_HI: *out = ew_hi(c); break;
    case DOMA_INFO_ROWS_PER_STEP:
      doma_choose_rows(c);
      *out = doma_sweep_uses_rows2(c) ? 2 : 1;
      break;
    case DOMA_INFO_SWEEP_BLOCKS: *out = c->last_blocks; break;
    case DOMA_INFO_TILE_X: *out = c->tile_x; break;
    default: snprintf(c->err, DOMA_MAX_ERR, "get_info: unknown item %d", what); return 4;
  }
  return 0;
}

int
doma_b200_set_sweep_variant(doma_b200_ctx * c, int32_t variant)
{
  ENTER(c);
  if (variant != DOMA_SWEEP_AUTO && variant != DOMA_SWEEP_FLAGS && variant != DOMA_SWEEP_LL)
  {
    snprintf(c->err, DOMA_MAX_ERR, "set_sweep_variant: %d is none of DOMA_SWEEP_AUTO, DOMA_SWEEP_FLAGS, DOMA_SWEEP_LL", variant);
    return 4;
  }
  c->variant = variant;
  return c->tile_auto ? doma_choose_tile(c) : 0; // the best tile depends on the protocol
}

int
doma_b200_set_rows_per_step(doma_b200_ctx * c, int32_t rows)
{
  ENTER(c);
  if (rows < 0 || rows > 2)
  {
    snprintf(c->err, DOMA_MAX_ERR, "set_rows_per_step: %d is none of 0 (automatic), 1, 2", rows);
    return 4;
  }
  c->rows2 = rows == 0 ? -1 : rows - 1;
  return 0;
}

int
doma_b200_set_pairing(doma_b200_ctx * c, int32_t on)
{
  ENTER(c);
  if ((on != 0) == c->pairing_enabled)
    return 0;
  c->pairing_enabled = on != 0;
  return doma_build_units(c);
}

int
doma_b200_set_check(doma_b200_ctx * c, int32_t on)
{
  ENTER(c);
  c->check = on != 0;
  return 0;
}

int
doma_b200_set_tile(doma_b200_ctx * c, int32_t tx, int32_t ty, int32_t tz)
{
  ENTER(c);
  if (ty == 0 && tz == 0)
    return doma_choose_tile(c); // back to the library's own choice
  if (tx != 8 && tx != 4)
  {
    snprintf(c->err, DOMA_MAX_ERR, "set_tile: the x extent of a tile is 8 or 4 (the row a warp holds in registers), not %d", tx);
    return 6;
  }
  return doma_sweep_setup_tiles(c, tx, ty, tz);
}

int
doma_b200_last_sweep_ms(doma_b200_ctx * c, float * out)
{
  ENTER(c);
  if (!c->have_sweep_time)
  {
    *out = 0.f;
    return 0;
  }
  DOMA_CUDA(c, cudaEventSynchronize(c->ev1));
  DOMA_CUDA(c, cudaEventElapsedTime(out, c->ev0, c->ev1));
  return 0;
}

// ---- text output (BrickMesh3D::dumpToTextFile / dumpDNPsToTextFile, src/BrickMesh3D.cpp:216-295; 2D: src/BrickMesh2D.cpp) ----
// The reference writes one value per line with `stream << std::setprecision(6) << value << std::endl`, i.e. "%.6g\n" and a
// flush (a write system call) per line, serially.  Same bytes here: the values are formatted by n_threads host threads
// into per-thread buffers and written with one fwrite per buffer.
namespace
{
int
write_values_text(const double * v, int64_t n, const char * path, int n_threads, char * err)
{
  if (n_threads < 1)
    n_threads = (int)std::max(1u, std::min(std::thread::hardware_concurrency(), 64u));
  if ((int64_t)n_threads > std::max<int64_t>(1, n / 4096))
    n_threads = (int)std::max<int64_t>(1, n / 4096);
  std::vector<std::vector<char>> bufs(n_threads);
  auto work = [&](int t) {
    const int64_t lo = n * t / n_threads, hi = n * (t + 1) / n_threads;
    std::vector<char> & b = bufs[t];
    b.resize((size_t)(hi - lo) * 26 + 32);
    size_t pos = 0;
    for (int64_t i = lo; i < hi; ++i)
      pos += (size_t)snprintf(b.data() + pos, 26, "%.6g\n", v[i]);
    b.resize(pos);
  };
  std::vector<std::thread> th;
  for (int t = 1; t < n_threads; ++t)
    th.emplace_back(work, t);
  work(0);
  for (auto & t : th)
    t.join();
  FILE * f = fopen(path, "wb");
  if (!f)
  {
    snprintf(err, DOMA_MAX_ERR, "write_values_text: cannot open %s", path);
    return 9;
  }
  bool ok = true;
  for (auto & b : bufs)
    ok = ok && fwrite(b.data(), 1, b.size(), f) == b.size();
  ok = (fclose(f) == 0) && ok;
  if (!ok)
  {
    snprintf(err, DOMA_MAX_ERR, "write_values_text: short write to %s", path);
    return 9;
  }
  return 0;
}
} // namespace

int
doma_b200_write_values_text(const double * values, int64_t n, const char * path, int32_t n_threads)
{
  if (!values || n < 0 || !path)
  {
    snprintf(g_create_err, DOMA_MAX_ERR, "write_values_text: invalid argument");
    return 1;
  }
  return write_values_text(values, n, path, n_threads, g_create_err);
}

// A dump = a host snapshot of the state (taken on the calling thread, so the solve may go on and overwrite the device
// arrays) + the formatting and writing of the files, which runs inline (doma_b200_dump_text) or on a worker thread
// that the next dump, doma_b200_dump_wait or doma_b200_destroy joins (doma_b200_dump_text_async).
struct DumpJob
{
  std::string base;
  int64_t N = 0;
  int n_threads = 0;
  std::vector<std::vector<double>> flux, dnps;
  std::thread worker;
  int rc = 0;
  char err[DOMA_MAX_ERR] = {0};
  void run()
  {
    if (base.empty()) // a rank that only takes part in the gather of a cell-sharded run
      flux.clear(), dnps.clear();
    for (size_t g = 0; g < flux.size() && !rc; ++g)
      rc = write_values_text(flux[g].data(), N, (base + "_g" + std::to_string(g) + "_flux.txt").c_str(), n_threads, err);
    for (size_t d = 0; d < dnps.size() && !rc; ++d)
      rc = write_values_text(dnps[d].data(), N, (base + "_d" + std::to_string(d) + "_dnps.txt").c_str(), n_threads, err);
    flux.clear(), dnps.clear();
  }
};

namespace
{
int
join_dump(doma_b200_ctx * c)
{
  DumpJob * j = (DumpJob *)c->dump_job;
  if (!j)
    return 0;
  if (j->worker.joinable())
    j->worker.join();
  const int rc = j->rc;
  if (rc)
    memcpy(c->err, j->err, DOMA_MAX_ERR);
  delete j;
  c->dump_job = nullptr;
  return rc;
}

int
snapshot_for_dump(doma_b200_ctx * c, DumpJob * j, const char * file_base, int with_dnps, int n_threads)
{
  DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  const size_t N = (size_t)c->N;
  j->base = file_base, j->N = c->N, j->n_threads = n_threads;
  j->flux.resize(c->G);
  for (int g = 0; g < c->G; ++g)
  {
    j->flux[g].resize(N);
    const double * src = nullptr;
    const int rc = whole_array(c, DOMA_ARR_PHI_TOT, c->phi_tot + (size_t)g * N, &src);
    if (rc)
      return rc;
    DOMA_CUDA(c, cudaMemcpyAsync(j->flux[g].data(), src, sizeof(double) * N, cudaMemcpyDeviceToHost, c->stream));
    DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  if (with_dnps && c->D > 0 && c->dnp_cur)
  {
    // dumpDNPsToTextFile: num_d_groups = the largest precursor count of any cell's material; a cell whose material has
    // fewer prints 0 for every precursor group (src/BrickMesh3D.cpp:268-289)
    std::vector<int32_t> mat(N), nd((size_t)c->M);
    DOMA_CUDA(c, cudaMemcpy(mat.data(), c->mat, sizeof(int32_t) * N, cudaMemcpyDeviceToHost));
    DOMA_CUDA(c, cudaMemcpy(nd.data(), c->n_delayed, sizeof(int32_t) * c->M, cudaMemcpyDeviceToHost));
    int dmax = 0;
    for (size_t i = 0; i < N; ++i)
      dmax = std::max(dmax, (int)nd[mat[i]]);
    j->dnps.resize(dmax);
    for (int d = 0; d < dmax; ++d)
    {
      j->dnps[d].resize(N);
      const double * src = nullptr;
      const int rc = whole_array(c, DOMA_ARR_DNP_CUR, c->dnp_cur + (size_t)d * N, &src);
      if (rc)
        return rc;
      DOMA_CUDA(c, cudaMemcpyAsync(j->dnps[d].data(), src, sizeof(double) * N, cudaMemcpyDeviceToHost, c->stream));
      DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
      for (size_t i = 0; i < N; ++i)
        if (nd[mat[i]] != dmax)
          j->dnps[d][i] = 0.0;
    }
  }
  return 0;
}

int
dump_impl(doma_b200_ctx * c, const char * file_base, int with_dnps, int n_threads, bool async)
{
  if (!file_base)
  {
    snprintf(c->err, DOMA_MAX_ERR, "dump_text: null file base");
    return 1;
  }
  int rc = join_dump(c); // one dump in flight per context; also surfaces the previous one's error
  if (rc)
    return rc;
  DumpJob * j = new DumpJob();
  if ((rc = snapshot_for_dump(c, j, file_base, with_dnps, n_threads)))
  {
    delete j;
    return rc;
  }
  c->dump_job = j;
  if (async)
  {
    j->worker = std::thread([j] { j->run(); });
    return 0;
  }
  j->run();
  return join_dump(c);
}
} // namespace

int
doma_b200_dump_text(doma_b200_ctx * c, const char * file_base, int32_t with_dnps, int32_t n_threads)
{
  ENTER(c);
  return dump_impl(c, file_base, with_dnps, n_threads, false);
}

int
doma_b200_dump_text_async(doma_b200_ctx * c, const char * file_base, int32_t with_dnps, int32_t n_threads)
{
  ENTER(c);
  return dump_impl(c, file_base, with_dnps, n_threads, true);
}

int
doma_b200_dump_wait(doma_b200_ctx * c)
{
  ENTER(c);
  return join_dump(c);
}

namespace
{
// measurement mode (DOMA_B200_COMM_TIMING): device time of one collective, in isolation (synchronises)
struct CommTimer
{
  doma_b200_ctx * c;
  int slot;
  cudaEvent_t a = nullptr, b = nullptr;
  CommTimer(doma_b200_ctx * ctx, int s) : c(ctx), slot(s)
  {
    if (!c->comm_timing)
      return;
    cudaEventCreate(&a), cudaEventCreate(&b);
    if (c->nccl_comm && slot < 2 && c->comm_scratch) // align the ranks first, or the time of the first collective after a host gap is rank skew
      g_nccl.AllReduce(c->comm_scratch, c->comm_scratch, 1, /*ncclDouble*/ 8, /*ncclSum*/ 0, c->nccl_comm, c->stream);
    cudaStreamSynchronize(c->stream);
    cudaEventRecord(a, c->stream);
  }
  ~CommTimer()
  {
    if (!c->comm_timing)
      return;
    cudaEventRecord(b, c->stream);
    cudaEventSynchronize(b);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    c->comm_ms[slot] += ms, c->comm_calls[slot] += 1;
    cudaEventDestroy(a), cudaEventDestroy(b);
  }
};
} // namespace

// ---- angle sharding across ranks: one in-place ncclAllReduce(phi_cur) per sweep ----
int
doma_b200_comm_unique_id(char * out, int32_t len)
{
  if (!out || len < 128)
    return 1;
  int rc = load_nccl(g_create_err);
  if (rc)
    return rc;
  NcclId id;
  const int r = g_nccl.GetUniqueId(&id);
  if (r)
  {
    snprintf(g_create_err, DOMA_MAX_ERR, "ncclGetUniqueId: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    return 22;
  }
  memcpy(out, id.b, 128);
  return 0;
}

int
doma_b200_comm_init(doma_b200_ctx * c, int32_t rank, int32_t nranks, const char * unique_id, int32_t len)
{
  ENTER(c);
  if (len < 128 || !unique_id || rank < 0 || rank >= nranks)
  {
    snprintf(c->err, DOMA_MAX_ERR, "comm_init: bad arguments");
    return 1;
  }
  int rc = load_nccl(c->err);
  if (rc)
    return rc;
  NcclId id;
  memcpy(id.b, unique_id, 128);
  typedef int (*init_fn)(void **, int, NcclId, int);
  const int r = ((init_fn)g_nccl.CommInitRank)(&c->nccl_comm, nranks, id, rank);
  if (r)
  {
    snprintf(c->err, DOMA_MAX_ERR, "ncclCommInitRank: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    return 23;
  }
  c->rank = rank;
  c->nranks = nranks;
  return nranks > 1 ? doma_b200_set_cell_sharding(c, 1) : 0;
}

// Cell sharding of everything that is not the sweep (ctx.h): on by default once a communicator of more than one rank
// exists; off = the r01 scheme (all-reduce of phi_cur, every rank repeats the elementwise work on all cells).
// Must be switched on all ranks alike, between solves (the state of the other cells is not moved).
int
doma_b200_set_cell_sharding(doma_b200_ctx * c, int32_t on)
{
  ENTER(c);
  if (!c->nccl_comm || c->nranks < 2)
  {
    if (on)
    {
      snprintf(c->err, DOMA_MAX_ERR, "set_cell_sharding: needs a communicator of at least two ranks (doma_b200_comm_init)");
      return 4;
    }
    return 0;
  }
  DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  if (!on)
  {
    if (c->cell_sharded)
    {
      // every array becomes whole on every rank again
      const int64_t GN = (int64_t)c->G * c->N;
      (void)GN;
      double * arrays[] = {c->phi_tot, c->phi_prev, c->phi_last};
      for (double * base : arrays)
        for (int g = 0; base && g < c->G; ++g)
        {
          int rc = gather_slices(c, base + (size_t)g * c->N);
          if (rc)
            return rc;
          DOMA_CUDA(c, cudaMemcpyAsync(base + (size_t)g * c->N, c->gather_tmp, sizeof(double) * c->N, cudaMemcpyDeviceToDevice, c->stream));
        }
      double * dn[] = {c->dnp_cur, c->dnp_last};
      for (double * base : dn)
        for (int d = 0; base && d < c->D; ++d)
        {
          int rc = gather_slices(c, base + (size_t)d * c->N);
          if (rc)
            return rc;
          DOMA_CUDA(c, cudaMemcpyAsync(base + (size_t)d * c->N, c->gather_tmp, sizeof(double) * c->N, cudaMemcpyDeviceToDevice, c->stream));
        }
      int rc = doma_comm_gather_q(c);
      if (rc)
        return rc;
      if (c->phi_cur_reduced)
      {
        if ((rc = gather_slices(c, c->phi_red - c->cell_lo)))
          return rc;
        DOMA_CUDA(c, cudaMemcpyAsync(c->phi_cur, c->gather_tmp, sizeof(double) * c->N, cudaMemcpyDeviceToDevice, c->stream));
      }
      DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    c->cell_sharded = false;
    return 0;
  }
  if (c->cell_sharded)
    return 0;
  const int64_t P = c->nranks;
  c->slice = (((c->N + P - 1) / P) + 1) & ~(int64_t)1; // even: 16-byte aligned slices
  if (P * c->slice > c->N + DOMA_PAD)
  {
    snprintf(c->err, DOMA_MAX_ERR, "set_cell_sharding: %lld ranks need more padding than the %d doubles the arrays carry", (long long)P, DOMA_PAD);
    return 4;
  }
  c->cell_lo = std::min<int64_t>(c->N, c->rank * c->slice);
  c->cell_hi = std::min<int64_t>(c->N, c->cell_lo + c->slice);
  if (!c->phi_red)
  {
    DOMA_CUDA(c, cudaMalloc(&c->phi_red, sizeof(double) * c->slice));
    DOMA_CUDA(c, cudaMalloc(&c->gather_tmp, sizeof(double) * P * c->slice));
    c->device_bytes += sizeof(double) * (P + 1) * c->slice;
  }
  // the increment of the own cells as it stands (phi_cur is whole and identical on every rank at this point)
  DOMA_CUDA(c, cudaMemcpyAsync(c->phi_red, c->phi_cur + c->cell_lo, sizeof(double) * (c->cell_hi - c->cell_lo), cudaMemcpyDeviceToDevice, c->stream));
  DOMA_CUDA(c, cudaMemsetAsync(c->phi_cur, 0, sizeof(double) * c->N, c->stream));
  DOMA_CUDA(c, cudaStreamSynchronize(c->stream));
  c->cell_sharded = true;
  c->q_slice_only = false;
  c->phi_cur_reduced = true;
  return 0;
}

// The sum of the sweep's increment over the ranks, on the context's stream.  Replicated mode: one in-place
// ncclAllReduce(phi_cur).  Cell-sharded mode: ncclReduceScatter(phi_cur) -> the own cells' sums in phi_red, and the
// local accumulator is cleared for the next sweep.
int
doma_b200_allreduce_phi_cur(doma_b200_ctx * c)
{
  ENTER(c);
  if (!c->nccl_comm)
    return 0; // single rank
  int r = 0;
  if (c->cell_sharded)
  {
    {
      CommTimer timer(c, 1);
      r = g_nccl.ReduceScatter(c->phi_cur, c->phi_red, (size_t)c->slice, /*ncclDouble*/ 8, /*ncclSum*/ 0, c->nccl_comm, c->stream);
    }
    if (!r)
    {
      CommTimer timer(c, 3);
      DOMA_CUDA(c, cudaMemsetAsync(c->phi_cur, 0, sizeof(double) * c->N, c->stream));
    }
    c->phi_cur_reduced = true;
  }
  else
  {
    CommTimer timer(c, 1);
    r = g_nccl.AllReduce(c->phi_cur, c->phi_cur, (size_t)c->N, /*ncclDouble*/ 8, /*ncclSum*/ 0, c->nccl_comm, c->stream);
  }
  if (r)
  {
    snprintf(c->err, DOMA_MAX_ERR, "%s: %s", c->cell_sharded ? "ncclReduceScatter" : "ncclAllReduce", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    return 24;
  }
  c->launches += 1;
  return 0;
}

} // extern "C"

// ---- collectives the other translation units use ----

int
doma_comm_gather_q(doma_b200_ctx * c)
{
  if (!c->cell_sharded || !c->q_slice_only)
    return 0;
  CommTimer timer(c, 0);
  const int r = g_nccl.AllGather(c->q + c->cell_lo, c->q, (size_t)c->slice, /*ncclDouble*/ 8, c->nccl_comm, c->stream);
  if (r)
  {
    snprintf(c->err, DOMA_MAX_ERR, "ncclAllGather(q): %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    return 24;
  }
  c->launches += 1;
  c->q_slice_only = false;
  return 0;
}

int
doma_comm_sum_pair(doma_b200_ctx * c)
{
  CommTimer timer(c, 2);
  const int r = g_nccl.AllReduce(c->red_out, c->red_out, 2, /*ncclDouble*/ 8, /*ncclSum*/ 0, c->nccl_comm, c->stream);
  if (r)
  {
    snprintf(c->err, DOMA_MAX_ERR, "ncclAllReduce(pair): %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    return 24;
  }
  c->launches += 1;
  return 0;
}
