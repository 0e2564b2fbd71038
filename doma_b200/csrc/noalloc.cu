// noalloc.cu - doma_b200_scatter_residual_noalloc: the scattering residual (computeScatteringResidual, 3D:547-563) for
// callers that cannot let the library allocate (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  It has no counterpart in the reference.
//
// The reduction's scratch - one pair of partial sums per thread block and the final pair - lives in a FILE-SCOPE
// __device__ array of fixed size instead of the context's cudaMalloc'ed red_partial / red_out:
//   * every GPU that loads the library holds DOMA_NOALLOC_SCRATCH_BYTES (128 KiB + 256 B) until the process exits;
//   * a block reduces DOMA_NOALLOC_CELLS_PER_BLOCK = 4 096 cells, the array has room for 8 192 blocks, so the entry
//     point takes at most DOMA_NOALLOC_MAX_CELLS = 33 554 432 cells (twice the 256^3 bench mesh) and rejects more;
//   * ALL calls on a device share the array.  Every slot a call reads it has written itself first (block b writes
//     slots b and MAX_BLOCKS + b, the closing kernel reads the first n_blocks of each half), so nothing has to start at
//     zero and the contents a previous call left behind are never read; two calls that OVERLAP on the GPU (different
//     streams, or graph branches that may run concurrently) race on it - serialise them.
// In host code `g_scratch` is a host-side stand-in for the symbol, not a device address: its address is taken once per
// device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels receive.
//
// The entry point calls no cudaMalloc / cudaFree / allocating library routine: two kernel launches, and - only when the
// caller asks for the residual on the host - one 16-byte cudaMemcpyAsync + cudaStreamSynchronize.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "doma_b200.h"

namespace
{
constexpr int NA_THREADS = 256;
constexpr int NA_CELLS_PER_THREAD = DOMA_NOALLOC_CELLS_PER_BLOCK / NA_THREADS;
constexpr int NA_MAX_BLOCKS = (int)(DOMA_NOALLOC_MAX_CELLS / DOMA_NOALLOC_CELLS_PER_BLOCK);
constexpr int NA_SCRATCH_DOUBLES = 2 * NA_MAX_BLOCKS + 32; // partial pairs + the final pair (padded to 256 bytes)
static_assert(NA_SCRATCH_DOUBLES * sizeof(double) == DOMA_NOALLOC_SCRATCH_BYTES, "header and kernel disagree on the scratch size");
static_assert(DOMA_NOALLOC_CELLS_PER_BLOCK % NA_THREADS == 0, "a block's chunk is a whole number of cells per thread");

// [0, MAX_BLOCKS): sum V cur^2 per block | [MAX_BLOCKS, 2 MAX_BLOCKS): sum V (tot + cur)^2 per block | [2 MAX_BLOCKS, +2): the final pair
static __device__ __align__(256) double g_scratch[NA_SCRATCH_DOUBLES];

// One block = one contiguous chunk of 4 096 cells; thread t takes cells t, t + 256, ... of the chunk (coalesced).
// Per-cell expressions as scatter_residual_kernel (elementwise.cu; this file is compiled with -fmad=false as well).
__global__ void __launch_bounds__(NA_THREADS)
noalloc_partial_kernel(int64_t n, int nx, int ny, int dims, const double * __restrict__ dx, const double * __restrict__ dy,
                       const double * __restrict__ dz, const double * __restrict__ phi_tot_g, const double * __restrict__ phi_cur,
                       double * __restrict__ scratch)
{
  __shared__ double sa[NA_THREADS / 32], sb[NA_THREADS / 32];
  double diff = 0.0, total = 0.0;
  const int64_t base = (int64_t)blockIdx.x * DOMA_NOALLOC_CELLS_PER_BLOCK + threadIdx.x;
#pragma unroll 4
  for (int r = 0; r < NA_CELLS_PER_THREAD; ++r)
  {
    const int64_t c = base + (int64_t)r * NA_THREADS;
    if (c < n)
    {
      const int i = (int)(c % nx);
      const int64_t row = c / nx;
      const int j = (int)(row % ny), k = (int)(row / ny);
      const double v = dims == 3 ? dx[i] * dy[j] * dz[k] : dx[i] * dy[j];
      const double cur = phi_cur[c], s = phi_tot_g[c] + cur;
      diff = diff + (cur * cur) * v;
      total = total + (s * s) * v;
    }
  }
  for (int o = 16; o > 0; o >>= 1)
  {
    diff += __shfl_down_sync(0xffffffffu, diff, o);
    total += __shfl_down_sync(0xffffffffu, total, o);
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0)
    sa[w] = diff, sb[w] = total;
  __syncthreads();
  if (threadIdx.x == 0)
  {
    double ta = 0.0, tb = 0.0;
    for (int q = 0; q < NA_THREADS / 32; ++q)
      ta += sa[q], tb += sb[q];
    scratch[blockIdx.x] = ta;
    scratch[NA_MAX_BLOCKS + blockIdx.x] = tb;
  }
}

// Closing kernel: warp 0 sums the first n_blocks partials of the first half, warp 1 of the second (lane-strided, then a
// fixed shuffle tree: deterministic); the pair goes to the scratch's tail and to the caller's output.
__global__ void
noalloc_final_kernel(int n_blocks, double * __restrict__ scratch, double * __restrict__ sums_out)
{
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double a = 0.0;
  for (int i = lane; i < n_blocks; i += 32)
    a += scratch[w * NA_MAX_BLOCKS + i];
  for (int o = 16; o > 0; o >>= 1)
    a += __shfl_down_sync(0xffffffffu, a, o);
  if (lane == 0)
  {
    scratch[2 * NA_MAX_BLOCKS + w] = a;
    sums_out[w] = a;
  }
}

// device address of g_scratch, taken once per device with that device current
constexpr int NA_MAX_DEVICES = 64;
std::mutex g_mutex;
double * g_scratch_ptr[NA_MAX_DEVICES] = {nullptr};

int
fail(char * err, int32_t err_len, int rc, const char * fmt, ...)
{
  if (err && err_len > 0)
  {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, (size_t)err_len, fmt, ap);
    va_end(ap);
  }
  return rc;
}
} // namespace

extern "C" int
doma_b200_scatter_residual_noalloc(const double * phi_tot_g_dev, const double * phi_cur_dev, const double * dx_dev, const double * dy_dev,
                                   const double * dz_dev, int32_t dims, int32_t nx, int32_t ny, int32_t nz, void * cuda_stream,
                                   double * sums_out_dev, double * residual_out_host, char * err_out, int32_t err_len)
{
  if (err_out && err_len > 0)
    err_out[0] = 0;
  if (dims != 2 && dims != 3)
    return fail(err_out, err_len, 1, "scatter_residual_noalloc: dims must be 2 or 3, not %d", dims);
  if (nx < 1 || ny < 1 || (dims == 3 && nz < 1))
    return fail(err_out, err_len, 1, "scatter_residual_noalloc: mesh %d x %d x %d has a non-positive extent", nx, ny, nz);
  if (!phi_tot_g_dev || !phi_cur_dev || !dx_dev || !dy_dev || (dims == 3 && !dz_dev) || !sums_out_dev)
    return fail(err_out, err_len, 1, "scatter_residual_noalloc: null device pointer (phi_tot_g, phi_cur, dx, dy, dz for 3D, sums_out are required)");
  const int64_t n = (int64_t)nx * ny * (dims == 3 ? nz : 1);
  if (n > (int64_t)DOMA_NOALLOC_MAX_CELLS)
    return fail(err_out, err_len, 40,
                "scatter_residual_noalloc: %lld cells exceed the %lld the static scratch is sized for (DOMA_NOALLOC_MAX_CELLS); use "
                "doma_b200_scatter_residual",
                (long long)n, (long long)DOMA_NOALLOC_MAX_CELLS);
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess)
    return fail(err_out, err_len, 100 + (int)e, "scatter_residual_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= NA_MAX_DEVICES)
    return fail(err_out, err_len, 1, "scatter_residual_noalloc: device ordinal %d out of range", dev);
  double * scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_scratch_ptr[dev])
    {
      // set-up, once per device, with that device current: the symbol's device address (the name g_scratch itself is
      // only a host-side stand-in here)
      void * p = nullptr;
      e = cudaGetSymbolAddress(&p, g_scratch);
      if (e != cudaSuccess || !p)
        return fail(err_out, err_len, 100 + (int)e, "scatter_residual_noalloc: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
      g_scratch_ptr[dev] = static_cast<double *>(p);
    }
    scratch = g_scratch_ptr[dev];
  }
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  const int n_blocks = (int)((n + DOMA_NOALLOC_CELLS_PER_BLOCK - 1) / DOMA_NOALLOC_CELLS_PER_BLOCK);
  noalloc_partial_kernel<<<n_blocks, NA_THREADS, 0, stream>>>(n, nx, ny, dims, dx_dev, dy_dev, dz_dev, phi_tot_g_dev, phi_cur_dev, scratch);
  if ((e = cudaGetLastError()) != cudaSuccess)
    return fail(err_out, err_len, 100 + (int)e, "scatter_residual_noalloc: partial-sum kernel: %s", cudaGetErrorString(e));
  noalloc_final_kernel<<<1, 64, 0, stream>>>(n_blocks, scratch, sums_out_dev);
  if ((e = cudaGetLastError()) != cudaSuccess)
    return fail(err_out, err_len, 100 + (int)e, "scatter_residual_noalloc: closing kernel: %s", cudaGetErrorString(e));
  if (residual_out_host)
  {
    double h[2] = {0.0, 0.0};
    e = cudaMemcpyAsync(h, sums_out_dev, sizeof(h), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess)
      e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess)
      return fail(err_out, err_len, 100 + (int)e, "scatter_residual_noalloc: reading the sums back: %s", cudaGetErrorString(e));
    *residual_out_host = h[1] > 1e-8 ? sqrt(h[0]) / sqrt(h[1]) : 0.0; // 3D:562
  }
  return 0;
}
