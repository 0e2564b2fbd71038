// cell_ceiling.cu - what can the SM issue of the sweep's own per-cell instruction stream when nothing but the FP64 pipe
// and the schedulers are in the way?  Runs cell_solve<3, TWDD, V> (the product's code, doma_b200/csrc/cell_solve.cuh)
// from registers, no memory, no wavefront, in two forms:
//   chained:     x, y, z of a cell feed the next cell of the same chain (the sweep's real dependency: one chain per
//                warp with V = 1, two with V = 2 = the R2 kernel's pair)
//   independent: every cell starts from fresh inputs (the chain is cut), so only issue and pipe capacity remain
// for W warps per SM.  Prints cells (= updates per lane) per clock per SM and the FP64-pipe share that corresponds to
// (55.75 FP64 instructions per update, 2 issue cycles each on a 16-lane pipe).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I include -I doma_b200/csrc -o /tmp/cell_ceiling scripts/cell_ceiling.cu
#include <cuda_runtime.h>
#include <stdio.h>

#include "doma_b200.h"

namespace
{
#include "cell_solve.cuh"
}

template <int V, bool CHAINED, int WARPS, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB)
ceiling_kernel(double * out, int iters, double q0, double s0, double m0, double e0, double z0)
{
  const int lane = threadIdx.x & 31;
  double q[V], sig[V], m[V], e[V], z_[V], x[V], y[V], z[V], psi[V], acc[V];
#pragma unroll
  for (int v = 0; v < V; ++v)
  {
    q[v] = q0 + 1e-3 * v;
    sig[v] = s0 + 1e-2 * lane;
    m[v] = m0 * (1.0 + 0.01 * lane);
    e[v] = e0 * (1.0 + 0.02 * lane);
    z_[v] = z0 * (1.0 + 0.03 * lane);
    x[v] = 0.3 + 0.001 * threadIdx.x + v;
    y[v] = 0.4;
    z[v] = 0.5;
    acc[v] = 0.0;
  }
  for (int it = 0; it < iters; ++it)
  {
#pragma unroll
    for (int r = 0; r < 8; ++r) // 8 cells = one row of the sweep's tile
    {
      if (!CHAINED)
      {
#pragma unroll
        for (int v = 0; v < V; ++v)
          x[v] = 0.3 + acc[v] * 1e-30, y[v] = 0.4, z[v] = 0.5; // cut the x -> x', y -> y', z -> z' chains (acc keeps the loop from being hoisted)
      }
      cell_solve<3, DOMA_SCHEME_TWDD, V>(q, sig, m, e, z_, x, y, z, psi);
#pragma unroll
      for (int v = 0; v < V; ++v)
        acc[v] += psi[v] + (CHAINED ? 0.0 : x[v] + y[v] + z[v]);
    }
  }
  double s = 0.0;
#pragma unroll
  for (int v = 0; v < V; ++v)
    s += acc[v] + x[v] + y[v] + z[v];
  if (s == 12345.678)
    out[0] = s;
}

template <int V, bool CHAINED, int WARPS, int MINB>
void
run(const char * what, int n_sms, double * d_out, int clock_khz)
{
  const int iters = 2048;
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  int per_sm = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ceiling_kernel<V, CHAINED, WARPS, MINB>, WARPS * 32, 0);
  const int blocks = n_sms * MINB;
  ceiling_kernel<V, CHAINED, WARPS, MINB><<<blocks, WARPS * 32>>>(d_out, 16, 1.0, 1.0, 0.3, 0.2, 0.1);
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  ceiling_kernel<V, CHAINED, WARPS, MINB><<<blocks, WARPS * 32>>>(d_out, iters, 1.0, 1.0, 0.3, 0.2, 0.1);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, a, b);
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, ceiling_kernel<V, CHAINED, WARPS, MINB>);
  const double warp_cells = (double)blocks * WARPS * iters * 8.0 * V;
  const double clk = ms * 1e-3 * clock_khz * 1e3;
  const double cells_per_clk_sm = warp_cells / n_sms / clk;
  // the FP64 pipe of an SM issues 2 warp-instructions per clock (4 sub-partitions x 16 lanes)
  printf("%-12s V %d warps/SM %2d (resident blocks/SM %d, %3d regs) : %.5f warp-cells/clk/SM  = %5.1f cycles per warp-cell per scheduler, FP64 pipe %.1f %%\n",
         what, V, WARPS * MINB, per_sm, fa.numRegs, cells_per_clk_sm, 4.0 / cells_per_clk_sm, 100.0 * cells_per_clk_sm * 55.75 / 2.0);
}

int
main()
{
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int clock_khz = 0;
  cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, 0);
  double * d_out;
  cudaMalloc(&d_out, 8);
  printf("device %s, %d SMs, %d MHz\n", p.name, p.multiProcessorCount, clock_khz / 1000);
  const int S = p.multiProcessorCount;
  run<1, true, 4, 1>("chained", S, d_out, clock_khz);
  run<1, true, 4, 2>("chained", S, d_out, clock_khz);
  run<1, true, 4, 4>("chained", S, d_out, clock_khz);
  run<1, true, 4, 5>("chained", S, d_out, clock_khz);
  run<1, true, 4, 6>("chained", S, d_out, clock_khz);
  run<1, true, 4, 8>("chained", S, d_out, clock_khz);
  run<1, true, 4, 12>("chained", S, d_out, clock_khz);
  run<1, true, 4, 16>("chained", S, d_out, clock_khz);
  run<2, true, 4, 2>("chained", S, d_out, clock_khz);
  run<2, true, 4, 3>("chained", S, d_out, clock_khz);
  run<2, true, 4, 4>("chained", S, d_out, clock_khz);
  run<2, true, 4, 6>("chained", S, d_out, clock_khz);
  run<4, true, 4, 2>("chained", S, d_out, clock_khz);
  run<4, true, 4, 3>("chained", S, d_out, clock_khz);
  run<1, false, 4, 4>("independent", S, d_out, clock_khz);
  run<1, false, 4, 8>("independent", S, d_out, clock_khz);
  run<1, false, 4, 16>("independent", S, d_out, clock_khz);
  run<2, false, 4, 4>("independent", S, d_out, clock_khz);
  run<4, false, 4, 2>("independent", S, d_out, clock_khz);
  return 0;
}
