// fp64_coissue.cu - does an FP64 instruction cost the scheduler one issue slot or two?  
// 16 warps per SM, 4 independent DFMA chains per thread (enough to saturate the pipe alone, profiles/r01_fp64_peak_b200.txt),
// plus K independent integer instructions (LOP3/IMAD on separate registers) per DFMA, interleaved in program order.
// If the FP64 pipe (2 cycles per warp instruction) leaves the other issue cycle free, DFMA throughput stays at ~2 per
// clock per SM up to K = 1 and total issue approaches 4 per clock per SM; if not, every integer instruction displaces FP64 work.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/fp64_coissue scripts/fp64_coissue.cu && /tmp/fp64_coissue
#include <cuda_runtime.h>
#include <stdio.h>

template <int K, int KIND>
__global__ void __launch_bounds__(512, 1)
mix_kernel(double * out, int iters, double a, double b, unsigned m)
{
  double x[4];
  unsigned n[4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
    x[i] = threadIdx.x * 1e-9 + i, n[i] = threadIdx.x + i;
  for (int it = 0; it < iters; ++it)
  {
#pragma unroll
    for (int r = 0; r < 16; ++r)
#pragma unroll
      for (int i = 0; i < 4; ++i)
      {
        x[i] = fma(x[i], a, b);
#pragma unroll
        for (int k = 0; k < K; ++k)
        {
          if (KIND == 0)
            n[(i + k) & 3] = (n[(i + k) & 3] ^ m) + 0x9E3779B9u; // LOP3 + IADD (2 integer instructions)
          else if (KIND == 1)
            n[(i + k) & 3] = n[(i + k) & 3] * m + 12345u; // IMAD
          else
            asm volatile("mov.b32 %0, %0;" : "+r"(n[(i + k) & 3])); // nothing the compiler can drop
        }
      }
  }
  double s = 0.0;
  unsigned t = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    s += x[i], t += n[i];
  if (s == 12345.678 || t == 0x12345u)
    out[0] = s + t;
}

template <int K, int KIND>
void
run(int n_sms, double * d_out, int clock_khz, const char * label, double int_per_k)
{
  const int iters = 2048, warps = 16;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  mix_kernel<K, KIND><<<n_sms, warps * 32>>>(d_out, 64, 0.999999, 1e-9, 0x5bd1e995u);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  mix_kernel<K, KIND><<<n_sms, warps * 32>>>(d_out, iters, 0.999999, 1e-9, 0x5bd1e995u);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  const double dfma = (double)warps * iters * 64.0, clk = ms * 1e-3 * clock_khz * 1e3;
  printf("%-6s K=%d  DFMA/clk/SM %.3f  integer/clk/SM %.3f  total %.3f\n", label, K, dfma / clk, dfma * K * int_per_k / clk,
         dfma * (1 + K * int_per_k) / clk);
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
  printf("device %s, %d SMs, %d MHz; 16 warps per SM, 4 DFMA chains per thread + K integer ops per DFMA\n", p.name, p.multiProcessorCount, clock_khz / 1000);
  run<0, 0>(p.multiProcessorCount, d_out, clock_khz, "none", 0);
  run<1, 1>(p.multiProcessorCount, d_out, clock_khz, "imad", 1);
  run<2, 1>(p.multiProcessorCount, d_out, clock_khz, "imad", 1);
  run<3, 1>(p.multiProcessorCount, d_out, clock_khz, "imad", 1);
  run<1, 0>(p.multiProcessorCount, d_out, clock_khz, "lop+add", 2);
  run<2, 0>(p.multiProcessorCount, d_out, clock_khz, "lop+add", 2);
  return 0;
}
