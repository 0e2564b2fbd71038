// fp64_peak.cu — measures the FP64 (DFMA) issue ceiling of the device, the denominator of the
// "FP64 pipe fraction" DESIGN.md reports beside the HBM-model roofline (the number is
// not in MEASURED_PEAKS.json and has to be measured on the box).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak scripts/fp64_peak.cu && ./fp64_peak
// Prints, for W warps per SM sub-partition and ILP independent DFMA chains per thread, the achieved
// DFMA rate in warp-instructions per clock per SM and TFLOP/s; the last line is the best.
#include <cuda_runtime.h>
#include <stdio.h>

// OP: 0 = DFMA, 1 = DADD, 2 = DMUL, 3 = DMUL then DADD alternating (same chain)
template <int ILP, int OP = 0>
__global__ void __launch_bounds__(1024, 1)
dfma_kernel(double * out, int iters, double a, double b)
{
  double x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i)
    x[i] = threadIdx.x * 1e-9 + i;
  for (int it = 0; it < iters; ++it)
  {
#pragma unroll
    for (int r = 0; r < 16; ++r)
#pragma unroll
      for (int i = 0; i < ILP; ++i)
        x[i] = OP == 0 ? fma(x[i], a, b) : OP == 1 ? __dadd_rn(x[i], b) : OP == 2 ? __dmul_rn(x[i], a) : ((r & 1) ? __dadd_rn(x[i], b) : __dmul_rn(x[i], a));
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < ILP; ++i)
    s += x[i];
  if (s == 12345.678)
    out[0] = s;
}

template <int ILP, int OP = 0>
double
run(int warps_per_sm, int n_sms, double * d_out, double * inst_per_clk_sm, int clock_khz)
{
  const int iters = 4096;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  dfma_kernel<ILP, OP><<<n_sms, warps_per_sm * 32>>>(d_out, 64, 0.999999, 1e-9);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  dfma_kernel<ILP, OP><<<n_sms, warps_per_sm * 32>>>(d_out, iters, 0.999999, 1e-9);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  const double warp_inst = (double)n_sms * warps_per_sm * iters * 16.0 * ILP;
  const double sec = ms * 1e-3;
  *inst_per_clk_sm = warp_inst / n_sms / (sec * clock_khz * 1e3);
  return warp_inst * 32 * 2 / sec / 1e12;
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
  printf("device %s, %d SMs, max clock %d MHz\n", p.name, p.multiProcessorCount, clock_khz / 1000);
  printf("%6s %4s %14s %10s\n", "warps", "ILP", "inst/clk/SM", "TFLOP/s");
  double best = 0.0;
  const int ws[] = {4, 8, 16, 20, 32};
  for (int w : ws)
  {
    double ipc, tf;
#define RUN(I)                                                     \
  tf = run<I>(w, p.multiProcessorCount, d_out, &ipc, clock_khz); \
  printf("%6d %4d %14.4f %10.3f\n", w, I, ipc, tf);              \
  if (tf > best)                                                   \
    best = tf;
    RUN(1) RUN(2) RUN(3) RUN(4) RUN(8)
  }
  // the same ceiling for the other FP64 instructions (warp-instructions per clock per SM; a "flop" column would
  // count DADD / DMUL as one): is a DFMA dearer than a DADD or DMUL on this pipe?
  const char * names[] = {"DFMA", "DADD", "DMUL", "DMUL/DADD"};
  for (int w : {4, 16})
  {
    double ipc[4][3];
    run<1, 0>(w, p.multiProcessorCount, d_out, &ipc[0][0], clock_khz); run<2, 0>(w, p.multiProcessorCount, d_out, &ipc[0][1], clock_khz); run<4, 0>(w, p.multiProcessorCount, d_out, &ipc[0][2], clock_khz);
    run<1, 1>(w, p.multiProcessorCount, d_out, &ipc[1][0], clock_khz); run<2, 1>(w, p.multiProcessorCount, d_out, &ipc[1][1], clock_khz); run<4, 1>(w, p.multiProcessorCount, d_out, &ipc[1][2], clock_khz);
    run<1, 2>(w, p.multiProcessorCount, d_out, &ipc[2][0], clock_khz); run<2, 2>(w, p.multiProcessorCount, d_out, &ipc[2][1], clock_khz); run<4, 2>(w, p.multiProcessorCount, d_out, &ipc[2][2], clock_khz);
    run<1, 3>(w, p.multiProcessorCount, d_out, &ipc[3][0], clock_khz); run<2, 3>(w, p.multiProcessorCount, d_out, &ipc[3][1], clock_khz); run<4, 3>(w, p.multiProcessorCount, d_out, &ipc[3][2], clock_khz);
    for (int o = 0; o < 4; ++o)
      printf("op %-10s warps %2d  inst/clk/SM at ILP 1/2/4: %.4f %.4f %.4f\n", names[o], w, ipc[o][0], ipc[o][1], ipc[o][2]);
  }
  printf("fp64_peak_tflops %.3f (nominal clock; DFMA = 2 flop)\n", best);
  return 0;
}
