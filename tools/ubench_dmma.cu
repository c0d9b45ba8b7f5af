// Micro-benchmark: FP64 mma.sync.m8n8k4 (DMMA) against plain DFMA on one B200, to decide how the regression
// kernel should form its weighted Gram.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_dmma ubench_dmma.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int ILP>
__global__ void k_dmma(double* out, int iters, double a0, double b0) {
  double c[ILP][2];
#pragma unroll
  for (int i = 0; i < ILP; ++i) c[i][0] = c[i][1] = 0.0;
  double a = a0 + threadIdx.x, b = b0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) dmma(c[i][0], c[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void k_dfma(double* out, int iters, double a0, double b0) {
  double c[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) c[i] = i;
  double a = a0, b = b0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) c[i] = fma(a, c[i], b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float timeit(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  f();
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  f();
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms;
}

int main() {
  double* out; cudaMalloc(&out, 148 * 1024 * 8 * 4);
  cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("%s SMs %d clock %d kHz\n", pr.name, pr.multiProcessorCount, clk);
  const int iters = 20000;
  const int sms = pr.multiProcessorCount;
  double best_dmma = 0.0, best_dfma = 0.0;
  for (int warps = 4; warps <= 32; warps *= 2) {
    {
      float ms = timeit([&] { k_dmma<1><<<sms, warps * 32>>>(out, iters, 1.0, 1.0); });
      double n = (double)iters * 1 * warps * sms;
      if (n * 512 / ms / 1e9 > best_dmma) best_dmma = n * 512 / ms / 1e9;
      printf("dmma ilp1 warps/SM %2d: %.3f ms  %.2f TFLOP/s  %.1f clk/instr/SMSP\n", warps, ms, n * 512 / ms / 1e9,
             ms * 1e-3 * 1.965e9 / ((double)iters * 1 * warps / 4));
    }
    {
      float ms = timeit([&] { k_dmma<4><<<sms, warps * 32>>>(out, iters, 1.0, 1.0); });
      double n = (double)iters * 4 * warps * sms;
      if (n * 512 / ms / 1e9 > best_dmma) best_dmma = n * 512 / ms / 1e9;
      printf("dmma ilp4 warps/SM %2d: %.3f ms  %.2f TFLOP/s  %.1f clk/instr/SMSP\n", warps, ms, n * 512 / ms / 1e9,
             ms * 1e-3 * 1.965e9 / ((double)iters * 4 * warps / 4));
    }
    {
      float ms = timeit([&] { k_dfma<1><<<sms, warps * 32>>>(out, iters, 1.0, 1.0); });
      double n = (double)iters * 1 * warps * sms;
      if (n * 64 / ms / 1e9 > best_dfma) best_dfma = n * 64 / ms / 1e9;
      printf("dfma ilp1 warps/SM %2d: %.3f ms  %.2f TFLOP/s  %.1f clk/instr/SMSP\n", warps, ms, n * 64 / ms / 1e9,
             ms * 1e-3 * 1.965e9 / ((double)iters * 1 * warps / 4));
    }
    {
      float ms = timeit([&] { k_dfma<8><<<sms, warps * 32>>>(out, iters, 1.0, 1.0); });
      double n = (double)iters * 8 * warps * sms;
      if (n * 64 / ms / 1e9 > best_dfma) best_dfma = n * 64 / ms / 1e9;
      printf("dfma ilp8 warps/SM %2d: %.3f ms  %.2f TFLOP/s  %.1f clk/instr/SMSP\n", warps, ms, n * 64 / ms / 1e9,
             ms * 1e-3 * 1.965e9 / ((double)iters * 8 * warps / 4));
    }
  }
  // one machine-readable line: the FP64 peaks of this GPU (MEASURED_PEAKS.json carries none)
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"fp64_dfma_tflops\": %.2f, \"fp64_dmma_tflops\": %.2f, "
         "\"how\": \"tools/ubench_dmma.cu: dependent-free DFMA / mma.sync.m8n8k4.f64 chains, 4..32 warps per SM, best of the sweep; 2 flop per FMA\"}\n",
         pr.name, sms, best_dfma, best_dmma);
  return 0;
}
