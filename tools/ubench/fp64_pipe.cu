// Micro-benchmark: what DFMA rate does one B200 SM sustain, alone and mixed with integer work?
// (answers whether "fp64 pipe 50 %" in the ncu capture of blr_mh2_kernel is a hardware roof)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_pipe fp64_pipe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int K, int MIX>
__global__ void __launch_bounds__(256) dfma_kernel(double *out, int iters, double a, double b, int *iout, int im) {
  double x[K];
  int z[K];
#pragma unroll
  for (int k = 0; k < K; ++k) { x[k] = threadIdx.x * 1e-3 + k; z[k] = threadIdx.x + k; }
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
      x[k] = fma(x[k], a, b);
      if (MIX >= 1) z[k] = z[k] * im + i;          // IMAD
      if (MIX >= 2) z[k] = (z[k] & 0xffff) ^ im;   // LOP3
    }
  }
  double s = 0.0; int t = 0;
#pragma unroll
  for (int k = 0; k < K; ++k) { s += x[k]; t += z[k]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  iout[blockIdx.x * blockDim.x + threadIdx.x] = t;
}

template <int K, int MIX>
void run(int ctas_per_sm, int sms, double *out, int *iout) {
  const int iters = 4096;
  int grid = sms * ctas_per_sm;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  dfma_kernel<K, MIX><<<grid, 256>>>(out, iters, 0.999999, 1e-9, iout, 3);
  cudaEventRecord(e0);
  dfma_kernel<K, MIX><<<grid, 256>>>(out, iters, 0.999999, 1e-9, iout, 3);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  double warp_dfma = (double)grid * 8 * iters * K;
  double cycles = ms * 1e-3 * clk * 1e3;
  printf("K=%d mix=%d ctas/SM=%d: %.3f ms, DFMA warp-instr/clk/SM = %.3f (thread-FMA/clk/SM = %.1f), %.2f TFLOP/s\n", K, MIX,
         ctas_per_sm, ms, warp_dfma / cycles / sms, 32 * warp_dfma / cycles / sms, 2 * 32 * warp_dfma / (ms * 1e-3) * 1e-12);
}

int main() {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("SMs %d, clock attr %d kHz\n", sms, clk);
  double *out; int *iout;
  cudaMalloc(&out, sizeof(double) * sms * 8 * 256); cudaMalloc(&iout, sizeof(int) * sms * 8 * 256);
  run<1, 0>(2, sms, out, iout); run<2, 0>(2, sms, out, iout); run<4, 0>(2, sms, out, iout); run<8, 0>(2, sms, out, iout);
  run<8, 0>(1, sms, out, iout); run<8, 0>(4, sms, out, iout); run<1, 0>(8, sms, out, iout);
  run<8, 1>(2, sms, out, iout); run<8, 2>(2, sms, out, iout); run<4, 2>(4, sms, out, iout);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
