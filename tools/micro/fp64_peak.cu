// Microbenchmark: vector DFMA and DMMA (mma.sync m8n8k4 f64) issue rates on one B200 -- design input for the register-tile kernels.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_dfma(double* out, int iters) {
  double a[16]; double x = threadIdx.x * 1e-9, y = 1.0000001;
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = i + x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = fma(a[i], y, x);
  }
  double s = 0; 
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmma(double* out, int iters) {
  double c[8][2]; double a = threadIdx.x * 1e-9, b = 1.0000001;
#pragma unroll
  for (int i = 0; i < 8; ++i) { c[i][0] = i; c[i][1] = -i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int sms = p.multiProcessorCount;
  double* out; cudaMalloc(&out, sizeof(double) * sms * 8 * 1024);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int threads : {128, 256, 512, 1024}) {
    for (int mode = 0; mode < 2; ++mode) {
      int iters = 20000; float ms = 0;
      for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        if (mode == 0) k_dfma<<<sms * 2, threads>>>(out, iters); else k_dmma<<<sms * 2, threads>>>(out, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
      }
      double fl = mode == 0 ? 2.0 * 16 * iters * threads * sms * 2 : 2.0 * 256 * 8 * iters * (threads / 32) * sms * 2;
      printf("%s threads/CTA=%4d x2 CTAs/SM: %.2f TFLOP/s (%.2f ms)\n", mode ? "DMMA m8n8k4" : "DFMA       ", threads, fl / ms / 1e9, ms);
    }
  }
  return 0;
}
