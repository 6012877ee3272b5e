// fp64_peak.cu -- what the FP64 pipes of one B200 deliver: vector DFMA (independent chains, ILP 8) and tensor DMMA
// (mma.sync.m8n8k4.f64) at several warps per SM; also the latency of a dependent DFMA chain.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak fp64_peak.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_dfma(double* out, int iters, double a, double b) {
  double x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = fma(x[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_chain(double* out, long long* clk, int iters, double a, double b) {
  double x = threadIdx.x * 1e-3;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) x = fma(x, a, b);
  }
  long long t1 = clock64();
  out[threadIdx.x] = x;
  if (threadIdx.x == 0) *clk = t1 - t0;
}

__global__ void k_dmma(double* out, int iters, double a, double b) {
  double c[4][2];
#pragma unroll
  for (int i = 0; i < 4; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int sms = p.multiProcessorCount; int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  printf("%s: %d SMs, max clock %.0f MHz\n", p.name, sms, khz / 1e3);
  double* out; cudaMalloc(&out, sizeof(double) * sms * 4 * 1024);
  long long* clk; cudaMalloc(&clk, 8);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int nt : {128, 256, 512, 1024}) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0); k_dfma<<<sms, nt>>>(out, iters, 1.0000001, 1e-9); cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      if (rep) printf("DFMA  %2d warps/SM: %7.2f TFLOP/s  (%.1f lanes/clk/SM at max clock)\n", nt / 32,
                      2.0 * iters * 8 * nt * sms / ms / 1e9, (double)iters * 8 * nt / (ms * 1e-3 * khz * 1e3));
    }
  }
  for (int nt : {128, 256, 512, 1024}) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0); k_dmma<<<sms, nt>>>(out, iters, 1.0000001, 1e-9); cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      // one m8n8k4 = 8*8*4 multiply-adds per warp
      if (rep) printf("DMMA  %2d warps/SM: %7.2f TFLOP/s\n", nt / 32, 2.0 * iters * 4 * 256 * (nt / 32) * sms / ms / 1e9);
    }
  }
  k_chain<<<1, 32>>>(out, clk, 1000, 1.0000001, 1e-9); cudaDeviceSynchronize();
  long long h; cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
  printf("dependent DFMA latency: %.1f clk\n", h / 8000.0);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
