// Microbenchmark: dependent-issue latencies (cycles) of the FP64 operations on the scalar critical paths of the
// eigen-solver kernels (implicit QL recurrence, pivot publication): DFMA, DMUL, sqrt, 1/x, rsqrt, shared-memory
// load-to-use, __syncthreads round trip.  One warp, one chain.  Design input only.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k_lat(double* out, long long* clk, int iters, double seed) {
  __shared__ double sm[64];
  double x = seed + threadIdx.x * 1e-9, y = 1.0000001, z = 0.5;
  sm[threadIdx.x & 63] = x;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (MODE == 0) x = fma(x, y, z);
      else if (MODE == 1) x = x * y;
      else if (MODE == 2) x = sqrt(x) + 1.0;
      else if (MODE == 3) x = 1.0 / x + 1.0;
      else if (MODE == 4) x = rsqrt(x) + 1.0;
      else if (MODE == 5) { sm[threadIdx.x & 63] = x; x = sm[(threadIdx.x + 0) & 63] + 1.0; }
      else if (MODE == 6) { __syncthreads(); x += 1.0; }
      else if (MODE == 7) x = x / y + 1.0;
      else if (MODE == 8) { x += __shfl_xor_sync(0xffffffffu, x, 1); }
    }
  }
  long long t1 = clock64();
  out[threadIdx.x] = x;
  if (threadIdx.x == 0) clk[0] = t1 - t0;
}
int main() {
  double* out; long long* clk; cudaMalloc(&out, 8192); cudaMalloc(&clk, 8);
  const char* names[] = {"DFMA", "DMUL", "sqrt(+DADD)", "1/x(+DADD)", "rsqrt(+DADD)", "STS->LDS(+DADD)", "__syncthreads 256thr(+DADD)", "x/y(+DADD)", "SHFL(+DADD)"};
  const int iters = 2000;
  for (int m = 0; m < 9; ++m) {
    long long h = 0;
    for (int rep = 0; rep < 2; ++rep) {
      int nt = (m == 6) ? 256 : 32;
      switch (m) {
        case 0: k_lat<0><<<1, nt>>>(out, clk, iters, 1.5); break;
        case 1: k_lat<1><<<1, nt>>>(out, clk, iters, 1.5); break;
        case 2: k_lat<2><<<1, nt>>>(out, clk, iters, 1.5); break;
        case 3: k_lat<3><<<1, nt>>>(out, clk, iters, 1.5); break;
        case 4: k_lat<4><<<1, nt>>>(out, clk, iters, 1.5); break;
        case 5: k_lat<5><<<1, nt>>>(out, clk, iters, 1.5); break;
        case 6: k_lat<6><<<1, nt>>>(out, clk, iters, 1.5); break;
        case 7: k_lat<7><<<1, nt>>>(out, clk, iters, 1.5); break;
        case 8: k_lat<8><<<1, nt>>>(out, clk, iters, 1.5); break;
      }
      cudaDeviceSynchronize();
      cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
    }
    printf("%-30s %8.1f cycles per dependent op\n", names[m], (double)h / (iters * 8.0));
  }
  return 0;
}
