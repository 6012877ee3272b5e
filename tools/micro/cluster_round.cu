// cluster_round.cu -- cost of one all-gather "round" inside a 4-CTA cluster on B200 (clocks per round), for the design of
// the cluster kernels (rcwa_cluster.cuh).  Variants:
//   0  st.async + mbarrier complete_tx, every publisher stores to all 4 CTAs (what CkPipe does)
//   1  st.shared::cluster + barrier.cluster.arrive.release / wait.acquire
//   2  st.async, remote CTAs only; own CTA by st.shared + a second (CTA-local) arrival
//   3  like 0 but one payload double per CTA (latency floor)
//   4  no communication: __syncthreads only (floor of the loop itself)
//   5  publishers write a local staging segment, __syncthreads, ONE thread sends it to the 3 other CTAs with
//      cp.async.bulk.shared::cluster.shared::cta (DSMEM bulk copy, complete_tx on the destination's mbarrier)
//   6  like 0 with 16-byte stores (st.async .v2.f64): half the store instructions
// Each round: NPUB publisher threads per CTA send VALS doubles each to every CTA, all threads wait, read one value, sync.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o cluster_round cluster_round.cu ; run: ./cluster_round
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int NCTA = 4, NT = 352;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t mapa(uint32_t a, uint32_t r) { uint32_t o; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(o) : "r"(a), "r"(r)); return o; }
__device__ __forceinline__ uint32_t ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void csync() { asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }

template <int VARIANT>
__global__ void __launch_bounds__(NT, 1) k_round(int rounds, int npub, int vals, long long* out, double* sink) {
  __shared__ __align__(16) double buf[2][512];
  __shared__ __align__(8) unsigned long long mbar[2];
  const int tid = threadIdx.x;
  const uint32_t rank = ctarank();
  uint32_t rb[NCTA], rm[NCTA];
  for (int r = 0; r < NCTA; ++r) { rb[r] = mapa(smem_u32(&buf[0][0]), r); rm[r] = mapa(smem_u32(&mbar[0]), r); }
  const uint32_t mbl = smem_u32(&mbar[0]);
  const int narr = (VARIANT == 2) ? 2 : 1;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbl), "r"(narr) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbl + 8u), "r"(narr) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = tid; i < 1024; i += NT) (&buf[0][0])[i] = 0.0;
  csync();
  double acc = 0.0;
  const bool pub = tid % 27 == 0 && tid / 27 < npub;          // publishers spread over the warps
  const int slot0 = ((int)rank * npub + tid / 27) * vals;
  const uint32_t bytes = (VARIANT == 2 || VARIANT == 5) ? 8u * (NCTA - 1) * npub * vals : (VARIANT == 3) ? 8u * NCTA : 8u * NCTA * npub * vals;
  long long t0 = clock64();
  for (int k = 0; k < rounds; ++k) {
    const int b = k & 1;
    const double v = acc + k;
    if (VARIANT == 0 || VARIANT == 3) {
      if (pub && (VARIANT == 0 || tid == 0)) {
        const int nv = (VARIANT == 3) ? 1 : vals;
        for (int q = 0; q < nv; ++q)
#pragma unroll
          for (int r = 0; r < NCTA; ++r)
            asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f64 [%0], %1, [%2];" ::"r"(rb[r] + 8u * (b * 512 + slot0 + q)), "d"(v), "r"(rm[r] + 8u * b) : "memory");
      }
    } else if (VARIANT == 2) {
      if (pub) {
        for (int q = 0; q < vals; ++q) {
          buf[b][slot0 + q] = v;
#pragma unroll
          for (int r = 0; r < NCTA; ++r)
            if (r != (int)rank) asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f64 [%0], %1, [%2];" ::"r"(rb[r] + 8u * (b * 512 + slot0 + q)), "d"(v), "r"(rm[r] + 8u * b) : "memory");
        }
      }
      __syncthreads();          // local part written
      if (tid == 32) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(mbl + 8u * b) : "memory");
    } else if (VARIANT == 5) {
      if (pub) for (int q = 0; q < vals; ++q) buf[b][slot0 + q] = v;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      if (tid < NCTA && tid != (int)rank) {
        const uint32_t off = 8u * (b * 512 + (int)rank * npub * vals);
        asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(rb[tid] + off),
                     "r"(smem_u32(&buf[0][0]) + off), "r"(8u * npub * vals), "r"(rm[tid] + 8u * b)
                     : "memory");
      }
    } else if (VARIANT == 6) {
      if (pub)
        for (int q = 0; q < vals; q += 2)
#pragma unroll
          for (int r = 0; r < NCTA; ++r)
            asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f64 [%0], {%1, %2}, [%3];" ::"r"(rb[r] + 8u * (b * 512 + slot0 + q)), "d"(v), "d"(v + 1.0), "r"(rm[r] + 8u * b) : "memory");
    } else if (VARIANT == 1) {
      if (pub)
        for (int q = 0; q < vals; ++q)
#pragma unroll
          for (int r = 0; r < NCTA; ++r) asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(rb[r] + 8u * (b * 512 + slot0 + q)), "d"(v) : "memory");
    }
    if (VARIANT == 1) csync();
    else if (VARIANT != 4) {
      if (tid == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbl + 8u * b), "r"(bytes) : "memory");
      const uint32_t parity = (k >> 1) & 1;
      uint32_t done = 0;
      int spins = 0;
      do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(mbl + 8u * b), "r"(parity) : "memory");
        if (!done && ++spins > (1 << 22)) __trap();
      } while (!done);
    }
    acc += buf[b][(tid * 7) & 255];
    __syncthreads();
  }
  long long t1 = clock64();
  csync();
  if (tid == 0 && rank == 0 && blockIdx.x == 0) out[0] = (t1 - t0) / rounds;
  if (acc == 12345.678) sink[0] = acc;
}

template <int V>
static void run(const char* name, int npub, int vals, long long* d_out, double* d_sink) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(NCTA * 33); cfg.blockDim = dim3(NT);
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = NCTA; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  const int rounds = 2000;
  for (int rep = 0; rep < 2; ++rep) {
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_round<V>, rounds, npub, vals, d_out, d_sink);
    if (e != cudaSuccess) { printf("%s: launch %s\n", name, cudaGetErrorString(e)); return; }
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
  }
  long long h = 0;
  cudaMemcpy(&h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-58s npub %2d vals %d : %5lld clk / round\n", name, npub, vals, h);
}

int main() {
  long long* d_out; double* d_sink;
  cudaMalloc(&d_out, 8); cudaMalloc(&d_sink, 8);
  run<4>("4 no communication (__syncthreads only)", 13, 4, d_out, d_sink);
  run<3>("3 st.async, one double per CTA (latency floor)", 13, 4, d_out, d_sink);
  run<0>("0 st.async to all 4 CTAs + mbarrier", 13, 4, d_out, d_sink);
  run<0>("0 st.async to all 4 CTAs + mbarrier", 13, 1, d_out, d_sink);
  run<0>("0 st.async to all 4 CTAs + mbarrier", 4, 4, d_out, d_sink);
  run<6>("6 st.async.v2.f64 to all 4 CTAs + mbarrier", 13, 4, d_out, d_sink);
  run<5>("5 local staging + cp.async.bulk to the 3 other CTAs", 13, 4, d_out, d_sink);
  run<5>("5 local staging + cp.async.bulk to the 3 other CTAs", 13, 2, d_out, d_sink);
  run<2>("2 st.async remote only, local st.shared + arrive", 13, 4, d_out, d_sink);
  run<1>("1 st.shared::cluster + barrier.cluster", 13, 4, d_out, d_sink);
  run<1>("1 st.shared::cluster + barrier.cluster", 13, 1, d_out, d_sink);
  run<1>("1 st.shared::cluster + barrier.cluster", 1, 1, d_out, d_sink);
  return 0;
}
