// rcwa_common.cuh -- kernel parameter block and small device helpers shared by every translation unit of the library.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

typedef double2 cplx;

struct KParams {
  int n_cells, N, ff, ld;     // ld: leading dimension of on-chip matrices (odd -> conflict-free columns)
  int B;
  double n_inc, n_exit, eps_g;
  const cplx* tw;             // [n_cells] exp(-2 pi i p / n_cells)
  const cplx* Dk;             // [2N+1]    up-sampling factor / n_cells
  const int8_t* patterns;     // [B][n_cells]
  const double* params;       // [B][4] or [1][4]
  int pstride;
  // chunk workspaces
  double* Cbuf;               // [B][ff][ff]
  double* Lbuf;               // [B][ff][ff]  unit-lower L' below the diagonal, D^-1/2 on it
  double* Wbuf;               // [B][ff][ff]  M0, then W
  double* tri;                // [B][2][ff]
  double* lam;                // [B][ff]
  cplx* rot_cs;               // [B][maxrot]
  uint16_t* sweeps;           // [B][maxsw]  QL sweep descriptors: first column | rotation count << 8
  int* nsw;                   // [B]
  int maxsw;
  int* nrot;                  // [B]
  int maxrot;
  int* wstatus;               // [B] working status bits
  int fallback_only;          // k_tail: process only structures flagged by rk_tail (PIRL_ST_TAIL_FALLBACK)
  unsigned long long* phase_clk;  // optional [32] per-phase clock64 accumulators (debug / profiling), may be null
  unsigned long long* trace;      // optional CTA trace (PIRL_TRACE=1): [0] = record counter, then records of 4 words
  double* scratch;            // per-CTA spill area for matrices that do not fit in shared memory
  size_t scratch_stride;      // doubles per CTA
  int smem_mats;              // how many (kernel-specific) matrices live in shared memory
  int* queue;                 // optional work counter (zeroed before the launch): persistent CTAs fetch structures dynamically
  double* Einv;               // optional [B][ff][ff]: E^-1 exported by the front kernels for the adjoint sweep (null otherwise)
  int pol;                    // 1 = TM (default), 0 = TE (_base.py:142-148): generic front kernels + k_smm only
  int n_layers;               // k_smm: layers per structure (entries b * n_layers + l of the per-layer workspaces)
  int layer_rev;              // k_smm: star products taken in reversed list order (the reference's 'SMM' loop, _base.py:141)
  // outputs
  double* eff; double* de_r3; double* de_t3; int* status;
};

#define PIRL_ST_EIG 1
#define PIRL_ST_NONFINITE_ 2
#define PIRL_ST_NPD 4

__device__ __forceinline__ cplx cmul(cplx a, cplx b) {
  return make_double2(fma(a.x, b.x, -a.y * b.y), fma(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ cplx cfma(cplx a, cplx b, cplx c) {  // a*b + c
  c.x = fma(a.x, b.x, c.x); c.x = fma(-a.y, b.y, c.x);
  c.y = fma(a.x, b.y, c.y); c.y = fma(a.y, b.x, c.y);
  return c;
}
__device__ __forceinline__ cplx cfnma(cplx a, cplx b, cplx c) {  // c - a*b
  c.x = fma(-a.x, b.x, c.x); c.x = fma(a.y, b.y, c.x);
  c.y = fma(-a.x, b.y, c.y); c.y = fma(-a.y, b.x, c.y);
  return c;
}
__device__ __forceinline__ cplx cinv(cplx a) {
  double d = 1.0 / fma(a.x, a.x, a.y * a.y);
  return make_double2(a.x * d, -a.y * d);
}
__device__ __forceinline__ cplx cadd(cplx a, cplx b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ cplx csub(cplx a, cplx b) { return make_double2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ cplx cscale(cplx a, double s) { return make_double2(a.x * s, a.y * s); }

// ---- CTA trace (poor man's timeline; there is no nsys on the box): kernel id, SM id, start / end in globaltimer ns ----
#define PIRL_TRACE_CAP (1 << 20)
__device__ __forceinline__ unsigned long long pirl_gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned pirl_smid() {
  unsigned s;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(s));
  return s;
}
struct CtaTrace {
  unsigned long long t0;
  __device__ __forceinline__ CtaTrace(const unsigned long long* tr) : t0(tr ? pirl_gtime() : 0ull) {}
  __device__ __forceinline__ void done(unsigned long long* tr, int kernel_id, int first) const {
    if (!tr || threadIdx.x != 0) return;
    const unsigned long long slot = atomicAdd(tr, 1ull);
    if (slot >= PIRL_TRACE_CAP) return;
    unsigned long long* r = tr + 1 + 4 * slot;
    r[0] = ((unsigned long long)kernel_id << 32) | pirl_smid();
    r[1] = ((unsigned long long)blockIdx.x << 32) | (unsigned)first;
    r[2] = t0;
    r[3] = pirl_gtime();
  }
};

// next structure of a persistent CTA: static grid stride without a queue, else first item = blockIdx.x, then
// gridDim.x + atomicAdd(queue, 1).  Dynamic fetching keeps late-starting CTAs (SM slots held by a co-running kernel) from
// becoming stragglers.  Must be reached by every thread of the CTA.
__device__ __forceinline__ int pirl_next_item(int* queue, int cur, int* slot) {
  if (!queue) return cur + (int)gridDim.x;
  __syncthreads();
  if (threadIdx.x == 0) *slot = (int)gridDim.x + atomicAdd(queue, 1);
  __syncthreads();
  return *slot;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ cplx warp_sum(cplx v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
    v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
  }
  return v;
}

// ---------------------------------------------------------------------------------------------------
// shared-memory layouts (host and device use the same functions)
// ---------------------------------------------------------------------------------------------------
__host__ __device__ constexpr int pirl_even(int x) { return (x + 1) & ~1; }
// k_front: cre, cim [2N+1 each], rowk, colk, dinv [ff each], pattern bytes
__host__ __device__ inline int front_vec_doubles(int ff, int N, int n_cells) {
  return pirl_even(2 * (2 * N + 1) + 3 * ff + (n_cells + 7) / 8);
}
// k_tridiag: v, pvec, yvec, dd, ee [ff each] + 8 scalars
__host__ __device__ inline int tridiag_vec_doubles(int ff) { return pirl_even(5 * ff + 8); }
// k_tail (in doubles): 11 complex vectors + 3 real rows + 2*16 reduction slots + ints
__host__ __device__ inline int tail_vec_doubles(int ff) { return pirl_even(22 * ff + 3 * ff + 64 + 8); }

// Toeplitz convolution matrix entry in the real basis: idx 2(p-1)=cos_p, 2(p-1)+1=sin_p, 2N=const.
__device__ __forceinline__ double basis_toeplitz(int i, int j, int N, const double* cre, const double* cim) {
  const int z = 2 * N;
  const double r2 = 1.4142135623730951;
  if (i == z && j == z) return cre[0];
  if (i == z || j == z) {
    int o = (i == z) ? j : i;
    int p = (o >> 1) + 1;
    return (o & 1) ? -r2 * cim[p] : r2 * cre[p];
  }
  int p = (i >> 1) + 1, q = (j >> 1) + 1;
  int d = p - q;
  int ad = d < 0 ? -d : d;
  double crd = cre[ad];
  double cid = d >= 0 ? cim[ad] : -cim[ad];
  double crs = cre[p + q], cis = cim[p + q];
  int si = i & 1, sj = j & 1;
  if (!si && !sj) return crd + crs;
  if (si && sj) return crd - crs;
  if (!si && sj) return -cis + cid;
  return -cis - cid;
}

