// rcwa_cluster.cuh -- thread-block-CLUSTER register-tile kernels for the large Fourier orders (sm_100a).
//
// Orders 61..103 (122 <= ff = 2N+1 <= 208; BASELINE config C5 is order 100, ff = 201) do not fit one SM: a real ff x ff
// matrix is 323 KB at ff = 201, more than the register file (256 KB) or the shared memory (227 KB) of an SM.  Here ONE
// structure is solved by a cluster of NCTA = 4 CTAs on four SMs:
//   * the FF x FF matrix (FF = 208, identity-padded) lives in the REGISTERS of the cluster: thread (ti, tj) of a TR x TC grid
//     (TR = 52 thread rows, 13 per CTA; TC = 26) owns the cyclic tile {(ti + TR a, tj + TC b) : a < 4, b < 8};
//   * the cluster is split by THREAD ROW, so a pivot COLUMN is produced by all four CTAs and is all-gathered through
//     distributed shared memory (st.shared::cluster into the same offset of every CTA, then barrier.cluster); every
//     factorisation is written in its symmetric form so that only columns are ever published (LDL^T, the symmetric sweep
//     operator for E^-1, Householder), row-side quantities (partial row sums, y = T v) never leave their CTA;
//   * the one genuinely row-oriented step (X <- L^-1 X of the reduction to standard form) is replaced by a column pass, a
//     transpose of the distributed tile through an L2-resident scratch, and a second column pass (C is symmetric);
//   * the unit-lower factor L of A is needed by every CTA and is kept as a packed triangle (174 KB) in each CTA's shared
//     memory; the Householder reflectors go to an L2-resident scratch and are prefetched one step ahead.
// Same mathematics as rk_front (rcwa_regtile.cuh); replaces convolution_matrix.py:89-114 (x2) and _base.py:151-155 of the
// reference up to the reduction of the eigenproblem.
#pragma once
#include "rcwa_common.cuh"

template <int FF_, int NCTA_, int TRL_, int TC_, int RT_, int CT_>
struct ClusterCfg {
  static constexpr int FF = FF_, NCTA = NCTA_, TRL = TRL_, TC = TC_, RT = RT_, CT = CT_;
  static constexpr int TR = NCTA * TRL;
  static constexpr int NTHR = TRL * TC;
  static constexpr int NT = ((NTHR + 31) / 32) * 32;
  static constexpr int KA_PER_KB = TR / TC;
  static constexpr int LROWS = TRL * RT;                 // matrix rows owned by one CTA
  static constexpr int FE = pirl_even(FF);
  static constexpr int RW = TC + 1;                      // row stride of the partial-sum buffer
  static constexpr int RS_LANES = 4;                     // lanes that share one row sum
  static constexpr int RS_THREADS = ((RS_LANES * LROWS + 31) / 32) * 32;
  static constexpr int RS_WARPS = RS_THREADS / 32;
  static_assert(TR * RT == FF && TC * CT == FF, "exact tiling only");
  static_assert(TR % TC == 0, "TR must be a multiple of TC");
  static_assert(RS_THREADS <= NT, "row sums need 4 lanes per local row");
  static_assert(NCTA * RS_WARPS <= 32, "one warp reduces the partial dot products of the cluster");
};

// ---- cluster primitives (PTX: mapa / st.shared::cluster / barrier.cluster) ----
__device__ __forceinline__ uint32_t ck_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t ck_mapa(uint32_t addr, uint32_t rank) {
  uint32_t o;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(o) : "r"(addr), "r"(rank));
  return o;
}
__device__ __forceinline__ void ck_st_cluster(uint32_t addr, double v) {
  asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ void ck_st_cluster_i32(uint32_t addr, int v) {
  asm volatile("st.shared::cluster.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ck_cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void ck_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// ---- cluster pipeline: data rounds signalled by mbarrier transaction counts instead of barrier.cluster ----
// A "round" is one all-gather (every CTA stores its part of a vector into the same offsets of every CTA).  The stores are
// st.async: each one carries its own completion (complete_tx of 8 bytes) to an mbarrier in the DESTINATION CTA, so a consumer
// waits on its local mbarrier for the byte count of the round -- one DSMEM store latency plus the mbarrier wake-up, instead
// of a release fence + barrier.cluster round trip (measured 1100 -> 400 clocks per elimination step).  Two mbarriers are used
// alternately (round r -> mbarrier r & 1, phase parity (r >> 1) & 1).
// Rule that makes two data buffers per kind sufficient: a CTA issues its stores of round r + 1 only after a CTA barrier
// that follows its threads' reads of round r.  (Round r + 2 data can only be sent by a CTA that has received all of round
// r + 1, so it never lands on a buffer that is still being read.)
template <int NCTA>
struct CkPipe {
  uint32_t rb[NCTA];      // shared::cluster address of sm[] in every CTA of the cluster
  uint32_t mbl;           // shared::cta address of this CTA's mbarrier pair
  uint32_t mbo;           // byte offset of the mbarrier pair from sm[]
  uint32_t rnd;           // rounds completed (uniform over the cluster)
  __device__ __forceinline__ void init(double* sm, int mbar_off_doubles, int tid) {
#pragma unroll
    for (int r = 0; r < NCTA; ++r) rb[r] = ck_mapa(ck_smem_u32(sm), (uint32_t)r);
    mbo = 8u * (uint32_t)mbar_off_doubles;
    mbl = ck_smem_u32(sm) + mbo;
    rnd = 0;
    if (tid == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbl) : "memory");
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbl + 8u) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  }
  // store v into sm[off] of every CTA, counted on the mbarrier of the round that the next wait() completes
  __device__ __forceinline__ void pub(int off, double v) const {
    const uint32_t mb = mbo + 8u * (rnd & 1u);
#pragma unroll
    for (int r = 0; r < NCTA; ++r)
      asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f64 [%0], %1, [%2];" ::"r"(rb[r] + 8u * (uint32_t)off),
                   "d"(v), "r"(rb[r] + mb)
                   : "memory");
  }
  // plain remote store (no completion): for data that a later barrier.cluster orders
  __device__ __forceinline__ void bcast(int off, double v) const {
#pragma unroll
    for (int r = 0; r < NCTA; ++r) ck_st_cluster(rb[r] + 8u * (uint32_t)off, v);
  }
  // wait until `bytes` of the current round have arrived in THIS CTA; every thread of the CTA calls it
  __device__ __forceinline__ void wait(uint32_t bytes, int tid) {
    const uint32_t mb = mbl + 8u * (rnd & 1u);
    if (tid == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(bytes) : "memory");
    const uint32_t parity = (rnd >> 1) & 1u;
    uint32_t done = 0;
    int spins = 0;
    do {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(mb), "r"(parity)
          : "memory");
      if (!done && ++spins > (1 << 24)) __trap();      // a lost round would otherwise hang the GPU (protocol error, never data dependent)
    } while (!done);
    ++rnd;
  }
};

// shared-memory carve-up of ck_front (in doubles; the same offsets in every CTA of the cluster)
template <class C>
struct CFrontSmem {
  static constexpr int FE = C::FE;
  static constexpr int cre = 0;
  static constexpr int cim = cre + FE;
  static constexpr int colbuf = cim + FE;                 // [2][FE] all-gathered pivot column
  static constexpr int c2buf = colbuf + 2 * FE;           // [2][FE] column k+1 (rows of this CTA only)
  static constexpr int pvec = c2buf + 2 * FE;             // [2][FE]
  static constexpr int pinvA = pvec + 2 * FE;
  static constexpr int disA = pinvA + FE;
  static constexpr int betaS = disA + FE;
  static constexpr int scal = betaS + FE;                 // [2][4] pivot scalars
  static constexpr int kkp = scal + 8;                    // [2][32] partial v.p of the row-sum warps of the cluster
  static constexpr int ints = kkp + 64;                   // [8] ints: next item, bad flags [NCTA]
  static constexpr int mbar = ints + 8;                   // two mbarriers (CkPipe)
  static constexpr int red = mbar + 2;                    // [LROWS][RW]
  static constexpr int LT = pirl_even(red + C::LROWS * C::RW);     // packed unit-lower factor, column major
  static constexpr int tw = LT + C::FF * (C::FF + 1) / 2;          // [n_cells] complex (runtime), then pattern bits
  static __host__ __device__ constexpr int doubles(int n_cells) { return pirl_even(tw) + 2 * n_cells + pirl_even((n_cells + 7) / 8); }
};

// per-cluster L2-resident scratch of ck_front (doubles): transpose buffer, reflectors
template <class C>
struct CFrontScratch {
  static constexpr size_t X = 0;
  static constexpr size_t V = (size_t)C::FF * C::FF;
  static constexpr size_t doubles = 2 * (size_t)C::FF * C::FF;
};

#define CK_PHASE(n)                                                                       \
  do {                                                                                    \
    if (P.phase_clk && tid == 0 && crank == 0) {                                          \
      const long long now_ = clock64();                                                   \
      atomicAdd(&P.phase_clk[n], (unsigned long long)(now_ - pc_));                       \
      pc_ = now_;                                                                         \
    }                                                                                     \
  } while (0)

template <class C>
__global__ void __launch_bounds__(C::NT, 1) ck_front(KParams P, double* __restrict__ cscratch) {
  constexpr int FF = C::FF, NCTA = C::NCTA, TRL = C::TRL, TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, NT = C::NT;
  constexpr int FE = C::FE, RW = C::RW, LROWS = C::LROWS;
  using S = CFrontSmem<C>;
  extern __shared__ __align__(16) double sm[];
  const int FFR = P.ff, N = P.N, z = 2 * N, nc = P.n_cells;
  double* cre = sm + S::cre;
  double* cim = sm + S::cim;
  double* pinvA = sm + S::pinvA;
  double* disA = sm + S::disA;
  double* betaS = sm + S::betaS;
  int* ints = reinterpret_cast<int*>(sm + S::ints);
  double* red = sm + S::red;
  double* LT = sm + S::LT;
  cplx* twS = reinterpret_cast<cplx*>(sm + pirl_even(S::tw));
  uint8_t* su = reinterpret_cast<uint8_t*>(sm + pirl_even(S::tw) + 2 * nc);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int crank = (int)ck_cluster_ctarank();
  const int cid = blockIdx.x / NCTA, nclusters = gridDim.x / NCTA;
  const bool act = tid < C::NTHR;
  const int til = act ? tid / TC : TRL - 1;
  const int tj = act ? tid % TC : TC - 1;
  const int ti = crank * TRL + til;
  CkPipe<NCTA> pipe;
  pipe.init(sm, S::mbar, tid);
  auto ltoff = [](int k) { return k * FF - (k * (k - 1)) / 2; };     // column k of the packed factor: entries i = k..FF-1
  double* Xg = cscratch + (size_t)cid * CFrontScratch<C>::doubles + CFrontScratch<C>::X;
  double* Vg = cscratch + (size_t)cid * CFrontScratch<C>::doubles + CFrontScratch<C>::V;

  for (int t = tid; t < nc; t += NT) twS[t] = P.tw[t];
  // padding rows / columns (>= ff) stay identity: unit pivots, zero columns of L (never written again)
  for (int i = tid; i < FF; i += NT) { pinvA[i] = 1.0; disA[i] = 1.0; betaS[i] = 0.0; }
  for (int i = tid; i < FF * (FF + 1) / 2; i += NT) LT[i] = 0.0;
  ck_cluster_sync();       // every CTA of the cluster is resident before the first remote store

  for (int sidx = cid; sidx < P.B;) {
    const double* pr = P.params + (size_t)sidx * P.pstride;
    const double sa = sin(pr[1] * (3.14159265358979323846 / 180.0));
    const double nr = pr[3];
    const double er = nr * nr, eg = P.eps_g;
    int bad = 0;
    long long pc_ = clock64();
    for (int p = tid; p < nc; p += NT) su[p] = P.patterns[(size_t)sidx * nc + p] > 0;
    __syncthreads();
    // ---- Fourier coefficients c_k, k = 0..2N (closed form of the up-sampled FFT): one thread per k, every CTA its own copy ----
    for (int k = tid; k <= 2 * N; k += NT) {
      double sx = 0.0, sy = 0.0;
      int idx = 0;
      const int km = k % nc;
      for (int p = 0; p < nc; ++p) {
        if (su[p]) { const cplx w = twS[idx]; sx += w.x; sy += w.y; }
        idx += km; if (idx >= nc) idx -= nc;
      }
      const cplx d = __ldg(&P.Dk[k]);
      cre[k] = sx * d.x - sy * d.y;
      cim[k] = sx * d.y + sy * d.x;
    }
    ck_cluster_sync();     // also: the previous structure's remote traffic into this CTA is complete before new publishes
    CK_PHASE(0);

    double t[RT][CT];
    // =========================== phase 1: A' = P A P, LDL^T (column k all-gathered) ===========================
    {
      const double da = 1.0 / er - 1.0 / eg, ia = 1.0 / eg;
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b) {
          const int i = ti + TR * a, j = tj + TC * b;
          const int pi = (i < z) ? (i ^ 1) : i, pj = (j < z) ? (j ^ 1) : j;
          if (i >= FFR || j >= FFR) t[a][b] = (i == j) ? 1.0 : 0.0;
          else t[a][b] = fma(da, basis_toeplitz(pi, pj, N, cre, cim), (i == j) ? ia : 0.0);
        }
    }
    // Pipelined rounds (CkPipe): column k + 1 is updated FIRST by its owners and sent off, then everybody applies the bulk of
    // step k while those stores are in flight.  publish_col(k): the owners of column k (tile column KBC) all-gather it into
    // colbuf[k & 1]; the owner of the pivot adds 1/d and d^-1/2.
#define CK_LDLT_PUBLISH(KBC, KK)                                                                              \
  do {                                                                                                        \
    const int kac_ = (KBC) / C::KA_PER_KB;                                                                    \
    const int cbo_ = S::colbuf + ((KK) & 1) * FE;                                                             \
    _Pragma("unroll") for (int a = 0; a < RT; ++a) if (a >= kac_) pipe.pub(cbo_ + ti + TR * a, t[a][KBC]);    \
    if (ti == (KK) - kac_ * TR) {                                                                             \
      const double piv_ = t[kac_][KBC];                                                                       \
      if (!(piv_ > 0.0)) bad = 1;                                                                             \
      pipe.pub(S::pinvA + (KK), 1.0 / piv_);                                                                  \
      pipe.pub(S::disA + (KK), rsqrt(piv_));                                                                  \
    }                                                                                                         \
  } while (0)
    if (act && tj == 0) CK_LDLT_PUBLISH(0, 0);
#pragma unroll
    for (int kb = 0; kb < CT; ++kb) {
      const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
      for (int kj = 0; kj < TC; ++kj) {
        const int k = kb * TC + kj;
        if (k >= FFR) break;
        pipe.wait(8u * (uint32_t)((RT - ka) * TR + 2), tid);
        const double* cb = sm + S::colbuf + (k & 1) * FE;
        const double p = pinvA[k];
        double li[RT], cj[CT];
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) { const int i = ti + TR * a; li[a] = (i > k) ? -cb[i] * p : 0.0; }
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= kb) { const int j = tj + TC * b; cj[b] = (j > k) ? cb[j] : 0.0; }
        // unit-lower column k into this CTA's packed copy of L
        for (int i = k + 1 + tid; i < FFR; i += NT) LT[ltoff(k) + (i - k)] = cb[i] * p;
        __syncthreads();       // every thread of this CTA has consumed round k: the next round may be sent
        if (k + 1 < FFR && act) {       // look-ahead: column k + 1 first, published before the bulk of the update
          if (kj + 1 < TC) {
            if (tj == kj + 1) {
#pragma unroll
              for (int a = 0; a < RT; ++a) if (a >= ka) t[a][kb] = fma(li[a], cj[kb], t[a][kb]);
              cj[kb] = 0.0;
              CK_LDLT_PUBLISH(kb, k + 1);
            }
          } else if (kb + 1 < CT) {
            if (tj == 0) {
#pragma unroll
              for (int a = 0; a < RT; ++a) if (a >= ka) t[a][(kb + 1 < CT) ? kb + 1 : kb] = fma(li[a], cj[(kb + 1 < CT) ? kb + 1 : kb], t[a][(kb + 1 < CT) ? kb + 1 : kb]);
              cj[(kb + 1 < CT) ? kb + 1 : kb] = 0.0;
              CK_LDLT_PUBLISH(((kb + 1 < CT) ? kb + 1 : kb), k + 1);
            }
          }
        }
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= kb) {
#pragma unroll
            for (int a = 0; a < RT; ++a)
              if (a >= ka) t[a][b] = fma(li[a], cj[b], t[a][b]);
          }
      }
    }
#undef CK_LDLT_PUBLISH
    ck_cluster_sync();     // every CTA has read the last column before phase 2 publishes into the same buffer
    CK_PHASE(1);
    // =========================== phase 2: E -> -E^-1 by symmetric sweeps (column k all-gathered), B' ===========================
    {
      const double de = er - eg;
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b) {
          const int i = ti + TR * a, j = tj + TC * b;
          if (i >= FFR || j >= FFR) t[a][b] = (i == j) ? 1.0 : 0.0;
          else t[a][b] = fma(de, basis_toeplitz(i, j, N, cre, cim), (i == j) ? eg : 0.0);
        }
    }
    // published column: d - 1 at the pivot (u_k = -(d - 1) p = p - 1 and v_k = d - 1 then need no special case on the reader
    // side); the pivot owner adds 1/d.  Same look-ahead pipeline as phase 1.
#define CK_SWEEP_PUBLISH(KBC, KK)                                                                             \
  do {                                                                                                        \
    const int kac_ = (KBC) / C::KA_PER_KB;                                                                    \
    const int cbo_ = S::colbuf + ((KK) & 1) * FE;                                                             \
    _Pragma("unroll") for (int a = 0; a < RT; ++a) {                                                          \
      const int i_ = ti + TR * a;                                                                             \
      pipe.pub(cbo_ + i_, (i_ == (KK)) ? t[a][KBC] - 1.0 : t[a][KBC]);                                        \
    }                                                                                                         \
    if (ti == (KK) - kac_ * TR) {                                                                             \
      const double piv_ = t[kac_][KBC];                                                                       \
      if (!(piv_ > 0.0)) bad = 1;                                                                             \
      pipe.pub(S::scal + ((KK) & 1) * 4, 1.0 / piv_);                                                         \
    }                                                                                                         \
  } while (0)
    if (act && tj == 0) CK_SWEEP_PUBLISH(0, 0);
#pragma unroll
    for (int kb = 0; kb < CT; ++kb) {
      const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
      for (int kj = 0; kj < TC; ++kj) {
        const int k = kb * TC + kj;
        if (k >= FFR) break;
        const int kti = k - ka * TR;
        pipe.wait(8u * (uint32_t)(FF + 1), tid);
        const double* cb = sm + S::colbuf + (k & 1) * FE;
        const double p = sm[S::scal + (k & 1) * 4];
        double u[RT], v[CT];
#pragma unroll
        for (int a = 0; a < RT; ++a) u[a] = -cb[ti + TR * a] * p;
#pragma unroll
        for (int b = 0; b < CT; ++b) v[b] = cb[tj + TC * b];
        __syncthreads();       // round k consumed by every thread of this CTA
        if (k + 1 < FFR && act) {
          if (kj + 1 < TC) {
            if (tj == kj + 1) {
#pragma unroll
              for (int a = 0; a < RT; ++a) t[a][kb] = fma(u[a], v[kb], t[a][kb]);
              v[kb] = 0.0;
              CK_SWEEP_PUBLISH(kb, k + 1);
            }
          } else if (kb + 1 < CT) {
            if (tj == 0) {
#pragma unroll
              for (int a = 0; a < RT; ++a) t[a][(kb + 1 < CT) ? kb + 1 : kb] = fma(u[a], v[(kb + 1 < CT) ? kb + 1 : kb], t[a][(kb + 1 < CT) ? kb + 1 : kb]);
              v[(kb + 1 < CT) ? kb + 1 : kb] = 0.0;
              CK_SWEEP_PUBLISH(((kb + 1 < CT) ? kb + 1 : kb), k + 1);
            }
          }
        }
#pragma unroll
        for (int b = 0; b < CT; ++b)
#pragma unroll
          for (int a = 0; a < RT; ++a) t[a][b] = fma(u[a], v[b], t[a][b]);
        if (ti == kti && tj == kj) t[ka][kb] = -p;
      }
    }
#undef CK_SWEEP_PUBLISH
    ck_cluster_sync();     // phase boundary: every CTA has left the pipelined rounds
    if (P.Einv && act) {   // adjoint sweep needs E^-1 (original harmonic order)
      double* Eg = P.Einv + (size_t)sidx * FFR * FFR;
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (ti + TR * a < FFR && tj + TC * b < FFR) Eg[(ti + TR * a) * FFR + tj + TC * b] = -t[a][b];
    }
    // B'_ij = s_i s_j Einv_ij - delta_ij with s = (swapped) normalised kx: even slot -> -p sin(angle), odd slot -> +p sin(angle)
    {
      double si[RT], sj[CT];
#pragma unroll
      for (int a = 0; a < RT; ++a) { const int i = ti + TR * a; const double kp = ((i >> 1) + 1) * sa; si[a] = (i >= z) ? 0.0 : ((i & 1) ? kp : -kp); }
#pragma unroll
      for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; const double kp = ((j >> 1) + 1) * sa; sj[b] = (j >= z) ? 0.0 : ((j & 1) ? kp : -kp); }
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b) {
          const int i = ti + TR * a, j = tj + TC * b;
          t[a][b] = -si[a] * sj[b] * t[a][b] - ((i == j) ? 1.0 : 0.0);
        }
    }
    CK_PHASE(2);
    // =========================== phase 3: C = D^-1/2 L^-1 B' L^-T D^-1/2 ===========================
    // two column passes X <- X L^-T (only this CTA's rows of the pivot column are needed: no cluster traffic) with a
    // transpose of the distributed tile through the L2-resident scratch in between (C is symmetric)
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
      __syncthreads();
#pragma unroll
      for (int kb = 0; kb < CT; ++kb) {
#pragma unroll 1
        for (int kj = 0; kj < TC; ++kj) {
          const int k = kb * TC + kj;
          if (k >= FFR) break;
          double* cb = sm + S::c2buf + (k & 1) * FE;
          if (act && tj == kj) {
#pragma unroll
            for (int a = 0; a < RT; ++a) cb[ti + TR * a] = t[a][kb];
          }
          __syncthreads();
          double lj[CT];
          const double* lcol = LT + ltoff(k) - k;
#pragma unroll
          for (int b = 0; b < CT; ++b)
            if (b >= kb) { const int j = tj + TC * b; lj[b] = (j > k) ? -lcol[j] : 0.0; }
#pragma unroll
          for (int a = 0; a < RT; ++a) {
            const double c = cb[ti + TR * a];
#pragma unroll
            for (int b = 0; b < CT; ++b)
              if (b >= kb) t[a][b] = fma(c, lj[b], t[a][b]);
          }
        }
      }
      if (pass == 0) {
        if (act) {
#pragma unroll
          for (int a = 0; a < RT; ++a)
#pragma unroll
            for (int b = 0; b < CT; ++b) __stcg(&Xg[(size_t)(ti + TR * a) * FF + tj + TC * b], t[a][b]);
        }
        ck_cluster_sync();
#pragma unroll
        for (int a = 0; a < RT; ++a)
#pragma unroll
          for (int b = 0; b < CT; ++b) t[a][b] = __ldcg(&Xg[(size_t)(tj + TC * b) * FF + ti + TR * a]);
      }
    }
    __syncthreads();
    {
      double di[RT], dj[CT];
#pragma unroll
      for (int a = 0; a < RT; ++a) di[a] = disA[ti + TR * a];
#pragma unroll
      for (int b = 0; b < CT; ++b) dj[b] = disA[tj + TC * b];
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b) t[a][b] *= di[a] * dj[b];
    }
    CK_PHASE(3);
    // =========================== phase 4: Householder tridiagonalisation ===========================
    double* tg = P.tri + (size_t)sidx * 2 * FFR;
    // Two pipelined rounds per step: (A) column k of the trailing matrix -> colbuf[k & 1] of every CTA, (B) p = beta C22 v and
    // the partial dot products v . p -> pvec[k & 1], kkp[k & 1].  After round B the owners of column k + 1 apply the rank-2
    // update to that column first and send round A of step k + 1 before the bulk of the update.
    // publish_x(k): owners of column k (tile column KBC) send rows of the blocks a >= ka(k); the diagonal goes to tri.
#define CK_HH_PUBLISH(KBC, KK)                                                                                \
  do {                                                                                                        \
    const int kac_ = (KBC) / C::KA_PER_KB;                                                                    \
    const int xbo_ = S::colbuf + ((KK) & 1) * FE;                                                             \
    _Pragma("unroll") for (int a = 0; a < RT; ++a) if (a >= kac_) pipe.pub(xbo_ + ti + TR * a, t[a][KBC]);    \
    if (ti == (KK) - kac_ * TR) tg[KK] = t[kac_][KBC];                                                        \
  } while (0)
    if (FFR >= 3 && act && tj == 0) CK_HH_PUBLISH(0, 0);
#pragma unroll
    for (int kb = 0; kb < CT; ++kb) {
      const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
      for (int kj = 0; kj < TC; ++kj) {
        const int k = kb * TC + kj;
        if (k > FFR - 3) break;
        double* c2 = sm + S::c2buf + (k & 1) * FE;      // column k+1 of the trailing matrix (rows > k of this CTA)
        if (act) {                               // column k+1 sits in this column block, or is the first of the next one
          const int kbn = (kb + 1 < CT) ? kb + 1 : kb;   // constant after unrolling
          if (kj + 1 < TC) {
            if (tj == kj + 1) {
#pragma unroll
              for (int a = 0; a < RT; ++a)
                if (a >= ka) { const int i = ti + TR * a; if (i > k) c2[i] = t[a][kb]; }
            }
          } else if (tj == 0) {
#pragma unroll
            for (int a = 0; a < RT; ++a)
              if (a >= ka) { const int i = ti + TR * a; if (i > k) c2[i] = t[a][kbn]; }
          }
        }
        pipe.wait(8u * (uint32_t)((RT - ka) * TR), tid);
        const double* xb = sm + S::colbuf + (k & 1) * FE;
        // (A) s = |x|^2 -> alpha, v0, beta, redundantly in every warp;  (B) partial row sums of C22 x with the UN-normalised
        // column: v = x - alpha e_{k+1} gives C22 v = C22 x - alpha C22[:, k+1], corrected by the row-sum threads
        double s = 0.0;
        {
          constexpr int NS = (FF + 31) / 32;
#pragma unroll
          for (int u = 0; u < NS; ++u) {
            const int i = k + 1 + lane + 32 * u;
            const double x = (i < FFR) ? xb[i] : 0.0;
            s = fma(x, x, s);
          }
        }
        double vi[RT], vj[CT];
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) { const int i = ti + TR * a; vi[a] = (i > k && i < FFR) ? xb[i] : 0.0; }
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= kb) { const int j = tj + TC * b; vj[b] = (j > k && j < FFR) ? xb[j] : 0.0; }
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) {
            double acc = 0.0;
#pragma unroll
            for (int b = 0; b < CT; ++b)
              if (b >= kb) acc = fma(t[a][b], vj[b], acc);
            if (act) red[(a * TRL + til) * RW + tj] = acc;
          }
        s = warp_sum(s);
        const double x0 = xb[k + 1];
        const double tail2 = s - x0 * x0;
        const bool skip = !(tail2 > 0.0);     // nothing to annihilate (uniform over the cluster)
        double alpha = 0.0, v0 = 0.0, beta = 0.0;
        if (!skip) {
          alpha = -copysign(sqrt(s), x0);
          v0 = x0 - alpha;
          beta = 1.0 / (s - alpha * x0);
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= ka) { if (ti + TR * a == k + 1) vi[a] = v0; }
#pragma unroll
          for (int b = 0; b < CT; ++b)
            if (b >= kb) { if (tj + TC * b == k + 1) vj[b] = v0; }
        }
        __syncthreads();       // partial sums visible; round A of this step consumed by every thread of this CTA
        double wi[RT], wj[CT];
        if (skip) {
          if (tid == 0) { betaS[k] = 0.0; if (crank == 0) tg[FFR + k] = x0; }
#pragma unroll
          for (int a = 0; a < RT; ++a) { wi[a] = 0.0; vi[a] = 0.0; }
#pragma unroll
          for (int b = 0; b < CT; ++b) { wj[b] = 0.0; vj[b] = 0.0; }
        } else {
          // rows of this CTA: p = beta C22 v (four lanes per row), all-gathered; partial v . p per warp -> every CTA
          const int pvo = S::pvec + (k & 1) * FE, kko = S::kkp + (k & 1) * 32;
          if (tid < C::RS_THREADS) {
            const int lr = tid >> 2, sub = tid & 3;
            double acc = 0.0;
            if (lr < LROWS) {
#pragma unroll
              for (int c = 0; c < (TC + 3) / 4; ++c) { const int cc = sub + 4 * c; if (cc < TC) acc += red[lr * RW + cc]; }
            }
            acc += __shfl_xor_sync(0xffffffffu, acc, 1);
            acc += __shfl_xor_sync(0xffffffffu, acc, 2);
            double contrib = 0.0;
            const int r = crank * TRL + (lr % TRL) + TR * (lr / TRL);
            if (lr < LROWS && sub == 0 && r > k && r < FFR) {
              acc = fma(-alpha, c2[r], acc);                    // C22 v = C22 x - alpha C22[:, k+1]
              const double vr = (r == k + 1) ? v0 : xb[r];
              const double pr_ = beta * acc;
              pipe.pub(pvo + r, pr_);
              contrib = pr_ * vr;
              __stcg(&Vg[(size_t)k * FF + r], vr);              // reflector -> L2-resident scratch (read back in phase 5)
            }
            contrib = warp_sum(contrib);
            if (lane == 0) pipe.pub(kko + crank * C::RS_WARPS + warp, contrib);
          }
          if (tid == 0) { betaS[k] = beta; if (crank == 0) tg[FFR + k] = alpha; }
          pipe.wait(8u * (uint32_t)((FFR - 1 - k) + NCTA * C::RS_WARPS), tid);
          const double* pvec = sm + pvo;
          double kk = (lane < NCTA * C::RS_WARPS) ? sm[kko + lane] : 0.0;
          kk = warp_sum(kk);
          kk *= 0.5 * beta;
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= ka) { const int i = ti + TR * a; wi[a] = (i > k && i < FFR) ? fma(-kk, vi[a], pvec[i]) : 0.0; }
#pragma unroll
          for (int b = 0; b < CT; ++b)
            if (b >= kb) { const int j = tj + TC * b; wj[b] = (j > k && j < FFR) ? fma(-kk, vj[b], pvec[j]) : 0.0; }
        }
        // look-ahead: column k + 1 gets its rank-2 update first and is sent as round A of step k + 1
        if (k + 1 <= FFR - 3 && act) {
          if (kj + 1 < TC) {
            if (tj == kj + 1) {
#pragma unroll
              for (int a = 0; a < RT; ++a) if (a >= ka) t[a][kb] = fma(-vi[a], wj[kb], fma(-wi[a], vj[kb], t[a][kb]));
              wj[kb] = 0.0; vj[kb] = 0.0;
              CK_HH_PUBLISH(kb, k + 1);
            }
          } else if (kb + 1 < CT) {
            if (tj == 0) {
#pragma unroll
              for (int a = 0; a < RT; ++a)
                if (a >= ka) t[a][(kb + 1 < CT) ? kb + 1 : kb] = fma(-vi[a], wj[(kb + 1 < CT) ? kb + 1 : kb], fma(-wi[a], vj[(kb + 1 < CT) ? kb + 1 : kb], t[a][(kb + 1 < CT) ? kb + 1 : kb]));
              wj[(kb + 1 < CT) ? kb + 1 : kb] = 0.0; vj[(kb + 1 < CT) ? kb + 1 : kb] = 0.0;
              CK_HH_PUBLISH(((kb + 1 < CT) ? kb + 1 : kb), k + 1);
            }
          }
        }
        if (!skip) {
#pragma unroll
          for (int b = 0; b < CT; ++b)
            if (b >= kb) {
#pragma unroll
              for (int a = 0; a < RT; ++a)
                if (a >= ka) t[a][b] = fma(-vi[a], wj[b], fma(-wi[a], vj[b], t[a][b]));
            }
        }
      }
    }
#undef CK_HH_PUBLISH
    ck_cluster_sync();     // phase boundary
    // trailing 2x2 block of the real harmonics
    if (act) {
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b) {
          const int i = ti + TR * a, j = tj + TC * b;
          if (i == FFR - 2 && j == FFR - 2) tg[FFR - 2] = t[a][b];
          if (i == FFR - 1 && j == FFR - 1) { tg[FFR - 1] = t[a][b]; tg[2 * FFR - 1] = 0.0; }
          if (i == FFR - 1 && j == FFR - 2) tg[2 * FFR - 2] = t[a][b];
        }
    }
    CK_PHASE(4);
    // =========================== phase 5: T = Q^T = H_{n-3} ... H_0 accumulated backwards (no cluster traffic) ===========================
#pragma unroll
    for (int a = 0; a < RT; ++a)
#pragma unroll
      for (int b = 0; b < CT; ++b) t[a][b] = (ti + TR * a == tj + TC * b) ? 1.0 : 0.0;
    ck_cluster_sync();       // reflectors of every CTA are visible in L2
    double* pvec = sm + S::pvec;
    {
      double vn[CT];         // reflector of the NEXT step, prefetched from L2 under the current one
      {
        const int k0 = FFR - 3;
#pragma unroll
        for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; vn[b] = (k0 >= 0 && j > k0 && j < FFR) ? __ldcg(&Vg[(size_t)k0 * FF + j]) : 0.0; }
      }
#pragma unroll
      for (int kb = CT - 1; kb >= 0; --kb) {
        const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
        for (int kj = TC - 1; kj >= 0; --kj) {
          const int k = kb * TC + kj;
          if (k > FFR - 3) continue;
          double vj[CT];
#pragma unroll
          for (int b = 0; b < CT; ++b) vj[b] = vn[b];
          if (k > 0) {
#pragma unroll
            for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; vn[b] = (j > k - 1 && j < FFR) ? __ldcg(&Vg[(size_t)(k - 1) * FF + j]) : 0.0; }
          }
          const double beta = betaS[k];
          if (beta == 0.0) continue;            // uniform
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= ka) {
              double acc = 0.0;
#pragma unroll
              for (int b = 0; b < CT; ++b)
                if (b >= kb) acc = fma(t[a][b], vj[b], acc);
              if (act) red[(a * TRL + til) * RW + tj] = acc;
            }
          __syncthreads();
          if (tid < C::RS_THREADS) {
            const int lr = tid >> 2, sub = tid & 3;
            double acc = 0.0;
            if (lr < LROWS) {
#pragma unroll
              for (int c = 0; c < (TC + 3) / 4; ++c) { const int cc = sub + 4 * c; if (cc < TC) acc += red[lr * RW + cc]; }
            }
            acc += __shfl_xor_sync(0xffffffffu, acc, 1);
            acc += __shfl_xor_sync(0xffffffffu, acc, 2);
            if (lr < LROWS && sub == 0) pvec[crank * TRL + (lr % TRL) + TR * (lr / TRL)] = -beta * acc;
          }
          __syncthreads();
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= ka) {
              const int i = ti + TR * a;
              const double y = (i > k) ? pvec[i] : 0.0;
#pragma unroll
              for (int b = 0; b < CT; ++b)
                if (b >= kb) t[a][b] = fma(y, vj[b], t[a][b]);
            }
        }
      }
    }
    __syncthreads();
    CK_PHASE(5);
    // =========================== phase 6: M0'^T = Q^T D^-1/2 L^-1 (no cluster traffic) ===========================
    {
      double dj[CT];
#pragma unroll
      for (int b = 0; b < CT; ++b) dj[b] = disA[tj + TC * b];
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b) t[a][b] *= dj[b];
    }
#pragma unroll
    for (int kb = CT - 1; kb >= 0; --kb) {
#pragma unroll 1
      for (int kj = TC - 1; kj >= 0; --kj) {
        const int k = kb * TC + kj;
        if (k == 0 || k >= FFR) continue;
        double* cb = sm + S::c2buf + (k & 1) * FE;
        if (act && tj == kj) {
#pragma unroll
          for (int a = 0; a < RT; ++a) cb[ti + TR * a] = t[a][kb];
        }
        __syncthreads();
        double lj[CT];
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b <= kb) { const int j = tj + TC * b; lj[b] = (j < k) ? -LT[ltoff(j) + (k - j)] : 0.0; }
#pragma unroll
        for (int a = 0; a < RT; ++a) {
          const double c = cb[ti + TR * a];
#pragma unroll
          for (int b = 0; b < CT; ++b)
            if (b <= kb) t[a][b] = fma(c, lj[b], t[a][b]);
        }
      }
    }
    CK_PHASE(6);
    // ---- store: t(i, j) = M0'[j][i] (swapped ordering);  layout Wbuf[row][mode] in ORIGINAL harmonic order ----
    if (act) {
      double* Wg = P.Wbuf + (size_t)sidx * FFR * FFR;
#pragma unroll
      for (int b = 0; b < CT; ++b) {
        const int j = tj + TC * b;
        const int row = (j < z) ? (j ^ 1) : j;
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (j < FFR && ti + TR * a < FFR) Wg[(size_t)row * FFR + ti + TR * a] = t[a][b];
      }
    }
    // ---- status, next structure ----
    if (__syncthreads_or(bad)) { if (tid == 0) ck_st_cluster_i32(pipe.rb[0] + 8u * S::ints + 4u * (1 + crank), 1); }
    else if (tid == 0) ck_st_cluster_i32(pipe.rb[0] + 8u * S::ints + 4u * (1 + crank), 0);
    ck_cluster_sync();
    if (crank == 0 && tid == 0) {
      int anybad = 0;
#pragma unroll
      for (int r = 0; r < NCTA; ++r) anybad |= ints[1 + r];
      P.wstatus[sidx] = anybad ? PIRL_ST_NPD : 0;
      const int nxt = P.queue ? nclusters + atomicAdd(P.queue, 1) : sidx + nclusters;
#pragma unroll
      for (int r = 0; r < NCTA; ++r) ck_st_cluster_i32(pipe.rb[r] + 8u * S::ints, nxt);
    }
    ck_cluster_sync();
    sidx = ints[0];
    CK_PHASE(7);
  }
}
