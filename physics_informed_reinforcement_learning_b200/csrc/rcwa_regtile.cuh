// rcwa_regtile.cuh -- register-tile kernels of the RCWA path (sm_100a).
//
// The ff x ff matrices of ONE structure live in the REGISTERS of one small CTA: thread (ti, tj) of a TR x TC grid owns
// the cyclic tile  { (ti + TR*a, tj + TC*b) : a < RT, b < CT }.  Every factorisation step is a rank-1 / rank-2 update
// whose pivot row / column is published through shared memory once (RT + CT shared loads per RT*CT DFMAs), so the
// kernels are bound by the FP64 pipe instead of shared-memory bandwidth (profiles/r01_baseline_*: 5-6 % FP64 activity
// for the shared-memory versions).  The step loops are unrolled over the tile-column block index so that all register
// indices are compile-time constants and finished slabs of the shrinking windows are skipped at compile time.
//
// rk_front : pattern -> Fourier coefficients -> A' = LDL^T (L kept in shared memory) -> E^-1 (Gauss-Jordan) -> B'
//            -> C = D^-1/2 L^-1 B' L^-T D^-1/2 -> Householder tridiagonalisation (reflectors kept in the upper triangle
//            of the same shared matrix) -> Q^T accumulated backwards -> M0^T = Q^T D^-1/2 L^-1 -> global.
//            (replaces convolution_matrix.py:89-114 x2, _base.py:151-155 up to the reduction of the eigenproblem)
//   The problem is solved in the cos<->sin swapped ordering (A' = P A P, B' = P B P): the swap that Kx E^-1 Kx needs
//   becomes an index relabelling, so no data ever moves between threads for it.
#pragma once
#include "rcwa_kernels.cuh"

// FFR_ (0 = FF_): number of REAL harmonics when the tile grid is padded (rk_front only): rows / columns >= FFR are identity
// padding that stays decoupled through every factorisation step and is dropped at the stores (e.g. 121 = 11^2 -> 126 = 42x3 = 14x9).
template <int FF_, int TR_, int TC_, int RT_, int CT_, int MINB_, int REGCAP_ = 255, int FFR_ = 0>
struct TileCfg {
  static constexpr int FF = FF_, TR = TR_, TC = TC_, RT = RT_, CT = CT_, MINB = MINB_;
  static constexpr int FFR = FFR_ > 0 ? FFR_ : FF_;
  static constexpr int N = (FFR - 1) / 2;
  static constexpr int NTHR = TR * TC;
  static constexpr int NT = ((NTHR + 31) / 32) * 32;
  // register cap for MINB CTAs per SM
  static constexpr int MAXREG = REGCAP_ < ((65536 / (MINB * NT)) / 8) * 8 ? REGCAP_ : ((65536 / (MINB * NT)) / 8) * 8;
  static constexpr int LD = (FF % 2) ? FF : FF + 1;   // odd leading dimension: conflict-free column walks
  static constexpr int KA_PER_KB = TR / TC;           // tile-row block index = kb / KA_PER_KB
  static_assert(TR * RT == FF && TC * CT == FF, "exact tiling only");
  static_assert(TR % TC == 0, "TR must be a multiple of TC");
  static_assert(NTHR >= FF, "row reductions use one thread per row");
};

// shared-memory carve-up of rk_front (in doubles)
template <class C>
struct FrontSmem {
  static constexpr int cre = 0;
  static constexpr int cim = cre + (2 * C::N + 1);
  static constexpr int Lsm = pirl_even(cim + (2 * C::N + 1));
  static constexpr int rowbuf = pirl_even(Lsm + C::FF * C::LD);       // [2][FF]
  static constexpr int colbuf = rowbuf + 2 * pirl_even(C::FF);         // [2][FF]
  static constexpr int scal = colbuf + 2 * pirl_even(C::FF);           // [2][4]
  static constexpr int pinvA = scal + 8;
  static constexpr int disA = pinvA + pirl_even(C::FF);
  static constexpr int red = disA + pirl_even(C::FF);                  // [FF][TC]
  static constexpr int pvec = red + pirl_even(C::FF * C::TC);
  static constexpr int dd = pvec + pirl_even(C::FF);
  static constexpr int ee = dd + pirl_even(C::FF);
  static constexpr int tw = ee + pirl_even(C::FF);                     // [n_cells] complex -> 2*n_cells doubles (runtime)
  static __host__ __device__ constexpr int doubles(int n_cells) { return tw + 2 * n_cells + pirl_even((n_cells + 7) / 8); }
};

#ifdef PIRL_HH_PHASES
#define HH_PHASE(n) RK_PHASE(n)
#else
#define HH_PHASE(n) do { } while (0)
#endif
#define RK_PHASE(n)                                                                       \
  do {                                                                                    \
    if (P.phase_clk && tid == 0) {                                                        \
      const long long now_ = clock64();                                                   \
      atomicAdd(&P.phase_clk[n], (unsigned long long)(now_ - pc_));                       \
      pc_ = now_;                                                                         \
    }                                                                                     \
  } while (0)

template <class C>
__global__ void __maxnreg__(C::MAXREG) rk_front(KParams P) {
  constexpr int FF = C::FF, TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, N = C::N, LD = C::LD, NT = C::NT;
  constexpr int FFR = C::FFR;            // real harmonics (== FF unless the tile grid is padded)
  constexpr bool PAD = FFR != FF;
  constexpr int z = 2 * N;
  using S = FrontSmem<C>;
  extern __shared__ __align__(16) double sm[];
  double* cre = sm + S::cre;
  double* cim = sm + S::cim;
  double* Lsm = sm + S::Lsm;
  double* rowbuf = sm + S::rowbuf;
  double* colbuf = sm + S::colbuf;
  double* scal = sm + S::scal;
  double* pinvA = sm + S::pinvA;
  double* disA = sm + S::disA;
  double* red = sm + S::red;
  double* pvec = sm + S::pvec;
  double* dd = sm + S::dd;
  double* ee = sm + S::ee;
  cplx* twS = reinterpret_cast<cplx*>(sm + S::tw);
  const int nc = P.n_cells;
  uint8_t* su = reinterpret_cast<uint8_t*>(sm + S::tw + 2 * nc);
  constexpr int FE = pirl_even(FF);

  const int tid = threadIdx.x, lane = tid & 31;
  const bool act = tid < C::NTHR;
  const int ti = act ? tid / TC : TR - 1;
  const int tj = act ? tid % TC : TC - 1;
  __shared__ int sbad;
  __shared__ int snext;
  const CtaTrace trc(P.trace);

  for (int t = tid; t < nc; t += NT) twS[t] = P.tw[t];

  for (int sidx = blockIdx.x; sidx < P.B; sidx = pirl_next_item(P.queue, sidx, &snext)) {
    __syncthreads();
    if (tid == 0) sbad = 0;
    const double* pr = P.params + (size_t)sidx * P.pstride;
    const double sa = sin(pr[1] * (3.14159265358979323846 / 180.0));
    const double nr = pr[3];
    const double er = nr * nr, eg = P.eps_g;
    int bad = 0;
    long long pc_ = clock64();
    for (int p = tid; p < nc; p += NT) su[p] = P.patterns[(size_t)sidx * nc + p] > 0;
    __syncthreads();
    // ---- Fourier coefficients c_k, k = 0..2N (closed form of the up-sampled FFT): one thread per k ----
    for (int k = tid; k <= 2 * N; k += NT) {
      double sx = 0.0, sy = 0.0;
      int idx = 0;
      const int km = k % nc;                     // harmonics >= n_cells alias (2 * order >= n_cells)
      for (int p = 0; p < nc; ++p) {
        if (su[p]) { const cplx w = twS[idx]; sx += w.x; sy += w.y; }
        idx += km; if (idx >= nc) idx -= nc;
      }
      const cplx d = __ldg(&P.Dk[k]);
      cre[k] = sx * d.x - sy * d.y;
      cim[k] = sx * d.y + sy * d.x;
    }
    __syncthreads();

    RK_PHASE(0);
    double t[RT][CT];
    // =========================== phase 1: A' = P A P, LDL^T ===========================
    {
      const double da = 1.0 / er - 1.0 / eg, ia = 1.0 / eg;
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b) {
          const int i = ti + TR * a, j = tj + TC * b;
          const int pi = (i < z) ? (i ^ 1) : i, pj = (j < z) ? (j ^ 1) : j;
          if (PAD && (i >= FFR || j >= FFR)) t[a][b] = (i == j) ? 1.0 : 0.0;
          else t[a][b] = fma(da, basis_toeplitz(pi, pj, N, cre, cim), (i == j) ? ia : 0.0);
        }
    }
#pragma unroll
    for (int kb = 0; kb < CT; ++kb) {
      const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
      for (int kj = 0; kj < TC; ++kj) {
        const int k = kb * TC + kj;
        const int kti = k - ka * TR;
        if (act && tj == kj) {
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= ka) { const int i = ti + TR * a; if (i >= k) Lsm[i * LD + k] = t[a][kb]; }
          if (ti == kti) {
            const double piv = t[ka][kb];
            if (!(piv > 0.0)) bad = 1;
            pinvA[k] = 1.0 / piv;
            disA[k] = rsqrt(piv);
          }
        }
        __syncthreads();
        const double p = pinvA[k];
        double li[RT];
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) { const int i = ti + TR * a; li[a] = (i > k) ? -Lsm[i * LD + k] * p : 0.0; }
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= kb) {
            const int j = tj + TC * b;
            const double cj = (j > k) ? Lsm[j * LD + k] : 0.0;
#pragma unroll
            for (int a = 0; a < RT; ++a)
              if (a >= ka) t[a][b] = fma(li[a], cj, t[a][b]);
          }
      }
    }
    __syncthreads();
    // raw columns (l_ik d_k) -> unit-lower L
    for (int idx = tid; idx < FF * FF; idx += NT) {
      const int i = idx / FF, k = idx - i * FF;
      if (i > k) Lsm[i * LD + k] *= pinvA[k];
    }
    RK_PHASE(1);
    // =========================== phase 2: E, Gauss-Jordan inverse, B' ===========================
    {
      const double de = er - eg;
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b) {
          const int i = ti + TR * a, j = tj + TC * b;
          if (PAD && (i >= FFR || j >= FFR)) t[a][b] = (i == j) ? 1.0 : 0.0;
          else t[a][b] = fma(de, basis_toeplitz(i, j, N, cre, cim), (i == j) ? eg : 0.0);
        }
    }
#pragma unroll
    for (int kb = 0; kb < CT; ++kb) {
      const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
      for (int kj = 0; kj < TC; ++kj) {
        const int k = kb * TC + kj;
        const int kti = k - ka * TR;
        double* rb = rowbuf + (k & 1) * FE;
        double* cb = colbuf + (k & 1) * FE;
        double* sc = scal + (k & 1) * 4;
        if (act) {
          if (ti == kti) {
#pragma unroll
            for (int b = 0; b < CT; ++b) rb[tj + TC * b] = t[ka][b];
          }
          if (tj == kj) {
#pragma unroll
            for (int a = 0; a < RT; ++a) cb[ti + TR * a] = t[a][kb];
            if (ti == kti) {
              const double piv = t[ka][kb];
              if (!(piv > 0.0)) bad = 1;
              sc[0] = 1.0 / piv;
              sc[1] = piv;
            }
          }
        }
        __syncthreads();
        const double p = sc[0], piv = sc[1];
        double colv[RT];
#pragma unroll
        for (int a = 0; a < RT; ++a) { const int i = ti + TR * a; colv[a] = (i == k) ? (p - 1.0) : -cb[i] * p; }
#pragma unroll
        for (int b = 0; b < CT; ++b) {
          const int j = tj + TC * b;
          const double rv = (j == k) ? (1.0 + piv) : rb[j];
#pragma unroll
          for (int a = 0; a < RT; ++a) t[a][b] = fma(colv[a], rv, t[a][b]);
        }
      }
    }
    __syncthreads();
    if (P.Einv && act) {   // adjoint sweep needs E^-1 (original harmonic order)
      double* Eg = P.Einv + (size_t)sidx * FFR * FFR;
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (!PAD || (ti + TR * a < FFR && tj + TC * b < FFR)) Eg[(ti + TR * a) * FFR + tj + TC * b] = t[a][b];
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
          t[a][b] = si[a] * sj[b] * t[a][b] - ((i == j) ? 1.0 : 0.0);
        }
    }
    RK_PHASE(2);
    // =========================== phase 3: C = D^-1/2 L^-1 B' L^-T D^-1/2 ===========================
#pragma unroll
    for (int kb = 0; kb < CT; ++kb) {      // rows: X <- L^-1 X
      const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
      for (int kj = 0; kj < TC; ++kj) {
        const int k = kb * TC + kj;
        const int kti = k - ka * TR;
        double* rb = rowbuf + (k & 1) * FE;
        if (act && ti == kti) {
#pragma unroll
          for (int b = 0; b < CT; ++b) rb[tj + TC * b] = t[ka][b];
        }
        __syncthreads();
        double li[RT];
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) { const int i = ti + TR * a; li[a] = (i > k) ? -Lsm[i * LD + k] : 0.0; }
#pragma unroll
        for (int b = 0; b < CT; ++b) {
          const double r = rb[tj + TC * b];
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= ka) t[a][b] = fma(li[a], r, t[a][b]);
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int kb = 0; kb < CT; ++kb) {      // columns: X <- X L^-T
      const int ka = kb / C::KA_PER_KB;
      (void)ka;
#pragma unroll 1
      for (int kj = 0; kj < TC; ++kj) {
        const int k = kb * TC + kj;
        double* cb = colbuf + (k & 1) * FE;
        if (act && tj == kj) {
#pragma unroll
          for (int a = 0; a < RT; ++a) cb[ti + TR * a] = t[a][kb];
        }
        __syncthreads();
        double lj[CT];
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= kb) { const int j = tj + TC * b; lj[b] = (j > k) ? -Lsm[j * LD + k] : 0.0; }
#pragma unroll
        for (int a = 0; a < RT; ++a) {
          const double c = cb[ti + TR * a];
#pragma unroll
          for (int b = 0; b < CT; ++b)
            if (b >= kb) t[a][b] = fma(c, lj[b], t[a][b]);
        }
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
    RK_PHASE(3);
    // =========================== phase 4: Householder tridiagonalisation ===========================
#pragma unroll
    for (int kb = 0; kb < CT; ++kb) {
      const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
      for (int kj = 0; kj < TC; ++kj) {
        const int k = kb * TC + kj;
        if (k > FF - 3) break;
        const int kti = k - ka * TR;
        double* xb = colbuf + (k & 1) * FE;
        double* c2 = rowbuf + (k & 1) * FE;      // column k+1 of the trailing matrix (rows > k)
        if (act && tj == kj) {
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= ka) { const int i = ti + TR * a; if (i > k) xb[i] = t[a][kb]; }
          if (ti == kti) dd[k] = t[ka][kb];
        }
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
        HH_PHASE(16);
        __syncthreads();
        HH_PHASE(17);
        // Two independent chains after the barrier, written as straight-line code so that they interleave:
        //  (A) s = |x|^2 -> alpha, v0, beta (loads, five shuffle levels, sqrt, division)
        //  (B) partial row sums of C22 x with the UN-normalised column x: v = x - alpha e_{k+1} gives
        //      C22 v = C22 x - alpha C22[:, k+1], and the correction is applied by the row-sum threads after the next barrier.
        double s = 0.0;
        {
          constexpr int NS = (FF + 31) / 32;
#pragma unroll
          for (int u = 0; u < NS; ++u) {
            const int i = k + 1 + lane + 32 * u;
            const double x = (i < FF) ? xb[i] : 0.0;
            s = fma(x, x, s);
          }
        }
        double vi[RT], vj[CT];
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) { const int i = ti + TR * a; vi[a] = (i > k) ? xb[i] : 0.0; }
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= kb) { const int j = tj + TC * b; vj[b] = (j > k) ? xb[j] : 0.0; }
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) {
            double acc = 0.0;
#pragma unroll
            for (int b = 0; b < CT; ++b)
              if (b >= kb) acc = fma(t[a][b], vj[b], acc);
            if (act) red[(ti + TR * a) * TC + tj] = acc;
          }
        s = warp_sum(s);
        const double x0 = xb[k + 1];
        const double tail2 = s - x0 * x0;
        if (!(tail2 > 0.0)) {                 // nothing to annihilate (uniform over the CTA)
          if (tid == 0) { ee[k] = x0; Lsm[k * LD + k] = 0.0; }
          continue;
        }
        const double alpha = -copysign(sqrt(s), x0);
        const double v0 = x0 - alpha;
        const double beta = 1.0 / (s - alpha * x0);
        HH_PHASE(18);
        // v differs from x only in entry k+1
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) { if (ti + TR * a == k + 1) vi[a] = v0; }
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= kb) { if (tj + TC * b == k + 1) vj[b] = v0; }
        HH_PHASE(19);
        __syncthreads();
        HH_PHASE(20);
        // rows: p = beta C22 v; the same (FF + 31) / 32 warps also reduce v . p (for kk), so that no second all-warp reduction
        // sits between the barrier and the rank-2 update
        constexpr int NWR = (FF + 31) / 32;
        if (tid < 32 * NWR) {
          const int r = tid;
          double contrib = 0.0;
          if (r < FF && r > k) {
            double acc = 0.0;
#pragma unroll
            for (int c = 0; c < TC; ++c) acc += red[r * TC + c];
            acc = fma(-alpha, c2[r], acc);                    // C22 v = C22 x - alpha C22[:, k+1]
            const double vr = (r == k + 1) ? v0 : xb[r];
            const double pr_ = beta * acc;
            pvec[r] = pr_;
            contrib = pr_ * vr;
            Lsm[k * LD + r] = vr;                             // reflector -> upper triangle of the shared matrix
          } else if (r == k) {
            Lsm[k * LD + k] = beta;
            ee[k] = alpha;
          }
          contrib = warp_sum(contrib);
          if (lane == 0) scal[tid >> 5] = contrib;
        }
        HH_PHASE(21);
        __syncthreads();
        HH_PHASE(22);
        double kk = 0.0;
#pragma unroll
        for (int w = 0; w < NWR; ++w) kk += scal[w];
        kk *= 0.5 * beta;
        HH_PHASE(23);
        double wi[RT];
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) { const int i = ti + TR * a; wi[a] = (i > k) ? fma(-kk, vi[a], pvec[i]) : 0.0; }
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= kb) {
            const int j = tj + TC * b;
            const double wjb = (j > k) ? fma(-kk, vj[b], pvec[j]) : 0.0;
#pragma unroll
            for (int a = 0; a < RT; ++a)
              if (a >= ka) t[a][b] = fma(-vi[a], wjb, fma(-wi[a], vj[b], t[a][b]));
          }
        HH_PHASE(24);
      }
    }
    // trailing 2x2 block
    if (act) {
      constexpr int a2 = (FF - 2) / TR, a1 = (FF - 1) / TR, b2 = (FF - 2) / TC, b1 = (FF - 1) / TC;
      if (ti == (FF - 2) % TR && tj == (FF - 2) % TC) dd[FF - 2] = t[a2][b2];
      if (ti == (FF - 1) % TR && tj == (FF - 1) % TC) { dd[FF - 1] = t[a1][b1]; ee[FF - 1] = 0.0; }
      if (ti == (FF - 1) % TR && tj == (FF - 2) % TC) ee[FF - 2] = t[a1][b2];
    }
    RK_PHASE(4);
    // =========================== phase 5: T = Q^T = H_{n-3} ... H_0 accumulated backwards ===========================
#pragma unroll
    for (int a = 0; a < RT; ++a)
#pragma unroll
      for (int b = 0; b < CT; ++b) t[a][b] = (ti + TR * a == tj + TC * b) ? 1.0 : 0.0;
    __syncthreads();
#pragma unroll
    for (int kb = CT - 1; kb >= 0; --kb) {
      const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
      for (int kj = TC - 1; kj >= 0; --kj) {
        const int k = kb * TC + kj;
        if (k > FF - 3) continue;
        const double beta = Lsm[k * LD + k];
        if (beta == 0.0) continue;            // uniform
        double vj[CT];
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= kb) { const int j = tj + TC * b; vj[b] = (j > k) ? Lsm[k * LD + j] : 0.0; }
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) {
            double acc = 0.0;
#pragma unroll
            for (int b = 0; b < CT; ++b)
              if (b >= kb) acc = fma(t[a][b], vj[b], acc);
            if (act) red[(ti + TR * a) * TC + tj] = acc;
          }
        __syncthreads();
        if (tid < FF && tid > k) {
          double acc = 0.0;
#pragma unroll
          for (int c = 0; c < TC; ++c) acc += red[tid * TC + c];
          pvec[tid] = -beta * acc;
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
    __syncthreads();
    RK_PHASE(5);
    // =========================== phase 6: M0'^T = Q^T D^-1/2 L^-1 ===========================
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
        if (k == 0) continue;
        double* cb = colbuf + (k & 1) * FE;
        if (act && tj == kj) {
#pragma unroll
          for (int a = 0; a < RT; ++a) cb[ti + TR * a] = t[a][kb];
        }
        __syncthreads();
        double lj[CT];
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b <= kb) { const int j = tj + TC * b; lj[b] = (j < k) ? -Lsm[k * LD + j] : 0.0; }
#pragma unroll
        for (int a = 0; a < RT; ++a) {
          const double c = cb[ti + TR * a];
#pragma unroll
          for (int b = 0; b < CT; ++b)
            if (b <= kb) t[a][b] = fma(c, lj[b], t[a][b]);
        }
      }
    }
    RK_PHASE(6);
    // ---- store: t(i, j) = M0'[j][i] (swapped ordering);  old layout Wbuf[row][mode] in ORIGINAL harmonic order ----
    if (act) {
      double* Wg = P.Wbuf + (size_t)sidx * FFR * FFR;
#pragma unroll
      for (int b = 0; b < CT; ++b) {
        const int j = tj + TC * b;
        const int row = (j < z) ? (j ^ 1) : j;
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (!PAD || (j < FFR && ti + TR * a < FFR)) Wg[row * FFR + ti + TR * a] = t[a][b];
      }
    }
    if (bad) sbad = 1;
    __syncthreads();
    double* tg = P.tri + (size_t)sidx * 2 * FFR;
    for (int q = tid; q < FFR; q += NT) { tg[q] = dd[q]; tg[FFR + q] = (q == FFR - 1) ? 0.0 : ee[q]; }
    if (tid == 0) P.wstatus[sidx] = sbad ? PIRL_ST_NPD : 0;
    RK_PHASE(7);
  }
  trc.done(P.trace, 5, P.B);
}

// ===================================================================================================================
// rk_tail : layer solve in the mode basis with the register-tile machinery (replaces _base.py:159,
//           transfer_method.py:42-75, solver.py:29-38).  Same equations as k_tail, reorganised so that O(n^3) work is
//           only spent where it matters (tools/proto_tail.py is the NumPy statement of exactly these steps):
//   * modes are sorted by DESCENDING eigenvalue: [insignificant evanescent (X = exp(-k0 q d) < 1e-12) | significant
//     evanescent | propagating].  For the first n_i modes X is zero to working precision, so top and bottom of the layer
//     decouple there: G_ii = (Q + H_II)_ii and S_ii = (Q + H_I)_ii; both are eliminated WITHOUT pivoting (numerical range
//     in the closed first quadrant), carrying the right-hand sides / output functionals along, so no inverse, no
//     back-substitution and no n^3 product is ever formed for them;
//   * the trailing m = n - n_i block is dense and small: P = (Schur complement of G)^-1 in shared memory, the coupling
//     term 2 (Hq X) P (X Q) is added to the m significant columns of S only, and the final m x m system is solved with
//     partial pivoting in shared memory.
//   If m > MMAX (very thin layers) the structure is flagged and the generic k_tail kernel finishes it.
// ===================================================================================================================
#define PIRL_ST_TAIL_FALLBACK 0x100   // working-status bit: structure left to the generic k_tail

template <class C, int MMAX_>
struct TailSmem {
  static constexpr int FF = C::FF, FE = pirl_even(C::FF), MMAX = MMAX_;
  // complex vectors (2 doubles per element)
  static constexpr int qv = 0;
  static constexpr int Xv = qv + 2 * FE;
  static constexpr int XQ = Xv + 2 * FE;
  static constexpr int omx2 = XQ + 2 * FE;
  static constexpr int DI = omx2 + 2 * FE;
  static constexpr int DII = DI + 2 * FE;
  static constexpr int rowbuf = DII + 2 * FE;       // [2][FE] complex
  static constexpr int colbuf = rowbuf + 4 * FE;    // [2][FE] complex
  static constexpr int scal = colbuf + 4 * FE;      // [2][8] complex
  static constexpr int gvec = scal + 32;            // [3][FE] complex
  static constexpr int frs = gvec + 6 * FE;         // [3][FE] complex
  static constexpr int cs = frs + 6 * FE;           // [FE] complex
  static constexpr int yv = cs + 2 * FE;            // [FE] complex
  static constexpr int ds = yv + 2 * FE;            // [FE] complex
  static constexpr int rhsS = ds + 2 * FE;          // [FE] complex  (trailing right-hand side of S)
  static constexpr int misc = rhsS + 2 * FE;        // corner[3] complex, red[8] complex, ints
  static constexpr int wr3 = misc + 64;             // [3][FE] real
  static constexpr int lamS = wr3 + 3 * FE;         // [FE] real
  static constexpr int pos = lamS + FE;             // [FE] ints (as doubles' storage)
  static constexpr int regA = pos + FE;             // Tsm: [MMAX][MMAX] complex
  static constexpr int regB = regA + 2 * MMAX * MMAX;
  static constexpr int regB_doubles =
      (C::FF * C::LD > 2 * C::FF * MMAX) ? ((C::FF * C::LD > 2 * MMAX * (MMAX + 1)) ? C::FF * C::LD : 2 * MMAX * (MMAX + 1))
                                         : ((2 * C::FF * MMAX > 2 * MMAX * (MMAX + 1)) ? 2 * C::FF * MMAX : 2 * MMAX * (MMAX + 1));
  static constexpr int doubles = regB + pirl_even(regB_doubles);
};

template <class C, int MMAX>
__global__ void __maxnreg__(C::MAXREG) rk_tail(KParams P) {
  constexpr int FF = C::FF, TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, N = C::N, LD = C::LD, NT = C::NT;
  constexpr int z = 2 * N;
  using S = TailSmem<C, MMAX>;
  constexpr int FE = S::FE;
  extern __shared__ __align__(16) double sm[];
  cplx* qv = reinterpret_cast<cplx*>(sm + S::qv);
  cplx* Xv = reinterpret_cast<cplx*>(sm + S::Xv);
  cplx* XQ = reinterpret_cast<cplx*>(sm + S::XQ);
  cplx* omx2 = reinterpret_cast<cplx*>(sm + S::omx2);
  cplx* DI = reinterpret_cast<cplx*>(sm + S::DI);
  cplx* DII = reinterpret_cast<cplx*>(sm + S::DII);
  cplx* rowbuf = reinterpret_cast<cplx*>(sm + S::rowbuf);
  cplx* colbuf = reinterpret_cast<cplx*>(sm + S::colbuf);
  cplx* scal = reinterpret_cast<cplx*>(sm + S::scal);
  cplx* gvec = reinterpret_cast<cplx*>(sm + S::gvec);
  cplx* frs = reinterpret_cast<cplx*>(sm + S::frs);
  cplx* cs = reinterpret_cast<cplx*>(sm + S::cs);
  cplx* yv = reinterpret_cast<cplx*>(sm + S::yv);
  cplx* dsv = reinterpret_cast<cplx*>(sm + S::ds);
  cplx* rhsS = reinterpret_cast<cplx*>(sm + S::rhsS);
  cplx* corner = reinterpret_cast<cplx*>(sm + S::misc);        // [4]
  cplx* red = corner + 4;                                        // [8]
  int* ints = reinterpret_cast<int*>(red + 8);                   // [8]
  double* wr3 = sm + S::wr3;
  double* lamS = sm + S::lamS;
  int* pos = reinterpret_cast<int*>(sm + S::pos);
  cplx* Tsm = reinterpret_cast<cplx*>(sm + S::regA);
  double* Wsm = sm + S::regB;
  cplx* HqS = reinterpret_cast<cplx*>(sm + S::regB);
  cplx* Ssm = reinterpret_cast<cplx*>(sm + S::regB);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = NT / 32;
  const bool act = tid < C::NTHR;
  const int ti = act ? tid / TC : TR - 1;
  const int tj = act ? tid % TC : TC - 1;
  const CtaTrace trc(P.trace);

  for (int sidx = blockIdx.x; sidx < P.B; sidx += gridDim.x) {
    __syncthreads();
    long long pc_ = clock64();
    const double* pr = P.params + (size_t)sidx * P.pstride;
    const double wl = pr[0], thick = pr[2];
    const double sa = sin(pr[1] * (3.14159265358979323846 / 180.0));
    const double k0d = 2.0 * 3.14159265358979323846 / wl * thick;
    const double nI = P.n_inc, nII = P.n_exit;
    const double* lamg = P.lam + (size_t)sidx * FF;
    const double* Wg = P.Wbuf + (size_t)sidx * FF * FF;
    for (int t = tid; t < FF; t += NT) lamS[t] = lamg[t];
    __syncthreads();
    // ---- sort modes by descending eigenvalue (rank by counting), per-mode scalars, external-medium diagonals ----
    if (tid < FF) {
      const double l = lamS[tid];
      int rk = 0;
      for (int j = 0; j < FF; ++j) { const double lj = lamS[j]; rk += (lj > l) || (lj == l && j < tid); }
      pos[tid] = rk;
      cplx q, X;
      if (l >= 0.0) { const double s = sqrt(l); q = make_double2(s, 0.0); X = make_double2(exp(-k0d * s), 0.0); }
      else { const double s = sqrt(-l); q = make_double2(0.0, s); double sn, cn; sincos(k0d * s, &sn, &cn); X = make_double2(cn, -sn); }
      const bool insig = (l > 0.0) && (X.x < 1e-12);
      if (insig) X = make_double2(0.0, 0.0);
      qv[rk] = q; Xv[rk] = X; XQ[rk] = cmul(X, q);
      const cplx x2 = cmul(X, X);
      omx2[rk] = make_double2(1.0 - x2.x, -x2.y);
      // harmonic tid (original order: 2(p-1) cos, 2(p-1)+1 sin, 2N const)
      const double kap = (tid < z) ? ((tid >> 1) + 1) * sa : 0.0;
      const double aI = nI * nI - kap * kap, aII = nII * nII - kap * kap;
      DI[tid] = (aI >= 0.0) ? make_double2(0.0, sqrt(aI) / (nI * nI)) : make_double2(sqrt(-aI) / (nI * nI), 0.0);
      DII[tid] = (aII >= 0.0) ? make_double2(0.0, sqrt(aII) / (nII * nII)) : make_double2(sqrt(-aII) / (nII * nII), 0.0);
    }
    if (tid == 0) ints[0] = 0;
    __syncthreads();
    if (tid < FF) { if (Xv[tid].x == 0.0 && Xv[tid].y == 0.0) atomicAdd(&ints[0], 1); }
    // ---- W -> shared, columns permuted into sorted order ----
    for (int idx = tid; idx < FF * FF; idx += NT) {
      const int r = idx / FF, j = idx - r * FF;
      Wsm[r * LD + pos[j]] = Wg[idx];
    }
    __syncthreads();
    const int ni = ints[0];            // insignificant modes are exactly the first ni sorted modes (X decreases with lambda)
    const int m = FF - ni;
    if (m > MMAX) {                    // uniform: leave this structure to the generic kernel
      if (tid == 0) atomicOr(&P.wstatus[sidx], PIRL_ST_TAIL_FALLBACK);
      continue;
    }
    for (int t = tid; t < FF; t += NT) {
      wr3[t] = Wsm[0 * LD + t];
      wr3[FE + t] = Wsm[1 * LD + t];
      wr3[2 * FE + t] = Wsm[z * LD + t];
    }
    RK_PHASE(8);

    double tre[RT][CT], tim[RT][CT];
    // =========================== G = Q + W^T D_II W ===========================
#pragma unroll
    for (int a = 0; a < RT; ++a)
#pragma unroll
      for (int b = 0; b < CT; ++b) { tre[a][b] = 0.0; tim[a][b] = 0.0; }
#pragma unroll 1
    for (int r = 0; r < FF; ++r) {
      const cplx d = DII[r];
      const double* wrow = Wsm + r * LD;
      double wi[RT], wj[CT];
#pragma unroll
      for (int a = 0; a < RT; ++a) wi[a] = wrow[ti + TR * a];
#pragma unroll
      for (int b = 0; b < CT; ++b) wj[b] = wrow[tj + TC * b];
      if (d.y == 0.0) {
#pragma unroll
        for (int a = 0; a < RT; ++a) { const double s = d.x * wi[a];
#pragma unroll
          for (int b = 0; b < CT; ++b) tre[a][b] = fma(s, wj[b], tre[a][b]); }
      } else {
#pragma unroll
        for (int a = 0; a < RT; ++a) { const double s = d.y * wi[a];
#pragma unroll
          for (int b = 0; b < CT; ++b) tim[a][b] = fma(s, wj[b], tim[a][b]); }
      }
    }
#pragma unroll
    for (int a = 0; a < RT; ++a)
#pragma unroll
      for (int b = 0; b < CT; ++b)
        if (ti + TR * a == tj + TC * b) { const cplx q = qv[ti + TR * a]; tre[a][b] += q.x; tim[a][b] += q.y; }
    // right-hand sides w_r (3 columns) on the threads tj < 3
    cplx rx[RT];
#pragma unroll
    for (int a = 0; a < RT; ++a) rx[a] = make_double2((tj < 3) ? wr3[tj * FE + ti + TR * a] : 0.0, 0.0);
    RK_PHASE(9);
    // ---- eliminate the first ni pivots (no pivoting) ----
    {
      bool done = false;
#pragma unroll
      for (int kb = 0; kb < CT; ++kb) {
        const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
        for (int kj = 0; kj < TC; ++kj) {
          const int k = kb * TC + kj;
          if (k >= ni) { done = true; break; }
          const int kti = k - ka * TR;
          cplx* rb = rowbuf + (k & 1) * FE;
          cplx* cb = colbuf + (k & 1) * FE;
          cplx* sc = scal + (k & 1) * 8;
          if (act) {
            if (ti == kti) {
#pragma unroll
              for (int b = 0; b < CT; ++b) if (b >= kb) rb[tj + TC * b] = make_double2(tre[ka][b], tim[ka][b]);
              if (tj < 3) sc[1 + tj] = rx[ka];
            }
            if (tj == kj) {
#pragma unroll
              for (int a = 0; a < RT; ++a) if (a >= ka) cb[ti + TR * a] = make_double2(tre[a][kb], tim[a][kb]);
              if (ti == kti) sc[0] = cinv(make_double2(tre[ka][kb], tim[ka][kb]));
            }
          }
          __syncthreads();
          const cplx pinv = sc[0];
          const cplx rk = sc[1 + (tj < 3 ? tj : 0)];
          cplx l[RT];
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= ka) {
              const int i = ti + TR * a;
              l[a] = (i > k) ? cmul(cb[i], pinv) : make_double2(0.0, 0.0);
              if (tj < 3) rx[a] = cfnma(l[a], rk, rx[a]);
            }
#pragma unroll
          for (int b = 0; b < CT; ++b)
            if (b >= kb) {
              const int j = tj + TC * b;
              const cplx r = (j > k) ? rb[j] : make_double2(0.0, 0.0);
#pragma unroll
              for (int a = 0; a < RT; ++a)
                if (a >= ka) {
                  tre[a][b] = fma(-l[a].x, r.x, tre[a][b]); tre[a][b] = fma(l[a].y, r.y, tre[a][b]);
                  tim[a][b] = fma(-l[a].x, r.y, tim[a][b]); tim[a][b] = fma(-l[a].y, r.x, tim[a][b]);
                }
            }
        }
        if (done) break;
      }
    }
    RK_PHASE(10);
    // ---- trailing block -> shared, P = inverse (Gauss-Jordan, no pivoting), g_r = P rhs_r ----
    if (act) {
#pragma unroll
      for (int a = 0; a < RT; ++a) {
        const int i = ti + TR * a;
        if (i >= ni) {
#pragma unroll
          for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; if (j >= ni) Tsm[(i - ni) * m + (j - ni)] = make_double2(tre[a][b], tim[a][b]); }
          if (tj < 3) frs[tj * FE + (i - ni)] = rx[a];      // (frs reused as temporary rhs storage)
        }
      }
    }
    for (int k = 0; k < m; ++k) {
      __syncthreads();
      cplx* rb = rowbuf; cplx* cb = colbuf;
      for (int t = tid; t < m; t += NT) { rb[t] = Tsm[k * m + t]; cb[t] = Tsm[t * m + k]; }
      __syncthreads();
      const cplx p = cinv(rb[k]);
      for (int idx = tid; idx < m * m; idx += NT) {
        const int i = idx / m, j = idx - i * m;
        cplx v;
        if (i == k) v = (j == k) ? p : cmul(rb[j], p);
        else { const cplx f = cmul(cb[i], p); v = (j == k) ? make_double2(-f.x, -f.y) : cfnma(f, rb[j], Tsm[idx]); }
        Tsm[idx] = v;
      }
    }
    __syncthreads();
    for (int idx = tid; idx < 3 * m; idx += NT) {
      const int r = idx / m, i = idx - r * m;
      cplx s = make_double2(0.0, 0.0);
      for (int j = 0; j < m; ++j) s = cfma(Tsm[i * m + j], frs[r * FE + j], s);
      gvec[r * FE + i] = s;
    }
    __syncthreads();
    // P' = P diag(2 X Q): the only form of P used from here on (coupling term of S, and d_s' = P' c_s')
    for (int idx = tid; idx < m * m; idx += NT) {
      const int j = idx % m;
      const cplx v = XQ[ni + j];
      Tsm[idx] = cmul(Tsm[idx], make_double2(2.0 * v.x, 2.0 * v.y));
    }
    RK_PHASE(11);
    // =========================== S = Hq (1 - X^2) + 2Q + [2 (Hq X) P (X Q)] on the significant columns ===========================
#pragma unroll
    for (int a = 0; a < RT; ++a)
#pragma unroll
      for (int b = 0; b < CT; ++b) { tre[a][b] = 0.0; tim[a][b] = 0.0; }
#pragma unroll 1
    for (int r = 0; r < FF; ++r) {
      const cplx d = DI[r];
      const double* wrow = Wsm + r * LD;
      double wi[RT], wj[CT];
#pragma unroll
      for (int a = 0; a < RT; ++a) wi[a] = wrow[ti + TR * a];
#pragma unroll
      for (int b = 0; b < CT; ++b) wj[b] = wrow[tj + TC * b];
      if (d.y == 0.0) {
#pragma unroll
        for (int a = 0; a < RT; ++a) { const double s = d.x * wi[a];
#pragma unroll
          for (int b = 0; b < CT; ++b) tre[a][b] = fma(s, wj[b], tre[a][b]); }
      } else {
#pragma unroll
        for (int a = 0; a < RT; ++a) { const double s = d.y * wi[a];
#pragma unroll
          for (int b = 0; b < CT; ++b) tim[a][b] = fma(s, wj[b], tim[a][b]); }
      }
    }
    __syncthreads();      // all reads of Wsm done: region B becomes HqS
    // Hq = H_I - Q ; publish (Hq X)[:, s'] ; then scale columns by (1 - X^2) and add 2Q
#pragma unroll
    for (int a = 0; a < RT; ++a)
#pragma unroll
      for (int b = 0; b < CT; ++b) {
        const int i = ti + TR * a, j = tj + TC * b;
        cplx h = make_double2(tre[a][b], tim[a][b]);
        const cplx q = qv[j];
        if (i == j) h = csub(h, q);
        if (j >= ni && act) HqS[i * m + (j - ni)] = cmul(h, Xv[j]);
        h = cmul(h, omx2[j]);
        if (i == j) { h.x += 2.0 * q.x; h.y += 2.0 * q.y; }
        tre[a][b] = h.x; tim[a][b] = h.y;
      }
    __syncthreads();
    {
      const int b0 = ni / TC;           // first tile-column block that holds significant columns
#pragma unroll 1
      for (int l = 0; l < m; ++l) {
        cplx hv[RT];
#pragma unroll
        for (int a = 0; a < RT; ++a) hv[a] = HqS[(ti + TR * a) * m + l];
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= b0) {
            const int j = tj + TC * b;
            const cplx pv = (j >= ni) ? Tsm[l * m + (j - ni)] : make_double2(0.0, 0.0);
#pragma unroll
            for (int a = 0; a < RT; ++a) {
              tre[a][b] = fma(hv[a].x, pv.x, tre[a][b]); tre[a][b] = fma(-hv[a].y, pv.y, tre[a][b]);
              tim[a][b] = fma(hv[a].x, pv.y, tim[a][b]); tim[a][b] = fma(hv[a].y, pv.x, tim[a][b]);
            }
          }
      }
    }
    // extra column b = r0 w_z (threads tj == 0), extra rows f_r = (1 - X^2) w_r (threads ti < 3), corner = 0
    cplx bx[RT], fr[CT];
    {
      const cplx r0 = make_double2(DI[z].x, DI[z].y + 1.0 / nI);
#pragma unroll
      for (int a = 0; a < RT; ++a) bx[a] = (tj == 0) ? cscale(r0, wr3[2 * FE + ti + TR * a]) : make_double2(0.0, 0.0);
#pragma unroll
      for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; fr[b] = (ti < 3) ? cscale(omx2[j], wr3[ti * FE + j]) : make_double2(0.0, 0.0); }
    }
    cplx corn = make_double2(0.0, 0.0);
    __syncthreads();      // HqS reads done: region B becomes Ssm later
    RK_PHASE(12);
    {
      bool done = false;
#pragma unroll
      for (int kb = 0; kb < CT; ++kb) {
        const int ka = kb / C::KA_PER_KB;
#pragma unroll 1
        for (int kj = 0; kj < TC; ++kj) {
          const int k = kb * TC + kj;
          if (k >= ni) { done = true; break; }
          const int kti = k - ka * TR;
          cplx* rb = rowbuf + (k & 1) * FE;
          cplx* cb = colbuf + (k & 1) * FE;
          cplx* sc = scal + (k & 1) * 8;
          if (act) {
            if (ti == kti) {
#pragma unroll
              for (int b = 0; b < CT; ++b) if (b >= kb) rb[tj + TC * b] = make_double2(tre[ka][b], tim[ka][b]);
              if (tj == 0) sc[1] = bx[ka];
            }
            if (tj == kj) {
#pragma unroll
              for (int a = 0; a < RT; ++a) if (a >= ka) cb[ti + TR * a] = make_double2(tre[a][kb], tim[a][kb]);
              if (ti == kti) sc[0] = cinv(make_double2(tre[ka][kb], tim[ka][kb]));
              if (ti < 3) sc[2 + ti] = fr[kb];
            }
          }
          __syncthreads();
          const cplx pinv = sc[0];
          const cplx bk = sc[1];
          cplx l[RT];
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= ka) {
              const int i = ti + TR * a;
              l[a] = (i > k) ? cmul(cb[i], pinv) : make_double2(0.0, 0.0);
              if (tj == 0) bx[a] = cfnma(l[a], bk, bx[a]);
            }
          const cplx lr = (ti < 3) ? cmul(sc[2 + ti], pinv) : make_double2(0.0, 0.0);
          if (ti < 3 && tj == 0) corn = cfnma(lr, bk, corn);
#pragma unroll
          for (int b = 0; b < CT; ++b)
            if (b >= kb) {
              const int j = tj + TC * b;
              const cplx r = (j > k) ? rb[j] : make_double2(0.0, 0.0);
              if (ti < 3) fr[b] = cfnma(lr, r, fr[b]);
#pragma unroll
              for (int a = 0; a < RT; ++a)
                if (a >= ka) {
                  tre[a][b] = fma(-l[a].x, r.x, tre[a][b]); tre[a][b] = fma(l[a].y, r.y, tre[a][b]);
                  tim[a][b] = fma(-l[a].x, r.y, tim[a][b]); tim[a][b] = fma(-l[a].y, r.x, tim[a][b]);
                }
            }
        }
        if (done) break;
      }
    }
    RK_PHASE(13);
    // ---- trailing m x m system with partial pivoting in shared memory ----
    const int ms = m + 1;                 // row stride (last column = right-hand side)
    if (act) {
#pragma unroll
      for (int a = 0; a < RT; ++a) {
        const int i = ti + TR * a;
        if (i >= ni) {
#pragma unroll
          for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; if (j >= ni) Ssm[(i - ni) * ms + (j - ni)] = make_double2(tre[a][b], tim[a][b]); }
          if (tj == 0) Ssm[(i - ni) * ms + m] = bx[a];
        }
      }
      if (ti < 3) {
#pragma unroll
        for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; if (j >= ni) frs[ti * FE + (j - ni)] = fr[b]; }
        if (tj == 0) corner[ti] = corn;
      }
    }
    for (int k = 0; k < m; ++k) {
      __syncthreads();
      if (warp == 0) {
        double best = -1.0; int bi = k;
        for (int i = k + lane; i < m; i += 32) {
          const cplx v = Ssm[i * ms + k];
          const double mg = fabs(v.x) + fabs(v.y);
          if (mg > best) { best = mg; bi = i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const double ob = __shfl_xor_sync(0xffffffffu, best, o);
          const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
          if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (lane == 0) ints[1] = bi;
      }
      __syncthreads();
      const int pv = ints[1];
      if (pv != k) {
        for (int j = k + tid; j < ms; j += NT) { const cplx t0 = Ssm[k * ms + j]; Ssm[k * ms + j] = Ssm[pv * ms + j]; Ssm[pv * ms + j] = t0; }
        __syncthreads();
      }
      const cplx pinv = cinv(Ssm[k * ms + k]);
      const int w = ms - (k + 1);
      for (int idx = tid; idx < (m - k - 1) * w; idx += NT) {
        const int i = k + 1 + idx / w, j = k + 1 + idx % w;
        const cplx f = cmul(Ssm[i * ms + k], pinv);
        Ssm[i * ms + j] = cfnma(f, Ssm[k * ms + j], Ssm[i * ms + j]);
      }
    }
    __syncthreads();
    if (warp == 0) {      // back substitution (column oriented) on the right-hand-side column
      for (int k = m - 1; k >= 0; --k) {
        const cplx xk = cmul(Ssm[k * ms + m], cinv(Ssm[k * ms + k]));
        __syncwarp();
        if (lane == 0) cs[k] = xk;
        for (int i = lane; i < k; i += 32) Ssm[i * ms + m] = cfnma(Ssm[i * ms + k], xk, Ssm[i * ms + m]);
        __syncwarp();
      }
    }
    __syncthreads();
    // ---- outputs: y = (q X c)_s', d_s' = 2 P y, T_r = 2 g_r.y, R_r = -corner_r + f_r.c_s' + (w_r X)_s'.d_s' ----
    for (int t = tid; t < m; t += NT) yv[t] = cmul(XQ[ni + t], cs[t]);
    for (int i = warp; i < m; i += NW) {
      cplx s = make_double2(0.0, 0.0);
      for (int j = lane; j < m; j += 32) s = cfma(Tsm[i * m + j], cs[j], s);
      s = warp_sum(s);
      if (lane == 0) dsv[i] = s;
    }
    __syncthreads();
    if (warp < 6) {
      const int r = warp % 3;
      cplx s = make_double2(0.0, 0.0);
      if (warp < 3) {
        for (int j = lane; j < m; j += 32) s = cfma(gvec[r * FE + j], yv[j], s);
        s = warp_sum(s);
        s = make_double2(2.0 * s.x, 2.0 * s.y);
      } else {
        for (int j = lane; j < m; j += 32) {
          s = cfma(frs[r * FE + j], cs[j], s);
          s = cfma(cscale(Xv[ni + j], wr3[r * FE + ni + j]), dsv[j], s);
        }
        s = warp_sum(s);
        s = csub(s, corner[r]);
      }
      if (lane == 0) red[warp] = s;
    }
    __syncthreads();
    if (tid == 0) {
      const double is2 = 0.7071067811865476;
      cplx Tc = red[0], Ts = red[1], T0 = red[2];
      cplx Rc = red[3], Rs = red[4], R0 = red[5];
      R0.x -= 1.0;
      cplx Tm = make_double2((Tc.x - Ts.y) * is2, (Tc.y + Ts.x) * is2);
      cplx Tp = make_double2((Tc.x + Ts.y) * is2, (Tc.y - Ts.x) * is2);
      cplx Rm = make_double2((Rc.x - Rs.y) * is2, (Rc.y + Rs.x) * is2);
      cplx Rp = make_double2((Rc.x + Rs.y) * is2, (Rc.y - Rs.x) * is2);
      const double a1I = nI * nI - sa * sa, a1II = nII * nII - sa * sa;
      const double k1I = a1I >= 0.0 ? sqrt(a1I) : 0.0;
      const double k1II = a1II >= 0.0 ? sqrt(a1II) : 0.0;
      double der[3], det[3];
      der[0] = (Rm.x * Rm.x + Rm.y * Rm.y) * k1I / nI;
      der[1] = (R0.x * R0.x + R0.y * R0.y);
      der[2] = (Rp.x * Rp.x + Rp.y * Rp.y) * k1I / nI;
      det[0] = (Tm.x * Tm.x + Tm.y * Tm.y) * k1II / (nII * nII) * nI;
      det[1] = (T0.x * T0.x + T0.y * T0.y) * nI / nII;
      det[2] = (Tp.x * Tp.x + Tp.y * Tp.y) * k1II / (nII * nII) * nI;
      int st = P.wstatus[sidx];
      if (!isfinite(det[0]) || !isfinite(det[1]) || !isfinite(der[1])) st |= PIRL_ST_NONFINITE_;
      P.eff[sidx] = det[0];
      if (P.de_r3) { P.de_r3[3 * sidx] = der[0]; P.de_r3[3 * sidx + 1] = der[1]; P.de_r3[3 * sidx + 2] = der[2]; }
      if (P.de_t3) { P.de_t3[3 * sidx] = det[0]; P.de_t3[3 * sidx + 1] = det[1]; P.de_t3[3 * sidx + 2] = det[2]; }
      if (P.status) P.status[sidx] = st;
    }
    RK_PHASE(14);
  }
  trc.done(P.trace, 6, P.B);
}
