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

template <int FF_, int TR_, int TC_, int RT_, int CT_, int MINB_>
struct TileCfg {
  static constexpr int FF = FF_, TR = TR_, TC = TC_, RT = RT_, CT = CT_, MINB = MINB_;
  static constexpr int N = (FF - 1) / 2;
  static constexpr int NTHR = TR * TC;
  static constexpr int NT = ((NTHR + 31) / 32) * 32;
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

template <class C>
__global__ void __launch_bounds__(C::NT, C::MINB) rk_front(KParams P) {
  constexpr int FF = C::FF, TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, N = C::N, LD = C::LD, NT = C::NT;
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

  for (int t = tid; t < nc; t += NT) twS[t] = P.tw[t];

  for (int sidx = blockIdx.x; sidx < P.B; sidx += gridDim.x) {
    __syncthreads();
    if (tid == 0) sbad = 0;
    const double* pr = P.params + (size_t)sidx * P.pstride;
    const double sa = sin(pr[1] * (3.14159265358979323846 / 180.0));
    const double nr = pr[3];
    const double er = nr * nr, eg = P.eps_g;
    int bad = 0;
    for (int p = tid; p < nc; p += NT) su[p] = P.patterns[(size_t)sidx * nc + p] > 0;
    __syncthreads();
    // ---- Fourier coefficients c_k, k = 0..2N (closed form of the up-sampled FFT): one thread per k ----
    for (int k = tid; k <= 2 * N; k += NT) {
      double sx = 0.0, sy = 0.0;
      int idx = 0;
      for (int p = 0; p < nc; ++p) {
        if (su[p]) { const cplx w = twS[idx]; sx += w.x; sy += w.y; }
        idx += k; if (idx >= nc) idx -= nc;
      }
      const cplx d = __ldg(&P.Dk[k]);
      cre[k] = sx * d.x - sy * d.y;
      cim[k] = sx * d.y + sy * d.x;
    }
    __syncthreads();

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
          t[a][b] = fma(da, basis_toeplitz(pi, pj, N, cre, cim), (i == j) ? ia : 0.0);
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
    // =========================== phase 2: E, Gauss-Jordan inverse, B' ===========================
    {
      const double de = er - eg;
#pragma unroll
      for (int a = 0; a < RT; ++a)
#pragma unroll
        for (int b = 0; b < CT; ++b) {
          const int i = ti + TR * a, j = tj + TC * b;
          t[a][b] = fma(de, basis_toeplitz(i, j, N, cre, cim), (i == j) ? eg : 0.0);
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
        if (act && tj == kj) {
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= ka) { const int i = ti + TR * a; if (i > k) xb[i] = t[a][kb]; }
          if (ti == kti) dd[k] = t[ka][kb];
        }
        __syncthreads();
        double s = 0.0;
        for (int i = k + 1 + lane; i < FF; i += 32) { const double x = xb[i]; s = fma(x, x, s); }
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
        double vi[RT], vj[CT];
#pragma unroll
        for (int a = 0; a < RT; ++a)
          if (a >= ka) { const int i = ti + TR * a; vi[a] = (i > k) ? ((i == k + 1) ? v0 : xb[i]) : 0.0; }
#pragma unroll
        for (int b = 0; b < CT; ++b)
          if (b >= kb) { const int j = tj + TC * b; vj[b] = (j > k) ? ((j == k + 1) ? v0 : xb[j]) : 0.0; }
        // partial row sums of C22 v
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
        if (tid < FF) {
          const int r = tid;
          if (r > k) {
            double acc = 0.0;
#pragma unroll
            for (int c = 0; c < TC; ++c) acc += red[r * TC + c];
            pvec[r] = beta * acc;
            Lsm[k * LD + r] = (r == k + 1) ? v0 : xb[r];     // reflector -> upper triangle of the shared matrix
          } else if (r == k) {
            Lsm[k * LD + k] = beta;
            ee[k] = alpha;
          }
        }
        __syncthreads();
        double kk = 0.0;
        for (int i = k + 1 + lane; i < FF; i += 32) kk = fma(pvec[i], (i == k + 1) ? v0 : xb[i], kk);
        kk = 0.5 * beta * warp_sum(kk);
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
      }
    }
    // trailing 2x2 block
    if (act) {
      constexpr int a2 = (FF - 2) / TR, a1 = (FF - 1) / TR, b2 = (FF - 2) / TC, b1 = (FF - 1) / TC;
      if (ti == (FF - 2) % TR && tj == (FF - 2) % TC) dd[FF - 2] = t[a2][b2];
      if (ti == (FF - 1) % TR && tj == (FF - 1) % TC) { dd[FF - 1] = t[a1][b1]; ee[FF - 1] = 0.0; }
      if (ti == (FF - 1) % TR && tj == (FF - 2) % TC) ee[FF - 2] = t[a1][b2];
    }
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
    // ---- store: t(i, j) = M0'[j][i] (swapped ordering);  old layout Wbuf[row][mode] in ORIGINAL harmonic order ----
    if (act) {
      double* Wg = P.Wbuf + (size_t)sidx * FF * FF;
#pragma unroll
      for (int b = 0; b < CT; ++b) {
        const int j = tj + TC * b;
        const int row = (j < z) ? (j ^ 1) : j;
#pragma unroll
        for (int a = 0; a < RT; ++a) Wg[row * FF + ti + TR * a] = t[a][b];
      }
    }
    if (bad) sbad = 1;
    __syncthreads();
    double* tg = P.tri + (size_t)sidx * 2 * FF;
    for (int q = tid; q < FF; q += NT) { tg[q] = dd[q]; tg[FF + q] = ee[q]; }
    if (tid == 0) P.wstatus[sidx] = sbad ? PIRL_ST_NPD : 0;
  }
}
