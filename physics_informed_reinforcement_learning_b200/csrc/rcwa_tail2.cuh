// rcwa_tail2.cuh -- layer solve in the mode basis, REAL-arithmetic bordered formulation (sm_100a register tiles).
//
// Replaces _base.py:159, transfer_method.py:42-75 and solver.py:29-38 of the reference (tools/proto_tail2.py is the NumPy
// statement of exactly these steps, checked against the oracle fixtures to 1.4e-11).
//
// Modes are sorted by DESCENDING eigenvalue: [I insignificant evanescent | S significant evanescent | P propagating].
// Unknowns are the top field u and the bottom field hb of the layer in the mode basis (impedance form, C = q coth(k0 d q),
// Sg = q csch(k0 d q)) for the evanescent modes, and (c, hb) in amplitude form -- no resonance singularity -- for the few
// propagating modes.  The external media enter as H = W^T D W with D diagonal: real (>= 0) on evanescent harmonics,
// purely imaginary on the (three) propagating harmonics.  The imaginary rank-3 parts become auxiliary unknowns
//     tau_r = i bI_r (w_r . h_t),   sigma_r = i bII_r (w_r . hb),
// so that everything attached to an evanescent mode is a REAL symmetric positive definite matrix
//     K = [[R_I + C, -Sg], [-Sg, R_II + C]],   R = W^T Re(D) W,
// eliminated without pivoting in real arithmetic (one real 81 x 81 register tile at a time):
//   top   : LDL^T steps on the I pivots, symmetric sweep (Gauss-Jordan) steps on the S pivots -> (Schur_S)^-1, parked in
//           shared memory; bottom: LDL^T steps on I, fold the parked block in (tile and cross columns), LDL^T steps on S.
// What is left is a complex system of size 3 n_p + n_hI + n_hII (about 24) for [h_t,P | tau | hb_P | sigma | c_P], solved
// with partial pivoting in shared memory; its tau / sigma entries ARE the outputs: R_r = tau_r / (i bI_r) - delta_r,const,
// T_r = sigma_r / (i bII_r).
#pragma once
#include "rcwa_regtile.cuh"

#define PIRL_T2_XTOL 1e-11        // modes with exp(-k0 d q) below this do not couple top and bottom of the layer

template <class C, int NSMAX_, int NPMAX_, int NHMAX_>
struct Tail2Smem {
  static constexpr int FF = C::FF, FE = pirl_even(C::FF), NSMAX = NSMAX_, NPMAX = NPMAX_, NHMAX = NHMAX_;
  static constexpr int NXB = 2 * C::TC;                       // extra columns of the bottom system (two per thread column)
  static constexpr int NBMAX = 3 * NPMAX + 2 * NHMAX;         // border unknowns
  static constexpr int NCT = NPMAX + NHMAX;                   // top corner size
  static constexpr int NCB = NPMAX + NXB;                     // bottom corner size
  // doubles
  static constexpr int lamS = 0;
  static constexpr int Cd = lamS + FE;
  static constexpr int Sg = Cd + FE;
  static constexpr int dre = Sg + FE;                         // [2][FE]  Re D_I, Re D_II per harmonic
  static constexpr int beta = dre + 2 * FE;                   // [2][NHMAX] Im D of the propagating harmonics
  static constexpr int qP = beta + 2 * NHMAX;                 // [NPMAX] |q| of the propagating modes
  static constexpr int XP = qP + pirl_even(NPMAX);            // [NPMAX] complex
  static constexpr int cb = XP + 2 * NPMAX;                   // [2][FE] published pivot column
  static constexpr int sc = cb + 2 * FE;                      // [2][NXB + 2] published row of the extra columns, 1/pivot, pivot
  static constexpr int Yt = sc + 2 * (NXB + 2);               // [FF][NHMAX]
  static constexpr int dt = Yt + pirl_even(C::FF * NHMAX);    // [FE]
  static constexpr int CT = dt + FE;                          // [NCT][NCT]
  static constexpr int CB = CT + pirl_even(NCT * NCT);        // [NCB][NCB]
  static constexpr int parkSS = CB + pirl_even(NCB * NCB);    // [NSMAX][NSMAX]
  static constexpr int parkX = parkSS + NSMAX * NSMAX;        // [NSMAX][NPMAX + NHMAX]
  static constexpr int ints = parkX + NSMAX * (NPMAX + NHMAX);// pos[FE] ints, pI[NHMAX], pII[NHMAX], counters
  static constexpr int Wsm = pirl_even(ints + FE / 2 + NHMAX + 16);   // [FF][LD]; later the border system
  static constexpr int zsys = 2 * NBMAX * (NBMAX + 1);        // complex border matrix incl. right-hand side
  // the W region is dead after the second product: it then holds the border system and the bottom pivot rows
  static constexpr int Yb = Wsm + zsys;                       // [FF][NXB]
  static constexpr int db = Yb + pirl_even(C::FF * NXB);      // [FE]
  static constexpr int wregion = (C::FF * C::LD > db + FE - Wsm) ? C::FF * C::LD : db + FE - Wsm;
  static constexpr int doubles = Wsm + wregion;
};

// ---------------------------------------------------------------------------------------------------------------------
// LDL^T steps on pivots [kbeg, kend) of the symmetric tile t (rows/cols in the cyclic layout of TileCfg), carrying the
// extra columns x0 (column tj) and x1 (column TC + tj).  Only the pivot COLUMN is published (the tile is symmetric).
// Saves the pivot and the pivot row of the extra columns (Ys[k][0..NXS), ds[k]) for the corner term -Y^T D^-1 Y.
// ---------------------------------------------------------------------------------------------------------------------
template <class C, bool TWO, int NXS>
__device__ __forceinline__ void t2_ldlt_range(double (&t)[C::RT][C::CT], double (&x0)[C::RT], double (&x1)[C::RT], int kbeg, int kend,
                                              double* cbuf, double* scbuf, double* Ys, double* ds, int nx, bool act, int ti, int tj,
                                              int tid, int& bad) {
  constexpr int TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, FE = pirl_even(C::FF);
  constexpr int SCW = 2 * TC + 2;
  bool done = false;
#pragma unroll
  for (int kb = 0; kb < CT; ++kb) {
    const int ka = kb / C::KA_PER_KB;
    if ((kb + 1) * TC <= kbeg) continue;          // uniform
#pragma unroll 1
    for (int kj = 0; kj < TC; ++kj) {
      const int k = kb * TC + kj;
      if (k < kbeg) continue;
      if (k >= kend) { done = true; break; }
      const int kti = k - ka * TR;
      double* cb = cbuf + (k & 1) * FE;
      double* sc = scbuf + (k & 1) * SCW;
      // publisher side does all the masking: finished rows (and the pivot itself) are published as zeros, the pivot
      // owner also publishes 1/d, so the readers' inner step is loads + multiply-adds only
      if (act) {
        if (tj == kj) {
#pragma unroll
          for (int a = 0; a < RT; ++a) if (a >= ka) { const int i = ti + TR * a; cb[i] = (i > k) ? t[a][kb] : 0.0; }
          if (ti == kti) { const double d = t[ka][kb]; sc[SCW - 2] = 1.0 / d; sc[SCW - 1] = d; }
        }
        if (ti == kti) {
          sc[tj] = x0[ka];
          if (TWO) sc[TC + tj] = x1[ka];
        }
      }
      __syncthreads();
      const double pinv = sc[SCW - 2];
      if (tid < nx) Ys[k * NXS + tid] = sc[tid];
      if (tid == 32) ds[k] = sc[SCW - 1];
      double l[RT];
#pragma unroll
      for (int a = 0; a < RT; ++a) if (a >= ka) l[a] = cb[ti + TR * a] * pinv;
      const double y0 = sc[tj];
#pragma unroll
      for (int a = 0; a < RT; ++a) if (a >= ka) x0[a] = fma(-l[a], y0, x0[a]);
      if (TWO) {
        const double y1 = sc[TC + tj];
#pragma unroll
        for (int a = 0; a < RT; ++a) if (a >= ka) x1[a] = fma(-l[a], y1, x1[a]);
      }
#pragma unroll
      for (int b = 0; b < CT; ++b)
        if (b >= kb) {
          const double r = cb[tj + TC * b];
#pragma unroll
          for (int a = 0; a < RT; ++a) if (a >= ka) t[a][b] = fma(-l[a], r, t[a][b]);
        }
    }
    if (done) break;
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Symmetric sweep steps on pivots [kbeg, kend) inside the window rows / columns >= kbeg:  after the sweeps
//   t[S,S] = -(Schur_S)^-1,  t[S,R] = (Schur_S)^-1 M_SR,  t[R,R] = Schur complement,  x0[S] = (Schur_S)^-1 x_S,  x0[R] = Schur.
// The tile stays symmetric, so again only the pivot column is published.
// ---------------------------------------------------------------------------------------------------------------------
template <class C, int NXS>
__device__ __forceinline__ void t2_sweep_range(double (&t)[C::RT][C::CT], double (&x0)[C::RT], int kbeg, int kend, double* cbuf,
                                               double* scbuf, double* Ys, double* ds, int nx, bool act, int ti, int tj, int tid,
                                               int& bad) {
  constexpr int TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, FE = pirl_even(C::FF);
  constexpr int SCW = 2 * TC + 2;
  const int a0 = kbeg / TR, b0 = kbeg / TC;     // first tile row / column blocks that intersect the window
  bool done = false;
#pragma unroll
  for (int kb = 0; kb < CT; ++kb) {
    const int ka = kb / C::KA_PER_KB;
    if ((kb + 1) * TC <= kbeg) continue;          // uniform
#pragma unroll 1
    for (int kj = 0; kj < TC; ++kj) {
      const int k = kb * TC + kj;
      if (k < kbeg) continue;
      if (k >= kend) { done = true; break; }
      const int kti = k - ka * TR;
      double* cb = cbuf + (k & 1) * FE;
      double* sc = scbuf + (k & 1) * SCW;
      // published column: zeros on finished rows, d - 1 at the pivot (so that u_k = -(d - 1) p = p - 1 and v_k = d - 1 need
      // no special case on the reader side); the pivot owner also publishes 1/d
      if (act) {
        if (tj == kj) {
#pragma unroll
          for (int a = 0; a < RT; ++a)
            if (a >= a0) { const int i = ti + TR * a; cb[i] = (i >= kbeg) ? ((i == k) ? t[a][kb] - 1.0 : t[a][kb]) : 0.0; }
          if (ti == kti) { const double d = t[ka][kb]; sc[SCW - 2] = 1.0 / d; sc[SCW - 1] = d; }
        }
        if (ti == kti) sc[tj] = x0[ka];
      }
      __syncthreads();
      const double p = sc[SCW - 2];
      if (tid < nx) Ys[k * NXS + tid] = sc[tid];
      if (tid == 32) ds[k] = sc[SCW - 1];
      double u[RT];
#pragma unroll
      for (int a = 0; a < RT; ++a) if (a >= a0) u[a] = -cb[ti + TR * a] * p;
      const double y0 = sc[tj];
#pragma unroll
      for (int a = 0; a < RT; ++a) if (a >= a0) x0[a] = fma(u[a], y0, x0[a]);
#pragma unroll
      for (int b = 0; b < CT; ++b)
        if (b >= b0) {
          const double v = cb[tj + TC * b];
#pragma unroll
          for (int a = 0; a < RT; ++a) if (a >= a0) t[a][b] = fma(u[a], v, t[a][b]);
        }
      if (ti == kti && tj == kj) t[ka][kb] = -p;
    }
    if (done) break;
  }
}

// accumulate t += W^T diag(dre) W (rows of Wsm are harmonics, columns are the sorted modes); harmonics with dre == 0
// (the propagating ones) are skipped -- uniform branch
template <class C>
__device__ __forceinline__ void t2_product(double (&t)[C::RT][C::CT], const double* Wsm, const double* dre, int ti, int tj) {
  constexpr int FF = C::FF, TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, LD = C::LD;
#pragma unroll
  for (int a = 0; a < RT; ++a)
#pragma unroll
    for (int b = 0; b < CT; ++b) t[a][b] = 0.0;
#pragma unroll 2
  for (int r = 0; r < FF; ++r) {
    const double d = dre[r];
    if (d == 0.0) continue;
    const double* wrow = Wsm + r * LD;
    double wi[RT], wj[CT];
#pragma unroll
    for (int a = 0; a < RT; ++a) wi[a] = d * wrow[ti + TR * a];
#pragma unroll
    for (int b = 0; b < CT; ++b) wj[b] = wrow[tj + TC * b];
#pragma unroll
    for (int a = 0; a < RT; ++a)
#pragma unroll
      for (int b = 0; b < CT; ++b) t[a][b] = fma(wi[a], wj[b], t[a][b]);
  }
}

template <class C, int NSMAX, int NPMAX, int NHMAX>
__global__ void __maxnreg__(C::MAXREG) rk_tail2(KParams P) {
  constexpr int FF = C::FF, TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, N = C::N, LD = C::LD, NT = C::NT;
  constexpr int FFR = C::FFR;            // real modes / harmonics; the FF - FFR padded ones are decoupled, strongly evanescent dummies
  constexpr bool PAD = FFR != FF;
  constexpr int z = 2 * N;
  using S = Tail2Smem<C, NSMAX, NPMAX, NHMAX>;
  constexpr int FE = S::FE, NXB = S::NXB, NCT = S::NCT, NCB = S::NCB, NPX = NPMAX + NHMAX;
  extern __shared__ __align__(16) double sm[];
  double* lamS = sm + S::lamS;
  double* Cd = sm + S::Cd;
  double* Sg = sm + S::Sg;
  double* dre = sm + S::dre;
  double* beta = sm + S::beta;
  double* qP = sm + S::qP;
  cplx* XP = reinterpret_cast<cplx*>(sm + S::XP);
  double* cbuf = sm + S::cb;
  double* scbuf = sm + S::sc;
  double* Yt = sm + S::Yt;
  double* dt = sm + S::dt;
  double* Yb = sm + S::Yb;
  double* db = sm + S::db;
  double* CTm = sm + S::CT;
  double* CBm = sm + S::CB;
  double* parkSS = sm + S::parkSS;
  double* parkX = sm + S::parkX;
  int* pos = reinterpret_cast<int*>(sm + S::ints);
  int* pI = pos + FE;
  int* pII = pI + NHMAX;
  int* cnt = pII + NHMAX;                 // [0] n_i, [1] n_e, [2] n_hI, [3] n_hII, [4] pivot row of the border solve
  double* Wsm = sm + S::Wsm;
  cplx* Zs = reinterpret_cast<cplx*>(sm + S::Wsm);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool act = tid < C::NTHR;
  const int ti = act ? tid / TC : TR - 1;
  const int tj = act ? tid % TC : TC - 1;
  const CtaTrace trc(P.trace);
  __shared__ int snext;

  for (int sidx = blockIdx.x; sidx < P.B; sidx = pirl_next_item(P.queue, sidx, &snext)) {
    __syncthreads();
    long long pc_ = clock64();
    const double* pr = P.params + (size_t)sidx * P.pstride;
    const double wl = pr[0], thick = pr[2];
    const double sa = sin(pr[1] * (3.14159265358979323846 / 180.0));
    const double k0d = 2.0 * 3.14159265358979323846 / wl * thick;
    const double nI = P.n_inc, nII = P.n_exit;
    const double* lamg = P.lam + (size_t)sidx * FFR;
    const double* Wg = P.Wbuf + (size_t)sidx * FFR * FFR;
    for (int t = tid; t < FF; t += NT) lamS[t] = (!PAD || t < FFR) ? lamg[t] : 1e12;   // padded modes: q = 1e6, X = 0
    if (tid < 8) cnt[tid] = 0;
    __syncthreads();
    // ---- sort modes by descending eigenvalue (rank by counting); slab scalars; external-medium diagonals ----
    int myrank = 0;
    double myl = 0.0;
    if (tid < FF) {
      const double l = lamS[tid];
      myl = l;
      int rk = 0;
      for (int j = 0; j < FF; ++j) { const double lj = lamS[j]; rk += (lj > l) || (lj == l && j < tid); }
      pos[tid] = rk;
      myrank = rk;
      double cdv = 0.0, sgv = 0.0;
      if (l > 0.0) {
        const double qq = sqrt(l), x = k0d * qq, X = exp(-x);
        if (X < PIRL_T2_XTOL) { cdv = qq; atomicAdd(&cnt[0], 1); }
        else { const double om = -expm1(-2.0 * x); cdv = qq * (1.0 + X * X) / om; sgv = 2.0 * qq * X / om; }
        atomicAdd(&cnt[1], 1);
      }
      Cd[rk] = cdv; Sg[rk] = sgv;
      // harmonic tid (original order: 2(p-1) cos_p, 2(p-1)+1 sin_p, 2N const)
      const bool padh = PAD && tid >= FFR;      // padded harmonic: no admittance, not a propagating order
      const double kap = (tid < z) ? ((tid >> 1) + 1) * sa : 0.0;
      const double aI = padh ? 0.0 : nI * nI - kap * kap, aII = padh ? 0.0 : nII * nII - kap * kap;
      dre[tid] = (aI < 0.0) ? sqrt(-aI) / (nI * nI) : 0.0;
      dre[FE + tid] = (aII < 0.0) ? sqrt(-aII) / (nII * nII) : 0.0;
      if (aI > 0.0 || aII > 0.0) {       // propagating harmonic: its slot in the border lists (ordered by index)
        int rI = 0, rII = 0;
        for (int r = 0; r < tid; ++r) {
          const double kr = (r < z) ? ((r >> 1) + 1) * sa : 0.0;
          rI += (nI * nI - kr * kr > 0.0); rII += (nII * nII - kr * kr > 0.0);
        }
        if (aI > 0.0) { if (rI < NHMAX) { pI[rI] = tid; beta[rI] = sqrt(aI) / (nI * nI); } atomicAdd(&cnt[2], 1); }
        if (aII > 0.0) { if (rII < NHMAX) { pII[rII] = tid; beta[NHMAX + rII] = sqrt(aII) / (nII * nII); } atomicAdd(&cnt[3], 1); }
      }
    }
    __syncthreads();
    const int ni = cnt[0], ne = cnt[1], nhI = cnt[2], nhII = cnt[3];
    const int np = FF - ne, ns = ne - ni;
    const int nxb = nhII + np + nhI;
    if (ns > NSMAX || np > NPMAX || nhI > NHMAX || nhII > NHMAX || nxb > NXB) {   // uniform: leave it to the generic kernel
      if (tid == 0) atomicOr(&P.wstatus[sidx], PIRL_ST_TAIL_FALLBACK);
      continue;
    }
    if (tid < FF && !(myl > 0.0)) {
      const int j = myrank - ne;
      const double s = sqrt(-myl);
      double sn, cn;
      sincos(k0d * s, &sn, &cn);
      qP[j] = s; XP[j] = make_double2(cn, -sn);
    }
    // ---- W -> shared, columns permuted into sorted order ----
    for (int idx = tid; idx < FF * FF; idx += NT) {
      const int r = idx / FF, j = idx - r * FF;
      if (!PAD) Wsm[r * LD + pos[j]] = Wg[idx];
      else Wsm[r * LD + pos[j]] = (r < FFR && j < FFR) ? Wg[r * FFR + j] : 0.0;
    }
    __syncthreads();
    RK_PHASE(8);
    int bad = 0;
    double t[RT][CT], x0[RT], x1[RT];
    // =========================== top: R_I + C, eliminate I (LDL^T), sweep S ===========================
    t2_product<C>(t, Wsm, dre, ti, tj);
#pragma unroll
    for (int a = 0; a < RT; ++a) {
      const int i = ti + TR * a;
#pragma unroll
      for (int b = 0; b < CT; ++b) if (i == tj + TC * b) t[a][b] += Cd[i];
      x0[a] = (tj < nhI) ? Wsm[pI[tj] * LD + i] : 0.0;
      x1[a] = 0.0;
    }
    RK_PHASE(9);
    t2_ldlt_range<C, false, NHMAX>(t, x0, x1, 0, ni, cbuf, scbuf, Yt, dt, nhI, act, ti, tj, tid, bad);
    RK_PHASE(10);
    t2_sweep_range<C, NHMAX>(t, x0, ni, ne, cbuf, scbuf, Yt, dt, nhI, act, ti, tj, tid, bad);
    RK_PHASE(11);
    // ---- park: Sg (Schur_S)^-1 Sg, Sg G_P, Sg G_w; top corner blocks ----
    if (act) {
#pragma unroll
      for (int a = 0; a < RT; ++a) {
        const int i = ti + TR * a;
        if (i >= ni) {
          const double sgi = Sg[i];
#pragma unroll
          for (int b = 0; b < CT; ++b) {
            const int j = tj + TC * b;
            if (i < ne) {
              if (j >= ne) parkX[(i - ni) * NPX + (j - ne)] = sgi * t[a][b];
              else if (j >= ni) parkSS[(i - ni) * NSMAX + (j - ni)] = -sgi * Sg[j] * t[a][b];
            } else if (j >= ne) CTm[(i - ne) * NCT + (j - ne)] = t[a][b];
          }
          if (tj < nhI) {
            if (i < ne) parkX[(i - ni) * NPX + np + tj] = sgi * x0[a];
            else { CTm[(i - ne) * NCT + np + tj] = x0[a]; CTm[(np + tj) * NCT + (i - ne)] = x0[a]; }
          }
        }
      }
    }
    // =========================== bottom: R_II + C, eliminate I, fold the top in, eliminate S ===========================
    t2_product<C>(t, Wsm, dre + FE, ti, tj);
#pragma unroll
    for (int a = 0; a < RT; ++a) {
      const int i = ti + TR * a;
#pragma unroll
      for (int b = 0; b < CT; ++b) if (i == tj + TC * b) t[a][b] += Cd[i];
      x0[a] = (tj < nhII) ? Wsm[pII[tj] * LD + i] : 0.0;
      x1[a] = (TC + tj < nhII) ? Wsm[pII[(TC + tj < NHMAX) ? TC + tj : 0] * LD + i] : 0.0;
    }
    RK_PHASE(12);
    t2_ldlt_range<C, true, NXB>(t, x0, x1, 0, ni, cbuf, scbuf, Yb, db, nxb, act, ti, tj, tid, bad);
    __syncthreads();      // parked blocks are visible; all reads of Wsm are done
    if (act) {
#pragma unroll
      for (int a = 0; a < RT; ++a) {
        const int i = ti + TR * a;
        if (i >= ni && i < ne) {
#pragma unroll
          for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; if (j >= ni && j < ne) t[a][b] -= parkSS[(i - ni) * NSMAX + (j - ni)]; }
          const int c0 = tj - nhII, c1 = TC + tj - nhII;
          if (c0 >= 0 && c0 < np + nhI) x0[a] += parkX[(i - ni) * NPX + c0];
          if (c1 >= 0 && c1 < np + nhI) x1[a] += parkX[(i - ni) * NPX + c1];
        }
      }
    }
    t2_ldlt_range<C, true, NXB>(t, x0, x1, ni, ne, cbuf, scbuf, Yb, db, nxb, act, ti, tj, tid, bad);
    RK_PHASE(13);
    // ---- bottom corner blocks ----
    if (act) {
#pragma unroll
      for (int a = 0; a < RT; ++a) {
        const int i = ti + TR * a;
        if (i >= ne) {
#pragma unroll
          for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; if (j >= ne) CBm[(i - ne) * NCB + (j - ne)] = t[a][b]; }
          CBm[(i - ne) * NCB + np + tj] = x0[a]; CBm[(np + tj) * NCB + (i - ne)] = x0[a];
          CBm[(i - ne) * NCB + np + TC + tj] = x1[a]; CBm[(np + TC + tj) * NCB + (i - ne)] = x1[a];
        }
      }
    }
    __syncthreads();
    // corner terms -Y^T D^-1 Y of the eliminated pivots
    double* rdt = cbuf;                    // reciprocals of the pivots (the published-column buffers are free now)
    double* rdb = cbuf + FE;
    for (int k = tid; k < ne; k += NT) {
      if (!(dt[k] > 0.0) || !(db[k] > 0.0)) bad = 1;
      rdt[k] = 1.0 / dt[k]; rdb[k] = 1.0 / db[k];
    }
    __syncthreads();
    for (int idx = tid; idx < nhI * nhI + nxb * nxb; idx += NT) {
      double s0 = 0.0, s1 = 0.0;            // two chains over the ~75 pivots
      if (idx < nhI * nhI) {
        const int a = idx / nhI, b = idx - a * nhI;
        int k = 0;
        for (; k + 1 < ne; k += 2) {
          s0 = fma(Yt[k * NHMAX + a] * rdt[k], Yt[k * NHMAX + b], s0);
          s1 = fma(Yt[(k + 1) * NHMAX + a] * rdt[k + 1], Yt[(k + 1) * NHMAX + b], s1);
        }
        if (k < ne) s0 = fma(Yt[k * NHMAX + a] * rdt[k], Yt[k * NHMAX + b], s0);
        CTm[(np + a) * NCT + np + b] = -(s0 + s1);
      } else {
        const int q = idx - nhI * nhI, a = q / nxb, b = q - a * nxb;
        int k = 0;
        for (; k + 1 < ne; k += 2) {
          s0 = fma(Yb[k * NXB + a] * rdb[k], Yb[k * NXB + b], s0);
          s1 = fma(Yb[(k + 1) * NXB + a] * rdb[k + 1], Yb[(k + 1) * NXB + b], s1);
        }
        if (k < ne) s0 = fma(Yb[k * NXB + a] * rdb[k], Yb[k * NXB + b], s0);
        CBm[(np + a) * NCB + np + b] = -(s0 + s1);
      }
    }
    if (bad) cnt[5] = 1;       // a non-positive pivot: K was not positive definite
    __syncthreads();
    RK_PHASE(14);
    // =========================== border system [h_t,P | tau | hb_P | sigma | c_P] ===========================
    const int nb = 3 * np + nhI + nhII, ms = nb + 1;
    const int o_tau = np, o_hb = np + nhI, o_sig = 2 * np + nhI, o_c = 2 * np + nhI + nhII;
    for (int idx = tid; idx < nb * ms; idx += NT) {
      const int r = idx / ms, c = idx - r * ms;
      cplx v = make_double2(0.0, 0.0);
      // decode: type 0 h_t,P  1 tau  2 hb_P  3 sigma  4 c_P ; index within the type
      const int tr = (r < o_tau) ? 0 : (r < o_hb) ? 1 : (r < o_sig) ? 2 : (r < o_c) ? 3 : 4;
      const int ir = r - ((tr == 0) ? 0 : (tr == 1) ? o_tau : (tr == 2) ? o_hb : (tr == 3) ? o_sig : o_c);
      if (c == nb) {        // right-hand side: the incident wave sits in the tau row of the constant harmonic
        if (tr == 1 && pI[ir] == z) v.x = (beta[ir] + 1.0 / nI) / beta[ir];
      } else {
        const int tc = (c < o_tau) ? 0 : (c < o_hb) ? 1 : (c < o_sig) ? 2 : (c < o_c) ? 3 : 4;
        const int ic = c - ((tc == 0) ? 0 : (tc == 1) ? o_tau : (tc == 2) ? o_hb : (tc == 3) ? o_sig : o_c);
        // real symmetric corners left by the eliminated evanescent unknowns
        const int rt = (tr == 0) ? ir : (tr == 1) ? np + ir : -1;
        const int ct = (tc == 0) ? ic : (tc == 1) ? np + ic : -1;
        if (rt >= 0 && ct >= 0) v.x += CTm[rt * NCT + ct];
        const int rb = (tr == 2) ? ir : (tr == 3) ? np + ir : (tr == 0) ? np + nhII + ir : (tr == 1) ? np + nhII + np + ir : -1;
        const int cbi = (tc == 2) ? ic : (tc == 3) ? np + ic : (tc == 0) ? np + nhII + ic : (tc == 1) ? np + nhII + np + ic : -1;
        if (rb >= 0 && cbi >= 0) v.x += CBm[rb * NCB + cbi];
        // slab terms of the propagating modes (q = i |q|), radiation terms of the propagating harmonics
        if (tr == 0 || tr == 2 || tr == 4) {
          const cplx X = XP[ir];
          const cplx q = make_double2(0.0, qP[ir]);
          const cplx X2 = cmul(X, X);
          if (tr == 0) {
            if (tc == 4 && ic == ir) v = cadd(v, cmul(q, make_double2(1.0 + X2.x, X2.y)));
            if (tc == 2 && ic == ir) v = csub(v, cmul(q, X));
          } else if (tr == 2) {
            if (tc == 2 && ic == ir) v = cadd(v, q);
            if (tc == 4 && ic == ir) v = csub(v, cscale(cmul(q, X), 2.0));
          } else {
            if (tc == 0 && ic == ir) v.x += 1.0;
            if (tc == 4 && ic == ir) v = csub(v, make_double2(1.0 - X2.x, -X2.y));
            if (tc == 2 && ic == ir) v = csub(v, X);
          }
        } else if (tr == tc && ir == ic) {
          v.y += 1.0 / beta[(tr == 1) ? ir : NHMAX + ir];
        }
      }
      Zs[idx] = v;
    }
    // ---- Gaussian elimination with partial pivoting on the border system ----
    for (int k = 0; k < nb; ++k) {
      __syncthreads();
      if (warp == 0) {
        double best = -1.0; int bi = k;
        for (int i = k + lane; i < nb; i += 32) {
          const cplx v = Zs[i * ms + k];
          const double mg = fabs(v.x) + fabs(v.y);
          if (mg > best) { best = mg; bi = i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const double ob = __shfl_xor_sync(0xffffffffu, best, o);
          const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
          if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (lane == 0) cnt[4] = bi;
      }
      __syncthreads();
      const int pv = cnt[4];
      if (pv != k) {
        for (int j = k + tid; j < ms; j += NT) { const cplx t0 = Zs[k * ms + j]; Zs[k * ms + j] = Zs[pv * ms + j]; Zs[pv * ms + j] = t0; }
        __syncthreads();
      }
      const cplx pinv = cinv(Zs[k * ms + k]);
      const int w = ms - (k + 1);
      for (int idx = tid; idx < (nb - k - 1) * w; idx += NT) {
        const int i = k + 1 + idx / w, j = k + 1 + idx % w;
        const cplx f = cmul(Zs[i * ms + k], pinv);
        Zs[i * ms + j] = cfnma(f, Zs[k * ms + j], Zs[i * ms + j]);
      }
    }
    __syncthreads();
    if (warp == 0) {      // back substitution (column oriented) on the right-hand-side column
      for (int k = nb - 1; k >= 0; --k) {
        const cplx xk = cmul(Zs[k * ms + nb], cinv(Zs[k * ms + k]));
        __syncwarp();
        if (lane == 0) Zs[k * ms + nb] = xk;
        for (int i = lane; i < k; i += 32) Zs[i * ms + nb] = cfnma(Zs[i * ms + k], xk, Zs[i * ms + nb]);
        __syncwarp();
      }
    }
    __syncthreads();
    if (tid == 0) {
      // R_r = tau_r / (i bI_r) (tau = tau~ + r0 on the constant harmonic), T_r = sigma_r / (i bII_r); zero on evanescent harmonics
      cplx Rh[3] = {make_double2(0.0, 0.0), make_double2(0.0, 0.0), make_double2(0.0, 0.0)};
      cplx Th[3] = {make_double2(0.0, 0.0), make_double2(0.0, 0.0), make_double2(0.0, 0.0)};
      for (int a = 0; a < nhI; ++a) {
        const int r = pI[a];
        const int slot = (r == 0) ? 0 : (r == 1) ? 1 : (r == z) ? 2 : -1;
        if (slot < 0) continue;
        cplx tau = Zs[(o_tau + a) * ms + nb];
        if (r == z) tau.y += beta[a] + 1.0 / nI;          // + r0, r0 = i (bI_z + 1 / n_I)
        Rh[slot] = make_double2(tau.y / beta[a], -tau.x / beta[a]);
      }
      for (int a = 0; a < nhII; ++a) {
        const int r = pII[a];
        const int slot = (r == 0) ? 0 : (r == 1) ? 1 : (r == z) ? 2 : -1;
        if (slot < 0) continue;
        const cplx sg = Zs[(o_sig + a) * ms + nb];
        Th[slot] = make_double2(sg.y / beta[NHMAX + a], -sg.x / beta[NHMAX + a]);
      }
      const double is2 = 0.7071067811865476;
      const cplx Tc = Th[0], Ts = Th[1], T0 = Th[2];
      const cplx Rc = Rh[0], Rs = Rh[1];
      cplx R0 = Rh[2];
      R0.x -= 1.0;
      const cplx Tm = make_double2((Tc.x - Ts.y) * is2, (Tc.y + Ts.x) * is2);
      const cplx Tp = make_double2((Tc.x + Ts.y) * is2, (Tc.y - Ts.x) * is2);
      const cplx Rm = make_double2((Rc.x - Rs.y) * is2, (Rc.y + Rs.x) * is2);
      const cplx Rp = make_double2((Rc.x + Rs.y) * is2, (Rc.y - Rs.x) * is2);
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
      if (cnt[5] || !isfinite(det[0]) || !isfinite(det[1]) || !isfinite(der[1])) st |= PIRL_ST_NONFINITE_;
      P.eff[sidx] = det[0];
      if (P.de_r3) { P.de_r3[3 * sidx] = der[0]; P.de_r3[3 * sidx + 1] = der[1]; P.de_r3[3 * sidx + 2] = der[2]; }
      if (P.de_t3) { P.de_t3[3 * sidx] = det[0]; P.de_t3[3 * sidx + 1] = det[1]; P.de_t3[3 * sidx + 2] = det[2]; }
      if (P.status) P.status[sidx] = st;
    }
    RK_PHASE(15);
  }
  trc.done(P.trace, 6, P.B);
}
