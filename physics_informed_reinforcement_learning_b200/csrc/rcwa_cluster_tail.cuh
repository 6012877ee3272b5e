// rcwa_cluster_tail.cuh -- layer solve in the mode basis on a 4-CTA thread-block cluster (sm_100a), large orders.
//
// Same REAL bordered formulation as rk_tail2 (rcwa_tail2.cuh; tools/proto_tail2.py is its NumPy statement): modes sorted by
// descending eigenvalue [I insignificant evanescent | S significant evanescent | P propagating]; the real symmetric positive
// definite matrices K_top = R_I + C and K_bot = R_II + C (R = W^T Re(D) W) are eliminated without pivoting on the register
// tile of the cluster (LDL^T on I, symmetric sweep on S, the extra columns of the propagating harmonics / modes carried
// along), leaving a complex border system of about 25 unknowns.  Replaces _base.py:159, transfer_method.py:42-75 and
// solver.py:29-38 of the reference.
// Cluster organisation (see rcwa_cluster.cuh): the tile is split by thread row; each elimination step all-gathers the pivot
// column and broadcasts the pivot row of the extra columns through distributed shared memory as one pipelined round
// (st.async + mbarrier transaction count, look-ahead publish of the next column: CkPipe).
// The parked blocks are written and read by the same thread row, so they stay in the owner's shared memory; the small
// corner blocks are broadcast and every CTA solves the border system redundantly (CTA 0 stores the result).  W is streamed
// from L2 in row chunks, permuted into sorted mode order while it is staged.
#pragma once
#include "rcwa_cluster.cuh"

#define PIRL_CT_XTOL 1e-11                 // modes with exp(-k0 d q) below this do not couple top and bottom of the layer
#define PIRL_ST_TAIL_FALLBACK_ 0x100       // working-status bit: structure left to the generic k_tail

template <class C, int NSMAX_, int NPMAX_, int NHMAX_, int RCH_>
struct CTailSmem {
  static constexpr int FF = C::FF, FE = C::FE, NSMAX = NSMAX_, NPMAX = NPMAX_, NHMAX = NHMAX_, RCH = RCH_;
  static constexpr int NXS = C::TC;                            // extra columns carried by the eliminations (one per thread column)
  static constexpr int SCW = NXS + 2;                          // published pivot row of the extra columns, 1/pivot, pivot
  static constexpr int NBMAX = 3 * NPMAX + 2 * NHMAX;          // border unknowns
  static constexpr int NCT = NPMAX + NHMAX;                    // top corner size
  static constexpr int NCB = NPMAX + NXS;                      // bottom corner size
  static constexpr int NPX = NPMAX + NHMAX;
  static constexpr int WLD = FF + 1;                           // row stride of the staged W chunk
  static constexpr int lamS = 0;
  static constexpr int Cd = lamS + FE;
  static constexpr int Sg = Cd + FE;
  static constexpr int dre = Sg + FE;                          // [2][FE]  Re D_I, Re D_II per harmonic
  static constexpr int beta = dre + 2 * FE;                    // [2][NHMAX] Im D of the propagating harmonics
  static constexpr int qP = beta + pirl_even(2 * NHMAX);       // [NPMAX] |q| of the propagating modes
  static constexpr int XP = qP + pirl_even(NPMAX);             // [NPMAX] complex
  static constexpr int cb = XP + 2 * NPMAX;                    // [2][FE] all-gathered pivot column
  static constexpr int sc = cb + 2 * FE;                       // [2][SCW]
  static constexpr int Yt = sc + pirl_even(2 * SCW);           // [FF][NHMAX]
  static constexpr int dt = Yt + pirl_even(FF * NHMAX);        // [FE]
  static constexpr int Yb = dt + FE;                           // [FF][NXS]
  static constexpr int db = Yb + pirl_even(FF * NXS);          // [FE]
  static constexpr int CT = db + FE;                           // [NCT][NCT]
  static constexpr int CB = CT + pirl_even(NCT * NCT);         // [NCB][NCB]
  static constexpr int parkSS = CB + pirl_even(NCB * NCB);     // [NSMAX][NSMAX]  (entries of this CTA's rows only)
  static constexpr int parkX = parkSS + NSMAX * NSMAX;         // [NSMAX][NPX]
  static constexpr int ints = parkX + pirl_even(NSMAX * NPX);  // pos[FE], ipos[FE], pI[8], pII[8], cnt[16] (ints)
  static constexpr int mbar = pirl_even(ints + FE + 16);       // two mbarriers (CkPipe)
  static constexpr int chunk = mbar + 2;                       // [RCH][WLD] staged rows of W; later the border system
  static constexpr int zsys = 2 * NBMAX * (NBMAX + 1);
  static constexpr int doubles = chunk + pirl_even((RCH * WLD > zsys) ? RCH * WLD : zsys);
};

// t = W^T diag(dre) W on the cluster tile; W rows streamed from L2 in chunks of RCH harmonics, columns permuted to sorted order
template <class C, class S>
__device__ __forceinline__ void ck_product(double (&t)[C::RT][C::CT], const double* __restrict__ Wg, int FFR, const int* pos,
                                           const double* dre, double* chunk, int ti, int tj, int tid) {
  constexpr int TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, NT = C::NT, WLD = S::WLD, RCH = S::RCH;
#pragma unroll
  for (int a = 0; a < RT; ++a)
#pragma unroll
    for (int b = 0; b < CT; ++b) t[a][b] = 0.0;
  for (int r0 = 0; r0 < FFR; r0 += RCH) {
    const int nr = min(RCH, FFR - r0);
    __syncthreads();
    for (int idx = tid; idx < nr * FFR; idx += NT) {
      const int rr = idx / FFR, j = idx - rr * FFR;
      chunk[rr * WLD + pos[j]] = __ldg(&Wg[(size_t)(r0 + rr) * FFR + j]);
    }
    __syncthreads();
#pragma unroll 2
    for (int rr = 0; rr < nr; ++rr) {
      const double d = dre[r0 + rr];
      if (d == 0.0) continue;                      // propagating harmonic (uniform)
      const double* wrow = chunk + rr * WLD;
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
}

// The same product on the FP64 tensor path (mma.sync.aligned.m8n8k4.f64; it wins in g_eigback,
// profiles/r02m_dmma_trial_g_eigback.txt, but LOSES here -- see CK_PRODUCT below; kept as the measured alternative).
// This CTA's rows of W^T D W form RT * ceil(TRL / 8) row tiles of 8 (each tile: eight consecutive thread rows of one
// tile-row block a), the FF columns (FF + 7) / 8 column tiles; a warp owns up to CPW column
// tiles for ALL row tiles, walks the staged harmonics four at a time (A fragment = d_r W[r][row], B fragment = W[r][col], one
// double per lane each) and keeps its accumulators in registers over all chunks.  The result is brought into the cyclic
// register-tile layout through the (then free) chunk buffer, two tile-row blocks at a time.
template <class C, class S>
__device__ __forceinline__ void ck_product_mma(double (&t)[C::RT][C::CT], const double* __restrict__ Wg, int FFR, const int* pos,
                                               const double* dre, double* chunk, int crank, int til, int tj, int tid) {
  constexpr int TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, NT = C::NT, TRL = C::TRL, FF = C::FF, WLD = S::WLD, RCH = S::RCH;
  constexpr int TPA = (TRL + 7) / 8;                 // row tiles per tile-row block
  constexpr int NRT = RT * TPA;                      // row tiles of this CTA
  constexpr int NCT = (FF + 7) / 8;                  // column tiles
  constexpr int NW = NT / 32;
  constexpr int CPW = (NCT + NW - 1) / NW;           // column tiles per warp
  static_assert(2 * TRL * WLD <= RCH * WLD, "the staging of two tile-row blocks must fit the chunk buffer");
  const int lane = tid & 31, warp = tid >> 5, g = lane >> 2, t4 = lane & 3;
  double acc[CPW][NRT][2];
  int arow[NRT];                                     // sorted-mode index of this lane's row in each row tile (-1: padding)
  int bcol[CPW];                                     // column of this lane in each of its column tiles (-1: none)
#pragma unroll
  for (int rt = 0; rt < NRT; ++rt) {
    const int a = rt / TPA, tl = (rt % TPA) * 8 + g;
    arow[rt] = (tl < TRL) ? crank * TRL + tl + TR * a : -1;
  }
#pragma unroll
  for (int c = 0; c < CPW; ++c) {
    const int ct = warp + NW * c, j = ct * 8 + g;
    bcol[c] = (ct < NCT && j < FF) ? j : -1;
#pragma unroll
    for (int rt = 0; rt < NRT; ++rt) { acc[c][rt][0] = 0.0; acc[c][rt][1] = 0.0; }
  }
  for (int r0 = 0; r0 < FFR; r0 += RCH) {
    const int nr = min(RCH, FFR - r0);
    __syncthreads();
    for (int idx = tid; idx < nr * FFR; idx += NT) {
      const int rr = idx / FFR, j = idx - rr * FFR;
      chunk[rr * WLD + pos[j]] = __ldg(&Wg[(size_t)(r0 + rr) * FFR + j]);
    }
    for (int idx = tid; idx < nr * (FF - FFR); idx += NT) {      // padded (dummy) modes sort first: zero columns of W
      const int rr = idx / (FF - FFR), j = idx - rr * (FF - FFR);
      chunk[rr * WLD + j] = 0.0;
    }
    __syncthreads();
    for (int k0 = 0; k0 < nr; k0 += 4) {
      const int r = k0 + t4;
      const bool rok = r < nr;
      const double d = rok ? dre[r0 + r] : 0.0;      // propagating harmonics have d = 0: they drop out
      const double* wrow = chunk + (rok ? r : 0) * WLD;
      double af[NRT], bf[CPW];
#pragma unroll
      for (int rt = 0; rt < NRT; ++rt) af[rt] = (arow[rt] >= 0) ? d * wrow[arow[rt]] : 0.0;
#pragma unroll
      for (int c = 0; c < CPW; ++c) bf[c] = (bcol[c] >= 0 && rok) ? wrow[bcol[c]] : 0.0;
#pragma unroll
      for (int c = 0; c < CPW; ++c) {
        if (warp + NW * c >= NCT) continue;          // uniform per warp
#pragma unroll
        for (int rt = 0; rt < NRT; ++rt)
          asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                       : "+d"(acc[c][rt][0]), "+d"(acc[c][rt][1])
                       : "d"(af[rt]), "d"(bf[c]));
      }
    }
  }
  // accumulators (C[row = g][col = 2 t4 + {0, 1}] of every tile) -> cyclic register tile, two tile-row blocks at a time
#pragma unroll
  for (int h = 0; h < (RT + 1) / 2; ++h) {
    __syncthreads();
#pragma unroll
    for (int c = 0; c < CPW; ++c) {
      const int ct = warp + NW * c;
      if (ct >= NCT) continue;
#pragma unroll
      for (int rt = 0; rt < NRT; ++rt) {
        const int a = rt / TPA, tl = (rt % TPA) * 8 + g;
        if (a / 2 != h || tl >= TRL) continue;
        double* dst = chunk + ((a & 1) * TRL + tl) * WLD + ct * 8 + 2 * t4;
        if (ct * 8 + 2 * t4 < FF) dst[0] = acc[c][rt][0];
        if (ct * 8 + 2 * t4 + 1 < FF) dst[1] = acc[c][rt][1];
      }
    }
    __syncthreads();
#pragma unroll
    for (int a = 0; a < RT; ++a)
      if (a / 2 == h) {
#pragma unroll
        for (int b = 0; b < CT; ++b) t[a][b] = chunk[((a & 1) * TRL + til) * WLD + tj + TC * b];
      }
  }
  __syncthreads();       // the chunk buffer is free again (padded sorted columns are re-zeroed by the caller's next structure)
}

// Elimination steps on pivots [kbeg, kend) of the symmetric cluster tile, carrying the extra columns x0 (column tj), as
// PIPELINED ROUNDS (CkPipe, rcwa_cluster.cuh): per step one all-gather of the pivot column (every CTA), the pivot row of the
// extra columns and 1/d, d (owner CTA), stored with st.async and counted on an mbarrier; the owners of column k + 1 and of
// row k + 1 apply step k to their entries FIRST and send round k + 1 before the bulk of step k is applied.
//   SWEEP = false: LDL^T steps (rows / columns > k);
//   SWEEP = true : symmetric sweep steps inside the window rows / columns >= kbeg; afterwards
//     t[S,S] = -(Schur_S)^-1,  t[S,R] = (Schur_S)^-1 M_SR,  t[R,R] = Schur complement,  x0[S] = (Schur_S)^-1 x_S,  x0[R] = Schur.
// Every CTA keeps the pivot and the pivot row of the extra columns (Ys[k][0..nx), ds[k]) for the corner term -Y^T D^-1 Y.
template <class C, class S, bool SWEEP>
__device__ __forceinline__ void ck_elim_range(double (&t)[C::RT][C::CT], double (&x0)[C::RT], int kbeg, int kend, double* sm,
                                              CkPipe<C::NCTA>& pipe, double* Ys, int ys, double* ds, int nx, bool act, int ti,
                                              int tj, int tid) {
  constexpr int TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, FE = C::FE, SCW = S::SCW, NXS = S::NXS;
  if (kbeg >= kend) return;
  const int a0 = SWEEP ? kbeg / TR : 0, b0 = SWEEP ? kbeg / TC : 0;     // sweep: first tile blocks that intersect the window
  // publish round KK: column KK (tile column KBC) by its owners, 1/d and d by the pivot owner, x0 of row KK by the row owners
#define CK_ELIM_PUBLISH_COL(KBC, KK)                                                                               \
  do {                                                                                                             \
    const int kac_ = (KBC) / C::KA_PER_KB;                                                                         \
    const int cbo_ = S::cb + ((KK) & 1) * FE, sco_ = S::sc + ((KK) & 1) * SCW;                                     \
    _Pragma("unroll") for (int a = 0; a < RT; ++a) {                                                               \
      const int i_ = ti + TR * a;                                                                                  \
      if (SWEEP) { if (a >= a0) pipe.pub(cbo_ + i_, (i_ >= kbeg) ? ((i_ == (KK)) ? t[a][KBC] - 1.0 : t[a][KBC]) : 0.0); } \
      else if (a >= kac_) pipe.pub(cbo_ + i_, (i_ > (KK)) ? t[a][KBC] : 0.0);                                      \
    }                                                                                                              \
    if (ti == (KK) - kac_ * TR) { const double d_ = t[kac_][KBC]; pipe.pub(sco_ + NXS, 1.0 / d_); pipe.pub(sco_ + NXS + 1, d_); } \
  } while (0)
#define CK_ELIM_PUBLISH_ROW(KAR, KK) pipe.pub(S::sc + ((KK) & 1) * SCW + tj, x0[KAR])
  // prologue: round kbeg (every thread of this CTA is past a CTA barrier that follows its reads of the previous round)
  if (act) {
#pragma unroll
    for (int kb = 0; kb < CT; ++kb) {
      const int ka = kb / C::KA_PER_KB;
      if (kbeg >= kb * TC && kbeg < (kb + 1) * TC) {
        if (tj == kbeg - kb * TC) CK_ELIM_PUBLISH_COL(kb, kbeg);
        if (ti == kbeg - ka * TR) CK_ELIM_PUBLISH_ROW(ka, kbeg);
      }
    }
  }
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
      pipe.wait(8u * (uint32_t)((SWEEP ? (RT - a0) : (RT - ka)) * TR + 2 + TC), tid);
      const double* cb = sm + S::cb + (k & 1) * FE;
      const double* sc = sm + S::sc + (k & 1) * SCW;
      const double p = sc[NXS];
      if (tid < nx) Ys[k * ys + tid] = sc[tid];
      if (tid == 32) ds[k] = sc[NXS + 1];
      double u[RT], v[CT];
#pragma unroll
      for (int a = 0; a < RT; ++a) if (a >= (SWEEP ? a0 : ka)) u[a] = -cb[ti + TR * a] * p;
#pragma unroll
      for (int b = 0; b < CT; ++b) if (b >= (SWEEP ? b0 : kb)) v[b] = cb[tj + TC * b];
      const double y0 = sc[tj];
      __syncthreads();       // round k consumed by every thread of this CTA: the next round may be sent
#pragma unroll
      for (int a = 0; a < RT; ++a) if (a >= (SWEEP ? a0 : ka)) x0[a] = fma(u[a], y0, x0[a]);
      if (k + 1 < kend && act) {       // look-ahead: column k + 1 and the extra-column entries of row k + 1 first
        const int kbn = (kj + 1 < TC) ? kb : ((kb + 1 < CT) ? kb + 1 : kb);     // constant per branch after unrolling
        const int kan = kbn / C::KA_PER_KB;
        if (kj + 1 < TC) {
          if (tj == kj + 1) {
#pragma unroll
            for (int a = 0; a < RT; ++a) if (a >= (SWEEP ? a0 : ka)) t[a][kb] = fma(u[a], v[kb], t[a][kb]);
            v[kb] = 0.0;
            CK_ELIM_PUBLISH_COL(kb, k + 1);
          }
          if (ti == k + 1 - ka * TR) CK_ELIM_PUBLISH_ROW(ka, k + 1);
        } else if (kb + 1 < CT) {
          if (tj == 0) {
#pragma unroll
            for (int a = 0; a < RT; ++a)
              if (a >= (SWEEP ? a0 : ka)) t[a][(kb + 1 < CT) ? kb + 1 : kb] = fma(u[a], v[(kb + 1 < CT) ? kb + 1 : kb], t[a][(kb + 1 < CT) ? kb + 1 : kb]);
            v[(kb + 1 < CT) ? kb + 1 : kb] = 0.0;
            CK_ELIM_PUBLISH_COL(((kb + 1 < CT) ? kb + 1 : kb), k + 1);
          }
          if (ti == k + 1 - kan * TR) CK_ELIM_PUBLISH_ROW((((kb + 1 < CT) ? kb + 1 : kb) / C::KA_PER_KB), k + 1);
        }
      }
#pragma unroll
      for (int b = 0; b < CT; ++b)
        if (b >= (SWEEP ? b0 : kb)) {
#pragma unroll
          for (int a = 0; a < RT; ++a) if (a >= (SWEEP ? a0 : ka)) t[a][b] = fma(u[a], v[b], t[a][b]);
        }
      if (SWEEP && ti == kti && tj == kj) t[ka][kb] = -p;
    }
    if (done) break;
  }
#undef CK_ELIM_PUBLISH_COL
#undef CK_ELIM_PUBLISH_ROW
}

// measured at order 100 (profiles/r02u_cluster_check_dmma_products.txt): the DMMA product costs 208k / 145k clocks per
// structure against 131k / 113k for the register-tiled DFMA product (4-way bank conflicts on the fragment loads at the odd
// row stride, 2.75 warps per scheduler) -> DFMA stays the default; -DPIRL_CK_PRODUCT_MMA builds the tensor variant
#ifdef PIRL_CK_PRODUCT_MMA
#define CK_PRODUCT(t_, W_, F_, p_, d_, c_) ck_product_mma<C, S>(t_, W_, F_, p_, d_, c_, crank, til, tj, tid)
#else
#define CK_PRODUCT(t_, W_, F_, p_, d_, c_) ck_product<C, S>(t_, W_, F_, p_, d_, c_, ti, tj, tid)
#endif

template <class C, int NSMAX, int NPMAX, int NHMAX, int RCH>
__global__ void __launch_bounds__(C::NT, 1) ck_tail2(KParams P) {
  constexpr int FF = C::FF, NCTA = C::NCTA, TRL = C::TRL, TR = C::TR, TC = C::TC, RT = C::RT, CT = C::CT, NT = C::NT;
  using S = CTailSmem<C, NSMAX, NPMAX, NHMAX, RCH>;
  constexpr int FE = C::FE, NXS = S::NXS, NCT = S::NCT, NCB = S::NCB, NPX = S::NPX;
  extern __shared__ __align__(16) double sm[];
  const int FFR = P.ff, N = P.N, z = 2 * N, npad = FF - FFR;
  double* lamS = sm + S::lamS;
  double* Cd = sm + S::Cd;
  double* Sg = sm + S::Sg;
  double* dre = sm + S::dre;
  double* beta = sm + S::beta;
  double* qP = sm + S::qP;
  cplx* XP = reinterpret_cast<cplx*>(sm + S::XP);
  double* Yt = sm + S::Yt;
  double* dt = sm + S::dt;
  double* Yb = sm + S::Yb;
  double* db = sm + S::db;
  double* CTm = sm + S::CT;
  double* CBm = sm + S::CB;
  double* parkSS = sm + S::parkSS;
  double* parkX = sm + S::parkX;
  int* pos = reinterpret_cast<int*>(sm + S::ints);
  int* ipos = pos + FE;
  int* pI = ipos + FE;
  int* pII = pI + 8;
  int* cnt = pII + 8;                     // [0] n_i, [1] n_e, [2] n_hI, [3] n_hII, [4] pivot row, [5] bad, [8] next item
  double* chunk = sm + S::chunk;
  cplx* Zs = reinterpret_cast<cplx*>(sm + S::chunk);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int crank = (int)ck_cluster_ctarank();
  const int cid = blockIdx.x / NCTA, nclusters = gridDim.x / NCTA;
  const bool act = tid < C::NTHR;
  const int til = act ? tid / TC : TRL - 1;
  const int tj = act ? tid % TC : TC - 1;
  const int ti = crank * TRL + til;
  CkPipe<NCTA> pipe;
  pipe.init(sm, S::mbar, tid);
  auto bcast = [&](int off, double v) { pipe.bcast(off, v); };
  const uint32_t ints_b = 8u * (uint32_t)S::ints + 4u * (uint32_t)(2 * FE + 16);     // byte offset of cnt[]
  ck_cluster_sync();

  for (int sidx = cid; sidx < P.B;) {
    long long pc_ = clock64();
    bool skip = false;
    const double* pr = P.params + (size_t)sidx * P.pstride;
    const double wl = pr[0], thick = pr[2];
    const double sa = sin(pr[1] * (3.14159265358979323846 / 180.0));
    const double k0d = 2.0 * 3.14159265358979323846 / wl * thick;
    const double nI = P.n_inc, nII = P.n_exit;
    const double* lamg = P.lam + (size_t)sidx * FFR;
    const double* Wg = P.Wbuf + (size_t)sidx * FFR * FFR;
    for (int t = tid; t < FF; t += NT) lamS[t] = (t < FFR) ? lamg[t] : 1e12;   // padded modes: q = 1e6, X = 0
    if (tid < 8) cnt[tid] = 0;
    for (int idx = tid; idx < RCH * S::WLD; idx += NT) chunk[idx] = 0.0;      // padded sorted columns of the staged W stay zero
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
      ipos[rk] = tid;
      myrank = rk;
      double cdv = 0.0, sgv = 0.0;
      if (l > 0.0) {
        const double qq = sqrt(l), x = k0d * qq, X = exp(-x);
        if (X < PIRL_CT_XTOL) { cdv = qq; atomicAdd(&cnt[0], 1); }
        else { const double om = -expm1(-2.0 * x); cdv = qq * (1.0 + X * X) / om; sgv = 2.0 * qq * X / om; }
        atomicAdd(&cnt[1], 1);
      }
      Cd[rk] = cdv; Sg[rk] = sgv;
      // harmonic tid (original order: 2(p-1) cos_p, 2(p-1)+1 sin_p, 2N const)
      const bool padh = tid >= FFR;             // padded harmonic: no admittance, not a propagating order
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
    if (ns > NSMAX || np > NPMAX || nhI > NHMAX || nhII > NHMAX || nxb > NXS) skip = true;   // uniform over the cluster
    if (skip) {
      if (tid == 0 && crank == 0) atomicOr(&P.wstatus[sidx], PIRL_ST_TAIL_FALLBACK_);
    } else {
      if (tid < FF && !(myl > 0.0)) {
        const int j = myrank - ne;
        const double s = sqrt(-myl);
        double sn, cn;
        sincos(k0d * s, &sn, &cn);
        qP[j] = s; XP[j] = make_double2(cn, -sn);
      }
      __syncthreads();
      CK_PHASE(8);
      double t[RT][CT], x0[RT];
      // sorted mode i -> column of W in global memory (-1: padded dummy mode)
      auto wglob = [&](int harm, int i) -> double { return (i >= npad) ? __ldg(&Wg[(size_t)harm * FFR + ipos[i]]) : 0.0; };
      // =========================== top: R_I + C, eliminate I (LDL^T), sweep S ===========================
      CK_PRODUCT(t, Wg, FFR, pos, dre, chunk);
#pragma unroll
      for (int a = 0; a < RT; ++a) {
        const int i = ti + TR * a;
#pragma unroll
        for (int b = 0; b < CT; ++b) if (i == tj + TC * b) t[a][b] += Cd[i];
        x0[a] = (tj < nhI) ? wglob(pI[tj], i) : 0.0;
      }
      CK_PHASE(9);
      ck_elim_range<C, S, false>(t, x0, 0, ni, sm, pipe, Yt, NHMAX, dt, nhI, act, ti, tj, tid);
      CK_PHASE(10);
      ck_elim_range<C, S, true>(t, x0, ni, ne, sm, pipe, Yt, NHMAX, dt, nhI, act, ti, tj, tid);
      CK_PHASE(11);
      // ---- park (this CTA's rows only): Sg (Schur_S)^-1 Sg, Sg G_P, Sg G_w; top corner blocks -> every CTA ----
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
              } else if (j >= ne) bcast(S::CT + (i - ne) * NCT + (j - ne), t[a][b]);
            }
            if (tj < nhI) {
              if (i < ne) parkX[(i - ni) * NPX + np + tj] = sgi * x0[a];
              else { bcast(S::CT + (i - ne) * NCT + np + tj, x0[a]); bcast(S::CT + (np + tj) * NCT + (i - ne), x0[a]); }
            }
          }
        }
      }
      ck_cluster_sync();    // every CTA has left the top sweep before the bottom eliminations reuse the published buffers
      // =========================== bottom: R_II + C, eliminate I, fold the top in, eliminate S ===========================
      CK_PRODUCT(t, Wg, FFR, pos, dre + FE, chunk);
#pragma unroll
      for (int a = 0; a < RT; ++a) {
        const int i = ti + TR * a;
#pragma unroll
        for (int b = 0; b < CT; ++b) if (i == tj + TC * b) t[a][b] += Cd[i];
        x0[a] = (tj < nhII) ? wglob(pII[tj], i) : 0.0;
      }
      CK_PHASE(12);
      ck_elim_range<C, S, false>(t, x0, 0, ni, sm, pipe, Yb, NXS, db, nxb, act, ti, tj, tid);
      __syncthreads();      // parked blocks of this CTA are visible
      if (act) {
#pragma unroll
        for (int a = 0; a < RT; ++a) {
          const int i = ti + TR * a;
          if (i >= ni && i < ne) {
#pragma unroll
            for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; if (j >= ni && j < ne) t[a][b] -= parkSS[(i - ni) * NSMAX + (j - ni)]; }
            const int c0 = tj - nhII;
            if (c0 >= 0 && c0 < np + nhI) x0[a] += parkX[(i - ni) * NPX + c0];
          }
        }
      }
      ck_elim_range<C, S, false>(t, x0, ni, ne, sm, pipe, Yb, NXS, db, nxb, act, ti, tj, tid);
      CK_PHASE(13);
      // ---- bottom corner blocks -> every CTA ----
      if (act) {
#pragma unroll
        for (int a = 0; a < RT; ++a) {
          const int i = ti + TR * a;
          if (i >= ne) {
#pragma unroll
            for (int b = 0; b < CT; ++b) { const int j = tj + TC * b; if (j >= ne) bcast(S::CB + (i - ne) * NCB + (j - ne), t[a][b]); }
            bcast(S::CB + (i - ne) * NCB + np + tj, x0[a]);
            bcast(S::CB + (np + tj) * NCB + (i - ne), x0[a]);
          }
        }
      }
      ck_cluster_sync();
      // corner terms -Y^T D^-1 Y of the eliminated pivots (every CTA holds all pivot rows)
      int bad = 0;
      double* rdt = sm + S::cb;            // reciprocals of the pivots (the published-column buffers are free now)
      double* rdb = sm + S::cb + FE;
      for (int k = tid; k < ne; k += NT) {
        if (!(dt[k] > 0.0) || !(db[k] > 0.0)) bad = 1;
        rdt[k] = 1.0 / dt[k]; rdb[k] = 1.0 / db[k];
      }
      __syncthreads();
      for (int idx = tid; idx < nhI * nhI + nxb * nxb; idx += NT) {
        double s0 = 0.0, s1 = 0.0;          // two chains: the sums run over ~200 pivots
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
            s0 = fma(Yb[k * NXS + a] * rdb[k], Yb[k * NXS + b], s0);
            s1 = fma(Yb[(k + 1) * NXS + a] * rdb[k + 1], Yb[(k + 1) * NXS + b], s1);
          }
          if (k < ne) s0 = fma(Yb[k * NXS + a] * rdb[k], Yb[k * NXS + b], s0);
          CBm[(np + a) * NCB + np + b] = -(s0 + s1);
        }
      }
      if (bad) cnt[5] = 1;       // a non-positive pivot: K was not positive definite
      __syncthreads();
      CK_PHASE(14);
      // =========================== border system [h_t,P | tau | hb_P | sigma | c_P] (every CTA, redundantly) ===========================
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
          const int rbi = (tr == 2) ? ir : (tr == 3) ? np + ir : (tr == 0) ? np + nhII + ir : (tr == 1) ? np + nhII + np + ir : -1;
          const int cbi = (tc == 2) ? ic : (tc == 3) ? np + ic : (tc == 0) ? np + nhII + ic : (tc == 1) ? np + nhII + np + ic : -1;
          if (rbi >= 0 && cbi >= 0) v.x += CBm[rbi * NCB + cbi];
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
      if (tid == 0 && crank == 0) {
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
      CK_PHASE(15);
    }
    // ---- next structure (uniform over the cluster) ----
    ck_cluster_sync();
    if (crank == 0 && tid == 0) {
      const int nxt = P.queue ? nclusters + atomicAdd(P.queue, 1) : sidx + nclusters;
#pragma unroll
      for (int r = 0; r < NCTA; ++r) ck_st_cluster_i32(pipe.rb[r] + ints_b + 4u * 8u, nxt);
    }
    ck_cluster_sync();
    sidx = cnt[8];
  }
}
