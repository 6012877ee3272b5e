// rcwa_smm.cuh -- scattering-matrix (algo = 'SMM') form of the layer solve: S_ref (*) S_layer (*) S_trn with Redheffer star
// products (sm_100a).  The reference's alternative to the enhanced transmittance matrix method:
//   scattering_1d_1/2/3   meent/on_numpy/scattering_method.py:12-73
//   A_B_matrices(_half_space), S_layer, S_RT, RedhefferStar, homogeneous_1D   meent/on_numpy/smm_util.py:15-39, 42-67, 70-85,
//                                                                             289-318, 136-182
// It runs behind the same front kernels (modes W, eigenvalues lambda from rk_front / ck_front + k_ql + k_apply) in place of
// the ETM tail, statement for statement in the operator sense: the same blocks A, B, S11..S22 are formed and combined with
// the same products and inverses (general complex matrices, Gauss-Jordan with partial pivoting), in the REAL harmonic basis
// {sqrt2 cos, sqrt2 sin, 1} -- a unitary change of basis that commutes with every diagonal medium matrix (they depend on
// kx^2 only), so every block is the reference's block conjugated by that unitary and the efficiencies are identical.
// One CTA per structure, blocks in an L2-resident per-CTA scratch.  This is the reference's slower variant (1.6x-2.9x its
// TMM) and is not tuned: it exists so that algo = 'SMM' is a real operator, and as the stable form for
// multilayer stacks: P.n_layers > 1 takes one more star product per layer (entries b * n_layers + l of the per-layer
// workspaces; list order, or reversed list order when P.layer_rev is set -- the reference's 'SMM' loop walks the list
// backwards, _base.py:141).  P.pol = 0 is the TE case (_base.py:142-148): orthonormal modes, V = W Q, no n^2 factors.
#pragma once
#include "rcwa_common.cuh"

constexpr int SMM_NT = 256;
constexpr int SMM_NMAT = 17;          // complex ff x ff blocks of per-CTA scratch

__host__ __device__ inline size_t smm_scratch_doubles(int ff) { return (size_t)SMM_NMAT * 2 * ff * ff + 2 * (size_t)ff * ff; }
// shared: cre, cim [2N+1], per-harmonic complex diagonals vg, vr, vt, kzr, kzt, q, X, the four diagonal S-blocks of a half
// space (d11, d12, d21, d22), the pivot column and the two pivot rows of the inverse [ff each], pattern bits, pivot slots,
// row permutation, then the matrix region: GEMM panels (2 x ff x SMM_KC complex) and, when it fits, one ff x ff complex
// matrix for the in-place inverse (the two share the region)
constexpr int SMM_NVEC = 14;
constexpr int SMM_KC = 16;            // GEMM panel depth
constexpr int SMM_MAXB = 4;           // 2 x 2 output blocks per thread and row panel
__host__ __device__ inline int smm_vec_doubles(int ff, int N, int n_cells) {
  return pirl_even(2 * (2 * N + 1)) + 2 * SMM_NVEC * ff + pirl_even((n_cells + 7) / 8) + 32 + pirl_even(ff / 2 + 1);
}
__host__ __device__ inline int smm_panel_doubles(int ff) { return 2 * (2 * ff * SMM_KC); }
// mats = 1: the region also holds the ff x ff complex matrix of the in-place inverse
__host__ __device__ inline int smm_smem_doubles(int ff, int N, int n_cells, int mats) {
  const int region = (mats && 2 * ff * ff > smm_panel_doubles(ff)) ? 2 * ff * ff : smm_panel_doubles(ff);
  return smm_vec_doubles(ff, N, n_cells) + region;
}

// C = A B  (n x n complex, row major; C must not alias A or B).  Each thread owns up to SMM_MAXB 2 x 2 output blocks of a row
// panel; A[panel rows, k0 : k0 + KC] and B[k0 : k0 + KC, :] are staged in shared memory (`tile`), so A is read from L2 once
// and B once per row panel instead of once per output block.
__device__ __noinline__ void smm_gemm(cplx* __restrict__ C, const cplx* __restrict__ A, const cplx* __restrict__ B, int n, int tid,
                                         cplx* __restrict__ tile) {
  const int nb = (n + 1) / 2;                                   // 2 x 2 blocks per row / column
  int prow_b = (SMM_NT * SMM_MAXB) / nb;                        // block rows per panel
  if (prow_b > nb) prow_b = nb;
  const int prow = 2 * prow_b;                                  // matrix rows per panel
  cplx* As = tile;                                              // [prow][KC]
  cplx* Bs = tile + (size_t)((prow < n) ? prow : n) * SMM_KC;   // [KC][n]
  for (int r0 = 0; r0 < n; r0 += prow) {
    const int rows = (n - r0 < prow) ? n - r0 : prow;
    const int nblk = ((rows + 1) / 2) * nb;
    cplx acc[SMM_MAXB][4];
#pragma unroll
    for (int q = 0; q < SMM_MAXB; ++q) { acc[q][0] = acc[q][1] = acc[q][2] = acc[q][3] = make_double2(0, 0); }
    for (int k0 = 0; k0 < n; k0 += SMM_KC) {
      const int kc = (n - k0 < SMM_KC) ? n - k0 : SMM_KC;
      __syncthreads();                                          // previous chunk consumed
      for (int idx = tid; idx < rows * SMM_KC; idx += SMM_NT) {
        const int i = idx / SMM_KC, kk = idx - i * SMM_KC;
        As[idx] = (kk < kc) ? A[(size_t)(r0 + i) * n + k0 + kk] : make_double2(0, 0);
      }
      for (int idx = tid; idx < SMM_KC * n; idx += SMM_NT) {
        const int kk = idx / n, j = idx - kk * n;
        Bs[idx] = (kk < kc) ? B[(size_t)(k0 + kk) * n + j] : make_double2(0, 0);
      }
      __syncthreads();
#pragma unroll
      for (int q = 0; q < SMM_MAXB; ++q) {
        const int blk = tid + q * SMM_NT;
        if (blk < nblk) {
          const int bi = blk / nb, bj = blk - bi * nb;
          const int i0 = 2 * bi, j0 = 2 * bj;
          const cplx* a0 = As + (size_t)i0 * SMM_KC;
          const cplx* a1 = As + (size_t)((i0 + 1 < rows) ? i0 + 1 : i0) * SMM_KC;
          const int j1 = (j0 + 1 < n) ? j0 + 1 : j0;
#pragma unroll
          for (int kk = 0; kk < SMM_KC; ++kk) {
            const cplx x0 = a0[kk], x1 = a1[kk];
            const cplx y0 = Bs[kk * n + j0], y1 = Bs[kk * n + j1];
            acc[q][0] = cfma(x0, y0, acc[q][0]); acc[q][1] = cfma(x0, y1, acc[q][1]);
            acc[q][2] = cfma(x1, y0, acc[q][2]); acc[q][3] = cfma(x1, y1, acc[q][3]);
          }
        }
      }
    }
#pragma unroll
    for (int q = 0; q < SMM_MAXB; ++q) {
      const int blk = tid + q * SMM_NT;
      if (blk < nblk) {
        const int bi = blk / nb, bj = blk - bi * nb;
        const int i0 = r0 + 2 * bi, j0 = 2 * bj;
        const bool i1 = i0 + 1 < r0 + rows, j1 = j0 + 1 < n;
        C[(size_t)i0 * n + j0] = acc[q][0];
        if (j1) C[(size_t)i0 * n + j0 + 1] = acc[q][1];
        if (i1) { C[(size_t)(i0 + 1) * n + j0] = acc[q][2]; if (j1) C[(size_t)(i0 + 1) * n + j0 + 1] = acc[q][3]; }
      }
    }
  }
  __syncthreads();
}

// C = alpha A + beta B (+ gamma I)
__device__ __forceinline__ void smm_axpby(cplx* C, double alpha, const cplx* A, double beta, const cplx* B, double gamma, int n, int tid) {
  for (int idx = tid; idx < n * n; idx += SMM_NT) {
    cplx v = make_double2(0, 0);
    if (A) { v.x += alpha * A[idx].x; v.y += alpha * A[idx].y; }
    if (B) { v.x += beta * B[idx].x; v.y += beta * B[idx].y; }
    if (gamma != 0.0 && idx / n == idx % n) v.x += gamma;
    C[idx] = v;
  }
  __syncthreads();
}

// C = diag(dl) A diag(dr)  (either diagonal may be null)
__device__ __forceinline__ void smm_scale(cplx* C, const cplx* dl, const cplx* A, const cplx* dr, int n, int tid) {
  for (int idx = tid; idx < n * n; idx += SMM_NT) {
    const int i = idx / n, j = idx - i * n;
    cplx v = A[idx];
    if (dl) v = cmul(dl[i], v);
    if (dr) v = cmul(v, dr[j]);
    C[idx] = v;
  }
  __syncthreads();
}

// pivot of column k among rows >= k (largest |re| + |im|, lowest row on ties) -> *slot (-1: the column is zero); warp 0 only
__device__ __forceinline__ void smm_pivot_search(const cplx* __restrict__ A, int n, int k, int lane, int* slot) {
  double best = -1.0; int bi = k;
  for (int i = k + lane; i < n; i += 32) {
    const cplx v = A[(size_t)i * n + k];
    const double mg = fabs(v.x) + fabs(v.y);
    if (mg > best) { best = mg; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  if (lane == 0) *slot = (best > 0.0) ? bi : -1;
}

// Ainv = A^-1 by Gauss-Jordan with partial pivoting on [A | I]; A is destroyed.  Returns (to every thread) whether a zero
// pivot was met.  `slot`: one shared int, `fcol`: n shared complex numbers.
// One thread owns one COLUMN of [A | I] through a step (row swap, pivot-row scaling and elimination touch that column only), so
// a step needs two barriers: the multipliers (column k) are staged in `fcol` first, and the owner of column k + 1 finds the
// next pivot while it eliminates.  Rows are processed eight at a time, loads before stores, to keep eight L2 requests in flight.
__device__ __noinline__ bool smm_inverse(cplx* __restrict__ Ainv, cplx* __restrict__ A, int n, int tid, int* slot,
                                            cplx* __restrict__ fcol) {
  bool singular = false;
  for (int idx = tid; idx < n * n; idx += SMM_NT) Ainv[idx] = make_double2((idx / n == idx % n) ? 1.0 : 0.0, 0.0);
  if (tid < 32) smm_pivot_search(A, n, 0, tid, slot);
  __syncthreads();
  for (int k = 0; k < n; ++k) {
    const int pv = *slot;
    if (pv >= 0) for (int i = tid; i < n; i += SMM_NT) fcol[i] = A[(size_t)i * n + k];
    __syncthreads();
    if (pv < 0) {                                       // zero column: flag it and look at the next one
      singular = true;
      if (tid < 32 && k + 1 < n) smm_pivot_search(A, n, k + 1, tid, slot);
      __syncthreads();
      continue;
    }
    const cplx pinv = cinv(fcol[pv]);
    const cplx fswap = fcol[k];                         // multiplier of the row that moves from k to pv
    for (int c = tid; c < 2 * n; c += SMM_NT) {
      const bool inA = c < n;
      if (inA && c <= k) continue;                      // finished columns of A are never read again
      cplx* col = inA ? A + c : Ainv + (c - n);
      const cplx a = col[(size_t)k * n];
      cplx pk;
      if (pv != k) { pk = cmul(col[(size_t)pv * n], pinv); col[(size_t)pv * n] = a; }
      else pk = cmul(a, pinv);
      col[(size_t)k * n] = pk;
      const bool track = inA && c == k + 1;             // this thread finds the pivot of the next step
      double best = -1.0; int bi = k + 1;
      for (int i0 = 0; i0 < n; i0 += 8) {
        cplx v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { const int i = i0 + u; if (i < n && i != k) v[u] = col[(size_t)i * n]; }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int i = i0 + u;
          if (i < n && i != k) {
            const cplx r = cfnma((i == pv) ? fswap : fcol[i], pk, v[u]);
            col[(size_t)i * n] = r;
            if (track && i > k) { const double mg = fabs(r.x) + fabs(r.y); if (mg > best) { best = mg; bi = i; } }
          }
        }
      }
      if (track) *slot = (best > 0.0) ? bi : -1;
    }
    __syncthreads();
  }
  return singular;
}

// Shared-memory work area of the inverse: multipliers, the two pivot rows, row permutation, pivot candidates
struct SmmInvSm {
  cplx* fcol; cplx* rowk; cplx* rowp; int* piv; double* candm; int* candi; int* slot;
  cplx* Ms;            // ff x ff complex matrix in shared memory, or null (matrix stays in the L2-resident scratch)
};

// In-place Gauss-Jordan inverse with partial pivoting of the n x n complex matrix Ms in SHARED memory; the result (columns
// permuted back) is written to the global Ainv.  Thread (g, c) owns rows i = g, g + G, ... of column c: every element is
// read and written by one thread per step; the multipliers (column k) and the two rows that swap are staged first, and the
// threads of column k + 1 leave pivot candidates for the next step.  Two barriers per step.
__device__ __noinline__ bool smm_inverse_smem(cplx* __restrict__ Ainv, int n, int tid, const SmmInvSm& W) {
  cplx* Ms = W.Ms;
  int G = SMM_NT / n; if (G > 8) G = 8; if (G < 1) G = 1;
  const int g = tid / n, c = tid - g * n;
  const bool active = g < G;                 // (n <= SMM_NT is guaranteed by the plan when Ms is present)
  bool singular = false;
  if (tid < G) {                             // pivot candidates of column 0
    double best = -1.0; int bi = 0;
    for (int i = tid; i < n; i += G) { const cplx v = Ms[(size_t)i * n]; const double mg = fabs(v.x) + fabs(v.y); if (mg > best) { best = mg; bi = i; } }
    W.candm[tid] = best; W.candi[tid] = bi;
  }
  __syncthreads();
  for (int k = 0; k < n; ++k) {
    double best = -1.0; int pv = k;
    for (int q = 0; q < G; ++q) { const double m = W.candm[q]; const int i = W.candi[q]; if (m > best || (m == best && i < pv)) { best = m; pv = i; } }
    if (!(best > 0.0)) pv = -1;
    if (pv >= 0) {
      for (int i = tid; i < n; i += SMM_NT) { W.fcol[i] = Ms[(size_t)i * n + k]; W.rowk[i] = Ms[(size_t)k * n + i]; W.rowp[i] = Ms[(size_t)pv * n + i]; }
    }
    if (tid == 0) W.piv[k] = (pv >= 0) ? pv : k;
    __syncthreads();
    if (pv < 0) {
      singular = true;
      if (tid < G) {
        double b2 = -1.0; int bi = k + 1;
        if (k + 1 < n) for (int i = k + 1 + tid; i < n; i += G) { const cplx v = Ms[(size_t)i * n + k + 1]; const double mg = fabs(v.x) + fabs(v.y); if (mg > b2) { b2 = mg; bi = i; } }
        W.candm[tid] = b2; W.candi[tid] = bi;
      }
      __syncthreads();
      continue;
    }
    if (active) {
      const cplx pinv = cinv(W.fcol[pv]);
      const cplx fswap = W.fcol[k];
      const bool special = c == k;                     // column k of the in-place form: starts from the unit vector e_k
      const cplx pk = special ? pinv : cmul(W.rowp[c], pinv);
      const cplx a = W.rowk[c];                        // old row k: lands in row pv
      const bool track = c == k + 1;
      double tb = -1.0; int ti = k + 1;
      for (int i = g; i < n; i += G) {
        if (i == k) { Ms[(size_t)k * n + c] = pk; continue; }
        cplx v = (i == pv) ? a : Ms[(size_t)i * n + c];
        if (special) v = make_double2(0.0, 0.0);
        const cplx r = cfnma((i == pv) ? fswap : W.fcol[i], pk, v);
        Ms[(size_t)i * n + c] = r;
        if (track && i > k) { const double mg = fabs(r.x) + fabs(r.y); if (mg > tb) { tb = mg; ti = i; } }
      }
      if (track) { W.candm[g] = tb; W.candi[g] = ti; }
    }
    __syncthreads();
  }
  // A^-1 = (P A)^-1 P: undo the row swaps as column swaps in reverse order, folded into one column map
  int* src = reinterpret_cast<int*>(W.rowk);          // rowk is free now: n ints fit in n complex numbers
  if (tid == 0) {
    for (int j = 0; j < n; ++j) src[j] = j;
    for (int k = n - 1; k >= 0; --k) { const int p2 = W.piv[k]; if (p2 != k) { const int t0 = src[k]; src[k] = src[p2]; src[p2] = t0; } }
  }
  __syncthreads();
  for (int idx = tid; idx < n * n; idx += SMM_NT) {
    const int i = idx / n, j = idx - i * n;
    Ainv[idx] = Ms[(size_t)i * n + src[j]];
  }
  __syncthreads();
  return singular;
}

// Ainv = A^-1 (A is destroyed): in shared memory when the plan found room for the matrix, else in the L2-resident scratch
__device__ __forceinline__ bool smm_invert(cplx* __restrict__ Ainv, cplx* __restrict__ A, int n, int tid, const SmmInvSm& W) {
  if (W.Ms == nullptr) return smm_inverse(Ainv, A, n, tid, W.slot, W.fcol);
  for (int idx = tid; idx < n * n; idx += SMM_NT) W.Ms[idx] = A[idx];
  __syncthreads();
  return smm_inverse_smem(Ainv, n, tid, W);
}

// C = dl (.) A (.) dr (+ add) (+ diag(dg)): row scaling by dl, column scaling by dr (either may be null), optional addend and
// diagonal
__device__ __forceinline__ void smm_scale_add(cplx* C, const cplx* dl, const cplx* A, const cplx* dr, const cplx* add, const cplx* dg,
                                              int n, int tid) {
  for (int idx = tid; idx < n * n; idx += SMM_NT) {
    const int i = idx / n, j = idx - i * n;
    cplx v = A[idx];
    if (dl) v = cmul(dl[i], v);
    if (dr) v = cmul(v, dr[j]);
    if (add) v = cadd(v, add[idx]);
    if (dg && i == j) v = cadd(v, dg[i]);
    C[idx] = v;
  }
  __syncthreads();
}

// RedhefferStar (smm_util.py:289-318): OUT = SA (*) SB.  Blocks are {S11, S12, S21, S22}; T0..T3 are temporaries.
// Returns whether an inverse met a zero pivot.  The reference inverts both I - SB11 SA22 (D) and I - SA22 SB11 (F); here
// F = I + SA22 D SB11 (push-through identity), one inverse per star product.
__device__ __forceinline__ bool smm_redheffer(cplx* const (&O)[4], cplx* const (&SA)[4], cplx* const (&SB)[4], cplx* T0, cplx* T1,
                                              cplx* T2, cplx* T3, int n, int tid, const SmmInvSm& W, cplx* tile) {
  bool sing = false;
  smm_gemm(T0, SB[0], SA[3], n, tid, tile);
  smm_axpby(T0, -1.0, T0, 0.0, nullptr, 1.0, n, tid);
  sing |= smm_invert(T1, T0, n, tid, W);             // T1 = D = inv(I - SB11 SA22)
  smm_gemm(T2, T1, SB[0], n, tid, tile);                             // T2 = D SB11
  smm_gemm(T3, T1, SB[1], n, tid, tile);                             // T3 = D SB12
  // S12 = SA12 D SB12 ; S11 = SA11 + SA12 D SB11 SA21
  smm_gemm(O[1], SA[1], T3, n, tid, tile);
  smm_gemm(T0, SA[1], T2, n, tid, tile);
  smm_gemm(T1, T0, SA[2], n, tid, tile);
  smm_axpby(O[0], 1.0, SA[0], 1.0, T1, 0.0, n, tid);
  // F = I + SA22 D SB11 ; S21 = SB21 F SA21 ; S22 = SB22 + SB21 F SA22 SB12
  smm_gemm(T0, SA[3], T2, n, tid, tile);
  smm_axpby(T0, 1.0, T0, 0.0, nullptr, 1.0, n, tid);           // T0 = F
  smm_gemm(T1, SB[2], T0, n, tid, tile);                             // T1 = SB21 F
  smm_gemm(O[2], T1, SA[2], n, tid, tile);
  smm_gemm(T3, T1, SA[3], n, tid, tile);
  smm_gemm(T0, T3, SB[1], n, tid, tile);
  smm_axpby(O[3], 1.0, SB[3], 1.0, T0, 0.0, n, tid);
  return sing;
}

// OUT = SA (*) SB where every block of SA is DIAGONAL (a half space, S_RT of smm_util.py:70-85): a11, a12, a21, a22 are its
// diagonals.  Products with a diagonal block are scalings: four GEMMs and one inverse.
__device__ __forceinline__ bool smm_redheffer_diag_left(cplx* const (&O)[4], const cplx* a11, const cplx* a12, const cplx* a21,
                                                        const cplx* a22, cplx* const (&SB)[4], cplx* T0, cplx* T1, cplx* T2, cplx* T3,
                                                        int n, int tid, const SmmInvSm& W, cplx* tile) {
  for (int idx = tid; idx < n * n; idx += SMM_NT) {           // T0 = I - SB11 a22
    const int i = idx / n, j = idx - i * n;
    const cplx v = cmul(SB[0][idx], a22[j]);
    T0[idx] = make_double2(((i == j) ? 1.0 : 0.0) - v.x, -v.y);
  }
  __syncthreads();
  const bool sing = smm_invert(T1, T0, n, tid, W);  // T1 = D
  smm_gemm(T2, T1, SB[0], n, tid, tile);                             // T2 = D SB11
  smm_gemm(T3, T1, SB[1], n, tid, tile);                             // T3 = D SB12
  smm_scale_add(O[0], a12, T2, a21, nullptr, a11, n, tid);     // S11 = a11 + a12 D SB11 a21
  smm_scale_add(O[1], a12, T3, nullptr, nullptr, nullptr, n, tid);   // S12 = a12 D SB12
  for (int idx = tid; idx < n * n; idx += SMM_NT) {           // T0 = F = I + a22 D SB11
    const int i = idx / n, j = idx - i * n;
    const cplx v = cmul(a22[i], T2[idx]);
    T0[idx] = make_double2(((i == j) ? 1.0 : 0.0) + v.x, v.y);
  }
  __syncthreads();
  smm_gemm(T1, SB[2], T0, n, tid, tile);                             // T1 = SB21 F
  smm_scale_add(O[2], nullptr, T1, a21, nullptr, nullptr, n, tid);   // S21 = SB21 F a21
  smm_scale_add(T0, nullptr, T1, a22, nullptr, nullptr, n, tid);     // T0 = SB21 F a22
  smm_gemm(T2, T0, SB[1], n, tid, tile);
  smm_axpby(O[3], 1.0, SB[3], 1.0, T2, 0.0, n, tid);           // S22 = SB22 + SB21 F a22 SB12
  return sing;
}

// Column `z` of S11 and S21 of SA (*) SB where every block of SB is DIAGONAL (b11, b12, b21, b22): all the last star product
// has to deliver (the incident field is the single harmonic z).  With D = inv(I - b11 SA22), u = b11 (.) SA21[:, z], v = D u:
//   S11[:, z] = SA11[:, z] + SA12 v            S21[:, z] = b21 (.) (SA21[:, z] + SA22 v)        (F x = x + SA22 D b11 x)
// One inverse and matrix-vector products.  `u`, `v`: n shared complex numbers each, `pa`, `pb`: three each (n may be as small
// as 3); results for `rows[0..3)` only.
__device__ __forceinline__ bool smm_redheffer_diag_right_col(cplx (&s11)[3], cplx (&s21)[3], const int (&rows)[3], cplx* const (&SA)[4],
                                                             const cplx* b11, const cplx* b21, int z, cplx* T0, cplx* T1, cplx* u,
                                                             cplx* v, cplx* pa, cplx* pb, int n, int tid, const SmmInvSm& W, cplx* tile) {
  for (int idx = tid; idx < n * n; idx += SMM_NT) {           // T0 = I - b11 SA22
    const int i = idx / n, j = idx - i * n;
    const cplx w = cmul(b11[i], SA[3][idx]);
    T0[idx] = make_double2(((i == j) ? 1.0 : 0.0) - w.x, -w.y);
  }
  for (int i = tid; i < n; i += SMM_NT) u[i] = cmul(b11[i], SA[2][(size_t)i * n + z]);
  __syncthreads();
  const bool sing = smm_invert(T1, T0, n, tid, W);  // T1 = D
  const int lane = tid & 31, warp = tid >> 5;
  for (int i = warp; i < n; i += SMM_NT / 32) {                // v = D u, one warp per row
    cplx acc = make_double2(0, 0);
    for (int j = lane; j < n; j += 32) acc = cfma(T1[(size_t)i * n + j], u[j], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { acc.x += __shfl_xor_sync(0xffffffffu, acc.x, o); acc.y += __shfl_xor_sync(0xffffffffu, acc.y, o); }
    if (lane == 0) v[i] = acc;
  }
  __syncthreads();
  if (warp < 6) {                                              // six dot products: rows[r] of SA12 v and of SA22 v
    const int r = warp % 3, i = rows[r];
    const cplx* Mrow = ((warp < 3) ? SA[1] : SA[3]) + (size_t)i * n;
    cplx acc = make_double2(0, 0);
    for (int j = lane; j < n; j += 32) acc = cfma(Mrow[j], v[j], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { acc.x += __shfl_xor_sync(0xffffffffu, acc.x, o); acc.y += __shfl_xor_sync(0xffffffffu, acc.y, o); }
    if (lane == 0) ((warp < 3) ? pa : pb)[r] = acc;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int i = rows[r];
    s11[r] = cadd(SA[0][(size_t)i * n + z], pa[r]);
    s21[r] = cmul(b21[i], cadd(SA[2][(size_t)i * n + z], pb[r]));
  }
  return sing;
}

// homogeneous_1D (smm_util.py:136-182) for one harmonic with normalised kx = kap: V = Q (-i Kz)^-1, Kz (times n^2 for TM when
// `pol`); exact zeros are replaced by the reference's 1e-20 (1 + 1j) perturbation.
__device__ __forceinline__ void smm_homogeneous(double kap, double n_index, bool pol, cplx* V, cplx* Kz) {
  const double er = n_index * n_index, x = er - kap * kap;
  cplx Q = make_double2(x, 0.0);
  if (x == 0.0) Q = make_double2(1e-20, -1e-20);                       // conj(perturbation)
  cplx kz = (x >= 0.0) ? make_double2(sqrt(x), 0.0) : make_double2(0.0, -sqrt(-x));   // conj(sqrt(x))
  if (kz.x == 0.0 && kz.y == 0.0) kz = make_double2(1e-20, 1e-20);
  if (pol) { kz.x *= er; kz.y *= er; }
  const cplx ev = make_double2(kz.y, -kz.x);                           // -i Kz
  *V = cmul(Q, cinv(ev));
  *Kz = kz;
}

__global__ void __launch_bounds__(SMM_NT) k_smm(KParams P, double* __restrict__ smm_scratch) {
  extern __shared__ __align__(16) double sm[];
  const int ff = P.ff, N = P.N, z = 2 * N, nc = P.n_cells, n = ff;
  const int L = P.n_layers > 0 ? P.n_layers : 1;
  const bool tm = P.pol != 0;
  const int tid = threadIdx.x;
  double* cre = sm;
  double* cim = cre + (2 * N + 1);
  cplx* vg = reinterpret_cast<cplx*>(sm + pirl_even(2 * (2 * N + 1)));
  cplx* vr = vg + ff;
  cplx* vt = vr + ff;
  cplx* kzr = vt + ff;
  cplx* kzt = kzr + ff;
  cplx* qv = kzt + ff;
  cplx* Xv = qv + ff;
  cplx* d11 = Xv + ff;                 // diagonal S-blocks of the half space being attached (reflection, then transmission)
  cplx* d12 = d11 + ff;
  cplx* d21 = d12 + ff;
  cplx* d22 = d21 + ff;
  cplx* fcol = d22 + ff;               // pivot column of the inverse
  cplx* rowk = fcol + ff;              // the two rows that swap in a step of the in-place inverse
  cplx* rowp = rowk + ff;
  uint8_t* su = reinterpret_cast<uint8_t*>(rowp + ff);
  double* misc = reinterpret_cast<double*>(su + 8 * pirl_even((nc + 7) / 8));      // [32]: pivot slot and candidates
  int* slot = reinterpret_cast<int*>(misc);
  cplx* tile = reinterpret_cast<cplx*>(sm + smm_vec_doubles(ff, N, nc));           // GEMM panels / matrix of the inverse
  SmmInvSm W;
  W.fcol = fcol; W.rowk = rowk; W.rowp = rowp; W.slot = slot; W.candm = misc + 8; W.candi = reinterpret_cast<int*>(misc + 16);
  W.piv = reinterpret_cast<int*>(misc + 32);
  W.Ms = P.smem_mats ? tile : nullptr;
  cplx* base = reinterpret_cast<cplx*>(smm_scratch + (size_t)blockIdx.x * smm_scratch_doubles(ff));
  const size_t m2 = (size_t)n * n;
  cplx* M[SMM_NMAT];
#pragma unroll
  for (int i = 0; i < SMM_NMAT; ++i) M[i] = base + i * m2;
  double* Amat = reinterpret_cast<double*>(base + SMM_NMAT * m2);      // [n][n] real: E^- convolution matrix in the real basis

  for (int b = blockIdx.x; b < P.B; b += gridDim.x) {
    __syncthreads();
    const double* pr0 = P.params + (size_t)b * L * P.pstride;          // (wavelength, angle) are those of the structure's first entry
    const double wl = pr0[0];
    const double sa = sin(pr0[1] * (3.14159265358979323846 / 180.0));
    const double nI = P.n_inc, nII = P.n_exit, eg = P.eps_g;
    // per-harmonic media (scattering_1d_1, scattering_method.py:12-31)
    for (int j = tid; j < ff; j += SMM_NT) {
      const double kap = (j < z) ? ((j >> 1) + 1) * sa : 0.0;
      cplx dummy;
      smm_homogeneous(kap, 1.0, false, &vg[j], &dummy);               // gap medium, pol = None
      smm_homogeneous(kap, nI, tm, &vr[j], &kzr[j]);
      smm_homogeneous(kap, nII, tm, &vt[j], &kzt[j]);
    }
    __syncthreads();
    cplx *A = M[0], *B = M[1], *Ai = M[2], *T0 = M[3], *T1 = M[4], *T2 = M[5], *T3 = M[6], *SL11 = M[7], *SL12 = M[8];
    cplx* G[4] = {M[9], M[10], M[11], M[12]};
    cplx* H[4] = {M[13], M[14], M[15], M[16]};
    // S_RT of the reflection region (A_B_matrices_half_space + S_RT, ref_mode = True): four diagonal blocks
    for (int i = tid; i < n; i += SMM_NT) {
      const cplx rr = cmul(cinv(vg[i]), vr[i]);
      const cplx a = make_double2(1.0 + rr.x, rr.y), bb = make_double2(1.0 - rr.x, -rr.y), ai = cinv(a);
      d11[i] = cmul(make_double2(-bb.x, -bb.y), ai);
      d12[i] = cscale(ai, 2.0);
      d21[i] = cscale(csub(a, cmul(cmul(bb, ai), bb)), 0.5);
      d22[i] = cmul(bb, ai);
    }
    __syncthreads();
    bool sing = false;
    int wst = 0;
    int cur = 0;                          // global matrix so far: 0 = R (reflection region only), 1 = G, 2 = H
    for (int step = 0; step < L; ++step) {
      const int l = P.layer_rev ? L - 1 - step : step;
      const size_t e = (size_t)b * L + l;                               // entry of this layer in the per-layer workspaces
      const double* pr = P.params + e * P.pstride;
      const double thick = pr[2], nr = pr[3], er = nr * nr;
      const double k0d = 2.0 * 3.14159265358979323846 / wl * thick;
      const double* lam = P.lam + e * ff;
      const double* Wg = P.Wbuf + e * ff * ff;
      wst |= P.wstatus[e];
      for (int j = tid; j < ff; j += SMM_NT) {                          // per-mode q and X (S_layer, smm_util.py:55)
        const double lm = lam[j];
        if (lm >= 0.0) { const double s = sqrt(lm); qv[j] = make_double2(s, 0.0); Xv[j] = make_double2(exp(-k0d * s), 0.0); }
        else { const double s = sqrt(-lm); double sn, cn; sincos(k0d * s, &sn, &cn); qv[j] = make_double2(0.0, s); Xv[j] = make_double2(cn, -sn); }
      }
      if (tm) {
        // E^- of this layer in the real basis (closed-form Fourier coefficients, convolution_matrix.py:101-114);
        // W^-1 = W^T E^- (the TM modes are E^- -orthonormal: W^T E^- W = I)
        for (int p = tid; p < nc; p += SMM_NT) su[p] = P.patterns[e * nc + p] > 0;
        __syncthreads();
        for (int k = tid; k <= 2 * N; k += SMM_NT) {
          double sx = 0.0, sy = 0.0;
          int idx = 0;
          const int km = k % nc;
          for (int p = 0; p < nc; ++p) {
            if (su[p]) { const cplx w = __ldg(&P.tw[idx]); sx += w.x; sy += w.y; }
            idx += km; if (idx >= nc) idx -= nc;
          }
          const cplx d = __ldg(&P.Dk[k]);
          cre[k] = sx * d.x - sy * d.y;
          cim[k] = sx * d.y + sy * d.x;
        }
        __syncthreads();
        const double da = 1.0 / er - 1.0 / eg, ia = 1.0 / eg;
        for (int idx = tid; idx < n * n; idx += SMM_NT) {
          const int i = idx / n, j = idx - i * n;
          Amat[idx] = fma(da, basis_toeplitz(i, j, N, cre, cim), (i == j) ? ia : 0.0);
        }
      }
      __syncthreads();
      // A_B_matrices (smm_util.py:24-39) with Wg = I: A = W^-1 + V^-1 Vg, B = W^-1 - V^-1 Vg, V^-1 = Q^-1 W^T (TM: V = E^- W Q
      // with W^T E^- W = I; TE: V = W Q with W^T W = I)
      for (int idx = tid; idx < n * n; idx += SMM_NT) {
        const int m = idx / n, j = idx - m * n;
        double winv = 0.0;
        if (tm) { for (int i = 0; i < n; ++i) winv = fma(Wg[(size_t)i * n + m], Amat[(size_t)i * n + j], winv); }
        else winv = Wg[(size_t)j * n + m];
        const cplx vi = cscale(cmul(cinv(qv[m]), vg[j]), Wg[(size_t)j * n + m]);
        A[idx] = make_double2(winv + vi.x, vi.y);
        B[idx] = make_double2(winv - vi.x, -vi.y);
      }
      __syncthreads();
      // S_layer (smm_util.py:42-67)
      smm_axpby(T0, 1.0, A, 0.0, nullptr, 0.0, n, tid);
      sing |= smm_invert(Ai, T0, n, tid, W);            // A_i
      smm_gemm(T0, B, Ai, n, tid, tile);
      smm_scale(T1, Xv, T0, Xv, n, tid);                          // T1 = X B A_i X
      smm_gemm(T2, T1, B, n, tid, tile);
      smm_axpby(T2, 1.0, A, -1.0, T2, 0.0, n, tid);               // A - X B A_i X B
      sing |= smm_invert(T3, T2, n, tid, W);            // term_i
      smm_gemm(T0, T1, A, n, tid, tile);
      smm_axpby(T0, 1.0, T0, -1.0, B, 0.0, n, tid);
      smm_gemm(SL11, T3, T0, n, tid, tile);                             // S11 = S22
      smm_gemm(T0, Ai, B, n, tid, tile);
      smm_gemm(T2, B, T0, n, tid, tile);
      smm_axpby(T2, 1.0, A, -1.0, T2, 0.0, n, tid);               // A - B A_i B
      smm_scale(T0, Xv, T2, nullptr, n, tid);
      smm_gemm(SL12, T3, T0, n, tid, tile);                             // S12 = S21
      cplx* Lm[4] = {SL11, SL12, SL12, SL11};
      // Sg = Sg (*) S_layer   (scattering_1d_2)
      if (cur == 0) { sing |= smm_redheffer_diag_left(G, d11, d12, d21, d22, Lm, T0, T1, T2, T3, n, tid, W, tile); cur = 1; }
      else if (cur == 1) { sing |= smm_redheffer(H, G, Lm, T0, T1, T2, T3, n, tid, W, tile); cur = 2; }
      else { sing |= smm_redheffer(G, H, Lm, T0, T1, T2, T3, n, tid, W, tile); cur = 1; }
    }
    // transmission region (ref_mode = False: blocks swapped -> b11 = s22, b21 = s12); Sg = Sg (*) S_trn (scattering_1d_3), of
    // which only column z (the incident harmonic) of S11 and S21 is used: rows {cos 1, sin 1, 0} = orders -1, +1, 0
    for (int i = tid; i < n; i += SMM_NT) {
      const cplx rr = cmul(cinv(vg[i]), vt[i]);
      const cplx a = make_double2(1.0 + rr.x, rr.y), bb = make_double2(1.0 - rr.x, -rr.y), ai = cinv(a);
      d11[i] = cmul(bb, ai);                                            // s22
      d21[i] = cscale(ai, 2.0);                                         // s12
    }
    __syncthreads();
    cplx c11[3], c21[3];
    const int rows[3] = {0, 1, z};
    if (cur == 1) sing |= smm_redheffer_diag_right_col(c11, c21, rows, G, d11, d21, z, T0, T1, qv, Xv, d12, d22, n, tid, W, tile);
    else sing |= smm_redheffer_diag_right_col(c11, c21, rows, H, d11, d21, z, T0, T1, qv, Xv, d12, d22, n, tid, W, tile);
    // fields of the incident zeroth order (c_inc = delta: W_r = I) and efficiencies of orders -1, 0, +1
    if (tid == 0) {
      const cplx Rc = c11[0], Rs = c11[1], R0 = c11[2];
      const cplx Tc = c21[0], Ts = c21[1], T0c = c21[2];
      const double is2 = 0.7071067811865476;
      const cplx Tm = make_double2((Tc.x - Ts.y) * is2, (Tc.y + Ts.x) * is2);
      const cplx Tp = make_double2((Tc.x + Ts.y) * is2, (Tc.y - Ts.x) * is2);
      const cplx Rm = make_double2((Rc.x - Rs.y) * is2, (Rc.y + Rs.x) * is2);
      const cplx Rp = make_double2((Rc.x + Rs.y) * is2, (Rc.y - Rs.x) * is2);
      // TM: de_ri = Re(Kzr) |r|^2 / Re(k_inc_z) / n_I^2 ; de_ti = Re(Kzt) |t|^2 / Re(k_inc_z) n_I^2 / n_II^4 ; TE: without the
      // index factors   (scattering_method.py:62-69)
      const double kinc = nI;
      const double fr = tm ? 1.0 / (nI * nI) : 1.0, ft = tm ? (nI * nI) / (nII * nII * nII * nII) : 1.0;
      double der[3], det[3];
      der[0] = kzr[0].x * (Rm.x * Rm.x + Rm.y * Rm.y) / kinc * fr;
      der[1] = kzr[z].x * (R0.x * R0.x + R0.y * R0.y) / kinc * fr;
      der[2] = kzr[0].x * (Rp.x * Rp.x + Rp.y * Rp.y) / kinc * fr;
      det[0] = kzt[0].x * (Tm.x * Tm.x + Tm.y * Tm.y) / kinc * ft;
      det[1] = kzt[z].x * (T0c.x * T0c.x + T0c.y * T0c.y) / kinc * ft;
      det[2] = kzt[0].x * (Tp.x * Tp.x + Tp.y * Tp.y) / kinc * ft;
      int st = wst;
      if (sing || !isfinite(det[0]) || !isfinite(det[1]) || !isfinite(der[1])) st |= PIRL_ST_NONFINITE_;
      P.eff[b] = det[0];
      if (P.de_r3) { P.de_r3[3 * b] = der[0]; P.de_r3[3 * b + 1] = der[1]; P.de_r3[3 * b + 2] = der[2]; }
      if (P.de_t3) { P.de_t3[3 * b] = det[0]; P.de_t3[3 * b + 1] = det[1]; P.de_t3[3 * b + 2] = det[2]; }
      if (P.status) P.status[b] = st;
    }
  }
}
