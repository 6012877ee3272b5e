// rcwa_grad.cuh -- adjoint (reverse-mode) gradient d eff / d s_p of the batched RCWA evaluation.
//
// The reference has no gradient code; the contract is the central finite difference of the
// reference forward solver w.r.t. a continuous pixel value (JLAB/solver.py:21-22 is affine in the cell value, the
// permittivity and its reciprocal are taken pixel by pixel before convolution_matrix.py:89-114).  The reverse sweep is
// validated step by step in tools/proto_grad.py.
//
// Two kernels after the forward front / QL / apply kernels have left W (A-orthonormal modes) and lambda in the workspaces:
//   g_tail     forward layer solve (same algebra as k_tail) keeping P, Hq, H_II and the pivoted LU of S, then the reverse
//              sweep of the layer solve.  Every adjoint of the layer solve is rank one or two (S-bar = -bb c+^H), so this is
//              a dozen matrix-vector products; it emits lambda-bar and a rank-15 factorisation of Pm = W^T W-bar.
//   g_eigback  reverse sweep of the generalised symmetric eigenproblem B w = lambda A w (three real GEMMs), of
//              B = S_K^T E^-1 S_K - I and E^-1 (two GEMMs), Toeplitz-diagonal sums, inverse DFT to the pixels.
//
// Convention for a real objective f and a complex intermediate z: zbar = df/dRe z + i df/dIm z, df = Re(conj(zbar) dz).
#pragma once
#include "rcwa_regtile.cuh"

#define PIRL_GRAD_RANK 16   // 15 used: 6 complex outer products (12 real) + 3 rows of W-bar
#define PIRL_GRAD_NV 20     // doubles per mode in Vf: g1a g1b g2a g2b bb (h-2c+) -hb Pw2 c+ (ybar-2g2a)

struct GParams {
  double* lamb;        // [B][ff]       d f / d lambda_j
  double* Lf;          // [B][ff][16]   Pm = Lf Rf^T
  double* Rf;          // [B][ff][16]
  double* Vf;          // [B][ff][20]   ten complex vectors: rank-one adjoints of the diagonal matrices q, X, 1-X^2, Xq
  double* grad;        // [B][n_cells]  output
  double* gscratch;    // per-CTA scratch (L2 resident)
  size_t gstride;      // doubles per CTA
  int smem_mats;       // matrices held in shared memory (kernel specific)
  int packed;          // g_tail, shared-memory variant: G_II is inverted as a packed lower triangle in shared memory
  int lu_k0;           // ... and the trailing block [lu_k0, ff)^2 of S is factorised in shared memory (ff: never)
  int use_mma;         // g_eigback: the six real GEMMs on FP64 tensor tiles (mma.sync.m8n8k4.f64) instead of register-tiled DFMA
};

__host__ __device__ inline int grad_tail_vec_doubles(int ff) { return tail_vec_doubles(ff) + pirl_even(ff / 2 + 2) + pirl_even(ff); }
// per-CTA scratch of g_tail (doubles): Hq, H_II complex [ff][ff]; two complex spill matrices [ff][ld]; 16 complex vectors
__host__ __device__ inline size_t grad_tail_scratch_doubles(int ff, int ld) {
  return 4 * (size_t)ff * ff + 4 * (size_t)ff * ld + 32 * (size_t)ff + 16;
}
// g_eigback shared vectors (doubles): lam, lamb, dpm, rowk, colk [ff]; cre, cim, gEr, gEi, gAr, gAi [2N+1 = ff]; q, X complex [ff]; pattern bits
__host__ __device__ inline int grad_eig_vec_doubles(int ff, int n_cells) { return pirl_even(15 * ff + (n_cells + 7) / 8 + 2); }
__host__ __device__ inline size_t grad_eig_scratch_doubles(int ff, int ld) { return 4 * (size_t)ff * ld + 16; }

__device__ __forceinline__ cplx cconj(cplx a) { return make_double2(a.x, -a.y); }

// 1 / z for the pivots of the elimination loops: hardware seed + two Newton steps (full double accuracy up to an ulp or
// two, no slow path; the IEEE division is ~4x longer and sits on the critical path of every elimination step)
__device__ __forceinline__ double rcp_newton(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  double e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  e = fma(-d, r, 1.0);
  return fma(r, e, r);
}
__device__ __forceinline__ cplx cinv_fast(cplx a) {
  const double d = rcp_newton(fma(a.x, a.x, a.y * a.y));
  return make_double2(a.x * d, -a.y * d);
}

// y_i = sum_j M[i][j] x_j (+ diag_i x_i): one warp per row.  M complex symmetric in every use, so this also serves M^T x.
template <int NT>
__device__ __forceinline__ void cta_symv(const cplx* M, int ld, const cplx* x, cplx* y, int ff, const cplx* diag = nullptr) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = warp; i < ff; i += NT / 32) {
    cplx s = make_double2(0.0, 0.0);
    for (int j = lane; j < ff; j += 32) s = cfma(M[(size_t)i * ld + j], x[j], s);
    s = warp_sum(s);
    if (lane == 0) {
      if (diag) s = cfma(diag[i], x[i], s);
      y[i] = s;
    }
  }
}


// ---------------------------------------------------------------------------------------------------
// Complex register tiles for the dense steps of g_tail (ff <= 96, 512 threads = 16 warps): warp w owns rows w + 16 a
// (a < RA), lane l owns columns l + 32 c (c < CC).  Per elimination step only the pivot row and column pass through shared
// memory (RA + CC complex loads per RA * CC complex FMAs) instead of the whole matrix.
// ---------------------------------------------------------------------------------------------------
template <int RA, int CC>
__device__ __forceinline__ void ctile_store(const cplx (&t)[RA][CC], cplx* M, int ld, int ff, int w, int lane) {
#pragma unroll
  for (int a = 0; a < RA; ++a) {
    const int i = w + 16 * a;
#pragma unroll
    for (int c = 0; c < CC; ++c) {
      const int j = lane + 32 * c;
      if (i < ff && j < ff) M[i * ld + j] = t[a][c];
    }
  }
}

// All tile indices below are compile-time constants after unrolling (the step index is split as k = 16 KB + kk with KB
// unrolled): a run-time row / column slot makes the compiler index the tile in local memory and re-store it every step.
//
// publish row k (owner warp) and column k (owner lanes) of the tile; KA = k >> 4 is a constant at every call site
template <int RA, int CC>
__device__ __forceinline__ void ctile_publish(const cplx (&t)[RA][CC], cplx* rowk, cplx* colk, int k, const int KA, int ff, int w, int lane) {
  const int KC = KA >> 1;
  if (w == (k & 15)) {
#pragma unroll
    for (int c = 0; c < CC; ++c) { const int j = lane + 32 * c; if (j < ff) rowk[j] = t[KA][c]; }
  }
  if (lane == (k & 31)) {
#pragma unroll
    for (int a = 0; a < RA; ++a) { const int i = w + 16 * a; if (i < ff) colk[i] = t[a][KC]; }
  }
}

// in-place inverse by Gauss-Jordan without pivoting (complex symmetric G with a dominant diagonal).  One barrier per step:
// the owners publish row / column k + 1 right after their update of step k into the other pair of buffers, and the owner of
// the next pivot computes its reciprocal one step ahead (pslot), off the critical path.
template <int RA, int CC>
__device__ __forceinline__ void ctile_gj_inverse(cplx (&t)[RA][CC], cplx* rc0, cplx* rc1, cplx* rc2, cplx* rc3, cplx* pslot, int ff, int w,
                                                 int lane) {
  __syncthreads();
  ctile_publish<RA, CC>(t, rc0, rc1, 0, 0, ff, w, lane);
  if (w == 0 && lane == 0) pslot[0] = cinv_fast(t[0][0]);
  __syncthreads();
  const int nra = (ff - w + 15) >> 4;              // row slots of this warp that hold a row (warp-uniform)
#pragma unroll
  for (int KB = 0; KB < RA; ++KB) {
#pragma unroll 1
    for (int kk = 0; kk < 16; ++kk) {
      const int k = 16 * KB + kk;
      if (k >= ff) break;
      const cplx* rowk = (k & 1) ? rc2 : rc0;
      const cplx* colk = (k & 1) ? rc3 : rc1;
      const cplx piv = rowk[k];
      const cplx p = pslot[k & 1];
      const int k1 = k + 1;
      if (k1 < ff && w == (k1 & 15) && lane == (k1 & 31)) {
        cplx told;
        if (kk < 15) told = t[KB][KB >> 1];
        else told = t[(KB + 1 < RA) ? KB + 1 : KB][(KB + 1 < RA) ? ((KB + 1) >> 1) : 0];
        const cplx cvn = cscale(cmul(colk[k1], p), -1.0);
        pslot[k1 & 1] = cinv_fast(cfma(cvn, rowk[k1], told));
      }
      cplx rv[CC];
#pragma unroll
      for (int c = 0; c < CC; ++c) {
        const int j = lane + 32 * c;
        rv[c] = (j < ff) ? ((j == k) ? make_double2(1.0 + piv.x, piv.y) : rowk[j]) : make_double2(0.0, 0.0);
      }
      // (loads first, see the S product in g_tail; row slots past the matrix work on a clamped row, their tile is never read)
      cplx ck[RA];
#pragma unroll
      for (int a = 0; a < RA; ++a) ck[a] = colk[min(w + 16 * a, ff - 1)];
#pragma unroll
      for (int a = 0; a < RA; ++a) {
        if (a == RA - 1 && nra < RA) continue;
        const int i = w + 16 * a;
        const cplx cv = (i == k) ? make_double2(1.0 - p.x, -p.y) : cmul(ck[a], p);     // minus the usual multiplier
#pragma unroll
        for (int c = 0; c < CC; ++c) t[a][c] = cfnma(cv, rv[c], t[a][c]);
      }
      if (k1 < ff) {
        if (kk < 15) ctile_publish<RA, CC>(t, (k & 1) ? rc0 : rc2, (k & 1) ? rc1 : rc3, k1, KB, ff, w, lane);
        else ctile_publish<RA, CC>(t, (k & 1) ? rc0 : rc2, (k & 1) ? rc1 : rc3, k1, (KB + 1 < RA) ? KB + 1 : KB, ff, w, lane);
      }
      __syncthreads();
    }
  }
}

// LU with implicit partial pivoting: rows are never moved; rperm[k] = pivot row of step k.  Multipliers are left in column k
// of the rows still unused, the pivot rows hold U.  The right-hand side bvec (shared) is eliminated alongside.
// Two barriers per step: column k + 1 and its magnitudes are published right after the update of step k (buffers alternate).
template <int RA, int CC>
__device__ __forceinline__ void ctile_lu_publish_col(const cplx (&t)[RA][CC], cplx* colk, double* magk, unsigned used, int k, const int KC,
                                                     int ff, int w, int lane) {
  if (lane == (k & 31)) {
#pragma unroll
    for (int a = 0; a < RA; ++a) {
      const int i = w + 16 * a;
      if (i < ff) { colk[i] = t[a][KC]; magk[i] = ((used >> a) & 1u) ? -1.0 : fabs(t[a][KC].x) + fabs(t[a][KC].y); }
    }
  }
}

template <int RA, int CC>
__device__ __forceinline__ void ctile_lu_implicit(cplx (&t)[RA][CC], cplx* rowk, cplx* colk0, cplx* colk1, double* magk0, double* magk1,
                                                  cplx* bvec, int* rperm, int ff, int w, int lane) {
  unsigned used = 0;
#pragma unroll
  for (int a = 0; a < RA; ++a) if (w + 16 * a >= ff) used |= 1u << a;
  __syncthreads();
  ctile_lu_publish_col<RA, CC>(t, colk0, magk0, used, 0, 0, ff, w, lane);
  __syncthreads();
#pragma unroll
  for (int KC = 0; KC < CC; ++KC) {
#pragma unroll 1
    for (int kk = 0; kk < 32; ++kk) {
      const int k = 32 * KC + kk;
      if (k >= ff) break;
      const cplx* colk = (k & 1) ? colk1 : colk0;
      const double* magk = (k & 1) ? magk1 : magk0;
      // every warp finds the same pivot (largest |re| + |im| among the unused rows, lowest row on ties)
      // (tried: integer keys with the row packed into the low bits and the reciprocal of each lane's best candidate computed
      // under the shuffle rounds -- 15 % slower per step than this plain search)
      double best = -2.0; int pv = 0;
      for (int i = lane; i < ff; i += 32) { const double m = magk[i]; if (m > best) { best = m; pv = i; } }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, pv, o);
        if (ob > best || (ob == best && oi < pv)) { best = ob; pv = oi; }
      }
      const int pa = pv >> 4;
      if (w == (pv & 15)) {
        // the pivot row sits in a run-time slot: chain of value selects over the slots (no run-time tile index)
        cplx rowv[CC];
#pragma unroll
        for (int c = 0; c < CC; ++c) rowv[c] = t[0][c];
#pragma unroll
        for (int a = 1; a < RA; ++a)
#pragma unroll
          for (int c = 0; c < CC; ++c) { rowv[c].x = (pa == a) ? t[a][c].x : rowv[c].x; rowv[c].y = (pa == a) ? t[a][c].y : rowv[c].y; }
#pragma unroll
        for (int c = 0; c < CC; ++c) { const int j = lane + 32 * c; if (j < ff) rowk[j] = rowv[c]; }
        used |= 1u << pa;
        if (lane == 0) rperm[k] = pv;
      }
      __syncthreads();
      const cplx pinv = cinv_fast(rowk[k]);
      const cplx bk = bvec[pv];
      cplx rv[CC];
#pragma unroll
      for (int c = 0; c < CC; ++c) { const int j = lane + 32 * c; rv[c] = (j < ff && j > k) ? rowk[j] : make_double2(0.0, 0.0); }
      cplx ck[RA];
#pragma unroll
      for (int a = 0; a < RA; ++a) ck[a] = colk[min(w + 16 * a, ff - 1)];
#pragma unroll
      for (int a = 0; a < RA; ++a) {
        if ((used >> a) & 1u) continue;          // warp-uniform
        const int i = w + 16 * a;
        const cplx f = cmul(ck[a], pinv);
#pragma unroll
        for (int c = 0; c < CC; ++c) t[a][c] = cfnma(f, rv[c], t[a][c]);
        if (lane == kk) t[a][KC] = f;
        if (lane == 0) bvec[i] = cfnma(f, bk, bvec[i]);
      }
      if (k + 1 < ff) {
        if (kk < 31) ctile_lu_publish_col<RA, CC>(t, (k & 1) ? colk0 : colk1, (k & 1) ? magk0 : magk1, used, k + 1, KC, ff, w, lane);
        else ctile_lu_publish_col<RA, CC>(t, (k & 1) ? colk0 : colk1, (k & 1) ? magk0 : magk1, used, k + 1, (KC + 1 < CC) ? KC + 1 : KC, ff, w, lane);
      }
      __syncthreads();
    }
  }
}

// The two elimination loops run as real (not inlined) calls: inside the call only the tile and a handful of pointers are
// live, so the 128-register budget of a 512-thread CTA holds the tile without spilling it every step (inlined into g_tail
// the two dozen long-lived pointers of the kernel pushed the tile into local memory: 18 stores + 14 loads per step).
template <int RA, int CC>
struct CTile { cplx v[RA][CC]; };

// Shared-memory operands of the non-inlined calls travel as offsets (in doubles) from the dynamic shared array: a pointer
// argument would be generic and every load in the callee a generic load.
template <class T>
__device__ __forceinline__ T* smem_at(int off) {
  extern __shared__ __align__(16) double sm[];
  return reinterpret_cast<T*>(sm + off);
}
__device__ __forceinline__ int smem_off(const void* p) {
  extern __shared__ __align__(16) double sm[];
  return (int)(reinterpret_cast<const double*>(p) - sm);
}

template <int RA, int CC>
__device__ __noinline__ CTile<RA, CC> ctile_gj_call(CTile<RA, CC> T, int rc0, int rc1, int rc2, int rc3, int pslot, int ff, int w, int lane) {
  cplx t[RA][CC];   // explicit register copy: the by-value parameter itself stays in param (local) memory
#pragma unroll
  for (int a = 0; a < RA; ++a)
#pragma unroll
    for (int c = 0; c < CC; ++c) t[a][c] = T.v[a][c];
  ctile_gj_inverse<RA, CC>(t, smem_at<cplx>(rc0), smem_at<cplx>(rc1), smem_at<cplx>(rc2), smem_at<cplx>(rc3), smem_at<cplx>(pslot), ff, w, lane);
  CTile<RA, CC> R;
#pragma unroll
  for (int a = 0; a < RA; ++a)
#pragma unroll
    for (int c = 0; c < CC; ++c) R.v[a][c] = t[a][c];
  return R;
}
template <int RA, int CC>
__device__ __noinline__ CTile<RA, CC> ctile_lu_call(CTile<RA, CC> T, int rowk, int colk0, int colk1, int magk0, int magk1, int bvec, int rperm,
                                                    int ff, int w, int lane) {
  cplx t[RA][CC];
#pragma unroll
  for (int a = 0; a < RA; ++a)
#pragma unroll
    for (int c = 0; c < CC; ++c) t[a][c] = T.v[a][c];
  ctile_lu_implicit<RA, CC>(t, smem_at<cplx>(rowk), smem_at<cplx>(colk0), smem_at<cplx>(colk1), smem_at<double>(magk0), smem_at<double>(magk1),
                            smem_at<cplx>(bvec), smem_at<int>(rperm), ff, w, lane);
  CTile<RA, CC> R;
#pragma unroll
  for (int a = 0; a < RA; ++a)
#pragma unroll
    for (int c = 0; c < CC; ++c) R.v[a][c] = t[a][c];
  return R;
}

// W^T Re(D_I) W and W^T Re(D_II) W (real parts of H_I, H_II) for the rows / columns of this thread: harmonic index outermost,
// all loads of a step issued before its multiply-adds; row slots past the matrix read a clamped row (never used afterwards)
template <int RA, int CC>
struct RTile2 { double a[RA][CC], b[RA][CC]; };

template <int RA, int CC>
__device__ __noinline__ RTile2<RA, CC> rtile_wdw_call(int oW, int oDI, int oDII, int ff, int ld, int w, int lane) {
  const double* Wr = smem_at<double>(oW);
  const cplx* DI = smem_at<cplx>(oDI);
  const cplx* DII = smem_at<cplx>(oDII);
  double aI[RA][CC], aII[RA][CC];
#pragma unroll
  for (int a = 0; a < RA; ++a)
#pragma unroll
    for (int c = 0; c < CC; ++c) { aI[a][c] = 0.0; aII[a][c] = 0.0; }
  int rix[RA], cix[CC];
#pragma unroll
  for (int a = 0; a < RA; ++a) rix[a] = min(w + 16 * a, ff - 1);
#pragma unroll
  for (int c = 0; c < CC; ++c) cix[c] = min(lane + 32 * c, ff - 1);
  const bool last_slot = w + 16 * (RA - 1) < ff;
#pragma unroll 2
  for (int m = 0; m < ff; ++m) {
    const double d1 = DI[m].x, d2 = DII[m].x;
    const double* wrow = Wr + m * ld;
    double wj[CC], wi[RA];
#pragma unroll
    for (int c = 0; c < CC; ++c) wj[c] = wrow[cix[c]];
#pragma unroll
    for (int a = 0; a < RA; ++a) wi[a] = wrow[rix[a]];
#pragma unroll
    for (int a = 0; a < RA - 1; ++a) {
      const double w1 = wi[a] * d1, w2 = wi[a] * d2;
#pragma unroll
      for (int c = 0; c < CC; ++c) { aI[a][c] = fma(w1, wj[c], aI[a][c]); aII[a][c] = fma(w2, wj[c], aII[a][c]); }
    }
    if (last_slot) {
      const double w1 = wi[RA - 1] * d1, w2 = wi[RA - 1] * d2;
#pragma unroll
      for (int c = 0; c < CC; ++c) { aI[RA - 1][c] = fma(w1, wj[c], aI[RA - 1][c]); aII[RA - 1][c] = fma(w2, wj[c], aII[RA - 1][c]); }
    }
  }
  RTile2<RA, CC> R;
#pragma unroll
  for (int a = 0; a < RA; ++a)
#pragma unroll
    for (int c = 0; c < CC; ++c) { R.a[a][c] = aI[a][c]; R.b[a][c] = aII[a][c]; }
  return R;
}

// (A P) for the rows / columns of this thread, A = Hq X and P complex [ff][ld] in shared memory
template <int RA, int CC>
__device__ __noinline__ CTile<RA, CC> ctile_ap_call(int oA, int oP, int ff, int ld, int w, int lane) {
  const cplx* Am = smem_at<cplx>(oA);
  const cplx* Pm = smem_at<cplx>(oP);
  cplx t[RA][CC];
#pragma unroll
  for (int a = 0; a < RA; ++a)
#pragma unroll
    for (int c = 0; c < CC; ++c) t[a][c] = make_double2(0.0, 0.0);
  int roff[RA], cix[CC];
#pragma unroll
  for (int a = 0; a < RA; ++a) roff[a] = min(w + 16 * a, ff - 1) * ld;
#pragma unroll
  for (int c = 0; c < CC; ++c) cix[c] = min(lane + 32 * c, ff - 1);
  const bool last_slot = w + 16 * (RA - 1) < ff;
#pragma unroll 1
  for (int k = 0; k < ff; ++k) {
    cplx pk[CC], f[RA];
#pragma unroll
    for (int c = 0; c < CC; ++c) pk[c] = Pm[k * ld + cix[c]];
#pragma unroll
    for (int a = 0; a < RA; ++a) f[a] = Am[roff[a] + k];
#pragma unroll
    for (int a = 0; a < RA - 1; ++a)
#pragma unroll
      for (int c = 0; c < CC; ++c) t[a][c] = cfma(f[a], pk[c], t[a][c]);
    if (last_slot) {
#pragma unroll
      for (int c = 0; c < CC; ++c) t[RA - 1][c] = cfma(f[RA - 1], pk[c], t[RA - 1][c]);
    }
  }
  CTile<RA, CC> R;
#pragma unroll
  for (int a = 0; a < RA; ++a)
#pragma unroll
    for (int c = 0; c < CC; ++c) R.v[a][c] = t[a][c];
  return R;
}

template <int RA, int CC>
__device__ __forceinline__ CTile<RA, CC> ctile_pack(const cplx (&t)[RA][CC]) {
  CTile<RA, CC> T;
#pragma unroll
  for (int a = 0; a < RA; ++a)
#pragma unroll
    for (int c = 0; c < CC; ++c) T.v[a][c] = t[a][c];
  return T;
}
template <int RA, int CC>
__device__ __forceinline__ void ctile_unpack(cplx (&t)[RA][CC], const CTile<RA, CC>& T) {
#pragma unroll
  for (int a = 0; a < RA; ++a)
#pragma unroll
    for (int c = 0; c < CC; ++c) t[a][c] = T.v[a][c];
}

// Triangular solve by one warp with the vector in registers (lane l holds x[l + 32 c]): per step one shuffle broadcast of
// x_k, one complex multiply and one multiply-add sit on the dependency chain; the matrix elements of the step are loaded
// ahead of it.  DESC: k runs downwards and rows i < k are updated (else upwards, i > k).  ROWACC: the element that couples
// x_i to x_k is M[k][i] (else M[i][k]); CONJ: its conjugate (and conj(dinv)); UNIT: unit diagonal.
template <int CC, bool DESC, bool ROWACC, bool CONJ, bool UNIT>
__device__ __forceinline__ void warp_trisolve(cplx (&x)[CC], const cplx* M, int ld, const cplx* dinv, int ff, int lane) {
#pragma unroll
  for (int KC0 = 0; KC0 < CC; ++KC0) {
    const int KC = DESC ? CC - 1 - KC0 : KC0;
#pragma unroll 1
    for (int kk0 = 0; kk0 < 32; ++kk0) {
      const int kk = DESC ? 31 - kk0 : kk0;
      const int k = 32 * KC + kk;
      if (k >= ff) { if (DESC) continue; else break; }
      cplx m[CC];
#pragma unroll
      for (int c = 0; c < CC; ++c) {
        const int i = min(lane + 32 * c, ff - 1);
        m[c] = ROWACC ? M[k * ld + i] : M[i * ld + k];
      }
      cplx xk = make_double2(__shfl_sync(0xffffffffu, x[KC].x, kk), __shfl_sync(0xffffffffu, x[KC].y, kk));
      if (!UNIT) {
        const cplx d = dinv[k];
        xk = cmul(xk, CONJ ? cconj(d) : d);
        if (lane == kk) x[KC] = xk;
      }
#pragma unroll
      for (int c = 0; c < CC; ++c) {
        const int i = lane + 32 * c;
        const bool on = DESC ? (i < k) : (i > k && i < ff);
        if (on) x[c] = cfnma(CONJ ? cconj(m[c]) : m[c], xk, x[c]);
      }
    }
  }
}

// ===================================================================================================
// g_tail : forward layer solve + its reverse sweep
// ===================================================================================================
template <int NT, int NCOL>
__global__ void __launch_bounds__(NT) g_tail(KParams P, GParams G) {
  extern __shared__ __align__(16) double sm[];
  const int ff = P.ff, ld = P.ld, N = P.N, z = 2 * N;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = NT / 32;
  cplx* qv = reinterpret_cast<cplx*>(sm);
  cplx* Xv = qv + ff;
  cplx* XQ = Xv + ff;
  cplx* omx2 = XQ + ff;
  cplx* DI = omx2 + ff;
  cplx* DII = DI + ff;
  cplx* rowk = DII + ff;
  cplx* colk = rowk + ff;
  cplx* bvec = colk + ff;
  cplx* yv = bvec + ff;
  cplx* dv = yv + ff;
  double* wr = reinterpret_cast<double*>(dv + ff);  // [3][ff]
  cplx* red = reinterpret_cast<cplx*>(wr + 3 * ff + ((3 * ff) & 1));
  int* ipiv = reinterpret_cast<int*>(red + 32);     // [ff + 1]
  double* magk = sm + tail_vec_doubles(ff) + pirl_even(ff / 2 + 2);   // [ff] pivot magnitudes; later the inverse row permutation
  constexpr bool REG = (NCOL == 3) && (NT == 512);  // complex register tiles for ff <= 96
  constexpr int RA = REG ? 6 : 1;
  double* mat0 = sm + grad_tail_vec_doubles(ff);
  const size_t csz = (size_t)ff * ld;
  double* scr = G.gscratch + (size_t)blockIdx.x * G.gstride;
  cplx* Hq = reinterpret_cast<cplx*>(scr);          // [ff][ff] global (L2 resident)
  cplx* HII = Hq + (size_t)ff * ff;
  cplx* spill = HII + (size_t)ff * ff;
  // the register-tile variant is only launched when both matrices fit in shared memory: plain shared pointers then (a pointer
  // that may be shared or global compiles to generic loads with 64-bit address arithmetic in the inner loops)
  cplx* Pm = (REG || G.smem_mats >= 1) ? reinterpret_cast<cplx*>(mat0) : spill;
  cplx* Sm = (REG || G.smem_mats >= 2) ? reinterpret_cast<cplx*>(mat0) + csz : spill + csz;
  // W staged in S's storage; in the packed mode of the shared-memory variant S lives in L2, W then takes the (still empty)
  // shared storage of the packed G_II
  double* Wr = (!REG && G.packed) ? mat0 : reinterpret_cast<double*>(Sm);
  cplx* gv = spill + 2 * csz;                       // 16 global vectors
  cplx* Pk = reinterpret_cast<cplx*>(mat0);         // packed lower triangle of G_II (G.packed)
  cplx* v_ell = gv;           cplx* v_x = gv + ff;       cplx* v_yb = gv + 2 * ff;  cplx* v_qb = gv + 3 * ff;
  cplx* v_Xb = gv + 4 * ff;   cplx* v_hb = gv + 5 * ff;  cplx* v_w2 = gv + 6 * ff;  cplx* v_pw2 = gv + 7 * ff;
  cplx* v_x2 = gv + 8 * ff;   cplx* v_pm = gv + 9 * ff;  cplx* v_in = gv + 10 * ff; /* 6 vectors: v_in .. v_in + 6 ff */

  for (int b = blockIdx.x; b < P.B; b += gridDim.x) {
    __syncthreads();
    const double* pr = P.params + (size_t)b * P.pstride;
    const double wl = pr[0], ang = pr[1], thick = pr[2];
    const double sa = sin(ang * (3.14159265358979323846 / 180.0));
    const double k0d = 2.0 * 3.14159265358979323846 / wl * thick;
    const double nI = P.n_inc, nII = P.n_exit;
    const double* lamg = P.lam + (size_t)b * ff;
    long long pc_ = clock64();
    for (int t = tid; t < ff; t += NT) {
      const double l = lamg[t];
      cplx q, X;
      if (l >= 0.0) { double s = sqrt(l); q = make_double2(s, 0.0); X = make_double2(exp(-k0d * s), 0.0); }
      else { double s = sqrt(-l); q = make_double2(0.0, s); double sn, cs; sincos(k0d * s, &sn, &cs); X = make_double2(cs, -sn); }
      qv[t] = q; Xv[t] = X; XQ[t] = cmul(X, q);
      cplx x2 = cmul(X, X);
      omx2[t] = make_double2(1.0 - x2.x, -x2.y);
      const double kap = (t < z) ? ((t >> 1) + 1) * sa : 0.0;
      const double aI = nI * nI - kap * kap, aII = nII * nII - kap * kap;
      DI[t] = (aI >= 0.0) ? make_double2(0.0, sqrt(aI) / (nI * nI)) : make_double2(sqrt(-aI) / (nI * nI), 0.0);
      DII[t] = (aII >= 0.0) ? make_double2(0.0, sqrt(aII) / (nII * nII)) : make_double2(sqrt(-aII) / (nII * nII), 0.0);
    }
    const double* Wg = P.Wbuf + (size_t)b * ff * ff;
    for (int i = warp; i < ff; i += NW)
      for (int j = lane; j < ff; j += 32) Wr[i * ld + j] = Wg[i * ff + j];
    for (int t = tid; t < ff; t += NT) {
      wr[t] = Wg[0 * ff + t];
      wr[ff + t] = Wg[1 * ff + t];
      wr[2 * ff + t] = Wg[(size_t)z * ff + t];
    }
    __syncthreads();
    const cplx r0 = make_double2(DI[z].x, DI[z].y + 1.0 / nI);
    RK_PHASE(8);    // (slots 8-13 belong to rk_tail2 in the forward path, which does not run in a gradient call)
    if constexpr (REG) {
      cplx tl[RA][NCOL];
      // ---- G_II -> register tile, Hq and H_II -> global ----
      // D is real on evanescent harmonics and imaginary on the few propagating ones: W^T Re(D) W is one real product with
      // the harmonic index outermost (nine loads feed 36 multiply-adds); the imaginary rank-one parts are added at the end
      int* lstI = ipiv;                                  // both lists are dead before the LU needs ipiv / magk
      int* lstII = reinterpret_cast<int*>(magk);
      if (warp == 0) {
        int c1 = 0, c2 = 0;
        for (int m0 = 0; m0 < ff; m0 += 32) {
          const int m = m0 + lane;
          const bool p1 = m < ff && DI[min(m, ff - 1)].y != 0.0, p2 = m < ff && DII[min(m, ff - 1)].y != 0.0;
          const unsigned b1 = __ballot_sync(0xffffffffu, p1), b2 = __ballot_sync(0xffffffffu, p2);
          const unsigned below = (1u << lane) - 1u;
          if (p1) lstI[1 + c1 + __popc(b1 & below)] = m;
          if (p2) lstII[1 + c2 + __popc(b2 & below)] = m;
          c1 += __popc(b1); c2 += __popc(b2);
        }
        if (lane == 0) { lstI[0] = c1; lstII[0] = c2; }
      }
      {
        const int nra = (ff - warp + 15) >> 4;             // row slots of this warp that hold a row (warp-uniform)
        const RTile2<RA, NCOL> R2_ = rtile_wdw_call<RA, NCOL>(smem_off(Wr), smem_off(DI), smem_off(DII), ff, ld, warp, lane);
        const double (&aI)[RA][NCOL] = R2_.a;
        const double (&aII)[RA][NCOL] = R2_.b;
        RK_PHASE(9);
        __syncthreads();                                   // lists visible
        // imaginary parts: sum over the propagating harmonics of Im(D_m) w_m w_m^T, first medium I (-> Hq), then II
#pragma unroll
        for (int pass = 0; pass < 2; ++pass) {
          const int* lst = pass ? lstII : lstI;
          const cplx* Dm = pass ? DII : DI;
          const int np = lst[0];
          double im[RA][NCOL];
#pragma unroll
          for (int a = 0; a < RA; ++a)
#pragma unroll
            for (int c = 0; c < NCOL; ++c) im[a][c] = 0.0;
          for (int e = 0; e < np; ++e) {
            const int m = lst[1 + e];
            const double beta = Dm[m].y;
            const double* wrow = Wr + m * ld;
            double wj[NCOL];
#pragma unroll
            for (int c = 0; c < NCOL; ++c) wj[c] = wrow[min(lane + 32 * c, ff - 1)];
#pragma unroll
            for (int a = 0; a < RA; ++a) {
              if (a >= nra) continue;
              const double wi = wrow[warp + 16 * a] * beta;
#pragma unroll
              for (int c = 0; c < NCOL; ++c) im[a][c] = fma(wi, wj[c], im[a][c]);
            }
          }
#pragma unroll
          for (int a = 0; a < RA; ++a) {
            const int i = warp + 16 * a;
#pragma unroll
            for (int c = 0; c < NCOL; ++c) {
              const int j = lane + 32 * c;
              if (pass == 0) {
                if (a < nra && j < ff) {
                  cplx h = make_double2(aI[a][c], im[a][c]);
                  if (i == j) h = csub(h, qv[i]);
                  Hq[(size_t)i * ff + j] = h;
                }
              } else {
                cplx g = make_double2(aII[a][c], im[a][c]);
                if (a < nra && j < ff) {
                  HII[(size_t)i * ff + j] = g;
                  if (i == j) g = cadd(g, qv[i]);
                }
                tl[a][c] = g;
              }
            }
          }
        }
      }
      RK_PHASE(16);
      { const CTile<RA, NCOL> R_ = ctile_gj_call<RA, NCOL>(ctile_pack<RA, NCOL>(tl), smem_off(rowk), smem_off(colk), smem_off(yv), smem_off(dv), smem_off(red), ff, warp, lane); ctile_unpack<RA, NCOL>(tl, R_); }
      ctile_store<RA, NCOL>(tl, Pm, ld, ff, warp, lane);
      // Hq X staged in S's storage (W is dead, the LU rows are written back much later)
      for (int i = warp; i < ff; i += NW)
        for (int k = lane; k < ff; k += 32) Sm[i * ld + k] = cmul(Hq[(size_t)i * ff + k], Xv[k]);
      __syncthreads();
      RK_PHASE(17);
      // ---- S = Hq (I - X^2) + 2Q + 2 (Hq X) P (X Q) -> register tile ----
      // k outermost: per k three shared-memory loads of P and six broadcast loads of Hq X feed 18 complex multiply-adds
      {
        const int nra = (ff - warp + 15) >> 4;
        const cplx* h0 = Hq + (size_t)warp * ff;           // row warp + 16 a starts at h0 + a * hstep
        const int hstep = 16 * ff;
        { const CTile<RA, NCOL> R_ = ctile_ap_call<RA, NCOL>(smem_off(Sm), smem_off(Pm), ff, ld, warp, lane); ctile_unpack<RA, NCOL>(tl, R_); }
        RK_PHASE(10);
#pragma unroll
        for (int a = 0; a < RA; ++a) {
          if (a >= nra) continue;
          const int i = warp + 16 * a;
#pragma unroll
          for (int c = 0; c < NCOL; ++c) {
            const int j = min(lane + 32 * c, ff - 1);
            cplx sv = cmul(tl[a][c], XQ[j]);
            sv = make_double2(2.0 * sv.x, 2.0 * sv.y);
            sv = cfma(h0[a * hstep + j], omx2[j], sv);
            if (i == j) { sv.x += 2.0 * qv[j].x; sv.y += 2.0 * qv[j].y; }
            tl[a][c] = sv;
          }
        }
      }
      RK_PHASE(18);
      for (int t = tid; t < ff; t += NT) bvec[t] = cscale(r0, wr[2 * ff + t]);
      { const CTile<RA, NCOL> R_ = ctile_lu_call<RA, NCOL>(ctile_pack<RA, NCOL>(tl), smem_off(rowk), smem_off(colk), smem_off(yv), smem_off(magk), smem_off(dv), smem_off(bvec), smem_off(ipiv), ff, warp, lane); ctile_unpack<RA, NCOL>(tl, R_); }
      RK_PHASE(11);
      // rows out in pivot order: P_r S = L U in the usual layout; b' = P_r b
      int* pos = reinterpret_cast<int*>(magk);
      for (int k = tid; k < ff; k += NT) { pos[ipiv[k]] = k; colk[k] = bvec[ipiv[k]]; }
      __syncthreads();
#pragma unroll
      for (int a = 0; a < RA; ++a) {
        const int i = warp + 16 * a;
        if (i < ff) {
          const int pr_ = pos[i];
#pragma unroll
          for (int c = 0; c < NCOL; ++c) { const int j = lane + 32 * c; if (j < ff) Sm[pr_ * ld + j] = tl[a][c]; }
        }
      }
      for (int k = tid; k < ff; k += NT) bvec[k] = colk[k];
    } else {
    // ---- G_II -> Pm, Hq and H_II -> global ----
    for (int i = warp; i < ff; i += NW) {
      cplx aI[NCOL], aII[NCOL];
#pragma unroll
      for (int c = 0; c < NCOL; ++c) { aI[c] = make_double2(0.0, 0.0); aII[c] = make_double2(0.0, 0.0); }
      for (int m = 0; m < ff; ++m) {
        const double wi = Wr[m * ld + i];
        const cplx d1 = DI[m], d2 = DII[m];
#pragma unroll
        for (int c = 0; c < NCOL; ++c) {
          const int j = lane + 32 * c;
          if (j < ff) {
            const double pr2 = wi * Wr[m * ld + j];
            aI[c].x = fma(d1.x, pr2, aI[c].x); aI[c].y = fma(d1.y, pr2, aI[c].y);
            aII[c].x = fma(d2.x, pr2, aII[c].x); aII[c].y = fma(d2.y, pr2, aII[c].y);
          }
        }
      }
#pragma unroll
      for (int c = 0; c < NCOL; ++c) {
        const int j = lane + 32 * c;
        if (j < ff) {
          cplx g = aII[c], h = aI[c];
          HII[(size_t)i * ff + j] = g;
          if (i == j) { g = cadd(g, qv[i]); h = csub(h, qv[i]); }
          Pm[i * ld + j] = g;
          Hq[(size_t)i * ff + j] = h;
        }
      }
    }
    if (G.packed) {   // W is dead: its storage takes the lower triangle of G_II
      __syncthreads();
      for (int i = warp; i < ff; i += NW)
        for (int j = lane; j <= i; j += 32) Pk[i * (i + 1) / 2 + j] = Pm[i * ld + j];
    }
    // ---- P = G_II^-1 : complex symmetric, no pivoting ----
    RK_PHASE(16);
    if (G.packed) {
      // Neither complex matrix fits in shared memory at this size, but the lower triangle of the symmetric G_II does:
      // symmetric sweep operator (A_kk <- -1/A_kk, A_ik <- A_ik / A_kk, A_ij <- A_ij - A_ik A_kj / A_kk keeps A symmetric and
      // ends at -A^-1), one staged column per step instead of a pass over the whole matrix through L2
      const int npair = (ff + 1) >> 1;
      for (int k = 0; k < ff; ++k) {
        __syncthreads();
        for (int t = tid; t < ff; t += NT) colk[t] = (t >= k) ? Pk[t * (t + 1) / 2 + k] : Pk[k * (k + 1) / 2 + t];
        __syncthreads();
        const cplx p = cinv(colk[k]);
        for (int pr = warp; pr < npair; pr += NW) {      // rows in pairs (r, ff - 1 - r): equal work per warp
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const int i = half ? ff - 1 - pr : pr;
            if (half && i == pr) continue;
            cplx* row = Pk + i * (i + 1) / 2;
            if (i == k) {
              for (int j = lane; j <= i; j += 32) row[j] = (j == k) ? make_double2(-p.x, -p.y) : cmul(colk[j], p);
            } else {
              const cplx f = cmul(colk[i], p);
              for (int j = lane; j <= i; j += 32) row[j] = (j == k) ? f : cfnma(f, colk[j], row[j]);
            }
          }
        }
      }
      __syncthreads();
      for (int i = warp; i < ff; i += NW)
        for (int j = lane; j < ff; j += 32) {
          const int a = max(i, j), b2 = min(i, j);
          const cplx v = Pk[a * (a + 1) / 2 + b2];
          Pm[i * ld + j] = make_double2(-v.x, -v.y);
        }
    } else
    for (int k = 0; k < ff; ++k) {
      __syncthreads();
      for (int t = tid; t < ff; t += NT) { rowk[t] = Pm[k * ld + t]; colk[t] = Pm[t * ld + k]; }
      __syncthreads();
      const cplx p = cinv(rowk[k]);
      for (int i = warp; i < ff; i += NW) {
        const cplx f = cmul(colk[i], p);
        if (i == k) {
          for (int j = lane; j < ff; j += 32) Pm[i * ld + j] = (j == k) ? p : cmul(rowk[j], p);
        } else {
          for (int j = lane; j < ff; j += 32)
            Pm[i * ld + j] = (j == k) ? make_double2(-f.x, -f.y) : cfnma(f, rowk[j], Pm[i * ld + j]);
        }
      }
    }
    __syncthreads();
    // ---- S = Hq (I - X^2) + 2Q + 2 (Hq X) P (X Q) ----
    RK_PHASE(17);
    for (int i = warp; i < ff; i += NW) {
      cplx acc[NCOL];
#pragma unroll
      for (int c = 0; c < NCOL; ++c) acc[c] = make_double2(0.0, 0.0);
      const cplx* hrow = Hq + (size_t)i * ff;
      for (int k = 0; k < ff; ++k) {
        const cplx f = cmul(hrow[k], Xv[k]);
#pragma unroll
        for (int c = 0; c < NCOL; ++c) {
          const int j = lane + 32 * c;
          if (j < ff) acc[c] = cfma(f, Pm[k * ld + j], acc[c]);   // (reading the packed P from shared memory instead: slower, gather)
        }
      }
#pragma unroll
      for (int c = 0; c < NCOL; ++c) {
        const int j = lane + 32 * c;
        if (j < ff) {
          cplx s = cmul(acc[c], XQ[j]);
          s = make_double2(2.0 * s.x, 2.0 * s.y);
          s = cfma(hrow[j], omx2[j], s);
          if (i == j) { s.x += 2.0 * qv[i].x; s.y += 2.0 * qv[i].y; }
          Sm[i * ld + j] = s;
        }
      }
    }
    RK_PHASE(18);
    for (int t = tid; t < ff; t += NT) bvec[t] = cscale(r0, wr[2 * ff + t]);
    // ---- P_r S = L U with partial pivoting; multipliers kept below the diagonal, pivots in ipiv ----
    // S does not fit in shared memory at the sizes of this variant, but its trailing block does after a few steps: from step
    // k0s on the block [k0s, ff)^2 is factorised in shared memory (the storage of the packed G_II, free by now) and written back
    // at the end; the L columns left of it stay in L2 and only take part in the row swaps.
    const int k0s = G.packed ? G.lu_k0 : ff;
    const int ldt = ff - k0s;
    cplx* Ts = reinterpret_cast<cplx*>(mat0);
    bool staged = false;
#define PIRL_SEL(i, j) (*((staged && (j) >= k0s) ? &Ts[((i) - k0s) * ldt + ((j) - k0s)] : &Sm[(i) * ld + (j)]))
    for (int k = 0; k < ff; ++k) {
      __syncthreads();
      if (k == k0s) {
        for (int i = k0s + warp; i < ff; i += NW)
          for (int j = k0s + lane; j < ff; j += 32) Ts[(i - k0s) * ldt + (j - k0s)] = Sm[i * ld + j];
        staged = true;
        __syncthreads();
      }
      if (warp == 0) {
        double best = -1.0; int bi = k;
        for (int i = k + lane; i < ff; i += 32) {
          cplx a = PIRL_SEL(i, k);
          double mg = fabs(a.x) + fabs(a.y);
          if (mg > best) { best = mg; bi = i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          double ob = __shfl_xor_sync(0xffffffffu, best, o);
          int oi = __shfl_xor_sync(0xffffffffu, bi, o);
          if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (lane == 0) ipiv[k] = bi;
      }
      __syncthreads();
      const int pv = ipiv[k];
      if (pv != k) {
        for (int j = tid; j < ff; j += NT) { cplx t = PIRL_SEL(k, j); PIRL_SEL(k, j) = PIRL_SEL(pv, j); PIRL_SEL(pv, j) = t; }
        if (tid == 0) { cplx t = bvec[k]; bvec[k] = bvec[pv]; bvec[pv] = t; }
        __syncthreads();
      }
      const cplx pinv = cinv(PIRL_SEL(k, k));
      const cplx bk = bvec[k];
      for (int i = k + 1 + warp; i < ff; i += NW) {
        const cplx f = cmul(PIRL_SEL(i, k), pinv);
        for (int j = k + 1 + lane; j < ff; j += 32) PIRL_SEL(i, j) = cfnma(f, PIRL_SEL(k, j), PIRL_SEL(i, j));
        if (lane == 0) { bvec[i] = cfnma(f, bk, bvec[i]); PIRL_SEL(i, k) = f; }
      }
    }
#undef PIRL_SEL
    if (staged) {
      __syncthreads();
      for (int i = k0s + warp; i < ff; i += NW)
        for (int j = k0s + lane; j < ff; j += 32) Sm[i * ld + j] = Ts[(i - k0s) * ldt + (j - k0s)];
    }
    }
    __syncthreads();
    RK_PHASE(12);
    for (int t = tid; t < ff; t += NT) colk[t] = cinv(Sm[t * ld + t]);   // 1 / U_kk, also used by the U^H solve of the reverse sweep
    __syncthreads();
    if (warp == 0) {   // back substitution U c+ = b'
      cplx xr[NCOL];
#pragma unroll
      for (int c = 0; c < NCOL; ++c) xr[c] = bvec[min(lane + 32 * c, ff - 1)];
      warp_trisolve<NCOL, true, false, false, false>(xr, Sm, ld, colk, ff, lane);
#pragma unroll
      for (int c = 0; c < NCOL; ++c) if (lane + 32 * c < ff) bvec[lane + 32 * c] = xr[c];
    }
    __syncthreads();
    // ---- y = Q X c+,  d = 2 P y ----
    RK_PHASE(19);
    for (int t = tid; t < ff; t += NT) yv[t] = cmul(XQ[t], bvec[t]);
    __syncthreads();
    cta_symv<NT>(Pm, ld, yv, dv, ff);
    __syncthreads();
    for (int t = tid; t < ff; t += NT) {
      dv[t] = make_double2(2.0 * dv[t].x, 2.0 * dv[t].y);
      v_ell[t] = make_double2(wr[t] * 0.7071067811865476, wr[ff + t] * 0.7071067811865476);   // order -1 = (cos1 + i sin1)/sqrt2
    }
    __syncthreads();
    if (warp < 3) {   // T on {cos1, sin1, const}
      cplx s = make_double2(0.0, 0.0);
      for (int j = lane; j < ff; j += 32) { double w = wr[warp * ff + j]; s.x = fma(w, dv[j].x, s.x); s.y = fma(w, dv[j].y, s.y); }
      s = warp_sum(s);
      if (lane == 0) red[warp] = s;
    }
    __syncthreads();
    const double is2 = 0.7071067811865476;
    const cplx Tc = red[0], Ts = red[1], T0 = red[2];
    const cplx Tm = make_double2((Tc.x - Ts.y) * is2, (Tc.y + Ts.x) * is2);
    const double a1II = nII * nII - sa * sa;
    const double k1II = a1II >= 0.0 ? sqrt(a1II) : 0.0;
    const double alpha = k1II / (nII * nII) * nI;
    const double eff = (Tm.x * Tm.x + Tm.y * Tm.y) * alpha;
    if (tid == 0) {
      int st = P.wstatus[b] & 0xff;
      const double det1 = (T0.x * T0.x + T0.y * T0.y) * nI / nII;
      if (!isfinite(eff) || !isfinite(det1)) st |= PIRL_ST_NONFINITE_;
      P.eff[b] = eff;
      if (P.status) P.status[b] = st;
    }
    // =============================== reverse sweep ===============================
    RK_PHASE(20);
    double* Lf = G.Lf + (size_t)b * ff * PIRL_GRAD_RANK;
    double* Rf = G.Rf + (size_t)b * ff * PIRL_GRAD_RANK;
    const cplx tb = make_double2(2.0 * alpha * Tm.x, 2.0 * alpha * Tm.y);
    for (int t = tid; t < ff; t += NT) {
      const cplx el = v_ell[t];
      v_x[t] = cmul(el, cconj(tb));                    // conj(d-bar),  d-bar = conj(ell) t-bar
      const cplx ellb = cmul(cconj(dv[t]), tb);
      Lf[t * PIRL_GRAD_RANK + 12] = wr[t];        Rf[t * PIRL_GRAD_RANK + 12] = ellb.x * is2;   // W-bar row cos1
      Lf[t * PIRL_GRAD_RANK + 13] = wr[ff + t];   Rf[t * PIRL_GRAD_RANK + 13] = ellb.y * is2;   // W-bar row sin1
      Lf[t * PIRL_GRAD_RANK + 15] = 0.0;          Rf[t * PIRL_GRAD_RANK + 15] = 0.0;
    }
    __syncthreads();
    cta_symv<NT>(Pm, ld, v_x, v_yb, ff);               // P conj(d-bar);  y-bar = 2 conj(.)
    __syncthreads();
    for (int t = tid; t < ff; t += NT) {
      const cplx yb = make_double2(2.0 * v_yb[t].x, -2.0 * v_yb[t].y);
      v_yb[t] = yb;
      const cplx cp = bvec[t];
      v_qb[t] = cmul(cconj(cmul(Xv[t], cp)), yb);
      v_Xb[t] = cmul(cconj(cmul(qv[t], cp)), yb);
      rowk[t] = cmul(cconj(XQ[t]), yb);                // c+-bar : right-hand side of S^H bb = c+-bar
    }
    __syncthreads();
    if (warp == 0) {
      // U^H w = c (forward), L^H v = w (backward, unit diagonal), bb = P_r^T v (swaps in reverse order)
      {
        cplx xr[NCOL];
#pragma unroll
        for (int c = 0; c < NCOL; ++c) xr[c] = rowk[min(lane + 32 * c, ff - 1)];
        warp_trisolve<NCOL, false, true, true, false>(xr, Sm, ld, colk, ff, lane);
        warp_trisolve<NCOL, true, true, true, true>(xr, Sm, ld, colk, ff, lane);
#pragma unroll
        for (int c = 0; c < NCOL; ++c) if (lane + 32 * c < ff) rowk[lane + 32 * c] = xr[c];
        __syncwarp();
      }
      if constexpr (REG) {        // bb[rperm[k]] = v[k]
        __syncwarp();
        for (int k = lane; k < ff; k += 32) colk[ipiv[k]] = rowk[k];
        __syncwarp();
        for (int k = lane; k < ff; k += 32) rowk[k] = colk[k];
      } else if (lane == 0) {
        for (int k = ff - 1; k >= 0; --k) {
          const int pv = ipiv[k];
          if (pv != k) { cplx t = rowk[k]; rowk[k] = rowk[pv]; rowk[pv] = t; }
        }
      }
    }
    __syncthreads();
    for (int t = tid; t < ff; t += NT) {
      const cplx bb = rowk[t];
      Lf[t * PIRL_GRAD_RANK + 14] = wr[2 * ff + t];
      Rf[t * PIRL_GRAD_RANK + 14] = bb.x * r0.x + bb.y * r0.y;    // Re(conj(bb) r0) : W-bar row const
      v_x[t] = cconj(bb);
      v_w2[t] = cscale(cmul(bvec[t], XQ[t]), 2.0);
    }
    __syncthreads();
    cta_symv<NT>(Hq, ff, v_x, v_hb, ff);               // Hq conj(bb);  hb = Hq^H bb = conj(.)
    cta_symv<NT>(Pm, ld, v_w2, v_pw2, ff);             // P w2
    __syncthreads();
    for (int t = tid; t < ff; t += NT) {
      const cplx hq = v_hb[t];                         // = conj(hb)
      v_x2[t] = cmul(Xv[t], hq);                       // conj(M1^H bb) = X conj(hb)
    }
    __syncthreads();
    cta_symv<NT>(Pm, ld, v_x2, v_pm, ff);              // pm = P conj(m1hb)
    __syncthreads();
    double* lambg = G.lamb + (size_t)b * ff;
    for (int t = tid; t < ff; t += NT) {
      const cplx cp = bvec[t], bb = rowk[t], X = Xv[t], q = qv[t];
      const cplx hb = cconj(v_hb[t]), pw2 = v_pw2[t], pm = v_pm[t], yb = v_yb[t];
      cplx qb = v_qb[t], Xb = v_Xb[t];
      const cplx ccp = cconj(cp);
      // S-bar = -bb c+^H.  Hq D1 term:
      const cplx D1b = cscale(cmul(ccp, hb), -1.0);
      Xb = cfma(cscale(cconj(X), -2.0), D1b, Xb);
      qb = cfma(cscale(bb, -2.0), ccp, qb);
      // 2 (Hq X) P (X q) term:
      const cplx xqb = cscale(cmul(ccp, cconj(pm)), -2.0);
      Xb = cfma(cconj(q), xqb, Xb);
      qb = cfma(cconj(X), xqb, qb);
      const cplx h = cfma(X, pw2, cmul(cp, omx2[t]));
      Xb = cfnma(hb, cconj(pw2), Xb);
      qb = cfma(bb, cconj(h), qb);
      // G-bar = g1a g1b^H + g2a g2b^H ;  g1a = -y-bar, g1b = d/2, g2a = conj(pm), g2b = pw2
      const cplx g1a = cscale(yb, -1.0), g1b = cscale(dv[t], 0.5), g2a = cconj(pm), g2b = pw2;
      qb = cfma(g1a, cconj(g1b), qb);
      qb = cfma(g2a, cconj(g2b), qb);
      qb = cfma(cscale(cconj(X), -k0d), Xb, qb);
      double lb = 0.0;
      if (q.y == 0.0) { if (q.x > 0.0) lb = qb.x / (2.0 * q.x); }
      else lb = -qb.y / (2.0 * q.y);
      lambg[t] = lb;
      // rank-one pairs (H, a, c): H-bar = a c^H.  (H_I: a = -bb, c = h), (H_II: g1a, g1b), (H_II: g2a, g2b)
      const cplx a1 = cscale(bb, -1.0);
      v_in[t] = cconj(a1); v_in[ff + t] = h;
      v_in[2 * ff + t] = cconj(g1a); v_in[3 * ff + t] = g1b;
      v_in[4 * ff + t] = cconj(g2a); v_in[5 * ff + t] = g2b;
      {   // adjoints of the diagonal matrices taken as full matrices (Daleckii-Krein form of the eigenvector sweep):
          // q-bar = g1a g1b^H + g2a g2b^H + bb (h - 2 c+)^H ; X-bar = -hb (P w2)^H ; (1-X^2)-bar = -hb c+^H ; (Xq)-bar = (y-bar - 2 g2a) c+^H
        cplx* V = reinterpret_cast<cplx*>(G.Vf + ((size_t)b * ff + t) * PIRL_GRAD_NV);
        V[0] = g1a; V[1] = g1b; V[2] = g2a; V[3] = g2b; V[4] = bb;
        V[5] = make_double2(h.x - 2.0 * cp.x, h.y - 2.0 * cp.y);
        V[6] = cscale(hb, -1.0); V[7] = pw2; V[8] = cp;
        V[9] = make_double2(yb.x - 2.0 * g2a.x, yb.y - 2.0 * g2a.y);
      }
      double* R = Rf + t * PIRL_GRAD_RANK;
      R[0] = h.x;    R[1] = h.y;    R[2] = a1.x;   R[3] = -a1.y;
      R[4] = g1b.x;  R[5] = g1b.y;  R[6] = g1a.x;  R[7] = -g1a.y;
      R[8] = g2b.x;  R[9] = g2b.y;  R[10] = g2a.x; R[11] = -g2a.y;
    }
    __syncthreads();
    RK_PHASE(21);
    // six products H x -> Lf columns (Re, -Im);  H_I = Hq + Q.  One pass over the rows for all six vectors.
    for (int i = warp; i < ff; i += NW) {
      cplx s6[6];
#pragma unroll
      for (int v = 0; v < 6; ++v) s6[v] = make_double2(0.0, 0.0);
      for (int j = lane; j < ff; j += 32) {
        const cplx m1 = Hq[(size_t)i * ff + j], m2 = HII[(size_t)i * ff + j];
#pragma unroll
        for (int v = 0; v < 6; ++v) s6[v] = cfma(v < 2 ? m1 : m2, v_in[(size_t)v * ff + j], s6[v]);
      }
#pragma unroll
      for (int v = 0; v < 6; ++v) s6[v] = warp_sum(s6[v]);
      if (lane == 0) {
        s6[0] = cfma(qv[i], v_in[i], s6[0]);
        s6[1] = cfma(qv[i], v_in[ff + i], s6[1]);
#pragma unroll
        for (int v = 0; v < 6; ++v) { Lf[i * PIRL_GRAD_RANK + 2 * v] = s6[v].x; Lf[i * PIRL_GRAD_RANK + 2 * v + 1] = -s6[v].y; }
      }
    }
    RK_PHASE(22);
  }
}

// ===================================================================================================
// CTA-wide real GEMM on ff x ff matrices (shared memory or L2-resident scratch): C = A * B or A * B^T.
// 16 x 16 threads, 6 x 6 register tile per thread (rows ty + 16 r, columns tx + 16 c) over 96 x 96 blocks of C.
// C must not alias A or B.
// ===================================================================================================
template <bool TRANS_B, int RT>
__device__ __forceinline__ void cta_gemm_rt(double* __restrict__ C, const double* __restrict__ A, const double* __restrict__ B,
                                            int ff, int ld) {
  constexpr int BLK = 16 * RT;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  for (int i0 = 0; i0 < ff; i0 += BLK)
    for (int j0 = 0; j0 < ff; j0 += BLK) {
      double acc[RT][RT];
      int ri[RT], cj[RT];
#pragma unroll
      for (int r = 0; r < RT; ++r) {
        ri[r] = min(i0 + ty + 16 * r, ff - 1) * ld;
        cj[r] = min(j0 + tx + 16 * r, ff - 1);
        if (TRANS_B) cj[r] *= ld;
#pragma unroll
        for (int c = 0; c < RT; ++c) acc[r][c] = 0.0;
      }
      for (int k = 0; k < ff; ++k) {
        double a[RT], bq[RT];
#pragma unroll
        for (int r = 0; r < RT; ++r) {
          a[r] = A[ri[r] + k];
          bq[r] = TRANS_B ? B[cj[r] + k] : B[k * ld + cj[r]];
        }
#pragma unroll
        for (int r = 0; r < RT; ++r)
#pragma unroll
          for (int c = 0; c < RT; ++c) acc[r][c] = fma(a[r], bq[c], acc[r][c]);
      }
#pragma unroll
      for (int r = 0; r < RT; ++r) {
        const int i = i0 + ty + 16 * r;
#pragma unroll
        for (int c = 0; c < RT; ++c) {
          const int j = j0 + tx + 16 * c;
          if (i < ff && j < ff) C[i * ld + j] = acc[r][c];
        }
      }
    }
}
// The same GEMM on the FP64 tensor path: mma.sync.aligned.m8n8k4.f64 (legacy warp-level DMMA; tcgen05 has no FP64 kind).
// 8 warps; C is cut in 8 x 8 tiles, a warp owns blocks of 3 x 3 tiles (ff = 81: 11 x 11 tiles -> 16 blocks, two per warp)
// and walks k in steps of 4: three A fragments + three B fragments (one double per lane each) feed nine DMMAs.
// Fragment layout (PTX ISA, m8n8k4 .f64): A[row = lane / 4][k = lane % 4], B[k = lane % 4][col = lane / 4],
// C[row = lane / 4][col = 2 (lane % 4) + {0, 1}].  Out-of-range rows / columns / k read as zero.
template <bool TRANS_B>
__device__ __forceinline__ void cta_gemm_mma(double* __restrict__ C, const double* __restrict__ A, const double* __restrict__ B,
                                             int ff, int ld) {
  constexpr int TB = 3;                              // tiles per block side
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int g = lane >> 2, t4 = lane & 3;
  const int nt = (ff + 7) / 8, nb = (nt + TB - 1) / TB;
  for (int blk = warp; blk < nb * nb; blk += nwarps) {
    const int bi = blk / nb, bj = blk - bi * nb;
    double acc[TB][TB][2];
    int arow[TB], bcol[TB];
    bool aok[TB], bok[TB];
#pragma unroll
    for (int r = 0; r < TB; ++r) {
      const int i = (bi * TB + r) * 8 + g, j = (bj * TB + r) * 8 + g;
      aok[r] = i < ff; bok[r] = j < ff;
      arow[r] = min(i, ff - 1) * ld;
      bcol[r] = TRANS_B ? min(j, ff - 1) * ld : min(j, ff - 1);
#pragma unroll
      for (int c = 0; c < TB; ++c) { acc[r][c][0] = 0.0; acc[r][c][1] = 0.0; }
    }
    for (int k0 = 0; k0 < ff; k0 += 4) {
      const int k = k0 + t4;
      const bool kok = k < ff;
      const int kc = min(k, ff - 1);
      double a[TB], b[TB];
#pragma unroll
      for (int r = 0; r < TB; ++r) {
        const double av = A[arow[r] + kc];
        const double bv = TRANS_B ? B[bcol[r] + kc] : B[kc * ld + bcol[r]];
        a[r] = (aok[r] && kok) ? av : 0.0;
        b[r] = (bok[r] && kok) ? bv : 0.0;
      }
#pragma unroll
      for (int r = 0; r < TB; ++r)
#pragma unroll
        for (int c = 0; c < TB; ++c)
          asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                       : "+d"(acc[r][c][0]), "+d"(acc[r][c][1])
                       : "d"(a[r]), "d"(b[c]));
    }
#pragma unroll
    for (int r = 0; r < TB; ++r) {
      const int i = (bi * TB + r) * 8 + g;
#pragma unroll
      for (int c = 0; c < TB; ++c) {
        const int j = (bj * TB + c) * 8 + 2 * t4;
        if (i < ff && j < ff) C[i * ld + j] = acc[r][c][0];
        if (i < ff && j + 1 < ff) C[i * ld + j + 1] = acc[r][c][1];
      }
    }
  }
}

// register tile 6 x 6 (96 x 96 per pass) up to ff = 96, 8 x 8 (128 x 128 per pass) beyond: ff = 121 is one pass instead of four
template <bool TRANS_B>
__device__ __forceinline__ void cta_gemm(double* __restrict__ C, const double* __restrict__ A, const double* __restrict__ B,
                                         int ff, int ld, int use_mma = 0) {
  if (use_mma) cta_gemm_mma<TRANS_B>(C, A, B, ff, ld);
  else if (ff <= 96) cta_gemm_rt<TRANS_B, 6>(C, A, B, ff, ld);
  else cta_gemm_rt<TRANS_B, 8>(C, A, B, ff, ld);
}

// q = sqrt(lambda) (principal branch), X = exp(-k0 d q)
__device__ __forceinline__ void mode_qx(double l, double k0d, cplx& q, cplx& X) {
  if (l >= 0.0) { const double s = sqrt(l); q = make_double2(s, 0.0); X = make_double2(exp(-k0d * s), 0.0); }
  else { const double s = sqrt(-l); q = make_double2(0.0, s); double sn, cs; sincos(k0d * s, &sn, &cs); X = make_double2(cs, -sn); }
}
// (exp(z) - 1)/z for Re z <= 0
__device__ __forceinline__ cplx cexpm1_over_z(cplx z) {
  if (z.x * z.x + z.y * z.y < 1e-4) {
    cplx r = make_double2(1.0 / 720.0, 0.0);
    r = cfma(r, z, make_double2(1.0 / 120.0, 0.0));
    r = cfma(r, z, make_double2(1.0 / 24.0, 0.0));
    r = cfma(r, z, make_double2(1.0 / 6.0, 0.0));
    r = cfma(r, z, make_double2(0.5, 0.0));
    r = cfma(r, z, make_double2(1.0, 0.0));
    return r;
  }
  double sn, cs;
  sincos(z.y, &sn, &cs);
  const double e = exp(z.x);
  return cmul(make_double2(e * cs - 1.0, e * sn), cinv(z));
}

// adjoint of basis_toeplitz: accumulate sgn * Tb[i][j] * dT_ij/dc into (gr, gi) (shared, atomics)
__device__ __forceinline__ void toeplitz_adjoint_add(int i, int j, int N, double v, double* gr, double* gi) {
  const int z = 2 * N;
  const double r2 = 1.4142135623730951;
  if (i == z && j == z) { atomicAdd(&gr[0], v); return; }
  if (i == z || j == z) {
    const int o = (i == z) ? j : i;
    const int p = (o >> 1) + 1;
    if (o & 1) atomicAdd(&gi[p], -r2 * v); else atomicAdd(&gr[p], r2 * v);
    return;
  }
  const int p = (i >> 1) + 1, q = (j >> 1) + 1;
  const int d = p - q, ad = d < 0 ? -d : d;
  const double sg = d >= 0 ? 1.0 : -1.0;
  const int si = i & 1, sj = j & 1;
  if (!si && !sj) { atomicAdd(&gr[ad], v); atomicAdd(&gr[p + q], v); }
  else if (si && sj) { atomicAdd(&gr[ad], v); atomicAdd(&gr[p + q], -v); }
  else if (!si && sj) { atomicAdd(&gi[p + q], -v); atomicAdd(&gi[ad], sg * v); }
  else { atomicAdd(&gi[p + q], -v); atomicAdd(&gi[ad], -sg * v); }
}

// ===================================================================================================
// g_eigback : reverse sweep through the generalised symmetric eigenproblem, B, E^-1, the Toeplitz maps and the DFT
// ===================================================================================================
template <int NT, bool SM_ALL>
__global__ void __launch_bounds__(NT) g_eigback(KParams P, GParams G) {
  static_assert(NT == 256, "cta_gemm assumes 16 x 16 threads");
  extern __shared__ __align__(16) double sm[];
  const int ff = P.ff, ld = P.ld, N = P.N, z = 2 * N, nc = P.n_cells;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = NT / 32;
  double* lam = sm;
  double* lamb = lam + ff;
  double* dpm = lamb + ff;
  double* gEr = dpm + 5 * ff;   // (four spare vectors)
  double* gEi = gEr + ff;
  double* gAr = gEi + ff;
  double* gAi = gAr + ff;
  cplx* qS = reinterpret_cast<cplx*>(gAi + ff + (ff & 1));   // ff odd: keep the complex vectors 16-byte aligned
  cplx* XS = qS + ff;
  uint8_t* su = reinterpret_cast<uint8_t*>(XS + ff);
  double* mat0 = sm + grad_eig_vec_doubles(ff, nc);
  const size_t msz = (size_t)ff * ld;
  double* scr = G.gscratch + (size_t)blockIdx.x * G.gstride;
  double* M[4];
  for (int m = 0; m < 4; ++m) M[m] = (SM_ALL || m < G.smem_mats) ? mat0 + m * msz : scr + m * msz;   // SM_ALL: plain shared pointers
  double *M0 = M[0], *M1 = M[1], *M2 = M[2], *M3 = M[3];

  for (int b = blockIdx.x; b < P.B; b += gridDim.x) {
    __syncthreads();
    const double* pr = P.params + (size_t)b * P.pstride;
    const double ang = pr[1], nr = pr[3];
    const double sa = sin(ang * (3.14159265358979323846 / 180.0));
    const double er = nr * nr, eg = P.eps_g;
    long long pc_ = clock64();
    for (int t = tid; t < ff; t += NT) {
      lam[t] = P.lam[(size_t)b * ff + t];
      lamb[t] = G.lamb[(size_t)b * ff + t];
      gEr[t] = 0.0; gEi[t] = 0.0; gAr[t] = 0.0; gAi[t] = 0.0;
    }
    for (int p = tid; p < nc; p += NT) su[p] = P.patterns[(size_t)b * nc + p] > 0;
    __syncthreads();
    // ---- M0 = W ; M1 = G_B ; M2 = G_A  (adjoints of B~ = W^T B W and A~ = W^T A W) ----
    // dW = W Cm with Cm_ij = (dB~_ij - lam_j dA~_ij)/(lam_j - lam_i): the only singular combination is
    // K_ij = (Pm_ij - Pm_ji)/(lam_j - lam_i).  The layer solve is invariant under W -> W R, diag(f(lam)) -> R^T diag(f) R,
    // which gives K_ij = sum_f Re[conj(Phibar_f,ij + Phibar_f,ji) (f_j - f_i)/(lam_j - lam_i)] with stable divided differences
    // (tools/proto_grad.py).  G_B = K/2 + diag(lambda-bar);  G_A = -(sym(Pm) + (lam_i + lam_j) K / 2)/2, diag -(lam lamb + Pm_jj/2).
    const double* Wg = P.Wbuf + (size_t)b * ff * ff;
    const double* Lf = G.Lf + (size_t)b * ff * PIRL_GRAD_RANK;
    const double* Rf = G.Rf + (size_t)b * ff * PIRL_GRAD_RANK;
    const double* Vf = G.Vf + (size_t)b * ff * PIRL_GRAD_NV;
    const double k0d = 2.0 * 3.14159265358979323846 / pr[0] * pr[2];
    constexpr int NVEC = 2 * PIRL_GRAD_RANK + PIRL_GRAD_NV;   // 52 doubles per mode
    for (int t = tid; t < ff; t += NT) mode_qx(lam[t], k0d, qS[t], XS[t]);
    // element (mode m, component r) of the factor arrays: base[m * sm_ + r * sr_].  Staged copies are component-major
    // (consecutive modes in consecutive banks); the global originals are mode-major.
    int fsm = PIRL_GRAD_RANK, fsr = 1, vsm = PIRL_GRAD_NV, vsr = 1;
    if (ld >= NVEC) {   // stage the per-mode vectors in M3's storage (free until the first GEMM)
      for (int idx = tid; idx < ff * PIRL_GRAD_RANK; idx += NT) {
        const int m = idx / PIRL_GRAD_RANK, r = idx - m * PIRL_GRAD_RANK;
        M3[r * ff + m] = Lf[idx];
        M3[ff * PIRL_GRAD_RANK + r * ff + m] = Rf[idx];
      }
      for (int idx = tid; idx < ff * PIRL_GRAD_NV; idx += NT) {
        const int m = idx / PIRL_GRAD_NV, r = idx - m * PIRL_GRAD_NV;
        M3[2 * ff * PIRL_GRAD_RANK + r * ff + m] = Vf[idx];
      }
      Lf = M3; Rf = M3 + ff * PIRL_GRAD_RANK; Vf = M3 + 2 * ff * PIRL_GRAD_RANK;
      fsm = 1; fsr = ff; vsm = 1; vsr = ff;
    }
    __syncthreads();
    for (int i = warp; i < ff; i += NW) {
      double li[PIRL_GRAD_RANK], ri[PIRL_GRAD_RANK];
      cplx vi[PIRL_GRAD_NV / 2];
#pragma unroll
      for (int r = 0; r < PIRL_GRAD_RANK; ++r) { li[r] = Lf[i * fsm + r * fsr]; ri[r] = Rf[i * fsm + r * fsr]; }
#pragma unroll
      for (int r = 0; r < PIRL_GRAD_NV / 2; ++r) vi[r] = make_double2(Vf[i * vsm + 2 * r * vsr], Vf[i * vsm + (2 * r + 1) * vsr]);
      const double lmi = lam[i];
      const cplx qi = qS[i], Xi = XS[i];
      for (int j = lane; j < ff; j += 32) M0[i * ld + j] = Wg[i * ff + j];
      // G_B and G_A are symmetric (K_ij = K_ji): the diagonal and the pairs j > i are computed, the mirror image is stored too
      for (int j = i + lane; j < ff; j += 32) {
        double pij = 0.0, pji = 0.0;
#pragma unroll
        for (int r = 0; r < PIRL_GRAD_RANK; ++r) {
          pij = fma(li[r], Rf[j * fsm + r * fsr], pij);
          pji = fma(ri[r], Lf[j * fsm + r * fsr], pji);
        }
        const double lmj = lam[j];
        if (i == j) {
          M1[i * ld + j] = lamb[i];
          M2[i * ld + j] = -fma(lmi, lamb[i], 0.5 * pij);
          continue;
        }
        cplx vj[PIRL_GRAD_NV / 2];   // (the staged copy is only 8-byte aligned: ff * ld is odd)
#pragma unroll
        for (int r = 0; r < PIRL_GRAD_NV / 2; ++r) vj[r] = make_double2(Vf[j * vsm + 2 * r * vsr], Vf[j * vsm + (2 * r + 1) * vsr]);
        const cplx qj = qS[j], Xj = XS[j];
        // S_f,ij = sum over pairs a_i conj(c_j) + a_j conj(c_i)
        cplx Sq = make_double2(0.0, 0.0);
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          Sq = cfma(vi[2 * r], cconj(vj[2 * r + 1]), Sq);
          Sq = cfma(vj[2 * r], cconj(vi[2 * r + 1]), Sq);
        }
        const cplx mhi = vi[6], mhj = vj[6];
        const cplx SX = cfma(mhi, cconj(vj[7]), cmul(mhj, cconj(vi[7])));
        const cplx SD = cfma(mhi, cconj(vj[8]), cmul(mhj, cconj(vi[8])));
        const cplx SXQ = cfma(vi[9], cconj(vj[8]), cmul(vj[9], cconj(vi[8])));
        // divided differences (f_j - f_i)/(lam_j - lam_i)
        const cplx dq = cinv(cadd(qi, qj));
        cplx dX;
        const double dl = lmj - lmi;
        if (fabs(dl) > 1e-2 * (fabs(lmi) + fabs(lmj))) {   // well separated: plain difference quotient (loses <= 2 digits)
          const double idl = 1.0 / dl;
          dX = make_double2((Xj.x - Xi.x) * idl, (Xj.y - Xi.y) * idl);
        } else {                                            // close pair: X_ref (exp(z) - 1)/z with Re z <= 0
          cplx zz = make_double2(-k0d * (qj.x - qi.x), -k0d * (qj.y - qi.y));
          cplx Xref = Xi;
          if (zz.x > 0.0) { zz = make_double2(-zz.x, -zz.y); Xref = Xj; }
          dX = cscale(cmul(cmul(Xref, cexpm1_over_z(zz)), dq), -k0d);
        }
        const cplx dD1 = cscale(cmul(cadd(Xi, Xj), dX), -1.0);
        const cplx dXQ = cfma(Xj, dq, cmul(qi, dX));
        cplx acc = cmul(cconj(Sq), dq);
        acc = cfma(cconj(SX), dX, acc);
        acc = cfma(cconj(SD), dD1, acc);
        acc = cfma(cconj(SXQ), dXQ, acc);
        const double K = acc.x;
        const double gb = 0.5 * K, ga = -0.5 * fma(0.5 * (lmi + lmj), K, 0.5 * (pij + pji));
        M1[i * ld + j] = gb; M1[j * ld + i] = gb;
        M2[i * ld + j] = ga; M2[j * ld + i] = ga;
      }
    }
    __syncthreads();
    RK_PHASE(24);
    cta_gemm<false>(M3, M0, M1, ff, ld, G.use_mma);                 // W G_B
    __syncthreads();
    cta_gemm<true>(M1, M3, M0, ff, ld, G.use_mma);                  // B-bar = W G_B W^T
    __syncthreads();
    cta_gemm<false>(M3, M0, M2, ff, ld, G.use_mma);                 // W G_A
    __syncthreads();
    cta_gemm<true>(M2, M3, M0, ff, ld, G.use_mma);                  // A-bar = W G_A W^T
    __syncthreads();
    RK_PHASE(25);
    for (int i = warp; i < ff; i += NW)
      for (int j = lane; j < ff; j += 32) toeplitz_adjoint_add(i, j, N, M2[i * ld + j], gAr, gAi);
    RK_PHASE(26);
    // ---- M0 = E^-1, exported by the front kernel of this call (P.Einv) ----
    {
      const double* Eg = P.Einv + (size_t)b * ff * ff;
      for (int i = warp; i < ff; i += NW)
        for (int j = lane; j < ff; j += 32) M0[i * ld + j] = Eg[i * ff + j];
    }
    RK_PHASE(27);
    // ---- M1 = (E^-1)-bar = S_K B-bar S_K^T  (2x2 cos/sin blocks swap and change sign; constant row / column drop out) ----
    for (int idx = tid; idx < N * N; idx += NT) {
      const int pp = idx / N, qq = idx - pp * N;
      const int ic = 2 * pp, is = ic + 1, jc = 2 * qq, js = jc + 1;
      const double kk = ((pp + 1) * sa) * ((qq + 1) * sa);
      M3[is * ld + js] = kk * M1[ic * ld + jc];
      M3[is * ld + jc] = -kk * M1[ic * ld + js];
      M3[ic * ld + js] = -kk * M1[is * ld + jc];
      M3[ic * ld + jc] = kk * M1[is * ld + js];
    }
    for (int t = tid; t < ff; t += NT) { M3[z * ld + t] = 0.0; M3[t * ld + z] = 0.0; }
    __syncthreads();
    cta_gemm<false>(M2, M0, M3, ff, ld, G.use_mma);
    __syncthreads();
    cta_gemm<false>(M1, M2, M0, ff, ld, G.use_mma);                 // -E-bar = E^-1 (E^-1)-bar E^-1
    __syncthreads();
    RK_PHASE(28);
    for (int i = warp; i < ff; i += NW)
      for (int j = lane; j < ff; j += 32) toeplitz_adjoint_add(i, j, N, -M1[i * ld + j], gEr, gEi);
    RK_PHASE(29);
    __syncthreads();
    // ---- g_k = conj(c-bar_k) D_k ; pixel gradient = Re sum_k g_k exp(-2 pi i k p / n) ----
    for (int k = tid; k <= 2 * N; k += NT) {
      const cplx d = __ldg(&P.Dk[k]);
      const cplx e = cmul(make_double2(gEr[k], -gEi[k]), d), a = cmul(make_double2(gAr[k], -gAi[k]), d);
      gEr[k] = e.x; gEi[k] = e.y; gAr[k] = a.x; gAi[k] = a.y;
    }
    __syncthreads();
    for (int p = tid; p < nc; p += NT) {
      double vE = 0.0, vA = 0.0;
      for (int k = 0; k <= 2 * N; ++k) {
        const cplx t = __ldg(&P.tw[(k * p) % nc]);
        vE += gEr[k] * t.x - gEi[k] * t.y;
        vA += gAr[k] * t.x - gAi[k] * t.y;
      }
      const double ep = su[p] ? er : eg;
      G.grad[(size_t)b * nc + p] = 0.5 * (er - eg) * (vE - vA / (ep * ep));
    }
    RK_PHASE(30);
  }
}
