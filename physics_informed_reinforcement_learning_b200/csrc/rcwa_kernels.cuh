// rcwa_kernels.cuh -- sm_100a kernels of the batched 1D TM RCWA path (one CTA per candidate structure).
//
// Math (validated on CPU by tools/proto_kernels.py against the oracle):
//   the reference's complex problem (_base.py:150-159, transfer_method.py:42-75) is carried into the real
//   basis {sqrt2 cos(pGx), sqrt2 sin(pGx), 1}.  At normal incidence (theta=0, the only case DeflectorGym
//   uses) the Toeplitz matrices E, A=E^- and B = Kx E^-1 Kx - I become REAL SYMMETRIC there, so the modes
//   come from a real generalised symmetric-definite eigenproblem  B w = lambda A w,  W^T A W = I.
//   Pipeline:  k_front   pattern -> Fourier coefficients -> E, A -> E^-1 -> B -> LDL^T(A) -> C = L^-1 B L^-T
//              k_tridiag C -> (d, e) by Householder,  M0 = L^-T Q_H
//              k_ql      implicit QL on (d,e), one thread per structure, emits plane rotations + eigenvalues
//              k_apply   W = M0 * rotations
//              k_tail    admittance-free layer solve in the mode basis -> R, T -> efficiencies
#pragma once
#include "rcwa_common.cuh"

// ===================================================================================================
// k_front : pattern -> C (standard-form matrix) and L (factor of A)
//   replaces convolution_matrix.py:89-114 (x2), _base.py:151-153 and the reduction of _base.py:155
// ===================================================================================================
template <int NT>
__global__ void __launch_bounds__(NT) k_front(KParams P) {
  extern __shared__ __align__(16) double sm[];
  const int ff = P.ff, ld = P.ld, N = P.N, z = 2 * N, nc = P.n_cells;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = NT / 32;
  double* cre = sm;
  double* cim = cre + (2 * N + 1);
  double* rowk = cim + (2 * N + 1);
  double* colk = rowk + ff;
  double* dinv = colk + ff;
  uint8_t* su = reinterpret_cast<uint8_t*>(dinv + ff);
  double* mat0 = sm + front_vec_doubles(ff, N, nc);
  const size_t msz = (size_t)ff * ld;
  double* scr = P.scratch + (size_t)blockIdx.x * P.scratch_stride;
  double* E = (P.smem_mats >= 1) ? mat0 : scr;
  double* A = (P.smem_mats >= 2) ? mat0 + msz : ((P.smem_mats >= 1) ? scr : scr + msz);

  __shared__ int sbad;
  for (int b = blockIdx.x; b < P.B; b += gridDim.x) {
    if (tid == 0) sbad = 0;
    __syncthreads();
    const double* pr = P.params + (size_t)b * P.pstride;
    const double ang = pr[1], nr = pr[3];
    const double sa = sin(ang * (3.14159265358979323846 / 180.0));
    const double er = nr * nr, eg = P.eps_g;
    int bad = 0;
    for (int p = tid; p < nc; p += NT) su[p] = P.patterns[(size_t)b * nc + p] > 0;
    __syncthreads();
    // ---- Fourier coefficients c_k, k = 0..2N, of the up-sampled pattern (closed form) ----
    for (int k = warp; k <= 2 * N; k += NW) {
      double sx = 0.0, sy = 0.0;
      for (int p = lane; p < nc; p += 32) {
        if (su[p]) {
          cplx t = __ldg(&P.tw[(k * p) % nc]);
          sx += t.x; sy += t.y;
        }
      }
      sx = warp_sum(sx); sy = warp_sum(sy);
      if (lane == 0) {
        cplx d = __ldg(&P.Dk[k]);
        cre[k] = sx * d.x - sy * d.y;
        cim[k] = sx * d.y + sy * d.x;
      }
    }
    __syncthreads();
    if (P.pol == 0) {
      // TE (_base.py:142-148): A = Kx^2 - E is real symmetric in the real basis -> a STANDARD symmetric eigenproblem:
      // C = Kx^2 - E goes straight to the tridiagonalisation, the factor is the identity (W orthonormal, V = W Q)
      const double de = er - eg;
      double* Cg = P.Cbuf + (size_t)b * ff * ff;
      double* Lg = P.Lbuf + (size_t)b * ff * ff;
      for (int i = warp; i < ff; i += NW) {
        const double kap = (i < z) ? ((i >> 1) + 1) * sa : 0.0;
        for (int j = lane; j < ff; j += 32) {
          const double dl = (i == j) ? 1.0 : 0.0;
          Cg[i * ff + j] = kap * kap * dl - fma(de, basis_toeplitz(i, j, N, cre, cim), eg * dl);
          Lg[i * ff + j] = dl;
        }
      }
      if (tid == 0) P.wstatus[b] = 0;
      continue;
    }
    // ---- E = eps_g I + (eps_r - eps_g) T(c),  A = I/eps_g + (1/eps_r - 1/eps_g) T(c) (real basis) ----
    {
      const double de = er - eg, da = 1.0 / er - 1.0 / eg, ia = 1.0 / eg;
      for (int i = warp; i < ff; i += NW)
        for (int j = lane; j < ff; j += 32) {
          double t = basis_toeplitz(i, j, N, cre, cim);
          double dl = (i == j) ? 1.0 : 0.0;
          E[i * ld + j] = fma(de, t, eg * dl);
          A[i * ld + j] = fma(da, t, ia * dl);
        }
    }
    // ---- E <- E^-1 : in-place Gauss-Jordan, no pivoting (E is symmetric positive definite) ----
    for (int k = 0; k < ff; ++k) {
      __syncthreads();
      for (int t = tid; t < ff; t += NT) { rowk[t] = E[k * ld + t]; colk[t] = E[t * ld + k]; }
      __syncthreads();
      const double piv = rowk[k];
      if (!(piv > 0.0)) bad = 1;
      const double p = 1.0 / piv;
      for (int i = warp; i < ff; i += NW) {
        const double f = colk[i] * p;
        if (i == k) {
          for (int j = lane; j < ff; j += 32) E[i * ld + j] = (j == k) ? p : rowk[j] * p;
        } else {
          for (int j = lane; j < ff; j += 32) E[i * ld + j] = (j == k) ? -f : fma(-f, rowk[j], E[i * ld + j]);
        }
      }
    }
    __syncthreads();
    if (P.Einv) {
      double* Eg = P.Einv + (size_t)b * ff * ff;
      for (int i = warp; i < ff; i += NW)
        for (int j = lane; j < ff; j += 32) Eg[i * ff + j] = E[i * ld + j];
      __syncthreads();
    }
    // ---- B = S_K^T E^-1 S_K - I, in place (2x2 cos/sin blocks swap and change sign) ----
    for (int idx = tid; idx < N * N; idx += NT) {
      int pp = idx / N, qq = idx - pp * N;
      int ic = 2 * pp, is = ic + 1, jc = 2 * qq, js = jc + 1;
      double e00 = E[ic * ld + jc], e01 = E[ic * ld + js], e10 = E[is * ld + jc], e11 = E[is * ld + js];
      double kk = ((pp + 1) * sa) * ((qq + 1) * sa);
      double dl = (pp == qq) ? 1.0 : 0.0;
      E[ic * ld + jc] = kk * e11 - dl;
      E[ic * ld + js] = -kk * e10;
      E[is * ld + jc] = -kk * e01;
      E[is * ld + js] = kk * e00 - dl;
    }
    for (int t = tid; t < ff; t += NT) {
      E[z * ld + t] = (t == z) ? -1.0 : 0.0;
      if (t != z) E[t * ld + z] = 0.0;
    }
    // ---- A = L' D L'^T (right-looking, no pivoting; A is SPD) ----
    for (int k = 0; k < ff - 1; ++k) {
      __syncthreads();
      const double ip = 1.0 / A[k * ld + k];
      for (int i = k + 1 + warp; i < ff; i += NW) {
        const double f = A[i * ld + k] * ip;
        for (int j = k + 1 + lane; j < ff; j += 32) A[i * ld + j] = fma(-f, A[j * ld + k], A[i * ld + j]);
      }
    }
    __syncthreads();
    for (int t = tid; t < ff; t += NT) {
      double dgl = A[t * ld + t];
      if (!(dgl > 0.0)) bad = 1;
      dinv[t] = 1.0 / dgl;
    }
    __syncthreads();
    for (int i = warp; i < ff; i += NW)
      for (int j = lane; j < i; j += 32) A[i * ld + j] *= dinv[j];
    // ---- C' = L'^-1 B L'^-T as two row-local passes X <- X L'^-T with a transpose in between
    //      (B is symmetric, so (B L'^-T)^T = L'^-1 B); each warp owns its rows: only warp-level syncs ----
    __syncthreads();
    for (int pass = 0; pass < 2; ++pass) {
      for (int k = 0; k < ff - 1; ++k) {
        for (int i = warp; i < ff; i += NW) {
          const double x = E[i * ld + k];
          for (int j = k + 1 + lane; j < ff; j += 32) E[i * ld + j] = fma(-x, A[j * ld + k], E[i * ld + j]);
        }
        __syncwarp();
      }
      __syncthreads();
      if (pass == 0) {
        for (int i = warp; i < ff; i += NW)
          for (int j = lane; j < i; j += 32) {
            const double a = E[i * ld + j], c = E[j * ld + i];
            E[i * ld + j] = c; E[j * ld + i] = a;
          }
        __syncthreads();
      }
    }
    // ---- C = D^-1/2 X D^-1/2 ; store C and the factor ----
    double* Cg = P.Cbuf + (size_t)b * ff * ff;
    double* Lg = P.Lbuf + (size_t)b * ff * ff;
    for (int i = warp; i < ff; i += NW) {
      const double si = sqrt(dinv[i]);
      for (int j = lane; j < ff; j += 32) {
        Cg[i * ff + j] = E[i * ld + j] * si * sqrt(dinv[j]);
        Lg[i * ff + j] = (j < i) ? A[i * ld + j] : ((j == i) ? si : 0.0);
      }
    }
    if (bad) sbad = 1;
    __syncthreads();
    if (tid == 0) P.wstatus[b] = sbad ? PIRL_ST_NPD : 0;
  }
}

// ===================================================================================================
// k_tridiag : C -> tridiagonal (d, e);  M0 = L^-T Q_H
// ===================================================================================================
template <int NT>
__global__ void __launch_bounds__(NT) k_tridiag(KParams P) {
  extern __shared__ __align__(16) double sm[];
  const int ff = P.ff, ld = P.ld;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = NT / 32;
  double* v = sm;
  double* pvec = v + ff;
  double* yvec = pvec + ff;
  double* dd = yvec + ff;
  double* ee = dd + ff;
  double* scal = ee + ff;
  double* mat0 = sm + tridiag_vec_doubles(ff);
  const size_t msz = (size_t)ff * ld;
  double* scr = P.scratch + (size_t)blockIdx.x * P.scratch_stride;
  double* Cm = (P.smem_mats >= 1) ? mat0 : scr;
  double* Mm = (P.smem_mats >= 2) ? mat0 + msz : ((P.smem_mats >= 1) ? scr : scr + msz);

  for (int b = blockIdx.x; b < P.B; b += gridDim.x) {
    __syncthreads();
    const double* Cg = P.Cbuf + (size_t)b * ff * ff;
    const double* Lg = P.Lbuf + (size_t)b * ff * ff;
    for (int i = warp; i < ff; i += NW)
      for (int j = lane; j < ff; j += 32) {
        Cm[i * ld + j] = Cg[i * ff + j];
        Mm[i * ld + j] = (i == j) ? 1.0 : 0.0;
      }
    for (int k = 0; k < ff - 2; ++k) {
      const int m = ff - k - 1, r0 = k + 1;
      __syncthreads();
      if (warp == 0) {
        double s = 0.0;
        for (int t = lane; t < m; t += 32) { double x = Cm[(r0 + t) * ld + k]; v[t] = x; s = fma(x, x, s); }
        s = warp_sum(s);
        if (lane == 0) {
          const double x0 = v[0];
          const double tail2 = s - x0 * x0;
          dd[k] = Cm[k * ld + k];
          if (!(tail2 > 0.0)) { ee[k] = x0; scal[0] = 0.0; }
          else {
            const double alpha = -copysign(sqrt(s), x0);
            v[0] = x0 - alpha;
            ee[k] = alpha;
            scal[0] = 1.0 / (s - alpha * x0);   // beta = 2 / v^T v
          }
        }
      }
      __syncthreads();
      const double beta = scal[0];
      if (beta == 0.0) continue;
      // p = beta * A22 v  (rows of the trailing block) and y = beta * M[:, r0:] v
      for (int row = warp; row < m + ff; row += NW) {
        const double* rp = (row < m) ? &Cm[(r0 + row) * ld + r0] : &Mm[(row - m) * ld + r0];
        double s = 0.0;
        for (int j = lane; j < m; j += 32) s = fma(rp[j], v[j], s);
        s = warp_sum(s);
        if (lane == 0) { if (row < m) pvec[row] = beta * s; else yvec[row - m] = beta * s; }
      }
      __syncthreads();
      double s = 0.0;
      for (int t = lane; t < m; t += 32) s = fma(pvec[t], v[t], s);
      s = warp_sum(s);
      const double K = 0.5 * beta * s;
      for (int i = warp; i < m; i += NW) {
        const double vi = v[i], wi = fma(-K, vi, pvec[i]);
        for (int j = lane; j < m; j += 32) {
          const double vj = v[j], wj = fma(-K, vj, pvec[j]);
          double a = Cm[(r0 + i) * ld + r0 + j];
          a = fma(-vi, wj, a);
          a = fma(-wi, vj, a);
          Cm[(r0 + i) * ld + r0 + j] = a;
        }
      }
      for (int i = warp; i < ff; i += NW) {
        const double yi = yvec[i];
        for (int j = lane; j < m; j += 32) Mm[i * ld + r0 + j] = fma(-yi, v[j], Mm[i * ld + r0 + j]);
      }
    }
    __syncthreads();
    if (tid == 0) {
      dd[ff - 2] = Cm[(ff - 2) * ld + ff - 2];
      dd[ff - 1] = Cm[(ff - 1) * ld + ff - 1];
      ee[ff - 2] = Cm[(ff - 1) * ld + ff - 2];
      ee[ff - 1] = 0.0;
    }
    // M0 = L'^-T D^-1/2 Q_H
    for (int i = warp; i < ff; i += NW) {
      const double si = Lg[i * ff + i];
      for (int j = lane; j < ff; j += 32) Mm[i * ld + j] *= si;
    }
    for (int k = ff - 1; k >= 1; --k) {
      __syncthreads();
      for (int i = warp; i < k; i += NW) {
        const double l = __ldg(&Lg[k * ff + i]);
        for (int j = lane; j < ff; j += 32) Mm[i * ld + j] = fma(-l, Mm[k * ld + j], Mm[i * ld + j]);
      }
    }
    __syncthreads();
    double* Wg = P.Wbuf + (size_t)b * ff * ff;
    for (int i = warp; i < ff; i += NW)
      for (int j = lane; j < ff; j += 32) Wg[i * ff + j] = Mm[i * ld + j];
    double* tg = P.tri + (size_t)b * 2 * ff;
    for (int t = tid; t < ff; t += NT) { tg[t] = dd[t]; tg[ff + t] = ee[t]; }
  }
}

// ===================================================================================================
// k_ql : implicit QL with Wilkinson shifts on the tridiagonal, ONE THREAD PER STRUCTURE.
//   Only the scalar recurrences run here (latency bound, all structures in flight at once); the O(n^3)
//   part -- applying the rotations to the n x n basis -- is k_apply.  d/e live in shared memory,
//   lane-interleaved so that divergent positions never bank-conflict.
//   Output stream: rotations (c, s) in application order, plus one descriptor per QL sweep
//   (first column hi, rotation count): a sweep rotates columns (hi, hi+1), (hi-1, hi), ... in that order.
// ===================================================================================================
__global__ void __launch_bounds__(32) k_ql(KParams P) {
  // Lock-step organisation: the 32 lanes (32 structures) run one QL SWEEP per pass of the outer loop, the rotation loop
  // has a warp-uniform trip count (longest sweep of the pass), and the search for negligible off-diagonals is folded into
  // the sweep (every new e[k] is tested against its final neighbours as it is produced).  A lane never waits for another
  // lane's extra iteration on the same eigenvalue, and the dependent chain of one rotation is rsqrt + 6 DFMA/DMUL.
  // The rotation sequence is exactly the one of the textbook loop nest (for l: repeat {scan, shift, sweep}).
  extern __shared__ __align__(16) double sm[];
  const int ff = P.ff, lane = threadIdx.x;
  const int b = blockIdx.x * 32 + lane;
  const CtaTrace trc(P.trace);
  const bool valid = b < P.B;
  const int bb_ = valid ? b : P.B - 1;      // out-of-range lanes shadow the last structure (no stores)
  double* d = sm + lane;              // d[i] at d[i*32]
  double* e = sm + (size_t)ff * 32 + lane;
  const double* tg = P.tri + (size_t)bb_ * 2 * ff;
  for (int i = 0; i < ff; ++i) { d[i * 32] = tg[i]; e[i * 32] = tg[ff + i]; }
  e[(ff - 1) * 32] = 0.0;
  cplx* rc = P.rot_cs + (size_t)bb_ * P.maxrot;
  uint16_t* sw = P.sweeps + (size_t)bb_ * P.maxsw;
  int nrot = 0, nsw = 0;
  bool ok = true;
  const double eps = 2.220446049250313e-16;
  constexpr int NONE = 1 << 20;
  int l = 0, it = 0, m = 0;
  bool need_scan = true;
  while (true) {
    bool active = ok && l < ff - 1;
    if (active && need_scan) {
      while (true) {       // first negligible off-diagonal at or after l; converged eigenvalues are stepped over
        m = l;
        while (m < ff - 1) {
          const double dsum = fabs(d[m * 32]) + fabs(d[(m + 1) * 32]);
          if (fabs(e[m * 32]) <= eps * dsum) break;
          ++m;
        }
        if (m != l) break;
        ++l; it = 0;
        if (l >= ff - 1) break;
      }
      need_scan = false;
      active = l < ff - 1;
    }
    if (active && ++it > 60) { ok = false; active = false; }
    if (!__any_sync(0xffffffffu, active)) break;
    const int len = active ? m - l : 0;
    const int maxlen = __reduce_max_sync(0xffffffffu, len);
    double g = 0.0, s = 1.0, c = 1.0, p = 0.0;
    if (active) {
      const double dl = d[l * 32], el = e[l * 32];
      g = (d[(l + 1) * 32] - dl) / (2.0 * el);
      const double r = sqrt(fma(g, g, 1.0));
      g = d[m * 32] - dl + el / (g + copysign(r, g));
    }
    int i = m - 1;
    const int hi = i;
    int cnt = 0, first = NONE, second = NONE;
    bool underflow = false;
    double dold = active ? d[m * 32] : 0.0;                 // old d[i+1]
    double ei = active ? e[i * 32] : 0.0;                   // old e[i]
    double di = active ? d[i * 32] : 0.0;                   // old d[i]
    double dn1 = 0.0;                                       // new d[i+2]
    for (int step = 0; step < maxlen; ++step) {
      if (step < len && !underflow) {
        // prefetch the next step's operands (independent of the rotation chain)
        const int in = (i > l) ? i - 1 : i;
        const double ein = e[in * 32], din = d[in * 32];
        const double f = s * ei, bb = c * ei;
        const double h = fma(f, f, g * g);
        if (h == 0.0) {
          e[(i + 1) * 32] = 0.0;
          d[(i + 1) * 32] = dold - p;
          e[m * 32] = 0.0;
          underflow = true;
        } else {
          const double rinv = rsqrt(h);      // one reciprocal square root instead of sqrt + divide on the critical path
          const double enew = h * rinv;
          e[(i + 1) * 32] = enew;
          s = f * rinv;
          c = g * rinv;
          g = dold - p;
          const double r = fma(di - g, s, 2.0 * c * bb);
          p = s * r;
          const double dn0 = g + p;
          d[(i + 1) * 32] = dn0;
          g = fma(c, r, -bb);
          if (nrot < P.maxrot && valid) rc[nrot] = make_double2(c, s);
          ++nrot; ++cnt;
          if (step > 0 && enew <= eps * (fabs(dn0) + fabs(dn1))) { second = first; first = i + 1; }
          dn1 = dn0;
          dold = di; ei = ein; di = din;
          --i;
        }
      }
    }
    if (active) {
      if (cnt) { if (nsw < P.maxsw && valid) sw[nsw] = (uint16_t)(hi | (cnt << 8)); ++nsw; }
      if (underflow) { need_scan = true; }
      else {
        const double dln = dold - p;          // dold is the old d[l] after the last step
        d[l * 32] = dln;
        e[l * 32] = g;
        e[m * 32] = 0.0;
        if (fabs(g) <= eps * (fabs(dln) + fabs(dn1))) { second = first; first = l; }
        const int mnew = first < m ? first : m;
        if (mnew > l) m = mnew;
        else {                                // eigenvalue l has converged
          ++l; it = 0;
          if (second != NONE) { if (second > l) m = second; else need_scan = true; }
          else if (m <= l) need_scan = true;
        }
      }
    }
  }
  if (nrot > P.maxrot) { ok = false; nrot = P.maxrot; }
  if (nsw > P.maxsw) { ok = false; nsw = 0; nrot = 0; }
  if (valid) {
    double* lg = P.lam + (size_t)b * ff;
    for (int i = 0; i < ff; ++i) lg[i] = d[i * 32];
    P.nrot[b] = nrot;
    P.nsw[b] = nsw;
    if (!ok) atomicOr(&P.wstatus[b], PIRL_ST_EIG);
  }
  __syncwarp();
  trc.done(P.trace, 2, b);
}

// ===================================================================================================
// k_apply : W = M0 * G_1 * G_2 * ...   Each thread owns one row of the basis in shared memory.
//   K consecutive QL sweeps are applied together as a systolic pipeline: stage j works two columns behind stage j-1, the
//   values between the stages stay in registers, so one step (K rotations) costs one shared load and one shared store of
//   the row element instead of K of each (the single-sweep version was bound by shared-memory wavefronts).
//   At step i stage j rotates columns (i + 2j, i + 2j + 1); where a sweep does not cover that column the stage gets the
//   identity rotation.  The rotations of a group are staged into shared memory in step-major order ([step][stage]) while
//   the previous group is being applied (double buffer).  Rows carry 2K-1 padding columns on both sides so that the hot
//   loop needs no bounds checks.
// ===================================================================================================
template <int K>
__host__ __device__ constexpr int apply_pad() { return 2 * K - 1; }
__host__ __device__ inline int apply_ld(int ff, int K) { int ld = ff + 2 * (2 * K - 1); return ld | 1; }
__host__ __device__ inline int apply_slots(int ff, int K) { return (ff + 2 * K) * K; }

template <int ROWS, int K, int FFMAX>
__global__ void __launch_bounds__(ROWS) k_apply(KParams P) {
  extern __shared__ __align__(16) double Z[];                // [rows][ld]
  // the rotation buffers are a separate (static) object: the compiler may then hoist their loads over the row stores
  constexpr int nslots = (FFMAX + 2 * K) * K;
  __shared__ cplx rbuf[2 * nslots];
  constexpr int SWS = 256;                                    // sweep descriptors kept in shared memory (rest: global)
  __shared__ uint16_t ssw[SWS];
  const int ff = P.ff, tid = threadIdx.x;
  const int ld = apply_ld(ff, K);
  constexpr int PAD = apply_pad<K>();
  const int b = blockIdx.x, r0 = blockIdx.y * ROWS;
  const int rows = min(ROWS, ff - r0);
  const CtaTrace trc(P.trace);
  double* Wg = P.Wbuf + (size_t)b * ff * ff + (size_t)r0 * ff;
  for (int idx = tid; idx < rows * ld; idx += ROWS) Z[idx] = 0.0;
  __syncthreads();
  for (int idx = tid; idx < rows * ff; idx += ROWS) {
    int rr = idx / ff, cc = idx - rr * ff;
    Z[rr * ld + PAD + cc] = Wg[idx];
  }
  const int nr = min(P.nrot[b], P.maxrot);
  const int nsw = min(P.nsw[b], P.maxsw);
  const cplx* rc = P.rot_cs + (size_t)b * P.maxrot;
  const uint16_t* sw = P.sweeps + (size_t)b * P.maxsw;
  for (int t = tid; t < min(nsw, SWS); t += ROWS) ssw[t] = sw[t];
  __syncthreads();
  double* zr = Z + (tid < rows ? tid : 0) * ld + PAD;
  constexpr int PRE = 4;                                      // slots per thread prefetched under the previous group
  // group descriptors (uniform): hi, lo, base of the K sweeps; step range of the group
  int ghi[K], glo[K], gbase[K];
  int istart = 0, iend = 0, base_next = 0;
  auto load_group = [&](int g) {
    istart = -(1 << 20); iend = 1 << 20;
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const int sidx = g * K + j;
      unsigned w = (sidx < nsw) ? (sidx < SWS ? ssw[sidx] : sw[sidx]) : 0u;
      int cnt = (int)(w >> 8);
      const int hi = (int)(w & 255u);
      if (base_next + cnt > nr) cnt = max(nr - base_next, 0);     // inconsistent stream (flagged by k_ql)
      ghi[j] = hi; glo[j] = hi - cnt + 1; gbase[j] = base_next;
      base_next += cnt;
      if (cnt > 0) { istart = max(istart, hi - 2 * j); iend = min(iend, hi - cnt + 1 - 2 * j); }
    }
    if (iend > istart) { istart = 0; iend = 1; }              // empty group: zero steps
  };
  // staging of the group in (ghi, glo, gbase, istart, iend): slot [t*K + j] = rotation of stage j at step t (or identity)
  auto fetch_slot = [&](int idx, int T) -> cplx {
    cplx v = make_double2(1.0, 0.0);
    if (idx < T * K) {
      const int t = idx / K, j = idx - t * K;
      int hi = ghi[0], lo = glo[0], bs = gbase[0];
#pragma unroll
      for (int jj = 1; jj < K; ++jj) if (j == jj) { hi = ghi[jj]; lo = glo[jj]; bs = gbase[jj]; }
      const int c = istart - t + 2 * j;
      if (c <= hi && c >= lo) v = rc[bs + (hi - c)];
    }
    return v;
  };
  cplx pre[PRE];
  auto stage_issue = [&]() {                                  // global loads of the first PRE*ROWS slots into registers
    const int T = istart - iend + 1;
#pragma unroll
    for (int q = 0; q < PRE; ++q) pre[q] = fetch_slot(tid + q * ROWS, T);
  };
  auto stage_commit = [&](cplx* dst) {                        // registers -> shared; remaining slots directly
    const int T = istart - iend + 1;
#pragma unroll
    for (int q = 0; q < PRE; ++q) { const int idx = tid + q * ROWS; if (idx < T * K) dst[idx] = pre[q]; }
    for (int idx = PRE * ROWS + tid; idx < T * K; idx += ROWS) dst[idx] = fetch_slot(idx, T);
  };
  const int ngroups = (nsw + K - 1) / K;
  if (ngroups > 0) { load_group(0); stage_issue(); stage_commit(rbuf); }
  __syncthreads();
  for (int g = 0; g < ngroups; ++g) {
    const cplx* cp = rbuf + (size_t)(g & 1) * nslots;
    const int is = istart, T = istart - iend + 1;
    // descriptors and staging of the next group (global loads in flight while this group is applied)
    if (g + 1 < ngroups) { load_group(g + 1); stage_issue(); }
    if (tid < rows && T > 0) {
      double carry[K], dl[K];
#pragma unroll
      for (int j = 0; j < K; ++j) { carry[j] = zr[is + 2 * j + 1]; dl[j] = zr[is + 2 * j + 2]; }
      double* zp = zr + is;
      // software pipeline: the rotations and the row element of step t+1 are loaded while step t is computed
      cplx cs[K];
#pragma unroll
      for (int j = 0; j < K; ++j) cs[j] = cp[j];
      double in0 = zp[0];
#pragma unroll 2
      for (int t = 0; t < T; ++t, --zp, cp += K) {
        cplx csn[K];
#pragma unroll
        for (int j = 0; j < K; ++j) csn[j] = cp[K + j];     // (one step past the group at the end: inside the buffer, unused)
        const double in0n = zp[-1];
        // last stage first: it consumes the value the previous stage produced one step ago
#pragma unroll
        for (int j = K - 1; j >= 0; --j) {
          const double in = (j == 0) ? in0 : dl[j - 1];
          const double out = fma(cs[j].y, in, cs[j].x * carry[j]);
          carry[j] = fma(-cs[j].y, carry[j], cs[j].x * in);
          if (j == K - 1) zp[2 * K - 1] = out; else dl[j] = out;
        }
#pragma unroll
        for (int j = 0; j < K; ++j) cs[j] = csn[j];
        in0 = in0n;
      }
      // flush: zp now points at column (i_end - 1)
#pragma unroll
      for (int j = 0; j < K; ++j) { zp[2 * j + 1] = carry[j]; if (j < K - 1) zp[2 * j + 2] = dl[j]; }
    }
    if (g + 1 < ngroups) stage_commit(rbuf + (size_t)((g + 1) & 1) * nslots);
    __syncthreads();
  }
  for (int idx = tid; idx < rows * ff; idx += ROWS) {
    int rr = idx / ff, cc = idx - rr * ff;
    Wg[idx] = Z[rr * ld + PAD + cc];
  }
  trc.done(P.trace, 3, b);
}

// ===================================================================================================
// k_tail : layer solve in the mode basis (replaces _base.py:159, transfer_method.py:42-75, solver.py:29-38)
//   G_II = Q + W^T D_II W ; P = G_II^-1 ; Hq = W^T D_I W - Q
//   S = Hq (I - X^2) + 2Q + 2 (Hq X) P (X Q) ;  S c+ = r0 W[const,:]^T
//   d = 2 P (Q X c+) ; u = (I - X^2) c+ + X d ;  T = W d ; R = W u - delta
// ===================================================================================================
template <int NT, int NCOL>
__global__ void __launch_bounds__(NT) k_tail(KParams P) {
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
  int* ipiv = reinterpret_cast<int*>(red + 32);
  double* mat0 = sm + tail_vec_doubles(ff);
  const size_t csz = (size_t)ff * ld;  // complex elements per matrix
  double* scr = P.scratch + (size_t)blockIdx.x * P.scratch_stride;
  cplx* Hq = reinterpret_cast<cplx*>(scr);                      // [ff][ff] always global (L2 resident)
  cplx* spill = Hq + (size_t)ff * ff;
  cplx* Pm = (P.smem_mats >= 1) ? reinterpret_cast<cplx*>(mat0) : spill;
  cplx* Sm = (P.smem_mats >= 2) ? reinterpret_cast<cplx*>(mat0) + csz : ((P.smem_mats >= 1) ? spill : spill + csz);
  double* Wr = reinterpret_cast<double*>(Sm);                   // W staged in S's storage (real, ld)

  for (int b = blockIdx.x; b < P.B; b += gridDim.x) {
    if (P.fallback_only && !(P.wstatus[b] & 0x100)) continue;   // uniform per CTA
    __syncthreads();
    const double* pr = P.params + (size_t)b * P.pstride;
    const double wl = pr[0], ang = pr[1], thick = pr[2];
    const double sa = sin(ang * (3.14159265358979323846 / 180.0));
    const double k0d = 2.0 * 3.14159265358979323846 / wl * thick;
    const double nI = P.n_inc, nII = P.n_exit;
    const double* lamg = P.lam + (size_t)b * ff;
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
      // D = i * conj(sqrt(a)) / n^2 : a>=0 -> (0, sqrt(a)/n^2) ; a<0 -> (sqrt(-a)/n^2, 0)
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
    // ---- G_II -> Pm, Hq -> global ----
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
          if (i == j) { g = cadd(g, qv[i]); h = csub(h, qv[i]); }
          Pm[i * ld + j] = g;
          Hq[i * ff + j] = h;
        }
      }
    }
    // ---- P = G_II^-1 : complex symmetric Gauss-Jordan, no pivoting ----
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
          if (j < ff) acc[c] = cfma(f, Pm[k * ld + j], acc[c]);
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
    {
      const cplx r0 = make_double2(DI[z].x, DI[z].y + 1.0 / nI);
      for (int t = tid; t < ff; t += NT) bvec[t] = cscale(r0, wr[2 * ff + t]);
    }
    // ---- solve S c+ = b : Gaussian elimination with partial pivoting ----
    for (int k = 0; k < ff; ++k) {
      __syncthreads();
      if (warp == 0) {
        double best = -1.0; int bi = k;
        for (int i = k + lane; i < ff; i += 32) {
          cplx a = Sm[i * ld + k];
          double mg = fabs(a.x) + fabs(a.y);
          if (mg > best) { best = mg; bi = i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          double ob = __shfl_xor_sync(0xffffffffu, best, o);
          int oi = __shfl_xor_sync(0xffffffffu, bi, o);
          if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (lane == 0) ipiv[0] = bi;
      }
      __syncthreads();
      const int pv = ipiv[0];
      if (pv != k) {
        for (int j = k + tid; j < ff; j += NT) { cplx t = Sm[k * ld + j]; Sm[k * ld + j] = Sm[pv * ld + j]; Sm[pv * ld + j] = t; }
        if (tid == 0) { cplx t = bvec[k]; bvec[k] = bvec[pv]; bvec[pv] = t; }
        __syncthreads();
      }
      const cplx pinv = cinv(Sm[k * ld + k]);
      const cplx bk = bvec[k];
      for (int i = k + 1 + warp; i < ff; i += NW) {
        const cplx f = cmul(Sm[i * ld + k], pinv);
        for (int j = k + 1 + lane; j < ff; j += 32) Sm[i * ld + j] = cfnma(f, Sm[k * ld + j], Sm[i * ld + j]);
        if (lane == 0) bvec[i] = cfnma(f, bk, bvec[i]);
      }
    }
    __syncthreads();
    if (warp == 0) {   // back substitution, column oriented, one warp
      for (int k = ff - 1; k >= 0; --k) {
        cplx xk = cmul(bvec[k], cinv(Sm[k * ld + k]));
        __syncwarp();
        if (lane == 0) bvec[k] = xk;
        for (int i = lane; i < k; i += 32) bvec[i] = cfnma(Sm[i * ld + k], xk, bvec[i]);
        __syncwarp();
      }
    }
    __syncthreads();
    // ---- d = 2 P (Q X c+),  u = (1 - X^2) c+ + X d ----
    for (int t = tid; t < ff; t += NT) yv[t] = cmul(XQ[t], bvec[t]);
    __syncthreads();
    for (int i = warp; i < ff; i += NW) {
      cplx s = make_double2(0.0, 0.0);
      for (int j = lane; j < ff; j += 32) s = cfma(Pm[i * ld + j], yv[j], s);
      s = warp_sum(s);
      if (lane == 0) dv[i] = make_double2(2.0 * s.x, 2.0 * s.y);
    }
    __syncthreads();
    for (int t = tid; t < ff; t += NT) yv[t] = cfma(Xv[t], dv[t], cmul(omx2[t], bvec[t]));   // u
    __syncthreads();
    if (warp < 6) {   // six projections: T and R on {cos1, sin1, const}
      const int r = warp % 3;
      const cplx* vec = (warp < 3) ? dv : yv;
      cplx s = make_double2(0.0, 0.0);
      for (int j = lane; j < ff; j += 32) { double w = wr[r * ff + j]; s.x = fma(w, vec[j].x, s.x); s.y = fma(w, vec[j].y, s.y); }
      s = warp_sum(s);
      if (lane == 0) red[warp] = s;
    }
    __syncthreads();
    if (tid == 0) {
      const double is2 = 0.7071067811865476;
      cplx Tc = red[0], Ts = red[1], T0 = red[2];
      cplx Rc = red[3], Rs = red[4], R0 = red[5];
      R0.x -= 1.0;
      // order -1: (c + i s)/sqrt2 ; order +1: (c - i s)/sqrt2
      cplx Tm = make_double2((Tc.x - Ts.y) * is2, (Tc.y + Ts.x) * is2);
      cplx Tp = make_double2((Tc.x + Ts.y) * is2, (Tc.y - Ts.x) * is2);
      cplx Rm = make_double2((Rc.x - Rs.y) * is2, (Rc.y + Rs.x) * is2);
      cplx Rp = make_double2((Rc.x + Rs.y) * is2, (Rc.y - Rs.x) * is2);
      const double a1I = nI * nI - sa * sa, a1II = nII * nII - sa * sa;
      const double k1I = a1I >= 0.0 ? sqrt(a1I) : 0.0;      // Re(k_I,z) of orders +-1 (normalised by k0)
      const double k1II = a1II >= 0.0 ? sqrt(a1II) : 0.0;
      double der[3], det[3];
      der[0] = (Rm.x * Rm.x + Rm.y * Rm.y) * k1I / nI;
      der[1] = (R0.x * R0.x + R0.y * R0.y);
      der[2] = (Rp.x * Rp.x + Rp.y * Rp.y) * k1I / nI;
      det[0] = (Tm.x * Tm.x + Tm.y * Tm.y) * k1II / (nII * nII) * nI;
      det[1] = (T0.x * T0.x + T0.y * T0.y) * nI / nII;
      det[2] = (Tp.x * Tp.x + Tp.y * Tp.y) * k1II / (nII * nII) * nI;
      int st = P.wstatus[b] & 0xff;
      if (!isfinite(det[0]) || !isfinite(det[1]) || !isfinite(der[1])) st |= PIRL_ST_NONFINITE_;
      P.eff[b] = det[0];
      if (P.de_r3) { P.de_r3[3 * b] = der[0]; P.de_r3[3 * b + 1] = der[1]; P.de_r3[3 * b + 2] = der[2]; }
      if (P.de_t3) { P.de_t3[3 * b] = det[0]; P.de_t3[3 * b + 1] = det[1]; P.de_t3[3 * b + 2] = det[2]; }
      if (P.status) P.status[b] = st;
    }
  }
}

// ===================================================================================================
// env kernels (bit-exact integer / float64 bookkeeping)
// ===================================================================================================
__global__ void k_env_flip(int8_t* patterns, const int64_t* actions, int64_t B, int n_cells) {
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  int64_t a = actions[b];
  if (a >= 0 && a <= (int64_t)n_cells - 1) {
    int8_t* p = patterns + b * n_cells + a;
    *p = (*p == -1) ? 1 : -1;
  }
}

__global__ void k_env_reward(const double* eff_new, double* eff_prev, double* reward, double* best_eff,
                             int8_t* best_patterns, const int8_t* patterns, int64_t B, int n_cells) {
  int64_t b = blockIdx.x;
  if (b >= B) return;
  const double en = eff_new[b];
  __shared__ int better;
  if (threadIdx.x == 0) {
    reward[b] = en - eff_prev[b];
    better = 0;
    if (best_eff && en > best_eff[b]) { better = 1; }
  }
  __syncthreads();
  if (better && best_patterns)
    for (int p = threadIdx.x; p < n_cells; p += blockDim.x) best_patterns[b * n_cells + p] = patterns[b * n_cells + p];
  __syncthreads();
  if (threadIdx.x == 0) {
    eff_prev[b] = en;
    if (better) best_eff[b] = en;
  }
}
