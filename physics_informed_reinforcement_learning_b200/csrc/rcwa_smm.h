// rcwa_smm.h -- host interface of the scattering-matrix layer solve (rcwa_smm.cu), used by rcwa_api.cu when cfg.algo == 1.
#pragma once
#include "rcwa_common.cuh"

struct SmmPlan {
  bool ok = false;
  int grid = 0;                    // persistent CTAs (each with its own scratch)
  size_t smem = 0;
  int mats = 0;                    // 1: the matrix of the in-place inverse lives in shared memory
  size_t scratch_doubles = 0;      // per CTA
};

cudaError_t smm_plan(int ff, int N, int n_cells, int sm_count, size_t smem_optin, SmmPlan* plan);
// Enqueues k_smm over P.B structures: reads P.Wbuf (modes) and P.lam (eigenvalues), writes eff / de_r3 / de_t3 / status.
cudaError_t smm_launch(const SmmPlan& plan, const KParams& P, double* scratch, cudaStream_t st);
