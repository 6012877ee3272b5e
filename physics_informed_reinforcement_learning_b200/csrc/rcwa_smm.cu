// rcwa_smm.cu -- launch of the scattering-matrix layer solve (rcwa_smm.cuh).
#include "rcwa_smm.cuh"
#include "rcwa_smm.h"

cudaError_t smm_plan(int ff, int N, int n_cells, int sm_count, size_t smem_optin, SmmPlan* plan) {
  *plan = SmmPlan();
  // the matrix of the in-place inverse goes to shared memory when it fits beside the vectors (ff <= SMM_NT: one thread per column)
  plan->mats = (ff <= SMM_NT && (size_t)smm_smem_doubles(ff, N, n_cells, 1) * 8 <= smem_optin) ? 1 : 0;
  plan->smem = (size_t)smm_smem_doubles(ff, N, n_cells, plan->mats) * 8;
  if (plan->smem > smem_optin) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(k_smm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan->smem);
  if (e != cudaSuccess) return e;
  int per_sm = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_smm, SMM_NT, plan->smem);
  if (e != cudaSuccess) return e;
  plan->grid = sm_count * (per_sm > 2 ? 2 : per_sm);
  plan->scratch_doubles = smm_scratch_doubles(ff);
  plan->ok = per_sm > 0;
  return cudaSuccess;
}

cudaError_t smm_launch(const SmmPlan& plan, const KParams& P, double* scratch, cudaStream_t st) {
  const int grid = P.B < plan.grid ? P.B : plan.grid;
  KParams Q = P;
  Q.smem_mats = plan.mats;
  k_smm<<<grid, SMM_NT, plan.smem, st>>>(Q, scratch);
  return cudaGetLastError();
}
