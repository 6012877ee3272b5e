// rcwa_cluster.cu -- instantiation and launch of the thread-block-cluster kernels (see rcwa_cluster.cuh).
#include "rcwa_cluster.cuh"
#include "rcwa_cluster_tail.cuh"
#include "rcwa_cluster.h"

#include <algorithm>

using Cfg208 = ClusterCfg<208, 4, 13, 26, 4, 8>;   // orders 61..103: 52 x 26 thread grid over four CTAs, 4 x 8 tile per thread
using Cfg84 = ClusterCfg<84, 4, 7, 14, 3, 6>;      // test variant (orders <= 41): exercises the same code on small matrices

template <class C>
static cudaError_t launch_cfg(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* at, int clusters, size_t smem, cudaStream_t st) {
  *cfg = cudaLaunchConfig_t{};
  cfg->gridDim = dim3((unsigned)(clusters * C::NCTA));
  cfg->blockDim = dim3(C::NT);
  cfg->dynamicSmemBytes = smem;
  cfg->stream = st;
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = C::NCTA;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg->attrs = at;
  cfg->numAttrs = 1;
  return cudaSuccess;
}

template <class C>
static cudaError_t plan_front(int n_cells, size_t smem_optin, ClusterPlan* plan) {
  plan->front_smem = (size_t)CFrontSmem<C>::doubles(n_cells) * 8;
  plan->front_scratch_doubles = CFrontScratch<C>::doubles;
  if (plan->front_smem > smem_optin) return cudaSuccess;       // does not fit: front_ok stays false
  cudaError_t e = cudaFuncSetAttribute(ck_front<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan->front_smem);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute at[1];
  launch_cfg<C>(&cfg, at, 1, plan->front_smem, nullptr);
  int n = 0;
  e = cudaOccupancyMaxActiveClusters(&n, ck_front<C>, &cfg);
  if (e != cudaSuccess) return e;
  plan->front_clusters = n;
  plan->front_ok = n > 0;
  return cudaSuccess;
}

#define CK_TAIL208 ck_tail2<Cfg208, 64, 10, 5, 32>
using TailSm208 = CTailSmem<Cfg208, 64, 10, 5, 32>;
#define CK_TAIL84 ck_tail2<Cfg84, 56, 8, 5, 28>
using TailSm84 = CTailSmem<Cfg84, 56, 8, 5, 28>;

template <class C, class K>
static cudaError_t plan_tail(K kern, size_t smem, size_t smem_optin, ClusterPlan* plan) {
  plan->tail_smem = smem;
  if (smem > smem_optin) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute at[1];
  launch_cfg<C>(&cfg, at, 1, smem, nullptr);
  int n = 0;
  e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
  if (e != cudaSuccess) return e;
  plan->tail_clusters = n;
  plan->tail_ok = n > 0;
  return cudaSuccess;
}

cudaError_t ck_plan(int ff, int n_cells, size_t smem_optin, bool small, ClusterPlan* plan) {
  *plan = ClusterPlan();
  plan->small = small;
  if (small) {
    if (ff > Cfg84::FF) return cudaSuccess;
    cudaError_t e = plan_front<Cfg84>(n_cells, smem_optin, plan);
    if (e != cudaSuccess) return e;
    return plan_tail<Cfg84>(CK_TAIL84, (size_t)TailSm84::doubles * 8, smem_optin, plan);
  }
  if (ff > Cfg208::FF) return cudaSuccess;
  cudaError_t e = plan_front<Cfg208>(n_cells, smem_optin, plan);
  if (e != cudaSuccess) return e;
  return plan_tail<Cfg208>(CK_TAIL208, (size_t)TailSm208::doubles * 8, smem_optin, plan);
}

cudaError_t ck_launch_front(const ClusterPlan& plan, const KParams& P, double* scratch, cudaStream_t st) {
  const int clusters = std::min(P.B, plan.front_clusters);
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute at[1];
  if (plan.small) {
    launch_cfg<Cfg84>(&cfg, at, clusters, plan.front_smem, st);
    return cudaLaunchKernelEx(&cfg, ck_front<Cfg84>, P, scratch);
  }
  launch_cfg<Cfg208>(&cfg, at, clusters, plan.front_smem, st);
  return cudaLaunchKernelEx(&cfg, ck_front<Cfg208>, P, scratch);
}

cudaError_t ck_launch_tail(const ClusterPlan& plan, const KParams& P, cudaStream_t st) {
  const int clusters = std::min(P.B, plan.tail_clusters);
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute at[1];
  if (plan.small) {
    launch_cfg<Cfg84>(&cfg, at, clusters, plan.tail_smem, st);
    return cudaLaunchKernelEx(&cfg, CK_TAIL84, P);
  }
  launch_cfg<Cfg208>(&cfg, at, clusters, plan.tail_smem, st);
  return cudaLaunchKernelEx(&cfg, CK_TAIL208, P);
}
