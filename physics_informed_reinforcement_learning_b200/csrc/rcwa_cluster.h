// rcwa_cluster.h -- host interface of the thread-block-cluster kernels (rcwa_cluster.cu), used by rcwa_api.cu.
#pragma once
#include "rcwa_common.cuh"

// largest number of harmonics the cluster kernels hold (tile grid of 208 x 208, padded with identity rows / columns)
constexpr int kClusterFF = 208;
constexpr int kClusterCtas = 4;

struct ClusterPlan {
  bool front_ok = false;          // ck_front can run on this device for this n_cells
  int front_clusters = 0;         // co-resident clusters (persistent grid = clusters * kClusterCtas CTAs)
  size_t front_smem = 0;
  size_t front_scratch_doubles = 0;   // per cluster
  bool tail_ok = false;           // ck_tail2 likewise
  int tail_clusters = 0;
  size_t tail_smem = 0;
  bool small = false;             // test variant: 84 x 84 tile grid (orders <= 41), PIRL_FORCE_CLUSTER=1
};

// Fills the plan for ff harmonics and n_cells cells; returns cudaSuccess or the first CUDA error.
cudaError_t ck_plan(int ff, int n_cells, size_t smem_optin, bool small, ClusterPlan* plan);
// Enqueues ck_front over P.B structures (P.queue must point to a zeroed counter when P.B > plan.front_clusters).
cudaError_t ck_launch_front(const ClusterPlan& plan, const KParams& P, double* scratch, cudaStream_t st);
// Enqueues ck_tail2 (layer solve) over P.B structures; structures it cannot take are flagged in P.wstatus (0x100) for k_tail.
cudaError_t ck_launch_tail(const ClusterPlan& plan, const KParams& P, cudaStream_t st);
