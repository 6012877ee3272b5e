// rcwa_api.cu -- C ABI (include/pirl_rcwa.h) over the kernels in rcwa_kernels.cuh.
// Host side only owns: the handle, its device workspaces and two tiny tables.  No torch, no C++ types
// in the exported signatures.
#include "rcwa_kernels.cuh"
#include "rcwa_regtile.cuh"
#include "rcwa_tail2.cuh"
#include "rcwa_grad.cuh"
#include "rcwa_cluster.h"
#include "rcwa_smm.h"
#include "../../include/pirl_rcwa.h"

#include <complex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <utility>
#include <vector>
#include <algorithm>

static thread_local std::string g_err;

static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
#define CUDA_TRY(expr)                                                                                 \
  do {                                                                                                 \
    cudaError_t _e = (expr);                                                                           \
    if (_e != cudaSuccess)                                                                             \
      return fail((int)_e, std::string(#expr) + ": " + cudaGetErrorString(_e));                        \
  } while (0)

struct pirl_handle {
  pirl_cfg cfg;
  int ff, ld, maxrot, sm_count;
  size_t smem_optin;
  // tables
  cplx* d_tw = nullptr;
  cplx* d_Dk = nullptr;
  // chunk workspaces
  double *Cbuf = nullptr, *Lbuf = nullptr, *Wbuf = nullptr, *tri = nullptr, *lam = nullptr;
  cplx* rot_cs = nullptr;
  uint16_t* sweeps = nullptr;
  int *nrot = nullptr, *nsw = nullptr, *wstatus = nullptr;
  int maxsw = 0;
  // sub-chunk pipeline: the latency-bound QL kernel of sub-chunk c runs on a side stream under the front kernels of c+1...
  static constexpr int kMaxSub = 16, kSide = 4;
  cudaStream_t side[kSide] = {nullptr, nullptr, nullptr, nullptr};
  cudaStream_t fstream[kSide] = {nullptr, nullptr, nullptr, nullptr};   // front kernels of the sub-chunks
  cudaEvent_t evStart = nullptr;
  int* queues = nullptr; unsigned qslot = 0;      // work counters of the persistent kernels (one per launch, round robin)
  static constexpr int kQueues = 64;
  cudaEvent_t evF[kMaxSub] = {}, evQ[kMaxSub] = {};
  int sub_size = 0;
  double* scratch = nullptr;
  size_t scratch_stride = 0;
  int scratch_ctas = 0;
  // launch plans
  int front_grid, front_mats; size_t front_smem;
  int tri_grid, tri_mats; size_t tri_smem;
  int tail_grid, tail_mats; size_t tail_smem;
  int apply_rows, apply_k; size_t apply_smem; int apply_variant = 0;   // variant > 0: experiment instances for ff > 96 (PIRL_APPLY)
  size_t ql_smem;
  unsigned long long* phase_clk = nullptr;   // [32] device counters (PIRL_PHASE_CLOCKS=1)
  unsigned long long* trace = nullptr;       // CTA trace records (PIRL_TRACE=1)
  bool no_rktail = false, old_rktail = false; int rk_tail2_grid = 0; size_t rk_tail2_smem = 0;
  bool regfront121 = false;   // ff = 121: padded register-tile front kernel, shared-memory kernels behind it
  // thread-block-cluster path (rcwa_cluster.cu): orders whose matrices do not fit one SM
  bool cluster = false; ClusterPlan cplan; double* cscratch = nullptr;
  double* stack_params = nullptr;                // [max_batch][4] per-layer parameter rows of pirl_rcwa_eval_stack
  SmmPlan smm; double* smm_scratch = nullptr;    // cfg.algo == 1: scattering-matrix layer solve (rcwa_smm.cu)
  bool regtile = false; int rk_front_grid = 0; size_t rk_front_smem = 0; int rk_tail_grid = 0; size_t rk_tail_smem = 0;
  int last_chunk = 0;
  long long launches = 0;
  // host staging for eval_host
  int8_t* hd_patterns = nullptr; double* hd_params = nullptr; double* hd_out = nullptr; int32_t* hd_status = nullptr;
  int64_t hd_cap = 0;
  cudaStream_t own_stream = nullptr;
  // the workspaces serve ONE evaluation at a time, but calls arrive on different streams (the caller's for the device entry
  // points, own_stream for the *_host ones): evBusy marks the end of the last call, busy_stream the stream it ran on
  cudaEvent_t evBusy = nullptr; cudaStream_t busy_stream = nullptr; bool busy_valid = false;
  // adjoint gradient (allocated on the first pirl_rcwa_eval_grad)
  double *g_lamb = nullptr, *g_Lf = nullptr, *g_Rf = nullptr, *g_Vf = nullptr, *g_Einv = nullptr, *g_scr_tail = nullptr, *g_scr_eig = nullptr;
  int g_eig_mma = 0;   // PIRL_GEIG_DMMA: GEMMs of g_eigback on mma.sync.m8n8k4.f64 tiles
  int g_tail_grid = 0, g_tail_mats = 0, g_eig_grid = 0, g_eig_mats = 0; bool g_tail_regtile = false; int g_tail_packed = 0, g_tail_lu_k0 = 0;
  size_t g_tail_smem = 0, g_eig_smem = 0, g_tail_stride = 0, g_eig_stride = 0;
  double* hd_grad = nullptr; int64_t hd_grad_cap = 0;
  // per-kernel timing (pirl_rcwa_set_stage_timing)
  bool timing = false;
  bool sync_launch = false; int sync_fail_stage = -1; cudaError_t sync_err = cudaSuccess;   // PIRL_SYNC_LAUNCH=1: synchronise after every kernel
  std::vector<cudaEvent_t> ev_free;
  std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> ev_pending;  // (stage, (start, stop))
};

static const char* kStageNames[] = {"k_front", "k_tridiag", "k_ql", "k_apply", "k_tail", "rk_front", "rk_tail", "g_tail", "g_eigback", "ck_front", "ck_tail", "k_smm"};
static constexpr int kNumStages = 12;

// register-tile configuration of the headline size (order 40 -> ff = 81 = 9 x 9 threads, 9 x 9 tile each)
#ifdef PIRL_FRONT_9x9
using Cfg81 = TileCfg<81, 9, 9, 9, 9, 3>;
#else
// rk_front is capped at 112 registers: 4 warps x 112 + one k_ql warp x 64 registers fill one SM sub-partition exactly, so a
// k_ql CTA can sit beside two rk_front CTAs (sub-chunk pipeline); ptxas reports no spills at this cap
#ifdef PIRL_FRONT3
using Cfg81 = TileCfg<81, 27, 9, 3, 9, 3, 80>;    // experiment: three CTAs per SM (spills)
#else
using Cfg81 = TileCfg<81, 27, 9, 3, 9, 2, 112>;
#endif
#endif
// order 60: 121 = 11^2 has no usable exact tiling -> 126 = 42 x 3 = 14 x 9 with five identity-padded rows / columns
using Cfg121 = TileCfg<126, 42, 14, 3, 9, 1, 96, 121>;   // 19 warps: 5 on one SM sub-partition -> at most 102 registers each
#define RK_TAIL2_121 rk_tail2<Cfg121, 64, 10, 5>
using Tail2Sm121 = Tail2Smem<Cfg121, 64, 10, 5>;   // one CTA per SM: room for 64 significant evanescent modes
using Cfg81T = TileCfg<81, 27, 9, 3, 9, 1>;   // tail: complex tiles, one CTA per SM
using Cfg81T2 = TileCfg<81, 27, 9, 3, 9, 2>;  // real-arithmetic tail: one real tile at a time, two CTAs per SM
#define RK_TAIL2 rk_tail2<Cfg81T2, 56, 10, 5>
using Tail2Sm = Tail2Smem<Cfg81T2, 56, 10, 5>;
static constexpr int kTailMmax = 64;   // 64 keeps rk_tail + one k_ql CTA within one SM's shared memory

// Takes the handle's workspaces over for one call on `st`: waits (on the device) for the previous call if that was issued on
// another stream, and marks the end of this call when it goes out of scope.  Same-stream sequences pay one event record.
struct StreamTurn {
  pirl_handle* h; cudaStream_t st;
  StreamTurn(pirl_handle* h_, cudaStream_t st_) : h(h_), st(st_) {
    if (h->evBusy && h->busy_valid && h->busy_stream != st) cudaStreamWaitEvent(st, h->evBusy, 0);
  }
  ~StreamTurn() {
    if (h->evBusy && cudaEventRecord(h->evBusy, st) == cudaSuccess) { h->busy_stream = st; h->busy_valid = true; }
  }
};

static cudaEvent_t take_event(pirl_handle* h) {
  if (!h->ev_free.empty()) { cudaEvent_t e = h->ev_free.back(); h->ev_free.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
struct StageTimer {   // brackets one kernel launch with events when timing is on
  pirl_handle* h; int stage; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr;
  StageTimer(pirl_handle* h_, int stage_, cudaStream_t st_) : h(h_), stage(stage_), st(st_) {
    if (h->timing) { a = take_event(h); b = take_event(h); cudaEventRecord(a, st); }
  }
  ~StageTimer() {
    if (h->timing) { cudaEventRecord(b, st); h->ev_pending.push_back({stage, {a, b}}); }
    if (h->sync_launch && h->sync_fail_stage < 0) {   // diagnostic: name the first kernel that faults
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { h->sync_fail_stage = stage; h->sync_err = e; }
    }
  }
};

extern "C" int pirl_abi_version(void) { return PIRL_ABI_VERSION; }
extern "C" const char* pirl_last_error(void) { return g_err.c_str(); }

static constexpr int FRONT_NT = 256;
static constexpr int TRI_NT = 256;
static constexpr int TAIL_NT = 512;

template <typename T>
static cudaError_t dmalloc(T** p, size_t n) { return cudaMalloc((void**)p, n * sizeof(T)); }

// every kernel asks for the maximum shared-memory carveout: kernels with different carveouts cannot share an SM, and the
// sub-chunk pipeline relies on k_ql CTAs running beside the front / back kernels
template <typename K>
static cudaError_t max_carveout(K kern) {
  return cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
}

// how many matrices of `mat_bytes` fit beside `vec_bytes` in `budget` (capped at want)
static int mats_that_fit(size_t budget, size_t vec_bytes, size_t mat_bytes, int want) {
  int k = 0;
  while (k < want && vec_bytes + (size_t)(k + 1) * mat_bytes <= budget) ++k;
  return k;
}

extern "C" int pirl_rcwa_destroy(pirl_handle* h);

// scattering-matrix kernel: planned at create for algo = 1 / TE handles, on the first layer-stack call otherwise
static int smm_prepare(pirl_handle* h) {
  if (h->smm.ok) return 0;
  CUDA_TRY(smm_plan(h->ff, h->cfg.order, h->cfg.n_cells, h->sm_count, h->smem_optin - 1024, &h->smm));
  if (!h->smm.ok) return fail(-9, "k_smm does not fit on this device");
  // one scratch block (36 ff^2 doubles: 11.6 MB at order 100) per persistent CTA; no launch ever has more than max_batch
  // structures, so a small handle (the JLABCode drop-in caches max_batch = 64) must not pay for 2 x sm_count of them
  h->smm.grid = std::min(h->smm.grid, h->cfg.max_batch);
  CUDA_TRY(dmalloc(&h->smm_scratch, (size_t)h->smm.grid * h->smm.scratch_doubles));
  return 0;
}

extern "C" int pirl_rcwa_create(const pirl_cfg* cfg, pirl_handle** out) {
  if (!cfg || !out) return fail(-1, "null argument");
  if (cfg->order < 1 || cfg->order > 126) return fail(-2, "order must be in [1,126] (rotation column index is 8 bit)");
  if (cfg->n_cells < 2 || cfg->n_cells > 65536) return fail(-3, "n_cells out of range");
  if (cfg->pol != 1 && cfg->pol != 0) return fail(-4, "pol must be 1 (TM, what DeflectorGym evaluates) or 0 (TE)");
  if (cfg->theta_deg != 0.0) return fail(-5, "only normal incidence (theta=0) is implemented (real-symmetric formulation)");
  if (cfg->max_batch < 1) return fail(-6, "max_batch must be >= 1");
  if (cfg->algo != 0 && cfg->algo != 1) return fail(-10, "algo must be 0 (ETM / TMM) or 1 (SMM)");
  if (cfg->precision != 0) return fail(-11, "precision 1 (complex64 fast mode) is not implemented: float64 only");
  for (int r : cfg->reserved) if (r != 0) return fail(-12, "reserved fields of pirl_cfg must be zero");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(e == cudaSuccess ? -7 : (int)e, "no CUDA device: this library has no CPU fallback");
  if (cfg->device < 0 || cfg->device >= ndev) return fail(-8, "bad device ordinal");
  CUDA_TRY(cudaSetDevice(cfg->device));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, cfg->device));

  pirl_handle* h = new pirl_handle();
  struct Guard { pirl_handle* h; ~Guard() { if (h) pirl_rcwa_destroy(h); } } guard{h};   // error paths below free everything
  h->cfg = *cfg;
  const int N = cfg->order, ff = 2 * N + 1, nc = cfg->n_cells;
  h->ff = ff;
  h->ld = ff;  // ff is odd -> column walks are bank-conflict free
  h->maxrot = 3 * ff * ff;
  h->sm_count = prop.multiProcessorCount;
  h->smem_optin = prop.sharedMemPerBlockOptin;

  // ---- tables (host, long double) ----
  {
    std::vector<cplx> tw(nc), Dk(2 * N + 1);
    const long double PI = 3.141592653589793238462643383279502884L;
    for (int p = 0; p < nc; ++p) {
      long double a = -2.0L * PI * p / nc;
      tw[p] = make_double2((double)cosl(a), (double)sinl(a));
    }
    const long double R = 4.0L * N + 2.0L, M = R * nc;
    for (int k = 0; k <= 2 * N; ++k) {
      // zero-order hold factor of the up-sampled pattern: 1 at k = 0, exactly 0 at the other multiples of n_cells
      // (harmonics that alias onto the mean: they occur when 2 * order >= n_cells)
      std::complex<long double> d(k == 0 ? 1.0L : 0.0L, 0.0L);
      if (k % nc != 0) {
        std::complex<long double> num = 1.0L - std::polar(1.0L, -2.0L * PI * k / nc);
        std::complex<long double> den = R * (1.0L - std::polar(1.0L, -2.0L * PI * k / M));
        d = num / den;
      }
      d /= (long double)nc;
      Dk[k] = make_double2((double)d.real(), (double)d.imag());
    }
    CUDA_TRY(dmalloc(&h->d_tw, nc));
    CUDA_TRY(dmalloc(&h->d_Dk, 2 * N + 1));
    CUDA_TRY(cudaMemcpy(h->d_tw, tw.data(), nc * sizeof(cplx), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(h->d_Dk, Dk.data(), (2 * N + 1) * sizeof(cplx), cudaMemcpyHostToDevice));
  }
  // ---- chunk workspaces ----
  const size_t Bc = cfg->max_batch, m2 = (size_t)ff * ff;
  CUDA_TRY(dmalloc(&h->Cbuf, Bc * m2));
  CUDA_TRY(dmalloc(&h->Lbuf, Bc * m2));
  CUDA_TRY(dmalloc(&h->Wbuf, Bc * m2));
  CUDA_TRY(dmalloc(&h->tri, Bc * 2 * ff));
  CUDA_TRY(dmalloc(&h->lam, Bc * ff));
  CUDA_TRY(dmalloc(&h->rot_cs, Bc * h->maxrot));
  h->maxsw = 16 * ff;
  CUDA_TRY(dmalloc(&h->sweeps, Bc * h->maxsw));
  CUDA_TRY(dmalloc(&h->nrot, Bc));
  CUDA_TRY(dmalloc(&h->nsw, Bc));
  CUDA_TRY(dmalloc(&h->wstatus, Bc));

  // ---- launch plans: matrices go to shared memory when they fit, else to a per-CTA L2-resident spill ----
  const size_t budget = h->smem_optin - 1024;   // static __shared__ of the kernels counts against the opt-in limit
  const size_t rmat = (size_t)ff * h->ld * sizeof(double), cmat = 2 * rmat;
  {
    size_t vec = (size_t)front_vec_doubles(ff, N, nc) * 8;
    h->front_mats = mats_that_fit(budget, vec, rmat, 2);
    h->front_smem = vec + h->front_mats * rmat;
    int per_sm = std::max<size_t>(1, std::min<size_t>(2048 / FRONT_NT, budget / std::max<size_t>(h->front_smem + 1024, 1)));
    h->front_grid = h->sm_count * per_sm;
    CUDA_TRY(cudaFuncSetAttribute(k_front<FRONT_NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
  }
  {
    size_t vec = (size_t)tridiag_vec_doubles(ff) * 8;
    h->tri_mats = mats_that_fit(budget, vec, rmat, 2);
    h->tri_smem = vec + h->tri_mats * rmat;
    int per_sm = std::max<size_t>(1, std::min<size_t>(2048 / TRI_NT, budget / std::max<size_t>(h->tri_smem + 1024, 1)));
    h->tri_grid = h->sm_count * per_sm;
    CUDA_TRY(cudaFuncSetAttribute(k_tridiag<TRI_NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
  }
  {
    size_t vec = (size_t)tail_vec_doubles(ff) * 8;
    h->tail_mats = mats_that_fit(budget, vec, cmat, 2);
    h->tail_smem = vec + h->tail_mats * cmat;
    int per_sm = std::max<size_t>(1, std::min<size_t>(2048 / TAIL_NT, budget / std::max<size_t>(h->tail_smem + 1024, 1)));
    h->tail_grid = h->sm_count * per_sm;
    if (ff <= 96) CUDA_TRY(cudaFuncSetAttribute(k_tail<TAIL_NT, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
    else if (ff <= 128) CUDA_TRY(cudaFuncSetAttribute(k_tail<TAIL_NT, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
    else CUDA_TRY(cudaFuncSetAttribute(k_tail<TAIL_NT, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
  }
  {
    // k_apply: thread per row; K sweeps per systolic pass (4 for the headline size, 2 for the larger orders)
    h->apply_rows = (ff <= 96) ? 96 : 64;
    h->apply_k = (ff <= 96) ? 4 : 2;
    auto apply_bytes = [&](int rows, int k) { return (size_t)std::min(rows, ff) * apply_ld(ff, k) * 8; };
    const size_t abudget = budget - 2 * (size_t)(256 + 2 * 4) * 4 * sizeof(cplx);   // static rotation buffers
    h->apply_smem = apply_bytes(h->apply_rows, h->apply_k);
    if (h->apply_smem > abudget) { h->apply_rows = 32; h->apply_smem = apply_bytes(32, h->apply_k); }
    if (h->apply_rows == 96 && ff <= 81) CUDA_TRY(cudaFuncSetAttribute(k_apply<96, 4, 81>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)abudget));
    else if (h->apply_rows == 96) CUDA_TRY(cudaFuncSetAttribute(k_apply<96, 4, 96>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)abudget));
    else if (h->apply_rows == 64) CUDA_TRY(cudaFuncSetAttribute(k_apply<64, 2, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)abudget));
    else CUDA_TRY(cudaFuncSetAttribute(k_apply<32, 2, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)abudget));
    // large orders (96 < ff <= 208): more sweeps per systolic pass and smaller row blocks -> more independent rotation chains
    // per SM (the 64-row / 2-sweep variant runs ONE 64-thread CTA per SM at ff = 201: profiles/r02k_ncu_full_c5_*.csv)
    if (ff > 96 && ff <= 208) {
      // measured at order 100, batch 512 (profiles/r02l_apply_variants.txt, r02p_apply_variants.txt): 0 (64 rows x 2 sweeps): 8.25 ms,
      // 1 (32 x 4): 7.32, 2 (64 x 4): 6.46, 3 (32 x 2): 6.95, 4 (67 x 4: three row blocks cover ff = 201 exactly): 6.01, 5 (67 x 6): 4.92
      int v = (ff > 134 && ff <= 201) ? 5 : 2;
      if (const char* e = getenv("PIRL_APPLY")) v = atoi(e);
      h->apply_variant = v;
      if (v == 1) { h->apply_rows = 32; h->apply_k = 4; h->apply_smem = apply_bytes(32, 4); CUDA_TRY(cudaFuncSetAttribute(k_apply<32, 4, 208>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->apply_smem)); CUDA_TRY(max_carveout(k_apply<32, 4, 208>)); }
      else if (v == 2) { h->apply_rows = 64; h->apply_k = 4; h->apply_smem = apply_bytes(64, 4); CUDA_TRY(cudaFuncSetAttribute(k_apply<64, 4, 208>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->apply_smem)); CUDA_TRY(max_carveout(k_apply<64, 4, 208>)); }
      else if (v == 3) { h->apply_rows = 32; h->apply_k = 2; h->apply_smem = apply_bytes(32, 2); CUDA_TRY(cudaFuncSetAttribute(k_apply<32, 2, 208>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->apply_smem)); CUDA_TRY(max_carveout(k_apply<32, 2, 208>)); }
      else if (v == 4) { h->apply_rows = 67; h->apply_k = 4; h->apply_smem = apply_bytes(67, 4); CUDA_TRY(cudaFuncSetAttribute(k_apply<67, 4, 208>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->apply_smem)); CUDA_TRY(max_carveout(k_apply<67, 4, 208>)); }
      else if (v == 5) { h->apply_rows = 67; h->apply_k = 6; h->apply_smem = apply_bytes(67, 6); CUDA_TRY(cudaFuncSetAttribute(k_apply<67, 6, 208>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->apply_smem)); CUDA_TRY(max_carveout(k_apply<67, 6, 208>)); }
      else h->apply_variant = 0;
    }
    if (ff <= 81) {      // experiment: more sweeps per pass at the headline size (PIRL_APPLY81 = 6 or 8)
      int k81 = 0;
      if (const char* e = getenv("PIRL_APPLY81")) k81 = atoi(e);
      if (k81 == 6) { h->apply_variant = 16; h->apply_rows = 96; h->apply_k = 6; h->apply_smem = apply_bytes(96, 6); CUDA_TRY(cudaFuncSetAttribute(k_apply<96, 6, 81>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->apply_smem)); CUDA_TRY(max_carveout(k_apply<96, 6, 81>)); }
      if (k81 == 8) { h->apply_variant = 18; h->apply_rows = 96; h->apply_k = 8; h->apply_smem = apply_bytes(96, 8); CUDA_TRY(cudaFuncSetAttribute(k_apply<96, 8, 81>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->apply_smem)); CUDA_TRY(max_carveout(k_apply<96, 8, 81>)); }
    }
    h->ql_smem = (size_t)2 * ff * 32 * 8;
    CUDA_TRY(cudaFuncSetAttribute(k_ql, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
  }
  // the register-tile front kernels keep the twiddle table and the pattern of n_cells entries in shared memory: very long
  // patterns (about 10^4 cells and more) do not fit and run the generic kernels
  const size_t rk_budget = h->smem_optin - 1024;
  const bool tm = cfg->pol == 1;      // TE runs the generic front kernels (standard symmetric eigenproblem) and k_smm
  if (tm && ff == Cfg81::FF && !getenv("PIRL_NO_REGTILE") && (size_t)FrontSmem<Cfg81>::doubles(nc) * 8 <= rk_budget) {
    h->regtile = true;
    h->rk_front_smem = (size_t)FrontSmem<Cfg81>::doubles(nc) * 8;
    CUDA_TRY(cudaFuncSetAttribute(rk_front<Cfg81>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
    int per_sm = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rk_front<Cfg81>, Cfg81::NT, h->rk_front_smem));
    if (per_sm < 1) return fail(-9, "rk_front does not fit on this device");
    if (const char* e = getenv("PIRL_FRONT_PER_SM")) per_sm = std::max(1, std::min(per_sm, atoi(e)));   // experiment knob
    h->rk_front_grid = h->sm_count * per_sm;
    h->rk_tail_smem = (size_t)TailSmem<Cfg81T, kTailMmax>::doubles * 8;
    CUDA_TRY(cudaFuncSetAttribute(rk_tail<Cfg81T, kTailMmax>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rk_tail<Cfg81T, kTailMmax>, Cfg81T::NT, h->rk_tail_smem));
    if (per_sm < 1) return fail(-9, "rk_tail does not fit on this device");
    h->rk_tail_grid = h->sm_count * per_sm;
    h->rk_tail2_smem = (size_t)Tail2Sm::doubles * 8;
    CUDA_TRY(cudaFuncSetAttribute(RK_TAIL2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
    CUDA_TRY(max_carveout(RK_TAIL2));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, RK_TAIL2, Cfg81T2::NT, h->rk_tail2_smem));
    if (per_sm < 1) return fail(-9, "rk_tail2 does not fit on this device");
    if (const char* e = getenv("PIRL_TAIL_PER_SM")) per_sm = std::max(1, std::min(per_sm, atoi(e)));   // experiment knob
    h->rk_tail2_grid = h->sm_count * per_sm;
    h->old_rktail = getenv("PIRL_OLD_RKTAIL") != nullptr;
  }
  if (tm && ff == Cfg121::FFR && !getenv("PIRL_NO_REGTILE") && (size_t)FrontSmem<Cfg121>::doubles(nc) * 8 <= rk_budget) {
    h->regfront121 = true;
    h->rk_front_smem = (size_t)FrontSmem<Cfg121>::doubles(nc) * 8;
    CUDA_TRY(cudaFuncSetAttribute(rk_front<Cfg121>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
    CUDA_TRY(max_carveout(rk_front<Cfg121>));
    int per_sm = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rk_front<Cfg121>, Cfg121::NT, h->rk_front_smem));
    if (per_sm < 1) return fail(-9, "rk_front (order 60) does not fit on this device");
    h->rk_front_grid = h->sm_count * per_sm;
    h->rk_tail2_smem = (size_t)Tail2Sm121::doubles * 8;
    CUDA_TRY(cudaFuncSetAttribute(RK_TAIL2_121, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(h->smem_optin - 1024)));
    CUDA_TRY(max_carveout(RK_TAIL2_121));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, RK_TAIL2_121, Cfg121::NT, h->rk_tail2_smem));
    if (per_sm < 1) return fail(-9, "rk_tail2 (order 60) does not fit on this device");
    h->rk_tail2_grid = h->sm_count * per_sm;
  }
  // cluster path: every order without a single-SM register-tile kernel whose padded matrix fits the 208 x 208 cluster tile
  // (PIRL_FORCE_CLUSTER=1: the 84 x 84 test variant for orders <= 41; PIRL_NO_CLUSTER=1: generic kernels)
  if (tm && !h->regtile && !h->regfront121 && !getenv("PIRL_NO_CLUSTER")) {
    const bool small = getenv("PIRL_FORCE_CLUSTER") != nullptr && ff <= 84;
    if (small || (ff > 81 && ff <= kClusterFF)) {
      CUDA_TRY(ck_plan(ff, nc, h->smem_optin, small, &h->cplan));
      if (h->cplan.front_ok) {
        h->cluster = true;
        CUDA_TRY(dmalloc(&h->cscratch, (size_t)h->cplan.front_clusters * h->cplan.front_scratch_doubles));
      }
    }
  }
  if (cfg->algo == 1 || !tm) { if (int rc = smm_prepare(h)) return rc; }
  CUDA_TRY(max_carveout(k_front<FRONT_NT>)); CUDA_TRY(max_carveout(k_tridiag<TRI_NT>)); CUDA_TRY(max_carveout(k_ql));
  CUDA_TRY(max_carveout(k_apply<96, 4, 81>)); CUDA_TRY(max_carveout(k_apply<96, 4, 96>)); CUDA_TRY(max_carveout(k_apply<64, 2, 256>)); CUDA_TRY(max_carveout(k_apply<32, 2, 256>));
  CUDA_TRY(max_carveout(k_tail<TAIL_NT, 3>)); CUDA_TRY(max_carveout(k_tail<TAIL_NT, 4>)); CUDA_TRY(max_carveout(k_tail<TAIL_NT, 8>));
  CUDA_TRY(max_carveout(rk_front<Cfg81>)); CUDA_TRY(max_carveout(rk_tail<Cfg81T, kTailMmax>));
  h->sync_launch = getenv("PIRL_SYNC_LAUNCH") != nullptr;
  if (getenv("PIRL_TRACE")) {
    CUDA_TRY(dmalloc(&h->trace, 1 + 4 * (size_t)PIRL_TRACE_CAP));
    CUDA_TRY(cudaMemset(h->trace, 0, sizeof(unsigned long long)));
  }
  if (getenv("PIRL_PHASE_CLOCKS")) {
    CUDA_TRY(dmalloc(&h->phase_clk, 32));
    CUDA_TRY(cudaMemset(h->phase_clk, 0, 32 * sizeof(unsigned long long)));
  }
  // spill area: sized for the hungriest kernel (tail: Hq + up to 2 complex matrices)
  h->scratch_ctas = std::max(std::max(h->front_grid, h->tri_grid), h->tail_grid);
  h->scratch_stride = 2 * m2 + 2 * (2 * (size_t)ff * h->ld) + 16;
  CUDA_TRY(dmalloc(&h->scratch, (size_t)h->scratch_ctas * h->scratch_stride));
  CUDA_TRY(dmalloc(&h->queues, pirl_handle::kQueues));
  CUDA_TRY(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
  {
    int lo = 0, hi = 0;
    CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    for (int i = 0; i < pirl_handle::kSide; ++i) CUDA_TRY(cudaStreamCreateWithPriority(&h->side[i], cudaStreamNonBlocking, hi));
    for (int i = 0; i < pirl_handle::kSide; ++i) CUDA_TRY(cudaStreamCreateWithPriority(&h->fstream[i], cudaStreamNonBlocking, lo));
    CUDA_TRY(cudaEventCreateWithFlags(&h->evStart, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&h->evBusy, cudaEventDisableTiming));
    for (int i = 0; i < pirl_handle::kMaxSub; ++i) {
      CUDA_TRY(cudaEventCreateWithFlags(&h->evF[i], cudaEventDisableTiming));
      CUDA_TRY(cudaEventCreateWithFlags(&h->evQ[i], cudaEventDisableTiming));
    }
    // sub-chunk pipeline (register-tile path): a chunk of at least 4 full waves is cut in two halves by default
    // (PIRL_SUBCHUNK=<structures per sub-chunk> overrides, 0 switches it off).  Measured with the CTA trace
    // (tools/cta_trace.py, profiles/r01p_cta_trace_subchunk2048.txt): 8.72 -> 8.23 ms per 4096.
    h->sub_size = -1;   // auto
    if (const char* e = getenv("PIRL_SUBCHUNK")) h->sub_size = atoi(e);
    h->no_rktail = getenv("PIRL_NO_RKTAIL") != nullptr;
  }
  guard.h = nullptr;
  *out = h;
  return 0;
}

extern "C" int pirl_rcwa_destroy(pirl_handle* h) {
  if (!h) return 0;
  cudaSetDevice(h->cfg.device);
  cudaFree(h->d_tw); cudaFree(h->d_Dk);
  cudaFree(h->Cbuf); cudaFree(h->Lbuf); cudaFree(h->Wbuf); cudaFree(h->tri); cudaFree(h->lam);
  cudaFree(h->rot_cs); cudaFree(h->sweeps); cudaFree(h->nrot); cudaFree(h->nsw); cudaFree(h->wstatus); cudaFree(h->scratch); cudaFree(h->phase_clk); cudaFree(h->trace); cudaFree(h->queues);
  cudaFree(h->hd_patterns); cudaFree(h->hd_params); cudaFree(h->hd_out); cudaFree(h->hd_status); cudaFree(h->cscratch); cudaFree(h->smm_scratch); cudaFree(h->stack_params);
  cudaFree(h->g_lamb); cudaFree(h->g_Lf); cudaFree(h->g_Rf); cudaFree(h->g_Vf); cudaFree(h->g_Einv); cudaFree(h->g_scr_tail); cudaFree(h->g_scr_eig); cudaFree(h->hd_grad);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  for (auto q : h->side) if (q) cudaStreamDestroy(q);
  for (auto q : h->fstream) if (q) cudaStreamDestroy(q);
  if (h->evStart) cudaEventDestroy(h->evStart);
  if (h->evBusy) cudaEventDestroy(h->evBusy);
  for (auto e : h->evF) if (e) cudaEventDestroy(e);
  for (auto e : h->evQ) if (e) cudaEventDestroy(e);
  for (auto& p : h->ev_pending) { cudaEventDestroy(p.second.first); cudaEventDestroy(p.second.second); }
  for (auto e : h->ev_free) cudaEventDestroy(e);
  delete h;
  return 0;
}

// kernel parameters of the sub-chunk [off, off + cnt) of a chunk whose first structure has index 0 in the workspaces
static KParams sub_params(const KParams& P0, int n_cells, int ff, int off, int cnt) {
  KParams P = P0;
  const size_t m2 = (size_t)ff * ff;
  P.B = cnt;
  P.patterns += (size_t)off * n_cells;
  P.params += (size_t)off * P.pstride;
  P.Cbuf += off * m2; P.Lbuf += off * m2; P.Wbuf += off * m2;
  P.tri += (size_t)off * 2 * ff; P.lam += (size_t)off * ff;
  P.rot_cs += (size_t)off * P.maxrot; P.sweeps += (size_t)off * P.maxsw;
  P.nrot += off; P.nsw += off; P.wstatus += off;
  P.eff += off;
  if (P.Einv) P.Einv += off * m2;
  if (P.de_r3) P.de_r3 += 3 * (size_t)off;
  if (P.de_t3) P.de_t3 += 3 * (size_t)off;
  if (P.status) P.status += off;
  return P;
}

static int* next_queue(pirl_handle* h, cudaStream_t st) {
  int* q = h->queues + (h->qslot++ % pirl_handle::kQueues);
  cudaMemsetAsync(q, 0, sizeof(int), st);
  return q;
}

static void launch_front(pirl_handle* h, KParams P, cudaStream_t st) {
  if (h->cluster) {
    if (P.B > h->cplan.front_clusters) P.queue = next_queue(h, st);
    StageTimer t(h, 9, st);
    cudaError_t e = ck_launch_front(h->cplan, P, h->cscratch, st);
    if (e != cudaSuccess && h->sync_fail_stage < 0) { h->sync_fail_stage = 9; h->sync_err = e; }
    h->launches += 1;
  } else if (h->regfront121) {
    if (P.B > h->rk_front_grid) P.queue = next_queue(h, st);
    StageTimer t(h, 5, st);
    rk_front<Cfg121><<<std::min(P.B, h->rk_front_grid), Cfg121::NT, h->rk_front_smem, st>>>(P);
    h->launches += 1;
  } else if (h->regtile) {
    if (P.B > h->rk_front_grid) P.queue = next_queue(h, st);
    StageTimer t(h, 5, st);
    rk_front<Cfg81><<<std::min(P.B, h->rk_front_grid), Cfg81::NT, h->rk_front_smem, st>>>(P);
    h->launches += 1;
  } else {
    P.smem_mats = h->front_mats;
    { StageTimer t(h, 0, st); k_front<FRONT_NT><<<std::min(P.B, h->front_grid), FRONT_NT, h->front_smem, st>>>(P); }
    P.smem_mats = h->tri_mats;
    { StageTimer t(h, 1, st); k_tridiag<TRI_NT><<<std::min(P.B, h->tri_grid), TRI_NT, h->tri_smem, st>>>(P); }
    h->launches += 2;
  }
}

static void launch_ql(pirl_handle* h, const KParams& P, cudaStream_t st) {
  StageTimer t(h, 2, st);
  k_ql<<<(P.B + 31) / 32, 32, h->ql_smem, st>>>(P);
  h->launches += 1;
}

static void launch_apply(pirl_handle* h, const KParams& P, cudaStream_t st) {
    StageTimer t(h, 3, st);
    dim3 g(P.B, (h->ff + h->apply_rows - 1) / h->apply_rows);
    // (the third template argument sizes the static rotation buffers: ff <= 81 keeps three CTAs per SM at the headline size)
    if (h->apply_variant == 1) k_apply<32, 4, 208><<<g, 32, h->apply_smem, st>>>(P);
    else if (h->apply_variant == 2) k_apply<64, 4, 208><<<g, 64, h->apply_smem, st>>>(P);
    else if (h->apply_variant == 3) k_apply<32, 2, 208><<<g, 32, h->apply_smem, st>>>(P);
    else if (h->apply_variant == 4) k_apply<67, 4, 208><<<g, 67, h->apply_smem, st>>>(P);
    else if (h->apply_variant == 5) k_apply<67, 6, 208><<<g, 67, h->apply_smem, st>>>(P);
    else if (h->apply_variant == 16) k_apply<96, 6, 81><<<g, 96, h->apply_smem, st>>>(P);
    else if (h->apply_variant == 18) k_apply<96, 8, 81><<<g, 96, h->apply_smem, st>>>(P);
    else if (h->apply_rows == 96 && h->ff <= 81) k_apply<96, 4, 81><<<g, 96, h->apply_smem, st>>>(P);
    else if (h->apply_rows == 96) k_apply<96, 4, 96><<<g, 96, h->apply_smem, st>>>(P);
    else if (h->apply_rows == 64) k_apply<64, 2, 256><<<g, 64, h->apply_smem, st>>>(P);
    else k_apply<32, 2, 256><<<g, 32, h->apply_smem, st>>>(P);
    h->launches += 1;
}

static void launch_tail(pirl_handle* h, KParams P, cudaStream_t st) {
  if (h->cfg.algo == 1 || h->cfg.pol == 0) {          // scattering-matrix form of the layer solve (the only tail for TE)
    StageTimer t(h, 11, st);
    cudaError_t e = smm_launch(h->smm, P, h->smm_scratch, st);
    if (e != cudaSuccess && h->sync_fail_stage < 0) { h->sync_fail_stage = 11; h->sync_err = e; }
    h->launches += 1;
    return;
  }
  if (h->cluster && h->cplan.tail_ok && !h->no_rktail) {
    if (P.B > h->cplan.tail_clusters) P.queue = next_queue(h, st);
    StageTimer t(h, 10, st);
    cudaError_t e = ck_launch_tail(h->cplan, P, st);
    if (e != cudaSuccess && h->sync_fail_stage < 0) { h->sync_fail_stage = 10; h->sync_err = e; }
    P.fallback_only = 1;
    P.queue = nullptr;
    h->launches += 1;
  } else if (h->regfront121 && !h->no_rktail) {
    if (P.B > h->rk_tail2_grid) P.queue = next_queue(h, st);
    StageTimer t(h, 6, st);
    RK_TAIL2_121<<<std::min(P.B, h->rk_tail2_grid), Cfg121::NT, h->rk_tail2_smem, st>>>(P);
    P.fallback_only = 1;
    P.queue = nullptr;
    h->launches += 1;
  } else if (h->regtile && !h->no_rktail) {
    if (!h->old_rktail && P.B > h->rk_tail2_grid) P.queue = next_queue(h, st);
    StageTimer t(h, 6, st);
    if (h->old_rktail) rk_tail<Cfg81T, kTailMmax><<<std::min(P.B, h->rk_tail_grid), Cfg81T::NT, h->rk_tail_smem, st>>>(P);
    else RK_TAIL2<<<std::min(P.B, h->rk_tail2_grid), Cfg81T2::NT, h->rk_tail2_smem, st>>>(P);
    P.fallback_only = 1;
    P.queue = nullptr;
    h->launches += 1;
  }
  P.smem_mats = h->tail_mats;
  const int tg = std::min(P.B, h->tail_grid);
  StageTimer t(h, 4, st);
  if (h->ff <= 96) k_tail<TAIL_NT, 3><<<tg, TAIL_NT, h->tail_smem, st>>>(P);
  else if (h->ff <= 128) k_tail<TAIL_NT, 4><<<tg, TAIL_NT, h->tail_smem, st>>>(P);
  else k_tail<TAIL_NT, 8><<<tg, TAIL_NT, h->tail_smem, st>>>(P);
  h->launches += 1;
}

static void launch_back(pirl_handle* h, const KParams& P, cudaStream_t st) {
  launch_apply(h, P, st);
  launch_tail(h, P, st);
}

static KParams chunk_params(pirl_handle* h, const int8_t* patterns, const double* params, int pstride, int Bc,
                            double* eff, double* de_r3, double* de_t3, int32_t* status) {
  KParams P;
  memset(&P, 0, sizeof(P));
  P.n_cells = h->cfg.n_cells; P.N = h->cfg.order; P.ff = h->ff; P.ld = h->ld; P.B = Bc;
  P.n_inc = h->cfg.n_inc; P.n_exit = h->cfg.n_exit; P.eps_g = h->cfg.n_groove * h->cfg.n_groove; P.pol = h->cfg.pol;
  P.tw = h->d_tw; P.Dk = h->d_Dk; P.patterns = patterns; P.params = params; P.pstride = pstride;
  P.Cbuf = h->Cbuf; P.Lbuf = h->Lbuf; P.Wbuf = h->Wbuf; P.tri = h->tri; P.lam = h->lam;
  P.rot_cs = h->rot_cs; P.sweeps = h->sweeps; P.nrot = h->nrot; P.nsw = h->nsw; P.maxrot = h->maxrot; P.maxsw = h->maxsw;
  P.wstatus = h->wstatus; P.phase_clk = h->phase_clk; P.trace = h->trace;
  P.scratch = h->scratch; P.scratch_stride = h->scratch_stride;
  P.eff = eff; P.de_r3 = de_r3; P.de_t3 = de_t3; P.status = status;
  return P;
}

static constexpr int GTAIL_NT = 512, GEIG_NT = 256;

// gradient workspaces of the sub-chunk [off, ...): same offsets as sub_params
static GParams sub_gparams(const GParams& G0, int n_cells, int ff, int off) {
  GParams G = G0;
  G.lamb += (size_t)off * ff;
  G.Lf += (size_t)off * ff * PIRL_GRAD_RANK; G.Rf += (size_t)off * ff * PIRL_GRAD_RANK; G.Vf += (size_t)off * ff * PIRL_GRAD_NV;
  G.grad += (size_t)off * n_cells;
  return G;
}

static void launch_gtail(pirl_handle* h, const KParams& P, GParams G, cudaStream_t st) {
  StageTimer t(h, 7, st);
  G.gscratch = h->g_scr_tail; G.gstride = h->g_tail_stride; G.smem_mats = h->g_tail_mats; G.packed = h->g_tail_packed; G.lu_k0 = h->g_tail_lu_k0;
  const int g = std::min(P.B, h->g_tail_grid);
  if (h->g_tail_regtile) g_tail<GTAIL_NT, 3><<<g, GTAIL_NT, h->g_tail_smem, st>>>(P, G);
  else if (h->ff <= 128) g_tail<GTAIL_NT, 4><<<g, GTAIL_NT, h->g_tail_smem, st>>>(P, G);
  else g_tail<GTAIL_NT, 8><<<g, GTAIL_NT, h->g_tail_smem, st>>>(P, G);
  h->launches += 1;
}

static void launch_geig(pirl_handle* h, const KParams& P, GParams G, cudaStream_t st) {
  StageTimer t(h, 8, st);
  G.gscratch = h->g_scr_eig; G.gstride = h->g_eig_stride; G.smem_mats = h->g_eig_mats; G.packed = 0; G.lu_k0 = 0;
  G.use_mma = h->g_eig_mma;
  if (h->g_eig_mats >= 4) g_eigback<GEIG_NT, true><<<std::min(P.B, h->g_eig_grid), GEIG_NT, h->g_eig_smem, st>>>(P, G);
  else g_eigback<GEIG_NT, false><<<std::min(P.B, h->g_eig_grid), GEIG_NT, h->g_eig_smem, st>>>(P, G);
  h->launches += 1;
}

// One chunk.  G0 == nullptr: forward path (tails = rk_tail2 / k_tail).  G0 != nullptr: adjoint-gradient path (tails = g_tail on
// each sub-chunk, then g_eigback on the whole chunk); P0 then carries P.Einv.
static int launch_chunk(pirl_handle* h, KParams P, int Bc, cudaStream_t st, const GParams* G0 = nullptr) {

  // Sub-chunk pipeline.  k_ql is one thread per structure and purely latency bound (about 3 ms whatever the batch), so a
  // chunk is cut into sub-chunks: all front kernels are issued first on the caller's stream, each sub-chunk's k_ql runs on
  // a high-priority side stream as soon as its front kernel is done (under the following front / back kernels), and the
  // back kernels (k_apply, tail) follow on the caller's stream, each waiting for its own k_ql.
  int sub = h->sub_size;
  if (!(h->regtile || h->regfront121)) sub = 0;   // the generic kernels share per-CTA scratch across launches; the cluster kernels fill the GPU
  if (sub < 0) sub = (Bc >= 4 * h->rk_front_grid) ? (Bc + 1) / 2 : 0;
  int nsub = (sub > 0) ? (Bc + sub - 1) / sub : 1;
  if (nsub > pirl_handle::kMaxSub) { sub = (Bc + pirl_handle::kMaxSub - 1) / pirl_handle::kMaxSub; nsub = (Bc + sub - 1) / sub; }
  if (nsub <= 1) {
    launch_front(h, P, st);
    launch_ql(h, P, st);
    if (!G0) launch_back(h, P, st);
    else { launch_apply(h, P, st); launch_gtail(h, P, *G0, st); launch_geig(h, P, *G0, st); }
  } else {
    // Order on the device: fronts of all sub-chunks back to back (each on its own stream so that the persistent CTAs of
    // sub-chunk c+1 backfill the SMs as those of c drain), k_ql(c) on a high-priority side stream as soon as front(c) is done
    // (one warp per SM, co-resident with two rk_front CTAs: that is why rk_front is capped at 112 registers -- 4 warps x 112
    // + one k_ql warp x 64 registers fill one SM sub-partition exactly), then all k_apply, then all tails on the caller's
    // stream: the last k_ql hides under the k_apply launches of the earlier sub-chunks.
    const int per = sub;
    KParams Ps[pirl_handle::kMaxSub];
    CUDA_TRY(cudaEventRecord(h->evStart, st));
    for (int c = 0; c < nsub; ++c) {
      const int off = c * per, cnt = std::min(per, Bc - off);
      Ps[c] = sub_params(P, h->cfg.n_cells, h->ff, off, cnt);
      cudaStream_t fs = (h->regtile || h->regfront121) ? h->fstream[c % pirl_handle::kSide] : h->fstream[0];   // the shared-memory front kernels share per-CTA scratch
      if (c < pirl_handle::kSide) CUDA_TRY(cudaStreamWaitEvent(fs, h->evStart, 0));
      launch_front(h, Ps[c], fs);
      CUDA_TRY(cudaEventRecord(h->evF[c], fs));
      cudaStream_t qs = h->side[c % pirl_handle::kSide];
      CUDA_TRY(cudaStreamWaitEvent(qs, h->evF[c], 0));
      launch_ql(h, Ps[c], qs);
      CUDA_TRY(cudaEventRecord(h->evQ[c], qs));
    }
    // caller's stream: k_apply of all but the last sub-chunk, ONE tail launch over them (the last k_ql hides under it; the
    // tail CTAs it displaces fetch less work), then k_apply + tail of the last sub-chunk
    for (int c = 0; c + 1 < nsub; ++c) {
      CUDA_TRY(cudaStreamWaitEvent(st, h->evQ[c], 0));
      launch_apply(h, Ps[c], st);
    }
    if (!G0) launch_tail(h, sub_params(P, h->cfg.n_cells, h->ff, 0, (nsub - 1) * per), st);
    else launch_gtail(h, sub_params(P, h->cfg.n_cells, h->ff, 0, (nsub - 1) * per), *G0, st);
    CUDA_TRY(cudaStreamWaitEvent(st, h->evQ[nsub - 1], 0));
    launch_apply(h, Ps[nsub - 1], st);
    if (!G0) launch_tail(h, Ps[nsub - 1], st);
    else {
      launch_gtail(h, Ps[nsub - 1], sub_gparams(*G0, h->cfg.n_cells, h->ff, (nsub - 1) * per), st);
      launch_geig(h, P, *G0, st);
    }
  }
  h->last_chunk = Bc;
  if (h->sync_fail_stage >= 0)
    return fail((int)h->sync_err, std::string("kernel ") + kStageNames[h->sync_fail_stage] + ": " + cudaGetErrorString(h->sync_err));
  CUDA_TRY(cudaGetLastError());
  return 0;
}

extern "C" int pirl_rcwa_set_subchunk(pirl_handle* h, int32_t structures_per_subchunk) {
  if (!h) return fail(-1, "null argument");
  h->sub_size = structures_per_subchunk;
  return 0;
}
extern "C" int pirl_rcwa_num_stages(void) { return kNumStages; }
extern "C" const char* pirl_rcwa_stage_name(int32_t s) { return (s >= 0 && s < kNumStages) ? kStageNames[s] : ""; }
extern "C" int pirl_rcwa_set_stage_timing(pirl_handle* h, int32_t enable) {
  if (!h) return fail(-1, "null argument");
  h->timing = enable != 0;
  return 0;
}
extern "C" int pirl_rcwa_stage_times(pirl_handle* h, double* ms, int64_t* launches) {
  if (!h) return fail(-1, "null argument");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  for (auto& p : h->ev_pending) {
    CUDA_TRY(cudaEventSynchronize(p.second.second));
    float t = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&t, p.second.first, p.second.second));
    if (ms) ms[p.first] += (double)t;
    if (launches) launches[p.first] += 1;
    h->ev_free.push_back(p.second.first);
    h->ev_free.push_back(p.second.second);
  }
  h->ev_pending.clear();
  return 0;
}

extern "C" int pirl_rcwa_eval(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                              int64_t B, double* eff, double* de_r3, double* de_t3, int32_t* status, void* stream) {
  if (B == 0 && h) return 0;   // empty batch: nothing to do (pointers may be null)
  if (!h || !patterns || !params || !eff) return fail(-1, "null argument");
  if (B < 0) return fail(-2, "negative batch");
  if (param_stride != 0 && param_stride != 4) return fail(-3, "param_stride must be 0 or 4");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  cudaStream_t st = (cudaStream_t)stream;
  StreamTurn turn(h, st);
  const int64_t mb = h->cfg.max_batch;
  for (int64_t off = 0; off < B; off += mb) {
    const int Bc = (int)std::min<int64_t>(mb, B - off);
    int rc = launch_chunk(h, chunk_params(h, patterns + off * h->cfg.n_cells, params + off * param_stride, param_stride, Bc,
                                          eff + off, de_r3 ? de_r3 + 3 * off : nullptr, de_t3 ? de_t3 + 3 * off : nullptr,
                                          status ? status + off : nullptr), Bc, st);
    if (rc) return rc;
  }
  return 0;
}

extern "C" int pirl_rcwa_eval_host(pirl_handle* h, const int8_t* patterns, const double* params,
                                   int32_t param_stride, int64_t B, double* eff, double* de_r3, double* de_t3,
                                   int32_t* status) {
  if (B == 0 && h) return 0;   // empty batch: nothing to do (pointers may be null)
  if (!h || !patterns || !params || !eff) return fail(-1, "null argument");
  if (B < 0) return fail(-2, "negative batch");
  if (param_stride != 0 && param_stride != 4) return fail(-3, "param_stride must be 0 or 4");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  if (B > h->hd_cap) {
    cudaFree(h->hd_patterns); cudaFree(h->hd_params); cudaFree(h->hd_out); cudaFree(h->hd_status);
    h->hd_patterns = nullptr; h->hd_params = nullptr; h->hd_out = nullptr; h->hd_status = nullptr; h->hd_cap = 0;
    CUDA_TRY(dmalloc(&h->hd_patterns, (size_t)B * h->cfg.n_cells));
    CUDA_TRY(dmalloc(&h->hd_params, (size_t)B * 4));
    CUDA_TRY(dmalloc(&h->hd_out, (size_t)B * 7));
    CUDA_TRY(dmalloc(&h->hd_status, (size_t)B));
    h->hd_cap = B;
  }
  cudaStream_t st = h->own_stream;
  const size_t np = param_stride ? (size_t)B * 4 : 4;
  CUDA_TRY(cudaMemcpyAsync(h->hd_patterns, patterns, (size_t)B * h->cfg.n_cells, cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaMemcpyAsync(h->hd_params, params, np * sizeof(double), cudaMemcpyHostToDevice, st));
  double* d_eff = h->hd_out;
  double* d_r3 = h->hd_out + B;
  double* d_t3 = h->hd_out + 4 * B;
  int rc = pirl_rcwa_eval(h, h->hd_patterns, h->hd_params, param_stride, B, d_eff, de_r3 ? d_r3 : nullptr,
                          de_t3 ? d_t3 : nullptr, h->hd_status, st);
  if (rc) return rc;
  CUDA_TRY(cudaMemcpyAsync(eff, d_eff, (size_t)B * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (de_r3) CUDA_TRY(cudaMemcpyAsync(de_r3, d_r3, (size_t)B * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (de_t3) CUDA_TRY(cudaMemcpyAsync(de_t3, d_t3, (size_t)B * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (status) CUDA_TRY(cudaMemcpyAsync(status, h->hd_status, (size_t)B * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// layer stacks (the reference's layer loop, _base.py:141): every (structure, layer) pair goes through the front kernels as
// an independent eigenproblem, k_smm combines the layers of a structure with one Redheffer star product each
// ---------------------------------------------------------------------------------------------------
__global__ void k_expand_params(const double* __restrict__ params, int pstride, const double* __restrict__ thick, int tstride,
                                int L, long long BL, double* __restrict__ out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= BL) return;
  const long long b = idx / L;
  const int l = (int)(idx - b * L);
  const double* p = params + b * pstride;
  out[4 * idx + 0] = p[0]; out[4 * idx + 1] = p[1]; out[4 * idx + 2] = thick[b * tstride + l]; out[4 * idx + 3] = p[3];
}

extern "C" int pirl_rcwa_eval_stack(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                                    const double* thickness, int32_t thickness_stride, int32_t n_layers, int64_t B, double* eff,
                                    double* de_r3, double* de_t3, int32_t* status, void* stream) {
  if (B == 0 && h) return 0;
  if (!h || !patterns || !params || !thickness || !eff) return fail(-1, "null argument");
  if (B < 0) return fail(-2, "negative batch");
  if (param_stride != 0 && param_stride != 4) return fail(-3, "param_stride must be 0 or 4");
  if (n_layers < 1 || n_layers > h->cfg.max_batch) return fail(-5, "n_layers must be in [1, max_batch]");
  if (thickness_stride != 0 && thickness_stride != n_layers) return fail(-6, "thickness_stride must be 0 or n_layers");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  if (int rc = smm_prepare(h)) return rc;
  if (!h->stack_params) CUDA_TRY(dmalloc(&h->stack_params, (size_t)h->cfg.max_batch * 4));
  cudaStream_t st = (cudaStream_t)stream;
  StreamTurn turn(h, st);
  const int L = n_layers, nc = h->cfg.n_cells;
  const int64_t per = h->cfg.max_batch / L;
  for (int64_t off = 0; off < B; off += per) {
    const int Bc = (int)std::min<int64_t>(per, B - off);
    const long long BL = (long long)Bc * L;
    k_expand_params<<<(unsigned)((BL + 127) / 128), 128, 0, st>>>(params + off * param_stride, param_stride,
                                                                  thickness + off * thickness_stride, thickness_stride, L, BL,
                                                                  h->stack_params);
    KParams P = chunk_params(h, patterns + off * L * nc, h->stack_params, 4, (int)BL, eff + off, nullptr, nullptr, nullptr);
    launch_front(h, P, st);
    launch_ql(h, P, st);
    launch_apply(h, P, st);
    KParams S = P;
    S.B = Bc; S.n_layers = L; S.layer_rev = (h->cfg.algo == 1) ? 1 : 0;
    S.de_r3 = de_r3 ? de_r3 + 3 * off : nullptr; S.de_t3 = de_t3 ? de_t3 + 3 * off : nullptr; S.status = status ? status + off : nullptr;
    {
      StageTimer t(h, 11, st);
      cudaError_t e = smm_launch(h->smm, S, h->smm_scratch, st);
      if (e != cudaSuccess && h->sync_fail_stage < 0) { h->sync_fail_stage = 11; h->sync_err = e; }
      h->launches += 2;
    }
    h->last_chunk = (int)BL;
    if (h->sync_fail_stage >= 0)
      return fail((int)h->sync_err, std::string("kernel ") + kStageNames[h->sync_fail_stage] + ": " + cudaGetErrorString(h->sync_err));
    CUDA_TRY(cudaGetLastError());
  }
  return 0;
}

extern "C" int pirl_rcwa_eval_stack_host(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                                         const double* thickness, int32_t thickness_stride, int32_t n_layers, int64_t B,
                                         double* eff, double* de_r3, double* de_t3, int32_t* status) {
  if (B == 0 && h) return 0;
  if (!h || !patterns || !params || !thickness || !eff) return fail(-1, "null argument");
  if (B < 0 || n_layers < 1) return fail(-2, "bad batch / layer count");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  const size_t nc = h->cfg.n_cells, L = n_layers;
  const size_t np = param_stride ? (size_t)B * 4 : 4, nt = thickness_stride ? (size_t)B * L : L;
  int8_t* d_pat = nullptr; double* d_par = nullptr; double* d_th = nullptr; double* d_out = nullptr; int32_t* d_st = nullptr;
  struct Free { int8_t*& a; double*& b; double*& c; double*& d; int32_t*& e;
                ~Free() { cudaFree(a); cudaFree(b); cudaFree(c); cudaFree(d); cudaFree(e); } } guard{d_pat, d_par, d_th, d_out, d_st};
  CUDA_TRY(dmalloc(&d_pat, (size_t)B * L * nc)); CUDA_TRY(dmalloc(&d_par, np)); CUDA_TRY(dmalloc(&d_th, nt));
  CUDA_TRY(dmalloc(&d_out, (size_t)B * 7)); CUDA_TRY(dmalloc(&d_st, (size_t)B));
  cudaStream_t st = h->own_stream;
  CUDA_TRY(cudaMemcpyAsync(d_pat, patterns, (size_t)B * L * nc, cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaMemcpyAsync(d_par, params, np * sizeof(double), cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaMemcpyAsync(d_th, thickness, nt * sizeof(double), cudaMemcpyHostToDevice, st));
  int rc = pirl_rcwa_eval_stack(h, d_pat, d_par, param_stride, d_th, thickness_stride, n_layers, B, d_out, d_out + B, d_out + 4 * B, d_st, st);
  if (rc) { cudaStreamSynchronize(st); return rc; }
  CUDA_TRY(cudaMemcpyAsync(eff, d_out, (size_t)B * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (de_r3) CUDA_TRY(cudaMemcpyAsync(de_r3, d_out + B, (size_t)B * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (de_t3) CUDA_TRY(cudaMemcpyAsync(de_t3, d_out + 4 * B, (size_t)B * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (status) CUDA_TRY(cudaMemcpyAsync(status, d_st, (size_t)B * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return 0;
}

extern "C" int pirl_env_flip(int8_t* patterns, const int64_t* actions, int64_t B, int32_t n_cells, void* stream) {
  if (!patterns || !actions) return fail(-1, "null argument");
  if (B <= 0) return 0;
  k_env_flip<<<(unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(patterns, actions, B, n_cells);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

extern "C" int pirl_env_reward(const double* eff_new, double* eff_prev, double* reward, double* best_eff,
                               int8_t* best_patterns, const int8_t* patterns, int64_t B, int32_t n_cells,
                               void* stream) {
  if (!eff_new || !eff_prev || !reward) return fail(-1, "null argument");
  if (best_eff && best_patterns && !patterns) return fail(-2, "patterns required for best tracking");
  if (B <= 0) return 0;
  k_env_reward<<<(unsigned)B, 64, 0, (cudaStream_t)stream>>>(eff_new, eff_prev, reward, best_eff, best_patterns,
                                                             patterns, B, n_cells);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

extern "C" int64_t pirl_rcwa_launch_count(pirl_handle* h) { return h ? h->launches : 0; }

extern "C" int pirl_rcwa_debug_fetch(pirl_handle* h, int32_t what, double* dst, int64_t max_elems) {
  if (!h || !dst) return fail(-1, "null argument");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  CUDA_TRY(cudaDeviceSynchronize());
  const size_t B = h->last_chunk, ff = h->ff;
  const double* src = nullptr;
  size_t n = 0;
  switch (what) {
    case 0: src = h->Cbuf; n = B * ff * ff; break;
    case 1: src = h->tri; n = B * 2 * ff; break;
    case 2: src = h->lam; n = B * ff; break;
    case 3: src = h->Wbuf; n = B * ff * ff; break;
    case 4: {
      std::vector<int> tmp(B);
      CUDA_TRY(cudaMemcpy(tmp.data(), h->nrot, B * sizeof(int), cudaMemcpyDeviceToHost));
      for (size_t i = 0; i < B && (int64_t)i < max_elems; ++i) dst[i] = (double)tmp[i];
      return 0;
    }
    case 7: {   // working status words of the last chunk (bit 0x100: layer solve left to the generic kernel)
      std::vector<int> tmp(B);
      CUDA_TRY(cudaMemcpy(tmp.data(), h->wstatus, B * sizeof(int), cudaMemcpyDeviceToHost));
      for (size_t i = 0; i < B && (int64_t)i < max_elems; ++i) dst[i] = (double)tmp[i];
      return 0;
    }
    case 5: {   // per-phase clock accumulators (32 values), then reset
      if (!h->phase_clk) return fail(-3, "phase clocks not enabled (PIRL_PHASE_CLOCKS=1)");
      unsigned long long tmp[32];
      CUDA_TRY(cudaMemcpy(tmp, h->phase_clk, sizeof(tmp), cudaMemcpyDeviceToHost));
      CUDA_TRY(cudaMemset(h->phase_clk, 0, sizeof(tmp)));
      for (int i = 0; i < 32 && i < max_elems; ++i) dst[i] = (double)tmp[i];
      return 0;
    }
    case 6: {   // CTA trace: raw 64-bit words (counter, then 4 words per record) copied into dst, then reset
      if (!h->trace) return fail(-3, "CTA trace not enabled (PIRL_TRACE=1)");
      unsigned long long cnt = 0;
      CUDA_TRY(cudaMemcpy(&cnt, h->trace, sizeof(cnt), cudaMemcpyDeviceToHost));
      if (cnt > PIRL_TRACE_CAP) cnt = PIRL_TRACE_CAP;
      size_t words = 1 + 4 * (size_t)cnt;
      if ((int64_t)words > max_elems) words = (size_t)max_elems;
      CUDA_TRY(cudaMemcpy(dst, h->trace, words * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
      CUDA_TRY(cudaMemset(h->trace, 0, sizeof(unsigned long long)));
      return 0;
    }
    default: return fail(-2, "unknown debug item");
  }
  if ((int64_t)n > max_elems) n = (size_t)max_elems;
  CUDA_TRY(cudaMemcpy(dst, src, n * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// adjoint gradient
// ---------------------------------------------------------------------------------------------------

template <typename K>
static int plan_grad_kernel(pirl_handle* h, K kern, int nt, size_t vec_bytes, size_t mat_bytes, int want, int* mats,
                            size_t* smem, int* grid) {
  const size_t budget = h->smem_optin - 1024;
  *mats = mats_that_fit(budget, vec_bytes, mat_bytes, want);
  *smem = vec_bytes + (size_t)*mats * mat_bytes;
  CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
  CUDA_TRY(max_carveout(kern));
  int per_sm = 0;
  CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, nt, *smem));
  if (per_sm < 1) return fail(-9, "gradient kernel does not fit on this device");
  *grid = h->sm_count * per_sm;
  return 0;
}

static int grad_prepare(pirl_handle* h) {
  if (h->g_lamb) return 0;
  const int ff = h->ff;
  const size_t Bc = h->cfg.max_batch;
  const size_t rmat = (size_t)ff * h->ld * sizeof(double), cmat = 2 * rmat;
  int rc;
  const size_t tvec = (size_t)grad_tail_vec_doubles(ff) * 8;
  // register-tile variant (shared pointers only): needs both complex matrices in shared memory, i.e. ff <= 81
  h->g_tail_regtile = ff <= 96 && mats_that_fit(h->smem_optin - 1024, tvec, cmat, 2) == 2;
  if (h->g_tail_regtile) rc = plan_grad_kernel(h, g_tail<GTAIL_NT, 3>, GTAIL_NT, tvec, cmat, 2, &h->g_tail_mats, &h->g_tail_smem, &h->g_tail_grid);
  else if (ff <= 128) rc = plan_grad_kernel(h, g_tail<GTAIL_NT, 4>, GTAIL_NT, tvec, cmat, 2, &h->g_tail_mats, &h->g_tail_smem, &h->g_tail_grid);
  else rc = plan_grad_kernel(h, g_tail<GTAIL_NT, 8>, GTAIL_NT, tvec, cmat, 2, &h->g_tail_mats, &h->g_tail_smem, &h->g_tail_grid);
  if (rc) return rc;
  {   // no matrix in shared memory (ff >= 115): the packed lower triangle of G_II may still fit
    const size_t pk = (size_t)ff * (ff + 1) / 2 * sizeof(cplx);
    if (!h->g_tail_regtile && h->g_tail_mats == 0 && tvec + pk <= h->smem_optin - 1024) {
      // ... and so does the trailing block of S after the first lu_k0 elimination steps (same storage, later)
      const size_t avail = h->smem_optin - 1024 - tvec;
      int side = ff;
      while (side > 0 && (size_t)side * side * sizeof(cplx) > avail) --side;
      h->g_tail_packed = 1;
      h->g_tail_lu_k0 = ff - side;
      h->g_tail_smem = tvec + std::max(pk, (size_t)side * side * sizeof(cplx));
    }
  }
  const size_t evec = (size_t)grad_eig_vec_doubles(ff, h->cfg.n_cells) * 8;
  if (mats_that_fit(h->smem_optin - 1024, evec, rmat, 4) == 4)   // all four in shared memory: variant with shared pointers only
    rc = plan_grad_kernel(h, g_eigback<GEIG_NT, true>, GEIG_NT, evec, rmat, 4, &h->g_eig_mats, &h->g_eig_smem, &h->g_eig_grid);
  else
    rc = plan_grad_kernel(h, g_eigback<GEIG_NT, false>, GEIG_NT, evec, rmat, 4, &h->g_eig_mats, &h->g_eig_smem, &h->g_eig_grid);
  if (rc) return rc;
  // measured (profiles/r02m_dmma_trial.txt): g_eigback 4.79 -> 3.85 ms per 4096 at order 40, 7.72 -> 6.74 ms per 2048 at order 60
  h->g_eig_mma = 1;
  if (const char* e = getenv("PIRL_GEIG_DMMA")) h->g_eig_mma = atoi(e);
  h->g_tail_stride = grad_tail_scratch_doubles(ff, h->ld);
  h->g_eig_stride = grad_eig_scratch_doubles(ff, h->ld);
  CUDA_TRY(dmalloc(&h->g_scr_tail, (size_t)h->g_tail_grid * h->g_tail_stride));
  CUDA_TRY(dmalloc(&h->g_scr_eig, (size_t)h->g_eig_grid * h->g_eig_stride));
  CUDA_TRY(dmalloc(&h->g_Lf, Bc * ff * PIRL_GRAD_RANK));
  CUDA_TRY(dmalloc(&h->g_Rf, Bc * ff * PIRL_GRAD_RANK));
  CUDA_TRY(dmalloc(&h->g_Vf, Bc * ff * PIRL_GRAD_NV));
  CUDA_TRY(dmalloc(&h->g_Einv, Bc * ff * ff));
  CUDA_TRY(dmalloc(&h->g_lamb, Bc * ff));
  return 0;
}

extern "C" int pirl_rcwa_eval_grad(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                                   int64_t B, double* eff, double* grad, int32_t* status, void* stream) {
  if (B == 0 && h) return 0;
  if (!h || !patterns || !params || !eff || !grad) return fail(-1, "null argument");
  if (B < 0) return fail(-2, "negative batch");
  if (param_stride != 0 && param_stride != 4) return fail(-3, "param_stride must be 0 or 4");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  if (h->cfg.algo != 0 || h->cfg.pol != 1) return fail(-4, "the adjoint gradient is implemented for algo = 0 (ETM), TM only");
  if (int rc = grad_prepare(h)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  StreamTurn turn(h, st);
  const int64_t mb = h->cfg.max_batch;
  const int ff = h->ff;
  for (int64_t off = 0; off < B; off += mb) {
    const int Bc = (int)std::min<int64_t>(mb, B - off);
    KParams P = chunk_params(h, patterns + off * h->cfg.n_cells, params + off * param_stride, param_stride, Bc, eff + off,
                             nullptr, nullptr, status ? status + off : nullptr);
    P.Einv = h->g_Einv;
    GParams G;
    memset(&G, 0, sizeof(G));
    G.lamb = h->g_lamb; G.Lf = h->g_Lf; G.Rf = h->g_Rf; G.Vf = h->g_Vf; G.grad = grad + off * h->cfg.n_cells;
    if (int rc = launch_chunk(h, P, Bc, st, &G)) return rc;
  }
  return 0;
}

extern "C" int pirl_rcwa_eval_grad_host(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                                        int64_t B, double* eff, double* grad, int32_t* status) {
  if (B == 0 && h) return 0;
  if (!h || !patterns || !params || !eff || !grad) return fail(-1, "null argument");
  if (B < 0) return fail(-2, "negative batch");
  if (param_stride != 0 && param_stride != 4) return fail(-3, "param_stride must be 0 or 4");
  CUDA_TRY(cudaSetDevice(h->cfg.device));
  const size_t nc = h->cfg.n_cells;
  if (B > h->hd_cap) {
    cudaFree(h->hd_patterns); cudaFree(h->hd_params); cudaFree(h->hd_out); cudaFree(h->hd_status);
    h->hd_patterns = nullptr; h->hd_params = nullptr; h->hd_out = nullptr; h->hd_status = nullptr; h->hd_cap = 0;
    CUDA_TRY(dmalloc(&h->hd_patterns, (size_t)B * nc));
    CUDA_TRY(dmalloc(&h->hd_params, (size_t)B * 4));
    CUDA_TRY(dmalloc(&h->hd_out, (size_t)B * 7));
    CUDA_TRY(dmalloc(&h->hd_status, (size_t)B));
    h->hd_cap = B;
  }
  if (B > h->hd_grad_cap) {
    cudaFree(h->hd_grad); h->hd_grad = nullptr; h->hd_grad_cap = 0;
    CUDA_TRY(dmalloc(&h->hd_grad, (size_t)B * nc));
    h->hd_grad_cap = B;
  }
  cudaStream_t st = h->own_stream;
  const size_t np = param_stride ? (size_t)B * 4 : 4;
  CUDA_TRY(cudaMemcpyAsync(h->hd_patterns, patterns, (size_t)B * nc, cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaMemcpyAsync(h->hd_params, params, np * sizeof(double), cudaMemcpyHostToDevice, st));
  int rc = pirl_rcwa_eval_grad(h, h->hd_patterns, h->hd_params, param_stride, B, h->hd_out, h->hd_grad, h->hd_status, st);
  if (rc) return rc;
  CUDA_TRY(cudaMemcpyAsync(eff, h->hd_out, (size_t)B * sizeof(double), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaMemcpyAsync(grad, h->hd_grad, (size_t)B * nc * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (status) CUDA_TRY(cudaMemcpyAsync(status, h->hd_status, (size_t)B * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// pirl_rcwa_eval_static: the forward evaluation for callers that must not allocate device memory.
//
// Who that is: an RL rollout worker that shares its GPU with a learner whose framework has already claimed (nearly) all of the
// device memory for its own pool (XLA-style preallocation, a torch process with a per-process memory fraction, an MPS client
// with a memory limit), and any loop that cannot afford the device-wide synchronisation a cudaMalloc / cudaFree implies.  For
// them pirl_rcwa_create -- a dozen cudaMallocs, gigabytes at large max_batch -- either fails or stalls the neighbour.
//
// This entry point takes no handle.  The workspaces that pirl_rcwa_create would cudaMalloc (the matrices handed from rk_front
// to k_ql, k_apply and rk_tail2, the rotation stream, the twiddle tables, the k_tail spill area, the work counters) live in
// ONE file-scope device array below, so the driver reserves them when the library's module is loaded and nothing is allocated
// afterwards.  The price is a fixed, small shape: the DeflectorGym setting only (TM, order 40, at most PIRL_STATIC_MAX_CELLS
// cells, at most PIRL_STATIC_MAX_BATCH structures per call), and about 18 MB held on every GPU that loads the library.
//
// The kernels are the ones pirl_rcwa_eval launches at order 40 (rk_front, k_ql, k_apply, rk_tail2, k_tail on the flagged
// structures), all on the caller's stream, no side streams and no events.
// ---------------------------------------------------------------------------------------------------
namespace pirl_static {
constexpr int kOrder = 40, kFF = 2 * kOrder + 1, kMaxB = 32, kMaxCells = 1024, kTailCtas = 8;
static_assert(kFF == Cfg81::FF, "the static entry point runs the order-40 register-tile kernels");
constexpr size_t al(size_t x) { return (x + 255) / 256 * 256; }
constexpr size_t kM2 = (size_t)kFF * kFF, kMaxRot = 3 * kM2, kMaxSw = 16 * kFF;
constexpr size_t kSpillStride = 2 * kM2 + 2 * (2 * kM2) + 16;            // doubles per k_tail CTA, as in pirl_rcwa_create
// byte offsets into the array (every region starts on a 256-byte boundary)
constexpr size_t oTw = 0;
constexpr size_t oDk = oTw + al(kMaxCells * sizeof(cplx));
constexpr size_t oC = oDk + al(kFF * sizeof(cplx));
constexpr size_t oL = oC + al(kMaxB * kM2 * 8);
constexpr size_t oW = oL + al(kMaxB * kM2 * 8);
constexpr size_t oTri = oW + al(kMaxB * kM2 * 8);
constexpr size_t oLam = oTri + al(kMaxB * 2 * kFF * 8);
constexpr size_t oRot = oLam + al(kMaxB * kFF * 8);
constexpr size_t oSw = oRot + al(kMaxB * kMaxRot * sizeof(cplx));
constexpr size_t oSpill = oSw + al(kMaxB * kMaxSw * sizeof(uint16_t));
constexpr size_t oCnt = oSpill + al(kTailCtas * kSpillStride * 8);      // nrot | nsw | wstatus | work counters: cleared per call
constexpr size_t kCntBytes = al((3 * kMaxB + pirl_handle::kQueues) * sizeof(int));
constexpr size_t kBytes = oCnt + kCntBytes;
static_assert(kBytes < (size_t)20 << 20, "static workspace is meant to stay modest (< 20 MB)");
}  // namespace pirl_static

// The workspace itself.  HOST code must never pass this name to a kernel or a memcpy as if it were a pointer: on the host the
// symbol is only a stand-in the runtime uses to look the array up; its device address comes from cudaGetSymbolAddress, taken
// once per device in static_device() below.
static __device__ __align__(256) unsigned char g_pirl_static_ws[pirl_static::kBytes];

namespace pirl_static {
struct Dev {
  bool ready = false;
  unsigned char* base = nullptr;   // device address of g_pirl_static_ws on this device
  int tab_cells = 0;               // n_cells the twiddle / Dirichlet tables currently hold (0: none)
  std::vector<cplx> tw, Dk;        // host copies of the tables (kept alive across the asynchronous upload)
  pirl_handle h;                   // launch plan only: every device pointer in it points INTO the static array
};
constexpr int kMaxDev = 64;
static Dev g_dev[kMaxDev];
static std::mutex g_mu;

// Once per device, with that device current: find the array's device address and fix the launch plan of the five kernels.
// Nothing here allocates device memory (attribute and occupancy queries only).
static int static_device(int device, Dev** out) {
  Dev& d = g_dev[device];
  *out = &d;
  if (d.ready) return 0;
  void* p = nullptr;
  CUDA_TRY(cudaGetSymbolAddress(&p, g_pirl_static_ws));
  if (!p || ((uintptr_t)p & 255)) return fail(-9, "static workspace: unexpected device address");
  d.base = (unsigned char*)p;
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  pirl_handle* h = &d.h;
  h->ff = kFF; h->ld = kFF; h->maxrot = (int)kMaxRot; h->maxsw = (int)kMaxSw;
  h->sm_count = prop.multiProcessorCount; h->smem_optin = prop.sharedMemPerBlockOptin;
  const int optin = (int)(h->smem_optin - 1024);
  const size_t budget = h->smem_optin - 1024, cmat = 2 * kM2 * sizeof(double);
  // same attributes and sizes as pirl_rcwa_create picks for order 40 (so a handle on this device keeps its behaviour)
  h->regtile = true; h->sub_size = 0;
  CUDA_TRY(cudaFuncSetAttribute(rk_front<Cfg81>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
  CUDA_TRY(cudaFuncSetAttribute(k_ql, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
  CUDA_TRY(cudaFuncSetAttribute(RK_TAIL2, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
  CUDA_TRY(cudaFuncSetAttribute(k_tail<TAIL_NT, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
  h->apply_rows = 96; h->apply_k = 4; h->apply_variant = 0;
  h->apply_smem = (size_t)std::min(96, kFF) * apply_ld(kFF, 4) * 8;
  CUDA_TRY(cudaFuncSetAttribute(k_apply<96, 4, 81>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(budget - 2 * (size_t)(256 + 2 * 4) * 4 * sizeof(cplx))));
  CUDA_TRY(max_carveout(rk_front<Cfg81>)); CUDA_TRY(max_carveout(k_ql)); CUDA_TRY(max_carveout(k_apply<96, 4, 81>));
  CUDA_TRY(max_carveout(RK_TAIL2)); CUDA_TRY(max_carveout(k_tail<TAIL_NT, 3>));
  h->ql_smem = (size_t)2 * kFF * 32 * 8;
  h->rk_tail2_smem = (size_t)Tail2Sm::doubles * 8;
  int per_sm = 0;
  CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, RK_TAIL2, Cfg81T2::NT, h->rk_tail2_smem));
  if (per_sm < 1) return fail(-9, "rk_tail2 does not fit on this device");
  h->rk_tail2_grid = h->sm_count * per_sm;
  h->rk_front_grid = h->sm_count;                 // rk_front_smem depends on n_cells: set per call
  const size_t vec = (size_t)tail_vec_doubles(kFF) * 8;
  h->tail_mats = mats_that_fit(budget, vec, cmat, 2);
  h->tail_smem = vec + h->tail_mats * cmat;
  h->tail_grid = kTailCtas;                        // the spill area below holds kTailCtas CTAs
  // carve the array
  h->d_tw = (cplx*)(d.base + oTw); h->d_Dk = (cplx*)(d.base + oDk);
  h->Cbuf = (double*)(d.base + oC); h->Lbuf = (double*)(d.base + oL); h->Wbuf = (double*)(d.base + oW);
  h->tri = (double*)(d.base + oTri); h->lam = (double*)(d.base + oLam);
  h->rot_cs = (cplx*)(d.base + oRot); h->sweeps = (uint16_t*)(d.base + oSw);
  h->scratch = (double*)(d.base + oSpill); h->scratch_stride = kSpillStride; h->scratch_ctas = kTailCtas;
  int* cnt = (int*)(d.base + oCnt);
  h->nrot = cnt; h->nsw = cnt + kMaxB; h->wstatus = cnt + 2 * kMaxB; h->queues = cnt + 3 * kMaxB;
  d.ready = true;
  return 0;
}

// twiddles exp(-2 pi i p / n_cells) and the zero-order-hold factor of the up-sampled pattern (the tables of pirl_rcwa_create)
static void host_tables(int nc, std::vector<cplx>& tw, std::vector<cplx>& Dk) {
  const long double PI = 3.141592653589793238462643383279502884L;
  tw.resize(nc); Dk.resize(kFF);
  for (int p = 0; p < nc; ++p) {
    const long double a = -2.0L * PI * p / nc;
    tw[p] = make_double2((double)cosl(a), (double)sinl(a));
  }
  const long double R = 4.0L * kOrder + 2.0L, M = R * nc;
  for (int k = 0; k < kFF; ++k) {
    std::complex<long double> v(k == 0 ? 1.0L : 0.0L, 0.0L);
    if (k % nc != 0)
      v = (1.0L - std::polar(1.0L, -2.0L * PI * k / nc)) / (R * (1.0L - std::polar(1.0L, -2.0L * PI * k / M)));
    v /= (long double)nc;
    Dk[k] = make_double2((double)v.real(), (double)v.imag());
  }
}
}  // namespace pirl_static

extern "C" int pirl_rcwa_eval_static(const pirl_cfg* cfg, const int8_t* patterns, const double* params, int32_t param_stride,
                                     int64_t B, double* eff, double* de_r3, double* de_t3, int32_t* status, void* stream) {
  using namespace pirl_static;
  if (!cfg) return fail(-1, "null argument");
  if (cfg->order != kOrder || cfg->pol != 1 || cfg->algo != 0 || cfg->theta_deg != 0.0 || cfg->precision != 0)
    return fail(-13, "pirl_rcwa_eval_static covers TM, order 40, theta 0, algo 0, float64 only (use a pirl_rcwa_create handle otherwise)");
  if (cfg->n_cells < 2 || cfg->n_cells > kMaxCells)
    return fail(-13, "pirl_rcwa_eval_static: n_cells must be in [2, " + std::to_string(kMaxCells) + "] (PIRL_STATIC_MAX_CELLS)");
  if (B < 0) return fail(-2, "negative batch");
  if (B > kMaxB)
    return fail(-13, "pirl_rcwa_eval_static: B = " + std::to_string(B) + " exceeds the static workspace of " + std::to_string(kMaxB) +
                         " structures (PIRL_STATIC_MAX_BATCH): split the batch or use a pirl_rcwa_create handle");
  if (param_stride != 0 && param_stride != 4) return fail(-3, "param_stride must be 0 or 4");
  for (int r : cfg->reserved) if (r != 0) return fail(-12, "reserved fields of pirl_cfg must be zero");
  if (B == 0) return 0;
  if (!patterns || !params || !eff) return fail(-1, "null argument");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) return fail(e == cudaSuccess ? -7 : (int)e, "no CUDA device: this library has no CPU fallback");
  if (cfg->device < 0 || cfg->device >= ndev || cfg->device >= kMaxDev) return fail(-8, "bad device ordinal");
  CUDA_TRY(cudaSetDevice(cfg->device));
  std::lock_guard<std::mutex> lock(g_mu);
  Dev* d = nullptr;
  if (int rc = static_device(cfg->device, &d)) return rc;
  pirl_handle* h = &d->h;
  cudaStream_t st = (cudaStream_t)stream;
  const int nc = cfg->n_cells;
  const size_t front_smem = (size_t)FrontSmem<Cfg81>::doubles(nc) * 8;
  if (front_smem > h->smem_optin - 1024) return fail(-13, "pirl_rcwa_eval_static: n_cells too large for rk_front's shared memory");
  h->cfg = *cfg; h->cfg.max_batch = kMaxB;
  h->rk_front_smem = front_smem;
  if (d->tab_cells != nc) {   // tables follow n_cells; uploaded in stream order, behind whatever an earlier call still runs
    host_tables(nc, d->tw, d->Dk);
    d->tab_cells = 0;
    CUDA_TRY(cudaMemcpyAsync(h->d_tw, d->tw.data(), (size_t)nc * sizeof(cplx), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(h->d_Dk, d->Dk.data(), (size_t)kFF * sizeof(cplx), cudaMemcpyHostToDevice, st));
    d->tab_cells = nc;
  }
  // the array keeps whatever the previous call left: rotation counts, sweep counts, status words and work counters start at zero
  CUDA_TRY(cudaMemsetAsync(d->base + oCnt, 0, kCntBytes, st));
  return launch_chunk(h, chunk_params(h, patterns, params, param_stride, (int)B, eff, de_r3, de_t3, status), (int)B, st);
}
