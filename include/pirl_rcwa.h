/*
 * pirl_rcwa.h -- C ABI of the B200-native batched RCWA backend for DeflectorGym.
 *
 * This is the drop-in boundary for ONE path of the reference
 * (jLabKAIST/Physics-Informed-Reinforcement-Learning): the per-step physics call
 *
 *     MeentBase.get_efficiency(struct)        deflector-gym/deflector_gym/envs/meent_env.py:26-46
 *       -> JLABCode(...).reproduce_acs_cell   deflector-gym/third_party/meent/JLAB/solver.py:16-38
 *       -> to_conv_mat (x2)                   .../meent/on_numpy/convolution_matrix.py:89-114
 *       -> RCWALight.solve -> solve_1d (TM)   .../meent/on_numpy/rcwa.py:21-34, _base.py:119-180
 *       -> transfer_1d_1/2/3                  .../meent/on_numpy/transfer_method.py:4-75
 *
 * and the discrete env transition around it (DeflectorBase.flip, base.py:37-42;
 * MeentIndex.step, meent_env.py:90-99).  Everything is plain C: pointers, sizes, no
 * torch / C++ types.  All functions return 0 on success, <0 for an argument error,
 * >0 for a CUDA runtime error code; pirl_last_error() returns a description.
 * No exception crosses this boundary.  There is NO CPU fallback: if no CUDA device
 * is usable, pirl_rcwa_create fails.
 */
#ifndef PIRL_RCWA_H
#define PIRL_RCWA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PIRL_ABI_VERSION 2

/* status bits reported per structure (0 == ok) */
#define PIRL_ST_EIG_NOT_CONVERGED 1  /* QL iteration cap / rotation buffer exceeded          */
#define PIRL_ST_NONFINITE 2          /* a non-finite efficiency was produced                 */
#define PIRL_ST_NOT_POSDEF 4         /* permittivity Toeplitz matrices not positive definite */

/* Constants the reference hard-codes per env (meent_env.py:36-44, base.py:16-20). */
typedef struct pirl_cfg {
  int32_t n_cells;     /* pixels per period (256 / 512)                                  */
  int32_t order;       /* Fourier order N; ff = 2N+1 harmonics                            */
  double n_inc;        /* n_I  = 1.45  (incidence medium)                                 */
  double n_exit;       /* n_II = 1.0                                                      */
  double n_groove;     /* 1.0                                                             */
  double theta_deg;    /* incidence angle; only 0 is supported (as DeflectorGym uses)     */
  int32_t pol;         /* 1 = TM (what DeflectorGym evaluates; register-tile / cluster kernels), 0 = TE
                          (_base.py:142-148: generic front kernels + the S-matrix layer solve)    */
  int32_t max_batch;   /* structures processed per internal chunk (workspace sizing)      */
  int32_t device;      /* CUDA device ordinal                                             */
  int32_t algo;        /* 0 = ETM / 'TMM' (transfer_method.py:4-75, the default of the reference),
                          1 = SMM: scattering matrices + Redheffer star products
                              (scattering_method.py:12-73, smm_util.py:15-85, 289-318)        */
  int32_t precision;   /* 0 = float64 / complex128 (reference precision).  1 is reserved for the
                          complex64 fast mode of SURVEY.md section 8b and is rejected by this version */
  int32_t reserved[7]; /* must be zero (room for later fields without another ABI bump)   */
} pirl_cfg;

typedef struct pirl_handle pirl_handle;

int pirl_abi_version(void);
const char* pirl_last_error(void);

/* Replaces RCWALight.__init__/_BaseRCWA.__init__ (rcwa.py:9-18, _base.py:81-117): allocates all
 * device workspaces and tables once (the reference re-reads its material tables every call). */
int pirl_rcwa_create(const pirl_cfg* cfg, pirl_handle** out);
int pirl_rcwa_destroy(pirl_handle* h);

/* Batched get_efficiency.  Replaces meent_env.py:26-46 + solver.py:16-38 for B structures.
 *   patterns   device, int8 [B][n_cells], values +1 / -1 (anything > 0 counts as ridge)
 *   params     device, double [B][4] or [1][4]: {wavelength_nm, deflection_angle_deg, thickness_nm, n_ridge}
 *              (n_ridge = host-interpolated n_Si(lambda), convolution_matrix.py:40-52)
 *   param_stride  4 for per-structure rows, 0 to broadcast row 0
 *   eff        device, double [B]      = de_ti[N-1]           (solver.py:38)
 *   de_r3      device, double [B][3] or NULL: de_ri[N-1], de_ri[N], de_ri[N+1]
 *   de_t3      device, double [B][3] or NULL: de_ti[N-1], de_ti[N], de_ti[N+1]
 *   status     device, int32 [B] or NULL
 *   stream     cudaStream_t (as void*); work is enqueued asynchronously, no host<->device copies.
 */
int pirl_rcwa_eval(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                   int64_t B, double* eff, double* de_r3, double* de_t3, int32_t* status, void* stream);

/* Same call with HOST buffers (pinned or pageable): H2D of patterns/params, eval, D2H of results,
 * synchronous.  This is the entry point a ctypes binding inside MeentBase.get_efficiency uses. */
int pirl_rcwa_eval_host(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                        int64_t B, double* eff, double* de_r3, double* de_t3, int32_t* status);

/* Forward evaluation for callers that cannot allocate device memory.  (No counterpart in the reference: its solver is NumPy
 * on the host.  Added for deployments of THIS library.)
 *
 * Who needs it: a rollout worker sharing its GPU with a learner whose framework has already taken (almost) all device memory
 * into its own pool -- XLA-style preallocation, a torch process with a memory fraction, an MPS client with a memory limit --
 * where the cudaMallocs of pirl_rcwa_create fail; and loops that cannot tolerate the device-wide synchronisation a
 * cudaMalloc / cudaFree implies.  This call takes NO handle and performs no cudaMalloc, cudaFree or other allocating call:
 * every workspace lives in one static device array inside the library (about 18 MB, reserved on every GPU that loads the
 * library, from module load until the process exits).
 *
 * Limits (anything beyond them is rejected with -13 and a message in pirl_last_error; use a handle instead):
 *   cfg->order == PIRL_STATIC_ORDER, pol 1 (TM), algo 0, theta_deg 0, precision 0; 2 <= cfg->n_cells <= PIRL_STATIC_MAX_CELLS;
 *   0 <= B <= PIRL_STATIC_MAX_BATCH structures per call (split larger batches into consecutive calls on one stream).
 *   cfg->max_batch is ignored; cfg->device selects the GPU (made current, like the other entry points).
 * Arguments and results as pirl_rcwa_eval (device pointers, work enqueued on `stream`); same kernels as a handle runs at
 * order 40, so results agree with pirl_rcwa_eval to rounding.
 *
 * Set-up: the FIRST call on each device looks up the array's device address and sets kernel attributes; a call with a new
 * n_cells re-uploads two small tables with cudaMemcpyAsync on `stream`.  Make the first call (per device, per n_cells) outside
 * any stream capture.
 * Sharing: all calls on a device use the same array and it keeps the previous call's contents; the call clears the counters
 * that must start at zero itself.  Two calls that overlap on the GPU (different streams, or different host threads' streams)
 * race on the array: issue them on one stream, or order the streams with events.  Host-side the call is serialised by a mutex. */
#define PIRL_STATIC_ORDER 40
#define PIRL_STATIC_MAX_CELLS 1024
#define PIRL_STATIC_MAX_BATCH 32
int pirl_rcwa_eval_static(const pirl_cfg* cfg, const int8_t* patterns, const double* params, int32_t param_stride,
                          int64_t B, double* eff, double* de_r3, double* de_t3, int32_t* status, void* stream);

/* Batched efficiency + adjoint gradient d eff / d s_p (p = 0..n_cells-1), s_p the cell value of solver.py:21-22
 * (u = (s+1)/2, eps = u (n_r^2 - n_g^2) + n_g^2, 1/eps taken cell by cell before convolution_matrix.py:89-114).
 * The reference has no gradient code; the contract is the central finite difference of its forward solver
 * (reproduce_acs_cell, solver.py:16-38) with respect to one continuous cell value.
 *   eff   device, double [B]            grad   device, double [B][n_cells]
 * Reverse-mode sweep through the same factorisations as pirl_rcwa_eval (about 3x the cost of a forward call). */
int pirl_rcwa_eval_grad(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                        int64_t B, double* eff, double* grad, int32_t* status, void* stream);
/* Same with HOST buffers (H2D, eval + gradient, D2H, synchronous). */
int pirl_rcwa_eval_grad_host(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                             int64_t B, double* eff, double* grad, int32_t* status);

/* Layer stacks: the reference's layer loop (_base.py:141, JLABCode with ucell of shape (L, 1, n_cells), JLAB/solver.py:16-38).
 *   patterns   device, int8 [B][n_layers][n_cells], layers in the reference's LIST order
 *   params     device, double [B][4] or [1][4]: {wavelength_nm, deflection_angle_deg, (ignored), n_ridge}
 *   thickness  device, double [B][n_layers] or [1][n_layers] (thickness_stride = n_layers or 0)
 * Every (structure, layer) pair is one eigenproblem of the front kernels (n_layers <= max_batch); the layers of a structure
 * are combined with Redheffer star products (smm_util.py:289-318).  cfg.algo keeps the reference's layer ORDER per
 * algorithm: 0 ('TMM') = list order from the incidence side; 1 ('SMM') = the reversed list -- the reference walks the list
 * backwards for both algorithms, which for its S-matrix form appends the layers in reversed order (its 'SMM' result of a
 * list equals its 'TMM' result of the reversed list; verified against the unmodified reference, tests/golden/
 * rcwa_stack_golden.npz).  Outputs as pirl_rcwa_eval. */
int pirl_rcwa_eval_stack(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                         const double* thickness, int32_t thickness_stride, int32_t n_layers, int64_t B, double* eff,
                         double* de_r3, double* de_t3, int32_t* status, void* stream);
/* Same with HOST buffers (H2D, solve, D2H, synchronous). */
int pirl_rcwa_eval_stack_host(pirl_handle* h, const int8_t* patterns, const double* params, int32_t param_stride,
                              const double* thickness, int32_t thickness_stride, int32_t n_layers, int64_t B, double* eff,
                              double* de_r3, double* de_t3, int32_t* status);

/* DeflectorBase.flip for B envs (base.py:37-42): patterns[b][a] = -patterns[b][a] iff 0 <= a <= n_cells-1. */
int pirl_env_flip(int8_t* patterns, const int64_t* actions, int64_t B, int32_t n_cells, void* stream);

/* MeentIndex.step bookkeeping (meent_env.py:90-99): reward = eff_new - eff_prev (this order, float64),
 * eff_prev <- eff_new, best tracking as BestRecorder (wrappers.py:10-16: strict >). best_* may be NULL. */
int pirl_env_reward(const double* eff_new, double* eff_prev, double* reward, double* best_eff,
                    int8_t* best_patterns, const int8_t* patterns, int64_t B, int32_t n_cells, void* stream);

/* Tuning knob of the internal sub-chunk pipeline (k_ql of one half of a chunk runs under the front kernel of the other half):
 * structures per sub-chunk; -1 = automatic (default: two halves for chunks of at least four waves), 0 = off (every kernel of
 * a chunk runs alone on the caller's stream -- what bench.py uses to time single kernels for the roofline).  Results are
 * bit-identical in every setting. */
int pirl_rcwa_set_subchunk(pirl_handle* h, int32_t structures_per_subchunk);

/* Number of kernels this library launched since the handle was created (bench.py's gpu_launches). */
int64_t pirl_rcwa_launch_count(pirl_handle* h);

/* Per-kernel device timing (bench.py's roofline): when enabled, every kernel launched by pirl_rcwa_eval is
 * bracketed by CUDA events on the caller's stream.  pirl_rcwa_stage_times synchronises those events, ADDS the
 * elapsed milliseconds per stage into ms[0..num_stages) and the launch counts into launches[0..num_stages)
 * (either may be NULL), and clears the internal record. */
int pirl_rcwa_num_stages(void);
const char* pirl_rcwa_stage_name(int32_t stage);
int pirl_rcwa_set_stage_timing(pirl_handle* h, int32_t enable);
int pirl_rcwa_stage_times(pirl_handle* h, double* ms, int64_t* launches);

/* Test hook: copy a stage intermediate of the LAST chunk evaluated to a host buffer.
 * what: 0 = reduced standard-form matrix C [B][ff][ff], 1 = tridiagonal d,e [B][2][ff],
 *       2 = eigenvalues [B][ff], 3 = modes W (A-orthonormal, real basis) [B][ff][ff],
 *       4 = rotation counts int64-as-double [B]                                               */
int pirl_rcwa_debug_fetch(pirl_handle* h, int32_t what, double* dst, int64_t max_elems);

#ifdef __cplusplus
}
#endif
#endif /* PIRL_RCWA_H */
