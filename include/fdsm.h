/* libfdsm -- C ABI of the B200-native FD score-matching hot path.
 *
 * The reference (taufikxu/FD-ScoreMatching) is pure Python/PyTorch and has NO FFI layer: its
 * boundary is the Python call `loss = f(net, samples, ...)`.  This header is the
 * C-ABI a maintainer binds (ctypes stub in INTEGRATION.md) to replace the torch op sequences
 * cited on each entry point (paths relative to the reference repo root).
 *
 * Conventions (all entry points):
 *   - plain pointers + sizes; every pointer is a DEVICE pointer unless named `h_*`;
 *   - the caller owns every buffer including workspaces; the library never allocates or frees
 *     device memory and keeps no pointer after return;
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); no host sync, no host
 *     read of device data => CUDA-graph capturable;
 *   - return 0 on success, <0 for argument errors, >0 = cudaError_t; message via
 *     fdsm_last_error_string() (thread-local);
 *   - row-major fp32 matrices with an explicit leading dimension `ld*` in elements;
 *   - 128-bit vector path when D%4==0, ld%4==0 and 16-byte aligned bases; scalar path otherwise
 *     (D=11 of config 1).
 */
#ifndef FDSM_H_
#define FDSM_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FDSM_ABI_VERSION 11

#define FDSM_OK 0
#define FDSM_ERR_INVALID_ARGUMENT (-1)
#define FDSM_ERR_UNSUPPORTED (-2)

typedef void* fdsm_stream_t; /* cudaStream_t */

int fdsm_version(void);
const char* fdsm_last_error_string(void);
/* 1 if the current device is compute capability 10.x (the only target), else 0. */
int fdsm_device_supported(void);

/* ------------------------------------------------------------------------------------------
 * Kernels (1)+(2): direction sampling + per-sample L2 normalisation + stacked-batch build.
 *
 * Replaces   u = randn_like(x); v = u / ||u||_2 * eps; cat([x, x+v, x-v])            or
 *            n = randn_like(x); xt = x + sigma_b*n; u = ...; cat([xt+v, xt-v])
 * at SSM/losses/efficient_sm.py:9-16,28-37,53-59,79-84,105-112; SSM/losses/dsm.py:124-132;
 * DeepEBM/ESM/loss_train.py:101-107,217-222; ncsn/losses/sliced_sm.py:175-188,211-224;
 * ncsnv2/losses/dsm.py:92-113.
 * ------------------------------------------------------------------------------------------ */
#define FDSM_ROW_X 1     /* emit base row          (x or xt)   */
#define FDSM_ROW_PLUS 2  /* emit base + v                       */
#define FDSM_ROW_MINUS 4 /* emit base - v                       */

typedef struct fdsm_stack_args {
  const float* x;        /* (B, D) samples; may be NULL iff row_mask == 0                       */
  int64_t ldx;
  const float* dir;      /* injected direction (B, D), or NULL => Philox stream 0                */
  int64_t lddir;
  int32_t normalize_dir; /* 1: v = dir/||dir||*dir_scale (dir is raw u); 0: v = dir as given     */
  float dir_scale;       /* eps (times sqrt(D) for dsm_fd, efficient_sm.py:33-35)                */
  const float* noise;    /* injected unit-Gaussian DSM noise n (B, D), or NULL => Philox stream 1;
                            only used when `sigma` != NULL                                       */
  int64_t ldnoise;
  const float* sigma;    /* per-sample sigma (B) -- or the sigma table when `labels` != NULL;
                            NULL => no DSM noise                                                 */
  const int64_t* labels; /* optional (B): sigma_b = sigma[labels[b]] (ncsn/ncsnv2)               */
  int32_t perturb_input; /* 1: base = x + sigma_b*n; 0: base = x (dsm_tracetrick_FD, dsm.py:131) */
  int32_t dot_mode;      /* dot_out: 0 = sum_d v*(sigma_b*n) (efficient_sm.py:41, dsm.py:137);
                                     1 = sum_d v*((x+sigma_b*n)-x) (loss_train.py:112)            */
  float* out;            /* (popcount(row_mask)*B, D) stacked rows in X,PLUS,MINUS order; may be
                            NULL iff row_mask == 0                                               */
  int64_t ldout;
  int32_t row_mask;
  float* v_out;          /* optional (B, D): the normalised direction v                          */
  int64_t ldv;
  float* dot_out;        /* optional (B)                                                         */
  uint64_t seed;         /* Philox key                                                           */
  uint64_t offset;       /* Philox draw counter (step)                                           */
  uint64_t row0;         /* global index of local row 0 (rank*B_local): shards draw what the
                            un-sharded batch would                                               */
  int64_t B, D;
  float* out_lo;         /* optional: when set, `out` receives the TF32 hi half and `out_lo` the lo half of every
                            stacked row (x = hi + lo), and columns [D, ldout) are written as the GEMM padding:
                            column D = 1 (folded bias), the rest 0 -- the operand format of fdsm_mlp_forward      */
  const uint64_t* offset_dev; /* optional DEVICE counter added to `offset` at run time (e.g. the optimizer step
                            counter), so a captured CUDA graph draws fresh directions on every replay             */
  int32_t pad_ones;      /* 1 (with out_lo == NULL): `out` rows are plain fp32 but columns [D, ldout) still get the GEMM
                            padding (column D = 1, rest 0): the operand format of fdsm_mlp_forward with presplit = 0 */
} fdsm_stack_args;

int fdsm_stack(const fdsm_stack_args* args, fdsm_stream_t stream);

/* Kernel (1) alone: v = eps_scale * u/||u||, u ~ Philox N(0,I) (dir == NULL) or injected. */
int fdsm_direction(const float* dir, int64_t lddir, float* v, int64_t ldv, int64_t B, int64_t D, float scale,
                   uint64_t seed, uint64_t offset, uint64_t row0, fdsm_stream_t stream);

/* Raw Philox N(0,1) draws of a stream (0 = direction, 1 = DSM noise), no normalisation: what the
 * fused kernels consume internally; exposed for the statistical / bit-exactness tests. */
int fdsm_philox_normal(float* out, int64_t ld, int64_t B, int64_t D, int32_t stream_id, uint64_t seed,
                       uint64_t offset, uint64_t row0, fdsm_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Kernel (4): central-difference combine, forward and analytic backward.
 * ------------------------------------------------------------------------------------------ */
/* Energy form.  E = energy(stacked) (3B) [mode 0] or (2B) [mode 1].
 *   mode 0: per_b = (E2-E3)^2/8 - E2 - E3 + c1*E1        efficient_sm.py:61-71,113-121;
 *                                                         loss_train.py:225-232
 *   mode 1: d = E1-E2; per_b = d^2 + 4d                   efficient_sm.py:21-22
 *   loss = scale * (mean ? mean_b : sum_b) per_b ; per_sample (optional) = scale*per_b
 * bwd: dE = gout[0] * dloss/dE   (gout: device scalar, NULL => 1)                           */
int fdsm_combine_energy_fwd(const float* E, int64_t B, int32_t mode, float c1, float scale, int32_t mean,
                            float* per_sample, float* loss, fdsm_stream_t stream);
int fdsm_combine_energy_bwd(const float* E, int64_t B, int32_t mode, float c1, float scale, int32_t mean,
                            const float* gout, float* dE, fdsm_stream_t stream);

/* DSM energy form.  E = energy([xt+v, xt-v]) (2B); logp = -E.
 *   r_b = alpha*(E[B+b]-E[b]) + beta*dot_b ; per_b = r_b^2 * cw_b ; loss = gamma*mean_b per_b
 *   cw_b = 1/sigma_b^2 when inv_sigma2 != NULL (sigma_b = inv_sigma2[b]), else 1.
 *   loss_train.py:109-117 (alpha .5, beta 1/sigma0^2, gamma loss_coe*D/eps^2);
 *   efficient_sm.py:39-43 (alpha sigma^2/2, beta 1, gamma 1/(2 eps^2));
 *   SSM/losses/dsm.py:132-139 (alpha sigma^2, beta 2, gamma D/(8 eps^2)).                    */
int fdsm_combine_dsm_fwd(const float* E, const float* dot, const float* sigma_b, int64_t B, float alpha, float beta,
                         float gamma, float* per_sample, float* loss, fdsm_stream_t stream);
int fdsm_combine_dsm_bwd(const float* E, const float* dot, const float* sigma_b, int64_t B, float alpha, float beta,
                         float gamma, const float* gout, float* dE, fdsm_stream_t stream);

/* Score-net form.  s = scorenet([x+v, x-v]) viewed (2B, D); v (B, D).
 *   projected=0: per_b = w_b * sum_d[(s1+s2)^2/8 + (s1-s2)*v*k]     efficient_sm.py:86-96,
 *                ncsn/losses/sliced_sm.py:192-205, ncsnv2/losses/dsm.py:118-131
 *   projected=1: per_b = w_b * sum_d[((s1+s2)*v)^2/8 + (s1-s2)*v/2] ncsn/losses/sliced_sm.py:234-238
 *   w_b = sigma[labels[b]]^power when sigma != NULL (labels NULL => sigma[b]), else 1.
 *   loss = mean_b per_b.  `ds` (2B, D, optional) = gout[0] * dloss/ds  (gout NULL => 1).
 *   Any of per_sample / loss may be NULL (backward-only call); ds NULL => forward only;
 *   both given => the fused fwd+bwd single pass (20*D bytes per sample).                     */
int fdsm_combine_score(const float* s, int64_t lds, const float* v, int64_t ldv, const float* sigma,
                       const int64_t* labels, float power, int64_t B, int64_t D, float k, int32_t projected,
                       const float* gout, float* per_sample, float* loss, float* ds, int64_t ldds,
                       fdsm_stream_t stream);
/* data[0..n) *= g[0] unless g[0] == 1 (device scalar; no host read, graph-capturable).  The autograd wrapper of the
 * fused pass above computes ds for an upstream gradient of 1 inside the FORWARD launch and applies the actual upstream
 * gradient here: for the usual `loss.backward()` on the returned loss every block reads the scalar and exits, so the
 * score-form combine moves 20*D bytes per sample instead of the 32*D of a forward launch plus a backward launch. */
int fdsm_scale_unless_one(float* data, int64_t n, const float* g, fdsm_stream_t stream);

/* The score-form combine for callers that cannot provide (or allocate) the per-sample workspace: code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
 * Same arithmetic, kernels and launch configuration as fdsm_combine_score
 * called with (per_sample = workspace, loss, ds): *loss (required) = mean_b per_b, `ds` (optional) = gout[0] * dloss/ds;
 * the per-sample terms the mean reads live in a file-scope `static __device__` array of
 * FDSM_COMBINE_SCORE_STATIC_MAX_B floats (256 KiB, present once on every GPU that loads the library, until exit) and
 * are not returned.  It allocates nothing: no cudaMalloc / cudaFree / allocating library call, on any path.
 *   - B > FDSM_COMBINE_SCORE_STATIC_MAX_B returns FDSM_ERR_INVALID_ARGUMENT (message names the limit); use
 *     fdsm_combine_score with a caller-owned buffer for larger batches.
 *   - The first call on each device resolves the array's device address (cudaGetSymbolAddress, that device current) and
 *     caches it; make that first call (B = 0 is enough) before starting a stream capture.
 *   - Every call on a device shares the one array and it keeps the previous call's contents (every entry a call reads is
 *     written by that call first): calls that overlap on the GPU (different streams, concurrent graph branches) RACE.
 *     Serialise them on one stream or use fdsm_combine_score.                                                          */
#define FDSM_COMBINE_SCORE_STATIC_MAX_B 65536
int fdsm_combine_score_static(const float* s, int64_t lds, const float* v, int64_t ldv, const float* sigma,
                              const int64_t* labels, float power, int64_t B, int64_t D, float k, int32_t projected,
                              const float* gout, float* loss, float* ds, int64_t ldds, fdsm_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Kernel (3) building blocks: TF32 hi/lo operand split and the tcgen05/TMEM GEMM fed by TMA.
 *
 * Replaces the cuBLAS fp32 SGEMMs behind nn.Linear in the energy MLPs
 * (DeepEBM/ESM/models.py:129-160,932-955; SSM/models/dkef.py:442-461).  tcgen05 has no fp32 MMA
 * kind: operands are pre-split x = hi + lo (both TF32-representable) and passes=3 issues
 * hi*hi + hi*lo + lo*hi per k-step (fp32-grade products, fp32 accumulation in TMEM).
 * ------------------------------------------------------------------------------------------ */
/* out_hi/out_lo[r,c] = split(in[r,c]) (c < cols); column `cols` := 1 if (ones_col & 1) (bias column of an
 * augmented operand) ; remaining columns up to out_cols := 0.  out_lo may be NULL (hi only).
 * ones_col & 2: no rounding -- out_hi receives the fp32 value itself (padded copy; out_lo must be NULL): the
 * raw-operand format of the in-kernel-split GEMM. */
int fdsm_tf32_split(const float* in, int64_t ldin, int64_t rows, int64_t cols, float* out_hi, float* out_lo,
                    int64_t ldo, int64_t out_cols, int32_t ones_col, fdsm_stream_t stream);

typedef struct fdsm_gemm_args {
  /* D[M x N] = sum_k A[m,k] * B[n,k]
   * major 0 (K-major):  operand stored [rows x K] row-major (K contiguous, as nn.Linear.weight)
   * major 1 (MN-major): operand stored [K x rows] row-major (rows contiguous)                  */
  const float* a_hi; const float* a_lo; int64_t lda; int32_t a_major;
  const float* b_hi; const float* b_lo; int64_t ldb; int32_t b_major;
  int64_t M, N, K;
  int32_t passes;        /* 1: hi*hi (plain TF32) ; 3: 3xTF32.  passes = 3 with a_lo == b_lo == NULL: the operands
                            are plain fp32 and the hi/lo split happens inside the kernel (shared memory)            */
  int32_t splits;        /* split-K: partial s is written at out + s*split_stride (caller reduces) */
  int32_t tile_n;        /* UMMA N (multiple of 16, <= 256); 0 = auto                              */
  float* out; int64_t ldo; int64_t split_stride;
} fdsm_gemm_args;

/* 16-byte aligned bases and ld % 4 == 0 required (TMA); out-of-range rows/columns read as zero. */
int fdsm_tc_gemm(const fdsm_gemm_args* args, fdsm_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Kernel (3): the energy MLP (forward + backward) as tcgen05 GEMM launches with fused epilogues.
 *
 * Replaces the torch modules MLP_Quadratic_MNIST / MLP_Quadratic (DeepEBM/ESM/models.py:129-160,
 * 932-955: AF(fc_k(.)) x n_hidden, E = ln1(h)*ln2(h) + lq(h^2)) and the config-1 3-layer Softplus
 * MLP with a Linear(H,1) head (SSM/models/dkef.py:442-461 shape), and their autograd backward.
 *
 * Parameters and gradients: ONE flat fp32 vector, tightly packed in the reference's
 * named_parameters() order (fc1.weight[H,D], fc1.bias[H], ..., then ln1.weight[H], ln1.bias,
 * ln2.weight[H], ln2.bias, lq.weight[H], lq.bias -- or out.weight[H], out.bias for head = 0).
 * The input rows arrive as the TF32 hi/lo halves of the augmented batch [rows x in_ld]
 * (column in_dim = 1, the folded bias; remaining padding 0): exactly what fdsm_stack writes when
 * `out_lo` is set -- or, with presplit = 0, as the plain fp32 rows fdsm_stack writes with `pad_ones`.  All scratch lives in the caller's workspace.
 * ------------------------------------------------------------------------------------------ */
typedef struct fdsm_mlp_desc {
  int32_t in_dim;    /* D                                                         */
  int32_t hidden;    /* H; H > 256 (MLP_Quadratic: 5000) runs with several 256-wide N tiles per GEMM */
  int32_t n_hidden;  /* number of fc layers (3 for MLP_Quadratic_MNIST)           */
  int32_t act;       /* 1 softplus, 2 softplus - ln2 (NormalizedSoftplus), 3 ELU  */
  int32_t head;      /* 1 quadratic head, 0 Linear(H,1)                           */
  int32_t passes;    /* 3 = 3xTF32 (fp32-grade), 1 = plain TF32                   */
  int32_t presplit;  /* 0 (default): activations travel through HBM as plain fp32 and every GEMM splits its operands
                        into TF32 hi/lo in shared memory (x_lo ignored, may be NULL; x_hi = the fp32 rows);
                        1: activations are stored pre-split (x_hi/x_lo both required), twice the HBM traffic  */
} fdsm_mlp_desc;

/* Sizes for `rows` stacked input rows: parameter count, workspace floats, padded input leading dim. */
int fdsm_mlp_sizes(const fdsm_mlp_desc* desc, int64_t rows, int64_t* n_params, int64_t* workspace_floats,
                   int64_t* in_ld);
/* Once per optimizer step: flat parameters -> augmented transposed TF32 hi/lo weights (in workspace). */
int fdsm_mlp_prepare(const fdsm_mlp_desc* desc, int64_t rows, const float* params, float* workspace,
                     fdsm_stream_t stream);
/* energy[rows] = E(x); keeps the activations needed by the backward in the workspace. */
int fdsm_mlp_forward(const fdsm_mlp_desc* desc, const float* x_hi, const float* x_lo, int64_t rows,
                     const float* params, float* workspace, float* energy, fdsm_stream_t stream);
/* grads (flat, same layout as params) (+)= dL/dparams given dE[rows] = dL/dE; needs the workspace of the
 * matching forward.  Deterministic: fixed-order split-K reduction, fixed-order fp64 head sums (no atomics). */
int fdsm_mlp_backward(const fdsm_mlp_desc* desc, const float* x_hi, const float* x_lo, int64_t rows,
                      const float* params, float* workspace, const float* dE, float* grads, int32_t accumulate,
                      fdsm_stream_t stream);
/* FD-SSM energy form (a1/a7: SSM/losses/efficient_sm.py:65-71, DeepEBM/ESM/loss_train.py:227-233) with the combine
 * folded into the backward: `energy` = the forward's output on the stacked batch [x | x+v | x-v] (3B rows).
 *   *loss = scale * mean_b|sum_b [ (E2-E3)^2/8 - E2 - E3 + c1 E1 ]   (same arithmetic as fdsm_combine_energy_fwd)
 *   grads (+)= dloss/dparams, dE taken analytically inside the head-backward kernel (as fdsm_combine_energy_bwd).
 * Replaces the launches combine_fwd + combine_bwd + backward(dE) of the un-fused sequence.  loss may be NULL. */
int fdsm_mlp_backward_fd_ssm(const fdsm_mlp_desc* desc, const float* x_hi, const float* x_lo, int64_t B,
                             const float* params, float* workspace, const float* energy, float c1, float scale,
                             int32_t mean, float* loss, float* grads, int32_t accumulate, fdsm_stream_t stream);

/* FD-DSM energy form (a9: DeepEBM/ESM/loss_train.py:109-117; a3/a6 through alpha/beta/gamma as in fdsm_combine_dsm_fwd)
 * with the combine folded into the backward: `energy` = the forward's output on [xt+v | xt-v] (2B rows), `dot` = the
 * per-sample v.(xt - x) (or v.noise) that fdsm_stack wrote, `sigma_b` = per-sample sigma (NULL: no 1/sigma^2 weight).
 *   *loss = gamma * mean_b (alpha (E[B+b] - E[b]) + beta dot_b)^2 / sigma_b^2   (same arithmetic as fdsm_combine_dsm_fwd)
 *   grads (+)= dloss/dparams, dE taken analytically inside the head-backward kernel (as fdsm_combine_dsm_bwd).
 * Replaces the launches combine_dsm_fwd + combine_dsm_bwd + backward(dE).  loss may be NULL. */
int fdsm_mlp_backward_fd_dsm(const fdsm_mlp_desc* desc, const float* x_hi, const float* x_lo, int64_t B,
                             const float* params, float* workspace, const float* energy, const float* dot,
                             const float* sigma_b, float alpha, float beta, float gamma, float* loss, float* grads,
                             int32_t accumulate, fdsm_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Kernel (3) in difference-propagating form: the whole FD-SSM energy-form loss (a1 / a5 / a7) AND its parameter gradient
 * for the energy MLPs above in one call -- direction sampling, forward, head + combine, backward.
 *
 * Instead of the stacked batch [x | x+v | x-v] the network carries, per sample, (h0, a, b) with h(x+-v) = h0 + b +- a
 * (a = odd part ~ eps, b = even part ~ eps^2; csrc/fd_jet.cuh): the same three function values the reference evaluates
 * (DeepEBM/ESM/loss_train.py:222-223, SSM/losses/efficient_sm.py:59-60), with rounding relative to the differences, so
 * single TF32 operands keep loss and gradient within ~1e-4 of the float64 value where the reference's own fp32 stacked
 * arithmetic is off by 1e-2 .. 1 at the BASELINE configs.  One launch (csrc/fd_chain.cuh) runs every layer forward, the head,
 * the per-sample loss and all data-gradient GEMMs for a tile of samples; the weight gradients follow as split-K GEMMs.
 *   *loss = scale * mean_b|sum_b [ (E+ - E-)^2/8 - (E+ + E- - 2 E0) + (c1 - 2) E0 ]
 *   grads (+)= dloss/dparams   (flat, the layout of fdsm_mlp_*)
 * desc->passes: 3 = every GEMM with its weights as TF32 hi + lo (two MMA passes; the input layer's weights travel once as
 * fp32 and lo is derived on chip); 1 = weights rounded to TF32, one pass everywhere; any other value (default, 2) = hi + lo
 * for the hidden layers, round-to-nearest TF32 weights for the input layer (config 2 against fp64: loss / gradient within
 * 2e-5 / 9e-5, passes = 3: 2e-6 / 6e-5, passes = 1: 1e-4 / 2e-4); 4 = as 2 with the input layer in kind::f16: fp16 copies of
 * W_0 and of the rows [x | v] carry TF32's 11 significant bits in half the bytes (the input layer is bound by that stream;
 * same measured errors as 2), valid while |x| and |W_0| stay below 65504 -- beyond that the loss turns NaN, never silently
 * wrong.  desc->presplit is ignored.  Supported: hidden <= 223, n_hidden <= 5 (fdsm_jet_supported).
 * Deterministic: fixed-order tile, split-K and head reductions.  Graph-capturable (uses internal side streams by
 * event fork/join).
 * ------------------------------------------------------------------------------------------ */
/* Optional fused tail of the single-GPU training step (DeepEBM/train_ssm.py:65-70: clip_grad_norm_ + Adam.step): the
 * fixed-order sums of the split-K partials and of the per-tile head records, the squared gradient norm, the clip
 * coefficient and the Adam update run as ONE launch (with a grid-wide barrier) after the weight-gradient GEMMs (instead of n_hidden
 * reduces + finish + fdsm_grad_sumsq + fdsm_adam_step).  `grads` is still written.  The arithmetic is fdsm_adam_step's;
 * *step_counter is read at entry and incremented by one at the end, as fdsm_grad_sumsq does. */
typedef struct fdsm_jet_update {
  float* exp_avg; float* exp_avg_sq;               /* Adam moments, n_params each                                         */
  float lr; const float* lr_dev;                   /* learning rate: *lr_dev if non-NULL (schedules write it), else lr    */
  float beta1, beta2, eps, max_norm;               /* max_norm <= 0: no clipping                                          */
  int64_t* step_counter;                           /* device scalar                                                       */
  float* scratch;                                  /* >= 1024 floats, zeroed once by the caller, then owned by the library */
} fdsm_jet_update;
typedef struct fdsm_jet_args {
  const float* x; int64_t ldx; int64_t B;   /* samples (B, D)                                                          */
  const float* dir; int64_t lddir;          /* injected direction (B, D) or NULL => Philox stream 0 (as fdsm_stack)     */
  int32_t normalize_dir; float dir_scale;   /* 1: v = u/||u|| * dir_scale ; 0: v = u as given (u = dir, or the Philox rows) */
  uint64_t seed, offset, row0;              /* Philox key / draw counter / global index of local row 0                  */
  const uint64_t* offset_dev;               /* optional device counter added to `offset` (graph replay)                 */
  const float* params; float* workspace;    /* flat parameters ; fdsm_jet_sizes() floats, 16-byte aligned                */
  float c1, scale; int32_t mean;
  float* loss; float* grads; int32_t accumulate;   /* loss may be NULL                                                  */
  void* ev_chain_start; void* ev_chain_stop;       /* optional cudaEvent_t pair recorded on `stream` around the chain kernel
                                                      (measurement: bench.py's roofline of the dominant kernel)         */
  void* trace;                                     /* optional debug buffer: tiles x 38 uint64 %globaltimer stamps of the chain
                                                      kernel's phases (scripts/chain_trace.py); NULL in production      */
  const uint8_t* x_u8; float x_u8_scale; float x_noise_std;
                                                   /* x_u8 != NULL replaces x (ldx = its row stride in BYTES): the samples are
                                                      x_u8 * x_u8_scale + x_noise_std * N(0,1) (Philox stream 1, row / offset as
                                                      the direction) -- torchvision's ToTensor (1/255) and the on-device noise
                                                      of DeepEBM/train_ssm.py:62 fused into the input build, so that the host
                                                      ships one byte per pixel                                           */
  int32_t mode;                                    /* 0: the finite-difference objective above.  1: EXACT sliced score matching
                                                      (SSM/losses/sliced_sm.py:7-34; DeepEBM/ESM/loss_train.py:173-189):
                                                        *loss = scale * mean_b|sum_b [ (u.grad E)^2 / 2 - u' Hess E u ]
                                                      with u = the direction as given (dir, or Philox N(0,1) rows when dir is
                                                      NULL; normalize_dir / dir_scale apply as above), evaluated in forward
                                                      mode: the three streams carry (h, dh/dtau, 1/2 d2h/dtau2) of
                                                      tau -> h(x + tau u) through the same kernel; c1 is ignored           */
  const fdsm_jet_update* update;                   /* NULL: loss + gradient only.  Non-NULL: the parameters (`params`, written
                                                      in place) take the clipped Adam step in the same call              */
} fdsm_jet_args;
int fdsm_jet_supported(const fdsm_mlp_desc* desc);
/* offsets (optional, >= 8 + 4 * n_hidden entries): test view of the workspace -- [0] samples per tile S, [1] padded rows per
 * stream Bp, [2] npad, [3] kpad0, [4] kpadh, [5] XV offset, [6] passes, [7] tiles, then per layer l: [8+4l] Wt_l,
 * [9+4l] H_l (l >= 1; [9] = CTAs per cluster of the chain kernel), [10+4l] dZ_l (float offsets), [11+4l] split-K count of
 * the weight-gradient GEMM. */
int fdsm_jet_sizes(const fdsm_mlp_desc* desc, int64_t B, int64_t* n_params, int64_t* workspace_floats, int64_t* offsets);
int fdsm_jet_fd_ssm(const fdsm_mlp_desc* desc, const fdsm_jet_args* args, fdsm_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Optimizer step on the flat parameter vector (the reference's host loop, DeepEBM/train_ssm.py:65-70:
 * clip_grad_norm_(params, clip) then Adam.step()).  Per-step scalars come from device memory so the
 * step replays from one CUDA graph.
 * ------------------------------------------------------------------------------------------ */
/* sumsq[0..FDSM_SUMSQ_PARTIALS) = partial sums of g^2 over fixed contiguous chunks (fixed order: bit-identical on every
 * rank); their sum is the squared norm.  *step_counter += 1 if given. */
#define FDSM_SUMSQ_PARTIALS 32
int fdsm_grad_sumsq(const float* g, int64_t n, float* sumsq, int64_t* step_counter, fdsm_stream_t stream);
/* torch.optim.Adam semantics (no weight decay, no amsgrad) at t = *step_counter; gradient pre-scaled by
 * min(1, max_norm/(sqrt(sumsq)+1e-6)) when max_norm > 0; lr_dev (device scalar) overrides lr when non-NULL. */
int fdsm_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                   const float* lr_dev, float beta1, float beta2, float eps, float max_norm, const float* sumsq,
                   const int64_t* step_counter, fdsm_stream_t stream);
/* The host's read of the step's result (`tloss.item()`, DeepEBM/train_ssm.py:84; every `loss.item()` of the runners):
 * n <= 4096 floats from device memory `src` stored by a kernel into `h_dst`, PINNED host memory (cudaHostAlloc /
 * torch pin_memory: device-mapped under unified addressing; anything else is rejected), ordered on `stream` -- the host
 * reads them after synchronising with the stream.  In place of cudaMemcpyAsync(D2H) of 4 bytes: a copy-engine operation
 * between two graph replays idles the stream on both sides of it (bench.py e2e: 36 us per step at config 2). */
int fdsm_publish_scalars(const float* src, float* h_dst, int64_t n, fdsm_stream_t stream);
/* Host plumbing of the end-to-end step (`x_real = itr.__next__().to(device)`, DeepEBM/train_ssm.py:61), one call each instead of
 * four stream / event operations from the host language -- on a busy host the step loop is otherwise host-bound:
 * fdsm_stage_h2d: copy_stream waits for the cudaEvent_t ev_free (may be NULL: the previous consumer of dst), copies nbytes from
 *   PINNED host memory h_src to dst, records the cudaEvent_t ev_ready; main_stream (where the step follows) waits for it.
 * fdsm_publish_scalars_record: records the cudaEvent_t ev (may be NULL) on stream, then fdsm_publish_scalars. */
int fdsm_stage_h2d(void* dst, const void* h_src, int64_t nbytes, fdsm_stream_t copy_stream, void* ev_free, void* ev_ready,
                   fdsm_stream_t main_stream);
int fdsm_publish_scalars_record(const float* src, float* h_dst, int64_t n, fdsm_stream_t stream, void* ev);

/* ------------------------------------------------------------------------------------------
 * The step's single exchange over NVLink peer memory: all-reduce of the flat fp32 gradient
 * (sum over ranks in rank order, times `scale`), fused with the squared-norm partials fdsm_adam_step reads and the
 * step-counter increment.  Replaces the reference's nn.DataParallel gradient reduction
 * (DeepEBM/train_ssm.py:33-34) and, in this package, grad.mul_(1/W) + ncclAllReduce + fdsm_grad_sumsq.
 *
 * peer_buffers[r] = address, valid on THIS device, of rank r's receive buffer (fdsm_peer_buffer_bytes(n, world)
 * bytes, zero-filled once before the first call on any rank; e.g. torch symmetric memory or cudaIpc mappings).
 * scratch: >= 1024 bytes of LOCAL device memory, zero-filled once (words [192, 204): debug phase timestamps of block 0).
 * scratch[2] != 0 after a launch means a peer did not arrive within 20 s: the kernel gives up instead of hanging the GPU,
 * publishes NaN squared-norm partials (a clipped update then yields NaN parameters rather than a silently partial sum)
 * and leaves the step counter alone; the status is sticky -- callers poll it (ddp.PeerAllReduce.timed_out) and abort.
 * Wire format: 16-byte lines {g[2w], epoch, g[2w+1], epoch}; each 8-byte half is validated by its own flag (the store
 * atomicity NCCL's LL protocol relies on), nothing is assumed about 16-byte stores as a unit.
 * out may alias grad.  Every rank must call it the same number of times with the same n.  Graph-capturable.
 * Results are bit-identical on every rank. */
#define FDSM_PEER_MAX_WORLD 8
#define FDSM_PEER_MAX_BLOCKS 160
int64_t fdsm_peer_buffer_bytes(int64_t n, int32_t world);
int fdsm_peer_allreduce(const float* grad, float* out, int64_t n, void* const* peer_buffers, int32_t rank,
                        int32_t world, float scale, float* sumsq, int64_t* step_counter, uint32_t* scratch,
                        fdsm_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * SURVEY.md §8f rows N3 / N4 (the callers either side of the hot path).
 * fdsm_langevin_step: out = (x + step_size * score) + noise * noise_scale, the reference's rounding order
 * (ncsnv2/models/__init__.py:34-39 with noise_scale = sqrt(2 step_size); DeepEBM/ESM/sampling_mdsm.py:63-96).
 * noise == NULL draws N(0,1) in-kernel from the Philox noise stream keyed by (seed, offset, row0 + row).  out may alias x.
 * fdsm_ema_update: shadow = one_minus_mu * params + mu * shadow (ncsnv2/models/ema.py:16-21; the caller rounds 1 - mu
 * from double, as Python does in the reference).
 * ------------------------------------------------------------------------------------------ */
int fdsm_langevin_step(const float* x, int64_t ldx, const float* score, int64_t lds, const float* noise, int64_t ldnoise,
                       float* out, int64_t ldo, int64_t B, int64_t D, float step_size, float noise_scale, uint64_t seed,
                       uint64_t offset, uint64_t row0, fdsm_stream_t stream);
int fdsm_ema_update(float* shadow, const float* params, int64_t n, float one_minus_mu, float mu, fdsm_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * SURVEY.md §8f row N1, first piece: InstanceNorm++ fused with the ELU that follows it, forward and backward
 * (InstanceNorm2dPlus.forward, ncsnv2/models/normalization.py:163-176; the conditional variant whose
 * [gamma | alpha | beta] rows come from an embedding of the noise level, ncsn/models/cond_refinenet_dilated.py:304-333
 * = ncsn/models/normalization.py ConditionalInstanceNorm2dPlus; `act(normalize(x))` of the residual blocks,
 * ncsnv2/models/layers.py:401-456).  The convolutions between them stay on cuDNN.
 *
 *   mu, var  = mean / biased variance of each (n, c) plane of HW elements;  rstd = 1/sqrt(var + 1e-5)
 *   m, v     = mean / UNBIASED variance over c of mu[n, :];                 mustd = (mu - m)/sqrt(v + 1e-5)
 *   out      = act(gamma * ((x - mu) * rstd + mustd * alpha) + beta),       act = identity (0) or ELU (1)
 *
 * x, out, gout, dx: contiguous (N, C, HW) fp32.  gamma / alpha / beta: rows of C floats, row n at `+ n * param_stride`
 * (param_stride = 0: one shared row, the unconditional module); beta may be NULL (bias=False).
 * stats: fdsm_inorm_plus_stats_floats(N, C) floats written by the forward ([mu | rstd | mustd | 1/sqrt(v + eps)]) and read
 * by the backward.  Backward: per-sample parameter gradients dgamma, dalpha, dbeta (N, C each; the caller sums over n for
 * a shared row), dx (may be NULL: parameters only), scratch = N * C floats.  `out` is read only when act = 1.
 * ------------------------------------------------------------------------------------------ */
int64_t fdsm_inorm_plus_stats_floats(int64_t N, int64_t C);
/* Kernel launches one call makes (two per chunk of samples; one chunk unless FDSM_NORM_CHUNK_MB > 0 asks for L2-sized
 * chunks -- measured slower on B200, see csrc/fd_norm.cu) -- for launch counters. */
int64_t fdsm_inorm_plus_launches(int64_t N, int64_t C, int64_t HW, int32_t backward, int32_t act, int32_t with_dx);
int fdsm_inorm_plus_fwd(const float* x, int64_t N, int64_t C, int64_t HW, const float* gamma, const float* alpha,
                        const float* beta, int64_t param_stride, int32_t act, float* out, float* stats, fdsm_stream_t stream);
int fdsm_inorm_plus_bwd(const float* x, const float* out, const float* gout, int64_t N, int64_t C, int64_t HW,
                        const float* gamma, const float* alpha, int64_t param_stride, int32_t act, const float* stats,
                        float* scratch, float* dx, float* dgamma, float* dalpha, float* dbeta, fdsm_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* FDSM_H_ */
