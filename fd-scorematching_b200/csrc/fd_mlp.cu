// Kernel (3), assembled: forward and backward of the reference's energy MLPs
//   MLP_Quadratic_MNIST / MLP_Quadratic (DeepEBM/ESM/models.py:129-160,932-955): AF(fc_k(.)) x L, E = ln1(h)*ln2(h) + lq(h^2)
//   3-layer Softplus MLP of config 1 (SSM/models/dkef.py:442-461 shape, out_features = 1): linear head
// as a sequence of tcgen05 GEMM launches (tc_gemm.cuh) plus three small CUDA-core kernels.
//
// Data layout (all fp32, caller-owned workspace; see Layout below):
//   parameters  flat, TIGHTLY packed in the reference's named_parameters order:
//               fc1.weight[H x D] fc1.bias[H] ... fcL.weight[H x H] fcL.bias[H] then the head
//               (ln1.weight[H] ln1.bias ln2.weight[H] ln2.bias lq.weight[H] lq.bias | out.weight[H] out.bias)
//   Wt_l        per layer, TF32 hi/lo halves of the AUGMENTED TRANSPOSED weight [Kpad_l x Npad]:
//               rows 0..K_l-1 = W_l^T, row K_l = bias_l, remaining rows/columns zero.  The layer input carries a
//               ones column at index K_l, so the bias add is folded into the GEMM and the weight-gradient GEMM
//               emits the bias gradient as row K_l.
//   activations hi/lo halves [rows x Kpad] written by the previous GEMM's epilogue (or by fdsm_stack for layer 1)
// Forward  layer l: A = input (K-major), B = Wt_l (MN-major)  -> act -> hi/lo (+ones column)
// Backward dgrad l: A = dZ_l (K-major over n), B = Wt_l (K-major: rows k, contiguous n) -> * act'(H_{l-1}) -> dZ_{l-1}
//          wgrad l: A = input_l (MN-major, M = K_l+1), B = dZ_l (MN-major), contraction over the rows, split-K
//                   partials -> deterministic fixed-order reduce + transpose into the flat gradient.
#include "tc_gemm_host.cuh"

namespace fdsm {
namespace mlp {

inline long long round_up(long long x, long long m) { return (x + m - 1) / m * m; }

struct Layout {
  int D, H, L, act, head, passes;
  bool raw;            // presplit == 0 and passes == 3: plain fp32 activations, split inside the GEMMs
  long long rows;
  int npad;            // UMMA N of every hidden layer: ceil16(H)
  int kpad0, kpadh;    // padded augmented input widths: ceil32(D+1), ceil32(H+1)
  // parameter offsets
  long long w_off[16], b_off[16], head_off, n_params;
  // workspace offsets (floats)
  long long wt_hi[16], wt_lo[16], wt_raw[16];
  long long h_hi[16], h_lo[16];     // hidden activations 1..L-1 (hi/lo, ld kpadh)
  long long h_last;                 // H_L fp32 [rows x npad]
  long long dz_hi[16], dz_lo[16];   // dZ_l per layer [rows x npad] (lo only in presplit mode): wgrad_l runs on a side
                                    // stream while dgrad_l..1 continue, so the buffers are not recycled
  long long l1, l2;                 // per-row head dot products
  long long head_acc;               // fp64 accumulators of the head-parameter gradients (3(H+1) doubles)
  long long partial[2];             // split-K partials: [0] input layer (main stream), [1] hidden layers (side stream)
  long long total;
  int splits[16];
};

inline int make_layout(const fdsm_mlp_desc* d, long long rows, Layout& y) {
  FDSM_REQUIRE(d != nullptr, "fdsm_mlp: desc is NULL");
  FDSM_REQUIRE(d->in_dim >= 1 && d->hidden >= 1 && d->n_hidden >= 1 && d->n_hidden <= 15, "fdsm_mlp: bad sizes");
  FDSM_REQUIRE(d->hidden <= 256, "fdsm_mlp: hidden=%d > 256 is not supported by the fused head epilogue yet", d->hidden);
  FDSM_REQUIRE(d->act >= 1 && d->act <= 3, "fdsm_mlp: act must be 1 (softplus), 2 (nsoftplus) or 3 (elu)");
  FDSM_REQUIRE(d->head == 0 || d->head == 1, "fdsm_mlp: head must be 0 (linear) or 1 (quadratic)");
  FDSM_REQUIRE(d->passes == 1 || d->passes == 3, "fdsm_mlp: passes must be 1 or 3");
  FDSM_REQUIRE(rows >= 0 && rows < (1ll << 31), "fdsm_mlp: bad row count");
  y.D = d->in_dim; y.H = d->hidden; y.L = d->n_hidden; y.act = d->act; y.head = d->head; y.passes = d->passes;
  y.raw = d->presplit == 0 && d->passes == 3;
  FDSM_REQUIRE(d->presplit == 0 || d->presplit == 1, "fdsm_mlp: presplit must be 0 or 1");
  y.rows = rows;
  y.npad = (int)round_up(y.H, 16);
  y.kpad0 = (int)round_up(y.D + 1, 32);
  y.kpadh = (int)round_up(y.H + 1, 32);
  long long o = 0;
  for (int l = 0; l < y.L; ++l) {
    const int K = l ? y.H : y.D;
    y.w_off[l] = o; o += (long long)y.H * K;
    y.b_off[l] = o; o += y.H;
  }
  y.head_off = o;
  o += y.head ? 3ll * (y.H + 1) : (y.H + 1);
  y.n_params = o;
  // workspace
  long long w = 0;
  auto take = [&](long long n) { long long r = w; w += round_up(n, 64); return r; };   // 256-byte aligned sections
  long long max_partial[2] = {0, 0};
  for (int l = 0; l < y.L; ++l) {
    const int kp = l ? y.kpadh : y.kpad0;
    y.wt_hi[l] = take((long long)kp * y.npad);
    y.wt_lo[l] = take((long long)kp * y.npad);
    y.wt_raw[l] = take((long long)kp * y.npad);
    // wgrad: M = K_l + 1 rows of dWt in tiles of 128, contraction over `rows` in k-blocks of 32
    const int mt = (int)((l ? y.H : y.D) + 1 + tc::BM - 1) / tc::BM;
    const long long kblocks = (rows + tc::BK - 1) / tc::BK;
    long long s = (2ll * sm_count() / 2) / mt;   // ~one CTA per SM
    if (s > kblocks) s = kblocks;
    if (s < 1) s = 1;
    y.splits[l] = (int)s;
    const long long ps = s * (long long)kp * y.npad;
    if (ps > max_partial[l ? 1 : 0]) max_partial[l ? 1 : 0] = ps;
  }
  for (int l = 1; l < y.L; ++l) {
    y.h_hi[l] = take(rows * y.kpadh);
    y.h_lo[l] = take(rows * y.kpadh);
  }
  y.h_last = take(rows * y.npad);
  for (int l = 0; l < y.L; ++l) {   // dZ_l = dL/d(pre-activation of layer l)
    y.dz_hi[l] = take(rows * y.npad);
    y.dz_lo[l] = y.raw ? y.dz_hi[l] : take(rows * y.npad);
  }
  y.l1 = take(rows);
  y.l2 = take(rows);
  y.head_acc = take(2ll * 3 * (y.H + 1));
  y.partial[0] = take(max_partial[0]);
  y.partial[1] = take(max_partial[1]);
  y.total = w;
  return FDSM_OK;
}

// ---- weight preparation: flat parameters -> augmented transposed TF32 hi/lo (all layers in one launch) ----------
struct PrepArgs {
  const float* W[16]; const float* b[16];
  float* hi[16]; float* lo[16]; float* raw[16];
  int K[16], kpad[16];
  int H, npad;
};
__global__ void __launch_bounds__(256) prep_weights_kernel(const PrepArgs a) {
  // 32x32 tile transpose through shared memory: coalesced reads of W[n][k] (k fastest) and writes of Wt[k][n]
  __shared__ float tile[32][33];
  const int l = blockIdx.z;
  const int K = a.K[l], kpad = a.kpad[l], H = a.H, npad = a.npad;
  const int k0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
  if (k0 >= kpad) return;   // uniform per block
  const float* __restrict__ W = a.W[l];
  const float* __restrict__ b = a.b[l];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  for (int i = ty; i < 32; i += 8) {
    const int n = n0 + i, k = k0 + tx;
    float v = 0.f;
    if (n < H) {
      if (k < K) v = W[(long long)n * K + k];
      else if (k == K) v = b[n];
    }
    tile[i][tx] = v;
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int k = k0 + i, n = n0 + tx;
    if (k < kpad && n < npad) {
      float hi, lo;
      ptx::tf32_split(tile[tx][i], hi, lo);
      a.hi[l][(long long)k * npad + n] = hi;
      a.lo[l][(long long)k * npad + n] = lo;
      a.raw[l][(long long)k * npad + n] = tile[tx][i];
    }
  }
}

// ---- head backward: dE -> dZ_L (hi/lo) and the head parameter gradients -----------------------------------
// dH[r,j] = g (l2 w1[j] + l1 w2[j] + 2 wq[j] h)   (quadratic)   |   g w[j]   (linear);   dZ = dH * act'(h)
// The head-parameter sums cancel heavily across the x / x+v / x-v rows of a sample (dE = 2c, c(q-1), c(-q-1)), so they
// are accumulated in fp64 (registers, then fp64 atomics into `acc`) and rounded to fp32 once by head_finish_kernel.
constexpr int HB_ROWS = 32;
__global__ void __launch_bounds__(256) head_bwd_kernel(const float* __restrict__ hlast, int ldh, const float* __restrict__ dE,
                                                       const float* __restrict__ l1v, const float* __restrict__ l2v,
                                                       const float* __restrict__ head_w, int H, int npad, long long rows,
                                                       int act, int head, float* __restrict__ dz_hi, float* __restrict__ dz_lo,
                                                       double* __restrict__ acc) {
  const int j = threadIdx.x;
  const long long r0 = (long long)blockIdx.x * HB_ROWS;
  const bool col = j < H;
  float w1 = 0.f, w2 = 0.f, wq = 0.f;
  if (col) {
    w1 = head_w[j];
    if (head) { w2 = head_w[H + 1 + j]; wq = head_w[2 * (H + 1) + j]; }
  }
  double a1 = 0., a2 = 0., aq = 0., sb1 = 0., sb2 = 0., sbq = 0.;
#pragma unroll 4
  for (int i = 0; i < HB_ROWS; ++i) {
    const long long r = r0 + i;
    if (r >= rows) break;
    const float g = __ldg(dE + r);
    float dz = 0.f;
    if (col) {
      const float h = __ldg(hlast + r * ldh + j);
      float dh;
      if (head) {
        const float ga = g * __ldg(l2v + r), gb = g * __ldg(l1v + r);
        dh = fmaf(ga, w1, fmaf(gb, w2, 2.f * g * wq * h));
        a1 += (double)ga * h; a2 += (double)gb * h; aq += (double)g * h * h;
        if (j == 0) { sb1 += ga; sb2 += gb; sbq += g; }
      } else {
        dh = g * w1;
        a1 += (double)g * h;
        if (j == 0) sb1 += g;
      }
      dz = dh * tc::act_bwd_from_h(act, h);
    }
    if (j < npad) {
      float hi, lo;
      ptx::tf32_split(dz, hi, lo);
      dz_hi[r * npad + j] = dz_lo ? hi : dz;
      if (dz_lo) dz_lo[r * npad + j] = lo;
    }
  }
  if (col) {
    atomicAdd(acc + j, a1);
    if (head) { atomicAdd(acc + H + 1 + j, a2); atomicAdd(acc + 2 * (H + 1) + j, aq); }
    if (j == 0) {
      atomicAdd(acc + H, sb1);
      if (head) { atomicAdd(acc + 2 * H + 1, sb2); atomicAdd(acc + 3 * H + 2, sbq); }
    }
  }
}

__global__ void __launch_bounds__(256) head_finish_kernel(const double* __restrict__ acc, int n, float* __restrict__ ghead,
                                                          int accumulate) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) ghead[i] = accumulate ? ghead[i] + (float)acc[i] : (float)acc[i];
}

// ---- split-K reduce: partial[s][k][n] -> gW[n][k] (+)= sum_s, gb[n] (+)= sum_s partial[s][K][n] --------------
// 8 (k) x 32 (n) outputs per 1024-thread block; 4 thread groups each sum a fixed quarter of the splits in fp64, the
// quarters are combined in a fixed order (deterministic), then a transposing store through shared memory.
constexpr int RW_GROUPS = 4;
__global__ void __launch_bounds__(1024) reduce_wgrad_kernel(const float* __restrict__ partial, int splits, int kpad, int npad,
                                                            int H, int K, float* __restrict__ gW, float* __restrict__ gb,
                                                            int accumulate) {
  __shared__ double part[RW_GROUPS][8][32];
  __shared__ float tile[8][33];
  const int k0 = blockIdx.x * 8, n0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = (threadIdx.x >> 5) & 7, tz = threadIdx.x >> 8;
  const long long stride = (long long)kpad * npad;
  {
    const int k = k0 + ty, n = n0 + tx;
    double acc = 0.;
    if (k <= K && n < H) {
      const int per = (splits + RW_GROUPS - 1) / RW_GROUPS;
      const int s0 = tz * per, s1 = min(splits, s0 + per);
      const float* p = partial + (long long)k * npad + n;
#pragma unroll 4
      for (int s = s0; s < s1; ++s) acc += (double)__ldg(p + s * stride);
    }
    part[tz][ty][tx] = acc;
  }
  __syncthreads();
  if (tz == 0) {
    double acc = part[0][ty][tx];
#pragma unroll
    for (int g = 1; g < RW_GROUPS; ++g) acc += part[g][ty][tx];
    tile[ty][tx] = (float)acc;
  }
  __syncthreads();
  if (threadIdx.x < 256) {
    // 32 n x 8 k, k fastest so the stores to gW[n][k] run along k
    const int kk = threadIdx.x & 7, nn = threadIdx.x >> 3;
    const int n = n0 + nn, k = k0 + kk;
    if (n < H) {
      const float v = tile[kk][nn];
      if (k < K) {
        float* o = gW + (long long)n * K + k;
        *o = accumulate ? *o + v : v;
      } else if (k == K) {
        gb[n] = accumulate ? gb[n] + v : v;
      }
    }
  }
}

// Per-host-thread side stream + fork/join events (created once, never destroyed; no device memory involved).
struct SideStream {
  cudaStream_t stream = nullptr;
  cudaEvent_t fork[2] = {nullptr, nullptr}, join = nullptr;
  bool ok() const { return stream && fork[0] && fork[1] && join; }
};
inline SideStream& side_stream() {
  static thread_local SideStream s;
  static thread_local bool tried = false;
  if (!tried) {
    tried = true;
    if (cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess) s.stream = nullptr;
    for (auto* e : {&s.fork[0], &s.fork[1], &s.join})
      if (cudaEventCreateWithFlags(e, cudaEventDisableTiming) != cudaSuccess) *e = nullptr;
    (void)cudaGetLastError();
  }
  return s;
}

}  // namespace mlp
}  // namespace fdsm

using namespace fdsm;
using mlp::Layout;

extern "C" int fdsm_mlp_sizes(const fdsm_mlp_desc* desc, int64_t rows, int64_t* n_params, int64_t* workspace_floats,
                              int64_t* in_ld) {
  Layout y;
  int rc = mlp::make_layout(desc, rows, y);
  if (rc) return rc;
  if (n_params) *n_params = y.n_params;
  if (workspace_floats) *workspace_floats = y.total;
  if (in_ld) *in_ld = y.kpad0;
  return FDSM_OK;
}

extern "C" int fdsm_mlp_prepare(const fdsm_mlp_desc* desc, int64_t rows, const float* params, float* workspace,
                                fdsm_stream_t stream) {
  Layout y;
  int rc = mlp::make_layout(desc, rows, y);
  if (rc) return rc;
  FDSM_REQUIRE(params && workspace, "fdsm_mlp_prepare: NULL pointer");
  cudaStream_t st = (cudaStream_t)stream;
  mlp::PrepArgs a;
  a.H = y.H; a.npad = y.npad;
  int kmax = 0;
  for (int l = 0; l < y.L; ++l) {
    a.W[l] = params + y.w_off[l]; a.b[l] = params + y.b_off[l];
    a.hi[l] = workspace + y.wt_hi[l]; a.lo[l] = workspace + y.wt_lo[l]; a.raw[l] = workspace + y.wt_raw[l];
    a.K[l] = l ? y.H : y.D; a.kpad[l] = l ? y.kpadh : y.kpad0;
    if (a.kpad[l] > kmax) kmax = a.kpad[l];
  }
  dim3 grid((kmax + 31) / 32, (y.npad + 31) / 32, y.L);
  mlp::prep_weights_kernel<<<grid, 256, 0, st>>>(a);
  FDSM_LAUNCH_CHECK("fdsm_mlp_prepare");
}

extern "C" int fdsm_mlp_forward(const fdsm_mlp_desc* desc, const float* x_hi, const float* x_lo, int64_t rows,
                                const float* params, float* workspace, float* energy, fdsm_stream_t stream) {
  Layout y;
  int rc = mlp::make_layout(desc, rows, y);
  if (rc) return rc;
  if (rows == 0) return FDSM_OK;
  FDSM_REQUIRE(x_hi && (x_lo || y.passes == 1 || y.raw) && params && workspace && energy, "fdsm_mlp_forward: NULL pointer");
  auto lo = [&](float* q) -> float* { return y.raw ? nullptr : q; };
  cudaStream_t st = (cudaStream_t)stream;
  float* ws = workspace;
  for (int l = 0; l < y.L; ++l) {
    const bool last = l == y.L - 1;
    const int K = (l ? y.H : y.D) + 1;
    tc::Operand A = l ? tc::Operand{ws + y.h_hi[l], lo(ws + y.h_lo[l]), y.kpadh, tc::MAJOR_K}
                      : tc::Operand{x_hi, y.raw ? nullptr : x_lo, y.kpad0, tc::MAJOR_K};
    tc::Operand B = y.raw ? tc::Operand{ws + y.wt_raw[l], nullptr, y.npad, tc::MAJOR_MN}
                          : tc::Operand{ws + y.wt_hi[l], ws + y.wt_lo[l], y.npad, tc::MAJOR_MN};
    tc::GemmParams p = {};
    p.act = y.act;
    if (!last) {
      p.epi = tc::EPI_ACT_SPLIT;
      p.out_hi = ws + y.h_hi[l + 1]; p.out_lo = lo(ws + y.h_lo[l + 1]); p.ldo = y.kpadh; p.out_cols = y.kpadh;
    } else {
      p.epi = tc::EPI_HEAD;
      p.out_hi = ws + y.h_last; p.ldo = y.npad; p.out_cols = y.npad;
      p.head_w = params + y.head_off; p.head_mode = y.head;
      p.e_out = energy; p.l1_out = ws + y.l1; p.l2_out = ws + y.l2;
    }
    rc = tc::launch_gemm(p, A, B, (int)rows, y.H, K, y.npad, y.passes, 1, st);
    if (rc) return rc;
  }
  return FDSM_OK;
}

extern "C" int fdsm_mlp_backward(const fdsm_mlp_desc* desc, const float* x_hi, const float* x_lo, int64_t rows,
                                 const float* params, float* workspace, const float* dE, float* grads, int32_t accumulate,
                                 fdsm_stream_t stream) {
  Layout y;
  int rc = mlp::make_layout(desc, rows, y);
  if (rc) return rc;
  FDSM_REQUIRE(params && workspace && grads, "fdsm_mlp_backward: NULL pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (rows == 0) {
    if (!accumulate) return cuda_status(cudaMemsetAsync(grads, 0, y.n_params * sizeof(float), st), "fdsm_mlp_backward");
    return FDSM_OK;
  }
  FDSM_REQUIRE(x_hi && (x_lo || y.passes == 1 || y.raw) && dE, "fdsm_mlp_backward: NULL pointer");
  float* ws = workspace;
  auto lo = [&](float* q) -> float* { return y.raw ? nullptr : q; };
  const int head_n = (int)(y.n_params - y.head_off);
  double* hacc = reinterpret_cast<double*>(ws + y.head_acc);
  rc = cuda_status(cudaMemsetAsync(hacc, 0, head_n * sizeof(double), st), "fdsm_mlp_backward(memset)");
  if (rc) return rc;
  // Two streams: the critical chain head_bwd -> dgrad_{L-1} -> ... -> dgrad_1 -> wgrad_0 stays on `stream`; the hidden
  // layers' weight gradients (wgrad_l + reduce_l, l >= 1) only need dZ_l and run on a side stream, filling the SMs the
  // 96-CTA dgrad launches leave idle and hiding their launch/fill/drain latencies.  Fork/join by events: capturable.
  mlp::SideStream& side = mlp::side_stream();
  const bool forked = y.L > 1 && side.ok();
  cudaStream_t sd = forked ? side.stream : st;
  const int L1 = y.L - 1;
  mlp::head_bwd_kernel<<<(unsigned)((rows + mlp::HB_ROWS - 1) / mlp::HB_ROWS), 256, 0, st>>>(
      ws + y.h_last, y.npad, dE, ws + y.l1, ws + y.l2, params + y.head_off, y.H, y.npad, rows, y.act, y.head,
      ws + y.dz_hi[L1], lo(ws + y.dz_lo[L1]), hacc);
  mlp::head_finish_kernel<<<(head_n + 255) / 256, 256, 0, st>>>(hacc, head_n, grads + y.head_off, accumulate);
  rc = cuda_status(cudaGetLastError(), "fdsm_mlp_backward(head)");
  if (rc) return rc;
  auto wgrad = [&](int l, cudaStream_t s) -> int {
    // dWt_l[K+1 x H] = In_l^T dZ_l: split-K partials, then a fixed-order reduce + transpose into the flat gradient
    const int K = l ? y.H : y.D, kp = l ? y.kpadh : y.kpad0;
    tc::Operand In = l ? tc::Operand{ws + y.h_hi[l], lo(ws + y.h_lo[l]), y.kpadh, tc::MAJOR_MN}
                       : tc::Operand{x_hi, y.raw ? nullptr : x_lo, y.kpad0, tc::MAJOR_MN};
    tc::Operand B{ws + y.dz_hi[l], lo(ws + y.dz_lo[l]), y.npad, tc::MAJOR_MN};
    float* partial = ws + y.partial[l ? 1 : 0];
    tc::GemmParams p = {};
    p.epi = tc::EPI_STORE;
    p.out_hi = partial; p.ldo = y.npad; p.split_stride = (long long)kp * y.npad;
    int r = tc::launch_gemm(p, In, B, K + 1, y.npad, (int)rows, y.npad, y.passes, y.splits[l], s);
    if (r) return r;
    dim3 grid((K + 1 + 7) / 8, (y.H + 31) / 32);
    mlp::reduce_wgrad_kernel<<<grid, 1024, 0, s>>>(partial, y.splits[l], kp, y.npad, y.H, K, grads + y.w_off[l],
                                                  grads + y.b_off[l], accumulate);
    return cuda_status(cudaGetLastError(), "fdsm_mlp_backward(reduce)");
  };
  for (int l = L1; l >= 1; --l) {
    if (forked) {   // side stream: wait until dZ_l is complete, then wgrad_l
      rc = cuda_status(cudaEventRecord(side.fork[l & 1], st), "fdsm_mlp_backward(fork)");
      if (!rc) rc = cuda_status(cudaStreamWaitEvent(sd, side.fork[l & 1], 0), "fdsm_mlp_backward(fork)");
      if (rc) return rc;
    }
    rc = wgrad(l, sd);
    if (rc) return rc;
    // dgrad: dZ_{l-1} = (dZ_l W_l) * act'(H_{l-1})
    tc::Operand A{ws + y.dz_hi[l], lo(ws + y.dz_lo[l]), y.npad, tc::MAJOR_K};
    tc::Operand B = y.raw ? tc::Operand{ws + y.wt_raw[l], nullptr, y.npad, tc::MAJOR_K}   // rows k (K_l of them), contiguous n
                          : tc::Operand{ws + y.wt_hi[l], ws + y.wt_lo[l], y.npad, tc::MAJOR_K};
    tc::GemmParams p = {};
    p.epi = tc::EPI_DACT_SPLIT;
    p.act = y.act;
    p.out_hi = ws + y.dz_hi[l - 1]; p.out_lo = lo(ws + y.dz_lo[l - 1]); p.ldo = y.npad; p.out_cols = y.npad;
    p.aux_hi = ws + y.h_hi[l]; p.aux_lo = lo(ws + y.h_lo[l]); p.ldaux = y.kpadh;
    rc = tc::launch_gemm(p, A, B, (int)rows, y.H, y.npad, y.npad, y.passes, 1, st);
    if (rc) return rc;
  }
  rc = wgrad(0, st);
  if (rc) return rc;
  if (forked) {
    rc = cuda_status(cudaEventRecord(side.join, sd), "fdsm_mlp_backward(join)");
    if (!rc) rc = cuda_status(cudaStreamWaitEvent(st, side.join, 0), "fdsm_mlp_backward(join)");
    if (rc) return rc;
  }
  return FDSM_OK;
}
