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
#include <type_traits>

#include "tc_gemm_host.cuh"
#include "fd_combine.cuh"
#include <cooperative_groups.h>
#include "fd_chain.cuh"

namespace fdsm {
namespace mlp {

inline long long round_up(long long x, long long m) { return (x + m - 1) / m * m; }
constexpr int HB_WARPS = 8;   // head_bwd_kernel: warps (= rows in flight) per block
constexpr int HB_BLOCKS_PER_SM = 2;   // resident blocks per SM the kernel is compiled for (<= 128 registers)
constexpr int HB_ROWS = 3;            // rows a warp has in flight per iteration
// number of per-block partial-sum records of the head backward: H <= 256: one warp per row, HB_BLOCKS_PER_SM blocks per SM;
// H > 256 (head_bwd_wide_kernel): row groups, so that column tiles x row groups ~ 2 blocks per SM
inline int head_blocks(long long rows, int H) {
  if (H <= 256) {
    long long b = (rows + HB_WARPS - 1) / HB_WARPS, cap = (long long)HB_BLOCKS_PER_SM * sm_count();
    return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
  }
  const long long ct = (H + 255) / 256;
  long long g = 2ll * sm_count() / ct;
  if (g < 1) g = 1;
  if (g > rows) g = rows < 1 ? 1 : rows;
  return (int)g;
}

struct Layout {
  int D, H, L, act, head, passes;
  bool raw;            // presplit == 0 and passes == 3: plain fp32 activations, split inside the GEMMs
  long long rows;
  int npad;            // padded hidden width: ceil16(H)
  int bn, nt;          // UMMA N per CTA (min(npad, 256)) and number of N tiles (> 1 when H > 256)
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
  long long head_part;              // nt > 1: per-tile partial head dots [nt][rows][3]
  long long head_acc;               // per-block fp64 partial sums of the head-parameter gradients
  long long partial[2];             // split-K partials: [0] input layer (main stream), [1] hidden layers (side stream)
  long long total;
  int splits[16];
};

inline int make_layout(const fdsm_mlp_desc* d, long long rows, Layout& y) {
  FDSM_REQUIRE(d != nullptr, "fdsm_mlp: desc is NULL");
  FDSM_REQUIRE(d->in_dim >= 1 && d->hidden >= 1 && d->n_hidden >= 1 && d->n_hidden <= 15, "fdsm_mlp: bad sizes");
  FDSM_REQUIRE(d->act >= 1 && d->act <= 3, "fdsm_mlp: act must be 1 (softplus), 2 (nsoftplus) or 3 (elu)");
  FDSM_REQUIRE(d->head == 0 || d->head == 1, "fdsm_mlp: head must be 0 (linear) or 1 (quadratic)");
  FDSM_REQUIRE(d->passes == 1 || d->passes == 3, "fdsm_mlp: passes must be 1 or 3");
  FDSM_REQUIRE(rows >= 0 && rows < (1ll << 31), "fdsm_mlp: bad row count");
  y.D = d->in_dim; y.H = d->hidden; y.L = d->n_hidden; y.act = d->act; y.head = d->head; y.passes = d->passes;
  y.raw = d->presplit == 0 && d->passes == 3;
  FDSM_REQUIRE(d->presplit == 0 || d->presplit == 1, "fdsm_mlp: presplit must be 0 or 1");
  y.rows = rows;
  y.npad = (int)round_up(y.H, 16);
  y.bn = y.npad < 256 ? y.npad : 256;
  y.nt = (y.npad + y.bn - 1) / y.bn;
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
    // raw-operand pipeline: only the fp32 transposed weight; pre-split pipeline: its TF32 hi/lo halves
    y.wt_hi[l] = y.raw ? 0 : take((long long)kp * y.npad);
    y.wt_lo[l] = y.raw ? 0 : take((long long)kp * y.npad);
    y.wt_raw[l] = y.raw ? take((long long)kp * y.npad) : 0;
    // wgrad: M = K_l + 1 rows of dWt in tiles of 128, contraction over `rows` in k-blocks of 32
    int mt = (int)((l ? y.H : y.D) + 1 + tc::BM - 1) / tc::BM;
    if (y.raw && tc::pair_mode() && mt >= 2) mt = (mt + 1) & ~1;   // CTA pairs: the grid's x extent is padded to even
    const long long kblocks = (rows + tc::BK - 1) / tc::BK;
    // ~one CTA per SM.  Hidden layers (l >= 1) run on the side stream next to a dgrad launch of ceil(rows/128)*nt CTAs
    // (one CTA per SM: ~200 KB of shared memory each): sized to the SMs that launch leaves free, so that neither grid
    // queues behind the other (FDSM_WGRAD_SIDE_CTAS / FDSM_WGRAD_MAIN_CTAS override the CTA budgets)
    long long budget = sm_count();
    if (y.L > 1) {
      static const int forced_side = [] { const char* e = getenv("FDSM_WGRAD_SIDE_CTAS"); return e ? atoi(e) : 0; }();
      static const int forced_main = [] { const char* e = getenv("FDSM_WGRAD_MAIN_CTAS"); return e ? atoi(e) : 0; }();
      const long long dgrad_ctas = ((rows + tc::BM - 1) / tc::BM) * y.nt, free_sms = sm_count() - dgrad_ctas;
      const bool share = free_sms >= sm_count() / 4 && dgrad_ctas >= sm_count() / 4;
      const long long unit_h = (long long)((y.H + 1 + tc::BM - 1) / tc::BM) * y.nt;   // CTAs per split of a hidden layer
      long long side = forced_side > 0 ? forced_side : (share ? 3 * free_sms / 4 : sm_count());
      if (side < unit_h) side = unit_h;
      // layer 0 (main stream, after the last dgrad) runs next to the side stream's wgrad_1 and is the longer of the two
      // (K_0 >> H): it gets every SM the side grid does not take (measured optimum at config 2: 36-44 side / 104-112 main)
      const long long side_ctas = side / unit_h * unit_h;
      if (l >= 1) budget = side;
      else budget = forced_main > 0 ? forced_main : (share && side_ctas < sm_count() ? sm_count() - side_ctas : budget);
    }
    long long s = budget / ((long long)mt * y.nt);
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
  y.head_part = y.nt > 1 ? take(3ll * y.nt * rows) : 0;
  y.head_acc = take(2ll * head_blocks(rows, y.H) * (y.head ? 3 : 1) * (y.H + 1));   // per-block fp64 partials
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
__device__ __forceinline__ void prep_weights_body(const PrepArgs& a, const int bx, const int by, const int l) {
  // 32x32 tile transpose through shared memory: coalesced reads of W[n][k] (k fastest) and writes of Wt[k][n]
  __shared__ float tile[32][33];
  const int K = a.K[l], kpad = a.kpad[l], H = a.H, npad = a.npad;
  const int k0 = bx * 32, n0 = by * 32;
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
      if (a.raw[l]) {
        a.raw[l][(long long)k * npad + n] = tile[tx][i];
      } else {
        float hi, lo;
        ptx::tf32_split(tile[tx][i], hi, lo);
        a.hi[l][(long long)k * npad + n] = hi;
        a.lo[l][(long long)k * npad + n] = lo;
      }
    }
  }
}
__global__ void __launch_bounds__(256) prep_weights_kernel(const PrepArgs a) { prep_weights_body(a, blockIdx.x, blockIdx.y, blockIdx.z); }

// ---- head backward: dE -> dZ_L and the head parameter gradients --------------------------------------------
// dH[r,j] = g (l2 w1[j] + l1 w2[j] + 2 wq[j] h)   (quadratic)   |   g w[j]   (linear);   dZ = dH * act'(h)
// g = dL/dE_r comes either from an explicit dE vector (generic autograd path) or -- FD-SSM energy form -- straight from
// the stacked energies (the combine backward folded in: no dE round trip, no extra launch).
// The head-parameter sums cancel heavily across the x / x+v / x-v rows of a sample (dE = 2c, c(q-1), c(-q-1)), so they
// are accumulated in fp64: registers -> fixed-order block reduce -> per-block partials -> fixed-order sum in
// head_finish_kernel (bit-reproducible, no atomics).
// One warp per row, lanes across the columns (coalesced 128-byte row segments), grid ~ 2 blocks per SM.
struct HeadGrad {
  const float* dE;                                    // explicit dL/dE, or NULL:
  const float* E; long long B; float c1, coef;        // energy-form combine, rows = 3B (dot == NULL)
  const float* dot; const float* sigma; float alpha, beta;   // DSM combine, rows = 2B (dot != NULL; coef = 2 alpha gamma / B)
};
__device__ __forceinline__ float head_g(const HeadGrad& s, long long r) {
  if (s.dE) return __ldg(s.dE + r);
  if (s.dot) {
    const bool second = r >= s.B;
    const float g = combine_dsm_grad(s.E, s.dot, s.sigma, s.B, s.alpha, s.beta, s.coef, second ? r - s.B : r);
    return second ? g : -g;
  }
  return combine_energy_grad_row(s.E, s.B, s.c1, s.coef, r);
}
template <int NI, bool HEAD>
__global__ void __launch_bounds__(HB_WARPS * 32, HB_BLOCKS_PER_SM) head_bwd_kernel(const float* __restrict__ hlast, int ldh, const HeadGrad src,
                                                                 const float* __restrict__ l1v, const float* __restrict__ l2v,
                                                                 const float* __restrict__ head_w, int H, int npad,
                                                                 long long rows, int act, float* __restrict__ dz_hi,
                                                                 float* __restrict__ dz_lo, double* __restrict__ partial) {
  constexpr int head = HEAD ? 1 : 0;
  __shared__ double red[HB_WARPS][NI * 32];
  __shared__ double redb[HB_WARPS][3];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long W = (long long)gridDim.x * HB_WARPS;
  // columns beyond H carry zero weights and zero activations: every sum below gets exact zeros from them and dZ = 0 is
  // what the padded columns must hold, so the inner loops need no column branches (bit i of the masks = column
  // lane + 32 i is a real / a stored one)
  // head weights live in shared memory (zero beyond H), the fp64 sums in registers: 3 resident blocks per SM
  __shared__ float sw[3][NI * 32];
  double a1[NI], a2[NI], aq[NI];
  unsigned hmask = 0, nmask = 0;
  for (int j = threadIdx.x; j < NI * 32; j += HB_WARPS * 32) {
    sw[0][j] = j < H ? __ldg(head_w + j) : 0.f;
    sw[1][j] = HEAD && j < H ? __ldg(head_w + H + 1 + j) : 0.f;
    sw[2][j] = HEAD && j < H ? __ldg(head_w + 2 * (H + 1) + j) : 0.f;
  }
#pragma unroll
  for (int i = 0; i < NI; ++i) {
    const int j = lane + 32 * i;
    a1[i] = a2[i] = aq[i] = 0.;
    if (j < H) hmask |= 1u << i;
    if (j < npad) nmask |= 1u << i;
  }
  __syncthreads();
  double sb1 = 0., sb2 = 0., sbq = 0.;
  pdl_launch_dependents();   // the dgrad GEMM sets up behind this grid
  pdl_wait();                // launched programmatically behind the last forward GEMM: head weights above, E / H_L below
  // HB_ROWS rows per iteration, every load of all of them issued before the arithmetic: the kernel is latency-bound
  // (each warp walks ~5 rows; one row at a time exposed two dependent L2 round trips per row)
  constexpr int U = HB_ROWS;
  for (long long r = (long long)blockIdx.x * HB_WARPS + warp; r < rows; r += U * W) {
    long long rr[U];
    bool okr[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { rr[u] = r + u * W; okr[u] = rr[u] < rows; }
    float g[U], ga[U], gb[U], h[U][NI];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      g[u] = okr[u] ? head_g(src, rr[u]) : 0.f;
      ga[u] = g[u]; gb[u] = 0.f;
      if (HEAD && okr[u]) { ga[u] = g[u] * __ldg(l2v + rr[u]); gb[u] = g[u] * __ldg(l1v + rr[u]); }
      const float* hrow = hlast + rr[u] * ldh + lane;
#pragma unroll
      for (int i = 0; i < NI; ++i) h[u][i] = (okr[u] && (hmask >> i & 1)) ? __ldg(hrow + 32 * i) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!okr[u]) continue;
      const double dga = (double)ga[u], dgb = (double)gb[u], dg = (double)g[u];
      const float g2 = 2.f * g[u];
      float* drow = dz_hi + rr[u] * npad + lane;
      float* lrow = dz_lo ? dz_lo + rr[u] * npad + lane : nullptr;
#pragma unroll
      for (int i = 0; i < NI; ++i) {
        const double hd = (double)h[u][i];
        float dh;
        if (HEAD) {
          dh = fmaf(ga[u], sw[0][lane + 32 * i], fmaf(gb[u], sw[1][lane + 32 * i], g2 * sw[2][lane + 32 * i] * h[u][i]));
          a1[i] += dga * hd; a2[i] += dgb * hd; aq[i] += dg * hd * hd;
        } else {
          dh = g[u] * sw[0][lane + 32 * i];
          a1[i] += dg * hd;
        }
        const float dz = dh * tc::act_bwd_from_h(act, h[u][i]);
        if (nmask >> i & 1) {
          if (lrow) {
            float hi, lo;
            ptx::tf32_split(dz, hi, lo);
            drow[32 * i] = hi;
            lrow[32 * i] = lo;
          } else {
            drow[32 * i] = dz;
          }
        }
      }
      sb1 += dga; sb2 += dgb; sbq += dg;
    }
  }
  // fixed-order block reduction of the per-warp sums, one accumulator family at a time
  const int head_n = head ? 3 * (H + 1) : (H + 1);
  double* out = partial + (long long)blockIdx.x * head_n;
  const int nacc = head ? 3 : 1;
  for (int q = 0; q < nacc; ++q) {
#pragma unroll
    for (int i = 0; i < NI; ++i) red[warp][lane + 32 * i] = q == 0 ? a1[i] : (q == 1 ? a2[i] : aq[i]);
    if (lane == 0) redb[warp][q] = q == 0 ? sb1 : (q == 1 ? sb2 : sbq);
    __syncthreads();
    for (int j = threadIdx.x; j <= H; j += HB_WARPS * 32) {
      double s = 0.;
      if (j < H) {
#pragma unroll
        for (int w = 0; w < HB_WARPS; ++w) s += red[w][j];
      } else {
#pragma unroll
        for (int w = 0; w < HB_WARPS; ++w) s += redb[w][q];
      }
      out[q * (H + 1) + j] = s;
    }
    __syncthreads();
  }
}

// H > 256: one thread per column, a block = 256 columns x one row group (blockIdx.y); the three fp64 sums live in
// registers, the per-block record goes to partial[blockIdx.y][...] (the same format head_finish_kernel reduces).
__global__ void __launch_bounds__(256) head_bwd_wide_kernel(const float* __restrict__ hlast, int ldh, const HeadGrad src,
                                                            const float* __restrict__ l1v, const float* __restrict__ l2v,
                                                            const float* __restrict__ head_w, int H, int npad, long long rows,
                                                            int act, int head, float* __restrict__ dz_hi,
                                                            float* __restrict__ dz_lo, double* __restrict__ partial) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  const long long per = (rows + gridDim.y - 1) / gridDim.y;
  const long long r0 = blockIdx.y * per, r1 = r0 + per < rows ? r0 + per : rows;
  const bool col = j < H;
  float w1 = 0.f, w2 = 0.f, wq = 0.f;
  if (col) {
    w1 = __ldg(head_w + j);
    if (head) { w2 = __ldg(head_w + H + 1 + j); wq = __ldg(head_w + 2 * (H + 1) + j); }
  }
  double a1 = 0., a2 = 0., aq = 0., sb1 = 0., sb2 = 0., sbq = 0.;
#pragma unroll 4
  for (long long r = r0; r < r1; ++r) {
    const float g = head_g(src, r);
    float ga = g, gb = 0.f;
    if (head) { ga = g * __ldg(l2v + r); gb = g * __ldg(l1v + r); }
    float dz = 0.f;
    if (col) {
      const float h = __ldg(hlast + r * ldh + j);
      float dh;
      if (head) {
        dh = fmaf(ga, w1, fmaf(gb, w2, 2.f * g * wq * h));
        a1 += (double)ga * h; a2 += (double)gb * h; aq += (double)g * h * h;
      } else {
        dh = g * w1;
        a1 += (double)g * h;
      }
      dz = dh * tc::act_bwd_from_h(act, h);
    }
    if (j < npad) {
      if (dz_lo) {
        float hi, lo;
        ptx::tf32_split(dz, hi, lo);
        dz_hi[r * npad + j] = hi;
        dz_lo[r * npad + j] = lo;
      } else {
        dz_hi[r * npad + j] = dz;
      }
    }
    if (j == 0) { sb1 += (double)ga; sb2 += (double)gb; sbq += (double)g; }
  }
  const int head_n = head ? 3 * (H + 1) : (H + 1);
  double* out = partial + (long long)blockIdx.y * head_n;
  if (col) {
    out[j] = a1;
    if (head) { out[H + 1 + j] = a2; out[2 * (H + 1) + j] = aq; }
  }
  if (j == 0) {
    out[H] = sb1;
    if (head) { out[2 * H + 1] = sb2; out[3 * H + 2] = sbq; }
  }
}

// H > 256 forward: E, l1, l2 from the per-tile partial dots of EPI_HEAD (fixed tile order), biases added here
__global__ void __launch_bounds__(256) head_energy_kernel(const float* __restrict__ part, int nt, long long rows,
                                                          const float* __restrict__ head_w, int H, int head,
                                                          float* __restrict__ e_out, float* __restrict__ l1_out,
                                                          float* __restrict__ l2_out) {
  const long long r = blockIdx.x * 256ll + threadIdx.x;
  if (r >= rows) return;
  float l1 = 0.f, l2 = 0.f, lq = 0.f;
  for (int t = 0; t < nt; ++t) {
    const float* p = part + ((long long)t * rows + r) * 3;
    l1 += p[0]; l2 += p[1]; lq += p[2];
  }
  l1 += __ldg(head_w + H);
  if (head) { l2 += __ldg(head_w + 2 * H + 1); lq += __ldg(head_w + 3 * H + 2); }
  e_out[r] = head ? fmaf(l1, l2, lq) : l1;
  l1_out[r] = l1;
  l2_out[r] = l2;
}

// ghead[j] (+)= sum over the head_bwd blocks of partial[b][j] in a fixed order; when `loss` is set, one extra block
// evaluates the energy-form combine forward (the step's loss) -- same code as fdsm_combine_energy_fwd.
struct HeadLoss {
  const float* E; long long B; float c1, scale; int mean; float* loss;
  const float* dot; const float* sigma; float alpha, beta, gamma;      // dot != NULL: the DSM combine forward instead
};
__global__ void __launch_bounds__(1024) head_finish_kernel(const double* __restrict__ partial, int nblk, int n,
                                                           float* __restrict__ ghead, int accumulate, const HeadLoss hl) {
  __shared__ double part[32][33];
  __shared__ float red[32];
  const int nb = (n + 31) / 32;
  if ((int)blockIdx.x >= nb) {
    if (hl.dot) combine_dsm_fwd_block(hl.E, hl.dot, hl.sigma, hl.B, hl.alpha, hl.beta, hl.gamma, nullptr, hl.loss, red);
    else combine_energy_fwd_block(hl.E, hl.B, 0, hl.c1, hl.scale, hl.mean, nullptr, hl.loss, red);
    return;
  }
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + tx;
  double s = 0.;
  if (j < n)
    for (int b = ty; b < nblk; b += 32) s += partial[(long long)b * n + j];
  part[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && j < n) {
    double t = 0.;
#pragma unroll
    for (int g = 0; g < 32; ++g) t += part[g][tx];
    ghead[j] = accumulate ? ghead[j] + (float)t : (float)t;
  }
}

// ---- split-K reduce: partial[s][k][n] -> gW[n][k] (+)= sum_s, gb[n] (+)= sum_s partial[s][K][n] --------------
// 8 (k) x 32 (n) outputs per 1024-thread block; 4 thread groups each sum a fixed quarter of the splits in fp64, the
// quarters are combined in a fixed order (deterministic), then a transposing store through shared memory.
// GROUPS = 4 (many splits: the hidden layers' 74): four thread groups each sum a fixed quarter of the splits, combined in
// a fixed order.  GROUPS = 1 (<= 24 splits: the input layer's 21): 256-thread blocks, every load of a thread in flight at
// once, the whole grid in one wave -- this launch sits on the step's critical path.
template <int GROUPS>
__global__ void __launch_bounds__(256 * GROUPS) reduce_wgrad_kernel(const float* __restrict__ partial, int splits, int kpad,
                                                                    int npad, int H, int K, float* __restrict__ gW,
                                                                    float* __restrict__ gb, int accumulate) {
  __shared__ double part[GROUPS][8][32];
  __shared__ float tile[8][33];
  pdl_launch_dependents();
  pdl_wait();   // no-ops unless launched programmatically (main-stream instance only: fd_mlp.cu wgrad())
  const int k0 = blockIdx.x * 8, n0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = (threadIdx.x >> 5) & 7, tz = threadIdx.x >> 8;
  const long long stride = (long long)kpad * npad;
  {
    const int k = k0 + ty, n = n0 + tx;
    double acc = 0.;
    if (k <= K && n < H) {
      const int per = (splits + GROUPS - 1) / GROUPS;
      const int s0 = tz * per, s1 = min(splits, s0 + per);
      const float* p = partial + (long long)k * npad + n;
      int s = s0;
      for (; s + 8 <= s1; s += 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = __ldg(p + (s + u) * stride);
#pragma unroll
        for (int u = 0; u < 8; ++u) acc += (double)v[u];
      }
      for (; s < s1; ++s) acc += (double)__ldg(p + s * stride);
    }
    part[tz][ty][tx] = acc;
  }
  __syncthreads();
  if (tz == 0) {
    double acc = part[0][ty][tx];
#pragma unroll
    for (int g = 1; g < GROUPS; ++g) acc += part[g][ty][tx];
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
    a.hi[l] = y.raw ? nullptr : workspace + y.wt_hi[l]; a.lo[l] = y.raw ? nullptr : workspace + y.wt_lo[l];
    a.raw[l] = y.raw ? workspace + y.wt_raw[l] : nullptr;
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
      p.head_part = y.nt > 1 ? ws + y.head_part : nullptr;
    }
    rc = tc::launch_gemm(p, A, B, (int)rows, y.H, K, y.bn, y.passes, 1, st);
    if (rc) return rc;
  }
  if (y.nt > 1) {   // several N tiles: combine the per-tile head dots (fixed order), add the biases, form the energy
    mlp::head_energy_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(ws + y.head_part, y.nt, rows, params + y.head_off,
                                                                             y.H, y.head, energy, ws + y.l1, ws + y.l2);
    return cuda_status(cudaGetLastError(), "fdsm_mlp_forward(head_energy)");
  }
  return FDSM_OK;
}

namespace fdsm {
namespace mlp {
static int backward_impl(const fdsm_mlp_desc* desc, const float* x_hi, const float* x_lo, int64_t rows, const float* params,
                         float* workspace, const HeadGrad& hg, const HeadLoss& hl, float* grads, int32_t accumulate,
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
  FDSM_REQUIRE(x_hi && (x_lo || y.passes == 1 || y.raw) && (hg.dE || hg.E), "fdsm_mlp_backward: NULL pointer");
  float* ws = workspace;
  auto lo = [&](float* q) -> float* { return y.raw ? nullptr : q; };
  const int head_n = (int)(y.n_params - y.head_off);
  double* hacc = reinterpret_cast<double*>(ws + y.head_acc);
  // Two streams: the critical chain head_bwd -> dgrad_{L-1} -> ... -> dgrad_1 -> wgrad_0 stays on `stream`; the hidden
  // layers' weight gradients (wgrad_l + reduce_l, l >= 1) only need dZ_l and run on a side stream, filling the SMs the
  // 96-CTA dgrad launches leave idle and hiding their launch/fill/drain latencies.  Fork/join by events: capturable.
  mlp::SideStream& side = mlp::side_stream();
  const bool forked = y.L > 1 && side.ok();
  cudaStream_t sd = forked ? side.stream : st;
  const int L1 = y.L - 1;
  {
    const int nblk = mlp::head_blocks(rows, y.H), ni = (y.npad + 31) / 32;
    if (y.H > 256) {
      dim3 grid((y.npad + 255) / 256, nblk);
      mlp::head_bwd_wide_kernel<<<grid, 256, 0, st>>>(ws + y.h_last, y.npad, hg, ws + y.l1, ws + y.l2, params + y.head_off, y.H,
                                                      y.npad, rows, y.act, y.head, ws + y.dz_hi[L1], lo(ws + y.dz_lo[L1]), hacc);
    } else {
    auto launch_head = [&](auto tag) {
      constexpr int NI = decltype(tag)::value;
      auto go = [&](auto kernel) {
        launch_pdl(kernel, dim3(nblk), dim3(mlp::HB_WARPS * 32), 0, st, (const float*)(ws + y.h_last), y.npad, hg,
                   (const float*)(ws + y.l1), (const float*)(ws + y.l2), params + y.head_off, y.H, y.npad, (long long)rows, y.act,
                   ws + y.dz_hi[L1], lo(ws + y.dz_lo[L1]), hacc);
      };
      if (y.head) go(mlp::head_bwd_kernel<NI, true>); else go(mlp::head_bwd_kernel<NI, false>);
    };
    switch (ni) {
      case 1: launch_head(std::integral_constant<int, 1>{}); break;
      case 2: launch_head(std::integral_constant<int, 2>{}); break;
      case 3: launch_head(std::integral_constant<int, 3>{}); break;
      case 4: launch_head(std::integral_constant<int, 4>{}); break;
      case 5: launch_head(std::integral_constant<int, 5>{}); break;
      case 6: launch_head(std::integral_constant<int, 6>{}); break;
      case 7: launch_head(std::integral_constant<int, 7>{}); break;
      default: launch_head(std::integral_constant<int, 8>{}); break;
    }
    }
    rc = cuda_status(cudaGetLastError(), "fdsm_mlp_backward(head)");
    if (rc) return rc;
    // the head-gradient sums (and the loss) are off the critical chain: side stream
    cudaStream_t sf = st;
    if (forked) {
      rc = cuda_status(cudaEventRecord(side.fork[(L1 + 1) & 1], st), "fdsm_mlp_backward(fork)");
      if (!rc) rc = cuda_status(cudaStreamWaitEvent(sd, side.fork[(L1 + 1) & 1], 0), "fdsm_mlp_backward(fork)");
      if (rc) return rc;
      sf = sd;
    }
    mlp::head_finish_kernel<<<(head_n + 31) / 32 + (hl.loss ? 1 : 0), 1024, 0, sf>>>(hacc, nblk, head_n, grads + y.head_off,
                                                                                      accumulate, hl);
    rc = cuda_status(cudaGetLastError(), "fdsm_mlp_backward(head_finish)");
    if (rc) return rc;
  }
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
    int r = tc::launch_gemm(p, In, B, K + 1, y.npad, (int)rows, y.bn, y.passes, y.splits[l], s);
    if (r) return r;
    dim3 grid((K + 1 + 7) / 8, (y.H + 31) / 32);
    // the main-stream instance (layer 0, the last kernel of the backward) is launched programmatically: its blocks are
    // resident when the wgrad GEMM drains; side-stream instances are not (idle blocks would take the SM slots the
    // main stream's dgrad CTAs want)
    const bool pdl = s == st && forked;
    cudaError_t e;
    if (y.splits[l] <= 24)
      e = pdl ? launch_pdl(mlp::reduce_wgrad_kernel<1>, grid, dim3(256), 0, s, (const float*)partial, y.splits[l], kp, y.npad, y.H,
                           K, grads + y.w_off[l], grads + y.b_off[l], (int)accumulate)
              : (mlp::reduce_wgrad_kernel<1><<<grid, 256, 0, s>>>(partial, y.splits[l], kp, y.npad, y.H, K, grads + y.w_off[l],
                                                                 grads + y.b_off[l], accumulate), cudaGetLastError());
    else
      e = pdl ? launch_pdl(mlp::reduce_wgrad_kernel<4>, grid, dim3(1024), 0, s, (const float*)partial, y.splits[l], kp, y.npad, y.H,
                           K, grads + y.w_off[l], grads + y.b_off[l], (int)accumulate)
              : (mlp::reduce_wgrad_kernel<4><<<grid, 1024, 0, s>>>(partial, y.splits[l], kp, y.npad, y.H, K, grads + y.w_off[l],
                                                                  grads + y.b_off[l], accumulate), cudaGetLastError());
    return cuda_status(e, "fdsm_mlp_backward(reduce)");
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
    rc = tc::launch_gemm(p, A, B, (int)rows, y.H, y.npad, y.bn, y.passes, 1, st);
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
}  // namespace mlp
}  // namespace fdsm

extern "C" int fdsm_mlp_backward(const fdsm_mlp_desc* desc, const float* x_hi, const float* x_lo, int64_t rows,
                                 const float* params, float* workspace, const float* dE, float* grads, int32_t accumulate,
                                 fdsm_stream_t stream) {
  FDSM_REQUIRE(rows == 0 || dE, "fdsm_mlp_backward: dE is NULL");
  mlp::HeadGrad hg = {dE, nullptr, 0, 0.f, 0.f, nullptr, nullptr, 0.f, 0.f};
  mlp::HeadLoss hl = {nullptr, 0, 0.f, 0.f, 0, nullptr, nullptr, nullptr, 0.f, 0.f, 0.f};
  return mlp::backward_impl(desc, x_hi, x_lo, rows, params, workspace, hg, hl, grads, accumulate, stream);
}

extern "C" int fdsm_mlp_backward_fd_ssm(const fdsm_mlp_desc* desc, const float* x_hi, const float* x_lo, int64_t B,
                                        const float* params, float* workspace, const float* energy, float c1, float scale,
                                        int32_t mean, float* loss, float* grads, int32_t accumulate, fdsm_stream_t stream) {
  FDSM_REQUIRE(B >= 0 && B < (1ll << 29), "fdsm_mlp_backward_fd_ssm: bad batch size");
  FDSM_REQUIRE(B == 0 || energy, "fdsm_mlp_backward_fd_ssm: energy is NULL");
  const float coef = mean && B ? scale / (float)B : scale;
  mlp::HeadGrad hg = {nullptr, energy, B, c1, coef, nullptr, nullptr, 0.f, 0.f};
  mlp::HeadLoss hl = {energy, B, c1, scale, mean, B ? loss : nullptr, nullptr, nullptr, 0.f, 0.f, 0.f};
  if (B == 0 && loss) {
    int rc = cuda_status(cudaMemsetAsync(loss, 0, sizeof(float), (cudaStream_t)stream), "fdsm_mlp_backward_fd_ssm");
    if (rc) return rc;
  }
  return mlp::backward_impl(desc, x_hi, x_lo, 3 * B, params, workspace, hg, hl, grads, accumulate, stream);
}

extern "C" int fdsm_mlp_backward_fd_dsm(const fdsm_mlp_desc* desc, const float* x_hi, const float* x_lo, int64_t B,
                                        const float* params, float* workspace, const float* energy, const float* dot,
                                        const float* sigma_b, float alpha, float beta, float gamma, float* loss, float* grads,
                                        int32_t accumulate, fdsm_stream_t stream) {
  FDSM_REQUIRE(B >= 0 && B < (1ll << 29), "fdsm_mlp_backward_fd_dsm: bad batch size");
  FDSM_REQUIRE(B == 0 || (energy && dot), "fdsm_mlp_backward_fd_dsm: energy / dot is NULL");
  const float c = B ? 2.f * alpha * gamma / (float)B : 0.f;      // as fdsm_combine_dsm_bwd with gout = 1
  mlp::HeadGrad hg = {nullptr, energy, B, 0.f, c, dot, sigma_b, alpha, beta};
  mlp::HeadLoss hl = {energy, B, 0.f, 0.f, 1, B ? loss : nullptr, dot, sigma_b, alpha, beta, gamma};
  if (B == 0 && loss) {
    int rc = cuda_status(cudaMemsetAsync(loss, 0, sizeof(float), (cudaStream_t)stream), "fdsm_mlp_backward_fd_dsm");
    if (rc) return rc;
  }
  return mlp::backward_impl(desc, x_hi, x_lo, 2 * B, params, workspace, hg, hl, grads, accumulate, stream);
}

#include "fd_jet_host.inc"
