// The chain kernel: forward of every layer, the head + FD combine, and the data-gradient GEMMs of the energy MLP in ONE
// launch, in the difference-propagating form of fd_jet.cuh.  A CTA owns S samples (S = 8..40): the 3 S rows
// [h0 | a | b] of its samples are one 128-row tcgen05 tile, so everything between the input and dZ_0 is local to the CTA
// -- the only cross-CTA step of the backward, the weight-gradient contraction over the batch, runs afterwards as split-K
// tcgen05 GEMMs over the activations / dZ this kernel leaves in HBM.
//
//   phase g = 0..L-1   forward layer g:  acc[128 x npad] = A_g W_g^T  (TMA-fed for g = 0, resident in shared memory after)
//                      epilogue: jets -> next layer's A operand written IN PLACE in the UMMA (128-byte swizzle, K-major)
//                      layout; last layer: head, per-sample loss, head gradient, activation backward -> dZ_{L-1}
//   phase g = L..2L-2  dgrad of layer l = 2L-1-g: acc = dZ_l W_l;  epilogue: jet backward with H_l (read back from L2) ->
//                      dZ_{l-1} in place
//
// The epilogue works one 32-column k-block at a time and publishes each finished block on an mbarrier (a_ready[j]): the
// MMA issuer starts the NEXT phase's k-block j -- and sends the block to HBM through a TMA store (H_l, dZ_l: operands of
// the weight-gradient GEMMs) -- while the epilogue warps are still on block j+1, so after phase 0 the tensor pipe and the
// weight stream run under the epilogue instead of between epilogues.
//
// What bounds the weight stream is the L2 -> SM delivery rate (every CTA walks all weights: ~3 MB per CTA as hi + lo,
// x 128 CTAs against the chip's ~6300 B/clk L2 throughput), so the weights travel ONCE, as plain fp32: tcgen05 kind::tf32
// ignores the 13 low mantissa bits, i.e. the raw tile IS the hi operand; once the hi MMAs of a slot have completed, the
// epilogue warps -- idle during phase 0 -- overwrite the slot IN PLACE with lo = x - trunc(x) and the lo MMAs follow, one
// k-block behind the hi MMAs.  Later phases stream under the epilogues, where the delivery rate is not the limit and the
// epilogue warps are busy: there the weights come as two TMA copies, hi and lo.  That is PASSES = 3 (every GEMM as hi + lo).
// PASSES = 2, the default, keeps hi + lo for the hidden layers and runs the input layer on round-to-nearest TF32 weights
// alone: the rounding of W_0 moves loss / gradient by 2e-5 / 9e-5 (config 2, against fp64), the rounding of the hidden
// weights by 1e-4 / 2e-4 (tests/tools/jet_accuracy_study.py), and phase 0 -- nothing to overlap it with -- is the one
// place where the second pass costs its full time.  PASSES = 1 issues hi MMAs only, everywhere.  With CL > 1 the CTAs of a cluster split each slot's TMA
// boxes and multicast them.
//
// Shared memory: A region NA x (3 S x 128 B) (NA = ceil32(H+1)/32 k-blocks; the TMA ring of phase 0, then the resident
// operand; the 128-row MMA reads past the 3 S rows of a block into the next one: those accumulator rows are never read)
// + W ring (nraw x 28 KB, one k-block each) + head weights / head-gradient records.
// TMEM: one 128 x npad fp32 accumulator.  Warps 0..15 run the epilogues (warp w owns samples w, w+16, ...; lane = column
// mod 32; two samples at a time for instruction-level parallelism); warp 16 lane 0 = TMA producer, warp 17 lane 0 = MMA +
// TMA-store issuer; warp 2 also allocates TMEM.
#pragma once
#include <cuda_fp16.h>
#include "fd_jet.cuh"

namespace fdsm {
namespace jet {

constexpr int MAXL = 5;              // hidden layers the chain supports (CUtensorMap budget of the parameter space)
constexpr int EPI_WARPS = 16, EPI_T = EPI_WARPS * 32;
constexpr int PROD_WARP = 16, MMA_WARP = 17;
constexpr int CT = EPI_T + 64;
constexpr int MAX_RAW = 5, WSLOT_BYTES = 28672;   // slot = 7 boxes of 32 x 32 fp32 (4 KB each)
constexpr int MAX_NA = 7;            // ceil32(H+1) <= 224
constexpr int MAX_S = 40, MAX_SPW = (MAX_S + EPI_WARPS - 1) / EPI_WARPS;   // samples per tile / per epilogue warp
constexpr int HWS = 224;             // head-weight row stride in shared memory
constexpr int HREC = 3 * HWS + 4;    // one head-gradient record: [gw1 | gw2 | gwq | gb1 gb2 gbq .]
constexpr int NCOEF = 12;            // per-sample head-gradient coefficients kept in shared memory between the two passes
constexpr int NBARS = 2 * MAX_NA + 4 * MAX_RAW + 1 + MAX_NA;
constexpr int CHAIN_MISC_BYTES = (3 * HWS + 4 * HREC + MAX_S * NCOEF) * 4 + NBARS * 8 + 16 + EPI_WARPS * 4 + 64;

struct alignas(64) ChainParams {
  CUtensorMap t_in;                  // XV [2][Bp][kpad0] as (32 k, row, stream, k-block): box {32, S, 3, AG} -- the third stream is
                                     //   out of bounds: the b rows of the input layer arrive as zeros        (load)
  CUtensorMap t_wf[MAXL];            // Wt_l = [hi rows | lo rows] [2 kpad_l x npad], MN-major, as (32 n, k, n-panel[, hi|lo]):
                                     //   l = 0, PASSES != 3: 3-D, box {32, 64, nbx} ; otherwise 4-D, box {32, 32, nbx, HL}
  CUtensorMap t_wb[MAXL];            // the same memory, K-major view for the dgrad: 3-D (n, k, hi|lo), box {32, 32 nbx, HL}
  CUtensorMap t_h[MAXL];             // H_l [3][Bp][kpadh], l = 1..L-1, as (column, row, stream)      (store, box {32, S, 3})
  CUtensorMap t_dz[MAXL];            // dZ_l [3][Bp][npad] (l = 0: 2 streams)                         (store, box {32, S, 3 | 2})
  const float* h[MAXL];              // H_l again, for the backward's plain loads
  const float* head_w;               // [w1(H) b1 | w2(H) b2 | wq(H) bq]  or  [w(H) b]
  float* loss_part;                  // [tiles]   sum over the tile's samples of the per-sample loss (unscaled)
  float* head_part;                  // [tiles][head_n]  head-parameter gradient of the tile
  int B, Bp, S, D, H, L, npad, kpad0, kpadh, head;
  int ablk, nraw;                    // bytes of one A k-block (3 S x 128), slots of the weight ring
  int ag;                            // k-blocks of phase 0's A operand per TMA instruction / barrier
  // input layer in kind::f16 (fdsm_mlp_desc.passes = 4): fp16 carries TF32's 11 significant bits in half the bytes, and phase 0 is
  // the L2 -> shared-memory stream of W_0 and [x | v].  Both operands K-major, k-blocks of 64:
  CUtensorMap t_in16;                // XV16 [2][Bp][kpad16] halfs as (64 k, row, stream, k-block): box {64, S, 3, AG}
  CUtensorMap t_w16;                 // W16 [npad][kpad16] halfs (W_0 | b_0 | 0) as (64 k, n, k-block): box {64, npad, 2} = one unit
  int in16, nkb16;
  float act_offset;                  // ln 2 for the normalised softplus
  float c1, coef;                    // loss_b = dE^2/8 - sE + (c1 - 2) E0 ; dL/dloss_b = coef (= scale / B or scale)
  unsigned long long* trace;         // debug: per-CTA %globaltimer stamps [tile][TRACE_N] (NULL in production)
  int trace_mode;                    // debug: what the free slots 22.. of a 3-layer trace hold (0: fwd1 k-block stamps of the MMA thread,
                                     //   1: per-warp end of fwd1's epilogue, 2: per-warp end of block 1 of fwd1's epilogue)
};
constexpr int TRACE_N = 2 + 4 * (2 * MAXL - 1);
__device__ __forceinline__ void stamp_now(const ChainParams& p, int slot) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  p.trace[(long long)blockIdx.x * TRACE_N + slot] = t;
}
__device__ __forceinline__ void stamp(const ChainParams& p, int slot) {   // epilogue-side stamps: one thread of warp 3
  if (p.trace && threadIdx.x == 96) stamp_now(p, slot);
}

// the last A block's 128-row MMA reads up to 16 KB from its base: the weight rings behind it cover the overrun
__host__ __device__ constexpr int chain_smem_bytes(int NA, int ablk, int nslots) {
  return 1024 + NA * ablk + nslots * WSLOT_BYTES + CHAIN_MISC_BYTES;
}

__device__ __forceinline__ void tma_store_3d(const void* tmap, uint32_t src, int32_t c0, int32_t c1, int32_t c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"((uint64_t)tmap), "r"(src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ float lds(const uint8_t* p) { return *reinterpret_cast<const float*>(p); }
__device__ __forceinline__ void sts(uint8_t* p, float v) { *reinterpret_cast<float*>(p) = v; }

// ---- epilogue bodies: TWO samples of one column at once (the epilogues are bound by instruction issue and by the latency of
// ---- one element's dependency chain; e[n] = address of sample n's h0-stream element in the block, the a / b streams sit
// ---- `sst` and 2 `sst` bytes further; v1 = the second sample exists -- otherwise e[1] = e[0] and nothing is stored for it).
// ---- Columns >= H need no masking: their pre-activations are exact zeros (zero-padded weight columns, zero-filled by the
// ---- dump beyond npad), everything computed from them is finite, and every consumer multiplies it by a zero weight row /
// ---- column or ignores it -- except column H of a hidden layer's output, the ones column that carries the next bias.
template <int ACT, int MODE>
__device__ __forceinline__ void epi_fwd_hidden(uint8_t* const (&e)[2], bool v1, int sst, bool ones, float off) {
  float z[2], P[2], Q[2];
  Jet o[2];
#pragma unroll
  for (int n = 0; n < 2; ++n) { z[n] = lds(e[n]); P[n] = lds(e[n] + sst); Q[n] = lds(e[n] + 2 * sst); }
  jet_fwd_n<ACT, MODE, 2>(z, P, Q, off, o);
#pragma unroll
  for (int n = 0; n < 2; ++n) {
    if (n == 0 || v1) {
      sts(e[n], ones ? 1.f : tf32r(o[n].h0));
      sts(e[n] + sst, tf32r(o[n].a));
      sts(e[n] + 2 * sst, tf32r(o[n].b));
    }
  }
}

// last forward layer, pass 1: jets (kept unrounded in place) and the head's nine sums per sample
template <int ACT, int MODE>
__device__ __forceinline__ void epi_head_pass1(uint8_t* const (&e0)[2], bool v1, int sst, int ablk, int NA, int lane, const float* hw,
                                               bool quad, float off, float (&sum)[2][9]) {
#pragma unroll
  for (int n = 0; n < 2; ++n)
#pragma unroll
    for (int q = 0; q < 9; ++q) sum[n][q] = 0.f;
#pragma unroll 1
  for (int j = 0; j < NA; ++j) {
    const int c = lane + 32 * j;
    const float w1 = hw[c], w2 = hw[HWS + c], wq = hw[2 * HWS + c];      // zero beyond H
    float z[2], P[2], Q[2];
    Jet o[2];
#pragma unroll
    for (int n = 0; n < 2; ++n) {
      const uint8_t* e = e0[n] + j * ablk;
      z[n] = lds(e); P[n] = lds(e + sst); Q[n] = lds(e + 2 * sst);
    }
    jet_fwd_n<ACT, MODE, 2>(z, P, Q, off, o);
#pragma unroll
    for (int n = 0; n < 2; ++n) {
      uint8_t* e = e0[n] + j * ablk;
      if (n == 0 || v1) { sts(e, o[n].h0); sts(e + sst, o[n].a); sts(e + 2 * sst, o[n].b); }   // for this lane, in pass 2
      sum[n][0] = fmaf(w1, o[n].h0, sum[n][0]); sum[n][1] = fmaf(w1, o[n].a, sum[n][1]); sum[n][2] = fmaf(w1, o[n].b, sum[n][2]);
      if (quad) {
        sum[n][3] = fmaf(w2, o[n].h0, sum[n][3]); sum[n][4] = fmaf(w2, o[n].a, sum[n][4]); sum[n][5] = fmaf(w2, o[n].b, sum[n][5]);
        if constexpr (MODE == JMODE_TAYLOR) {
          sum[n][6] = fmaf(wq, o[n].a * o[n].h0, sum[n][6]);
          sum[n][7] = fmaf(wq, fmaf(o[n].a, o[n].a, 2.f * o[n].h0 * o[n].b), sum[n][7]);
        } else {
          sum[n][6] = fmaf(wq, o[n].a * (o[n].b + o[n].h0), sum[n][6]);
          sum[n][7] = fmaf(wq, fmaf(o[n].b, o[n].b, fmaf(o[n].a, o[n].a, 2.f * o[n].h0 * o[n].b)), sum[n][7]);
        }
        sum[n][8] = fmaf(wq, o[n].h0 * o[n].h0, sum[n][8]);
      }
    }
  }
#pragma unroll
  for (int n = 0; n < 2; ++n)
#pragma unroll
    for (int q = 0; q < 9; ++q)
      if (q < 3 || quad) sum[n][q] = warp_sum(sum[n][q]);
}

// from the nine sums of a sample: its loss term and the coefficients of the head backward (oracle/jet.py quadratic_head_jet)
//   cs[0..3] = G1h G1a G1b gd   cs[4..7] = G2h G2a G2b gs   cs[8] = ge
template <int MODE>
__device__ __forceinline__ float head_sample(const float (&sum)[9], bool quad, float b1, float b2, float bq, float c1, float cf,
                                             bool lane0, float* cs, float& gb1, float& gb2, float& gbq) {
  const float l1 = sum[0] + b1, A1 = sum[1], B1 = sum[2], l2 = sum[3] + b2, A2 = sum[4], B2 = sum[5];
  if constexpr (MODE == JMODE_TAYLOR) {
    // exact SSM: d1 = E', d2h = E''/2;  loss_b = d1^2 / 2 - 2 d2h;  gd = dL/dd1, gs = dL/dd2h, no E0 term
    float d1, d2h;
    if (quad) {
      d1 = A1 * l2 + l1 * A2 + 2.f * sum[6];
      d2h = B1 * l2 + A1 * A2 + l1 * B2 + sum[7];
    } else {
      d1 = A1; d2h = B1;
    }
    const float gd = cf * d1, gs = -2.f * cf;
    float G1h, G2h, G1a, G2a, G1b, G2b;
    if (quad) {
      G1h = gd * A2 + gs * B2; G2h = gd * A1 + gs * B1;
      G1a = gd * l2 + gs * A2; G2a = gd * l1 + gs * A1;
      G1b = gs * l2;           G2b = gs * l1;
    } else {
      G1h = 0.f; G1a = gd; G1b = gs; G2h = G2a = G2b = 0.f;
    }
    gb1 += G1h; gb2 += G2h;
    if (lane0) {
      float4* c4 = reinterpret_cast<float4*>(cs);
      c4[0] = make_float4(G1h, G1a, G1b, gd);
      c4[1] = make_float4(G2h, G2a, G2b, gs);
      c4[2] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    return 0.5f * d1 * d1 - 2.f * d2h;
  }
  float dE, sE, E0;
  if (quad) {
    dE = 2.f * (A1 * (l2 + B2) + A2 * (l1 + B1)) + 4.f * sum[6];
    sE = 2.f * (l1 * B2 + B1 * l2 + B1 * B2 + A1 * A2) + 2.f * sum[7];
    E0 = fmaf(l1, l2, sum[8]) + bq;
  } else {
    dE = 2.f * A1; sE = 2.f * B1; E0 = l1;
  }
  const float gd = cf * dE * 0.25f, gs = -cf, ge = cf * (c1 - 2.f);
  float G1h, G2h, G1a, G2a, G1b, G2b;
  if (quad) {
    G1h = 2.f * (gd * A2 + gs * B2) + ge * l2; G2h = 2.f * (gd * A1 + gs * B1) + ge * l1;
    G1a = 2.f * (gd * (l2 + B2) + gs * A2);     G2a = 2.f * (gd * (l1 + B1) + gs * A1);
    G1b = 2.f * (gd * A2 + gs * (l2 + B2));     G2b = 2.f * (gd * A1 + gs * (l1 + B1));
  } else {
    G1h = ge; G1a = 2.f * gd; G1b = 2.f * gs; G2h = G2a = G2b = 0.f;
  }
  gb1 += G1h; gb2 += G2h; gbq += ge;
  if (lane0) {
    float4* c4 = reinterpret_cast<float4*>(cs);
    c4[0] = make_float4(G1h, G1a, G1b, gd);
    c4[1] = make_float4(G2h, G2a, G2b, gs);
    c4[2] = make_float4(ge, 0.f, 0.f, 0.f);
  }
  return dE * dE * 0.125f - sE + (c1 - 2.f) * E0;
}

// last forward layer, pass 2, one block: gradient w.r.t. the jets, head-weight gradient terms, activation backward -> dZ
template <int ACT, int MODE>
__device__ __forceinline__ void epi_head_pass2(uint8_t* const (&e)[2], bool v1, int sst, float w1, float w2, float wq,
                                               const float* const (&cs)[2], bool quad, float off, float& gw1, float& gw2, float& gwq) {
  Jet o[2];
  float g0[2], ga[2], gb[2], dz0[2], dP[2], dQ[2];
#pragma unroll
  for (int n = 0; n < 2; ++n) {
    const float4 k1 = reinterpret_cast<const float4*>(cs[n])[0], k2 = reinterpret_cast<const float4*>(cs[n])[1];
    const float ge = cs[n][8];
    const float G1h = k1.x, G1a = k1.y, G1b = k1.z, gd = k1.w, G2h = k2.x, G2a = k2.y, G2b = k2.z, gs = k2.w;
    o[n] = Jet{lds(e[n]), lds(e[n] + sst), lds(e[n] + 2 * sst)};
    g0[n] = G1h * w1; ga[n] = G1a * w1; gb[n] = G1b * w1;
    const float use = (n == 0 || v1) ? 1.f : 0.f;      // the stand-in second sample adds nothing to the weight gradients
    gw1 += use * (G1h * o[n].h0 + G1a * o[n].a + G1b * o[n].b);
    if (quad) {
      gw2 += use * (G2h * o[n].h0 + G2a * o[n].a + G2b * o[n].b);
      if constexpr (MODE == JMODE_TAYLOR) {
        g0[n] += G2h * w2 + wq * 2.f * (gd * o[n].a + gs * o[n].b);
        ga[n] += G2a * w2 + wq * 2.f * (gd * o[n].h0 + gs * o[n].a);
        gb[n] += G2b * w2 + wq * 2.f * gs * o[n].h0;
        gwq += use * (2.f * gd * o[n].h0 * o[n].a + gs * fmaf(o[n].a, o[n].a, 2.f * o[n].h0 * o[n].b));
      } else {
        g0[n] += G2h * w2 + wq * (4.f * gd * o[n].a + 4.f * gs * o[n].b + 2.f * ge * o[n].h0);
        ga[n] += G2a * w2 + wq * 4.f * (gd * (o[n].b + o[n].h0) + gs * o[n].a);
        gb[n] += G2b * w2 + wq * 4.f * (gd * o[n].a + gs * (o[n].b + o[n].h0));
        gwq += use * (4.f * gd * o[n].a * (o[n].b + o[n].h0) +
                      2.f * gs * fmaf(o[n].b, o[n].b, fmaf(o[n].a, o[n].a, 2.f * o[n].h0 * o[n].b)) + ge * o[n].h0 * o[n].h0);
      }
    }
  }
  jet_bwd_n<ACT, MODE, 2>(o, off, g0, ga, gb, dz0, dP, dQ);
#pragma unroll
  for (int n = 0; n < 2; ++n)
    if (n == 0 || v1) { sts(e[n], tf32r(dz0[n])); sts(e[n] + sst, tf32r(dP[n])); sts(e[n] + 2 * sst, tf32r(dQ[n])); }
}

// dgrad epilogue, one block: rows hold dL/dH_l as (g0, ga, gb); H_l (hp[n]: the sample's h0 element in HBM / L2; the a / b
// streams `hstride` floats further) -> dZ_{l-1}
template <int ACT, int MODE>
__device__ __forceinline__ void epi_dgrad(uint8_t* const (&e)[2], bool v1, int sst, const float* const (&hp)[2], long long hstride,
                                          float off) {
  Jet o[2];
  float g0[2], ga[2], gb[2], dz0[2], dP[2], dQ[2];
#pragma unroll
  for (int n = 0; n < 2; ++n) o[n] = Jet{__ldcg(hp[n]), __ldcg(hp[n] + hstride), __ldcg(hp[n] + 2 * hstride)};
#pragma unroll
  for (int n = 0; n < 2; ++n) { g0[n] = lds(e[n]); ga[n] = lds(e[n] + sst); gb[n] = lds(e[n] + 2 * sst); }
  jet_bwd_n<ACT, MODE, 2>(o, off, g0, ga, gb, dz0, dP, dQ);
#pragma unroll
  for (int n = 0; n < 2; ++n)
    if (n == 0 || v1) { sts(e[n], tf32r(dz0[n])); sts(e[n] + sst, tf32r(dP[n])); sts(e[n] + 2 * sst, tf32r(dQ[n])); }
}

template <int ACT, int PASSES, int CL, int MODE>
__global__ void __launch_bounds__(CT, 1) chain_kernel(const __grid_constant__ ChainParams p) {
  using namespace ptx;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - raw);
  const int NA = p.kpadh >> 5, ablk = p.ablk, nraw = p.nraw;
  const uint32_t a_base = base;
  const uint32_t w_base = base + NA * ablk;
  uint8_t* a_gen = gen;
  uint8_t* w_gen = gen + NA * ablk;
  float* hw = reinterpret_cast<float*>(w_gen + nraw * WSLOT_BYTES);    // [3][HWS]    head weights
  float* hrec = hw + 3 * HWS;                                                  // [4][HREC]   head-gradient records
  float* coef_s = hrec + 4 * HREC;                                             // [MAX_S][NCOEF]
  uint8_t* bars_gen = reinterpret_cast<uint8_t*>(coef_s + MAX_S * NCOEF);
  const uint32_t bars = w_base + nraw * WSLOT_BYTES + (3 * HWS + 4 * HREC + MAX_S * NCOEF) * 4;
  auto a_full = [&](int s) { return bars + 8u * s; };
  auto a_empty = [&](int s) { return bars + 8u * (MAX_NA + s); };
  auto w_full = [&](int s) { return bars + 8u * (2 * MAX_NA + s); };
  auto w_empty = [&](int s) { return bars + 8u * (2 * MAX_NA + MAX_RAW + s); };
  const uint32_t acc_full = bars + 8u * (2 * MAX_NA + 4 * MAX_RAW);
  auto a_ready = [&](int j) { return bars + 8u * (2 * MAX_NA + 4 * MAX_RAW + 1 + j); };
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(bars_gen + 8 * NBARS);
  float* red = reinterpret_cast<float*>(bars_gen + 8 * NBARS + 16);            // [EPI_WARPS] loss partials

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x;
  const int S = p.S, H = p.H, L = p.L, npad = p.npad;
  const int nbx = (npad + 31) >> 5;                 // 32-wide blocks of the layer's outputs: boxes per W k-block (forward),
                                                    // k-blocks of a dgrad GEMM, blocks of a dZ store
  constexpr bool in16 = PASSES == 4;      // input layer in kind::f16 (hidden layers as PASSES = 2)
  const int nkb0 = in16 ? p.nkb16 : p.kpad0 >> 5;
  const int nphase = 2 * L - 1;
  const uint32_t crank = CL > 1 ? cluster_ctarank() : 0u;
  constexpr uint16_t CMASK = (uint16_t)((1u << CL) - 1u);

  if (p.trace && threadIdx.x == 96) stamp_now(p, 0);
  pdl_launch_dependents();           // the weight-gradient launch behind this kernel is processed while it runs
  if (threadIdx.x == 0) {
    for (int s = 0; s < MAX_NA; ++s) { mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1); mbar_init(a_ready(s), EPI_WARPS); }
    for (int s = 0; s < MAX_RAW; ++s) {
      mbar_init(w_full(s), 1); mbar_init(w_empty(s), CL);
    }
    mbar_init(acc_full, 2);          // the MMAs' commit + the store thread (earlier TMA stores have left shared memory)
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(smem_u32((const void*)tmem_ptr_smem), 256);
  // zero the A region: rows the TMA boxes never touch (the b stream of the input layer) must read as zero
  for (int i = threadIdx.x; i < NA * ablk / 16; i += CT) reinterpret_cast<float4*>(a_gen)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  {  // head weights, zero beyond H
    const int nh = p.head ? 3 : 1;
    for (int i = threadIdx.x; i < 3 * HWS; i += CT) {
      const int h = i / HWS, c = i % HWS;
      hw[i] = (h < nh && c < H) ? __ldg(p.head_w + h * (H + 1) + c) : 0.f;
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();    // every CTA's barriers exist before a peer's multicast or commit can reach them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  stamp(p, 1);

  // Weight stream: ONE TMA instruction per "unit" (measured, scripts/tma_ingest_bench.cu: what an SM can pull through TMA is
  // set by the number of barrier phases / instructions, ~650 clk per ring slot + ~70 per box whatever their size, not by
  // bytes -- 7 boxes of 4 KB per slot deliver 24 B/clk, one 56 KB box 85 B/clk).  A unit fills a double slot (2 x 28 KB):
  //   phase 0, PASSES != 3:  two k-blocks of W_0 (TF32 hi)          3-D box {32 n, 64 k, nbx}      [nblk][64 k][32 n]
  //   phase 0, PASSES == 3:  one k-block of W_0, hi and lo          4-D box {32 n, 32 k, nbx, 2}   [hi|lo][nblk][32 k][32 n]
  //   forward l >= 1:        one k-block of W_l, hi (and lo)        4-D box {32 n, 32 k, nbx, HL}
  //   dgrad l:               one k-block (32 outputs) of W_l        3-D box {32 n, 32 nbx rows k, HL}  [hi|lo][k][32 n]
  constexpr int HL = PASSES >= 2 ? 2 : 1;            // copies of a hidden-layer unit: hi (+ lo)   (PASSES = 4: as 2)
  constexpr int WK0 = PASSES == 3 ? 1 : 2;           // k-blocks of W_0 per unit
  const int nunits = nraw >> 1;                      // ring depth in units
  const uint32_t unit0_bytes = in16 ? (uint32_t)npad * 256u                // fp16 input layer: two k-blocks of npad rows x 128 B
                                    : (uint32_t)nbx * 4096u * 2u;          // both layouts of phase 0 fill 2 x nbx boxes
  const uint32_t unith_bytes = (uint32_t)nbx * 4096u * (uint32_t)HL;
  // phase 0's A operand arrives in groups of AG k-blocks, one 4-D instruction and one barrier phase each (the same
  // per-barrier cost): group slot g covers the A-region blocks [AG g, AG g + AG); k-blocks past the end and the b stream are
  // out of bounds of the tensor map, i.e. land as zeros
  const int AG = p.ag, nag = NA / AG;

  if (warp == PROD_WARP) {
    // ===== TMA producer (one thread): the weight stream of every phase back to back, plus phase 0's A operand ============
    if (lane == 0) {
      pdl_wait();        // XV comes from the kernel before this one on the stream (programmatic launch: see launch_chain_cl)
      int wu = 0; uint32_t wph = 0; int useq = 0;
      auto next_unit = [&]() -> uint32_t {          // waits for the ring slot, arms its barrier; returns its address (0: a peer issues)
        mbar_wait_parked(w_empty(2 * wu), wph ^ 1u);                 // every CTA of the cluster has finished with this unit
        return w_base + 2 * wu * WSLOT_BYTES;
      };
      auto advance = [&]() { if (++wu == nunits) { wu = 0; wph ^= 1u; } ++useq; };
      for (int g = 0; g < nphase; ++g) {
        const bool fwd = g < L;
        const int layer = fwd ? g : 2 * L - 1 - g;
        const int nkb = fwd ? (g == 0 ? nkb0 : NA) : nbx;
        const CUtensorMap* tw = fwd ? &p.t_wf[layer] : &p.t_wb[layer];
        for (int kb = 0; kb < nkb; ++kb) {
          if (g == 0) {
            if (kb % AG == 0) {
              const int gs = (kb / AG) % nag;
              mbar_wait_parked(a_empty(gs), ((kb / AG / nag) & 1) ^ 1u);
              mbar_arrive_expect_tx(a_full(gs), (uint32_t)(AG * ablk));
              tma_load_4d(a_base + gs * AG * ablk, in16 ? &p.t_in16 : &p.t_in, a_full(gs), 0, tile * S, 0, kb);
            }
            if (kb % WK0 != 0) continue;
          }
          const uint32_t dst = next_unit();
          const uint32_t bytes = g == 0 ? unit0_bytes : unith_bytes;
          mbar_arrive_expect_tx(w_full(2 * wu), bytes);
          if (CL == 1 || (uint32_t)(useq % CL) == crank) {          // clusters: the CTAs take turns and multicast the unit
            if (g == 0 && in16) {
              if (CL > 1) tma_load_3d_mc(dst, &p.t_w16, w_full(2 * wu), 0, 0, kb, CMASK);
              else tma_load_3d(dst, &p.t_w16, w_full(2 * wu), 0, 0, kb);
            } else if (g == 0 && PASSES != 3) {
              if (CL > 1) tma_load_3d_mc(dst, tw, w_full(2 * wu), 0, 32 * kb, 0, CMASK);
              else tma_load_3d(dst, tw, w_full(2 * wu), 0, 32 * kb, 0);
            } else if (fwd) {
              if (CL > 1) tma_load_4d_mc(dst, tw, w_full(2 * wu), 0, 32 * kb, 0, 0, CMASK);
              else tma_load_4d(dst, tw, w_full(2 * wu), 0, 32 * kb, 0, 0);
            } else {
              if (CL > 1) tma_load_3d_mc(dst, tw, w_full(2 * wu), 32 * kb, 0, 0, CMASK);
              else tma_load_3d(dst, tw, w_full(2 * wu), 32 * kb, 0, 0);
            }
          }
          advance();
        }
      }
    }
  } else if (warp == MMA_WARP) {
    // ===== MMA issuer + TMA-store issuer (one thread) =====================================================================
    if (lane == 0) {
      int wu = 0; uint32_t wph = 0;
      // block kb of the output of epilogue `ge` leaves for HBM (operands of the weight-gradient GEMMs / of the backward)
      auto store_block = [&](int ge, int kb) {
        const CUtensorMap* tm;          // (the map's box carries the stream count: 3, or 2 for dZ_0)
        if (ge < L - 1) tm = &p.t_h[ge + 1];
        else if (ge == L - 1) tm = &p.t_dz[L - 1];
        else tm = &p.t_dz[2 * L - 2 - ge];
        tma_store_3d(tm, a_base + kb * ablk, kb * 32, tile * S, 0);
      };
      for (int g = 0; g < nphase; ++g) {
        const bool fwd = g < L;
        const int nkb = fwd ? (g == 0 ? nkb0 : NA) : nbx;
        const uint32_t idesc = make_idesc_tf32(128, npad, tc::MAJOR_K, fwd ? tc::MAJOR_MN : tc::MAJOR_K);
        // forward (MN-major B): n-blocks of a unit are 4 KB apart -- 8 KB in phase 0's two-k-block units
        const uint32_t b_lbo = fwd ? ((g == 0 && WK0 == 2) ? 8192u : 4096u) : 16u;
        const uint32_t b_sbo = fwd ? 512u : 1024u, b_step = fwd ? 1024u : 32u, b_lay = fwd ? 1u : 2u;
        // contraction length that carries data: input columns + the ones column (forward), the layer's outputs (dgrad);
        // 8-column steps of a k-block that lie entirely in the zero padding are not issued
        const int kvalid = fwd ? (g == 0 ? p.D + 1 : H + 1) : H;
        auto mmas = [&](int as_, uint32_t bbase, bool first, int kb_) {
          const int n8 = min(4, (kvalid - 32 * kb_ + 7) >> 3);
#pragma unroll
          for (int k8 = 0; k8 < 4; ++k8) {
            if (k8 >= n8) break;
            const uint64_t da = make_smem_desc(a_base + as_ * ablk + k8 * 32, 16u, 1024u, 2u);
            const uint64_t db = make_smem_desc(bbase + k8 * b_step, b_lbo, b_sbo, b_lay);
            mma_tf32_ss(tmem_base, da, db, idesc, (first && k8 == 0) ? 0u : 1u);
          }
        };
        // fp16 input layer (PASSES = 4, phase 0): both operands K-major (the data-gradient phases' descriptor form), k-blocks
        // of 64 columns, 16 per MMA
        auto mmas16 = [&](int as_, uint32_t bbase, bool first, int kb_) {
          const uint32_t idesc16 = make_idesc_f16(128, npad, tc::MAJOR_K, tc::MAJOR_K);
          const int n16 = min(4, (kvalid - 64 * kb_ + 15) >> 4);
#pragma unroll
          for (int k16 = 0; k16 < 4; ++k16) {
            if (k16 >= n16) break;
            const uint64_t da = make_smem_desc(a_base + as_ * ablk + k16 * 32, 16u, 1024u, 2u);
            const uint64_t db = make_smem_desc(bbase + k16 * 32, 16u, 1024u, 2u);
            mma_f16_ss(tmem_base, da, db, idesc16, (first && k16 == 0) ? 0u : 1u);
          }
        };
        auto release = [&]() {
          if (CL > 1) tc_commit_mc(w_empty(2 * wu), CMASK);
          else tc_commit(w_empty(2 * wu));
          if (++wu == nunits) { wu = 0; wph ^= 1u; }
        };
        // the second copy (lo) of a unit follows the first: nbx boxes (forward) / 32 nbx rows of 128 bytes (dgrad) further
        const uint32_t lo_off = (uint32_t)nbx * 4096u;
        for (int kb = 0; kb < nkb; ++kb) {
          const uint32_t ubase = w_base + 2 * wu * WSLOT_BYTES;
          if (g == 0) {
            const int gs = (kb / AG) % nag, as = gs * AG + kb % AG;
            if (kb % AG == 0) mbar_wait_parked(a_full(gs), (kb / AG / nag) & 1);
            if (kb % WK0 == 0) mbar_wait_parked(w_full(2 * wu), wph);
            if (p.trace && p.trace_mode == 3 && L == 3 && kb % 2 == 0 && kb < 32) stamp_now(p, 22 + kb / 2);   // (debug)
            tc_fence_after();
            if (in16) {
              mmas16(as, ubase + (kb & 1) * (uint32_t)npad * 128u, kb == 0, kb);
            } else if (WK0 == 2) {
              mmas(as, ubase + (kb & 1) * 4096u, kb == 0, kb);
            } else {
              mmas(as, ubase, kb == 0, kb);
              mmas(as, ubase + lo_off, false, kb);
            }
            if (kb % AG == AG - 1 || kb == nkb - 1) tc_commit(a_empty(gs));
            if (kb % WK0 == WK0 - 1 || kb == nkb - 1) release();
          } else {
            mbar_wait_parked(a_ready(kb), (g - 1) & 1);      // the epilogue has finished this k-block of the operand
            if (p.trace && p.trace_mode == 0 && g == 1 && L == 3 && kb < 8) stamp_now(p, 22 + 2 * kb);      // (debug: free slots of a 3-layer trace)
            store_block(g - 1, kb);
            mbar_wait_parked(w_full(2 * wu), wph);
            if (p.trace && p.trace_mode == 0 && g == 1 && L == 3 && kb < 8) stamp_now(p, 23 + 2 * kb);
            tc_fence_after();
            mmas(kb, ubase, kb == 0, kb);
            if (HL == 2) mmas(kb, ubase + lo_off, false, kb);
            release();
          }
        }
        if (p.trace) stamp_now(p, 5 + 4 * g);            // phase g's MMAs issued
        tc_commit(acc_full);
        if (g > 0) {
          // before the epilogue overwrites the A region, the stores of the previous epilogue must have read it; once,
          // before the backward re-reads H_l through plain loads, they must also have LANDED (full completion)
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          if (g == L - 1) {
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            asm volatile("fence.proxy.async;" ::: "memory");
          } else {
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          }
        }
        mbar_arrive(acc_full);
      }
      for (int kb = 0; kb < nbx; ++kb) {                  // output of the last epilogue (dZ_0, or dZ_{L-1} when L = 1)
        mbar_wait_parked(a_ready(kb), (nphase - 1) & 1);
        store_block(nphase - 1, kb);
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
  } else {
    // ===== epilogue warps ==================================================================================================
    const float off = p.act_offset;
    const int sst = S * 128;                            // byte distance between the h0 / a / b rows of a sample
    const int spw = (S - warp + EPI_WARPS - 1) / EPI_WARPS;   // samples of this warp: i = warp + 16 k, k < spw (0..3)
    // element (sample i, column lane + 32 j) of the A region: eb[k] + j * ablk.  S is a multiple of 8, so the three rows of
    // a sample share the swizzle term (row & 7 = i & 7)
    auto ebase = [&](int i) { return a_gen + i * 128 + ((((lane >> 2) ^ (i & 7)) << 4) | ((lane & 3) << 2)); };
    auto publish = [&](int j) {      // this warp's part of k-block j is final: visible to the tensor core and the TMA unit
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(a_ready(j));
    };
    float gw1[MAX_NA], gw2[MAX_NA], gwq[MAX_NA];        // this warp's head-gradient partial sums (set in phase L - 1)
    float gb1 = 0.f, gb2 = 0.f, gbq = 0.f, loss_acc = 0.f;
    for (int g = 0; g < nphase; ++g) {
      const bool fwd = g < L;
      const int layer = fwd ? g : 2 * L - 1 - g;
      mbar_wait_parked(acc_full, g & 1);
      tc_fence_after();
      stamp(p, 2 + 4 * g);         // accumulator complete
      // -- dump: accumulator rows -> A region (raw pre-activations / data gradients, fp32, swizzled K-major layout) --------
      {
        const int q = warp & 3, part = warp >> 2;
        const int row = q * 32 + lane;
        const int nch = npad >> 3, per = (nch + 3) >> 2;
        const int c0 = part * per, c1 = min(nch, c0 + per);
        const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
        uint8_t* rbase = a_gen + row * 128;
        for (int ch = c0; ch < c1; ++ch) {
          float v[8];
          tmem_ld_32x8(trow + (uint32_t)(ch * 8), v);
          if (row < 3 * S) {
            uint8_t* bp = rbase + (ch >> 2) * ablk;
            const int k16 = (ch & 3) * 2;               // 16-byte chunk of the 128-byte row
            *reinterpret_cast<float4*>(bp + ((k16 ^ (row & 7)) << 4)) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(bp + (((k16 + 1) ^ (row & 7)) << 4)) = make_float4(v[4], v[5], v[6], v[7]);
          }
        }
      }
      if (warp < 4 && threadIdx.x < 3 * S)            // columns [npad, kpadh): exact zeros (see the epilogue bodies)
        for (int c = npad; c < p.kpadh; c += 4)
          *reinterpret_cast<float4*>(a_gen + (c >> 5) * ablk + threadIdx.x * 128 + ((((c & 31) >> 2) ^ (threadIdx.x & 7)) << 4)) =
              make_float4(0.f, 0.f, 0.f, 0.f);
      tc_fence_before();
      named_bar_sync(1, EPI_T);
      stamp(p, 3 + 4 * g);         // dump done
      if (fwd && g < L - 1) {
        // -- hidden layer: jets, block by block -----------------------------------------------------------------------------
#pragma unroll 1
        for (int j = 0; j < NA; ++j) {
          const bool ones = lane + 32 * j == H;
#pragma unroll 1
          for (int k0 = 0; k0 < spw; k0 += 2) {
            const bool v1 = k0 + 1 < spw;
            uint8_t* e0 = ebase(warp + EPI_WARPS * k0) + j * ablk;
            uint8_t* e[2] = {e0, v1 ? e0 + EPI_WARPS * 128 : e0};     // 16 rows further: same row & 7, same swizzle term
            epi_fwd_hidden<ACT, MODE>(e, v1, sst, ones, off);
          }
          publish(j);
          if (p.trace && p.trace_mode == 2 && g == 1 && L == 3 && j == 1 && lane == 0) stamp_now(p, 22 + warp);
        }
        if (p.trace && p.trace_mode == 1 && g == 1 && L == 3 && lane == 0) stamp_now(p, 22 + warp);
      } else if (fwd) {
        // -- last forward layer: jets, head, per-sample loss and its gradient, activation backward -> dZ_{L-1} ---------------
        const bool quad = p.head != 0;
        const float b1 = __ldg(p.head_w + H), b2 = quad ? __ldg(p.head_w + 2 * H + 1) : 0.f, bq = quad ? __ldg(p.head_w + 3 * H + 2) : 0.f;
        auto sample_done = [&](const float (&sum)[9], int i) {
          const bool valid = tile * S + i < p.B;
          const float l = head_sample<MODE>(sum, quad, b1, b2, bq, p.c1, valid ? p.coef : 0.f, lane == 0, coef_s + i * NCOEF, gb1, gb2, gbq);
          if (valid) loss_acc += l;
        };
#pragma unroll 1
        for (int k0 = 0; k0 < spw; k0 += 2) {
          const bool v1 = k0 + 1 < spw;
          const int i0 = warp + EPI_WARPS * k0;
          uint8_t* e0 = ebase(i0);
          uint8_t* e[2] = {e0, v1 ? e0 + EPI_WARPS * 128 : e0};
          float sum[2][9];
          epi_head_pass1<ACT, MODE>(e, v1, sst, ablk, NA, lane, hw, quad, off, sum);
          sample_done(sum[0], i0);
          if (v1) sample_done(sum[1], i0 + EPI_WARPS);
        }
        __syncwarp();
        // pass 2, block by block (the per-column head-weight sums go to a runtime-indexed, i.e. local-memory, array; the tile
        // sums over the warps wait until the last phase is over: see below the phase loop)
#pragma unroll 1
        for (int j = 0; j < NA; ++j) {
          const int c = lane + 32 * j;
          const float w1 = hw[c], w2 = hw[HWS + c], wq = hw[2 * HWS + c];
          float s1 = 0.f, s2 = 0.f, sq = 0.f;
#pragma unroll 1
          for (int k0 = 0; k0 < spw; k0 += 2) {
            const bool v1 = k0 + 1 < spw;
            const int i0 = warp + EPI_WARPS * k0;
            uint8_t* e0 = ebase(i0) + j * ablk;
            uint8_t* e[2] = {e0, v1 ? e0 + EPI_WARPS * 128 : e0};
            const float* c0 = coef_s + i0 * NCOEF;
            const float* cs[2] = {c0, v1 ? c0 + EPI_WARPS * NCOEF : c0};
            epi_head_pass2<ACT, MODE>(e, v1, sst, w1, w2, wq, cs, quad, off, s1, s2, sq);
          }
          gw1[j] = s1; gw2[j] = s2; gwq[j] = sq;
          publish(j);
        }
      } else {
        // -- dgrad epilogue of layer `layer`: rows hold dL/dH_layer as (g0, ga, gb); H_layer comes back from L2 --------------
        const long long ldh = p.kpadh, hstride = (long long)p.Bp * ldh;
        const float* hcol = p.h[layer] + (long long)(tile * S) * ldh + lane;
#pragma unroll 1
        for (int j = 0; j < NA; ++j) {
#pragma unroll 1
          for (int k0 = 0; k0 < spw; k0 += 2) {
            const bool v1 = k0 + 1 < spw;
            const int i0 = warp + EPI_WARPS * k0;
            uint8_t* e0 = ebase(i0) + j * ablk;
            uint8_t* e[2] = {e0, v1 ? e0 + EPI_WARPS * 128 : e0};
            const float* h0 = hcol + i0 * ldh + 32 * j;
            const float* hp[2] = {h0, v1 ? h0 + EPI_WARPS * ldh : h0};
            epi_dgrad<ACT, MODE>(e, v1, sst, hp, hstride, off);
          }
          publish(j);
        }
      }
      stamp(p, 4 + 4 * g);         // this warp's part of the epilogue done
    }
    // -- head-parameter gradient and loss of the tile, behind the last phase instead of between the head epilogue and the
    //    first data-gradient epilogue (4.5 us there, in front of every warp's next accumulator dump).  The weight ring is idle by
    //    now (every slot was consumed by MMAs that completed before the last accumulator was handed over; nraw >= 2 slots =
    //    56 KB >= 16 records): each warp parks its partial sums in a record of its own, then every output adds the sixteen in
    //    the fixed order of the former four rounds: ((w + w4) + w8) + w12 for q = w & 3, then (q0 + q1) + (q2 + q3).
    {
      fence_proxy_async();
      float* park = reinterpret_cast<float*>(w_gen) + warp * HREC;
#pragma unroll 1
      for (int j = 0; j < NA; ++j) {
        const int c = lane + 32 * j;
        park[c] = gw1[j]; park[HWS + c] = gw2[j]; park[2 * HWS + c] = gwq[j];
      }
      if (lane == 0) { park[3 * HWS] = gb1; park[3 * HWS + 1] = gb2; park[3 * HWS + 2] = gbq; park[3 * HWS + 3] = loss_acc; }
      named_bar_sync(1, EPI_T);
      const float* pk = reinterpret_cast<const float*>(w_gen);
      const int nh = p.head != 0 ? 3 : 1, head_n = nh * (H + 1);
      for (int o = threadIdx.x; o < head_n; o += EPI_T) {
        const int h = o / (H + 1), c = o % (H + 1);
        const int src = c < H ? h * HWS + c : 3 * HWS + h;
        float r[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
          r[q] = ((pk[q * HREC + src] + pk[(q + 4) * HREC + src]) + pk[(q + 8) * HREC + src]) + pk[(q + 12) * HREC + src];
        p.head_part[(long long)tile * head_n + o] = (r[0] + r[1]) + (r[2] + r[3]);
      }
      if (threadIdx.x == 0) {
        float sl = 0.f;
#pragma unroll
        for (int w = 0; w < EPI_WARPS; ++w) sl += pk[w * HREC + 3 * HWS + 3];
        p.loss_part[tile] = sl;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();    // no CTA leaves while a peer may still multicast into it / arrive on its barriers
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

}  // namespace jet
}  // namespace fdsm

namespace fdsm {
namespace jet {

// ---- the chain's input: XV = [x | 1 | 0.. ; v | 0..] as TF32-rounded rows, v = eps u/||u|| (Philox stream 0 or injected) ----
// One group of 128 threads per sample.
struct InputArgs {
  const float* x; long long ldx;
  const unsigned char* x8; float x8_scale, x_noise;   // x8 != NULL: x = x8 * x8_scale + x_noise * N(0,1)  (Philox STREAM_NOISE)
  const float* dir; long long lddir; int normalize; float scale;
  unsigned long long seed, offset, row0; const unsigned long long* offset_dev;
  float* xv; int B, Bp, D, kpad0;
  __half* xv16; int kpad16;          // optional fp16 copy of the rows for the chain's kind::f16 input layer (NULL: none)
};
// FAST = compile-time specialisation of the two production configurations (the mode tests fold away and the compiler
// interleaves the Philox chains of a thread's chunks, as for the stack kernel's MODE_ENERGY3): 1 = fp32 samples, 2 = uint8
// pixels + Gaussian noise; both: Philox direction, normalised, 16-byte aligned rows, D <= 1024.  0 = everything else.
// Bit-identical outputs (tests/test_jet_gpu.py compares against FDSM_JET_INPUTS_GENERAL=1).
template <int FAST, bool IN16 = false>
__device__ __forceinline__ void jet_inputs_body(InputArgs a, const int block, const int nblocks) {
  constexpr int G = 128, MAXC = 2;
  __shared__ float red[256 / 32];
  if (a.offset_dev) a.offset += __ldg(a.offset_dev);
  if (FAST) { a.dir = nullptr; a.normalize = 1; }
  if (FAST == 1) a.x8 = nullptr;
  if (FAST == 2) a.x = nullptr;
  const int gid = threadIdx.x / G, t = threadIdx.x % G;
  const int D = a.D, nchunk = (D + 3) >> 2;
  const bool xvec = FAST ? true : a.x8 ? ((D & 3) == 0 && (a.ldx & 3) == 0 && (reinterpret_cast<uintptr_t>(a.x8) & 3) == 0)
                                       : ((D & 3) == 0 && (a.ldx & 3) == 0 && aligned16(a.x));
  const bool dvec = !FAST && a.dir && (D & 3) == 0 && (a.lddir & 3) == 0 && aligned16(a.dir);
  const long long ngroups = (a.Bp + 1) / 2;
  for (long long rg = block; rg < ngroups; rg += nblocks) {
    const long long b = rg * 2 + gid;
    const bool active = b < a.B;
    auto raw = [&](int j) -> float4 {
      const int e = 4 * j;
      if (!FAST && a.dir) {
        const float* r = a.dir + b * a.lddir;
        if (dvec) return ldg4(r + e);
        return make_float4(e < D ? __ldg(r + e) : 0.f, e + 1 < D ? __ldg(r + e + 1) : 0.f, e + 2 < D ? __ldg(r + e + 2) : 0.f,
                           e + 3 < D ? __ldg(r + e + 3) : 0.f);
      }
      float4 n = philox_normal4(a.seed, a.offset, STREAM_DIRECTION, a.row0 + (unsigned long long)b, (uint32_t)j);
      if (!FAST) {       // (FAST: D % 4 == 0, every chunk is whole)
        if (e + 1 >= D) n.y = 0.f;
        if (e + 2 >= D) n.z = 0.f;
        if (e + 3 >= D) n.w = 0.f;
      }
      return n;
    };
    // the sample row: fp32, or uint8 pixels scaled (torchvision ToTensor: / 255) plus the on-device noise of
    // DeepEBM/train_ssm.py:62 (x + ssm_eps * randn) drawn from Philox stream 1 -- the host ships 1 byte per pixel
    auto load_x = [&](int j) -> float4 {
      const int e = 4 * j;
      if (FAST == 2 || (!FAST && a.x8)) {
        const unsigned char* r = a.x8 + b * a.ldx;
        uchar4 q;
        if (xvec) q = *reinterpret_cast<const uchar4*>(r + e);
        else q = make_uchar4(e < D ? r[e] : 0, e + 1 < D ? r[e + 1] : 0, e + 2 < D ? r[e + 2] : 0, e + 3 < D ? r[e + 3] : 0);
        float4 xx = make_float4(q.x * a.x8_scale, q.y * a.x8_scale, q.z * a.x8_scale, q.w * a.x8_scale);
        if (FAST == 2 || a.x_noise != 0.f) {
          const float4 n = philox_normal4(a.seed, a.offset, STREAM_NOISE, a.row0 + (unsigned long long)b, (uint32_t)j);
          xx.x = fmaf(a.x_noise, n.x, xx.x); xx.y = fmaf(a.x_noise, n.y, xx.y); xx.z = fmaf(a.x_noise, n.z, xx.z);
          xx.w = fmaf(a.x_noise, n.w, xx.w);
        }
        if (!FAST) {
          if (e + 1 >= D) xx.y = 0.f;
          if (e + 2 >= D) xx.z = 0.f;
          if (e + 3 >= D) xx.w = 0.f;
        }
        return xx;
      }
      const float* xrow = a.x + b * a.ldx;
      if (xvec) return ldg4(xrow + e);
      return make_float4(e < D ? __ldg(xrow + e) : 0.f, e + 1 < D ? __ldg(xrow + e + 1) : 0.f, e + 2 < D ? __ldg(xrow + e + 2) : 0.f,
                         e + 3 < D ? __ldg(xrow + e + 3) : 0.f);
    };
    // the draw stays in registers between the norm pass and the write pass when the row fits (D <= 4 MAXC G = 1024);
    // longer rows regenerate it (counter-based: identical numbers)
    float4 dreg[MAXC], xreg[MAXC];
    float ss = 0.f;
    if (active) {
      // the sample row is requested BEFORE the Philox / Box-Muller arithmetic: the HBM latency hides under it
#pragma unroll
      for (int k = 0; k < MAXC; ++k) {
        const int j = t + k * G;
        xreg[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (j < nchunk) xreg[k] = load_x(j);
      }
#pragma unroll
      for (int k = 0; k < MAXC; ++k) {
        const int j = t + k * G;
        dreg[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (j < nchunk) {
          dreg[k] = raw(j);
          ss += dreg[k].x * dreg[k].x + dreg[k].y * dreg[k].y + dreg[k].z * dreg[k].z + dreg[k].w * dreg[k].w;
        }
      }
      if (!FAST && a.normalize)
        for (int j = t + MAXC * G; j < nchunk; j += G) {
          const float4 d = raw(j);
          ss += d.x * d.x + d.y * d.y + d.z * d.z + d.w * d.w;
        }
    }
    float nrm = 1.f;
    if (a.normalize) nrm = sqrtf(group_sum<G>(ss, red));     // uniform
    if (b >= a.Bp) continue;
    float* xr = a.xv + b * (long long)a.kpad0;
    float* vr = a.xv + ((long long)a.Bp + b) * a.kpad0;
    __half* xr16 = IN16 ? a.xv16 + b * (long long)a.kpad16 : nullptr;
    __half* vr16 = IN16 ? a.xv16 + ((long long)a.Bp + b) * a.kpad16 : nullptr;
    if (!active) {          // rows of the padded tail: zeros (finite everywhere downstream, masked out of every sum)
      for (int c = t; c < a.kpad0; c += G) { xr[c] = 0.f; vr[c] = 0.f; }
      if (IN16) for (int c = t; c < a.kpad16; c += G) { xr16[c] = __float2half_rn(0.f); vr16[c] = __float2half_rn(0.f); }
      continue;
    }
    const bool exact = !FAST && a.dir != nullptr;     // injected draws keep the reference's rounding sequence u / ||u|| * eps
    const float inv = a.normalize && !exact ? __fdiv_rn(a.scale, nrm) : 1.f;
    for (int j = t, k = 0; j < nchunk; j += G, ++k) {
      const int e = 4 * j;
      float4 d, xx;
      if (k == 0) d = dreg[0];
      else if (k == 1) d = dreg[MAXC - 1];
      else d = raw(j);
      xx = k == 0 ? xreg[0] : k == 1 ? xreg[MAXC - 1] : load_x(j);
      float4 v;
      if (!a.normalize) v = d;
      else if (exact) v = make_float4(__fmul_rn(__fdiv_rn(d.x, nrm), a.scale), __fmul_rn(__fdiv_rn(d.y, nrm), a.scale),
                                      __fmul_rn(__fdiv_rn(d.z, nrm), a.scale), __fmul_rn(__fdiv_rn(d.w, nrm), a.scale));
      else v = make_float4(d.x * inv, d.y * inv, d.z * inv, d.w * inv);
      // kpad0 is a multiple of 32 and the buffer is 16-byte aligned: the chunk store is always in range and aligned.
      // A partial last chunk (D % 4 != 0) carries the first padding columns itself: x gets the ones column at D, v zeros
      if (!FAST) {
        if (D - e == 1) xx.y = 1.f;
        if (D - e == 2) xx.z = 1.f;
        if (D - e == 3) xx.w = 1.f;
      }
      *reinterpret_cast<float4*>(xr + e) = make_float4(tf32r(xx.x), tf32r(xx.y), tf32r(xx.z), tf32r(xx.w));
      *reinterpret_cast<float4*>(vr + e) = make_float4(tf32r(v.x), tf32r(v.y), tf32r(v.z), tf32r(v.w));
      if (IN16) {           // (kpad16 is a multiple of 64: the 8-byte chunk store is in range and aligned)
        const __half2 x01 = __floats2half2_rn(xx.x, xx.y), x23 = __floats2half2_rn(xx.z, xx.w);
        const __half2 v01 = __floats2half2_rn(v.x, v.y), v23 = __floats2half2_rn(v.z, v.w);
        uint2 px, pv;
        px.x = *reinterpret_cast<const unsigned*>(&x01); px.y = *reinterpret_cast<const unsigned*>(&x23);
        pv.x = *reinterpret_cast<const unsigned*>(&v01); pv.y = *reinterpret_cast<const unsigned*>(&v23);
        *reinterpret_cast<uint2*>(xr16 + e) = px;
        *reinterpret_cast<uint2*>(vr16 + e) = pv;
      }
    }
    for (int c = 4 * nchunk + t; c < a.kpad0; c += G) { xr[c] = c == D ? 1.f : 0.f; vr[c] = 0.f; }
    if (IN16) for (int c = 4 * nchunk + t; c < a.kpad16; c += G) { xr16[c] = __float2half_rn(c == D ? 1.f : 0.f); vr16[c] = __float2half_rn(0.f); }
  }
}

template <int FAST>
__global__ void __launch_bounds__(256, 4) jet_inputs_kernel(InputArgs a) {
  pdl_launch_dependents();      // the chain kernel behind this grid sets up (barriers, TMEM, head weights) as blocks here retire
  jet_inputs_body<FAST>(a, blockIdx.x, gridDim.x);
}

// ---- after the chain: loss and head-parameter gradient from the per-tile records (fixed order: deterministic) ------------
// grid = ceil(head_n / 32) blocks of 32 x 32 threads: thread (x, y) adds column 32 blockIdx + x over tiles y, y + 32, ...;
// the 32 partial sums of a column are then added in order.  Block 0 also reduces the loss.
__global__ void __launch_bounds__(1024) jet_finish_kernel(const float* __restrict__ loss_part, const float* __restrict__ head_part,
                                                          int tiles, int head_n, float loss_scale, float* __restrict__ loss,
                                                          float* __restrict__ ghead, int accumulate) {
  __shared__ float part[32][33];
  const int x = threadIdx.x & 31, y = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + x;
  float s = 0.f;
  if (j < head_n)
    for (int t = y; t < tiles; t += 32) s += __ldcg(head_part + (long long)t * head_n + j);
  part[y][x] = s;
  __syncthreads();
  if (y == 0 && j < head_n) {
    double d = 0.;
#pragma unroll
    for (int k = 0; k < 32; ++k) d += (double)part[k][x];
    ghead[j] = accumulate ? ghead[j] + (float)d : (float)d;
  }
  if (loss && blockIdx.x == 0 && y == 1) {
    double d = 0.;
    for (int t = x; t < tiles; t += 32) d += (double)__ldcg(loss_part + t);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if (x == 0) *loss = (float)(d * (double)loss_scale);
  }
}

}  // namespace jet
}  // namespace fdsm
