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
// Weights reach the CTAs by cluster multicast (CL = 1, 2 or 4 CTAs per cluster): every CTA of a cluster walks the same
// weight stream, so each loads 1/CL of a ring slot and multicasts it to all -- the L2 read traffic of the weight stream,
// which bounds the un-clustered kernel (every CTA re-reads every W), drops by CL.
//
// Shared memory: A region NA x 16 KB (NA = ceil32(H+1)/32 k-blocks; the TMA ring of phase 0, then the resident operand)
// + W ring 3 x 28 KB (one k-block of W_hi or W_lo per slot) + head weights / head-gradient records.  TMEM: one
// 128 x npad fp32 accumulator.  Warps 0..15 run the epilogues (warp w owns samples w, w+16, ...; lane = column mod 32);
// warp 16 lane 0 = TMA producer, warp 17 lane 0 = MMA + TMA-store issuer; warp 2 also allocates TMEM.
#pragma once
#include "fd_jet.cuh"

namespace fdsm {
namespace jet {

constexpr int MAXL = 5;              // hidden layers the chain supports (CUtensorMap budget of the parameter space)
constexpr int EPI_WARPS = 16, EPI_T = EPI_WARPS * 32;
constexpr int CT = EPI_T + 64;       // + producer warp + MMA/store warp
constexpr int WSLOTS = 3, WSLOT_BYTES = 28672;   // 7 boxes of 32 x 32 fp32 (4 KB each)
constexpr int MAX_NA = 7;            // ceil32(H+1) <= 224
constexpr int MAX_S = 40, MAX_SPW = (MAX_S + EPI_WARPS - 1) / EPI_WARPS;   // samples per tile / per epilogue warp
constexpr int A_KB_BYTES = 128 * 128;   // one k-block of the A operand: 128 rows x 32 fp32
constexpr int HWS = 224;             // head-weight row stride in shared memory
constexpr int HREC = 3 * HWS + 4;    // one head-gradient record: [gw1 | gw2 | gwq | gb1 gb2 gbq .]
constexpr int NCOEF = 12;            // per-sample head-gradient coefficients kept in shared memory between the two passes
constexpr int NBARS = 2 * MAX_NA + 2 * WSLOTS + 1 + MAX_NA;

struct alignas(64) ChainParams {
  CUtensorMap t_in;                  // [2 Bp x kpad0]  rows [x | v]                          (load, box {32, S})
  CUtensorMap t_wf[MAXL];            // Wt_l [2 kpad_l x npad] (hi rows, then lo rows), MN-major view (load, box {32, 32})
  CUtensorMap t_wb[MAXL];            // the same memory, K-major view for the dgrad          (load, box {32, 32})
  CUtensorMap t_h[MAXL];             // H_l [3 Bp x kpadh], l = 1..L-1                       (store, box {32, S})
  CUtensorMap t_dz[MAXL];            // dZ_l [3 Bp x npad] (l = 0: 2 Bp rows)                (store, box {32, S})
  const float* h[MAXL];              // H_l again, for the backward's plain loads
  const float* head_w;               // [w1(H) b1 | w2(H) b2 | wq(H) bq]  or  [w(H) b]
  float* loss_part;                  // [tiles]   sum over the tile's samples of the per-sample loss (unscaled)
  float* head_part;                  // [tiles][head_n]  head-parameter gradient of the tile
  int B, Bp, S, H, L, npad, kpad0, kpadh, head;
  float act_offset;                  // ln 2 for the normalised softplus
  float c1, coef;                    // loss_b = dE^2/8 - sE + (c1 - 2) E0 ; dL/dloss_b = coef (= scale / B or scale)
  unsigned long long* trace;         // debug: per-CTA %globaltimer stamps [tile][TRACE_N] (NULL in production)
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

__host__ __device__ constexpr int chain_smem_bytes(int NA) {
  return 1024 + NA * A_KB_BYTES + WSLOTS * WSLOT_BYTES + 3 * HWS * 4 + 4 * HREC * 4 + MAX_S * NCOEF * 4 + NBARS * 8 + 16 + EPI_WARPS * 4 + 64;
}

// ---- shared-memory addressing of the A operand (K-major, 128-byte swizzle): element (row, col) ---------------------------
__device__ __forceinline__ uint32_t a_off(int row, int col) {
  return (uint32_t)((col >> 5) * A_KB_BYTES + row * 128 + (((((col & 31) >> 2) ^ (row & 7)) << 4) | ((col & 3) << 2)));
}

__device__ __forceinline__ void tma_store_2d(const void* tmap, uint32_t src, int32_t c0, int32_t c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"((uint64_t)tmap), "r"(src), "r"(c0), "r"(c1)
               : "memory");
}

template <int ACT, int PASSES, int CL>
__global__ void __launch_bounds__(CT, 1) chain_kernel(const __grid_constant__ ChainParams p) {
  using namespace ptx;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - raw);
  const int NA = p.kpadh >> 5;
  const uint32_t a_base = base;
  const uint32_t w_base = base + NA * A_KB_BYTES;
  uint8_t* a_gen = gen;
  uint8_t* w_gen = gen + NA * A_KB_BYTES;
  float* hw = reinterpret_cast<float*>(w_gen + WSLOTS * WSLOT_BYTES);          // [3][HWS]    head weights
  float* hrec = hw + 3 * HWS;                                                  // [4][HREC]   head-gradient records
  float* coef_s = hrec + 4 * HREC;                                             // [MAX_S][NCOEF]
  uint8_t* bars_gen = reinterpret_cast<uint8_t*>(coef_s + MAX_S * NCOEF);
  const uint32_t bars = w_base + WSLOTS * WSLOT_BYTES + (3 * HWS + 4 * HREC + MAX_S * NCOEF) * 4;
  auto a_full = [&](int s) { return bars + 8u * s; };
  auto a_empty = [&](int s) { return bars + 8u * (MAX_NA + s); };
  auto w_full = [&](int s) { return bars + 8u * (2 * MAX_NA + s); };
  auto w_empty = [&](int s) { return bars + 8u * (2 * MAX_NA + WSLOTS + s); };
  const uint32_t acc_full = bars + 8u * (2 * MAX_NA + 2 * WSLOTS);
  auto a_ready = [&](int j) { return bars + 8u * (2 * MAX_NA + 2 * WSLOTS + 1 + j); };
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(bars_gen + 8 * NBARS);
  float* red = reinterpret_cast<float*>(bars_gen + 8 * NBARS + 16);            // [EPI_WARPS] loss partials

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x;
  const int S = p.S, H = p.H, L = p.L, npad = p.npad;
  const int nbx = (npad + 31) >> 5;                 // 32-wide blocks of the layer's outputs: boxes per W k-block (forward),
                                                    // k-blocks of a dgrad GEMM, blocks of a dZ store
  const int nkb0 = p.kpad0 >> 5;
  const int nphase = 2 * L - 1;
  const uint32_t crank = CL > 1 ? cluster_ctarank() : 0u;
  constexpr uint16_t CMASK = (uint16_t)((1u << CL) - 1u);

  if (p.trace && threadIdx.x == 96) stamp_now(p, 0);
  if (threadIdx.x == 0) {
    for (int s = 0; s < MAX_NA; ++s) { mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1); mbar_init(a_ready(s), EPI_WARPS); }
    for (int s = 0; s < WSLOTS; ++s) { mbar_init(w_full(s), 1); mbar_init(w_empty(s), CL); }
    mbar_init(acc_full, 2);          // the MMAs' commit + the store thread (earlier TMA stores have left shared memory)
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(smem_u32((const void*)tmem_ptr_smem), 256);
  // zero the A region: rows the TMA boxes never touch (the b stream of the input layer, rows >= 3S) must read as zero
  for (int i = threadIdx.x; i < NA * A_KB_BYTES / 16; i += CT) reinterpret_cast<float4*>(a_gen)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
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

  if (warp == EPI_WARPS) {
    // ===== TMA producer (one thread): the weight stream of every phase back to back, plus phase 0's A operand ============
    if (lane == 0) {
      int ws = 0; uint32_t wph = 0;
      for (int g = 0; g < nphase; ++g) {
        const bool fwd = g < L;
        const int layer = fwd ? g : 2 * L - 1 - g;
        const int nkb = fwd ? (g == 0 ? nkb0 : NA) : nbx;
        const int kpad_l = layer == 0 ? p.kpad0 : p.kpadh;
        for (int kb = 0; kb < nkb; ++kb) {
          if (g == 0) {
            const int as = kb % NA;
            mbar_wait(a_empty(as), ((kb / NA) & 1) ^ 1u);
            mbar_arrive_expect_tx(a_full(as), (uint32_t)(2 * S * 128));
            tma_load_2d(a_base + as * A_KB_BYTES, &p.t_in, a_full(as), kb * 32, tile * S);
            tma_load_2d(a_base + as * A_KB_BYTES + S * 128, &p.t_in, a_full(as), kb * 32, p.Bp + tile * S);
          }
          for (int ps = 0; ps < PASSES; ++ps) {
            mbar_wait(w_empty(ws), wph ^ 1u);           // every CTA of the cluster has finished reading this slot
            mbar_arrive_expect_tx(w_full(ws), (uint32_t)nbx * 4096u);
            const uint32_t dst = w_base + ws * WSLOT_BYTES;
            for (int i = 0; i < nbx; ++i) {
              if (CL > 1 && (uint32_t)(i % CL) != crank) continue;
              // forward (W MN-major: box = 32 outputs x 32 k):  (32 i, k-block);  dgrad (K-major: 32 k x 32 rows): (k-block, 32 i)
              const int c0 = fwd ? 32 * i : 32 * kb, c1 = ps * kpad_l + (fwd ? 32 * kb : 32 * i);
              const CUtensorMap* tw = fwd ? &p.t_wf[layer] : &p.t_wb[layer];
              if (CL > 1) tma_load_2d_mc(dst + i * 4096, tw, w_full(ws), c0, c1, CMASK);
              else tma_load_2d(dst + i * 4096, tw, w_full(ws), c0, c1);
            }
            if (++ws == WSLOTS) { ws = 0; wph ^= 1u; }
          }
        }
      }
    }
  } else if (warp == EPI_WARPS + 1) {
    // ===== MMA issuer + TMA-store issuer (one thread) =====================================================================
    if (lane == 0) {
      int ws = 0; uint32_t wph = 0;
      // block kb of the output of epilogue `ge` leaves for HBM (operands of the weight-gradient GEMMs / of the backward)
      auto store_block = [&](int ge, int kb) {
        const CUtensorMap* tm;
        int nstream = 3;
        if (ge < L - 1) tm = &p.t_h[ge + 1];
        else if (ge == L - 1) { tm = &p.t_dz[L - 1]; nstream = L == 1 ? 2 : 3; }
        else { const int lay = 2 * L - 1 - ge; tm = &p.t_dz[lay - 1]; nstream = lay == 1 ? 2 : 3; }
        for (int s = 0; s < nstream; ++s) tma_store_2d(tm, a_base + kb * A_KB_BYTES + s * S * 128, kb * 32, s * p.Bp + tile * S);
      };
      for (int g = 0; g < nphase; ++g) {
        const bool fwd = g < L;
        const int nkb = fwd ? (g == 0 ? nkb0 : NA) : nbx;
        const uint32_t idesc = make_idesc_tf32(128, npad, tc::MAJOR_K, fwd ? tc::MAJOR_MN : tc::MAJOR_K);
        const uint32_t b_lbo = fwd ? 4096u : 16u, b_sbo = fwd ? 512u : 1024u, b_step = fwd ? 1024u : 32u, b_lay = fwd ? 1u : 2u;
        for (int kb = 0; kb < nkb; ++kb) {
          const int as = g == 0 ? kb % NA : kb;
          if (g == 0) {
            mbar_wait(a_full(as), (kb / NA) & 1);
          } else {
            mbar_wait(a_ready(kb), (g - 1) & 1);        // the epilogue has finished this k-block of the operand
            store_block(g - 1, kb);
          }
          tc_fence_after();
          for (int ps = 0; ps < PASSES; ++ps) {
            mbar_wait(w_full(ws), wph);
            tc_fence_after();
            const uint32_t bsl = w_base + ws * WSLOT_BYTES;
#pragma unroll
            for (int k8 = 0; k8 < 4; ++k8) {
              const uint64_t da = make_smem_desc(a_base + as * A_KB_BYTES + k8 * 32, 16u, 1024u, 2u);
              const uint64_t db = make_smem_desc(bsl + k8 * b_step, b_lbo, b_sbo, b_lay);
              mma_tf32_ss(tmem_base, da, db, idesc, (kb > 0 || ps > 0 || k8 > 0) ? 1u : 0u);
            }
            if (CL > 1) tc_commit_mc(w_empty(ws), CMASK);
            else tc_commit(w_empty(ws));
            if (++ws == WSLOTS) { ws = 0; wph ^= 1u; }
          }
          if (g == 0) tc_commit(a_empty(as));
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
        mbar_wait(a_ready(kb), (nphase - 1) & 1);
        store_block(nphase - 1, kb);
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
  } else {
    // ===== epilogue warps ==================================================================================================
    const float off = p.act_offset;
    auto A = [&](int row, int col) -> float& { return *reinterpret_cast<float*>(a_gen + a_off(row, col)); };
    auto publish = [&](int j) {      // this warp's part of k-block j is final: visible to the tensor core and the TMA unit
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(a_ready(j));
    };
    for (int g = 0; g < nphase; ++g) {
      const bool fwd = g < L;
      const int layer = fwd ? g : 2 * L - 1 - g;
      mbar_wait(acc_full, g & 1);
      tc_fence_after();
      stamp(p, 2 + 4 * g);         // accumulator complete
      // -- dump: accumulator rows -> A region (raw pre-activations / data gradients, fp32, swizzled K-major layout) --------
      {
        const int q = warp & 3, part = warp >> 2;
        const int row = q * 32 + lane;
        const int nch = npad >> 3, per = (nch + 3) >> 2;
        const int c0 = part * per, c1 = min(nch, c0 + per);
        const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
        for (int ch = c0; ch < c1; ++ch) {
          float v[8];
          tmem_ld_32x8(trow + (uint32_t)(ch * 8), v);
          if (row < 3 * S) {
            const int col = ch * 8;
            *reinterpret_cast<float4*>(a_gen + a_off(row, col)) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(a_gen + a_off(row, col + 4)) = make_float4(v[4], v[5], v[6], v[7]);
          }
        }
      }
      tc_fence_before();
      named_bar_sync(1, EPI_T);
      stamp(p, 3 + 4 * g);         // dump done
      if (fwd && g < L - 1) {
        // -- hidden layer: jets, block by block -----------------------------------------------------------------------------
        for (int j = 0; j < NA; ++j) {
          const int c = lane + 32 * j;
#pragma unroll
          for (int k = 0; k < MAX_SPW; ++k) {
            const int i = warp + EPI_WARPS * k;
            if (i < S) {
              Jet o = {0.f, 0.f, 0.f};
              if (c < H) {
                o = jet_fwd<ACT>(A(i, c), A(S + i, c), A(2 * S + i, c), off);
                o.h0 = tf32r(o.h0); o.a = tf32r(o.a); o.b = tf32r(o.b);
              } else if (c == H) {
                o.h0 = 1.f;             // the ones column: folds the next layer's bias into its GEMM (h0 stream only)
              }
              A(i, c) = o.h0; A(S + i, c) = o.a; A(2 * S + i, c) = o.b;
            }
          }
          publish(j);
        }
      } else if (fwd) {
        // -- last forward layer: jets, head, per-sample loss and its gradient, activation backward -> dZ_{L-1} ---------------
        const bool quad = p.head != 0;
        const float b1 = __ldg(p.head_w + H), b2 = quad ? __ldg(p.head_w + 2 * H + 1) : 0.f, bq = quad ? __ldg(p.head_w + 3 * H + 2) : 0.f;
        float gb1 = 0.f, gb2 = 0.f, gbq = 0.f, loss_acc = 0.f;
        // pass 1, sample by sample: jets (kept unrounded in place), the head's nine sums, the per-sample coefficients
        for (int i = warp; i < S; i += EPI_WARPS) {
          float l1 = 0.f, l2 = 0.f, A1 = 0.f, A2 = 0.f, B1 = 0.f, B2 = 0.f, qha = 0.f, qs = 0.f, q0 = 0.f;
          for (int j = 0; j < NA; ++j) {
            const int c = lane + 32 * j;
            Jet o = {0.f, 0.f, 0.f};
            if (c < H) o = jet_fwd<ACT>(A(i, c), A(S + i, c), A(2 * S + i, c), off);
            A(i, c) = o.h0; A(S + i, c) = o.a; A(2 * S + i, c) = o.b;
            const float w1 = hw[c], w2 = hw[HWS + c], wq = hw[2 * HWS + c];
            l1 = fmaf(w1, o.h0, l1); A1 = fmaf(w1, o.a, A1); B1 = fmaf(w1, o.b, B1);
            if (quad) {
              l2 = fmaf(w2, o.h0, l2); A2 = fmaf(w2, o.a, A2); B2 = fmaf(w2, o.b, B2);
              qha = fmaf(wq, o.a * (o.b + o.h0), qha);
              qs = fmaf(wq, fmaf(o.b, o.b, fmaf(o.a, o.a, 2.f * o.h0 * o.b)), qs);
              q0 = fmaf(wq, o.h0 * o.h0, q0);
            }
          }
          l1 = warp_sum(l1); A1 = warp_sum(A1); B1 = warp_sum(B1);
          if (quad) {
            l2 = warp_sum(l2); A2 = warp_sum(A2); B2 = warp_sum(B2);
            qha = warp_sum(qha); qs = warp_sum(qs); q0 = warp_sum(q0);
          }
          l1 += b1; l2 += b2;
          float dE, sE, E0;
          if (quad) {
            dE = 2.f * (A1 * (l2 + B2) + A2 * (l1 + B1)) + 4.f * qha;
            sE = 2.f * (l1 * B2 + B1 * l2 + B1 * B2 + A1 * A2) + 2.f * qs;
            E0 = fmaf(l1, l2, q0) + bq;
          } else {
            dE = 2.f * A1; sE = 2.f * B1; E0 = l1;
          }
          const bool valid = tile * S + i < p.B;
          const float cf = valid ? p.coef : 0.f;
          if (valid) loss_acc += dE * dE * 0.125f - sE + (p.c1 - 2.f) * E0;
          const float gd = cf * dE * 0.25f, gs = -cf, ge = cf * (p.c1 - 2.f);
          // coefficients of w1 / w2 in dL/dh0, dL/da, dL/db (quadratic head) -- see oracle/jet.py quadratic_head_jet
          float G1h, G2h, G1a, G2a, G1b, G2b;
          if (quad) {
            G1h = 2.f * (gd * A2 + gs * B2) + ge * l2; G2h = 2.f * (gd * A1 + gs * B1) + ge * l1;
            G1a = 2.f * (gd * (l2 + B2) + gs * A2);     G2a = 2.f * (gd * (l1 + B1) + gs * A1);
            G1b = 2.f * (gd * A2 + gs * (l2 + B2));     G2b = 2.f * (gd * A1 + gs * (l1 + B1));
          } else {
            G1h = ge; G1a = 2.f * gd; G1b = 2.f * gs; G2h = G2a = G2b = 0.f;
          }
          gb1 += G1h; gb2 += G2h; gbq += ge;
          if (lane == 0) {
            float4* cs = reinterpret_cast<float4*>(coef_s + i * NCOEF);
            cs[0] = make_float4(G1h, G1a, G1b, gd);
            cs[1] = make_float4(G2h, G2a, G2b, gs);
            cs[2] = make_float4(ge, 0.f, 0.f, 0.f);
          }
        }
        __syncwarp();
        // pass 2, block by block: gradient w.r.t. the jets, head-weight gradients, activation backward
        float gw1[MAX_NA], gw2[MAX_NA], gwq[MAX_NA];
#pragma unroll
        for (int j = 0; j < MAX_NA; ++j) {
          gw1[j] = gw2[j] = gwq[j] = 0.f;
          if (j < NA) {
            const int c = lane + 32 * j;
            const float w1 = hw[c], w2 = hw[HWS + c], wq = hw[2 * HWS + c];
#pragma unroll
            for (int k = 0; k < MAX_SPW; ++k) {
              const int i = warp + EPI_WARPS * k;
              if (i < S) {
                const float4* cs = reinterpret_cast<const float4*>(coef_s + i * NCOEF);
                const float4 k1 = cs[0], k2 = cs[1];
                const float G1h = k1.x, G1a = k1.y, G1b = k1.z, gd = k1.w, G2h = k2.x, G2a = k2.y, G2b = k2.z, gs = k2.w;
                const float ge = coef_s[i * NCOEF + 8];
                const Jet o = {A(i, c), A(S + i, c), A(2 * S + i, c)};
                float g0 = G1h * w1, ga = G1a * w1, gb = G1b * w1;
                gw1[j] += G1h * o.h0 + G1a * o.a + G1b * o.b;
                if (quad) {
                  g0 += G2h * w2 + wq * (4.f * gd * o.a + 4.f * gs * o.b + 2.f * ge * o.h0);
                  ga += G2a * w2 + wq * 4.f * (gd * (o.b + o.h0) + gs * o.a);
                  gb += G2b * w2 + wq * 4.f * (gd * o.a + gs * (o.b + o.h0));
                  gw2[j] += G2h * o.h0 + G2a * o.a + G2b * o.b;
                  gwq[j] += 4.f * gd * o.a * (o.b + o.h0) + 2.f * gs * fmaf(o.b, o.b, fmaf(o.a, o.a, 2.f * o.h0 * o.b)) + ge * o.h0 * o.h0;
                }
                float dz0 = 0.f, dP = 0.f, dQ = 0.f;
                if (c < H) jet_bwd<ACT>(o, off, g0, ga, gb, dz0, dP, dQ);
                A(i, c) = tf32r(dz0); A(S + i, c) = tf32r(dP); A(2 * S + i, c) = tf32r(dQ);
              }
            }
            publish(j);
          }
        }
        // tile sums in a fixed order: four rounds of four warps add their partial sums into record (warp & 3)
        if (lane == 0) red[warp] = loss_acc;
        for (int r = 0; r < EPI_WARPS / 4; ++r) {
          if ((warp >> 2) == r) {
            float* rec = hrec + (warp & 3) * HREC;
#pragma unroll
            for (int j = 0; j < MAX_NA; ++j) {
              if (j < NA) {
                const int c = lane + 32 * j;
                if (r == 0) { rec[c] = gw1[j]; rec[HWS + c] = gw2[j]; rec[2 * HWS + c] = gwq[j]; }
                else { rec[c] += gw1[j]; rec[HWS + c] += gw2[j]; rec[2 * HWS + c] += gwq[j]; }
              }
            }
            if (lane == 0) {
              if (r == 0) { rec[3 * HWS] = gb1; rec[3 * HWS + 1] = gb2; rec[3 * HWS + 2] = gbq; }
              else { rec[3 * HWS] += gb1; rec[3 * HWS + 1] += gb2; rec[3 * HWS + 2] += gbq; }
            }
          }
          named_bar_sync(1, EPI_T);
        }
        const int nh = quad ? 3 : 1, head_n = nh * (H + 1);
        for (int o = threadIdx.x; o < head_n; o += EPI_T) {
          const int h = o / (H + 1), c = o % (H + 1);
          const int src = c < H ? h * HWS + c : 3 * HWS + h;
          p.head_part[(long long)tile * head_n + o] = (hrec[src] + hrec[HREC + src]) + (hrec[2 * HREC + src] + hrec[3 * HREC + src]);
        }
        if (threadIdx.x == 0) {
          float s = 0.f;
#pragma unroll
          for (int w = 0; w < EPI_WARPS; ++w) s += red[w];
          p.loss_part[tile] = s;
        }
      } else {
        // -- dgrad epilogue of layer `layer`: rows hold dL/dH_layer as (g0, ga, gb); H_layer comes back from L2 --------------
        const float* hsrc = p.h[layer];
        const long long ldh = p.kpadh, sstride = (long long)p.Bp * ldh;
        for (int j = 0; j < NA; ++j) {
          const int c = lane + 32 * j;
          Jet o[MAX_SPW];
#pragma unroll
          for (int k = 0; k < MAX_SPW; ++k) {
            const int i = warp + EPI_WARPS * k;
            o[k] = Jet{0.f, 0.f, 0.f};
            if (i < S && c < H) {
              const float* hp = hsrc + (long long)(tile * S + i) * ldh + c;
              o[k] = Jet{__ldcg(hp), __ldcg(hp + sstride), __ldcg(hp + 2 * sstride)};
            }
          }
#pragma unroll
          for (int k = 0; k < MAX_SPW; ++k) {
            const int i = warp + EPI_WARPS * k;
            if (i < S) {
              float dz0 = 0.f, dP = 0.f, dQ = 0.f;
              if (c < H) jet_bwd<ACT>(o[k], off, A(i, c), A(S + i, c), A(2 * S + i, c), dz0, dP, dQ);
              A(i, c) = tf32r(dz0); A(S + i, c) = tf32r(dP); A(2 * S + i, c) = tf32r(dQ);
            }
          }
          publish(j);
        }
      }
      stamp(p, 4 + 4 * g);         // this warp's part of the epilogue done
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
  const float* dir; long long lddir; int normalize; float scale;
  unsigned long long seed, offset, row0; const unsigned long long* offset_dev;
  float* xv; int B, Bp, D, kpad0;
};
__global__ void __launch_bounds__(256) jet_inputs_kernel(InputArgs a) {
  constexpr int G = 128, MAXC = 2;
  __shared__ float red[256 / 32];
  if (a.offset_dev) a.offset += __ldg(a.offset_dev);
  const int gid = threadIdx.x / G, t = threadIdx.x % G;
  const int D = a.D, nchunk = (D + 3) >> 2;
  const bool xvec = (D & 3) == 0 && (a.ldx & 3) == 0 && aligned16(a.x);
  const bool dvec = a.dir && (D & 3) == 0 && (a.lddir & 3) == 0 && aligned16(a.dir);
  const long long ngroups = (a.Bp + 1) / 2;
  for (long long rg = blockIdx.x; rg < ngroups; rg += gridDim.x) {
    const long long b = rg * 2 + gid;
    const bool active = b < a.B;
    auto raw = [&](int j) -> float4 {
      const int e = 4 * j;
      if (a.dir) {
        const float* r = a.dir + b * a.lddir;
        if (dvec) return ldg4(r + e);
        return make_float4(e < D ? __ldg(r + e) : 0.f, e + 1 < D ? __ldg(r + e + 1) : 0.f, e + 2 < D ? __ldg(r + e + 2) : 0.f,
                           e + 3 < D ? __ldg(r + e + 3) : 0.f);
      }
      float4 n = philox_normal4(a.seed, a.offset, STREAM_DIRECTION, a.row0 + (unsigned long long)b, (uint32_t)j);
      if (e + 1 >= D) n.y = 0.f;
      if (e + 2 >= D) n.z = 0.f;
      if (e + 3 >= D) n.w = 0.f;
      return n;
    };
    // the draw stays in registers between the norm pass and the write pass when the row fits (D <= 4 MAXC G = 1024);
    // longer rows regenerate it (counter-based: identical numbers)
    float4 dreg[MAXC];
    float ss = 0.f;
    if (active) {
#pragma unroll
      for (int k = 0; k < MAXC; ++k) {
        const int j = t + k * G;
        dreg[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (j < nchunk) {
          dreg[k] = raw(j);
          ss += dreg[k].x * dreg[k].x + dreg[k].y * dreg[k].y + dreg[k].z * dreg[k].z + dreg[k].w * dreg[k].w;
        }
      }
      if (a.normalize)
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
    if (!active) {          // rows of the padded tail: zeros (finite everywhere downstream, masked out of every sum)
      for (int c = t; c < a.kpad0; c += G) { xr[c] = 0.f; vr[c] = 0.f; }
      continue;
    }
    const bool exact = a.dir != nullptr;     // injected draws keep the reference's rounding sequence u / ||u|| * eps
    const float inv = a.normalize && !exact ? __fdiv_rn(a.scale, nrm) : 1.f;
    const float* xrow = a.x + b * a.ldx;
    for (int j = t, k = 0; j < nchunk; j += G, ++k) {
      const int e = 4 * j;
      float4 d, xx;
      if (k == 0) d = dreg[0];
      else if (k == 1) d = dreg[MAXC - 1];
      else d = raw(j);
      if (xvec) xx = ldg4(xrow + e);
      else xx = make_float4(e < D ? __ldg(xrow + e) : 0.f, e + 1 < D ? __ldg(xrow + e + 1) : 0.f, e + 2 < D ? __ldg(xrow + e + 2) : 0.f,
                            e + 3 < D ? __ldg(xrow + e + 3) : 0.f);
      float4 v;
      if (!a.normalize) v = d;
      else if (exact) v = make_float4(__fmul_rn(__fdiv_rn(d.x, nrm), a.scale), __fmul_rn(__fdiv_rn(d.y, nrm), a.scale),
                                      __fmul_rn(__fdiv_rn(d.z, nrm), a.scale), __fmul_rn(__fdiv_rn(d.w, nrm), a.scale));
      else v = make_float4(d.x * inv, d.y * inv, d.z * inv, d.w * inv);
      // kpad0 is a multiple of 32 and the buffer is 16-byte aligned: the chunk store is always in range and aligned.
      // A partial last chunk (D % 4 != 0) carries the first padding columns itself: x gets the ones column at D, v zeros
      if (D - e == 1) xx.y = 1.f;
      if (D - e == 2) xx.z = 1.f;
      if (D - e == 3) xx.w = 1.f;
      *reinterpret_cast<float4*>(xr + e) = make_float4(tf32r(xx.x), tf32r(xx.y), tf32r(xx.z), tf32r(xx.w));
      *reinterpret_cast<float4*>(vr + e) = make_float4(tf32r(v.x), tf32r(v.y), tf32r(v.z), tf32r(v.w));
    }
    for (int c = 4 * nchunk + t; c < a.kpad0; c += G) { xr[c] = c == D ? 1.f : 0.f; vr[c] = 0.f; }
  }
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
