// Kernel (3): the energy-MLP dense layers as tcgen05/TMEM GEMMs fed by TMA (sm_100a).
//
//   D[M x N] = sum_k A[m,k] * B[n,k]        (fp32 in, fp32 accumulate in TMEM)
//
// The reference's nn.Linear is true-fp32 SGEMM (DeepEBM/ESM/models.py:932-955 under torch's
// allow_tf32=False); tcgen05 has no fp32 MMA kind, so operands arrive PRE-SPLIT into TF32 hi/lo
// halves (x = hi + lo, written by the producing kernel's epilogue) and each k-step issues
// hi*hi + hi*lo + lo*hi ("3xTF32", ~fp32 product accuracy).  passes=1 issues hi*hi only.
//
// One 128 x BN output tile per CTA (BN <= 256), optional split-K over gridDim.z.  Warp roles:
// warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM allocator, warps 4..7 = epilogue
// (TMEM -> registers -> fused epilogue -> global).  BK = 32 fp32 = one 128-byte swizzle row.
#pragma once
#include <cuda.h>

#include "fdsm_common.cuh"
#include "tc_ptx.cuh"

namespace fdsm {
namespace tc {

constexpr int BM = 128;            // UMMA M
constexpr int BK = 32;             // fp32 elements per k-block (128 B)
constexpr int UMMA_K = 8;          // tf32: 32 B per instruction
constexpr int A_TILE_BYTES = BM * BK * 4;    // 16 KB
constexpr int B_TILE_BYTES = 256 * BK * 4;   // 32 KB (max BN)
// 16 warps: warp 0 TMA, warp 1 MMA issue, warp 2 TMEM allocation (raw kernel: warps 2-15 convert), ALL warps run the
// epilogue.  The epilogue is ALU-issue bound with dependent chains (activation polynomials): 8 warps left each scheduler
// with 2 warps and half its issue slots empty (4.7 us per 128x208 activation tile); 16 warps = 4 per scheduler.
constexpr int kThreads = 512;
constexpr int EPI_PARTS = kThreads / 128;   // warps per 32-lane TMEM quarter
constexpr int TMEM_COLS = 256;

enum Major : int { MAJOR_K = 0, MAJOR_MN = 1 };
enum Act : int { ACT_NONE = 0, ACT_SOFTPLUS = 1, ACT_NSOFTPLUS = 2, ACT_ELU = 3 };
enum Epi : int {
  EPI_STORE = 0,       // out_hi[z][m][n] = D           (plain fp32; split-K partials indexed by blockIdx.z)
  EPI_ACT_SPLIT = 1,   // h = act(D) -> hi/lo ; col N := 1 (bias column of the next layer) ; cols (N, out_cols) := 0
  EPI_HEAD = 2,        // h = act(D) -> out_hi (fp32) ; E = (h.w1+b1)(h.w2+b2) + h^2.wq + bq -> e_out, l1_out, l2_out
  EPI_DACT_SPLIT = 3,  // g = D * act'(aux) -> hi/lo ; cols [N, out_cols) := 0
};

struct alignas(64) GemmParams {
  CUtensorMap ta_hi, ta_lo, tb_hi, tb_lo;
  int M, N, K;              // logical sizes (rows of D, valid columns of D, contraction length)
  int BN;                   // UMMA N (multiple of 16, <= 256, >= N - n0 for the last tile)
  int a_major, b_major;
  int passes;               // 1 or 3
  int raw_slots, lo_slots, slot_bytes;  // gemm_raw_kernel only: raw / lo ring depths and bytes per slot
  unsigned long long* trace;  // debug: per-CTA phase timestamps (8 per CTA), NULL in production
  int epi, act;
  float* out_hi; float* out_lo; long long ldo; long long split_stride; int out_cols;
  const float* aux_hi; const float* aux_lo; long long ldaux;     // EPI_DACT_SPLIT: H of the previous layer
  const float* head_w;      // EPI_HEAD: [w1(N) | b1 | w2(N) | b2 | wq(N) | bq] = the reference's ln1, ln2, lq parameters
  int head_mode;            // 1: quadratic head ; 0: linear head [w(N) | b]
  float* e_out; float* l1_out; float* l2_out;
  float* head_part;         // EPI_HEAD with more than one N tile (H > 256): partial dots [tile][M][3] (no biases), summed by
                            // mlp::head_energy_kernel; NULL = single tile, the epilogue finishes the energy itself
};

__device__ __forceinline__ float act_fwd(int act, float z) {
  switch (act) {
    case ACT_SOFTPLUS: return z > 20.f ? z : log1pf(expf(z));                       // nn.Softplus(beta=1, threshold=20)
    case ACT_NSOFTPLUS: return (z > 20.f ? z : log1pf(expf(z))) - 0.69314718055994531f;  // models.py:19-28
    case ACT_ELU: return z > 0.f ? z : expm1f(z);
    default: return z;
  }
}
// derivative expressed through the stored activation h = act(z); e^-h on the MUFU pipe (ex2.approx, |error| < 1 ulp of
// the result: the same arithmetic as the GEMM epilogues' act_bwd_t)
__device__ __forceinline__ float act_bwd_from_h(int act, float h) {
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(h * -1.4426950408889634f));
  switch (act) {
    case ACT_SOFTPLUS: return 1.f - e;               // sigmoid(z) = 1 - exp(-softplus(z))
    case ACT_NSOFTPLUS: return fmaf(-0.5f, e, 1.f);  // softplus = h + ln2
    case ACT_ELU: return h > 0.f ? 1.f : h + 1.f;
    default: return 1.f;
  }
}

}  // namespace tc
}  // namespace fdsm
#include "tc_epilogue.cuh"
namespace fdsm {
namespace tc {

struct SmemLayout {
  // dynamic smem (1024-byte aligned): per stage [A_hi | A_lo | B_hi | B_lo]
  static __host__ __device__ constexpr int stage_bytes(int passes) {
    return (passes == 3 ? 2 : 1) * (A_TILE_BYTES + B_TILE_BYTES);
  }
  static __host__ __device__ constexpr int stages(int passes) { return passes == 3 ? 2 : 4; }
  static __host__ __device__ constexpr int total(int passes) { return stages(passes) * stage_bytes(passes) + 1024 + 256; }
};

template <int PASSES, int EPI>
__global__ void __launch_bounds__(kThreads, 1) gemm_kernel(const __grid_constant__ GemmParams p) {
  using namespace ptx;
  // EPI_HEAD: head weights [w1 | b1], [w2 | b2], [wq | bq] staged once in shared memory (N <= 256)
  __shared__ __align__(16) float head_smem[EPI == EPI_HEAD ? 3 * HW_STRIDE : 4];
  if constexpr (EPI == EPI_HEAD) {
    const int nh = p.head_mode ? 3 : 1;
    // weights [0, N), zeros up to HW_BIAS, the bias at HW_BIAS: columns beyond N then contribute nothing without a select
    for (int i = threadIdx.x; i < nh * HW_STRIDE; i += kThreads) {
      const int h = i / HW_STRIDE, j = i % HW_STRIDE;
      const int col = blockIdx.y * p.BN + j;   // this tile's columns [n0, n0 + BN)
      head_smem[i] = (j < p.BN && col < p.N) ? __ldg(p.head_w + h * (p.N + 1) + col)
                                             : (j == HW_BIAS ? __ldg(p.head_w + h * (p.N + 1) + p.N) : 0.f);
    }
  }
  constexpr int STAGES = SmemLayout::stages(PASSES);
  constexpr int STAGE_BYTES = SmemLayout::stage_bytes(PASSES);
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - raw);
  const uint32_t bars = base + STAGES * STAGE_BYTES;              // full[STAGES], empty[STAGES], tmem_full, tmem_ptr
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(gen + STAGES * STAGE_BYTES + 8 * (2 * STAGES + 1));
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
  const uint32_t tmem_full_bar = bars + 8u * (2 * STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * p.BN;
  const int kblocks = (p.K + BK - 1) / BK;
  const int kb0 = (int)((long long)kblocks * blockIdx.z / gridDim.z);
  const int kb1 = (int)((long long)kblocks * (blockIdx.z + 1) / gridDim.z);

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&p.ta_hi);
    prefetch_tensormap(&p.tb_hi);
    if (PASSES == 3) {
      prefetch_tensormap(&p.ta_lo);
      prefetch_tensormap(&p.tb_lo);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(smem_u32((const void*)tmem_ptr_smem), TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===== TMA producer ===========================================================================
    if (lane == 0) {
      const int b_boxes_mn = (p.BN + 31) / 32;
      const uint32_t tx = (uint32_t)(PASSES == 3 ? 2 : 1) *
                          (uint32_t)(A_TILE_BYTES + (p.b_major == MAJOR_K ? p.BN * BK * 4 : b_boxes_mn * 32 * BK * 4));
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t sa = base + stage * STAGE_BYTES;
        const uint32_t fb = full_bar(stage);
        mbar_arrive_expect_tx(fb, tx);
        const int k0 = kb * BK;
#pragma unroll
        for (int h = 0; h < (PASSES == 3 ? 2 : 1); ++h) {
          const void* ta = h ? (const void*)&p.ta_lo : (const void*)&p.ta_hi;
          const void* tb = h ? (const void*)&p.tb_lo : (const void*)&p.tb_hi;
          const uint32_t a_dst = sa + h * A_TILE_BYTES;
          const uint32_t b_dst = sa + (PASSES == 3 ? 2 : 1) * A_TILE_BYTES + h * B_TILE_BYTES;
          if (p.a_major == MAJOR_K) {
            tma_load_2d(a_dst, ta, fb, k0, m0);                     // box {32 k, 128 rows}
          } else {
            for (int i = 0; i < BM / 32; ++i) tma_load_2d(a_dst + i * 4096, ta, fb, m0 + 32 * i, k0);  // box {32 m, 32 k}
          }
          if (p.b_major == MAJOR_K) {
            tma_load_2d(b_dst, tb, fb, k0, n0);                     // box {32 k, BN rows}
          } else {
            for (int i = 0; i < b_boxes_mn; ++i) tma_load_2d(b_dst + i * 4096, tb, fb, n0 + 32 * i, k0);
          }
        }
        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) ================================================================
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(BM, p.BN, p.a_major, p.b_major);
      // per-UMMA_K advance of the descriptor start address: 32 B inside the swizzle row (K-major) or
      // one 8-row k-group = 1024 B (MN-major)
      const uint32_t a_lbo = p.a_major == MAJOR_K ? 16u : 4096u, b_lbo = p.b_major == MAJOR_K ? 16u : 4096u;
      const uint32_t a_step = p.a_major == MAJOR_K ? 32u : 1024u, b_step = p.b_major == MAJOR_K ? 32u : 1024u;
      const uint32_t a_sbo = p.a_major == MAJOR_K ? 1024u : 512u, b_sbo = p.b_major == MAJOR_K ? 1024u : 512u;
      const uint32_t a_lay = p.a_major == MAJOR_K ? 2u : 1u, b_lay = p.b_major == MAJOR_K ? 2u : 1u;
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        const uint32_t sa = base + stage * STAGE_BYTES;
        const uint32_t a_hi = sa, a_lo = sa + A_TILE_BYTES;
        const uint32_t b_hi = sa + (PASSES == 3 ? 2 : 1) * A_TILE_BYTES, b_lo = b_hi + B_TILE_BYTES;
#pragma unroll
        for (int k8 = 0; k8 < BK / UMMA_K; ++k8) {
          const uint64_t dah = make_smem_desc(a_hi + k8 * a_step, a_lbo, a_sbo, a_lay);
          const uint64_t dbh = make_smem_desc(b_hi + k8 * b_step, b_lbo, b_sbo, b_lay);
          mma_tf32_ss(tmem_base, dah, dbh, idesc, (kb > kb0 || k8 > 0) ? 1u : 0u);
          if (PASSES == 3) {
            const uint64_t dal = make_smem_desc(a_lo + k8 * a_step, a_lbo, a_sbo, a_lay);
            const uint64_t dbl = make_smem_desc(b_lo + k8 * b_step, b_lbo, b_sbo, b_lay);
            mma_tf32_ss(tmem_base, dah, dbl, idesc, 1u);
            mma_tf32_ss(tmem_base, dal, dbh, idesc, 1u);
          }
        }
        tc_commit(empty_bar(stage));                    // smem slot free once these MMAs have read it
        if (kb == kb1 - 1) tc_commit(tmem_full_bar);    // accumulator complete
        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  }
  __syncwarp();
  {
    // ===== epilogue (all 8 warps once their role is done): TMEM -> registers -> fused op -> global ==========
    // warp w may read TMEM lanes 32*(w%4)..+31; the EPI_PARTS warps of a quarter split the columns
    const int q = warp & 3;
    const int part = warp >> 2;   // EPI_PARTS warps share each 32-lane quarter: each takes a contiguous range of column steps
    const int row = m0 + q * 32 + lane;
    const bool row_ok = row < p.M;
    const bool have_acc = kb1 > kb0;
    if (have_acc) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    float l1 = 0.f, l2 = 0.f, lq = 0.f;
    const bool last_tile = n0 + p.BN >= p.N;   // the tile that also writes the padding columns [N, out_cols)
    const int ncols_out = (EPI == EPI_STORE || EPI == EPI_HEAD) ? min(p.N - n0, p.BN) : (last_tile ? p.out_cols - n0 : p.BN);
    const int nsteps = (ncols_out + CW - 1) / CW, sper = (nsteps + EPI_PARTS - 1) / EPI_PARTS;
    const int c_begin = min(part * sper, nsteps) * CW, c_end = min((part + 1) * sper, nsteps) * CW;
    epilogue_dispatch<EPI>(p, tmem_base + ((uint32_t)(q * 32) << 16), have_acc, row, row_ok, n0, c_begin, c_end, head_smem, l1,
                           l2, lq);
    if constexpr (EPI == EPI_HEAD) {   // combine the two column halves of each row
      __shared__ float hsum[EPI_PARTS - 1][BM][3];
      if (part > 0) { hsum[part - 1][q * 32 + lane][0] = l1; hsum[part - 1][q * 32 + lane][1] = l2; hsum[part - 1][q * 32 + lane][2] = lq; }
      __syncthreads();
      if (part == 0 && row_ok) {
#pragma unroll
        for (int h = 0; h < EPI_PARTS - 1; ++h) {   // fixed order: deterministic
          l1 += hsum[h][q * 32 + lane][0]; l2 += hsum[h][q * 32 + lane][1]; lq += hsum[h][q * 32 + lane][2];
        }
        if (p.head_part) {   // several N tiles: leave the partial dots, mlp::head_energy_kernel adds biases and combines
          float* hp = p.head_part + ((long long)blockIdx.y * p.M + row) * 3;
          hp[0] = l1; hp[1] = l2; hp[2] = lq;
          l1 = l2 = lq = 0.f;
        }
        l1 += head_smem[HW_BIAS];
        l2 += p.head_mode ? head_smem[HW_STRIDE + HW_BIAS] : 0.f;
        lq += p.head_mode ? head_smem[2 * HW_STRIDE + HW_BIAS] : 0.f;
        // quadratic head ln1(h)*ln2(h) + lq(h^2) (models.py:338,954) or a plain Linear(H,1) head (head_mode 0)
        p.e_out[row] = p.head_mode ? fmaf(l1, l2, lq) : l1;
        p.l1_out[row] = l1;
        p.l2_out[row] = l2;
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}


// ---- weight-gradient GEMM, wide stages -----------------------------------------------------------------------------------
// dWt[m][n] = sum_r In[r][m] dZ[r][n] over the batch rows r: both operands MN-major (the contraction index is the row of the
// stored matrices), one pass of plain TF32, split-K partials (EPI_STORE).  What an SM can pull through TMA is set by the
// number of instructions / barrier phases (~650 clk per pipeline stage + ~70 per box, whatever their size:
// scripts/tma_ingest_bench.cu), so gemm_kernel's stage of 11 boxes of 32 x 32 (one 32-row k-block, 43 KB) costs as much as
// this kernel's stage of TWO 3-D boxes covering 64 rows (86 KB): {32 m, 64 r, 4 panels} of In and {32 n, 64 r, BN/32 panels}
// of dZ, landing as [panel][64 r][32] -- the k-blocks of a stage are 4 KB apart inside each 8 KB panel.
constexpr int WIDE_BK = 64;
constexpr int WIDE_A_BYTES = BM * WIDE_BK * 4;       // 32 KB
constexpr int WIDE_B_BYTES = 256 * WIDE_BK * 4;      // 64 KB (max BN)
constexpr int WIDE_STAGES = 2;
constexpr int WIDE_SMEM = WIDE_STAGES * (WIDE_A_BYTES + WIDE_B_BYTES) + 1024 + 256;

// CS > 1: the CTAs of a cluster (CS consecutive splits of one output tile, cluster dimension z) add their accumulators
// through distributed shared memory before anything goes to HBM -- every CTA stages its 128 x BN tile in its own shared
// memory (the pipeline's buffers, free by then), and CTA r of the cluster sums rows [128 r / CS, 128 (r + 1) / CS) over the
// CS tiles in rank order (fixed order: deterministic) and writes them as partial blockIdx.z / CS: 1 / CS of the split-K
// partial traffic, written as coalesced 16-byte row segments, and 1 / CS of what the reduce / update kernel reads back.
// one CTA's job: output tile `m_tile` of dWt, split `split` of `nsplits` along the contraction
struct WideJob {
  const CUtensorMap* ta; const CUtensorMap* tb;      // (param space)
  int M, N, K, BN;
  float* out; long long ldo, split_stride;
  unsigned long long* trace; int cta;
  int m_tile, split, nsplits;
};
template <int CS>
__device__ __forceinline__ void gemm_wide_body(const WideJob& p) {
  using namespace ptx;
  constexpr int STAGE_BYTES = WIDE_A_BYTES + WIDE_B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - raw);
  const uint32_t bars = base + WIDE_STAGES * STAGE_BYTES;
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(gen + WIDE_STAGES * STAGE_BYTES + 8 * (2 * WIDE_STAGES + 1));
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (WIDE_STAGES + s); };
  const uint32_t tmem_full_bar = bars + 8u * (2 * WIDE_STAGES);

  auto stamp = [&](int slot) {      // debug (scripts/wgrad_trace.py): %globaltimer stamps, 8 per CTA
    if (p.trace) {
      const int cta = p.cta;
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      if (cta < 160) p.trace[cta * 8 + slot] = t;
    }
  };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = p.m_tile * BM, n0 = 0;
  griddep_launch_dependents();      // the kernel behind this one (launched programmatically) may become resident and wait
  if (threadIdx.x == 0) stamp(0);
  const int kblocks = (p.K + WIDE_BK - 1) / WIDE_BK;          // 64-row blocks of the contraction
  const int kb0 = (int)((long long)kblocks * p.split / p.nsplits);
  const int kb1 = (int)((long long)kblocks * (p.split + 1) / p.nsplits);
  const int nbp = (p.BN + 31) / 32;                           // 32-column panels of dZ per stage

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(p.ta);
    prefetch_tensormap(p.tb);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < WIDE_STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(smem_u32((const void*)tmem_ptr_smem), TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  griddep_wait();      // (no-op unless launched programmatically: everything above overlaps the previous kernel's drain)
  if (threadIdx.x == 0) stamp(1);

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t tx = (uint32_t)(WIDE_A_BYTES + nbp * 32 * WIDE_BK * 4);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        const uint32_t sa = base + stage * STAGE_BYTES;
        mbar_arrive_expect_tx(full_bar(stage), tx);
        tma_load_3d(sa, p.ta, full_bar(stage), 0, kb * WIDE_BK, m0 >> 5);
        tma_load_3d(sa + WIDE_A_BYTES, p.tb, full_bar(stage), 0, kb * WIDE_BK, n0 >> 5);
        if (++stage == WIDE_STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(BM, p.BN, MAJOR_MN, MAJOR_MN);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        if (kb == kb0) stamp(2);
        if (kb == kb1 - 1) stamp(3);
        const uint32_t sa = base + stage * STAGE_BYTES, sb = sa + WIDE_A_BYTES;
        // rows beyond K (a ragged last block) arrive as zeros: issuing their MMAs is harmless
#pragma unroll
        for (int k8 = 0; k8 < WIDE_BK / UMMA_K; ++k8) {
          // MN-major, 128-byte swizzle with 32-byte atoms: 8 rows of the contraction = 1 KB; panels 8 KB apart
          const uint64_t da = make_smem_desc(sa + k8 * 1024u, 8192u, 512u, 1u);
          const uint64_t db = make_smem_desc(sb + k8 * 1024u, 8192u, 512u, 1u);
          mma_tf32_ss(tmem_base, da, db, idesc, (kb > kb0 || k8 > 0) ? 1u : 0u);
        }
        tc_commit(empty_bar(stage));
        if (kb == kb1 - 1) tc_commit(tmem_full_bar);
        if (++stage == WIDE_STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  }
  __syncwarp();
  {
    const int q = warp & 3, part = warp >> 2;
    const int row = m0 + q * 32 + lane;
    const bool row_ok = row < p.M;
    const bool have_acc = kb1 > kb0;
    if (have_acc) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    if (threadIdx.x == 128) stamp(4);
    const int ncols_out = min(p.N - n0, p.BN);
    const int nsteps = (ncols_out + CW - 1) / CW, sper = (nsteps + EPI_PARTS - 1) / EPI_PARTS;
    const int c_begin = min(part * sper, nsteps) * CW, c_end = min((part + 1) * sper, nsteps) * CW;
    (void)row; (void)row_ok;
    // stage: accumulator row (q, lane) -> shared [128][ldt] (ldt = BN + 4: the 16-byte row pieces of a warp's 32 rows fall
    // on all eight bank groups); un-staged, every warp-wide 16-byte store touches 32 different rows and the epilogue of a
    // 128 x 208 tile takes 4.5 us (profiles/r02_wgrad_trace_v1.txt)
    const int ldt = p.BN + 4;
    float* stage = reinterpret_cast<float*>(gen);
    {
      const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
      float* srow = stage + (q * 32 + lane) * ldt;
      for (int c = c_begin; c < c_end; c += CW) {
        float v[8];
        if (have_acc) {
          tmem_ld_32x8(trow + (uint32_t)c, v);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = 0.f;
        }
        *reinterpret_cast<float4*>(srow + c) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(srow + c + 4) = make_float4(v[4], v[5], v[6], v[7]);
      }
    }
    if constexpr (CS == 1) {
      // coalesced row copies shared -> global (one asynchronous bulk copy per row when the rows are 16-byte aligned)
      const bool vec = ((p.ldo | n0) & 3) == 0 && aligned16(p.out) && (p.split_stride & 3) == 0;
      __syncthreads();       // the tile is complete (every thread of the CTA runs this epilogue)
      stage_store_tile(stage, ldt, p.out + (long long)p.split * p.split_stride + (long long)m0 * p.ldo + n0, p.ldo,
                       min(BM, p.M - m0), ncols_out, vec);
    } else {
      tc_fence_before();
      cluster_sync_all();                       // every tile of the cluster is staged (release / acquire at cluster scope)
      const uint32_t crank = cluster_ctarank();
      constexpr int RPC = BM / CS;              // rows this CTA reduces
      const int nc4 = (ncols_out + 3) >> 2;
      const uint32_t stage_s = smem_u32(stage);
      uint32_t peer[CS];
#pragma unroll
      for (int c = 0; c < CS; ++c)
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(peer[c]) : "r"(stage_s), "r"(c));
      float* outp = p.out + (long long)(p.split / CS) * p.split_stride;
      for (int i = threadIdx.x; i < RPC * nc4; i += kThreads) {
        const int r = crank * RPC + i / nc4, c4 = i % nc4;
        if (m0 + r >= p.M) continue;
        const uint32_t off = (uint32_t)(r * ldt + 4 * c4) * 4u;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int c = 0; c < CS; ++c) {
          float4 v;
          asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(peer[c] + off));
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        float* o = outp + (long long)(m0 + r) * p.ldo + n0 + 4 * c4;
        if (4 * c4 + 3 < ncols_out && aligned16(o)) {
          *reinterpret_cast<float4*>(o) = acc;
        } else {
          if (4 * c4 < ncols_out) o[0] = acc.x;
          if (4 * c4 + 1 < ncols_out) o[1] = acc.y;
          if (4 * c4 + 2 < ncols_out) o[2] = acc.z;
          if (4 * c4 + 3 < ncols_out) o[3] = acc.w;
        }
      }
      cluster_sync_all();                       // no CTA leaves while a peer still reads its tile
    }
    if (threadIdx.x == 128) stamp(5);
    tc_fence_before();
  }
  __syncthreads();
  if (threadIdx.x == 0) stamp(6);
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int CS>
__global__ void __launch_bounds__(kThreads, 1) gemm_wide_kernel(const __grid_constant__ GemmParams p) {
  WideJob j{&p.ta_hi, &p.tb_hi, p.M, p.N, p.K, p.BN, p.out_hi, p.ldo, p.split_stride, p.trace,
            (int)(blockIdx.x + gridDim.x * blockIdx.z), (int)blockIdx.x, (int)blockIdx.z, (int)gridDim.z};
  gemm_wide_body<CS>(j);
}

// Several weight-gradient GEMMs (the layers of one backward pass) as ONE launch: CTA b serves job l with
// cta0[l] <= b < cta0[l + 1]; inside a job the M tiles vary fastest.  One graph node instead of one per layer on as many
// streams (no fork / join around them, and a single stream predecessor for the kernel that follows).
constexpr int WIDE_MAX_JOBS = 5;
struct alignas(64) WideBatchParams {
  CUtensorMap ta[WIDE_MAX_JOBS], tb[WIDE_MAX_JOBS];
  int M[WIDE_MAX_JOBS], K[WIDE_MAX_JOBS], mt[WIDE_MAX_JOBS], splits[WIDE_MAX_JOBS], cta0[WIDE_MAX_JOBS + 1];
  float* out[WIDE_MAX_JOBS]; long long split_stride[WIDE_MAX_JOBS];
  int N, BN, njobs; long long ldo;
  unsigned long long* trace;
};
template <int VARIANT>      // (a template only so that the header may be included by several translation units)
__global__ void __launch_bounds__(kThreads, 1) gemm_wide_batch_kernel(const __grid_constant__ WideBatchParams p) {
  int l = 0;
  while (l + 1 < p.njobs && (int)blockIdx.x >= p.cta0[l + 1]) ++l;
  const int local = (int)blockIdx.x - p.cta0[l];
  WideJob j{&p.ta[l], &p.tb[l], p.M[l], p.N, p.K[l], p.BN, p.out[l], p.ldo, p.split_stride[l], p.trace,
            (int)blockIdx.x, local % p.mt[l], local / p.mt[l], p.splits[l]};
  gemm_wide_body<1>(j);
}

}  // namespace tc
}  // namespace fdsm
