// Kernel (3), raw-operand flavour: the same 3xTF32 tcgen05 GEMM as tc_gemm.cuh, but the operands arrive as
// PLAIN fp32 and the TF32 hi/lo split happens inside the kernel, in shared memory.
//
// Why: an SM ingests ~43-46 B/clk from L2 (B300_MICROARCH.md, "LTS throughput cap" / 148), and the pre-split
// kernel moves hi AND lo of both operands through that port: 84 KB per k-block against 1248 clk of MMA work
// (12 x 128x208x8 kind::tf32) -> load-bound at ~65 GB/s per SM (measured, profiles/r01_gemm_sweep.txt).  Loading
// the fp32 value once halves the ingest (42 KB per k-block < the MMA time) and halves the HBM traffic of every
// producer (stack kernel, activation epilogues).
//
// Pipeline (per k-block i; r = i % R raw slots, l = i % 2 lo slots):
//   warp 0   TMA: waits raw_empty[r], loads raw A|B tiles into raw slot r (raw_full[r], expect-tx)
//   warps 2-7 converters: wait raw_full[r] and lo_empty[l]; the raw fp32 tile IS the hi operand (kind::tf32 ignores the
//            13 low mantissa bits: hi = trunc_tf32(x), nothing to write back); lo = rna_tf32(x - trunc_tf32(x)) into lo
//            slot l (elementwise, so the TMA swizzle is irrelevant); fence.proxy.async; arrive conv[l]
//            (-DFDSM_RAW_REWRITE_HI=1: hi = rna_tf32(x) rewritten in place -- bit-identical to the pre-split kernel but
//            42 KB more shared-memory traffic per k-block; the main loop is shared-memory-port bound: 297 -> 255 KB
//            per k-block, measured 28.6 -> 24.1 us on the layer-1 forward GEMM)
//   warp 1   MMA: waits conv[l]; hi*hi + hi*lo + lo*hi per UMMA_K; tcgen05.commit -> raw_empty[r], lo_empty[l]
// Same MMA order as the pre-split kernel; the split differs in the last bit of hi (truncation instead of round-to-
// nearest), so results agree with it to fp32 rounding, not bit for bit (tests/test_tc_gemm_gpu.py).
#pragma once
#include "tc_gemm.cuh"

namespace fdsm {
namespace tc {

constexpr int MAX_LO_SLOTS = 4;   // lo ring depth is a launch parameter (GemmParams::lo_slots), barriers are laid out for the maximum
#ifndef FDSM_A_COLLECTOR
#define FDSM_A_COLLECTOR 1
#endif
#ifndef FDSM_RAW_REWRITE_HI
#define FDSM_RAW_REWRITE_HI 0
#endif
constexpr int CONV_WARP0 = 2;                       // warps [2, 8) convert
constexpr int CONV_WARPS = kThreads / 32 - CONV_WARP0;
constexpr int MAX_RAW_SLOTS = 4;

__host__ __device__ inline int raw_b_tile_bytes(int BN, int b_major) {
  return b_major == MAJOR_K ? BN * BK * 4 : ((BN + 31) / 32) * 32 * BK * 4;
}
__host__ __device__ inline int raw_slot_bytes(int BN, int b_major) {
  return A_TILE_BYTES + ((raw_b_tile_bytes(BN, b_major) + 1023) & ~1023);
}
__host__ inline int raw_smem_total(int raw_slots, int lo_slots, int slot_bytes) { return (raw_slots + lo_slots) * slot_bytes + 1024 + 256; }

__device__ __forceinline__ void trace_stamp(const GemmParams& p, int slot) {
  if (p.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    const unsigned cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    if (cta < 160) p.trace[cta * 8 + slot] = t;
  }
}

// PAIR = true: launched as clusters of 2 CTAs along x (cta_group::2).  The pair computes a 256 x BN tile with ONE MMA stream
// issued by the leader (cluster rank 0): each CTA stages its own 128 rows of A and HALF of the B tile (BN/2 columns) and
// converts them locally; the tensor cores of both SMs read both halves of B.  Per SM and k-block the shared-memory port
// moves 16+13 KB (TMA) + 29 (converter reads) + 29 (lo writes) + 12 x 7.3 KB (MMA operand reads) = 175 KB instead of
// 255 KB: the main loop is bound by that port.
template <int EPI, bool PAIR>
__global__ void __launch_bounds__(kThreads, 1) gemm_raw_kernel(const __grid_constant__ GemmParams p) {
  using namespace ptx;
  const uint32_t crank = PAIR ? cluster_ctarank() : 0u;     // 0 = leader
  const int BNL = PAIR ? p.BN / 2 : p.BN;                   // B-tile columns staged by THIS CTA
  __shared__ __align__(16) float head_smem[EPI == EPI_HEAD ? 3 * HW_STRIDE : 4];
  if (threadIdx.x == 0) trace_stamp(p, 0);   // CTA start
  const int R = p.raw_slots;
  const int SLOT = p.slot_bytes;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - raw);
  const uint32_t lo_base = base + R * SLOT;
  const int LO_SLOTS = p.lo_slots;
  const uint32_t bars = base + (R + LO_SLOTS) * SLOT;
  // barriers: raw_full[4] raw_empty[4] conv[2] lo_empty[2] tmem_full ; then the TMEM base pointer
  auto raw_full = [&](int s) { return bars + 8u * s; };
  auto raw_empty = [&](int s) { return bars + 8u * (MAX_RAW_SLOTS + s); };
  auto conv_bar = [&](int s) { return bars + 8u * (2 * MAX_RAW_SLOTS + s); };
  auto lo_empty = [&](int s) { return bars + 8u * (2 * MAX_RAW_SLOTS + MAX_LO_SLOTS + s); };
  const uint32_t tmem_full_bar = bars + 8u * (2 * MAX_RAW_SLOTS + 2 * MAX_LO_SLOTS);
  volatile uint32_t* tmem_ptr_smem =
      reinterpret_cast<volatile uint32_t*>(gen + (R + LO_SLOTS) * SLOT + 8 * (2 * MAX_RAW_SLOTS + 2 * MAX_LO_SLOTS + 1));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * p.BN;
  const int kblocks = (p.K + BK - 1) / BK;
  const int kb0 = (int)((long long)kblocks * blockIdx.z / gridDim.z);
  const int kb1 = (int)((long long)kblocks * (blockIdx.z + 1) / gridDim.z);
  const int b_bytes = raw_b_tile_bytes(BNL, p.b_major);
  const int nb0 = n0 + (int)crank * BNL;                    // first B column staged by this CTA

  // TMA producer state; the first R k-blocks are requested BEFORE the CTA-wide setup barrier (TMEM allocation takes
  // ~0.5 us and the first load ~2 us: they overlap instead of adding up).
  const int b_boxes_mn = (BNL + 31) / 32;
  auto issue_loads = [&](int kb, int r) {
    const uint32_t a_dst = base + r * SLOT, b_dst = a_dst + A_TILE_BYTES;
    const uint32_t fb = raw_full(r);
    mbar_arrive_expect_tx(fb, (uint32_t)(A_TILE_BYTES + b_bytes));
    const int k0 = kb * BK;
    if (p.a_major == MAJOR_K) {
      tma_load_2d(a_dst, &p.ta_hi, fb, k0, m0);
    } else {
      for (int i = 0; i < BM / 32; ++i) tma_load_2d(a_dst + i * 4096, &p.ta_hi, fb, m0 + 32 * i, k0);
    }
    if (p.b_major == MAJOR_K) {
      tma_load_2d(b_dst, &p.tb_hi, fb, k0, nb0);
    } else {
      for (int i = 0; i < b_boxes_mn; ++i) tma_load_2d(b_dst + i * 4096, &p.tb_hi, fb, nb0 + 32 * i, k0);
    }
  };
  const int pre = min(R, kb1 - kb0);
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < MAX_RAW_SLOTS; ++s) {
      mbar_init(raw_full(s), 1);
      mbar_init(raw_empty(s), 1);
    }
    for (int s = 0; s < MAX_LO_SLOTS; ++s) {
      mbar_init(conv_bar(s), PAIR ? 2 * CONV_WARPS : CONV_WARPS);   // pair: the peer's converters arrive on the leader's
      mbar_init(lo_empty(s), 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    if constexpr (PAIR) tmem_alloc_pair(smem_u32((const void*)tmem_ptr_smem), TMEM_COLS);
    else tmem_alloc(smem_u32((const void*)tmem_ptr_smem), TMEM_COLS);
  }
  // Everything above is independent of the previous kernel's output and overlaps its tail when this grid was launched
  // with programmatic stream serialization; from here on its global-memory writes are visible.
  griddep_wait();
  if (warp == 0 && lane == 0)
    for (int i = 0; i < pre; ++i) issue_loads(kb0 + i, i);   // first loads in flight before the CTA-wide barrier
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
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all();   // both CTAs' barriers are initialised before any remote arrive / multicast commit
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0) trace_stamp(p, 1);   // setup done

  if (warp == 0) {
    // ===== TMA producer ===========================================================================
    if (lane == 0) {
      int r = pre % R;
      uint32_t phase = pre == R ? 1u : 0u;
      for (int kb = kb0 + pre; kb < kb1; ++kb) {
        mbar_wait(raw_empty(r), phase ^ 1u);
        issue_loads(kb, r);
        if (++r == R) { r = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread; pair: of the leader CTA only) ===================================
    if (lane == 0 && crank == 0) {
      const uint32_t idesc = make_idesc_tf32(PAIR ? 2 * BM : BM, p.BN, p.a_major, p.b_major);
      const uint32_t a_lbo = p.a_major == MAJOR_K ? 16u : 4096u, b_lbo = p.b_major == MAJOR_K ? 16u : 4096u;
      const uint32_t a_step = p.a_major == MAJOR_K ? 32u : 1024u, b_step = p.b_major == MAJOR_K ? 32u : 1024u;
      const uint32_t a_sbo = p.a_major == MAJOR_K ? 1024u : 512u, b_sbo = p.b_major == MAJOR_K ? 1024u : 512u;
      const uint32_t a_lay = p.a_major == MAJOR_K ? 2u : 1u, b_lay = p.b_major == MAJOR_K ? 2u : 1u;
      int r = 0, l = 0;
      uint32_t lphase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        if constexpr (PAIR) mbar_wait_cluster(conv_bar(l), lphase);
        else mbar_wait(conv_bar(l), lphase);
        tc_fence_after();
        if (kb == kb0) trace_stamp(p, 2);      // first k-block converted: MMA starts
        if (kb == kb1 - 1) trace_stamp(p, 3);  // last k-block converted
        const uint32_t a_hi = base + r * SLOT, b_hi = a_hi + A_TILE_BYTES;
        const uint32_t a_lo = lo_base + l * SLOT, b_lo = a_lo + A_TILE_BYTES;
#pragma unroll
        for (int k8 = 0; k8 < BK / UMMA_K; ++k8) {
          const uint64_t dah = make_smem_desc(a_hi + k8 * a_step, a_lbo, a_sbo, a_lay);
          const uint64_t dbh = make_smem_desc(b_hi + k8 * b_step, b_lbo, b_sbo, b_lay);
          const uint64_t dal = make_smem_desc(a_lo + k8 * a_step, a_lbo, a_sbo, a_lay);
          const uint64_t dbl = make_smem_desc(b_lo + k8 * b_step, b_lbo, b_sbo, b_lay);
          if constexpr (PAIR) {
            mma_tf32_ss_pair(tmem_base, dah, dbh, idesc, (kb > kb0 || k8 > 0) ? 1u : 0u);
            mma_tf32_ss_pair(tmem_base, dah, dbl, idesc, 1u);
            mma_tf32_ss_pair(tmem_base, dal, dbh, idesc, 1u);
          } else {
#if FDSM_A_COLLECTOR
            mma_tf32_ss_afill(tmem_base, dah, dbh, idesc, (kb > kb0 || k8 > 0) ? 1u : 0u);
            mma_tf32_ss_alastuse(tmem_base, dah, dbl, idesc, 1u);
#else
            mma_tf32_ss(tmem_base, dah, dbh, idesc, (kb > kb0 || k8 > 0) ? 1u : 0u);
            mma_tf32_ss(tmem_base, dah, dbl, idesc, 1u);
#endif
            mma_tf32_ss(tmem_base, dal, dbh, idesc, 1u);
          }
        }
        if constexpr (PAIR) {   // free the stage / publish the accumulator in BOTH CTAs
          tc_commit_pair(raw_empty(r));
          tc_commit_pair(lo_empty(l));
          if (kb == kb1 - 1) tc_commit_pair(tmem_full_bar);
        } else {
          tc_commit(raw_empty(r));
          tc_commit(lo_empty(l));
          if (kb == kb1 - 1) tc_commit(tmem_full_bar);
        }
        if (++r == R) r = 0;
        if (++l == LO_SLOTS) { l = 0; lphase ^= 1u; }
      }
    }
  } else {
    // ===== converters (warps 2..7): raw fp32 -> TF32 hi (in place) + lo ==================================
    const int ct = threadIdx.x - CONV_WARP0 * 32;
    const int nvec = (A_TILE_BYTES + b_bytes) >> 4;   // A and B tiles are adjacent inside a slot
    int r = 0, l = 0;
    uint32_t rphase = 0, lphase = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(raw_full(r), rphase);
      mbar_wait(lo_empty(l), lphase ^ 1u);
      float4* src = reinterpret_cast<float4*>(gen + r * SLOT);
      float4* dst = reinterpret_cast<float4*>(gen + (R + l) * SLOT);
      constexpr int CT = CONV_WARPS * 32, UNR = 4;
      for (int i0 = ct; i0 < nvec; i0 += CT * UNR) {
        float4 x[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u)
          if (i0 + u * CT < nvec) x[u] = src[i0 + u * CT];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          if (i0 + u * CT < nvec) {
            float4 hi, lo;
#if FDSM_RAW_REWRITE_HI
            tf32_split_alu(x[u].x, hi.x, lo.x); tf32_split_alu(x[u].y, hi.y, lo.y);
            tf32_split_alu(x[u].z, hi.z, lo.z); tf32_split_alu(x[u].w, hi.w, lo.w);
            src[i0 + u * CT] = hi;
#else
            // the tensor core ignores the 13 low mantissa bits of a kind::tf32 operand (truncation): the raw value
            // already IS the hi half; lo = rna_tf32(x - trunc(x)), no write-back
            tf32_split_trunc(x[u].x, hi.x, lo.x); tf32_split_trunc(x[u].y, hi.y, lo.y);
            tf32_split_trunc(x[u].z, hi.z, lo.z); tf32_split_trunc(x[u].w, hi.w, lo.w);
#endif
            dst[i0 + u * CT] = lo;
          }
        }
      }
      fence_proxy_async();   // generic-proxy stores -> visible to the tensor core's async-proxy reads
      __syncwarp();
      if (lane == 0) {
        if constexpr (PAIR) mbar_arrive_remote(conv_bar(l), 0u);   // the leader's barrier (also from the leader itself)
        else mbar_arrive(conv_bar(l));
      }
      if (++r == R) { r = 0; rphase ^= 1u; }
      if (++l == LO_SLOTS) { l = 0; lphase ^= 1u; }
    }
  }
  __syncwarp();
  {
    // ===== epilogue: TMEM -> registers -> fused op -> shared staging tile -> coalesced global stores ==========
    // warp w may read TMEM lanes 32*(w%4)..+31; the EPI_PARTS warps of a quarter split the columns.
    const int q = warp & 3;
    const int part = warp >> 2;   // EPI_PARTS warps share each 32-lane quarter: each takes a contiguous range of column steps
    const int rloc = q * 32 + lane;
    const int row = m0 + rloc;
    const bool row_ok = row < p.M;
    const bool have_acc = kb1 > kb0;
    if (have_acc) {
      mbar_wait(tmem_full_bar, 0);   // every MMA has completed: the pipeline's shared memory is free from here on
      tc_fence_after();
    }
    if (threadIdx.x == 128) trace_stamp(p, 4);   // accumulator complete
    griddep_launch_dependents();   // main loop done: let the next kernel's CTAs take the idle SMs and set up
    float l1 = 0.f, l2 = 0.f, lq = 0.f;
    const bool last_tile = n0 + p.BN >= p.N;
    const int ncols_out = (EPI == EPI_STORE || EPI == EPI_HEAD) ? min(p.N - n0, p.BN) : (last_tile ? p.out_cols - n0 : p.BN);
    const int nsteps = (ncols_out + CW - 1) / CW, sper = (nsteps + EPI_PARTS - 1) / EPI_PARTS;
    const int c_begin = min(part * sper, nsteps) * CW, c_end = min((part + 1) * sper, nsteps) * CW;
    float* stage = reinterpret_cast<float*>(gen);
    const int ldt = nsteps * CW + 4;   // +4 floats: 16-byte row stores of 8 consecutive rows fall into distinct banks
    const int rows_valid = min(BM, p.M - m0);
    if constexpr (EPI == EPI_DACT_SPLIT) {
      __syncthreads();   // all warps are past the pipeline before the tile is overwritten
      const bool avec = ((p.ldaux | n0) & 3) == 0 && aligned16(p.aux_hi);
      stage_load_tile(stage, ldt, p.aux_hi + (long long)m0 * p.ldaux + n0, p.ldaux, rows_valid, min(p.N - n0, p.BN), avec);
    }
    __syncthreads();
    epilogue_dispatch<EPI, true>(p, tmem_base + ((uint32_t)(q * 32) << 16), have_acc, row, row_ok, n0, c_begin, c_end, head_smem,
                                 l1, l2, lq, stage + rloc * ldt);
    if (threadIdx.x == 128) trace_stamp(p, 5);   // epilogue math of warp 4 done
    if constexpr (EPI == EPI_HEAD) {
      // partial head sums of parts 1..: in the (free) pipeline memory behind the staging tile, not in static shared memory
      float (*hsum)[BM][3] = reinterpret_cast<float (*)[BM][3]>(stage + BM * ldt);
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
        p.e_out[row] = p.head_mode ? fmaf(l1, l2, lq) : l1;
        p.l1_out[row] = l1;
        p.l2_out[row] = l2;
      }
    } else {
      __syncthreads();
    }
    {
      float* gout = p.out_hi + (long long)blockIdx.z * p.split_stride + (long long)m0 * p.ldo + n0;
      const bool ovec = ((p.ldo | n0) & 3) == 0 && aligned16(p.out_hi) && ((blockIdx.z ? p.split_stride : 0) & 3) == 0;
      const int cols_store = EPI == EPI_HEAD ? min(nsteps * CW, p.out_cols - n0) : ncols_out;
      stage_store_tile(stage, ldt, gout, p.ldo, rows_valid, cols_store, ovec);
    }
    tc_fence_before();
  }
  if constexpr (PAIR) cluster_sync_all();    // both CTAs are done with their accumulators before the paired dealloc
  else __syncthreads();
  if (threadIdx.x == 0) trace_stamp(p, 6);   // all warps done
  if (warp == 2) {
    tc_fence_after();
    if constexpr (PAIR) tmem_dealloc_pair(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace tc
}  // namespace fdsm
