// Fused epilogues of the tcgen05 GEMM (included by tc_gemm.cuh): TMEM -> registers -> op -> global, 8 accumulator
// columns at a time inside a NON-unrolled loop.  (The first version unrolled 32 columns x 4 epilogue kinds x runtime
// activation switches into ~190 KB of straight-line SASS and was instruction-fetch bound: ncu showed 5.9 M warp
// instructions per launch but ~85 us of wall time; profiles/r01_ncu_gemm_epilogue_icache.txt.)
#pragma once

namespace fdsm {
namespace tc {

// Activation arithmetic of the epilogues.  The 128 x 208 tile of one CTA is 26 624 activations on 256 threads, so the
// epilogue is ALU-issue bound: libdevice's log1pf(expf(z)) costs ~45 instructions per element = 4.8 us per tile, more
// than a K=200 main loop (profiles/r01_step_trace_gemm_timeline.txt: epi-math 7.3 us).
//   0: libdevice (log1pf(expf(z)), expm1f, expf)
//   1: MUFU only (ex2.approx / lg2.approx): ~10 instructions, absolute error ~2e-7 (3 ulp at O(1))
//   2 (default): ex2.approx for e^-|z| and log1p(t) = 2 atanh(t / (2 + t)) as an odd polynomial in s = t/(2+t) <= 1/3
//      (Newton-refined reciprocal, 6 FMAs): ~19 instructions, |error| <= 5 ulp, the same grade as 0 (3 ulp; simulated
//      against float64 in fp32 arithmetic over [-25, 25]); derivative 1 - c e^-h with ex2.approx (|error| < 1 ulp)
#ifndef FDSM_FAST_ACT
#define FDSM_FAST_ACT 2
#endif
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// softplus(z) = max(z, 0) + log1p(e^-|z|)
__device__ __forceinline__ float softplus_poly(float z) {
  const float t = ex2_approx(fabsf(z) * -1.4426950408889634f);   // e^-|z| in (0, 1]
  const float d = 2.f + t;
  const float r0 = rcp_approx(d);
  const float s0 = t * r0;
  const float s = fmaf(fmaf(-d, s0, t), r0, s0);                  // t / (2 + t), one Newton step on the quotient
  const float u = s * s;                                         // <= 1/9
  float q = 0.076923077f;                                        // 1/13
  q = fmaf(q, u, 0.090909091f);
  q = fmaf(q, u, 0.11111111f);
  q = fmaf(q, u, 0.14285715f);
  q = fmaf(q, u, 0.2f);
  q = fmaf(q, u, 0.33333334f);
  const float s2 = s + s;
  return fmaxf(z, 0.f) + fmaf(s2 * u, q, s2);                    // 2 atanh(s) = 2s (1 + u/3 + u^2/5 + ...)
}
constexpr int CW = 8;   // columns per epilogue step
constexpr int HW_STRIDE = 260;   // shared-memory row stride of the staged head weights: w[0, N), zeros, bias at HW_BIAS
constexpr int HW_BIAS = HW_STRIDE - 1;

template <int ACT>
__device__ __forceinline__ float act_fwd_t(float z) {
  if constexpr (ACT == ACT_SOFTPLUS || ACT == ACT_NSOFTPLUS) {
    // nn.Softplus(beta=1, threshold=20): z > 20 ? z : log1p(exp(z)); evaluated branch-free (exp overflow -> inf is
    // discarded by the select)
#if FDSM_FAST_ACT == 2
    const float r = softplus_poly(z);
#elif FDSM_FAST_ACT == 1
    // softplus(z) = max(z, 0) + log1p(exp(-|z|)) on the MUFU pipe (ex2.approx / lg2.approx): absolute error ~2e-7 at
    // O(1) values (3 fp32 ulp) for ~10 instructions instead of ~45 -- the precise form made the epilogue, not the MMA,
    // the critical path of the K=200 layers.  z > 20 falls out naturally (log1p(e^-20) < 2^-24 z).
    const float t = __expf(-fabsf(z));
    const float r = fmaxf(z, 0.f) + __logf(1.f + t);
#else
    const float sp = log1pf(expf(z));
    const float r = z > 20.f ? z : sp;
#endif
    return ACT == ACT_NSOFTPLUS ? r - 0.69314718055994531f : r;   // models.py:19-28
  } else if constexpr (ACT == ACT_ELU) {
#if FDSM_FAST_ACT == 1
    return z > 0.f ? z : __expf(z) - 1.f;   // absolute error ~1e-7, branch-free
#else
    return z > 0.f ? z : expm1f(z);
#endif
  } else {
    return z;
  }
}
template <int ACT>
__device__ __forceinline__ float act_bwd_t(float h) {   // derivative through the stored activation h = act(z)
#if FDSM_FAST_ACT
  if constexpr (ACT == ACT_SOFTPLUS) return 1.f - ex2_approx(h * -1.4426950408889634f);
  else if constexpr (ACT == ACT_NSOFTPLUS) return fmaf(-0.5f, ex2_approx(h * -1.4426950408889634f), 1.f);
#else
  if constexpr (ACT == ACT_SOFTPLUS) return 1.f - expf(-h);
  else if constexpr (ACT == ACT_NSOFTPLUS) return 1.f - 0.5f * expf(-h);
#endif
  else if constexpr (ACT == ACT_ELU) return h > 0.f ? 1.f : h + 1.f;
  else return 1.f;
}

__device__ __forceinline__ void st8(float* o, const float (&v)[CW]) {
  *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// One thread = one output row; columns [c_begin, c_end) of this CTA's tile in steps of CW.
// STAGED: results go to this thread's row of a shared-memory tile (`srow`, columns relative to n0) and are written to
// global memory afterwards by coalesced row copies (stage_store_tile); EPI_DACT_SPLIT then also finds its aux operand in
// the same tile (stage_load_tile), which it overwrites in place.  Un-staged, every warp-wide 16-byte store touches 32
// different rows = 32 L1 wavefronts, and the epilogue of a 128x208 tile took 4.7 us (profiles/r01_gemm_trace.txt).
template <int EPI, int ACT, bool STAGED = false>
__device__ __forceinline__ void epilogue_loop(const GemmParams& p, uint32_t tmem_row, bool have_acc, int row, bool row_ok,
                                              int n0, int c_begin, int c_end, const float* __restrict__ hw, float& l1,
                                              float& l2, float& lq, float* srow = nullptr) {
  const bool vec_ok = ((p.ldo | n0) & 3) == 0 && aligned16(p.out_hi) && ((blockIdx.z ? p.split_stride : 0) & 3) == 0;
#pragma unroll 1
  for (int c = c_begin; c < c_end; c += CW) {
    float v[CW];
    if (have_acc && c < p.BN) {
      ptx::tmem_ld_32x8(tmem_row + (uint32_t)c, v);
    } else {
#pragma unroll
      for (int i = 0; i < CW; ++i) v[i] = 0.f;
    }
    const int nvalid = p.N - n0 - c;   // columns [0, nvalid) of this step are real outputs
    if constexpr (STAGED && EPI == EPI_STORE) {
      st8(srow + c, v);
    } else if constexpr (EPI == EPI_STORE) {
      if (row_ok) {
        float* o = p.out_hi + (long long)blockIdx.z * p.split_stride + (long long)row * p.ldo + n0 + c;
        if (nvalid >= CW && vec_ok) {
          st8(o, v);
        } else {
#pragma unroll
          for (int i = 0; i < CW; ++i)
            if (i < nvalid) o[i] = v[i];
        }
      }
    } else if constexpr (EPI == EPI_HEAD) {
      // head weights of these 8 columns: 128-bit broadcast loads (zero beyond N, so no per-element select on them)
      const float* hwc = hw + c;   // staged per tile: index = column within the tile
      float wa[CW], wb[CW], wq[CW];
      {
        const float4 a0 = *reinterpret_cast<const float4*>(hwc), a1 = *reinterpret_cast<const float4*>(hwc + 4);
        wa[0] = a0.x; wa[1] = a0.y; wa[2] = a0.z; wa[3] = a0.w; wa[4] = a1.x; wa[5] = a1.y; wa[6] = a1.z; wa[7] = a1.w;
        if (p.head_mode) {
          const float4 b0 = *reinterpret_cast<const float4*>(hwc + HW_STRIDE), b1 = *reinterpret_cast<const float4*>(hwc + HW_STRIDE + 4);
          const float4 q0 = *reinterpret_cast<const float4*>(hwc + 2 * HW_STRIDE), q1 = *reinterpret_cast<const float4*>(hwc + 2 * HW_STRIDE + 4);
          wb[0] = b0.x; wb[1] = b0.y; wb[2] = b0.z; wb[3] = b0.w; wb[4] = b1.x; wb[5] = b1.y; wb[6] = b1.z; wb[7] = b1.w;
          wq[0] = q0.x; wq[1] = q0.y; wq[2] = q0.z; wq[3] = q0.w; wq[4] = q1.x; wq[5] = q1.y; wq[6] = q1.z; wq[7] = q1.w;
        }
      }
#pragma unroll
      for (int i = 0; i < CW; ++i) {
        // branch-free: the 8 columns of a step must overlap in the pipes; with per-element branches each activation
        // ran alone, ~100 clk per element (profiles/r01_step_trace.txt)
        const float h = act_fwd_t<ACT>(v[i]);     // finite for padding columns too (accumulator 0)
        l1 = fmaf(h, wa[i], l1);
        if (p.head_mode) {
          l2 = fmaf(h, wb[i], l2);
          lq = fmaf(h * h, wq[i], lq);
        }
        v[i] = i < nvalid ? h : 0.f;
      }
      if constexpr (STAGED) st8(srow + c, v);
      else if (row_ok && c < p.out_cols) st8(p.out_hi + (long long)row * p.ldo + n0 + c, v);
    } else {
      // EPI_ACT_SPLIT / EPI_DACT_SPLIT: TF32 hi/lo rows of the next GEMM's operand, padded to out_cols
      float hi[CW], lo[CW], a[CW];
      if constexpr (EPI == EPI_DACT_SPLIT) {
#pragma unroll
        for (int i = 0; i < CW; ++i) a[i] = 0.f;
        if constexpr (STAGED) {
          const float4 x0 = *reinterpret_cast<const float4*>(srow + c), x1 = *reinterpret_cast<const float4*>(srow + c + 4);
          a[0] = x0.x; a[1] = x0.y; a[2] = x0.z; a[3] = x0.w; a[4] = x1.x; a[5] = x1.y; a[6] = x1.z; a[7] = x1.w;
        } else if (row_ok && nvalid > 0) {
          const float* ah = p.aux_hi + (long long)row * p.ldaux + n0 + c;   // ldaux % 4 == 0, 16-byte aligned base
          const float4 x0 = ldg4(ah), x1 = ldg4(ah + 4);
          a[0] = x0.x; a[1] = x0.y; a[2] = x0.z; a[3] = x0.w; a[4] = x1.x; a[5] = x1.y; a[6] = x1.z; a[7] = x1.w;
          if (p.aux_lo) {
            const float* al = p.aux_lo + (long long)row * p.ldaux + n0 + c;
            const float4 y0 = ldg4(al), y1 = ldg4(al + 4);
            a[0] += y0.x; a[1] += y0.y; a[2] += y0.z; a[3] += y0.w; a[4] += y1.x; a[5] += y1.y; a[6] += y1.z; a[7] += y1.w;
          }
        }
      }
#pragma unroll
      for (int i = 0; i < CW; ++i) {
        float y;   // branch-free, see EPI_HEAD
        if constexpr (EPI == EPI_ACT_SPLIT) {
          const float h = act_fwd_t<ACT>(v[i]);
          y = i < nvalid ? h : (i == nvalid ? 1.f : 0.f);   // column N: bias column of the next layer's augmented input
        } else {
          const float g = v[i] * act_bwd_t<ACT>(a[i]);
          y = i < nvalid ? g : 0.f;
        }
        if (p.out_lo) ptx::tf32_split(y, hi[i], lo[i]);
        else hi[i] = y;   // raw-operand pipeline: the consumer GEMM splits in shared memory
      }
      if constexpr (STAGED) {
        st8(srow + c, hi);   // raw-operand pipeline only: out_lo == NULL
      } else if (row_ok) {
        st8(p.out_hi + (long long)row * p.ldo + n0 + c, hi);
        if (p.out_lo) st8(p.out_lo + (long long)row * p.ldo + n0 + c, lo);
      }
    }
  }
}

template <int EPI, bool STAGED = false>
__device__ __forceinline__ void epilogue_dispatch(const GemmParams& p, uint32_t tmem_row, bool have_acc, int row, bool row_ok,
                                                  int n0, int c_begin, int c_end, const float* hw, float& l1, float& l2,
                                                  float& lq, float* srow = nullptr) {
  if constexpr (EPI == EPI_STORE) {
    epilogue_loop<EPI, ACT_NONE, STAGED>(p, tmem_row, have_acc, row, row_ok, n0, c_begin, c_end, hw, l1, l2, lq, srow);
  } else {
    switch (p.act) {   // warp-uniform, outside the column loop
      case ACT_SOFTPLUS: epilogue_loop<EPI, ACT_SOFTPLUS, STAGED>(p, tmem_row, have_acc, row, row_ok, n0, c_begin, c_end, hw, l1, l2, lq, srow); break;
      case ACT_NSOFTPLUS: epilogue_loop<EPI, ACT_NSOFTPLUS, STAGED>(p, tmem_row, have_acc, row, row_ok, n0, c_begin, c_end, hw, l1, l2, lq, srow); break;
      case ACT_ELU: epilogue_loop<EPI, ACT_ELU, STAGED>(p, tmem_row, have_acc, row, row_ok, n0, c_begin, c_end, hw, l1, l2, lq, srow); break;
      default: epilogue_loop<EPI, ACT_NONE, STAGED>(p, tmem_row, have_acc, row, row_ok, n0, c_begin, c_end, hw, l1, l2, lq, srow); break;
    }
  }
}

// Coalesced copies between a [rows x ld] global tile and the [BM x ldt] shared staging tile: warp w takes rows w, w+8, ...;
// the lanes sweep a row in 16-byte pieces (scalar when the global side is not 16-byte aligned).
__device__ __forceinline__ void stage_store_tile(const float* __restrict__ stage, int ldt, float* __restrict__ g, long long ld,
                                                 int rows, int cols, bool vec) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (vec && (cols & 3) == 0 && (ldt & 3) == 0 && cols > 0) {
    // one asynchronous bulk copy per row (shared -> global through the TMA unit): the threads only wait until the
    // shared-memory reads are done, the global writes drain behind the CTA
    ptx::fence_proxy_async();   // this thread's generic-proxy tile writes -> visible to the async proxy
    __syncthreads();
    for (int r = threadIdx.x; r < rows; r += kThreads) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                   ::"l"(g + (long long)r * ld), "r"(ptx::smem_u32(stage + r * ldt)), "r"(cols * 4)
                   : "memory");
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    return;
  }
  const int c4 = vec ? (cols & ~3) : 0;
  for (int r = warp; r < rows; r += kThreads / 32) {
    const float* s = stage + r * ldt;
    float* o = g + (long long)r * ld;
    for (int c = 4 * lane; c < c4; c += 128) *reinterpret_cast<float4*>(o + c) = *reinterpret_cast<const float4*>(s + c);
    for (int c = c4 + lane; c < cols; c += 32) o[c] = s[c];
  }
}
__device__ __forceinline__ void stage_load_tile(float* __restrict__ stage, int ldt, const float* __restrict__ g, long long ld,
                                                int rows, int cols, bool vec) {
  // the tile as a flat list of 16-byte pieces, 8 independent loads in flight per thread (a row-by-row loop exposed one
  // global-memory latency per row: 14 us for a 128x200 tile, profiles/r01_step_trace.txt)
  if (vec) {
    const int c4n = cols >> 2, total = rows * c4n;
    constexpr int UNR = 8;
    for (int i0 = threadIdx.x; i0 < total; i0 += kThreads * UNR) {
      float4 x[UNR];
      int off[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int i = i0 + u * kThreads;
        if (i < total) {
          const int r = i / c4n, c = (i - r * c4n) << 2;
          off[u] = r * ldt + c;
          x[u] = ldg4(g + (long long)r * ld + c);
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u)
        if (i0 + u * kThreads < total) *reinterpret_cast<float4*>(stage + off[u]) = x[u];
    }
  }
  const int c4 = vec ? (cols & ~3) : 0;
  if (c4 < cols) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = warp; r < rows; r += kThreads / 32)
      for (int c = c4 + lane; c < cols; c += 32) stage[r * ldt + c] = __ldg(g + (long long)r * ld + c);
  }
}

}  // namespace tc
}  // namespace fdsm
