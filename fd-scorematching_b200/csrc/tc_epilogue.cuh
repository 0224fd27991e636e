// Fused epilogues of the tcgen05 GEMM (included by tc_gemm.cuh): TMEM -> registers -> op -> global, 8 accumulator
// columns at a time inside a NON-unrolled loop.  (The first version unrolled 32 columns x 4 epilogue kinds x runtime
// activation switches into ~190 KB of straight-line SASS and was instruction-fetch bound: ncu showed 5.9 M warp
// instructions per launch but ~85 us of wall time; profiles/r01_ncu_gemm_epilogue_icache.txt.)
#pragma once

namespace fdsm {
namespace tc {

#ifndef FDSM_FAST_ACT
#define FDSM_FAST_ACT 0
#endif
constexpr int CW = 8;   // columns per epilogue step
constexpr int HW_STRIDE = 260;   // shared-memory row stride of the staged head weights [w(N) | b]

template <int ACT>
__device__ __forceinline__ float act_fwd_t(float z) {
  if constexpr (ACT == ACT_SOFTPLUS || ACT == ACT_NSOFTPLUS) {
    // nn.Softplus(beta=1, threshold=20): z > 20 ? z : log1p(exp(z)); evaluated branch-free (exp overflow -> inf is
    // discarded by the select)
#if FDSM_FAST_ACT
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
    return z > 0.f ? z : expm1f(z);
  } else {
    return z;
  }
}
template <int ACT>
__device__ __forceinline__ float act_bwd_t(float h) {   // derivative through the stored activation h = act(z)
#if FDSM_FAST_ACT
  if constexpr (ACT == ACT_SOFTPLUS) return 1.f - __expf(-h);
  else if constexpr (ACT == ACT_NSOFTPLUS) return 1.f - 0.5f * __expf(-h);
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
template <int EPI, int ACT>
__device__ __forceinline__ void epilogue_loop(const GemmParams& p, uint32_t tmem_row, bool have_acc, int row, bool row_ok,
                                              int n0, int c_begin, int c_end, const float* __restrict__ hw, float& l1,
                                              float& l2, float& lq) {
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
    if constexpr (EPI == EPI_STORE) {
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
#pragma unroll
      for (int i = 0; i < CW; ++i) {
        float h = 0.f;
        if (i < nvalid) {
          h = act_fwd_t<ACT>(v[i]);
          const int cc = n0 + c + i;
          l1 = fmaf(h, hw[cc], l1);
          if (p.head_mode) {
            l2 = fmaf(h, hw[HW_STRIDE + cc], l2);
            lq = fmaf(h * h, hw[2 * HW_STRIDE + cc], lq);
          }
        }
        v[i] = h;
      }
      if (row_ok && c < p.out_cols) st8(p.out_hi + (long long)row * p.ldo + n0 + c, v);
    } else {
      // EPI_ACT_SPLIT / EPI_DACT_SPLIT: TF32 hi/lo rows of the next GEMM's operand, padded to out_cols
      float hi[CW], lo[CW], a[CW];
      if constexpr (EPI == EPI_DACT_SPLIT) {
#pragma unroll
        for (int i = 0; i < CW; ++i) a[i] = 0.f;
        if (row_ok && nvalid > 0) {
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
        float y = 0.f;
        if (i < nvalid) {
          if constexpr (EPI == EPI_ACT_SPLIT) y = act_fwd_t<ACT>(v[i]);
          else y = v[i] * act_bwd_t<ACT>(a[i]);
        } else if (EPI == EPI_ACT_SPLIT && i == nvalid) {
          y = 1.f;   // bias column of the next layer's augmented input
        }
        if (p.out_lo) ptx::tf32_split(y, hi[i], lo[i]);
        else hi[i] = y;   // raw-operand pipeline: the consumer GEMM splits in shared memory
      }
      if (row_ok) {
        st8(p.out_hi + (long long)row * p.ldo + n0 + c, hi);
        if (p.out_lo) st8(p.out_lo + (long long)row * p.ldo + n0 + c, lo);
      }
    }
  }
}

template <int EPI>
__device__ __forceinline__ void epilogue_dispatch(const GemmParams& p, uint32_t tmem_row, bool have_acc, int row, bool row_ok,
                                                  int n0, int c_begin, int c_end, const float* hw, float& l1, float& l2,
                                                  float& lq) {
  if constexpr (EPI == EPI_STORE) {
    epilogue_loop<EPI, ACT_NONE>(p, tmem_row, have_acc, row, row_ok, n0, c_begin, c_end, hw, l1, l2, lq);
  } else {
    switch (p.act) {   // warp-uniform, outside the column loop
      case ACT_SOFTPLUS: epilogue_loop<EPI, ACT_SOFTPLUS>(p, tmem_row, have_acc, row, row_ok, n0, c_begin, c_end, hw, l1, l2, lq); break;
      case ACT_NSOFTPLUS: epilogue_loop<EPI, ACT_NSOFTPLUS>(p, tmem_row, have_acc, row, row_ok, n0, c_begin, c_end, hw, l1, l2, lq); break;
      case ACT_ELU: epilogue_loop<EPI, ACT_ELU>(p, tmem_row, have_acc, row, row_ok, n0, c_begin, c_end, hw, l1, l2, lq); break;
      default: epilogue_loop<EPI, ACT_NONE>(p, tmem_row, have_acc, row, row_ok, n0, c_begin, c_end, hw, l1, l2, lq); break;
    }
  }
}

}  // namespace tc
}  // namespace fdsm
