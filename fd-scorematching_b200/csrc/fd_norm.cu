// SURVEY.md §8f row N1, first piece: InstanceNorm++ with the ELU that follows it in every block of the score networks, forward
// and backward, as HBM-bound kernels (the reference runs it as ~10 ATen launches per direction:
// ncsnv2/models/normalization.py:150-176, ncsn/models/cond_refinenet_dilated.py:304-333; ELU at ncsnv2/models/layers.py:443-448).
//
//   mu[n,c], var[n,c] over the HW plane (biased, eps 1e-5: F.instance_norm);  m[n], v[n] = mean / UNBIASED variance of mu[n,:]
//   mustd = (mu - m) / sqrt(v + 1e-5);   y = gamma * ((x - mu) * rstd + mustd * alpha) + beta;   out = act ? elu(y) : y
//
// Layout: x, out, gout, dx are contiguous NCHW fp32; a plane (n, c) is HW consecutive floats and is owned by a group of LANES
// lanes (LANES = 4..32 by plane size: whole warps for 32x32 and 16x16 planes, sub-warp groups below), 128-bit accesses when
// HW % 4 == 0 and the bases are 16-byte aligned.  Forward = plane statistics (read x) + apply (read x, write out): 12 B per
// element against the 8 B algorithmic; backward = plane sums (read gout, out, x) + apply (read the three, write dx): 28 B
// against 16 B.  The per-sample channel statistics (C values) are recomputed by every group from the plane statistics (L1/L2
// hits) instead of a third launch.  All reductions are fixed-order shuffles: deterministic.
#include "fdsm_common.cuh"

namespace fdsm {
namespace {

constexpr float kEps = 1e-5f;
constexpr int kThreads = 256;

template <int LANES>
__device__ __forceinline__ float lanes_sum(float v) {
#pragma unroll
  for (int o = LANES / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int VEC> struct Pack;
template <> struct Pack<4> {
  float4 v;
  __device__ __forceinline__ void load(const float* p, long long q) { v = __ldg(reinterpret_cast<const float4*>(p) + q); }
  __device__ __forceinline__ void store(float* p, long long q) const { *(reinterpret_cast<float4*>(p) + q) = v; }
  __device__ __forceinline__ float& at(int i) { return (&v.x)[i]; }
};
template <> struct Pack<1> {
  float v;
  __device__ __forceinline__ void load(const float* p, long long q) { v = __ldg(p + q); }
  __device__ __forceinline__ void store(float* p, long long q) const { p[q] = v; }
  __device__ __forceinline__ float& at(int) { return v; }
};

// exp(y) - 1 as ATen's CUDA ELU forms it (exp, then the subtraction), on the MUFU pipe: expm1f's ~25 instructions made the apply
// kernel issue-bound (ncu: sm throughput 70 %, DRAM 48 %; profiles/r02e_norm_fwd_ncu_full_summary.csv)
__device__ __forceinline__ float elu_fwd(float y) { return y > 0.f ? y : __expf(y) - 1.f; }
// d elu / dy from the OUTPUT: 1 where out > 0, else exp(y) = out + 1
__device__ __forceinline__ float elu_bwd_from_out(float o) { return o > 0.f ? 1.f : o + 1.f; }

struct NormDims {
  long long planes;   // N * C
  int C, HW;
  long long p0, p1;   // the planes of this launch (a whole number of samples): see the chunking note at the entry points
};

// stats = [mu (N*C) | rstd (N*C) | mustd (N*C) | s (N)]
// ---- forward 1: plane mean and 1/sqrt(var + eps) -----------------------------------------------------------------------------
template <int LANES, int VEC>
__global__ void __launch_bounds__(kThreads, 4) inorm_stats_kernel(const float* __restrict__ x, NormDims d, float* __restrict__ stats) {
  constexpr int GROUPS = kThreads / LANES, HELD = 8;
  const int lane = threadIdx.x % LANES, grp = threadIdx.x / LANES;
  const int nq = d.HW / VEC;
  const float inv = 1.f / (float)d.HW;
  pdl_wait();
  for (long long p = d.p0 + (long long)blockIdx.x * GROUPS + grp; p < d.p1; p += (long long)gridDim.x * GROUPS) {
    const float* xp = x + p * d.HW;
    float mean, ss = 0.f;
    if (nq <= LANES * HELD) {          // the plane stays in registers between the two passes
      Pack<VEC> r[HELD];
      float sum = 0.f;
#pragma unroll
      for (int u = 0; u < HELD; ++u) {
        const int q = lane + u * LANES;
        if (q < nq) {
          r[u].load(xp, q);
#pragma unroll
          for (int i = 0; i < VEC; ++i) sum += r[u].at(i);
        }
      }
      mean = lanes_sum<LANES>(sum) * inv;
#pragma unroll
      for (int u = 0; u < HELD; ++u) {
        const int q = lane + u * LANES;
        if (q < nq) {
#pragma unroll
          for (int i = 0; i < VEC; ++i) { const float t = r[u].at(i) - mean; ss = fmaf(t, t, ss); }
        }
      }
    } else {                           // long planes: second pass re-reads (L1 / L2)
      float sum = 0.f;
      for (int q = lane; q < nq; q += LANES) {
        Pack<VEC> r; r.load(xp, q);
#pragma unroll
        for (int i = 0; i < VEC; ++i) sum += r.at(i);
      }
      mean = lanes_sum<LANES>(sum) * inv;
      for (int q = lane; q < nq; q += LANES) {
        Pack<VEC> r; r.load(xp, q);
#pragma unroll
        for (int i = 0; i < VEC; ++i) { const float t = r.at(i) - mean; ss = fmaf(t, t, ss); }
      }
    }
    const float var = lanes_sum<LANES>(ss) * inv;
    if (lane == 0) {
      stats[p] = mean;
      stats[d.planes + p] = 1.f / sqrtf(var + kEps);
    }
  }
}

// mean and 1/sqrt(unbiased var + eps) of the C plane means of sample n, by one group (every lane gets both)
template <int LANES>
__device__ __forceinline__ void sample_stats(const float* __restrict__ mu_n, int C, int lane, float& m, float& s) {
  float sum = 0.f;
  for (int k = lane; k < C; k += LANES) sum += __ldcg(mu_n + k);
  m = lanes_sum<LANES>(sum) / (float)C;
  float ss = 0.f;
  for (int k = lane; k < C; k += LANES) { const float t = __ldcg(mu_n + k) - m; ss = fmaf(t, t, ss); }
  const float v = lanes_sum<LANES>(ss) / (float)(C - 1);      // C == 1: 0/0 = NaN, as torch.var of one element
  s = 1.f / sqrtf(v + kEps);
}

// ---- forward 2: out = act(gamma * ((x - mu) * rstd + mustd * alpha) + beta) -------------------------------------------------------
template <int LANES, int VEC>
__global__ void __launch_bounds__(kThreads) inorm_apply_kernel(const float* __restrict__ x, NormDims d, const float* __restrict__ gamma,
                                                               const float* __restrict__ alpha, const float* __restrict__ beta,
                                                               long long pstride, int act, float* __restrict__ stats,
                                                               float* __restrict__ out) {
  constexpr int GROUPS = kThreads / LANES;
  const int lane = threadIdx.x % LANES, grp = threadIdx.x / LANES;
  const int nq = d.HW / VEC;
  pdl_wait();
  for (long long p = d.p0 + (long long)blockIdx.x * GROUPS + grp; p < d.p1; p += (long long)gridDim.x * GROUPS) {
    const long long n = p / d.C;
    const int c = (int)(p - n * d.C);
    float m, s;
    sample_stats<LANES>(stats + n * d.C, d.C, lane, m, s);
    const float mu = __ldcg(stats + p), rstd = __ldcg(stats + d.planes + p);
    const float mustd = (mu - m) * s;
    if (lane == 0) {
      stats[2 * d.planes + p] = mustd;
      if (c == 0) stats[3 * d.planes + n] = s;
    }
    const float g = __ldg(gamma + n * pstride + c), shift = mustd * __ldg(alpha + n * pstride + c);
    const float b = beta ? __ldg(beta + n * pstride + c) : 0.f;
    const float* xp = x + p * d.HW;
    float* op = out + p * d.HW;
#pragma unroll 4
    for (int q = lane; q < nq; q += LANES) {
      Pack<VEC> r; r.load(xp, q);
#pragma unroll
      for (int i = 0; i < VEC; ++i) {
        const float y = fmaf(g, fmaf(r.at(i) - mu, rstd, shift), b);
        r.at(i) = act ? elu_fwd(y) : y;
      }
      r.store(op, q);
    }
  }
}

// ---- backward 1: S1 = sum g, S2 = sum g * xhat per plane (g = gout * elu'(y)); per-sample parameter gradients --------------------
template <int LANES, int VEC>
__global__ void __launch_bounds__(kThreads) inorm_bwd_sums_kernel(const float* __restrict__ x, const float* __restrict__ out,
                                                                  const float* __restrict__ gout, NormDims d,
                                                                  const float* __restrict__ gamma, const float* __restrict__ alpha,
                                                                  long long pstride, int act, const float* __restrict__ stats,
                                                                  float* __restrict__ s2, float* __restrict__ dgamma,
                                                                  float* __restrict__ dalpha, float* __restrict__ dbeta) {
  constexpr int GROUPS = kThreads / LANES;
  const int lane = threadIdx.x % LANES, grp = threadIdx.x / LANES;
  const int nq = d.HW / VEC;
  pdl_wait();
  for (long long p = d.p0 + (long long)blockIdx.x * GROUPS + grp; p < d.p1; p += (long long)gridDim.x * GROUPS) {
    const float mu = __ldg(stats + p), rstd = __ldg(stats + d.planes + p);
    const float* xp = x + p * d.HW;
    const float* op = out + p * d.HW;
    const float* gp = gout + p * d.HW;
    float a1 = 0.f, a2 = 0.f;
#pragma unroll 4
    for (int q = lane; q < nq; q += LANES) {
      Pack<VEC> rx, rg, ro;
      rx.load(xp, q); rg.load(gp, q);
      if (act) ro.load(op, q);
#pragma unroll
      for (int i = 0; i < VEC; ++i) {
        const float g = act ? rg.at(i) * elu_bwd_from_out(ro.at(i)) : rg.at(i);
        a1 += g;
        a2 = fmaf(g, (rx.at(i) - mu) * rstd, a2);
      }
    }
    a1 = lanes_sum<LANES>(a1);
    a2 = lanes_sum<LANES>(a2);
    if (lane == 0) {
      const long long n = p / d.C;
      const int c = (int)(p - n * d.C);
      const float mustd = __ldg(stats + 2 * d.planes + p);
      const float ga = __ldg(gamma + n * pstride + c), al = __ldg(alpha + n * pstride + c);
      s2[p] = a2;
      dbeta[p] = a1;                               // (= S1: the apply kernel reads it back)
      dgamma[p] = fmaf(mustd * al, a1, a2);        // sum g * (xhat + mustd * alpha)
      dalpha[p] = ga * mustd * a1;
    }
  }
}

// ---- backward 2: dx = gamma * rstd * (g - S1/HW - xhat * S2/HW) + dmu/HW, dmu through the channel standardisation -----------------
template <int LANES, int VEC>
__global__ void __launch_bounds__(kThreads) inorm_bwd_apply_kernel(const float* __restrict__ x, const float* __restrict__ out,
                                                                   const float* __restrict__ gout, NormDims d,
                                                                   const float* __restrict__ gamma, const float* __restrict__ alpha,
                                                                   long long pstride, int act, const float* __restrict__ stats,
                                                                   const float* __restrict__ s1, const float* __restrict__ s2,
                                                                   float* __restrict__ dx) {
  constexpr int GROUPS = kThreads / LANES;
  const int lane = threadIdx.x % LANES, grp = threadIdx.x / LANES;
  const int nq = d.HW / VEC;
  const float inv = 1.f / (float)d.HW;
  pdl_wait();
  for (long long p = d.p0 + (long long)blockIdx.x * GROUPS + grp; p < d.p1; p += (long long)gridDim.x * GROUPS) {
    const long long n = p / d.C;
    const int c = (int)(p - n * d.C);
    // d_k = dL/dmustd_k = gamma_k alpha_k S1_k;  D1 = mean_k d_k,  D2 = sum_k d_k mustd_k
    float t1 = 0.f, t2 = 0.f;
    for (int k = lane; k < d.C; k += LANES) {
      const float dk = __ldg(gamma + n * pstride + k) * __ldg(alpha + n * pstride + k) * __ldcg(s1 + n * d.C + k);
      t1 += dk;
      t2 = fmaf(dk, __ldg(stats + 2 * d.planes + n * d.C + k), t2);
    }
    const float D1 = lanes_sum<LANES>(t1) / (float)d.C, D2 = lanes_sum<LANES>(t2);
    const float mu = __ldg(stats + p), rstd = __ldg(stats + d.planes + p), mustd = __ldg(stats + 2 * d.planes + p);
    const float s = __ldg(stats + 3 * d.planes + n);
    const float ga = __ldg(gamma + n * pstride + c), al = __ldg(alpha + n * pstride + c);
    const float S1 = __ldcg(s1 + p), S2 = __ldcg(s2 + p);
    const float dmu = s * (ga * al * S1 - D1 - mustd * D2 / (float)(d.C - 1));
    const float k1 = ga * rstd;
    const float c0 = (dmu - k1 * S1) * inv, c2 = -k1 * S2 * inv;
    const float* xp = x + p * d.HW;
    const float* op = out + p * d.HW;
    const float* gp = gout + p * d.HW;
    float* dp = dx + p * d.HW;
#pragma unroll 4
    for (int q = lane; q < nq; q += LANES) {
      Pack<VEC> rx, rg, ro;
      rx.load(xp, q); rg.load(gp, q);
      if (act) ro.load(op, q);
#pragma unroll
      for (int i = 0; i < VEC; ++i) {
        const float g = act ? rg.at(i) * elu_bwd_from_out(ro.at(i)) : rg.at(i);
        rx.at(i) = fmaf(k1, g, fmaf(c2, (rx.at(i) - mu) * rstd, c0));
      }
      rx.store(dp, q);
    }
  }
}

// group width by plane size: the smallest power of two >= min(32, items per plane), at least 4
inline int pick_lanes(long long nq) { return nq >= 32 ? 32 : nq > 8 ? 16 : nq > 4 ? 8 : 4; }
inline unsigned grid_for(long long planes, int lanes) {
  const long long groups = kThreads / lanes, want = (planes + groups - 1) / groups, cap = (long long)sm_count() * 32;
  return (unsigned)(want < cap ? (want > 0 ? want : 1) : cap);
}

// Samples per launch pair (FDSM_NORM_CHUNK_MB > 0; default 0 = the whole tensor in one pair).  The two kernels of a direction read
// the same activations twice, and samples are independent, so the tensor could be walked in chunks small enough for the second
// kernel to find the first one's reads in L2.  Measured on B200 (profiles/r02_norm_kernels_chunk_sweep.txt), it does not pay:
// at (256, 128, 32, 32) the forward takes 87 us unchunked, 105 / 108 / 126 / 143 / 189 us with 96 / 64 / 48 / 32 / 16 MB
// chunks -- the short launches lose more to ramps and tails than the L2 hits return.  The knob stays for other shapes.
inline long long chunk_samples(long long N, long long C, long long HW, int tensors) {
  static const long long mb = [] { const char* e = getenv("FDSM_NORM_CHUNK_MB"); return e ? atoll(e) : 0ll; }();
  if (mb <= 0) return N;
  const long long per_sample = C * HW * 4 * tensors;
  long long n = (mb << 20) / (per_sample > 0 ? per_sample : 1);
  return n < 1 ? 1 : (n > N ? N : n);
}

#define FDSM_NORM_DISPATCH(KERNEL, ...)                                                                                     \
  do {                                                                                                                      \
    cudaError_t e_;                                                                                                         \
    const dim3 g_(grid_for(d.p1 - d.p0, lanes)), b_(kThreads);                                                                 \
    if (vec == 4) {                                                                                                         \
      if (lanes == 32) e_ = launch_pdl(KERNEL<32, 4>, g_, b_, 0, st, __VA_ARGS__);                                          \
      else if (lanes == 16) e_ = launch_pdl(KERNEL<16, 4>, g_, b_, 0, st, __VA_ARGS__);                                     \
      else if (lanes == 8) e_ = launch_pdl(KERNEL<8, 4>, g_, b_, 0, st, __VA_ARGS__);                                       \
      else e_ = launch_pdl(KERNEL<4, 4>, g_, b_, 0, st, __VA_ARGS__);                                                       \
    } else {                                                                                                                \
      if (lanes == 32) e_ = launch_pdl(KERNEL<32, 1>, g_, b_, 0, st, __VA_ARGS__);                                          \
      else if (lanes == 16) e_ = launch_pdl(KERNEL<16, 1>, g_, b_, 0, st, __VA_ARGS__);                                     \
      else if (lanes == 8) e_ = launch_pdl(KERNEL<8, 1>, g_, b_, 0, st, __VA_ARGS__);                                       \
      else e_ = launch_pdl(KERNEL<4, 1>, g_, b_, 0, st, __VA_ARGS__);                                                       \
    }                                                                                                                       \
    rc = cuda_status(e_, #KERNEL);                                                                                          \
  } while (0)

}  // namespace
}  // namespace fdsm

using namespace fdsm;

extern "C" int64_t fdsm_inorm_plus_stats_floats(int64_t N, int64_t C) { return N < 0 || C < 0 ? -1 : 3 * N * C + N; }

extern "C" int64_t fdsm_inorm_plus_launches(int64_t N, int64_t C, int64_t HW, int32_t backward, int32_t act, int32_t with_dx) {
  if (N <= 0 || C <= 0 || HW <= 0) return 0;
  const long long step = backward ? (with_dx ? chunk_samples(N, C, HW, act ? 4 : 3) : N) : chunk_samples(N, C, HW, 2);
  return (N + step - 1) / step * ((backward && !with_dx) ? 1 : 2);
}

extern "C" int fdsm_inorm_plus_fwd(const float* x, int64_t N, int64_t C, int64_t HW, const float* gamma, const float* alpha,
                                   const float* beta, int64_t param_stride, int32_t act, float* out, float* stats,
                                   fdsm_stream_t stream) {
  FDSM_REQUIRE(N >= 0 && C >= 1 && HW >= 1 && C <= INT32_MAX && HW <= INT32_MAX, "fdsm_inorm_plus_fwd: bad sizes");
  FDSM_REQUIRE(param_stride == 0 || param_stride >= C, "fdsm_inorm_plus_fwd: param_stride must be 0 (shared rows) or >= C");
  FDSM_REQUIRE(act == 0 || act == 1, "fdsm_inorm_plus_fwd: act must be 0 (none) or 1 (ELU)");
  if (N == 0) return FDSM_OK;
  FDSM_REQUIRE(x && gamma && alpha && out && stats, "fdsm_inorm_plus_fwd: NULL pointer");
  NormDims d{N * C, (int)C, (int)HW, 0, 0};
  const int vec = (HW % 4 == 0 && aligned16(x) && aligned16(out)) ? 4 : 1;
  const int lanes = pick_lanes(HW / vec);
  cudaStream_t st = (cudaStream_t)stream;
  const long long step = chunk_samples(N, C, HW, 2);      // x + out
  int rc = FDSM_OK;
  for (long long n0 = 0; n0 < N && !rc; n0 += step) {
    d.p0 = n0 * C; d.p1 = (n0 + step < N ? n0 + step : N) * C;
    FDSM_NORM_DISPATCH(inorm_stats_kernel, x, d, stats);
    if (rc) return rc;
    FDSM_NORM_DISPATCH(inorm_apply_kernel, x, d, gamma, alpha, beta, (long long)param_stride, (int)act, stats, out);
  }
  return rc;
}

extern "C" int fdsm_inorm_plus_bwd(const float* x, const float* out, const float* gout, int64_t N, int64_t C, int64_t HW,
                                   const float* gamma, const float* alpha, int64_t param_stride, int32_t act, const float* stats,
                                   float* scratch, float* dx, float* dgamma, float* dalpha, float* dbeta, fdsm_stream_t stream) {
  FDSM_REQUIRE(N >= 0 && C >= 1 && HW >= 1 && C <= INT32_MAX && HW <= INT32_MAX, "fdsm_inorm_plus_bwd: bad sizes");
  FDSM_REQUIRE(param_stride == 0 || param_stride >= C, "fdsm_inorm_plus_bwd: param_stride must be 0 (shared rows) or >= C");
  FDSM_REQUIRE(act == 0 || act == 1, "fdsm_inorm_plus_bwd: act must be 0 (none) or 1 (ELU)");
  if (N == 0) return FDSM_OK;
  FDSM_REQUIRE(x && gout && gamma && alpha && stats && scratch && dgamma && dalpha && dbeta && (out || !act),
               "fdsm_inorm_plus_bwd: NULL pointer (out is needed when act = 1; dx alone may be NULL)");
  NormDims d{N * C, (int)C, (int)HW, 0, 0};
  const int vec = (HW % 4 == 0 && aligned16(x) && aligned16(gout) && (!act || aligned16(out)) && (!dx || aligned16(dx))) ? 4 : 1;
  const int lanes = pick_lanes(HW / vec);
  cudaStream_t st = (cudaStream_t)stream;
  const float* o = act ? out : x;      // never read when act = 0
  const long long step = dx ? chunk_samples(N, C, HW, act ? 4 : 3) : N;      // x, gout, (out,) dx
  int rc = FDSM_OK;
  for (long long n0 = 0; n0 < N && !rc; n0 += step) {
    d.p0 = n0 * C; d.p1 = (n0 + step < N ? n0 + step : N) * C;
    FDSM_NORM_DISPATCH(inorm_bwd_sums_kernel, x, o, gout, d, gamma, alpha, (long long)param_stride, (int)act, stats, scratch, dgamma,
                       dalpha, dbeta);
    if (rc || !dx) continue;
    FDSM_NORM_DISPATCH(inorm_bwd_apply_kernel, x, o, gout, d, gamma, alpha, (long long)param_stride, (int)act, stats,
                       (const float*)dbeta, (const float*)scratch, dx);
  }
  return rc;
}
