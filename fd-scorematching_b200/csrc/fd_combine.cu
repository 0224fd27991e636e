// Kernel (4): the central-difference combine and its analytic backward.
//   energy / DSM forms: O(B) data -> one 1024-thread block, fixed-order (deterministic) reduction.
//   score-net form: O(B*D) streaming, HBM-bound; fused fwd+bwd pass = 20*D bytes per sample
//   (R s1,s2,v ; W ds1,ds2).  Row sums by warp shuffles + shared-memory staging, then a tiny
//   fixed-order mean kernel.
#include "fd_combine.cuh"

#include <mutex>

namespace fdsm {
namespace {

constexpr int kBlock = 256;

// ---- energy form (device helpers in fd_combine.cuh) ----------------------------------------------
__global__ void __launch_bounds__(1024) combine_energy_fwd_kernel(const float* __restrict__ E, int64_t B, int mode,
                                                                  float c1, float scale, int mean,
                                                                  float* __restrict__ per_sample,
                                                                  float* __restrict__ loss) {
  __shared__ float red[32];
  combine_energy_fwd_block(E, B, mode, c1, scale, mean, per_sample, loss, red);
}

__global__ void __launch_bounds__(256) combine_energy_bwd_kernel(const float* __restrict__ E, int64_t B, int mode,
                                                                 float c1, float coef, const float* __restrict__ gout,
                                                                 float* __restrict__ dE) {
  const float c = coef * (gout ? __ldg(gout) : 1.f);
  for (int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; b < B; b += (int64_t)gridDim.x * blockDim.x) {
    if (mode == 0) {
      const float q = (E[B + b] - E[2 * B + b]) * 0.25f;
      dE[b] = c * c1;
      dE[B + b] = c * (q - 1.f);
      dE[2 * B + b] = c * (-q - 1.f);
    } else {
      const float g = c * (2.f * (E[b] - E[B + b]) + 4.f);
      dE[b] = g;
      dE[B + b] = -g;
    }
  }
}

// ---- DSM energy form ---------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) combine_dsm_fwd_kernel(const float* __restrict__ E, const float* __restrict__ dot,
                                                               const float* __restrict__ sigma_b, int64_t B, float alpha,
                                                               float beta, float gamma, float* __restrict__ per_sample,
                                                               float* __restrict__ loss) {
  __shared__ float red[32];
  combine_dsm_fwd_block(E, dot, sigma_b, B, alpha, beta, gamma, per_sample, loss, red);
}

__global__ void __launch_bounds__(256) combine_dsm_bwd_kernel(const float* __restrict__ E, const float* __restrict__ dot,
                                                              const float* __restrict__ sigma_b, int64_t B, float alpha,
                                                              float beta, float gamma, const float* __restrict__ gout,
                                                              float* __restrict__ dE) {
  const float c = 2.f * alpha * gamma / (float)B * (gout ? __ldg(gout) : 1.f);
  for (int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; b < B; b += (int64_t)gridDim.x * blockDim.x) {
    const float g = combine_dsm_grad(E, dot, sigma_b, B, alpha, beta, c, b);
    dE[b] = -g;
    dE[B + b] = g;
  }
}

// ---- score-net form -----------------------------------------------------------------------------
struct ScoreArgs {
  const float* s; int64_t lds;
  const float* v; int64_t ldv;
  const float* sigma; const int64_t* labels; float power;
  int64_t B; int D; float k; int projected;
  const float* gout;
  float* per_sample; float* ds; int64_t ldds;
};

template <bool VEC>
__device__ __forceinline__ float4 ld4(const float* row, int j, int D) {
  if constexpr (VEC) return ldg4(row + 4 * j);
  const int e = 4 * j;
  return make_float4(e < D ? __ldg(row + e) : 0.f, e + 1 < D ? __ldg(row + e + 1) : 0.f,
                     e + 2 < D ? __ldg(row + e + 2) : 0.f, e + 3 < D ? __ldg(row + e + 3) : 0.f);
}
template <bool VEC>
__device__ __forceinline__ void st4(float* row, int j, int D, float4 v) {
  if constexpr (VEC) { stg4(row + 4 * j, v); return; }
  const int e = 4 * j;
  if (e < D) row[e] = v.x;
  if (e + 1 < D) row[e + 1] = v.y;
  if (e + 2 < D) row[e + 2] = v.z;
  if (e + 3 < D) row[e + 3] = v.w;
}

template <int G, bool VEC>
__global__ void __launch_bounds__(kBlock) combine_score_kernel(const ScoreArgs a) {
  constexpr int RPB = kBlock / G;
  __shared__ float red[kBlock / 32];
  const int gid = threadIdx.x / G, t = threadIdx.x % G;
  const int D = a.D, nchunk = (D + 3) >> 2;
  const int64_t ngroups = (a.B + RPB - 1) / RPB;
  const float g0 = a.gout ? __ldg(a.gout) : 1.f;
  for (int64_t rg = blockIdx.x; rg < ngroups; rg += gridDim.x) {
    const int64_t b = rg * RPB + gid;
    const bool active = b < a.B;
    float acc = 0.f, w = 1.f;
    if (active) {
      if (a.sigma) {
        const float sg = a.labels ? __ldg(a.sigma + __ldg(a.labels + b)) : __ldg(a.sigma + b);
        w = (a.power == 2.f) ? sg * sg : (a.power == 1.f ? sg : powf(sg, a.power));
      }
      const float gw = g0 * w / (float)a.B;
      const float* s1 = a.s + b * a.lds;
      const float* s2 = a.s + (a.B + b) * a.lds;
      const float* vr = a.v + b * a.ldv;
      float* d1 = a.ds ? a.ds + b * a.ldds : nullptr;
      float* d2 = a.ds ? a.ds + (a.B + b) * a.ldds : nullptr;
      // 2-deep software unroll: independent 128-bit loads in flight per thread
      for (int j = t; j < nchunk; j += G) {
        const float4 p = ld4<VEC>(s1, j, D), q = ld4<VEC>(s2, j, D), v = ld4<VEC>(vr, j, D);
        float4 o1, o2;
#define FDSM_SCORE_ELT(c)                                                         \
  {                                                                               \
    const float g1 = p.c + q.c, g2 = p.c - q.c;                                   \
    if (a.projected) {                                                            \
      const float gv = g1 * v.c;                                                  \
      acc += gv * gv * 0.125f + g2 * v.c * 0.5f;                                  \
      const float t1 = gv * v.c * 0.25f;                                          \
      o1.c = gw * (t1 + 0.5f * v.c);                                              \
      o2.c = gw * (t1 - 0.5f * v.c);                                              \
    } else {                                                                      \
      acc += g1 * g1 * 0.125f + g2 * v.c * a.k;                                   \
      const float t1 = g1 * 0.25f, t2 = a.k * v.c;                                \
      o1.c = gw * (t1 + t2);                                                      \
      o2.c = gw * (t1 - t2);                                                      \
    }                                                                             \
  }
        FDSM_SCORE_ELT(x) FDSM_SCORE_ELT(y) FDSM_SCORE_ELT(z) FDSM_SCORE_ELT(w)
#undef FDSM_SCORE_ELT
        if (d1) {
          st4<VEC>(d1, j, D, o1);
          st4<VEC>(d2, j, D, o2);
        }
      }
    }
    if (a.per_sample) {  // uniform
      acc = group_sum<G>(acc, red);
      if (active && t == 0) a.per_sample[b] = acc * w;
    }
  }
}

__global__ void __launch_bounds__(1024) mean_kernel(const float* __restrict__ per, int64_t B, float* __restrict__ out) {
  __shared__ float red[32];
  float acc = 0.f;
  for (int64_t b = threadIdx.x; b < B; b += blockDim.x) acc += per[b];
  acc = block_sum_1024(acc, red);
  if (threadIdx.x == 0) *out = acc / (float)B;
}

// data *= *g unless *g == 1 (the usual root-loss case: every block reads the scalar and leaves -- no traffic)
__global__ void __launch_bounds__(256) scale_unless_one_kernel(float* __restrict__ data, int64_t n, const float* __restrict__ g) {
  const float c = __ldg(g);
  if (c == 1.f) return;
  const int64_t n4 = aligned16(data) ? n >> 2 : 0;
  float4* d4 = reinterpret_cast<float4*>(data);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 v = d4[i];
    v.x *= c; v.y *= c; v.z *= c; v.w *= c;
    d4[i] = v;
  }
  for (int64_t i = 4 * n4 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) data[i] *= c;
}

template <int G>
void launch_score(const ScoreArgs& a, bool vec, int grid, cudaStream_t st) {
  if (vec) combine_score_kernel<G, true><<<grid, kBlock, 0, st>>>(a);
  else combine_score_kernel<G, false><<<grid, kBlock, 0, st>>>(a);
}

}  // namespace
}  // namespace fdsm

using namespace fdsm;

extern "C" int fdsm_combine_energy_fwd(const float* E, int64_t B, int32_t mode, float c1, float scale, int32_t mean,
                                       float* per_sample, float* loss, fdsm_stream_t stream) {
  FDSM_REQUIRE(B >= 0, "fdsm_combine_energy_fwd: B < 0");
  FDSM_REQUIRE(mode == 0 || mode == 1, "fdsm_combine_energy_fwd: mode must be 0 or 1");
  FDSM_REQUIRE(B == 0 || E, "fdsm_combine_energy_fwd: E is NULL");
  FDSM_REQUIRE(per_sample || loss, "fdsm_combine_energy_fwd: nothing to write");
  combine_energy_fwd_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(E, B, mode, c1, scale, mean, per_sample, loss);
  FDSM_LAUNCH_CHECK("fdsm_combine_energy_fwd");
}

extern "C" int fdsm_combine_energy_bwd(const float* E, int64_t B, int32_t mode, float c1, float scale, int32_t mean,
                                       const float* gout, float* dE, fdsm_stream_t stream) {
  FDSM_REQUIRE(B >= 0, "fdsm_combine_energy_bwd: B < 0");
  FDSM_REQUIRE(mode == 0 || mode == 1, "fdsm_combine_energy_bwd: mode must be 0 or 1");
  if (B == 0) return FDSM_OK;
  FDSM_REQUIRE(E && dE, "fdsm_combine_energy_bwd: NULL pointer");
  const float coef = mean ? scale / (float)B : scale;
  const int grid = (int)((B + 255) / 256 < 1184 ? (B + 255) / 256 : 1184);
  combine_energy_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(E, B, mode, c1, coef, gout, dE);
  FDSM_LAUNCH_CHECK("fdsm_combine_energy_bwd");
}

extern "C" int fdsm_combine_dsm_fwd(const float* E, const float* dot, const float* sigma_b, int64_t B, float alpha,
                                    float beta, float gamma, float* per_sample, float* loss, fdsm_stream_t stream) {
  FDSM_REQUIRE(B >= 0, "fdsm_combine_dsm_fwd: B < 0");
  FDSM_REQUIRE(B == 0 || (E && dot), "fdsm_combine_dsm_fwd: NULL pointer");
  FDSM_REQUIRE(per_sample || loss, "fdsm_combine_dsm_fwd: nothing to write");
  combine_dsm_fwd_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(E, dot, sigma_b, B, alpha, beta, gamma, per_sample, loss);
  FDSM_LAUNCH_CHECK("fdsm_combine_dsm_fwd");
}

extern "C" int fdsm_combine_dsm_bwd(const float* E, const float* dot, const float* sigma_b, int64_t B, float alpha,
                                    float beta, float gamma, const float* gout, float* dE, fdsm_stream_t stream) {
  FDSM_REQUIRE(B >= 0, "fdsm_combine_dsm_bwd: B < 0");
  if (B == 0) return FDSM_OK;
  FDSM_REQUIRE(E && dot && dE, "fdsm_combine_dsm_bwd: NULL pointer");
  const int grid = (int)((B + 255) / 256 < 1184 ? (B + 255) / 256 : 1184);
  combine_dsm_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(E, dot, sigma_b, B, alpha, beta, gamma, gout, dE);
  FDSM_LAUNCH_CHECK("fdsm_combine_dsm_bwd");
}

extern "C" int fdsm_combine_score(const float* s, int64_t lds, const float* v, int64_t ldv, const float* sigma,
                                  const int64_t* labels, float power, int64_t B, int64_t D, float k, int32_t projected,
                                  const float* gout, float* per_sample, float* loss, float* ds, int64_t ldds,
                                  fdsm_stream_t stream) {
  FDSM_REQUIRE(B >= 0 && D >= 0 && D < (1ll << 31), "fdsm_combine_score: bad shape");
  FDSM_REQUIRE(per_sample || ds, "fdsm_combine_score: nothing to write (loss needs per_sample as workspace)");
  FDSM_REQUIRE(!loss || per_sample, "fdsm_combine_score: loss needs the per_sample workspace");
  FDSM_REQUIRE(!labels || sigma, "fdsm_combine_score: labels without sigma table");
  cudaStream_t st = (cudaStream_t)stream;
  if (B > 0 && D > 0) {
    FDSM_REQUIRE(s && v, "fdsm_combine_score: NULL input");
    FDSM_REQUIRE(lds >= D && ldv >= D && (!ds || ldds >= D), "fdsm_combine_score: leading dimension < D");
    ScoreArgs a{s, lds, v, ldv, sigma, labels, power, B, (int)D, k, projected, gout, per_sample, ds, ldds};
    const bool vec = D % 4 == 0 && lds % 4 == 0 && ldv % 4 == 0 && aligned16(s) && aligned16(v) &&
                     (!ds || (aligned16(ds) && ldds % 4 == 0));
    const int64_t nchunk = (D + 3) / 4;
    int G = 32;
    while (G < kBlock && (int64_t)G * 4 < nchunk) G <<= 1;
    const int64_t ngroups = (B + (kBlock / G) - 1) / (kBlock / G);
    const int64_t cap = (int64_t)sm_count() * 8;
    const int grid = (int)(ngroups < cap ? ngroups : cap);
    switch (G) {
      case 32: launch_score<32>(a, vec, grid, st); break;
      case 64: launch_score<64>(a, vec, grid, st); break;
      case 128: launch_score<128>(a, vec, grid, st); break;
      default: launch_score<256>(a, vec, grid, st); break;
    }
    int rc = cuda_status(cudaGetLastError(), "fdsm_combine_score");
    if (rc) return rc;
  }
  if (loss) {
    mean_kernel<<<1, 1024, 0, st>>>(per_sample, B, loss);
    FDSM_LAUNCH_CHECK("fdsm_combine_score(mean)");
  }
  return FDSM_OK;
}

// ---- no-allocation variant of the score-form combine (include/fdsm.h: fdsm_combine_score_static) ---------------------
// The per-sample terms that the mean kernel sums are the kernel's scratch; here they live in a file-scope device array
// (one copy per GPU that loads the library, held until exit) instead of a caller buffer.  In host code the array's NAME
// is a host-side stand-in for the symbol, not a device address: the address is resolved once per device, with that
// device current, by cudaGetSymbolAddress and cached.
static __device__ __align__(256) float g_score_static_scratch[FDSM_COMBINE_SCORE_STATIC_MAX_B];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
float* g_static_ptr[kMaxDevices] = {};

int static_scratch(float** out) {
  int dev = -1;
  int rc = fdsm::cuda_status(cudaGetDevice(&dev), "fdsm_combine_score_static(cudaGetDevice)");
  if (rc) return rc;
  if (dev < 0 || dev >= kMaxDevices) {
    fdsm::set_error("fdsm_combine_score_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
    return FDSM_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[dev]) {
    void* p = nullptr;
    rc = fdsm::cuda_status(cudaGetSymbolAddress(&p, g_score_static_scratch), "fdsm_combine_score_static(cudaGetSymbolAddress)");
    if (rc) return rc;
    if (!p) {
      fdsm::set_error("fdsm_combine_score_static: cudaGetSymbolAddress returned a null address");
      return FDSM_ERR_UNSUPPORTED;
    }
    g_static_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_static_ptr[dev];
  return FDSM_OK;
}
}  // namespace

extern "C" int fdsm_combine_score_static(const float* s, int64_t lds, const float* v, int64_t ldv, const float* sigma,
                                         const int64_t* labels, float power, int64_t B, int64_t D, float k,
                                         int32_t projected, const float* gout, float* loss, float* ds, int64_t ldds,
                                         fdsm_stream_t stream) {
  FDSM_REQUIRE(B >= 0 && D >= 0 && D < (1ll << 31), "fdsm_combine_score_static: bad shape");
  FDSM_REQUIRE(B <= FDSM_COMBINE_SCORE_STATIC_MAX_B,
               "fdsm_combine_score_static: B = %lld exceeds the static scratch (FDSM_COMBINE_SCORE_STATIC_MAX_B = %d samples); "
               "use fdsm_combine_score with a caller-owned per_sample buffer",
               (long long)B, FDSM_COMBINE_SCORE_STATIC_MAX_B);
  FDSM_REQUIRE(!labels || sigma, "fdsm_combine_score_static: labels without sigma table");
  float* per = nullptr;
  int rc = static_scratch(&per);      // set-up = the first call on this device (B = 0 will do: warm up before a capture)
  if (rc) return rc;
  FDSM_REQUIRE(loss, "fdsm_combine_score_static: loss is NULL");
  cudaStream_t st = (cudaStream_t)stream;
  if (B > 0 && D > 0) {
    FDSM_REQUIRE(s && v, "fdsm_combine_score_static: NULL input");
    FDSM_REQUIRE(lds >= D && ldv >= D && (!ds || ldds >= D), "fdsm_combine_score_static: leading dimension < D");
    ScoreArgs a{s, lds, v, ldv, sigma, labels, power, B, (int)D, k, projected, gout, per, ds, ldds};
    const bool vec = D % 4 == 0 && lds % 4 == 0 && ldv % 4 == 0 && aligned16(s) && aligned16(v) &&
                     (!ds || (aligned16(ds) && ldds % 4 == 0));
    const int64_t nchunk = (D + 3) / 4;
    int G = 32;
    while (G < kBlock && (int64_t)G * 4 < nchunk) G <<= 1;
    const int64_t ngroups = (B + (kBlock / G) - 1) / (kBlock / G);
    const int64_t cap = (int64_t)sm_count() * 8;
    const int grid = (int)(ngroups < cap ? ngroups : cap);
    switch (G) {
      case 32: launch_score<32>(a, vec, grid, st); break;
      case 64: launch_score<64>(a, vec, grid, st); break;
      case 128: launch_score<128>(a, vec, grid, st); break;
      default: launch_score<256>(a, vec, grid, st); break;
    }
    rc = cuda_status(cudaGetLastError(), "fdsm_combine_score_static");
    if (rc) return rc;
  } else if (B > 0) {
    // D == 0: every per-sample term is an empty sum, but the scratch keeps the previous call's values: clear what is read
    rc = cuda_status(cudaMemsetAsync(per, 0, (size_t)B * sizeof(float), st), "fdsm_combine_score_static(memset)");
    if (rc) return rc;
  }
  mean_kernel<<<1, 1024, 0, st>>>(per, B, loss);
  FDSM_LAUNCH_CHECK("fdsm_combine_score_static(mean)");
}

extern "C" int fdsm_scale_unless_one(float* data, int64_t n, const float* g, fdsm_stream_t stream) {
  FDSM_REQUIRE(n >= 0, "fdsm_scale_unless_one: n < 0");
  if (n == 0) return FDSM_OK;
  FDSM_REQUIRE(data && g, "fdsm_scale_unless_one: NULL pointer");
  const int64_t want = (n / 4 + 255) / 256 + 1, cap = (int64_t)sm_count() * 8;
  scale_unless_one_kernel<<<(int)(want < cap ? want : cap), 256, 0, (cudaStream_t)stream>>>(data, n, g);
  FDSM_LAUNCH_CHECK("fdsm_scale_unless_one");
}
