// the elementwise pieces either side of the hot path, on the same primitives.
//   fdsm_langevin_step: one annealed-Langevin update  x <- x + step_size * score + sqrt(2 step_size) * z
//     (ncsnv2/models/__init__.py:34-39, ncsn/runners/anneal_runner.py sampling loop, DeepEBM/ESM/sampling_mdsm.py:63-96),
//     z injected or drawn in-kernel from the Philox noise stream.  HBM-bound: R x, R score (8D) + W x (4D) = 12*D B/sample
//     (the reference: randn_like + 2 multiplies + 2 adds = 5 launches, 40*D B/sample).
//   fdsm_ema_update: shadow <- (1 - mu) * p + mu * shadow  (ncsnv2/models/ema.py:16-21), one pass, 12 B/parameter.
#include "fdsm_common.cuh"

namespace fdsm {
namespace {

constexpr int kBlock = 256;

template <bool VEC>
__global__ void __launch_bounds__(kBlock) langevin_kernel(const float* __restrict__ x, long long ldx, const float* __restrict__ g,
                                                          long long ldg, const float* __restrict__ z, long long ldz,
                                                          float* __restrict__ out, long long ldo, long long B, int D,
                                                          float step_size, float noise_scale, unsigned long long seed,
                                                          unsigned long long offset, unsigned long long row0) {
  const int nchunk = (D + 3) >> 2;
  const long long total = B * nchunk;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / nchunk;
    const int j = (int)(i - b * nchunk), e = 4 * j;
    float4 xv, gv, zv;
    if constexpr (VEC) {
      xv = ldg4(x + b * ldx + e);
      gv = ldg4(g + b * ldg + e);
      zv = z ? ldg4(z + b * ldz + e) : philox_normal4(seed, offset, STREAM_NOISE, row0 + (unsigned long long)b, (uint32_t)j);
    } else {
      auto ld = [&](const float* p, int k) { return e + k < D ? __ldg(p + e + k) : 0.f; };
      xv = make_float4(ld(x + b * ldx, 0), ld(x + b * ldx, 1), ld(x + b * ldx, 2), ld(x + b * ldx, 3));
      gv = make_float4(ld(g + b * ldg, 0), ld(g + b * ldg, 1), ld(g + b * ldg, 2), ld(g + b * ldg, 3));
      if (z) zv = make_float4(ld(z + b * ldz, 0), ld(z + b * ldz, 1), ld(z + b * ldz, 2), ld(z + b * ldz, 3));
      else zv = philox_normal4(seed, offset, STREAM_NOISE, row0 + (unsigned long long)b, (uint32_t)j);
    }
    // (x + step_size * grad) + noise * sqrt(2 step_size): the reference's four roundings, in its order
    auto upd = [&](float xx, float gg, float zz) {
      return __fadd_rn(__fadd_rn(xx, __fmul_rn(step_size, gg)), __fmul_rn(zz, noise_scale));
    };
    const float4 r = make_float4(upd(xv.x, gv.x, zv.x), upd(xv.y, gv.y, zv.y), upd(xv.z, gv.z, zv.z), upd(xv.w, gv.w, zv.w));
    float* o = out + b * ldo + e;
    if constexpr (VEC) {
      stg4(o, r);
    } else {
      if (e < D) o[0] = r.x;
      if (e + 1 < D) o[1] = r.y;
      if (e + 2 < D) o[2] = r.z;
      if (e + 3 < D) o[3] = r.w;
    }
  }
}

__global__ void __launch_bounds__(kBlock) ema_kernel(float* __restrict__ shadow, const float* __restrict__ p, long long n, float om, float mu) {
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  // (1 - mu) * p + mu * shadow, the reference's rounding order (two products, one sum)
  if (aligned16(shadow) && aligned16(p)) {
    const long long nv = n >> 2;
    for (long long i = tid; i < nv; i += nth) {
      float4 s = reinterpret_cast<float4*>(shadow)[i];
      const float4 q = ldg4(p + 4 * i);
      s.x = __fadd_rn(__fmul_rn(om, q.x), __fmul_rn(mu, s.x)); s.y = __fadd_rn(__fmul_rn(om, q.y), __fmul_rn(mu, s.y));
      s.z = __fadd_rn(__fmul_rn(om, q.z), __fmul_rn(mu, s.z)); s.w = __fadd_rn(__fmul_rn(om, q.w), __fmul_rn(mu, s.w));
      reinterpret_cast<float4*>(shadow)[i] = s;
    }
    for (long long i = (nv << 2) + tid; i < n; i += nth) shadow[i] = __fadd_rn(__fmul_rn(om, p[i]), __fmul_rn(mu, shadow[i]));
  } else {
    for (long long i = tid; i < n; i += nth) shadow[i] = __fadd_rn(__fmul_rn(om, p[i]), __fmul_rn(mu, shadow[i]));
  }
}

}  // namespace
}  // namespace fdsm

using namespace fdsm;

extern "C" int fdsm_langevin_step(const float* x, int64_t ldx, const float* score, int64_t lds, const float* noise,
                                  int64_t ldnoise, float* out, int64_t ldo, int64_t B, int64_t D, float step_size,
                                  float noise_scale, uint64_t seed, uint64_t offset, uint64_t row0, fdsm_stream_t stream) {
  FDSM_REQUIRE(B >= 0 && D >= 0 && D < (1ll << 31), "fdsm_langevin_step: bad shape");
  if (B == 0 || D == 0) return FDSM_OK;
  FDSM_REQUIRE(x && score && out, "fdsm_langevin_step: NULL pointer");
  FDSM_REQUIRE(ldx >= D && lds >= D && ldo >= D && (!noise || ldnoise >= D), "fdsm_langevin_step: leading dimension < D");
  auto ok = [&](const float* p, int64_t ld) { return !p || (aligned16(p) && ld % 4 == 0); };
  const bool vec = D % 4 == 0 && ok(x, ldx) && ok(score, lds) && ok(noise, ldnoise) && ok(out, ldo);
  const int64_t total = B * ((D + 3) / 4), nb = (total + kBlock - 1) / kBlock, cap = (int64_t)sm_count() * 8;
  const int grid = (int)(nb < cap ? nb : cap);
  cudaStream_t st = (cudaStream_t)stream;
  if (vec) langevin_kernel<true><<<grid, kBlock, 0, st>>>(x, ldx, score, lds, noise, ldnoise, out, ldo, B, (int)D, step_size,
                                                          noise_scale, seed, offset, row0);
  else langevin_kernel<false><<<grid, kBlock, 0, st>>>(x, ldx, score, lds, noise, ldnoise, out, ldo, B, (int)D, step_size,
                                                       noise_scale, seed, offset, row0);
  FDSM_LAUNCH_CHECK("fdsm_langevin_step");
}

extern "C" int fdsm_ema_update(float* shadow, const float* params, int64_t n, float one_minus_mu, float mu, fdsm_stream_t stream) {
  FDSM_REQUIRE(n >= 0, "fdsm_ema_update: n < 0");
  if (n == 0) return FDSM_OK;
  FDSM_REQUIRE(shadow && params, "fdsm_ema_update: NULL pointer");
  const int64_t nb = (n / 4 + kBlock - 1) / kBlock + 1, cap = (int64_t)sm_count() * 8;
  ema_kernel<<<(int)(nb < cap ? nb : cap), kBlock, 0, (cudaStream_t)stream>>>(shadow, params, n, one_minus_mu, mu);
  FDSM_LAUNCH_CHECK("fdsm_ema_update");
}
