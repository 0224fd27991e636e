// Shared device/host helpers for libfdsm (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/fdsm.h"

namespace fdsm {

// ---- error plumbing (thread-local message, int return codes; include/fdsm.h) -------------
void set_error(const char* fmt, ...);
int cuda_status(cudaError_t e, const char* what);

#define FDSM_REQUIRE(cond, ...)            \
  do {                                     \
    if (!(cond)) {                         \
      ::fdsm::set_error(__VA_ARGS__);      \
      return FDSM_ERR_INVALID_ARGUMENT;    \
    }                                      \
  } while (0)

#define FDSM_LAUNCH_CHECK(what) return ::fdsm::cuda_status(cudaGetLastError(), what)

__host__ __device__ static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
int sm_count();

// ---- programmatic dependent launch (PDL) for the small kernels of the step's critical chain: a kernel launched through
// launch_pdl may become resident while its stream predecessor drains and must execute pdl_wait() before it touches
// anything the predecessor wrote; pdl_launch_dependents() lets the NEXT kernel of the stream (if it was launched with
// the attribute) do the same.  Both instructions are no-ops in a grid launched without the attribute.
// FDSM_GEMM_PDL=0 switches every programmatic launch of the library off. ----
inline bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("FDSM_GEMM_PDL"); return !(e && e[0] == '0'); }();
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

// ---- Philox4x32-10 (Salmon et al. 2011), counter-based --------------------------------------
// counter = (chunk index within the row, global row lo, global row hi | stream<<28, draw offset)
// key     = (seed lo, seed hi)
// One call yields 4 uint32 -> 4 N(0,1) through two Box-Muller pairs.  Keyed by GLOBAL row so a
// batch sharded over ranks draws exactly what the un-sharded batch would.
struct Philox {
  static constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  static constexpr uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  __host__ __device__ static inline void round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
    // one 32x32->64 multiply per lane pair (IMAD.WIDE.U32) instead of separate mulhi + mullo
    const uint64_t p0 = (uint64_t)M0 * c[0], p1 = (uint64_t)M1 * c[2];
    const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0, hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
    c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
  }
  __host__ __device__ static inline void gen(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      round(c, k0, k1);
      k0 += W0; k1 += W1;
    }
  }
};

enum : uint32_t { STREAM_DIRECTION = 0, STREAM_NOISE = 1 };

#ifdef __CUDACC__
#ifndef FDSM_FAST_BOXMULLER
#define FDSM_FAST_BOXMULLER 1
#endif
// sqrt(-2 ln u), u in (0, 1)
__device__ __forceinline__ float box_muller_radius(float u) {
  const float w = 1.0f - u;                                                          // exact for u >= 0.5
  const float q_series = w * (2.0f + w * (1.0f + w * (0.6666667f + w * 0.5f)));      // -2 ln(1 - w), w < 1/64: rel 2e-9
  const float q_mufu = -1.3862943611198906f * __log2f(u);                            // -2 ln2 log2(u)
  const float q = w < 0.015625f ? q_series : q_mufu;
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(q));
  return r;
}

// 4 standard normals for (row, chunk) of a stream.  u in (0,1): (24-bit mantissa + 0.5) * 2^-24.
__device__ __forceinline__ float4 philox_normal4(uint64_t seed, uint64_t offset, uint32_t stream, uint64_t row,
                                                 uint32_t chunk) {
  uint32_t c[4] = {chunk, (uint32_t)row, ((uint32_t)(row >> 32) & 0x0FFFFFFFu) | (stream << 28),
                   (uint32_t)offset ^ (uint32_t)(offset >> 32) * 0x9E3779B9u};
  Philox::gen(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  const float k = 5.9604644775390625e-8f;  // 2^-24
  float u0 = ((float)(c[0] >> 8) + 0.5f) * k, u1 = ((float)(c[1] >> 8) + 0.5f) * k;
  float u2 = ((float)(c[2] >> 8) + 0.5f) * k, u3 = ((float)(c[3] >> 8) + 0.5f) * k;
#if FDSM_FAST_BOXMULLER
  // Box-Muller on the MUFU pipe.  The libdevice forms (logf, sqrtf, sincospif) cost ~330 instructions per 4 normals
  // and made the stack kernel issue-bound at 34% of HBM bandwidth; these cost ~40 and stay within 1e-5 (absolute)
  // of the float64 oracle: lg2.approx loses RELATIVE accuracy near 1, so u > 63/64 takes the log1p series on the
  // exactly representable w = 1 - u; sin/cos.approx get an exactly reduced angle in (-pi, pi].
  float s0, c0, s1, c1;
  const float r0 = box_muller_radius(u0), r1 = box_muller_radius(u2);
  __sincosf(6.283185307179586f * (u1 > 0.5f ? u1 - 1.0f : u1), &s0, &c0);
  __sincosf(6.283185307179586f * (u3 > 0.5f ? u3 - 1.0f : u3), &s1, &c1);
#else
  float r0 = sqrtf(-2.0f * logf(u0)), r1 = sqrtf(-2.0f * logf(u2));
  float s0, c0, s1, c1;
  sincospif(2.0f * u1, &s0, &c0);
  sincospif(2.0f * u3, &s1, &c1);
#endif
  return make_float4(r0 * c0, r0 * s0, r1 * c1, r1 * s1);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum over a group of G consecutive threads (G power of two, 32 <= G <= blockDim.x, group
// aligned to warps).  `red` = shared scratch of blockDim.x/32 floats.  All threads of the block
// must call it (contains __syncthreads when G > 32).  Result broadcast to every thread of the group.
template <int G>
__device__ __forceinline__ float group_sum(float v, float* red) {
  v = warp_sum(v);
  if constexpr (G > 32) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();  // protect `red` reuse across consecutive calls
    if (lane == 0) red[warp] = v;
    __syncthreads();
    constexpr int WPG = G / 32;
    const int g0 = (warp / WPG) * WPG;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < WPG; ++i) s += red[g0 + i];  // fixed order -> deterministic
    v = s;
  }
  return v;
}

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
// streaming store: the stacked batch is consumed by the next kernel out of L2, never re-read here
__device__ __forceinline__ void stg4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
#endif

}  // namespace fdsm
