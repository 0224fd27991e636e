// Host-loop pieces of the training step on the flat parameter vector (DeepEBM/train_ssm.py:65-70):
// clip_grad_norm_ (global L2 norm, coefficient clip/(norm+1e-6) clamped to 1) and the Adam update, as two
// launches with every scalar that changes per step (step count, learning rate) read from device memory, so the
// whole step replays from one CUDA graph.
#include "fdsm_common.cuh"

namespace fdsm {
namespace {

// FDSM_SUMSQ_PARTIALS blocks, each over a fixed contiguous chunk, each with a fixed summation order: the partials (and
// their fixed-order sum in adam_kernel) are bit-identical across ranks and runs.
constexpr int kPartials = FDSM_SUMSQ_PARTIALS;
__global__ void __launch_bounds__(1024) sumsq_kernel(const float* __restrict__ g, long long n, float* __restrict__ out,
                                                     long long* __restrict__ step) {
  __shared__ float red[32];
  pdl_launch_dependents();
  pdl_wait();
  const long long chunk = ((n + kPartials - 1) / kPartials + 3) & ~3ll;
  const long long lo = blockIdx.x * chunk, hi = lo + chunk < n ? lo + chunk : n;
  float acc = 0.f;
  if (aligned16(g)) {
    const long long nv = hi > lo ? (hi - lo) >> 2 : 0;
    const float4* g4 = reinterpret_cast<const float4*>(g + lo);
    for (long long i = threadIdx.x; i < nv; i += blockDim.x) {
      const float4 v = __ldg(g4 + i);
      acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    for (long long i = lo + (nv << 2) + threadIdx.x; i < hi; i += blockDim.x) acc += g[i] * g[i];
  } else {
    for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) acc += g[i] * g[i];
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += red[i];
    out[blockIdx.x] = s;
    if (step && blockIdx.x == 0) *step += 1;
  }
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, long long n, float lr, const float* __restrict__ lr_dev,
                                                   float b1, float b2, float eps, float max_norm,
                                                   const float* __restrict__ sumsq, const long long* __restrict__ step) {
  pdl_wait();
  const float t = (float)__ldg(step);
  const float lrv = lr_dev ? __ldg(lr_dev) : lr;
  float coef = 1.f;
  if (max_norm > 0.f) {
    float ss = 0.f;
    for (int i = 0; i < kPartials; ++i) ss += __ldg(sumsq + i);
    coef = fminf(max_norm / (sqrtf(ss) + 1e-6f), 1.f);
  }
  const float bc1 = 1.f - powf(b1, t), bc2s = sqrtf(1.f - powf(b2, t));
  const float step_size = lrv / bc1;
  auto upd = [&](float& pi, float gi, float& mi, float& vi) {
    gi *= coef;
    mi = b1 * mi + (1.f - b1) * gi;
    vi = b2 * vi + (1.f - b2) * gi * gi;
    pi -= step_size * (mi / (sqrtf(vi) / bc2s + eps));
  };
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  if (aligned16(p) && aligned16(g) && aligned16(m) && aligned16(v)) {   // 128-bit path + scalar tail
    const long long nv = n >> 2;
    for (long long i = tid; i < nv; i += nth) {
      float4 p4 = reinterpret_cast<float4*>(p)[i], m4 = reinterpret_cast<float4*>(m)[i], v4 = reinterpret_cast<float4*>(v)[i];
      const float4 g4 = reinterpret_cast<const float4*>(g)[i];
      upd(p4.x, g4.x, m4.x, v4.x); upd(p4.y, g4.y, m4.y, v4.y); upd(p4.z, g4.z, m4.z, v4.z); upd(p4.w, g4.w, m4.w, v4.w);
      reinterpret_cast<float4*>(p)[i] = p4; reinterpret_cast<float4*>(m)[i] = m4; reinterpret_cast<float4*>(v)[i] = v4;
    }
    for (long long i = (nv << 2) + tid; i < n; i += nth) upd(p[i], g[i], m[i], v[i]);
  } else {
    for (long long i = tid; i < n; i += nth) upd(p[i], g[i], m[i], v[i]);
  }
}

// The step's result leaves for the host as a plain store into mapped pinned memory: a copy-engine D2H of 4 bytes between two
// graph replays idles the stream on both sides of it (fdsm_publish_scalars).
__global__ void publish_scalars_kernel(const float* __restrict__ src, volatile float* h_dst, int n) {
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) h_dst[i] = __ldcg(src + i);
  __threadfence_system();
}

}  // namespace
}  // namespace fdsm

using namespace fdsm;

// Host plumbing of the end-to-end step in ONE call each (the step loop is host-bound on a busy host long before it is
// link- or device-bound when every stream / event operation is a separate Python call): the batch's way in ...
extern "C" int fdsm_stage_h2d(void* dst, const void* h_src, int64_t nbytes, fdsm_stream_t copy_stream, void* ev_free, void* ev_ready,
                              fdsm_stream_t main_stream) {
  FDSM_REQUIRE(nbytes >= 0 && (nbytes == 0 || (dst && h_src)) && ev_ready, "fdsm_stage_h2d: bad arguments");
  cudaStream_t cs = (cudaStream_t)copy_stream, ms = (cudaStream_t)main_stream;
  int rc;
  if (ev_free && (rc = cuda_status(cudaStreamWaitEvent(cs, (cudaEvent_t)ev_free, 0), "fdsm_stage_h2d(wait)"))) return rc;
  if (nbytes && (rc = cuda_status(cudaMemcpyAsync(dst, h_src, (size_t)nbytes, cudaMemcpyHostToDevice, cs), "fdsm_stage_h2d(copy)"))) return rc;
  if ((rc = cuda_status(cudaEventRecord((cudaEvent_t)ev_ready, cs), "fdsm_stage_h2d(record)"))) return rc;
  return cuda_status(cudaStreamWaitEvent(ms, (cudaEvent_t)ev_ready, 0), "fdsm_stage_h2d(join)");
}

extern "C" int fdsm_publish_scalars(const float* src, float* h_dst, int64_t n, fdsm_stream_t stream);
// ... and the result's way out: `ev` (may be NULL) is recorded on `stream` first -- the consumer of the staged batch is done
extern "C" int fdsm_publish_scalars_record(const float* src, float* h_dst, int64_t n, fdsm_stream_t stream, void* ev) {
  if (ev) {
    const int rc = cuda_status(cudaEventRecord((cudaEvent_t)ev, (cudaStream_t)stream), "fdsm_publish_scalars_record");
    if (rc) return rc;
  }
  return fdsm_publish_scalars(src, h_dst, n, stream);
}

extern "C" int fdsm_publish_scalars(const float* src, float* h_dst, int64_t n, fdsm_stream_t stream) {
  FDSM_REQUIRE(n >= 0 && n <= 4096, "fdsm_publish_scalars: n must be in [0, 4096] (a step's scalars, not a tensor copy)");
  if (n == 0) return FDSM_OK;
  FDSM_REQUIRE(src && h_dst, "fdsm_publish_scalars: NULL pointer");
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, h_dst);
  if (e != cudaSuccess || at.type != cudaMemoryTypeHost || at.devicePointer == nullptr) {
    (void)cudaGetLastError();
    set_error("fdsm_publish_scalars: h_dst must be pinned (page-locked, device-mapped) host memory");
    return FDSM_ERR_INVALID_ARGUMENT;
  }
  return cuda_status(launch_pdl(publish_scalars_kernel, dim3((unsigned)((n + 127) / 128)), dim3(128), 0, (cudaStream_t)stream, src,
                                (volatile float*)at.devicePointer, (int)n), "fdsm_publish_scalars");
}

extern "C" int fdsm_grad_sumsq(const float* g, int64_t n, float* sumsq, int64_t* step_counter, fdsm_stream_t stream) {
  FDSM_REQUIRE(n >= 0 && sumsq && (g || n == 0), "fdsm_grad_sumsq: bad arguments");
  return cuda_status(launch_pdl(sumsq_kernel, dim3(kPartials), dim3(256), 0, (cudaStream_t)stream, g, (long long)n, sumsq,
                                (long long*)step_counter), "fdsm_grad_sumsq");
}

extern "C" int fdsm_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                              const float* lr_dev, float beta1, float beta2, float eps, float max_norm, const float* sumsq,
                              const int64_t* step_counter, fdsm_stream_t stream) {
  FDSM_REQUIRE(n >= 0, "fdsm_adam_step: n < 0");
  if (n == 0) return FDSM_OK;
  FDSM_REQUIRE(params && grads && exp_avg && exp_avg_sq && step_counter, "fdsm_adam_step: NULL pointer");
  FDSM_REQUIRE(max_norm <= 0.f || sumsq, "fdsm_adam_step: clipping needs the squared gradient norm");
  const long long nb = (n / 4 + 255) / 256 + 1;
  const long long cap = (long long)sm_count() * 8;
  return cuda_status(launch_pdl(adam_kernel, dim3((unsigned)(nb < cap ? nb : cap)), dim3(256), 0, (cudaStream_t)stream, params, grads,
                                exp_avg, exp_avg_sq, (long long)n, lr, lr_dev, beta1, beta2, eps, max_norm, sumsq,
                                (const long long*)step_counter), "fdsm_adam_step");
}
