// Device helpers of kernel (4), energy form, shared by fd_combine.cu and the fused head kernels of fd_mlp.cu
// (which fold the combine forward/backward into the MLP backward: same arithmetic, same reduction order).
#pragma once
#include "fdsm_common.cuh"

namespace fdsm {

__device__ __forceinline__ float block_sum_1024(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float s = 0.f;
  const int nw = blockDim.x >> 5;
  for (int i = 0; i < nw; ++i) s += red[i];
  return s;
}

__device__ __forceinline__ float energy_per(const float* E, int64_t B, int64_t b, int mode, float c1) {
  if (mode == 0) {
    const float e1 = E[b], e2 = E[B + b], e3 = E[2 * B + b];
    const float d = e2 - e3;
    // (d*d)/8 - e2 - e3 + c1*e1 in the reference's left-to-right order (efficient_sm.py:65-67)
    return __fadd_rn(__fmul_rn(__fmul_rn(d, d), 0.125f), __fadd_rn(__fsub_rn(-e2, e3), __fmul_rn(c1, e1)));
  }
  const float d = E[b] - E[B + b];
  return __fadd_rn(__fmul_rn(d, d), __fmul_rn(4.f, d));
}


// The whole forward of the energy-form combine for ONE 1024-thread block (fixed-order reduction -> deterministic).
__device__ __forceinline__ void combine_energy_fwd_block(const float* __restrict__ E, int64_t B, int mode, float c1,
                                                         float scale, int mean, float* __restrict__ per_sample,
                                                         float* __restrict__ loss, float* red) {
  float acc = 0.f;
  for (int64_t b = threadIdx.x; b < B; b += blockDim.x) {
    const float p = energy_per(E, B, b, mode, c1);
    if (per_sample) per_sample[b] = p * scale;
    acc += p;
  }
  acc = block_sum_1024(acc, red);
  if (threadIdx.x == 0 && loss) *loss = (mean ? acc / (float)B : acc) * scale;
}

// dL/dE of row r of the stacked energies (mode 0: [E1 | E2 | E3], coef = scale/B or scale)
__device__ __forceinline__ float combine_energy_grad_row(const float* __restrict__ E, int64_t B, float c1, float coef, int64_t r) {
  const int which = (r >= B) + (r >= 2 * B);   // rows = 3B: no 64-bit division (it cost ~80 instructions per row)
  const int64_t b = r - which * B;
  if (which == 0) return coef * c1;
  const float q = (__ldg(E + B + b) - __ldg(E + 2 * B + b)) * 0.25f;
  return which == 1 ? coef * (q - 1.f) : coef * (-q - 1.f);
}

// ---- DSM energy form (E = energy([xt+v | xt-v]), 2B rows): r_b = alpha (E[B+b] - E[b]) + beta dot_b,
// per_b = r_b^2 / sigma_b^2, loss = gamma mean_b per_b.  Shared by fd_combine.cu and the fused head kernels of fd_mlp.cu. ----
__device__ __forceinline__ void combine_dsm_fwd_block(const float* __restrict__ E, const float* __restrict__ dot,
                                                      const float* __restrict__ sigma_b, int64_t B, float alpha, float beta,
                                                      float gamma, float* __restrict__ per_sample, float* __restrict__ loss,
                                                      float* red) {
  float acc = 0.f;
  for (int64_t b = threadIdx.x; b < B; b += blockDim.x) {
    const float r = alpha * (E[B + b] - E[b]) + beta * dot[b];
    float p = r * r;
    if (sigma_b) {
      const float s = sigma_b[b];
      p = p / (s * s);
    }
    if (per_sample) per_sample[b] = p * gamma;
    acc += p;
  }
  acc = block_sum_1024(acc, red);
  if (threadIdx.x == 0 && loss) *loss = acc / (float)B * gamma;
}

// dL/dE of sample b's SECOND row (row B + b); the first row (row b) gets the negative.  c = 2 alpha gamma / B (x gout).
__device__ __forceinline__ float combine_dsm_grad(const float* __restrict__ E, const float* __restrict__ dot,
                                                  const float* __restrict__ sigma_b, int64_t B, float alpha, float beta,
                                                  float c, int64_t b) {
  const float r = alpha * (__ldg(E + B + b) - __ldg(E + b)) + beta * __ldg(dot + b);
  float g = c * r;
  if (sigma_b) {
    const float s = __ldg(sigma_b + b);
    g = g / (s * s);
  }
  return g;
}

}  // namespace fdsm
