// Kernels (1)+(2): Philox/injected direction sampling, per-sample L2 normalisation and the fused
// build of the stacked batch [x, x+v, x-v] / [xt+v, xt-v].  HBM-bound streaming kernel:
// algorithmic traffic 16*D B/sample (energy form, Philox) or 12*D (noised 2-row form); one pass
// over x, the direction lives in registers between the norm reduction and the stores.
//
// Mapping: a group of G threads (G = 32..256, power of two) owns one sample row; a 256-thread block
// owns 256/G rows.  Each thread keeps <= MAXC float4 chunks of the direction in registers; rows
// longer than 4*MAXC*256 elements take the re-generate path (Philox is counter-based, so pass 2
// recomputes exactly the numbers pass 1 reduced).
#include "fdsm_common.cuh"
#include "tc_ptx.cuh"

namespace fdsm {
namespace {

constexpr int kBlock = 256;
constexpr int MAXC = 4;

template <bool VEC>
__device__ __forceinline__ float4 load4(const float* row, int j, int D) {
  if constexpr (VEC) {
    return ldg4(row + 4 * j);
  } else {
    const int e = 4 * j;
    float4 r;
    r.x = e + 0 < D ? __ldg(row + e + 0) : 0.f;
    r.y = e + 1 < D ? __ldg(row + e + 1) : 0.f;
    r.z = e + 2 < D ? __ldg(row + e + 2) : 0.f;
    r.w = e + 3 < D ? __ldg(row + e + 3) : 0.f;
    return r;
  }
}

template <bool VEC>
__device__ __forceinline__ void store4(float* row, int j, int D, float4 v) {
  if constexpr (VEC) {
    stg4(row + 4 * j, v);
  } else {
    const int e = 4 * j;
    if (e + 0 < D) row[e + 0] = v.x;
    if (e + 1 < D) row[e + 1] = v.y;
    if (e + 2 < D) row[e + 2] = v.z;
    if (e + 3 < D) row[e + 3] = v.w;
  }
}

__device__ __forceinline__ float4 mask_tail(float4 v, int j, int D) {
  const int e = 4 * j;
  if (e + 1 >= D) v.y = 0.f;
  if (e + 2 >= D) v.z = 0.f;
  if (e + 3 >= D) v.w = 0.f;
  return v;
}

template <bool VEC>
__device__ __forceinline__ float4 raw_direction(const fdsm_stack_args& a, int64_t b, int j) {
  if (a.dir) return load4<VEC>(a.dir + b * a.lddir, j, (int)a.D);
  const float4 n = philox_normal4(a.seed, a.offset, STREAM_DIRECTION, a.row0 + (uint64_t)b, (uint32_t)j);
  if constexpr (VEC) return n;          // the vector path requires D % 4 == 0: no partial chunk
  return mask_tail(n, j, (int)a.D);
}

template <bool VEC>
__device__ __forceinline__ float4 raw_noise(const fdsm_stack_args& a, int64_t b, int j) {
  if (a.noise) return load4<VEC>(a.noise + b * a.ldnoise, j, (int)a.D);
  const float4 n = philox_normal4(a.seed, a.offset, STREAM_NOISE, a.row0 + (uint64_t)b, (uint32_t)j);
  if constexpr (VEC) return n;
  return mask_tail(n, j, (int)a.D);
}

// one stacked row chunk: plain fp32, or TF32 hi/lo halves when the batch feeds the tensor-core MLP
template <bool VEC>
__device__ __forceinline__ void store_row(const fdsm_stack_args& a, int64_t r, int j, int D, float4 v) {
  if (a.out_lo) {
    float4 hi, lo;
    ptx::tf32_split(v.x, hi.x, lo.x); ptx::tf32_split(v.y, hi.y, lo.y);
    ptx::tf32_split(v.z, hi.z, lo.z); ptx::tf32_split(v.w, hi.w, lo.w);
    store4<VEC>(a.out + r * a.ldout, j, D, hi);
    store4<VEC>(a.out_lo + r * a.ldout, j, D, lo);
  } else {
    store4<VEC>(a.out + r * a.ldout, j, D, v);
  }
}

__device__ __forceinline__ float sumsq(float4 d) { return d.x * d.x + d.y * d.y + d.z * d.z + d.w * d.w; }

// v = d / nrm * scale with the reference's rounding sequence (division, then multiply).
__device__ __forceinline__ float4 normalise(float4 d, float nrm, float scale) {
  return make_float4(__fmul_rn(__fdiv_rn(d.x, nrm), scale), __fmul_rn(__fdiv_rn(d.y, nrm), scale),
                     __fmul_rn(__fdiv_rn(d.z, nrm), scale), __fmul_rn(__fdiv_rn(d.w, nrm), scale));
}

// MODE: compile-time specialisations of the two production configurations; the runtime mode tests below fold away and
// the row loop becomes straight-line code the compiler can interleave across the chunks (config-2 launch: 17.2 -> 11.5 us).
// Same arithmetic in the same order as the general instantiation (MODE 0): the outputs are bit-identical.
//   1 = energy-form training step: Philox direction, rows [x | x+v | x-v], no noise / injected draws / hi-lo split /
//       v / dot outputs;
//   2 = noised 2-row form (a3/a9/a11/a13 in production mode): Philox direction AND Philox noise, sigma per sample or
//       through a label-indexed table, rows [xt+v | xt-v]; v / dot outputs, perturb_input, dot_mode stay runtime.
constexpr int MODE_GENERAL = 0, MODE_ENERGY3 = 1, MODE_NOISED2 = 2;
template <int G, bool VEC, bool CACHE, int MODE = MODE_GENERAL>
__global__ void __launch_bounds__(kBlock) stack_kernel(fdsm_stack_args a) {
  if constexpr (MODE == MODE_ENERGY3) {
    a.dir = nullptr; a.noise = nullptr; a.sigma = nullptr; a.labels = nullptr; a.out_lo = nullptr; a.v_out = nullptr;
    a.dot_out = nullptr; a.row_mask = 7; a.normalize_dir = 1;
  }
  if constexpr (MODE == MODE_NOISED2) {
    a.dir = nullptr; a.noise = nullptr; a.out_lo = nullptr; a.row_mask = 6; a.normalize_dir = 1;
    __builtin_assume(a.sigma != nullptr);
    __builtin_assume(a.x != nullptr);
  }
  pdl_launch_dependents();   // the GEMM that consumes the stacked rows sets up (barriers, TMEM) while this grid runs
  if (a.offset_dev) a.offset += __ldg(reinterpret_cast<const unsigned long long*>(a.offset_dev));
  constexpr int RPB = kBlock / G;
  __shared__ float red[kBlock / 32];
  const int gid = threadIdx.x / G, t = threadIdx.x % G;
  const int D = (int)a.D;
  const int nchunk = (D + 3) >> 2;
  const int64_t ngroups = (a.B + RPB - 1) / RPB;
  const int nrows = __popc(a.row_mask & 7);
  const bool need_norm = (a.dir == nullptr) || a.normalize_dir;

  for (int64_t rg = blockIdx.x; rg < ngroups; rg += gridDim.x) {
    const int64_t b = rg * RPB + gid;
    const bool active = b < a.B;

    // ---- pass 1: direction chunks -> registers, sum of squares -> group reduction -------------
    float4 d[MAXC];
    float4 xb[MAXC];   // CACHE: the sample row is requested BEFORE the Philox/Box-Muller arithmetic so that the HBM
                       // latency hides under it (the kernel was alternating compute-only and memory-only phases)
    float ss = 0.f;
    if (active) {
      if constexpr (CACHE) {
        if (a.x) {
          const float* xrow0 = a.x + b * a.ldx;
#pragma unroll
          for (int c = 0; c < MAXC; ++c) {
            const int j = t + c * G;
            xb[c] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (j < nchunk) xb[c] = load4<VEC>(xrow0, j, D);
          }
        }
#pragma unroll
        for (int c = 0; c < MAXC; ++c) {
          const int j = t + c * G;
          d[c] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (j < nchunk) {
            d[c] = raw_direction<VEC>(a, b, j);
            ss += sumsq(d[c]);
          }
        }
      } else if (need_norm) {
        for (int j = t; j < nchunk; j += G) ss += sumsq(raw_direction<VEC>(a, b, j));
      }
    }
    float nrm = 1.f;
    if (need_norm) nrm = sqrtf(group_sum<G>(ss, red));  // uniform branch: all threads take it
    // Injected directions (parity with the reference's torch ops) keep its rounding sequence u / ||u|| * eps with IEEE
    // division per element; for in-kernel Philox draws there is no reference bit pattern to match, so the row uses one
    // division (eps / ||u||) and a multiply per element: ~40 fewer instructions per float4 (the kernel is issue-bound).
#ifndef FDSM_FAST_NORMALISE
#define FDSM_FAST_NORMALISE 1
#endif
    const bool exact_div = !FDSM_FAST_NORMALISE || a.dir != nullptr;
    const float inv = need_norm && !exact_div ? __fdiv_rn(a.dir_scale, nrm) : 0.f;

    // ---- pass 2: build rows -----------------------------------------------------------------
    float dot = 0.f;
    if (active) {
      float sig = 0.f;
      if (a.sigma) sig = a.labels ? __ldg(a.sigma + __ldg(a.labels + b)) : __ldg(a.sigma + b);
      const float* xrow = a.x ? a.x + b * a.ldx : nullptr;
      auto emit = [&](int j, float4 dc, float4 base) {
        const float4 v = !need_norm ? dc
                         : exact_div ? normalise(dc, nrm, a.dir_scale)
                                     : make_float4(dc.x * inv, dc.y * inv, dc.z * inv, dc.w * inv);
        if (a.sigma) {
          const float4 n = raw_noise<VEC>(a, b, j);
          const float4 sn = make_float4(__fmul_rn(sig, n.x), __fmul_rn(sig, n.y), __fmul_rn(sig, n.z),
                                        __fmul_rn(sig, n.w));
          const float4 xn = make_float4(__fadd_rn(base.x, sn.x), __fadd_rn(base.y, sn.y), __fadd_rn(base.z, sn.z),
                                        __fadd_rn(base.w, sn.w));
          if (a.dot_mode == 1) {
            dot += v.x * __fsub_rn(xn.x, base.x) + v.y * __fsub_rn(xn.y, base.y) + v.z * __fsub_rn(xn.z, base.z) +
                   v.w * __fsub_rn(xn.w, base.w);
          } else {
            dot += v.x * sn.x + v.y * sn.y + v.z * sn.z + v.w * sn.w;
          }
          if (a.perturb_input) base = xn;
        }
        int r = 0;
        if (a.row_mask & FDSM_ROW_X) {
          store_row<VEC>(a, (int64_t)r * a.B + b, j, D, base);
          ++r;
        }
        if (a.row_mask & FDSM_ROW_PLUS) {
          store_row<VEC>(a, (int64_t)r * a.B + b, j, D,
                         make_float4(__fadd_rn(base.x, v.x), __fadd_rn(base.y, v.y), __fadd_rn(base.z, v.z),
                                     __fadd_rn(base.w, v.w)));
          ++r;
        }
        if (a.row_mask & FDSM_ROW_MINUS) {
          store_row<VEC>(a, (int64_t)r * a.B + b, j, D,
                         make_float4(__fsub_rn(base.x, v.x), __fsub_rn(base.y, v.y), __fsub_rn(base.z, v.z),
                                     __fsub_rn(base.w, v.w)));
        }
        if (a.v_out) store4<VEC>(a.v_out + b * a.ldv, j, D, v);
      };
      if constexpr (CACHE) {
#pragma unroll
        for (int c = 0; c < MAXC; ++c) {
          const int j = t + c * G;
          if (j < nchunk) emit(j, d[c], xrow ? xb[c] : make_float4(0.f, 0.f, 0.f, 0.f));
        }
      } else {
        for (int j = t; j < nchunk; j += G)
          emit(j, raw_direction<VEC>(a, b, j), xrow ? load4<VEC>(xrow, j, D) : make_float4(0.f, 0.f, 0.f, 0.f));
      }
    }
    if ((a.out_lo || a.pad_ones) && active) {   // GEMM padding columns [D, ldout): ones column then zeros
      for (int r = 0; r < nrows; ++r) {
        const int64_t row = (int64_t)r * a.B + b;
        for (int c = D + t; c < (int)a.ldout; c += G) {
          a.out[row * a.ldout + c] = c == D ? 1.f : 0.f;
          if (a.out_lo) a.out_lo[row * a.ldout + c] = 0.f;
        }
      }
    }
    if (a.dot_out) {  // uniform branch
      dot = group_sum<G>(dot, red);
      if (active && t == 0) a.dot_out[b] = dot;
    }
  }
}

template <bool VEC>
__global__ void __launch_bounds__(kBlock) philox_normal_kernel(float* out, int64_t ld, int64_t B, int D, uint32_t stream_id,
                                                              uint64_t seed, uint64_t offset, uint64_t row0) {
  const int nchunk = (D + 3) >> 2;
  const int64_t total = B * nchunk;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / nchunk;
    const int j = (int)(i - b * nchunk);
    store4<VEC>(out + b * ld, j, D, philox_normal4(seed, offset, stream_id, row0 + (uint64_t)b, (uint32_t)j));
  }
}

template <int G>
int launch_g(const fdsm_stack_args& a, bool vec, bool cache, int grid, cudaStream_t st) {
  // FDSM_STACK_SPECIALISE=0 (read once) routes everything through the general instantiation: A/B timing, bisecting
  static const bool specialise = [] { const char* e = getenv("FDSM_STACK_SPECIALISE"); return !(e && e[0] == '0'); }();
  if (specialise && vec && cache && !a.dir && !a.out_lo && a.x) {
    if (!a.sigma && !a.v_out && !a.dot_out && a.row_mask == 7) {
      stack_kernel<G, true, true, MODE_ENERGY3><<<grid, kBlock, 0, st>>>(a);
      FDSM_LAUNCH_CHECK("fdsm_stack");
    }
    if (a.sigma && !a.noise && a.row_mask == 6) {
      stack_kernel<G, true, true, MODE_NOISED2><<<grid, kBlock, 0, st>>>(a);
      FDSM_LAUNCH_CHECK("fdsm_stack");
    }
  }
  if (vec) {
    if (cache) stack_kernel<G, true, true><<<grid, kBlock, 0, st>>>(a);
    else stack_kernel<G, true, false><<<grid, kBlock, 0, st>>>(a);
  } else {
    if (cache) stack_kernel<G, false, true><<<grid, kBlock, 0, st>>>(a);
    else stack_kernel<G, false, false><<<grid, kBlock, 0, st>>>(a);
  }
  FDSM_LAUNCH_CHECK("fdsm_stack");
}

}  // namespace
}  // namespace fdsm

extern "C" int fdsm_stack(const fdsm_stack_args* args, fdsm_stream_t stream) {
  using namespace fdsm;
  FDSM_REQUIRE(args != nullptr, "fdsm_stack: args is NULL");
  const fdsm_stack_args& a = *args;
  FDSM_REQUIRE(a.B >= 0 && a.D >= 0, "fdsm_stack: negative size (B=%lld, D=%lld)", (long long)a.B, (long long)a.D);
  FDSM_REQUIRE(a.D < (1ll << 31), "fdsm_stack: D=%lld too large", (long long)a.D);
  FDSM_REQUIRE((a.row_mask & ~7) == 0, "fdsm_stack: bad row_mask %d", a.row_mask);
  if (a.B == 0 || a.D == 0) return FDSM_OK;  // empty batch: nothing to do (torch.cat of empties)
  const int rows = __builtin_popcount(a.row_mask);
  FDSM_REQUIRE(rows == 0 || (a.x && a.out), "fdsm_stack: x/out must be non-NULL when row_mask != 0");
  FDSM_REQUIRE(rows > 0 || a.v_out || a.dot_out, "fdsm_stack: nothing to write");
  FDSM_REQUIRE(!a.x || a.ldx >= a.D, "fdsm_stack: ldx < D");
  FDSM_REQUIRE(!a.out || a.ldout >= a.D, "fdsm_stack: ldout < D");
  FDSM_REQUIRE(!a.dir || a.lddir >= a.D, "fdsm_stack: lddir < D");
  FDSM_REQUIRE(!a.v_out || a.ldv >= a.D, "fdsm_stack: ldv < D");
  FDSM_REQUIRE(!(a.sigma && a.noise) || a.ldnoise >= a.D, "fdsm_stack: ldnoise < D");
  FDSM_REQUIRE(!a.dot_out || a.sigma, "fdsm_stack: dot_out needs sigma (DSM noise)");
  FDSM_REQUIRE(!a.labels || a.sigma, "fdsm_stack: labels given without a sigma table");
  FDSM_REQUIRE(!(a.out_lo || a.pad_ones) || (a.out && a.ldout > a.D),
               "fdsm_stack: out_lo / pad_ones need out and ldout > D (room for the ones column)");

  bool vec = (a.D % 4 == 0);
  auto ok = [&](const float* p, int64_t ld) { return !p || (aligned16(p) && ld % 4 == 0); };
  vec = vec && ok(a.x, a.ldx) && ok(a.dir, a.lddir) && ok(a.out, a.ldout) && ok(a.out_lo, a.ldout) && ok(a.v_out, a.ldv) &&
        (!a.sigma || ok(a.noise, a.ldnoise));

  const int64_t nchunk = (a.D + 3) / 4;
  int G = 32;
  while (G < kBlock && (int64_t)G * MAXC < nchunk) G <<= 1;
  const bool cache = (int64_t)G * MAXC >= nchunk;
  const int64_t ngroups = (a.B + (kBlock / G) - 1) / (kBlock / G);
  const int64_t cap = (int64_t)sm_count() * 8;
  const int grid = (int)(ngroups < cap ? ngroups : cap);
  cudaStream_t st = (cudaStream_t)stream;
  switch (G) {
    case 32: return launch_g<32>(a, vec, cache, grid, st);
    case 64: return launch_g<64>(a, vec, cache, grid, st);
    case 128: return launch_g<128>(a, vec, cache, grid, st);
    default: return launch_g<256>(a, vec, cache, grid, st);
  }
}

extern "C" int fdsm_direction(const float* dir, int64_t lddir, float* v, int64_t ldv, int64_t B, int64_t D,
                              float scale, uint64_t seed, uint64_t offset, uint64_t row0, fdsm_stream_t stream) {
  FDSM_REQUIRE(v != nullptr, "fdsm_direction: v is NULL");
  fdsm_stack_args a = {};
  a.dir = dir; a.lddir = lddir; a.normalize_dir = 1; a.dir_scale = scale;
  a.v_out = v; a.ldv = ldv; a.row_mask = 0;
  a.seed = seed; a.offset = offset; a.row0 = row0; a.B = B; a.D = D;
  return fdsm_stack(&a, stream);
}

extern "C" int fdsm_philox_normal(float* out, int64_t ld, int64_t B, int64_t D, int32_t stream_id, uint64_t seed,
                                  uint64_t offset, uint64_t row0, fdsm_stream_t stream) {
  using namespace fdsm;
  FDSM_REQUIRE(out != nullptr, "fdsm_philox_normal: out is NULL");
  FDSM_REQUIRE(B >= 0 && D >= 0 && D < (1ll << 31) && ld >= D, "fdsm_philox_normal: bad shape");
  FDSM_REQUIRE(stream_id == 0 || stream_id == 1, "fdsm_philox_normal: stream_id must be 0 or 1");
  if (B == 0 || D == 0) return FDSM_OK;
  const bool vec = (D % 4 == 0) && (ld % 4 == 0) && aligned16(out);
  const int64_t total = B * ((D + 3) / 4);
  const int64_t cap = (int64_t)sm_count() * 8;
  const int64_t nb = (total + kBlock - 1) / kBlock;
  const int grid = (int)(nb < cap ? nb : cap);
  cudaStream_t st = (cudaStream_t)stream;
  if (vec) philox_normal_kernel<true><<<grid, kBlock, 0, st>>>(out, ld, B, (int)D, (uint32_t)stream_id, seed, offset, row0);
  else philox_normal_kernel<false><<<grid, kBlock, 0, st>>>(out, ld, B, (int)D, (uint32_t)stream_id, seed, offset, row0);
  FDSM_LAUNCH_CHECK("fdsm_philox_normal");
}
