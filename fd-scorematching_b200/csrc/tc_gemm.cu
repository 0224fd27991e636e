// C ABI of the tensor-core building blocks: TF32 hi/lo operand split and the generic tcgen05 GEMM.
#include "tc_gemm_host.cuh"

namespace fdsm {
namespace {

// out_hi/out_lo[r, c] = split(in[r, c]) for c < cols ; column `cols` := ones_col ? 1 : 0 ; columns (cols, out_cols) := 0
__global__ void __launch_bounds__(256) tf32_split_kernel(const float* __restrict__ in, long long ldin, long long rows, int cols,
                                                         float* __restrict__ out_hi, float* __restrict__ out_lo,
                                                         long long ldo, int out_cols, int ones_col) {
  const long long total = rows * out_cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / out_cols;
    const int c = (int)(i - r * out_cols);
    float x = 0.f;
    if (c < cols) x = __ldg(in + r * ldin + c);
    else if (c == cols && (ones_col & 1)) x = 1.f;
    float hi, lo;
    ptx::tf32_split(x, hi, lo);
    if (ones_col & 2) hi = x;   // raw-operand format: padded fp32 copy
    out_hi[r * ldo + c] = hi;
    if (out_lo) out_lo[r * ldo + c] = lo;
  }
}

}  // namespace
}  // namespace fdsm

using namespace fdsm;

extern "C" int fdsm_tf32_split(const float* in, int64_t ldin, int64_t rows, int64_t cols, float* out_hi, float* out_lo,
                               int64_t ldo, int64_t out_cols, int32_t ones_col, fdsm_stream_t stream) {
  FDSM_REQUIRE(rows >= 0 && cols >= 0 && out_cols >= cols + (ones_col & 1), "fdsm_tf32_split: bad shape");
  FDSM_REQUIRE(!(ones_col & 2) || !out_lo, "fdsm_tf32_split: the raw (unrounded) format has no lo half");
  FDSM_REQUIRE(ldo >= out_cols && (cols == 0 || ldin >= cols), "fdsm_tf32_split: leading dimension too small");
  if (rows == 0 || out_cols == 0) return FDSM_OK;
  FDSM_REQUIRE((in || cols == 0) && out_hi, "fdsm_tf32_split: NULL pointer");
  const long long total = rows * out_cols;
  const long long nb = (total + 255) / 256;
  const long long cap = (long long)sm_count() * 8;
  tf32_split_kernel<<<(int)(nb < cap ? nb : cap), 256, 0, (cudaStream_t)stream>>>(in, ldin, rows, (int)cols, out_hi, out_lo,
                                                                                   ldo, (int)out_cols, ones_col);
  FDSM_LAUNCH_CHECK("fdsm_tf32_split");
}

extern "C" int fdsm_tc_gemm(const fdsm_gemm_args* g, fdsm_stream_t stream) {
  FDSM_REQUIRE(g != nullptr, "fdsm_tc_gemm: args is NULL");
  FDSM_REQUIRE(g->M >= 0 && g->N >= 0 && g->K >= 0 && g->M < (1ll << 31) && g->N < (1ll << 31) && g->K < (1ll << 31),
               "fdsm_tc_gemm: bad sizes");
  FDSM_REQUIRE(g->out != nullptr, "fdsm_tc_gemm: out is NULL");
  FDSM_REQUIRE((g->a_major | 1) == 1 && (g->b_major | 1) == 1, "fdsm_tc_gemm: major must be 0 (K) or 1 (MN)");
  if (g->M == 0 || g->N == 0) return FDSM_OK;
  FDSM_REQUIRE(g->K > 0, "fdsm_tc_gemm: K must be > 0");
  FDSM_REQUIRE(g->ldo >= g->N, "fdsm_tc_gemm: ldo < N");
  tc::GemmParams p = {};
  p.epi = tc::EPI_STORE;
  p.act = tc::ACT_NONE;
  p.out_hi = g->out;
  p.ldo = g->ldo;
  p.split_stride = g->split_stride;
  int BN = g->tile_n > 0 ? g->tile_n : (int)((g->N + 15) / 16 * 16);
  if (BN > 256) BN = 256;
  tc::Operand A{g->a_hi, g->a_lo, g->lda, g->a_major}, B{g->b_hi, g->b_lo, g->ldb, g->b_major};
  return tc::launch_gemm(p, A, B, (int)g->M, (int)g->N, (int)g->K, BN, g->passes, g->splits > 0 ? g->splits : 1,
                         (cudaStream_t)stream);
}

/* Debug only (not in include/fdsm.h): per-CTA phase timestamps of the raw-operand GEMM, see scripts/gemm_trace.py. */
extern "C" void fdsm_debug_set_gemm_trace(unsigned long long* device_buffer) {
  tc::trace_buffer() = device_buffer;   // launch i (since this call) writes at device_buffer + i * 8 * 160
  tc::trace_launch_index() = 0;
}
