// Host side of the tcgen05 GEMM: TMA tensor-map encoding and launch.
#pragma once
#include <cudaTypedefs.h>

#include <cstdlib>
#include <mutex>

#include "tc_gemm_raw.cuh"

namespace fdsm {
namespace tc {

// cuTensorMapEncodeTiled through the runtime's driver-entry-point lookup (no -lcuda link dependency)
inline PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  });
  return fn;
}

// Row-major fp32 matrix [rows x cols] with leading dimension ld (elements); box = {32 cols, box_rows}; 128-byte swizzle;
// out-of-bounds elements read as zero.
// `mn_major` operands (contraction index = row) use the 32-byte-atom flavour of the 128-byte swizzle: the only
// shared-memory layout tcgen05 accepts for MN-major 32-bit operands.
inline int make_tmap(CUtensorMap* m, const float* ptr, long long rows, long long cols, long long ld, int box_rows,
                     bool mn_major, const char* what) {
  auto fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return FDSM_ERR_UNSUPPORTED;
  }
  FDSM_REQUIRE(ptr && aligned16(ptr), "%s: operand pointer must be non-NULL and 16-byte aligned", what);
  FDSM_REQUIRE(ld % 4 == 0 && ld >= cols, "%s: leading dimension %lld must be a multiple of 4 and >= %lld", what, ld, cols);
  FDSM_REQUIRE(box_rows >= 1 && box_rows <= 256, "%s: box rows %d out of range", what, box_rows);
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
  cuuint32_t box[2] = {32, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("%s: cuTensorMapEncodeTiled failed (CUresult %d)", what, (int)r);
    return FDSM_ERR_INVALID_ARGUMENT;
  }
  return FDSM_OK;
}

// General tiled map over fp32 data: dims / box innermost first (dims[0] contiguous), strides in BYTES for dims 1.. (need not be
// monotonic: a later dimension may step by less than an earlier one -- e.g. the 32-column panels of a row-major matrix as
// dimension 2 with a 128-byte stride, so that ONE box covers several panels in the [panel][row][32] order tcgen05 wants).
inline int make_tmap_nd(CUtensorMap* m, const void* ptr, int rank, const long long* dims, const long long* strides_bytes,
                        const int* box, bool mn_major, const char* what, bool f16 = false) {
  auto fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return FDSM_ERR_UNSUPPORTED;
  }
  FDSM_REQUIRE(ptr && aligned16(ptr), "%s: operand pointer must be non-NULL and 16-byte aligned", what);
  FDSM_REQUIRE(rank >= 2 && rank <= 5, "%s: rank %d out of range", what, rank);
  cuuint64_t d[5]; cuuint64_t st[4]; cuuint32_t bx[5]; cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    FDSM_REQUIRE(dims[i] >= 1 && box[i] >= 1 && box[i] <= 256, "%s: dimension %d: size %lld, box %d", what, i, dims[i], box[i]);
    d[i] = (cuuint64_t)dims[i]; bx[i] = (cuuint32_t)box[i]; es[i] = 1;
    if (i) {
      FDSM_REQUIRE(strides_bytes[i - 1] > 0 && strides_bytes[i - 1] % 16 == 0, "%s: stride %lld of dimension %d", what, strides_bytes[i - 1], i);
      st[i - 1] = (cuuint64_t)strides_bytes[i - 1];
    }
  }
  CUresult r = fn(m, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(ptr), d, st, bx,
                  es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("%s: cuTensorMapEncodeTiled (rank %d) failed (CUresult %d)", what, rank, (int)r);
    return FDSM_ERR_INVALID_ARGUMENT;
  }
  return FDSM_OK;
}

// debug hook (scripts/gemm_trace.py): when set, every raw-operand launch writes 8 %globaltimer stamps per CTA
inline unsigned long long*& trace_buffer() {
  static unsigned long long* buf = nullptr;
  return buf;
}
constexpr int TRACE_STRIDE = 8 * 160;   // stamps per launch slot (<= 160 CTAs traced)
inline int& trace_launch_index() {
  static int idx = 0;
  return idx;
}

// CTA pairs (cta_group::2) are OPT-IN (FDSM_GEMM_PAIR=1 in the environment, read once).  Measured on B200
// (profiles/r01_gemm_pair_vs_single.txt): correct, and 7% faster on a deep-K 4096^3 problem, but every kind::tf32 MMA is
// only K=8 deep and the paired instruction carries a fixed cross-SM cost: the config-2 shapes (25 / 7 k-blocks per CTA)
// run at ~1.3 us per k-block against 0.96 us for the single-CTA kernel, so the single-CTA kernel stays the default.
inline bool pair_mode() {
  static const bool on = [] {
    const char* e = getenv("FDSM_GEMM_PAIR");
    return e && e[0] == '1';
  }();
  return on;
}

// FDSM_GEMM_PDL=0 in the environment launches the GEMMs without programmatic stream serialization (A/B measurements)
inline bool pdl_mode() {
  static const bool on = [] {
    const char* e = getenv("FDSM_GEMM_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

inline int lo_slots_override() {
  static const int v = [] {
    const char* e = getenv("FDSM_GEMM_LO_SLOTS");
    const int x = e ? atoi(e) : 0;
    return x >= 1 && x <= MAX_LO_SLOTS ? x : 0;
  }();
  return v;
}

struct Operand {
  const float* hi; const float* lo; long long ld; int major;   // MAJOR_K: [R x K] row-major; MAJOR_MN: [K x R] row-major
};

// Fills the tensor maps + sizes of `p` and launches.  `p` must already carry the epilogue fields.
inline int launch_gemm(GemmParams& p, const Operand& A, const Operand& B, int M, int N, int K, int BN, int passes,
                       int splits, cudaStream_t st) {
  FDSM_REQUIRE(passes == 1 || passes == 3, "tc_gemm: passes must be 1 or 3");
  FDSM_REQUIRE(M > 0 && N > 0 && K > 0, "tc_gemm: empty problem (M=%d N=%d K=%d)", M, N, K);
  FDSM_REQUIRE(BN % 16 == 0 && BN >= 16 && BN <= 256, "tc_gemm: BN=%d must be a multiple of 16 in [16, 256]", BN);
  const bool raw = passes == 3 && !A.lo && !B.lo;   // plain fp32 operands: hi/lo split inside the kernel
  FDSM_REQUIRE(passes == 1 || raw || (A.lo && B.lo),
               "tc_gemm: 3-pass mode needs the lo halves of both operands, or of neither (in-kernel split)");
  const int kblocks = (K + BK - 1) / BK;
  FDSM_REQUIRE(splits >= 1 && splits <= kblocks, "tc_gemm: splits=%d must be in [1, %d]", splits, kblocks);
  p.M = M; p.N = N; p.K = K; p.BN = BN; p.a_major = A.major; p.b_major = B.major; p.passes = passes;
  int rc;
  for (int h = 0; h < (passes == 3 && !raw ? 2 : 1); ++h) {
    CUtensorMap* ta = h ? &p.ta_lo : &p.ta_hi;
    CUtensorMap* tb = h ? &p.tb_lo : &p.tb_hi;
    if (A.major == MAJOR_K) rc = make_tmap(ta, h ? A.lo : A.hi, M, K, A.ld, BM, false, "tc_gemm A");
    else rc = make_tmap(ta, h ? A.lo : A.hi, K, M, A.ld, BK, true, "tc_gemm A");
    if (rc) return rc;
    if (B.major == MAJOR_K) rc = make_tmap(tb, h ? B.lo : B.hi, N, K, B.ld, BN, false, "tc_gemm B");
    else rc = make_tmap(tb, h ? B.lo : B.hi, K, N, B.ld, BK, true, "tc_gemm B");
    if (rc) return rc;
  }
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, splits);
  if (raw) {
    static std::once_flag once_raw;
    std::call_once(once_raw, [] {
#define FDSM_SET_SMEM_RAW(E)                                                                                                 \
  cudaFuncSetAttribute(gemm_raw_kernel<E, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 5 * 1024);       \
  cudaFuncSetAttribute(gemm_raw_kernel<E, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 5 * 1024)
      FDSM_SET_SMEM_RAW(EPI_STORE); FDSM_SET_SMEM_RAW(EPI_ACT_SPLIT); FDSM_SET_SMEM_RAW(EPI_HEAD); FDSM_SET_SMEM_RAW(EPI_DACT_SPLIT);
#undef FDSM_SET_SMEM_RAW
    });
    // CTA pairs (cta_group::2) whenever there are at least two M tiles; an odd tile count is padded by one CTA whose rows
    // are all out of range (TMA zero fill, nothing stored)
    const bool pair = pair_mode() && grid.x >= 2;
    if (pair) {
      grid.x = (grid.x + 1) & ~1u;
      if (B.major == MAJOR_K) {   // this CTA stages BN/2 rows of B: rebuild the map with the half-height box
        rc = make_tmap(&p.tb_hi, B.hi, N, K, B.ld, BN / 2, false, "tc_gemm B");
        if (rc) return rc;
      }
    }
    p.trace = trace_buffer() ? trace_buffer() + (size_t)TRACE_STRIDE * trace_launch_index()++ : nullptr;
    p.slot_bytes = raw_slot_bytes(pair ? BN / 2 : BN, B.major);
    // ring depths: a CTA pair hands every k-block across two SMs (remote arrive -> MMA -> multicast commit), so it wants
    // a deeper lo ring (3) than the single CTA (2); the raw ring takes what is left (>= 2: TMA latency)
    const int budget = 227 * 1024 - 5 * 1024;
    p.lo_slots = lo_slots_override() > 0 ? lo_slots_override() : (pair ? 3 : 2);
    p.raw_slots = MAX_RAW_SLOTS;
    while (p.raw_slots > 1 && raw_smem_total(p.raw_slots, p.lo_slots, p.slot_bytes) > budget) --p.raw_slots;
    while (p.raw_slots < 2 && p.lo_slots > 2) {
      --p.lo_slots;
      p.raw_slots = MAX_RAW_SLOTS;
      while (p.raw_slots > 1 && raw_smem_total(p.raw_slots, p.lo_slots, p.slot_bytes) > budget) --p.raw_slots;
    }
    const int smem_raw = raw_smem_total(p.raw_slots, p.lo_slots, p.slot_bytes);
    // the epilogue stages the 128 x BN output tile in the pipeline's shared memory
    FDSM_REQUIRE(BM * (((BN + CW - 1) / CW) * CW + 4) * 4 + (EPI_PARTS - 1) * BM * 3 * 4 <= (p.raw_slots + p.lo_slots) * p.slot_bytes,
                 "tc_gemm: output staging tile does not fit the pipeline's shared memory (BN=%d)", BN);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = smem_raw; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = pair ? 2 : 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    // programmatic dependent launch: this grid's CTAs may start (barrier init, TMEM allocation) while the previous kernel
    // of the stream drains; the kernel executes griddepcontrol.wait before it touches global memory
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl_mode() ? 2 : 1;
    cudaError_t e;
#define FDSM_LAUNCH_RAW(E) \
  e = pair ? cudaLaunchKernelEx(&cfg, gemm_raw_kernel<E, true>, p) : cudaLaunchKernelEx(&cfg, gemm_raw_kernel<E, false>, p)
    switch (p.epi) {
      case EPI_STORE: FDSM_LAUNCH_RAW(EPI_STORE); break;
      case EPI_ACT_SPLIT: FDSM_LAUNCH_RAW(EPI_ACT_SPLIT); break;
      case EPI_HEAD: FDSM_LAUNCH_RAW(EPI_HEAD); break;
      default: FDSM_LAUNCH_RAW(EPI_DACT_SPLIT); break;
    }
#undef FDSM_LAUNCH_RAW
    return cuda_status(e, "tc_gemm(raw)");
  }
  const int smem = SmemLayout::total(passes);
  static std::once_flag once;
  std::call_once(once, [] {
#define FDSM_SET_SMEM(P, E) cudaFuncSetAttribute(gemm_kernel<P, E>, cudaFuncAttributeMaxDynamicSharedMemorySize, SmemLayout::total(P))
    FDSM_SET_SMEM(1, EPI_STORE); FDSM_SET_SMEM(1, EPI_ACT_SPLIT); FDSM_SET_SMEM(1, EPI_HEAD); FDSM_SET_SMEM(1, EPI_DACT_SPLIT);
    FDSM_SET_SMEM(3, EPI_STORE); FDSM_SET_SMEM(3, EPI_ACT_SPLIT); FDSM_SET_SMEM(3, EPI_HEAD); FDSM_SET_SMEM(3, EPI_DACT_SPLIT);
#undef FDSM_SET_SMEM
  });
#define FDSM_LAUNCH(P, E) gemm_kernel<P, E><<<grid, kThreads, smem, st>>>(p)
#define FDSM_LAUNCH_EPI(P)                                   \
  switch (p.epi) {                                           \
    case EPI_STORE: FDSM_LAUNCH(P, EPI_STORE); break;        \
    case EPI_ACT_SPLIT: FDSM_LAUNCH(P, EPI_ACT_SPLIT); break;\
    case EPI_HEAD: FDSM_LAUNCH(P, EPI_HEAD); break;          \
    default: FDSM_LAUNCH(P, EPI_DACT_SPLIT); break;          \
  }
  if (passes == 3) { FDSM_LAUNCH_EPI(3) } else { FDSM_LAUNCH_EPI(1) }
#undef FDSM_LAUNCH_EPI
#undef FDSM_LAUNCH
  FDSM_LAUNCH_CHECK("tc_gemm");
}


// Co-resident CTAs of gemm_wide_kernel when launched as clusters of `cs` along the split dimension (one CTA per SM by shared
// memory; clusters must fit inside a GPC, so this is a little below the SM count for cs = 4).  0: not launchable.
template <int CS>
inline int wgrad_cluster_capacity_of() {
  cudaFuncSetAttribute(gemm_wide_kernel<CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, WIDE_SMEM);
  if (CS == 1) return sm_count();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(1, 1, (unsigned)(sm_count() / CS * CS)); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = WIDE_SMEM;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = CS;
  cfg.attrs = at; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, gemm_wide_kernel<CS>, &cfg) != cudaSuccess) { (void)cudaGetLastError(); n = 0; }
  return n * CS;
}
inline int wgrad_cluster_capacity(int cs) {
  static const int c1 = wgrad_cluster_capacity_of<1>(), c2 = wgrad_cluster_capacity_of<2>(), c4 = wgrad_cluster_capacity_of<4>();
  return cs == 4 ? c4 : cs == 2 ? c2 : c1;
}

// dWt[M x N] partials = In^T dZ: In [rows x lda] (M of its columns), dZ [rows x ldb] (N <= 256 of its columns), both row-major
// with the contraction over `rows`.  `cluster` (1, 2, 4; splits % cluster == 0) consecutive splits of an output tile add their
// accumulators on chip (distributed shared memory): partial j = splits [cluster j, cluster (j + 1)) is written at
// out + j * split_stride with leading dimension ldo -- splits / cluster partials in all.
// FDSM_WGRAD_WIDE=0 routes the launch through the general kernel (A/B measurements; cluster must be 1 then).
inline int launch_wgrad_wide(const float* In, long long lda, const float* dZ, long long ldb, int M, int N, long long rows, int splits,
                             int cluster, float* out, long long ldo, long long split_stride, cudaStream_t st, bool pdl = false) {
  static const bool wide = [] { const char* e = getenv("FDSM_WGRAD_WIDE"); return !(e && e[0] == '0'); }();
  if (!wide || N > 256) {
    FDSM_REQUIRE(cluster == 1, "wgrad: the general kernel has no cluster reduction");
    GemmParams gp = {};
    gp.epi = EPI_STORE;
    gp.out_hi = out; gp.ldo = ldo; gp.split_stride = split_stride;
    Operand A{In, nullptr, lda, MAJOR_MN}, B{dZ, nullptr, ldb, MAJOR_MN};
    return launch_gemm(gp, A, B, M, N, (int)rows, N, 1, splits, st);
  }
  FDSM_REQUIRE(M >= 1 && N >= 1 && rows >= 1 && splits >= 1, "wgrad: bad shape");
  FDSM_REQUIRE((cluster == 1 || cluster == 2 || cluster == 4) && splits % cluster == 0, "wgrad: %d splits in clusters of %d", splits, cluster);
  GemmParams p = {};
  p.M = M; p.N = N; p.K = (int)rows; p.BN = (N + 15) / 16 * 16; p.a_major = MAJOR_MN; p.b_major = MAJOR_MN; p.passes = 1;
  p.epi = EPI_STORE; p.out_hi = out; p.ldo = ldo; p.split_stride = split_stride;
  {
    // (32 columns, row, 32-column panel): a box of 4 (In) / BN/32 (dZ) panels x 64 rows lands as [panel][64 rows][32]
    const long long da[3] = {32, rows, (lda + 31) / 32}, sa[2] = {4 * lda, 128};
    const int ba[3] = {32, WIDE_BK, BM / 32};
    int rc = make_tmap_nd(&p.ta_hi, In, 3, da, sa, ba, true, "wgrad In");
    if (rc) return rc;
    const int nbp = (p.BN + 31) / 32;
    const long long db[3] = {32, rows, nbp}, sb[2] = {4 * ldb, 128};
    const int bb[3] = {32, WIDE_BK, nbp};
    rc = make_tmap_nd(&p.tb_hi, dZ, 3, db, sb, bb, true, "wgrad dZ");
    if (rc) return rc;
  }
  p.trace = trace_buffer() ? trace_buffer() + (size_t)TRACE_STRIDE * trace_launch_index()++ : nullptr;
  (void)wgrad_cluster_capacity(1);     // (sets the shared-memory attribute of every instantiation, once)
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((M + BM - 1) / BM, 1, splits); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = WIDE_SMEM; cfg.stream = st;
  cudaLaunchAttribute at[2];
  int na = 0;
  if (cluster > 1) {
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = 1; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = cluster;
    ++na;
  }
  // programmatic launch only where the stream predecessor is a kernel of ours that the caller vouches for (`pdl`)
  if (pdl && pdl_enabled()) {
    at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = at; cfg.numAttrs = na;
  cudaError_t e = cluster == 4 ? cudaLaunchKernelEx(&cfg, gemm_wide_kernel<4>, p)
                : cluster == 2 ? cudaLaunchKernelEx(&cfg, gemm_wide_kernel<2>, p)
                               : cudaLaunchKernelEx(&cfg, gemm_wide_kernel<1>, p);
  return cuda_status(e, "tc_gemm(wgrad wide)");
}

// The same GEMMs for several layers in one launch (see gemm_wide_batch_kernel).  Job l: In_l [rows_l x lda_l] (M_l of its
// columns), dZ_l [rows_l x ldb] (N of its columns, the same N and ldb for every job), splits_l partials at out_l.
struct WgradJob { const float* In; long long lda; const float* dZ; int M; long long rows; int splits; float* out; long long split_stride; };
inline int launch_wgrad_batch(const WgradJob* jobs, int njobs, int N, long long ldb, long long ldo, cudaStream_t st, bool pdl) {
  FDSM_REQUIRE(njobs >= 1 && njobs <= WIDE_MAX_JOBS && N >= 1 && N <= 256, "wgrad batch: %d jobs, N = %d", njobs, N);
  WideBatchParams p = {};
  p.N = N; p.BN = (N + 15) / 16 * 16; p.njobs = njobs; p.ldo = ldo;
  const int nbp = (p.BN + 31) / 32;
  int ctas = 0;
  for (int l = 0; l < njobs; ++l) {
    const WgradJob& j = jobs[l];
    FDSM_REQUIRE(j.M >= 1 && j.rows >= 1 && j.splits >= 1, "wgrad batch: bad shape in job %d", l);
    const long long da[3] = {32, j.rows, (j.lda + 31) / 32}, sa[2] = {4 * j.lda, 128};
    const int ba[3] = {32, WIDE_BK, BM / 32};
    int rc = make_tmap_nd(&p.ta[l], j.In, 3, da, sa, ba, true, "wgrad In");
    if (rc) return rc;
    const long long db[3] = {32, j.rows, nbp}, sb[2] = {4 * ldb, 128};
    const int bb[3] = {32, WIDE_BK, nbp};
    rc = make_tmap_nd(&p.tb[l], j.dZ, 3, db, sb, bb, true, "wgrad dZ");
    if (rc) return rc;
    p.M[l] = j.M; p.K[l] = (int)j.rows; p.mt[l] = (j.M + BM - 1) / BM; p.splits[l] = j.splits;
    p.out[l] = j.out; p.split_stride[l] = j.split_stride;
    p.cta0[l] = ctas; ctas += p.mt[l] * j.splits;
  }
  p.cta0[njobs] = ctas;
  p.trace = trace_buffer() ? trace_buffer() + (size_t)TRACE_STRIDE * trace_launch_index()++ : nullptr;
  static std::once_flag once;
  std::call_once(once, [] { cudaFuncSetAttribute(gemm_wide_batch_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, WIDE_SMEM); });
  if (pdl) return cuda_status(launch_pdl(gemm_wide_batch_kernel<0>, dim3(ctas), dim3(kThreads), (size_t)WIDE_SMEM, st, p), "tc_gemm(wgrad batch)");
  gemm_wide_batch_kernel<0><<<ctas, kThreads, WIDE_SMEM, st>>>(p);
  FDSM_LAUNCH_CHECK("tc_gemm(wgrad batch)");
}

}  // namespace tc
}  // namespace fdsm
