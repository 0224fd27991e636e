// Microbenchmark (measurement tool, not product code): how fast can ONE SM pull an L2-resident operand stream through TMA
// into a shared-memory ring, as a function of the box shape, the ring depth and the number of CTAs pulling at once?
// Both ingest-bound kernels of the step (csrc/fd_chain.cuh phases, tc::gemm_kernel weight-gradient GEMMs) stream
// 32 x 32 fp32 boxes (4 KB, 128-byte rows) and sit at ~25-29 B/clk per SM; this tool separates "bytes in flight / latency"
// from "boxes per second".
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scripts/_bin/tma_ingest_bench scripts/tma_ingest_bench.cu -lcuda
//   scripts/_bin/tma_ingest_bench
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                         \
  do {                                                                                \
    cudaError_t e_ = (x);                                                             \
    if (e_ != cudaSuccess) {                                                          \
      fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));      \
      exit(1);                                                                        \
    }                                                                                 \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ int g_poll = 0;      // 0: mbarrier.try_wait (may suspend the thread), 1: mbarrier.test_wait (pure polling)
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  if (g_poll) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  }
  return ok != 0;
}
// bounded wait: a broken configuration ends the kernel instead of hanging the GPU
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
  for (long long i = 0; i < 20000000ll; ++i)
    if (mbar_try(bar, parity)) return true;
  return false;
}
__device__ __forceinline__ void tma2d(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
               "l"((uint64_t)tmap), "r"(bar), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma3d(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
               "l"((uint64_t)tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

__device__ __forceinline__ void tma4d(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
               "l"((uint64_t)tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

struct alignas(64) Params {
  CUtensorMap map;
  int mode;          // 0: `nbox` 2-D boxes {32, rows} per slot, box i at inner coordinate 32 i ; 1: one 3-D box {32, rows, nbox}
                     // 2: `nbox` 2-D boxes, one issuing thread per box (warp 0 lanes 0..nbox-1)
  int rows, nbox;    // box rows, boxes per slot
  int depth;         // ring slots
  int kblocks;       // k-blocks of one pass over the operand (outer coordinate advances by `rows`)
  int passes;        // passes over the operand
  int consume_clk;   // consumer holds a slot this long (emulates the MMAs)
  int* fail;
};

__global__ void __launch_bounds__(128, 1) ingest_kernel(const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t slot_bytes = (uint32_t)p.nbox * p.rows * 128u;
  const uint32_t bars = base + p.depth * slot_bytes;
  auto full = [&](int s) { return bars + 8u * s; };
  auto empty = [&](int s) { return bars + 8u * (16 + s); };
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.depth; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int total = p.kblocks * p.passes;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    int s = 0; uint32_t ph = 0;
    for (int it = 0; it < total; ++it) {
      const int kb = it % p.kblocks;
      bool ok = true;
      if (lane == 0) {
        ok = mbar_wait(empty(s), ph ^ 1u);
        if (ok) mbar_expect_tx(full(s), slot_bytes);
      }
      ok = __shfl_sync(0xffffffffu, ok, 0);
      if (!ok) { if (lane == 0) atomicAdd(p.fail, 1); return; }
      const uint32_t dst = base + s * slot_bytes;
      if (p.mode == 0) {
        if (lane == 0)
          for (int i = 0; i < p.nbox; ++i) tma2d(dst + i * p.rows * 128, &p.map, full(s), 32 * i, kb * p.rows);
      } else if (p.mode == 1) {
        if (lane == 0) tma3d(dst, &p.map, full(s), 0, kb * p.rows, 0);
      } else if (p.mode == 3) {      // 4-D {32, rows, nbox / 2, 2}: two row panels (hi | lo halves of the operand) in one instruction
        if (lane == 0) tma4d(dst, &p.map, full(s), 0, kb * p.rows, 0, 0);
      } else if (p.mode == 4) {      // two 3-D instructions per slot (nbox / 2 boxes each), issued by lanes 0 and 1
        if (lane < 2) tma3d(dst + lane * (p.nbox / 2) * p.rows * 128, &p.map, full(s), 0, kb * p.rows, lane * (p.nbox / 2));
      } else {
        if (lane < p.nbox) tma2d(dst + lane * p.rows * 128, &p.map, full(s), 32 * lane, kb * p.rows);
      }
      if (++s == p.depth) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 1 && lane == 0) {
    int s = 0; uint32_t ph = 0;
    for (int it = 0; it < total; ++it) {
      if (p.consume_clk >= 0 && !mbar_wait(full(s), ph)) { atomicAdd(p.fail, 1); return; }
      if (p.consume_clk > 0) {
        const long long t0 = clock64();
        while (clock64() - t0 < p.consume_clk) {}
      }
      mbar_arrive(empty(s));
      if (++s == p.depth) { s = 0; ph ^= 1u; }
    }
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  CK(cudaSetDevice(0));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  const double ghz = prop.clockRate * 1e-6;
  printf("%s, %d SMs, %.3f GHz (max)\n", prop.name, prop.multiProcessorCount, ghz);
  void* fnp = nullptr;
  cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fnp, cudaEnableDefault, &qr));
  EncodeFn encode = (EncodeFn)fnp;
  // the operand: a weight-like matrix [K x N] fp32, N = 224 contiguous (7 boxes of 32), K = 800: 700 KB, L2-resident
  const int K = 800, N = 224;
  float* w;
  CK(cudaMalloc(&w, (size_t)K * N * 4));
  CK(cudaMemset(w, 0, (size_t)K * N * 4));
  int* fail;
  CK(cudaMalloc(&fail, 4));
  CK(cudaMemset(fail, 0, 4));
  CK(cudaFuncSetAttribute(ingest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));

  struct Case { const char* name; int mode, rows, nbox, depth, grid, consume; int poll = 0; };
  std::vector<Case> cases;
  const int grids[] = {1, 32, 128, 148};
  for (int g : grids) {
    cases.push_back({"7 x {32x32} boxes, 1 thread", 0, 32, 7, 4, g, 0});
    cases.push_back({"7 x {32x32} boxes, 7 threads", 2, 32, 7, 4, g, 0});
    cases.push_back({"1 x 3-D {32x32x7} box", 1, 32, 7, 4, g, 0});
  }
  for (int d : {2, 3, 6}) cases.push_back({"7 x {32x32} boxes, 1 thread", 0, 32, 7, d, 128, 0});
  for (int d : {2, 3, 6}) cases.push_back({"1 x 3-D {32x32x7} box", 1, 32, 7, d, 128, 0});
  cases.push_back({"7 x {32x64} boxes (56 KB slots)", 0, 64, 7, 3, 128, 0});
  cases.push_back({"1 x 3-D {32x64x7} box (56 KB)", 1, 64, 7, 3, 128, 0});
  cases.push_back({"1 x {32x224}-row box (K-major 28 KB)", 0, 224, 1, 4, 128, 0});
  cases.push_back({"1 x 4-D {32x32x7x2} box (56 KB)", 3, 32, 14, 3, 128, 0});
  cases.push_back({"1 x 4-D {32x32x7x2} box, depth 2", 3, 32, 14, 2, 128, 0});
  cases.push_back({"1 x 3-D {32x32x4} box (16 KB)", 1, 32, 4, 4, 128, 0});
  cases.push_back({"1 x 3-D {32x32x2} box (8 KB)", 1, 32, 2, 4, 128, 0});
  cases.push_back({"1 x 3-D {32x32x1} box (4 KB)", 1, 32, 1, 4, 128, 0});
  cases.push_back({"1 x 3-D {32x128x7} box (112 KB)", 1, 128, 7, 1, 128, 0});
  cases.push_back({"3 x {32x32} boxes (12 KB)", 0, 32, 3, 4, 128, 0});
  cases.push_back({"3 x {32x64} boxes (24 KB)", 0, 64, 3, 4, 128, 0});
  cases.push_back({"3 x {32x128} boxes (48 KB)", 0, 128, 3, 3, 128, 0});
  cases.push_back({"3 x {32x256} boxes (96 KB)", 0, 256, 3, 2, 128, 0});
  cases.push_back({"7 x {32x96} boxes (84 KB)", 0, 96, 7, 2, 128, 0});
  cases.push_back({"1 x 3-D {32x96x7} box (84 KB)", 1, 96, 7, 2, 128, 0});
  cases.push_back({"7 x {32x32}, test_wait polling", 0, 32, 7, 4, 128, 0, 1});
  cases.push_back({"1 x 3-D {32x32x7}, test_wait polling", 1, 32, 7, 4, 128, 0, 1});
  cases.push_back({"7 x {32x32}, no flow control", 0, 32, 7, 4, 128, -1, 0});
  cases.push_back({"1 x 3-D {32x32x7}, no flow control", 1, 32, 7, 4, 128, -1, 0});
  cases.push_back({"1 x {32x32} box/slot, no flow control", 0, 32, 1, 4, 128, -1, 0});
  cases.push_back({"7 x {32x32} boxes, consumer 440 clk", 0, 32, 7, 4, 128, 440});
  cases.push_back({"1 x 3-D {32x32x7}, consumer 440 clk", 1, 32, 7, 4, 128, 440});

  printf("%-40s %5s %5s %6s | %9s %9s %9s\n", "pattern", "depth", "CTAs", "slotKB", "us", "GB/s/SM", "B/clk/SM");
  for (const Case& c : cases) {
    Params p = {};
    CUresult r;
    if (c.mode == 3) {
      cuuint64_t dims[4] = {32, (cuuint64_t)(K / 2), (cuuint64_t)(N / 32), 2};
      cuuint64_t strides[3] = {(cuuint64_t)N * 4, 128, (cuuint64_t)(K / 2) * N * 4};
      cuuint32_t box[4] = {32, (cuuint32_t)c.rows, (cuuint32_t)(c.nbox / 2), 2};
      cuuint32_t es[4] = {1, 1, 1, 1};
      r = encode(&p.map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, w, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else if (c.mode == 1 || c.mode == 4) {
      cuuint64_t dims[3] = {32, (cuuint64_t)K, (cuuint64_t)(N / 32)};
      cuuint64_t strides[2] = {(cuuint64_t)N * 4, 128};
      cuuint32_t box[3] = {32, (cuuint32_t)c.rows, (cuuint32_t)c.nbox};
      cuuint32_t es[3] = {1, 1, 1};
      r = encode(&p.map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, w, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else if (c.nbox == 1) {   // K-major view: [N' = 5600 rows x 32] -- rows of 128 B, contiguous
      cuuint64_t dims[2] = {32, (cuuint64_t)((size_t)K * N / 32)};
      cuuint64_t strides[1] = {128};
      cuuint32_t box[2] = {32, (cuuint32_t)c.rows};
      cuuint32_t es[2] = {1, 1};
      r = encode(&p.map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, w, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
      cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)K};
      cuuint64_t strides[1] = {(cuuint64_t)N * 4};
      cuuint32_t box[2] = {32, (cuuint32_t)c.rows};
      cuuint32_t es[2] = {1, 1};
      r = encode(&p.map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, w, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    if (r != CUDA_SUCCESS) { printf("%-40s encode failed (%d)\n", c.name, (int)r); continue; }
    p.mode = c.mode; p.rows = c.rows; p.nbox = c.nbox; p.depth = c.depth;
    p.kblocks = c.nbox == 1 ? (K * N / 32) / c.rows : (c.mode == 3 ? K / 2 : K) / c.rows;
    p.passes = 40; p.consume_clk = c.consume; p.fail = fail;
    CK(cudaMemcpyToSymbol(g_poll, &c.poll, sizeof(int)));
    const size_t slot = (size_t)c.nbox * c.rows * 128;
    const size_t smem = 1024 + c.depth * slot + 512;
    ingest_kernel<<<c.grid, 128, smem>>>(p);   // warm-up (L2 fill)
    CK(cudaGetLastError());
    CK(cudaEventRecord(e0));
    ingest_kernel<<<c.grid, 128, smem>>>(p);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    int f = 0;
    CK(cudaMemcpy(&f, fail, 4, cudaMemcpyDeviceToHost));
    const double bytes = (double)slot * p.kblocks * p.passes;
    const double gbs = bytes / (ms * 1e-3) / 1e9;
    printf("%-40s %5d %5d %6.1f | %9.1f %9.1f %9.1f%s\n", c.name, c.depth, c.grid, slot / 1024.0, ms * 1e3, gbs, gbs / ghz,
           f ? "  (TIMED OUT)" : "");
    if (f) CK(cudaMemset(fail, 0, 4));
  }
  return 0;
}
