// The step's single exchange, as ONE kernel over NVLink peer memory: all-reduce (mean) of the flat fp32
// gradient fused with the global-norm partial sums of clip_grad_norm_ and the optimizer's step-counter increment.
// Replaces  grad.mul_(1/W) -> ncclAllReduce -> sumsq  (3 launches, the NCCL one with ~20-30 us of latency at 1 MB).
//
// Push design, no grid-wide or end-of-kernel barrier:
//   every rank owns a peer-mapped receive buffer   slot[parity][src rank][n_pad]   +   flag[parity][block][src rank]
//   block i:  (1) stores chunk i of the local gradient into slot[parity][me] of EVERY peer (posted NVLink writes),
//             (2) fence.sys, then sets flag[parity][i][me] = epoch on every peer (st.release.sys),
//             (3) spins (ld.acquire.sys) until its own flag[parity][i][r] reached epoch for every r,
//             (4) sums chunk i over the ranks in rank order 0..W-1 (own chunk from the local gradient) -> every rank
//                 computes bit-identical sums; scales by 1/W; accumulates the squared norm of its chunk.
//   Block i only ever depends on block i of the peers, so the kernel cannot deadlock on residency.
//   parity = epoch & 1 double-buffers the receive slots: a peer can run at most one step ahead (its step t+1 push needs
//   nothing from me, its step t+2 push needs my step t+1 flags, which I set after finishing step t), so a slot is never
//   overwritten while it is being read.  The epoch lives in device memory -> the kernel replays from a CUDA graph.
//   The last block to finish (ticket counter) folds the per-block squared norms into the FDSM_SUMSQ_PARTIALS slots the
//   Adam kernel reads (fixed order -> deterministic) and advances epoch and step counter.
#include "fdsm_common.cuh"

namespace fdsm {
namespace {

constexpr int kPeerThreads = 256;
constexpr int kMaxWorld = FDSM_PEER_MAX_WORLD;
constexpr int kMaxBlocks = FDSM_PEER_MAX_BLOCKS;

struct PeerArgs {
  const float* grad; float* out; long long n, n_pad;
  float* slot[kMaxWorld];        // receive-slot base of every rank (peer-mapped addresses valid on THIS device)
  unsigned* flag[kMaxWorld];     // flag base of every rank
  int rank, world;
  float scale;
  float* sumsq; long long* step;
  unsigned* scratch;             // local: [0] epoch, [1] ticket, [2] status, [4..4+kMaxBlocks) per-block squared norms
  unsigned long long timeout_ns;
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void __launch_bounds__(kPeerThreads) peer_allreduce_kernel(const PeerArgs a) {
  __shared__ float red[kPeerThreads / 32];
  __shared__ int is_last;
  const int tid = threadIdx.x, blk = blockIdx.x, nblk = gridDim.x, W = a.world, me = a.rank;
  const unsigned epoch = *reinterpret_cast<volatile unsigned*>(a.scratch) + 1u;
  const int parity = (int)(epoch & 1u);
  // chunk i of the vector, in units of 4 floats
  const long long nv = a.n_pad >> 2, per = (nv + nblk - 1) / nblk;
  const long long v0 = blk * per, v1 = v0 + per < nv ? v0 + per : nv;
  const bool vec = aligned16(a.grad) && aligned16(a.out);
  auto load4 = [&](const float* base, long long v) -> float4 {
    const long long e = v << 2;
    if (vec && e + 3 < a.n) return *reinterpret_cast<const float4*>(base + e);
    return make_float4(e < a.n ? base[e] : 0.f, e + 1 < a.n ? base[e + 1] : 0.f, e + 2 < a.n ? base[e + 2] : 0.f,
                       e + 3 < a.n ? base[e + 3] : 0.f);
  };
  // (1) push my chunk into every peer's slot[parity][me]
  const long long my_slot = ((long long)(parity * W + me)) * a.n_pad;
  for (long long v = v0 + tid; v < v1; v += kPeerThreads) {
    const float4 x = load4(a.grad, v);
    for (int p = 0; p < W; ++p)
      if (p != me) *reinterpret_cast<float4*>(a.slot[p] + my_slot + (v << 2)) = x;
  }
  // (2) publish: every thread's stores are ordered before the flag by fence + barrier + release store
  __threadfence_system();
  __syncthreads();
  const long long fidx = ((long long)(parity * kMaxBlocks + blk)) * kMaxWorld;
  if (tid < W && tid != me) st_release_sys(a.flag[tid] + fidx + me, epoch);
  // (3) wait for chunk i of every peer
  if (tid < W && tid != me) {
    const unsigned* f = a.flag[me] + fidx + tid;
    // tight volatile poll; the timeout clock is read once per 4096 polls only (%globaltimer is slow to read)
    const volatile unsigned* fv = f;
    unsigned long long t0 = 0;
    bool arrived = false;
    while (!arrived) {
      for (int i = 0; i < 4096 && !arrived; ++i) arrived = (int)(*fv - epoch) >= 0;
      if (arrived) break;
      const unsigned long long t = globaltimer();
      if (t0 == 0) t0 = t;
      if (t - t0 > a.timeout_ns) {   // a peer died: do not hang the GPU, flag the error and go on
        atomicExch(a.scratch + 2, 1u);
        break;
      }
    }
    (void)ld_acquire_sys(f);         // acquire: the peer's slot stores happen-before the reads below
  }
  __syncthreads();
  // (4) reduce in rank order (bit-identical on every rank), scale, squared norm
  float acc2 = 0.f;
  for (long long v = v0 + tid; v < v1; v += kPeerThreads) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int p = 0; p < W; ++p) {
      float4 x;
      if (p == me) x = load4(a.grad, v);
      else x = __ldcg(reinterpret_cast<const float4*>(a.slot[me] + ((long long)(parity * W + p)) * a.n_pad + (v << 2)));
      s.x += x.x; s.y += x.y; s.z += x.z; s.w += x.w;
    }
    s.x *= a.scale; s.y *= a.scale; s.z *= a.scale; s.w *= a.scale;
    const long long e = v << 2;
    if (vec && e + 3 < a.n) {
      *reinterpret_cast<float4*>(a.out + e) = s;
    } else {
      if (e < a.n) a.out[e] = s.x; else s.x = 0.f;
      if (e + 1 < a.n) a.out[e + 1] = s.y; else s.y = 0.f;
      if (e + 2 < a.n) a.out[e + 2] = s.z; else s.z = 0.f;
      if (e + 3 < a.n) a.out[e + 3] = s.w; else s.w = 0.f;
    }
    acc2 += s.x * s.x + s.y * s.y + s.z * s.z + s.w * s.w;
  }
  acc2 = warp_sum(acc2);
  if ((tid & 31) == 0) red[tid >> 5] = acc2;
  __syncthreads();
  if (tid == 0) {
    float s = 0.f;
    for (int i = 0; i < kPeerThreads / 32; ++i) s += red[i];
    reinterpret_cast<float*>(a.scratch + 4)[blk] = s;
    __threadfence();
    is_last = atomicAdd(a.scratch + 1, 1u) == (unsigned)(nblk - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    const volatile float* part = reinterpret_cast<const volatile float*>(a.scratch + 4);
    if (tid < FDSM_SUMSQ_PARTIALS) {
      float s = 0.f;
      for (int j = tid; j < nblk; j += FDSM_SUMSQ_PARTIALS) s += part[j];
      if (a.sumsq) a.sumsq[tid] = s;
    }
    if (tid == 0) {
      a.scratch[1] = 0u;          // ticket for the next launch
      a.scratch[0] = epoch;       // every block of this launch has read the old value long ago
      if (a.step) *a.step += 1;
    }
  }
}

}  // namespace
}  // namespace fdsm

using namespace fdsm;

extern "C" int64_t fdsm_peer_buffer_bytes(int64_t n, int32_t world) {
  if (n < 0 || world < 1 || world > kMaxWorld) return -1;
  const int64_t n_pad = (n + 3) / 4 * 4;
  return 2 * (int64_t)world * n_pad * 4 + 2ll * kMaxBlocks * kMaxWorld * 4;
}

extern "C" int fdsm_peer_allreduce(const float* grad, float* out, int64_t n, void* const* peer_buffers, int32_t rank,
                                   int32_t world, float scale, float* sumsq, int64_t* step_counter, uint32_t* scratch,
                                   fdsm_stream_t stream) {
  FDSM_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, "fdsm_peer_allreduce: bad rank/world");
  FDSM_REQUIRE(n >= 0 && (n == 0 || (grad && out)) && peer_buffers && scratch, "fdsm_peer_allreduce: bad arguments");
  PeerArgs a = {};
  a.grad = grad; a.out = out; a.n = n; a.n_pad = (n + 3) / 4 * 4;
  for (int p = 0; p < world; ++p) {
    FDSM_REQUIRE(peer_buffers[p] && aligned16(peer_buffers[p]), "fdsm_peer_allreduce: peer buffer %d NULL or misaligned", p);
    a.slot[p] = static_cast<float*>(peer_buffers[p]);
    a.flag[p] = reinterpret_cast<unsigned*>(a.slot[p] + 2ll * world * a.n_pad);
  }
  a.rank = rank; a.world = world; a.scale = scale; a.sumsq = sumsq; a.step = (long long*)step_counter; a.scratch = scratch;
  a.timeout_ns = 20ull * 1000000000ull;
  // the grid is a function of n only: it must be identical on every rank (flags are per block)
  long long nblk = (a.n_pad / 4 + 511) / 512;
  if (nblk < 1) nblk = 1;
  if (nblk > kMaxBlocks) nblk = kMaxBlocks;
  peer_allreduce_kernel<<<(int)nblk, kPeerThreads, 0, (cudaStream_t)stream>>>(a);
  FDSM_LAUNCH_CHECK("fdsm_peer_allreduce");
}
