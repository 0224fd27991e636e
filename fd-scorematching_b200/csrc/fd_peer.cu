// The step's single exchange, as ONE kernel over NVLink peer memory: all-reduce (mean) of the flat fp32
// gradient fused with the global-norm partial sums of clip_grad_norm_ and the optimizer's step-counter increment.
// Replaces  grad.mul_(1/W) -> ncclAllReduce -> sumsq  (3 launches, the NCCL one with ~20-30 us of latency at 1 MB).
//
// One-shot push with the flag INSIDE the data, in 8-byte units ("LL", the layout of NCCL's low-latency protocol): the
// gradient travels as 16-byte lines {g[2w], epoch, g[2w+1], epoch}.  The guarantee relied upon is the one NCCL LL relies on:
// an aligned 8-byte store is observed whole, so a receiver that sees the epoch in a half sees that half's float; the two
// halves are validated independently, nothing is assumed about the 16-byte store as a unit (round 1 shipped
// {g, g, g, epoch} lines, which did assume it).  No membar.sys, no separate flag round trip (the first version -- plain
// stores, __threadfence_system, flag store, flag poll -- spent 7.5 us in the fence and 5 us waiting for the peer's
// fence; profiles/r01_peer_allreduce.txt).  tests/test_ddp_gpu.py replays the kernel 10^4 times from a CUDA graph against
// ncclAllReduce, and bench.py --gpus N prints the same cross-check ("ddp_check") on every multi-GPU run.
//   every rank owns a peer-mapped receive buffer   slot[parity][src rank][n2]   of 16-byte lines, n2 = ceil(n / 2)
//   thread t, word w:  (1) stores word w of the local gradient into slot[parity][me][w] of EVERY peer (posted NVLink writes),
//                      (2) polls its own slot[parity][r][w] (volatile 16-byte loads, L1 bypassed) until the epoch shows up,
//                      (3) sums the ranks in rank order 0..W-1 (own value from registers) -> bit-identical on every rank;
//                          scales by 1/W; writes out[2w..2w+1]; accumulates the squared norm.
//   A thread only ever depends on the same word of the peers, so the kernel cannot deadlock on residency.
//   parity = epoch & 1 double-buffers the receive slots: a peer can run at most one step ahead (its epoch t+2 push comes
//   after its epoch t+1 wait, which needs my epoch t+1 push, which I issue after finishing epoch t), so a word is never
//   overwritten before it is read, and a stale word (epoch t-2) never passes for a fresh one.  The epoch lives in device
//   memory -> the kernel replays from a CUDA graph.
//   The last block to finish (ticket counter) folds the per-block squared norms into the FDSM_SUMSQ_PARTIALS slots the
//   Adam kernel reads (fixed order -> deterministic) and advances epoch and step counter.
//   A poll that exceeds the timeout (a peer died) sets the sticky status word; the launch then publishes NaN squared norms
//   (the clipped update turns the parameters into NaN instead of silently applying a partial sum) and does NOT advance the
//   step counter; ddp.PeerAllReduce.timed_out() / FDSSMTrainer raise on it.
#include <cstdlib>

#include "fdsm_common.cuh"

namespace fdsm {
namespace {

constexpr int kPeerThreads = 1024;   // one 16-byte line per thread at config 2 (119 k lines on <= 160 blocks): every push is in flight before the first poll
constexpr int kMaxWorld = FDSM_PEER_MAX_WORLD;
constexpr int kMaxBlocks = FDSM_PEER_MAX_BLOCKS;

struct PeerArgs {
  const float* grad; float* out; long long n, n3;      // n3: 16-byte lines = ceil(n / 2) (the name predates the 2-float line)
  long long nown;                // two-shot: words owned per rank = ceil(n3 / world)
  uint4* slot[kMaxWorld];        // receive-slot base of every rank (peer-mapped addresses valid on THIS device)
  int rank, world;
  float scale;
  float* sumsq; long long* step;
  unsigned* scratch;             // local: [0] epoch, [1] ticket, [2] status, [4..4+kMaxBlocks) per-block squared norms
  unsigned long long timeout_ns;
};

__device__ __forceinline__ unsigned long long globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ uint4 ld_volatile_16(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_16(uint4* p, uint4 v) {
  asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ uint4 poll_word(const uint4* src, unsigned epoch, unsigned long long& t0, const PeerArgs& a) {
  uint4 x = ld_volatile_16(src);
  for (unsigned spin = 1; x.y != epoch || x.w != epoch; ++spin) {     // each 8-byte half carries its own flag
    if ((spin & 1023u) == 0) {   // the timeout clock is read once per 1024 polls (%globaltimer is slow to read)
      const unsigned long long t = globaltimer();
      if (t0 == 0) t0 = t;
      if (t - t0 > a.timeout_ns) {
        atomicExch(a.scratch + 2, 1u);   // a peer died: do not hang the GPU; the launch poisons its result (see the tail)
        break;
      }
    }
    x = ld_volatile_16(src);
  }
  return x;
}

__global__ void __launch_bounds__(kPeerThreads) peer_allreduce_kernel(const PeerArgs a) {
  __shared__ float red[kPeerThreads / 32];
  __shared__ int is_last;
  pdl_launch_dependents();     // the optimizer kernel behind this one is launched while it runs
  pdl_wait();                  // launched programmatically itself: the gradient comes from the kernel before it on the stream
  const int tid = threadIdx.x, blk = blockIdx.x, nblk = gridDim.x, W = a.world, me = a.rank;
  const unsigned epoch = *reinterpret_cast<volatile unsigned*>(a.scratch) + 1u;
  // phase stamps of block 0 (debug, u64 at scratch[192..]): start, pushed, reduced, done
  unsigned long long* stamps = reinterpret_cast<unsigned long long*>(a.scratch + 192);
  const bool stamp = blk == 0 && tid == 0;
  if (stamp) stamps[0] = globaltimer();
  const long long parity_off = (long long)(epoch & 1u) * W * a.n3;
  const long long per = (a.n3 + nblk - 1) / nblk;
  const long long w0 = blk * per, w1 = w0 + per < a.n3 ? w0 + per : a.n3;
  // (1) push: word w of my gradient, tagged with the epoch, into slot[parity][me][w] of every peer
  for (long long w = w0 + tid; w < w1; w += kPeerThreads) {
    const long long e = 2 * w;
    uint4 x;
    x.x = __float_as_uint(a.grad[e]);
    x.y = epoch;
    x.z = e + 1 < a.n ? __float_as_uint(a.grad[e + 1]) : 0u;
    x.w = epoch;
    for (int p = 0; p < W; ++p)
      if (p != me) st_16(a.slot[p] + parity_off + (long long)me * a.n3 + w, x);
  }
  if (stamp) stamps[1] = globaltimer();
  // (2)+(3) wait per word, reduce in rank order (bit-identical on every rank), scale, squared norm
  float acc2 = 0.f;
  unsigned long long t0 = 0;
  for (long long w = w0 + tid; w < w1; w += kPeerThreads) {
    const long long e = 2 * w;
    const float g0 = a.grad[e], g1 = e + 1 < a.n ? a.grad[e + 1] : 0.f;
    float s0 = 0.f, s1 = 0.f;
    for (int p = 0; p < W; ++p) {
      if (p == me) {
        s0 += g0; s1 += g1;
        continue;
      }
      const uint4 x = poll_word(a.slot[me] + parity_off + (long long)p * a.n3 + w, epoch, t0, a);
      s0 += __uint_as_float(x.x); s1 += __uint_as_float(x.z);
    }
    s0 *= a.scale; s1 *= a.scale;
    a.out[e] = s0;
    if (e + 1 < a.n) a.out[e + 1] = s1; else s1 = 0.f;
    acc2 += s0 * s0 + s1 * s1;
  }
  if (stamp) stamps[2] = globaltimer();
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
  if (stamp) stamps[3] = globaltimer();
  if (is_last) {
    __threadfence();
    const volatile float* part = reinterpret_cast<const volatile float*>(a.scratch + 4);
    const bool failed = *reinterpret_cast<volatile unsigned*>(a.scratch + 2) != 0u;     // sticky: this or an earlier launch
    if (tid < FDSM_SUMSQ_PARTIALS) {
      float s = 0.f;
      for (int j = tid; j < nblk; j += FDSM_SUMSQ_PARTIALS) s += part[j];
      if (a.sumsq) a.sumsq[tid] = failed ? __int_as_float(0x7fc00000) : s;
    }
    if (tid == 0) {
      a.scratch[1] = 0u;          // ticket for the next launch
      a.scratch[0] = epoch;       // every block of this launch has read the old value long ago
      if (a.step && !failed) *a.step += 1;
    }
  }
}


// Two-shot flavour for large worlds (the one-shot push moves (W-1) x n words in and out of every GPU: 12 us of NVLink
// time at W = 8).  Rank o owns the words [o*nown, (o+1)*nown):
//   (1) every rank sends each word to its OWNER only            -> slotA[parity][src][w - o*nown]   (reduce-scatter)
//   (2) the owner polls the W-1 contributions, sums in rank order, scales, and sends the RESULT word to every peer
//                                                               -> slotB[parity][w]                 (all-gather)
//   (3) every rank polls the result words it does not own.
// Traffic per GPU: 2 x (W-1)/W x n words instead of (W-1) x n; one more NVLink hop of latency.  Same flag-in-data lines,
// same parity double buffering; every element is summed by exactly one rank, so all ranks hold identical bits.
// Receive buffer layout (16-byte words): A = [2][W][nown], then B = [2][W * nown].
__global__ void __launch_bounds__(kPeerThreads) peer_allreduce2_kernel(const PeerArgs a) {
  __shared__ float red[kPeerThreads / 32];
  __shared__ int is_last;
  pdl_launch_dependents();     // the optimizer kernel behind this one is launched while it runs
  pdl_wait();                  // launched programmatically itself: the gradient comes from the kernel before it on the stream
  const int tid = threadIdx.x, blk = blockIdx.x, nblk = gridDim.x, W = a.world, me = a.rank;
  const unsigned epoch = *reinterpret_cast<volatile unsigned*>(a.scratch) + 1u;
  unsigned long long* stamps = reinterpret_cast<unsigned long long*>(a.scratch + 192);
  const bool stamp = blk == 0 && tid == 0;
  if (stamp) stamps[0] = globaltimer();
  const long long nown = a.nown, gsz = (long long)nblk * kPeerThreads, gtid = (long long)blk * kPeerThreads + tid;
  const long long offA = (long long)(epoch & 1u) * W * nown;                    // slotA[parity]
  const long long offB = 2ll * W * nown + (long long)(epoch & 1u) * W * nown;   // slotB[parity]
  auto load_word = [&](long long w) -> uint4 {
    const long long e = 2 * w;
    uint4 x;
    x.x = e < a.n ? __float_as_uint(a.grad[e]) : 0u;
    x.y = epoch;
    x.z = e + 1 < a.n ? __float_as_uint(a.grad[e + 1]) : 0u;
    x.w = epoch;
    return x;
  };
  // (1) reduce-scatter sends: the ranges of the other owners, the nearest owner first
  for (int d = 1; d < W; ++d) {
    const int o = (me + d) % W;
    for (long long i = gtid; i < nown; i += gsz) {
      const long long w = (long long)o * nown + i;
      if (w < a.n3) st_16(a.slot[o] + offA + (long long)me * nown + i, load_word(w));
    }
  }
  if (stamp) stamps[1] = globaltimer();
  // The squared norm must come out bit-identical on every rank (it sets the clip coefficient of the update: replicas whose
  // norms differ in the last bit drift apart).  A thread visits the same words on every rank -- index i of every owner's
  // range -- but in a rank-dependent ORDER (its own range first): it therefore keeps one partial per owner range and adds
  // them in owner order at the end.
  float acc_o[kMaxWorld];
#pragma unroll
  for (int q = 0; q < kMaxWorld; ++q) acc_o[q] = 0.f;
  unsigned long long t0 = 0;
  auto emit = [&](int o, long long w, float s0, float s1) {
    const long long e = 2 * w;
    if (e < a.n) a.out[e] = s0; else s0 = 0.f;
    if (e + 1 < a.n) a.out[e + 1] = s1; else s1 = 0.f;
    const float v = s0 * s0 + s1 * s1;
#pragma unroll
    for (int q = 0; q < kMaxWorld; ++q)      // (compile-time indices: the partials stay in registers)
      if (q == o) acc_o[q] += v;
  };
  // (2) my range: gather, reduce in rank order, broadcast the result.  The W - 1 contributions of a word are requested
  //     together and only those that have not arrived yet are polled: at 8 ranks a thread otherwise pays seven dependent
  //     round trips here and seven more in (3) (first 8-GPU run of round 2's end state: 20.7 us for the exchange, 9.1 at 2 ranks)
  for (long long i = gtid; i < nown; i += gsz) {
    const long long w = (long long)me * nown + i;
    if (w >= a.n3) break;
    const uint4 mine = load_word(w);
    uint4 pre[kMaxWorld];
#pragma unroll
    for (int p = 0; p < kMaxWorld; ++p)
      if (p < W && p != me) pre[p] = ld_volatile_16(a.slot[me] + offA + (long long)p * nown + i);
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int p = 0; p < kMaxWorld; ++p) {
      if (p < W) {
        uint4 x = mine;
        if (p != me) {
          x = pre[p];
          if (x.y != epoch || x.w != epoch) x = poll_word(a.slot[me] + offA + (long long)p * nown + i, epoch, t0, a);
        }
        s0 += __uint_as_float(x.x); s1 += __uint_as_float(x.z);
      }
    }
    s0 *= a.scale; s1 *= a.scale;
    uint4 r;
    r.x = __float_as_uint(s0); r.y = epoch; r.z = __float_as_uint(s1); r.w = epoch;
    for (int d = 1; d < W; ++d) st_16(a.slot[(me + d) % W] + offB + w, r);
    emit(me, w, s0, s1);
  }
  // (3) the other ranges: wait for the owners' results -- index i of every other owner's range requested together, then taken
  //     in the order of the owners' distance (each owner's partial still sees its words in increasing i: same sums as before)
  for (long long i = gtid; i < nown; i += gsz) {
    uint4 pre[kMaxWorld];
#pragma unroll
    for (int d = 1; d < kMaxWorld; ++d) {
      if (d < W) {
        const long long w = (long long)((me + d) % W) * nown + i;
        if (w < a.n3) pre[d] = ld_volatile_16(a.slot[me] + offB + w);
      }
    }
#pragma unroll
    for (int d = 1; d < kMaxWorld; ++d) {
      if (d < W) {
        const int o = (me + d) % W;
        const long long w = (long long)o * nown + i;
        if (w < a.n3) {
          uint4 x = pre[d];
          if (x.y != epoch || x.w != epoch) x = poll_word(a.slot[me] + offB + w, epoch, t0, a);
          emit(o, w, __uint_as_float(x.x), __uint_as_float(x.z));
        }
      }
    }
  }
  if (stamp) stamps[2] = globaltimer();
  float acc2 = 0.f;
#pragma unroll
  for (int q = 0; q < kMaxWorld; ++q) acc2 += acc_o[q];
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
  if (stamp) stamps[3] = globaltimer();
  if (is_last) {
    __threadfence();
    const volatile float* part = reinterpret_cast<const volatile float*>(a.scratch + 4);
    const bool failed = *reinterpret_cast<volatile unsigned*>(a.scratch + 2) != 0u;     // sticky: this or an earlier launch
    if (tid < FDSM_SUMSQ_PARTIALS) {
      float s = 0.f;
      for (int j = tid; j < nblk; j += FDSM_SUMSQ_PARTIALS) s += part[j];
      if (a.sumsq) a.sumsq[tid] = failed ? __int_as_float(0x7fc00000) : s;
    }
    if (tid == 0) {
      a.scratch[1] = 0u;
      a.scratch[0] = epoch;
      if (a.step && !failed) *a.step += 1;
    }
  }
}

}  // namespace
}  // namespace fdsm

using namespace fdsm;

extern "C" int64_t fdsm_peer_buffer_bytes(int64_t n, int32_t world) {
  if (n < 0 || world < 1 || world > kMaxWorld) return -1;
  const int64_t n3 = (n + 1) / 2;
  const int64_t nown = (n3 + world - 1) / world;
  const int64_t one_shot = 2 * (int64_t)world * (n3 > 0 ? n3 : 1);           // [2][W][n3]
  const int64_t two_shot = 4 * (int64_t)world * (nown > 0 ? nown : 1);       // A [2][W][nown] + B [2][W * nown]
  return (one_shot > two_shot ? one_shot : two_shot) * 16;                   // either protocol fits
}

extern "C" int fdsm_peer_allreduce(const float* grad, float* out, int64_t n, void* const* peer_buffers, int32_t rank,
                                   int32_t world, float scale, float* sumsq, int64_t* step_counter, uint32_t* scratch,
                                   fdsm_stream_t stream) {
  FDSM_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, "fdsm_peer_allreduce: bad rank/world");
  FDSM_REQUIRE(n >= 0 && (n == 0 || (grad && out)) && peer_buffers && scratch, "fdsm_peer_allreduce: bad arguments");
  PeerArgs a = {};
  a.grad = grad; a.out = out; a.n = n; a.n3 = (n + 1) / 2;
  for (int p = 0; p < world; ++p) {
    FDSM_REQUIRE(peer_buffers[p] && aligned16(peer_buffers[p]), "fdsm_peer_allreduce: peer buffer %d NULL or misaligned", p);
    a.slot[p] = static_cast<uint4*>(peer_buffers[p]);
  }
  a.nown = (a.n3 + world - 1) / world;
  a.rank = rank; a.world = world; a.scale = scale; a.sumsq = sumsq; a.step = (long long*)step_counter; a.scratch = scratch;
  static const unsigned long long timeout_ms = [] {     // FDSM_PEER_TIMEOUT_MS: tests shorten the 20 s a dead peer is waited for
    const char* e = getenv("FDSM_PEER_TIMEOUT_MS");
    const long long v = e ? atoll(e) : 0;
    return v > 0 ? (unsigned long long)v : 20000ull;
  }();
  a.timeout_ns = timeout_ms * 1000000ull;
  // the grid is a function of n only (one 16-byte word per thread where possible)
  long long nblk = (a.n3 + kPeerThreads - 1) / kPeerThreads;
  if (nblk < 1) nblk = 1;
  if (nblk > kMaxBlocks) nblk = kMaxBlocks;
  // one-shot up to 4 ranks (one NVLink hop); two-shot beyond (the one-shot push is NVLink-bandwidth bound at 8);
  // FDSM_PEER_TWO_SHOT=0/1 overrides -- it must be set identically on every rank
  static const int two_shot_env = [] {
    const char* e = getenv("FDSM_PEER_TWO_SHOT");
    return e ? (e[0] == '1' ? 1 : 0) : -1;
  }();
  const bool two_shot = world > 1 && (two_shot_env >= 0 ? two_shot_env == 1 : world > 4);
  const cudaError_t e = two_shot ? launch_pdl(peer_allreduce2_kernel, dim3((unsigned)nblk), dim3(kPeerThreads), 0, (cudaStream_t)stream, a)
                                 : launch_pdl(peer_allreduce_kernel, dim3((unsigned)nblk), dim3(kPeerThreads), 0, (cudaStream_t)stream, a);
  return cuda_status(e, "fdsm_peer_allreduce");
}
