// Peer-to-peer collectives over NVLink-mapped memory: barrier, one-shot / two-shot all-reduce
// (fused with the 1/world_size averaging and fp32<->bf16 handling), broadcast.  No NCCL.
//
// Reference path being replaced: /root/reference/dpp.py:21,39 -> torch DDP -> ProcessGroupNCCL
// all-reduce of every gradient bucket, the constructor-time parameter
// broadcast (K2) and the per-forward BatchNorm-buffer broadcast (K3).
//
// Memory model.  Every rank owns one "symmetric arena" (a cudaMalloc'd block mapped into all
// peers through CUDA IPC); an object lives at the same offset in every rank's arena.  A rank
// addresses peer p's copy as bufs[p] + offset: plain ld/st on those addresses travel over
// NVLink 5 / NVSwitch.  Synchronisation is a flag exchange in the arenas' signal pads:
//
//     flag(dst, channel, block, src)  <- st.release.sys (epoch)       written by src into dst's pad
//     wait                            <- ld.acquire.sys  >= epoch     dst spins on its own pad
//
// `epoch` is a per-(channel, block) counter kept in device memory and only ever incremented,
// so flags are never reset (no reset race) and a CUDA graph can replay the kernels; comparisons
// are wrap-safe (signed difference), so the counters may roll over 2^32.
//
// Failure handling (the watchdog/abort a NCCL communicator brings with it).  A wait
// is bounded by wall time (%globaltimer, `timeout_ns` in the context, 10 minutes by default like
// the NCCL process-group timeout).  A block whose wait expires writes the error word (which the
// host keeps in pinned, host-visible memory and polls every step), performs NONE of its loads /
// stores of peer data, and overwrites its share of the LOCAL result with NaN: a collective that
// did not complete can never pass for a completed one.
//
// Algorithms (S = message bytes, N = ranks, in place on the arena object):
//   one-shot : barrier | every rank loads all N copies, sums in rank order 0..N-1 (bit-identical
//              everywhere), holds the result in registers | barrier | stores its own copy.
//              Ingress (N-1)*S per GPU, 2 barriers: best for small messages.
//   two-shot : barrier | rank r owns slice r: loads it from all N ranks, sums, scales, and PUSHES
//              the result into slice r of all N arenas | barrier.  One pass, no mid barrier (a
//              thread overwrites only what it has already read).  Per GPU and direction the links
//              carry (N-1)/N*S of loads plus (N-1)/N*S of pushes, so t >= 2(N-1)/N * S / B_link,
//              i.e. bus bandwidth S/t*2(N-1)/N is bounded by the link bandwidth.
//   push     : small messages (<= 256 KB), ONE flag exchange: every rank stores its data into a
//              per-source staging slot of every peer, releases a flag, waits for the N-1 peer flags
//              and sums the staged copies in rank order (bit-identical everywhere).  Staging is
//              double-buffered on the epoch parity, so no exit barrier is needed.
//   wire16   : two-shot for fp32 buckets with a 16-bit wire format (bf16 / fp16 compress hooks):
//              local cast fp32 -> 16 bit | barrier | owner pulls 16-bit slices, accumulates in fp32,
//              pushes the 16-bit result | barrier | local cast back to fp32.  Half the NVLink bytes.
//   broadcast: barrier | every rank pulls its 1/N slice from the root and pushes it to everyone |
//              barrier (scatter + all-gather fused, root egress S instead of (N-1)*S).
//
// Test hook: with `emulate` set, blockIdx.y plays the rank and all "peers" are buffers on one
// GPU, launched cooperatively so the spinning blocks are co-resident (the profiling guide
// forbids mutually-waiting separate launches on one GPU).
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <string.h>

#include "common.cuh"
#include "b200ddp.h"

namespace b200 {

constexpr int kMaxRanks = 8;
constexpr int kMaxCommBlocks = 128;          // flag rows per channel
constexpr int kNumChannels = 4;              // independent flag/epoch sets (one per stream role)
constexpr unsigned long long kDefaultTimeoutNs = 600ull * 1000ull * 1000ull * 1000ull;   // 10 min (NCCL PG default)
constexpr size_t kPushMaxBytes = 256u << 10;   // per-source staging slot of the push all-reduce

struct CommCtx {
  unsigned char* bufs[kMaxRanks];   // arena base of every rank, as mapped in THIS process
  uint32_t* flags[kMaxRanks];       // signal pad of every rank  [channel][block][src]
  uint32_t* epochs;                 // local: [rank (emulation only)][channel][block]
  uint32_t* error;                  // local error word
  int rank, world, channel, emulate;
  unsigned char* mc;                // NVLS: multicast mapping of the arenas (0 if the box has none)
  unsigned long long timeout_ns;    // bound on every wait (0: kDefaultTimeoutNs)
  unsigned long long* stamps;       // optional: [0] = %globaltimer when block 0 starts, [1] = when it ends (device-timed logger)
};

__device__ __forceinline__ int ctx_rank(const CommCtx& c) { return c.emulate ? (int)blockIdx.y : c.rank; }

__device__ __forceinline__ uint32_t* flag_slot(const CommCtx& c, int dst, int block, int src) {
  return c.flags[dst] + ((size_t)c.channel * kMaxCommBlocks + block) * kMaxRanks + src;
}

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Device-side timestamps for the logger: free of extra launches, valid under CUDA-graph replay.
__device__ __forceinline__ void stamp(const CommCtx& c, int which) {
  if (c.stamps != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) c.stamps[which] = global_ns();
}

__device__ __forceinline__ void raise_error(const CommCtx& c, uint32_t code) {
  // plain system-scope store: the word may live in pinned host memory (no PCIe atomics needed)
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(c.error), "r"(code) : "memory");
  __threadfence_system();
}

// Spin until *flag >= epoch (wrap-safe) or the context's time bound expires.  false = timed out.
__device__ __forceinline__ bool wait_flag(const CommCtx& c, const uint32_t* flag, uint32_t epoch, int peer) {
  if ((int32_t)(ld_acquire_sys(flag) - epoch) >= 0) return true;
  const unsigned long long limit = c.timeout_ns ? c.timeout_ns : kDefaultTimeoutNs;
  const unsigned long long t0 = global_ns();
  unsigned spins = 0;
  while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
    if ((++spins & 63u) == 0 && global_ns() - t0 > limit) {
      raise_error(c, 0xDEAD0000u | (uint32_t)peer);       // peer never arrived
      return false;
    }
  }
  return true;
}

__device__ __forceinline__ uint32_t next_epoch(const CommCtx& c, int rank) {
  __shared__ uint32_t epoch_s;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* e = c.epochs + ((size_t)(c.emulate ? rank : 0) * kNumChannels + c.channel) * kMaxCommBlocks + blockIdx.x;
    epoch_s = *e + 1u;
    *e = epoch_s;
  }
  __syncthreads();
  return epoch_s;
}

// All ranks' block `blockIdx.x` meet.  Acts as a system-scope release (before) / acquire (after).
// Returns false (uniformly over the block) when a peer did not arrive in time.
__device__ __forceinline__ bool block_barrier(const CommCtx& c, int rank) {
  __shared__ int ok_s;
  const uint32_t epoch = next_epoch(c, rank);
  if (threadIdx.x == 0) ok_s = 1;
  __syncthreads();
  if ((int)threadIdx.x < c.world) {
    const int p = threadIdx.x;
    // st.release.sys is cumulative: it publishes this block's earlier peer stores (ordered
    // before it by the bar.sync above) without a separate fence round trip
    st_release_sys(flag_slot(c, p, blockIdx.x, rank), epoch);
    if (!wait_flag(c, flag_slot(c, rank, blockIdx.x, p), epoch, p)) ok_s = 0;
  }
  __syncthreads();
  return ok_s != 0;
}

// A collective that cannot complete marks its share of the local result: NaN in every 16-byte vector.
__device__ __forceinline__ void poison_local(const CommCtx& c, int rank, size_t offset, size_t nvec) {
  const uint4 nan4 = make_uint4(0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u);   // NaN as fp32 and as bf16 pairs
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x)
    *reinterpret_cast<uint4*>(c.bufs[rank] + offset + i * 16) = nan4;
}

// ---- 16-byte vector helpers ----------------------------------------------------------------
__device__ __forceinline__ uint4 ld_peer(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer(void* p, uint4 v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};"
               ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <typename T> struct Vec;
template <> struct Vec<float> {
  __device__ static void acc(float (&a)[8], uint4 v) {
    a[0] += __uint_as_float(v.x); a[1] += __uint_as_float(v.y);
    a[2] += __uint_as_float(v.z); a[3] += __uint_as_float(v.w);
  }
  __device__ static uint4 pack(const float (&a)[8], float s) {
    return make_uint4(__float_as_uint(a[0] * s), __float_as_uint(a[1] * s), __float_as_uint(a[2] * s),
                      __float_as_uint(a[3] * s));
  }
};
template <> struct Vec<__nv_bfloat16> {
  __device__ static void acc(float (&a)[8], uint4 v) {     // bf16 pairs, accumulate in fp32
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      a[2 * i + 0] += __uint_as_float(w[i] << 16);
      a[2 * i + 1] += __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static uint4 pack(const float (&a)[8], float s) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(a[2 * i] * s, a[2 * i + 1] * s);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

// ---- two-shot (reduce-scatter + push all-gather in one pass) -----------------------------------
// NR = compile-time world size (0: runtime, up to kMaxRanks), U = vectors per thread per trip.
// A thread issues all U*(N) 16-byte loads of a trip before consuming any, so U*(N-1) remote
// requests per thread are in flight: NVLink latency (~2 us) is covered by memory-level
// parallelism rather than by occupying more SMs.
template <typename T, int NR, int U>
__global__ void __launch_bounds__(512) allreduce_twoshot_kernel(const CommCtx c, size_t offset, size_t nvec,
                                                                float scale) {
  constexpr int NMAX = NR ? NR : kMaxRanks;
  const int rank = ctx_rank(c), N = NR ? NR : c.world;
  stamp(c, 0);
  if (!block_barrier(c, rank)) return poison_local(c, rank, offset, nvec);
  const size_t per = (nvec + N - 1) / N;
  const size_t begin = min(nvec, (size_t)rank * per), end = min(nvec, begin + per);
  const size_t trip = (size_t)gridDim.x * blockDim.x * U;
  for (size_t base = begin + (size_t)blockIdx.x * blockDim.x * U + threadIdx.x; base < end; base += trip) {
    uint4 v[U][NMAX];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * blockDim.x;
#pragma unroll
      for (int p = 0; p < NMAX; ++p)
        if (p < N && i < end) v[u][p] = ld_peer(c.bufs[p] + offset + i * 16);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * blockDim.x;
      if (i >= end) continue;
      float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
      for (int p = 0; p < NMAX; ++p)
        if (p < N) Vec<T>::acc(a, v[u][p]);
      const uint4 r = Vec<T>::pack(a, scale);
#pragma unroll
      for (int p = 0; p < NMAX; ++p)
        if (p < N) st_peer(c.bufs[p] + offset + i * 16, r);
    }
  }
  if (!block_barrier(c, rank)) poison_local(c, rank, offset, nvec);
  stamp(c, 1);
}

// ---- NVLS (in-switch reduction through the NVLink multicast mapping) ---------------------------------
// multimem.ld_reduce on the multicast address returns the SUM over every rank's copy, computed
// inside NVSwitch; multimem.st writes all copies at once.  Rank r handles slice r: ingress per GPU
// drops from 2(N-1)/N*S to ~S/N*... (one reduced stream in, one broadcast stream out) and only a few
// SMs are needed.  Requires the arena to be bound to an NVLink multicast object (csrc/runtime/vmm_arena.cpp).
template <typename T> struct Multimem;
template <> struct Multimem<float> {
  __device__ static uint4 ld_reduce(const void* p) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
  }
  __device__ static void st(void* p, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
};
template <> struct Multimem<__nv_bfloat16> {
  __device__ static uint4 ld_reduce(const void* p) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
  }
  __device__ static void st(void* p, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};"
                 ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
};

template <typename T>
__global__ void __launch_bounds__(512) allreduce_nvls_kernel(const CommCtx c, size_t offset, size_t nvec, float scale) {
  const int rank = ctx_rank(c), N = c.world;
  stamp(c, 0);
  if (!block_barrier(c, rank)) return poison_local(c, rank, offset, nvec);
  const size_t per = (nvec + N - 1) / N;
  const size_t begin = min(nvec, (size_t)rank * per), end = min(nvec, begin + per);
  constexpr int U = 4;
  const size_t trip = (size_t)gridDim.x * blockDim.x * U;
  for (size_t base = begin + (size_t)blockIdx.x * blockDim.x * U + threadIdx.x; base < end; base += trip) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * blockDim.x;
      if (i < end) v[u] = Multimem<T>::ld_reduce(c.mc + offset + i * 16);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * blockDim.x;
      if (i >= end) continue;
      float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      Vec<T>::acc(a, v[u]);
      Multimem<T>::st(c.mc + offset + i * 16, Vec<T>::pack(a, scale));
    }
  }
  __threadfence_system();
  if (!block_barrier(c, rank)) poison_local(c, rank, offset, nvec);
  stamp(c, 1);
}

// ---- one-shot ------------------------------------------------------------------------------------
constexpr int kOneShotVecPerThread = 4;
template <typename T>
__global__ void __launch_bounds__(512) allreduce_oneshot_kernel(const CommCtx c, size_t offset, size_t nvec,
                                                                float scale) {
  const int rank = ctx_rank(c), N = c.world;
  stamp(c, 0);
  if (!block_barrier(c, rank)) return poison_local(c, rank, offset, nvec);
  uint4 res[kOneShotVecPerThread];
  const size_t base = (size_t)blockIdx.x * blockDim.x * kOneShotVecPerThread + threadIdx.x;
#pragma unroll
  for (int k = 0; k < kOneShotVecPerThread; ++k) {
    const size_t i = base + (size_t)k * blockDim.x;
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (i < nvec) {
      uint4 v[kMaxRanks];
#pragma unroll
      for (int p = 0; p < kMaxRanks; ++p)
        if (p < N) v[p] = ld_peer(c.bufs[p] + offset + i * 16);
#pragma unroll
      for (int p = 0; p < kMaxRanks; ++p)          // fixed order 0..N-1: same bits on every rank
        if (p < N) Vec<T>::acc(a, v[p]);
    }
    res[k] = Vec<T>::pack(a, scale);
  }
  if (!block_barrier(c, rank)) return poison_local(c, rank, offset, nvec);   // else: everyone has read everything
#pragma unroll
  for (int k = 0; k < kOneShotVecPerThread; ++k) {
    const size_t i = base + (size_t)k * blockDim.x;
    if (i < nvec) *reinterpret_cast<uint4*>(c.bufs[rank] + offset + i * 16) = res[k];
  }
  stamp(c, 1);
}

// ---- broadcast (scatter from root + push all-gather) ------------------------------------------------
__global__ void __launch_bounds__(512) broadcast_kernel(const CommCtx c, size_t offset, size_t nvec, int root) {
  constexpr int U = 8;
  const int rank = ctx_rank(c), N = c.world;
  stamp(c, 0);
  if (!block_barrier(c, rank)) {
    if (rank != root) poison_local(c, rank, offset, nvec);
    return;
  }
  const size_t per = (nvec + N - 1) / N;
  const size_t begin = min(nvec, (size_t)rank * per), end = min(nvec, begin + per);
  const size_t trip = (size_t)gridDim.x * blockDim.x * U;
  for (size_t base = begin + (size_t)blockIdx.x * blockDim.x * U + threadIdx.x; base < end; base += trip) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * blockDim.x;
      if (i < end) v[u] = ld_peer(c.bufs[root] + offset + i * 16);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * blockDim.x;
      if (i >= end) continue;
#pragma unroll
      for (int p = 0; p < kMaxRanks; ++p)
        if (p < N && p != root) st_peer(c.bufs[p] + offset + i * 16, v[u]);
    }
  }
  if (!block_barrier(c, rank) && rank != root) poison_local(c, rank, offset, nvec);
  stamp(c, 1);
}

__global__ void __launch_bounds__(32) barrier_kernel(const CommCtx c) { block_barrier(c, ctx_rank(c)); }

// ---- push all-reduce: one flag exchange, double-buffered staging ---------------------------------------
// Staging object at `stage_off` in every arena: [parity 2][source rank kMaxRanks][kPushMaxBytes].  Block b owns
// vectors [b*chunk, (b+1)*chunk) on EVERY rank, so block b's flag from peer p covers exactly the data block b reads.
// Parity = epoch & 1 of this (channel, block): two pushes in a row alternate slots; if another collective ran on
// the block in between, that collective's barrier already proves the peers are done with the previous push.
constexpr int kPushVecPerThread = 1;      // small messages: spread the (N-1) remote stores per vector over many threads / blocks
template <typename T>
__global__ void __launch_bounds__(512) allreduce_push_kernel(const CommCtx c, size_t offset, size_t nvec, float scale,
                                                             size_t stage_off) {
  __shared__ int ok_s;
  const int rank = ctx_rank(c), N = c.world;
  stamp(c, 0);
  const uint32_t epoch = next_epoch(c, rank);
  const size_t parity_off = stage_off + (size_t)(epoch & 1u) * kMaxRanks * kPushMaxBytes;
  const size_t base = (size_t)blockIdx.x * blockDim.x * kPushVecPerThread + threadIdx.x;
  uint4 mine[kPushVecPerThread];
#pragma unroll
  for (int k = 0; k < kPushVecPerThread; ++k) {
    const size_t i = base + (size_t)k * blockDim.x;
    if (i < nvec) {
      mine[k] = *reinterpret_cast<const uint4*>(c.bufs[rank] + offset + i * 16);
#pragma unroll
      for (int p = 0; p < kMaxRanks; ++p)
        if (p < N && p != rank) st_peer(c.bufs[p] + parity_off + (size_t)rank * kPushMaxBytes + i * 16, mine[k]);
    }
  }
  if (threadIdx.x == 0) ok_s = 1;
  __syncthreads();
  if ((int)threadIdx.x < N && (int)threadIdx.x != rank) {
    const int p = threadIdx.x;
    st_release_sys(flag_slot(c, p, blockIdx.x, rank), epoch);          // publishes the stores above (cumulative)
    if (!wait_flag(c, flag_slot(c, rank, blockIdx.x, p), epoch, p)) ok_s = 0;
  }
  __syncthreads();
  const bool ok = ok_s != 0;
#pragma unroll
  for (int k = 0; k < kPushVecPerThread; ++k) {
    const size_t i = base + (size_t)k * blockDim.x;
    if (i >= nvec) continue;
    uint4 out = make_uint4(0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u);
    if (ok) {
      float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
      for (int p = 0; p < kMaxRanks; ++p) {            // fixed order 0..N-1: same bits on every rank
        if (p >= N) continue;
        const uint4 v = p == rank ? mine[k]
                                  : ld_peer(c.bufs[rank] + parity_off + (size_t)p * kPushMaxBytes + i * 16);
        Vec<T>::acc(a, v);
      }
      out = Vec<T>::pack(a, scale);
    }
    *reinterpret_cast<uint4*>(c.bufs[rank] + offset + i * 16) = out;
  }
  stamp(c, 1);
}

// ---- two-shot with a 16-bit wire format for fp32 data ----------------------------------------------------
// Unit of work: 8 floats = 32 bytes of fp32 = one 16-byte vector on the wire.  Staging object at `stage_off`:
// [nw * 16 bytes].  W = __nv_bfloat16 or __half.
template <typename W> struct Wire;
template <> struct Wire<__nv_bfloat16> {
  __device__ static uint32_t pack2(float a, float b) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
  }
  __device__ static float2 unpack2(uint32_t w) { return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u)); }
};
template <> struct Wire<__half> {
  __device__ static uint32_t pack2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
  }
  __device__ static float2 unpack2(uint32_t w) { return __half22float2(*reinterpret_cast<const __half2*>(&w)); }
};

template <typename W>
__global__ void __launch_bounds__(512) allreduce_wire16_kernel(const CommCtx c, size_t offset, size_t nw, float scale,
                                                               size_t stage_off) {
  const int rank = ctx_rank(c), N = c.world;
  const size_t stride = (size_t)gridDim.x * blockDim.x, tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  float* data = reinterpret_cast<float*>(c.bufs[rank] + offset);
  stamp(c, 0);
  // 1. local cast into the staging object (pre-scaled by 1/world like the stock compress hooks: keeps fp16 in range)
  for (size_t i = tid; i < nw; i += stride) {
    const float4 lo = *reinterpret_cast<const float4*>(data + i * 8), hi = *reinterpret_cast<const float4*>(data + i * 8 + 4);
    const uint4 w = make_uint4(Wire<W>::pack2(lo.x * scale, lo.y * scale), Wire<W>::pack2(lo.z * scale, lo.w * scale),
                               Wire<W>::pack2(hi.x * scale, hi.y * scale), Wire<W>::pack2(hi.z * scale, hi.w * scale));
    *reinterpret_cast<uint4*>(c.bufs[rank] + stage_off + i * 16) = w;
  }
  __threadfence_system();
  if (!block_barrier(c, rank)) return poison_local(c, rank, offset, nw * 2);
  // 2. owner of slice `rank`: pull the 16-bit slices, accumulate in fp32 in rank order, push the 16-bit result
  const size_t per = (nw + N - 1) / N;
  const size_t begin = min(nw, (size_t)rank * per), end = min(nw, begin + per);
  for (size_t i = begin + tid; i < end; i += stride) {
    uint4 v[kMaxRanks];
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p)
      if (p < N) v[p] = ld_peer(c.bufs[p] + stage_off + i * 16);
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p) {
      if (p >= N) continue;
      const uint32_t w[4] = {v[p].x, v[p].y, v[p].z, v[p].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = Wire<W>::unpack2(w[j]);
        a[2 * j] += f.x;
        a[2 * j + 1] += f.y;
      }
    }
    const uint4 r = make_uint4(Wire<W>::pack2(a[0], a[1]), Wire<W>::pack2(a[2], a[3]), Wire<W>::pack2(a[4], a[5]),
                               Wire<W>::pack2(a[6], a[7]));
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p)
      if (p < N) st_peer(c.bufs[p] + stage_off + i * 16, r);
  }
  if (!block_barrier(c, rank)) return poison_local(c, rank, offset, nw * 2);
  // 3. local cast back
  for (size_t i = tid; i < nw; i += stride) {
    const uint4 w = *reinterpret_cast<const uint4*>(c.bufs[rank] + stage_off + i * 16);
    const float2 f0 = Wire<W>::unpack2(w.x), f1 = Wire<W>::unpack2(w.y), f2 = Wire<W>::unpack2(w.z), f3 = Wire<W>::unpack2(w.w);
    *reinterpret_cast<float4*>(data + i * 8) = make_float4(f0.x, f0.y, f1.x, f1.y);
    *reinterpret_cast<float4*>(data + i * 8 + 4) = make_float4(f2.x, f2.y, f3.x, f3.y);
  }
  stamp(c, 1);
}

template <typename K, typename... Args>
static int launch_comm(K kernel, const CommCtx& c, int blocks, int threads, cudaStream_t st, Args... args) {
  if (c.world < 1 || c.world > kMaxRanks || blocks < 1 || blocks > kMaxCommBlocks || c.channel < 0 ||
      c.channel >= kNumChannels)
    return static_cast<int>(cudaErrorInvalidValue);
  if (c.emulate) {
    CommCtx cc = c;
    void* params[] = {(void*)&cc, (void*)&args...};
    return static_cast<int>(cudaLaunchCooperativeKernel((const void*)kernel, dim3(blocks, c.world), dim3(threads),
                                                        params, 0, st));
  }
  kernel<<<blocks, threads, 0, st>>>(c, args...);
  return static_cast<int>(cudaPeekAtLastError());
}

static int pick_blocks(size_t nvec_per_rank, int max_blocks) {
  size_t b = (nvec_per_rank + 512 * 8 - 1) / (512 * 8);
  if (b < 1) b = 1;
  if (b > (size_t)max_blocks) b = max_blocks;
  return (int)b;
}

}  // namespace b200

using namespace b200;

B200_API int b200_comm_ctx_bytes(void) { return (int)sizeof(CommCtx); }
B200_API int b200_comm_max_blocks(void) { return kMaxCommBlocks; }
B200_API size_t b200_comm_signal_bytes(void) {
  return (size_t)kNumChannels * kMaxCommBlocks * kMaxRanks * sizeof(uint32_t);
}
B200_API size_t b200_comm_epoch_bytes(void) {
  return (size_t)kMaxRanks * kNumChannels * kMaxCommBlocks * sizeof(uint32_t);
}
B200_API size_t b200_allreduce_oneshot_max_bytes(void) {
  return (size_t)kMaxCommBlocks * 512 * kOneShotVecPerThread * 16;
}

// dtype: 0 = fp32, 1 = bf16.  algo: 1 = one-shot, 2 = two-shot, 3 = NVLS multimem.  `nbytes` % 16 == 0.
B200_API int b200_allreduce(const void* ctx, size_t offset, size_t nbytes, int dtype, float scale, int algo,
                            int max_blocks, cudaStream_t stream) {
  const CommCtx& c = *static_cast<const CommCtx*>(ctx);
  if (nbytes % 16 != 0 || offset % 16 != 0) return static_cast<int>(cudaErrorInvalidValue);
  const size_t nvec = nbytes / 16;
  if (nvec == 0) return 0;
  max_blocks = max(1, min(max_blocks, kMaxCommBlocks));
  if (algo == 1) {
    const size_t per_block = (size_t)512 * kOneShotVecPerThread;
    const size_t blocks = (nvec + per_block - 1) / per_block;
    if (blocks > (size_t)kMaxCommBlocks) return static_cast<int>(cudaErrorInvalidValue);   // too big for one-shot
    if (dtype == 0)
      return launch_comm(allreduce_oneshot_kernel<float>, c, (int)blocks, 512, stream, offset, nvec, scale);
    return launch_comm(allreduce_oneshot_kernel<__nv_bfloat16>, c, (int)blocks, 512, stream, offset, nvec, scale);
  }
  if (algo == 3) {          // NVLS
    if (c.mc == nullptr || c.emulate) return static_cast<int>(cudaErrorNotSupported);
    const int nb = pick_blocks((nvec + c.world - 1) / c.world, max_blocks);      // the caller caps it (24 by default)
    if (dtype == 0) return launch_comm(allreduce_nvls_kernel<float>, c, nb, 512, stream, offset, nvec, scale);
    return launch_comm(allreduce_nvls_kernel<__nv_bfloat16>, c, nb, 512, stream, offset, nvec, scale);
  }
  const int blocks = pick_blocks((nvec + c.world - 1) / c.world, max_blocks);
#define B200_TWOSHOT(T)                                                                                          \
  switch (c.world) {                                                                                             \
    case 2: return launch_comm(allreduce_twoshot_kernel<T, 2, 8>, c, blocks, 512, stream, offset, nvec, scale);  \
    case 4: return launch_comm(allreduce_twoshot_kernel<T, 4, 4>, c, blocks, 512, stream, offset, nvec, scale);  \
    case 8: return launch_comm(allreduce_twoshot_kernel<T, 8, 2>, c, blocks, 512, stream, offset, nvec, scale);  \
    default: return launch_comm(allreduce_twoshot_kernel<T, 0, 1>, c, blocks, 512, stream, offset, nvec, scale); \
  }
  if (dtype == 0) { B200_TWOSHOT(float) }
  B200_TWOSHOT(__nv_bfloat16)
#undef B200_TWOSHOT
}

B200_API size_t b200_allreduce_push_max_bytes(void) { return kPushMaxBytes; }
B200_API size_t b200_allreduce_push_stage_bytes(void) { return (size_t)2 * kMaxRanks * kPushMaxBytes; }

// Small-message all-reduce with a single flag exchange.  `stage_off`: arena offset of a staging object of
// b200_allreduce_push_stage_bytes() bytes (same offset on every rank, used by one channel only).
B200_API int b200_allreduce_push(const void* ctx, size_t offset, size_t nbytes, int dtype, float scale, size_t stage_off,
                                 cudaStream_t stream) {
  const CommCtx& c = *static_cast<const CommCtx*>(ctx);
  if (nbytes % 16 != 0 || offset % 16 != 0 || stage_off % 16 != 0 || nbytes > kPushMaxBytes)
    return static_cast<int>(cudaErrorInvalidValue);
  const size_t nvec = nbytes / 16;
  if (nvec == 0) return 0;
  const size_t per_block = (size_t)512 * kPushVecPerThread;
  const int blocks = (int)((nvec + per_block - 1) / per_block);
  if (blocks > kMaxCommBlocks) return static_cast<int>(cudaErrorInvalidValue);
  if (dtype == 0) return launch_comm(allreduce_push_kernel<float>, c, blocks, 512, stream, offset, nvec, scale, stage_off);
  return launch_comm(allreduce_push_kernel<__nv_bfloat16>, c, blocks, 512, stream, offset, nvec, scale, stage_off);
}

// fp32 data, 16-bit wire (wire: 1 = bf16, 2 = fp16), fp32 accumulation.  `nbytes` % 32 == 0; the staging object at
// `stage_off` needs nbytes / 2 bytes.
B200_API int b200_allreduce_wire16(const void* ctx, size_t offset, size_t nbytes, int wire, float scale, size_t stage_off,
                                   int max_blocks, cudaStream_t stream) {
  const CommCtx& c = *static_cast<const CommCtx*>(ctx);
  if (nbytes % 32 != 0 || offset % 16 != 0 || stage_off % 16 != 0 || (wire != 1 && wire != 2))
    return static_cast<int>(cudaErrorInvalidValue);
  const size_t nw = nbytes / 32;
  if (nw == 0) return 0;
  const int blocks = pick_blocks((nw + c.world - 1) / c.world, max(1, min(max_blocks, kMaxCommBlocks)));
  if (wire == 1) return launch_comm(allreduce_wire16_kernel<__nv_bfloat16>, c, blocks, 512, stream, offset, nw, scale, stage_off);
  return launch_comm(allreduce_wire16_kernel<__half>, c, blocks, 512, stream, offset, nw, scale, stage_off);
}

B200_API int b200_broadcast(const void* ctx, size_t offset, size_t nbytes, int root, int max_blocks,
                            cudaStream_t stream) {
  const CommCtx& c = *static_cast<const CommCtx*>(ctx);
  if (nbytes % 16 != 0 || offset % 16 != 0 || root < 0 || root >= c.world)
    return static_cast<int>(cudaErrorInvalidValue);
  const size_t nvec = nbytes / 16;
  if (nvec == 0) return 0;
  const int blocks = pick_blocks((nvec + c.world - 1) / c.world, max(1, min(max_blocks, kMaxCommBlocks)));
  return launch_comm(broadcast_kernel, c, blocks, 512, stream, offset, nvec, root);
}

// 1-thread kernel that writes %globaltimer to *dst: the main-stream marks of the device-timed logger
__global__ void stamp_kernel(unsigned long long* dst) {
  pdl_launch_dependents();
  pdl_wait();
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  *dst = t;
}
B200_API int b200_stamp(unsigned long long* dst, cudaStream_t stream) {
  return static_cast<int>(launch_pdl(stamp_kernel, dim3(1), dim3(1), 0, stream, dst));
}

B200_API int b200_barrier(const void* ctx, cudaStream_t stream) {
  const CommCtx& c = *static_cast<const CommCtx*>(ctx);
  return launch_comm(barrier_kernel, c, 1, 32, stream);
}

// ---- arena plumbing (cudaMalloc + CUDA IPC): control-plane helpers, called once at set-up ------------
B200_API int b200_arena_alloc(size_t bytes, void** out) {
  cudaError_t e = cudaMalloc(out, bytes);
  if (e != cudaSuccess) return static_cast<int>(e);
  return static_cast<int>(cudaMemset(*out, 0, bytes));
}
B200_API int b200_arena_free(void* p) { return static_cast<int>(cudaFree(p)); }
B200_API int b200_ipc_get_handle(void* p, unsigned char out[64]) {
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) return static_cast<int>(e);
  static_assert(sizeof(h) == 64, "IPC handle size");
  memcpy(out, &h, 64);
  return 0;
}
B200_API int b200_ipc_open_handle(const unsigned char in[64], void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, in, 64);
  return static_cast<int>(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
}
B200_API int b200_ipc_close_handle(void* p) { return static_cast<int>(cudaIpcCloseMemHandle(p)); }
