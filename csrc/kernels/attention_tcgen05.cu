// Fused causal self-attention for sm_100a (head dimension 64, bf16): softmax(Q K^T / sqrt(d)) V and
// its backward, with the score matrix kept on chip (TMEM + shared memory) instead of the
// [B*H, T, T] round trip through HBM of the six-GEMM formulation in ops/native_tf.py.
//
// No counterpart on the reference's own path (/root/reference/dpp.py trains a ResNet); this serves the
// driver's GPT-2 DDP configuration (BASELINE.json configs[3]) where stock PyTorch
// calls a flash-attention library kernel.
//
// Operands are read IN PLACE from the fused projection output qkv[B*T, 3*C] (q | k | v, head h at
// columns h*64) through one TMA descriptor (box 64 x 128, SWIZZLE_128B), so a [128 x 64] tile lands
// in shared memory as  row r -> 128 bytes at r*128, 16-byte chunk c stored at chunk c ^ (r & 7).
// That one image serves two UMMA views:
//   K-major  (rows = M or N, 64 reduction elements per row): desc(addr + 32*k16, LBO 16, SBO 1024)
//   MN-major (rows = reduction index, 64 M/N elements per row): desc(addr + 2048*k16, LBO = distance
//             between 64-element MN blocks, SBO 1024)
// so Q, K, V, dO are each loaded once and used by every product that needs them.
//
// Forward, one CTA = one (batch, head, 128-query block), two CTAs per SM:
//   warp 0  TMA producer : Q once, then K_j, V_j through a 3-slot ring
//   warp 1  MMA issuer   : S = Q K_j^T (TMEM cols 0..127), O_j = P_j V_j (cols 128..191, fresh each block)
//   warps 2-5 softmax    : thread = query row; two passes over S in TMEM (max, then exp2 / sum / bf16 P into
//                          shared memory in the K-major image), running output kept in registers:
//                          O = O * alpha + O_j, so no TMEM read-modify-write is needed
// Backward, one CTA = one (batch, head, 128-key block) looping over the query blocks at or below it:
//   S = Q_i K_j^T, dP = dO_i V_j^T -> P = exp2(S*c - L), dS = P (dP - D) / sqrt(d)  (bf16 into shared memory)
//   dV += P^T dO_i, dK += dS^T Q_i accumulate in TMEM across the loop; dQ_i = dS K_j is read back each
//   iteration and added to an fp32 buffer with vector reductions (red.global.add.v4.f32).
#include <cuda.h>
#include <cuda_bf16.h>
#include <math.h>

#include <type_traits>

#include "common.cuh"
#include "tcgen05.cuh"
#include "b200ddp.h"

namespace b200 {
namespace attn {
using namespace b200::tc;

constexpr int TILE_BYTES = 128 * 128;            // [128 rows x 64 bf16]
constexpr uint32_t SW128 = 2u;                   // descriptor layout type

__device__ __forceinline__ uint64_t desc_k(uint32_t tile, int k16) {          // K-major view, k16-th 16-element slice
  return make_desc(tile + k16 * 32, 16, 1024, SW128);
}
__device__ __forceinline__ uint64_t desc_mn(uint32_t tile, int k16, uint32_t mn_block_stride) {   // MN-major view
  return make_desc(tile + k16 * 2048, mn_block_stride, 1024, SW128);
}
// address of 16-byte chunk c8 (8 bf16 columns) of row r in a [128 x 128] bf16 operand held as two
// [128 x 64] K-major tiles side by side
__device__ __forceinline__ uint32_t p_chunk(uint32_t base, int r, int c8) {
  return base + (c8 >> 3) * TILE_BYTES + r * 128 + (((c8 & 7) ^ (r & 7)) << 4);
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------------
struct FwdArgs {
  __nv_bfloat16* out;      // [B*T, C]
  float* lse;              // [B*H, T]  log2-domain: max + log2(sum)
  int B, H, T, C;
  float scale_log2;        // log2(e) / sqrt(d)
  unsigned* error;
};

constexpr int FWD_THREADS = 192;
constexpr int FWD_SMEM = TILE_BYTES /*Q*/ + 3 * TILE_BYTES /*K,V ring*/ + 2 * TILE_BYTES /*P*/ + 1024;

__global__ void __launch_bounds__(FWD_THREADS, 2)
attn_fwd_kernel(const FwdArgs a, const __grid_constant__ CUtensorMap map_qkv) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t q_full, kv_full[3], kv_empty[3], s_full, p_full, o_full;
  __shared__ uint32_t tmem_base_s;

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int nq = (a.T + 127) >> 7;
  const int qb = nq - 1 - (int)blockIdx.x;                 // longest rows of work are scheduled first
  const int bh = blockIdx.y, b = bh / a.H, h = bh - b * a.H;
  const int nkv = qb + 1;
  const int row0 = b * a.T;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base, sRing = base + TILE_BYTES, sP = base + 4 * TILE_BYTES;

  if (t == 32) tma_prefetch_desc(&map_qkv);
  if (t == 0) {
    mbar_init(&q_full, 1);
    for (int s = 0; s < 3; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 1);
    }
    mbar_init(&s_full, 1);
    mbar_init(&p_full, 4);
    mbar_init(&o_full, 1);
    mbar_init_fence();
  }
  if (warp == 0) tmem_alloc(&tmem_base_s, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tS = tmem_base_s, tO = tmem_base_s + 128;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(&q_full, TILE_BYTES);
      tma_load_2d(sQ, &map_qkv, h * 64, row0 + qb * 128, &q_full);
      for (int idx = 0; idx < 2 * nkv; ++idx) {            // K_0, V_0, K_1, V_1, ...
        const int slot = idx % 3;
        if (idx >= 3) mbar_wait(&kv_empty[slot], ((idx / 3) - 1) & 1, a.error);
        mbar_expect_tx(&kv_full[slot], TILE_BYTES);
        tma_load_2d(sRing + slot * TILE_BYTES, &map_qkv, (1 + (idx & 1)) * a.C + h * 64, row0 + (idx >> 1) * 128,
                    &kv_full[slot]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idS = make_idesc_mn(true, false, false, 128, 128);
      constexpr uint32_t idO = make_idesc_mn(true, false, true, 128, 64);
      mbar_wait(&q_full, 0, a.error);
      auto issue_s = [&](int j) {
        const int idx = 2 * j, slot = idx % 3;
        mbar_wait(&kv_full[slot], (idx / 3) & 1, a.error);
        tc_fence_after();
        const uint32_t sK = sRing + slot * TILE_BYTES;
#pragma unroll
        for (int k = 0; k < 4; ++k) umma<true>(tS, desc_k(sQ, k), desc_k(sK, k), idS, k > 0);
        umma_commit(&kv_empty[slot]);
        umma_commit(&s_full);
      };
      issue_s(0);
      for (int j = 0; j < nkv; ++j) {
        mbar_wait(&p_full, j & 1, a.error);                // P_j is in shared memory, S_j has been consumed
        const int idx = 2 * j + 1, slot = idx % 3;
        mbar_wait(&kv_full[slot], (idx / 3) & 1, a.error);
        tc_fence_after();
        const uint32_t sV = sRing + slot * TILE_BYTES;
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma<true>(tO, desc_k(sP + (k >> 2) * TILE_BYTES, k & 3), desc_mn(sV, k, 8192), idO, k > 0);
        umma_commit(&kv_empty[slot]);
        umma_commit(&o_full);
        if (j + 1 < nkv) issue_s(j + 1);
      }
    }
  } else {
    const int lg = warp & 3;                               // TMEM lane quarter this warp may read
    const int row = lg * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(lg * 32) << 16;
    float m = -INFINITY, l = 0.f, alpha_prev = 1.f;
    float o[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) o[i] = 0.f;
    uint32_t r[32];

    // one key block; DIAG (compile time) = the block on the diagonal, the only one that needs the causal mask
    auto block = [&](int j, auto diag_tag) {
      constexpr bool DIAG = decltype(diag_tag)::value;
      mbar_wait(&s_full, j & 1, a.error);
      tc_fence_after();
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        tmem_ld32_nowait(tS + lane_addr + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float v = __uint_as_float(r[i]);
          if (!DIAG || c * 32 + i <= row) mx = fmaxf(mx, v);
        }
      }
      const float m_new = fmaxf(m, mx * a.scale_log2);
      const float alpha = exp2f(m - m_new);
      if (j > 0) {                                         // fold in the previous block's P V
        mbar_wait(&o_full, (j - 1) & 1, a.error);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          tmem_ld32_nowait(tO + lane_addr + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[c * 32 + i] = o[c * 32 + i] * alpha_prev + __uint_as_float(r[i]);
        }
      }
      float sum = 0.f;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        tmem_ld32_nowait(tS + lane_addr + c * 32, r);
        tmem_ld_wait();
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = exp2f(__uint_as_float(r[i]) * a.scale_log2 - m_new);
          float p1 = exp2f(__uint_as_float(r[i + 1]) * a.scale_log2 - m_new);
          if (DIAG && c * 32 + i > row) p0 = 0.f;
          if (DIAG && c * 32 + i + 1 > row) p1 = 0.f;
          sum += p0 + p1;
          w[i >> 1] = pack_bf16x2(p0, p1);
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
          st_shared_v4(p_chunk(sP, row, c * 4 + q), w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
      }
      l = l * alpha + sum;
      m = m_new;
      alpha_prev = alpha;
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full);
    };
    for (int j = 0; j < nkv - 1; ++j) block(j, std::false_type{});
    block(nkv - 1, std::true_type{});
    mbar_wait(&o_full, (nkv - 1) & 1, a.error);
    tc_fence_after();
    const float inv = 1.f / l;
    const int qrow = qb * 128 + row;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      tmem_ld32_nowait(tO + lane_addr + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[c * 32 + i] = (o[c * 32 + i] * alpha_prev + __uint_as_float(r[i])) * inv;
    }
    if (qrow < a.T) {
      uint4* dst = reinterpret_cast<uint4*>(a.out + (size_t)(row0 + qrow) * a.C + h * 64);
#pragma unroll
      for (int q = 0; q < 8; ++q)
        dst[q] = make_uint4(pack_bf16x2(o[8 * q], o[8 * q + 1]), pack_bf16x2(o[8 * q + 2], o[8 * q + 3]),
                            pack_bf16x2(o[8 * q + 4], o[8 * q + 5]), pack_bf16x2(o[8 * q + 6], o[8 * q + 7]));
      a.lse[(size_t)bh * a.T + qrow] = m + log2f(l);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    __syncwarp();
    tmem_dealloc(tmem_base_s, 256);
  }
}

// ------------------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------------------
struct BwdArgs {
  const float* lse;        // [B*H, T]
  const float* delta;      // [B*H, T]  rowsum(dO * O)
  float* dq32;             // [B*T, C]  fp32 accumulator (zeroed by the caller)
  __nv_bfloat16* dqkv;     // [B*T, 3C] dK and dV are written here directly
  int B, H, T, C;
  float scale_log2, scale;
  unsigned* error;
};

constexpr int BWD_THREADS = 320;                 // warps 0-7 compute, 8 TMA producer, 9 MMA issuer
constexpr int BWD_SMEM = 2 * TILE_BYTES /*K,V*/ + 4 * TILE_BYTES /*Q,dO x 2 stages*/ + 4 * TILE_BYTES /*P,dS*/ + 1024;

__device__ __forceinline__ void red_add_v4(float* addr, float x, float y, float z, float w) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}

__global__ void __launch_bounds__(BWD_THREADS, 1)
attn_bwd_kernel(const BwdArgs a, const __grid_constant__ CUtensorMap map_qkv, const __grid_constant__ CUtensorMap map_do) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t kv_full, qdo_full[2], qdo_empty[2], sdp_full, pds_full, dq_full;
  __shared__ uint32_t tmem_base_s;

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int nq = (a.T + 127) >> 7;
  const int jb = blockIdx.x;                               // key block; block 0 has the most query blocks
  const int bh = blockIdx.y, b = bh / a.H, h = bh - b * a.H;
  const int n_it = nq - jb;
  const int row0 = b * a.T;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sK = base, sV = base + TILE_BYTES;
  const uint32_t sQ0 = base + 2 * TILE_BYTES;              // stage st: Q at sQ0 + st*2*TILE, dO right after it
  const uint32_t sP = base + 6 * TILE_BYTES, sdS = base + 8 * TILE_BYTES;

  if (t == 32) {
    tma_prefetch_desc(&map_qkv);
    tma_prefetch_desc(&map_do);
  }
  if (t == 0) {
    mbar_init(&kv_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&qdo_full[s], 1);
      mbar_init(&qdo_empty[s], 1);
    }
    mbar_init(&sdp_full, 1);
    mbar_init(&pds_full, 8);
    mbar_init(&dq_full, 1);
    mbar_init_fence();
  }
  if (warp == 8) tmem_alloc(&tmem_base_s, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tS = tmem_base_s, tdP = tS + 128, tdV = tS + 256, tdK = tS + 320, tdQ = tS + 384;

  if (warp == 8) {
    if (lane == 0) {
      mbar_expect_tx(&kv_full, 2 * TILE_BYTES);
      tma_load_2d(sK, &map_qkv, a.C + h * 64, row0 + jb * 128, &kv_full);
      tma_load_2d(sV, &map_qkv, 2 * a.C + h * 64, row0 + jb * 128, &kv_full);
      for (int it = 0; it < n_it; ++it) {
        const int st = it & 1;
        if (it >= 2) mbar_wait(&qdo_empty[st], ((it >> 1) - 1) & 1, a.error);
        mbar_expect_tx(&qdo_full[st], 2 * TILE_BYTES);
        const uint32_t sQ = sQ0 + st * 2 * TILE_BYTES;
        tma_load_2d(sQ, &map_qkv, h * 64, row0 + (jb + it) * 128, &qdo_full[st]);
        tma_load_2d(sQ + TILE_BYTES, &map_do, h * 64, row0 + (jb + it) * 128, &qdo_full[st]);
      }
    }
  } else if (warp == 9) {
    if (lane == 0) {
      constexpr uint32_t idSP = make_idesc_mn(true, false, false, 128, 128);
      constexpr uint32_t idKV = make_idesc_mn(true, true, true, 128, 64);
      constexpr uint32_t idQ = make_idesc_mn(true, false, true, 128, 64);
      mbar_wait(&kv_full, 0, a.error);
      auto issue_sdp = [&](int it) {
        const int st = it & 1;
        mbar_wait(&qdo_full[st], (it >> 1) & 1, a.error);
        tc_fence_after();
        const uint32_t sQ = sQ0 + st * 2 * TILE_BYTES, sdO = sQ + TILE_BYTES;
#pragma unroll
        for (int k = 0; k < 4; ++k) umma<true>(tS, desc_k(sQ, k), desc_k(sK, k), idSP, k > 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma<true>(tdP, desc_k(sdO, k), desc_k(sV, k), idSP, k > 0);
        umma_commit(&sdp_full);
      };
      issue_sdp(0);
      for (int it = 0; it < n_it; ++it) {
        const int st = it & 1;
        const uint32_t sQ = sQ0 + st * 2 * TILE_BYTES, sdO = sQ + TILE_BYTES;
        mbar_wait(&pds_full, it & 1, a.error);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k)       // dV[key, d] += sum_q P[q, key] dO[q, d]
          umma<true>(tdV, desc_mn(sP, k, TILE_BYTES), desc_mn(sdO, k, 8192), idKV, (it > 0 || k > 0) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 8; ++k)       // dK[key, d] += sum_q dS[q, key] Q[q, d]
          umma<true>(tdK, desc_mn(sdS, k, TILE_BYTES), desc_mn(sQ, k, 8192), idKV, (it > 0 || k > 0) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 8; ++k)       // dQ[q, d] = sum_key dS[q, key] K[key, d]
          umma<true>(tdQ, desc_k(sdS + (k >> 2) * TILE_BYTES, k & 3), desc_mn(sK, k, 8192), idQ, k > 0);
        umma_commit(&qdo_empty[st]);
        umma_commit(&dq_full);
        if (it + 1 < n_it) issue_sdp(it + 1);
      }
    }
  } else {
    const int lg = warp & 3, hf = warp >> 2;               // lane quarter, column half
    const int row = lg * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(lg * 32) << 16;
    uint32_t r[32];

    auto flush_dq = [&](int it) {                          // dQ of iteration `it` -> fp32 accumulator
      mbar_wait(&dq_full, it & 1, a.error);
      tc_fence_after();
      tmem_ld32_nowait(tdQ + lane_addr + hf * 32, r);
      tmem_ld_wait();
      const int qrow = (jb + it) * 128 + row;
      if (qrow < a.T) {
        float* dst = a.dq32 + (size_t)(row0 + qrow) * a.C + h * 64 + hf * 32;
#pragma unroll
        for (int q = 0; q < 8; ++q)
          red_add_v4(dst + 4 * q, __uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]), __uint_as_float(r[4 * q + 2]),
                     __uint_as_float(r[4 * q + 3]));
      }
    };

    auto block = [&](int it, auto diag_tag) {
      constexpr bool DIAG = decltype(diag_tag)::value;
      const int qrow = (jb + it) * 128 + row;
      const bool valid = qrow < a.T;
      const float L = valid ? __ldg(a.lse + (size_t)bh * a.T + qrow) : INFINITY;
      const float Ds = valid ? __ldg(a.delta + (size_t)bh * a.T + qrow) * a.scale : 0.f;
      if (it > 0) flush_dq(it - 1);
      mbar_wait(&sdp_full, it & 1, a.error);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        const int col0 = hf * 64 + c * 32;
        float p[32];
        tmem_ld32_nowait(tS + lane_addr + col0, r);
        tmem_ld_wait();
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          p[i] = exp2f(__uint_as_float(r[i]) * a.scale_log2 - L);
          if (DIAG && col0 + i > row) p[i] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) w[i] = pack_bf16x2(p[2 * i], p[2 * i + 1]);
#pragma unroll
        for (int q = 0; q < 4; ++q)
          st_shared_v4(p_chunk(sP, row, (col0 >> 3) + q), w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
        tmem_ld32_nowait(tdP + lane_addr + col0, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i)                   // dS = P * (dP - D) / sqrt(d), the scale folded into one FMA
          w[i] = pack_bf16x2(p[2 * i] * (__uint_as_float(r[2 * i]) * a.scale - Ds),
                             p[2 * i + 1] * (__uint_as_float(r[2 * i + 1]) * a.scale - Ds));
#pragma unroll
        for (int q = 0; q < 4; ++q)
          st_shared_v4(p_chunk(sdS, row, (col0 >> 3) + q), w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
      }
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&pds_full);
    };
    block(0, std::true_type{});                            // the first query block is the diagonal one
    for (int it = 1; it < n_it; ++it) block(it, std::false_type{});
    flush_dq(n_it - 1);                                    // also means every MMA of this CTA has completed
    const int krow = jb * 128 + row;
#pragma unroll
    for (int which = 0; which < 2; ++which) {              // 0: dK, 1: dV
      tmem_ld32_nowait((which ? tdV : tdK) + lane_addr + hf * 32, r);
      tmem_ld_wait();
      if (krow < a.T) {
        uint4* dst = reinterpret_cast<uint4*>(a.dqkv + (size_t)(row0 + krow) * 3 * a.C + (1 + which) * a.C + h * 64 + hf * 32);
#pragma unroll
        for (int q = 0; q < 4; ++q)
          dst[q] = make_uint4(pack_bf16x2(__uint_as_float(r[8 * q]), __uint_as_float(r[8 * q + 1])),
                              pack_bf16x2(__uint_as_float(r[8 * q + 2]), __uint_as_float(r[8 * q + 3])),
                              pack_bf16x2(__uint_as_float(r[8 * q + 4]), __uint_as_float(r[8 * q + 5])),
                              pack_bf16x2(__uint_as_float(r[8 * q + 6]), __uint_as_float(r[8 * q + 7])));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    __syncwarp();
    tmem_dealloc(tmem_base_s, 512);
  }
}

// delta[b, h, t] = sum_d dO[b, t, h*64 + d] * O[b, t, h*64 + d]; 8 lanes per (token, head), 16 bytes each
__global__ void attn_delta_kernel(const __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ dout,
                                  float* __restrict__ delta, int B, int H, int T) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t pair = gid >> 3;
  const int sub = (int)(gid & 7);
  const size_t rows = (size_t)B * T;
  float acc = 0.f;
  const bool ok = pair < rows * H;
  size_t row = 0;
  int h = 0;
  if (ok) {
    row = pair / H;
    h = (int)(pair - row * H);
    const size_t off = (row * H + h) * 64 + sub * 8;
    const uint4 x = *reinterpret_cast<const uint4*>(out + off), y = *reinterpret_cast<const uint4*>(dout + off);
    const __nv_bfloat162* xs = reinterpret_cast<const __nv_bfloat162*>(&x);
    const __nv_bfloat162* ys = reinterpret_cast<const __nv_bfloat162*>(&y);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 fx = __bfloat1622float2(xs[i]), fy = __bfloat1622float2(ys[i]);
      acc += fx.x * fy.x + fx.y * fy.y;
    }
  }
  acc += __shfl_xor_sync(0xffffffffu, acc, 4);
  acc += __shfl_xor_sync(0xffffffffu, acc, 2);
  acc += __shfl_xor_sync(0xffffffffu, acc, 1);
  if (ok && sub == 0) {
    const size_t bb = row / T, tt = row - bb * T;
    delta[(bb * H + h) * T + tt] = acc;
  }
}

// dqkv[row, 0:C] = bf16(dq32[row, 0:C])
__global__ void attn_dq_cast_kernel(const float* __restrict__ dq32, __nv_bfloat16* __restrict__ dqkv, size_t rows, int C) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int per_row = C >> 3;
  if (gid >= rows * per_row) return;
  const size_t row = gid / per_row;
  const int c = (int)(gid - row * per_row) * 8;
  const float4 x = *reinterpret_cast<const float4*>(dq32 + row * C + c), y = *reinterpret_cast<const float4*>(dq32 + row * C + c + 4);
  *reinterpret_cast<uint4*>(dqkv + row * 3 * C + c) =
      make_uint4(pack_bf16x2(x.x, x.y), pack_bf16x2(x.z, x.w), pack_bf16x2(y.x, y.y), pack_bf16x2(y.z, y.w));
}

static int set_smem(const void* fn, int bytes, bool* done) {
  if (*done) return 0;
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return static_cast<int>(e);
  *done = true;
  return 0;
}

}  // namespace attn
}  // namespace b200

using namespace b200::attn;

// out[B*T, H*64] = causal softmax(Q K^T / sqrt(64)) V per head, lse[B*H, T] saved for the backward.
// Returns cudaErrorNotSupported (801) if the TMA descriptor cannot be built (caller falls back).
B200_API int b200_attention_fwd(const void* qkv, void* out, float* lse, int B, int H, int T, unsigned* error, cudaStream_t st) {
  const int C = H * 64;
  CUtensorMap map{};
  if (!b200::tc::tma_make_map_2d(&map, qkv, true, 3ll * C, (long long)B * T, 3ll * C, 64, 128)) return 801;
  static bool cfg = false;
  if (int rc = set_smem(reinterpret_cast<const void*>(attn_fwd_kernel), FWD_SMEM, &cfg)) return rc;
  FwdArgs a{};
  a.out = static_cast<__nv_bfloat16*>(out);
  a.lse = lse;
  a.B = B; a.H = H; a.T = T; a.C = C;
  a.scale_log2 = 1.4426950408889634f / 8.0f;
  a.error = error;
  dim3 grid((T + 127) / 128, B * H);
  attn_fwd_kernel<<<grid, FWD_THREADS, FWD_SMEM, st>>>(a, map);
  return static_cast<int>(cudaPeekAtLastError());
}

// dqkv[B*T, 3*H*64] from dout[B*T, H*64]; delta [B*H*T] and dq32 [B*T*H*64] are caller-provided fp32 scratch.
B200_API int b200_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                                float* dq32, void* dqkv, int B, int H, int T, unsigned* error, cudaStream_t st) {
  const int C = H * 64;
  const size_t rows = (size_t)B * T;
  CUtensorMap map_qkv{}, map_do{};
  if (!b200::tc::tma_make_map_2d(&map_qkv, qkv, true, 3ll * C, (long long)rows, 3ll * C, 64, 128)) return 801;
  if (!b200::tc::tma_make_map_2d(&map_do, dout, true, C, (long long)rows, C, 64, 128)) return 801;
  static bool cfg = false;
  if (int rc = set_smem(reinterpret_cast<const void*>(attn_bwd_kernel), BWD_SMEM, &cfg)) return rc;
  cudaError_t e = cudaMemsetAsync(dq32, 0, rows * C * sizeof(float), st);
  if (e != cudaSuccess) return static_cast<int>(e);
  {
    const size_t threads = rows * H * 8;
    attn_delta_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(
        static_cast<const __nv_bfloat16*>(out), static_cast<const __nv_bfloat16*>(dout), delta, B, H, T);
  }
  BwdArgs a{};
  a.lse = lse; a.delta = delta; a.dq32 = dq32;
  a.dqkv = static_cast<__nv_bfloat16*>(dqkv);
  a.B = B; a.H = H; a.T = T; a.C = C;
  a.scale = 0.125f;
  a.scale_log2 = 1.4426950408889634f / 8.0f;
  a.error = error;
  dim3 grid((T + 127) / 128, B * H);
  attn_bwd_kernel<<<grid, BWD_THREADS, BWD_SMEM, st>>>(a, map_qkv, map_do);
  {
    const size_t threads = rows * (C / 8);
    attn_dq_cast_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(dq32, static_cast<__nv_bfloat16*>(dqkv), rows, C);
  }
  return static_cast<int>(cudaPeekAtLastError());
}
