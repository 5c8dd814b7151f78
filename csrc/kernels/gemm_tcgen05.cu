// tcgen05 / TMEM GEMM core for sm_100a with software operand gathers: dense GEMMs (Linear
// forward / dgrad / wgrad, fp32-as-TF32 and bf16) and implicit-GEMM convolution
// (fprop / dgrad / wgrad on NHWC activations and KRSC filters).
//
// Replaces cuDNN convolution and cuBLAS GEMM on the reference's path
// (/root/reference/dpp.py:50-52 -> ResNet-18 convs, fc; SURVEY.md K7/K8/K9/K16, work items
// W9-W12).  The FFMA kernels in conv.cu stay as the fallback for shapes this kernel does not
// take (channel counts that are not a multiple of the 16-byte vector, e.g. the RGB stem).
//
// Kernel anatomy (one CTA = one 128 x BN output tile, optionally one split of K):
//
//   warps 0-7  producers : gather 16-byte chunks of A and B with cp.async (zero-fill for
//                          padding / out-of-range) straight into the UMMA canonical shared-memory
//                          layouts (K-major and 16-bit MN-major: no-swizzle core matrices; 32-bit
//                          MN-major: SWIZZLE_128B_BASE32B), 4-stage ring, then
//                          fence.proxy.async + mbarrier arrive on full[stage]
//   warp  8    MMA issuer: one elected thread waits full[stage], issues 4 x tcgen05.mma
//                          (kind::tf32 K=8 or kind::f16 K=16, M=128, N=BN, accumulator in TMEM),
//                          tcgen05.commit -> empty[stage]; after the last k-block commit -> accum
//   warps 0-7  epilogue  : tcgen05.ld the accumulator (lane = row), then either park the
//                          partial tile for split-K (last split to arrive reduces in order) or
//                          store with optional bias / BatchNorm statistics / transposed output.
//
// Operand "kinds" (how a 16-byte chunk (mn, k) is found in global memory):
//   DENSE_K      ptr[mn*ld + k]                       k contiguous      (Linear x, W; conv filters in fprop)
//   IM2COL_K     X[n, p*st-pad+r, q*st-pad+s, c]      k=(r,s,c)         (conv fprop activations)
//   IM2COL_T_K   dY[n,(h+pad-r)/st,(w+pad-s)/st, k']  k=(r,s,k')        (conv dgrad)
//   DENSE_MN     ptr[k*ld + mn]                       mn contiguous     (Linear wgrad / dgrad, conv wgrad dY)
//   WEIGHT_T_MN  Wt[(k'*R*S + rs)*C + mn]             k=(rs,k')         (conv dgrad filters)
//   IM2COL_MN    X[pixel k shifted by (r,s)][c]       mn=(r,s,c)        (conv wgrad activations)
// K-kinds land in shared memory K-major, MN-kinds MN-major; the instruction descriptor's
// a_major / b_major bits tell the tensor core which.
//
// Small-M convolutions (deep layers of a 32x32-input ResNet: 32..512 output pixels) run
// "swapped": D^T = W * im2col(X)^T, so the 128-row MMA dimension is filled by output channels
// and the few pixels become N; the epilogue stores transposed.
#include <cuda_bf16.h>
#include <stdio.h>

#include "common.cuh"
#include "b200ddp.h"

namespace b200 {
namespace tc {

constexpr int BM = 128;
constexpr int STAGES = 4;
constexpr int PRODUCERS = 256;              // 8 warps
constexpr int THREADS = PRODUCERS + 32;     // + MMA warp
constexpr int ROW_BYTES = 128;              // bytes of K per tile row per stage (32 fp32 / 64 bf16)
constexpr unsigned kWaitLimit = 1u << 28;

enum Kind { DENSE_K = 0, IM2COL_K = 1, IM2COL_T_K = 2, DENSE_MN = 3, WEIGHT_T_MN = 4, IM2COL_MN = 5 };
__host__ __device__ constexpr bool is_mn(int kind) { return kind >= DENSE_MN; }

struct Geom {
  int Nb, H, W, C, K, R, S, P, Q, stride, pad;
};

struct Args {
  const void* a;
  const void* b;
  float* out;
  long long lda, ldb, ldo;
  int M, N, Kd;
  Geom g;
  int splits, k_per_split;     // k_per_split in elements, multiple of the stage depth
  float* workspace;
  unsigned* counters;
  const float* bias;           // per output CHANNEL (column, or row when transposed)
  float* stats;                // [2 * channels]
  int accumulate;
  int transpose_out;           // out[col * ldo + row]
  unsigned* error;
};

// ---- PTX wrappers ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, unsigned* error) {
  unsigned spins = 0;
  while (!mbar_try(bar, parity)) {
    if (++spins > kWaitLimit) {          // never hang the GPU: flag and fall through
      if (error) atomicExch(error, 0xBAD0BAD0u);
      break;
    }
  }
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int n = valid ? 16 : 0;           // src-size 0 -> 16 bytes of zeros
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols));
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
template <bool BF16>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  if (BF16) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
  }
}
// 32 lanes x 16 consecutive 32-bit columns -> 16 registers (thread = lane = accumulator row)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor, SWIZZLE_NONE canonical layout (all offsets in bytes, /16 encoded).
// layout_type: 0 = SWIZZLE_NONE, 1 = SWIZZLE_128B_BASE32B (the only MN-major layout for 32-bit operands)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout_type = 0) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46) | ((uint64_t)layout_type << 61);
}
__host__ __device__ constexpr uint32_t make_idesc(bool bf16, bool a_mn, bool b_mn, int n) {
  return (1u << 4) | ((bf16 ? 1u : 2u) << 7) | ((bf16 ? 1u : 2u) << 10) | ((a_mn ? 1u : 0u) << 15) |
         ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

// ---- operand chunk addressing ------------------------------------------------------------------
// Returns the global address of the 16-byte chunk starting at element (mn, k) and whether it is
// in range (false -> the chunk is zero-filled).  ES = element size in bytes.
template <int KIND, int ES>
__device__ __forceinline__ const char* chunk_src(const void* base, long long ld, const Geom& g, int mn, int k,
                                                 int mn_limit, int k_limit, bool& valid) {
  const char* p = static_cast<const char*>(base);
  valid = (mn < mn_limit) && (k < k_limit);
  if (!valid) return p;
  if (KIND == DENSE_K) return p + ((long long)mn * ld + k) * ES;
  if (KIND == DENSE_MN) return p + ((long long)k * ld + mn) * ES;
  if (KIND == IM2COL_K) {
    const int q = mn % g.Q, pp = (mn / g.Q) % g.P, n = mn / (g.Q * g.P);
    const int c = k % g.C, rs = k / g.C, s = rs % g.S, r = rs / g.S;
    const int h = pp * g.stride - g.pad + r, w = q * g.stride - g.pad + s;
    valid = h >= 0 && h < g.H && w >= 0 && w < g.W;
    return valid ? p + (((long long)(n * g.H + h) * g.W + w) * g.C + c) * ES : p;
  }
  if (KIND == IM2COL_T_K) {
    const int w = mn % g.W, h = (mn / g.W) % g.H, n = mn / (g.W * g.H);
    const int co = k % g.K, rs = k / g.K, s = rs % g.S, r = rs / g.S;
    const int hh = h + g.pad - r, ww = w + g.pad - s;
    const int pp = hh / g.stride, q = ww / g.stride;
    valid = hh >= 0 && ww >= 0 && (hh % g.stride) == 0 && (ww % g.stride) == 0 && pp < g.P && q < g.Q;
    return valid ? p + (((long long)(n * g.P + pp) * g.Q + q) * g.K + co) * ES : p;
  }
  if (KIND == WEIGHT_T_MN) {
    const int co = k % g.K, rs = k / g.K;
    return p + (((long long)co * g.R * g.S + rs) * g.C + mn) * ES;
  }
  // IM2COL_MN: k = output pixel, mn = (r, s, c)
  {
    const int q = k % g.Q, pp = (k / g.Q) % g.P, n = k / (g.Q * g.P);
    const int c = mn % g.C, rs = mn / g.C, s = rs % g.S, r = rs / g.S;
    const int h = pp * g.stride - g.pad + r, w = q * g.stride - g.pad + s;
    valid = h >= 0 && h < g.H && w >= 0 && w < g.W;
    return valid ? p + (((long long)(n * g.H + h) * g.W + w) * g.C + c) * ES : p;
  }
}

// Fill one operand tile of one stage.  ROWS = tile extent along mn (128 for A, BN for B).
// `k0` = first reduction index of the stage, `k_end` = end of this split's range.
template <int KIND, int ES, int ROWS>
__device__ __forceinline__ void load_tile(uint32_t stage_smem, const void* base, long long ld, const Geom& g, int mn0,
                                          int mn_limit, int k0, int k_end, int t) {
  constexpr int T = 16 / ES;                  // elements per 16-byte chunk
  constexpr int KELEMS = ROW_BYTES / ES;      // reduction elements per stage (32 fp32 / 64 bf16)
  constexpr int CHUNKS = ROWS * 8;            // 16-byte chunks per stage
  if (!is_mn(KIND)) {
    // K-major canonical layout: chunk (row, kc) at kc*(ROWS*16) + (row/8)*128 + (row%8)*16
    for (int idx = t; idx < CHUNKS; idx += PRODUCERS) {
      const int row = idx % ROWS, kc = idx / ROWS;
      bool valid;
      const char* src = chunk_src<KIND, ES>(base, ld, g, mn0 + row, k0 + kc * T, mn_limit, k_end, valid);
      cp_async16(stage_smem + kc * (ROWS * 16) + (row >> 3) * 128 + (row & 7) * 16, src, valid);
    }
  } else {
    constexpr int MNC = ROWS / T;
    for (int idx = t; idx < CHUNKS; idx += PRODUCERS) {
      const int mnc = idx % MNC, kk = idx / MNC;
      if (kk >= KELEMS) break;
      bool valid;
      const char* src = chunk_src<KIND, ES>(base, ld, g, mn0 + mnc * T, k0 + kk, mn_limit, k_end, valid);
      uint32_t off;
      if (ES == 2) {
        // 16-bit MN-major, no swizzle: chunk (mnc, k) at (k/8)*((ROWS/T)*128) + mnc*128 + (k%8)*16
        off = (kk >> 3) * (MNC * 128) + mnc * 128 + (kk & 7) * 16;
      } else {
        // 32-bit MN-major exists only as SWIZZLE_128B_BASE32B: atom = 32 mn (one 128-byte row) x 4 k
        // rows; inside a row the 32-byte units are XOR-ed with the row index (byte-address bits
        // [5,7) ^= [7,9)).  Atoms tile MN first (512 B apart), then K.  (Layout measured on B200.)
        const int c = mnc & 7, kr = kk & 3;
        off = (mnc >> 3) * 512 + (kk >> 2) * ((ROWS / 32) * 512) + kr * 128 + ((((c >> 1) ^ kr) << 5) | ((c & 1) << 4));
      }
      cp_async16(stage_smem + off, src, valid);
    }
  }
}

template <int AK, int BK_, int BN, bool BF16>
__global__ void __launch_bounds__(THREADS, 1) gemm_kernel(const Args a) {
  constexpr int ES = BF16 ? 2 : 4;
  constexpr int T = 16 / ES;
  constexpr int KELEMS = ROW_BYTES / ES;            // per stage
  constexpr int UMMA_K = 32 / ES;                   // 8 (tf32) / 16 (bf16)
  constexpr int A_STAGE = BM * ROW_BYTES, B_STAGE = BN * ROW_BYTES;
  constexpr uint32_t TMEM_COLS = BN < 32 ? 32 : BN;
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES], accum_bar;
  __shared__ uint32_t tmem_base_s;
  __shared__ unsigned ticket_s;

  const int t = threadIdx.x, warp = t >> 5;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int k_begin = blockIdx.z * a.k_per_split;
  const int k_end = min(a.Kd, k_begin + a.k_per_split);
  const int num_kb = (k_end - k_begin + KELEMS - 1) / KELEMS;
  const uint32_t smem_a = (smem_u32(smem) + 1023u) & ~1023u, smem_b = smem_a + STAGES * A_STAGE;   // swizzle is address based

  if (t == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], PRODUCERS);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) tmem_alloc(&tmem_base_s, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = tmem_base_s;

  if (warp < 8) {
    // ================= producers =================
    int issued = 0;
    for (int kb = 0; kb < num_kb; ++kb) {
      const int s = kb % STAGES;
      if (kb >= STAGES) mbar_wait(&empty_bar[s], ((kb / STAGES) - 1) & 1, a.error);
      const int k0 = k_begin + kb * KELEMS;
      load_tile<AK, ES, BM>(smem_a + s * A_STAGE, a.a, a.lda, a.g, m0, a.M, k0, k_end, t);
      load_tile<BK_, ES, BN>(smem_b + s * B_STAGE, a.b, a.ldb, a.g, n0, a.N, k0, k_end, t);
      cp_async_commit();
      ++issued;
      if (issued >= STAGES - 1) {
        // the oldest in-flight stage has landed: make it visible to the tensor core and publish it
        cp_async_wait<STAGES - 2>();
        fence_proxy_async();
        mbar_arrive(&full_bar[(kb - (STAGES - 2)) % STAGES]);
      }
    }
    // drain: publish the last min(num_kb, STAGES-2) stages
    cp_async_wait<0>();
    fence_proxy_async();
    const int first_unpublished = max(0, num_kb - (STAGES - 2));
    for (int kb = first_unpublished; kb < num_kb; ++kb) mbar_arrive(&full_bar[kb % STAGES]);
  } else if ((t & 31) == 0) {
    // ================= MMA issuer (one thread) =================
    constexpr uint32_t idesc = make_idesc(BF16, is_mn(AK), is_mn(BK_), BN);
    // canonical layouts: K-major no-swizzle (LBO = k-chunk stride, SBO = 8-row-group stride);
    // bf16 MN-major no-swizzle (LBO = 8-k-group stride, SBO = mn-chunk stride);
    // fp32 MN-major 128B_BASE32B (LBO = mn-atom stride, SBO = 4-k-atom stride).
    constexpr bool A_SW = !BF16 && is_mn(AK), B_SW = !BF16 && is_mn(BK_);
    constexpr uint32_t A_LBO = A_SW ? 512 : (is_mn(AK) ? (BM / T) * 128 : BM * 16);
    constexpr uint32_t B_LBO = B_SW ? 512 : (is_mn(BK_) ? (BN / T) * 128 : BN * 16);
    constexpr uint32_t A_SBO = A_SW ? (BM / 32) * 512 : 128, B_SBO = B_SW ? (BN / 32) * 512 : 128;
    constexpr uint32_t A_STEP = A_SW ? 2 * A_SBO : (is_mn(AK) ? (UMMA_K / 8) * A_LBO : 2 * A_LBO);
    constexpr uint32_t B_STEP = B_SW ? 2 * B_SBO : (is_mn(BK_) ? (UMMA_K / 8) * B_LBO : 2 * B_LBO);
    for (int kb = 0; kb < num_kb; ++kb) {
      const int s = kb % STAGES;
      mbar_wait(&full_bar[s], (kb / STAGES) & 1, a.error);
      tc_fence_after();
#pragma unroll
      for (int j = 0; j < KELEMS / UMMA_K; ++j) {
        const uint64_t da = make_desc(smem_a + s * A_STAGE + j * A_STEP, A_LBO, A_SBO, A_SW ? 1u : 0u);
        const uint64_t db = make_desc(smem_b + s * B_STAGE + j * B_STEP, B_LBO, B_SBO, B_SW ? 1u : 0u);
        umma<BF16>(tmem_d, da, db, idesc, (kb | j) ? 1u : 0u);
      }
      umma_commit(&empty_bar[s]);          // frees the stage when these MMAs have read it
    }
    umma_commit(&accum_bar);               // accumulator complete
  }

  // ================= epilogue (warps 0-7) =================
  if (warp < 8) {
    mbar_wait(&accum_bar, 0, a.error);
    tc_fence_after();
    const int lane_grp = warp & 3, col_half = warp >> 2;
    const int row = lane_grp * 32 + (t & 31);                   // accumulator row == TMEM lane
    constexpr int COLS_PER_WARP = BN / 2;
    constexpr int NCH = (COLS_PER_WARP + 15) / 16;
    const int col_base = col_half * COLS_PER_WARP;
    const int tiles = gridDim.x * gridDim.y;
    const int tile_id = blockIdx.y * gridDim.x + blockIdx.x;
    bool writer = true;
    float acc[NCH][16];
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      if (COLS_PER_WARP >= 16 || col_half == 0)
        tmem_ld16(tmem_d + ((uint32_t)(lane_grp * 32) << 16) + (uint32_t)(col_base + c * 16), acc[c]);
    }
    if (a.splits > 1) {
      float* mine = a.workspace + ((size_t)blockIdx.z * tiles + tile_id) * (BM * BN);
#pragma unroll
      for (int c = 0; c < NCH; ++c)
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int col = col_base + c * 16 + j;
          if (col < BN && (COLS_PER_WARP >= 16 || col_half == 0)) mine[(size_t)col * BM + row] = acc[c][j];
        }
      __threadfence();
      asm volatile("bar.sync 1, 256;" ::: "memory");           // producers/epilogue warps only
      if (t == 0) ticket_s = atomicAdd(&a.counters[tile_id], 1u);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      writer = ticket_s == (unsigned)(a.splits - 1);
      if (writer) {
        __threadfence();
#pragma unroll
        for (int c = 0; c < NCH; ++c)
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[c][j] = 0.f;
        for (int s = 0; s < a.splits; ++s) {
          const float* part = a.workspace + ((size_t)s * tiles + tile_id) * (BM * BN);
#pragma unroll
          for (int c = 0; c < NCH; ++c)
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int col = col_base + c * 16 + j;
              if (col < BN && (COLS_PER_WARP >= 16 || col_half == 0)) acc[c][j] += __ldcg(part + (size_t)col * BM + row);
            }
        }
        if (t == 0) a.counters[tile_id] = 0u;
      }
    }
    if (writer) {
      const int grow = m0 + row;
      const bool row_ok = grow < a.M;
      float rsum = 0.f, rsq = 0.f;
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        if (!(COLS_PER_WARP >= 16 || col_half == 0)) continue;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int col = col_base + c * 16 + j;
          const int gcol = n0 + col;
          const bool ok = row_ok && col < BN && gcol < a.N;
          float v = acc[c][j];
          if (a.bias != nullptr && ok) v += __ldg(a.bias + (a.transpose_out ? grow : gcol));
          if (ok) {
            float* dst = a.transpose_out ? a.out + (size_t)gcol * a.ldo + grow : a.out + (size_t)grow * a.ldo + gcol;
            *dst = a.accumulate ? *dst + v : v;
          }
          if (a.stats != nullptr) {
            const float u = ok ? v : 0.f;
            if (a.transpose_out) {            // channel == row: accumulate along this thread's columns
              rsum += u;
              rsq += u * u;
            } else {                          // channel == column: reduce over the warp's 32 rows
              const float cs = warp_sum(u), cq = warp_sum(u * u);
              if ((t & 31) == 0 && col < BN && gcol < a.N) {
                atomicAdd(a.stats + gcol, cs);
                atomicAdd(a.stats + a.N + gcol, cq);
              }
            }
          }
        }
      }
      if (a.stats != nullptr && a.transpose_out && row_ok) {
        atomicAdd(a.stats + grow, rsum);
        atomicAdd(a.stats + a.M + grow, rsq);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem_d, TMEM_COLS);
}

// ---- host side ------------------------------------------------------------------------------------
template <int AK, int BK_, int BN, bool BF16>
static int launch(const Args& a, dim3 grid, cudaStream_t st) {
  constexpr int smem = STAGES * (BM + BN) * ROW_BYTES + 1024;
  static bool configured = false;     // per template instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_kernel<AK, BK_, BN, BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  gemm_kernel<AK, BK_, BN, BF16><<<grid, THREADS, smem, st>>>(a);
  return static_cast<int>(cudaPeekAtLastError());
}

template <int AK, int BK_, bool BF16>
static int dispatch_bn(Args a, size_t workspace_bytes, int max_counters, int force_bn, cudaStream_t st) {
  constexpr int ES = BF16 ? 2 : 4;
  constexpr int KELEMS = ROW_BYTES / ES;
  int bn = force_bn ? force_bn : (a.N <= 32 ? 32 : (a.N <= 64 ? 64 : 128));
  const int tm = ceil_div(a.M, BM), tn = ceil_div(a.N, bn);
  const int tiles = tm * tn;
  const int k_blocks = ceil_div(a.Kd, KELEMS);
  int splits = 1;
  if (tiles < kNumSMs && a.workspace && a.counters && tiles <= max_counters && k_blocks >= 4) {
    splits = min(min(kNumSMs / tiles, k_blocks / 2), 32);
    const size_t per_split = (size_t)tiles * BM * bn * sizeof(float);
    if ((size_t)splits * per_split > workspace_bytes) splits = (int)(workspace_bytes / per_split);
    splits = max(splits, 1);
  }
  const int kps = ceil_div(k_blocks, splits) * KELEMS;
  a.splits = ceil_div(a.Kd, kps);
  a.k_per_split = kps;
  dim3 grid(tm, tn, a.splits);
  if (bn == 32) return launch<AK, BK_, 32, BF16>(a, grid, st);
  if (bn == 64) return launch<AK, BK_, 64, BF16>(a, grid, st);
  return launch<AK, BK_, 128, BF16>(a, grid, st);
}

}  // namespace tc
}  // namespace b200

using namespace b200;
using namespace b200::tc;

static bool ptr16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Dense GEMM, fp32 (TF32 math) or bf16 inputs, fp32 output:
//   mode 0 (NT): out[M,N] = A[M,K]  * B[N,K]^T      (Linear forward:  x, W)
//   mode 1 (NN): out[M,N] = A[M,K]  * B[K,N]        (Linear dgrad:    dy, W)
//   mode 2 (TN): out[M,N] = A[K,M]^T * B[K,N]       (Linear wgrad:    dy^T, x)
B200_API int b200_gemm_tc(const void* A, const void* B, float* out, const float* bias, int M, int N, int K, int mode,
                          int bf16, int accumulate, float* workspace, size_t workspace_bytes, unsigned* counters,
                          int max_counters, unsigned* error, cudaStream_t stream) {
  const int vec = bf16 ? 8 : 4;
  if (!ptr16(A) || !ptr16(B) || M <= 0 || N <= 0 || K <= 0) return static_cast<int>(cudaErrorInvalidValue);
  Args a{};
  a.a = A; a.b = B; a.out = out; a.bias = bias; a.M = M; a.N = N; a.Kd = K; a.ldo = N; a.accumulate = accumulate;
  a.workspace = workspace; a.counters = counters; a.error = error;
  if (mode == 0) {
    if (K % vec) return static_cast<int>(cudaErrorInvalidValue);
    a.lda = K; a.ldb = K;
    return bf16 ? dispatch_bn<DENSE_K, DENSE_K, true>(a, workspace_bytes, max_counters, 0, stream)
                : dispatch_bn<DENSE_K, DENSE_K, false>(a, workspace_bytes, max_counters, 0, stream);
  }
  if (mode == 1) {
    if (K % vec || N % vec) return static_cast<int>(cudaErrorInvalidValue);
    a.lda = K; a.ldb = N;
    return bf16 ? dispatch_bn<DENSE_K, DENSE_MN, true>(a, workspace_bytes, max_counters, 0, stream)
                : dispatch_bn<DENSE_K, DENSE_MN, false>(a, workspace_bytes, max_counters, 0, stream);
  }
  if (mode == 2) {
    if (M % vec || N % vec) return static_cast<int>(cudaErrorInvalidValue);
    a.lda = M; a.ldb = N;
    return bf16 ? dispatch_bn<DENSE_MN, DENSE_MN, true>(a, workspace_bytes, max_counters, 0, stream)
                : dispatch_bn<DENSE_MN, DENSE_MN, false>(a, workspace_bytes, max_counters, 0, stream);
  }
  return static_cast<int>(cudaErrorInvalidValue);
}

static Geom make_geom(int Nb, int H, int W, int C, int K, int R, int S, int stride, int pad) {
  Geom g{};
  g.Nb = Nb; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.pad = pad;
  g.P = (H + 2 * pad - R) / stride + 1;
  g.Q = (W + 2 * pad - S) / stride + 1;
  return g;
}

// Returns cudaErrorNotSupported (801) when the shape needs the FFMA fallback.
B200_API int b200_conv2d_fprop_tc(const float* x, const float* w, float* y, float* stats, int Nb, int H, int W, int C,
                                  int K, int R, int S, int stride, int pad, float* workspace, size_t workspace_bytes,
                                  unsigned* counters, int max_counters, unsigned* error, cudaStream_t stream) {
  if (C % 4 || !ptr16(x) || !ptr16(w)) return static_cast<int>(cudaErrorNotSupported);
  Args a{};
  a.g = make_geom(Nb, H, W, C, K, R, S, stride, pad);
  const int pixels = Nb * a.g.P * a.g.Q;
  a.Kd = R * S * C; a.stats = stats; a.workspace = workspace; a.counters = counters; a.error = error; a.out = y;
  a.ldo = K;
  if (pixels >= 1024 || K < 128) {       // pixels fill the 128-row MMA dimension
    a.a = x; a.b = w; a.ldb = a.Kd; a.M = pixels; a.N = K;
    return dispatch_bn<IM2COL_K, DENSE_K, false>(a, workspace_bytes, max_counters, 0, stream);
  }
  a.a = w; a.b = x; a.lda = a.Kd; a.M = K; a.N = pixels; a.transpose_out = 1;   // swapped: channels are rows
  return dispatch_bn<DENSE_K, IM2COL_K, false>(a, workspace_bytes, max_counters, 0, stream);
}

B200_API int b200_conv2d_dgrad_tc(const float* dy, const float* w, float* dx, int Nb, int H, int W, int C, int K,
                                  int R, int S, int stride, int pad, float* workspace, size_t workspace_bytes,
                                  unsigned* counters, int max_counters, unsigned* error, cudaStream_t stream) {
  if (C % 4 || K % 4 || !ptr16(dy) || !ptr16(w)) return static_cast<int>(cudaErrorNotSupported);
  Args a{};
  a.g = make_geom(Nb, H, W, C, K, R, S, stride, pad);
  const int pixels = Nb * H * W;
  a.Kd = R * S * K; a.workspace = workspace; a.counters = counters; a.error = error; a.out = dx; a.ldo = C;
  if (pixels >= 1024 || C < 128) {
    a.a = dy; a.b = w; a.M = pixels; a.N = C;
    return dispatch_bn<IM2COL_T_K, WEIGHT_T_MN, false>(a, workspace_bytes, max_counters, 0, stream);
  }
  a.a = w; a.b = dy; a.M = C; a.N = pixels; a.transpose_out = 1;
  return dispatch_bn<WEIGHT_T_MN, IM2COL_T_K, false>(a, workspace_bytes, max_counters, 0, stream);
}

B200_API int b200_conv2d_wgrad_tc(const float* dy, const float* x, float* dw, int accumulate, int Nb, int H, int W,
                                  int C, int K, int R, int S, int stride, int pad, float* workspace,
                                  size_t workspace_bytes, unsigned* counters, int max_counters, unsigned* error,
                                  cudaStream_t stream) {
  if (C % 4 || K % 4 || !ptr16(dy) || !ptr16(x)) return static_cast<int>(cudaErrorNotSupported);
  Args a{};
  a.g = make_geom(Nb, H, W, C, K, R, S, stride, pad);
  a.a = dy; a.lda = K; a.b = x; a.M = K; a.N = R * S * C; a.Kd = Nb * a.g.P * a.g.Q;
  a.workspace = workspace; a.counters = counters; a.error = error; a.out = dw; a.ldo = a.N; a.accumulate = accumulate;
  return dispatch_bn<DENSE_MN, IM2COL_MN, false>(a, workspace_bytes, max_counters, 0, stream);
}
