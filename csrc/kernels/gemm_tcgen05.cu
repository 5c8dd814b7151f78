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
//                          layouts (K-major: SWIZZLE_128B; 16-bit MN-major: no-swizzle core
//                          matrices; 32-bit MN-major: SWIZZLE_128B_BASE32B), 4-stage ring, then
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
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdio.h>

#include "common.cuh"
#include "tcgen05.cuh"
#include "b200ddp.h"

namespace b200 {
namespace tc {

constexpr int BM = 128;
// Ring depth per tile width.  128-wide tiles run 3 stages (96 KB) so that two CTAs -- e.g. a weight-gradient kernel
// on a side stream next to the dgrad kernel of the following layer -- fit on one SM; the deep-layer convolutions this
// kernel serves are latency bound, and 3 x 32 KB per CTA on ~128 CTAs already covers HBM latency x bandwidth.
__host__ __device__ constexpr int stages_for(int bn) { return bn == 128 ? 3 : 4; }
// dynamic shared memory: the operand ring, re-used by the epilogue as [BM][BN+1] staging + a [64][BN+1] reduced slice
__host__ __device__ constexpr int smem_for(int bn) {
  return (stages_for(bn) * (128 + bn) * 128 > (128 + 64) * (bn + 1) * 4 ? stages_for(bn) * (128 + bn) * 128
                                                                        : (128 + 64) * (bn + 1) * 4) + 1024;
}
constexpr int PRODUCERS = 256;              // 8 warps
constexpr int THREADS = PRODUCERS + 32;     // + MMA warp
constexpr int ROW_BYTES = 128;              // bytes of K per tile row per stage (32 fp32 / 64 bf16)

enum Kind { DENSE_K = 0, IM2COL_K = 1, IM2COL_T_K = 2, DENSE_MN = 3, WEIGHT_T_MN = 4, IM2COL_MN = 5 };
__host__ __device__ constexpr bool is_mn(int kind) { return kind >= DENSE_MN; }

// Division by a launch-time constant as multiply-high + shift (valid for numerators < 2^31):
// the operand gathers decode (pixel, filter tap, channel) for every 16-byte chunk.
struct FastDiv {
  int d;
  unsigned mul, shr;
  __device__ __forceinline__ int div(int n) const { return d == 1 ? n : (int)(__umulhi((unsigned)n, mul) >> shr); }
  __device__ __forceinline__ void divmod(int n, int& q, int& r) const {
    q = div(n);
    r = n - q * d;
  }
};
static FastDiv make_fastdiv(int d) {
  FastDiv f{};
  f.d = d < 1 ? 1 : d;
  if (f.d > 1) {
    int lg = 0;
    while ((1ll << lg) < f.d) ++lg;                       // ceil(log2 d)
    const unsigned p = 31 + lg;
    f.mul = (unsigned)(((1ull << p) + (unsigned)f.d - 1) / (unsigned)f.d);
    f.shr = p - 32;
  }
  return f;
}

struct Geom {
  int Nb, H, W, C, K, R, S, P, Q, stride, pad;
  FastDiv fQ, fP, fC, fS, fW, fH, fK, fStride;
};

// GEMM whose epilogue IS the cross-GPU gradient reduction (north-star fusion):
// every rank computes its partial of each output tile; a tile's owner (tile_id % world) receives
// the peers' partials in its staging area over NVLink, sums them in rank order, scales by 1/world
// and pushes the finished tile into EVERY rank's copy of the output (the gradient-bucket slot).
// Tiles of different owners are in flight at once, so the transfer of tile i overlaps the MMA
// of tile i+1 on every GPU.  All pointers are peer mappings of symmetric-arena objects.
struct PeerReduce {
  float* out[8];          // each rank's output matrix (same offset in every arena)
  float* stage[8];        // each rank's staging area  [src rank][tile][BN][BM] fp32
  uint32_t* flags[8];     // each rank's flags         [tile][8] "partial from src arrived", then [tile] "tile final"
  uint32_t* epoch;        // local launch counter (bumped by peer_reduce_wait_kernel)
  int rank, world;
  float scale;
  int pad;
};

// GEMM + in-switch all-reduce in ONE persistent kernel (the NVLS form of the fusion above; bf16 or fp32 output
// that lives at the same offset of every rank's multicast-bound arena, e.g. a gradient-bucket slot):
//   phase A  every rank computes its partial of every 256 x BN tile into ITS copy of the output and, per 128-row half
//            tile, releases a flag on the half tile's owner rank (half % world);
//   phase B  when a CTA has run out of tiles, ALL its warps turn to the half tiles this rank owns (32-row chunks,
//            round-robin over the grid): wait for the world flags of the half tile, multimem.ld_reduce the chunk
//            through the multicast mapping (the switch adds the ranks' partials, fp32 accumulate), scale by 1/world,
//            multimem.st the result into every rank's copy, then release a "final" flag on every rank;
//   phase C  every CTA waits for the final flags of its share of ALL half tiles, so kernel completion means this
//            rank's copy is complete.  The last CTA to leave advances the launch epoch.
// One writer per element and one fixed reduction order in the switch: replicas end bit-identical.  No staging, no
// separate all-reduce launch, no entry barrier; early tiles are reduced while late tiles are still being computed.
struct NvlsReduce {
  unsigned char* mc_out;     // multicast mapping of the output matrix (element (r, c) at (r * ldo + c) * esize)
  uint32_t* flags[8];        // each rank's flag object as mapped here: [half tile][8] partial-in-place epochs,
                             // followed at word offset `final_off` by [half tile] final epochs
  uint32_t* epoch;           // local: [0] launch epoch, [1] CTA exit ticket
  unsigned long long timeout_ns;
  int rank, world;
  float scale;
  int final_off;
};

struct Args {
  const void* a;
  const void* b;
  float* out;
  long long lda, ldb, ldo;
  int M, N, Kd;
  Geom g;
  int splits, k_per_split;     // k_per_split in elements, multiple of the stage depth
  int cluster_reduce;          // the splits of a tile form one thread-block cluster and reduce through distributed smem
  float* workspace;
  unsigned* counters;
  const float* bias;           // per output CHANNEL (column, or row when transposed)
  float* stats;                // [2 * channels]
  int accumulate;
  int transpose_out;           // out[col * ldo + row]
  unsigned* error;
  int out_bf16, bias_bf16;     // output / bias element type (default fp32)
  // batched GEMM (attention): blockIdx.z = batch index z, element offsets (z / inner) * s0 + (z % inner) * s1
  int batch, inner;
  long long a_s0, a_s1, b_s0, b_s1, o_s0, o_s1;
  int tma_a, tma_b;            // operand is fetched by TMA: 1 tiled, 2 im2col (K-major), 3 fp32 MN-major tiled,
                               // 4 im2col fp32 MN-major (wgrad), 5 filter as {C, RS, K} boxes with flipped taps (dgrad)
  const void* filter3d;        // != nullptr: WEIGHT_T_MN operand is the original KRSC filter, fetched by mode 5
  int filter_k;                // its K (number of filters); its C is the tile's mn extent
  // window offset (r, s) of the running convolution -> tap of the ORIGINAL filter:
  //   ((f_r0 + f_rstep * r) * f_S + (f_s0 + f_sstep * s)); filter_rs = R*S of the original filter
  int f_r0, f_rstep, f_s0, f_sstep, f_S, filter_rs;
  PeerReduce peer;             // world > 1: fused all-reduce epilogue (needs splits == 1, row-major out)
  NvlsReduce nvls;             // world > 1: NVLS all-reduce fused into the persistent 2-CTA kernel
  // causal attention structure (batched calls): 1 = output tile is entirely above the diagonal -> skip it
  // (S = QK^T, dP = dO V^T); 2 = reduction index is the key, only keys <= query row matter (P V, dS K);
  // 3 = reduction index is the query, only queries >= key row matter (P^T dO, dS^T Q)
  int causal;
  // fused epilogue (persistent 2-CTA kernel, bf16 output; aux / aux_out are bf16 [M, N] with pitch ldo):
  //   1: aux_out = v, out = gelu(v)      2: out = v * gelu'(aux)      3: out = v + aux       (v = acc + bias)
  int epi;
  const void* aux;
  void* aux_out;
};

__device__ __forceinline__ void store_out(const Args& a, float* out, size_t idx, float v) {
  if (a.out_bf16) {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out) + idx;
    *o = __float2bfloat16_rn(a.accumulate ? __bfloat162float(*o) + v : v);
  } else {
    out[idx] = a.accumulate ? out[idx] + v : v;
  }
}
__device__ __forceinline__ float load_bias(const Args& a, int i) {
  return a.bias_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(a.bias)[i]) : __ldg(a.bias + i);
}

__host__ __device__ constexpr uint32_t make_idesc(bool bf16, bool a_mn, bool b_mn, int n) {
  return (1u << 4) | ((bf16 ? 1u : 2u) << 7) | ((bf16 ? 1u : 2u) << 10) | ((a_mn ? 1u : 0u) << 15) |
         ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

__device__ __forceinline__ float gelu_tanh(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  const float t = tanhf(u);
  const float du = 0.7978845608028654f * (1.f + 3.f * 0.044715f * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}
__device__ __forceinline__ void load_bf16x16(const __nv_bfloat16* p, float (&v)[16]) {
  const uint4 lo = reinterpret_cast<const uint4*>(p)[0], hi = reinterpret_cast<const uint4*>(p)[1];
  const __nv_bfloat162* a = reinterpret_cast<const __nv_bfloat162*>(&lo);
  const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&hi);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 x = __bfloat1622float2(a[j]), y = __bfloat1622float2(b[j]);
    v[2 * j] = x.x; v[2 * j + 1] = x.y; v[8 + 2 * j] = y.x; v[8 + 2 * j + 1] = y.y;
  }
}
__device__ __forceinline__ void unpack_bf16x16(const uint4& lo, const uint4& hi, float (&v)[16]) {
  const __nv_bfloat162* a = reinterpret_cast<const __nv_bfloat162*>(&lo);
  const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&hi);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 x = __bfloat1622float2(a[j]), y = __bfloat1622float2(b[j]);
    v[2 * j] = x.x; v[2 * j + 1] = x.y; v[8 + 2 * j] = y.x; v[8 + 2 * j + 1] = y.y;
  }
}
__device__ __forceinline__ void store_bf16x16(__nv_bfloat16* p, const float (&v)[16]) {
  uint32_t w[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
    w[j] = *reinterpret_cast<const uint32_t*>(&h);
  }
  reinterpret_cast<uint4*>(p)[0] = make_uint4(w[0], w[1], w[2], w[3]);
  reinterpret_cast<uint4*>(p)[1] = make_uint4(w[4], w[5], w[6], w[7]);
}

// ---- operand gathers ---------------------------------------------------------------------------
// One producer thread always works on the same tile row (K-major kinds) or the same 16-byte
// mn-chunk (MN-major kinds), so everything that depends on it -- pixel (n,p,q) or filter tap
// (r,s,c) -- is decoded ONCE in init(); per chunk only the reduction index is decoded, with
// multiply-high divisions.  ROWS = tile extent along mn (128 for A, BN for B); ES = element bytes.
template <int KIND, int ES, int ROWS>
struct TileLoader {
  static constexpr int T = 16 / ES;                       // elements per 16-byte chunk
  static constexpr int MNC = ROWS / T;                    // mn chunks per tile row (MN-major)
  static constexpr int PER = ROWS * 8 / PRODUCERS;        // chunks per thread per stage
  // K-major : 8 consecutive lanes read the 8 chunks (128 contiguous bytes) of one tile row, a thread
  //           keeps its chunk column kc = t%8 and owns rows t/8 + 32*i        -> 4 lines per warp load
  // MN-major: consecutive lanes read consecutive mn chunks of one k row, a thread keeps its mn chunk
  //           and owns k rows kfirst + KSTEP*i                                -> 1-4 lines per warp load
  static constexpr int KSTEP = PRODUCERS / MNC;
  static_assert(PER >= 1 && ROWS * 8 % PRODUCERS == 0, "tile too small for the producer count");

  const char* base;
  long long ld;
  int kc;                    // K-major: fixed 16-byte chunk column
  int slot, kfirst;          // MN-major: fixed mn chunk, first k row
  int mn[PER];
  bool ok[PER];
  int d0[PER], d1[PER], d2[PER];   // decoded row context (see init)

  __device__ __forceinline__ void init(const void* base_, long long ld_, const Geom& g, int mn0, int mn_limit, int t) {
    base = static_cast<const char*>(base_);
    ld = ld_;
    kc = t & 7;
    slot = t % MNC;
    kfirst = t / MNC;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      mn[i] = is_mn(KIND) ? mn0 + slot * T : mn0 + (t >> 3) + i * (PRODUCERS / 8);
      ok[i] = mn[i] < mn_limit;
      d0[i] = d1[i] = d2[i] = 0;
      if (KIND == IM2COL_K) {               // d0 = image, d1 = p*stride - pad, d2 = q*stride - pad
        int t1, q, pp;
        g.fQ.divmod(mn[i], t1, q);
        g.fP.divmod(t1, d0[i], pp);
        d1[i] = pp * g.stride - g.pad;
        d2[i] = q * g.stride - g.pad;
      } else if (KIND == IM2COL_T_K) {      // d0 = image, d1 = h + pad, d2 = w + pad
        int t1, w, h;
        g.fW.divmod(mn[i], t1, w);
        g.fH.divmod(t1, d0[i], h);
        d1[i] = h + g.pad;
        d2[i] = w + g.pad;
      } else if (KIND == IM2COL_MN) {       // d0 = channel, d1 = r - pad, d2 = s - pad
        int rs, r, sx;
        g.fC.divmod(mn[i], rs, d0[i]);
        g.fS.divmod(rs, r, sx);
        d1[i] = r - g.pad;
        d2[i] = sx - g.pad;
      }
    }
  }

  __device__ __forceinline__ void load(uint32_t stage_smem, const Geom& g, int k0, int k_end) const {
    // K-major: the reduction index is the same for all of this thread's chunks -> decode once
    int kr_s = 0, kr_c = 0, kr_r = 0, kr_sx = 0;
    const int kk0 = k0 + kc * T;
    if (KIND == IM2COL_K) {
      g.fC.divmod(kk0, kr_s, kr_c);
      g.fS.divmod(kr_s, kr_r, kr_sx);
    } else if (KIND == IM2COL_T_K) {
      g.fK.divmod(kk0, kr_s, kr_c);
      g.fS.divmod(kr_s, kr_r, kr_sx);
    }
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const int kidx = kfirst + i * KSTEP;                // MN-major: k row of this chunk
      const int k = is_mn(KIND) ? k0 + kidx : kk0;
      bool valid = ok[i] && k < k_end;
      const char* src = base;
      if (valid) {
        if (KIND == DENSE_K) {
          src = base + ((long long)mn[i] * ld + k) * ES;
        } else if (KIND == DENSE_MN) {
          src = base + ((long long)k * ld + mn[i]) * ES;
        } else if (KIND == IM2COL_K) {
          const int h = d1[i] + kr_r, w = d2[i] + kr_sx;
          valid = h >= 0 && h < g.H && w >= 0 && w < g.W;
          if (valid) src = base + (((long long)(d0[i] * g.H + h) * g.W + w) * g.C + kr_c) * ES;
        } else if (KIND == IM2COL_T_K) {
          const int hh = d1[i] - kr_r, ww = d2[i] - kr_sx;
          valid = hh >= 0 && ww >= 0;
          if (valid) {
            int pp, q, rh, rw;
            g.fStride.divmod(hh, pp, rh);
            g.fStride.divmod(ww, q, rw);
            valid = rh == 0 && rw == 0 && pp < g.P && q < g.Q;
            if (valid) src = base + (((long long)(d0[i] * g.P + pp) * g.Q + q) * g.K + kr_c) * ES;
          }
        } else if (KIND == WEIGHT_T_MN) {
          int rs, co;
          g.fK.divmod(k, rs, co);
          src = base + (((long long)co * g.R * g.S + rs) * g.C + mn[i]) * ES;
        } else {                          // IM2COL_MN: k = output pixel
          int t1, q, nn, pp;
          g.fQ.divmod(k, t1, q);
          g.fP.divmod(t1, nn, pp);
          const int h = pp * g.stride + d1[i], w = q * g.stride + d2[i];
          valid = h >= 0 && h < g.H && w >= 0 && w < g.W;
          if (valid) src = base + (((long long)(nn * g.H + h) * g.W + w) * g.C + d0[i]) * ES;
        }
      }
      uint32_t off;
      if (!is_mn(KIND)) {
        // K-major SWIZZLE_128B: a tile row is 128 contiguous bytes, 8 rows form a 1024-byte atom,
        // the 16-byte chunk index is XOR-ed with (row % 8) (address bits [4,7) ^= [7,10))
        const int row = (int)(threadIdx.x >> 3) + i * (PRODUCERS / 8);
        off = (row >> 3) * 1024 + (row & 7) * 128 + ((kc ^ (row & 7)) << 4);
      } else if (ES == 2) {
        // 16-bit MN-major, no swizzle: chunk (mnc, k) at (k/8)*(MNC*128) + mnc*128 + (k%8)*16
        off = (kidx >> 3) * (MNC * 128) + slot * 128 + (kidx & 7) * 16;
      } else {
        // 32-bit MN-major exists only as SWIZZLE_128B_BASE32B: atom = 32 mn (one 128-byte row) x 4 k
        // rows; inside a row the 32-byte units are XOR-ed with the row index (byte-address bits
        // [5,7) ^= [7,9)).  Atoms tile MN first (512 B apart), then K.  (Layout measured on B200.)
        const int c = slot & 7, kr = kidx & 3;
        off = (slot >> 3) * 512 + (kidx >> 2) * ((ROWS / 32) * 512) + kr * 128 + ((((c >> 1) ^ kr) << 5) | ((c & 1) << 4));
      }
      cp_async16(stage_smem + off, src, valid);
    }
  }
};

template <int AK, int BK_, int BN, bool BF16>
__global__ void __launch_bounds__(THREADS, 1) gemm_kernel(const Args a, const __grid_constant__ CUtensorMap map_a,
                                                          const __grid_constant__ CUtensorMap map_b) {
  constexpr int ES = BF16 ? 2 : 4;
  constexpr int T = 16 / ES;
  constexpr int KELEMS = ROW_BYTES / ES;            // per stage
  constexpr int UMMA_K = 32 / ES;                   // 8 (tf32) / 16 (bf16)
  constexpr int A_STAGE = BM * ROW_BYTES, B_STAGE = BN * ROW_BYTES;
  constexpr int STAGES = stages_for(BN);
  constexpr uint32_t TMEM_COLS = BN < 32 ? 32 : BN;          // 32 / 64 / 128 / 256: powers of two
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES], accum_bar, cl_ready, cl_done;
  __shared__ uint32_t tmem_base_s;
  __shared__ unsigned ticket_s;
  __shared__ int peer_ok_s;          // fused peer reduction: 0 once a peer's partial tile failed to arrive in time

  pdl_launch_dependents();            // the next kernel in the stream may set up while this one runs (common.cuh)
  const int t = threadIdx.x, warp = t >> 5;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int zb = a.batch > 1 ? (int)blockIdx.z : 0;                 // batch index (no split-K when batched)
  const int split_z = a.batch > 1 ? 0 : (int)blockIdx.z;
  const long long zo = zb / a.inner, zi = zb % a.inner;
  const char* a_base = static_cast<const char*>(a.a) + (zo * a.a_s0 + zi * a.a_s1) * ES;
  const char* b_base = static_cast<const char*>(a.b) + (zo * a.b_s0 + zi * a.b_s1) * ES;
  float* out_base = reinterpret_cast<float*>(reinterpret_cast<char*>(a.out) +
                                             (zo * a.o_s0 + zi * a.o_s1) * (a.out_bf16 ? 2 : 4));
  if (a.causal == 1 && n0 >= m0 + BM) {                            // whole tile masked: nothing reads it
    pdl_wait();                                                    // (still a link of the completion chain)
    return;
  }
  int k_begin = split_z * a.k_per_split;
  int k_end = min(a.Kd, k_begin + a.k_per_split);
  if (a.causal == 2) k_end = min(k_end, m0 + BM);                   // keys beyond the last query row of the tile are masked
  if (a.causal == 3) k_begin = max(k_begin, (m0 / KELEMS) * KELEMS);  // queries before the first key row see nothing
  const int num_kb = (k_end - k_begin + KELEMS - 1) / KELEMS;
  const uint32_t smem_a = (smem_u32(smem) + 1023u) & ~1023u, smem_b = smem_a + STAGES * A_STAGE;   // swizzle is address based

  if (t == 32) {
    if (a.tma_a) tma_prefetch_desc(&map_a);
    if (a.tma_b) tma_prefetch_desc(&map_b);
  }
  if (t == 0) {
    peer_ok_s = 1;
    for (int s = 0; s < STAGES; ++s) {
      // arrivals per stage: the 256 gather threads (if any operand is gathered) + the TMA issuer
      mbar_init(&full_bar[s], ((a.tma_a && a.tma_b) ? 0 : PRODUCERS) + ((a.tma_a || a.tma_b) ? 1 : 0));
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&accum_bar, 1);
    if (a.cluster_reduce) {
      mbar_init(&cl_ready, a.splits);
      mbar_init(&cl_done, a.splits);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) tmem_alloc(&tmem_base_s, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  if (a.cluster_reduce) cluster_sync_all();                 // every CTA's cl_* barriers exist before anyone signals them
  tc_fence_after();
  const uint32_t tmem_d = tmem_base_s;
  // everything above (barrier init, TMEM allocation, descriptor prefetch) touched no global memory and may have
  // overlapped the previous kernel; from here on its results are needed
  pdl_wait();

  if (warp < 8) {
    // ================= producers =================
    TileLoader<AK, ES, BM> la;
    TileLoader<BK_, ES, BN> lb;
    const bool gather_a = !a.tma_a, gather_b = !a.tma_b;
    if (gather_a) la.init(a_base, a.lda, a.g, m0, a.M, t);
    if (gather_b) lb.init(b_base, a.ldb, a.g, n0, a.N, t);
    const bool any_gather = gather_a || gather_b;
    // TMA boxes: K-major operand = one box [ROWS x 128 B]; bf16 MN-major operand = ROWS/64 boxes of
    // [64 k rows x 128 B of mn], 8 KB each.  Both land in the SWIZZLE_128B layout the MMA expects.
    // wgrad activations by im2col-mode TMA (tma_b == 4): one [32 pixels x 32 channels] box per 32 output columns
    // (= 32 channels of one filter tap); boxes past the last column are not fetched
    const int b_boxes = a.tma_b == 4 ? min(BN / 32, (a.N - n0 + 31) / 32) : 0;
    const uint32_t tx_bytes = (a.tma_a ? (uint32_t)A_STAGE : 0u) +
                              (a.tma_b == 4 ? (uint32_t)b_boxes * 4096u : (a.tma_b ? (uint32_t)B_STAGE : 0u));
    // im2col-mode TMA: window corner (w, h, n) of the tile's first output pixel
    int im_w_a = 0, im_h_a = 0, im_n_a = 0, im_w_b = 0, im_h_b = 0, im_n_b = 0;
    if (t == 0 && a.tma_a == 2) {
      int t1, q, p;
      a.g.fQ.divmod(m0, t1, q);
      a.g.fP.divmod(t1, im_n_a, p);
      im_w_a = q * a.g.stride - a.g.pad;
      im_h_a = p * a.g.stride - a.g.pad;
    }
    if (t == 0 && a.tma_b == 2) {
      int t1, q, p;
      a.g.fQ.divmod(n0, t1, q);
      a.g.fP.divmod(t1, im_n_b, p);
      im_w_b = q * a.g.stride - a.g.pad;
      im_h_b = p * a.g.stride - a.g.pad;
    }
    if (any_gather || t == 0) {
      int issued = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES;
        if (kb >= STAGES) mbar_wait(&empty_bar[s], ((kb / STAGES) - 1) & 1, a.error);
        const int k0 = k_begin + kb * KELEMS;
        if (t == 0 && tx_bytes) {
          mbar_expect_tx(&full_bar[s], tx_bytes);
          if (a.tma_a == 2 || a.tma_b == 2 || a.tma_a == 5 || a.tma_b == 5) {   // k-block -> (filter tap, first channel)
            int rs, c0, r, sx;
            a.g.fC.divmod(k0, rs, c0);
            a.g.fS.divmod(rs, r, sx);
            // dgrad as a forward convolution of dY: the filter tap that meets window offset (r, s) is the mirrored one,
            // and the reduction "channel" c0 is the filter index k'
            const int rs_flip = (a.f_r0 + a.f_rstep * r) * a.f_S + (a.f_s0 + a.f_sstep * sx);
            if (a.tma_a == 5)
              for (int bx = 0; bx < BM / 32; ++bx) tma_load_3d(smem_a + s * A_STAGE + bx * 4096, &map_a, m0 + bx * 32, rs_flip, c0, &full_bar[s]);
            if (a.tma_b == 5)
              for (int bx = 0; bx < BN / 32; ++bx) tma_load_3d(smem_b + s * B_STAGE + bx * 4096, &map_b, n0 + bx * 32, rs_flip, c0, &full_bar[s]);
            if (a.tma_a == 2)
              tma_load_im2col_4d(smem_a + s * A_STAGE, &map_a, c0, im_w_a, im_h_a, im_n_a, (unsigned short)sx, (unsigned short)r, &full_bar[s]);
            if (a.tma_b == 2)
              tma_load_im2col_4d(smem_b + s * B_STAGE, &map_b, c0, im_w_b, im_h_b, im_n_b, (unsigned short)sx, (unsigned short)r, &full_bar[s]);
          }
          // fp32 MN-major by TMA: boxes of [32 k rows x 32 mn] (4 KB) in the SWIZZLE_128B_ATOM_32B pattern, which is
          // the 128B_BASE32B layout the tensor core reads (32-byte units XOR-ed with the k row, bits [5,7) ^= [7,9))
          if (a.tma_b == 4) {                              // k0 = first output pixel of this k-block
            int t1, q, p, img;
            a.g.fQ.divmod(k0, t1, q);
            a.g.fP.divmod(t1, img, p);
            const int w0 = q * a.g.stride - a.g.pad, h0 = p * a.g.stride - a.g.pad;
            for (int bx = 0; bx < b_boxes; ++bx) {
              int rs, c0, r, sx;
              a.g.fC.divmod(n0 + bx * 32, rs, c0);
              a.g.fS.divmod(rs, r, sx);
              tma_load_im2col_4d(smem_b + s * B_STAGE + bx * 4096, &map_b, c0, w0, h0, img, (unsigned short)sx, (unsigned short)r,
                                 &full_bar[s]);
            }
          }
          if (a.tma_a == 3)
            for (int bx = 0; bx < BM / 32; ++bx) tma_load_2d(smem_a + s * A_STAGE + bx * 4096, &map_a, m0 + bx * 32, k0, &full_bar[s]);
          if (a.tma_b == 3)
            for (int bx = 0; bx < BN / 32; ++bx) tma_load_2d(smem_b + s * B_STAGE + bx * 4096, &map_b, n0 + bx * 32, k0, &full_bar[s]);
          if (a.tma_a == 1) {
            if (!is_mn(AK)) tma_load_2d(smem_a + s * A_STAGE, &map_a, k0, m0, &full_bar[s]);
            else
              for (int bx = 0; bx < BM / 64; ++bx) tma_load_2d(smem_a + s * A_STAGE + bx * 8192, &map_a, m0 + bx * 64, k0, &full_bar[s]);
          }
          if (a.tma_b == 1) {
            if (!is_mn(BK_)) tma_load_2d(smem_b + s * B_STAGE, &map_b, k0, n0, &full_bar[s]);
            else
              for (int bx = 0; bx < BN / 64; ++bx) tma_load_2d(smem_b + s * B_STAGE + bx * 8192, &map_b, n0 + bx * 64, k0, &full_bar[s]);
          }
        }
        if (!any_gather) continue;
        if (gather_a) la.load(smem_a + s * A_STAGE, a.g, k0, k_end);
        if (gather_b) lb.load(smem_b + s * B_STAGE, a.g, k0, k_end);
        cp_async_commit();
        ++issued;
        if (issued >= STAGES - 1) {
          // the oldest in-flight stage has landed: make it visible to the tensor core and publish it
          cp_async_wait<STAGES - 2>();
          fence_proxy_async();
          mbar_arrive(&full_bar[(kb - (STAGES - 2)) % STAGES]);
        }
      }
      if (any_gather) {
        // drain: publish the last min(num_kb, STAGES-2) stages
        cp_async_wait<0>();
        fence_proxy_async();
        const int first_unpublished = max(0, num_kb - (STAGES - 2));
        for (int kb = first_unpublished; kb < num_kb; ++kb) mbar_arrive(&full_bar[kb % STAGES]);
      }
    }
  } else if ((t & 31) == 0) {
    // ================= MMA issuer (one thread) =================
    constexpr uint32_t idesc = make_idesc(BF16, is_mn(AK), is_mn(BK_), BN);
    // canonical layouts: K-major SWIZZLE_128B (SBO = 8-row atom stride, start address walks 32 B per MMA);
    // bf16 MN-major no-swizzle (LBO = 8-k-group stride, SBO = mn-chunk stride);
    // fp32 MN-major 128B_BASE32B (LBO = mn-atom stride, SBO = 4-k-atom stride).
    // layout type: K-major -> SWIZZLE_128B (2); fp32 MN-major -> 128B_BASE32B (1); bf16 MN-major -> none (0)
    constexpr bool A_SW = !BF16 && is_mn(AK), B_SW = !BF16 && is_mn(BK_);
    constexpr uint32_t A_TYPE = !is_mn(AK) ? 2u : (A_SW ? 1u : 0u), B_TYPE = !is_mn(BK_) ? 2u : (B_SW ? 1u : 0u);
    constexpr uint32_t A_LBO = !is_mn(AK) ? 16 : (A_SW ? 512 : (BM / T) * 128);
    constexpr uint32_t B_LBO = !is_mn(BK_) ? 16 : (B_SW ? 512 : (BN / T) * 128);
    constexpr uint32_t A_SBO = !is_mn(AK) ? 1024 : (A_SW ? (BM / 32) * 512 : 128);
    constexpr uint32_t B_SBO = !is_mn(BK_) ? 1024 : (B_SW ? (BN / 32) * 512 : 128);
    // start-address advance per MMA (K = 32 bytes of reduction)
    constexpr uint32_t A_STEP = !is_mn(AK) ? 32 : (A_SW ? 2 * A_SBO : (UMMA_K / 8) * A_LBO);
    constexpr uint32_t B_STEP = !is_mn(BK_) ? 32 : (B_SW ? 2 * B_SBO : (UMMA_K / 8) * B_LBO);
    for (int kb = 0; kb < num_kb; ++kb) {
      const int s = kb % STAGES;
      mbar_wait(&full_bar[s], (kb / STAGES) & 1, a.error);
      tc_fence_after();
#pragma unroll
      for (int j = 0; j < KELEMS / UMMA_K; ++j) {
        // a TMA-fetched bf16 MN-major operand sits in the SWIZZLE_128B MN-major layout instead:
        // 64-mn atoms 8 KB apart (LBO), 8-k-row atoms 1 KB apart (SBO), 2 k-atoms per MMA
        const bool a_mn_tma = is_mn(AK) && a.tma_a == 1, b_mn_tma = is_mn(BK_) && a.tma_b == 1;
        const uint64_t da = a.tma_a >= 3 ? make_desc(smem_a + s * A_STAGE + j * 1024, 4096, 512, 1u)
                          : a_mn_tma ? make_desc(smem_a + s * A_STAGE + j * 2048, 8192, 1024, 2u)
                                     : make_desc(smem_a + s * A_STAGE + j * A_STEP, A_LBO, A_SBO, A_TYPE);
        const uint64_t db = a.tma_b >= 3 ? make_desc(smem_b + s * B_STAGE + j * 1024, 4096, 512, 1u)
                          : b_mn_tma ? make_desc(smem_b + s * B_STAGE + j * 2048, 8192, 1024, 2u)
                                     : make_desc(smem_b + s * B_STAGE + j * B_STEP, B_LBO, B_SBO, B_TYPE);
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
    // the accumulator is streamed out of TMEM 16 columns at a time (keeps the epilogue at 16 live
    // accumulator registers even for 256-column tiles)
    const uint32_t tmem_row = tmem_d + ((uint32_t)(lane_grp * 32) << 16) + (uint32_t)col_base;
    if (a.splits > 1 && a.cluster_reduce) {
      // ---- split-K inside a thread-block cluster: partial tiles stay in shared memory; every split CTA sums its
      // slice of rows straight out of its peers' shared memory (DSMEM) -- no workspace round trip through L2, no
      // global fence, no ticket spin.
      constexpr int LDS = BN + 1;
      float* stage = reinterpret_cast<float*>(smem + (smem_a - smem_u32(smem)));      // my partial [BM][LDS]
      float* red = stage + BM * LDS;                                                 // my reduced slice [<= 64][LDS]
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        float acc[16];
        tmem_ld16(tmem_row + c * 16, acc);
#pragma unroll
        for (int j = 0; j < 16; ++j) stage[row * LDS + col_base + c * 16 + j] = acc[j];
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (t < a.splits) mbar_arrive_release_cluster(map_to_cta(smem_u32(&cl_ready), (uint32_t)t));
      mbar_wait_cluster(&cl_ready, 0, a.error);
      const int rows_per = (BM + a.splits - 1) / a.splits;
      const int r0 = split_z * rows_per;
      const int nrows = max(0, min(BM, r0 + rows_per) - r0);
      const uint32_t stage_u32 = smem_u32(stage);
      for (int idx = t; idx < nrows * BN; idx += PRODUCERS) {
        const int rr = idx / BN, col = idx - rr * BN;             // lanes walk columns: contiguous remote reads
        const uint32_t local = stage_u32 + (uint32_t)(((r0 + rr) * LDS + col) * 4);
        float v = 0.f;
        for (int sp = 0; sp < a.splits; ++sp) v += ld_dsmem_f32(map_to_cta(local, (uint32_t)sp));   // fixed order
        red[rr * LDS + col] = v;
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (t < a.splits) mbar_arrive_release_cluster(map_to_cta(smem_u32(&cl_done), (uint32_t)t));   // done reading peer t
      const int lane = t & 31;
      if (!a.transpose_out) {
        for (int rr = warp; rr < nrows; rr += 8) {
          const int gr = m0 + r0 + rr;
          if (gr >= a.M) break;
          for (int col = lane; col < BN; col += 32) {
            const int gc = n0 + col;
            if (gc < a.N) {
              float v = red[rr * LDS + col];
              if (a.bias != nullptr) v += load_bias(a, gc);
              store_out(a, out_base, (size_t)gr * a.ldo + gc, v);
            }
          }
        }
        if (a.stats != nullptr && t < BN && n0 + t < a.N) {
          const int rows = min(nrows, a.M - m0 - r0);
          float cs = 0.f, cq = 0.f;
          const float bias = a.bias != nullptr ? load_bias(a, n0 + t) : 0.f;
          for (int rr = 0; rr < rows; ++rr) {
            const float v = red[rr * LDS + t] + bias;
            cs += v;
            cq += v * v;
          }
          if (rows > 0) {
            atomicAdd(a.stats + n0 + t, cs);
            atomicAdd(a.stats + a.N + n0 + t, cq);
          }
        }
      } else {
        for (int idx = t; idx < nrows * BN; idx += PRODUCERS) {
          const int rr = idx % nrows, col = idx / nrows;
          const int gr = m0 + r0 + rr, gc = n0 + col;
          if (gr < a.M && gc < a.N) {
            float v = red[rr * LDS + col];
            if (a.bias != nullptr) v += load_bias(a, gr);
            store_out(a, out_base, (size_t)gc * a.ldo + gr, v);
          }
        }
        if (a.stats != nullptr && t < nrows && m0 + r0 + t < a.M) {
          const int cols = min(BN, a.N - n0);
          const float bias = a.bias != nullptr ? load_bias(a, m0 + r0 + t) : 0.f;
          float cs = 0.f, cq = 0.f;
          for (int col = 0; col < cols; ++col) {
            const float v = red[t * LDS + col] + bias;
            cs += v;
            cq += v * v;
          }
          atomicAdd(a.stats + m0 + r0 + t, cs);
          atomicAdd(a.stats + a.M + m0 + r0 + t, cq);
        }
      }
      mbar_wait_cluster(&cl_done, 0, a.error);               // my shared memory is no longer being read: safe to leave
      writer = false;
    } else if (a.splits > 1) {
      // ---- split-K: park the partial tile, then ALL splits of this tile reduce it together ----
      // Every split CTA sums one slice of rows over all parked partials and writes that slice of
      // the output, so the reduction cost per CTA is one tile of reads no matter how many splits
      // there are (a lone "last CTA reduces everything" step would serialise splits x 64 KB).
      // Needs the splits of a tile co-resident: the host keeps tiles*splits <= 128 CTAs (1 per SM).
      float* mine = a.workspace + ((size_t)split_z * tiles + tile_id) * (BM * BN);
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        float acc[16];
        tmem_ld16(tmem_row + c * 16, acc);
#pragma unroll
        for (int j = 0; j < 16; ++j) mine[(size_t)(col_base + c * 16 + j) * BM + row] = acc[j];
      }
      __threadfence();
      asm volatile("bar.sync 1, 256;" ::: "memory");           // producer/epilogue warps only
      unsigned* arrive = a.counters + 2 * tile_id;
      if (t == 0) {
        atomicAdd(arrive, 1u);
        unsigned spins = 0, seen;
        do {
          asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(arrive) : "memory");
          if (++spins > kWaitLimit) {
            if (a.error) atomicExch(a.error, 0xBAD05B17u);
            break;
          }
        } while (seen < (unsigned)a.splits);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const int rows_per = (BM + a.splits - 1) / a.splits;
      const int r0 = split_z * rows_per;
      const int nrows = max(0, min(BM, r0 + rows_per) - r0);
      float* stage = reinterpret_cast<float*>(smem + (smem_a - smem_u32(smem)));
      constexpr int LDS = BN + 1;
      const float* ws_tile = a.workspace + (size_t)tile_id * (BM * BN);
      const size_t split_stride = (size_t)tiles * (BM * BN);
      for (int idx = t; idx < nrows * BN; idx += PRODUCERS) {
        const int rr = idx % nrows, col = idx / nrows;
        const float* src = ws_tile + (size_t)col * BM + r0 + rr;
        float v = 0.f;
        for (int sp = 0; sp < a.splits; ++sp) v += __ldcg(src + sp * split_stride);
        stage[rr * LDS + col] = v;
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const int lane = t & 31;
      if (!a.transpose_out) {
        for (int rr = warp; rr < nrows; rr += 8) {
          const int gr = m0 + r0 + rr;
          if (gr >= a.M) break;
          for (int col = lane; col < BN; col += 32) {
            const int gc = n0 + col;
            if (gc < a.N) {
              float v = stage[rr * LDS + col];
              if (a.bias != nullptr) v += load_bias(a, gc);
              store_out(a, out_base, (size_t)gr * a.ldo + gc, v);
            }
          }
        }
        if (a.stats != nullptr && t < BN && n0 + t < a.N) {
          const int rows = min(nrows, a.M - m0 - r0);
          float cs = 0.f, cq = 0.f;
          const float bias = a.bias != nullptr ? load_bias(a, n0 + t) : 0.f;
          for (int rr = 0; rr < rows; ++rr) {
            const float v = stage[rr * LDS + t] + bias;
            cs += v;
            cq += v * v;
          }
          if (rows > 0) {
            atomicAdd(a.stats + n0 + t, cs);
            atomicAdd(a.stats + a.N + n0 + t, cq);
          }
        }
      } else {
        for (int idx = t; idx < nrows * BN; idx += PRODUCERS) {
          const int rr = idx % nrows, col = idx / nrows;
          const int gr = m0 + r0 + rr, gc = n0 + col;
          if (gr < a.M && gc < a.N) {
            float v = stage[rr * LDS + col];
            if (a.bias != nullptr) v += load_bias(a, gr);
            store_out(a, out_base, (size_t)gc * a.ldo + gr, v);
          }
        }
        if (a.stats != nullptr && t < nrows && m0 + r0 + t < a.M) {
          const int cols = min(BN, a.N - n0);
          const float bias = a.bias != nullptr ? load_bias(a, m0 + r0 + t) : 0.f;
          float cs = 0.f, cq = 0.f;
          for (int col = 0; col < cols; ++col) {
            const float v = stage[t * LDS + col] + bias;
            cs += v;
            cq += v * v;
          }
          atomicAdd(a.stats + m0 + r0 + t, cs);
          atomicAdd(a.stats + a.M + m0 + r0 + t, cq);
        }
      }
      // the last split to leave re-arms the tile's counters for the next launch
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (t == 0) {
        const unsigned left = atomicAdd(arrive + 1, 1u);
        if (left == (unsigned)(a.splits - 1)) {
          arrive[0] = 0u;
          arrive[1] = 0u;
        }
      }
      writer = false;
    }
    if (a.peer.world > 1) {
      // ---- fused cross-GPU reduction of this tile ------------------------------------------------
      const PeerReduce& pr = a.peer;
      const int owner = tile_id % pr.world;
      const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(pr.epoch) + 1u;
      float* stage = reinterpret_cast<float*>(smem + (smem_a - smem_u32(smem)));
      constexpr int LDS = BN + 1;
      const int lane = t & 31;
      if (pr.rank != owner) {
        // partial tile -> owner's staging area ([col][row]: lanes write consecutive rows, 128 B per warp store)
        float* dst = pr.stage[owner] + ((size_t)pr.rank * tiles + tile_id) * (BM * BN);
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          float acc[16];
          tmem_ld16(tmem_row + c * 16, acc);
#pragma unroll
          for (int j = 0; j < 16; ++j) dst[(size_t)(col_base + c * 16 + j) * BM + row] = acc[j];
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (t == 0) st_release_sys(pr.flags[owner] + (size_t)tile_id * 8 + pr.rank, epoch);
      } else {
        // own partial -> shared memory, then add the peers' partials as they arrive
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          float acc[16];
          tmem_ld16(tmem_row + c * 16, acc);
#pragma unroll
          for (int j = 0; j < 16; ++j) stage[row * LDS + col_base + c * 16 + j] = acc[j];
        }
        if (t < pr.world && t != pr.rank) {
          const uint32_t* f = pr.flags[pr.rank] + (size_t)tile_id * 8 + t;
          unsigned spins = 0;
          while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
            if (++spins > kWaitLimit) {
              if (a.error) atomicExch(a.error, 0xBAD0FEEDu);
              peer_ok_s = 0;                                    // the tile is marked below instead of summed from stale data
              break;
            }
          }
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const bool peers_ok = peer_ok_s != 0;
        const float* mine = pr.stage[pr.rank] + (size_t)tile_id * (BM * BN);
        const size_t src_stride = (size_t)tiles * (BM * BN);
        for (int idx = t; idx < BM * BN; idx += PRODUCERS) {
          const int r = idx % BM, c = idx / BM;                 // lanes walk rows: coalesced staged reads
          float v = 0.f;
          for (int p = 0; p < pr.world; ++p)                    // fixed rank order
            v += (p == pr.rank) ? stage[r * LDS + c] : __ldcg(mine + p * src_stride + (size_t)c * BM + r);
          stage[r * LDS + c] = peers_ok ? v * pr.scale : __int_as_float(0x7fc00000);
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        // finished tile -> every rank's output, 128-byte row segments
        for (int rr = warp; rr < BM; rr += 8) {
          const int gr = m0 + rr;
          if (gr >= a.M) break;
          for (int col = lane; col < BN; col += 32) {
            const int gc = n0 + col;
            if (gc < a.N) {
              const float v = stage[rr * LDS + col];
              const size_t o = (size_t)gr * a.ldo + gc;
#pragma unroll
              for (int p = 0; p < 8; ++p)
                if (p < pr.world) {
                  if (a.out_bf16) reinterpret_cast<__nv_bfloat16*>(pr.out[p])[o] = __float2bfloat16_rn(v);
                  else pr.out[p][o] = v;
                }
            }
          }
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (t < pr.world) st_release_sys(pr.flags[t] + (size_t)tiles * 8 + tile_id, epoch);   // "tile final" on every rank
      }
      writer = false;
    }
    if (writer) {
      const int grow = m0 + row;
      const bool row_ok = grow < a.M;
      if (a.transpose_out) {
        // swapped orientation: rows are output channels, lanes walk consecutive channels, so the
        // stores of one accumulator column are already coalesced (out[col * ldo + row])
        float rsum = 0.f, rsq = 0.f;
        const float bias = (a.bias != nullptr && row_ok) ? load_bias(a, grow) : 0.f;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          float acc[16];
          tmem_ld16(tmem_row + c * 16, acc);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int gcol = n0 + col_base + c * 16 + j;
            if (row_ok && gcol < a.N) {
              const float v = acc[j] + bias;
              store_out(a, out_base, (size_t)gcol * a.ldo + grow, v);
              rsum += v;
              rsq += v * v;
            }
          }
        }
        if (a.stats != nullptr && row_ok) {
          atomicAdd(a.stats + grow, rsum);
          atomicAdd(a.stats + a.M + grow, rsq);
        }
      } else {
        // row-major output: stage the tile through (now idle) pipeline smem so that global stores
        // are 128-byte row segments and the BatchNorm column sums come from conflict-free smem reads
        float* stage = reinterpret_cast<float*>(smem + (smem_a - smem_u32(smem)));
        constexpr int LDS = BN + 1;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          float acc[16];
          tmem_ld16(tmem_row + c * 16, acc);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int col = col_base + c * 16 + j;
            const int gcol = n0 + col;
            float v = acc[j];
            if (a.bias != nullptr && gcol < a.N) v += load_bias(a, gcol);
            stage[row * LDS + col] = v;
          }
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const int lane = t & 31;
#pragma unroll 4
        for (int rr = 0; rr < BM / 8; ++rr) {
          const int r = warp * (BM / 8) + rr, gr = m0 + r;
          if (gr >= a.M) break;
#pragma unroll
          for (int col = lane; col < BN; col += 32) {
            const int gc = n0 + col;
            if (gc < a.N) store_out(a, out_base, (size_t)gr * a.ldo + gc, stage[r * LDS + col]);
          }
        }
        if (a.stats != nullptr && t < BN && n0 + t < a.N) {
          const int rows = min(BM, a.M - m0);
          float cs = 0.f, cq = 0.f;
          for (int r = 0; r < rows; ++r) {
            const float v = stage[r * LDS + t];
            cs += v;
            cq += v * v;
          }
          atomicAdd(a.stats + n0 + t, cs);
          atomicAdd(a.stats + a.N + n0 + t, cq);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem_d, TMEM_COLS);
}

// After a fused-reduction GEMM: wait until every tile of the output has been finalised in THIS rank's
// copy (owners signal "tile final" on all ranks), then advance the launch counter.
__global__ void __launch_bounds__(256) peer_reduce_wait_kernel(PeerReduce pr, int tiles, unsigned* error) {
  const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(pr.epoch) + 1u;
  for (int tile = threadIdx.x; tile < tiles; tile += blockDim.x) {
    const uint32_t* f = pr.flags[pr.rank] + (size_t)tiles * 8 + tile;
    unsigned spins = 0;
    while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
      if (++spins > kWaitLimit) {
        if (error) atomicExch(error, 0xBAD0D07Eu);
        break;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) *pr.epoch = epoch;
}


// =====================================================================================================
// 2-CTA (cta_group::2) dense GEMM: a cluster of two CTAs on one TPC computes a 256 x BN tile.
// Each CTA TMA-loads ITS 128 rows of A and ITS half of B (BN/2 rows); the leader CTA's single MMA
// thread issues tcgen05.mma.cta_group::2 (M = 256), which reads both halves of B through the pair, so
// per flop each SM pulls half as many operand bytes from L2 as the 1-CTA kernel -- the lever against
// the L2-bandwidth bound of 128-wide tiles.  Both CTAs' TMA loads complete on the LEADER's full
// barrier; tcgen05.commit multicasts "stage free" / "accumulator ready" to both CTAs.  Each CTA runs
// the epilogue for its own 128 rows out of its own TMEM.  TMA-only (dense operands), no split-K.
// =====================================================================================================
constexpr int T2_THREADS = 256;      // warp 0: TMA, warp 1: MMA (leader), warp 2: TMEM alloc, warps 4-7: epilogue
constexpr int T2P_THREADS = 384;     // persistent variant: warps 4-11 are the epilogue (two per TMEM lane quarter)

__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar_cluster) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar_cluster) : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols));
}
__device__ __forceinline__ void umma_commit2(uint64_t* bar) {   // arrive on this barrier in BOTH CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
template <bool BF16>
__device__ __forceinline__ void umma2(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  if (BF16) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
  }
}

template <int AK, int BK_, int BN, bool BF16, int ST>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(T2_THREADS, 1)
gemm2_kernel(const Args a, const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b) {
  constexpr int ES = BF16 ? 2 : 4;
  constexpr int KELEMS = ROW_BYTES / ES;
  constexpr int UMMA_K = 32 / ES;
  constexpr int BH = BN / 2;                                  // this CTA's share of the B tile
  constexpr int A_STAGE = BM * ROW_BYTES, B_STAGE = BH * ROW_BYTES;
  constexpr uint32_t TMEM_COLS = BN;
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t full_bar[ST], empty_bar[ST], accum_bar;
  __shared__ uint32_t tmem_base_s;

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const uint32_t cta = cluster_ctarank();                     // 0 = leader
  const int tile_m = blockIdx.x >> 1;
  const int m0 = tile_m * 2 * BM + (int)cta * BM, n0 = blockIdx.y * BN;
  const int num_kb = (a.Kd + KELEMS - 1) / KELEMS;
  const uint32_t smem_a = (smem_u32(smem) + 1023u) & ~1023u, smem_b = smem_a + ST * A_STAGE;

  if (t == 0) {
    for (int s = 0; s < ST; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) tmem_alloc2(&tmem_base_s, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                         // peer barriers exist before anyone signals them
  tc_fence_after();
  const uint32_t tmem_d = tmem_base_s;

  if (warp == 0 && lane == 0) {
    // ---- TMA producer (both CTAs): my A rows, my half of B; completion lands on the leader's barrier ----
    for (int kb = 0; kb < num_kb; ++kb) {
      const int s = kb % ST;
      if (kb >= ST) mbar_wait(&empty_bar[s], ((kb / ST) - 1) & 1, a.error);
      const int k0 = kb * KELEMS;
      const uint32_t bar = map_to_cta(smem_u32(&full_bar[s]), 0);
      if (cta == 0) mbar_expect_tx(&full_bar[s], 2u * (uint32_t)(A_STAGE + B_STAGE));
      if (!is_mn(AK)) tma_load_2d_2sm(smem_a + s * A_STAGE, &map_a, k0, m0, bar);
      else
        for (int bx = 0; bx < BM / 64; ++bx) tma_load_2d_2sm(smem_a + s * A_STAGE + bx * 8192, &map_a, m0 + bx * 64, k0, bar);
      const int nb0 = n0 + (int)cta * BH;
      if (!is_mn(BK_)) tma_load_2d_2sm(smem_b + s * B_STAGE, &map_b, k0, nb0, bar);
      else
        for (int bx = 0; bx < BH / 64; ++bx) tma_load_2d_2sm(smem_b + s * B_STAGE + bx * 8192, &map_b, nb0 + bx * 64, k0, bar);
    }
  } else if (warp == 1 && lane == 0 && cta == 0) {
    // ---- MMA issuer: one thread of the leader CTA drives both SMs ----
    constexpr uint32_t idesc = (1u << 4) | ((BF16 ? 1u : 2u) << 7) | ((BF16 ? 1u : 2u) << 10) |
                               ((is_mn(AK) ? 1u : 0u) << 15) | ((is_mn(BK_) ? 1u : 0u) << 16) |
                               ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
    for (int kb = 0; kb < num_kb; ++kb) {
      const int s = kb % ST;
      mbar_wait(&full_bar[s], (kb / ST) & 1, a.error);
      tc_fence_after();
#pragma unroll
      for (int j = 0; j < KELEMS / UMMA_K; ++j) {
        const uint64_t da = is_mn(AK) ? make_desc(smem_a + s * A_STAGE + j * 2048, 8192, 1024, 2u)
                                      : make_desc(smem_a + s * A_STAGE + j * 32, 16, 1024, 2u);
        const uint64_t db = is_mn(BK_) ? make_desc(smem_b + s * B_STAGE + j * 2048, 8192, 1024, 2u)
                                       : make_desc(smem_b + s * B_STAGE + j * 32, 16, 1024, 2u);
        umma2<BF16>(tmem_d, da, db, idesc, (kb | j) ? 1u : 0u);
      }
      umma_commit2(&empty_bar[s]);
    }
    umma_commit2(&accum_bar);
  } else if (warp >= 4) {
    // ---- epilogue (both CTAs, 4 warps = my 128 accumulator rows) ----
    mbar_wait(&accum_bar, 0, a.error);
    tc_fence_after();
    const int lane_grp = warp & 3;
    const int row = lane_grp * 32 + lane;
    float* stage = reinterpret_cast<float*>(smem + (smem_a - smem_u32(smem)));
    constexpr int LDS = BN + 1;
    const uint32_t tmem_row = tmem_d + ((uint32_t)(lane_grp * 32) << 16);
#pragma unroll 1
    for (int c = 0; c < BN / 16; ++c) {
      float acc[16];
      tmem_ld16(tmem_row + c * 16, acc);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int gcol = n0 + c * 16 + j;
        float v = acc[j];
        if (a.bias != nullptr && gcol < a.N) v += load_bias(a, gcol);
        stage[row * LDS + c * 16 + j] = v;
      }
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    float* out_base = a.out;
    for (int rr = 0; rr < BM / 4; ++rr) {
      const int r = lane_grp * (BM / 4) + rr, gr = m0 + r;
      if (gr >= a.M) break;
#pragma unroll
      for (int col = lane; col < BN; col += 32) {
        const int gc = n0 + col;
        if (gc < a.N) store_out(a, out_base, (size_t)gr * a.ldo + gc, stage[r * LDS + col]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                         // nobody leaves while the pair still talks to it
  if (warp == 2) tmem_dealloc2(tmem_d, TMEM_COLS);
}


// -----------------------------------------------------------------------------------------------------
// Persistent variant of the 2-CTA kernel: one cluster per TPC loops over output tiles; the accumulator
// is double-buffered in TMEM (2 x BN columns) so the epilogue of tile i (TMEM -> registers -> global,
// no shared-memory staging) overlaps the TMA/MMA main loop of tile i+1.
//   full/empty[ST]   : smem ring, continuous across tiles (as in gemm2_kernel)
//   tmem_full[2]     : MMA -> epilogue, multicast commit to both CTAs
//   tmem_empty[2]    : epilogue -> MMA, lives in the LEADER, 4 warps x 2 CTAs arrive (peer arrives remotely)
// -----------------------------------------------------------------------------------------------------


// Phases B and C of the NVLS-fused GEMM (see NvlsReduce): executed by every thread of every CTA of the persistent
// kernel once the CTA has no tiles left.
template <int BN>
__device__ __forceinline__ void nvls_reduce_phases(const Args& a, int tiles_n, int num_tiles, uint32_t epoch) {
  const NvlsReduce& nv = a.nvls;
  __shared__ int nv_ok_s;
  constexpr int CH = 32, QN = 128 / CH, U = 4;             // rows per work item, items per half tile, vectors in flight
  const int t = threadIdx.x, nthr = blockDim.x;
  const int halves = num_tiles * 2;
  const int esize = a.out_bf16 ? 2 : 4;
  const int owned = nv.rank < halves ? (halves - nv.rank + nv.world - 1) / nv.world : 0;
  const unsigned long long limit = nv.timeout_ns ? nv.timeout_ns : 600ull * 1000ull * 1000ull * 1000ull;
  for (int it = blockIdx.x; it < owned * QN; it += gridDim.x) {
    const int h = nv.rank + (it / QN) * nv.world, q = it % QN;
    if (t == 0) nv_ok_s = 1;
    __syncthreads();
    if (t < nv.world && !wait_flag_ns(nv.flags[nv.rank] + (size_t)h * 8 + t, epoch, limit)) {
      nv_ok_s = 0;
      if (a.error) atomicExch(a.error, 0xDEAD1000u | (unsigned)t);
    }
    __syncthreads();
    const bool ok = nv_ok_s != 0;
    const int tile = h >> 1;
    const int m0 = (tile / tiles_n) * 2 * BM + (h & 1) * BM + q * CH, n0 = (tile % tiles_n) * BN;
    const int vpr = min(BN, a.N - n0) * esize / 16;          // 16-byte vectors per row of this tile
    const int total = CH * vpr;
    for (int base = t; base < total; base += nthr * U) {
      uint4 v[U];
      size_t off[U];
      bool live[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = base + u * nthr;
        const int r = i / vpr, c = i - r * vpr;
        live[u] = i < total && m0 + r < a.M;
        off[u] = ((size_t)(m0 + r) * a.ldo + n0) * esize + (size_t)c * 16;
        if (live[u] && ok) v[u] = a.out_bf16 ? multimem_ld_reduce_bf16x8(nv.mc_out + off[u]) : multimem_ld_reduce_f32x4(nv.mc_out + off[u]);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!live[u]) continue;
        if (!ok) {          // a peer never delivered: mark the local result instead of passing a partial sum off as final
          *reinterpret_cast<uint4*>(reinterpret_cast<unsigned char*>(a.out) + off[u]) =
              make_uint4(0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u);
          continue;
        }
        uint4 r4;
        if (a.out_bf16) {
          const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
          uint32_t o[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const __nv_bfloat162 hh = __floats2bfloat162_rn(__uint_as_float(w[j] << 16) * nv.scale,
                                                            __uint_as_float(w[j] & 0xffff0000u) * nv.scale);
            o[j] = *reinterpret_cast<const uint32_t*>(&hh);
          }
          r4 = make_uint4(o[0], o[1], o[2], o[3]);
          multimem_st_bf16x8(nv.mc_out + off[u], r4);
        } else {
          r4 = make_uint4(__float_as_uint(__uint_as_float(v[u].x) * nv.scale), __float_as_uint(__uint_as_float(v[u].y) * nv.scale),
                          __float_as_uint(__uint_as_float(v[u].z) * nv.scale), __float_as_uint(__uint_as_float(v[u].w) * nv.scale));
          multimem_st_f32x4(nv.mc_out + off[u], r4);
        }
      }
    }
    __threadfence_system();
    __syncthreads();
    if (t < nv.world) st_release_sys(nv.flags[t] + nv.final_off + (size_t)h * QN + q, epoch);     // "chunk final" on every rank
  }
  // phase C: this rank's copy is complete when every chunk of every half tile has been finalised by its owner
  bool all = true;
  for (int c = blockIdx.x + t * gridDim.x; c < halves * QN; c += gridDim.x * nthr)
    if (!wait_flag_ns(nv.flags[nv.rank] + nv.final_off + c, epoch, limit)) all = false;
  if (!all && a.error) atomicExch(a.error, 0xDEAD2000u);
  __syncthreads();
  if (t == 0) {
    __threadfence();
    const unsigned k = atomicAdd(nv.epoch + 1, 1u);
    if (k == gridDim.x - 1) {            // last CTA out: every CTA has read the epoch, nobody needs the old value
      nv.epoch[1] = 0u;
      __threadfence();
      nv.epoch[0] = epoch;
    }
  }
}

template <int AK, int BK_, int BN, bool BF16, int ST>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(T2P_THREADS, 1)
gemm2p_kernel(const Args a, const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
              int tiles_n, int num_tiles) {
  constexpr int ES = BF16 ? 2 : 4;
  constexpr int KELEMS = ROW_BYTES / ES;
  constexpr int UMMA_K = 32 / ES;
  constexpr int BH = BN / 2;
  constexpr int A_STAGE = BM * ROW_BYTES, B_STAGE = BH * ROW_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BN;                      // two accumulator buffers
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t full_bar[ST], empty_bar[ST], tmem_full[2], tmem_empty[2];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float bias_s[2][BN];             // this tile's bias, staged once per tile by the epilogue warps

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const uint32_t cta = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int num_kb = (a.Kd + KELEMS - 1) / KELEMS;
  const uint32_t smem_a = (smem_u32(smem) + 1023u) & ~1023u, smem_b = smem_a + ST * A_STAGE;
  const uint32_t nv_epoch = a.nvls.world > 1 ? *reinterpret_cast<volatile uint32_t*>(a.nvls.epoch) + 1u : 0u;

  if (t == 96) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
  }
  if (t == 0) {
    for (int s = 0; s < ST; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full[b], 1);
      mbar_init(&tmem_empty[b], 16);                         // 8 epilogue warps in each of the 2 CTAs
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) tmem_alloc2(&tmem_base_s, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_d = tmem_base_s;

  if (warp == 0 && lane == 0) {
    // ---- TMA producer ----
    int kc = 0;                                               // running k-block counter across tiles
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int m0 = (tile / tiles_n) * 2 * BM + (int)cta * BM, n0 = (tile % tiles_n) * BN;
      for (int kb = 0; kb < num_kb; ++kb, ++kc) {
        const int s = kc % ST;
        if (kc >= ST) mbar_wait(&empty_bar[s], ((kc / ST) - 1) & 1, a.error);
        const int k0 = kb * KELEMS;
        const uint32_t bar = map_to_cta(smem_u32(&full_bar[s]), 0);
        if (cta == 0) mbar_expect_tx(&full_bar[s], 2u * (uint32_t)(A_STAGE + B_STAGE));
        if (!is_mn(AK)) tma_load_2d_2sm(smem_a + s * A_STAGE, &map_a, k0, m0, bar);
        else
          for (int bx = 0; bx < BM / 64; ++bx) tma_load_2d_2sm(smem_a + s * A_STAGE + bx * 8192, &map_a, m0 + bx * 64, k0, bar);
        const int nb0 = n0 + (int)cta * BH;
        if (!is_mn(BK_)) tma_load_2d_2sm(smem_b + s * B_STAGE, &map_b, k0, nb0, bar);
        else
          for (int bx = 0; bx < BH / 64; ++bx) tma_load_2d_2sm(smem_b + s * B_STAGE + bx * 8192, &map_b, nb0 + bx * 64, k0, bar);
      }
    }
  } else if (warp == 1 && lane == 0 && cta == 0) {
    // ---- MMA issuer (leader) ----
    constexpr uint32_t idesc = (1u << 4) | ((BF16 ? 1u : 2u) << 7) | ((BF16 ? 1u : 2u) << 10) |
                               ((is_mn(AK) ? 1u : 0u) << 15) | ((is_mn(BK_) ? 1u : 0u) << 16) |
                               ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
    int kc = 0, it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++it) {
      const int buf = it & 1;
      if (it >= 2) mbar_wait(&tmem_empty[buf], ((it >> 1) - 1) & 1, a.error);   // epilogues drained this buffer
      tc_fence_after();
      const uint32_t acc = tmem_d + (uint32_t)(buf * BN);
      for (int kb = 0; kb < num_kb; ++kb, ++kc) {
        const int s = kc % ST;
        mbar_wait(&full_bar[s], (kc / ST) & 1, a.error);
        tc_fence_after();
#pragma unroll
        for (int j = 0; j < KELEMS / UMMA_K; ++j) {
          const uint64_t da = is_mn(AK) ? make_desc(smem_a + s * A_STAGE + j * 2048, 8192, 1024, 2u)
                                        : make_desc(smem_a + s * A_STAGE + j * 32, 16, 1024, 2u);
          const uint64_t db = is_mn(BK_) ? make_desc(smem_b + s * B_STAGE + j * 2048, 8192, 1024, 2u)
                                         : make_desc(smem_b + s * B_STAGE + j * 32, 16, 1024, 2u);
          umma2<BF16>(acc, da, db, idesc, (kb | j) ? 1u : 0u);
        }
        umma_commit2(&empty_bar[s]);
      }
      umma_commit2(&tmem_full[buf]);
    }
  } else if (warp >= 4) {
    // ---- epilogue: TMEM -> registers -> global, 16 columns at a time.  8 warps: warp w reads TMEM lane
    // quarter w % 4 (hardware restriction) and the column half (w - 4) / 4 of the tile. ----
    constexpr int NCH = BN / 32;                              // 16-column chunks per warp
    const int lane_grp = warp & 3, half = (warp - 4) >> 2;
    const int row = lane_grp * 32 + lane;
    const uint32_t empty_leader0 = map_to_cta(smem_u32(&tmem_empty[0]), 0);
    const uint32_t empty_leader1 = map_to_cta(smem_u32(&tmem_empty[1]), 0);
    int it = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++it) {
      const int buf = it & 1;
      const int m0 = (tile / tiles_n) * 2 * BM + (int)cta * BM, n0 = (tile % tiles_n) * BN + half * (BN / 2);
      const int gr = m0 + row;
      if (a.bias != nullptr) {
        const int nt0 = (tile % tiles_n) * BN;
        for (int i = (warp - 4) * 32 + lane; i < BN; i += 256) bias_s[buf][i] = nt0 + i < a.N ? load_bias(a, nt0 + i) : 0.f;
        asm volatile("bar.sync 1, 256;" ::: "memory");       // the 8 epilogue warps only
      }
      // operands of the fused epilogue do not depend on the accumulator: fetch them while the MMAs run
      uint4 ax[2 * NCH];
      if (a.epi >= 2) {
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          const int gc0 = n0 + c * 16;
          if (gr < a.M && gc0 < a.N) {
            const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(a.aux) + (size_t)gr * a.ldo + gc0);
            ax[2 * c] = src[0];
            ax[2 * c + 1] = src[1];
          }
        }
      }
      mbar_wait(&tmem_full[buf], (it >> 1) & 1, a.error);
      tc_fence_after();
      const uint32_t tmem_row = tmem_d + ((uint32_t)(lane_grp * 32) << 16) + (uint32_t)(buf * BN + half * (BN / 2));
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        float acc[16];
        tmem_ld16(tmem_row + c * 16, acc);
        const int gc0 = n0 + c * 16;
        if (gr >= a.M || gc0 >= a.N) continue;
        if (a.bias != nullptr) {
          const float4* bs = reinterpret_cast<const float4*>(&bias_s[buf][half * (BN / 2) + c * 16]);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 b4 = bs[j];
            acc[4 * j] += b4.x; acc[4 * j + 1] += b4.y; acc[4 * j + 2] += b4.z; acc[4 * j + 3] += b4.w;
          }
        }
        const size_t o = (size_t)gr * a.ldo + gc0;
        const bool full = gc0 + 15 < a.N && !a.accumulate;
        if (a.epi != 0) {                                  // host guarantees bf16 output, N % 16 == 0, ldo % 8 == 0
          if (a.epi == 1) {
            store_bf16x16(reinterpret_cast<__nv_bfloat16*>(a.aux_out) + o, acc);
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] = gelu_tanh(acc[j]);
          } else {
            float x[16];
            unpack_bf16x16(ax[2 * c], ax[2 * c + 1], x);
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] = a.epi == 2 ? acc[j] * gelu_tanh_grad(x[j]) : acc[j] + x[j];
          }
          store_bf16x16(reinterpret_cast<__nv_bfloat16*>(a.out) + o, acc);
        } else if (full && a.out_bf16 && ((o & 7) == 0)) {
          store_bf16x16(reinterpret_cast<__nv_bfloat16*>(a.out) + o, acc);
        } else if (full && !a.out_bf16 && ((o & 3) == 0)) {
          float4* dst = reinterpret_cast<float4*>(a.out + o);
#pragma unroll
          for (int j = 0; j < 4; ++j) dst[j] = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (gc0 + j < a.N) store_out(a, a.out, o + j, acc[j]);
        }
      }
      // this warp is done reading buffer `buf`: tell the leader's MMA thread
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(buf ? empty_leader1 : empty_leader0);
      if (a.nvls.world > 1) {
        // NVLS fusion, phase A: this CTA's 128 rows of the tile are in this rank's copy of the output -- tell the owner
        __threadfence_system();
        asm volatile("bar.sync 2, 256;" ::: "memory");       // the 8 epilogue warps
        if (warp == 4 && lane == 0) {
          const int h = tile * 2 + (int)cta;
          st_release_sys(a.nvls.flags[h % a.nvls.world] + (size_t)h * 8 + a.nvls.rank, nv_epoch);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc2(tmem_d, TMEM_COLS);
  if (a.nvls.world > 1) nvls_reduce_phases<BN>(a, tiles_n, num_tiles, nv_epoch);
}

// wt[c][R-1-r][S-1-s][k] = w[k][r][s][c]: the filter of the convolution that computes dX from dY (stride 1).
// One 32 x 32 (k, c) tile per block through shared memory, both sides coalesced.
__global__ void __launch_bounds__(256) flip_transpose_filter_kernel(const float* __restrict__ w, float* __restrict__ wt, int K,
                                                                    int C, int RS) {
  __shared__ float tile[32][33];
  const int rs = blockIdx.z, k0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int i = ty; i < 32; i += 8) {
    const int k = k0 + i, c = c0 + tx;
    if (k < K && c < C) tile[i][tx] = w[((size_t)k * RS + rs) * C + c];
  }
  __syncthreads();
  const int rs_f = RS - 1 - rs;                 // (R-1-r) * S + (S-1-s)
  for (int i = ty; i < 32; i += 8) {
    const int c = c0 + i, k = k0 + tx;
    if (k < K && c < C) wt[((size_t)c * RS + rs_f) * K + k] = tile[tx][i];
  }
}

// Stride-2 dgrad by output parity: class (a, b) = input pixels (2i + a, 2j + b) is a stride-1 convolution of dY
// with the taps r = a + pad (mod 2), s = b + pad (mod 2) only.  Each class is computed into a compact
// [Nb, H2, W2, C] buffer by the forward kernels and interleaved into dX here (float4 per thread).
// tmp == nullptr: the class has no tap at all (e.g. three of the four classes of a 1x1 stride-2 filter) -> zeros.
__global__ void __launch_bounds__(256) scatter_parity_kernel(const float* __restrict__ tmp, float* __restrict__ dx, int Nb, int H,
                                                             int W, int C, int H2, int W2, int pa, int pb) {
  const int c4n = C >> 2;
  const size_t total = (size_t)Nb * H2 * W2 * c4n;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % c4n);
    size_t pix = i / c4n;
    const int j = (int)(pix % W2);
    pix /= W2;
    const int ii = (int)(pix % H2), n = (int)(pix / H2);
    const float4 v = tmp != nullptr ? reinterpret_cast<const float4*>(tmp)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    reinterpret_cast<float4*>(dx + (((size_t)n * H + 2 * ii + pa) * W + 2 * j + pb) * C)[c4] = v;
  }
}

// ---- host side ------------------------------------------------------------------------------------
// 2-D row-major global matrix [outer x inner] with `pitch` elements per row; box [box_outer x box_inner];
// 128-byte swizzle; out-of-range parts of a box are zero-filled.
static bool make_map(CUtensorMap* m, const void* base, bool bf16, long long inner, long long outer, long long pitch,
                     int box_inner, int box_outer) {
  return tma_make_map_2d(m, base, bf16, inner, outer, pitch, box_inner, box_outer);
}

static bool g_use_tma = true;
static bool g_cluster_reduce = false;  // split-K partials reduced through distributed shared memory (tuner-selected)
static int g_split_cap = 32;          // upper bound on the split-K factor (per-shape tuning knob, see ops/native.py)
static bool g_use_im2col_tma = true;
static bool g_use_s2_dgrad = true;    // stride-2 dgrad of large images by output parity (four stride-1 problems + interleave)
static bool g_use_filter3d = true;    // stride-1 dgrad reads the KRSC filter through a 3-D TMA descriptor (no re-layout kernel)
static bool g_use_mn32_tma = true;    // fp32 MN-major operands by TMA (SWIZZLE_128B_ATOM_32B)
static bool make_im2col_map(CUtensorMap* m, const void* x, bool bf16, const Geom& g, int channels, int pixels) {
  return tma_make_map_im2col(m, x, bf16, g.Nb, g.H, g.W, g.C, g.R, g.S, g.P, g.Q, g.stride, g.pad, channels, pixels);
}

template <int AK, int BK_, int BN, bool BF16>
static int launch(const Args& a_in, dim3 grid, cudaStream_t st) {
  constexpr int smem = smem_for(BN);
  constexpr int KELEMS = ROW_BYTES / (BF16 ? 2 : 4);
  Args a = a_in;
  CUtensorMap map_a{}, map_b{};
  a.tma_a = a.tma_b = 0;
  if (g_use_tma && a.batch <= 1) {
    // dense K-major: matrix [rows x K]; dense bf16 MN-major: matrix [K x mn], boxes of 64 mn x KELEMS k
    if (AK == DENSE_K) a.tma_a = make_map(&map_a, a.a, BF16, a.Kd, a.M, a.lda, KELEMS, BM);
    else if (AK == DENSE_MN && BF16) a.tma_a = make_map(&map_a, a.a, true, a.M, a.Kd, a.lda, 64, KELEMS);
    else if (AK == DENSE_MN && g_use_mn32_tma &&
             tma_make_map_2d(&map_a, a.a, false, a.M, a.Kd, a.lda, 32, KELEMS, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)) a.tma_a = 3;
    if (BK_ == DENSE_K) a.tma_b = make_map(&map_b, a.b, BF16, a.Kd, a.N, a.ldb, KELEMS, BN);
    else if (BK_ == DENSE_MN && BF16 && BN >= 64) a.tma_b = make_map(&map_b, a.b, true, a.N, a.Kd, a.ldb, 64, KELEMS);
    else if (BK_ == DENSE_MN && !BF16 && g_use_mn32_tma &&
             tma_make_map_2d(&map_b, a.b, false, a.N, a.Kd, a.ldb, 32, KELEMS, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)) a.tma_b = 3;
    if (!BF16 && a.filter3d != nullptr) {                   // required: the gather code cannot serve this operand
      const int rs = a.filter_rs;
      if (AK == WEIGHT_T_MN) { if (!tma_make_map_filter_3d(&map_a, a.filter3d, a.filter_k, rs, a.M)) return 801; a.tma_a = 5; }
      if (BK_ == WEIGHT_T_MN) { if (!tma_make_map_filter_3d(&map_b, a.filter3d, a.filter_k, rs, a.N)) return 801; a.tma_b = 5; }
    }
    if (BK_ == IM2COL_MN && !BF16 && g_use_im2col_tma && g_use_mn32_tma && a.g.C % 32 == 0 &&
        tma_make_map_im2col(&map_b, a.b, false, a.g.Nb, a.g.H, a.g.W, a.g.C, a.g.R, a.g.S, a.g.P, a.g.Q, a.g.stride, a.g.pad, 32,
                            KELEMS, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B))
      a.tma_b = 4;
    // convolution activations: im2col-mode TMA when a k-block never straddles a filter tap
    if (g_use_im2col_tma && a.g.C % KELEMS == 0) {
      if (AK == IM2COL_K && make_im2col_map(&map_a, a.a, BF16, a.g, KELEMS, BM)) a.tma_a = 2;
      if (BK_ == IM2COL_K && make_im2col_map(&map_b, a.b, BF16, a.g, KELEMS, BN)) a.tma_b = 2;
    }
  }
  if (a.filter3d != nullptr && !((a.tma_a == 5 && a.tma_b == 2) || (a.tma_a == 2 && a.tma_b == 5))) return 801;
  static bool configured = false;     // per template instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_kernel<AK, BK_, BN, BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  if (a.cluster_reduce) {
    // the splits of one output tile (grid.z) form a cluster
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = 1;
    attr.val.clusterDim.y = 1;
    attr.val.clusterDim.z = grid.z;
    cudaLaunchAttribute attrs[2];
    attrs[0] = attr;
    attrs[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attrs;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    static bool big_clusters = false;   // per template instantiation
    if (grid.z > 8 && !big_clusters) {
      cudaError_t e2 = cudaFuncSetAttribute(gemm_kernel<AK, BK_, BN, BF16>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
      if (e2 != cudaSuccess) return static_cast<int>(e2);
      big_clusters = true;
    }
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_kernel<AK, BK_, BN, BF16>, a, map_a, map_b);
    if (e != cudaSuccess) return static_cast<int>(e);
    return static_cast<int>(cudaPeekAtLastError());
  }
  {
    cudaError_t e = launch_pdl(gemm_kernel<AK, BK_, BN, BF16>, grid, dim3(THREADS), (size_t)smem, st, a, map_a, map_b);
    if (e != cudaSuccess) return static_cast<int>(e);
  }
  return static_cast<int>(cudaPeekAtLastError());
}


static int g_use_2cta = 2;       // 0 off, 1 one tile per cluster, 2 persistent with double-buffered TMEM

// 2-CTA launch: only for TMA-able dense operands, plain row-major output, no split-K / batching.
template <int AK, int BK_, int BN, bool BF16>
static int launch2(const Args& a_in, cudaStream_t st) {
  constexpr int ES = BF16 ? 2 : 4;
  constexpr int KELEMS = ROW_BYTES / ES;
  constexpr int BH = BN / 2;
  constexpr int ST = BN == 256 ? 6 : 8;                        // 32 KB resp. 24 KB per stage
  constexpr int smem = ST * (BM + BH) * ROW_BYTES + 1024;
  static_assert(smem <= 227 * 1024, "pipeline does not fit");
  static_assert(BM * (BN + 1) * 4 <= ST * (BM + BH) * ROW_BYTES, "epilogue staging does not fit");
  Args a = a_in;
  CUtensorMap map_a{}, map_b{};
  bool ok = true;
  if (AK == DENSE_K) ok &= make_map(&map_a, a.a, BF16, a.Kd, a.M, a.lda, KELEMS, BM);
  else ok &= make_map(&map_a, a.a, true, a.M, a.Kd, a.lda, 64, KELEMS);
  if (BK_ == DENSE_K) ok &= make_map(&map_b, a.b, BF16, a.Kd, a.N, a.ldb, KELEMS, BH);
  else ok &= make_map(&map_b, a.b, true, a.N, a.Kd, a.ldb, 64, KELEMS);
  if (!ok) return -1;
  static bool configured = false;
  auto kern = gemm2_kernel<AK, BK_, BN, BF16, ST>;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return static_cast<int>(e);
    configured = true;
  }
  if (g_use_2cta >= 2) {
    static bool configured_p = false;
    auto kp = gemm2p_kernel<AK, BK_, BN, BF16, ST>;
    if (!configured_p) {
      cudaError_t e = cudaFuncSetAttribute(kp, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (e != cudaSuccess) return static_cast<int>(e);
      configured_p = true;
    }
    const int tiles_n = ceil_div(a.N, BN), num_tiles = ceil_div(a.M, 2 * BM) * tiles_n;
    const int clusters = min(num_tiles, kNumSMs / 2);
    kp<<<dim3(2 * clusters), T2P_THREADS, smem, st>>>(a, map_a, map_b, tiles_n, num_tiles);
    return static_cast<int>(cudaPeekAtLastError());
  }
  dim3 grid(2 * ceil_div(a.M, 2 * BM), ceil_div(a.N, BN));
  kern<<<grid, T2_THREADS, smem, st>>>(a, map_a, map_b);
  return static_cast<int>(cudaPeekAtLastError());
}

template <int AK, int BK_, bool BF16>
static int try_2cta(const Args& a, cudaStream_t st) {
  constexpr bool a_ok = AK == DENSE_K || (AK == DENSE_MN && BF16);
  constexpr bool b_ok = BK_ == DENSE_K || (BK_ == DENSE_MN && BF16);
  if constexpr (a_ok && b_ok) {
    if (a.epi != 0 && (g_use_2cta < 2 || !a.out_bf16 || a.N % 16 || a.ldo % 8 || a.accumulate)) return -1;
    if (!g_use_2cta || !g_use_tma || a.batch > 1 || a.transpose_out || a.stats != nullptr || a.peer.world > 1) return -1;
    if (a.nvls.world > 1 && (g_use_2cta < 2 || a.epi != 0 || a.accumulate || a.bias != nullptr)) return -1;
    if (a.M < 512 || a.N < 128) return -1;
    // 256 x 256 or 256 x 128 cluster tiles: whichever needs less time on 74 clusters (rounds x tile width);
    // ties go to the wider tile (half the A traffic per flop)
    const long long rows = ceil_div(a.M, 2 * BM), t256 = rows * ceil_div(a.N, 256), t128 = rows * ceil_div(a.N, 128);
    const long long clusters = kNumSMs / 2;
    // (a 128-wide tile costs ~0.8 of a 256-wide one: twice the A traffic per flop runs into L2 bandwidth)
    const long long c256 = ((t256 + clusters - 1) / clusters) * 256, c128 = ((t128 + clusters - 1) / clusters) * 200;
    const int bn = (a.N >= 256 && c256 <= c128) ? 256 : 128;
    if ((bn == 256 ? t256 : t128) < 24) return -1;           // too few cluster tiles: keep 1-CTA + split-K
    if (bn == 256) return launch2<AK, BK_, 256, BF16>(a, st);
    return launch2<AK, BK_, 128, BF16>(a, st);
  } else {
    return -1;
  }
}

template <int AK, int BK_, bool BF16>
static int dispatch_bn(Args a, size_t workspace_bytes, int max_counters, int force_bn, cudaStream_t st) {
  constexpr int ES = BF16 ? 2 : 4;
  constexpr int KELEMS = ROW_BYTES / ES;
  if (a.batch < 1) a.batch = 1;
  if (a.inner < 1) a.inner = 1;
  if (!force_bn) {
    const int rc2 = try_2cta<AK, BK_, BF16>(a, st);
    if (rc2 >= 0) return rc2;
  }
  if (a.epi != 0) return static_cast<int>(cudaErrorNotSupported);   // fused epilogues exist only in the persistent kernel
  // 256-wide tiles halve the A traffic per flop (L2 bandwidth is what bounds 128x128 tiles); use them
  // when the problem is big enough to still fill the chip
  int bn = force_bn ? force_bn : (a.N <= 32 ? 32 : (a.N <= 64 ? 64 : 128));
  if (!force_bn && a.batch <= 1 && a.N >= 256 && (long long)ceil_div(a.M, BM) * ceil_div(a.N, 256) >= 2 * kNumSMs) bn = 256;
  const int tm = ceil_div(a.M, BM), tn = ceil_div(a.N, bn);
  const int tiles = tm * tn;
  const int k_blocks = ceil_div(a.Kd, KELEMS);
  int splits = 1;
  if (a.batch < 1) a.batch = 1;
  if (a.inner < 1) a.inner = 1;
  if (a.peer.world <= 1 && a.batch == 1 && tiles < 64 && a.workspace && a.counters && 2 * tiles <= max_counters && k_blocks >= 4) {
    // One CTA keeps at most 3 stages (96 KB) of loads in flight, i.e. ~60-100 GB/s at DRAM/L2
    // latency: streaming the operands fast needs MANY CTAs, so K is split until ~128 CTAs run
    // (all co-resident, one per SM: the splits of a tile wait for each other in the epilogue).
    splits = min(min(min(128 / tiles, k_blocks / 2), 32), g_split_cap);
    const size_t per_split = (size_t)tiles * BM * bn * sizeof(float);
    if ((size_t)splits * per_split > workspace_bytes) splits = (int)(workspace_bytes / per_split);
    splits = max(splits, 1);
  }
  const bool want_cluster = g_cluster_reduce && splits > 1 && bn <= 128 && a.peer.world <= 1;
  if (want_cluster) splits = min(splits, g_split_cap >= 16 ? 16 : 8);   // 8 = portable cluster size, 16 = hardware maximum
  const int kps = ceil_div(k_blocks, splits) * KELEMS;
  a.splits = ceil_div(a.Kd, kps);
  a.k_per_split = kps;
  a.cluster_reduce = want_cluster && a.splits > 1 && a.splits <= 16;
  dim3 grid(tm, tn, a.batch > 1 ? a.batch : a.splits);
  if (bn == 32) return launch<AK, BK_, 32, BF16>(a, grid, st);
  if (bn == 64) return launch<AK, BK_, 64, BF16>(a, grid, st);
  if (bn == 256) return launch<AK, BK_, 256, BF16>(a, grid, st);
  return launch<AK, BK_, 128, BF16>(a, grid, st);
}

}  // namespace tc
}  // namespace b200

using namespace b200;
using namespace b200::tc;

static bool ptr16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Switch the TMA operand path on/off (default on); off = every operand is gathered with cp.async.
B200_API void b200_gemm_tc_use_tma(int enabled) { b200::tc::g_use_tma = enabled != 0; }
// 2-CTA (cta_group::2, 256-row cluster tile) kernels for large dense GEMMs.
B200_API void b200_gemm_tc_use_s2_dgrad(int enabled) { b200::tc::g_use_s2_dgrad = enabled != 0; }
B200_API void b200_gemm_tc_use_filter3d(int enabled) { b200::tc::g_use_filter3d = enabled != 0; }
B200_API void b200_gemm_tc_use_mn32_tma(int enabled) { b200::tc::g_use_mn32_tma = enabled != 0; }
B200_API void b200_gemm_tc_split_cap(int cap) { b200::tc::g_split_cap = cap < 1 ? 1 : cap; }
B200_API void b200_gemm_tc_cluster_reduce(int enabled) { b200::tc::g_cluster_reduce = enabled != 0; }
B200_API void b200_gemm_tc_use_im2col_tma(int enabled) { b200::tc::g_use_im2col_tma = enabled != 0; }
B200_API void b200_gemm_tc_use_2cta(int mode) { b200::tc::g_use_2cta = mode; }   // 0 off, 1 simple, 2 persistent

// Dense (optionally batched) GEMM; inputs fp32 (TF32 math) or bf16, output fp32 or bf16:
//   mode 0 (NT): out[M,N] = A[M,K]  * B[N,K]^T      (Linear forward:  x, W)
//   mode 1 (NN): out[M,N] = A[M,K]  * B[K,N]        (Linear dgrad:    dy, W)
//   mode 2 (TN): out[M,N] = A[K,M]^T * B[K,N]       (Linear wgrad:    dy^T, x)
// desc = {lda, ldb, ldo, batch, inner, a_s0, a_s1, b_s0, b_s1, o_s0, o_s1} (elements); batch z addresses
// operand X at X + (z / inner) * x_s0 + (z % inner) * x_s1.  Leading dimensions must keep 16-byte
// alignment of every row (ld % 4 == 0 for fp32, % 8 for bf16).  flags: bit0 inputs bf16, bit1 output
// bf16, bit2 bias bf16, bit3 accumulate into out, bits 4-5 causal structure (see Args::causal).
B200_API int b200_gemm_tc_ex(const void* A, const void* B, void* out, const void* bias, int M, int N, int K, int mode,
                             int flags, const long long* desc, float* workspace, size_t workspace_bytes,
                             unsigned* counters, int max_counters, unsigned* error, cudaStream_t stream) {
  const bool bf16 = flags & 1;
  const int vec = bf16 ? 8 : 4;
  if (!ptr16(A) || !ptr16(B) || M <= 0 || N <= 0 || K <= 0) return static_cast<int>(cudaErrorInvalidValue);
  Args a{};
  a.a = A; a.b = B; a.out = static_cast<float*>(out); a.bias = static_cast<const float*>(bias);
  a.M = M; a.N = N; a.Kd = K;
  a.lda = desc[0]; a.ldb = desc[1]; a.ldo = desc[2];
  a.batch = (int)desc[3]; a.inner = (int)desc[4];
  a.a_s0 = desc[5]; a.a_s1 = desc[6]; a.b_s0 = desc[7]; a.b_s1 = desc[8]; a.o_s0 = desc[9]; a.o_s1 = desc[10];
  a.out_bf16 = (flags >> 1) & 1; a.bias_bf16 = (flags >> 2) & 1; a.accumulate = (flags >> 3) & 1;
  a.causal = (flags >> 4) & 3;
  a.workspace = workspace; a.counters = counters; a.error = error;
  if (a.lda % vec || a.ldb % vec) return static_cast<int>(cudaErrorInvalidValue);
  if (a.a_s0 % vec || a.a_s1 % vec || a.b_s0 % vec || a.b_s1 % vec) return static_cast<int>(cudaErrorInvalidValue);
#define B200_GEMM_MODE(AK, BKIND)                                                                    \
  return bf16 ? dispatch_bn<AK, BKIND, true>(a, workspace_bytes, max_counters, 0, stream)            \
              : dispatch_bn<AK, BKIND, false>(a, workspace_bytes, max_counters, 0, stream)
  if (mode == 0) { B200_GEMM_MODE(DENSE_K, DENSE_K); }
  if (mode == 1) { B200_GEMM_MODE(DENSE_K, DENSE_MN); }
  if (mode == 2) { B200_GEMM_MODE(DENSE_MN, DENSE_MN); }
#undef B200_GEMM_MODE
  return static_cast<int>(cudaErrorInvalidValue);
}

// b200_gemm_tc_ex with a fused epilogue (see Args::epi): epi 1 = GELU forward (aux_out receives the pre-activation),
// 2 = GELU backward (out = result * gelu'(aux)), 3 = residual add (out = result + aux).  bf16 output only; returns
// cudaErrorNotSupported (801) when the shape does not run on the persistent 2-CTA kernel -- the caller then uses the
// unfused kernels.
B200_API int b200_gemm_tc_epi(const void* A, const void* B, void* out, const void* bias, const void* aux, void* aux_out,
                              int epi, int M, int N, int K, int mode, int flags, const long long* desc, float* workspace,
                              size_t workspace_bytes, unsigned* counters, int max_counters, unsigned* error,
                              cudaStream_t stream) {
  const bool bf16 = flags & 1;
  if (!bf16 || !((flags >> 1) & 1) || epi < 1 || epi > 3) return static_cast<int>(cudaErrorNotSupported);
  if (!ptr16(A) || !ptr16(B) || !ptr16(out) || (epi == 1 ? !ptr16(aux_out) || !aux_out : !ptr16(aux) || !aux))
    return static_cast<int>(cudaErrorInvalidValue);
  Args a{};
  a.a = A; a.b = B; a.out = static_cast<float*>(out); a.bias = static_cast<const float*>(bias);
  a.M = M; a.N = N; a.Kd = K;
  a.lda = desc[0]; a.ldb = desc[1]; a.ldo = desc[2];
  a.batch = 1; a.inner = 1;
  a.out_bf16 = 1; a.bias_bf16 = (flags >> 2) & 1;
  a.workspace = workspace; a.counters = counters; a.error = error;
  a.epi = epi; a.aux = aux; a.aux_out = aux_out;
  if (a.lda % 8 || a.ldb % 8) return static_cast<int>(cudaErrorInvalidValue);
  if (mode == 0) return dispatch_bn<DENSE_K, DENSE_K, true>(a, workspace_bytes, max_counters, 0, stream);
  if (mode == 1) return dispatch_bn<DENSE_K, DENSE_MN, true>(a, workspace_bytes, max_counters, 0, stream);
  if (mode == 2) return dispatch_bn<DENSE_MN, DENSE_MN, true>(a, workspace_bytes, max_counters, 0, stream);
  return static_cast<int>(cudaErrorInvalidValue);
}

// Convenience form: contiguous operands, fp32 output.
B200_API int b200_gemm_tc(const void* A, const void* B, float* out, const float* bias, int M, int N, int K, int mode,
                          int bf16, int accumulate, float* workspace, size_t workspace_bytes, unsigned* counters,
                          int max_counters, unsigned* error, cudaStream_t stream) {
  long long desc[11] = {0, 0, N, 1, 1, 0, 0, 0, 0, 0, 0};
  if (mode == 0) { desc[0] = K; desc[1] = K; }
  else if (mode == 1) { desc[0] = K; desc[1] = N; }
  else { desc[0] = M; desc[1] = N; }
  return b200_gemm_tc_ex(A, B, out, bias, M, N, K, mode, (bf16 ? 1 : 0) | (accumulate ? 8 : 0), desc, workspace,
                         workspace_bytes, counters, max_counters, error, stream);
}

static Geom make_geom(int Nb, int H, int W, int C, int K, int R, int S, int stride, int pad) {
  Geom g{};
  g.Nb = Nb; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.pad = pad;
  g.P = (H + 2 * pad - R) / stride + 1;
  g.Q = (W + 2 * pad - S) / stride + 1;
  g.fQ = make_fastdiv(g.Q); g.fP = make_fastdiv(g.P); g.fC = make_fastdiv(C); g.fS = make_fastdiv(S);
  g.fW = make_fastdiv(W); g.fH = make_fastdiv(H); g.fK = make_fastdiv(K); g.fStride = make_fastdiv(stride);
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
  const bool pointwise = R == 1 && S == 1 && stride == 1 && pad == 0;   // a plain GEMM: X is [pixels, C] dense
  if (pixels >= 1024 || K < 128) {       // pixels fill the 128-row MMA dimension
    a.a = x; a.b = w; a.lda = C; a.ldb = a.Kd; a.M = pixels; a.N = K;
    if (pointwise) return dispatch_bn<DENSE_K, DENSE_K, false>(a, workspace_bytes, max_counters, 0, stream);
    return dispatch_bn<IM2COL_K, DENSE_K, false>(a, workspace_bytes, max_counters, 0, stream);
  }
  a.a = w; a.b = x; a.lda = a.Kd; a.ldb = C; a.M = K; a.N = pixels; a.transpose_out = 1;   // swapped: channels are rows
  if (pointwise) return dispatch_bn<DENSE_K, DENSE_K, false>(a, workspace_bytes, max_counters, 0, stream);
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
  const bool pointwise = R == 1 && S == 1 && stride == 1 && pad == 0;   // dX = dY[pixels,K] * W[K,C]
  // Stride-1 filters: dX is itself a convolution of dY (padding R-1-pad) with the flipped, transposed filter.
  // The filter is re-laid-out into the tail of the workspace (a few MB at most), after which the forward kernels
  // -- im2col-mode TMA for dY, tiled TMA for the filter -- do the work; no gather code runs on the SMs.
  if (!pointwise && g_use_im2col_tma && g_use_mn32_tma && g_use_tma && g_use_filter3d && stride == 1 && R == S &&
      R - 1 - pad >= 0 && K % 32 == 0 && C % 4 == 0) {
    // no re-layout at all: the original KRSC filter is fetched as {32 c, mirrored tap, 32 k'} boxes (MN-major operand)
    Args f{};
    f.g = make_geom(Nb, a.g.P, a.g.Q, K, C, R, S, 1, R - 1 - pad);
    f.Kd = R * S * K; f.workspace = workspace; f.counters = counters; f.error = error; f.out = dx; f.ldo = C;
    f.filter3d = w; f.filter_k = K;
    f.f_r0 = R - 1; f.f_rstep = -1; f.f_s0 = S - 1; f.f_sstep = -1; f.f_S = S; f.filter_rs = R * S;   // mirrored taps
    int rc;
    if (pixels >= 1024 || C < 128) {
      f.a = dy; f.b = w; f.lda = K; f.ldb = C; f.M = pixels; f.N = C;
      rc = dispatch_bn<IM2COL_K, WEIGHT_T_MN, false>(f, workspace_bytes, max_counters, 0, stream);
    } else {
      f.a = w; f.b = dy; f.lda = C; f.ldb = K; f.M = C; f.N = pixels; f.transpose_out = 1;
      rc = dispatch_bn<WEIGHT_T_MN, IM2COL_K, false>(f, workspace_bytes, max_counters, 0, stream);
    }
    if (rc != 801) return rc;
  }
  // Stride-2 filters on large images: the gather kernel multiplies three zero taps out of four.  Input pixels of one parity
  // class (h % 2, w % 2) only meet the filter taps r = a + pad, s = b + pad (mod 2), and for those dX is a stride-1
  // convolution of dY -- so each class goes through the TMA-fed forward kernels into a compact buffer at the head of
  // the workspace and is interleaved into dX afterwards.
  if (g_use_s2_dgrad && g_use_im2col_tma && g_use_mn32_tma && g_use_tma && g_use_filter3d && stride == 2 && R == S &&
      K % 32 == 0 && C % 4 == 0 && pixels >= 16384 && workspace != nullptr) {
    const int P = a.g.P, Q = a.g.Q;
    bool ok = true;
    // per class and axis: taps r0, r0 - 2, ... (n of them) reading dY index i + t, t = t_min .. t_min + n - 1
    struct Axis { int n, r_first, t_min, ext; } ax[2][2];            // [axis h/w][parity]
    for (int axis = 0; axis < 2 && ok; ++axis)
      for (int par = 0; par < 2; ++par) {
        const int dim = axis == 0 ? H : W;
        Axis& e = ax[axis][par];
        e.ext = (dim - par + 1) / 2;                                 // pixels of this parity along the axis
        e.n = 0; e.r_first = 0; e.t_min = 0;
        int t_max = 0;
        for (int r = 0; r < R; ++r) {
          if (((par + pad - r) & 1) != 0) continue;
          const int t = (par + pad - r) / 2;                         // exact: numerator is even
          if (e.n == 0) { e.t_min = t; t_max = t; }
          e.t_min = min(e.t_min, t);
          t_max = max(t_max, t);
          ++e.n;
        }
        if (e.n > 0) e.r_first = par + pad - 2 * e.t_min;            // tap that meets window offset 0
        (void)t_max;
      }
    const size_t tmp_bytes = (size_t)Nb * ax[0][0].ext * ax[1][0].ext * C * sizeof(float);   // parity (0,0) is the largest
    const size_t ws_left = workspace_bytes > tmp_bytes ? ((workspace_bytes - tmp_bytes) & ~(size_t)255) : 0;
    for (int pa = 0; pa < 2 && ok; ++pa)
      for (int pb = 0; pb < 2; ++pb)
        if (ax[0][pa].n > 0 && ax[1][pb].n > 0 && (ax[0][pa].t_min != ax[1][pb].t_min || ax[0][pa].t_min > 0))
          ok = false;                                                 // Geom carries one (non-negative) padding
    if (ok && tmp_bytes > 0 && ws_left >= (1u << 20)) {
      float* tmp = workspace;
      float* ws2 = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + ((tmp_bytes + 255) & ~(size_t)255));
      bool wrote = false, give_up = false;
      // classes that need a GEMM first, so that an unsupported configuration is detected before dX is touched
      for (int pass = 0; pass < 2 && !give_up; ++pass)
      for (int pa = 0; pa < 2 && !give_up; ++pa)
        for (int pb = 0; pb < 2; ++pb) {
          if ((pass == 0) != (ax[0][pa].n > 0 && ax[1][pb].n > 0)) continue;
          const Axis &eh = ax[0][pa], &ew = ax[1][pb];
          if (eh.ext == 0 || ew.ext == 0) continue;
          const size_t cells = (size_t)Nb * eh.ext * ew.ext * (C / 4);
          const size_t want = (cells + 255) / 256, cap = (size_t)(8 * kNumSMs);
          const int sg = (int)(want < cap ? want : cap);
          if (eh.n == 0 || ew.n == 0) {                               // no tap reaches this class: zeros
            scatter_parity_kernel<<<sg, 256, 0, stream>>>(nullptr, dx, Nb, H, W, C, eh.ext, ew.ext, pa, pb);
            continue;
          }
          Args f{};
          f.g = make_geom(Nb, P, Q, K, C, eh.n, ew.n, 1, -eh.t_min);  // dY is the input; eh.n x ew.n taps, stride 1
          f.g.P = eh.ext; f.g.Q = ew.ext;                             // output grid = this parity class
          f.g.fP = make_fastdiv(f.g.P); f.g.fQ = make_fastdiv(f.g.Q);
          f.Kd = eh.n * ew.n * K; f.workspace = ws2; f.counters = counters; f.error = error; f.out = tmp; f.ldo = C;
          f.filter3d = w; f.filter_k = K;
          f.f_r0 = eh.r_first; f.f_rstep = -2; f.f_s0 = ew.r_first; f.f_sstep = -2; f.f_S = S; f.filter_rs = R * S;
          f.a = dy; f.b = w; f.lda = K; f.ldb = C; f.M = Nb * eh.ext * ew.ext; f.N = C;
          const int rc = dispatch_bn<IM2COL_K, WEIGHT_T_MN, false>(f, ws_left, max_counters, 0, stream);
          if (rc == 801 && !wrote) { give_up = true; break; }         // descriptors unavailable: nothing written yet
          if (rc != 0) return rc;
          scatter_parity_kernel<<<sg, 256, 0, stream>>>(tmp, dx, Nb, H, W, C, eh.ext, ew.ext, pa, pb);
          wrote = true;
        }
      if (!give_up) return static_cast<int>(cudaPeekAtLastError());
    }
  }
  const size_t wt_bytes = (size_t)C * R * S * K * sizeof(float);
  if (!pointwise && g_use_im2col_tma && g_use_tma && stride == 1 && R == S && R - 1 - pad >= 0 && K % 32 == 0 && workspace &&
      2 * wt_bytes <= workspace_bytes) {
    const size_t ws_left = (workspace_bytes - wt_bytes) & ~(size_t)255;
    float* wt = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + ws_left);
    dim3 tg((C + 31) / 32, (K + 31) / 32, R * S);
    flip_transpose_filter_kernel<<<tg, 256, 0, stream>>>(w, wt, K, C, R * S);
    Args f{};
    f.g = make_geom(Nb, a.g.P, a.g.Q, K, C, R, S, 1, R - 1 - pad);        // input = dY [N,P,Q,K], C output channels
    f.Kd = R * S * K; f.workspace = workspace; f.counters = counters; f.error = error; f.out = dx; f.ldo = C;
    if (pixels >= 1024 || C < 128) {
      f.a = dy; f.b = wt; f.lda = K; f.ldb = f.Kd; f.M = pixels; f.N = C;
      return dispatch_bn<IM2COL_K, DENSE_K, false>(f, ws_left, max_counters, 0, stream);
    }
    f.a = wt; f.b = dy; f.lda = f.Kd; f.ldb = K; f.M = C; f.N = pixels; f.transpose_out = 1;
    return dispatch_bn<DENSE_K, IM2COL_K, false>(f, ws_left, max_counters, 0, stream);
  }
  if (pixels >= 1024 || C < 128) {
    a.a = dy; a.b = w; a.lda = K; a.ldb = C; a.M = pixels; a.N = C;
    if (pointwise) return dispatch_bn<DENSE_K, DENSE_MN, false>(a, workspace_bytes, max_counters, 0, stream);
    return dispatch_bn<IM2COL_T_K, WEIGHT_T_MN, false>(a, workspace_bytes, max_counters, 0, stream);
  }
  a.a = w; a.b = dy; a.lda = C; a.ldb = K; a.M = C; a.N = pixels; a.transpose_out = 1;
  if (pointwise) return dispatch_bn<DENSE_MN, DENSE_K, false>(a, workspace_bytes, max_counters, 0, stream);
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
  if (R == 1 && S == 1 && stride == 1 && pad == 0) {      // dW[K,C] = dY[pixels,K]^T * X[pixels,C]
    a.ldb = C;
    return dispatch_bn<DENSE_MN, DENSE_MN, false>(a, workspace_bytes, max_counters, 0, stream);
  }
  return dispatch_bn<DENSE_MN, IM2COL_MN, false>(a, workspace_bytes, max_counters, 0, stream);
}

// Fused "weight-gradient GEMM + all-reduce (average)" (SURVEY W12).  Same operand conventions as
// b200_gemm_tc_ex; `peer` is a host copy of PeerReduce filled with this process's mappings of every
// rank's output / staging / flag objects.  The output must be row-major (no split-K, no batching);
// staging needs world * tiles * 128 * BN * 4 bytes per rank, flags tiles * 9 words, where
// tiles = ceil(M/128) * ceil(N/BN) and BN is returned by b200_gemm_tc_peer_bn(N).  A trailing
// one-block kernel waits for the tiles other owners push into this rank's copy.
B200_API int b200_gemm_tc_peer_bn(int N) { return N <= 32 ? 32 : (N <= 64 ? 64 : 128); }

B200_API int b200_gemm_tc_peer(const void* A, const void* B, const void* bias, int M, int N, int K, int mode, int flags,
                               const long long* desc, const void* peer, unsigned* error, cudaStream_t stream) {
  const bool bf16 = flags & 1;
  const int vec = bf16 ? 8 : 4;
  if (!ptr16(A) || !ptr16(B) || M <= 0 || N <= 0 || K <= 0) return static_cast<int>(cudaErrorInvalidValue);
  Args a{};
  a.peer = *static_cast<const PeerReduce*>(peer);
  if (a.peer.world < 2 || a.peer.world > 8) return static_cast<int>(cudaErrorInvalidValue);
  a.a = A; a.b = B; a.out = a.peer.out[a.peer.rank]; a.bias = static_cast<const float*>(bias);
  a.M = M; a.N = N; a.Kd = K;
  a.lda = desc[0]; a.ldb = desc[1]; a.ldo = desc[2];
  a.batch = 1; a.inner = 1;
  a.out_bf16 = (flags >> 1) & 1; a.bias_bf16 = (flags >> 2) & 1;
  a.error = error;
  if (a.lda % vec || a.ldb % vec) return static_cast<int>(cudaErrorInvalidValue);
  const int bn = b200_gemm_tc_peer_bn(N);
  int rc;
#define B200_GEMM_MODE(AK, BKIND)                                                        \
  rc = bf16 ? dispatch_bn<AK, BKIND, true>(a, 0, 0, bn, stream) : dispatch_bn<AK, BKIND, false>(a, 0, 0, bn, stream)
  if (mode == 0) { B200_GEMM_MODE(DENSE_K, DENSE_K); }
  else if (mode == 1) { B200_GEMM_MODE(DENSE_K, DENSE_MN); }
  else if (mode == 2) { B200_GEMM_MODE(DENSE_MN, DENSE_MN); }
  else return static_cast<int>(cudaErrorInvalidValue);
#undef B200_GEMM_MODE
  if (rc != 0) return rc;
  const int tiles = ceil_div(M, BM) * ceil_div(N, bn);
  peer_reduce_wait_kernel<<<1, 256, 0, stream>>>(a.peer, tiles, error);
  return static_cast<int>(cudaPeekAtLastError());
}

B200_API int b200_peer_reduce_bytes(void) { return (int)sizeof(PeerReduce); }

// GEMM whose epilogue is the NVLS all-reduce (average) of its output: see NvlsReduce.  Same operand conventions as
// b200_gemm_tc_ex (mode 0/1/2, desc = {lda, ldb, ldo}); flags bit 0 = bf16 operands, bit 1 = bf16 output.  `nvls` is a
// host copy of NvlsReduce.  The output must live at the same offset of every rank's multicast-bound arena.
// Flag object per rank: b200_gemm_tc_nvls_flag_words(M, N) 32-bit words, zero-initialised once.
// Returns cudaErrorNotSupported (801) when the shape does not run on the persistent 2-CTA kernel (then call the
// plain GEMM followed by an all-reduce).
B200_API int b200_nvls_reduce_bytes(void) { return (int)sizeof(NvlsReduce); }
B200_API int b200_gemm_tc_nvls_flag_words(int M, int N) {
  const int halves = 2 * ceil_div(M, 2 * BM) * ceil_div(N, 128);        // upper bound over both tile widths
  return halves * 8 + halves * 4;
}
B200_API int b200_gemm_tc_nvls_final_off(int M, int N) { return 2 * ceil_div(M, 2 * BM) * ceil_div(N, 128) * 8; }

B200_API int b200_gemm_tc_nvls(const void* A, const void* B, void* out_local, int M, int N, int K, int mode, int flags,
                               const long long* desc, const void* nvls, unsigned* error, cudaStream_t stream) {
  const bool bf16 = flags & 1;
  const int vec = bf16 ? 8 : 4;
  if (!ptr16(A) || !ptr16(B) || !ptr16(out_local) || M <= 0 || N <= 0 || K <= 0) return static_cast<int>(cudaErrorInvalidValue);
  Args a{};
  a.nvls = *static_cast<const NvlsReduce*>(nvls);
  if (a.nvls.world < 2 || a.nvls.world > 8 || a.nvls.mc_out == nullptr) return static_cast<int>(cudaErrorInvalidValue);
  a.a = A; a.b = B; a.out = static_cast<float*>(out_local);
  a.M = M; a.N = N; a.Kd = K;
  a.lda = desc[0]; a.ldb = desc[1]; a.ldo = desc[2];
  a.batch = 1; a.inner = 1;
  a.out_bf16 = (flags >> 1) & 1;
  a.error = error;
  const int ovec = a.out_bf16 ? 8 : 4;
  if (a.lda % vec || a.ldb % vec || a.ldo % ovec || N % ovec) return static_cast<int>(cudaErrorInvalidValue);
  int rc = -1;
  if (bf16) {
    if (mode == 0) rc = try_2cta<DENSE_K, DENSE_K, true>(a, stream);
    else if (mode == 1) rc = try_2cta<DENSE_K, DENSE_MN, true>(a, stream);
    else if (mode == 2) rc = try_2cta<DENSE_MN, DENSE_MN, true>(a, stream);
  } else if (mode == 0) {
    rc = try_2cta<DENSE_K, DENSE_K, false>(a, stream);
  }
  return rc < 0 ? static_cast<int>(cudaErrorNotSupported) : rc;
}
