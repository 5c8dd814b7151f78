// Implicit-GEMM convolution (fprop / dgrad / wgrad) on NHWC activations and KRSC filters,
// fp32 storage, fp32 accumulate.  Portable CUDA-core (FFMA) path: it is the correctness
// baseline and the fallback for shapes the tcgen05 kernels (gemm_tcgen05.cu) do not take.
//
// Replaces, on the reference's path (/root/reference/dpp.py:50-52 -> torchvision ResNet-18),
// the cuDNN convolution kernels SURVEY.md §2.5(b) lists as K7 (fprop), K8 (dgrad), K9 (wgrad),
// and -- as a 1x1 "convolution" over a [M,1,1,K] tensor -- the cuBLAS calls behind
// nn.Linear (K16).
//
// One GEMM core, three operand gathers:
//   fprop : Y[m, k]   = sum_{r,s,c} X[n, p*st-pad+r, q*st-pad+s, c] * Wt[k, r, s, c]     M=N*P*Q  N=K      Kd=R*S*C
//   dgrad : dX[m, c]  = sum_{r,s,k} dY[n, (h+pad-r)/st, (w+pad-s)/st, k] * Wt[k, r, s, c]  M=N*H*W  N=C      Kd=R*S*K
//   wgrad : dW[k, j]  = sum_{m}     dY[m, k] * im2col(X)[m, j]   (j = (r,s,c))              M=K      N=R*S*C  Kd=N*P*Q
//
// Parallelism for the tiny, deep layers of a 32x32-input ResNet (GEMM-M as small as 32 with
// K up to 4608) comes from split-K: each split writes its partial tile to a scratch workspace,
// takes a ticket on a per-tile counter, and the last split to arrive sums the partials in a
// fixed order (deterministic) and runs the epilogue.  The epilogue can also accumulate the
// per-channel sum / sum-of-squares BatchNorm needs (fused statistics: saves one pass over Y).
#include <stdio.h>

#include "common.cuh"
#include "b200ddp.h"

namespace b200 {

constexpr int BM = 64, BN = 64, BK = 16, NT = 256;
constexpr int TILE_ELEMS = BM * BN;

struct ConvArgs {
  const float* a;      // fprop: X     dgrad: dY    wgrad: dY
  const float* b;      // fprop: Wt    dgrad: Wt    wgrad: X
  float* out;          // fprop: Y     dgrad: dX    wgrad: dW
  const float* bias;   // optional, fprop only: added per output column
  float* stats;        // optional, [2*N]: column sum, column sum of squares (atomicAdd)
  float* workspace;    // split-K partial tiles
  unsigned* counters;  // split-K tickets, one per output tile, zero between launches
  int Nb, H, W, C, K, R, S, P, Q, stride, pad;
  int M, N, Kd;
  int splits, k_per_split;
  int accumulate;      // wgrad: add to `out` instead of overwriting (gradient accumulation)
};

enum { FPROP = 0, DGRAD = 1, WGRAD = 2 };

__device__ __forceinline__ float4 zero4() { return make_float4(0.f, 0.f, 0.f, 0.f); }

// ---- operand gathers -------------------------------------------------------------------
// A element (row m, reduction index k) for each mode; scalar form (used when the vector
// preconditions do not hold, e.g. the 3-channel stem).
template <int MODE>
__device__ __forceinline__ float load_a_scalar(const ConvArgs& g, int m, int k) {
  if (m >= g.M || k >= g.Kd) return 0.f;
  if (MODE == FPROP) {
    const int q = m % g.Q, p = (m / g.Q) % g.P, n = m / (g.Q * g.P);
    const int c = k % g.C, rs = k / g.C, s = rs % g.S, r = rs / g.S;
    const int h = p * g.stride - g.pad + r, w = q * g.stride - g.pad + s;
    if (h < 0 || h >= g.H || w < 0 || w >= g.W) return 0.f;
    return __ldg(g.a + ((size_t)(n * g.H + h) * g.W + w) * g.C + c);
  } else if (MODE == DGRAD) {
    const int w = m % g.W, h = (m / g.W) % g.H, n = m / (g.W * g.H);
    const int co = k % g.K, rs = k / g.K, s = rs % g.S, r = rs / g.S;
    const int hh = h + g.pad - r, ww = w + g.pad - s;
    if (hh < 0 || ww < 0 || hh % g.stride || ww % g.stride) return 0.f;
    const int p = hh / g.stride, q = ww / g.stride;
    if (p >= g.P || q >= g.Q) return 0.f;
    return __ldg(g.a + ((size_t)(n * g.P + p) * g.Q + q) * g.K + co);
  } else {
    return __ldg(g.a + (size_t)k * g.K + m);   // dY[pixel k][channel m]
  }
}

template <int MODE>
__device__ __forceinline__ float load_b_scalar(const ConvArgs& g, int k, int n) {
  if (n >= g.N || k >= g.Kd) return 0.f;
  if (MODE == FPROP) {
    return __ldg(g.b + (size_t)n * g.Kd + k);
  } else if (MODE == DGRAD) {
    const int co = k % g.K, rs = k / g.K;
    return __ldg(g.b + ((size_t)co * g.R * g.S + rs) * g.C + n);
  } else {
    const int q = k % g.Q, p = (k / g.Q) % g.P, img = k / (g.Q * g.P);
    const int c = n % g.C, rs = n / g.C, s = rs % g.S, r = rs / g.S;
    const int h = p * g.stride - g.pad + r, w = q * g.stride - g.pad + s;
    if (h < 0 || h >= g.H || w < 0 || w >= g.W) return 0.f;
    return __ldg(g.b + ((size_t)(img * g.H + h) * g.W + w) * g.C + c);
  }
}

// Vector forms: 4 consecutive elements along the memory-contiguous GEMM axis.
//   fprop A: along k (channels c)        fprop B: along k
//   dgrad A: along k (channels co)       dgrad B: along n (channels ci)
//   wgrad A: along m (channels co)       wgrad B: along n (channels c)
template <int MODE>
__device__ __forceinline__ float4 load_a_vec(const ConvArgs& g, int m, int k) {
  if (MODE == FPROP) {
    if (m >= g.M || k >= g.Kd) return zero4();
    const int q = m % g.Q, p = (m / g.Q) % g.P, n = m / (g.Q * g.P);
    const int c = k % g.C, rs = k / g.C, s = rs % g.S, r = rs / g.S;
    const int h = p * g.stride - g.pad + r, w = q * g.stride - g.pad + s;
    if (h < 0 || h >= g.H || w < 0 || w >= g.W) return zero4();
    return ldg4(g.a + ((size_t)(n * g.H + h) * g.W + w) * g.C + c);
  } else if (MODE == DGRAD) {
    if (m >= g.M || k >= g.Kd) return zero4();
    const int w = m % g.W, h = (m / g.W) % g.H, n = m / (g.W * g.H);
    const int co = k % g.K, rs = k / g.K, s = rs % g.S, r = rs / g.S;
    const int hh = h + g.pad - r, ww = w + g.pad - s;
    if (hh < 0 || ww < 0 || hh % g.stride || ww % g.stride) return zero4();
    const int p = hh / g.stride, q = ww / g.stride;
    if (p >= g.P || q >= g.Q) return zero4();
    return ldg4(g.a + ((size_t)(n * g.P + p) * g.Q + q) * g.K + co);
  } else {
    if (m >= g.M || k >= g.Kd) return zero4();
    return ldg4(g.a + (size_t)k * g.K + m);
  }
}

template <int MODE>
__device__ __forceinline__ float4 load_b_vec(const ConvArgs& g, int k, int n) {
  if (n >= g.N || k >= g.Kd) return zero4();
  if (MODE == FPROP) {
    return ldg4(g.b + (size_t)n * g.Kd + k);
  } else if (MODE == DGRAD) {
    const int co = k % g.K, rs = k / g.K;
    return ldg4(g.b + ((size_t)co * g.R * g.S + rs) * g.C + n);
  } else {
    const int q = k % g.Q, p = (k / g.Q) % g.P, img = k / (g.Q * g.P);
    const int c = n % g.C, rs = n / g.C, s = rs % g.S, r = rs / g.S;
    const int h = p * g.stride - g.pad + r, w = q * g.stride - g.pad + s;
    if (h < 0 || h >= g.H || w < 0 || w >= g.W) return zero4();
    return ldg4(g.b + ((size_t)(img * g.H + h) * g.W + w) * g.C + c);
  }
}

// Is operand A / B of this mode gathered with its vector along the reduction axis (needs a
// transposing store into the [k][row] shared tile) or along the row axis (direct store)?
template <int MODE> struct Layout {
  static constexpr bool a_vec_along_k = (MODE != WGRAD);
  static constexpr bool b_vec_along_k = (MODE == FPROP);
};

template <int MODE, bool VEC_A, bool VEC_B>
__global__ void __launch_bounds__(NT) conv_gemm_kernel(const ConvArgs g) {
  __shared__ __align__(16) float As[2][BK][BM + 4];
  __shared__ __align__(16) float Bs[2][BK][BN + 4];
  __shared__ float col_sum[BN], col_sq[BN];
  __shared__ unsigned ticket_s;

  pdl_launch_dependents();
  pdl_wait();
  const int t = threadIdx.x;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int split = blockIdx.z;
  const int k_begin = split * g.k_per_split;
  const int k_end = min(g.Kd, k_begin + g.k_per_split);

  // static per-thread gather coordinates
  constexpr bool A_K = Layout<MODE>::a_vec_along_k;
  constexpr bool B_K = Layout<MODE>::b_vec_along_k;
  const int a_row = A_K ? (t >> 2) : ((t & 15) * 4);   // row (m) offset inside the tile
  const int a_k = A_K ? ((t & 3) * 4) : (t >> 4);      // k offset inside the tile
  const int b_row = B_K ? (t >> 2) : ((t & 15) * 4);
  const int b_k = B_K ? ((t & 3) * 4) : (t >> 4);

  float4 ra, rb;
  auto gload = [&](int kb) {
    if (VEC_A) {
      ra = load_a_vec<MODE>(g, m0 + a_row, kb + a_k);
    } else if (A_K) {
      ra.x = load_a_scalar<MODE>(g, m0 + a_row, kb + a_k + 0);
      ra.y = load_a_scalar<MODE>(g, m0 + a_row, kb + a_k + 1);
      ra.z = load_a_scalar<MODE>(g, m0 + a_row, kb + a_k + 2);
      ra.w = load_a_scalar<MODE>(g, m0 + a_row, kb + a_k + 3);
    } else {
      ra.x = load_a_scalar<MODE>(g, m0 + a_row + 0, kb + a_k);
      ra.y = load_a_scalar<MODE>(g, m0 + a_row + 1, kb + a_k);
      ra.z = load_a_scalar<MODE>(g, m0 + a_row + 2, kb + a_k);
      ra.w = load_a_scalar<MODE>(g, m0 + a_row + 3, kb + a_k);
    }
    if (VEC_B) {
      rb = load_b_vec<MODE>(g, kb + b_k, n0 + b_row);
    } else if (B_K) {
      rb.x = load_b_scalar<MODE>(g, kb + b_k + 0, n0 + b_row);
      rb.y = load_b_scalar<MODE>(g, kb + b_k + 1, n0 + b_row);
      rb.z = load_b_scalar<MODE>(g, kb + b_k + 2, n0 + b_row);
      rb.w = load_b_scalar<MODE>(g, kb + b_k + 3, n0 + b_row);
    } else {
      rb.x = load_b_scalar<MODE>(g, kb + b_k, n0 + b_row + 0);
      rb.y = load_b_scalar<MODE>(g, kb + b_k, n0 + b_row + 1);
      rb.z = load_b_scalar<MODE>(g, kb + b_k, n0 + b_row + 2);
      rb.w = load_b_scalar<MODE>(g, kb + b_k, n0 + b_row + 3);
    }
  };
  // elements past k_end must not leak into this split's sum
  auto mask_tail = [&](int kb) {
    if (A_K) {
      const int k = kb + a_k;
      if (k + 0 >= k_end) ra.x = 0.f;
      if (k + 1 >= k_end) ra.y = 0.f;
      if (k + 2 >= k_end) ra.z = 0.f;
      if (k + 3 >= k_end) ra.w = 0.f;
    } else if (kb + a_k >= k_end) {
      ra = zero4();
    }
  };
  auto sstore = [&](int buf) {
    if (A_K) {
      As[buf][a_k + 0][a_row] = ra.x;
      As[buf][a_k + 1][a_row] = ra.y;
      As[buf][a_k + 2][a_row] = ra.z;
      As[buf][a_k + 3][a_row] = ra.w;
    } else {
      *reinterpret_cast<float4*>(&As[buf][a_k][a_row]) = ra;
    }
    if (B_K) {
      Bs[buf][b_k + 0][b_row] = rb.x;
      Bs[buf][b_k + 1][b_row] = rb.y;
      Bs[buf][b_k + 2][b_row] = rb.z;
      Bs[buf][b_k + 3][b_row] = rb.w;
    } else {
      *reinterpret_cast<float4*>(&Bs[buf][b_k][b_row]) = rb;
    }
  };

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int ty = t >> 4, tx = t & 15;
  int buf = 0;
  if (k_begin < k_end) {
    gload(k_begin);
    mask_tail(k_begin);
    sstore(0);
  }
  __syncthreads();
  for (int kb = k_begin; kb < k_end; kb += BK) {
    const bool more = kb + BK < k_end;
    if (more) {
      gload(kb + BK);
      mask_tail(kb + BK);
    }
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w};
      const float bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (more) sstore(buf ^ 1);
    __syncthreads();
    buf ^= 1;
  }

  // ---- split-K: park the partial tile, last arriver reduces in split order ---------------
  const int tile_id = blockIdx.y * gridDim.x + blockIdx.x;
  if (g.splits > 1) {
    float* mine = g.workspace + ((size_t)split * gridDim.x * gridDim.y + tile_id) * TILE_ELEMS;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) mine[(i * 4 + j) * NT + t] = acc[i][j];
    __threadfence();
    __syncthreads();
    if (t == 0) ticket_s = atomicAdd(&g.counters[tile_id], 1u);
    __syncthreads();
    if (ticket_s != (unsigned)(g.splits - 1)) return;
    __threadfence();
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int s = 0; s < g.splits; ++s) {
      const float* part = g.workspace + ((size_t)s * gridDim.x * gridDim.y + tile_id) * TILE_ELEMS;
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] += __ldcg(part + (i * 4 + j) * NT + t);
    }
    if (t == 0) g.counters[tile_id] = 0u;   // ready for the next launch
  }

  // ---- epilogue ----------------------------------------------------------------------------
  const int row0 = m0 + ty * 4, col0 = n0 + tx * 4;
  if (g.bias != nullptr) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float bj = (col0 + j < g.N) ? __ldg(g.bias + col0 + j) : 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i][j] += bj;
    }
  }
  const bool vec_out = (g.N % 4 == 0) && (col0 + 3 < g.N) && (reinterpret_cast<uintptr_t>(g.out) & 15u) == 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int row = row0 + i;
    if (row >= g.M) continue;
    float* dst = g.out + (size_t)row * g.N + col0;
    if (vec_out) {
      float4 v = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      if (g.accumulate) {
        const float4 o = *reinterpret_cast<const float4*>(dst);
        v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
      }
      *reinterpret_cast<float4*>(dst) = v;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (col0 + j < g.N) dst[j] = g.accumulate ? dst[j] + acc[i][j] : acc[i][j];
    }
  }
  if (g.stats != nullptr) {
    if (t < BN) { col_sum[t] = 0.f; col_sq[t] = 0.f; }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float s = 0.f, sq = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (row0 + i < g.M) { s += acc[i][j]; sq += acc[i][j] * acc[i][j]; }
      }
      atomicAdd(&col_sum[tx * 4 + j], s);
      atomicAdd(&col_sq[tx * 4 + j], sq);
    }
    __syncthreads();
    if (t < BN && n0 + t < g.N) {
      atomicAdd(g.stats + n0 + t, col_sum[t]);
      atomicAdd(g.stats + g.N + n0 + t, col_sq[t]);
    }
  }
}

template <int MODE>
static int launch_mode(const ConvArgs& g, dim3 grid, bool vec_a, bool vec_b, cudaStream_t st) {
  if (vec_a && vec_b) (void)launch_pdl(conv_gemm_kernel<MODE, true, true>, grid, dim3(NT), 0, st, g);
  else if (vec_a) (void)launch_pdl(conv_gemm_kernel<MODE, true, false>, grid, dim3(NT), 0, st, g);
  else if (vec_b) (void)launch_pdl(conv_gemm_kernel<MODE, false, true>, grid, dim3(NT), 0, st, g);
  else (void)launch_pdl(conv_gemm_kernel<MODE, false, false>, grid, dim3(NT), 0, st, g);
  return static_cast<int>(cudaPeekAtLastError());
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int conv_dispatch(int mode, ConvArgs g, size_t workspace_bytes, int max_counters, cudaStream_t st) {
  if (mode == FPROP) { g.M = g.Nb * g.P * g.Q; g.N = g.K; g.Kd = g.R * g.S * g.C; }
  else if (mode == DGRAD) { g.M = g.Nb * g.H * g.W; g.N = g.C; g.Kd = g.R * g.S * g.K; }
  else { g.M = g.K; g.N = g.R * g.S * g.C; g.Kd = g.Nb * g.P * g.Q; }
  if (g.M <= 0 || g.N <= 0 || g.Kd <= 0) return static_cast<int>(cudaErrorInvalidValue);
  const int tm = ceil_div(g.M, BM), tn = ceil_div(g.N, BN);
  const int tiles = tm * tn;
  // split-K: aim for ~2 CTAs per SM when the output grid alone cannot fill the chip
  int splits = 1;
  const int k_tiles = ceil_div(g.Kd, BK);
  if (tiles < kNumSMs && g.workspace != nullptr && g.counters != nullptr && tiles <= max_counters) {
    splits = min(min(2 * kNumSMs / tiles, k_tiles / 2), 64);
    const size_t per_split = (size_t)tiles * TILE_ELEMS * sizeof(float);
    const size_t fit = workspace_bytes / per_split;
    if ((size_t)max(splits, 1) > fit) splits = (int)fit;
    splits = max(splits, 1);
  }
  int kps = ceil_div(k_tiles, splits) * BK;
  splits = ceil_div(g.Kd, kps);
  g.splits = splits;
  g.k_per_split = kps;
  bool vec_a, vec_b;
  const bool al = aligned16(g.a) && aligned16(g.b);
  if (mode == FPROP) { vec_a = al && g.C % 4 == 0; vec_b = al && g.Kd % 4 == 0; }
  else if (mode == DGRAD) { vec_a = al && g.K % 4 == 0; vec_b = al && g.C % 4 == 0; }
  else { vec_a = al && g.K % 4 == 0; vec_b = al && g.C % 4 == 0; }
  dim3 grid(tm, tn, splits);
  if (mode == FPROP) return launch_mode<FPROP>(g, grid, vec_a, vec_b, st);
  if (mode == DGRAD) return launch_mode<DGRAD>(g, grid, vec_a, vec_b, st);
  return launch_mode<WGRAD>(g, grid, vec_a, vec_b, st);
}

}  // namespace b200

using namespace b200;

B200_API int b200_conv2d_fprop(const float* x, const float* w, const float* bias, float* y, float* stats,
                               int Nb, int H, int W, int C, int K, int R, int S, int stride, int pad,
                               float* workspace, size_t workspace_bytes, unsigned* counters, int max_counters,
                               cudaStream_t stream) {
  ConvArgs g{};
  g.a = x; g.b = w; g.out = y; g.bias = bias; g.stats = stats; g.workspace = workspace; g.counters = counters;
  g.Nb = Nb; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.pad = pad;
  g.P = (H + 2 * pad - R) / stride + 1;
  g.Q = (W + 2 * pad - S) / stride + 1;
  return conv_dispatch(FPROP, g, workspace_bytes, max_counters, stream);
}

B200_API int b200_conv2d_dgrad(const float* dy, const float* w, float* dx,
                               int Nb, int H, int W, int C, int K, int R, int S, int stride, int pad,
                               float* workspace, size_t workspace_bytes, unsigned* counters, int max_counters,
                               cudaStream_t stream) {
  ConvArgs g{};
  g.a = dy; g.b = w; g.out = dx; g.workspace = workspace; g.counters = counters;
  g.Nb = Nb; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.pad = pad;
  g.P = (H + 2 * pad - R) / stride + 1;
  g.Q = (W + 2 * pad - S) / stride + 1;
  return conv_dispatch(DGRAD, g, workspace_bytes, max_counters, stream);
}

B200_API int b200_conv2d_wgrad(const float* dy, const float* x, float* dw, int accumulate,
                               int Nb, int H, int W, int C, int K, int R, int S, int stride, int pad,
                               float* workspace, size_t workspace_bytes, unsigned* counters, int max_counters,
                               cudaStream_t stream) {
  ConvArgs g{};
  g.a = dy; g.b = x; g.out = dw; g.workspace = workspace; g.counters = counters; g.accumulate = accumulate;
  g.Nb = Nb; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.pad = pad;
  g.P = (H + 2 * pad - R) / stride + 1;
  g.Q = (W + 2 * pad - S) / stride + 1;
  return conv_dispatch(WGRAD, g, workspace_bytes, max_counters, stream);
}

// ---------------------------------------------------------------------------------------------
// Forward convolution for callers that must not allocate
// ---------------------------------------------------------------------------------------------
// Some hosts of this library cannot call cudaMalloc (or anything that calls it for them) at the
// point where they need a convolution: code that runs while another stream of the process is
// being captured into a CUDA graph (an allocation there invalidates the capture), plug-ins that
// are handed a device and a stream by an embedding runtime that owns the memory budget, and
// latency-critical loops that have already frozen their allocations.  b200_conv2d_fprop expects
// such a caller to bring its own split-K scratch; this variant needs none, because its scratch
// is a fixed array that is part of the library image.
//
// Layout of the built-in scratch: [tickets: 1024 x u32][partial tiles: 4 MiB].
// Cost: every GPU that loads libb200ddp_kernels.so keeps these 4 MiB + 4 KiB until the process
// exits, whether or not this entry point is ever called.
namespace b200 {

constexpr int kNoAllocMaxTiles = 1024;                         // 64x64 output tiles with a ticket each
constexpr size_t kNoAllocTicketBytes = kNoAllocMaxTiles * sizeof(unsigned);
constexpr size_t kNoAllocPartialBytes = (size_t)4 << 20;       // split-K partial tiles
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocTicketBytes + kNoAllocPartialBytes];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address; the
// address each device maps it to is looked up once per device and cached here.
static unsigned char* g_noalloc_base[kNoAllocMaxDevices] = {};
static thread_local char g_noalloc_error[320] = "";

static int noalloc_fail(cudaError_t code, const char* fmt, long long a = 0, long long b = 0) {
  snprintf(g_noalloc_error, sizeof(g_noalloc_error), fmt, a, b);
  return static_cast<int>(code);
}

}  // namespace b200

B200_API const char* b200_conv2d_fprop_noalloc_error(void) { return g_noalloc_error; }

B200_API size_t b200_conv2d_fprop_noalloc_scratch_bytes(void) { return kNoAllocTicketBytes + kNoAllocPartialBytes; }

B200_API int b200_conv2d_fprop_noalloc(const float* x, const float* w, const float* bias, float* y, float* stats,
                                       int Nb, int H, int W, int C, int K, int R, int S, int stride, int pad,
                                       cudaStream_t stream) {
  g_noalloc_error[0] = '\0';
  if (Nb <= 0 || H <= 0 || W <= 0 || C <= 0 || K <= 0 || R <= 0 || S <= 0 || stride <= 0 || pad < 0)
    return noalloc_fail(cudaErrorInvalidValue, "b200_conv2d_fprop_noalloc: non-positive dimension");
  const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
  if (P <= 0 || Q <= 0)
    return noalloc_fail(cudaErrorInvalidValue, "b200_conv2d_fprop_noalloc: empty output (P=%lld, Q=%lld)", P, Q);
  const long long tiles = ceil_div_ll((long long)Nb * P * Q, BM) * ceil_div_ll(K, BN);
  if (tiles > kNoAllocMaxTiles)
    return noalloc_fail(cudaErrorInvalidValue,
                        "b200_conv2d_fprop_noalloc: output needs %lld tiles of 64x64 but the built-in scratch has "
                        "tickets for %lld; use b200_conv2d_fprop with caller-provided scratch",
                        tiles, (long long)kNoAllocMaxTiles);

  // Once per device, with that device current: resolve the symbol to this device's address.
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return noalloc_fail(err, "b200_conv2d_fprop_noalloc: cudaGetDevice failed (%lld)", err);
  if (dev < 0 || dev >= kNoAllocMaxDevices)
    return noalloc_fail(cudaErrorInvalidDevice, "b200_conv2d_fprop_noalloc: device ordinal %lld not supported", dev);
  unsigned char* base = g_noalloc_base[dev];
  if (base == nullptr) {
    void* p = nullptr;
    err = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (err != cudaSuccess || p == nullptr)
      return noalloc_fail(err != cudaSuccess ? err : cudaErrorInvalidSymbol,
                          "b200_conv2d_fprop_noalloc: cudaGetSymbolAddress failed (%lld) on device %lld", err, dev);
    base = g_noalloc_base[dev] = static_cast<unsigned char*>(p);
  }
  unsigned* tickets = reinterpret_cast<unsigned*>(base);
  float* partials = reinterpret_cast<float*>(base + kNoAllocTicketBytes);

  // The array keeps whatever the previous call left: tickets must start at zero (a launch that
  // was interrupted could have left some behind); partial tiles are fully overwritten before use.
  err = cudaMemsetAsync(tickets, 0, (size_t)tiles * sizeof(unsigned), stream);
  if (err != cudaSuccess) return noalloc_fail(err, "b200_conv2d_fprop_noalloc: clearing tickets failed (%lld)", err);

  ConvArgs g{};
  g.a = x; g.b = w; g.out = y; g.bias = bias; g.stats = stats; g.workspace = partials; g.counters = tickets;
  g.Nb = Nb; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.pad = pad; g.P = P; g.Q = Q;
  const int rc = conv_dispatch(FPROP, g, kNoAllocPartialBytes, kNoAllocMaxTiles, stream);
  if (rc != 0) noalloc_fail(static_cast<cudaError_t>(rc), "b200_conv2d_fprop_noalloc: kernel launch failed (%lld)", rc);
  return rc;
}
