// Transformer (GPT-2) building blocks: LayerNorm, GELU, causal softmax, token+position embedding,
// all with fp32 math on fp32 or bf16 storage.  The matrix products (QKV / projection / MLP / LM head
// and the batched QK^T, PV of attention) run on the tcgen05 GEMM in gemm_tcgen05.cu.
//
// Not on the reference's own path (/root/reference/dpp.py trains a CNN); these cover the
// driver-specified "GPT-2 medium DDP" configuration (BASELINE.json config 4).
#include <cuda_bf16.h>
#include <float.h>

#include "common.cuh"
#include "b200ddp.h"

namespace b200 {

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// ---- LayerNorm: one warp per row ---------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) layernorm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ gamma,
                                                            const T* __restrict__ beta, T* __restrict__ y,
                                                            float* __restrict__ mean, float* __restrict__ rstd, int M,
                                                            int D, float eps) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= M) return;
  const T* xr = x + (size_t)warp * D;
  float s = 0.f;
  for (int j = lane; j < D; j += 32) s += to_f(xr[j]);
  const float mu = warp_sum(s) / D;
  float q = 0.f;
  for (int j = lane; j < D; j += 32) {
    const float d = to_f(xr[j]) - mu;
    q += d * d;
  }
  const float rs = rsqrtf(warp_sum(q) / D + eps);
  T* yr = y + (size_t)warp * D;
  for (int j = lane; j < D; j += 32) yr[j] = from_f<T>((to_f(xr[j]) - mu) * rs * to_f(gamma[j]) + to_f(beta[j]));
  if (lane == 0) {
    mean[warp] = mu;
    rstd[warp] = rs;
  }
}

// dx per row; dgamma/dbeta accumulated into fp32 scratch (zeroed by the caller) with one atomic per
// (block, column): rows of a block are first combined in shared memory.
template <typename T>
__global__ void __launch_bounds__(256) layernorm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                            const T* __restrict__ gamma, const float* __restrict__ mean,
                                                            const float* __restrict__ rstd, T* __restrict__ dx,
                                                            float* __restrict__ dgamma_acc, float* __restrict__ dbeta_acc,
                                                            int M, int D) {
  extern __shared__ float sh[];            // [2*D]
  for (int j = threadIdx.x; j < 2 * D; j += blockDim.x) sh[j] = 0.f;
  __syncthreads();
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (D <= 1024) {
    // a lane always serves columns lane, lane+32, ...: keep its dgamma/dbeta partials in registers over
    // all rows of the warp and touch shared memory once at the end
    float pg[32], pb[32];
#pragma unroll
    for (int q = 0; q < 32; ++q) pg[q] = pb[q] = 0.f;
    for (int row = blockIdx.x * warps + w; row < M; row += gridDim.x * warps) {
      const T* xr = x + (size_t)row * D;
      const T* gr = dy + (size_t)row * D;
      const float mu = mean[row], rs = rstd[row];
      float a = 0.f, b = 0.f;
      for (int j = lane; j < D; j += 32) {
        const float g = to_f(gr[j]) * to_f(gamma[j]);
        a += g;
        b += g * (to_f(xr[j]) - mu) * rs;
      }
      a = warp_sum(a) / D;
      b = warp_sum(b) / D;
      T* dr = dx + (size_t)row * D;
#pragma unroll
      for (int q = 0; q < 32; ++q) {
        const int j = lane + q * 32;
        if (j < D) {
          const float go = to_f(gr[j]);
          const float xh = (to_f(xr[j]) - mu) * rs;
          dr[j] = from_f<T>(rs * (go * to_f(gamma[j]) - a - xh * b));
          pg[q] += go * xh;
          pb[q] += go;
        }
      }
    }
#pragma unroll
    for (int q = 0; q < 32; ++q) {
      const int j = lane + q * 32;
      if (j < D) {
        atomicAdd(&sh[j], pg[q]);
        atomicAdd(&sh[D + j], pb[q]);
      }
    }
  } else
  for (int row = blockIdx.x * warps + w; row < M; row += gridDim.x * warps) {
    const T* xr = x + (size_t)row * D;
    const T* gr = dy + (size_t)row * D;
    const float mu = mean[row], rs = rstd[row];
    float a = 0.f, b = 0.f;               // sum(g*gamma), sum(g*gamma*xhat)
    for (int j = lane; j < D; j += 32) {
      const float g = to_f(gr[j]) * to_f(gamma[j]);
      const float xh = (to_f(xr[j]) - mu) * rs;
      a += g;
      b += g * xh;
    }
    a = warp_sum(a) / D;
    b = warp_sum(b) / D;
    T* dr = dx + (size_t)row * D;
    for (int j = lane; j < D; j += 32) {
      const float go = to_f(gr[j]);
      const float xh = (to_f(xr[j]) - mu) * rs;
      dr[j] = from_f<T>(rs * (go * to_f(gamma[j]) - a - xh * b));
      atomicAdd(&sh[j], go * xh);
      atomicAdd(&sh[D + j], go);
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < D; j += blockDim.x) {
    atomicAdd(dgamma_acc + j, sh[j]);
    atomicAdd(dbeta_acc + j, sh[D + j]);
  }
}

// ---- bf16 LayerNorm backward, one pass, everything fused -------------------------------------------------
// D = NV * 256: lane l of a warp owns columns v*256 + l*8 .. +8 (16-byte loads, a warp reads 512 contiguous bytes).
// Per row: dx = rstd * (dy*gamma - mean(dy*gamma) - xhat * mean(dy*gamma*xhat)) [+ dres, the gradient arriving on
// the residual branch that bypasses the LayerNorm].  dgamma / dbeta partials stay in registers over all rows of a
// warp, are combined per block in shared memory, added to a zero-on-entry fp32 accumulator, and the LAST block
// (ticket) writes the bf16 results and re-zeroes accumulator and ticket -- no memset / cast launches around it.
template <int NV>
__global__ void __launch_bounds__(256, 1)
layernorm_bwd_bf16_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                          const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean,
                          const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dres,
                          __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dgamma, __nv_bfloat16* __restrict__ dbeta,
                          float* __restrict__ acc, unsigned* __restrict__ ticket, int M) {
  constexpr int D = NV * 256;
  __shared__ float sh[2 * D];
  __shared__ float gs[D];
  __shared__ unsigned last_s;
  for (int j = threadIdx.x; j < 2 * D; j += 256) sh[j] = 0.f;
  // gamma is kept "lane-major" (column v*256 + l*8 + e at gs[(v*8 + e)*32 + l]) so that the 32 lanes of a warp,
  // which read the same (v, e) of their own 8-column groups, hit 32 different banks
  for (int j = threadIdx.x; j < D; j += 256) gs[((j >> 8) * 8 + (j & 7)) * 32 + ((j >> 3) & 31)] = __bfloat162float(gamma[j]);
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float pg[NV * 8], pb[NV * 8];
#pragma unroll
  for (int i = 0; i < NV * 8; ++i) pg[i] = pb[i] = 0.f;
  const float inv_d = 1.f / (float)D;
  // two rows per iteration: all loads of both rows are issued before anything is consumed, which doubles the bytes in
  // flight per warp (the kernel is latency bound otherwise: 8 warps per SM, one 4 KB row each)
  const int stride = gridDim.x * 8;
  for (int row0 = blockIdx.x * 8 + w; row0 < M; row0 += 2 * stride) {
    uint4 xv[2][NV], dv[2][NV], rv[2][NV];
    float mu[2], rs[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int row = row0 + u * stride;
      if (row < M) {
        const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * D);
        const uint4* gr = reinterpret_cast<const uint4*>(dy + (size_t)row * D);
#pragma unroll
        for (int v = 0; v < NV; ++v) {
          xv[u][v] = xr[v * 32 + lane];
          dv[u][v] = gr[v * 32 + lane];
          rv[u][v] = dres ? reinterpret_cast<const uint4*>(dres + (size_t)row * D)[v * 32 + lane] : make_uint4(0, 0, 0, 0);
        }
        mu[u] = mean[row];
        rs[u] = rstd[row];
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int row = row0 + u * stride;
      if (row >= M) break;
      float a = 0.f, b = 0.f;
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        const __nv_bfloat162* xh2 = reinterpret_cast<const __nv_bfloat162*>(&xv[u][v]);
        const __nv_bfloat162* dh2 = reinterpret_cast<const __nv_bfloat162*>(&dv[u][v]);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 xf = __bfloat1622float2(xh2[e]), df = __bfloat1622float2(dh2[e]);
          const float g0 = df.x * gs[(v * 8 + 2 * e) * 32 + lane], g1 = df.y * gs[(v * 8 + 2 * e + 1) * 32 + lane];
          a += g0 + g1;
          b += g0 * ((xf.x - mu[u]) * rs[u]) + g1 * ((xf.y - mu[u]) * rs[u]);
        }
      }
      a = warp_sum(a) * inv_d;
      b = warp_sum(b) * inv_d;
      uint4* dr = reinterpret_cast<uint4*>(dx + (size_t)row * D);
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        const __nv_bfloat162* xh2 = reinterpret_cast<const __nv_bfloat162*>(&xv[u][v]);
        const __nv_bfloat162* dh2 = reinterpret_cast<const __nv_bfloat162*>(&dv[u][v]);
        const __nv_bfloat162* rh2 = reinterpret_cast<const __nv_bfloat162*>(&rv[u][v]);
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 xf = __bfloat1622float2(xh2[e]), df = __bfloat1622float2(dh2[e]), rf = __bfloat1622float2(rh2[e]);
          const float h0 = (xf.x - mu[u]) * rs[u], h1 = (xf.y - mu[u]) * rs[u];
          const float o0 = rs[u] * (df.x * gs[(v * 8 + 2 * e) * 32 + lane] - a - h0 * b) + rf.x;
          const float o1 = rs[u] * (df.y * gs[(v * 8 + 2 * e + 1) * 32 + lane] - a - h1 * b) + rf.y;
          pg[v * 8 + 2 * e] += df.x * h0;
          pg[v * 8 + 2 * e + 1] += df.y * h1;
          pb[v * 8 + 2 * e] += df.x;
          pb[v * 8 + 2 * e + 1] += df.y;
          const __nv_bfloat162 ob = __floats2bfloat162_rn(o0, o1);
          o[e] = *reinterpret_cast<const uint32_t*>(&ob);
        }
        dr[v * 32 + lane] = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
  }
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      atomicAdd(&sh[(v * 8 + e) * 32 + lane], pg[v * 8 + e]);          // lane-major like gs: conflict-free
      atomicAdd(&sh[D + (v * 8 + e) * 32 + lane], pb[v * 8 + e]);
    }
  __syncthreads();
  for (int j = threadIdx.x; j < D; j += 256) {               // back to column order for the global accumulator
    const int t2 = ((j >> 8) * 8 + (j & 7)) * 32 + ((j >> 3) & 31);
    atomicAdd(acc + j, sh[t2]);
    atomicAdd(acc + D + j, sh[D + t2]);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last_s = atomicAdd(ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
  __syncthreads();
  if (last_s) {                                            // every block's additions are visible: finish up
    __threadfence();
    float vg[NV], vb[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) {                           // all loads first: one L2 round trip, not NV
      vg[v] = __ldcg(acc + v * 256 + threadIdx.x);
      vb[v] = __ldcg(acc + D + v * 256 + threadIdx.x);
    }
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const int j = v * 256 + threadIdx.x;
      dgamma[j] = __float2bfloat16_rn(vg[v]);
      dbeta[j] = __float2bfloat16_rn(vb[v]);
      acc[j] = 0.f;
      acc[D + j] = 0.f;
    }
    if (threadIdx.x == 0) *ticket = 0u;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) scale_any_kernel(T* __restrict__ x, size_t n, const float* __restrict__ scale_dev) {
  const float sc = __ldg(scale_dev);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    x[i] = from_f<T>(to_f(x[i]) * sc);
}

template <typename T>
__global__ void __launch_bounds__(256) cast_from_f32_kernel(const float* __restrict__ in, T* __restrict__ out, size_t n,
                                                            int accumulate) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = from_f<T>(accumulate ? to_f(out[i]) + in[i] : in[i]);
}

// ---- GELU (tanh approximation, as GPT-2) -------------------------------------------------------------
__device__ __forceinline__ float gelu_f(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float gelu_grad_f(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  const float t = tanhf(u);
  const float du = 0.7978845608028654f * (1.f + 3.f * 0.044715f * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}
template <typename T>
__global__ void __launch_bounds__(256) gelu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    y[i] = from_f<T>(gelu_f(to_f(x[i])));
}
template <typename T>
__global__ void __launch_bounds__(256) gelu_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy,
                                                       T* __restrict__ dx, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dx[i] = from_f<T>(to_f(dy[i]) * gelu_grad_f(to_f(x[i])));
}

// ---- causal softmax over the rows of [Z, T, T] score matrices, in place -------------------------------
// forward : P[i, j] = softmax_j( scale * S[i, j] ) for j <= i, 0 for j > i
// backward: dS[i, j] = scale * P[i, j] * (dP[i, j] - sum_k P[i, k] dP[i, k]),   written over dP
template <typename T>
__global__ void __launch_bounds__(256) causal_softmax_fwd_kernel(T* __restrict__ s, int T_len, size_t rows, float scale) {
  const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const int i = (int)(warp % T_len);
  T* r = s + warp * T_len;
  float mx = -FLT_MAX;
  for (int j = lane; j <= i; j += 32) mx = fmaxf(mx, to_f(r[j]) * scale);
  mx = warp_max(mx);
  float se = 0.f;
  for (int j = lane; j <= i; j += 32) se += __expf(to_f(r[j]) * scale - mx);
  const float inv = 1.f / warp_sum(se);
  // columns past the diagonal 128-tile are never read (the P.V / dS.K GEMMs stop at the tile edge)
  const int j_end = min(T_len, (i / 128 + 1) * 128);
  for (int j = lane; j < j_end; j += 32) r[j] = from_f<T>(j <= i ? __expf(to_f(r[j]) * scale - mx) * inv : 0.f);
}
template <typename T>
__global__ void __launch_bounds__(256) causal_softmax_bwd_kernel(const T* __restrict__ p, T* __restrict__ dp, int T_len,
                                                                 size_t rows, float scale) {
  const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const int i = (int)(warp % T_len);
  const T* pr = p + warp * T_len;
  T* dr = dp + warp * T_len;
  float dot = 0.f;
  for (int j = lane; j <= i; j += 32) dot += to_f(pr[j]) * to_f(dr[j]);
  dot = warp_sum(dot);
  const int j_end = min(T_len, (i / 128 + 1) * 128);
  for (int j = lane; j < j_end; j += 32) dr[j] = from_f<T>(j <= i ? scale * to_f(pr[j]) * (to_f(dr[j]) - dot) : 0.f);
}

// ---- embedding: out[b, t, :] = wte[idx[b, t], :] + wpe[t, :] ---------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) embedding_fwd_kernel(const long long* __restrict__ idx, const T* __restrict__ wte,
                                                            const T* __restrict__ wpe, T* __restrict__ out, size_t rows,
                                                            int T_len, int D) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows * D; i += (size_t)gridDim.x * blockDim.x) {
    const size_t row = i / D;
    const int d = (int)(i % D);
    out[i] = from_f<T>(to_f(wte[(size_t)idx[row] * D + d]) + to_f(wpe[(size_t)(row % T_len) * D + d]));
  }
}
// gradients accumulate in fp32 scratch tables (token rows collide), converted afterwards
template <typename T>
__global__ void __launch_bounds__(256) embedding_bwd_kernel(const long long* __restrict__ idx, const T* __restrict__ dout,
                                                            float* __restrict__ dwte_acc, float* __restrict__ dwpe_acc,
                                                            size_t rows, int T_len, int D) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows * D; i += (size_t)gridDim.x * blockDim.x) {
    const size_t row = i / D;
    const int d = (int)(i % D);
    const float g = to_f(dout[i]);
    atomicAdd(dwte_acc + (size_t)idx[row] * D + d, g);
    atomicAdd(dwpe_acc + (size_t)(row % T_len) * D + d, g);
  }
}

// ---- softmax cross-entropy on bf16 / fp32 logits with a large vocabulary ------------------------------------
// One block per row; writes dlogits in the logits' dtype; row losses are averaged by mean_rows.
template <typename T>
__global__ void __launch_bounds__(256) xent_rows_kernel(const T* __restrict__ logits, const long long* __restrict__ target,
                                                        T* __restrict__ dlogits, float* __restrict__ row_loss, int B,
                                                        int V, int ld) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  const T* z = logits + (size_t)row * ld;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  float mx = -FLT_MAX;
  for (int j = threadIdx.x; j < V; j += blockDim.x) mx = fmaxf(mx, to_f(z[j]));
  mx = warp_max(mx);
  if (lane == 0) red[w] = mx;
  __syncthreads();
  mx = warp_max(lane < nw ? red[lane] : -FLT_MAX);
  __syncthreads();
  float se = 0.f;
  for (int j = threadIdx.x; j < V; j += blockDim.x) se += __expf(to_f(z[j]) - mx);
  se = warp_sum(se);
  if (lane == 0) red[w] = se;
  __syncthreads();
  se = warp_sum(lane < nw ? red[lane] : 0.f);
  const int t = (int)target[row];
  const float zt = to_f(z[t]);             // read before the write pass: dlogits may alias logits
  __syncthreads();
  const float inv = 1.f / se, inv_b = 1.f / (float)B;
  T* d = dlogits + (size_t)row * ld;
  for (int j = threadIdx.x; j < V; j += blockDim.x)
    d[j] = from_f<T>((__expf(to_f(z[j]) - mx) * inv - (j == t ? 1.f : 0.f)) * inv_b);
  for (int j = V + threadIdx.x; j < ld; j += blockDim.x) d[j] = from_f<T>(0.f);   // keep the padding columns zero
  if (threadIdx.x == 0) row_loss[row] = mx + __logf(se) - zt;
}

// bf16 rows up to ~100 K columns: the row is read from global memory ONCE with 16-byte loads and kept in
// shared memory for the max / sum / write passes (so dlogits may alias logits).  ld % 8 == 0.
__global__ void __launch_bounds__(512) xent_rows_smem_kernel(const __nv_bfloat16* __restrict__ logits,
                                                             const long long* __restrict__ target,
                                                             __nv_bfloat16* __restrict__ dlogits,
                                                             float* __restrict__ row_loss, int B, int V, int ld) {
  extern __shared__ uint4 rowbuf[];
  __shared__ float red[32];
  const int row = blockIdx.x, nvec = ld >> 3;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const uint4* z = reinterpret_cast<const uint4*>(logits + (size_t)row * ld);
  float mx = -FLT_MAX;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 v = z[i];
    rowbuf[i] = v;
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(h[e]);
      if (i * 8 + 2 * e < V) mx = fmaxf(mx, f.x);
      if (i * 8 + 2 * e + 1 < V) mx = fmaxf(mx, f.y);
    }
  }
  mx = warp_max(mx);
  if (lane == 0) red[w] = mx;
  __syncthreads();
  mx = warp_max(lane < nw ? red[lane] : -FLT_MAX);
  __syncthreads();
  float se = 0.f;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 v = rowbuf[i];
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(h[e]);
      if (i * 8 + 2 * e < V) se += __expf(f.x - mx);
      if (i * 8 + 2 * e + 1 < V) se += __expf(f.y - mx);
    }
  }
  se = warp_sum(se);
  if (lane == 0) red[w] = se;
  __syncthreads();
  se = warp_sum(lane < nw ? red[lane] : 0.f);
  const int t = (int)target[row];
  const float zt = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(rowbuf)[t]);
  const float inv = 1.f / se, inv_b = 1.f / (float)B;
  uint4* d = reinterpret_cast<uint4*>(dlogits + (size_t)row * ld);
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 v = rowbuf[i];
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(h[e]);
      const int c = i * 8 + 2 * e;
      const float g0 = c < V ? (__expf(f.x - mx) * inv - (c == t ? 1.f : 0.f)) * inv_b : 0.f;
      const float g1 = c + 1 < V ? (__expf(f.y - mx) * inv - (c + 1 == t ? 1.f : 0.f)) * inv_b : 0.f;
      const __nv_bfloat162 r = __floats2bfloat162_rn(g0, g1);
      o[e] = *reinterpret_cast<const uint32_t*>(&r);
    }
    d[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
  if (threadIdx.x == 0) row_loss[row] = mx + __logf(se) - zt;
}

__global__ void __launch_bounds__(1024) mean_rows2_kernel(const float* __restrict__ row_loss, float* __restrict__ loss,
                                                          int B) {
  __shared__ float sh[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < B; i += blockDim.x) s += row_loss[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = sh[threadIdx.x];
    v = warp_sum(v);
    if (threadIdx.x == 0) *loss = v / (float)B;
  }
}

static int grid1d(size_t n, int per_block = 256) {
  size_t g = (n + per_block - 1) / per_block;
  return (int)(g < 1 ? 1 : (g > 16 * kNumSMs ? 16 * kNumSMs : g));
}

}  // namespace b200

using namespace b200;
using bf16 = __nv_bfloat16;

#define B200_DISPATCH(bf, CALL_F32, CALL_BF16) \
  do {                                          \
    if (bf) { CALL_BF16; } else { CALL_F32; }   \
  } while (0)

B200_API int b200_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                                int M, int D, float eps, int is_bf16, cudaStream_t st) {
  const int grid = ceil_div(M, 8);
  B200_DISPATCH(is_bf16,
                (layernorm_fwd_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (const float*)gamma, (const float*)beta, (float*)y, mean, rstd, M, D, eps)),
                (layernorm_fwd_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)x, (const bf16*)gamma, (const bf16*)beta, (bf16*)y, mean, rstd, M, D, eps)));
  B200_LAUNCH_CHECK();
}

// dgamma_acc / dbeta_acc: fp32 [D] scratch, zero on entry.
B200_API int b200_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                                void* dx, float* dgamma_acc, float* dbeta_acc, int M, int D, int is_bf16,
                                cudaStream_t st) {
  const int grid = min(ceil_div(M, 8), 2 * kNumSMs);
  const size_t sm = 2 * D * sizeof(float);
  B200_DISPATCH(is_bf16,
                (layernorm_bwd_kernel<float><<<grid, 256, sm, st>>>((const float*)dy, (const float*)x, (const float*)gamma, mean, rstd, (float*)dx, dgamma_acc, dbeta_acc, M, D)),
                (layernorm_bwd_kernel<bf16><<<grid, 256, sm, st>>>((const bf16*)dy, (const bf16*)x, (const bf16*)gamma, mean, rstd, (bf16*)dx, dgamma_acc, dbeta_acc, M, D)));
  B200_LAUNCH_CHECK();
}

B200_API int b200_cast_from_f32(const float* in, void* out, size_t n, int accumulate, int is_bf16, cudaStream_t st) {
  B200_DISPATCH(is_bf16, (cast_from_f32_kernel<float><<<grid1d(n), 256, 0, st>>>(in, (float*)out, n, accumulate)),
                (cast_from_f32_kernel<bf16><<<grid1d(n), 256, 0, st>>>(in, (bf16*)out, n, accumulate)));
  B200_LAUNCH_CHECK();
}

// bf16, D in {768, 1024, 1280}: dx (+ dres), dgamma, dbeta in ONE launch.  acc: fp32 [2*D] and ticket: one word,
// both zero on entry and left zero on exit (persistent scratch owned by the caller, one per stream).
// Returns cudaErrorNotSupported (801) for other widths (use b200_layernorm_bwd).
B200_API int b200_layernorm_bwd_fused(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                                      const void* dres, void* dx, void* dgamma, void* dbeta, float* acc, unsigned* ticket,
                                      int M, int D, cudaStream_t st) {
  const int grid = min(ceil_div(M, 8), kNumSMs);          // ~170 registers: one 8-warp block per SM
#define B200_LN_BWD(NV)                                                                                               \
  layernorm_bwd_bf16_kernel<NV><<<grid, 256, 0, st>>>((const bf16*)dy, (const bf16*)x, (const bf16*)gamma, mean, rstd, \
                                                      (const bf16*)dres, (bf16*)dx, (bf16*)dgamma, (bf16*)dbeta, acc, ticket, M)
  if (D == 768) B200_LN_BWD(3);
  else if (D == 1024) B200_LN_BWD(4);
  else if (D == 1280) B200_LN_BWD(5);
  else return static_cast<int>(cudaErrorNotSupported);
#undef B200_LN_BWD
  B200_LAUNCH_CHECK();
}

B200_API int b200_gelu_fwd(const void* x, void* y, size_t n, int is_bf16, cudaStream_t st) {
  B200_DISPATCH(is_bf16, (gelu_fwd_kernel<float><<<grid1d(n), 256, 0, st>>>((const float*)x, (float*)y, n)),
                (gelu_fwd_kernel<bf16><<<grid1d(n), 256, 0, st>>>((const bf16*)x, (bf16*)y, n)));
  B200_LAUNCH_CHECK();
}

B200_API int b200_gelu_bwd(const void* x, const void* dy, void* dx, size_t n, int is_bf16, cudaStream_t st) {
  B200_DISPATCH(is_bf16, (gelu_bwd_kernel<float><<<grid1d(n), 256, 0, st>>>((const float*)x, (const float*)dy, (float*)dx, n)),
                (gelu_bwd_kernel<bf16><<<grid1d(n), 256, 0, st>>>((const bf16*)x, (const bf16*)dy, (bf16*)dx, n)));
  B200_LAUNCH_CHECK();
}

B200_API int b200_causal_softmax_fwd(void* s, int T_len, size_t rows, float scale, int is_bf16, cudaStream_t st) {
  const int grid = (int)((rows + 7) / 8);
  B200_DISPATCH(is_bf16, (causal_softmax_fwd_kernel<float><<<grid, 256, 0, st>>>((float*)s, T_len, rows, scale)),
                (causal_softmax_fwd_kernel<bf16><<<grid, 256, 0, st>>>((bf16*)s, T_len, rows, scale)));
  B200_LAUNCH_CHECK();
}

B200_API int b200_causal_softmax_bwd(const void* p, void* dp, int T_len, size_t rows, float scale, int is_bf16,
                                     cudaStream_t st) {
  const int grid = (int)((rows + 7) / 8);
  B200_DISPATCH(is_bf16, (causal_softmax_bwd_kernel<float><<<grid, 256, 0, st>>>((const float*)p, (float*)dp, T_len, rows, scale)),
                (causal_softmax_bwd_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)p, (bf16*)dp, T_len, rows, scale)));
  B200_LAUNCH_CHECK();
}

B200_API int b200_embedding_fwd(const long long* idx, const void* wte, const void* wpe, void* out, size_t rows,
                                int T_len, int D, int is_bf16, cudaStream_t st) {
  B200_DISPATCH(is_bf16,
                (embedding_fwd_kernel<float><<<grid1d(rows * D), 256, 0, st>>>(idx, (const float*)wte, (const float*)wpe, (float*)out, rows, T_len, D)),
                (embedding_fwd_kernel<bf16><<<grid1d(rows * D), 256, 0, st>>>(idx, (const bf16*)wte, (const bf16*)wpe, (bf16*)out, rows, T_len, D)));
  B200_LAUNCH_CHECK();
}

B200_API int b200_embedding_bwd(const long long* idx, const void* dout, float* dwte_acc, float* dwpe_acc, size_t rows,
                                int T_len, int D, int is_bf16, cudaStream_t st) {
  B200_DISPATCH(is_bf16,
                (embedding_bwd_kernel<float><<<grid1d(rows * D), 256, 0, st>>>(idx, (const float*)dout, dwte_acc, dwpe_acc, rows, T_len, D)),
                (embedding_bwd_kernel<bf16><<<grid1d(rows * D), 256, 0, st>>>(idx, (const bf16*)dout, dwte_acc, dwpe_acc, rows, T_len, D)));
  B200_LAUNCH_CHECK();
}

// `ld` = row pitch of logits / dlogits in elements (>= V; lets the LM head pad the vocabulary).
B200_API int b200_xent_rows(const void* logits, const long long* target, void* dlogits, float* row_loss, float* loss,
                            int B, int V, int ld, int is_bf16, cudaStream_t st) {
  const size_t row_bytes = (size_t)ld * 2;
  if (is_bf16 && ld % 8 == 0 && row_bytes <= 200 * 1024 && ((uintptr_t)logits % 16 == 0) && ((uintptr_t)dlogits % 16 == 0)) {
    static bool configured = false;
    if (!configured) {
      cudaError_t e = cudaFuncSetAttribute(xent_rows_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) return static_cast<int>(e);
      configured = true;
    }
    xent_rows_smem_kernel<<<B, 512, row_bytes, st>>>((const bf16*)logits, target, (bf16*)dlogits, row_loss, B, V, ld);
    mean_rows2_kernel<<<1, 1024, 0, st>>>(row_loss, loss, B);
    B200_LAUNCH_CHECK();
  }
  B200_DISPATCH(is_bf16,
                (xent_rows_kernel<float><<<B, 256, 0, st>>>((const float*)logits, target, (float*)dlogits, row_loss, B, V, ld)),
                (xent_rows_kernel<bf16><<<B, 256, 0, st>>>((const bf16*)logits, target, (bf16*)dlogits, row_loss, B, V, ld)));
  mean_rows2_kernel<<<1, 1024, 0, st>>>(row_loss, loss, B);
  B200_LAUNCH_CHECK();
}

// x *= *scale_dev (in place); scale_dev is a device fp32 scalar (e.g. the incoming loss gradient)
B200_API int b200_scale_inplace(void* x, size_t n, const float* scale_dev, int is_bf16, cudaStream_t st) {
  B200_DISPATCH(is_bf16, (scale_any_kernel<float><<<grid1d(n), 256, 0, st>>>((float*)x, n, scale_dev)),
                (scale_any_kernel<bf16><<<grid1d(n), 256, 0, st>>>((bf16*)x, n, scale_dev)));
  B200_LAUNCH_CHECK();
}
