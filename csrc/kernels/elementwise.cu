// Small elementwise kernels: input pipeline, scaling, accumulation.
#include <cuda_bf16.h>
#include <stdlib.h>

#include "common.cuh"
#include "b200ddp.h"

namespace b200 {

static int g_pdl = -1;       // -1: read B200DDP_PDL on first use (default on)
bool pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("B200DDP_PDL");
    g_pdl = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return g_pdl != 0;
}

// uint8 NHWC -> fp32 NHWC with torchvision's ToTensor + Normalize(mean, std) arithmetic
// (/root/reference/dpp.py:32): (x / 255 - mean) / std, same operation order, so values match
// the reference's host-side transform bit for bit.
__global__ void __launch_bounds__(256) preprocess_u8_kernel(const uint8_t* __restrict__ in, float* __restrict__ out,
                                                            size_t n, float mean, float stdv) {
  pdl_launch_dependents();
  pdl_wait();
  const size_t n4 = n / 4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const uchar4 u = reinterpret_cast<const uchar4*>(in)[i];
    float4 o;
    o.x = __fdiv_rn(__fdiv_rn((float)u.x, 255.f) - mean, stdv);
    o.y = __fdiv_rn(__fdiv_rn((float)u.y, 255.f) - mean, stdv);
    o.z = __fdiv_rn(__fdiv_rn((float)u.z, 255.f) - mean, stdv);
    o.w = __fdiv_rn(__fdiv_rn((float)u.w, 255.f) - mean, stdv);
    reinterpret_cast<float4*>(out)[i] = o;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const size_t i = n4 * 4 + threadIdx.x;
    out[i] = __fdiv_rn(__fdiv_rn((float)in[i], 255.f) - mean, stdv);
  }
}

// out = a * (*scale_dev or scale_host)
__global__ void __launch_bounds__(256) scale_kernel(const float* __restrict__ a, float* __restrict__ out, size_t n,
                                                    const float* __restrict__ scale_dev, float scale_host) {
  pdl_launch_dependents();
  pdl_wait();
  const float s = scale_dev ? __ldg(scale_dev) * scale_host : scale_host;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = a[i] * s;
}

// out = a + b
__global__ void __launch_bounds__(256) add_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                  float* __restrict__ out, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = a[i] + b[i];
}

// column sums of a [M, N] matrix (bias gradients): out[n] (+)= sum_m a[m, n]
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ a, float* __restrict__ out, int M, int N,
                                                     int accumulate) {
  const int n = blockIdx.x * 32 + (threadIdx.x & 31);
  const int r0 = threadIdx.x >> 5;   // 8 row lanes
  __shared__ float sh[8][33];
  float s = 0.f;
  if (n < N)
    for (int m = r0; m < M; m += 8) s += __ldg(a + (size_t)m * N + n);
  sh[r0][threadIdx.x & 31] = s;
  __syncthreads();
  if (r0 == 0 && n < N) {
    float tot = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) tot += sh[r][threadIdx.x & 31];
    out[n] = accumulate ? out[n] + tot : tot;
  }
}

// column sums of a tall bf16 matrix into fp32 accumulators (pre-zeroed): rows are split over
// blockIdx.y so thousands of rows are summed by many CTAs; one atomic per (CTA, column)
__global__ void __launch_bounds__(256) colsum_bf16_acc_kernel(const __nv_bfloat16* __restrict__ a, float* __restrict__ acc,
                                                              int M, int N, int rows_per_block) {
  const int n = blockIdx.x * 64 + (threadIdx.x & 31) * 2;          // two adjacent columns per lane (bf16x2 loads)
  const int r0 = blockIdx.y * rows_per_block + (threadIdx.x >> 5);
  const int r1 = min(M, (blockIdx.y + 1) * rows_per_block);
  __shared__ float sh[8][64];
  float s0 = 0.f, s1 = 0.f;
  if (n + 1 < N) {
    for (int m = r0; m < r1; m += 8) {
      const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(a + (size_t)m * N + n);
      s0 += __bfloat162float(v.x);
      s1 += __bfloat162float(v.y);
    }
  } else if (n < N) {
    for (int m = r0; m < r1; m += 8) s0 += __bfloat162float(a[(size_t)m * N + n]);
  }
  sh[threadIdx.x >> 5][(threadIdx.x & 31) * 2] = s0;
  sh[threadIdx.x >> 5][(threadIdx.x & 31) * 2 + 1] = s1;
  __syncthreads();
  if (threadIdx.x < 64 && blockIdx.x * 64 + threadIdx.x < N) {
    float tot = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) tot += sh[r][threadIdx.x];
    atomicAdd(acc + blockIdx.x * 64 + threadIdx.x, tot);
  }
}

// bias gradient of a bf16 Linear in one launch: column sums of a tall [M, N] bf16 matrix -> bf16 [N].
// Rows are split over blockIdx.y; partial sums go to a zero-on-entry fp32 accumulator; the last block (ticket)
// converts, writes the result and re-zeroes accumulator and ticket.
__global__ void __launch_bounds__(256) colsum_bf16_fused_kernel(const __nv_bfloat16* __restrict__ a, __nv_bfloat16* __restrict__ out,
                                                                float* __restrict__ acc, unsigned* __restrict__ ticket, int M,
                                                                int N, int rows_per_block) {
  // block = 256 columns x rows_per_block rows; a lane reads 8 adjacent columns (16 bytes) per row, a warp takes every
  // 8th row, four rows in flight per lane
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int n = blockIdx.x * 256 + lane * 8;
  const int r0 = blockIdx.y * rows_per_block + w;
  const int r1 = min(M, (blockIdx.y + 1) * rows_per_block);
  __shared__ float sh[8][256];
  __shared__ unsigned last_s;
  float s[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) s[e] = 0.f;
  if (n + 7 < N) {
    int m = r0;
    for (; m + 24 < r1; m += 32) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = *reinterpret_cast<const uint4*>(a + (size_t)(m + 8 * u) * N + n);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[u]);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = __bfloat1622float2(h[e]);
          s[2 * e] += f.x;
          s[2 * e + 1] += f.y;
        }
      }
    }
    for (; m < r1; m += 8) {
      const uint4 v = *reinterpret_cast<const uint4*>(a + (size_t)m * N + n);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(h[e]);
        s[2 * e] += f.x;
        s[2 * e + 1] += f.y;
      }
    }
  } else {
    for (int m = r0; m < r1; m += 8)
#pragma unroll
      for (int e = 0; e < 8; ++e)
        if (n + e < N) s[e] += __bfloat162float(a[(size_t)m * N + n + e]);
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) sh[w][lane * 8 + e] = s[e];
  __syncthreads();
  {
    const int c = threadIdx.x;
    if (blockIdx.x * 256 + c < N) {
      float tot = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r) tot += sh[r][c];
      atomicAdd(acc + blockIdx.x * 256 + c, tot);
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last_s = atomicAdd(ticket, 1u) == gridDim.x * gridDim.y - 1 ? 1u : 0u;
  __syncthreads();
  if (last_s) {
    __threadfence();
    for (int j0 = 0; j0 < N; j0 += 256 * 8) {              // eight independent L2 loads in flight per thread
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int j = j0 + u * 256 + threadIdx.x;
        v[u] = j < N ? __ldcg(acc + j) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int j = j0 + u * 256 + threadIdx.x;
        if (j < N) {
          out[j] = __float2bfloat16_rn(v[u]);
          acc[j] = 0.f;
        }
      }
    }
    if (threadIdx.x == 0) *ticket = 0u;
  }
}

// bf16 variant of colsum (bias gradients of bf16 Linear layers)
__global__ void __launch_bounds__(256) colsum_bf16_kernel(const __nv_bfloat16* __restrict__ a, __nv_bfloat16* __restrict__ out,
                                                          int M, int N, int accumulate) {
  const int n = blockIdx.x * 32 + (threadIdx.x & 31);
  const int r0 = threadIdx.x >> 5;
  __shared__ float sh[8][33];
  float s = 0.f;
  if (n < N)
    for (int m = r0; m < M; m += 8) s += __bfloat162float(a[(size_t)m * N + n]);
  sh[r0][threadIdx.x & 31] = s;
  __syncthreads();
  if (r0 == 0 && n < N) {
    float tot = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) tot += sh[r][threadIdx.x & 31];
    out[n] = __float2bfloat16_rn(accumulate ? __bfloat162float(out[n]) + tot : tot);
  }
}

static int grid1d(size_t n, int per_thread = 4) {
  size_t g = (n + 256 * per_thread - 1) / (256 * per_thread);
  return (int)(g < 1 ? 1 : (g > 8 * kNumSMs ? 8 * kNumSMs : g));
}

}  // namespace b200

using namespace b200;

// Programmatic dependent launch on/off for every kernel of the library (default on; see common.cuh).
B200_API void b200_set_pdl(int enabled) { b200::g_pdl = enabled ? 1 : 0; }

B200_API int b200_preprocess_u8(const uint8_t* in, float* out, size_t n, float mean, float stdv, cudaStream_t stream) {
  return static_cast<int>(launch_pdl(preprocess_u8_kernel, dim3(grid1d(n / 4 + 1, 2)), dim3(256), 0, stream, in, out, n, mean, stdv));
}

B200_API int b200_scale(const float* a, float* out, size_t n, const float* scale_dev, float scale_host,
                        cudaStream_t stream) {
  return static_cast<int>(launch_pdl(scale_kernel, dim3(grid1d(n)), dim3(256), 0, stream, a, out, n, scale_dev, scale_host));
}

B200_API int b200_add(const float* a, const float* b, float* out, size_t n, cudaStream_t stream) {
  add_kernel<<<grid1d(n), 256, 0, stream>>>(a, b, out, n);
  B200_LAUNCH_CHECK();
}

B200_API int b200_colsum(const float* a, float* out, int M, int N, int accumulate, cudaStream_t stream) {
  colsum_kernel<<<ceil_div(N, 32), 256, 0, stream>>>(a, out, M, N, accumulate);
  B200_LAUNCH_CHECK();
}

B200_API int b200_colsum_bf16(const void* a, void* out, int M, int N, int accumulate, cudaStream_t stream) {
  colsum_bf16_kernel<<<ceil_div(N, 32), 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(a),
                                                          static_cast<__nv_bfloat16*>(out), M, N, accumulate);
  B200_LAUNCH_CHECK();
}

// acc: fp32 [N], zero on entry; N must be even.  Follow with b200_cast_from_f32 to get a bf16 gradient.
B200_API int b200_colsum_bf16_acc(const void* a, float* acc, int M, int N, cudaStream_t stream) {
  if (N % 2) return static_cast<int>(cudaErrorInvalidValue);
  const int rows_per_block = 256;
  dim3 grid(ceil_div(N, 64), ceil_div(M, rows_per_block));
  colsum_bf16_acc_kernel<<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(a), acc, M, N, rows_per_block);
  B200_LAUNCH_CHECK();
}

// acc: fp32 [>= N], ticket: one word; both zero on entry, left zero on exit.  N must be even.
B200_API int b200_colsum_bf16_fused(const void* a, void* out, float* acc, unsigned* ticket, int M, int N, cudaStream_t stream) {
  if (N % 8 || (reinterpret_cast<uintptr_t>(a) & 15)) return static_cast<int>(cudaErrorInvalidValue);
  const int col_blocks = ceil_div(N, 256);
  int rows_per_block = 256;
  while (rows_per_block > 32 && col_blocks * ceil_div(M, rows_per_block) < 2 * kNumSMs) rows_per_block >>= 1;
  dim3 grid(col_blocks, ceil_div(M, rows_per_block));
  colsum_bf16_fused_kernel<<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(a), static_cast<__nv_bfloat16*>(out), acc,
                                                     ticket, M, N, rows_per_block);
  B200_LAUNCH_CHECK();
}
