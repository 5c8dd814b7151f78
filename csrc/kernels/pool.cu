// Pooling on NHWC tensors: 3x3/s2/p1 max pool (ResNet stem) and global average pool.
// Replaces ATen max_pool2d_with_indices / mean.dim and their backward twins
// (reached from /root/reference/dpp.py:50-52).
#include <float.h>

#include "common.cuh"
#include "b200ddp.h"

namespace b200 {

// Window scan order (r then s) and the strict '>' test match ATen, so ties (frequent after
// ReLU: many zeros) route the gradient to the same element as the reference stack.
__global__ void __launch_bounds__(256) maxpool3x3s2_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                               uint8_t* __restrict__ idx, int Nb, int H, int W, int C,
                                                               int P, int Q) {
  pdl_launch_dependents();
  pdl_wait();
  const int c4n = C >> 2;
  const size_t total = (size_t)Nb * P * Q * c4n;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % c4n) * 4;
    size_t pix = i / c4n;
    const int q = (int)(pix % Q); pix /= Q;
    const int p = (int)(pix % P);
    const int n = (int)(pix / P);
    float4 best = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX);
    uchar4 bi = make_uchar4(255, 255, 255, 255);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = p * 2 - 1 + r;
      if (h < 0 || h >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int w = q * 2 - 1 + s;
        if (w < 0 || w >= W) continue;
        const float4 v = ldg4(x + ((size_t)(n * H + h) * W + w) * C + c);
        const uint8_t code = (uint8_t)(r * 3 + s);
        if (v.x > best.x || bi.x == 255) { best.x = v.x; bi.x = code; }
        if (v.y > best.y || bi.y == 255) { best.y = v.y; bi.y = code; }
        if (v.z > best.z || bi.z == 255) { best.z = v.z; bi.z = code; }
        if (v.w > best.w || bi.w == 255) { best.w = v.w; bi.w = code; }
      }
    }
    const size_t o = ((size_t)(n * P + p) * Q + q) * C + c;
    st4(y + o, best);
    *reinterpret_cast<uchar4*>(idx + o) = bi;
  }
}

// Gather form of the backward: every input element looks at the (at most 2x2) windows that
// contain it -- no atomics, deterministic.
__global__ void __launch_bounds__(256) maxpool3x3s2_bwd_kernel(const float* __restrict__ dy,
                                                               const uint8_t* __restrict__ idx, float* __restrict__ dx,
                                                               int Nb, int H, int W, int C, int P, int Q) {
  pdl_launch_dependents();
  pdl_wait();
  const int c4n = C >> 2;
  const size_t total = (size_t)Nb * H * W * c4n;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % c4n) * 4;
    size_t pix = i / c4n;
    const int w = (int)(pix % W); pix /= W;
    const int h = (int)(pix % H);
    const int n = (int)(pix / H);
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    // windows p with p*2-1 <= h <= p*2+1  ->  p in [(h-1)/2 (ceil) , (h+1)/2 (floor)]
    const int p_lo = max(0, (h) / 2), p_hi = min(P - 1, (h + 1) / 2);
    const int q_lo = max(0, (w) / 2), q_hi = min(Q - 1, (w + 1) / 2);
    for (int p = p_lo; p <= p_hi; ++p) {
      const int r = h - (p * 2 - 1);
      if (r < 0 || r > 2) continue;
      for (int q = q_lo; q <= q_hi; ++q) {
        const int s = w - (q * 2 - 1);
        if (s < 0 || s > 2) continue;
        const size_t o = ((size_t)(n * P + p) * Q + q) * C + c;
        const uchar4 bi = *reinterpret_cast<const uchar4*>(idx + o);
        const float4 d = ldg4(dy + o);
        const uint8_t code = (uint8_t)(r * 3 + s);
        if (bi.x == code) g.x += d.x;
        if (bi.y == code) g.y += d.y;
        if (bi.z == code) g.z += d.z;
        if (bi.w == code) g.w += d.w;
      }
    }
    st4(dx + ((size_t)(n * H + h) * W + w) * C + c, g);
  }
}

__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int Nb,
                                                          int HW, int C) {
  const size_t total = (size_t)Nb * C;
  const float inv = 1.f / (float)HW;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C), n = (int)(i / C);
    float s = 0.f;
    for (int p = 0; p < HW; ++p) s += __ldg(x + ((size_t)n * HW + p) * C + c);
    y[i] = s * inv;
  }
}

__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int Nb,
                                                          int HW, int C) {
  const size_t total = (size_t)Nb * HW * C;
  const float inv = 1.f / (float)HW;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int n = (int)(i / ((size_t)HW * C));
    dx[i] = __ldg(dy + (size_t)n * C + c) * inv;
  }
}

static int grid1d(size_t n) {
  size_t g = (n + 255) / 256;
  return (int)(g < 1 ? 1 : (g > 8 * kNumSMs ? 8 * kNumSMs : g));
}

}  // namespace b200

using namespace b200;

B200_API int b200_maxpool3x3s2_fwd(const float* x, float* y, uint8_t* idx, int Nb, int H, int W, int C,
                                   cudaStream_t stream) {
  if (C % 4) return static_cast<int>(cudaErrorInvalidValue);
  const int P = (H + 2 - 3) / 2 + 1, Q = (W + 2 - 3) / 2 + 1;
  return static_cast<int>(launch_pdl(maxpool3x3s2_fwd_kernel, dim3(grid1d((size_t)Nb * P * Q * C / 4)), dim3(256), 0, stream,
                                     x, y, idx, Nb, H, W, C, P, Q));
}

B200_API int b200_maxpool3x3s2_bwd(const float* dy, const uint8_t* idx, float* dx, int Nb, int H, int W, int C,
                                   cudaStream_t stream) {
  if (C % 4) return static_cast<int>(cudaErrorInvalidValue);
  const int P = (H + 2 - 3) / 2 + 1, Q = (W + 2 - 3) / 2 + 1;
  return static_cast<int>(launch_pdl(maxpool3x3s2_bwd_kernel, dim3(grid1d((size_t)Nb * H * W * C / 4)), dim3(256), 0, stream,
                                     dy, idx, dx, Nb, H, W, C, P, Q));
}

B200_API int b200_avgpool_fwd(const float* x, float* y, int Nb, int HW, int C, cudaStream_t stream) {
  avgpool_fwd_kernel<<<grid1d((size_t)Nb * C), 256, 0, stream>>>(x, y, Nb, HW, C);
  B200_LAUNCH_CHECK();
}

B200_API int b200_avgpool_bwd(const float* dy, float* dx, int Nb, int HW, int C, cudaStream_t stream) {
  avgpool_bwd_kernel<<<grid1d((size_t)Nb * HW * C), 256, 0, stream>>>(dy, dx, Nb, HW, C);
  B200_LAUNCH_CHECK();
}
