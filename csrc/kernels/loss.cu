// Fused softmax cross-entropy (mean reduction): one launch produces the scalar loss AND the
// logits gradient (softmax - onehot) / B, replacing _log_softmax + nll_loss_forward and their
// two backward kernels (nn.CrossEntropyLoss at /root/reference/dpp.py:40,51).
#include <float.h>

#include "common.cuh"
#include "b200ddp.h"

namespace b200 {

// One warp per row; rows are short (10 classes for dpp.py) or long (50 257 for GPT-2).
// Single block: rows are strided over the block's warps, the loss is block-reduced without
// atomics so the result is deterministic.
__global__ void __launch_bounds__(1024) softmax_xent_kernel(const float* __restrict__ logits,
                                                            const long long* __restrict__ target,
                                                            float* __restrict__ loss, float* __restrict__ dlogits,
                                                            float* __restrict__ row_loss, int B, int V,
                                                            long long ignore_index) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const float inv_b = 1.f / (float)B;
  for (int row = blockIdx.x * nwarps + warp; row < B; row += gridDim.x * nwarps) {
    const float* z = logits + (size_t)row * V;
    float mx = -FLT_MAX;
    for (int j = lane; j < V; j += 32) mx = fmaxf(mx, z[j]);
    mx = warp_max(mx);
    float se = 0.f;
    for (int j = lane; j < V; j += 32) se += __expf(z[j] - mx);
    se = warp_sum(se);
    const float lse = mx + __logf(se);
    const long long t = target[row];
    const bool live = t != ignore_index;
    const float inv_se = 1.f / se;
    if (dlogits != nullptr) {
      float* d = dlogits + (size_t)row * V;
      for (int j = lane; j < V; j += 32) {
        const float p = __expf(z[j] - mx) * inv_se;
        d[j] = live ? (p - (j == (int)t ? 1.f : 0.f)) * inv_b : 0.f;
      }
    }
    if (lane == 0) row_loss[row] = live ? (lse - z[t]) : 0.f;
  }
}

__global__ void __launch_bounds__(1024) mean_rows_kernel(const float* __restrict__ row_loss, float* __restrict__ loss,
                                                         int B) {
  __shared__ float sh[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < B; i += blockDim.x) s += row_loss[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) *loss = v / (float)B;
  }
}

// Small-batch fast path: B <= 32 rows handled by ONE block, loss reduced in the same launch.
__global__ void __launch_bounds__(1024) softmax_xent_small_kernel(const float* __restrict__ logits,
                                                                  const long long* __restrict__ target,
                                                                  float* __restrict__ loss,
                                                                  float* __restrict__ dlogits, int B, int V) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float sh[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float my = 0.f;
  if (warp < B) {
    const float* z = logits + (size_t)warp * V;
    float mx = -FLT_MAX;
    for (int j = lane; j < V; j += 32) mx = fmaxf(mx, z[j]);
    mx = warp_max(mx);
    float se = 0.f;
    for (int j = lane; j < V; j += 32) se += __expf(z[j] - mx);
    se = warp_sum(se);
    const int t = (int)target[warp];
    const float inv_se = 1.f / se, inv_b = 1.f / (float)B;
    float* d = dlogits + (size_t)warp * V;
    for (int j = lane; j < V; j += 32) d[j] = (__expf(z[j] - mx) * inv_se - (j == t ? 1.f : 0.f)) * inv_b;
    my = mx + __logf(se) - z[t];
  }
  if (lane == 0) sh[warp] = warp < B ? my : 0.f;
  __syncthreads();
  if (warp == 0) {
    float v = sh[lane];
    v = warp_sum(v);
    if (lane == 0) *loss = v / (float)B;
  }
}

}  // namespace b200

using namespace b200;

B200_API int b200_softmax_xent(const float* logits, const long long* target, float* loss, float* dlogits,
                               float* row_loss_scratch, int B, int V, cudaStream_t stream) {
  if (B <= 32 && dlogits != nullptr) {
    (void)launch_pdl(softmax_xent_small_kernel, dim3(1), dim3(1024), 0, stream, logits, target, loss, dlogits, B, V);
    B200_LAUNCH_CHECK();
  }
  const int warps = 8;
  const int grid = min(ceil_div(B, warps), 8 * kNumSMs);
  softmax_xent_kernel<<<grid, warps * 32, 0, stream>>>(logits, target, loss, dlogits, row_loss_scratch, B, V, -100);
  mean_rows_kernel<<<1, 1024, 0, stream>>>(row_loss_scratch, loss, B);
  B200_LAUNCH_CHECK();
}
