// Training-mode BatchNorm2d on NHWC tensors, fused with the residual add and ReLU that
// follow it in every ResNet block.
//
// Reference path: /root/reference/dpp.py:50-52 runs torchvision's BasicBlock, which issues
// per conv: batch_norm (+ `num_batches_tracked += 1`), relu_, and for the block output add_
// + relu_.  Here:
//
//   forward  : [conv epilogue accumulates per-channel sum / sum^2]  ->  bn_fwd_apply
//              (normalise + gamma/beta + residual + ReLU + running-stat update + counter bump)
//   backward : bn_bwd_reduce (sum dy, sum dy*xhat, ReLU mask applied on the fly)  ->  bn_bwd_apply
//              (dx, residual gradient, dgamma/dbeta written to their gradient-bucket slots)
//
// Layout: x is [M, C] with C contiguous (M = N*H*W).  Threads own float4 channel groups.
// The statistic accumulators are self-cleaning: the forward apply zeroes the backward sums
// and vice versa, so no memset launches are needed in steady state.
#include "common.cuh"
#include "b200ddp.h"

namespace b200 {

// ---- per-channel sum / sum of squares (used when the producer did not fuse them) ---------
__global__ void __launch_bounds__(256) bn_stats_kernel(const float* __restrict__ x, float* __restrict__ sums,
                                                       int M, int C) {
  // blockIdx.y picks a chunk of up to 256 float4 channel groups; inside the chunk a thread owns
  // one channel group and walks rows, so partial sums stay in registers.
  __shared__ float sh[2048];
  const int tid = threadIdx.x, lanes = blockDim.x;
  const int c4n = C >> 2;
  const int cbase4 = blockIdx.y * 256;
  const int cw = min(c4n - cbase4, 256);
  for (int i = tid; i < 2048; i += lanes) sh[i] = 0.f;
  __syncthreads();
  const int rows_per_iter = max(1, lanes / cw);
  const int my_l = tid % cw, my_r = tid / cw;
  if (my_r < rows_per_iter) {
    const int c = (cbase4 + my_l) * 4;
    float4 s = make_float4(0, 0, 0, 0), q = make_float4(0, 0, 0, 0);
    for (int r = blockIdx.x * rows_per_iter + my_r; r < M; r += gridDim.x * rows_per_iter) {
      const float4 v = ldg4(x + (size_t)r * C + c);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      q.x += v.x * v.x; q.y += v.y * v.y; q.z += v.z * v.z; q.w += v.w * v.w;
    }
    atomicAdd(&sh[my_l * 4 + 0], s.x); atomicAdd(&sh[my_l * 4 + 1], s.y);
    atomicAdd(&sh[my_l * 4 + 2], s.z); atomicAdd(&sh[my_l * 4 + 3], s.w);
    atomicAdd(&sh[1024 + my_l * 4 + 0], q.x); atomicAdd(&sh[1024 + my_l * 4 + 1], q.y);
    atomicAdd(&sh[1024 + my_l * 4 + 2], q.z); atomicAdd(&sh[1024 + my_l * 4 + 3], q.w);
  }
  __syncthreads();
  for (int i = tid; i < cw * 4; i += lanes) {
    atomicAdd(sums + cbase4 * 4 + i, sh[i]);
    atomicAdd(sums + C + cbase4 * 4 + i, sh[1024 + i]);
  }
}

// ---- forward apply -------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bn_fwd_apply_kernel(
    const float* __restrict__ x, const float* __restrict__ sums, const float* __restrict__ gamma,
    const float* __restrict__ beta, const float* __restrict__ residual, float* __restrict__ out,
    float* running_mean, float* running_var, long long* num_batches_tracked, float* __restrict__ save_mean,
    float* __restrict__ save_invstd, float* zero_after, int zero_count, int M, int C, float momentum, float eps,
    int relu, int training) {
  const float inv_m = 1.f / (float)M;
  const int c4n = C >> 2;
  // the LAST block only does the per-channel bookkeeping (saved statistics, running statistics, clearing the other
  // accumulator), concurrently with the blocks that normalise: its dependent L2 round trips no longer extend the kernel
  const unsigned work_blocks = gridDim.x - 1;
  if (blockIdx.x == work_blocks) {
    if (blockIdx.y != 0) return;
    if (training) {
      for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float m = sums[c] * inv_m;
        const float var = fmaxf(sums[C + c] * inv_m - m * m, 0.f);
        save_mean[c] = m;
        save_invstd[c] = rsqrtf(var + eps);
        if (running_mean != nullptr) {
          const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
          running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
          running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
        }
      }
      if (threadIdx.x == 0 && num_batches_tracked != nullptr) *num_batches_tracked += 1;
    }
    for (int i = threadIdx.x; i < zero_count; i += blockDim.x) zero_after[i] = 0.f;
    return;
  }
  // A thread keeps ONE channel quad for the whole kernel: scale / shift are derived once and the loop over rows is
  // pure streaming (no per-element index arithmetic or constant reloads), four rows in flight per thread.
  const int cw = min(c4n, 256);                               // channel quads served by this block column
  const int rows_per_iter = 256 / cw;
  const int my_l = threadIdx.x % cw, my_r = threadIdx.x / cw;
  const int cq = blockIdx.y * 256 + my_l;
  if (my_r >= rows_per_iter || cq >= c4n) return;
  const int c = cq * 4;
  float sc[4], sh[4];
  {
    const float4 g = ldg4(gamma + c), b = ldg4(beta + c);
    const float gg[4] = {g.x, g.y, g.z, g.w}, bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float m, is;
      if (training) {
        m = __ldg(sums + c + j) * inv_m;
        const float var = fmaxf(__ldg(sums + C + c + j) * inv_m - m * m, 0.f);
        is = rsqrtf(var + eps);
      } else {
        m = running_mean[c + j];
        is = rsqrtf(running_var[c + j] + eps);
      }
      sc[j] = is * gg[j];
      sh[j] = bb[j] - m * sc[j];
    }
  }
  const int row_stride = (int)work_blocks * rows_per_iter;
  auto one = [&](float4 v, float4 r) {
    float4 o;
    o.x = v.x * sc[0] + sh[0] + r.x;
    o.y = v.y * sc[1] + sh[1] + r.y;
    o.z = v.z * sc[2] + sh[2] + r.z;
    o.w = v.w * sc[3] + sh[3] + r.w;
    if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
    return o;
  };
  int r = blockIdx.x * rows_per_iter + my_r;
  for (; r + 3 * row_stride < M; r += 4 * row_stride) {
    float4 v[4], q[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t off = (size_t)(r + u * row_stride) * C + c;
      v[u] = ldg4(x + off);
      q[u] = residual != nullptr ? ldg4(residual + off) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) st4(out + (size_t)(r + u * row_stride) * C + c, one(v[u], q[u]));
  }
  for (; r < M; r += row_stride) {
    const size_t off = (size_t)r * C + c;
    const float4 q = residual != nullptr ? ldg4(residual + off) : make_float4(0.f, 0.f, 0.f, 0.f);
    st4(out + off, one(ldg4(x + off), q));
  }
}

// ---- backward reduce: sums[0:C] = sum dy_eff, sums[C:2C] = sum dy_eff * xhat ---------------------
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(
    const float* __restrict__ dy, const float* __restrict__ out, const float* __restrict__ x,
    const float* __restrict__ save_mean, const float* __restrict__ save_invstd, float* __restrict__ sums,
    int M, int C, int relu) {
  __shared__ float sh[2048];
  const int tid = threadIdx.x, lanes = blockDim.x;
  const int c4n = C >> 2;
  const int cbase4 = blockIdx.y * 256;
  const int cw = min(c4n - cbase4, 256);
  for (int i = tid; i < 2048; i += lanes) sh[i] = 0.f;
  __syncthreads();
  const int rows_per_iter = max(1, lanes / cw);
  const int my_l = tid % cw, my_r = tid / cw;
  if (my_r < rows_per_iter) {
    const int c = (cbase4 + my_l) * 4;
    const float4 mu = ldg4(save_mean + c), is = ldg4(save_invstd + c);
    float4 s = make_float4(0, 0, 0, 0), q = make_float4(0, 0, 0, 0);
    auto add_row = [&](float4 g, float4 o, float4 v) {
      if (relu) {
        g.x = o.x > 0.f ? g.x : 0.f; g.y = o.y > 0.f ? g.y : 0.f;
        g.z = o.z > 0.f ? g.z : 0.f; g.w = o.w > 0.f ? g.w : 0.f;
      }
      s.x += g.x; s.y += g.y; s.z += g.z; s.w += g.w;
      q.x += g.x * (v.x - mu.x) * is.x; q.y += g.y * (v.y - mu.y) * is.y;
      q.z += g.z * (v.z - mu.z) * is.z; q.w += g.w * (v.w - mu.w) * is.w;
    };
    const int row_stride = gridDim.x * rows_per_iter;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    int r = blockIdx.x * rows_per_iter + my_r;
    for (; r + 2 * row_stride < M; r += 3 * row_stride) {        // three rows (up to nine 16-byte loads) in flight
      float4 g[3], o[3], v[3];
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const size_t off = (size_t)(r + u * row_stride) * C + c;
        g[u] = ldg4(dy + off);
        o[u] = relu ? ldg4(out + off) : zero4;
        v[u] = ldg4(x + off);
      }
#pragma unroll
      for (int u = 0; u < 3; ++u) add_row(g[u], o[u], v[u]);
    }
    for (; r < M; r += row_stride) {
      const size_t off = (size_t)r * C + c;
      add_row(ldg4(dy + off), relu ? ldg4(out + off) : zero4, ldg4(x + off));
    }
    atomicAdd(&sh[my_l * 4 + 0], s.x); atomicAdd(&sh[my_l * 4 + 1], s.y);
    atomicAdd(&sh[my_l * 4 + 2], s.z); atomicAdd(&sh[my_l * 4 + 3], s.w);
    atomicAdd(&sh[1024 + my_l * 4 + 0], q.x); atomicAdd(&sh[1024 + my_l * 4 + 1], q.y);
    atomicAdd(&sh[1024 + my_l * 4 + 2], q.z); atomicAdd(&sh[1024 + my_l * 4 + 3], q.w);
  }
  __syncthreads();
  for (int i = tid; i < cw * 4; i += lanes) {
    atomicAdd(sums + cbase4 * 4 + i, sh[i]);
    atomicAdd(sums + C + cbase4 * 4 + i, sh[1024 + i]);
  }
}

// ---- backward apply -----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(
    const float* __restrict__ dy, const float* __restrict__ out, const float* __restrict__ x,
    const float* __restrict__ gamma, const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
    const float* __restrict__ sums, float* __restrict__ dx, float* __restrict__ dres, float* __restrict__ dgamma,
    float* __restrict__ dbeta, int accumulate_param_grads, float* zero_after, int zero_count, int M, int C,
    int relu) {
  const float inv_m = 1.f / (float)M;
  const size_t total4 = (size_t)M * C / 4;
  const int c4n = C >> 2;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % c4n) * 4;
    const float4 mu = ldg4(save_mean + c), is = ldg4(save_invstd + c), ga = ldg4(gamma + c);
    const float4 sd = ldg4(sums + c), sx = ldg4(sums + C + c);
    float4 g = ldg4(dy + i * 4);
    if (relu) {
      const float4 o = ldg4(out + i * 4);
      g.x = o.x > 0.f ? g.x : 0.f; g.y = o.y > 0.f ? g.y : 0.f;
      g.z = o.z > 0.f ? g.z : 0.f; g.w = o.w > 0.f ? g.w : 0.f;
    }
    if (dres != nullptr) st4(dres + i * 4, g);
    const float4 v = ldg4(x + i * 4);
    float4 d;
    d.x = ga.x * is.x * (g.x - sd.x * inv_m - (v.x - mu.x) * is.x * sx.x * inv_m);
    d.y = ga.y * is.y * (g.y - sd.y * inv_m - (v.y - mu.y) * is.y * sx.y * inv_m);
    d.z = ga.z * is.z * (g.z - sd.z * inv_m - (v.z - mu.z) * is.z * sx.z * inv_m);
    d.w = ga.w * is.w * (g.w - sd.w * inv_m - (v.w - mu.w) * is.w * sx.w * inv_m);
    st4(dx + i * 4, d);
  }
  if (blockIdx.x == 0) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      if (accumulate_param_grads) {
        dbeta[c] += sums[c];
        dgamma[c] += sums[C + c];
      } else {
        dbeta[c] = sums[c];
        dgamma[c] = sums[C + c];
      }
    }
    for (int i = threadIdx.x; i < zero_count; i += blockDim.x) zero_after[i] = 0.f;
  }
}

// ---- backward, single launch: reduce -> grid barrier -> apply ------------------------------------
// All blocks are co-resident (grid <= 120, one or more per SM), so after publishing its partial
// sums a block may wait for the others on a ticket counter and then apply dx with the complete
// sums.  Saves one launch per BatchNorm layer per step and keeps dy / x hot in L2 between phases.
__global__ void __launch_bounds__(256) bn_bwd_fused_kernel(
    const float* __restrict__ dy, const float* __restrict__ out, const float* __restrict__ x,
    const float* __restrict__ gamma, const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
    float* __restrict__ sums, float* __restrict__ dx, float* __restrict__ dres, float* __restrict__ dgamma,
    float* __restrict__ dbeta, int accumulate_param_grads, float* zero_after, int zero_count, unsigned* tickets,
    int M, int C, int relu) {
  __shared__ float sh[2048];
  const int tid = threadIdx.x, lanes = blockDim.x;
  const int c4n = C >> 2;
  // ---- phase 1: partial sums (every block walks all channel chunks for its row stripe) ----
  for (int cbase4 = 0; cbase4 < c4n; cbase4 += 256) {
    const int cw = min(c4n - cbase4, 256);
    for (int i = tid; i < 2048; i += lanes) sh[i] = 0.f;
    __syncthreads();
    const int rows_per_iter = max(1, lanes / cw);
    const int my_l = tid % cw, my_r = tid / cw;
    if (my_r < rows_per_iter) {
      const int c = (cbase4 + my_l) * 4;
      const float4 mu = ldg4(save_mean + c), is = ldg4(save_invstd + c);
      float4 s = make_float4(0, 0, 0, 0), q = make_float4(0, 0, 0, 0);
      for (int r = blockIdx.x * rows_per_iter + my_r; r < M; r += gridDim.x * rows_per_iter) {
        const size_t off = (size_t)r * C + c;
        float4 g = ldg4(dy + off);
        if (relu) {
          const float4 o = ldg4(out + off);
          g.x = o.x > 0.f ? g.x : 0.f; g.y = o.y > 0.f ? g.y : 0.f;
          g.z = o.z > 0.f ? g.z : 0.f; g.w = o.w > 0.f ? g.w : 0.f;
        }
        const float4 v = ldg4(x + off);
        s.x += g.x; s.y += g.y; s.z += g.z; s.w += g.w;
        q.x += g.x * (v.x - mu.x) * is.x; q.y += g.y * (v.y - mu.y) * is.y;
        q.z += g.z * (v.z - mu.z) * is.z; q.w += g.w * (v.w - mu.w) * is.w;
      }
      atomicAdd(&sh[my_l * 4 + 0], s.x); atomicAdd(&sh[my_l * 4 + 1], s.y);
      atomicAdd(&sh[my_l * 4 + 2], s.z); atomicAdd(&sh[my_l * 4 + 3], s.w);
      atomicAdd(&sh[1024 + my_l * 4 + 0], q.x); atomicAdd(&sh[1024 + my_l * 4 + 1], q.y);
      atomicAdd(&sh[1024 + my_l * 4 + 2], q.z); atomicAdd(&sh[1024 + my_l * 4 + 3], q.w);
    }
    __syncthreads();
    for (int i = tid; i < cw * 4; i += lanes) {
      atomicAdd(sums + cbase4 * 4 + i, sh[i]);
      atomicAdd(sums + C + cbase4 * 4 + i, sh[1024 + i]);
    }
    __syncthreads();
  }
  // ---- grid barrier ----
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    atomicAdd(tickets, 1u);
    unsigned seen, spins = 0;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(tickets) : "memory");
    } while (seen < gridDim.x && ++spins < (1u << 28));
  }
  __syncthreads();
  // ---- phase 2: apply ----
  const float inv_m = 1.f / (float)M;
  const size_t total4 = (size_t)M * C / 4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % c4n) * 4;
    const float4 mu = ldg4(save_mean + c), is = ldg4(save_invstd + c), ga = ldg4(gamma + c);
    const float4 sd = __ldcg(reinterpret_cast<const float4*>(sums + c));
    const float4 sx = __ldcg(reinterpret_cast<const float4*>(sums + C + c));
    float4 g = ldg4(dy + i * 4);
    if (relu) {
      const float4 o = ldg4(out + i * 4);
      g.x = o.x > 0.f ? g.x : 0.f; g.y = o.y > 0.f ? g.y : 0.f;
      g.z = o.z > 0.f ? g.z : 0.f; g.w = o.w > 0.f ? g.w : 0.f;
    }
    if (dres != nullptr) st4(dres + i * 4, g);
    const float4 v = ldg4(x + i * 4);
    float4 d;
    d.x = ga.x * is.x * (g.x - sd.x * inv_m - (v.x - mu.x) * is.x * sx.x * inv_m);
    d.y = ga.y * is.y * (g.y - sd.y * inv_m - (v.y - mu.y) * is.y * sx.y * inv_m);
    d.z = ga.z * is.z * (g.z - sd.z * inv_m - (v.z - mu.z) * is.z * sx.z * inv_m);
    d.w = ga.w * is.w * (g.w - sd.w * inv_m - (v.w - mu.w) * is.w * sx.w * inv_m);
    st4(dx + i * 4, d);
  }
  if (blockIdx.x == 0) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const float sb = __ldcg(sums + c), sg = __ldcg(sums + C + c);
      if (accumulate_param_grads) { dbeta[c] += sb; dgamma[c] += sg; }
      else { dbeta[c] = sb; dgamma[c] = sg; }
    }
    for (int i = threadIdx.x; i < zero_count; i += blockDim.x) zero_after[i] = 0.f;
  }
  // last block to leave re-arms the barrier
  __syncthreads();
  if (tid == 0) {
    const unsigned left = atomicAdd(tickets + 1, 1u);
    if (left == gridDim.x - 1) { tickets[0] = 0u; tickets[1] = 0u; }
  }
}

static int grid_for(size_t work_items, int per_block, int cap) {
  size_t g = (work_items + per_block - 1) / per_block;
  if (g < 1) g = 1;
  if (g > (size_t)cap) g = cap;
  return (int)g;
}

}  // namespace b200

using namespace b200;

B200_API int b200_bn_stats(const float* x, float* sums, int M, int C, cudaStream_t stream) {
  if (C % 4 != 0) return static_cast<int>(cudaErrorInvalidValue);
  const int chunks = ceil_div(C / 4, 256);
  const int rows_per_block = max(1, 256 / min(C / 4, 256));
  const int gx = grid_for((size_t)M, rows_per_block * 8, max(1, 2 * kNumSMs / chunks));
  bn_stats_kernel<<<dim3(gx, chunks), 256, 0, stream>>>(x, sums, M, C);
  B200_LAUNCH_CHECK();
}

B200_API int b200_bn_fwd_apply(const float* x, const float* sums, const float* gamma, const float* beta,
                               const float* residual, float* out, float* running_mean, float* running_var,
                               long long* num_batches_tracked, float* save_mean, float* save_invstd,
                               float* zero_after, int zero_count, int M, int C, float momentum, float eps,
                               int relu, int training, cudaStream_t stream) {
  if (C % 4 != 0) return static_cast<int>(cudaErrorInvalidValue);
  const int chunks = ceil_div(C / 4, 256);                  // block columns of 256 channel quads
  const int rows_per_block = max(1, 256 / min(C / 4, 256));
  // ~2 rows per thread on small tensors (latency), up to 4 blocks per SM on large ones (bandwidth)
  const int gx = grid_for((size_t)M, rows_per_block * 2, max(1, 4 * kNumSMs / chunks)) + 1;   // + the bookkeeping block
  bn_fwd_apply_kernel<<<dim3(gx, chunks), 256, 0, stream>>>(x, sums, gamma, beta, residual, out, running_mean, running_var,
                                                num_batches_tracked, save_mean, save_invstd, zero_after, zero_count,
                                                M, C, momentum, eps, relu, training);
  B200_LAUNCH_CHECK();
}

B200_API int b200_bn_bwd_reduce(const float* dy, const float* out, const float* x, const float* save_mean,
                                const float* save_invstd, float* sums, int M, int C, int relu,
                                cudaStream_t stream) {
  if (C % 4 != 0) return static_cast<int>(cudaErrorInvalidValue);
  const int chunks = ceil_div(C / 4, 256);
  const int rows_per_block = max(1, 256 / min(C / 4, 256));
  const int gx = grid_for((size_t)M, rows_per_block * 2, max(1, 4 * kNumSMs / chunks));   // <= 2 rows per thread
  bn_bwd_reduce_kernel<<<dim3(gx, chunks), 256, 0, stream>>>(dy, out, x, save_mean, save_invstd, sums, M, C, relu);
  B200_LAUNCH_CHECK();
}

B200_API int b200_bn_bwd_apply(const float* dy, const float* out, const float* x, const float* gamma,
                               const float* save_mean, const float* save_invstd, const float* sums, float* dx,
                               float* dres, float* dgamma, float* dbeta, int accumulate_param_grads,
                               float* zero_after, int zero_count, int M, int C, int relu, cudaStream_t stream) {
  if (C % 4 != 0) return static_cast<int>(cudaErrorInvalidValue);
  const int grid = grid_for((size_t)M * C / 4, 256 * 2, 4 * kNumSMs);
  bn_bwd_apply_kernel<<<grid, 256, 0, stream>>>(dy, out, x, gamma, save_mean, save_invstd, sums, dx, dres, dgamma,
                                                dbeta, accumulate_param_grads, zero_after, zero_count, M, C, relu);
  B200_LAUNCH_CHECK();
}

B200_API int b200_bn_bwd_fused(const float* dy, const float* out, const float* x, const float* gamma,
                               const float* save_mean, const float* save_invstd, float* sums, float* dx, float* dres,
                               float* dgamma, float* dbeta, int accumulate_param_grads, float* zero_after,
                               int zero_count, unsigned* tickets, int M, int C, int relu, cudaStream_t stream) {
  if (C % 4 != 0) return static_cast<int>(cudaErrorInvalidValue);
  // all blocks must be resident at once (they wait for each other): <= 120 blocks of 256 threads
  const int grid = grid_for((size_t)M * C / 4, 256 * 2, 120);
  bn_bwd_fused_kernel<<<grid, 256, 0, stream>>>(dy, out, x, gamma, save_mean, save_invstd, sums, dx, dres, dgamma,
                                                dbeta, accumulate_param_grads, zero_after, zero_count, tickets, M, C,
                                                relu);
  B200_LAUNCH_CHECK();
}
