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
#include "tcgen05.cuh"
#include "b200ddp.h"

namespace b200 {

// ---- per-channel sum / sum of squares (used when the producer did not fuse them) ---------
__global__ void __launch_bounds__(256) bn_stats_kernel(const float* __restrict__ x, float* __restrict__ sums,
                                                       int M, int C) {
  pdl_launch_dependents();
  pdl_wait();
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
  pdl_launch_dependents();
  pdl_wait();
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
  pdl_launch_dependents();
  const int tid = threadIdx.x, lanes = blockDim.x;
  const int c4n = C >> 2;
  const int cbase4 = blockIdx.y * 256;
  const int cw = min(c4n - cbase4, 256);
  for (int i = tid; i < 2048; i += lanes) sh[i] = 0.f;
  __syncthreads();
  pdl_wait();
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
  pdl_launch_dependents();
  pdl_wait();
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


// =================================================================================================================
// Cluster BatchNorm: ONE kernel per direction for activation tensors of up to 131 072 elements (every BatchNorm of
// the dpp.py model behind the stem: 2048x64, 512x128, 128x256, 32x512).
//
// One thread-block cluster of 8 CTAs x 512 threads holds the whole tensor in registers (8 float4 per thread, all of
// the same channel quad).  Per-channel sums are reduced inside a CTA through shared memory and across the 8 CTAs
// through distributed shared memory (ld.shared::cluster), in a FIXED order -- results are bit-reproducible, unlike
// the float-atomic accumulators of the streaming kernels -- and the inputs are read exactly once.
//   backward : reduce (sum dy, sum dy*xhat) + apply in one launch instead of two (both phases sit on the critical
//              chain of the backward pass);
//   forward  : exact two-pass statistics (mean first, then sum (x-mean)^2 from the registers: no E[x^2]-E[x]^2
//              cancellation for large-mean activations) + normalise + residual + ReLU + running-stat update.
// =================================================================================================================
constexpr int kClCtas = 8, kClThreads = 512, kClVec = 8;
constexpr int kClMaxElems = kClCtas * kClThreads * kClVec * 4;      // 131072

struct ClusterPlan {
  int quad, row0, row_stride;      // this thread's channel quad, first row, row step
  int same;                        // threads per CTA that share a quad
};
__device__ __forceinline__ ClusterPlan cluster_plan(int C) {
  const int cq = C >> 2;
  const int g = (int)tc::cluster_ctarank() * kClThreads + (int)threadIdx.x;
  ClusterPlan p;
  p.quad = g % cq;
  p.row0 = g / cq;
  p.row_stride = kClCtas * kClThreads / cq;
  p.same = kClThreads / cq > 0 ? kClThreads / cq : 1;
  return p;
}

// Sum `vals[NV]` (per thread, attached to channel quad `quad`) over the whole cluster in a fixed order.
// part: smem [C/4 quads][NV] partial of this CTA;  scratch: smem [kClThreads][NV];  tot: smem [C/4][NV] result.
template <int NV>
__device__ __forceinline__ void cluster_channel_sum(const float (&vals)[NV], int C, float* scratch, float* part, float* tot) {
  const int cq = C >> 2, tid = threadIdx.x;
#pragma unroll
  for (int j = 0; j < NV; ++j) scratch[tid * NV + j] = vals[j];
  __syncthreads();
  // threads q, q + cq, q + 2cq ... of a CTA share channel quad q (cq divides the block size)
  for (int q = tid; q < cq; q += kClThreads) {
    float acc[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) acc[j] = 0.f;
    for (int t = q; t < kClThreads; t += cq)
#pragma unroll
      for (int j = 0; j < NV; ++j) acc[j] += scratch[t * NV + j];
#pragma unroll
    for (int j = 0; j < NV; ++j) part[q * NV + j] = acc[j];
  }
  tc::cluster_sync_all();                                   // every CTA's partials are in place
  const uint32_t part_u32 = tc::smem_u32(part);
  for (int i = tid; i < cq * NV; i += kClThreads) {
    float acc = 0.f;
#pragma unroll
    for (int r = 0; r < kClCtas; ++r) acc += tc::ld_dsmem_f32(tc::map_to_cta(part_u32 + (uint32_t)i * 4u, (uint32_t)r));
    tot[i] = acc;
  }
  tc::cluster_sync_all();                                   // nobody still reads my `part`; `tot` visible block-wide
}

__global__ void __launch_bounds__(kClThreads, 1) bn_bwd_cluster_kernel(
    const float* __restrict__ dy, const float* __restrict__ out, const float* __restrict__ x,
    const float* __restrict__ gamma, const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
    float* __restrict__ dx, float* __restrict__ dres, float* __restrict__ dgamma, float* __restrict__ dbeta,
    int accumulate_param_grads, float* zero_after, int zero_count, int M, int C, int relu) {
  extern __shared__ float sm[];
  const int cq = C >> 2;
  float* scratch = sm;                          // [512][8]
  float* part = scratch + kClThreads * 8;       // [cq][8]
  float* tot = part + cq * 8;                   // [cq][8]
  pdl_launch_dependents();
  const ClusterPlan p = cluster_plan(C);
  pdl_wait();
  const int c = p.quad * 4;
  const bool active = p.row0 < M && p.quad < cq;
  const float4 mu = ldg4(save_mean + c), is = ldg4(save_invstd + c);
  float4 g[kClVec], xh[kClVec];
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};       // sum g (4), sum g*xhat (4)
#pragma unroll
  for (int k = 0; k < kClVec; ++k) {
    const int r = p.row0 + k * p.row_stride;
    g[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    xh[k] = g[k];
    if (active && r < M) {
      const size_t off = (size_t)r * C + c;
      g[k] = ldg4(dy + off);
      const float4 v = ldg4(x + off);
      if (relu) {
        const float4 o = ldg4(out + off);
        g[k].x = o.x > 0.f ? g[k].x : 0.f; g[k].y = o.y > 0.f ? g[k].y : 0.f;
        g[k].z = o.z > 0.f ? g[k].z : 0.f; g[k].w = o.w > 0.f ? g[k].w : 0.f;
      }
      xh[k] = make_float4((v.x - mu.x) * is.x, (v.y - mu.y) * is.y, (v.z - mu.z) * is.z, (v.w - mu.w) * is.w);
    }
  }
#pragma unroll
  for (int k = 0; k < kClVec; ++k) {
    acc[0] += g[k].x; acc[1] += g[k].y; acc[2] += g[k].z; acc[3] += g[k].w;
    acc[4] += g[k].x * xh[k].x; acc[5] += g[k].y * xh[k].y; acc[6] += g[k].z * xh[k].z; acc[7] += g[k].w * xh[k].w;
  }
  cluster_channel_sum<8>(acc, C, scratch, part, tot);
  const float inv_m = 1.f / (float)M;
  const float4 ga = ldg4(gamma + c);
  const float* tq = tot + p.quad * 8;
  const float sd[4] = {tq[0] * inv_m, tq[1] * inv_m, tq[2] * inv_m, tq[3] * inv_m};
  const float sx[4] = {tq[4] * inv_m, tq[5] * inv_m, tq[6] * inv_m, tq[7] * inv_m};
  const float gs[4] = {ga.x * is.x, ga.y * is.y, ga.z * is.z, ga.w * is.w};
#pragma unroll
  for (int k = 0; k < kClVec; ++k) {
    const int r = p.row0 + k * p.row_stride;
    if (active && r < M) {
      const size_t off = (size_t)r * C + c;
      if (dres != nullptr) st4(dres + off, g[k]);
      float4 d;
      d.x = gs[0] * (g[k].x - sd[0] - xh[k].x * sx[0]);
      d.y = gs[1] * (g[k].y - sd[1] - xh[k].y * sx[1]);
      d.z = gs[2] * (g[k].z - sd[2] - xh[k].z * sx[2]);
      d.w = gs[3] * (g[k].w - sd[3] - xh[k].w * sx[3]);
      st4(dx + off, d);
    }
  }
  if (tc::cluster_ctarank() == 0) {
    for (int ch = threadIdx.x; ch < C; ch += kClThreads) {
      const float db = tot[(ch >> 2) * 8 + (ch & 3)], dg = tot[(ch >> 2) * 8 + 4 + (ch & 3)];
      if (accumulate_param_grads) {
        dbeta[ch] += db;
        dgamma[ch] += dg;
      } else {
        dbeta[ch] = db;
        dgamma[ch] = dg;
      }
    }
    for (int i = threadIdx.x; i < zero_count; i += kClThreads) zero_after[i] = 0.f;
  }
}

__global__ void __launch_bounds__(kClThreads, 1) bn_fwd_cluster_kernel(
    const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
    const float* __restrict__ residual, float* __restrict__ out, float* running_mean, float* running_var,
    long long* num_batches_tracked, float* __restrict__ save_mean, float* __restrict__ save_invstd, float* zero_after,
    int zero_count, int M, int C, float momentum, float eps, int relu) {
  extern __shared__ float sm[];
  const int cq = C >> 2;
  float* scratch = sm;                          // [512][4]
  float* part = scratch + kClThreads * 4;       // [cq][4]
  float* tot = part + cq * 4;                   // [cq][4]
  pdl_launch_dependents();
  const ClusterPlan p = cluster_plan(C);
  pdl_wait();
  const int c = p.quad * 4;
  const bool active = p.row0 < M;
  float4 v[kClVec];
  float acc[4] = {0, 0, 0, 0};
#pragma unroll
  for (int k = 0; k < kClVec; ++k) {
    const int r = p.row0 + k * p.row_stride;
    v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active && r < M) v[k] = ldg4(x + (size_t)r * C + c);
    acc[0] += v[k].x; acc[1] += v[k].y; acc[2] += v[k].z; acc[3] += v[k].w;
  }
  cluster_channel_sum<4>(acc, C, scratch, part, tot);
  const float inv_m = 1.f / (float)M;
  const float mean[4] = {tot[p.quad * 4 + 0] * inv_m, tot[p.quad * 4 + 1] * inv_m, tot[p.quad * 4 + 2] * inv_m,
                         tot[p.quad * 4 + 3] * inv_m};
  float sq[4] = {0, 0, 0, 0};
#pragma unroll
  for (int k = 0; k < kClVec; ++k) {
    const int r = p.row0 + k * p.row_stride;
    if (active && r < M) {
      const float a = v[k].x - mean[0], b = v[k].y - mean[1], cc = v[k].z - mean[2], d = v[k].w - mean[3];
      sq[0] += a * a; sq[1] += b * b; sq[2] += cc * cc; sq[3] += d * d;
    }
  }
  __syncthreads();                               // everyone has read `tot` (means) before it is overwritten
  cluster_channel_sum<4>(sq, C, scratch, part, tot);
  const float4 g4 = ldg4(gamma + c), b4 = ldg4(beta + c);
  const float gg[4] = {g4.x, g4.y, g4.z, g4.w}, bb[4] = {b4.x, b4.y, b4.z, b4.w};
  float sc[4], sh[4], var[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    var[j] = tot[p.quad * 4 + j] * inv_m;        // biased variance, exact two-pass
    const float is = rsqrtf(var[j] + eps);
    sc[j] = is * gg[j];
    sh[j] = bb[j] - mean[j] * sc[j];
  }
#pragma unroll
  for (int k = 0; k < kClVec; ++k) {
    const int r = p.row0 + k * p.row_stride;
    if (active && r < M) {
      const size_t off = (size_t)r * C + c;
      const float4 q = residual != nullptr ? ldg4(residual + off) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 o;
      o.x = v[k].x * sc[0] + sh[0] + q.x;
      o.y = v[k].y * sc[1] + sh[1] + q.y;
      o.z = v[k].z * sc[2] + sh[2] + q.z;
      o.w = v[k].w * sc[3] + sh[3] + q.w;
      if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
      st4(out + off, o);
    }
  }
  // per-channel bookkeeping: the thread of CTA 0 that owns row 0 of a quad writes its four channels
  if (tc::cluster_ctarank() == 0 && p.row0 == 0) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      save_mean[c + j] = mean[j];
      save_invstd[c + j] = rsqrtf(var[j] + eps);
      if (running_mean != nullptr) {
        const float unbiased = M > 1 ? var[j] * ((float)M / (float)(M - 1)) : var[j];
        running_mean[c + j] = (1.f - momentum) * running_mean[c + j] + momentum * mean[j];
        running_var[c + j] = (1.f - momentum) * running_var[c + j] + momentum * unbiased;
      }
    }
  }
  if (tc::cluster_ctarank() == 0) {
    if (threadIdx.x == 0 && num_batches_tracked != nullptr) *num_batches_tracked += 1;
    for (int i = threadIdx.x; i < zero_count; i += kClThreads) zero_after[i] = 0.f;
  }
}

template <typename K, typename... Args>
static int launch_cluster_bn(K kernel, size_t smem, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(kClCtas);
  cfg.blockDim = dim3(kClThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = kClCtas;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, args...);
  if (e != cudaSuccess) return static_cast<int>(e);
  return static_cast<int>(cudaPeekAtLastError());
}

static bool cluster_bn_ok(int M, int C) {
  const int cq = C / 4;
  // a quad's threads inside a CTA are tid, tid + cq, ...: needs cq to divide the block size
  return C % 4 == 0 && cq >= 1 && M >= 1 && (long long)M * C <= kClMaxElems && cq <= kClThreads && kClThreads % cq == 0;
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
  return static_cast<int>(launch_pdl(bn_stats_kernel, dim3(gx, chunks), dim3(256), 0, stream, x, sums, M, C));
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
  return static_cast<int>(launch_pdl(bn_fwd_apply_kernel, dim3(gx, chunks), dim3(256), 0, stream, x, sums, gamma, beta,
                                     residual, out, running_mean, running_var, num_batches_tracked, save_mean,
                                     save_invstd, zero_after, zero_count, M, C, momentum, eps, relu, training));
}

B200_API int b200_bn_bwd_reduce(const float* dy, const float* out, const float* x, const float* save_mean,
                                const float* save_invstd, float* sums, int M, int C, int relu,
                                cudaStream_t stream) {
  if (C % 4 != 0) return static_cast<int>(cudaErrorInvalidValue);
  const int chunks = ceil_div(C / 4, 256);
  const int rows_per_block = max(1, 256 / min(C / 4, 256));
  const int gx = grid_for((size_t)M, rows_per_block * 2, max(1, 4 * kNumSMs / chunks));   // <= 2 rows per thread
  return static_cast<int>(launch_pdl(bn_bwd_reduce_kernel, dim3(gx, chunks), dim3(256), 0, stream, dy, out, x, save_mean,
                                     save_invstd, sums, M, C, relu));
}

B200_API int b200_bn_bwd_apply(const float* dy, const float* out, const float* x, const float* gamma,
                               const float* save_mean, const float* save_invstd, const float* sums, float* dx,
                               float* dres, float* dgamma, float* dbeta, int accumulate_param_grads,
                               float* zero_after, int zero_count, int M, int C, int relu, cudaStream_t stream) {
  if (C % 4 != 0) return static_cast<int>(cudaErrorInvalidValue);
  const int grid = grid_for((size_t)M * C / 4, 256 * 2, 4 * kNumSMs);
  return static_cast<int>(launch_pdl(bn_bwd_apply_kernel, dim3(grid), dim3(256), 0, stream, dy, out, x, gamma, save_mean,
                                     save_invstd, sums, dx, dres, dgamma, dbeta, accumulate_param_grads, zero_after,
                                     zero_count, M, C, relu));
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

/* One-launch BatchNorm backward / forward for tensors of at most 131 072 elements (deterministic cluster reduction).
 * Return cudaErrorNotSupported (801) when the shape does not qualify: callers fall back to the streaming kernels. */
B200_API int b200_bn_bwd_cluster(const float* dy, const float* out, const float* x, const float* gamma,
                                 const float* save_mean, const float* save_invstd, float* dx, float* dres, float* dgamma,
                                 float* dbeta, int accumulate_param_grads, float* zero_after, int zero_count, int M, int C,
                                 int relu, cudaStream_t stream) {
  if (!cluster_bn_ok(M, C)) return static_cast<int>(cudaErrorNotSupported);
  const size_t smem = ((size_t)kClThreads * 8 + 2 * (size_t)(C / 4) * 8) * sizeof(float);
  return launch_cluster_bn(bn_bwd_cluster_kernel, smem, stream, dy, out, x, gamma, save_mean, save_invstd, dx, dres, dgamma,
                           dbeta, accumulate_param_grads, zero_after, zero_count, M, C, relu);
}

B200_API int b200_bn_fwd_cluster(const float* x, const float* gamma, const float* beta, const float* residual, float* out,
                                 float* running_mean, float* running_var, long long* num_batches_tracked, float* save_mean,
                                 float* save_invstd, float* zero_after, int zero_count, int M, int C, float momentum,
                                 float eps, int relu, cudaStream_t stream) {
  if (!cluster_bn_ok(M, C)) return static_cast<int>(cudaErrorNotSupported);
  const size_t smem = ((size_t)kClThreads * 4 + 2 * (size_t)(C / 4) * 4) * sizeof(float);
  return launch_cluster_bn(bn_fwd_cluster_kernel, smem, stream, x, gamma, beta, residual, out, running_mean, running_var,
                           num_batches_tracked, save_mean, save_invstd, zero_after, zero_count, M, C, momentum, eps, relu);
}
