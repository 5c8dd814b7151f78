// Fused multi-tensor optimizers: ONE launch updates every parameter tensor of a param group.
//
// Replaces torch's foreach SGD (`_foreach_add_(params, grads, alpha=-lr)`, SURVEY.md K20,
// reached from optimizer.step() at /root/reference/dpp.py:53) and fused Adam.
//
// A device-resident table describes the tensors (pointers + element counts) and a second
// table maps each CTA-sized chunk to (tensor, offset), so the launch geometry is fixed and
// the kernel is CUDA-graph friendly.  Hyper-parameters are read from device memory (`hyper`)
// so a captured graph follows learning-rate changes without re-capture.
#include <cuda_bf16.h>

#include "common.cuh"
#include "b200ddp.h"

namespace b200 {

constexpr int kChunk = 4096;   // elements per CTA (256 threads x 4 float4)

struct __align__(16) TensorRec {
  float* p;
  float* g;
  float* s0;      // SGD: momentum buffer   Adam: exp_avg
  float* s1;      // Adam: exp_avg_sq
  long long n;
  long long dtype;   // 0: p/g fp32   1: p/g bf16 (optimizer state stays fp32)
};

struct ChunkRec { int tensor; int chunk; };

// hyper layout (floats): SGD  [lr, momentum, dampening, weight_decay, nesterov, maximize, first_step, grad_scale]
//                        Adam [lr, beta1, beta2, eps, weight_decay, decoupled, maximize, grad_scale, step]
__global__ void __launch_bounds__(256) sgd_kernel(const TensorRec* __restrict__ recs, const ChunkRec* __restrict__ chunks,
                                                  const float* __restrict__ hyper) {
  pdl_launch_dependents();
  pdl_wait();
  const ChunkRec ch = chunks[blockIdx.x];
  const TensorRec r = recs[ch.tensor];
  const float lr = hyper[0], mom = hyper[1], damp = hyper[2], wd = hyper[3];
  const bool nesterov = hyper[4] != 0.f, maximize = hyper[5] != 0.f, first = hyper[6] != 0.f;
  const float gscale = hyper[7];
  const long long base = (long long)ch.chunk * kChunk;
  const long long end = min(r.n, base + kChunk);
  const bool vec = ((reinterpret_cast<uintptr_t>(r.p) | reinterpret_cast<uintptr_t>(r.g) |
                     reinterpret_cast<uintptr_t>(r.s0)) & 15u) == 0;
  auto upd = [&](float p, float g, float& buf) -> float {
    g *= gscale;
    if (maximize) g = -g;
    if (wd != 0.f) g = fmaf(wd, p, g);
    if (mom != 0.f) {
      buf = first ? g : fmaf(mom, buf, (1.f - damp) * g);
      g = nesterov ? fmaf(mom, buf, g) : buf;
    }
    return fmaf(-lr, g, p);
  };
  if (r.dtype == 1 && vec && (end - base) % 8 == 0 && base % 8 == 0) {
    // bf16 parameters / gradients, 8 per thread (16-byte loads and stores), fp32 momentum
    __nv_bfloat16* pb = reinterpret_cast<__nv_bfloat16*>(r.p);
    const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(r.g);
    for (long long i = base + threadIdx.x * 8; i < end; i += 256 * 8) {
      uint4 pv = *reinterpret_cast<const uint4*>(pb + i);
      const uint4 gv = *reinterpret_cast<const uint4*>(gb + i);
      float b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (r.s0) {
        const float4 b0 = *reinterpret_cast<const float4*>(r.s0 + i), b1 = *reinterpret_cast<const float4*>(r.s0 + i + 4);
        b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
      }
      __nv_bfloat162* p2 = reinterpret_cast<__nv_bfloat162*>(&pv);
      const __nv_bfloat162* g2 = reinterpret_cast<const __nv_bfloat162*>(&gv);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 pf = __bfloat1622float2(p2[e]), gf = __bfloat1622float2(g2[e]);
        p2[e] = __floats2bfloat162_rn(upd(pf.x, gf.x, b[2 * e]), upd(pf.y, gf.y, b[2 * e + 1]));
      }
      *reinterpret_cast<uint4*>(pb + i) = pv;
      if (r.s0) {
        *reinterpret_cast<float4*>(r.s0 + i) = make_float4(b[0], b[1], b[2], b[3]);
        *reinterpret_cast<float4*>(r.s0 + i + 4) = make_float4(b[4], b[5], b[6], b[7]);
      }
    }
  } else if (r.dtype == 1) {
    __nv_bfloat16* pb = reinterpret_cast<__nv_bfloat16*>(r.p);
    const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(r.g);
    for (long long i = base + threadIdx.x; i < end; i += 256) {
      float b = r.s0 ? r.s0[i] : 0.f;
      pb[i] = __float2bfloat16_rn(upd(__bfloat162float(pb[i]), __bfloat162float(gb[i]), b));
      if (r.s0) r.s0[i] = b;
    }
  } else if (vec && (end - base) % 4 == 0) {
    for (long long i = base + threadIdx.x * 4; i < end; i += 256 * 4) {
      float4 p = *reinterpret_cast<const float4*>(r.p + i);
      const float4 g = *reinterpret_cast<const float4*>(r.g + i);
      float4 b = make_float4(0, 0, 0, 0);
      if (r.s0) b = *reinterpret_cast<const float4*>(r.s0 + i);
      p.x = upd(p.x, g.x, b.x); p.y = upd(p.y, g.y, b.y); p.z = upd(p.z, g.z, b.z); p.w = upd(p.w, g.w, b.w);
      *reinterpret_cast<float4*>(r.p + i) = p;
      if (r.s0) *reinterpret_cast<float4*>(r.s0 + i) = b;
    }
  } else {
    for (long long i = base + threadIdx.x; i < end; i += 256) {
      float b = r.s0 ? r.s0[i] : 0.f;
      r.p[i] = upd(r.p[i], r.g[i], b);
      if (r.s0) r.s0[i] = b;
    }
  }
}

__global__ void __launch_bounds__(256) adam_kernel(const TensorRec* __restrict__ recs, const ChunkRec* __restrict__ chunks,
                                                   float* __restrict__ hyper, float* const* __restrict__ step_ptrs,
                                                   int num_tensors) {
  const ChunkRec ch = chunks[blockIdx.x];
  const TensorRec r = recs[ch.tensor];
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4];
  const bool decoupled = hyper[5] != 0.f, maximize = hyper[6] != 0.f;
  const float gscale = hyper[7];
  const float step = hyper[8] + 1.f;           // hyper[8] is bumped by adam_step_kernel afterwards
  const float bc1 = 1.f - __powf(b1, step), bc2 = 1.f - __powf(b2, step);
  const float step_size = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
  const long long base = (long long)ch.chunk * kChunk;
  const long long end = min(r.n, base + kChunk);
  const bool bf = r.dtype == 1;
  __nv_bfloat16* pb = reinterpret_cast<__nv_bfloat16*>(r.p);
  const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(r.g);
  for (long long i = base + threadIdx.x; i < end; i += 256) {
    float p = bf ? __bfloat162float(pb[i]) : r.p[i], g = (bf ? __bfloat162float(gb[i]) : r.g[i]) * gscale;
    if (maximize) g = -g;
    if (wd != 0.f) {
      if (decoupled) p *= (1.f - lr * wd);
      else g = fmaf(wd, p, g);
    }
    const float m = fmaf(b1, r.s0[i], (1.f - b1) * g);
    const float v = fmaf(b2, r.s1[i], (1.f - b2) * g * g);
    r.s0[i] = m;
    r.s1[i] = v;
    const float pn = p - step_size * m / (sqrtf(v) * inv_sqrt_bc2 + eps);
    if (bf) pb[i] = __float2bfloat16_rn(pn);
    else r.p[i] = pn;
  }
}

// bump the shared step counter and every per-tensor `step` scalar (torch state layout)
__global__ void adam_step_kernel(float* hyper, float* const* step_ptrs, int num_tensors) {
  const float s = hyper[8] + 1.f;
  for (int i = threadIdx.x; i < num_tensors; i += blockDim.x)
    if (step_ptrs[i]) *step_ptrs[i] = s;
  __syncthreads();
  if (threadIdx.x == 0) hyper[8] = s;
}

}  // namespace b200

using namespace b200;

B200_API int b200_optim_chunk_elems() { return kChunk; }
B200_API int b200_optim_rec_bytes() { return (int)sizeof(TensorRec); }

B200_API int b200_fused_sgd(const void* recs, const void* chunks, int num_chunks, const float* hyper,
                            cudaStream_t stream) {
  if (num_chunks <= 0) return 0;
  return static_cast<int>(launch_pdl(sgd_kernel, dim3(num_chunks), dim3(256), 0, stream, static_cast<const TensorRec*>(recs),
                                     static_cast<const ChunkRec*>(chunks), hyper));
}

B200_API int b200_fused_adam(const void* recs, const void* chunks, int num_chunks, float* hyper,
                             float* const* step_ptrs, int num_tensors, cudaStream_t stream) {
  if (num_chunks <= 0) return 0;
  adam_kernel<<<num_chunks, 256, 0, stream>>>(static_cast<const TensorRec*>(recs), static_cast<const ChunkRec*>(chunks), hyper,
                                              step_ptrs, num_tensors);
  adam_step_kernel<<<1, 256, 0, stream>>>(hyper, step_ptrs, num_tensors);
  B200_LAUNCH_CHECK();
}
