// Shared device/host helpers for the sm_100a kernel library.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#define B200_API extern "C" __attribute__((visibility("default")))

// Every launcher returns cudaError_t cast to int (0 == success) so ctypes callers can check.
#define B200_LAUNCH_CHECK() return static_cast<int>(cudaPeekAtLastError())

namespace b200 {

constexpr int kNumSMs = 148;

__host__ __device__ __forceinline__ int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ __forceinline__ long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

// system-scope release / acquire used by the cross-GPU flag protocol
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Spin until *flag >= epoch (wrap-safe signed difference) or `timeout_ns` of wall time (%globaltimer) has passed.
__device__ __forceinline__ bool wait_flag_ns(const uint32_t* flag, uint32_t epoch, unsigned long long timeout_ns) {
  if ((int32_t)(ld_acquire_sys(flag) - epoch) >= 0) return true;
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  unsigned spins = 0;
  while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
    if ((++spins & 63u) == 0) {
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > timeout_ns) return false;
    }
  }
  return true;
}

// NVLink multicast (NVLS) vector accesses: ld_reduce returns the SUM over all replicas (added in the switch, fp32
// accumulation), st writes every replica.
__device__ __forceinline__ uint4 multimem_ld_reduce_f32x4(const void* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f32x4(void* p, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_bf16x8(void* p, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};"
               ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- programmatic dependent launch (PDL) --------------------------------------------------------------------
// The training step is a chain of ~140 short kernels; between two dependent kernels the GPU otherwise idles for the
// launch latency plus the next kernel's prologue (barrier init, TMEM allocation, descriptor prefetch).  Kernels
// launched through launch_pdl() may become resident while their predecessor in the stream is still running:
//   * pdl_launch_dependents() at the top of a kernel lets the NEXT kernel's CTAs be scheduled as soon as all of
//     this kernel's CTAs have started;
//   * pdl_wait() blocks until the PREVIOUS kernel has completed and its memory is visible.  Every kernel launched
//     with the attribute must execute it before its first global-memory access (reads AND writes): that also keeps
//     completion transitive along the chain.
// Both are no-ops for a kernel launched the ordinary way.  CUDA graphs capture the attribute as a programmatic edge.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();      // defined in elementwise.cu (B200DDP_PDL=0 or b200_set_pdl(0) turns it off)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace b200
