// Shared helpers for the sm_100a kernels of the B200 data-parallel framework.
// Everything here is header-only device/host glue: error reporting for the C ABI,
// memory-ordering PTX wrappers for the cross-GPU flag protocol, Philox RNG, and
// warp reductions.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

// ---------------------------------------------------------------------------
// Error reporting for the C ABI. Every exported function returns 0 on success
// or a non-zero code; b200_last_error() gives the text.
// ---------------------------------------------------------------------------
extern "C" void b200_set_error(const char* fmt, ...);

#define B200_CUDA_OK(expr)                                                     \
  do {                                                                         \
    cudaError_t _e = (expr);                                                   \
    if (_e != cudaSuccess) {                                                   \
      b200_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr,        \
                     cudaGetErrorString(_e));                                  \
      return (int)_e;                                                          \
    }                                                                          \
  } while (0)

#define B200_REQUIRE(cond, ...)                                                \
  do {                                                                         \
    if (!(cond)) {                                                             \
      b200_set_error(__VA_ARGS__);                                             \
      return -1;                                                               \
    }                                                                          \
  } while (0)

#define B200_LAUNCH_OK()                                                       \
  do {                                                                         \
    cudaError_t _e = cudaGetLastError();                                       \
    if (_e != cudaSuccess) {                                                   \
      b200_set_error("%s:%d: kernel launch failed: %s", __FILE__, __LINE__,    \
                     cudaGetErrorString(_e));                                  \
      return (int)_e;                                                          \
    }                                                                          \
  } while (0)

#ifdef __CUDACC__

// ---------------------------------------------------------------------------
// Programmatic dependent launch (PDL). A kernel launched with b200_launch_pdl may start
// (its CTAs become resident) while its predecessor in the stream is still running; it must
// call pdl_wait() before touching anything the predecessor writes. pdl_trigger() lets the
// NEXT kernel start launching. Used by the two kernels of the fused step so the ~2 us
// launch latency at each kernel boundary overlaps the previous kernel's execution.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void pdl_trigger() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

template <typename... KArgs, typename... Args>
inline cudaError_t b200_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                   cudaStream_t stream, bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------------------
// Warp helpers
// ---------------------------------------------------------------------------
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

// ---------------------------------------------------------------------------
// System-scope memory ordering for flags that live in peer-mapped memory.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add_u32(uint32_t* p, uint32_t v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_sys() {
  asm volatile("fence.acq_rel.sys;" ::: "memory");
}
// 16-byte loads/stores that never allocate in (possibly stale) L1: used on
// buffers that a peer GPU writes or reads during the same kernel.
__device__ __forceinline__ float4 ld_volatile_f4(const float4* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_f4(float4* p, float4 v) {
  asm volatile("st.volatile.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ uint4 ld_volatile_u4(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_u4(uint4* p, uint4 v) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// ---------------------------------------------------------------------------
// Philox4x32-10 counter-based RNG (Salmon et al. 2011). Keyed by (seed) and a
// 128-bit counter the kernels build from (step, sample, layer, element), so a
// dropout mask is a pure function of those and replays identically.
// ---------------------------------------------------------------------------
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2,
                                                       uint32_t c3, uint32_t k0, uint32_t k1,
                                                       uint32_t out[4]) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)M0 * c0;
    uint64_t p1 = (uint64_t)M1 * c2;
    uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
    uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    uint32_t n0 = hi1 ^ c1 ^ k0;
    uint32_t n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// Bernoulli(keep = 1-p) scaled mask value for one element: 0 or 1/(1-p).
__device__ __forceinline__ float dropout_mask_value(uint64_t seed, uint64_t step, uint32_t sample,
                                                    uint32_t layer_elem, float p) {
  uint32_t r[4];
  philox4x32_10((uint32_t)step, (uint32_t)(step >> 32), sample, layer_elem, (uint32_t)seed,
                (uint32_t)(seed >> 32), r);
  // 24-bit uniform in [0,1)
  float u = (float)(r[0] >> 8) * (1.0f / 16777216.0f);
  return (u >= p) ? 1.0f / (1.0f - p) : 0.0f;
}

#endif  // __CUDACC__
