// Shared helpers for the sm_100a kernels of libmyfyp_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>

#define B2_MAX_K 64          // models per K-way launch (pointer table lives in kernel-argument space)
#define B2_SMS_FALLBACK 148  // B200

namespace b2 {

// ---------------------------------------------------------------------------------------------- errors
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);  // records the message, returns non-zero

#define B2_CUDA(expr)                                             \
  do {                                                            \
    cudaError_t _e = (expr);                                      \
    if (_e != cudaSuccess) return b2::cuda_fail(_e, #expr);       \
  } while (0)

#define B2_LAUNCH_CHECK(name)                                     \
  do {                                                            \
    cudaError_t _e = cudaGetLastError();                          \
    if (_e != cudaSuccess) return b2::cuda_fail(_e, name);        \
  } while (0)

int sm_count();  // SMs of the current device (cached per device)

// ---------------------------------------------------------------------------------------------- pointer tables
struct PtrTable {
  const float* p[B2_MAX_K];
};
struct WeightTable {
  float w[B2_MAX_K];
};

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---------------------------------------------------------------------------------------------- device helpers
#ifdef __CUDACC__
// streaming 128-bit load: data is read once, keep it out of L1.  Works on local HBM and on peer-mapped
// (NVLink) addresses alike - the aperture decides the route.
__device__ __forceinline__ float4 ld_stream(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float ld_stream1(const float* p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream(float4* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
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
#endif

// grid for a streaming elementwise kernel: enough CTAs to fill the chip several times over, capped
static inline int stream_grid(int64_t work_items, int threads, int max_ctas_per_sm = 8) {
  int64_t need = (work_items + threads - 1) / threads;
  int64_t cap = (int64_t)sm_count() * max_ctas_per_sm;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

}  // namespace b2
