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

// ---------------------------------------------------------------------------------------------- programmatic dependent launch
// The kernels of a local training step form one long dependency chain of short launches (10-40 us each, replayed from a
// CUDA graph).  Launched with the programmatic-stream-serialization attribute, kernel N+1 becomes resident and runs its
// prologue (barrier init, TMEM allocation, tensor-map prefetch, index arithmetic) while kernel N drains, and blocks in
// griddepcontrol.wait until N has completed and its writes are visible.  Every kernel launched through launch_pdl() MUST
// execute grid_dep_wait() before its first global access; grid_dep_launch() at the top lets the successor in as soon as
// the last CTA of this grid has started.  Measured on B200 (CNN / MLP / ResNet-18 step graphs): no gain - the chains are
// bound by the kernels' own critical paths, not by launch gaps - so the attribute is OFF by default (B2_PDL=1 turns it on).
bool pdl_enabled();

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
__device__ __forceinline__ void grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_sync() {
  grid_dep_launch();
  grid_dep_wait();
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

#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
static inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  (void)cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);  // errors surface through B2_LAUNCH_CHECK (cudaGetLastError)
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
