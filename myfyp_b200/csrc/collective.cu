// A2: two-shot FedAvg across the GPUs of a box as ONE kernel per GPU - reduce-scatter + all-gather over peer memory,
// fused with the sample-weighted sum, synchronised by flags in peer-mapped memory (no NCCL, no host round trip).
//
//   rank r of G:   phase 0  announce "my input is complete" to every peer (stream order on r guarantees it)
//                  phase 1  for the coordinate shard r:  out_r[shard_r] = sum_g w_g * src_g[shard_r]      (G-way peer pull)
//                  signal   last CTA to finish phase 1 -> reduced[r] = epoch on every peer  (st.release.sys)
//                  phase 2  for every peer g: wait reduced[g] == epoch (ld.acquire.sys), pull out_g[shard_g] -> out_r
//                  phase 3  tell every peer "I have pulled your shard"; leave only when every peer has pulled mine,
//                           so the caller may overwrite src / out as soon as the kernel is done.
// NVLink bytes per GPU: (G-1)/G * 4N in phase 1 + (G-1)/G * 4N in phase 2, against (G-1) * 4N for the one-shot kernel.
//
// The grid is persistent (<= resident CTAs): CTAs that spin in phase 2 must not keep phase-1 CTAs from being scheduled.
// The kernel reads its arguments from a device array indexed by blockIdx.y, so the SAME kernel also runs all G ranks in
// one launch on one GPU (gridDim.y = G, every pointer local) - that is how it is tested on a single-GPU box, where
// separate launches that wait on each other would not be guaranteed to be co-resident.
//
// Reference path: there is none (the reference exchanges pickles pairwise and averages with NumPy on the host,
// fedavg.py:61-69); the competitor is an NCCL all-reduce of pre-weighted tensors.
#include "../../include/myfyp_b200.h"
#include "common.cuh"

namespace b2 {

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// peer data written during this kernel: must not come from a stale L1 line
__device__ __forceinline__ float4 ld_volatile4(const float4* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void wait_flag(const unsigned int* flag, unsigned int epoch) {
  if (threadIdx.x == 0) {
    while (ld_acquire_sys(flag) < epoch) __nanosleep(64);
  }
  __syncthreads();
}

struct TwoShotWeights {
  float w[B2_TWOSHOT_MAX_RANKS];
};

// The address table is static per communicator (uploaded once); what changes per call - the weights and the epoch -
// travels in kernel-argument space, so a call is exactly one launch with no host-to-device copy in front of it.
__global__ void __launch_bounds__(512) fedavg_twoshot_kernel(const b2_twoshot_args* __restrict__ all_args, const __grid_constant__ TwoShotWeights wts,
                                                             unsigned int epoch) {
  const b2_twoshot_args& a = all_args[blockIdx.y];
  const int G = a.world, r = a.rank;
  unsigned int* my_flags = static_cast<unsigned int*>(a.flags[r]);  // [3][B2_TWOSHOT_MAX_RANKS] + counters
  const int64_t n4 = a.n / 4;
  const int64_t shard = (n4 + G - 1) / G;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  __shared__ int s_last;

  // ---- phase 0: my inputs are complete (this kernel is stream-ordered after whatever produced them)
  if (blockIdx.x == 0 && threadIdx.x < G) st_release_sys(static_cast<unsigned int*>(a.flags[threadIdx.x]) + 0 * B2_TWOSHOT_MAX_RANKS + r, epoch);
  for (int g = 0; g < G; ++g) wait_flag(my_flags + 0 * B2_TWOSHOT_MAX_RANKS + g, epoch);

  // ---- phase 1: reduce my shard from every rank's input
  {
    const int64_t lo = (int64_t)r * shard, hi = (lo + shard < n4) ? lo + shard : n4;
    float4* out = static_cast<float4*>(a.outs[r]);
    int64_t i = lo + tid0;
    // two output vectors per trip: 2*G independent 128-bit peer loads in flight per thread (NVLink latency ~2 us)
    for (; i + stride < hi; i += 2 * stride) {
      float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
#pragma unroll 4
      for (int g = 0; g < G; ++g) {
        const float4 v0 = ld_stream(static_cast<const float4*>(a.srcs[g]) + i);
        const float4 v1 = ld_stream(static_cast<const float4*>(a.srcs[g]) + i + stride);
        const float w = wts.w[g];
        a0.x = fmaf(w, v0.x, a0.x); a0.y = fmaf(w, v0.y, a0.y); a0.z = fmaf(w, v0.z, a0.z); a0.w = fmaf(w, v0.w, a0.w);
        a1.x = fmaf(w, v1.x, a1.x); a1.y = fmaf(w, v1.y, a1.y); a1.z = fmaf(w, v1.z, a1.z); a1.w = fmaf(w, v1.w, a1.w);
      }
      out[i] = a0;
      out[i + stride] = a1;
    }
    for (; i < hi; i += stride) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
      for (int g = 0; g < G; ++g) {
        const float4 v = ld_stream(static_cast<const float4*>(a.srcs[g]) + i);
        const float w = wts.w[g];
        acc.x = fmaf(w, v.x, acc.x); acc.y = fmaf(w, v.y, acc.y); acc.z = fmaf(w, v.z, acc.z); acc.w = fmaf(w, v.w, acc.w);
      }
      out[i] = acc;
    }
  }
  // ---- signal: the last CTA of this rank publishes "shard r is reduced"
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* counter = my_flags + 3 * B2_TWOSHOT_MAX_RANKS;
    s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    if (threadIdx.x == 0) my_flags[3 * B2_TWOSHOT_MAX_RANKS] = 0u;  // counter ready for the next call
    __threadfence_system();
    if (threadIdx.x < G) st_release_sys(static_cast<unsigned int*>(a.flags[threadIdx.x]) + 1 * B2_TWOSHOT_MAX_RANKS + r, epoch);
  }

  // ---- phase 2: gather the other shards (start with my right-hand neighbour to spread the load over the links)
  float4* out = static_cast<float4*>(a.outs[r]);
  for (int d = 1; d < G; ++d) {
    const int g = (r + d) % G;
    wait_flag(my_flags + 1 * B2_TWOSHOT_MAX_RANKS + g, epoch);
    const int64_t lo = (int64_t)g * shard, hi = (lo + shard < n4) ? lo + shard : n4;
    const float4* src = static_cast<const float4*>(a.outs[g]);
    int64_t i = lo + tid0;
    for (; i + 3 * stride < hi; i += 4 * stride) {  // four independent peer loads in flight per thread
      const float4 v0 = ld_volatile4(src + i), v1 = ld_volatile4(src + i + stride), v2 = ld_volatile4(src + i + 2 * stride), v3 = ld_volatile4(src + i + 3 * stride);
      out[i] = v0;
      out[i + stride] = v1;
      out[i + 2 * stride] = v2;
      out[i + 3 * stride] = v3;
    }
    for (; i < hi; i += stride) out[i] = ld_volatile4(src + i);
  }

  // ---- phase 3: everybody has pulled my shard before I return (the caller may then reuse src / out)
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* counter = my_flags + 3 * B2_TWOSHOT_MAX_RANKS + 1;
    s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    if (threadIdx.x == 0) my_flags[3 * B2_TWOSHOT_MAX_RANKS + 1] = 0u;
    __threadfence_system();
    if (threadIdx.x < G) st_release_sys(static_cast<unsigned int*>(a.flags[threadIdx.x]) + 2 * B2_TWOSHOT_MAX_RANKS + r, epoch);
    for (int g = 0; g < G; ++g) wait_flag(my_flags + 2 * B2_TWOSHOT_MAX_RANKS + g, epoch);
  }
}

}  // namespace b2

extern "C" {

int64_t b2_twoshot_flag_bytes(void) { return (3 * B2_TWOSHOT_MAX_RANKS + 8) * (int64_t)sizeof(unsigned int); }

int b2_fedavg_twoshot(const b2_twoshot_args* dev_args, int ranks_in_launch, int64_t n, const float* weights, int world, unsigned int epoch, int max_ctas,
                      void* stream) {
  if (world < 1 || world > B2_TWOSHOT_MAX_RANKS || weights == nullptr) {
    b2::set_error("twoshot: world=%d outside [1,%d] or NULL weights", world, B2_TWOSHOT_MAX_RANKS);
    return 1;
  }
  b2::TwoShotWeights wts;
  for (int g = 0; g < B2_TWOSHOT_MAX_RANKS; ++g) wts.w[g] = g < world ? weights[g] : 0.f;
  if (ranks_in_launch < 1 || ranks_in_launch > B2_TWOSHOT_MAX_RANKS) {
    b2::set_error("twoshot: %d ranks in one launch (max %d)", ranks_in_launch, B2_TWOSHOT_MAX_RANKS);
    return 1;
  }
  if (n <= 0 || (n & 3)) {
    b2::set_error("twoshot: n=%lld must be a positive multiple of 4", (long long)n);
    return 1;
  }
  // persistent grid: every CTA of every rank in this launch must be resident at once (they wait on each other)
  int per_rank = b2::sm_count() / ranks_in_launch;
  if (ranks_in_launch == 1) per_rank = b2::sm_count();
  if (max_ctas > 0 && per_rank > max_ctas) per_rank = max_ctas;
  const int64_t need = (n / 4 + 511) / 512;
  if (need < per_rank) per_rank = (int)need;
  if (per_rank < 1) per_rank = 1;
  dim3 grid(per_rank, ranks_in_launch);
  b2::fedavg_twoshot_kernel<<<grid, 512, 0, static_cast<cudaStream_t>(stream)>>>(dev_args, wts, epoch);
  B2_LAUNCH_CHECK("fedavg_twoshot_kernel");
  return 0;
}

}  // extern "C"
