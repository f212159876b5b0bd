// A6 + A7: Krum.  Not in the reference; oracle = fp64 direct pairwise distances.
//
// A6  Gram matrix of the centred models on the 5th-gen tensor cores.
//       G[i][j] = sum_c (theta_i[c]-theta_0[c]) * (theta_j[c]-theta_0[c]),   d_ij = G_ii + G_jj - 2 G_ij
//     The K models are K separately allocated slabs (local HBM or peer GPUs), so the operand tile is assembled by
//     the producer warps: 128-bit loads (the loads are the NVLink transfer), subtract the reference model, split
//     into tf32 hi / lo parts, store into the canonical SWIZZLE_128B K-major layout.  One tile serves as BOTH MMA
//     operands (A = B).  One elected thread issues tcgen05.mma kind::tf32 (M=128, N=Kpad, K=8); 3 MMAs per k-step
//     (hi*hi + hi*lo + lo*hi) give fp32-class accuracy.  The accumulator lives in TMEM across ALL coordinate
//     chunks of the CTA (split-K over coordinates across CTAs), is read once with tcgen05.ld, spilled to the
//     workspace, and the last CTA to finish reduces the partial Grams in a fixed order.
//     The MMA pipeline accumulates in fp32 with a small systematic drift per accumulation step, so one chain over
//     thousands of k-steps loses a few 1e-4 relative on the (all-positive) diagonal.  The stages therefore rotate over
//     8 independent TMEM accumulators (8 x 64 columns = the SM's whole tensor memory); the epilogue adds them in
//     registers.  Chains are 8x shorter, the drift 8x smaller.
//     smem tile per stage: 64 rows x 128 coords fp32 (4 swizzle atoms) for hi and for lo = 64 KB; 3 stages.
//     The MMA reads 128 A-rows: rows 64..127 alias the following atom - they only produce accumulator rows
//     64..127, which nobody reads.
// A7  scores + ranking in one tiny CTA.
#include <mutex>

#include "../../include/myfyp_b200.h"
#include "common.cuh"
#include "tcgen05.cuh"

namespace b2 {

constexpr int GR_ROWS = 64;                         // max models
constexpr int GR_CC = 128;                          // coordinates per stage
constexpr int GR_ATOMS = GR_CC / 32;                // 32 fp32 = 128 B per swizzle row
constexpr int GR_ATOM_BYTES = GR_ROWS * 128;        // 8 KB
constexpr int GR_TILE_BYTES = GR_ATOMS * GR_ATOM_BYTES;  // 32 KB (hi) ; lo follows
constexpr int GR_STAGE_BYTES = 2 * GR_TILE_BYTES;   // 64 KB
constexpr int GR_STAGES = 3;
constexpr int GR_PRODUCERS = 256;
constexpr int GR_THREADS = GR_PRODUCERS + 32;       // + the MMA / TMEM warp
constexpr int GR_SMEM = GR_STAGES * GR_STAGE_BYTES + GR_ATOM_BYTES /*over-read pad*/ + 1024 /*align*/;
constexpr int GR_ACCS = 8;                          // TMEM accumulators used round-robin (8 x 64 columns = all 512)
constexpr int GR_MAX_CTAS = 296;
constexpr int GR_HEADER = 256;

// UPT = float4 items per producer thread and stage = ceil(K / 8): a thread's items are the SAME 4 coordinates of UPT different
// models (item q = u * 256 + tid -> row q / 32, coordinate group q % 32 = tid % 32), so the reference chunk is loaded once.
template <int UPT>
__global__ void __launch_bounds__(GR_THREADS, 1) krum_gram_kernel(float* __restrict__ gram, const __grid_constant__ PtrTable srcs, int K, int64_t n, int passes,
                                                                   unsigned int* __restrict__ counter, float* __restrict__ partials) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[GR_STAGES], empty_bar[GR_STAGES], done_bar;
  __shared__ uint32_t tmem_base_s;
  __shared__ int is_last_s;

  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5;
  const int kpad = (K + 15) & ~15;  // MMA N (multiple of 16 for M=128)

  if (tid == 0) {
    for (int s = 0; s < GR_STAGES; ++s) {
      tc::mbar_init(&full_bar[s], GR_PRODUCERS / 32);  // one arrival per producer warp
      tc::mbar_init(&empty_bar[s], 1);
    }
    tc::mbar_init(&done_bar, 1);
    tc::mbar_fence_init();
  }
  if (warp == 8) tc::tmem_alloc<64 * GR_ACCS>(&tmem_base_s);
  // rows K..63 are never written by the producers: zero the whole tile area once
  for (int i = tid; i < (GR_STAGES * GR_STAGE_BYTES + GR_ATOM_BYTES) / 16; i += GR_THREADS) reinterpret_cast<float4*>(tiles)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  tc::fence_proxy_async();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = tmem_base_s;

  const int64_t n_chunks = (n + GR_CC - 1) / GR_CC;
  const int64_t my_chunks = (n_chunks > blockIdx.x) ? (n_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (warp < 8) {
    // ================================ producers: pull, centre, split, swizzle-store
    // The loads of stage it+1 are issued BEFORE stage it is converted and stored (two register buffers): with one stage of loads
    // in flight per CTA the kernel ran at the latency of a DRAM round trip per 128 coordinates (1.4 TB/s at K = 32).
    const float* ref = srcs.p[0];
    const int j = tid & 31, row0 = tid >> 5;
    auto issue = [&](int64_t it, float4 (&v)[UPT], float4& r) {
      const int64_t c = (blockIdx.x + it * gridDim.x) * (int64_t)GR_CC + 4 * j;
      const bool in = c < n;  // n % 4 == 0 is enforced by the host wrapper
      r = in ? __ldg(reinterpret_cast<const float4*>(ref + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int u = 0; u < UPT; ++u) {
        const int row = row0 + 8 * u;
        v[u] = (in && row < K) ? ld_stream(reinterpret_cast<const float4*>(srcs.p[row] + c)) : r;  // r - r = 0 for rows that do not exist
      }
    };
    auto convert = [&](int s, const float4 (&v)[UPT], const float4& r) {
      uint8_t* hi_tile = tiles + s * GR_STAGE_BYTES;
      uint8_t* lo_tile = hi_tile + GR_TILE_BYTES;
#pragma unroll
      for (int u = 0; u < UPT; ++u) {
        const int row = row0 + 8 * u;
        if (row < K) {
          const float d[4] = {v[u].x - r.x, v[u].y - r.y, v[u].z - r.z, v[u].w - r.w};
          float4 h, l;
          float* hp = &h.x;
          float* lp = &l.x;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            hp[e] = __uint_as_float(__float_as_uint(d[e]) & 0xffffe000u);  // tf32 keeps 10 mantissa bits
            lp[e] = d[e] - hp[e];
          }
          const uint32_t off = (uint32_t)(j >> 3) * GR_ATOM_BYTES + tc::sw128_offset((uint32_t)row, (uint32_t)(j & 7));
          *reinterpret_cast<float4*>(hi_tile + off) = h;
          *reinterpret_cast<float4*>(lo_tile + off) = l;
        }
      }
    };
    auto publish = [&](int64_t it, const float4 (&v)[UPT], const float4& r) {
      const int s = (int)(it % GR_STAGES);
      tc::mbar_wait(&empty_bar[s], (uint32_t)((it / GR_STAGES) & 1) ^ 1u);
      convert(s, v, r);
      // every lane publishes its stores to the async proxy, the warp meets, ONE lane arrives: 8 arrivals per stage instead of
      // 256 serialised updates of the same barrier word
      tc::fence_proxy_async();
      __syncwarp();
      if ((tid & 31) == 0) tc::mbar_arrive(&full_bar[s]);
    };
    float4 va[UPT], vb[UPT], ra, rb;
    if (my_chunks > 0) issue(0, va, ra);
    for (int64_t it = 0; it < my_chunks; it += 2) {
      if (it + 1 < my_chunks) issue(it + 1, vb, rb);
      publish(it, va, ra);
      if (it + 1 < my_chunks) {
        if (it + 2 < my_chunks) issue(it + 2, va, ra);
        publish(it + 1, vb, rb);
      }
    }
  } else if (tc::elect_one()) {
    // ================================ MMA issuer (one thread)
    const uint32_t idesc = tc::make_idesc(tc::FMT_TF32, 128, (uint32_t)kpad);
    for (int64_t it = 0; it < my_chunks; ++it) {
      const int s = (int)(it % GR_STAGES);
      const uint32_t ph = (uint32_t)((it / GR_STAGES) & 1);
      tc::mbar_wait(&full_bar[s], ph);
      tc::fence_after_sync();
      const uint32_t hi_addr = tc::smem_u32(tiles + s * GR_STAGE_BYTES);
      const uint32_t lo_addr = hi_addr + GR_TILE_BYTES;
      const uint32_t d_tmem = tmem_base + (uint32_t)(it % GR_ACCS) * 64u;  // this stage's accumulator
      uint32_t accumulate = it >= GR_ACCS ? 1u : 0u;                       // first visit overwrites
#pragma unroll
      for (int a = 0; a < GR_ATOMS; ++a) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {  // 8 tf32 = 32 bytes per k-step inside the 128-byte swizzle row
          const uint32_t o = (uint32_t)a * GR_ATOM_BYTES + (uint32_t)ks * 32u;
          const uint64_t dh = tc::smem_desc_sw128(hi_addr + o), dl = tc::smem_desc_sw128(lo_addr + o);
          tc::mma_tf32(d_tmem, dh, dh, idesc, accumulate);
          accumulate = 1;
          if (passes >= 3) {
            tc::mma_tf32(d_tmem, dh, dl, idesc, 1);
            tc::mma_tf32(d_tmem, dl, dh, idesc, 1);
          }
        }
      }
      tc::mma_commit(&empty_bar[s]);  // stage reusable once these MMAs have read it
    }
    tc::mma_commit(&done_bar);
  }

  // ================================ epilogue: TMEM -> workspace partial (rows < 64 live in warps 0,1)
  __syncwarp();
  float* mine = partials + (size_t)blockIdx.x * (GR_ROWS * GR_ROWS);
  if (warp < 2) {
    tc::mbar_wait(&done_bar, 0);
    tc::fence_after_sync();
    const int row = warp * 32 + (tid & 31);
#pragma unroll
    for (int cb = 0; cb < 2; ++cb) {
      float v[32];
#pragma unroll
      for (int e = 0; e < 32; ++e) v[e] = 0.f;
      if (cb * 32 < kpad) {
        const int used = (int)(my_chunks < GR_ACCS ? my_chunks : GR_ACCS);  // accumulators that were written at all
        for (int a = 0; a < used; ++a) {
          float t[32];
          tc::tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(a * 64 + cb * 32), t);
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] += t[e];
        }
      }
#pragma unroll
      for (int e = 0; e < 32; e += 4) *reinterpret_cast<float4*>(mine + row * GR_ROWS + cb * 32 + e) = make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]);
    }
    tc::fence_before_sync();
  }
  __threadfence();
  __syncthreads();
  if (warp == 8) tc::tmem_dealloc<64 * GR_ACCS>(tmem_base);
  if (tid == 0) {
    const unsigned int prev = atomicAdd(counter, 1u);
    is_last_s = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last_s) {
    // fixed summation order (deterministic), 16 bytes per load, four partials in flight per thread, and only the K x K corner is
    // read at all: the plain "one float, one partial at a time" loop cost 0.35 ms on its own (148 dependent L2 round trips x 14)
    __threadfence();
    for (int o4 = tid; o4 < GR_ROWS * GR_ROWS / 4; o4 += GR_THREADS) {
      const int i = (4 * o4) / GR_ROWS, j = (4 * o4) % GR_ROWS;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < K && j < K) {
        const float4* p = reinterpret_cast<const float4*>(partials) + o4;
        const size_t stride4 = GR_ROWS * GR_ROWS / 4;
        float4 a[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        unsigned int b = 0;
        for (; b + 4 <= gridDim.x; b += 4) {
          float4 t[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) t[u] = __ldcg(p + (size_t)(b + u) * stride4);
#pragma unroll
          for (int u = 0; u < 4; ++u) a[u].x += t[u].x, a[u].y += t[u].y, a[u].z += t[u].z, a[u].w += t[u].w;
        }
        for (; b < gridDim.x; ++b) {
          const float4 t = __ldcg(p + (size_t)b * stride4);
          a[0].x += t.x, a[0].y += t.y, a[0].z += t.z, a[0].w += t.w;
        }
        acc.x = (a[0].x + a[1].x) + (a[2].x + a[3].x);
        acc.y = (a[0].y + a[1].y) + (a[2].y + a[3].y);
        acc.z = (a[0].z + a[1].z) + (a[2].z + a[3].z);
        acc.w = (a[0].w + a[1].w) + (a[2].w + a[3].w);
        if (j + 1 >= K) acc.y = 0.f;
        if (j + 2 >= K) acc.z = 0.f;
        if (j + 3 >= K) acc.w = 0.f;
      }
      reinterpret_cast<float4*>(gram)[o4] = acc;
    }
    if (tid == 0) *counter = 0u;  // ready for the next call
  }
}

// ------------------------------------------------------------------------------------------------ A7
__global__ void __launch_bounds__(64) krum_select_kernel(const float* __restrict__ dist, int K, int f, float* __restrict__ scores, int* __restrict__ order) {
  __shared__ float s_score[B2_MAX_K];
  const int i = threadIdx.x;
  if (i < K) {
    float d[B2_MAX_K];
    int m = 0;
    int finite = 0;
    for (int j = 0; j < K; ++j) {
      if (j == i) continue;
      float x = dist[i * K + j];
      if (isnan(x)) x = INFINITY;  // a model with non-finite parameters is infinitely far from everybody: ranks stay a total order
      finite += isinf(x) ? 0 : 1;
      int p = m++;
      while (p > 0 && d[p - 1] > x) { d[p] = d[p - 1]; --p; }  // insertion sort (K <= 64)
      d[p] = x;
    }
    int closest = K - f - 2;
    if (closest < 1) closest = 1;
    if (closest > m) closest = m;
    // fewer finite neighbours than the rule asks for (more broken models than f allows): score over those there are, so that
    // healthy models still rank ahead of broken ones; a model without any finite neighbour scores +inf
    if (closest > finite) closest = finite;
    if (finite == 0 && K > 1) closest = 1;
    float sc = 0.f;
    for (int j = 0; j < closest; ++j) sc += d[j];
    s_score[i] = sc;
    scores[i] = sc;
  }
  __syncthreads();
  if (i < K) {
    const float me = s_score[i];
    int rank = 0;
    for (int j = 0; j < K; ++j) rank += (s_score[j] < me || (s_score[j] == me && j < i)) ? 1 : 0;
    order[rank] = i;
  }
}


static void launch_gram(int grid, cudaStream_t st, float* gram, const PtrTable& pt, int K, int64_t n, int passes, unsigned int* counter, float* partials) {
  if (K <= 8) krum_gram_kernel<1><<<grid, GR_THREADS, GR_SMEM, st>>>(gram, pt, K, n, passes, counter, partials);
  else if (K <= 16) krum_gram_kernel<2><<<grid, GR_THREADS, GR_SMEM, st>>>(gram, pt, K, n, passes, counter, partials);
  else if (K <= 32) krum_gram_kernel<4><<<grid, GR_THREADS, GR_SMEM, st>>>(gram, pt, K, n, passes, counter, partials);
  else krum_gram_kernel<8><<<grid, GR_THREADS, GR_SMEM, st>>>(gram, pt, K, n, passes, counter, partials);
}

}  // namespace b2

using namespace b2;

extern "C" {

int64_t b2_krum_gram_workspace_bytes(void) { return GR_HEADER + (int64_t)GR_MAX_CTAS * GR_ROWS * GR_ROWS * (int64_t)sizeof(float); }

int b2_krum_gram(float* gram, const void* const* srcs, int K, int64_t n, void* workspace, int passes, void* stream) {
  if (K < 1 || K > GR_ROWS) {
    set_error("krum_gram: K=%d outside [1,%d]", K, GR_ROWS);
    return 1;
  }
  if (n <= 0 || (n & 3)) {
    set_error("krum_gram: n=%lld must be a positive multiple of 4", (long long)n);
    return 1;
  }
  if (workspace == nullptr) {
    set_error("krum_gram: workspace is NULL");
    return 1;
  }
  PtrTable pt;
  for (int k = 0; k < B2_MAX_K; ++k) pt.p[k] = static_cast<const float*>(srcs[k < K ? k : 0]);
  for (int k = 0; k < K; ++k)
    if (!aligned16(srcs[k])) {
      set_error("krum_gram: slab %d is not 16-byte aligned", k);
      return 1;
    }
  static bool attr_set[64] = {false};
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    B2_CUDA(cudaFuncSetAttribute(krum_gram_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GR_SMEM));
    B2_CUDA(cudaFuncSetAttribute(krum_gram_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GR_SMEM));
    B2_CUDA(cudaFuncSetAttribute(krum_gram_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, GR_SMEM));
    B2_CUDA(cudaFuncSetAttribute(krum_gram_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, GR_SMEM));
    attr_set[dev] = true;
  }
  const int64_t chunks = (n + GR_CC - 1) / GR_CC;
  int grid = sm_count();
  if (grid > GR_MAX_CTAS) grid = GR_MAX_CTAS;
  if (chunks < grid) grid = (int)chunks;
  unsigned int* counter = static_cast<unsigned int*>(workspace);
  float* partials = reinterpret_cast<float*>(static_cast<char*>(workspace) + GR_HEADER);
  launch_gram(grid, static_cast<cudaStream_t>(stream), gram, pt, K, n, passes, counter, partials);
  B2_LAUNCH_CHECK("krum_gram_kernel");
  return 0;
}

// ------------------------------------------------------------------------------------------------ allocation-free variant
//
// b2_krum_gram_static: the same Gram kernel, for callers that must not allocate device memory.
//
// Who that is: a C/C++ host that runs the Krum scoring inside somebody else's process - an aggregation service thread
// next to a trainer whose caching allocator has already claimed the HBM, a plugin inside a runtime that owns every byte
// of device memory, a latency-critical control loop.  Such a caller has no torch allocator to borrow a workspace from,
// cudaMalloc may simply fail there, and cudaMalloc / cudaFree synchronise the whole device, which would stall the other
// nodes' training streams that share the GPU.  So the split-K scratch (completion counter + one partial Gram per CTA)
// lives in a file-scope __device__ array that is part of the library image:
//   * size: 256 B header + B2_STATIC_GRAM_CTAS (64) partials of 64 x 64 fp32 = 1 048 832 bytes, held by every GPU that
//     loads the library until the process exits;
//   * limits: K <= 64 models, n <= B2_STATIC_GRAM_MAX_N (2^26 coordinates = 256 MiB per model).  With at most 64 CTAs
//     the kernel cannot stream a larger model at a sensible fraction of the bandwidth; bigger problems belong to
//     b2_krum_gram with a caller-sized workspace.  Anything beyond the limits is rejected with an error, never clamped;
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress with that device current -
//     in host code the array's name is only a host-side handle, not a device pointer) and sets the kernel's
//     shared-memory attribute; make that first call before any stream capture begins;
//   * all calls share the one array and it keeps whatever the previous call left: the counter block is cleared on the
//     stream before every launch, the partials are fully overwritten by the CTAs that run.  Two calls that overlap on
//     the GPU (different streams, no dependency between them) race on the array - serialise them.
// Inputs and the 64 x 64 output are still the caller's.  No cudaMalloc / cudaFree / allocating library call is made.
#define B2_STATIC_GRAM_CTAS 64
#define B2_STATIC_GRAM_MAX_N (int64_t(1) << 26)
static __device__ __align__(256) unsigned char g_static_gram_scratch[GR_HEADER + B2_STATIC_GRAM_CTAS * GR_ROWS * GR_ROWS * sizeof(float)];

int64_t b2_krum_gram_static_max_n(void) { return B2_STATIC_GRAM_MAX_N; }

int b2_krum_gram_static(float* gram, const void* const* srcs, int K, int64_t n, int passes, void* stream) {
  if (K < 1 || K > GR_ROWS) {
    set_error("krum_gram_static: K=%d outside [1,%d]", K, GR_ROWS);
    return 1;
  }
  if (n <= 0 || (n & 3)) {
    set_error("krum_gram_static: n=%lld must be a positive multiple of 4", (long long)n);
    return 1;
  }
  if (n > B2_STATIC_GRAM_MAX_N) {
    set_error("krum_gram_static: n=%lld exceeds the built-in scratch limit of %lld coordinates; use b2_krum_gram with a workspace", (long long)n,
              (long long)B2_STATIC_GRAM_MAX_N);
    return 1;
  }
  if (gram == nullptr || srcs == nullptr) {
    set_error("krum_gram_static: NULL argument");
    return 1;
  }
  PtrTable pt;
  for (int k = 0; k < B2_MAX_K; ++k) pt.p[k] = static_cast<const float*>(srcs[k < K ? k : 0]);
  for (int k = 0; k < K; ++k)
    if (!aligned16(srcs[k])) {
      set_error("krum_gram_static: slab %d is not 16-byte aligned", k);
      return 1;
    }
  // once per device, with that device current: resolve the device address of the scratch symbol
  static void* scratch_ptr[64] = {nullptr};
  static std::mutex setup_mutex;
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error("krum_gram_static: device ordinal %d not supported", dev);
    return 1;
  }
  {
    std::lock_guard<std::mutex> lock(setup_mutex);
    if (scratch_ptr[dev] == nullptr) {
      void* p = nullptr;
      B2_CUDA(cudaGetSymbolAddress(&p, g_static_gram_scratch));
      B2_CUDA(cudaFuncSetAttribute(krum_gram_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GR_SMEM));
      B2_CUDA(cudaFuncSetAttribute(krum_gram_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GR_SMEM));
      B2_CUDA(cudaFuncSetAttribute(krum_gram_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, GR_SMEM));
      B2_CUDA(cudaFuncSetAttribute(krum_gram_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, GR_SMEM));
      scratch_ptr[dev] = p;
    }
  }
  unsigned char* base = static_cast<unsigned char*>(scratch_ptr[dev]);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t chunks = (n + GR_CC - 1) / GR_CC;
  int grid = sm_count();
  if (grid > B2_STATIC_GRAM_CTAS) grid = B2_STATIC_GRAM_CTAS;
  if (chunks < grid) grid = (int)chunks;
  B2_CUDA(cudaMemsetAsync(base, 0, GR_HEADER, s));  // the completion counter must start at zero
  launch_gram(grid, s, gram, pt, K, n, passes, reinterpret_cast<unsigned int*>(base), reinterpret_cast<float*>(base + GR_HEADER));
  B2_LAUNCH_CHECK("krum_gram_kernel(static scratch)");
  return 0;
}

int b2_krum_select(const float* dist, int K, int f, float* scores, int* order, void* stream) {
  if (K < 1 || K > B2_MAX_K) {
    set_error("krum_select: K=%d outside [1,%d]", K, B2_MAX_K);
    return 1;
  }
  krum_select_kernel<<<1, 64, 0, static_cast<cudaStream_t>(stream)>>>(dist, K, f, scores, order);
  B2_LAUNCH_CHECK("krum_select_kernel");
  return 0;
}

}  // extern "C"
