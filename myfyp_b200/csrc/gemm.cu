// T1 / T2: bf16 GEMM on tcgen05 tensor cores fed by TMA, with the MLP's fused epilogues.
//
//   D[M,N] = epilogue( A[M,K] * B[N,K]^T )        A, B bf16, both K-major (row-major, K contiguous); fp32 accumulate
//
// One kernel serves the three GEMMs of a linear layer because every activation and
// weight is kept in BOTH orientations by the epilogue that produces it (the transposed copy is a coalesced store
// there), so A and B are always K-major and a single shared-memory / descriptor layout is enough:
//   forward  Y  = relu(X W^T + b)      A = X  [B,in]     B = W   [out,in]   -> Y bf16 [B,out] and Y^T [out,B]
//   dgrad    dX = (dY W) * (Xact > 0)  A = dY [B,out]    B = W^T [in,out]   -> dX bf16 [B,in] and dX^T [in,B], colsum -> db
//   wgrad    dW = dY^T X               A = dY^T [out,B]  B = X^T [in,B]     -> dW fp32 [out,in]
//
// Structure (warp specialised, one 128 x 64 output tile per CTA):
//   warp 0     : TMA producer  - cp.async.bulk.tensor.2d (128B-swizzled boxes) into a 4-stage ring, mbarrier expect_tx
//   warp 1     : MMA issuer    - one elected thread, tcgen05.mma kind::f16 M=128 N=64 K=16, accumulator in TMEM
//   warps 2..5 : epilogue      - tcgen05.ld (32 lanes x 32 columns per warp), bias / mask / ReLU, stores, column sums
// TMA zero-fills out-of-bounds rows and the K tail, so M, N, K need no padding (only 16-byte row pitches).
//
// Operand majors.  Either operand may instead be stored "MN-major" = transposed, i.e. as [K, M] (resp. [K, N]) row-major
// (flags a_mn / b_mn).  TMA then fetches 64(K) x 64(MN) boxes (two side by side for the 128-row A tile) and the UMMA
// descriptor describes the canonical MN-major SWIZZLE_128B layout (LBO = distance between 64-element MN blocks, SBO =
// distance between 8-row K groups, major bit set in the instruction descriptor).  With that, the backward GEMMs of a
// linear / convolution layer read the SAME buffers as the forward one - no transposed copies:
//   wgrad  dW[out,in] = dY^T X   : A = dY [B,out] as MN-major, B = X [B,in] as MN-major
//   dgrad  dX[B,in]   = dY W     : A = dY [B,out] K-major,     B = W [out,in] as MN-major
// Split-K (grid.z): each z-slice reduces a range of k-blocks and atomically adds into the (pre-zeroed / accumulated) fp32
// output - used by wgrad where K = batch x pixels is long and M x N is small.
#include <cuda.h>
#include <cuda_bf16.h>

#include "../../include/myfyp_b200.h"
#include "common.cuh"
#include "tcgen05.cuh"

namespace b2 {

constexpr int GM_BM = 128, GM_BN = 64, GM_BK = 64;
constexpr int GM_STAGES = 4;
constexpr int GM_A_BYTES = GM_BM * GM_BK * 2;  // 16 KB
constexpr int GM_B_BYTES = GM_BN * GM_BK * 2;  //  8 KB
constexpr int GM_STAGE_BYTES = GM_A_BYTES + GM_B_BYTES;
constexpr int GM_SMEM = GM_STAGES * GM_STAGE_BYTES + 1024;
constexpr int GM_THREADS = 192;

struct GemmEpilogue {
  float* out_f32;
  __nv_bfloat16* out_bf16;
  __nv_bfloat16* out_bf16_t;
  const float* bias;
  const __nv_bfloat16* mask;
  float* colsum;
  int ldo, ldo16, ldt, ldm;
  int relu, accumulate;
};

__device__ __forceinline__ uint64_t smem_desc_sw128_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
  // MN-major canonical SWIZZLE_128B: rows are K (128 bytes = 64 MN elements each), 8-row groups 1024 bytes apart (SBO),
  // further 64-element MN blocks `lbo_bytes` apart (LBO)
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GM_THREADS, 1) gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, int M, int N, int K,
                                                                   int kb_per_split, GemmEpilogue ep) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[GM_STAGES], empty_bar[GM_STAGES], done_bar;
  __shared__ uint32_t tmem_base_s;

  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.y * GM_BM, n0 = blockIdx.x * GM_BN;
  const int total_k = (K + GM_BK - 1) / GM_BK;
  const int kb0 = blockIdx.z * kb_per_split;
  const int num_k = (total_k - kb0 < kb_per_split) ? (total_k - kb0 > 0 ? total_k - kb0 : 0) : kb_per_split;
  const bool split = gridDim.z > 1;

  if (tid == 0) {
    tc::tma_prefetch_desc(&tmap_a);
    tc::tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < GM_STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    tc::mbar_init(&done_bar, 1);
    tc::mbar_fence_init();
  }
  if (warp == 1) tc::tmem_alloc<64>(&tmem_base_s);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    if (tc::elect_one()) {
      for (int kb = 0; kb < num_k; ++kb) {
        const int s = kb % GM_STAGES;
        const uint32_t ph = (uint32_t)((kb / GM_STAGES) & 1);
        tc::mbar_wait(&empty_bar[s], ph ^ 1u);
        uint8_t* a_dst = tiles + s * GM_STAGE_BYTES;
        uint8_t* b_dst = a_dst + GM_A_BYTES;
        tc::mbar_arrive_expect_tx(&full_bar[s], GM_STAGE_BYTES);
        const int k0 = (kb0 + kb) * GM_BK;
        if constexpr (A_MN) {  // stored [K, M]: two 64(K) x 64(M) boxes side by side
          tc::tma_load_2d(a_dst, &tmap_a, &full_bar[s], m0, k0);
          tc::tma_load_2d(a_dst + GM_BK * 128, &tmap_a, &full_bar[s], m0 + 64, k0);
        } else {
          tc::tma_load_2d(a_dst, &tmap_a, &full_bar[s], k0, m0);
        }
        if constexpr (B_MN) {
          tc::tma_load_2d(b_dst, &tmap_b, &full_bar[s], n0, k0);
        } else {
          tc::tma_load_2d(b_dst, &tmap_b, &full_bar[s], k0, n0);
        }
      }
    }
  } else if (warp == 1) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc(tc::FMT_BF16, GM_BM, GM_BN) | (A_MN ? (1u << 15) : 0u) | (B_MN ? (1u << 16) : 0u);
      for (int kb = 0; kb < num_k; ++kb) {
        const int s = kb % GM_STAGES;
        const uint32_t ph = (uint32_t)((kb / GM_STAGES) & 1);
        tc::mbar_wait(&full_bar[s], ph);
        tc::fence_after_sync();
        const uint32_t a_addr = tc::smem_u32(tiles + s * GM_STAGE_BYTES);
        const uint32_t b_addr = a_addr + GM_A_BYTES;
#pragma unroll
        for (int ks = 0; ks < GM_BK / 16; ++ks) {  // UMMA_K = 16 bf16 = 32 bytes inside the 128-byte swizzle row
          // K-major: 16 elements = 32 bytes along the swizzle row; MN-major: 16 K-rows = 2048 bytes down the tile
          const uint64_t da = A_MN ? smem_desc_sw128_mn(a_addr + ks * 2048, GM_BK * 128) : tc::smem_desc_sw128(a_addr + ks * 32);
          const uint64_t db = B_MN ? smem_desc_sw128_mn(b_addr + ks * 2048, GM_BK * 128) : tc::smem_desc_sw128(b_addr + ks * 32);
          tc::mma_bf16(tmem_base, da, db, idesc, (kb | ks) != 0 ? 1u : 0u);
        }
        tc::mma_commit(&empty_bar[s]);
      }
      tc::mma_commit(&done_bar);
    }
  } else {
    // ================================ epilogue: this warp owns TMEM lanes [32q, 32q+32), q = warp % 4
    const int q = warp & 3;
    const int row = m0 + q * 32 + lane;
    tc::mbar_wait(&done_bar, 0);
    tc::fence_after_sync();
#pragma unroll
    for (int cb = 0; cb < GM_BN / 32; ++cb) {
      float v[32];
      if (num_k > 0) {
        tc::tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(cb * 32), v);
      } else {
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = 0.f;
      }
      const int nb = n0 + cb * 32;
      const bool row_ok = row < M;
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        const int n = nb + e;
        float x = v[e];
        if (ep.bias != nullptr && n < N) x += __ldg(ep.bias + n);
        if (ep.relu) x = fmaxf(x, 0.f);
        if (ep.mask != nullptr && row_ok && n < N) x = (__bfloat162float(ep.mask[(int64_t)row * ep.ldm + n]) > 0.f) ? x : 0.f;
        v[e] = (row_ok && n < N) ? x : 0.f;
      }
      if (ep.out_f32 != nullptr && row_ok) {
        float* o = ep.out_f32 + (int64_t)row * ep.ldo + nb;
        if (split) {
#pragma unroll
          for (int e = 0; e < 32; ++e)
            if (nb + e < N && v[e] != 0.f) atomicAdd(o + e, v[e]);
        } else if (nb + 32 <= N && (ep.ldo & 3) == 0) {
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            float4 w = make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]);
            if (ep.accumulate) {
              const float4 old = *reinterpret_cast<const float4*>(o + e);
              w.x += old.x; w.y += old.y; w.z += old.z; w.w += old.w;
            }
            *reinterpret_cast<float4*>(o + e) = w;
          }
        } else {
#pragma unroll
          for (int e = 0; e < 32; ++e)
            if (nb + e < N) o[e] = ep.accumulate ? o[e] + v[e] : v[e];
        }
      }
      if (ep.out_bf16 != nullptr && row_ok) {
        __nv_bfloat16* o = ep.out_bf16 + (int64_t)row * ep.ldo16 + nb;
        if (nb + 32 <= N && (ep.ldo16 & 7) == 0) {
#pragma unroll
          for (int e = 0; e < 32; e += 8) {
            __nv_bfloat162 p0 = __floats2bfloat162_rn(v[e], v[e + 1]), p1 = __floats2bfloat162_rn(v[e + 2], v[e + 3]);
            __nv_bfloat162 p2 = __floats2bfloat162_rn(v[e + 4], v[e + 5]), p3 = __floats2bfloat162_rn(v[e + 6], v[e + 7]);
            uint4 w;
            w.x = *reinterpret_cast<uint32_t*>(&p0); w.y = *reinterpret_cast<uint32_t*>(&p1);
            w.z = *reinterpret_cast<uint32_t*>(&p2); w.w = *reinterpret_cast<uint32_t*>(&p3);
            *reinterpret_cast<uint4*>(o + e) = w;
          }
        } else {
#pragma unroll
          for (int e = 0; e < 32; ++e)
            if (nb + e < N) o[e] = __float2bfloat16_rn(v[e]);
        }
      }
      if (ep.out_bf16_t != nullptr) {
        // transposed copy: for a fixed column the 32 lanes hold 32 consecutive rows -> one 64-byte row segment
#pragma unroll
        for (int e = 0; e < 32; ++e)
          if (nb + e < N && row_ok) ep.out_bf16_t[(int64_t)(nb + e) * ep.ldt + row] = __float2bfloat16_rn(v[e]);
      }
      if (ep.colsum != nullptr) {
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float sum = warp_sum(v[e]);
          if (lane == 0 && nb + e < N && sum != 0.f) atomicAdd(ep.colsum + nb + e, sum);
        }
      }
    }
    tc::fence_before_sync();
  }
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc<64>(tmem_base);
}

// ---------------------------------------------------------------------------------------------- host: tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = (EncodeTiledFn)p;
  }
  return fn;
}

// bf16 [rows, ld] row-major, box = 64 (inner) x box_rows, 128-byte swizzle
static int make_tmap(CUtensorMap* tm, const void* base, int rows, int cols, int ld, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return 1;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || (ld & 7)) {
    set_error("gemm operand must be 16-byte aligned with a row pitch that is a multiple of 8 elements (ld=%d)", ld);
    return 1;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)GM_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%d cols=%d ld=%d)", (int)r, rows, cols, ld);
    return 1;
  }
  return 0;
}

}  // namespace b2

using namespace b2;

extern "C" int b2_gemm_bf16(const b2_gemm_args* a, void* stream) {
  if (a == nullptr || a->M <= 0 || a->N <= 0 || a->K < 0) {
    set_error("gemm: bad shape");
    return 1;
  }
  CUtensorMap ta, tb;
  // K-major operand: [M, K] rows = M; MN-major operand: stored [K, M], rows = K, 64 x 64 boxes
  if (a->a_mn ? make_tmap(&ta, a->A, a->K, a->M, a->lda, GM_BK) : make_tmap(&ta, a->A, a->M, a->K, a->lda, GM_BM)) return 1;
  if (a->b_mn ? make_tmap(&tb, a->B, a->K, a->N, a->ldb, GM_BK) : make_tmap(&tb, a->B, a->N, a->K, a->ldb, GM_BN)) return 1;
  static bool attr_set[64] = {false};
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    B2_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, GM_SMEM));
    B2_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GM_SMEM));
    B2_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, GM_SMEM));
    B2_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GM_SMEM));
    attr_set[dev] = true;
  }
  const int total_k = (a->K + GM_BK - 1) / GM_BK;
  int splits = a->split_k > 1 ? a->split_k : 1;
  if (splits > total_k) splits = total_k > 0 ? total_k : 1;
  if (splits > 1 && (a->out_f32 == nullptr || a->out_bf16 != nullptr || a->out_bf16_t != nullptr || a->bias != nullptr || a->relu || a->mask != nullptr)) {
    set_error("gemm: split_k only combines with a plain fp32 (accumulating) output");
    return 1;
  }
  const int kb_per_split = (total_k + splits - 1) / splits;
  splits = kb_per_split > 0 ? (total_k + kb_per_split - 1) / kb_per_split : 1;
  GemmEpilogue ep{a->out_f32, static_cast<__nv_bfloat16*>(a->out_bf16), static_cast<__nv_bfloat16*>(a->out_bf16_t), a->bias, static_cast<const __nv_bfloat16*>(a->mask),
                  a->colsum, a->ldo, a->ldo16, a->ldt, a->ldm, a->relu, a->accumulate};
  dim3 grid((a->N + GM_BN - 1) / GM_BN, (a->M + GM_BM - 1) / GM_BM, splits > 0 ? splits : 1);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int kps = kb_per_split > 0 ? kb_per_split : 1;
  if (a->a_mn && a->b_mn) gemm_bf16_kernel<true, true><<<grid, GM_THREADS, GM_SMEM, st>>>(ta, tb, a->M, a->N, a->K, kps, ep);
  else if (a->a_mn) gemm_bf16_kernel<true, false><<<grid, GM_THREADS, GM_SMEM, st>>>(ta, tb, a->M, a->N, a->K, kps, ep);
  else if (a->b_mn) gemm_bf16_kernel<false, true><<<grid, GM_THREADS, GM_SMEM, st>>>(ta, tb, a->M, a->N, a->K, kps, ep);
  else gemm_bf16_kernel<false, false><<<grid, GM_THREADS, GM_SMEM, st>>>(ta, tb, a->M, a->N, a->K, kps, ep);
  B2_LAUNCH_CHECK("gemm_bf16_kernel");
  return 0;
}
