// fp32 -> bf16 staging for the tensor-core GEMMs: one pass writes the row-major bf16 copy, the transposed bf16
// copy (so the tensor stays K-major when it is used "the other way round" in dgrad / wgrad) and, optionally,
// the column sums (bias gradient of the last layer).  32 x 32 tiles through shared memory, coalesced both ways.
#include <cuda_bf16.h>

#include "../../include/myfyp_b200.h"
#include "common.cuh"

namespace b2 {

__global__ void __launch_bounds__(256) cast_transpose_kernel(const float* __restrict__ src, int R, int C, int lds, __nv_bfloat16* __restrict__ dst, int ldd,
                                                             __nv_bfloat16* __restrict__ dst_t, int ldt, float* __restrict__ colsum) {
  grid_dep_sync();
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = r0 + ty + 8 * i, c = c0 + tx;
    const float v = (r < R && c < C) ? src[(int64_t)r * lds + c] : 0.f;
    tile[ty + 8 * i][tx] = v;
    if (dst != nullptr && r < R && c < C) dst[(int64_t)r * ldd + c] = __float2bfloat16_rn(v);
  }
  __syncthreads();
  if (dst_t != nullptr) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = c0 + ty + 8 * i, r = r0 + tx;
      if (c < C && r < R) dst_t[(int64_t)c * ldt + r] = __float2bfloat16_rn(tile[tx][ty + 8 * i]);
    }
  }
  if (colsum != nullptr && ty == 0) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) s += tile[i][tx];
    if (c0 + tx < C && s != 0.f) atomicAdd(colsum + c0 + tx, s);
  }
}

}  // namespace b2

extern "C" int b2_cast_transpose(const float* src, int R, int C, int lds, void* dst_bf16, int ldd, void* dst_bf16_t, int ldt, float* colsum, void* stream) {
  if (R <= 0 || C <= 0) return 0;
  dim3 grid((C + 31) / 32, (R + 31) / 32);
  b2::launch_pdl(b2::cast_transpose_kernel, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), src, R, C, lds, static_cast<__nv_bfloat16*>(dst_bf16), ldd,
                                                                                static_cast<__nv_bfloat16*>(dst_bf16_t), ldt, colsum);
  B2_LAUNCH_CHECK("cast_transpose_kernel");
  return 0;
}
