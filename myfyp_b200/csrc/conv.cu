// T4 / T5: convolution as implicit GEMM in two steps (patch gather + the tcgen05 GEMM of gemm.cu), max-pooling and the
// small data-movement kernels of the backward pass.  Activations are NHWC bf16, i.e. exactly the [pixels, channels]
// row-major matrices the GEMM reads and writes, so a conv layer is
//     forward   col = im2col(X)                      Y = relu(col W^T + b)           (GEMM epilogue: bias + ReLU, bf16 out)
//     wgrad     dW  = dZ^T col                       (both operands MN-major, split-K over the pixels, fp32 atomics into the gradient slab)
//     dgrad     dcol = dZ W                          dX = col2im(dcol)               (gather form: no atomics)
// The patch (K) index is ordered (c, kh, kw) - PyTorch's own [Cout, Cin, KH, KW] weight layout flattened - so the weights
// need no permutation in either direction and dW lands directly in the slab.  A fully connected layer after a feature
// map is the same thing with a kernel as large as the map, which also reproduces nn.Flatten's NCHW feature order.
// Reference path: torch conv2d / max_pool2d through autograd (cuDNN); the reference itself ships no CNN.
#include <cuda_bf16.h>

#include "../../include/myfyp_b200.h"
#include "common.cuh"

namespace b2 {

struct ConvGeom {
  int B, H, W, C, KH, KW, stride, pad, OH, OW;
};

// one thread per (output pixel m, input channel c): KH*KW taps -> KH*KW consecutive columns of row m
template <bool NCHW_F32>
__global__ void __launch_bounds__(256) im2col_kernel(const void* __restrict__ xin, __nv_bfloat16* __restrict__ col, ConvGeom g, int ldc) {
  const int64_t total = (int64_t)g.B * g.OH * g.OW * g.C;
  const int KK = g.KH * g.KW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % g.C);
    const int64_t m = i / g.C;
    const int ow = (int)(m % g.OW);
    const int oh = (int)((m / g.OW) % g.OH);
    const int b = (int)(m / ((int64_t)g.OW * g.OH));
    __nv_bfloat16* dst = col + m * ldc + (int64_t)c * KK;
    for (int kh = 0; kh < g.KH; ++kh) {
      const int ih = oh * g.stride - g.pad + kh;
      for (int kw = 0; kw < g.KW; ++kw) {
        const int iw = ow * g.stride - g.pad + kw;
        __nv_bfloat16 v = __float2bfloat16_rn(0.f);
        if (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W) {
          if constexpr (NCHW_F32) {
            v = __float2bfloat16_rn(static_cast<const float*>(xin)[(((int64_t)b * g.C + c) * g.H + ih) * g.W + iw]);
          } else {
            v = static_cast<const __nv_bfloat16*>(xin)[(((int64_t)b * g.H + ih) * g.W + iw) * g.C + c];
          }
        }
        dst[kh * g.KW + kw] = v;
      }
    }
  }
}

// dx[b,h,w,c] = sum over the taps (kh,kw) whose output pixel exists of dcol[(b,oh,ow), c*KK + kh*KW + kw]
__global__ void __launch_bounds__(256) col2im_kernel(const __nv_bfloat16* __restrict__ dcol, __nv_bfloat16* __restrict__ dx, ConvGeom g, int ldc) {
  const int64_t total = (int64_t)g.B * g.H * g.W * g.C;
  const int KK = g.KH * g.KW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % g.C);
    const int64_t p = i / g.C;
    const int w = (int)(p % g.W);
    const int h = (int)((p / g.W) % g.H);
    const int b = (int)(p / ((int64_t)g.W * g.H));
    float acc = 0.f;
    for (int kh = 0; kh < g.KH; ++kh) {
      const int th = h + g.pad - kh;
      if (th < 0 || th % g.stride) continue;
      const int oh = th / g.stride;
      if (oh >= g.OH) continue;
      for (int kw = 0; kw < g.KW; ++kw) {
        const int tw = w + g.pad - kw;
        if (tw < 0 || tw % g.stride) continue;
        const int ow = tw / g.stride;
        if (ow >= g.OW) continue;
        const int64_t m = ((int64_t)b * g.OH + oh) * g.OW + ow;
        acc += __bfloat162float(dcol[m * ldc + (int64_t)c * KK + kh * g.KW + kw]);
      }
    }
    dx[i] = __float2bfloat16_rn(acc);
  }
}

// 2x2 / stride 2 max pooling over NHWC, remembering which of the four inputs won (first maximum, like PyTorch)
__global__ void __launch_bounds__(256) maxpool2_fwd_kernel(const __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ p, uint8_t* __restrict__ idx, int B, int H, int W,
                                                           int C) {
  const int OH = H / 2, OW = W / 2;
  const int64_t total = (int64_t)B * OH * OW * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int64_t q = i / C;
    const int ow = (int)(q % OW), oh = (int)((q / OW) % OH), b = (int)(q / ((int64_t)OW * OH));
    const __nv_bfloat16* src = y + (((int64_t)b * H + 2 * oh) * W + 2 * ow) * C + c;
    float best = __bfloat162float(src[0]);
    int bi = 0;
    const float v1 = __bfloat162float(src[C]), v2 = __bfloat162float(src[(int64_t)W * C]), v3 = __bfloat162float(src[(int64_t)W * C + C]);
    if (v1 > best) { best = v1; bi = 1; }
    if (v2 > best) { best = v2; bi = 2; }
    if (v3 > best) { best = v3; bi = 3; }
    p[i] = __float2bfloat16_rn(best);
    idx[i] = (uint8_t)bi;
  }
}

// dz[b,h,w,c] = dp[b,h/2,w/2,c] where (h,w) is the remembered maximum and the pre-pool activation is positive (ReLU)
__global__ void __launch_bounds__(256) maxpool2_relu_bwd_kernel(const __nv_bfloat16* __restrict__ dp, const uint8_t* __restrict__ idx, const __nv_bfloat16* __restrict__ y,
                                                                __nv_bfloat16* __restrict__ dz, int B, int H, int W, int C) {
  const int OH = H / 2, OW = W / 2;
  const int64_t total = (int64_t)B * H * W * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int64_t q = i / C;
    const int w = (int)(q % W), h = (int)((q / W) % H), b = (int)(q / ((int64_t)W * H));
    float g = 0.f;
    if ((h >> 1) < OH && (w >> 1) < OW) {
      const int64_t o = (((int64_t)b * OH + (h >> 1)) * OW + (w >> 1)) * C + c;
      const int pos = ((h & 1) << 1) | (w & 1);
      if (idx[o] == pos && __bfloat162float(y[i]) > 0.f) g = __bfloat162float(dp[o]);
    }
    dz[i] = __float2bfloat16_rn(g);
  }
}

// Fused forward tail of a conv block: 2x2 max-pool of Y (remembering the winner) AND the patch gather of the NEXT layer in
// one pass - the pooled value is scattered straight into the next layer's col matrix (up to KH*KW positions), the
// pooled activation itself is never materialised.  `g` is the NEXT layer's geometry over the pooled map [B, g.H, g.W, g.C].
// Entries of col that correspond to padding taps are never written and keep the zero they were allocated with.
__global__ void __launch_bounds__(256) pool_im2col_kernel(const __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ idx, __nv_bfloat16* __restrict__ col, ConvGeom g, int ldc) {
  const int H2 = g.H * 2, W2 = g.W * 2, C = g.C, KK = g.KH * g.KW;
  const int64_t total = (int64_t)g.B * g.H * g.W * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int64_t q = i / C;
    const int pw = (int)(q % g.W), ph = (int)((q / g.W) % g.H), b = (int)(q / ((int64_t)g.W * g.H));
    const __nv_bfloat16* src = y + (((int64_t)b * H2 + 2 * ph) * W2 + 2 * pw) * C + c;
    __nv_bfloat16 bv = src[0];
    float best = __bfloat162float(bv);
    int bi = 0;
    const __nv_bfloat16 r1 = src[C], r2 = src[(int64_t)W2 * C], r3 = src[(int64_t)W2 * C + C];
    if (__bfloat162float(r1) > best) { best = __bfloat162float(r1); bv = r1; bi = 1; }
    if (__bfloat162float(r2) > best) { best = __bfloat162float(r2); bv = r2; bi = 2; }
    if (__bfloat162float(r3) > best) { best = __bfloat162float(r3); bv = r3; bi = 3; }
    idx[i] = (uint8_t)bi;
    for (int kh = 0; kh < g.KH; ++kh) {
      const int th = ph + g.pad - kh;
      if (th < 0 || th % g.stride) continue;
      const int oh = th / g.stride;
      if (oh >= g.OH) continue;
      for (int kw = 0; kw < g.KW; ++kw) {
        const int tw = pw + g.pad - kw;
        if (tw < 0 || tw % g.stride) continue;
        const int ow = tw / g.stride;
        if (ow >= g.OW) continue;
        col[(((int64_t)b * g.OH + oh) * g.OW + ow) * ldc + (int64_t)c * KK + kh * g.KW + kw] = bv;
      }
    }
  }
}

// Fused backward head of a conv block: col2im of the NEXT layer's dcol (gather over its taps) -> gradient of the pooled
// activation -> routed to the remembered max position, times the ReLU mask -> dZ [B, 2H, 2W, C], plus the bias gradient
// (column sums of dZ) reduced per block in shared memory.  One pass, no dP / P buffers.
__global__ void __launch_bounds__(256) unpool_col2im_kernel(const __nv_bfloat16* __restrict__ dcol, const uint8_t* __restrict__ idx, const __nv_bfloat16* __restrict__ y,
                                                            __nv_bfloat16* __restrict__ dz, float* __restrict__ bias_grad, ConvGeom g, int ldc) {
  extern __shared__ float s_db[];  // C floats
  const int H2 = g.H * 2, W2 = g.W * 2, C = g.C, KK = g.KH * g.KW;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s_db[c] = 0.f;
  __syncthreads();
  const int64_t total = (int64_t)g.B * g.H * g.W * C;
  const __nv_bfloat16 zero = __float2bfloat16_rn(0.f);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int64_t q = i / C;
    const int pw = (int)(q % g.W), ph = (int)((q / g.W) % g.H), b = (int)(q / ((int64_t)g.W * g.H));
    float acc = 0.f;
    for (int kh = 0; kh < g.KH; ++kh) {
      const int th = ph + g.pad - kh;
      if (th < 0 || th % g.stride) continue;
      const int oh = th / g.stride;
      if (oh >= g.OH) continue;
      for (int kw = 0; kw < g.KW; ++kw) {
        const int tw = pw + g.pad - kw;
        if (tw < 0 || tw % g.stride) continue;
        const int ow = tw / g.stride;
        if (ow >= g.OW) continue;
        acc += __bfloat162float(dcol[(((int64_t)b * g.OH + oh) * g.OW + ow) * ldc + (int64_t)c * KK + kh * g.KW + kw]);
      }
    }
    const int win = idx[i];
    const int64_t base = (((int64_t)b * H2 + 2 * ph) * W2 + 2 * pw) * C + c;
    const int64_t offs[4] = {0, C, (int64_t)W2 * C, (int64_t)W2 * C + C};
    const __nv_bfloat16 gq = __float2bfloat16_rn(acc);  // the value the unfused path stores as dP
    const bool live = __bfloat162float(y[base + offs[win]]) > 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) dz[base + offs[k]] = (k == win && live) ? gq : zero;
    if (live && bias_grad != nullptr) atomicAdd(&s_db[c], __bfloat162float(gq));
  }
  __syncthreads();
  if (bias_grad != nullptr)
    for (int c = threadIdx.x; c < C; c += blockDim.x)
      if (s_db[c] != 0.f) atomicAdd(bias_grad + c, s_db[c]);
}

// out[c] += sum_m z[m, c]   (bias gradient of a conv layer): each block reduces a slab of rows in shared memory
__global__ void __launch_bounds__(256) colsum_bf16_kernel(const __nv_bfloat16* __restrict__ z, int M, int C, int ld, float* __restrict__ out, int rows_per_block) {
  extern __shared__ float s_acc[];  // C floats
  for (int c = threadIdx.x; c < C; c += blockDim.x) s_acc[c] = 0.f;
  __syncthreads();
  const int m0 = blockIdx.x * rows_per_block;
  const int m1 = min(M, m0 + rows_per_block);
  const int cols_per_pass = blockDim.x < C ? blockDim.x : C;  // threads over columns (coalesced), rows striped over thread groups
  const int groups = blockDim.x / cols_per_pass;
  const int gid = threadIdx.x / cols_per_pass, lane = threadIdx.x % cols_per_pass;
  if (gid < groups) {
    for (int c = lane; c < C; c += cols_per_pass) {
      float acc = 0.f;
      for (int m = m0 + gid; m < m1; m += groups) acc += __bfloat162float(z[(int64_t)m * ld + c]);
      atomicAdd(&s_acc[c], acc);
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x)
    if (s_acc[c] != 0.f) atomicAdd(out + c, s_acc[c]);
}

}  // namespace b2

using namespace b2;

extern "C" {

int b2_im2col(const void* x, int x_is_nchw_f32, void* col_bf16, int ldc, int B, int H, int W, int C, int KH, int KW, int stride, int pad, void* stream) {
  ConvGeom g{B, H, W, C, KH, KW, stride, pad, (H + 2 * pad - KH) / stride + 1, (W + 2 * pad - KW) / stride + 1};
  if (g.OH <= 0 || g.OW <= 0 || ldc < C * KH * KW) {
    set_error("im2col: bad geometry (OH=%d OW=%d ldc=%d < %d)", g.OH, g.OW, ldc, C * KH * KW);
    return 1;
  }
  const int64_t total = (int64_t)B * g.OH * g.OW * C;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (x_is_nchw_f32) im2col_kernel<true><<<stream_grid(total, 256), 256, 0, s>>>(x, static_cast<__nv_bfloat16*>(col_bf16), g, ldc);
  else im2col_kernel<false><<<stream_grid(total, 256), 256, 0, s>>>(x, static_cast<__nv_bfloat16*>(col_bf16), g, ldc);
  B2_LAUNCH_CHECK("im2col_kernel");
  return 0;
}

int b2_col2im(const void* dcol_bf16, int ldc, void* dx_bf16, int B, int H, int W, int C, int KH, int KW, int stride, int pad, void* stream) {
  ConvGeom g{B, H, W, C, KH, KW, stride, pad, (H + 2 * pad - KH) / stride + 1, (W + 2 * pad - KW) / stride + 1};
  const int64_t total = (int64_t)B * H * W * C;
  col2im_kernel<<<stream_grid(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(dcol_bf16), static_cast<__nv_bfloat16*>(dx_bf16), g, ldc);
  B2_LAUNCH_CHECK("col2im_kernel");
  return 0;
}

int b2_maxpool2_fwd(const void* y_bf16, void* p_bf16, void* idx_u8, int B, int H, int W, int C, void* stream) {
  const int64_t total = (int64_t)B * (H / 2) * (W / 2) * C;
  maxpool2_fwd_kernel<<<stream_grid(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(y_bf16), static_cast<__nv_bfloat16*>(p_bf16),
                                                                                             static_cast<uint8_t*>(idx_u8), B, H, W, C);
  B2_LAUNCH_CHECK("maxpool2_fwd_kernel");
  return 0;
}

int b2_maxpool2_relu_bwd(const void* dp_bf16, const void* idx_u8, const void* y_bf16, void* dz_bf16, int B, int H, int W, int C, void* stream) {
  const int64_t total = (int64_t)B * H * W * C;
  maxpool2_relu_bwd_kernel<<<stream_grid(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dp_bf16), static_cast<const uint8_t*>(idx_u8), static_cast<const __nv_bfloat16*>(y_bf16), static_cast<__nv_bfloat16*>(dz_bf16), B, H, W, C);
  B2_LAUNCH_CHECK("maxpool2_relu_bwd_kernel");
  return 0;
}

int b2_pool_im2col(const void* y_bf16, void* idx_u8, void* col_bf16, int ldc, int B, int PH, int PW, int C, int KH, int KW, int stride, int pad, void* stream) {
  ConvGeom g{B, PH, PW, C, KH, KW, stride, pad, (PH + 2 * pad - KH) / stride + 1, (PW + 2 * pad - KW) / stride + 1};
  if (g.OH <= 0 || g.OW <= 0 || ldc < C * KH * KW) {
    set_error("pool_im2col: bad geometry");
    return 1;
  }
  const int64_t total = (int64_t)B * PH * PW * C;
  pool_im2col_kernel<<<stream_grid(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(y_bf16), static_cast<uint8_t*>(idx_u8),
                                                                                            static_cast<__nv_bfloat16*>(col_bf16), g, ldc);
  B2_LAUNCH_CHECK("pool_im2col_kernel");
  return 0;
}

int b2_unpool_col2im(const void* dcol_bf16, int ldc, const void* idx_u8, const void* y_bf16, void* dz_bf16, float* bias_grad, int B, int PH, int PW, int C, int KH, int KW,
                     int stride, int pad, void* stream) {
  ConvGeom g{B, PH, PW, C, KH, KW, stride, pad, (PH + 2 * pad - KH) / stride + 1, (PW + 2 * pad - KW) / stride + 1};
  if (C > 8192) {
    set_error("unpool_col2im: C=%d too wide", C);
    return 1;
  }
  const int64_t total = (int64_t)B * PH * PW * C;
  unpool_col2im_kernel<<<stream_grid(total, 256, 2), 256, C * sizeof(float), static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dcol_bf16), static_cast<const uint8_t*>(idx_u8), static_cast<const __nv_bfloat16*>(y_bf16), static_cast<__nv_bfloat16*>(dz_bf16), bias_grad, g,
      ldc);
  B2_LAUNCH_CHECK("unpool_col2im_kernel");
  return 0;
}

int b2_colsum_bf16(const void* z_bf16, int M, int C, int ld, float* out, void* stream) {
  if (M <= 0 || C <= 0) return 0;
  if (C > 8192) {
    set_error("colsum: C=%d too wide", C);
    return 1;
  }
  int blocks = sm_count() * 2;
  int rows_per_block = (M + blocks - 1) / blocks;
  if (rows_per_block < 32) rows_per_block = 32;
  blocks = (M + rows_per_block - 1) / rows_per_block;
  colsum_bf16_kernel<<<blocks, 256, C * sizeof(float), static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(z_bf16), M, C, ld, out, rows_per_block);
  B2_LAUNCH_CHECK("colsum_bf16_kernel");
  return 0;
}

}  // extern "C"
