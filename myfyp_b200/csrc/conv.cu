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
__global__ void __launch_bounds__(256) col2im_kernel(const __nv_bfloat16* __restrict__ dcol, __nv_bfloat16* __restrict__ dx, ConvGeom g, int ldc,
                                                     const __nv_bfloat16* __restrict__ add) {
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
    if (add != nullptr) acc += __bfloat162float(add[i]);  // second consumer of the same activation (residual / shortcut branch)
    dx[i] = __float2bfloat16_rn(acc);
  }
}

// ------------------------------------------------------------------------------------------------ T5: batch normalisation
// Training statistics come for free from the producing GEMM's epilogue (column sums and sums of squares of the fp32
// accumulators, `stats` = [sum[C], sumsq[C]]).  mean / invstd are recomputed from them wherever needed (2 loads + rsqrt).
struct BnArgs {
  const float* stats;     // [2C] batch sums (training) - nullptr in eval mode
  const float* gamma;     // [C]
  const float* beta;      // [C]
  float* running_mean;    // [C] (updated by block 0 in training, read in eval)
  float* running_var;     // [C]
  float* num_batches;     // fp32 mirror slot of num_batches_tracked (may be null)
  float eps, momentum;
  int M, C, relu;
};

// All three kernels use the same thread layout: a thread owns 8 consecutive channels (one 16-byte bf16 vector) and walks rows;
// a block covers `rows_per_iter = 256 / (C / 8)` rows per iteration (C % 8 == 0, C / 8 <= 256 -> C <= 2048).
struct Bf16x8 {
  uint4 raw;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { raw = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void store(__nv_bfloat16* p) const { *reinterpret_cast<uint4*>(p) = raw; }
  __device__ __forceinline__ void unpack(float (&f)[8]) const {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 t = __bfloat1622float2(h[i]);
      f[2 * i] = t.x;
      f[2 * i + 1] = t.y;
    }
  }
  __device__ __forceinline__ void pack(const float (&f)[8]) {
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  }
};

// y = act(gamma * (z - mean) * invstd + beta (+ residual))
__global__ void __launch_bounds__(256) bn_apply_kernel(const __nv_bfloat16* __restrict__ z, const __nv_bfloat16* __restrict__ residual, __nv_bfloat16* __restrict__ y, BnArgs a) {
  const int groups = a.C >> 3, rows_per_iter = blockDim.x / groups;
  const int cg = threadIdx.x % groups, r = threadIdx.x / groups;
  const float invM = 1.f / (float)a.M;
  if (r < rows_per_iter) {
    float scale[8], shift[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = cg * 8 + j;
      float mean, var;
      if (a.stats != nullptr) {
        mean = a.stats[c] * invM;
        var = fmaxf(a.stats[a.C + c] * invM - mean * mean, 0.f);
      } else {
        mean = a.running_mean[c];
        var = a.running_var[c];
      }
      scale[j] = rsqrtf(var + a.eps) * a.gamma[c];
      shift[j] = a.beta[c] - mean * scale[j];
    }
    for (int m = blockIdx.x * rows_per_iter + r; m < a.M; m += gridDim.x * rows_per_iter) {
      const int64_t i = (int64_t)m * a.C + cg * 8;
      Bf16x8 v;
      float f[8];
      v.load(z + i);
      v.unpack(f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = f[j] * scale[j] + shift[j];
      if (residual != nullptr) {
        Bf16x8 rv;
        float rf[8];
        rv.load(residual + i);
        rv.unpack(rf);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] += rf[j];
      }
      if (a.relu) {
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
      }
      v.pack(f);
      v.store(y + i);
    }
  }
  if (a.stats != nullptr && blockIdx.x == 0) {  // running statistics (PyTorch semantics: unbiased variance, momentum)
    for (int c = threadIdx.x; c < a.C; c += blockDim.x) {
      const float mean = a.stats[c] * invM;
      const float var = fmaxf(a.stats[a.C + c] * invM - mean * mean, 0.f);
      const float unbiased = a.M > 1 ? var * (float)a.M / (float)(a.M - 1) : var;
      a.running_mean[c] = (1.f - a.momentum) * a.running_mean[c] + a.momentum * mean;
      a.running_var[c] = (1.f - a.momentum) * a.running_var[c] + a.momentum * unbiased;
    }
    if (threadIdx.x == 0 && a.num_batches != nullptr) a.num_batches[0] += 1.f;
  }
}

// pass 1 of the backward: g = dy * (y > 0 if relu);  red[c] += sum g, red[C + c] += sum g * xhat;  g is stored for pass 2
// (and is at the same time the gradient that flows into the residual branch)
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ z,
                                                            __nv_bfloat16* __restrict__ gbuf, float* __restrict__ red, BnArgs a) {
  extern __shared__ float s_red[];  // 2C
  for (int c = threadIdx.x; c < 2 * a.C; c += blockDim.x) s_red[c] = 0.f;
  __syncthreads();
  const int groups = a.C >> 3, rows_per_iter = blockDim.x / groups;
  const int cg = threadIdx.x % groups, r = threadIdx.x / groups;
  const float invM = 1.f / (float)a.M;
  if (r < rows_per_iter) {
    float mean[8], invstd[8], sg[8], sgx[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = cg * 8 + j;
      mean[j] = a.stats[c] * invM;
      invstd[j] = rsqrtf(fmaxf(a.stats[a.C + c] * invM - mean[j] * mean[j], 0.f) + a.eps);
      sg[j] = sgx[j] = 0.f;
    }
    for (int m = blockIdx.x * rows_per_iter + r; m < a.M; m += gridDim.x * rows_per_iter) {
      const int64_t i = (int64_t)m * a.C + cg * 8;
      Bf16x8 vd, vz;
      float g[8], zf[8];
      vd.load(dy + i);
      vz.load(z + i);
      vd.unpack(g);
      vz.unpack(zf);
      if (a.relu) {
        Bf16x8 vy;
        float yf[8];
        vy.load(y + i);
        vy.unpack(yf);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (!(yf[j] > 0.f)) g[j] = 0.f;
        vd.pack(g);
      }
      vd.store(gbuf + i);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        sg[j] += g[j];
        sgx[j] += g[j] * (zf[j] - mean[j]) * invstd[j];
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(&s_red[cg * 8 + j], sg[j]);
      atomicAdd(&s_red[a.C + cg * 8 + j], sgx[j]);
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * a.C; c += blockDim.x)
    if (s_red[c] != 0.f) atomicAdd(red + c, s_red[c]);
}

// pass 2: dz = gamma * invstd * (g - mean(g) - xhat * mean(g * xhat));  block 0 adds dgamma / dbeta into the gradient slab
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ gbuf, const __nv_bfloat16* __restrict__ z, const float* __restrict__ red,
                                                           __nv_bfloat16* __restrict__ dz, float* __restrict__ dgamma, float* __restrict__ dbeta, BnArgs a) {
  const int groups = a.C >> 3, rows_per_iter = blockDim.x / groups;
  const int cg = threadIdx.x % groups, r = threadIdx.x / groups;
  const float invM = 1.f / (float)a.M;
  if (r < rows_per_iter) {
    // dz = k1 * g + k2 * z + k3   with xhat = (z - mean) * invstd
    float k1[8], k2[8], k3[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = cg * 8 + j;
      const float mean = a.stats[c] * invM;
      const float invstd = rsqrtf(fmaxf(a.stats[a.C + c] * invM - mean * mean, 0.f) + a.eps);
      const float gi = a.gamma[c] * invstd, mg = red[c] * invM, mgx = red[a.C + c] * invM;
      k1[j] = gi;
      k2[j] = -gi * mgx * invstd;
      k3[j] = -gi * mg + gi * mgx * invstd * mean;
    }
    for (int m = blockIdx.x * rows_per_iter + r; m < a.M; m += gridDim.x * rows_per_iter) {
      const int64_t i = (int64_t)m * a.C + cg * 8;
      Bf16x8 vg, vz;
      float g[8], zf[8];
      vg.load(gbuf + i);
      vz.load(z + i);
      vg.unpack(g);
      vz.unpack(zf);
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] = k1[j] * g[j] + k2[j] * zf[j] + k3[j];
      vg.pack(g);
      vg.store(dz + i);
    }
  }
  if (blockIdx.x == 0)
    for (int c = threadIdx.x; c < a.C; c += blockDim.x) {
      dbeta[c] += red[c];
      dgamma[c] += red[a.C + c];
    }
}

// global average pooling over the H*W pixels of an NHWC map, and its backward (broadcast / HW)
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int B, int HW, int C, int ldo) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * C; i += gridDim.x * blockDim.x) {
    const int c = i % C, b = i / C;
    float acc = 0.f;
    for (int p = 0; p < HW; ++p) acc += __bfloat162float(x[((int64_t)b * HW + p) * C + c]);
    out[(int64_t)b * ldo + c] = __float2bfloat16_rn(acc / (float)HW);
  }
}
__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dout, int ldo, __nv_bfloat16* __restrict__ dx, int B, int HW, int C) {
  const int64_t total = (int64_t)B * HW * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int b = (int)(i / ((int64_t)HW * C));
    dx[i] = __float2bfloat16_rn(__bfloat162float(dout[(int64_t)b * ldo + c]) / (float)HW);
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

// conv weights fp32 [cout][cin][taps] -> bf16 [cout][taps][cin]; blockIdx.y = layer
__global__ void repack_conv_weights_kernel(const float* __restrict__ slab, __nv_bfloat16* __restrict__ dst, const int64_t* __restrict__ table) {
  const int64_t* t = table + 5 * blockIdx.y;
  const float* src = slab + t[0];
  __nv_bfloat16* out = dst + t[1];
  const int64_t cin = t[3], taps = t[4], n = t[2] * cin * taps;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c = i % cin, r = i / cin, tp = r % taps, co = r / taps;
    out[i] = __float2bfloat16_rn(src[(co * cin + c) * taps + tp]);
  }
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

int b2_col2im(const void* dcol_bf16, int ldc, void* dx_bf16, const void* add_bf16, int B, int H, int W, int C, int KH, int KW, int stride, int pad, void* stream) {
  ConvGeom g{B, H, W, C, KH, KW, stride, pad, (H + 2 * pad - KH) / stride + 1, (W + 2 * pad - KW) / stride + 1};
  const int64_t total = (int64_t)B * H * W * C;
  col2im_kernel<<<stream_grid(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(dcol_bf16), static_cast<__nv_bfloat16*>(dx_bf16), g, ldc,
                                                                                       static_cast<const __nv_bfloat16*>(add_bf16));
  B2_LAUNCH_CHECK("col2im_kernel");
  return 0;
}

// blocks of 256 threads covering 256 / (C / 8) rows per iteration; enough blocks to fill the chip, at least ~4 rows per thread
static int bn_grid(int M, int C, int ctas_per_sm = 8) {
  const int rows_per_iter = 256 / (C >> 3) > 0 ? 256 / (C >> 3) : 1;
  int64_t need = ((int64_t)M + rows_per_iter * 4 - 1) / (rows_per_iter * 4);
  const int64_t cap = (int64_t)sm_count() * ctas_per_sm;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

static BnArgs bn_args(const float* stats, const float* gamma, const float* beta, float* rm, float* rv, float* nb, float eps, float momentum, int M, int C, int relu) {
  return BnArgs{stats, gamma, beta, rm, rv, nb, eps, momentum, M, C, relu};
}

int b2_bn_apply(const void* z_bf16, const void* residual_bf16, void* y_bf16, const float* stats, const float* gamma, const float* beta, float* running_mean, float* running_var,
                float* num_batches, float eps, float momentum, int M, int C, int relu, void* stream) {
  if (M <= 0 || C <= 0) return 0;
  if ((C & 7) || C > 2048) {
    set_error("bn: C=%d must be a multiple of 8, at most 2048", C);
    return 1;
  }
  bn_apply_kernel<<<bn_grid(M, C), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(z_bf16), static_cast<const __nv_bfloat16*>(residual_bf16), static_cast<__nv_bfloat16*>(y_bf16),
      bn_args(stats, gamma, beta, running_mean, running_var, num_batches, eps, momentum, M, C, relu));
  B2_LAUNCH_CHECK("bn_apply_kernel");
  return 0;
}

int b2_bn_backward(const void* dy_bf16, const void* y_bf16, const void* z_bf16, void* g_bf16, void* dz_bf16, float* red2c, const float* stats, const float* gamma, float* dgamma,
                   float* dbeta, float eps, int M, int C, int relu, int red_is_zero, void* stream) {
  if (M <= 0 || C <= 0) return 0;
  if ((C & 7) || C > 2048) {
    set_error("bn_backward: C=%d must be a multiple of 8, at most 2048", C);
    return 1;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  BnArgs a = bn_args(stats, gamma, nullptr, nullptr, nullptr, nullptr, eps, 0.f, M, C, relu);
  if (!red_is_zero) B2_CUDA(cudaMemsetAsync(red2c, 0, 2 * C * sizeof(float), s));
  // the reduce pass ends in 2C atomics per block: fewer, longer blocks than the pure streaming passes
  bn_bwd_reduce_kernel<<<bn_grid(M, C, 2), 256, 2 * C * sizeof(float), s>>>(static_cast<const __nv_bfloat16*>(dy_bf16), static_cast<const __nv_bfloat16*>(y_bf16),
                                                                            static_cast<const __nv_bfloat16*>(z_bf16), static_cast<__nv_bfloat16*>(g_bf16), red2c, a);
  B2_LAUNCH_CHECK("bn_bwd_reduce_kernel");
  bn_bwd_apply_kernel<<<bn_grid(M, C), 256, 0, s>>>(static_cast<const __nv_bfloat16*>(g_bf16), static_cast<const __nv_bfloat16*>(z_bf16), red2c,
                                                    static_cast<__nv_bfloat16*>(dz_bf16), dgamma, dbeta, a);
  B2_LAUNCH_CHECK("bn_bwd_apply_kernel");
  return 0;
}

int b2_avgpool_fwd(const void* x_bf16, void* out_bf16, int ldo, int B, int HW, int C, void* stream) {
  avgpool_fwd_kernel<<<stream_grid((int64_t)B * C, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(x_bf16), static_cast<__nv_bfloat16*>(out_bf16), B,
                                                                                                     HW, C, ldo);
  B2_LAUNCH_CHECK("avgpool_fwd_kernel");
  return 0;
}

int b2_avgpool_bwd(const void* dout_bf16, int ldo, void* dx_bf16, int B, int HW, int C, void* stream) {
  avgpool_bwd_kernel<<<stream_grid((int64_t)B * HW * C, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(dout_bf16), ldo,
                                                                                                          static_cast<__nv_bfloat16*>(dx_bf16), B, HW, C);
  B2_LAUNCH_CHECK("avgpool_bwd_kernel");
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

extern "C" int b2_repack_conv_weights(const float* slab, void* dst_bf16, const int64_t* table, int n_layers, int64_t max_elems, void* stream) {
  if (n_layers <= 0) return 0;
  int64_t blocks = (max_elems + 255) / 256;
  if (blocks > 592) blocks = 592;
  if (blocks < 1) blocks = 1;
  repack_conv_weights_kernel<<<dim3((unsigned)blocks, (unsigned)n_layers), 256, 0, static_cast<cudaStream_t>(stream)>>>(slab, static_cast<__nv_bfloat16*>(dst_bf16), table);
  B2_LAUNCH_CHECK("repack_conv_weights_kernel");
  return 0;
}
