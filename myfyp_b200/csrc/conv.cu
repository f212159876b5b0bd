// T4 / T5: convolution as implicit GEMM in two steps (patch gather + the tcgen05 GEMM of gemm.cu), max-pooling and the
// small data-movement kernels of the backward pass.  Activations are NHWC bf16, i.e. exactly the [pixels, channels]
// row-major matrices the GEMM reads and writes, so a conv layer is
//     forward   col = im2col(X)                      Y = relu(col W^T + b)           (GEMM epilogue: bias + ReLU, bf16 out)
//     wgrad     dW  = dZ^T col                       (both operands MN-major, split-K over the pixels, fp32 atomics into the gradient slab)
//     dgrad     dcol = dZ W                          dX = col2im(dcol)               (gather form: no atomics)
// The patch (K) index is ordered (kh, kw, c): a tap's channels are contiguous, exactly like the NHWC activations, so every
// gather / scatter below moves 16-byte vectors of 8 channels with fully coalesced warps (the PyTorch order (c, kh, kw) would
// put a tap's channels 2*KH*KW bytes apart).  The price is paid on the small side: weights are repacked fp32 [Cout, Cin, taps]
// -> bf16 [Cout, taps, Cin] once per step for all layers (b2_repack_conv_weights) and the weight-gradient GEMM's epilogue
// scatters back into PyTorch's layout (gemm.cu, perm_cin).  A fully connected layer after a feature map is the same thing
// with a kernel as large as the map (nn.Flatten's NCHW feature order falls out of the same repack).
// Reference path: torch conv2d / max_pool2d through autograd (cuDNN); the reference itself ships no CNN.
#include <cuda_bf16.h>

#include "../../include/myfyp_b200.h"
#include "common.cuh"

namespace b2 {

struct ConvGeom {
  int B, H, W, C, KH, KW, stride, pad, OH, OW;
};

// V consecutive channels (V = 8: one 16-byte vector, V = 1: scalar fallback for channel counts that are not multiples of 8)
template <int V>
struct ChanVec;
template <>
struct ChanVec<8> {
  uint4 raw;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { raw = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void store(__nv_bfloat16* p) const { *reinterpret_cast<uint4*>(p) = raw; }
  __device__ __forceinline__ void zero() { raw = make_uint4(0u, 0u, 0u, 0u); }
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
template <>
struct ChanVec<1> {
  __nv_bfloat16 raw;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { raw = *p; }
  __device__ __forceinline__ void store(__nv_bfloat16* p) const { *p = raw; }
  __device__ __forceinline__ void zero() { raw = __float2bfloat16_rn(0.f); }
  __device__ __forceinline__ void unpack(float (&f)[1]) const { f[0] = __bfloat162float(raw); }
  __device__ __forceinline__ void pack(const float (&f)[1]) { raw = __float2bfloat16_rn(f[0]); }
};

// one thread per (output pixel m, tap, channel group): col[m, tap * C + c] = x[pixel + tap, c]  (zero outside the image)
template <int V>
__global__ void __launch_bounds__(256) im2col_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, ConvGeom g, int ldc) {
  grid_dep_sync();
  const int KK = g.KH * g.KW, CG = g.C / V;
  const int64_t total = (int64_t)g.B * g.OH * g.OW * KK * CG;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % CG);
    const int t = (int)((i / CG) % KK);
    const int64_t m = i / ((int64_t)CG * KK);
    const int ow = (int)(m % g.OW), oh = (int)((m / g.OW) % g.OH), b = (int)(m / ((int64_t)g.OW * g.OH));
    const int ih = oh * g.stride - g.pad + t / g.KW, iw = ow * g.stride - g.pad + t % g.KW;
    ChanVec<V> v;
    if (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W) v.load(x + (((int64_t)b * g.H + ih) * g.W + iw) * g.C + cg * V);
    else v.zero();
    v.store(col + m * ldc + (int64_t)t * g.C + cg * V);
  }
}

// the input layer: NCHW fp32 images -> the same (tap, c) patch rows
__global__ void __launch_bounds__(256) im2col_nchw_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ col, ConvGeom g, int ldc) {
  grid_dep_sync();
  const int KK = g.KH * g.KW;
  const int64_t total = (int64_t)g.B * g.OH * g.OW * KK * g.C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % g.C);
    const int t = (int)((i / g.C) % KK);
    const int64_t m = i / ((int64_t)g.C * KK);
    const int ow = (int)(m % g.OW), oh = (int)((m / g.OW) % g.OH), b = (int)(m / ((int64_t)g.OW * g.OH));
    const int ih = oh * g.stride - g.pad + t / g.KW, iw = ow * g.stride - g.pad + t % g.KW;
    float v = 0.f;
    if (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W) v = x[(((int64_t)b * g.C + c) * g.H + ih) * g.W + iw];
    col[m * ldc + (int64_t)t * g.C + c] = __float2bfloat16_rn(v);
  }
}

// input layer with a short patch (KH*KW*C <= 32, row pitch exactly 32 elements, e.g. 3x3x3 -> 27): one thread builds the whole
// 64-byte patch row of an output pixel in registers and writes it with four 16-byte stores (neighbouring threads read
// neighbouring pixels of the same channel plane, so the fp32 loads coalesce too)
__global__ void __launch_bounds__(256) im2col_nchw_row32_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ col, ConvGeom g) {
  grid_dep_sync();
  const int64_t total = (int64_t)g.B * g.OH * g.OW;
  const int KK = g.KH * g.KW;
  for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < total; m += (int64_t)gridDim.x * blockDim.x) {
    const int ow = (int)(m % g.OW), oh = (int)((m / g.OW) % g.OH), b = (int)(m / ((int64_t)g.OW * g.OH));
    __align__(16) __nv_bfloat16 row[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) row[j] = __float2bfloat16_rn(0.f);
    const float* img = x + (int64_t)b * g.C * g.H * g.W;
#pragma unroll 1
    for (int t = 0; t < KK; ++t) {
      const int ih = oh * g.stride - g.pad + t / g.KW, iw = ow * g.stride - g.pad + t % g.KW;
      if (ih < 0 || ih >= g.H || iw < 0 || iw >= g.W) continue;
      for (int c = 0; c < g.C; ++c) row[t * g.C + c] = __float2bfloat16_rn(__ldg(img + ((int64_t)c * g.H + ih) * g.W + iw));
    }
    uint4* dst = reinterpret_cast<uint4*>(col + m * 32);
    const uint4* src = reinterpret_cast<const uint4*>(row);
#pragma unroll
    for (int j = 0; j < 4; ++j) dst[j] = src[j];
  }
}

// dx[b,h,w,c] = sum over the taps (kh,kw) whose output pixel exists of dcol[(b,oh,ow), (kh*KW + kw) * C + c]  (gather form, no atomics)
template <int V>
__global__ void __launch_bounds__(256) col2im_kernel(const __nv_bfloat16* __restrict__ dcol, __nv_bfloat16* __restrict__ dx, ConvGeom g, int ldc,
                                                     const __nv_bfloat16* __restrict__ add) {
  grid_dep_sync();
  const int CG = g.C / V;
  const int64_t total = (int64_t)g.B * g.H * g.W * CG;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % CG);
    const int64_t p = i / CG;
    const int w = (int)(p % g.W), h = (int)((p / g.W) % g.H), b = (int)(p / ((int64_t)g.W * g.H));
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
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
        ChanVec<V> v;
        float f[V];
        v.load(dcol + (((int64_t)b * g.OH + oh) * g.OW + ow) * ldc + (int64_t)(kh * g.KW + kw) * g.C + cg * V);
        v.unpack(f);
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] += f[j];
      }
    }
    ChanVec<V> o;
    if (add != nullptr) {  // second consumer of the same activation (residual / shortcut branch)
      float f[V];
      o.load(add + p * g.C + cg * V);
      o.unpack(f);
#pragma unroll
      for (int j = 0; j < V; ++j) acc[j] += f[j];
    }
    o.pack(acc);
    o.store(dx + p * g.C + cg * V);
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
  grid_dep_sync();
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
  grid_dep_sync();
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
  grid_dep_sync();
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
  grid_dep_sync();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * C; i += gridDim.x * blockDim.x) {
    const int c = i % C, b = i / C;
    float acc = 0.f;
    for (int p = 0; p < HW; ++p) acc += __bfloat162float(x[((int64_t)b * HW + p) * C + c]);
    out[(int64_t)b * ldo + c] = __float2bfloat16_rn(acc / (float)HW);
  }
}
__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dout, int ldo, __nv_bfloat16* __restrict__ dx, int B, int HW, int C) {
  grid_dep_sync();
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
  grid_dep_sync();
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
  grid_dep_sync();
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
template <int V>
__global__ void __launch_bounds__(256) pool_im2col_kernel(const __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ idx, __nv_bfloat16* __restrict__ col, ConvGeom g, int ldc) {
  grid_dep_sync();
  const int H2 = g.H * 2, W2 = g.W * 2, C = g.C, CG = C / V;
  const int64_t total = (int64_t)g.B * g.H * g.W * CG;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % CG);
    const int64_t q = i / CG;
    const int pw = (int)(q % g.W), ph = (int)((q / g.W) % g.H), b = (int)(q / ((int64_t)g.W * g.H));
    const __nv_bfloat16* src = y + (((int64_t)b * H2 + 2 * ph) * W2 + 2 * pw) * C + cg * V;
    const int64_t offs[4] = {0, C, (int64_t)W2 * C, (int64_t)W2 * C + C};
    float best[V];
    uint8_t bi[V];
    ChanVec<V> v;
    v.load(src);
    v.unpack(best);
#pragma unroll
    for (int j = 0; j < V; ++j) bi[j] = 0;
#pragma unroll
    for (int k = 1; k < 4; ++k) {
      float f[V];
      v.load(src + offs[k]);
      v.unpack(f);
#pragma unroll
      for (int j = 0; j < V; ++j)
        if (f[j] > best[j]) {
          best[j] = f[j];
          bi[j] = (uint8_t)k;
        }
    }
    v.pack(best);  // exact: the maxima are bf16 values
    if constexpr (V == 8) {
      uint2 packed;
      packed.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
      packed.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
      *reinterpret_cast<uint2*>(idx + q * C + cg * 8) = packed;
    } else {
      idx[q * C + cg] = bi[0];
    }
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
        v.store(col + (((int64_t)b * g.OH + oh) * g.OW + ow) * ldc + (int64_t)(kh * g.KW + kw) * C + cg * V);
      }
    }
  }
}

// Fused backward head of a conv block: col2im of the NEXT layer's dcol (gather over its taps) -> gradient of the pooled
// activation -> routed to the remembered max position, times the ReLU mask -> dZ [B, 2H, 2W, C], plus the bias gradient
// (column sums of dZ) reduced per block in shared memory.  One pass, no dP / P buffers.
template <int V>
__global__ void __launch_bounds__(256) unpool_col2im_kernel(const __nv_bfloat16* __restrict__ dcol, const uint8_t* __restrict__ idx, const __nv_bfloat16* __restrict__ y,
                                                            __nv_bfloat16* __restrict__ dz, float* __restrict__ bias_grad, ConvGeom g, int ldc) {
  grid_dep_sync();
  extern __shared__ float s_db[];  // C floats
  const int H2 = g.H * 2, W2 = g.W * 2, C = g.C, CG = C / V;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s_db[c] = 0.f;
  __syncthreads();
  const int64_t total = (int64_t)g.B * g.H * g.W * CG;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % CG);
    const int64_t q = i / CG;
    const int pw = (int)(q % g.W), ph = (int)((q / g.W) % g.H), b = (int)(q / ((int64_t)g.W * g.H));
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
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
        ChanVec<V> v;
        float f[V];
        v.load(dcol + (((int64_t)b * g.OH + oh) * g.OW + ow) * ldc + (int64_t)(kh * g.KW + kw) * C + cg * V);
        v.unpack(f);
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] += f[j];
      }
    }
    uint8_t win[V];
    if constexpr (V == 8) {
      const uint2 packed = *reinterpret_cast<const uint2*>(idx + q * C + cg * 8);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        win[j] = (packed.x >> (8 * j)) & 0xFF;
        win[4 + j] = (packed.y >> (8 * j)) & 0xFF;
      }
    } else {
      win[0] = idx[q * C + cg];
    }
    ChanVec<V> gq;
    gq.pack(acc);  // the value the unfused path stores as dP
    gq.unpack(acc);
    const int64_t base = (((int64_t)b * H2 + 2 * ph) * W2 + 2 * pw) * C + cg * V;
    const int64_t offs[4] = {0, C, (int64_t)W2 * C, (int64_t)W2 * C + C};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      ChanVec<V> yv, o;
      float yf[V], of[V];
      yv.load(y + base + offs[k]);
      yv.unpack(yf);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const bool live = (win[j] == k) && yf[j] > 0.f;
        of[j] = live ? acc[j] : 0.f;
        if (live && bias_grad != nullptr) atomicAdd(&s_db[cg * V + j], acc[j]);
      }
      o.pack(of);
      o.store(dz + base + offs[k]);
    }
  }
  __syncthreads();
  if (bias_grad != nullptr)
    for (int c = threadIdx.x; c < C; c += blockDim.x)
      if (s_db[c] != 0.f) atomicAdd(bias_grad + c, s_db[c]);
}

// out[c] += sum_m z[m, c]   (bias gradient of a conv layer): each block reduces a slab of rows in shared memory
__global__ void __launch_bounds__(256) colsum_bf16_kernel(const __nv_bfloat16* __restrict__ z, int M, int C, int ld, float* __restrict__ out, int rows_per_block) {
  grid_dep_sync();
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

// conv weights fp32 [cout][cin][taps] -> bf16 [cout][taps][cin] (row pitch ld >= taps * cin); blockIdx.y = layer
__global__ void repack_conv_weights_kernel(const float* __restrict__ slab, __nv_bfloat16* __restrict__ dst, const int64_t* __restrict__ table) {
  grid_dep_sync();
  const int64_t* t = table + 6 * blockIdx.y;
  const float* src = slab + t[0];
  __nv_bfloat16* out = dst + t[1];
  const int64_t cin = t[3], taps = t[4], ld = t[5], n = t[2] * cin * taps;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c = i % cin, r = i / cin, tp = r % taps, co = r / taps;
    out[co * ld + tp * cin + c] = __float2bfloat16_rn(src[(co * cin + c) * taps + tp]);
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
  const int64_t total = (int64_t)B * g.OH * g.OW * KH * KW * C;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  __nv_bfloat16* col = static_cast<__nv_bfloat16*>(col_bf16);
  const bool vec = (C % 8 == 0) && (ldc % 8 == 0) && aligned16(x) && aligned16(col_bf16);
  if (x_is_nchw_f32 && KH * KW * C <= 32 && ldc == 32 && aligned16(col_bf16))
    launch_pdl(im2col_nchw_row32_kernel, dim3(stream_grid((int64_t)B * g.OH * g.OW, 256)), dim3(256), 0, s, static_cast<const float*>(x), col, g);
  else if (x_is_nchw_f32) launch_pdl(im2col_nchw_kernel, dim3(stream_grid(total, 256)), dim3(256), 0, s, static_cast<const float*>(x), col, g, ldc);
  else if (vec) launch_pdl(im2col_kernel<8>, dim3(stream_grid(total / 8, 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x), col, g, ldc);
  else launch_pdl(im2col_kernel<1>, dim3(stream_grid(total, 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x), col, g, ldc);
  B2_LAUNCH_CHECK("im2col_kernel");
  return 0;
}

int b2_col2im(const void* dcol_bf16, int ldc, void* dx_bf16, const void* add_bf16, int B, int H, int W, int C, int KH, int KW, int stride, int pad, void* stream) {
  ConvGeom g{B, H, W, C, KH, KW, stride, pad, (H + 2 * pad - KH) / stride + 1, (W + 2 * pad - KW) / stride + 1};
  const int64_t total = (int64_t)B * H * W * C;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16 *dcol = static_cast<const __nv_bfloat16*>(dcol_bf16), *add = static_cast<const __nv_bfloat16*>(add_bf16);
  __nv_bfloat16* dx = static_cast<__nv_bfloat16*>(dx_bf16);
  const bool vec = (C % 8 == 0) && (ldc % 8 == 0) && aligned16(dcol_bf16) && aligned16(dx_bf16) && aligned16(add_bf16);
  if (vec) launch_pdl(col2im_kernel<8>, dim3(stream_grid(total / 8, 256)), dim3(256), 0, s, dcol, dx, g, ldc, add);
  else launch_pdl(col2im_kernel<1>, dim3(stream_grid(total, 256)), dim3(256), 0, s, dcol, dx, g, ldc, add);
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
  launch_pdl(bn_apply_kernel, dim3(bn_grid(M, C)), dim3(256), 0, static_cast<cudaStream_t>(stream), 
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
  launch_pdl(bn_bwd_reduce_kernel, dim3(bn_grid(M, C, 2)), dim3(256), 2 * C * sizeof(float), s, static_cast<const __nv_bfloat16*>(dy_bf16), static_cast<const __nv_bfloat16*>(y_bf16),
                                                                            static_cast<const __nv_bfloat16*>(z_bf16), static_cast<__nv_bfloat16*>(g_bf16), red2c, a);
  B2_LAUNCH_CHECK("bn_bwd_reduce_kernel");
  launch_pdl(bn_bwd_apply_kernel, dim3(bn_grid(M, C)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(g_bf16), static_cast<const __nv_bfloat16*>(z_bf16), red2c,
                                                    static_cast<__nv_bfloat16*>(dz_bf16), dgamma, dbeta, a);
  B2_LAUNCH_CHECK("bn_bwd_apply_kernel");
  return 0;
}

int b2_avgpool_fwd(const void* x_bf16, void* out_bf16, int ldo, int B, int HW, int C, void* stream) {
  launch_pdl(avgpool_fwd_kernel, dim3(stream_grid((int64_t)B * C, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x_bf16), static_cast<__nv_bfloat16*>(out_bf16), B,
                                                                                                     HW, C, ldo);
  B2_LAUNCH_CHECK("avgpool_fwd_kernel");
  return 0;
}

int b2_avgpool_bwd(const void* dout_bf16, int ldo, void* dx_bf16, int B, int HW, int C, void* stream) {
  launch_pdl(avgpool_bwd_kernel, dim3(stream_grid((int64_t)B * HW * C, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(dout_bf16), ldo,
                                                                                                          static_cast<__nv_bfloat16*>(dx_bf16), B, HW, C);
  B2_LAUNCH_CHECK("avgpool_bwd_kernel");
  return 0;
}

int b2_maxpool2_fwd(const void* y_bf16, void* p_bf16, void* idx_u8, int B, int H, int W, int C, void* stream) {
  const int64_t total = (int64_t)B * (H / 2) * (W / 2) * C;
  launch_pdl(maxpool2_fwd_kernel, dim3(stream_grid(total, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(y_bf16), static_cast<__nv_bfloat16*>(p_bf16),
                                                                                             static_cast<uint8_t*>(idx_u8), B, H, W, C);
  B2_LAUNCH_CHECK("maxpool2_fwd_kernel");
  return 0;
}

int b2_maxpool2_relu_bwd(const void* dp_bf16, const void* idx_u8, const void* y_bf16, void* dz_bf16, int B, int H, int W, int C, void* stream) {
  const int64_t total = (int64_t)B * H * W * C;
  launch_pdl(maxpool2_relu_bwd_kernel, dim3(stream_grid(total, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), 
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
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* y = static_cast<const __nv_bfloat16*>(y_bf16);
  uint8_t* idx = static_cast<uint8_t*>(idx_u8);
  __nv_bfloat16* col = static_cast<__nv_bfloat16*>(col_bf16);
  const bool vec = (C % 8 == 0) && (ldc % 8 == 0) && aligned16(y_bf16) && aligned16(col_bf16) && aligned16(idx_u8);
  if (vec) launch_pdl(pool_im2col_kernel<8>, dim3(stream_grid(total / 8, 256)), dim3(256), 0, s, y, idx, col, g, ldc);
  else launch_pdl(pool_im2col_kernel<1>, dim3(stream_grid(total, 256)), dim3(256), 0, s, y, idx, col, g, ldc);
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
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16 *dcol = static_cast<const __nv_bfloat16*>(dcol_bf16), *y = static_cast<const __nv_bfloat16*>(y_bf16);
  const uint8_t* idx = static_cast<const uint8_t*>(idx_u8);
  __nv_bfloat16* dz = static_cast<__nv_bfloat16*>(dz_bf16);
  const bool vec = (C % 8 == 0) && (ldc % 8 == 0) && aligned16(dcol_bf16) && aligned16(y_bf16) && aligned16(dz_bf16) && aligned16(idx_u8);
  if (vec) launch_pdl(unpool_col2im_kernel<8>, dim3(stream_grid(total / 8, 256, 2)), dim3(256), C * sizeof(float), s, dcol, idx, y, dz, bias_grad, g, ldc);
  else launch_pdl(unpool_col2im_kernel<1>, dim3(stream_grid(total, 256, 2)), dim3(256), C * sizeof(float), s, dcol, idx, y, dz, bias_grad, g, ldc);
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
  launch_pdl(colsum_bf16_kernel, dim3(blocks), dim3(256), C * sizeof(float), static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(z_bf16), M, C, ld, out, rows_per_block);
  B2_LAUNCH_CHECK("colsum_bf16_kernel");
  return 0;
}

}  // extern "C"

extern "C" int b2_repack_conv_weights(const float* slab, void* dst_bf16, const int64_t* table, int n_layers, int64_t max_elems, void* stream) {
  if (n_layers <= 0) return 0;
  int64_t blocks = (max_elems + 255) / 256;
  if (blocks > 592) blocks = 592;
  if (blocks < 1) blocks = 1;
  launch_pdl(repack_conv_weights_kernel, dim3(dim3((unsigned)blocks, (unsigned)n_layers)), dim3(256), 0, static_cast<cudaStream_t>(stream), slab, static_cast<__nv_bfloat16*>(dst_bf16), table);
  B2_LAUNCH_CHECK("repack_conv_weights_kernel");
  return 0;
}
