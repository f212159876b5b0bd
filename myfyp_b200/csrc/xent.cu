// T3 + T8: fused log-softmax + NLL loss, its gradient, and the argmax-accuracy count - one kernel, one warp per
// row (C is small: 10 for MNIST/CIFAR).  Reference path: F.cross_entropy on log_softmax output
// (lightning_model.py:178,189) + torchmetrics Accuracy (:204) = several kernels and a softmax round trip.
//   loss_out   += scale * sum_rows ( logsumexp(row) - row[label] )        (atomic, one add per CTA)
//   dlogits[r] = scale * (softmax(row) - onehot(label))                   (optional)
//   correct    += #rows with argmax == label                              (optional; first max wins, like torch)
#include <cuda_bf16.h>

#include "../../include/myfyp_b200.h"
#include "common.cuh"

namespace b2 {

__global__ void __launch_bounds__(256) xent_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, float* __restrict__ loss_out,
                                                   float* __restrict__ dlogits, float* __restrict__ correct, int B, int C, float scale) {
  grid_dep_sync();
  __shared__ float s_loss[8], s_corr[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float wl = 0.f, wc = 0.f;
  for (int row = blockIdx.x * 8 + warp; row < B; row += gridDim.x * 8) {
    const float* x = logits + (int64_t)row * C;
    const int lab = (int)labels[row];
    float mx = -INFINITY;
    int amax = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      const float v = x[c];
      if (v > mx) { mx = v; amax = c; }
    }
    // warp arg-max (lowest index among equal maxima)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, amax, o);
      if (om > mx || (om == mx && oa < amax)) { mx = om; amax = oa; }
    }
    float se = 0.f;
    for (int c = lane; c < C; c += 32) se += __expf(x[c] - mx);
    se = warp_sum(se);
    const float lse = mx + __logf(se);
    if (dlogits != nullptr) {
      const float inv = scale / se;
      float* d = dlogits + (int64_t)row * C;
      for (int c = lane; c < C; c += 32) d[c] = __expf(x[c] - mx) * inv - (c == lab ? scale : 0.f);
    }
    if (lane == 0) {
      wl += lse - x[lab];
      wc += (amax == lab) ? 1.f : 0.f;
    }
  }
  if (lane == 0) { s_loss[warp] = wl; s_corr[warp] = wc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float tl = 0.f, tc = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { tl += s_loss[w]; tc += s_corr[w]; }
    if (tl != 0.f) atomicAdd(loss_out, tl * scale);
    if (correct != nullptr && tc != 0.f) atomicAdd(correct, tc);
  }
}

// Training variant, ONE block: loss written (not accumulated), dlogits emitted directly as the bf16 operand of the backward
// GEMMs (row pitch ld16), bias gradient of the last layer (column sums of dlogits) added - replaces memset + xent +
// cast + column-sum (4 launches) by one.  B <= a few thousand rows (a batch).
__global__ void __launch_bounds__(256) xent_train_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, float* __restrict__ loss_out,
                                                         __nv_bfloat16* __restrict__ dl16, int ld16, float* __restrict__ bias_grad, int B, int C, float scale) {
  grid_dep_sync();
  extern __shared__ float s_mem[];  // C column sums + 8 warp losses
  float* s_db = s_mem;
  float* s_loss = s_mem + C;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s_db[c] = 0.f;
  __syncthreads();
  float wl = 0.f;
  for (int row = warp; row < B; row += 8) {
    const float* x = logits + (int64_t)row * C;
    const int lab = (int)labels[row];
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, x[c]);
    mx = warp_max(mx);
    float se = 0.f;
    for (int c = lane; c < C; c += 32) se += __expf(x[c] - mx);
    se = warp_sum(se);
    const float inv = scale / se;
    for (int c = lane; c < C; c += 32) {
      const float d = __expf(x[c] - mx) * inv - (c == lab ? scale : 0.f);
      dl16[(int64_t)row * ld16 + c] = __float2bfloat16_rn(d);
      atomicAdd(&s_db[c], d);
    }
    if (lane == 0) wl += mx + __logf(se) - x[lab];
  }
  if (lane == 0) s_loss[warp] = wl;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += s_loss[w];
    *loss_out = t * scale;
  }
  if (bias_grad != nullptr)
    for (int c = threadIdx.x; c < C; c += blockDim.x) bias_grad[c] += s_db[c];
}

// confusion[label * C + argmax] += 1 : the sufficient statistic for accuracy, per-class precision / recall / F1
// (the fork author's extended MLP logs test_f1 / test_precision / test_recall, mlp_pytorch.txt)
__global__ void __launch_bounds__(256) confusion_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, float* __restrict__ conf, int B, int C) {
  grid_dep_sync();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int row = blockIdx.x * 8 + warp; row < B; row += gridDim.x * 8) {
    const float* x = logits + (int64_t)row * C;
    float mx = -INFINITY;
    int amax = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      const float v = x[c];
      if (v > mx) { mx = v; amax = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, amax, o);
      if (om > mx || (om == mx && oa < amax)) { mx = om; amax = oa; }
    }
    if (lane == 0) atomicAdd(conf + (int64_t)labels[row] * C + amax, 1.f);
  }
}

}  // namespace b2

extern "C" int b2_xent_train(const float* logits, const int64_t* labels, float* loss_out, void* dlogits_bf16, int ld16, float* bias_grad, int B, int C, float scale,
                             void* stream) {
  if (B <= 0 || C <= 0) return 0;
  if (C > 8192) {
    b2::set_error("xent_train: C=%d too wide for the single-block variant", C);
    return 1;
  }
  b2::launch_pdl(b2::xent_train_kernel, dim3(1), dim3(256), (C + 8) * sizeof(float), static_cast<cudaStream_t>(stream), logits, labels, loss_out, static_cast<__nv_bfloat16*>(dlogits_bf16), ld16, bias_grad,
                                                                                                B, C, scale);
  B2_LAUNCH_CHECK("xent_train_kernel");
  return 0;
}

extern "C" int b2_confusion(const float* logits, const int64_t* labels, float* conf, int B, int C, void* stream) {
  if (B <= 0 || C <= 0) return 0;
  int grid = (B + 7) / 8;
  const int cap = b2::sm_count() * 8;
  if (grid > cap) grid = cap;
  b2::launch_pdl(b2::confusion_kernel, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), logits, labels, conf, B, C);
  B2_LAUNCH_CHECK("confusion_kernel");
  return 0;
}

extern "C" int b2_xent_fwd_bwd(const float* logits, const int64_t* labels, float* loss_out, float* dlogits, float* correct, int B, int C, float scale, void* stream) {
  if (B <= 0 || C <= 0) return 0;
  int grid = (B + 7) / 8;
  const int cap = b2::sm_count() * 8;
  if (grid > cap) grid = cap;
  b2::launch_pdl(b2::xent_kernel, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), logits, labels, loss_out, dlogits, correct, B, C, scale);
  B2_LAUNCH_CHECK("xent_kernel");
  return 0;
}
