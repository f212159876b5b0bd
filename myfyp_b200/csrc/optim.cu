// T6: fused multi-tensor optimizers over the flat slab (Adam, SGD+momentum) with the FedProx proximal term and
// the Scaffold control-variate term folded into the gradient, and the gradient slab cleared on the way out.
//   g' = g + wd*p + mu*(p - p_ref) + cv_scale*(c - c_i)
// Reference path: torch.optim.Adam(lr=1e-3) per Lightning fit (lightning_model.py:181-183) plus a Python
// callback for Scaffold (scaffold_callback.py:90-110); FedProx does not exist there.
#include <cuda_bf16.h>

#include "../../include/myfyp_b200.h"
#include "common.cuh"

namespace b2 {

struct FusedTerms {
  const float4* prox_ref;
  const float4* cv_global;
  const float4* cv_local;
  float prox_mu, cv_scale, wd;
};

__device__ __forceinline__ float4 effective_grad(const float4 g, const float4 p, const FusedTerms& t, int64_t i) {
  float4 r = g;
  if (t.wd != 0.f) { r.x = fmaf(t.wd, p.x, r.x); r.y = fmaf(t.wd, p.y, r.y); r.z = fmaf(t.wd, p.z, r.z); r.w = fmaf(t.wd, p.w, r.w); }
  if (t.prox_ref != nullptr) {
    const float4 q = t.prox_ref[i];
    r.x = fmaf(t.prox_mu, p.x - q.x, r.x); r.y = fmaf(t.prox_mu, p.y - q.y, r.y); r.z = fmaf(t.prox_mu, p.z - q.z, r.z); r.w = fmaf(t.prox_mu, p.w - q.w, r.w);
  }
  if (t.cv_global != nullptr) {
    const float4 c = t.cv_global[i], ci = t.cv_local[i];
    r.x = fmaf(t.cv_scale, c.x - ci.x, r.x); r.y = fmaf(t.cv_scale, c.y - ci.y, r.y); r.z = fmaf(t.cv_scale, c.z - ci.z, r.z); r.w = fmaf(t.cv_scale, c.w - ci.w, r.w);
  }
  return r;
}

// state = {step, 1-beta1^step, 1-beta2^step}: the step counter lives on the device so that a captured CUDA graph of a
// whole local epoch replays with the right bias corrections (a host-side step would be frozen into the graph).
__global__ void adam_prepare_kernel(float* __restrict__ state, float b1, float b2) {
  grid_dep_sync();
  const float step = state[0] + 1.f;
  state[0] = step;
  state[1] = 1.f - powf(b1, step);
  state[2] = 1.f - powf(b2, step);
}

__global__ void adam_advance_kernel(float* __restrict__ state, float n) {
  grid_dep_sync(); state[0] += n; }

__global__ void __launch_bounds__(256) adam_kernel(float4* __restrict__ p, float4* __restrict__ g, float4* __restrict__ m, float4* __restrict__ v, int64_t n4, float lr,
                                                   float b1, float b2, float eps, float inv_bc1, float inv_sqrt_bc2, const float* __restrict__ state, float step_offset,
                                                   __nv_bfloat16* __restrict__ p16, FusedTerms t, int zero_grad) {
  grid_dep_sync();
  if (state != nullptr) {
    if (step_offset > 0.f) {  // step = (steps finished before this graph replay, on the device) + (index inside the graph, baked)
      const float step = state[0] + step_offset;
      inv_bc1 = 1.f / (1.f - powf(b1, step));
      inv_sqrt_bc2 = rsqrtf(1.f - powf(b2, step));
    } else {
      inv_bc1 = 1.f / state[1];
      inv_sqrt_bc2 = rsqrtf(state[2]);
    }
  }
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pv = p[i], mv = m[i], vv = v[i];
    const float4 gv = effective_grad(g[i], pv, t, i);
#define B2_ADAM1(c)                                                       \
  mv.c = fmaf(b1, mv.c, (1.f - b1) * gv.c);                               \
  vv.c = fmaf(b2, vv.c, (1.f - b2) * gv.c * gv.c);                        \
  pv.c -= lr * (mv.c * inv_bc1) / (sqrtf(vv.c) * inv_sqrt_bc2 + eps);
    B2_ADAM1(x) B2_ADAM1(y) B2_ADAM1(z) B2_ADAM1(w)
#undef B2_ADAM1
    p[i] = pv; m[i] = mv; v[i] = vv;
    if (p16 != nullptr) {  // bf16 mirror of the parameters = the tensor-core operands of the next step (no separate cast kernels)
      __nv_bfloat162 lo = __floats2bfloat162_rn(pv.x, pv.y), hi = __floats2bfloat162_rn(pv.z, pv.w);
      uint2 w;
      w.x = *reinterpret_cast<uint32_t*>(&lo);
      w.y = *reinterpret_cast<uint32_t*>(&hi);
      reinterpret_cast<uint2*>(p16)[i] = w;
    }
    if (zero_grad) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

__global__ void __launch_bounds__(256) cast_slab_kernel(const float4* __restrict__ p, __nv_bfloat16* __restrict__ p16, int64_t n4) {
  grid_dep_sync();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 pv = p[i];
    __nv_bfloat162 lo = __floats2bfloat162_rn(pv.x, pv.y), hi = __floats2bfloat162_rn(pv.z, pv.w);
    uint2 w;
    w.x = *reinterpret_cast<uint32_t*>(&lo);
    w.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(p16)[i] = w;
  }
}

__global__ void __launch_bounds__(256) sgd_kernel(float4* __restrict__ p, float4* __restrict__ g, float4* __restrict__ buf, int64_t n4, float lr, float momentum,
                                                  FusedTerms t, int zero_grad) {
  grid_dep_sync();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pv = p[i];
    float4 gv = effective_grad(g[i], pv, t, i);
    if (momentum != 0.f) {
      float4 b = buf[i];
      b.x = fmaf(momentum, b.x, gv.x); b.y = fmaf(momentum, b.y, gv.y); b.z = fmaf(momentum, b.z, gv.z); b.w = fmaf(momentum, b.w, gv.w);
      buf[i] = b;
      gv = b;
    }
    pv.x -= lr * gv.x; pv.y -= lr * gv.y; pv.z -= lr * gv.z; pv.w -= lr * gv.w;
    p[i] = pv;
    if (zero_grad) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

static int check_slab(const void* a, const void* b, const void* c, const void* d, int64_t n) {
  if ((n & 3) || !aligned16(a) || !aligned16(b) || (c && !aligned16(c)) || (d && !aligned16(d))) {
    set_error("optimizer slabs must be 16-byte aligned with n %% 4 == 0 (the slab layout pads to 64 elements)");
    return 1;
  }
  return 0;
}

}  // namespace b2

using namespace b2;

extern "C" {

int b2_adam_prepare(float* state, float beta1, float beta2, void* stream) {
  launch_pdl(adam_prepare_kernel, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), state, beta1, beta2);
  B2_LAUNCH_CHECK("adam_prepare_kernel");
  return 0;
}

int b2_adam_advance(float* state, float n_steps, void* stream) {
  launch_pdl(adam_advance_kernel, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), state, n_steps);
  B2_LAUNCH_CHECK("adam_advance_kernel");
  return 0;
}

int b2_cast_slab(const float* p, void* p_bf16, int64_t n, void* stream) {
  if (n <= 0) return 0;
  if ((n & 3) || !aligned16(p) || (reinterpret_cast<uintptr_t>(p_bf16) & 7u)) {
    set_error("cast_slab: n %% 4 == 0 and aligned slabs required");
    return 1;
  }
  launch_pdl(cast_slab_kernel, dim3(stream_grid(n / 4, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), reinterpret_cast<const float4*>(p), static_cast<__nv_bfloat16*>(p_bf16), n / 4);
  B2_LAUNCH_CHECK("cast_slab_kernel");
  return 0;
}

int b2_adam_step(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps, float weight_decay, float bias_corr1, float bias_corr2,
                 const float* dev_state, float step_offset, void* p_bf16, float prox_mu, const float* prox_ref, const float* cv_global, const float* cv_local, float cv_scale,
                 int zero_grad, void* stream) {
  if (n <= 0) return 0;
  if (check_slab(p, g, m, v, n) || check_slab(p, p, prox_ref, cv_global, n) || check_slab(p, p, cv_local, nullptr, n)) return 1;
  FusedTerms t{reinterpret_cast<const float4*>(prox_ref), (cv_global && cv_local) ? reinterpret_cast<const float4*>(cv_global) : nullptr,
               reinterpret_cast<const float4*>(cv_local), prox_mu, cv_scale, weight_decay};
  launch_pdl(adam_kernel, dim3(stream_grid(n / 4, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), reinterpret_cast<float4*>(p), reinterpret_cast<float4*>(g), reinterpret_cast<float4*>(m),
                                                                                      reinterpret_cast<float4*>(v), n / 4, lr, beta1, beta2, eps, 1.f / bias_corr1,
                                                                                      rsqrtf(bias_corr2), dev_state, step_offset, static_cast<__nv_bfloat16*>(p_bf16), t,
                                                                                      zero_grad);
  B2_LAUNCH_CHECK("adam_kernel");
  return 0;
}

int b2_sgd_step(float* p, float* g, float* buf, int64_t n, float lr, float momentum, float weight_decay, float prox_mu, const float* prox_ref, const float* cv_global,
                const float* cv_local, float cv_scale, int zero_grad, void* stream) {
  if (n <= 0) return 0;
  if (check_slab(p, g, buf, prox_ref, n) || check_slab(p, p, cv_global, cv_local, n)) return 1;
  FusedTerms t{reinterpret_cast<const float4*>(prox_ref), (cv_global && cv_local) ? reinterpret_cast<const float4*>(cv_global) : nullptr,
               reinterpret_cast<const float4*>(cv_local), prox_mu, cv_scale, weight_decay};
  launch_pdl(sgd_kernel, dim3(stream_grid(n / 4, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), reinterpret_cast<float4*>(p), reinterpret_cast<float4*>(g), reinterpret_cast<float4*>(buf),
                                                                                     n / 4, lr, momentum, t, zero_grad);
  B2_LAUNCH_CHECK("sgd_kernel");
  return 0;
}

}  // extern "C"
