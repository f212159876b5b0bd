// Aggregation / diffusion kernels.
//
// Every source pointer may be local HBM or a peer GPU's slab (peer access / CUDA-IPC mapping): the loads are
// the transfer.  All kernels are single-pass streaming kernels over flat fp32 slabs: 128-bit loads, the K
// loads of one output vector issued back to back (memory-level parallelism), grid-stride over the slab.
#include "../../include/myfyp_b200.h"
#include "common.cuh"

namespace b2 {

// ------------------------------------------------------------------------------------------------ A1
// out = (accumulate ? out : 0) + sum_k w_k * src_k          reference math: fedavg.py:61-69
template <int UNROLL>
__global__ void __launch_bounds__(256) fedavg_kernel(float4* __restrict__ out, const __grid_constant__ PtrTable srcs, const __grid_constant__ WeightTable wts,
                                                     int K, int64_t n4, int accumulate) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 acc = accumulate ? out[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    int k = 0;
    for (; k + UNROLL <= K; k += UNROLL) {
      float4 v[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) v[u] = ld_stream(reinterpret_cast<const float4*>(srcs.p[k + u]) + i);
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const float w = wts.w[k + u];
        acc.x = fmaf(w, v[u].x, acc.x);
        acc.y = fmaf(w, v[u].y, acc.y);
        acc.z = fmaf(w, v[u].z, acc.z);
        acc.w = fmaf(w, v[u].w, acc.w);
      }
    }
    for (; k < K; ++k) {
      const float4 v = ld_stream(reinterpret_cast<const float4*>(srcs.p[k]) + i);
      const float w = wts.w[k];
      acc.x = fmaf(w, v.x, acc.x);
      acc.y = fmaf(w, v.y, acc.y);
      acc.z = fmaf(w, v.z, acc.z);
      acc.w = fmaf(w, v.w, acc.w);
    }
    st_stream(out + i, acc);
  }
}

__global__ void fedavg_scalar_kernel(float* __restrict__ out, const __grid_constant__ PtrTable srcs, const __grid_constant__ WeightTable wts, int K, int64_t begin,
                                     int64_t n, int accumulate) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float acc = accumulate ? out[i] : 0.f;
    for (int k = 0; k < K; ++k) acc = fmaf(wts.w[k], ld_stream1(srcs.p[k] + i), acc);
    out[i] = acc;
  }
}

// ------------------------------------------------------------------------------------------------ A3 / A4
// acc = beta*acc + w*src   (beta == 0: acc is not read => plain (scaled) pull copy)
__global__ void __launch_bounds__(256) axpby_kernel(float4* __restrict__ acc, const float4* __restrict__ src, float w, float beta, int64_t n4) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  // two independent vectors per trip
  for (; i + stride < n4; i += 2 * stride) {
    float4 a = ld_stream(src + i), b = ld_stream(src + i + stride);
    if (beta != 0.f) {
      const float4 x = acc[i], y = acc[i + stride];
      a = make_float4(fmaf(w, a.x, beta * x.x), fmaf(w, a.y, beta * x.y), fmaf(w, a.z, beta * x.z), fmaf(w, a.w, beta * x.w));
      b = make_float4(fmaf(w, b.x, beta * y.x), fmaf(w, b.y, beta * y.y), fmaf(w, b.z, beta * y.z), fmaf(w, b.w, beta * y.w));
    } else if (w != 1.f) {
      a = make_float4(w * a.x, w * a.y, w * a.z, w * a.w);
      b = make_float4(w * b.x, w * b.y, w * b.z, w * b.w);
    }
    st_stream(acc + i, a);
    st_stream(acc + i + stride, b);
  }
  for (; i < n4; i += stride) {
    float4 a = ld_stream(src + i);
    if (beta != 0.f) {
      const float4 x = acc[i];
      a = make_float4(fmaf(w, a.x, beta * x.x), fmaf(w, a.y, beta * x.y), fmaf(w, a.z, beta * x.z), fmaf(w, a.w, beta * x.w));
    } else if (w != 1.f) {
      a = make_float4(w * a.x, w * a.y, w * a.z, w * a.w);
    }
    st_stream(acc + i, a);
  }
}

__global__ void axpby_scalar_kernel(float* __restrict__ acc, const float* __restrict__ src, float w, float beta, int64_t begin, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) acc[i] = (beta != 0.f ? beta * acc[i] : 0.f) + w * ld_stream1(src + i);
}

// ------------------------------------------------------------------------------------------------ A5
// Coordinate-wise median / trimmed mean.  One coordinate per thread, its K values in registers, bitonic
// sorting network with compile-time indices (no local memory), +inf padding up to KPAD.
template <int KPAD>
__device__ __forceinline__ void bitonic_sort(float (&v)[KPAD]) {
#pragma unroll
  for (int size = 2; size <= KPAD; size <<= 1) {
#pragma unroll
    for (int strd = size >> 1; strd > 0; strd >>= 1) {
#pragma unroll
      for (int i = 0; i < KPAD; ++i) {
        const int j = i ^ strd;
        if (j > i) {
          const bool up = ((i & size) == 0);
          const float lo = fminf(v[i], v[j]), hi = fmaxf(v[i], v[j]);
          v[i] = up ? lo : hi;
          v[j] = up ? hi : lo;
        }
      }
    }
  }
}

template <int KPAD>
__device__ __forceinline__ float pick_median(const float (&v)[KPAD], int K, int trim) {
  float r = 0.f;
  if (trim < 0) {
    const int a = (K - 1) >> 1, b = K >> 1;
#pragma unroll
    for (int k = 0; k < KPAD; ++k) r += (k == a ? 0.5f * v[k] : 0.f) + (k == b ? 0.5f * v[k] : 0.f);
  } else {
    const int lo = trim, hi = K - trim;
#pragma unroll
    for (int k = 0; k < KPAD; ++k) r += (k >= lo && k < hi) ? v[k] : 0.f;
    r /= (float)(hi - lo);
  }
  return r;
}

// Four coordinates per thread (KPAD <= 16): one 16-byte load per model instead of four 4-byte ones - 4x the bytes in flight per
// thread - and four independent sorting networks for the scheduler to interleave.  Needs n % 4 == 0 and 16-byte aligned slabs.
template <int KPAD>
__global__ void __launch_bounds__(256) median4_kernel(float4* __restrict__ out, const __grid_constant__ PtrTable srcs, int K, int trim, int64_t n4) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float a[KPAD], b[KPAD], c[KPAD], d[KPAD];
#pragma unroll
    for (int k = 0; k < KPAD; ++k) {
      if (k < K) {
        const float4 t = ld_stream(reinterpret_cast<const float4*>(srcs.p[k]) + i);
        a[k] = t.x; b[k] = t.y; c[k] = t.z; d[k] = t.w;
      } else {
        a[k] = b[k] = c[k] = d[k] = __int_as_float(0x7f800000);
      }
    }
    bitonic_sort<KPAD>(a);
    bitonic_sort<KPAD>(b);
    bitonic_sort<KPAD>(c);
    bitonic_sort<KPAD>(d);
    out[i] = make_float4(pick_median<KPAD>(a, K, trim), pick_median<KPAD>(b, K, trim), pick_median<KPAD>(c, K, trim), pick_median<KPAD>(d, K, trim));
  }
}

// Two coordinates per thread (KPAD = 32): 8-byte loads, 64 values in registers.
template <int KPAD>
__global__ void __launch_bounds__(256) median2_kernel(float2* __restrict__ out, const __grid_constant__ PtrTable srcs, int K, int trim, int64_t n2) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
    float a[KPAD], b[KPAD];
#pragma unroll
    for (int k = 0; k < KPAD; ++k) {
      if (k < K) {
        float2 t;
        asm volatile("ld.global.L1::no_allocate.v2.f32 {%0, %1}, [%2];" : "=f"(t.x), "=f"(t.y) : "l"(reinterpret_cast<const float2*>(srcs.p[k]) + i));
        a[k] = t.x; b[k] = t.y;
      } else {
        a[k] = b[k] = __int_as_float(0x7f800000);
      }
    }
    bitonic_sort<KPAD>(a);
    bitonic_sort<KPAD>(b);
    out[i] = make_float2(pick_median<KPAD>(a, K, trim), pick_median<KPAD>(b, K, trim));
  }
}

template <int KPAD>
__global__ void __launch_bounds__(KPAD >= 64 ? 128 : 256) median_kernel(float* __restrict__ out, const __grid_constant__ PtrTable srcs, int K, int trim, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float v[KPAD];
#pragma unroll
    for (int k = 0; k < KPAD; ++k) v[k] = (k < K) ? ld_stream1(srcs.p[k] + i) : __int_as_float(0x7f800000);
    // (Batcher's odd-even merge network has 20 % fewer compare-exchanges at 32 / 64 inputs, but nvcc does not resolve its index
    // arithmetic at compile time: v[] went to local memory and K = 64 ran 4.6x slower than this bitonic network.)
    bitonic_sort<KPAD>(v);
    out[i] = pick_median<KPAD>(v, K, trim);
  }
}

// ------------------------------------------------------------------------------------------------ A8
// x += sum_k wy_k * dy_k ;  c += wc * sum_k dc_k            reference math: scaffold.py:70-111
__global__ void __launch_bounds__(256) scaffold_server_kernel(float4* __restrict__ x, float4* __restrict__ c, const __grid_constant__ PtrTable dys,
                                                              const __grid_constant__ PtrTable dcs, const __grid_constant__ WeightTable wy, float wc, int K,
                                                              int64_t n4) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 ax = x[i], ac = make_float4(0.f, 0.f, 0.f, 0.f);
    int k = 0;
    for (; k + 4 <= K; k += 4) {
      float4 vy[4], vc[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        vy[u] = ld_stream(reinterpret_cast<const float4*>(dys.p[k + u]) + i);
        vc[u] = ld_stream(reinterpret_cast<const float4*>(dcs.p[k + u]) + i);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float w = wy.w[k + u];
        ax.x = fmaf(w, vy[u].x, ax.x); ax.y = fmaf(w, vy[u].y, ax.y); ax.z = fmaf(w, vy[u].z, ax.z); ax.w = fmaf(w, vy[u].w, ax.w);
        ac.x += vc[u].x; ac.y += vc[u].y; ac.z += vc[u].z; ac.w += vc[u].w;
      }
    }
    for (; k < K; ++k) {
      const float4 vy = ld_stream(reinterpret_cast<const float4*>(dys.p[k]) + i);
      const float4 vc = ld_stream(reinterpret_cast<const float4*>(dcs.p[k]) + i);
      const float w = wy.w[k];
      ax.x = fmaf(w, vy.x, ax.x); ax.y = fmaf(w, vy.y, ax.y); ax.z = fmaf(w, vy.z, ax.z); ax.w = fmaf(w, vy.w, ax.w);
      ac.x += vc.x; ac.y += vc.y; ac.z += vc.z; ac.w += vc.w;
    }
    float4 cc = c[i];
    cc.x = fmaf(wc, ac.x, cc.x); cc.y = fmaf(wc, ac.y, cc.y); cc.z = fmaf(wc, ac.z, cc.z); cc.w = fmaf(wc, ac.w, cc.w);
    x[i] = ax;
    c[i] = cc;
  }
}

// ------------------------------------------------------------------------------------------------ T7
// adj = (x - y) * inv_klr ; dc = adj ; c_i += adj ; dy = y - x      reference: scaffold_callback.py:124-140
__global__ void __launch_bounds__(256) scaffold_finalize_kernel(float* __restrict__ c_i, float* __restrict__ dy, float* __restrict__ dc, const float* __restrict__ x,
                                                                const float* __restrict__ y, float inv_klr, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float xv = x[i], yv = y[i];
    const float adj = (xv - yv) * inv_klr;
    dc[i] = adj;
    c_i[i] += adj;
    dy[i] = yv - xv;
  }
}

static int fill_tables(PtrTable& pt, WeightTable* wt, const void* const* srcs, const float* w, int K, bool* all_aligned) {
  if (K < 1 || K > B2_MAX_K) {
    set_error("K=%d outside [1,%d]", K, B2_MAX_K);
    return 1;
  }
  for (int k = 0; k < K; ++k) {
    pt.p[k] = static_cast<const float*>(srcs[k]);
    if (wt) wt->w[k] = w[k];
    if (!aligned16(srcs[k])) *all_aligned = false;
  }
  for (int k = K; k < B2_MAX_K; ++k) {
    pt.p[k] = pt.p[0];
    if (wt) wt->w[k] = 0.f;
  }
  return 0;
}

}  // namespace b2

using namespace b2;

extern "C" {

int b2_fedavg(float* out, const void* const* srcs, const float* weights, int K, int64_t n, int accumulate, void* stream) {
  PtrTable pt;
  WeightTable wt;
  bool al = aligned16(out);
  if (fill_tables(pt, &wt, srcs, weights, K, &al)) return 1;
  if (n <= 0) return 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t n4 = al ? n / 4 : 0;
  if (n4 > 0) {
    fedavg_kernel<8><<<stream_grid(n4, 256), 256, 0, s>>>(reinterpret_cast<float4*>(out), pt, wt, K, n4, accumulate);
    B2_LAUNCH_CHECK("fedavg_kernel");
  }
  if (n4 * 4 < n) {
    fedavg_scalar_kernel<<<stream_grid(n - n4 * 4, 256), 256, 0, s>>>(out, pt, wt, K, n4 * 4, n, accumulate);
    B2_LAUNCH_CHECK("fedavg_scalar_kernel");
  }
  return 0;
}

int b2_accumulate(float* acc, const float* src, float w, float beta, int64_t n, void* stream) {
  if (n <= 0) return 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool al = aligned16(acc) && aligned16(src);
  const int64_t n4 = al ? n / 4 : 0;
  if (n4 > 0) {
    axpby_kernel<<<stream_grid((n4 + 1) / 2, 256), 256, 0, s>>>(reinterpret_cast<float4*>(acc), reinterpret_cast<const float4*>(src), w, beta, n4);
    B2_LAUNCH_CHECK("axpby_kernel");
  }
  if (n4 * 4 < n) {
    axpby_scalar_kernel<<<stream_grid(n - n4 * 4, 256), 256, 0, s>>>(acc, src, w, beta, n4 * 4, n);
    B2_LAUNCH_CHECK("axpby_scalar_kernel");
  }
  return 0;
}

int b2_pull_copy(float* dst, const float* src, int64_t n, void* stream) { return b2_accumulate(dst, src, 1.f, 0.f, n, stream); }

int b2_median(float* out, const void* const* srcs, int K, int trim, int64_t n, void* stream) {
  PtrTable pt;
  bool al = true;
  if (fill_tables(pt, nullptr, srcs, nullptr, K, &al)) return 1;
  if (trim >= 0 && 2 * trim >= K) {
    set_error("trim=%d leaves nothing of K=%d", trim, K);
    return 1;
  }
  if (n <= 0) return 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool vec4 = al && (n & 3) == 0 && aligned16(out);
#define B2_MEDIAN4_CASE(KP)                                                                                                            \
  if (vec4 && K <= KP) {                                                                                                               \
    median4_kernel<KP><<<stream_grid(n / 4, 256, 8), 256, 0, s>>>(reinterpret_cast<float4*>(out), pt, K, trim, n / 4);                 \
    B2_LAUNCH_CHECK("median4_kernel");                                                                                                 \
    return 0;                                                                                                                          \
  }
  B2_MEDIAN4_CASE(2)
  B2_MEDIAN4_CASE(4)
  B2_MEDIAN4_CASE(8)
  B2_MEDIAN4_CASE(16)
#undef B2_MEDIAN4_CASE
  if (vec4 && K <= 32) {
    median2_kernel<32><<<stream_grid(n / 2, 256, 8), 256, 0, s>>>(reinterpret_cast<float2*>(out), pt, K, trim, n / 2);
    B2_LAUNCH_CHECK("median2_kernel");
    return 0;
  }
#define B2_MEDIAN_CASE(KP, TH)                                                             \
  if (K <= KP) {                                                                           \
    median_kernel<KP><<<stream_grid(n, TH, 8), TH, 0, s>>>(out, pt, K, trim, n);           \
    B2_LAUNCH_CHECK("median_kernel");                                                      \
    return 0;                                                                              \
  }
  B2_MEDIAN_CASE(2, 256)
  B2_MEDIAN_CASE(4, 256)
  B2_MEDIAN_CASE(8, 256)
  B2_MEDIAN_CASE(16, 256)
  B2_MEDIAN_CASE(32, 256)
  B2_MEDIAN_CASE(64, 128)
#undef B2_MEDIAN_CASE
  return 1;
}

int b2_scaffold_server(float* x, float* c, const void* const* dys, const void* const* dcs, const float* wy, float wc, int K, int64_t n, void* stream) {
  PtrTable py, pc;
  WeightTable wt;
  bool al = aligned16(x) && aligned16(c);
  if (fill_tables(py, &wt, dys, wy, K, &al) || fill_tables(pc, nullptr, dcs, nullptr, K, &al)) return 1;
  if (!al || (n & 3)) {
    set_error("scaffold_server: slabs must be 16-byte aligned and n a multiple of 4 (slab layout guarantees both)");
    return 1;
  }
  if (n <= 0) return 0;
  scaffold_server_kernel<<<stream_grid(n / 4, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<float4*>(x), reinterpret_cast<float4*>(c), py, pc, wt, wc,
                                                                                                 K, n / 4);
  B2_LAUNCH_CHECK("scaffold_server_kernel");
  return 0;
}

int b2_scaffold_finalize(float* c_i, float* dy, float* dc, const float* x, const float* y, float inv_klr, int64_t n, void* stream) {
  if (n <= 0) return 0;
  scaffold_finalize_kernel<<<stream_grid(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(c_i, dy, dc, x, y, inv_klr, n);
  B2_LAUNCH_CHECK("scaffold_finalize_kernel");
  return 0;
}

}  // extern "C"
