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
// Structure (persistent, warp specialised; 128 x BN output tiles, BN = 64 or 256, one CTA per SM walking the tile list):
//   warp 0     : TMA producer  - cp.async.bulk.tensor.2d (128B-swizzled boxes) into a 4-6 stage ring, mbarrier expect_tx
//   warp 1     : MMA issuer    - one elected thread, tcgen05.mma kind::f16 M=128 N=BN K=16, two accumulators in TMEM
//   warps 2..5 : epilogue      - tcgen05.ld (32 lanes x 32 columns per warp), bias / mask / ReLU, stores, column sums;
//                                drains accumulator i while the tensor core fills accumulator i+1
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
//
// Implicit-GEMM convolution (template parameter CONV; entry point b2_conv_gemm; stride 1, "same" padding, channels % 64 == 0).
// The patch matrix is never materialised: the producer warp fetches, per (filter tap, 64-channel block), a 4-D TMA box
// {64 channels, tile_w, tile_h, tile_n} straight out of the NHWC activation tensor, shifted by the tap offset - TMA's
// out-of-bounds zero fill IS the padding - and the box lands in shared memory in exactly the 128B-swizzled [pixels, 64]
// layout the UMMA descriptors already describe.  Weights are kept as bf16 [Cout, (kh, kw, cin)] (b2_repack_conv_weights).
//   CONV=1  fwd    z[pix, co]  = sum_{tap, ci} x[pix + tap, ci]  w[co, tap, ci]      A = 4-D boxes of x,  B = w K-major
//           dgrad  dx[pix, ci] = sum_{tap, co} dz[pix - tap, co] w[co, tap, ci]      A = 4-D boxes of dz, B = w MN-major
//   CONV=2  wgrad  dw[co, tap, ci] = sum_pix dz[pix, co] x[pix + tap, ci]            A = dz MN-major, B = 4-D boxes of x (MN-major),
//                  split-K over pixel blocks; the epilogue scatters into PyTorch's [Cout, Cin, KH, KW] gradient layout
#include <cuda.h>
#include <cuda_bf16.h>

#include "../../include/myfyp_b200.h"
#include "common.cuh"
#include "tcgen05.cuh"

namespace b2 {

constexpr int GM_BM = 128, GM_BK = 64;
constexpr int GM_A_BYTES = GM_BM * GM_BK * 2;  // 16 KB
constexpr int GM_THREADS = 192;
constexpr int GM_ACC_STAGES = 2;  // TMEM accumulators: the epilogue drains one while the tensor core fills the other

template <int BN>
struct GemmCfg {
  static constexpr int B_BYTES = BN * GM_BK * 2;
  static constexpr int STAGE_BYTES = GM_A_BYTES + B_BYTES;
  static constexpr int STAGES = 4;
  static constexpr int CTAS_PER_SM = (BN >= 256) ? 1 : 2;  // small tiles: two CTAs per SM double the epilogue warps per SM
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024;
  static constexpr uint32_t TMEM_COLS = (uint32_t)(GM_ACC_STAGES * BN) < 32u ? 32u : (uint32_t)(GM_ACC_STAGES * BN);
};

struct GemmEpilogue {
  float* out_f32;
  __nv_bfloat16* out_bf16;
  __nv_bfloat16* out_bf16_t;
  const float* bias;
  const __nv_bfloat16* mask;
  float* colsum;
  float* colsumsq;
  int ldo, ldo16, ldt, ldm;
  int relu, accumulate;
  const __nv_bfloat16* add16;  // optional bf16 [M, ldadd] added to the result (second consumer of an activation: residual branch gradient)
  int ldadd;
  long long* trace;  // debug (tools/gemm_trace.py): per-phase %globaltimer stamps of CTA 0, nullptr in production
  int perm_cin, perm_taps;  // wgrad of a convolution: column j = tap * cin + c is stored at c * taps + tap (PyTorch's [Cout, Cin, KH, KW])
};

// Geometry of the 4-D activation operand of an implicit-GEMM convolution (stride 1, pad = k / 2)
struct ConvGeom {
  int cblocks;    // channels of the 4-D tensor / 64
  int kw;         // filter width (taps = kh * kw)
  int off, sign;  // input coordinate = output coordinate + off + sign * tap index   (fwd / wgrad: -pad, +1; dgrad: +pad, -1)
  int hw_tiles;   // pixel tiles per image (tile = tile_h full rows), or 0 when a tile spans tile_n whole images
  int tile_h, tile_n;
  int ncols;      // CONV=1 with B MN-major (dgrad): columns per tap of the weight matrix (= cin); CONV=2: cin
};

__device__ __forceinline__ void trace_stamp(long long* trace, int slot) {
  if (trace != nullptr && blockIdx.x == 0) {
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    trace[slot] = t;
  }
}

__device__ __forceinline__ void conv_pixel_tile(const ConvGeom& cg, int pt, int& n, int& h0) {
  if (cg.hw_tiles > 0) {
    n = pt / cg.hw_tiles;
    h0 = (pt % cg.hw_tiles) * cg.tile_h;
  } else {
    n = pt * cg.tile_n;
    h0 = 0;
  }
}

// 32 consecutive bf16 of one row -> fp32 registers: four 16-byte loads when the row segment is complete and 16-byte aligned,
// else independent scalar loads (all issued before any use); columns >= n_valid read as 0
__device__ __forceinline__ void load_row32_bf16(const __nv_bfloat16* p, bool vec_ok, int n_valid, float (&out)[32]) {
  if (vec_ok && (reinterpret_cast<uintptr_t>(p) & 15u) == 0) {
    uint4 raw[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) raw[j] = *reinterpret_cast<const uint4*>(p + 8 * j);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw[j]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 t = __bfloat1622float2(h[i]);
        out[8 * j + 2 * i] = t.x;
        out[8 * j + 2 * i + 1] = t.y;
      }
    }
  } else {
    __nv_bfloat16 raw[32];
#pragma unroll
    for (int e = 0; e < 32; ++e) raw[e] = e < n_valid ? p[e] : __float2bfloat16_rn(0.f);
#pragma unroll
    for (int e = 0; e < 32; ++e) out[e] = __bfloat162float(raw[e]);
  }
}

// v[e] = value of column e in this lane's row; returns the sum over the 32 lanes (rows) of column `lane` (of squares when `sq`)
__device__ __forceinline__ float warp_colsum32(const float (&v)[32], int lane, bool sq) {
  float w[16];
  {
    const bool hi = lane & 16;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float a = v[i], b = v[i + 16];
      if (sq) {
        a *= a;
        b *= b;
      }
      w[i] = (hi ? b : a) + __shfl_xor_sync(0xffffffffu, hi ? a : b, 16);
    }
  }
#pragma unroll
  for (int half = 8; half >= 1; half >>= 1) {
    const bool hi = lane & half;
#pragma unroll
    for (int i = 0; i < half; ++i) w[i] = (hi ? w[i + half] : w[i]) + __shfl_xor_sync(0xffffffffu, hi ? w[i] : w[i + half], half);
  }
  return w[0];
}

__device__ __forceinline__ uint64_t smem_desc_sw128_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
  // MN-major canonical SWIZZLE_128B: rows are K (128 bytes = 64 MN elements each), 8-row groups 1024 bytes apart (SBO),
  // further 64-element MN blocks `lbo_bytes` apart (LBO)
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// Persistent, warp-specialised: every CTA walks work items w = blockIdx.x, blockIdx.x + gridDim.x, ...  where
// w = (output tile, k-split).  Three pipelines: smem ring (TMA -> MMA), TMEM accumulator ring (MMA -> epilogue), tile loop.
template <int BN, bool A_MN, bool B_MN, int CONV = 0>
__global__ void __launch_bounds__(GM_THREADS, GemmCfg<BN>::CTAS_PER_SM) gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, int M, int N, int K,
                                                                   int tiles_m, int tiles_n, int splits, int kb_per_split, GemmEpilogue ep, ConvGeom cg) {
  using Cfg = GemmCfg<BN>;
  if (threadIdx.x == 0) trace_stamp(ep.trace, 0);  // kernel entry
  grid_dep_launch();  // the next kernel of the step may become resident and run its own prologue while this one works
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[Cfg::STAGES], empty_bar[Cfg::STAGES], acc_full[GM_ACC_STAGES], acc_empty[GM_ACC_STAGES];
  __shared__ uint32_t tmem_base_s;

  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int total_k = (K + GM_BK - 1) / GM_BK;
  const int n_work = tiles_m * tiles_n * splits;
  const bool split = splits > 1;

  if (tid == 0) {
    tc::tma_prefetch_desc(&tmap_a);
    tc::tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < Cfg::STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < GM_ACC_STAGES; ++a) {
      tc::mbar_init(&acc_full[a], 1);
      tc::mbar_init(&acc_empty[a], 4);  // one arrival per epilogue warp
    }
    tc::mbar_fence_init();
  }
  if (warp == 1) tc::tmem_alloc<Cfg::TMEM_COLS>(&tmem_base_s);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = tmem_base_s;
  if (threadIdx.x == 0) trace_stamp(ep.trace, 1);  // prologue done
  grid_dep_wait();  // barriers, TMEM and tensor maps are ready: now wait for the producer of our operands to complete

  auto work_coords = [&](int w, int& m0, int& n0, int& kb0, int& nk) {
    const int z = w % splits;
    const int t = w / splits;
    m0 = (t / tiles_n) * GM_BM;  // consecutive work items share the A row block
    n0 = (t % tiles_n) * BN;
    kb0 = z * kb_per_split;
    nk = total_k - kb0;
    nk = nk < 0 ? 0 : (nk < kb_per_split ? nk : kb_per_split);
  };

  if (warp == 0) {
    // ================================ TMA producer
    if (tc::elect_one()) {
      uint32_t it = 0;
      for (int w = blockIdx.x; w < n_work; w += gridDim.x) {
        int m0, n0, kb0, nk;
        work_coords(w, m0, n0, kb0, nk);
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % Cfg::STAGES;
          tc::mbar_wait(&empty_bar[s], ((it / Cfg::STAGES) & 1u) ^ 1u);
          uint8_t* a_dst = tiles + s * Cfg::STAGE_BYTES;
          uint8_t* b_dst = a_dst + GM_A_BYTES;
          tc::mbar_arrive_expect_tx(&full_bar[s], Cfg::STAGE_BYTES);
          const int k0 = (kb0 + kb) * GM_BK;
          if constexpr (CONV == 1) {
            // A = activation patches: k-block -> (tap, channel block); one 4-D box {64 ch, W, tile_h, tile_n} shifted by the tap
            const int t = (kb0 + kb) / cg.cblocks, cb = (kb0 + kb) % cg.cblocks;
            int n_img, h0;
            conv_pixel_tile(cg, m0 / GM_BM, n_img, h0);
            tc::tma_load_4d(a_dst, &tmap_a, &full_bar[s], cb * GM_BK, cg.off + cg.sign * (t % cg.kw), h0 + cg.off + cg.sign * (t / cg.kw), n_img);
            if constexpr (B_MN) {  // dgrad: weights [Cout, taps * cin] read as [K = co, N = ci] of this tap
              tc::tma_load_2d(b_dst, &tmap_b, &full_bar[s], t * cg.ncols + n0, cb * GM_BK);
            } else {
              tc::tma_load_2d(b_dst, &tmap_b, &full_bar[s], k0, n0);
            }
          } else if constexpr (CONV == 2) {
            // wgrad: A = dz [pixels, co] MN-major; B = activation patches [64 pixels, 64 channels of tap t] MN-major
            tc::tma_load_2d(a_dst, &tmap_a, &full_bar[s], m0, k0);
            tc::tma_load_2d(a_dst + GM_BK * 128, &tmap_a, &full_bar[s], m0 + 64, k0);
            const int t = n0 / cg.ncols, c0 = n0 % cg.ncols;
            int n_img, h0;
            conv_pixel_tile(cg, kb0 + kb, n_img, h0);
            tc::tma_load_4d(b_dst, &tmap_b, &full_bar[s], c0, cg.off + cg.sign * (t % cg.kw), h0 + cg.off + cg.sign * (t / cg.kw), n_img);
          } else {
            if constexpr (A_MN) {  // stored [K, M]: 64(K) x 64(M) boxes side by side
              tc::tma_load_2d(a_dst, &tmap_a, &full_bar[s], m0, k0);
              tc::tma_load_2d(a_dst + GM_BK * 128, &tmap_a, &full_bar[s], m0 + 64, k0);
            } else {
              tc::tma_load_2d(a_dst, &tmap_a, &full_bar[s], k0, m0);
            }
            if constexpr (B_MN) {
#pragma unroll
              for (int j = 0; j < BN / 64; ++j) tc::tma_load_2d(b_dst + j * GM_BK * 128, &tmap_b, &full_bar[s], n0 + 64 * j, k0);
            } else {
              tc::tma_load_2d(b_dst, &tmap_b, &full_bar[s], k0, n0);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (one thread)
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc(tc::FMT_BF16, GM_BM, BN) | (A_MN ? (1u << 15) : 0u) | (B_MN ? (1u << 16) : 0u);
      uint32_t it = 0, acc_it = 0;
      for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++acc_it) {
        int m0, n0, kb0, nk;
        work_coords(w, m0, n0, kb0, nk);
        const uint32_t as = acc_it % GM_ACC_STAGES;
        tc::mbar_wait(&acc_empty[as], ((acc_it / GM_ACC_STAGES) & 1u) ^ 1u);  // epilogue has drained this accumulator
        tc::fence_after_sync();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % Cfg::STAGES;
          tc::mbar_wait(&full_bar[s], (it / Cfg::STAGES) & 1u);
          tc::fence_after_sync();
          if (it == 0) trace_stamp(ep.trace, 2);  // first operand tile landed
          const uint32_t a_addr = tc::smem_u32(tiles + s * Cfg::STAGE_BYTES);
          const uint32_t b_addr = a_addr + GM_A_BYTES;
#pragma unroll
          for (int ks = 0; ks < GM_BK / 16; ++ks) {
            // K-major: 16 elements = 32 bytes along the swizzle row; MN-major: 16 K-rows = 2048 bytes down the tile
            const uint64_t da = A_MN ? smem_desc_sw128_mn(a_addr + ks * 2048, GM_BK * 128) : tc::smem_desc_sw128(a_addr + ks * 32);
            const uint64_t db = B_MN ? smem_desc_sw128_mn(b_addr + ks * 2048, GM_BK * 128) : tc::smem_desc_sw128(b_addr + ks * 32);
            tc::mma_bf16(d_tmem, da, db, idesc, (kb | ks) != 0 ? 1u : 0u);
          }
          tc::mma_commit(&empty_bar[s]);
        }
        tc::mma_commit(&acc_full[as]);
        if (acc_it == 0) trace_stamp(ep.trace, 3);  // all MMAs of the first tile issued
      }
    }
  } else {
    // ================================ epilogue: this warp owns TMEM lanes [32q, 32q+32), q = warp % 4
    const int q = warp & 3;
    uint32_t acc_it = 0;
    for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++acc_it) {
      int m0, n0, kb0, nk;
      work_coords(w, m0, n0, kb0, nk);
      const uint32_t as = acc_it % GM_ACC_STAGES;
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < M;
      tc::mbar_wait(&acc_full[as], (acc_it / GM_ACC_STAGES) & 1u);
      tc::fence_after_sync();
      if (acc_it == 0 && q == 0 && lane == 0) trace_stamp(ep.trace, 4);  // first accumulator complete
#pragma unroll 1
      for (int cb = 0; cb < BN / 32; ++cb) {
        const int nb = n0 + cb * 32;
        if (nb >= N) break;  // warp-uniform
        float v[32];
        if (nk > 0) {
          tc::tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + (uint32_t)(cb * 32), v);
        } else {
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = 0.f;
        }
        // Side inputs are fetched as whole 32-column rows into their own registers BEFORE they are consumed (16-byte vector
        // loads on the common aligned path): per-element predicated loads compile into one serial load->use chain per column.
        const bool full = nb + 32 <= N;
        if (ep.bias != nullptr) {
          float bv[32];
          if (full && (reinterpret_cast<uintptr_t>(ep.bias + nb) & 15u) == 0) {
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              const float4 t = __ldg(reinterpret_cast<const float4*>(ep.bias + nb + e));
              bv[e] = t.x; bv[e + 1] = t.y; bv[e + 2] = t.z; bv[e + 3] = t.w;
            }
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) bv[e] = (nb + e < N) ? __ldg(ep.bias + nb + e) : 0.f;
          }
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] += bv[e];
        }
        if (ep.add16 != nullptr && row_ok) {
          float av[32];
          load_row32_bf16(ep.add16 + (int64_t)row * ep.ldadd + nb, full && (ep.ldadd & 7) == 0, N - nb, av);
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] += av[e];
        }
        if (ep.relu) {
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = fmaxf(v[e], 0.f);
        }
        if (ep.mask != nullptr && row_ok) {
          float mv[32];
          load_row32_bf16(ep.mask + (int64_t)row * ep.ldm + nb, full && (ep.ldm & 7) == 0, N - nb, mv);
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = mv[e] > 0.f ? v[e] : 0.f;
        }
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = (row_ok && nb + e < N) ? v[e] : 0.f;
        if (ep.out_f32 != nullptr && row_ok) {
          float* o = ep.out_f32 + (int64_t)row * ep.ldo + nb;
          if (ep.perm_cin > 0) {
            // column j = tap * cin + c goes to [row][c][tap]; with cin % 32 == 0 the 32 columns of this block share one tap
            const bool one_tap = (ep.perm_cin & 31) == 0;
            float* op = ep.out_f32 + (int64_t)row * ep.ldo + (int64_t)(nb % ep.perm_cin) * ep.perm_taps + nb / ep.perm_cin;
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              if (nb + e < N) {
                float* o1 = one_tap ? op + e * ep.perm_taps : ep.out_f32 + (int64_t)row * ep.ldo + (int64_t)((nb + e) % ep.perm_cin) * ep.perm_taps + (nb + e) / ep.perm_cin;
                if (split) {
                  if (v[e] != 0.f) atomicAdd(o1, v[e]);
                } else {
                  *o1 = ep.accumulate ? *o1 + v[e] : v[e];
                }
              }
            }
          } else if (split) {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (nb + e < N && v[e] != 0.f) atomicAdd(o + e, v[e]);
          } else if (nb + 32 <= N && (ep.ldo & 3) == 0) {
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              float4 wv = make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]);
              if (ep.accumulate) {
                const float4 old = *reinterpret_cast<const float4*>(o + e);
                wv.x += old.x; wv.y += old.y; wv.z += old.z; wv.w += old.w;
              }
              *reinterpret_cast<float4*>(o + e) = wv;
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
              uint4 wv;
              wv.x = *reinterpret_cast<uint32_t*>(&p0); wv.y = *reinterpret_cast<uint32_t*>(&p1);
              wv.z = *reinterpret_cast<uint32_t*>(&p2); wv.w = *reinterpret_cast<uint32_t*>(&p3);
              *reinterpret_cast<uint4*>(o + e) = wv;
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
        if (ep.colsum != nullptr || ep.colsumsq != nullptr) {
          // column sums (bias gradient / batch-norm statistics) of this warp's 32 x 32 block: a transposing butterfly leaves the
          // total of column `lane` in lane `lane` (31 shuffles instead of 32 full reductions), then ONE coalesced 32-wide atomic
          if (ep.colsum != nullptr) {
            const float t = warp_colsum32(v, lane, false);
            if (nb + lane < N && t != 0.f) atomicAdd(ep.colsum + nb + lane, t);
          }
          if (ep.colsumsq != nullptr) {
            const float t = warp_colsum32(v, lane, true);
            if (nb + lane < N && t != 0.f) atomicAdd(ep.colsumsq + nb + lane, t);
          }
        }
      }
      // every tcgen05.ld of this warp has completed (tmem_ld32 waits): hand the accumulator back to the MMA warp
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&acc_empty[as]);
    }
  }
  if (threadIdx.x == 64) trace_stamp(ep.trace, 5);  // epilogue warp 2 done
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
  if (threadIdx.x == 32) trace_stamp(ep.trace, 6);  // exit
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
  cuuint32_t box[2] = {(cuuint32_t)GM_BK, (cuuint32_t)box_rows};  // box_rows <= 256 (TMA limit)
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%d cols=%d ld=%d)", (int)r, rows, cols, ld);
    return 1;
  }
  return 0;
}

// bf16 NHWC activation [Nimg, H, W, C] as a 4-D tensor {C, W, H, N}; box = {64, W, tile_h, tile_n}, 128-byte swizzle, zero fill outside
static int make_tmap_nhwc(CUtensorMap* tm, const void* base, int n_img, int h, int w, int c, int tile_h, int tile_n) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return 1;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || (c & 63)) {
    set_error("conv_gemm: activation must be 16-byte aligned with channels %% 64 == 0 (c=%d)", c);
    return 1;
  }
  cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n_img};
  cuuint64_t strides[3] = {(cuuint64_t)c * 2, (cuuint64_t)w * c * 2, (cuuint64_t)h * w * c * 2};
  cuuint32_t box[4] = {(cuuint32_t)GM_BK, (cuuint32_t)w, (cuuint32_t)tile_h, (cuuint32_t)tile_n};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (4-D) failed with CUresult %d (n=%d h=%d w=%d c=%d box %dx%dx%d)", (int)r, n_img, h, w, c, w, tile_h, tile_n);
    return 1;
  }
  return 0;
}

// pixel tile of `pixels` (128 or 64) output positions = full rows of one image, or whole images
static int conv_tile_shape(int h, int w, int pixels, int* tile_h, int* tile_n, int* hw_tiles) {
  if (w <= 0 || h <= 0 || pixels % w != 0) return 1;
  int th = pixels / w;
  if (th <= h) {
    if (h % th != 0) return 1;
    *tile_h = th;
    *tile_n = 1;
    *hw_tiles = h / th;
    return 0;
  }
  if (pixels % (h * w) != 0) return 1;
  *tile_h = h;
  *tile_n = pixels / (h * w);
  *hw_tiles = 0;
  return 0;
}

}  // namespace b2

using namespace b2;

static long long* g_gemm_trace = nullptr;
/* debug hook (tools/gemm_trace.py): device buffer of 8 int64 that CTA 0 of every following b2_gemm_bf16 launch stamps with %globaltimer */
extern "C" void b2_gemm_set_trace(void* dev_buf) { g_gemm_trace = static_cast<long long*>(dev_buf); }

template <int BN>
static int launch_gemm(const b2_gemm_args* a, const CUtensorMap& ta, const CUtensorMap& tb, int splits, int kb_per_split, const GemmEpilogue& ep, cudaStream_t st) {
  using Cfg = GemmCfg<BN>;
  static bool attr_set[64] = {false};
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    B2_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<BN, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    B2_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<BN, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    B2_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<BN, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    B2_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<BN, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    attr_set[dev] = true;
  }
  const int tiles_m = (a->M + GM_BM - 1) / GM_BM, tiles_n = (a->N + BN - 1) / BN;
  const int n_work = tiles_m * tiles_n * splits;
  int grid = sm_count() * Cfg::CTAS_PER_SM;
  if (n_work < grid) grid = n_work;
  if (a->a_mn && a->b_mn) launch_pdl(gemm_bf16_kernel<BN, true, true>, dim3(grid), dim3(GM_THREADS), Cfg::SMEM, st, ta, tb, a->M, a->N, a->K, tiles_m, tiles_n, splits, kb_per_split, ep, ConvGeom{});
  else if (a->a_mn) launch_pdl(gemm_bf16_kernel<BN, true, false>, dim3(grid), dim3(GM_THREADS), Cfg::SMEM, st, ta, tb, a->M, a->N, a->K, tiles_m, tiles_n, splits, kb_per_split, ep, ConvGeom{});
  else if (a->b_mn) launch_pdl(gemm_bf16_kernel<BN, false, true>, dim3(grid), dim3(GM_THREADS), Cfg::SMEM, st, ta, tb, a->M, a->N, a->K, tiles_m, tiles_n, splits, kb_per_split, ep, ConvGeom{});
  else launch_pdl(gemm_bf16_kernel<BN, false, false>, dim3(grid), dim3(GM_THREADS), Cfg::SMEM, st, ta, tb, a->M, a->N, a->K, tiles_m, tiles_n, splits, kb_per_split, ep, ConvGeom{});
  B2_LAUNCH_CHECK("gemm_bf16_kernel");
  return 0;
}

extern "C" int b2_gemm_bf16(const b2_gemm_args* a, void* stream) {
  if (a == nullptr || a->M <= 0 || a->N <= 0 || a->K < 0) {
    set_error("gemm: bad shape");
    return 1;
  }
  // tile width: 256 columns when the problem is big enough to fill the chip with such tiles (halves the shared-memory
  // traffic per MMA, the limiter at N=64), 64 columns for the small / skinny GEMMs of federated local training
  const int64_t big_tiles = (int64_t)((a->M + GM_BM - 1) / GM_BM) * ((a->N + 255) / 256);
  const int bn = (a->N >= 256 && big_tiles >= sm_count() && !a->b_mn) ? 256 : 64;
  CUtensorMap ta, tb;
  // K-major operand: [M, K] rows = M; MN-major operand: stored [K, M], rows = K, 64 x 64 boxes
  if (a->a_mn ? make_tmap(&ta, a->A, a->K, a->M, a->lda, GM_BK) : make_tmap(&ta, a->A, a->M, a->K, a->lda, GM_BM)) return 1;
  if (a->b_mn ? make_tmap(&tb, a->B, a->K, a->N, a->ldb, GM_BK) : make_tmap(&tb, a->B, a->N, a->K, a->ldb, bn)) return 1;
  const int total_k = (a->K + GM_BK - 1) / GM_BK;
  int splits = a->split_k > 1 ? a->split_k : 1;
  if (splits > total_k) splits = total_k > 0 ? total_k : 1;
  if (splits > 1 && (a->out_f32 == nullptr || a->out_bf16 != nullptr || a->out_bf16_t != nullptr || a->bias != nullptr || a->relu || a->mask != nullptr)) {
    set_error("gemm: split_k only combines with a plain fp32 (accumulating) output");
    return 1;
  }
  int kb_per_split = (total_k + splits - 1) / splits;
  if (kb_per_split < 1) kb_per_split = 1;
  splits = total_k > 0 ? (total_k + kb_per_split - 1) / kb_per_split : 1;
  GemmEpilogue ep{a->out_f32, static_cast<__nv_bfloat16*>(a->out_bf16), static_cast<__nv_bfloat16*>(a->out_bf16_t), a->bias, static_cast<const __nv_bfloat16*>(a->mask),
                  a->colsum, a->colsumsq, a->ldo, a->ldo16, a->ldt, a->ldm, a->relu, a->accumulate, nullptr, 0, g_gemm_trace, a->perm_cin, a->perm_taps};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (bn == 256) return launch_gemm<256>(a, ta, tb, splits, kb_per_split, ep, st);
  return launch_gemm<64>(a, ta, tb, splits, kb_per_split, ep, st);
}

// ---------------------------------------------------------------------------------------------- implicit-GEMM convolution
template <bool A_MN, bool B_MN, int CONV>
static int launch_conv(int M, int N, int K, const CUtensorMap& ta, const CUtensorMap& tb, int splits, int kb_per_split, const GemmEpilogue& ep, const ConvGeom& cg, cudaStream_t st) {
  using Cfg = GemmCfg<64>;
  static bool attr_set[64] = {false};
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    B2_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<64, A_MN, B_MN, CONV>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    attr_set[dev] = true;
  }
  const int tiles_m = (M + GM_BM - 1) / GM_BM, tiles_n = (N + 63) / 64;
  const int n_work = tiles_m * tiles_n * splits;
  int grid = sm_count() * Cfg::CTAS_PER_SM;
  if (n_work < grid) grid = n_work;
  launch_pdl(gemm_bf16_kernel<64, A_MN, B_MN, CONV>, dim3(grid), dim3(GM_THREADS), Cfg::SMEM, st, ta, tb, M, N, K, tiles_m, tiles_n, splits, kb_per_split, ep, cg);
  B2_LAUNCH_CHECK("gemm_bf16_kernel (conv)");
  return 0;
}

extern "C" int b2_conv_gemm(const b2_conv_gemm_args* a, void* stream) {
  if (a == nullptr || a->batch <= 0 || a->h <= 0 || a->w <= 0 || a->cin <= 0 || a->cout <= 0 || a->ksize <= 0 || (a->ksize & 1) == 0) {
    set_error("conv_gemm: bad shape");
    return 1;
  }
  if ((a->cin & 63) || (a->mode == B2_CONV_DGRAD && (a->cout & 63))) {
    set_error("conv_gemm: channel counts must be multiples of 64 (cin=%d cout=%d)", a->cin, a->cout);
    return 1;
  }
  const int taps = a->ksize * a->ksize, pad = a->ksize / 2;
  const int pixels = a->batch * a->h * a->w;
  const int kdim = taps * a->cin;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CUtensorMap ta, tb;
  ConvGeom cg{};
  cg.kw = a->ksize;
  GemmEpilogue ep{};
  ep.out_f32 = a->out_f32;
  ep.out_bf16 = static_cast<__nv_bfloat16*>(a->out_bf16);
  ep.bias = a->bias;
  ep.colsum = a->colsum;
  ep.colsumsq = a->colsumsq;
  ep.relu = a->relu;
  ep.accumulate = a->accumulate;
  ep.add16 = static_cast<const __nv_bfloat16*>(a->add_bf16);
  const int tile_pixels = a->mode == B2_CONV_WGRAD ? GM_BK : GM_BM;
  if (conv_tile_shape(a->h, a->w, tile_pixels, &cg.tile_h, &cg.tile_n, &cg.hw_tiles)) {
    set_error("conv_gemm: a %d-pixel tile must be whole rows of one image or whole images (h=%d w=%d)", tile_pixels, a->h, a->w);
    return 1;
  }
  if (cg.tile_h > 256 || cg.tile_n > 256 || a->w > 256) {
    set_error("conv_gemm: TMA box limit (w=%d)", a->w);
    return 1;
  }
  if (a->mode == B2_CONV_FWD) {
    // z[pixels, cout] = patches(x)[pixels, taps * cin] * w[cout, taps * cin]^T
    cg.cblocks = a->cin / 64;
    cg.off = -pad;
    cg.sign = 1;
    if (make_tmap_nhwc(&ta, a->act, a->batch, a->h, a->w, a->cin, cg.tile_h, cg.tile_n)) return 1;
    if (make_tmap(&tb, a->weights, a->cout, kdim, kdim, 64)) return 1;
    ep.ldo = ep.ldo16 = a->cout;
    ep.ldadd = a->cout;
    return launch_conv<false, false, 1>(pixels, a->cout, kdim, ta, tb, 1, taps * cg.cblocks, ep, cg, st);
  }
  if (a->mode == B2_CONV_DGRAD) {
    // dx[pixels, cin] = patches'(dz)[pixels, taps * cout] * w'[cin, taps * cout]^T, taps mirrored
    cg.cblocks = a->cout / 64;
    cg.off = pad;
    cg.sign = -1;
    cg.ncols = a->cin;
    if (make_tmap_nhwc(&ta, a->act, a->batch, a->h, a->w, a->cout, cg.tile_h, cg.tile_n)) return 1;
    if (make_tmap(&tb, a->weights, a->cout, kdim, kdim, GM_BK)) return 1;
    ep.ldo = ep.ldo16 = a->cin;
    ep.ldadd = a->cin;
    return launch_conv<false, true, 1>(pixels, a->cin, taps * a->cout, ta, tb, 1, taps * cg.cblocks, ep, cg, st);
  }
  if (a->mode == B2_CONV_WGRAD) {
    // dw[cout, taps * cin] = dz[pixels, cout]^T * patches(x)[pixels, taps * cin]; K = pixels in 64-pixel blocks
    if (a->out_f32 == nullptr || a->out_bf16 != nullptr || a->bias != nullptr || a->relu || a->add_bf16 != nullptr || a->colsum != nullptr || a->colsumsq != nullptr) {
      set_error("conv_gemm: wgrad writes a plain fp32 gradient");
      return 1;
    }
    cg.cblocks = a->cin / 64;
    cg.off = -pad;
    cg.sign = 1;
    cg.ncols = a->cin;
    const int n_img_padded = cg.hw_tiles > 0 ? a->batch : (a->batch + cg.tile_n - 1) / cg.tile_n * cg.tile_n;
    const int kpix = n_img_padded * a->h * a->w;  // pixel blocks past the batch are zero-filled by TMA on both operands
    if (make_tmap(&ta, a->dz, pixels, a->cout, a->cout, GM_BK)) return 1;
    if (make_tmap_nhwc(&tb, a->act, a->batch, a->h, a->w, a->cin, cg.tile_h, cg.tile_n)) return 1;
    const int total_k = kpix / GM_BK;
    int splits = a->split_k > 1 ? a->split_k : 1;
    if (splits > total_k) splits = total_k;
    int kb_per_split = (total_k + splits - 1) / splits;
    splits = (total_k + kb_per_split - 1) / kb_per_split;
    ep.ldo = kdim;
    ep.perm_cin = a->cin;
    ep.perm_taps = taps;
    return launch_conv<true, true, 2>(a->cout, kdim, kpix, ta, tb, splits, kb_per_split, ep, cg, st);
  }
  set_error("conv_gemm: unknown mode %d", a->mode);
  return 1;
}
