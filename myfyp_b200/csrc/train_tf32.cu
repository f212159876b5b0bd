// Node-batched TF32 training step for the CNN family (conv3x3-ReLU-maxpool2 blocks, Flatten, Linear-ReLU-Linear).
//
// Why: federated nodes are tiny workloads (0.55 M parameters, batch 64: ~4 GFLOP per step = 3 us of tensor-core time), so a
// step is a chain of latency-bound launches.  The reference answers "many simulated nodes per box" with one Ray actor PROCESS
// per learner (p2pfl/learning/frameworks/simulation/actor_pool.py:101-128, virtual_learner.py:101-137); here the nodes that
// share a GPU train in the SAME launches: every kernel takes the node ("slot") from blockIdx.y, the slot's tensor maps, shapes
// and epilogue pointers live in a device table built once (b2_cnn_plan_create), the node's own pointers (parameter slab, Adam
// state, data set, permutation) in a bind table rewritten before a CUDA-graph replay.  One graph per GPU per epoch.
//
// Precision = the reference's: torch trains this model in fp32 storage with TF32 tensor-core convolutions (cuDNN default) and
// fp32 nn.Linear (pytorch/lightning_learner.py:78-100 builds a default fp32 Trainer).  Here: fp32 storage, convolutions as
// tcgen05.mma kind::tf32 on operands rounded to tf32 (round-to-nearest, as cuDNN does), linear layers as 3xTF32 (hi/lo split,
// hi*hi + lo*hi + hi*lo: ~2^-21 relative, fp32-class), head (last linear layer, softmax, loss) in plain fp32 FMA.
//
// Launches per step (13; the three weight-gradient GEMMs run on a side stream = a parallel branch of the graph):
//   conv0_fwd      first layer (1-3 input channels): producer warps gather the batch by permutation index straight from the
//                  NCHW data set, build the 27(32)-wide patch rows in swizzled shared memory; tcgen05 MMA; epilogue = bias +
//                  ReLU + 2x2 max-pool + winner index, all in registers / warp shuffles (pool partners share a warp)
//   conv_fwd   xL  implicit GEMM: 4-D TMA boxes of the pooled NHWC activation per filter tap (zero fill = padding), same epilogue
//   fc1_fwd        3xTF32, split-K, fp32 atomics into z1
//   head           h1 = relu(z1 + b1), logits, softmax, loss, dlogits, dW2, db2, dh1 = (dlogits W2) * (h1 > 0) as hi/lo, db1
//   fc1_wgrad *    3xTF32, both operands MN-major (the forward buffers, nothing transposed)
//   fc1_dgrad      3xTF32; epilogue routes each pooled gradient to its 2x2 winner (un-pool), applies the ReLU mask, adds db
//   conv_wgrad *xL implicit GEMM over pixels (split-K), activation boxes as the MN-major B operand
//   conv_dgrad xL  implicit GEMM with mirrored taps, same un-pool epilogue (the last one only masks: conv0_wgrad un-pools itself)
//   conv0_wgrad    producer-built operands again (patches of x, un-pooled dz), one TMEM accumulator across all tiles of a CTA
//   adam           fused Adam (+FedProx / Scaffold terms) over the slab; reads the gradient mirror through the layer's index
//                  permutation, clears it, and writes the next step's tf32 / hi-lo weight mirrors - no separate repack / cast
// Shared-memory operand layouts: K-major = canonical SWIZZLE_128B (one 128-byte row = 32 fp32, 16-byte chunks XOR row & 7);
// MN-major = the 32-byte-atom variant of it (32-byte chunks XOR row & 3), the only one the tensor core accepts for 32-bit types.
// TMA loads here read this GPU's memory only and none uses cluster multicast.
#include <cuda.h>

#include <cstdlib>
#include <mutex>
#include <vector>

#include "../../include/myfyp_b200.h"
#include "common.cuh"
#include "tcgen05.cuh"

namespace b2 {
namespace t32 {

constexpr int BM = 128;
constexpr int BK = 32;              // fp32 elements per k-block = one 128-byte swizzle row
constexpr int A_BYTES = BM * 128;   // 16 KB
constexpr int GG_THREADS = 192;     // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue
constexpr int ACC_STAGES = 2;

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// A lane of an epilogue warp owns one output row and writes 32 consecutive floats of it.  With 16-byte stores every warp instruction
// touches 32 half-filled 32-byte sectors; the 32-byte form (STG.256, sm_100) writes whole sectors with half the requests.
__device__ __forceinline__ void stg256(void* p, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t a4, uint32_t a5, uint32_t a6, uint32_t a7) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(a4), "r"(a5), "r"(a6), "r"(a7) : "memory");
}
__device__ __forceinline__ void store_row32(float* o, const float (&v)[32]) {
  if ((reinterpret_cast<uintptr_t>(o) & 31) == 0) {
#pragma unroll
    for (int e = 0; e < 32; e += 8)
      stg256(o + e, __float_as_uint(v[e]), __float_as_uint(v[e + 1]), __float_as_uint(v[e + 2]), __float_as_uint(v[e + 3]), __float_as_uint(v[e + 4]), __float_as_uint(v[e + 5]),
             __float_as_uint(v[e + 6]), __float_as_uint(v[e + 7]));
  } else {
#pragma unroll
    for (int e = 0; e < 32; e += 4) *reinterpret_cast<float4*>(o + e) = make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]);
  }
}
__device__ __forceinline__ void store_row8(float* o, const uint32_t (&b)[8]) {
  if ((reinterpret_cast<uintptr_t>(o) & 31) == 0) {
    stg256(o, b[0], b[1], b[2], b[3], b[4], b[5], b[6], b[7]);
  } else {
    reinterpret_cast<uint4*>(o)[0] = make_uint4(b[0], b[1], b[2], b[3]);
    reinterpret_cast<uint4*>(o)[1] = make_uint4(b[4], b[5], b[6], b[7]);
  }
}
// column block `pos` (8 of the 32 columns) of a register array, without dynamic indexing
__device__ __forceinline__ uint32_t pick8(const uint32_t (&a)[32], int pos, int j) {
  const uint32_t lo = pos & 1 ? a[8 + j] : a[j], hi = pos & 1 ? a[24 + j] : a[16 + j];
  return pos & 2 ? hi : lo;
}
__device__ __forceinline__ void store_codes32(uint8_t* ip, const uint32_t (&b)[8]) {
  if ((reinterpret_cast<uintptr_t>(ip) & 31) == 0) {
    stg256(ip, b[0], b[1], b[2], b[3], b[4], b[5], b[6], b[7]);
  } else {
    reinterpret_cast<uint4*>(ip)[0] = make_uint4(b[0], b[1], b[2], b[3]);
    reinterpret_cast<uint4*>(ip)[1] = make_uint4(b[4], b[5], b[6], b[7]);
  }
}
enum { EPI_STORE = 0, EPI_ATOMIC = 1, EPI_POOL = 2, EPI_UNPOOL = 3, EPI_MASK = 4 };

struct alignas(128) Problem {
  CUtensorMap ta, tb, ta_lo, tb_lo;
  int M, N, K, batch;
  int tiles_m, tiles_n, splits, kb_per_split;
  int cblocks, kw, off, sign, hw_tiles, tile_h, tile_n, ncols;
  float* out;
  float* out_lo;
  unsigned char* idx;
  float* colsum;
  long long bias_off;
  int ldo, relu, img_h, img_w, chans, pad0;
};

__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
// cvt.rna.tf32.f32 is emulated on sm_100 (NaN test + add + mask); after a ReLU the operand is finite-or-inf and non-negative
// (fmaxf(NaN, 0) = 0), where add + mask alone is the same rounding.  Returns the bits.
__device__ __forceinline__ uint32_t tf32_rna_bits_nonneg(float x) { return (__float_as_uint(x) + 0x1000u) & 0x7fffe000u; }
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

__device__ __forceinline__ bool slot_live(const b2_cnn_bind& bd, int step, int batch) { return bd.active != 0 && (long long)step * batch < (long long)bd.n_samples; }

__device__ __forceinline__ uint64_t desc_k(uint32_t addr) { return tc::smem_desc_sw128(addr); }
// MN-major operands of 32-bit types have ONE legal shared-memory layout: 128-byte rows (a row = one K index, 32 fp32 of the MN
// extent) swizzled in 32-byte chunks with a period of 4 rows (layout type 1, "128B swizzle with 32B atoms"; TMA writes it with
// CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B).  4-row groups are 512 B apart (SBO); the next 32 MN elements `lbo` bytes further (LBO).
__device__ __forceinline__ uint64_t desc_mn(uint32_t addr, uint32_t lbo) {
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | (32ull << 32) | (1ull << 46) | (1ull << 61);
}
// byte offset of (row, 16-byte chunk) in that layout
__device__ __forceinline__ uint32_t sw32_offset(uint32_t row, uint32_t chunk16) { return row * 128u + ((((chunk16 >> 1) ^ (row & 3u)) << 5) | ((chunk16 & 1u) << 4)); }

__device__ __forceinline__ void pixel_tile(const Problem& P, int pt, int& n, int& h0) {
  if (P.hw_tiles > 0) {
    n = pt / P.hw_tiles;
    h0 = (pt % P.hw_tiles) * P.tile_h;
  } else {
    n = pt * P.tile_n;
    h0 = 0;
  }
}

// sum over the 32 lanes (rows) of column `lane`
__device__ __forceinline__ float warp_colsum32(const float (&v)[32], int lane) {
  float w[16];
  {
    const bool hi = lane & 16;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float a = v[i], b = v[i + 16];
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

__host__ __device__ constexpr int cfg_mode(int epi) { return epi == EPI_ATOMIC ? 1 : ((epi == EPI_UNPOOL || epi == EPI_MASK) ? 2 : 0); }
// MODE 0: 4 stages, 2 CTAs per SM.  1 (split-K kernels with the atomic epilogue): 2 stages, 4 CTAs.  2 (convolutions with the
// un-pool / mask epilogues): 3 stages, 3 CTAs.  The TMA-fed kernels here live on the number of resident CTAs per SM.
template <bool X3, int BN, int MODE = 0>
struct Cfg {
  static constexpr bool LEAN = MODE == 1;
  static constexpr int B_BYTES = BN * 128;
  static constexpr int OPS = X3 ? 2 : 1;
  static constexpr int STAGE_BYTES = OPS * (A_BYTES + B_BYTES);
  // LEAN (kernels with a light epilogue): 3 stages, 3 CTAs per SM - these kernels are bound by the latency of their small TMA
  // boxes, not by bandwidth, so more resident CTAs is what buys throughput
  static constexpr int STAGES = LEAN ? (X3 ? 3 : 2) : ((MODE == 2 && !X3) ? 3 : 4);
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024;
  static constexpr int CTAS_PER_SM = X3 ? 1 : (LEAN ? 4 : (MODE == 2 ? 3 : 2));
  static constexpr uint32_t ACC_COLS = (uint32_t)(ACC_STAGES * BN);
  static constexpr uint32_t TMEM_COLS = ACC_COLS <= 32u ? 32u : ACC_COLS <= 64u ? 64u : ACC_COLS <= 128u ? 128u : ACC_COLS <= 256u ? 256u : 512u;
};

// ------------------------------------------------------------------------------------------------ grouped tf32 GEMM / implicit conv
// CONV 0: plain GEMM   1: A = 4-D activation boxes (fwd, or dgrad with B_MN)   2: wgrad (A = dz MN-major, B = activation boxes MN-major)
template <int CONV, bool A_MN, bool B_MN, bool X3, int EPI, int BN>
__global__ void __launch_bounds__(GG_THREADS, Cfg<X3, BN, cfg_mode(EPI)>::CTAS_PER_SM) ggemm_kernel(const Problem* __restrict__ probs, const b2_cnn_bind* __restrict__ binds, int step) {
  using C = Cfg<X3, BN, cfg_mode(EPI)>;
  const Problem& P = probs[blockIdx.y];
  const b2_cnn_bind& bd = binds[blockIdx.y];
  if (!slot_live(bd, step, P.batch)) return;  // whole CTA leaves before any barrier / TMEM allocation

  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[C::STAGES], empty_bar[C::STAGES], acc_full[ACC_STAGES], acc_empty[ACC_STAGES];
  __shared__ uint32_t tmem_base_s;
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int M = P.M, N = P.N;
  const int total_k = (P.K + BK - 1) / BK;
  const int tiles_n = P.tiles_n, splits = P.splits, kb_per_split = P.kb_per_split;
  const int n_work = P.tiles_m * tiles_n * splits;

  if (tid == 0) {
    tc::tma_prefetch_desc(&P.ta);
    tc::tma_prefetch_desc(&P.tb);
    if (X3) {
      tc::tma_prefetch_desc(&P.ta_lo);
      tc::tma_prefetch_desc(&P.tb_lo);
    }
    for (int s = 0; s < C::STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < ACC_STAGES; ++a) {
      tc::mbar_init(&acc_full[a], 1);
      tc::mbar_init(&acc_empty[a], 4);
    }
    tc::mbar_fence_init();
  }
  if (warp == 1) tc::tmem_alloc<C::TMEM_COLS>(&tmem_base_s);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = tmem_base_s;

  auto work_coords = [&](int w, int& m0, int& n0, int& kb0, int& nk) {
    const int z = w % splits;
    const int t = w / splits;
    m0 = (t / tiles_n) * BM;
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
          const int s = it % C::STAGES;
          tc::mbar_wait(&empty_bar[s], ((it / C::STAGES) & 1u) ^ 1u);
          uint8_t* a_dst = tiles + s * C::STAGE_BYTES;
          uint8_t* b_dst = a_dst + C::OPS * A_BYTES;
          const int k0 = (kb0 + kb) * BK;
          if constexpr (CONV == 1) {
            tc::mbar_arrive_expect_tx(&full_bar[s], C::STAGE_BYTES);
            const int t = (kb0 + kb) / P.cblocks, cb = (kb0 + kb) % P.cblocks;
            int n_img, h0;
            pixel_tile(P, m0 / BM, n_img, h0);
            tc::tma_load_4d(a_dst, &P.ta, &full_bar[s], cb * BK, P.off + P.sign * (t % P.kw), h0 + P.off + P.sign * (t / P.kw), n_img);
            if constexpr (B_MN) {  // dgrad: weights [cout, taps * cin] read as [K = co, N = ci] of tap t
#pragma unroll
              for (int j = 0; j < BN / 32; ++j) tc::tma_load_2d(b_dst + j * (BK * 128), &P.tb, &full_bar[s], t * P.ncols + n0 + 32 * j, cb * BK);
            } else {
              tc::tma_load_2d(b_dst, &P.tb, &full_bar[s], k0, n0);
            }
          } else if constexpr (CONV == 2) {
            // B blocks beyond the last (tap, channel block) are not fetched: their output columns are never stored
            int nblk = 0;
#pragma unroll
            for (int j = 0; j < BN / 32; ++j) nblk += (n0 + 32 * j < N) ? 1 : 0;
            // likewise the 32-row blocks of A beyond M (cout = 64: half of the 128-row tile): whatever the shared memory holds there
            // only reaches accumulator rows that are never stored
            int ablk = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) ablk += (m0 + 32 * j < M) ? 1 : 0;
            tc::mbar_arrive_expect_tx(&full_bar[s], (ablk + nblk) * (BK * 128));
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (m0 + 32 * j < M) tc::tma_load_2d(a_dst + j * (BK * 128), &P.ta, &full_bar[s], m0 + 32 * j, k0);
            int n_img, h0;
            pixel_tile(P, kb0 + kb, n_img, h0);
#pragma unroll
            for (int j = 0; j < BN / 32; ++j) {
              const int col = n0 + 32 * j;
              if (col < N) {
                const int t = col / P.ncols, c0 = col % P.ncols;
                tc::tma_load_4d(b_dst + j * (BK * 128), &P.tb, &full_bar[s], c0, P.off + P.sign * (t % P.kw), h0 + P.off + P.sign * (t / P.kw), n_img);
              }
            }
          } else {
            tc::mbar_arrive_expect_tx(&full_bar[s], C::STAGE_BYTES);
#pragma unroll
            for (int o = 0; o < C::OPS; ++o) {
              const CUtensorMap* ta = o == 0 ? &P.ta : &P.ta_lo;
              const CUtensorMap* tb = o == 0 ? &P.tb : &P.tb_lo;
              uint8_t* ad = a_dst + o * A_BYTES;
              uint8_t* bdst = b_dst + o * C::B_BYTES;
              if constexpr (A_MN) {
#pragma unroll
                for (int j = 0; j < 4; ++j) tc::tma_load_2d(ad + j * (BK * 128), ta, &full_bar[s], m0 + 32 * j, k0);
              } else {
                tc::tma_load_2d(ad, ta, &full_bar[s], k0, m0);
              }
              if constexpr (B_MN) {
#pragma unroll
                for (int j = 0; j < BN / 32; ++j) tc::tma_load_2d(bdst + j * (BK * 128), tb, &full_bar[s], n0 + 32 * j, k0);
              } else {
                tc::tma_load_2d(bdst, tb, &full_bar[s], k0, n0);
              }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (one thread)
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc(tc::FMT_TF32, BM, BN) | (A_MN ? (1u << 15) : 0u) | (B_MN ? (1u << 16) : 0u);
      uint32_t it = 0, acc_it = 0;
      for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++acc_it) {
        int m0, n0, kb0, nk;
        work_coords(w, m0, n0, kb0, nk);
        const uint32_t as = acc_it % ACC_STAGES;
        tc::mbar_wait(&acc_empty[as], ((acc_it / ACC_STAGES) & 1u) ^ 1u);
        tc::fence_after_sync();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % C::STAGES;
          tc::mbar_wait(&full_bar[s], (it / C::STAGES) & 1u);
          tc::fence_after_sync();
          const uint32_t a_addr = tc::smem_u32(tiles + s * C::STAGE_BYTES);
          const uint32_t b_addr = a_addr + C::OPS * A_BYTES;
#pragma unroll
          for (int ks = 0; ks < BK / 8; ++ks) {
            // K-major: 8 tf32 = 32 bytes along the swizzle row; MN-major: 8 K-rows = 1024 bytes down the tile
            const uint64_t da = A_MN ? desc_mn(a_addr + ks * 1024, BK * 128) : desc_k(a_addr + ks * 32);
            const uint64_t db = B_MN ? desc_mn(b_addr + ks * 1024, BK * 128) : desc_k(b_addr + ks * 32);
            tc::mma_tf32(d_tmem, da, db, idesc, (kb | ks) != 0 ? 1u : 0u);
            if constexpr (X3) {
              const uint64_t dal = A_MN ? desc_mn(a_addr + A_BYTES + ks * 1024, BK * 128) : desc_k(a_addr + A_BYTES + ks * 32);
              const uint64_t dbl = B_MN ? desc_mn(b_addr + C::B_BYTES + ks * 1024, BK * 128) : desc_k(b_addr + C::B_BYTES + ks * 32);
              tc::mma_tf32(d_tmem, dal, db, idesc, 1u);
              tc::mma_tf32(d_tmem, da, dbl, idesc, 1u);
            }
          }
          tc::mma_commit(&empty_bar[s]);
        }
        tc::mma_commit(&acc_full[as]);
      }
    }
  } else {
    // ================================ epilogue: this warp owns TMEM lanes [32q, 32q+32)
    const int q = warp & 3;
    const float* bias = P.bias_off >= 0 ? bd.slab + P.bias_off : nullptr;
    uint32_t acc_it = 0;
    for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++acc_it) {
      int m0, n0, kb0, nk;
      work_coords(w, m0, n0, kb0, nk);
      const uint32_t as = acc_it % ACC_STAGES;
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < M;
      tc::mbar_wait(&acc_full[as], (acc_it / ACC_STAGES) & 1u);
      tc::fence_after_sync();
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
        if (bias != nullptr) {
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(bias + nb + e));
            v[e] += t.x; v[e + 1] += t.y; v[e + 2] += t.z; v[e + 3] += t.w;
          }
        }
        if (P.relu) {
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = fmaxf(v[e], 0.f);
        }
        if constexpr (EPI == EPI_STORE) {
          if (row_ok) {
            store_row32(P.out + (int64_t)row * P.ldo + nb, v);
          }
        } else if constexpr (EPI == EPI_ATOMIC) {
          if (row_ok) {
            float* o = P.out + (int64_t)row * P.ldo + nb;
            // a lane owns a row, so every reduction of a warp instruction lands in a different 32-byte sector of L2: the 16-byte
            // vector form sends a quarter of the sector operations (measured: the scalar form cost conv2's weight gradient 25 %)
            if (nb + 32 <= N && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
#pragma unroll
              for (int e = 0; e < 32; e += 4)
                if (v[e] != 0.f || v[e + 1] != 0.f || v[e + 2] != 0.f || v[e + 3] != 0.f) red_add_v4(o + e, v[e], v[e + 1], v[e + 2], v[e + 3]);
            } else {
#pragma unroll
              for (int e = 0; e < 32; ++e)
                if (nb + e < N && v[e] != 0.f) atomicAdd(o + e, v[e]);
            }
          }
        } else if constexpr (EPI == EPI_POOL) {
          // rows of a tile are pixels in (image, h, w) order and W divides 16: the 2x2 window of a pixel lives in this warp at
          // lane ^ 1 (w) and lane ^ W (h).  Every lane of a window ends up with the window maximum and its position.
          const int W = P.img_w, HW = P.img_h * W;
          const int n_img = row / HW, rem = row - n_img * HW, h = rem / W, wq = rem - h * W;
          const int pos = ((h & 1) << 1) | (wq & 1);
          // Phase 1: per column the window maximum and the winner's position, known to all four lanes of the window afterwards.
          // Phase 2: the four lanes share the work - the lane at window position p converts and stores columns 8p .. 8p+7
          // (one 32-byte sector of the pooled row, 8 winner codes).
          uint32_t val[32];        // bits of the winning value (rounded variant: low 2 bits = 3 - winner position)
          uint32_t pk[2] = {0u, 0u};  // other variants: 2 bits of winner position per column
          const bool keyed = P.relu && P.out_lo == nullptr;
          if (keyed) {
            // non-negative, stored tf32-rounded (13 low bits zero): value and position travel as ONE unsigned key, bits | (3 - pos);
            // a 2x2 max with "lowest position wins a tie" is two shuffles and two integer maxima per column
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              uint32_t k = tf32_rna_bits_nonneg(v[e]) | (uint32_t)(3 - pos);
              k = max(k, __shfl_xor_sync(0xffffffffu, k, 1));
              val[e] = max(k, __shfl_xor_sync(0xffffffffu, k, W));
            }
          } else if (P.relu) {
            // non-negative, kept exact (the consumer splits hi / lo): unsigned order is float order; the winner is the lowest
            // position that holds the maximum
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const uint32_t k = __float_as_uint(v[e]) & 0x7fffffffu;
              uint32_t mx = max(k, __shfl_xor_sync(0xffffffffu, k, 1));
              mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, W));
              uint32_t cand = k == mx ? (uint32_t)pos : 4u;
              cand = min(cand, __shfl_xor_sync(0xffffffffu, cand, 1));
              cand = min(cand, __shfl_xor_sync(0xffffffffu, cand, W));
              val[e] = mx;
              pk[e >> 4] |= cand << (2 * (e & 15));
            }
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              float a = v[e];
              int pa = pos;
              const float o1 = __shfl_xor_sync(0xffffffffu, a, 1);
              bool take = (o1 > a) || (o1 == a && (pos ^ 1) < pa);
              a = take ? o1 : a;
              pa = take ? (pos ^ 1) : pa;
              const float o2 = __shfl_xor_sync(0xffffffffu, a, W);
              const int p2 = __shfl_xor_sync(0xffffffffu, pa, W);
              take = (o2 > a) || (o2 == a && p2 < pa);
              a = take ? o2 : a;
              pa = take ? p2 : pa;
              val[e] = __float_as_uint(a);
              pk[e >> 4] |= (uint32_t)pa << (2 * (e & 15));
            }
          }
          if (row_ok) {
            const int64_t pp = (int64_t)n_img * (HW >> 2) + (h >> 1) * (W >> 1) + (wq >> 1);
            const int64_t at = pp * P.ldo + nb + 8 * pos;
            const uint32_t mypk = (pk[pos >> 1] >> (16 * (pos & 1))) & 0xffffu;
            uint32_t hi[8], lo[8], cw[2] = {0u, 0u};
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint32_t k = pick8(val, pos, j);
              uint32_t code;
              if (keyed) {
                hi[j] = k & ~3u;
                code = (3u - (k & 3u)) | (k > 3u ? 4u : 0u);
              } else {
                const float a = __uint_as_float(k);
                code = ((mypk >> (2 * j)) & 3u) | (a > 0.f ? 4u : 0u);
                if (P.out_lo != nullptr) {  // consumer is a 3xTF32 GEMM: exact hi / lo split
                  hi[j] = k & 0xffffe000u;
                  lo[j] = __float_as_uint(tf32_rna(a - __uint_as_float(hi[j])));
                } else {
                  hi[j] = __float_as_uint(tf32_rna(a));
                }
              }
              cw[j >> 2] |= code << (8 * (j & 3));
            }
            store_row8(P.out + at, hi);
            if (P.out_lo != nullptr) store_row8(P.out_lo + at, lo);
            *reinterpret_cast<uint2*>(P.idx + at) = make_uint2(cw[0], cw[1]);
          }
        } else {
          // EPI_UNPOOL / EPI_MASK: this row's 32 values are gradients of pooled outputs (pooled pixel pr, channels cbase..+32)
          const int Cn = P.chans, per_row = N / Cn;
          const int64_t pr = (int64_t)row * per_row + nb / Cn;
          const int cbase = nb % Cn;
          uint32_t code[8];
          if (row_ok) {
            const uint4* ip = reinterpret_cast<const uint4*>(P.idx + pr * Cn + cbase);
            const uint4 i0 = ip[0], i1 = ip[1];
            code[0] = i0.x; code[1] = i0.y; code[2] = i0.z; code[3] = i0.w;
            code[4] = i1.x; code[5] = i1.y; code[6] = i1.z; code[7] = i1.w;
          } else {
#pragma unroll
            for (int g = 0; g < 8; ++g) code[g] = 0u;
          }
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            const uint32_t cde = (code[e >> 2] >> (8 * (e & 3))) & 0xffu;
            v[e] = (cde & 4u) ? tf32_rna(v[e]) : 0.f;  // ReLU mask of the pooled output (relu o max = max o relu)
          }
          if (P.colsum != nullptr) {
            const float t = warp_colsum32(v, lane);
            if (t != 0.f) atomicAdd(P.colsum + cbase + lane, t);
          }
          if (row_ok) {
            if constexpr (EPI == EPI_MASK) {
              store_row32(P.out + pr * Cn + cbase, v);
            } else {
              const int FW = P.img_w, FH = P.img_h, PW = FW >> 1, PHW = (FH >> 1) * PW;
              const int n_img = (int)(pr / PHW), rem = (int)(pr - (int64_t)n_img * PHW), ph = rem / PW, pw = rem - ph * PW;
              const int64_t fr = ((int64_t)n_img * FH + 2 * ph) * FW + 2 * pw;
#pragma unroll
              for (int ps = 0; ps < 4; ++ps) {
                float wv[32];
#pragma unroll
                for (int e = 0; e < 32; ++e) wv[e] = (((code[e >> 2] >> (8 * (e & 3))) & 3u) == (uint32_t)ps) ? v[e] : 0.f;
                store_row32(P.out + (fr + (ps >> 1) * FW + (ps & 1)) * Cn + cbase, wv);
              }
            }
          }
        }
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&acc_empty[as]);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ conv weight gradient, wide tiles
// dW[co, (tap, ci)] = sum_pixels dz[pixel, co] * x[pixel + tap, ci], one CTA per (cout tile, group of 9 column blocks, K split).
// The generic kernel above computes 64 output columns per pass over the pixels and therefore re-reads dz once per 64 columns
// (4.5x for a 32-channel layer, 9x for a 64-channel one); its pipeline is bound by the bytes that cross from L2 into shared
// memory.  Here a stage holds the dz tile ONCE plus nine activation boxes (= all nine taps of a 32-channel block: the boxes are
// the same pixels shifted by one, L2 serves them), the accumulator is 288 TMEM columns wide, and three N = 96 MMAs per k-step
// consume the stage.  Same operand layouts as CONV = 2 (both MN-major, 32-byte-atom swizzle).
constexpr int WW_NB = 9;                          // 32-column blocks per CTA
constexpr int WW_B_BYTES = WW_NB * BK * 128;      // 36 KB
constexpr int WW_STAGE = A_BYTES + WW_B_BYTES;    // 52 KB
constexpr int WW_STAGES = 4;
constexpr int WW_SMEM = WW_STAGES * WW_STAGE + 1024;

__global__ void __launch_bounds__(GG_THREADS, 1) wgrad_wide_kernel(const Problem* __restrict__ probs, const b2_cnn_bind* __restrict__ binds, int step) {
  const Problem& P = probs[blockIdx.y];
  const b2_cnn_bind& bd = binds[blockIdx.y];
  if (!slot_live(bd, step, P.batch)) return;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[WW_STAGES], empty_bar[WW_STAGES], acc_full, acc_empty;
  __shared__ uint32_t tmem_base_s;
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int M = P.M, N = P.N;
  const int total_k = (P.K + BK - 1) / BK;
  const int groups = (N + 32 * WW_NB - 1) / (32 * WW_NB);
  const int splits = P.splits, kb_per_split = P.kb_per_split;
  const int n_work = P.tiles_m * groups * splits;

  if (tid == 0) {
    tc::tma_prefetch_desc(&P.ta);
    tc::tma_prefetch_desc(&P.tb);
    for (int s = 0; s < WW_STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    tc::mbar_init(&acc_full, 1);
    tc::mbar_init(&acc_empty, 4);
    tc::mbar_fence_init();
  }
  if (warp == 1) tc::tmem_alloc<512>(&tmem_base_s);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = tmem_base_s;

  auto work_coords = [&](int w, int& m0, int& n0, int& kb0, int& nk) {
    const int z = w % splits;
    const int t = w / splits;
    m0 = (t / groups) * BM;
    n0 = (t % groups) * (32 * WW_NB);
    kb0 = z * kb_per_split;
    nk = total_k - kb0;
    nk = nk < 0 ? 0 : (nk < kb_per_split ? nk : kb_per_split);
  };

  if (warp == 0) {
    if (tc::elect_one()) {
      uint32_t it = 0;
      for (int w = blockIdx.x; w < n_work; w += gridDim.x) {
        int m0, n0, kb0, nk;
        work_coords(w, m0, n0, kb0, nk);
        int nblk = 0;
        for (int j = 0; j < WW_NB; ++j) nblk += (n0 + 32 * j < N) ? 1 : 0;
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % WW_STAGES;
          tc::mbar_wait(&empty_bar[s], ((it / WW_STAGES) & 1u) ^ 1u);
          uint8_t* a_dst = tiles + s * WW_STAGE;
          uint8_t* b_dst = a_dst + A_BYTES;
          const int k0 = (kb0 + kb) * BK;
          tc::mbar_arrive_expect_tx(&full_bar[s], A_BYTES + nblk * (BK * 128));
#pragma unroll
          for (int j = 0; j < 4; ++j) tc::tma_load_2d(a_dst + j * (BK * 128), &P.ta, &full_bar[s], m0 + 32 * j, k0);
          int n_img, h0;
          pixel_tile(P, kb0 + kb, n_img, h0);
          for (int j = 0; j < nblk; ++j) {
            const int col = n0 + 32 * j;
            const int t = col / P.ncols, c0 = col % P.ncols;
            tc::tma_load_4d(b_dst + j * (BK * 128), &P.tb, &full_bar[s], c0, P.off + P.sign * (t % P.kw), h0 + P.off + P.sign * (t / P.kw), n_img);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc = tc::make_idesc(tc::FMT_TF32, BM, 96) | (1u << 15) | (1u << 16);
      uint32_t it = 0, acc_it = 0;
      for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++acc_it) {
        int m0, n0, kb0, nk;
        work_coords(w, m0, n0, kb0, nk);
        tc::mbar_wait(&acc_empty, (acc_it & 1u) ^ 1u);
        tc::fence_after_sync();
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % WW_STAGES;
          tc::mbar_wait(&full_bar[s], (it / WW_STAGES) & 1u);
          tc::fence_after_sync();
          const uint32_t a_addr = tc::smem_u32(tiles + s * WW_STAGE);
          const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
          for (int ks = 0; ks < BK / 8; ++ks) {
            const uint64_t da = desc_mn(a_addr + ks * 1024, BK * 128);
#pragma unroll
            for (int g3 = 0; g3 < 3; ++g3) tc::mma_tf32(tmem_base + g3 * 96, da, desc_mn(b_addr + g3 * 3 * (BK * 128) + ks * 1024, BK * 128), idesc, (kb | ks) != 0 ? 1u : 0u);
          }
          tc::mma_commit(&empty_bar[s]);
        }
        tc::mma_commit(&acc_full);
      }
    }
  } else {
    const int q = warp & 3;
    uint32_t acc_it = 0;
    for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++acc_it) {
      int m0, n0, kb0, nk;
      work_coords(w, m0, n0, kb0, nk);
      const int row = m0 + q * 32 + lane;
      tc::mbar_wait(&acc_full, acc_it & 1u);
      tc::fence_after_sync();
      if (nk > 0) {
#pragma unroll 1
        for (int cb = 0; cb < WW_NB; ++cb) {
          const int nb = n0 + cb * 32;
          if (nb >= N) break;
          float v[32];
          tc::tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(cb * 32), v);
          if (row < M) {
            float* o = P.out + (int64_t)row * P.ldo + nb;
            if (nb + 32 <= N && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
#pragma unroll
              for (int e = 0; e < 32; e += 4)
                if (v[e] != 0.f || v[e + 1] != 0.f || v[e + 2] != 0.f || v[e + 3] != 0.f) red_add_v4(o + e, v[e], v[e + 1], v[e + 2], v[e + 3]);
            } else {
#pragma unroll
              for (int e = 0; e < 32; ++e)
                if (nb + e < N && v[e] != 0.f) atomicAdd(o + e, v[e]);
            }
          }
        }
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&acc_empty);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ first conv layer (1-3 input channels)
struct Conv0Prob {
  float* out;            // pooled output [B, H/2, W/2, cout] (tf32-rounded)
  unsigned char* idx;    // winner codes, same shape
  const float* wm;       // weight mirror [cout, 9 * cin] (tf32-rounded)
  float* gw;             // gradient mirror of the weight, same layout
  const float* dpm;      // masked gradient of the pooled output (conv0_wgrad)
  long long bias_off;
  int batch, H, cout, tiles;
};

constexpr int C0_THREADS = 288;  // warps 0-3 producers, warp 4 MMA, warps 5-8 epilogue
constexpr int C0_W = 32;         // image width this kernel is laid out for (a tile = 4 image rows = 128 pixels)
constexpr int C0_ROW = 48;       // staged row stride (floats): columns 4..35 hold the image row (16-byte aligned), 3 and 36 the zero halo;
                                 // 48 = 16 mod 32 keeps the two image rows a warp reads in disjoint banks
constexpr int C0_SIN = 6 * C0_ROW;  // staged input rows per channel: 4 + halo

__device__ __forceinline__ void named_bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }

// tile row r <-> pixel (hh, w) of the 4 x 32 tile such that a 2x2 pooling window sits inside ONE warp at lanes l, l^1, l^16
__device__ __forceinline__ void c0_row_pixel(int r, int& hh, int& w) {
  const int q = r >> 5, l = r & 31;
  hh = ((q >> 1) << 1) | (l >> 4);
  w = ((q & 1) << 4) | (l & 15);
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, bool valid) {
  const uint32_t sz = valid ? 16u : 0u;  // src-size 0: the destination is zero-filled, nothing is read
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(tc::smem_u32(smem_dst)), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Asynchronous copy (cp.async) of the tile's input rows h0-1 .. h0+4 of every channel into the staging buffer (interior columns
// 4..35; the border columns 3 and 36 stay zero from the prologue).  The batch row -> data-set sample lookup goes through the permutation.
__device__ __forceinline__ long long c0_sample(const b2_cnn_bind& bd, int step, int batch, int b) {
  long long posn = (long long)step * batch + b;
  if (posn >= bd.n_samples) posn = bd.n_samples - 1;  // rows past the data set repeat the last sample; the head ignores them
  return bd.perm != nullptr ? __ldg(bd.perm + posn) : posn;
}
// `sample` comes from c0_sample(), which the callers issue one tile earlier than the copies: the permutation lookup is a dependent
// global load in front of every address below
template <int CIN>
__device__ __forceinline__ void c0_stage_input(float* sin, const float* x, long long sample, int H, int h0, int ptid) {
  const float* xs = x + sample * (long long)(CIN * H * C0_W);
  for (int i = ptid; i < CIN * 6 * 8; i += 128) {  // 16 bytes per copy: 8 per image row
    const int c = i / 48, rr = (i % 48) >> 3, w4 = (i & 7) << 2;
    const int h = h0 - 1 + rr;
    const bool ok = h >= 0 && h < H;
    cp_async16(sin + c * C0_SIN + rr * C0_ROW + 4 + w4, ok ? xs + ((long long)c * H + h) * C0_W + w4 : xs, ok);
  }
  cp_async_commit();
}
__device__ __forceinline__ void c0_zero_borders(float* sin2, int cin, int tid, int nthreads) {
  for (int i = tid; i < 2 * cin * 12; i += nthreads) {
    const int st = i / (cin * 12), j = i % (cin * 12);
    const int c = j / 12, rr = (j % 12) >> 1, side = j & 1;
    sin2[st * (cin * C0_SIN) + c * C0_SIN + rr * C0_ROW + (side ? 36 : 3)] = 0.f;
  }
}

// one pixel's 3x3xCIN patch in (tap, channel) order, zero padded to 32, tf32-rounded, into row r of a SWIZZLE_128B tile
template <int CIN, bool MN>
__device__ __forceinline__ void c0_build_patch_row(uint8_t* tile, const float* sin, int r) {
  int hh, w;
  c0_row_pixel(r, hh, w);
  float pv[32];
#pragma unroll
  for (int k = 0; k < 32; ++k) pv[k] = (k == 9 * CIN || k == 9 * CIN + 1) ? 1.f : 0.f;  // the two ones meet the bias columns of conv0_fwd
#pragma unroll
  for (int kh = 0; kh < 3; ++kh)
#pragma unroll
    for (int kw = 0; kw < 3; ++kw)
#pragma unroll
      for (int c = 0; c < CIN; ++c) pv[(kh * 3 + kw) * CIN + c] = tf32_rna(sin[c * C0_SIN + (hh + kh) * C0_ROW + w + kw + 3]);
#pragma unroll
  for (int ch = 0; ch < 8; ++ch)
    *reinterpret_cast<float4*>(tile + (MN ? sw32_offset((uint32_t)r, (uint32_t)ch) : tc::sw128_offset((uint32_t)r, (uint32_t)ch))) = make_float4(pv[4 * ch], pv[4 * ch + 1], pv[4 * ch + 2], pv[4 * ch + 3]);
}

template <int CIN>
__global__ void __launch_bounds__(C0_THREADS, 2) conv0_fwd_kernel(const Conv0Prob* __restrict__ probs, const b2_cnn_bind* __restrict__ binds, int step) {
  const Conv0Prob& P = probs[blockIdx.y];
  const b2_cnn_bind& bd = binds[blockIdx.y];
  if (!slot_live(bd, step, P.batch)) return;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[2], empty_bar[2], acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_base_s;
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_tiles = base;                   // 2 x 16 KB
  uint8_t* b_tile = base + 2 * A_BYTES;      // cout x 128 B (<= 8 KB)
  float* sin = reinterpret_cast<float*>(base + 2 * A_BYTES + 8192);  // 2 x CIN x 6 x 34
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int cout = P.cout, H = P.H, tiles_per_img = H / 4;
  constexpr int KD = 9 * CIN;

  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&full_bar[s], 4);  // one arrival per producer warp
      tc::mbar_init(&empty_bar[s], 1);
      tc::mbar_init(&acc_full[s], 1);
      tc::mbar_init(&acc_empty[s], 4);
    }
    tc::mbar_fence_init();
  }
  if (warp == 4) tc::tmem_alloc<128>(&tmem_base_s);
  c0_zero_borders(sin, CIN, tid, C0_THREADS);
  // weights [cout, KD] -> K-major swizzled rows padded to 32.  Columns KD and KD+1 carry the bias (tf32 hi / lo parts) against two
  // columns of ones in every patch row: the tensor core adds the bias, to fp32 accuracy, and the epilogue does not load it.
  for (int i = tid; i < cout * 32; i += C0_THREADS) {
    const int rowi = i >> 5, k = i & 31;
    float val = 0.f;
    if (k < KD) {
      val = P.wm[rowi * KD + k];
    } else if (k < KD + 2) {
      const float bv = bd.slab[P.bias_off + rowi], bh = tf32_hi(bv);
      val = k == KD ? bh : tf32_rna(bv - bh);
    }
    *reinterpret_cast<float*>(b_tile + tc::sw128_offset((uint32_t)rowi, (uint32_t)(k >> 2)) + (k & 3) * 4) = val;
  }
  tc::fence_proxy_async();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = tmem_base_s;

  if (warp < 4) {
    // ================================ producers: gather the batch rows by permutation index, build patch rows
    // the input rows of tile i+1 are in flight (cp.async) while the patch rows of tile i are built
    uint32_t it = 0;
    const float* xin = bd.x;
    long long smp_next = 0;
    if ((int)blockIdx.x < P.tiles) {
      c0_stage_input<CIN>(sin, xin, c0_sample(bd, step, P.batch, blockIdx.x / tiles_per_img), H, (blockIdx.x % tiles_per_img) * 4, tid);
      if ((int)(blockIdx.x + gridDim.x) < P.tiles) smp_next = c0_sample(bd, step, P.batch, (blockIdx.x + gridDim.x) / tiles_per_img);
    }
    for (int t = blockIdx.x; t < P.tiles; t += gridDim.x, ++it) {
      const int s = it & 1;
      float* sn = sin + s * (CIN * C0_SIN);
      cp_async_wait_all();
      named_bar_sync(1, 128);  // everybody's copies have landed, and everybody is done reading the other staging buffer
      const int tn = t + gridDim.x, tnn = tn + gridDim.x;
      if (tn < P.tiles) c0_stage_input<CIN>(sin + (s ^ 1) * (CIN * C0_SIN), xin, smp_next, H, (tn % tiles_per_img) * 4, tid);
      if (tnn < P.tiles) smp_next = c0_sample(bd, step, P.batch, tnn / tiles_per_img);
      tc::mbar_wait(&empty_bar[s], ((it >> 1) & 1u) ^ 1u);
      c0_build_patch_row<CIN, false>(a_tiles + s * A_BYTES, sn, tid);
      // every lane publishes its stores to the async proxy, the warp meets, one lane arrives (4 barrier updates per tile, not 128)
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&full_bar[s]);
    }
  } else if (warp == 4) {
    if (tc::elect_one()) {
      const uint32_t idesc = tc::make_idesc(tc::FMT_TF32, BM, (uint32_t)cout);
      const uint32_t b_addr = tc::smem_u32(b_tile);
      uint32_t it = 0;
      for (int t = blockIdx.x; t < P.tiles; t += gridDim.x, ++it) {
        const int s = it & 1;
        tc::mbar_wait(&acc_empty[s], ((it >> 1) & 1u) ^ 1u);
        tc::mbar_wait(&full_bar[s], (it >> 1) & 1u);
        tc::fence_after_sync();
        const uint32_t a_addr = tc::smem_u32(a_tiles + s * A_BYTES);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) tc::mma_tf32(tmem_base + s * 64, desc_k(a_addr + ks * 32), desc_k(b_addr + ks * 32), idesc, ks != 0 ? 1u : 0u);
        tc::mma_commit(&empty_bar[s]);
        tc::mma_commit(&acc_full[s]);
      }
    }
  } else {
    // ================================ epilogue: bias + ReLU + 2x2 max-pool + winner code
    const int q = warp & 3;
    const int pos = ((lane >> 4) << 1) | (lane & 1);
    uint32_t it = 0;
    for (int t = blockIdx.x; t < P.tiles; t += gridDim.x, ++it) {
      const int s = it & 1;
      const int b = t / tiles_per_img, h0 = (t % tiles_per_img) * 4;
      tc::mbar_wait(&acc_full[s], (it >> 1) & 1u);
      tc::fence_after_sync();
      for (int cb = 0; cb < cout / 32; ++cb) {
        float v[32];
        tc::tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + s * 64 + cb * 32, v);
        // The bias came in through the GEMM (two spare K columns, see the prologue).  The stored activation is tf32-rounded (13 low
        // bits zero) and non-negative after the ReLU, so value and window position travel as ONE unsigned key: bits | (3 - pos).
        // A 2x2 max with "lowest position wins a tie" is then two shuffles and two integer maxima per column.
        uint32_t key[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          uint32_t k = tf32_rna_bits_nonneg(fmaxf(v[e], 0.f)) | (uint32_t)(3 - pos);
          k = max(k, __shfl_xor_sync(0xffffffffu, k, 1));
          key[e] = max(k, __shfl_xor_sync(0xffffffffu, k, 16));
        }
        // the four lanes of a window share the rest: the lane at position p stores columns 8p .. 8p+7 of the pooled row
        {
          const int ph = (h0 >> 1) + (q >> 1), pw = ((q & 1) << 3) + ((lane & 15) >> 1);
          const int64_t at = (((int64_t)b * (H >> 1) + ph) * (C0_W >> 1) + pw) * cout + cb * 32 + 8 * pos;
          uint32_t r[8], cw[2] = {0u, 0u};
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t k = pick8(key, pos, j);
            r[j] = k & ~3u;
            cw[j >> 2] |= ((3u - (k & 3u)) | (k > 3u ? 4u : 0u)) << (8 * (j & 3));
          }
          store_row8(P.out + at, r);
          *reinterpret_cast<uint2*>(P.idx + at) = make_uint2(cw[0], cw[1]);
        }
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&acc_empty[s]);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 4) tc::tmem_dealloc<128>(tmem_base);
}

// dW0[co, k] = sum_pixels dz[pixel, co] * patch[pixel, k]:  M = 64 (co; MN-major A, the block beyond cout stays zero), N = 32 (k),
// K = pixels.  With M = 64 the accumulator row m sits in TMEM lane 32 * (m / 16) + m % 16: epilogue warp q reads rows 16q .. 16q+15.
constexpr int C0W_A_STAGE = 2 * A_BYTES;  // 2 MN blocks x 128 pixel rows x 128 B
struct C0wPre {  // next tile's pooled gradients + winner codes of this thread's pixel, fetched one tile ahead
  float4 d[8];
  uint4 i0, i1;
};
template <int CIN, int CB>
__global__ void __launch_bounds__(C0_THREADS, 2) conv0_wgrad_kernel(const Conv0Prob* __restrict__ probs, const b2_cnn_bind* __restrict__ binds, int step) {
  const Conv0Prob& P = probs[blockIdx.y];
  const b2_cnn_bind& bd = binds[blockIdx.y];
  if (!slot_live(bd, step, P.batch)) return;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[2], empty_bar[2], done_bar;
  __shared__ uint32_t tmem_base_s;
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_tiles = base;                          // 2 x 32 KB
  uint8_t* b_tiles = base + 2 * C0W_A_STAGE;        // 2 x 16 KB
  float* sin = reinterpret_cast<float*>(base + 2 * C0W_A_STAGE + 2 * A_BYTES);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int H = P.H, tiles_per_img = H / 4;
  constexpr int cout = 32 * CB;
  constexpr int KD = 9 * CIN;

  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&full_bar[s], 4);  // one arrival per producer warp
      tc::mbar_init(&empty_bar[s], 1);
    }
    tc::mbar_init(&done_bar, 1);
    tc::mbar_fence_init();
  }
  if (warp == 4) tc::tmem_alloc<128>(&tmem_base_s);
  for (int i = tid; i < 2 * C0W_A_STAGE / 16; i += C0_THREADS) reinterpret_cast<float4*>(a_tiles)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  c0_zero_borders(sin, CIN, tid, C0_THREADS);
  tc::fence_proxy_async();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = tmem_base_s;

  if (warp < 4) {
    int hh, w;
    c0_row_pixel(tid, hh, w);
    const int pos = ((hh & 1) << 1) | (w & 1);
    auto fetch = [&](int t, C0wPre (&pre)[CB]) {
      const int b = t / tiles_per_img, h0 = (t % tiles_per_img) * 4;
      const int64_t pp = ((int64_t)b * (H >> 1) + ((h0 + hh) >> 1)) * (C0_W >> 1) + (w >> 1);
#pragma unroll
      for (int cb = 0; cb < CB; ++cb) {
        const float4* dp = reinterpret_cast<const float4*>(P.dpm + pp * cout + cb * 32);
        const uint4* ip = reinterpret_cast<const uint4*>(P.idx + pp * cout + cb * 32);
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) pre[cb].d[ch] = __ldg(dp + ch);
        pre[cb].i0 = __ldg(ip);
        pre[cb].i1 = __ldg(ip + 1);
      }
    };
    C0wPre cur[CB];
    uint32_t it = 0;
    const float* xin = bd.x;
    long long smp_next = 0;
    if ((int)blockIdx.x < P.tiles) {
      c0_stage_input<CIN>(sin, xin, c0_sample(bd, step, P.batch, blockIdx.x / tiles_per_img), H, (blockIdx.x % tiles_per_img) * 4, tid);
      if ((int)(blockIdx.x + gridDim.x) < P.tiles) smp_next = c0_sample(bd, step, P.batch, (blockIdx.x + gridDim.x) / tiles_per_img);
      fetch(blockIdx.x, cur);
    }
    for (int t = blockIdx.x; t < P.tiles; t += gridDim.x, ++it) {
      const int s = it & 1;
      float* sn = sin + s * (CIN * C0_SIN);
      cp_async_wait_all();
      named_bar_sync(1, 128);
      const int tn = t + gridDim.x, tnn = tn + gridDim.x;
      if (tn < P.tiles) c0_stage_input<CIN>(sin + (s ^ 1) * (CIN * C0_SIN), xin, smp_next, H, (tn % tiles_per_img) * 4, tid);
      if (tnn < P.tiles) smp_next = c0_sample(bd, step, P.batch, tnn / tiles_per_img);
      tc::mbar_wait(&empty_bar[s], ((it >> 1) & 1u) ^ 1u);
      // this pixel's dz row: the (already ReLU-masked) pooled gradient goes to the window's winner only
      uint8_t* arow = a_tiles + s * C0W_A_STAGE;
#pragma unroll
      for (int cb = 0; cb < CB; ++cb) {
        const uint32_t code[8] = {cur[cb].i0.x, cur[cb].i0.y, cur[cb].i0.z, cur[cb].i0.w, cur[cb].i1.x, cur[cb].i1.y, cur[cb].i1.z, cur[cb].i1.w};
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          const float4 d = cur[cb].d[ch];
          float4 o;
          o.x = (((code[ch] >> 0) & 3u) == (uint32_t)pos) ? d.x : 0.f;
          o.y = (((code[ch] >> 8) & 3u) == (uint32_t)pos) ? d.y : 0.f;
          o.z = (((code[ch] >> 16) & 3u) == (uint32_t)pos) ? d.z : 0.f;
          o.w = (((code[ch] >> 24) & 3u) == (uint32_t)pos) ? d.w : 0.f;
          *reinterpret_cast<float4*>(arow + cb * A_BYTES + sw32_offset((uint32_t)tid, (uint32_t)ch)) = o;
        }
      }
      if (tn < P.tiles) fetch(tn, cur);  // in flight while the patch row is built and the next input rows arrive
      c0_build_patch_row<CIN, true>(b_tiles + s * A_BYTES, sn, tid);
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&full_bar[s]);
    }
  } else if (warp == 4) {
    if (tc::elect_one()) {
      const uint32_t idesc = tc::make_idesc(tc::FMT_TF32, 64, 32) | (1u << 15) | (1u << 16);
      uint32_t it = 0;
      for (int t = blockIdx.x; t < P.tiles; t += gridDim.x, ++it) {
        const int s = it & 1;
        tc::mbar_wait(&full_bar[s], (it >> 1) & 1u);
        tc::fence_after_sync();
        const uint32_t a_addr = tc::smem_u32(a_tiles + s * C0W_A_STAGE), b_addr = tc::smem_u32(b_tiles + s * A_BYTES);
        // the tiles rotate over 4 accumulators (summed in the epilogue): 4x shorter fp32 accumulation chains in the tensor core
        const uint32_t d_tmem = tmem_base + (it & 3u) * 32u;
#pragma unroll
        for (int ks = 0; ks < 16; ++ks) tc::mma_tf32(d_tmem, desc_mn(a_addr + ks * 1024, A_BYTES), desc_mn(b_addr + ks * 1024, A_BYTES), idesc, (it >= 4u || ks != 0) ? 1u : 0u);
        tc::mma_commit(&empty_bar[s]);
      }
      tc::mma_commit(&done_bar);
    }
  } else {
    const int q = warp & 3;
    if (16 * q < cout) {  // warp-uniform; lanes 16-31 of the quarter hold nothing for M = 64
      tc::mbar_wait(&done_bar, 0);
      tc::fence_after_sync();
      const int my_tiles = (P.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
      float v[32];
      tc::tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16), v);
      for (int a = 1; a < 4 && a < my_tiles; ++a) {
        float t[32];
        tc::tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(a * 32), t);
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] += t[e];
      }
      if (lane < 16) {
        float* g = P.gw + (int64_t)(q * 16 + lane) * KD;
#pragma unroll
        for (int k = 0; k < KD; ++k)
          if (v[k] != 0.f) atomicAdd(g + k, v[k]);
      }
      tc::fence_before_sync();
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 4) tc::tmem_dealloc<128>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ head: fc2 + loss + their gradients, fp32
struct HeadProb {
  float* z1;       // [B, hidden] fc1 pre-activations (accumulated by the split-K GEMM; cleared here for the next step)
  float* dh_hi;    // [B, hidden] gradient w.r.t. z1, tf32 hi / lo parts
  float* dh_lo;
  float* gm;       // gradient mirror (slab offsets)
  long long off_fc1_b, off_fc2_w, off_fc2_b;
  int batch, hidden, classes, pad0;
};
constexpr int HD_ROWS = 16;
constexpr int HD_THREADS = 256;

__global__ void __launch_bounds__(HD_THREADS) head_kernel(const HeadProb* __restrict__ probs, const b2_cnn_bind* __restrict__ binds, int step, int train) {
  const HeadProb& P = probs[blockIdx.y];
  const b2_cnn_bind& bd = binds[blockIdx.y];
  if (!slot_live(bd, step, P.batch)) return;
  extern __shared__ float hsm[];
  const int Hd = P.hidden, Cn = P.classes;
  float* h1 = hsm;                       // [16][Hd]
  float* w2 = h1 + HD_ROWS * Hd;         // [Cn][Hd + 1]
  float* dl = w2 + Cn * (Hd + 1);        // [16][16]
  __shared__ float s_red[2];
  const int tid = threadIdx.x;
  const int s0 = blockIdx.x * HD_ROWS;
  const float* b1 = bd.slab + P.off_fc1_b;
  const float* W2 = bd.slab + P.off_fc2_w;
  const float* b2 = bd.slab + P.off_fc2_b;
  if (tid < 2) s_red[tid] = 0.f;
  // all global loads of a trip are issued before the first use: a load -> store -> load chain (z1 is cleared in place) would
  // cost one memory latency per element
  {
    float* zrow = P.z1 + (int64_t)s0 * Hd;  // the 16 rows of this CTA are contiguous
    for (int i0 = 0; i0 < HD_ROWS * Hd; i0 += 8 * HD_THREADS) {
      float zv[8], bv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * HD_THREADS + tid;
        const bool ok = i < HD_ROWS * Hd;
        zv[u] = ok ? zrow[i] : 0.f;
        bv[u] = ok ? __ldg(b1 + i % Hd) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * HD_THREADS + tid;
        if (i < HD_ROWS * Hd) {
          h1[i] = fmaxf(zv[u] + bv[u], 0.f);
          zrow[i] = 0.f;
        }
      }
    }
    for (int i0 = 0; i0 < Cn * Hd; i0 += 8 * HD_THREADS) {
      float wv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * HD_THREADS + tid;
        wv[u] = i < Cn * Hd ? __ldg(W2 + i) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * HD_THREADS + tid;
        if (i < Cn * Hd) w2[(i / Hd) * (Hd + 1) + (i % Hd)] = wv[u];
      }
    }
  }
  __syncthreads();
  // logits: thread -> (sample, class); 16-lane groups share a sample
  const int s = tid >> 4, c = tid & 15;
  float logit = -INFINITY;
  if (c < Cn) {
    float acc = __ldg(b2 + c);
    const float* hr = h1 + s * Hd;
    const float* wr = w2 + c * (Hd + 1);
    for (int j = 0; j < Hd; ++j) acc = fmaf(hr[j], wr[j], acc);
    logit = acc;
  }
  float mx = logit;
  int amax = c < Cn ? c : 0x7fffffff;
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const int oa = __shfl_xor_sync(0xffffffffu, amax, o);
    if (om > mx || (om == mx && oa < amax)) { mx = om; amax = oa; }
  }
  const float ex = c < Cn ? __expf(logit - mx) : 0.f;
  float se = ex;
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
  const long long posn = (long long)step * P.batch + s0 + s;
  const bool valid = posn < bd.n_samples;
  int label = -1;
  if (valid) {
    const long long sample = bd.perm != nullptr ? bd.perm[posn] : posn;
    label = (int)bd.y[sample];
  }
  const float lab_logit = __shfl_sync(0xffffffffu, logit, (tid & 16) + (label >= 0 && label < 16 ? label : 0));
  const long long left = (long long)bd.n_samples - (long long)step * P.batch;
  const float scale = 1.f / (float)(left < P.batch ? left : P.batch);
  if (c == 0 && valid) {
    const float nll = mx + __logf(se) - lab_logit;
    if (train) {
      atomicAdd(&s_red[0], nll * scale);
    } else {
      atomicAdd(&s_red[0], nll);
      atomicAdd(&s_red[1], amax == label ? 1.f : 0.f);
    }
  }
  if (train) dl[s * 16 + c] = (valid && c < Cn) ? (ex / se - (c == label ? 1.f : 0.f)) * scale : 0.f;
  __syncthreads();
  if (tid == 0) {
    if (train) {
      if (s_red[0] != 0.f) atomicAdd(bd.losses + step, s_red[0]);
    } else {
      atomicAdd(bd.losses, s_red[0]);
      atomicAdd(bd.losses + 1, s_red[1]);
    }
  }
  if (!train) return;
  if (tid < Cn) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < HD_ROWS; ++r) t += dl[r * 16 + tid];
    atomicAdd(P.gm + P.off_fc2_b + tid, t);
  }
  for (int j = tid; j < Hd; j += HD_THREADS) {
    float hcol[HD_ROWS];
#pragma unroll
    for (int r = 0; r < HD_ROWS; ++r) hcol[r] = h1[r * Hd + j];
    for (int cc = 0; cc < Cn; ++cc) {
      float acc = 0.f;
#pragma unroll
      for (int r = 0; r < HD_ROWS; ++r) acc = fmaf(dl[r * 16 + cc], hcol[r], acc);
      atomicAdd(P.gm + P.off_fc2_w + (int64_t)cc * Hd + j, acc);
    }
    float colsum = 0.f;
#pragma unroll
    for (int r = 0; r < HD_ROWS; ++r) {
      float d = 0.f;
      for (int cc = 0; cc < Cn; ++cc) d = fmaf(dl[r * 16 + cc], w2[cc * (Hd + 1) + j], d);
      d = hcol[r] > 0.f ? d : 0.f;
      const float hi = tf32_hi(d);
      P.dh_hi[(int64_t)(s0 + r) * Hd + j] = hi;
      P.dh_lo[(int64_t)(s0 + r) * Hd + j] = tf32_rna(d - hi);
      colsum += d;
    }
    atomicAdd(P.gm + P.off_fc1_b + j, colsum);
  }
}

// ------------------------------------------------------------------------------------------------ Adam + mirror maintenance
// The slab keeps PyTorch's layouts ([co][ci][tap]); the GEMMs read / write mirrors in [co][tap][ci] order.  Adam walks the slab
// row by row (one output channel / unit = one row): the row's gradients come in mirror order with coalesced loads, are turned
// through shared memory, p / m / v are updated in slab order with coalesced accesses, and the new parameters go back through
// shared memory into the mirror (tf32-rounded, or split hi / lo for the 3xTF32 layers).  Everything else (biases, last layer,
// padding) is processed in 1024-element chunks.
struct TensorEnt {
  long long begin;   // first element inside the slab
  long long lo_off;  // kind 2: offset of this tensor inside the lo mirror
  int rows, ct;      // rows of ct elements (kind 0: chunks of ADAM_CHUNK elements, the last one may be short)
  int kind;          // 0 plain, 1 permuted + tf32 mirror, 2 permuted + hi / lo mirror
  int cin, taps;
  int len;           // kind 0: elements in the region
};
constexpr int ADAM_CHUNK = 1024;
constexpr int ADAM_MAX_ENT = 12;
struct AdamTab {
  int n, total_rows, max_ct, pad0;
  TensorEnt e[ADAM_MAX_ENT];
};
struct SlotBufs {
  float* wm;
  float* wlo;
  float* gm;
};

__device__ __forceinline__ int pad32(int x) { return x + (x >> 5); }

struct AdamCoef {
  float lr, wd, b1, b2, eps, inv_bc1, inv_sqrt_bc2;
  // the slot's pointers and scalars, read from the bind table once (the table lives in global memory)
  float *slab, *m, *v;
  const float *prox, *cvg, *cvl;
  float mu, cvs, mom;
  int sgd;
};
// one element of the optimiser: FedProx / Scaffold terms, then SGD (+ momentum) or Adam.  m and v are updated in place.
__device__ __forceinline__ float adam_update(const AdamCoef& c, float p, float g, float& m, float& v, float q, float cg, float cl) {
  if (c.wd != 0.f) g = fmaf(c.wd, p, g);
  if (c.prox != nullptr) g = fmaf(c.mu, p - q, g);
  if (c.cvg != nullptr) g = fmaf(c.cvs, cg - cl, g);
  if (c.sgd) {
    if (c.mom != 0.f) {
      m = fmaf(c.mom, m, g);
      g = m;
    }
    return p - c.lr * g;
  }
  m = fmaf(c.b1, m, (1.f - c.b1) * g);
  v = fmaf(c.b2, v, (1.f - c.b2) * g * g);
  return p - c.lr * (m * c.inv_bc1) / (sqrtf(v) * c.inv_sqrt_bc2 + c.eps);
}
__device__ __forceinline__ float adam_scalar(const AdamCoef& c, int64_t i, float g) {
  float m = (!c.sgd || c.mom != 0.f) ? c.m[i] : 0.f, v = c.sgd ? 0.f : c.v[i];
  const float q = c.prox != nullptr ? c.prox[i] : 0.f, cg = c.cvg != nullptr ? c.cvg[i] : 0.f, cl = c.cvg != nullptr ? c.cvl[i] : 0.f;
  const float p = adam_update(c, c.slab[i], g, m, v, q, cg, cl);
  if (!c.sgd || c.mom != 0.f) c.m[i] = m;
  if (!c.sgd) c.v[i] = v;
  c.slab[i] = p;
  return p;
}
__device__ __forceinline__ float4 ldg4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void stg4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

__global__ void __launch_bounds__(256) adam_grouped_kernel(const SlotBufs* __restrict__ bufs, const b2_cnn_bind* __restrict__ binds, const __grid_constant__ AdamTab tab, int step, int batch,
                                                           float b1, float b2, float eps, int refresh_only) {
  const b2_cnn_bind& bd = binds[blockIdx.y];
  if (refresh_only ? bd.active == 0 : !slot_live(bd, step, batch)) return;
  extern __shared__ float asm_[];
  float* sg = asm_;
  float* sp = asm_ + pad32(tab.max_ct) + 1;
  const SlotBufs sb = bufs[blockIdx.y];
  const bool cv = bd.cv_global != nullptr && bd.cv_local != nullptr;
  AdamCoef c{bd.lr, bd.weight_decay, b1, b2, eps, 1.f, 1.f, bd.slab, bd.m, bd.v, bd.prox_ref, cv ? bd.cv_global : nullptr, cv ? bd.cv_local : nullptr,
             bd.prox_mu, bd.cv_scale, bd.momentum, bd.sgd};
  if (!refresh_only) {
    const float st = bd.adam_state[0] + (float)(step + 1);
    c.inv_bc1 = 1.f / (1.f - powf(b1, st));
    c.inv_sqrt_bc2 = rsqrtf(1.f - powf(b2, st));
  }
  const int tid = threadIdx.x;
  for (int row = blockIdx.x; row < tab.total_rows; row += gridDim.x) {
    int t = 0, r = row;
    while (t + 1 < tab.n && r >= tab.e[t].rows) {
      r -= tab.e[t].rows;
      ++t;
    }
    const TensorEnt& en = tab.e[t];
    if (en.kind == 0) {
      const int64_t i0 = en.begin + (int64_t)r * ADAM_CHUNK;
      int cnt = en.len - r * ADAM_CHUNK;
      cnt = cnt > ADAM_CHUNK ? ADAM_CHUNK : cnt;
      for (int l = tid; l < cnt; l += 256) {
        const int64_t i = i0 + l;
        if (!refresh_only) adam_scalar(c, i, sb.gm[i]);
        sb.gm[i] = 0.f;
      }
      continue;
    }
    const int ct = en.ct, taps = en.taps, cin = en.cin;
    const int64_t base = en.begin + (int64_t)r * ct;
    if (((en.begin | en.lo_off | (long long)ct) & 3) == 0) {
      // ---- 16-byte path: every thread owns 4 consecutive slab elements (one float4 of p / m / v and of the mirrors)
      for (int l = 4 * tid; l < ct; l += 4 * 256) {
        if (!refresh_only) {
          const float4 g = ldg4(sb.gm + base + l);
          float* d = sg + pad32(l);  // l .. l+3 share one 32-element block: contiguous after padding
          d[0] = g.x, d[1] = g.y, d[2] = g.z, d[3] = g.w;
        }
        stg4(sb.gm + base + l, make_float4(0.f, 0.f, 0.f, 0.f));
      }
      __syncthreads();
      for (int l = 4 * tid; l < ct; l += 4 * 256) {
        const int64_t i = base + l;
        float4 p4 = ldg4(c.slab + i), m4, v4, q4, cg4, cl4;
        m4 = v4 = q4 = cg4 = cl4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!refresh_only) {
          if (!c.sgd || c.mom != 0.f) m4 = ldg4(c.m + i);
          if (!c.sgd) v4 = ldg4(c.v + i);
          if (c.prox != nullptr) q4 = ldg4(c.prox + i);
          if (c.cvg != nullptr) cg4 = ldg4(c.cvg + i), cl4 = ldg4(c.cvl + i);
        }
        int ci = l / taps, tap = l - ci * taps;
        float* pp = &p4.x;
        float *mp = &m4.x, *vp = &v4.x;
        const float *qp = &q4.x, *cgp = &cg4.x, *clp = &cl4.x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j = pad32(tap * cin + ci);
          if (!refresh_only) pp[e] = adam_update(c, pp[e], sg[j], mp[e], vp[e], qp[e], cgp[e], clp[e]);
          sp[j] = pp[e];
          if (++tap == taps) tap = 0, ++ci;
        }
        if (!refresh_only) {
          if (!c.sgd || c.mom != 0.f) stg4(c.m + i, m4);
          if (!c.sgd) stg4(c.v + i, v4);
          stg4(c.slab + i, p4);
        }
      }
      __syncthreads();
      float* lo = sb.wlo + en.lo_off + (int64_t)r * ct;
      for (int l = 4 * tid; l < ct; l += 4 * 256) {
        const float* s4 = sp + pad32(l);
        const float p0 = s4[0], p1 = s4[1], p2 = s4[2], p3 = s4[3];
        if (en.kind == 1) {
          stg4(sb.wm + base + l, make_float4(tf32_rna(p0), tf32_rna(p1), tf32_rna(p2), tf32_rna(p3)));
        } else {
          const float4 hi = make_float4(tf32_hi(p0), tf32_hi(p1), tf32_hi(p2), tf32_hi(p3));
          stg4(sb.wm + base + l, hi);
          stg4(lo + l, make_float4(tf32_rna(p0 - hi.x), tf32_rna(p1 - hi.y), tf32_rna(p2 - hi.z), tf32_rna(p3 - hi.w)));
        }
      }
      __syncthreads();
      continue;
    }
    // ---- element-wise path (rows whose length or offset is not a multiple of 4: the first convolution)
    if (!refresh_only) {
      for (int l = tid; l < ct; l += 256) sg[pad32(l)] = sb.gm[base + l];
    }
    for (int l = tid; l < ct; l += 256) sb.gm[base + l] = 0.f;
    __syncthreads();
    for (int l = tid; l < ct; l += 256) {
      const int ci = l / taps, tap = l - ci * taps;
      const int j = pad32(tap * cin + ci);
      sp[j] = refresh_only ? c.slab[base + l] : adam_scalar(c, base + l, sg[j]);
    }
    __syncthreads();
    if (en.kind == 1) {
      for (int l = tid; l < ct; l += 256) sb.wm[base + l] = tf32_rna(sp[pad32(l)]);
    } else {
      float* lo = sb.wlo + en.lo_off + (int64_t)r * ct;
      for (int l = tid; l < ct; l += 256) {
        const float p = sp[pad32(l)], hi = tf32_hi(p);
        sb.wm[base + l] = hi;
        lo[l] = tf32_rna(p - hi);
      }
    }
    __syncthreads();
  }
}

__global__ void advance_kernel(const b2_cnn_bind* __restrict__ binds, int steps, int batch) {
  const b2_cnn_bind& bd = binds[threadIdx.x];
  if (bd.active == 0) return;
  const int mine = (bd.n_samples + batch - 1) / batch;
  bd.adam_state[0] += (float)(mine < steps ? mine : steps);
}

// ------------------------------------------------------------------------------------------------ host: plan
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

// fp32 [rows, ld] row-major; box = 32 (inner, one swizzle row) x box_rows
static int tmap2d(CUtensorMap* tm, const void* base, int rows, int cols, int ld, int box_rows, bool mn = false) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return 1;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || (ld & 3)) {
    set_error("tf32 operand must be 16-byte aligned with a row pitch that is a multiple of 4 elements (ld=%d)", ld);
    return 1;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  mn ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (fp32 2-D) failed with CUresult %d (rows=%d cols=%d ld=%d box_rows=%d)", (int)r, rows, cols, ld, box_rows);
    return 1;
  }
  return 0;
}

// fp32 NHWC activation as {C, W, H, N}; box = {32, W, tile_h, tile_n}
static int tmap_nhwc(CUtensorMap* tm, const void* base, int n_img, int h, int w, int c, int tile_h, int tile_n, bool mn = false) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return 1;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || (c & 31)) {
    set_error("tf32 conv: activation must be 16-byte aligned with channels %% 32 == 0 (c=%d)", c);
    return 1;
  }
  cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n_img};
  cuuint64_t strides[3] = {(cuuint64_t)c * 4, (cuuint64_t)w * c * 4, (cuuint64_t)h * w * c * 4};
  cuuint32_t box[4] = {(cuuint32_t)BK, (cuuint32_t)w, (cuuint32_t)tile_h, (cuuint32_t)tile_n};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  mn ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (fp32 4-D) failed with CUresult %d (n=%d h=%d w=%d c=%d tile %dx%d)", (int)r, n_img, h, w, c, tile_h, tile_n);
    return 1;
  }
  return 0;
}

static int tile_shape(int h, int w, int pixels, int* tile_h, int* tile_n, int* hw_tiles) {
  if (w <= 0 || h <= 0 || pixels % w != 0) return 1;
  const int th = pixels / w;
  if (th <= h) {
    if (h % th != 0) return 1;
    *tile_h = th; *tile_n = 1; *hw_tiles = h / th;
    return 0;
  }
  if (pixels % (h * w) != 0) return 1;
  *tile_h = h; *tile_n = pixels / (h * w); *hw_tiles = 0;
  return 0;
}

static inline int64_t up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

enum { BUF_ACT = 100, BUF_IDX = 110, BUF_DZ = 120, BUF_AHI = 1, BUF_ALO = 2, BUF_Z1 = 3, BUF_DHHI = 4, BUF_DHLO = 5, BUF_DPM0 = 6, BUF_WM = 7, BUF_WLO = 8, BUF_GM = 9 };
enum { RUN_REFRESH = 0, RUN_CONV0_FWD = 1, RUN_FC1_FWD = 2, RUN_HEAD_TRAIN = 3, RUN_FC1_WGRAD = 4, RUN_FC1_DGRAD = 5, RUN_CONV0_WGRAD = 6, RUN_ADAM = 7, RUN_HEAD_EVAL = 8,
       RUN_CONV_FWD = 10, RUN_CONV_WGRAD = 20, RUN_CONV_DGRAD = 30 };

struct SlotLayout {  // byte offsets inside one slot's workspace
  int64_t act[B2_CNN_MAX_CONV], idx[B2_CNN_MAX_CONV], dz[B2_CNN_MAX_CONV];
  int64_t a_hi, a_lo, z1, dh_hi, dh_lo, dpm0, wm, wlo, gm, total;
};

struct Plan {
  b2_cnn_desc d;
  int G, nl;
  int H[B2_CNN_MAX_CONV], Cin[B2_CNN_MAX_CONV], Cout[B2_CNN_MAX_CONV];  // conv l works on H[l] x H[l] images
  int feat, fh;                                                          // flattened features, final map height
  SlotLayout L;
  uint8_t* ws;
  int64_t slot_base;   // byte offset of slot 0 inside ws
  // device tables
  Problem* d_prob;     // [n_prob][G]
  Conv0Prob* d_c0;     // [G]
  HeadProb* d_head;    // [G]
  SlotBufs* d_bufs;    // [G]
  b2_cnn_bind* d_bind; // [G]
  b2_cnn_bind* h_bind; // pinned staging
  std::vector<Problem> h_prob;
  int n_prob;
  AdamTab tab;
  cudaStream_t side;
  cudaEvent_t ev_fork[B2_CNN_MAX_CONV + 1], ev_join[B2_CNN_MAX_CONV + 1];
  int sms;
  uint8_t* slot_ptr(int g, int64_t off) const { return ws + slot_base + (int64_t)g * L.total + off; }
  // problem indices
  int p_conv_fwd(int l) const { return l - 1; }
  int p_conv_wgrad(int l) const { return (nl - 1) + (l - 1); }
  int p_conv_dgrad(int l) const { return 2 * (nl - 1) + (l - 1); }
  int p_fc1_fwd() const { return 3 * (nl - 1); }
  int p_fc1_wgrad() const { return 3 * (nl - 1) + 1; }
  int p_fc1_dgrad() const { return 3 * (nl - 1) + 2; }
};

static int describe(const b2_cnn_desc* d, Plan* p) {
  if (d == nullptr || d->groups < 1 || d->groups > 64 || d->batch < 16 || (d->batch % HD_ROWS) || d->n_conv < 2 || d->n_conv > B2_CNN_MAX_CONV) {
    set_error("cnn plan: unsupported shape (groups in [1,64], batch %% 16 == 0, 2..4 conv blocks)");
    return 1;
  }
  if (d->image != C0_W || (d->in_ch != 1 && d->in_ch != 3)) {
    set_error("cnn plan: the first layer is laid out for 32 x 32 images with 1 or 3 channels (got %d, %d)", d->image, d->in_ch);
    return 1;
  }
  if (d->classes < 2 || d->classes > 16 || d->hidden < 64 || d->hidden > 1024 || (d->hidden % 64) || (d->slab_elems & 3)) {
    set_error("cnn plan: classes in [2,16], hidden %% 64 == 0 and <= 1024 required");
    return 1;
  }
  p->d = *d;
  p->G = d->groups;
  p->nl = d->n_conv;
  int h = d->image;
  for (int l = 0; l < d->n_conv; ++l) {
    p->H[l] = h;
    p->Cin[l] = l == 0 ? d->in_ch : d->conv_ch[l - 1];
    p->Cout[l] = d->conv_ch[l];
    if (d->conv_ch[l] % 32 || d->conv_ch[l] > 256 || (l == 0 && d->conv_ch[l] > 64)) {
      set_error("cnn plan: conv widths must be multiples of 32 (first layer <= 64, others <= 256)");
      return 1;
    }
    if (l > 0 && (h > 16 || h < 4 || ((int64_t)d->batch * h * h) % BM)) {
      set_error("cnn plan: layer %d works on %d x %d maps; 4..16 and batch * h * w %% 128 == 0 required", l, h, h);
      return 1;
    }
    h /= 2;
  }
  p->fh = h;
  p->feat = h * h * d->conv_ch[d->n_conv - 1];
  if (p->feat % 64) {
    set_error("cnn plan: flattened feature count %d must be a multiple of 64", p->feat);
    return 1;
  }
  // slot layout
  const int64_t B = d->batch;
  int64_t o = 0;
  auto take = [&](int64_t bytes) { const int64_t at = o; o = up(o + bytes, 1024); return at; };
  SlotLayout& L = p->L;
  for (int l = 0; l < d->n_conv; ++l) {
    const int64_t pooled = B * (p->H[l] / 2) * (p->H[l] / 2) * p->Cout[l];
    L.act[l] = l + 1 < d->n_conv ? take(pooled * 4) : -1;
    L.idx[l] = take(pooled);
    L.dz[l] = l > 0 ? take(B * p->H[l] * p->H[l] * p->Cout[l] * 4) : -1;
  }
  L.a_hi = take(B * p->feat * 4);
  L.a_lo = take(B * p->feat * 4);
  L.z1 = take(B * d->hidden * 4);
  L.dh_hi = take(B * d->hidden * 4);
  L.dh_lo = take(B * d->hidden * 4);
  L.dpm0 = take(B * (p->H[0] / 2) * (p->H[0] / 2) * p->Cout[0] * 4);
  L.wm = take(d->slab_elems * 4);
  L.wlo = take((int64_t)d->hidden * p->feat * 4);
  L.gm = take(d->slab_elems * 4);
  L.total = o;
  p->n_prob = 3 * (d->n_conv - 1) + 3;
  return 0;
}

static int64_t tables_bytes(const Plan& p) {
  int64_t o = 0;
  o = up(o + (int64_t)p.n_prob * p.G * sizeof(Problem), 1024);
  o = up(o + (int64_t)p.G * sizeof(Conv0Prob), 1024);
  o = up(o + (int64_t)p.G * sizeof(HeadProb), 1024);
  o = up(o + (int64_t)p.G * sizeof(SlotBufs), 1024);
  o = up(o + (int64_t)p.G * sizeof(b2_cnn_bind), 1024);
  return o;
}

template <int CONV, bool A_MN, bool B_MN, bool X3, int EPI, int BN>
static int launch_gg(const Plan& p, int pi, int step, cudaStream_t st) {
  using C = Cfg<X3, BN, cfg_mode(EPI)>;
  auto kern = ggemm_kernel<CONV, A_MN, B_MN, X3, EPI, BN>;
  static bool attr_set[64] = {false};
  static std::mutex mu;
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  {
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
      attr_set[dev] = true;
    }
  }
  const Problem& hp = p.h_prob[(size_t)pi * p.G];
  const int n_work = hp.tiles_m * hp.tiles_n * hp.splits;
  int per_slot = (p.sms * C::CTAS_PER_SM + p.G - 1) / p.G;
  if (per_slot < 8) per_slot = 8;
  const int gx = n_work < per_slot ? n_work : per_slot;
  kern<<<dim3(gx, p.G), GG_THREADS, C::SMEM, st>>>(p.d_prob + (size_t)pi * p.G, p.d_bind, step);
  B2_LAUNCH_CHECK("ggemm_kernel (tf32)");
  return 0;
}

template <int CIN, int CB>
static void launch_conv0_wgrad_t(const Plan& p, int gx, int smem, int step, cudaStream_t st) {
  conv0_wgrad_kernel<CIN, CB><<<dim3(gx, p.G), C0_THREADS, smem, st>>>(p.d_c0, p.d_bind, step);
}

static int launch_conv0(const Plan& p, bool wgrad, int step, cudaStream_t st) {
  static bool attr_set[64] = {false};
  static std::mutex mu;
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  const int sin_bytes = 2 * 3 * C0_SIN * 4;
  const int smem_f = 1024 + 2 * A_BYTES + 8192 + sin_bytes;
  const int smem_w = 1024 + 2 * C0W_A_STAGE + 2 * A_BYTES + sin_bytes;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      B2_CUDA(cudaFuncSetAttribute(conv0_fwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_f));
      B2_CUDA(cudaFuncSetAttribute(conv0_fwd_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_f));
      B2_CUDA(cudaFuncSetAttribute(conv0_wgrad_kernel<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_w));
      B2_CUDA(cudaFuncSetAttribute(conv0_wgrad_kernel<1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_w));
      B2_CUDA(cudaFuncSetAttribute(conv0_wgrad_kernel<3, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_w));
      B2_CUDA(cudaFuncSetAttribute(conv0_wgrad_kernel<3, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_w));
      attr_set[dev] = true;
    }
  }
  const int tiles = p.d.batch * p.H[0] / 4;
  int per_slot = (p.sms * 2 + p.G - 1) / p.G;
  if (per_slot < 8) per_slot = 8;
  if (wgrad && per_slot > 96) per_slot = 96;  // every CTA ends in 27 x cout atomics on the same few hundred addresses
  const int gx = tiles < per_slot ? tiles : per_slot;
  if (wgrad) {
    const int cb = p.Cout[0] / 32;
    if (p.d.in_ch == 3 && cb == 1) launch_conv0_wgrad_t<3, 1>(p, gx, smem_w, step, st);
    else if (p.d.in_ch == 3) launch_conv0_wgrad_t<3, 2>(p, gx, smem_w, step, st);
    else if (cb == 1) launch_conv0_wgrad_t<1, 1>(p, gx, smem_w, step, st);
    else launch_conv0_wgrad_t<1, 2>(p, gx, smem_w, step, st);
  } else {
    if (p.d.in_ch == 3) conv0_fwd_kernel<3><<<dim3(gx, p.G), C0_THREADS, smem_f, st>>>(p.d_c0, p.d_bind, step);
    else conv0_fwd_kernel<1><<<dim3(gx, p.G), C0_THREADS, smem_f, st>>>(p.d_c0, p.d_bind, step);
  }
  B2_LAUNCH_CHECK("conv0 kernel (tf32)");
  return 0;
}

static int launch_head(const Plan& p, int step, int train, cudaStream_t st) {
  const int smem = (HD_ROWS * p.d.hidden + p.d.classes * (p.d.hidden + 1) + HD_ROWS * 16) * 4;
  static bool attr_set[64] = {false};
  static std::mutex mu;
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  {
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      B2_CUDA(cudaFuncSetAttribute(head_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
      attr_set[dev] = true;
    }
  }
  head_kernel<<<dim3(p.d.batch / HD_ROWS, p.G), HD_THREADS, smem, st>>>(p.d_head, p.d_bind, step, train);
  B2_LAUNCH_CHECK("head_kernel");
  return 0;
}

static int launch_adam(const Plan& p, int step, int refresh_only, cudaStream_t st) {
  const int smem = (2 * (p.tab.max_ct + p.tab.max_ct / 32 + 2)) * 4;
  static bool attr_set[64] = {false};
  static std::mutex mu;
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  {
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      B2_CUDA(cudaFuncSetAttribute(adam_grouped_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
      attr_set[dev] = true;
    }
  }
  int gx = p.tab.total_rows;
  int cap = (p.sms * 6 + p.G - 1) / p.G;
  if (cap < 16) cap = 16;
  if (gx > cap) gx = cap;
  adam_grouped_kernel<<<dim3(gx, p.G), 256, smem, st>>>(p.d_bufs, p.d_bind, p.tab, step, p.d.batch, 0.9f, 0.999f, 1e-8f, refresh_only);
  B2_LAUNCH_CHECK("adam_grouped_kernel");
  return 0;
}

static int run_conv_fwd(const Plan& p, int l, int step, cudaStream_t st) { return launch_gg<1, false, false, false, EPI_POOL, 64>(p, p.p_conv_fwd(l), step, st); }
static int run_conv_wgrad(const Plan& p, int l, int step, cudaStream_t st) {
  // Default: the generic kernel's 64-column tiles.  The wide kernel (B2_T32_WIDE_WGRAD=1) moves half the bytes but measured
  // SLOWER on B200 (conv1 145 us vs 90 us, conv2 145 us vs 54 us for 8 nodes): thirteen small TMA boxes per stage and a single
  // resident CTA per SM; kept for the comparison (profiles/t32_step.md).
  static const bool wide = getenv("B2_T32_WIDE_WGRAD") != nullptr;
  if (!wide) return launch_gg<2, true, true, false, EPI_ATOMIC, 64>(p, p.p_conv_wgrad(l), step, st);
  static bool attr_set[64] = {false};
  static std::mutex mu;
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  {
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      B2_CUDA(cudaFuncSetAttribute(wgrad_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WW_SMEM));
      attr_set[dev] = true;
    }
  }
  const Problem& hp = p.h_prob[(size_t)p.p_conv_wgrad(l) * p.G];
  const int groups = (hp.N + 32 * WW_NB - 1) / (32 * WW_NB);
  const int n_work = hp.tiles_m * groups * hp.splits;
  int per_slot = (p.sms + p.G - 1) / p.G;
  if (per_slot < 4) per_slot = 4;
  const int gx = n_work < per_slot ? n_work : per_slot;
  wgrad_wide_kernel<<<dim3(gx, p.G), GG_THREADS, WW_SMEM, st>>>(p.d_prob + (size_t)p.p_conv_wgrad(l) * p.G, p.d_bind, step);
  B2_LAUNCH_CHECK("wgrad_wide_kernel");
  return 0;
}
static int run_conv_dgrad(const Plan& p, int l, int step, cudaStream_t st) {
  if (l == 1) {
    if (p.Cout[0] == 32) return launch_gg<1, false, true, false, EPI_MASK, 32>(p, p.p_conv_dgrad(l), step, st);
    return launch_gg<1, false, true, false, EPI_MASK, 64>(p, p.p_conv_dgrad(l), step, st);
  }
  return launch_gg<1, false, true, false, EPI_UNPOOL, 64>(p, p.p_conv_dgrad(l), step, st);
}

}  // namespace t32
}  // namespace b2

using namespace b2;
using namespace b2::t32;

extern "C" {

int64_t b2_cnn_workspace_bytes(const b2_cnn_desc* d) {
  Plan p{};
  if (describe(d, &p)) return -1;
  return up(tables_bytes(p), 1024) + (int64_t)p.G * p.L.total + 1024;
}

void* b2_cnn_plan_create(const b2_cnn_desc* d, void* workspace, int64_t bytes, void* stream) {
  Plan* p = new Plan();
  if (describe(d, p)) {
    delete p;
    return nullptr;
  }
  const int64_t need = up(tables_bytes(*p), 1024) + (int64_t)p->G * p->L.total + 1024;
  if (workspace == nullptr || bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 255u)) {
    set_error("cnn plan: workspace must be 256-byte aligned and hold %lld bytes (got %lld)", (long long)need, (long long)bytes);
    delete p;
    return nullptr;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  p->sms = sm_count();
  p->ws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 1023) & ~uintptr_t(1023));
  int64_t o = 0;
  p->d_prob = reinterpret_cast<Problem*>(p->ws + o); o = up(o + (int64_t)p->n_prob * p->G * sizeof(Problem), 1024);
  p->d_c0 = reinterpret_cast<Conv0Prob*>(p->ws + o); o = up(o + (int64_t)p->G * sizeof(Conv0Prob), 1024);
  p->d_head = reinterpret_cast<HeadProb*>(p->ws + o); o = up(o + (int64_t)p->G * sizeof(HeadProb), 1024);
  p->d_bufs = reinterpret_cast<SlotBufs*>(p->ws + o); o = up(o + (int64_t)p->G * sizeof(SlotBufs), 1024);
  p->d_bind = reinterpret_cast<b2_cnn_bind*>(p->ws + o); o = up(o + (int64_t)p->G * sizeof(b2_cnn_bind), 1024);
  p->slot_base = o;
  const int G = p->G, nl = p->nl, B = d->batch;
  const SlotLayout& L = p->L;
  if (cudaMemsetAsync(p->ws + p->slot_base, 0, (size_t)G * L.total, st) != cudaSuccess) {
    set_error("cnn plan: clearing the workspace failed");
    delete p;
    return nullptr;
  }
  p->h_prob.assign((size_t)p->n_prob * G, Problem{});
  std::vector<Conv0Prob> c0(G);
  std::vector<HeadProb> hd(G);
  std::vector<SlotBufs> sb(G);
  bool bad = false;
  for (int g = 0; g < G && !bad; ++g) {
    auto F = [&](int64_t off) { return reinterpret_cast<float*>(p->slot_ptr(g, off)); };
    float* wm = F(L.wm);
    float* wlo = F(L.wlo);
    float* gm = F(L.gm);
    sb[g] = SlotBufs{wm, wlo, gm};
    c0[g] = Conv0Prob{F(L.act[0]), p->slot_ptr(g, L.idx[0]), wm + d->off_conv_w[0], gm + d->off_conv_w[0], F(L.dpm0), d->off_conv_b[0], B, p->H[0], p->Cout[0], B * p->H[0] / 4};
    hd[g] = HeadProb{F(L.z1), F(L.dh_hi), F(L.dh_lo), gm, d->off_fc1_b, d->off_fc2_w, d->off_fc2_b, B, d->hidden, d->classes, 0};
    for (int l = 1; l < nl && !bad; ++l) {
      const int h = p->H[l], cin = p->Cin[l], cout = p->Cout[l], kdim = 9 * cin, pixels = B * h * h;
      const bool last = l + 1 == nl;
      float* x_in = F(L.act[l - 1]);                 // pooled output of layer l-1 = [B, h, h, cin]
      float* dz = F(L.dz[l]);                        // [B*h*h, cout]
      const float* w_m = wm + d->off_conv_w[l];      // [cout, 9*cin]
      {  // ---- forward
        Problem& P = p->h_prob[(size_t)p->p_conv_fwd(l) * G + g];
        P.M = pixels; P.N = cout; P.K = kdim; P.batch = B;
        P.tiles_m = pixels / BM; P.tiles_n = (cout + 63) / 64; P.splits = 1; P.kb_per_split = 9 * (cin / 32);
        P.cblocks = cin / 32; P.kw = 3; P.off = -1; P.sign = 1;
        bad |= tile_shape(h, h, BM, &P.tile_h, &P.tile_n, &P.hw_tiles) != 0;
        bad |= tmap_nhwc(&P.ta, x_in, B, h, h, cin, P.tile_h, P.tile_n) != 0;
        bad |= tmap2d(&P.tb, w_m, cout, kdim, kdim, 64) != 0;
        P.out = last ? F(L.a_hi) : F(L.act[l]);
        P.out_lo = last ? F(L.a_lo) : nullptr;
        P.idx = p->slot_ptr(g, L.idx[l]);
        P.bias_off = d->off_conv_b[l];
        P.ldo = cout; P.relu = 1; P.img_h = h; P.img_w = h;
      }
      {  // ---- weight gradient: dw[co, tap*cin + ci] = sum_pix dz[pix, co] x[pix + tap, ci]
        Problem& P = p->h_prob[(size_t)p->p_conv_wgrad(l) * G + g];
        P.M = cout; P.N = kdim; P.K = pixels; P.batch = B;
        P.tiles_m = (cout + BM - 1) / BM; P.tiles_n = (kdim + 63) / 64;
        const int total_k = pixels / BK;
        // K splits: as many as fill this slot's share of the chip in ONE wave (3 resident CTAs per SM), at least 4 k-blocks each.
        // (96-column tiles - 9 * cin is a whole number of them, dz is re-read 3x instead of 4.5x - fit only 2 CTAs per SM and
        // measured 1.9x SLOWER: these kernels live on the number of resident CTAs, not on bytes.)
        const int out_tiles = ((cout + BM - 1) / BM) * ((kdim + 63) / 64);
        int splits = (p->sms * 4 / G) / out_tiles;
        if (getenv("B2_T32_WIDE_WGRAD") != nullptr) {  // one CTA per (cout tile, 288-column group, K split)
          const int wgroups = (kdim + 32 * WW_NB - 1) / (32 * WW_NB) * ((cout + BM - 1) / BM);
          int want = (p->sms + G - 1) / G;
          if (want < 8) want = 8;
          splits = (want + wgroups - 1) / wgroups;
        }
        if (splits > total_k / 4) splits = total_k / 4;
        if (splits < 1) splits = 1;
        if (splits > 64) splits = 64;
        P.kb_per_split = (total_k + splits - 1) / splits;
        P.splits = (total_k + P.kb_per_split - 1) / P.kb_per_split;
        P.cblocks = cin / 32; P.kw = 3; P.off = -1; P.sign = 1; P.ncols = cin;
        bad |= tile_shape(h, h, BK, &P.tile_h, &P.tile_n, &P.hw_tiles) != 0;
        bad |= tmap2d(&P.ta, dz, pixels, cout, cout, BK, true) != 0;
        bad |= tmap_nhwc(&P.tb, x_in, B, h, h, cin, P.tile_h, P.tile_n, true) != 0;
        P.out = gm + d->off_conv_w[l];
        P.bias_off = -1; P.ldo = kdim;
      }
      {  // ---- input gradient: dx[pix, ci] = sum_{tap, co} dz[pix - tap, co] w[co, tap*cin + ci]  -> un-pooled dz of layer l-1
        Problem& P = p->h_prob[(size_t)p->p_conv_dgrad(l) * G + g];
        const int bn = (l == 1 && p->Cout[0] == 32) ? 32 : 64;
        P.M = pixels; P.N = cin; P.K = 9 * cout; P.batch = B;
        P.tiles_m = pixels / BM; P.tiles_n = (cin + bn - 1) / bn; P.splits = 1; P.kb_per_split = 9 * (cout / 32);
        P.cblocks = cout / 32; P.kw = 3; P.off = 1; P.sign = -1; P.ncols = cin;
        bad |= tile_shape(h, h, BM, &P.tile_h, &P.tile_n, &P.hw_tiles) != 0;
        bad |= tmap_nhwc(&P.ta, dz, B, h, h, cout, P.tile_h, P.tile_n) != 0;
        bad |= tmap2d(&P.tb, w_m, cout, kdim, kdim, BK, true) != 0;
        P.idx = p->slot_ptr(g, L.idx[l - 1]);
        P.colsum = gm + d->off_conv_b[l - 1];
        P.bias_off = -1; P.chans = cin;
        if (l == 1) {
          P.out = F(L.dpm0);
        } else {
          P.out = F(L.dz[l - 1]);
          P.img_h = p->H[l - 1]; P.img_w = p->H[l - 1];
        }
      }
    }
    const int feat = p->feat, hid = d->hidden, cl = p->Cout[nl - 1];
    const float* w1 = wm + d->off_fc1_w;  // [hid, feat] in (pixel, channel) feature order, tf32 hi
    {  // ---- fc1 forward: z1[B, hid] += a[B, feat] W1^T, 3xTF32, split-K
      Problem& P = p->h_prob[(size_t)p->p_fc1_fwd() * G + g];
      P.M = B; P.N = hid; P.K = feat; P.batch = B;
      P.tiles_m = (B + BM - 1) / BM; P.tiles_n = hid / 64;
      const int total_k = feat / BK;
      int splits = total_k / 8;
      if (splits < 1) splits = 1;
      if (splits > 16) splits = 16;
      P.kb_per_split = (total_k + splits - 1) / splits;
      P.splits = (total_k + P.kb_per_split - 1) / P.kb_per_split;
      bad |= tmap2d(&P.ta, F(L.a_hi), B, feat, feat, BM) != 0;
      bad |= tmap2d(&P.ta_lo, F(L.a_lo), B, feat, feat, BM) != 0;
      bad |= tmap2d(&P.tb, w1, hid, feat, feat, 64) != 0;
      bad |= tmap2d(&P.tb_lo, wlo, hid, feat, feat, 64) != 0;
      P.out = F(L.z1); P.ldo = hid; P.bias_off = -1;
    }
    {  // ---- fc1 weight gradient: dW1[hid, feat] = dh^T a  (both operands MN-major)
      Problem& P = p->h_prob[(size_t)p->p_fc1_wgrad() * G + g];
      P.M = hid; P.N = feat; P.K = B; P.batch = B;
      P.tiles_m = (hid + BM - 1) / BM; P.tiles_n = feat / 64; P.splits = 1; P.kb_per_split = (B + BK - 1) / BK;
      bad |= tmap2d(&P.ta, F(L.dh_hi), B, hid, hid, BK, true) != 0;
      bad |= tmap2d(&P.ta_lo, F(L.dh_lo), B, hid, hid, BK, true) != 0;
      bad |= tmap2d(&P.tb, F(L.a_hi), B, feat, feat, BK, true) != 0;
      bad |= tmap2d(&P.tb_lo, F(L.a_lo), B, feat, feat, BK, true) != 0;
      P.out = gm + d->off_fc1_w; P.ldo = feat; P.bias_off = -1;
    }
    {  // ---- fc1 input gradient: da[B, feat] = dh W1, un-pooled into dz of the last conv layer
      Problem& P = p->h_prob[(size_t)p->p_fc1_dgrad() * G + g];
      P.M = B; P.N = feat; P.K = hid; P.batch = B;
      P.tiles_m = (B + BM - 1) / BM; P.tiles_n = feat / 64; P.splits = 1; P.kb_per_split = hid / BK;
      bad |= tmap2d(&P.ta, F(L.dh_hi), B, hid, hid, BM) != 0;
      bad |= tmap2d(&P.ta_lo, F(L.dh_lo), B, hid, hid, BM) != 0;
      bad |= tmap2d(&P.tb, w1, hid, feat, feat, BK, true) != 0;
      bad |= tmap2d(&P.tb_lo, wlo, hid, feat, feat, BK, true) != 0;
      P.out = F(L.dz[nl - 1]);
      P.idx = p->slot_ptr(g, L.idx[nl - 1]);
      P.colsum = gm + d->off_conv_b[nl - 1];
      P.bias_off = -1; P.chans = cl; P.img_h = p->H[nl - 1]; P.img_w = p->H[nl - 1];
    }
  }
  if (bad) {
    delete p;
    return nullptr;
  }
  // Adam / mirror table: the permuted tensors row by row, everything between them in chunks
  AdamTab& tab = p->tab;
  tab = AdamTab{};
  struct Perm { int64_t begin, numel; int rows, ct, kind, cin, taps; };
  std::vector<Perm> perms;
  for (int l = 0; l < nl; ++l) perms.push_back(Perm{d->off_conv_w[l], (int64_t)p->Cout[l] * p->Cin[l] * 9, p->Cout[l], p->Cin[l] * 9, 1, p->Cin[l], 9});
  perms.push_back(Perm{d->off_fc1_w, (int64_t)d->hidden * p->feat, d->hidden, p->feat, 2, p->Cout[nl - 1], p->fh * p->fh});
  for (size_t a = 0; a < perms.size(); ++a)  // slab order
    for (size_t b = a + 1; b < perms.size(); ++b)
      if (perms[b].begin < perms[a].begin) std::swap(perms[a], perms[b]);
  int64_t cursor = 0;
  auto plain = [&](int64_t from, int64_t to) {
    if (to <= from || tab.n >= ADAM_MAX_ENT) return;
    TensorEnt& e = tab.e[tab.n++];
    e = TensorEnt{};
    e.begin = from; e.kind = 0; e.len = (int)(to - from); e.rows = (e.len + ADAM_CHUNK - 1) / ADAM_CHUNK; e.ct = ADAM_CHUNK;
  };
  for (const Perm& q : perms) {
    plain(cursor, q.begin);
    if (tab.n >= ADAM_MAX_ENT) break;
    TensorEnt& e = tab.e[tab.n++];
    e = TensorEnt{};
    e.begin = q.begin; e.rows = q.rows; e.ct = q.ct; e.kind = q.kind; e.cin = q.cin; e.taps = q.taps; e.lo_off = 0;
    if (q.ct > tab.max_ct) tab.max_ct = q.ct;
    cursor = q.begin + q.numel;
  }
  plain(cursor, d->slab_elems);
  for (int i = 0; i < tab.n; ++i) tab.total_rows += tab.e[i].rows;
  if (tab.max_ct > 15000 || (perms.size() * 2 + 1 > (size_t)ADAM_MAX_ENT)) {
    set_error("cnn plan: layer rows of %d elements exceed the optimizer kernel's shared memory", tab.max_ct);
    delete p;
    return nullptr;
  }
  bool fail = false;
  fail |= cudaMemcpyAsync(p->d_prob, p->h_prob.data(), p->h_prob.size() * sizeof(Problem), cudaMemcpyHostToDevice, st) != cudaSuccess;
  fail |= cudaMemcpyAsync(p->d_c0, c0.data(), c0.size() * sizeof(Conv0Prob), cudaMemcpyHostToDevice, st) != cudaSuccess;
  fail |= cudaMemcpyAsync(p->d_head, hd.data(), hd.size() * sizeof(HeadProb), cudaMemcpyHostToDevice, st) != cudaSuccess;
  fail |= cudaMemcpyAsync(p->d_bufs, sb.data(), sb.size() * sizeof(SlotBufs), cudaMemcpyHostToDevice, st) != cudaSuccess;
  fail |= cudaMemsetAsync(p->d_bind, 0, (size_t)G * sizeof(b2_cnn_bind), st) != cudaSuccess;
  fail |= cudaStreamSynchronize(st) != cudaSuccess;  // the host vectors above go out of scope
  fail |= cudaMallocHost(reinterpret_cast<void**>(&p->h_bind), (size_t)G * sizeof(b2_cnn_bind)) != cudaSuccess;
  fail |= cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking) != cudaSuccess;
  for (int i = 0; i <= B2_CNN_MAX_CONV; ++i) {
    fail |= cudaEventCreateWithFlags(&p->ev_fork[i], cudaEventDisableTiming) != cudaSuccess;
    fail |= cudaEventCreateWithFlags(&p->ev_join[i], cudaEventDisableTiming) != cudaSuccess;
  }
  if (fail) {
    cuda_fail(cudaGetLastError(), "cnn plan: table upload / stream set-up");
    delete p;
    return nullptr;
  }
  return p;
}

int b2_cnn_plan_destroy(void* plan) {
  Plan* p = static_cast<Plan*>(plan);
  if (p == nullptr) return 0;
  cudaFreeHost(p->h_bind);
  cudaStreamDestroy(p->side);
  for (int i = 0; i <= B2_CNN_MAX_CONV; ++i) {
    cudaEventDestroy(p->ev_fork[i]);
    cudaEventDestroy(p->ev_join[i]);
  }
  delete p;
  return 0;
}

int b2_cnn_plan_bind(void* plan, const b2_cnn_bind* host_binds, void* stream) {
  Plan* p = static_cast<Plan*>(plan);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // the previous copy out of the pinned staging buffer must have left it before it is overwritten
  B2_CUDA(cudaStreamSynchronize(st));
  memcpy(p->h_bind, host_binds, (size_t)p->G * sizeof(b2_cnn_bind));
  B2_CUDA(cudaMemcpyAsync(p->d_bind, p->h_bind, (size_t)p->G * sizeof(b2_cnn_bind), cudaMemcpyHostToDevice, st));
  return 0;
}

int b2_cnn_plan_refresh(void* plan, void* stream) {
  Plan* p = static_cast<Plan*>(plan);
  return launch_adam(*p, 0, 1, static_cast<cudaStream_t>(stream));
}

int b2_cnn_plan_advance(void* plan, int steps, void* stream) {
  Plan* p = static_cast<Plan*>(plan);
  advance_kernel<<<1, p->G, 0, static_cast<cudaStream_t>(stream)>>>(p->d_bind, steps, p->d.batch);
  B2_LAUNCH_CHECK("advance_kernel");
  return 0;
}

static int forward(Plan& p, int step, cudaStream_t st) {
  if (launch_conv0(p, false, step, st)) return -1;
  for (int l = 1; l < p.nl; ++l)
    if (run_conv_fwd(p, l, step, st)) return -1;
  if (launch_gg<0, false, false, true, EPI_ATOMIC, 64>(p, p.p_fc1_fwd(), step, st)) return -1;
  return p.nl + 1;
}

int b2_cnn_plan_eval_step(void* plan, int step, void* stream) {
  Plan* p = static_cast<Plan*>(plan);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int n = forward(*p, step, st);
  if (n < 0) return -1;
  if (launch_head(*p, step, 0, st)) return -1;
  return n + 1;
}

int b2_cnn_plan_train_step(void* plan, int step, void* stream) {
  Plan* pp = static_cast<Plan*>(plan);
  Plan& p = *pp;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int n = forward(p, step, st);
  if (n < 0) return -1;
  if (launch_head(p, step, 1, st)) return -1;
  ++n;
  // weight-gradient GEMMs on the side stream: a parallel branch next to the input-gradient chain.  B2_T32_WGRAD_BRANCH=0 makes it one
  // chain: on random data at G = 8 the chain looked better (336 vs 351 us per step, tools/t32_step_profile.py), inside the federation
  // the branch wins (epoch graph 34.9 vs 36.2 ms) - the branch stays.
  static const char* branch_env = getenv("B2_T32_WGRAD_BRANCH");
  const bool branch = branch_env != nullptr ? branch_env[0] != '0' : true;
  cudaStream_t ws = branch ? p.side : st;
  auto fork = [&](int i) -> int {
    if (!branch) return 0;
    if (cudaEventRecord(p.ev_fork[i], st) != cudaSuccess || cudaStreamWaitEvent(p.side, p.ev_fork[i], 0) != cudaSuccess) return cuda_fail(cudaGetLastError(), "fork to the wgrad stream");
    return 0;
  };
  if (fork(0)) return -1;
  if (launch_gg<0, true, true, true, EPI_STORE, 64>(p, p.p_fc1_wgrad(), step, ws)) return -1;
  if (launch_gg<0, false, true, true, EPI_UNPOOL, 64>(p, p.p_fc1_dgrad(), step, st)) return -1;
  n += 2;
  for (int l = p.nl - 1; l >= 1; --l) {
    if (fork(l)) return -1;
    if (run_conv_wgrad(p, l, step, ws)) return -1;
    if (run_conv_dgrad(p, l, step, st)) return -1;
    n += 2;
  }
  if (launch_conv0(p, true, step, st)) return -1;
  if (branch && (cudaEventRecord(p.ev_join[0], p.side) != cudaSuccess || cudaStreamWaitEvent(st, p.ev_join[0], 0) != cudaSuccess)) {
    cuda_fail(cudaGetLastError(), "join of the wgrad stream");
    return -1;
  }
  if (launch_adam(p, step, 0, st)) return -1;
  return n + 2;
}

int b2_cnn_plan_run(void* plan, int which, int step, void* stream) {
  Plan* pp = static_cast<Plan*>(plan);
  Plan& p = *pp;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (which == RUN_REFRESH) return launch_adam(p, 0, 1, st);
  if (which == RUN_CONV0_FWD) return launch_conv0(p, false, step, st);
  if (which == RUN_FC1_FWD) return launch_gg<0, false, false, true, EPI_ATOMIC, 64>(p, p.p_fc1_fwd(), step, st);
  if (which == RUN_HEAD_TRAIN) return launch_head(p, step, 1, st);
  if (which == RUN_HEAD_EVAL) return launch_head(p, step, 0, st);
  if (which == RUN_FC1_WGRAD) return launch_gg<0, true, true, true, EPI_STORE, 64>(p, p.p_fc1_wgrad(), step, st);
  if (which == RUN_FC1_DGRAD) return launch_gg<0, false, true, true, EPI_UNPOOL, 64>(p, p.p_fc1_dgrad(), step, st);
  if (which == RUN_CONV0_WGRAD) return launch_conv0(p, true, step, st);
  if (which == RUN_ADAM) return launch_adam(p, step, 0, st);
  if (which > RUN_CONV_FWD && which < RUN_CONV_FWD + p.nl) return run_conv_fwd(p, which - RUN_CONV_FWD, step, st);
  if (which > RUN_CONV_WGRAD && which < RUN_CONV_WGRAD + p.nl) return run_conv_wgrad(p, which - RUN_CONV_WGRAD, step, st);
  if (which > RUN_CONV_DGRAD && which < RUN_CONV_DGRAD + p.nl) return run_conv_dgrad(p, which - RUN_CONV_DGRAD, step, st);
  set_error("cnn plan: unknown kernel id %d", which);
  return 1;
}

int b2_cnn_plan_query(void* plan, int slot, int buf, void** ptr, int64_t* bytes) {
  Plan* p = static_cast<Plan*>(plan);
  if (p == nullptr || slot < 0 || slot >= p->G) {
    set_error("cnn plan: bad slot %d", slot);
    return 1;
  }
  const SlotLayout& L = p->L;
  const int64_t B = p->d.batch;
  int64_t off = -1, sz = 0;
  auto pooled = [&](int l) { return B * (p->H[l] / 2) * (p->H[l] / 2) * p->Cout[l]; };
  if (buf >= BUF_ACT && buf < BUF_ACT + p->nl - 1) { off = L.act[buf - BUF_ACT]; sz = pooled(buf - BUF_ACT) * 4; }
  else if (buf >= BUF_IDX && buf < BUF_IDX + p->nl) { off = L.idx[buf - BUF_IDX]; sz = pooled(buf - BUF_IDX); }
  else if (buf > BUF_DZ && buf < BUF_DZ + p->nl) { const int l = buf - BUF_DZ; off = L.dz[l]; sz = B * p->H[l] * p->H[l] * p->Cout[l] * 4; }
  else if (buf == BUF_AHI) { off = L.a_hi; sz = B * p->feat * 4; }
  else if (buf == BUF_ALO) { off = L.a_lo; sz = B * p->feat * 4; }
  else if (buf == BUF_Z1) { off = L.z1; sz = B * p->d.hidden * 4; }
  else if (buf == BUF_DHHI) { off = L.dh_hi; sz = B * p->d.hidden * 4; }
  else if (buf == BUF_DHLO) { off = L.dh_lo; sz = B * p->d.hidden * 4; }
  else if (buf == BUF_DPM0) { off = L.dpm0; sz = pooled(0) * 4; }
  else if (buf == BUF_WM) { off = L.wm; sz = p->d.slab_elems * 4; }
  else if (buf == BUF_WLO) { off = L.wlo; sz = (int64_t)p->d.hidden * p->feat * 4; }
  else if (buf == BUF_GM) { off = L.gm; sz = p->d.slab_elems * 4; }
  if (off < 0) {
    set_error("cnn plan: unknown buffer id %d", buf);
    return 1;
  }
  *ptr = p->slot_ptr(slot, off);
  *bytes = sz;
  return 0;
}

}  // extern "C"
