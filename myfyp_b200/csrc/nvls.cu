// A2 (NVLS variant): FedAvg across the GPUs of a box through NVSwitch multicast - the reduction happens INSIDE the switch.
//
// Every rank binds one physical buffer [flags | in | out] to a shared multicast object and maps it twice: a unicast address of its
// own memory and the multicast address of everybody's.  One kernel per GPU and call:
//     barrier      multimem.red.add on the multicast flag: one instruction bumps the flag on every GPU; spin on the local copy
//     reduce       rank r owns the coordinate shard r:  v = multimem.ld_reduce.add.v4.f32 [mc_in + i]    (the switch returns the
//                  SUM over all GPUs of that address - the pre-weighted partial averages n_g/N * theta_g - in one load)
//     broadcast    multimem.st.v4.f32 [mc_out + i], v                                                   (one store lands on every GPU)
//     barrier      as above; after it every GPU's `out` holds the complete average
// NVLink traffic per GPU: 4N/G read-reduced + 4N/G multicast-stored, against 2(G-1)/G * 4N for the peer-load two-shot kernel
// (collective.cu) and (G-1) * 4N for the one-shot pull.  Waits are bounded (a rank that never arrives makes every kernel give up
// and report it; the caller falls back to the gossip path).
//
// Set-up needs the VMM driver API (cuMemCreate, cuMulticast*): one process creates the multicast object and hands its POSIX file
// descriptor to the others; all devices are added before anyone binds.  Availability is probed by b2_multicast_probe (runtime.cu).
// Reference path: none - the reference averages pickled NumPy arrays on the host (p2pfl/learning/aggregators/fedavg.py:61-69).
#include <cuda.h>

#include <mutex>

#include "../../include/myfyp_b200.h"
#include "common.cuh"

namespace b2 {
namespace nvls {

struct Api {
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  bool ok;
};

static const Api& api() {
  static Api a{};
  static std::once_flag once;
  std::call_once(once, [] {
    auto get = [](const char* name) -> void* {
      void* p = nullptr;
      cudaDriverEntryPointQueryResult q;
      if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
      return p;
    };
    a.MulticastCreate = (decltype(a.MulticastCreate))get("cuMulticastCreate");
    a.MulticastAddDevice = (decltype(a.MulticastAddDevice))get("cuMulticastAddDevice");
    a.MulticastBindMem = (decltype(a.MulticastBindMem))get("cuMulticastBindMem");
    a.MulticastUnbind = (decltype(a.MulticastUnbind))get("cuMulticastUnbind");
    a.MulticastGetGranularity = (decltype(a.MulticastGetGranularity))get("cuMulticastGetGranularity");
    a.MemCreate = (decltype(a.MemCreate))get("cuMemCreate");
    a.MemRelease = (decltype(a.MemRelease))get("cuMemRelease");
    a.MemAddressReserve = (decltype(a.MemAddressReserve))get("cuMemAddressReserve");
    a.MemAddressFree = (decltype(a.MemAddressFree))get("cuMemAddressFree");
    a.MemMap = (decltype(a.MemMap))get("cuMemMap");
    a.MemUnmap = (decltype(a.MemUnmap))get("cuMemUnmap");
    a.MemSetAccess = (decltype(a.MemSetAccess))get("cuMemSetAccess");
    a.MemExport = (decltype(a.MemExport))get("cuMemExportToShareableHandle");
    a.MemImport = (decltype(a.MemImport))get("cuMemImportFromShareableHandle");
    a.MemGetAllocationGranularity = (decltype(a.MemGetAllocationGranularity))get("cuMemGetAllocationGranularity");
    a.ok = a.MulticastCreate && a.MulticastAddDevice && a.MulticastBindMem && a.MulticastGetGranularity && a.MemCreate && a.MemRelease && a.MemAddressReserve && a.MemMap &&
           a.MemSetAccess && a.MemExport && a.MemImport && a.MemUnmap && a.MemAddressFree && a.MemGetAllocationGranularity;
  });
  return a;
}

static int cu_fail(CUresult r, const char* what) {
  set_error("%s failed with CUresult %d", what, (int)r);
  return 1;
}
#define B2_CU(expr)                                      \
  do {                                                   \
    CUresult _r = (expr);                                \
    if (_r != CUDA_SUCCESS) return cu_fail(_r, #expr);   \
  } while (0)

constexpr int FLAG_BYTES = 4096;
constexpr unsigned int TIMEOUT_NS_SHIFT = 33;  // ~8.6 s: a first call may find its peers still capturing graphs

struct Comm {
  int world, rank, dev;
  size_t bytes;   // bound bytes (granularity multiple)
  CUmemGenericAllocationHandle mc, mem;
  CUdeviceptr uc_va, mc_va;
  bool added, bound;
  unsigned int epoch;
};

// one store that lands in every GPU's copy of the flag word
__device__ __forceinline__ void mm_st_flag(unsigned int* mc_addr, unsigned int v) {
  asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(mc_addr), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 mm_ld_reduce(const float4* mc_addr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc_addr) : "memory");
  return v;
}
__device__ __forceinline__ void mm_st(float4* mc_addr, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// all CTAs: wait until every rank's slot flags[0..world) has reached `epoch` (slots only grow; an epoch that some rank skipped
// leaves nothing behind that a later epoch could trip over); false after ~8.6 s (some rank never arrived)
__device__ __forceinline__ bool wait_all(const unsigned int* flags, int world, unsigned int epoch, int* s_ok) {
  if (threadIdx.x == 0) *s_ok = 1;
  __syncthreads();
  if ((int)threadIdx.x < world) {
    const unsigned long long t0 = globaltimer();
    while ((int)(ld_acquire_sys(flags + threadIdx.x) - epoch) < 0) {
      __nanosleep(100);
      if (globaltimer() - t0 > (1ull << TIMEOUT_NS_SHIFT)) {
        *s_ok = 0;
        break;
      }
    }
  }
  __syncthreads();
  return *s_ok != 0;
}

// flags (unsigned; local copy at uc, multicast alias at mc): [0..31] "input of rank g is complete", [32..63] "shard of rank g is
// everywhere", [64] CTA counter, [72] epoch of the last call that gave up waiting
__global__ void __launch_bounds__(512) nvls_allreduce_kernel(unsigned char* uc, unsigned char* mc, int64_t off_in, int64_t off_out, int64_t n4, int world, int rank, unsigned int epoch,
                                                              float scale) {
  unsigned int* lflags = reinterpret_cast<unsigned int*>(uc);
  unsigned int* mflags = reinterpret_cast<unsigned int*>(mc);
  __shared__ int s_ok, s_last;
  // ---- barrier A: every rank's input is complete (this kernel is stream-ordered behind its producer)
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    __threadfence_system();
    mm_st_flag(mflags + rank, epoch);
  }
  if (!wait_all(lflags, world, epoch, &s_ok)) {
    if (threadIdx.x == 0) lflags[72] = epoch;
    return;
  }
  // ---- reduce my shard in the switch, broadcast it to everybody
  const int64_t shard = (n4 + world - 1) / world;
  const int64_t lo = (int64_t)rank * shard, hi = (lo + shard < n4) ? lo + shard : n4;
  const float4* in = reinterpret_cast<const float4*>(mc + off_in);
  float4* out = reinterpret_cast<float4*>(mc + off_out);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 7 * stride < hi; i += 8 * stride) {  // eight switch reductions in flight per thread (a multicast load takes several us)
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = mm_ld_reduce(in + i + u * stride);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (scale != 1.f) { v[u].x *= scale; v[u].y *= scale; v[u].z *= scale; v[u].w *= scale; }
      mm_st(out + i + u * stride, v[u]);
    }
  }
  for (; i < hi; i += stride) {
    float4 v = mm_ld_reduce(in + i);
    if (scale != 1.f) { v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale; }
    mm_st(out + i, v);
  }
  // ---- barrier B: the last CTA of this rank announces "my shard is everywhere"
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(lflags + 64, 1u) == gridDim.x - 1);
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    lflags[64] = 0u;
    __threadfence_system();
    mm_st_flag(mflags + 32 + rank, epoch);
  }
  if (!wait_all(lflags + 32, world, epoch, &s_ok)) {
    if (threadIdx.x == 0) lflags[72] = epoch;
  }
}

}  // namespace nvls
}  // namespace b2

using namespace b2;
using namespace b2::nvls;

extern "C" {

int64_t b2_nvls_granularity(int world) {
  const Api& a = api();
  if (!a.ok) {
    set_error("nvls: the driver does not export the multicast / VMM entry points");
    return -1;
  }
  CUmulticastObjectProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.numDevices = (unsigned)world;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  prop.size = 2u << 20;
  size_t g = 0;
  if (a.MulticastGetGranularity(&g, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) return -1;
  return (int64_t)g;
}

/* Root only: create the multicast object for `world` devices x `bytes` and export it as a POSIX file descriptor. */
int b2_nvls_root_create(int world, int64_t bytes, int* out_fd) {
  const Api& a = api();
  if (!a.ok) {
    set_error("nvls: the driver does not export the multicast / VMM entry points");
    return 1;
  }
  B2_CUDA(cudaFree(0));
  CUmulticastObjectProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.numDevices = (unsigned)world;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  prop.size = (size_t)bytes;
  CUmemGenericAllocationHandle mc = 0;
  B2_CU(a.MulticastCreate(&mc, &prop));
  int fd = -1;
  CUresult r = a.MemExport(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  a.MemRelease(mc);  // the descriptor keeps the object alive; every rank (root included) imports it
  if (r != CUDA_SUCCESS) return cu_fail(r, "cuMemExportToShareableHandle(multicast)");
  *out_fd = fd;
  return 0;
}

/* Every rank, step 1: import the object and add this rank's (current) device.  ALL ranks must finish this before any binds. */
void* b2_nvls_join(int fd, int world, int rank, int64_t bytes) {
  const Api& a = api();
  if (!a.ok) {
    set_error("nvls: the driver does not export the multicast / VMM entry points");
    return nullptr;
  }
  cudaFree(0);
  Comm* c = new Comm();
  memset(c, 0, sizeof(*c));
  c->world = world; c->rank = rank; c->bytes = (size_t)bytes;
  cudaGetDevice(&c->dev);
  CUresult r = a.MemImport(&c->mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    cu_fail(r, "cuMemImportFromShareableHandle(multicast)");
    delete c;
    return nullptr;
  }
  r = a.MulticastAddDevice(c->mc, c->dev);
  if (r != CUDA_SUCCESS) {
    cu_fail(r, "cuMulticastAddDevice");
    a.MemRelease(c->mc);
    delete c;
    return nullptr;
  }
  c->added = true;
  return c;
}

/* Every rank, step 2 (after a barrier): allocate this rank's physical buffer, bind it, map the unicast and the multicast view. */
int b2_nvls_bind(void* comm) {
  const Api& a = api();
  Comm* c = static_cast<Comm*>(comm);
  B2_CUDA(cudaSetDevice(c->dev));
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = c->dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  B2_CU(a.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (gran == 0 || c->bytes % gran) {
    set_error("nvls: %zu bytes is not a multiple of the allocation granularity %zu", c->bytes, gran);
    return 1;
  }
  B2_CU(a.MemCreate(&c->mem, c->bytes, &prop, 0));
  B2_CU(a.MulticastBindMem(c->mc, 0, c->mem, 0, c->bytes, 0));
  c->bound = true;
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = c->dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  B2_CU(a.MemAddressReserve(&c->uc_va, c->bytes, gran, 0, 0));
  B2_CU(a.MemMap(c->uc_va, c->bytes, 0, c->mem, 0));
  B2_CU(a.MemSetAccess(c->uc_va, c->bytes, &acc, 1));
  B2_CU(a.MemAddressReserve(&c->mc_va, c->bytes, gran, 0, 0));
  B2_CU(a.MemMap(c->mc_va, c->bytes, 0, c->mc, 0));
  B2_CU(a.MemSetAccess(c->mc_va, c->bytes, &acc, 1));
  B2_CUDA(cudaMemset(reinterpret_cast<void*>(c->uc_va), 0, c->bytes));
  B2_CUDA(cudaDeviceSynchronize());
  return 0;
}

void* b2_nvls_local_ptr(void* comm) { return reinterpret_cast<void*>(static_cast<Comm*>(comm)->uc_va); }
void* b2_nvls_mc_ptr(void* comm) { return reinterpret_cast<void*>(static_cast<Comm*>(comm)->mc_va); }
int64_t b2_nvls_flag_bytes(void) { return FLAG_BYTES; }

/* out[all ranks] = scale * sum_ranks in[rank];  offsets in bytes from the start of the region (>= b2_nvls_flag_bytes(), 16-byte
 * aligned), n elements (multiple of 4).  `epoch` = 1, 2, 3... identical on every rank.  One launch, no host synchronisation. */
int b2_nvls_allreduce(void* comm, int64_t off_in, int64_t off_out, int64_t n, unsigned int epoch, float scale, int max_ctas, void* stream) {
  Comm* c = static_cast<Comm*>(comm);
  if (c == nullptr || !c->bound) {
    set_error("nvls: communicator is not bound");
    return 1;
  }
  if (c->world > 32) {
    set_error("nvls: at most 32 ranks");
    return 1;
  }
  if (n <= 0 || (n & 3) || (off_in & 15) || (off_out & 15) || off_in < FLAG_BYTES || off_out < FLAG_BYTES || (size_t)(off_in + 4 * n) > c->bytes || (size_t)(off_out + 4 * n) > c->bytes) {
    set_error("nvls: bad buffer geometry (n=%lld off_in=%lld off_out=%lld region=%zu)", (long long)n, (long long)off_in, (long long)off_out, c->bytes);
    return 1;
  }
  // persistent: every CTA must be resident (they wait on other GPUs).  One CTA per SM by default; a caller that owns the whole GPU
  // for the duration of the call may ask for up to two (50 registers x 512 threads: two fit), which helps above ~64 MB
  int grid = sm_count();
  if (max_ctas > 0) grid = max_ctas < 2 * grid ? max_ctas : 2 * grid;
  const int64_t need = (n / 4 / c->world + 511) / 512;
  if (need < grid) grid = (int)(need < 1 ? 1 : need);
  nvls_allreduce_kernel<<<grid, 512, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<unsigned char*>(c->uc_va), reinterpret_cast<unsigned char*>(c->mc_va), off_in, off_out, n / 4,
                                                                             c->world, c->rank, epoch, scale);
  B2_LAUNCH_CHECK("nvls_allreduce_kernel");
  return 0;
}

/* After the stream has been synchronised: the epoch of the last call whose wait timed out on this rank (0 = none). */
int b2_nvls_last_timeout(void* comm, unsigned int* epoch_out) {
  Comm* c = static_cast<Comm*>(comm);
  B2_CUDA(cudaMemcpy(epoch_out, reinterpret_cast<unsigned int*>(c->uc_va) + 72, sizeof(unsigned int), cudaMemcpyDeviceToHost));
  return 0;
}

int b2_nvls_destroy(void* comm) {
  const Api& a = api();
  Comm* c = static_cast<Comm*>(comm);
  if (c == nullptr) return 0;
  cudaSetDevice(c->dev);
  cudaDeviceSynchronize();
  if (c->mc_va) {
    a.MemUnmap(c->mc_va, c->bytes);
    a.MemAddressFree(c->mc_va, c->bytes);
  }
  if (c->uc_va) {
    a.MemUnmap(c->uc_va, c->bytes);
    a.MemAddressFree(c->uc_va, c->bytes);
  }
  if (c->bound && a.MulticastUnbind) a.MulticastUnbind(c->mc, c->dev, 0, c->bytes);
  if (c->mem) a.MemRelease(c->mem);
  if (c->mc) a.MemRelease(c->mc);
  delete c;
  return 0;
}

}  // extern "C"
