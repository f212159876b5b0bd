// Library runtime: error string, device queries, peer access, IPC arenas (the native side of the fabric).
#include <cstdarg>
#include <cstdlib>
#include <mutex>

#include <cuda.h>

#include "../../include/myfyp_b200.h"
#include "common.cuh"

namespace b2 {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
  return (int)e ? (int)e : -1;
}

int sm_count() {
  static int cache[64];
  static bool have[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return B2_SMS_FALLBACK;
  if (!have[dev]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = B2_SMS_FALLBACK;
    cache[dev] = n;
    have[dev] = true;
  }
  return cache[dev];
}
static int g_pdl = -1;
bool pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("B2_PDL");
    g_pdl = (e != nullptr && e[0] == '1') ? 1 : 0;
  }
  return g_pdl == 1;
}
}  // namespace b2

extern "C" {

/* 1 / 0: launch the training-step kernels with / without programmatic dependent launch (default: off; B2_PDL=1) */
void b2_set_pdl(int on) { b2::g_pdl = on ? 1 : 0; }

const char* b2_last_error(void) { return b2::g_err; }

int b2_version(void) { return 100; }

int b2_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int b2_enable_peer_access_all(void) {
  int n = b2_device_count(), ok = 0, cur = 0;
  if (n < 2) return 0;
  cudaGetDevice(&cur);
  for (int i = 0; i < n; ++i) {
    cudaSetDevice(i);
    for (int j = 0; j < n; ++j) {
      if (i == j) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, i, j);
      if (!can) continue;
      cudaError_t e = cudaDeviceEnablePeerAccess(j, 0);
      if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) ++ok;
      cudaGetLastError();
    }
  }
  cudaSetDevice(cur);
  return ok;
}

// Peer access among the listed GPUs only (both directions).  Unlike b2_enable_peer_access_all this creates no context on a GPU
// the process does not use.
int b2_enable_peer_access(const int* devs, int n) {
  int ok = 0, cur = 0;
  if (n < 2 || devs == nullptr) return 0;
  cudaGetDevice(&cur);
  for (int a = 0; a < n; ++a) {
    cudaSetDevice(devs[a]);
    for (int b = 0; b < n; ++b) {
      if (devs[a] == devs[b]) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, devs[a], devs[b]);
      if (!can) continue;
      cudaError_t e = cudaDeviceEnablePeerAccess(devs[b], 0);
      if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) ++ok;
      cudaGetLastError();
    }
  }
  cudaSetDevice(cur);
  return ok;
}

void* b2_arena_alloc(int64_t bytes) {
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, (size_t)bytes);
  if (e != cudaSuccess) {
    b2::cuda_fail(e, "cudaMalloc(arena)");
    return nullptr;
  }
  cudaMemset(p, 0, (size_t)bytes);
  return p;
}

int b2_arena_free(void* p) {
  B2_CUDA(cudaFree(p));
  return 0;
}

int b2_ipc_export(void* base, char* out_handle64) {
  cudaIpcMemHandle_t h;
  B2_CUDA(cudaIpcGetMemHandle(&h, base));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(out_handle64, &h, sizeof(h));
  return 0;
}

void* b2_ipc_open(const char* handle64) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    b2::cuda_fail(e, "cudaIpcOpenMemHandle");
    return nullptr;
  }
  return p;
}

int b2_ipc_close(void* p) {
  B2_CUDA(cudaIpcCloseMemHandle(p));
  return 0;
}

/* Interprocess events: the publisher records one after the copy into an arena slot, a consumer in another process makes its
 * stream wait for it before it dereferences the slot - cross-process ordering without a host synchronisation on either side. */
void* b2_ipc_event_create(void) {
  cudaEvent_t ev = nullptr;
  cudaError_t e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming | cudaEventInterprocess);
  if (e != cudaSuccess) {
    b2::cuda_fail(e, "cudaEventCreateWithFlags(interprocess)");
    return nullptr;
  }
  return ev;
}

int b2_ipc_event_export(void* ev, char* out_handle64) {
  cudaIpcEventHandle_t h;
  B2_CUDA(cudaIpcGetEventHandle(&h, static_cast<cudaEvent_t>(ev)));
  static_assert(sizeof(h) == 64, "cudaIpcEventHandle_t is 64 bytes");
  memcpy(out_handle64, &h, sizeof(h));
  return 0;
}

void* b2_ipc_event_open(const char* handle64) {
  cudaIpcEventHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  cudaEvent_t ev = nullptr;
  cudaError_t e = cudaIpcOpenEventHandle(&ev, h);
  if (e != cudaSuccess) {
    b2::cuda_fail(e, "cudaIpcOpenEventHandle");
    return nullptr;
  }
  return ev;
}

int b2_event_record(void* ev, void* stream) {
  B2_CUDA(cudaEventRecord(static_cast<cudaEvent_t>(ev), static_cast<cudaStream_t>(stream)));
  return 0;
}

int b2_stream_wait_event(void* stream, void* ev) {
  B2_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), static_cast<cudaEvent_t>(ev), 0));
  return 0;
}

int b2_event_destroy(void* ev) {
  B2_CUDA(cudaEventDestroy(static_cast<cudaEvent_t>(ev)));
  return 0;
}

/* 1 when the CURRENT device reports CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED (no context is created on any other GPU). */
int b2_multicast_supported(void) {
  int dev = 0, v = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuDeviceGetAttribute", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return 0;
  typedef CUresult (*GetAttrFn)(int*, CUdevice_attribute, CUdevice);
  if (((GetAttrFn)p)(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return 0;
  return v ? 1 : 0;
}

/* NVLS probe: does this box expose NVSwitch multicast objects (the basis of multimem.ld_reduce / multimem.st) to this process?
 * Reports, per device, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED and, when at least two devices say yes, the outcome of actually
 * creating a multicast object over all of them (cuMulticastGetGranularity + cuMulticastCreate + cuMulticastAddDevice); the object
 * is released again.  Text goes to `out`; returns 1 when a multicast object could be created, 0 when not, < 0 on a driver error. */
int b2_multicast_probe(char* out, int cap) {
  auto put = [&](const char* fmt, auto... a) {
    const int used = (int)strlen(out);
    if (used < cap - 1) snprintf(out + used, cap - used, fmt, a...);
  };
  if (cap > 0) out[0] = 0;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    put("no CUDA device\n");
    return -1;
  }
  auto entry = [&](const char* name) -> void* {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
    return p;
  };
  typedef CUresult (*GetAttrFn)(int*, CUdevice_attribute, CUdevice);
  typedef CUresult (*McGranFn)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  typedef CUresult (*McCreateFn)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  typedef CUresult (*McAddFn)(CUmemGenericAllocationHandle, CUdevice);
  typedef CUresult (*ReleaseFn)(CUmemGenericAllocationHandle);
  GetAttrFn get_attr = (GetAttrFn)entry("cuDeviceGetAttribute");
  McGranFn mc_gran = (McGranFn)entry("cuMulticastGetGranularity");
  McCreateFn mc_create = (McCreateFn)entry("cuMulticastCreate");
  McAddFn mc_add = (McAddFn)entry("cuMulticastAddDevice");
  ReleaseFn release = (ReleaseFn)entry("cuMemRelease");
  if (get_attr == nullptr) {
    put("cuDeviceGetAttribute not resolvable\n");
    return -1;
  }
  int yes = 0, cur = 0;
  cudaGetDevice(&cur);
  for (int d = 0; d < n; ++d) {
    int v = 0, fab = 0;
    const CUresult r = get_attr(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d);
    get_attr(&fab, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, d);
    put("device %d: MULTICAST_SUPPORTED=%d (CUresult %d), HANDLE_TYPE_FABRIC_SUPPORTED=%d\n", d, v, (int)r, fab);
    yes += v ? 1 : 0;
  }
  cudaSetDevice(cur);
  if (yes < 2 || mc_gran == nullptr || mc_create == nullptr || mc_add == nullptr || release == nullptr) {
    put("multicast object: not attempted (%d device(s) report support; entry points %s)\n", yes, mc_create ? "present" : "missing");
    return 0;
  }
  CUmulticastObjectProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.numDevices = (unsigned)yes;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  CUresult r = mc_gran(&gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  put("cuMulticastGetGranularity: CUresult %d, recommended granularity %zu bytes\n", (int)r, gran);
  if (r != CUDA_SUCCESS) return 0;
  prop.size = gran;
  CUmemGenericAllocationHandle h = 0;
  r = mc_create(&h, &prop);
  put("cuMulticastCreate(%u devices, %zu bytes): CUresult %d\n", prop.numDevices, prop.size, (int)r);
  if (r != CUDA_SUCCESS) return 0;
  int added = 0;
  for (int d = 0; d < n && added < yes; ++d) {
    r = mc_add(h, d);
    put("cuMulticastAddDevice(%d): CUresult %d\n", d, (int)r);
    added += r == CUDA_SUCCESS ? 1 : 0;
  }
  release(h);
  return added == yes ? 1 : 0;
}

}  // extern "C"
