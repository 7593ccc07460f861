// Error plumbing for the C ABI: every entry point returns 0 on success, a DFM_ERR_* code otherwise,
// and leaves a human-readable message retrievable through dfm_last_error().
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <cuda_runtime.h>

static thread_local char g_err[1024] = "";

extern "C" void dfm_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* dfm_last_error() { return g_err; }

extern "C" int dfm_abi_version() { return 1; }

// 1 when the current device is Blackwell-class (compute capability 10.x), 0 otherwise, <0 on error.
extern "C" int dfm_device_is_sm100() {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return -1;
  return major == 10 ? 1 : 0;
}

// Programmatic dependent launch switch (default on; DFM_PDL=0 in the environment or dfm_set_pdl(0) turns it off).
#include <stdlib.h>
static int g_pdl = -1;
extern "C" int dfm_pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("DFM_PDL");
    g_pdl = (e && e[0] == '0') ? 0 : 1;
  }
  return g_pdl;
}
extern "C" void dfm_set_pdl(int on) { g_pdl = on ? 1 : 0; }

// Page-lock an existing host range (the parameter-server store in /dev/shm) so H2D/D2H copies from it are truly
// asynchronous. Returns 0 on success; on failure the CUDA error state is cleared so later calls are not poisoned.
extern "C" int dfm_host_register(void* ptr, unsigned long long bytes, void** dev_ptr) {
  cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
  if (e != cudaSuccess) {
    cudaGetLastError();
    dfm_set_error("cudaHostRegister failed: %s", cudaGetErrorString(e));
    return 1;
  }
  if (dev_ptr) {  // device-visible alias of the range, for kernels that read the store directly (dfm_ps_gather)
    *dev_ptr = nullptr;
    e = cudaHostGetDevicePointer(dev_ptr, ptr, 0);
    if (e != cudaSuccess) {
      cudaGetLastError();
      *dev_ptr = nullptr;
    }
  }
  return 0;
}
extern "C" int dfm_host_unregister(void* ptr) {
  cudaError_t e = cudaHostUnregister(ptr);
  if (e != cudaSuccess) cudaGetLastError();
  return e == cudaSuccess ? 0 : 1;
}

// Persisting-L2 window over the training state (the 126 MB L2 makes keeping the optimizer state resident pay).
static cudaAccessPolicyWindow g_l2w;
static int g_l2w_on = 0;
extern "C" const cudaAccessPolicyWindow* dfm_l2_window() { return g_l2w_on ? &g_l2w : nullptr; }
// Returns the number of bytes the device agreed to set aside (0 = unsupported / disabled); bytes == 0 turns it off.
extern "C" long long dfm_set_l2_persist(void* base, unsigned long long bytes) {
  g_l2w_on = 0;
  if (!base || bytes == 0) return 0;
  int dev = 0, max_persist = 0, max_window = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
  cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev);
  if (max_persist <= 0 || max_window <= 0) { cudaGetLastError(); return 0; }
  unsigned long long set_aside = bytes < (unsigned long long)max_persist ? bytes : (unsigned long long)max_persist;
  if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, set_aside) != cudaSuccess) { cudaGetLastError(); return 0; }
  unsigned long long win = bytes < (unsigned long long)max_window ? bytes : (unsigned long long)max_window;
  g_l2w.base_ptr = base;
  g_l2w.num_bytes = win;
  g_l2w.hitRatio = (float)((double)set_aside / (double)win);
  if (g_l2w.hitRatio > 1.f) g_l2w.hitRatio = 1.f;
  g_l2w.hitProp = cudaAccessPropertyPersisting;
  g_l2w.missProp = cudaAccessPropertyStreaming;
  g_l2w_on = 1;
  return (long long)set_aside;
}

// ---- kernel timeline buffer (trace build only, see common.cuh). The entry points exist in both builds so the ABI is the
// same; in the normal build dfm_trace_enabled() returns 0 and the others do nothing.
#ifdef DFM_KTRACE
namespace dfm { enum { kTraceWords = 4 * 32 }; __device__ unsigned long long g_dfm_trace[kTraceWords]; }
#endif
extern "C" int dfm_trace_enabled() {
#ifdef DFM_KTRACE
  return 1;
#else
  return 0;
#endif
}
extern "C" int dfm_trace_words() { return 4 * 32; }
// Slots [0],[1] of every id are minima (reset to ~0), [2],[3] maxima (reset to 0).
extern "C" int dfm_trace_reset() {
#ifdef DFM_KTRACE
  unsigned long long h[dfm::kTraceWords];
  for (int i = 0; i < dfm::kTraceWords; ++i) h[i] = (i & 3) < 2 ? ~0ull : 0ull;
  if (cudaMemcpyToSymbol(dfm::g_dfm_trace, h, sizeof(h)) != cudaSuccess) { cudaGetLastError(); return 1; }
#endif
  return 0;
}
extern "C" int dfm_trace_read(unsigned long long* out) {
#ifdef DFM_KTRACE
  if (cudaMemcpyFromSymbol(out, dfm::g_dfm_trace, sizeof(unsigned long long) * dfm::kTraceWords) != cudaSuccess) {
    cudaGetLastError();
    return 1;
  }
  return 0;
#else
  (void)out;
  return 1;
#endif
}

// ---- flag-wait watchdog record (csrc/common.cuh wait_flag_geq): 8 words of pinned, device-mapped host memory. A kernel
// that gives up waiting for a peer writes (valid, missing rank, barrier id, expected seq, seen value) here and traps; the
// record survives the dead context because it is host memory.
static unsigned int* g_wd_host = nullptr;
extern "C" unsigned int* dfm_watchdog_record() {
  if (!g_wd_host) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, 8 * sizeof(unsigned int), cudaHostAllocPortable | cudaHostAllocMapped) != cudaSuccess) {
      cudaGetLastError();
      return nullptr;
    }
    memset(p, 0, 8 * sizeof(unsigned int));
    g_wd_host = (unsigned int*)p;
  }
  return g_wd_host;
}
// Writes the watchdog's message into buf and returns 1 if a flag wait has timed out in this process, else returns 0.
extern "C" int dfm_watchdog_message(char* buf, int cap) {
  const volatile unsigned int* r = g_wd_host;
  if (!r || r[0] == 0u || !buf || cap <= 0) return 0;
  snprintf(buf, (size_t)cap, "deepfm_b200: timed out waiting for rank %u (barrier %u, seq %u, saw %u)", r[1], r[2], r[3], r[4]);
  return 1;
}
