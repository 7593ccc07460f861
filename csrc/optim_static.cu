// dfm_flat_opt_step_static — the flat optimizer step for callers that cannot allocate device memory.
//
// Who needs this: code that runs after the GPU's memory has been handed to someone else — a serving or training
// host whose arena allocator already owns every free byte, a plugin loaded into such a process, a job held under
// a hard memory cap — and code on a latency path where cudaMalloc/cudaFree are unacceptable because both
// serialise the whole device. dfm_flat_opt_step() asks such a caller for a zeroed `done_counter` word of device
// scratch; this variant asks for nothing beyond the tensors it reads and writes.
//
// The scratch lives in the library image instead: one file-scope `static __device__` array of
// DFM_STATIC_OPT_SCRATCH_BYTES (256 B of counters + 64 Ki fp32 partial sums = 262,400 bytes). Every GPU that
// loads libdeepfm_b200.so carries that array until the process exits, whether or not this entry point is used.
//
// Differences from dfm_flat_opt_step (same update rules, same ranges, same side effects on w/g/s0/s1/shadow/step):
//   * one CTA per 1024 elements and no grid cap, so the size is bounded by the scratch: n <= 64 Mi elements
//     (DFM_STATIC_OPT_MAX_ELEMS); larger n is rejected with DFM_ERR_ARG and a message, nothing is launched;
//   * the L2 loss value is reduced through per-CTA partials in the scratch and summed in a fixed order by the last
//     CTA, then added to *reg_loss (the caller-allocated variant uses floating-point atomics).
//
// Rules for callers
//   * The first call on each device performs the one-time set-up for that device (it resolves the device address
//     of the array with cudaGetSymbolAddress while that device is current). Make that first call before starting
//     any stream capture.
//   * All calls on a device share the one array and it keeps whatever the previous call left. The entry point
//     itself re-zeroes the completion counter (a 4-byte stream-ordered memset, not an allocation) before each
//     launch; the partial sums are fully overwritten. Two calls that overlap on the GPU (different streams, or
//     different host threads feeding different streams) race on the array: serialise them.
//   * No cudaMalloc, cudaFree or allocating library routine is called, here or in anything this calls.
#include "common.cuh"
#include "optim_math.cuh"

#include <mutex>

#define DFM_STATIC_OPT_CHUNK 1024                 /* elements per CTA (256 threads x float4) */
#define DFM_STATIC_OPT_MAX_CHUNKS 65536           /* partial-sum slots */
#define DFM_STATIC_OPT_MAX_ELEMS ((long long)DFM_STATIC_OPT_CHUNK * DFM_STATIC_OPT_MAX_CHUNKS)
#define DFM_STATIC_OPT_SCRATCH_BYTES (256 + 4 * DFM_STATIC_OPT_MAX_CHUNKS)

namespace {

// [0,4): CTA completion counter; [256, 256 + 4*chunks): per-CTA partial sums of |w|^2 over the decay range.
// In HOST code this name is only a handle for cudaGetSymbolAddress — it is not a device address.
static __device__ __align__(256) unsigned char g_static_opt_scratch[DFM_STATIC_OPT_SCRATCH_BYTES];

struct StaticOptArgs {
  float* w;
  float* g;
  float* s0;
  float* s1;
  long long n;
  dfm::OptHyper h;
  float grad_scale;
  long long decay_begin, decay_end;
  float l2;
  long long shadow_begin, shadow_end;
  __nv_bfloat16* shadow;
  unsigned long long* step_ptr;
  float* reg_loss;
  unsigned char* scratch;  // device address of g_static_opt_scratch, resolved on the host at set-up
};

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads) flat_opt_static_kernel(const StaticOptArgs a) {
  using namespace dfm;
  __shared__ float s_lr;
  __shared__ float s_red[kThreads];
  __shared__ bool s_last;
  unsigned int* done = reinterpret_cast<unsigned int*>(a.scratch);
  float* partial = reinterpret_cast<float*>(a.scratch + 256);

  const unsigned long long t = *a.step_ptr + 1ull;
  if (threadIdx.x == 0) s_lr = opt_step_lr(a.h, t);
  __syncthreads();
  const float lr_t = s_lr;
  const int nslots = opt_num_slots(a.h.opt);
  float reg = 0.f;
  const long long i = (long long)blockIdx.x * kThreads + threadIdx.x;
  if (i < (a.n >> 2)) {
    const long long e = i << 2;
    float4 w = reinterpret_cast<float4*>(a.w)[i];
    float4 g = reinterpret_cast<float4*>(a.g)[i];
    float4 s0 = nslots > 0 ? reinterpret_cast<float4*>(a.s0)[i] : make_float4(0, 0, 0, 0);
    float4 s1 = nslots > 1 ? reinterpret_cast<float4*>(a.s1)[i] : make_float4(0, 0, 0, 0);
    g.x *= a.grad_scale; g.y *= a.grad_scale; g.z *= a.grad_scale; g.w *= a.grad_scale;
    if (e >= a.decay_begin && e < a.decay_end) {
      reg = w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w;
      g.x += a.l2 * w.x; g.y += a.l2 * w.y; g.z += a.l2 * w.z; g.w += a.l2 * w.w;
    }
    opt_update(a.h, lr_t, g.x, w.x, s0.x, s1.x);
    opt_update(a.h, lr_t, g.y, w.y, s0.y, s1.y);
    opt_update(a.h, lr_t, g.z, w.z, s0.z, s1.z);
    opt_update(a.h, lr_t, g.w, w.w, s0.w, s1.w);
    reinterpret_cast<float4*>(a.w)[i] = w;
    reinterpret_cast<float4*>(a.g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (nslots > 0) reinterpret_cast<float4*>(a.s0)[i] = s0;
    if (nslots > 1) reinterpret_cast<float4*>(a.s1)[i] = s1;
    if (e >= a.shadow_begin && e < a.shadow_end) {
      uint2 pk;
      pk.x = pack_bf16x2(w.x, w.y);
      pk.y = pack_bf16x2(w.z, w.w);
      *reinterpret_cast<uint2*>(a.shadow + (e - a.shadow_begin)) = pk;
    }
  }
  // fixed-order block reduction -> this CTA's partial
  s_red[threadIdx.x] = reg;
  __syncthreads();
  for (int o = kThreads / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) s_red[threadIdx.x] += s_red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    partial[blockIdx.x] = s_red[0];
    __threadfence();
    s_last = (atomicAdd(done, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  // last CTA: every partial is visible; sum them in a fixed order
  __threadfence();
  float acc = 0.f;
  for (unsigned int b = threadIdx.x; b < gridDim.x; b += kThreads) acc += __ldcg(partial + b);
  s_red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = kThreads / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) s_red[threadIdx.x] += s_red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    if (a.reg_loss) *a.reg_loss += 0.5f * a.l2 * s_red[0];
    *a.step_ptr = t;
    *done = 0u;
  }
}

// One-time set-up per device: resolve the device address of the static array (host code cannot use the symbol
// name as a pointer). Returns nullptr and sets the error on failure.
unsigned char* static_scratch_for_current_device() {
  static std::mutex mu;
  static unsigned char* ptrs[64] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    dfm_set_error("flat_opt_static: cudaGetDevice failed or device index out of range (%d): %s", dev,
                  cudaGetErrorString(e));
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_opt_scratch);
    if (e != cudaSuccess || p == nullptr) {
      dfm_set_error("flat_opt_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return nullptr;
    }
    ptrs[dev] = static_cast<unsigned char*>(p);
  }
  return ptrs[dev];
}

}  // namespace

extern "C" {

long long dfm_flat_opt_static_max_elems(void) { return DFM_STATIC_OPT_MAX_ELEMS; }
long long dfm_flat_opt_static_scratch_bytes(void) { return DFM_STATIC_OPT_SCRATCH_BYTES; }

int dfm_flat_opt_step_static(float* w, float* g, float* s0, float* s1, long long n, int opt, float lr, float beta1,
                             float beta2, float eps, float momentum, float grad_scale, long long decay_begin,
                             long long decay_end, float l2, long long shadow_begin, long long shadow_end, void* shadow,
                             unsigned long long* step_ptr, float* reg_loss, void* stream) {
  DFM_REQUIRE(n > 0 && n % 4 == 0, "flat_opt_static: n must be a positive multiple of 4");
  DFM_REQUIRE(n <= DFM_STATIC_OPT_MAX_ELEMS,
              "flat_opt_static: n = %lld exceeds the built-in scratch (max %lld elements = %d CTAs x %d); "
              "use dfm_flat_opt_step with caller-provided scratch for larger buffers",
              n, (long long)DFM_STATIC_OPT_MAX_ELEMS, DFM_STATIC_OPT_MAX_CHUNKS, DFM_STATIC_OPT_CHUNK);
  DFM_REQUIRE(decay_begin % 4 == 0 && decay_end % 4 == 0 && shadow_begin % 4 == 0 && shadow_end % 4 == 0,
              "flat_opt_static: ranges must be multiples of 4");
  DFM_REQUIRE(opt >= 0 && opt <= 4, "flat_opt_static: unknown optimizer id %d", opt);
  unsigned char* scratch = static_scratch_for_current_device();
  if (scratch == nullptr) return DFM_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  // the completion counter must start at zero even if an earlier launch died half-way
  DFM_CHECK(cudaMemsetAsync(scratch, 0, sizeof(unsigned int), st));
  StaticOptArgs a;
  a.w = w; a.g = g; a.s0 = s0; a.s1 = s1; a.n = n;
  a.h.opt = opt; a.h.lr = lr; a.h.beta1 = beta1; a.h.beta2 = beta2; a.h.eps = eps; a.h.momentum = momentum;
  a.grad_scale = grad_scale; a.decay_begin = decay_begin; a.decay_end = decay_end; a.l2 = l2;
  a.shadow_begin = shadow_begin; a.shadow_end = shadow_end; a.shadow = (__nv_bfloat16*)shadow;
  a.step_ptr = step_ptr; a.reg_loss = reg_loss; a.scratch = scratch;
  const unsigned int blocks = (unsigned int)dfm::ceil_div_ll(n / 4, kThreads);
  flat_opt_static_kernel<<<blocks, kThreads, 0, st>>>(a);
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

}  // extern "C"
