// Shared device helpers for the deepfm_b200 sm_100a kernels: error handling, PTX wrappers for
// mbarrier / TMA / tcgen05 / TMEM, vector loads/stores, reductions.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define DFM_OK 0
#define DFM_ERR_CUDA 1
#define DFM_ERR_ARG 2
#define DFM_ERR_UNSUPPORTED 3

extern "C" void dfm_set_error(const char* fmt, ...);

#define DFM_CHECK(expr)                                                                         \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      dfm_set_error("%s:%d CUDA error %s: %s", __FILE__, __LINE__, cudaGetErrorName(_e),        \
                    cudaGetErrorString(_e));                                                    \
      return DFM_ERR_CUDA;                                                                      \
    }                                                                                           \
  } while (0)

#define DFM_REQUIRE(cond, ...)                                                                  \
  do {                                                                                          \
    if (!(cond)) {                                                                              \
      dfm_set_error(__VA_ARGS__);                                                               \
      return DFM_ERR_ARG;                                                                       \
    }                                                                                           \
  } while (0)

extern "C" int dfm_pdl_enabled();
// Optional persisting-L2 window (parameters + gradients + optimizer slots, re-read every step): set once by
// dfm_set_l2_persist(); launch_k attaches it to every launch so it is captured into the CUDA graph's kernel nodes.
extern "C" const cudaAccessPolicyWindow* dfm_l2_window();

namespace dfm {

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// ---- kernel timeline (trace build only: `python -m deepfm_b200.build --trace` -> libdeepfm_b200_trace.so, -DDFM_KTRACE
// -rdc=true; scripts/step_trace.py). ncu serialises kernels, so it cannot show what overlaps what inside a step that runs
// on three graph branches (and across GPUs); %globaltimer stamps can. Per trace id: [0] first CTA entered the kernel,
// [1] first CTA got past griddepcontrol.wait (= every prerequisite grid has finished), [2] last CTA reached trace_end(),
// [3] free phase stamp (max). Ids: see scripts/step_trace.py. The normal build compiles all of this away.
enum { TR_EMB_FWD = 0, TR_FWD0 = 1, TR_LOSS = 4, TR_DGRAD0 = 5, TR_WGRAD0 = 8, TR_EMB_BWD = 11, TR_OPT = 12, TR_BITMAP = 13,
       TR_SWEEP = 14, TR_COLSUM = 15, TR_BAR_A = 16, TR_ALLREDUCE = 17, TR_BM_XCHG = 18, TR_AR_DENSE = 20, TR_AR_ROWS = 21,
       TR_AR_FENCE = 22, TR_OPT_DENSE = 24, TR_OPT_ROWS = 25, TR_OPT_FINAL = 26, TR_ROWLIST = 19, TR_MAX = 32 };
#ifdef DFM_KTRACE
extern __device__ unsigned long long g_dfm_trace[4 * TR_MAX];
__device__ __forceinline__ void trace_min(int id, int slot) {
  if (id >= 0 && threadIdx.x == 0) atomicMin(&g_dfm_trace[4 * id + slot], global_timer_ns());
}
__device__ __forceinline__ void trace_max(int id, int slot) {
  if (id >= 0 && threadIdx.x == 0) atomicMax(&g_dfm_trace[4 * id + slot], global_timer_ns());
}
#else
__device__ __forceinline__ void trace_min(int, int) {}
__device__ __forceinline__ void trace_max(int, int) {}
#endif
__device__ __forceinline__ void trace_end(int id) { trace_max(id, 2); }

// ---- programmatic dependent launch (PDL): a kernel launched with the attribute may begin while its predecessor in the
// stream is still running; pdl_wait() blocks until every prerequisite grid has completed and flushed, so everything
// before it (barrier init, TMEM allocation, descriptor prefetch, index math) overlaps the predecessor's tail.
// Without the launch attribute both instructions are no-ops. EVERY kernel launched through launch_k calls pdl_wait()
// before touching global data, which keeps the ordering transitive along the chain.
__device__ __forceinline__ void pdl_wait(int trace_id = -1) {
  trace_min(trace_id, 0);
  asm volatile("griddepcontrol.wait;" ::: "memory");
  trace_min(trace_id, 1);
}
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Kernels that run BESIDE the tcgen05 GEMMs on another graph branch (the untouched-row sweep, the bitmap kernels) ask for
// the LARGEST shared-memory carve-out although they need ~1 KB. An SM changes its L1/shared split only when it is empty:
// with the default preference such a kernel put its SMs into a small-carve-out configuration and every GEMM CTA (100+ KB of
// stages) waited until those SMs had drained — measured with scripts/step_trace.py: mlp_fwd0 started 26 us late behind a
// 148-CTA sweep. (Not applied to every kernel: gathers and epilogues lose L1 hit rate with a 28 KB L1, +10-30 % each.)
inline void dfm_prefer_max_shared(const void* fn) {
  constexpr int kCap = 256;
  static const void* seen[kCap];
  static int nseen = 0;
  for (int i = 0; i < nseen; ++i)
    if (seen[i] == fn) return;  // racy read of an append-only table: worst case the attribute is set twice
  cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
  cudaGetLastError();
  if (nseen < kCap) seen[nseen++] = fn;
}

// launch_k attaches the persisting-L2 window (if one is set) to the launch; streaming kernels that touch every byte once
// (the untouched-row sweep) switch it off for themselves: their lines must not push the step's working set out of L2.
inline bool& dfm_launch_no_window() {
  static thread_local bool off = false;
  return off;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  int na = dfm_pdl_enabled() ? 1 : 0;
  const cudaAccessPolicyWindow* w = dfm_launch_no_window() ? nullptr : dfm_l2_window();
  if (w) {
    attr[na].id = cudaLaunchAttributeAccessPolicyWindow;
    attr[na].val.accessPolicyWindow = *w;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// Tensor map of a bf16 row-major matrix [rows, cols] with leading dimension ld (elements), 128B swizzle, box = {64 inner,
// box_rows}, out-of-bounds elements read as zero (csrc/gemm_tcgen05.cu).
int make_tmap(CUtensorMap* tm, const void* base, int rows, int cols, int ld, int box_rows);

__host__ __device__ constexpr int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ constexpr long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

// ----------------------------------------------------------------------------------- warp utils
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
      : "=r"(pred));
  return pred;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ----------------------------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// ----------------------------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 2-D tiled load global -> shared, completion on an mbarrier (complete_tx::bytes).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

// ----------------------------------------------------------------------------------- tcgen05 / TMEM
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
               "n"(COLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 inputs, fp32 accumulate, one CTA.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// 32 lanes x 16 consecutive fp32 columns: thread i of the warp receives lane (base+i), columns [c, c+16).
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor for tcgen05.mma (sm_100 format: version field = 1).
//   layout_type: 0 none, 2 = 128B swizzle.   lbo/sbo in bytes.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}

// Instruction descriptor, kind::f16: BF16 x BF16 -> FP32, M x N tile, operand majors (0 = K-major, 1 = MN-major).
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4)                        // D format = F32
         | (1u << 7)                      // A format = BF16
         | (1u << 10)                     // B format = BF16
         | ((uint32_t)a_mn_major << 15)   // A major
         | ((uint32_t)b_mn_major << 16)   // B major
         | ((uint32_t)(N >> 3) << 17)     // N / 8
         | ((uint32_t)(M >> 4) << 24);    // M / 16
}

// ----------------------------------------------------------------------------------- cross-GPU flag waits
// Watchdog for every spin on a flag another GPU writes (reference analogue: Horovod's stall check, PDF p21): if the
// peer does not show up within DFM_FLAG_TIMEOUT_NS the kernel reports and traps, turning a silent hang (e.g. one rank
// died, or the ranks disagree on the number of steps, PDF p22) into a CUDA error the host can act on.
#ifndef DFM_FLAG_TIMEOUT_NS
#define DFM_FLAG_TIMEOUT_NS 20000000000ull
#endif
// What the watchdog saw goes into a record in pinned HOST memory (error.cu: dfm_watchdog_record), not through printf: a
// device printf gives the kernel a stack frame (its argument buffer), and no kernel of this library has one (see pick8
// below). The host can still read the record after the trap has killed the context (dfm_watchdog_message; printed by
// ops/lib.py when a native call or a synchronisation fails). The pointer is a per-translation-unit device variable (the
// production build has no relocatable device code), bound by wd_bind() from each file's host wrappers.
static __device__ unsigned int* g_wd_rec = nullptr;
extern "C" unsigned int* dfm_watchdog_record();
static inline void wd_bind(cudaStream_t st) {
  static bool done = false;  // static function: one flag per translation unit, like g_wd_rec
  if (done) return;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return;  // not inside a capture
  unsigned int* rec = dfm_watchdog_record();
  if (rec && cudaMemcpyToSymbol(g_wd_rec, &rec, sizeof(rec)) == cudaSuccess) done = true;
}
__device__ __forceinline__ void wait_flag_geq(const uint32_t* flag, uint32_t seq, int peer, int what) {
  unsigned long long t0 = 0;
  unsigned int spins = 0;
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if ((int32_t)(v - seq) >= 0) return;
    if ((++spins & 0x3FFu) == 0) {
      const unsigned long long now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > DFM_FLAG_TIMEOUT_NS) {
        unsigned int* rec = g_wd_rec;
        if (rec) {  // [0] valid, [1] missing rank, [2] barrier id, [3] expected sequence number, [4] last value seen
          rec[1] = (unsigned)peer; rec[2] = (unsigned)what; rec[3] = seq; rec[4] = v;
          __threadfence_system();
          rec[0] = 1u;
          __threadfence_system();
        }
        __trap();
      }
      __nanosleep(200);
    }
  }
}

// Element i of an 8-entry table that lives in the KERNEL PARAMETERS, without indexing it dynamically. A dynamic index
// makes the compiler copy the whole table into a per-thread stack frame (local memory), and a kernel with a stack frame
// costs far more than its loads: measured on the single-GPU step, 96 bytes of stack in emb_bwd added 8 us per step
// (cuobjdump -res-usage shows STACK:0 for every kernel of the library; keep it that way).
template <typename T>
__device__ __forceinline__ T pick8(T const (&arr)[8], int i) {
  T v = arr[0];
#pragma unroll
  for (int k = 1; k < 8; ++k) v = (i == k) ? arr[k] : v;
  return v;
}

// Cross-GPU barrier at the ENTRY of a kernel, whatever its grid: CTA 0 alone exchanges flags with the peers (threads
// 0..world-1: tell peer p "rank `rank` reached seq", wait until p said the same), then releases a LOCAL go word the other
// CTAs of this grid wait for. CTA 0 needs no other CTA to make progress, so nothing depends on how many CTAs are resident
// or on the order in which they are dispatched — a per-CTA flag pairing across ranks (the first version) did.
// flags[p] = rank p's flag array (peer-mapped); slot = this barrier's first word in it (8 words, one per rank);
// go = a word of THIS rank's array no peer writes. Sequence numbers only grow.
__device__ __forceinline__ void grid_entry_barrier(uint32_t* const (&flags)[8], size_t slot, uint32_t* go, int rank, int world,
                                                   uint32_t seq, int what) {
  if (world <= 1) return;
  const bool first = blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
  if (first) {
    if (threadIdx.x < (unsigned)world) {
      const int peer = threadIdx.x;
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pick8(flags, peer) + slot + rank), "r"(seq) : "memory");
      wait_flag_geq(pick8(flags, rank) + slot + peer, seq, peer, what);
    }
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(go), "r"(seq) : "memory");
  } else {
    if (threadIdx.x == 0) {
      uint32_t v;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(go) : "memory");
      } while ((int32_t)(v - seq) < 0);
    }
    __syncthreads();
  }
}

// ----------------------------------------------------------------------------------- vector memory ops
__device__ __forceinline__ void red_add_f32x4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

}  // namespace dfm
