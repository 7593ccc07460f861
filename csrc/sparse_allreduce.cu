// Data-parallel gradient exchange for REPLICATED embedding tables that ships only what differs between ranks
// (replaces Horovod's dense NCCL all-reduce of the 16-19 MB fusion buffer, HVD:262).
//
// The reference's embedding gradient is dense only because of the whole-table L2 term — and that term is
// IDENTICAL on every replica. The data gradient is non-zero only in the rows some rank looked up this step. So:
//
//   bitmap_exchange   (side stream, right after touched_bitmap)  every rank stores its touched-row bitmap (N/8 bytes) into
//                     slot [rank] of every peer's inbox with P2P stores (or one multimem.st per word) + flag barrier C.
//                     union_rowlist (csrc/split_opt.cu) then ORs the W bitmaps and lists the touched rows (same list on
//                     every rank); sweep_untouched updates every row NO rank touched, locally and identically on every
//                     replica (optimizer slots are replicated), off the critical path.
//   barrier A         (1-CTA kernel after emb_bwd) also deposits this rank's log-loss sum in every peer (global-mean loss)
//   sparse_allreduce  in-place all-reduce of [fm_bias block | MLP block | rows in the union bitmap]: item i is reduced by
//                     exactly one rank — multimem.ld_reduce.add in the switch (NVLS) or peer loads summed in rank order
//                     (P2P) — and the sum is stored into EVERY rank's gradient buffer (multimem.st / peer stores), so all
//                     replicas read bit-identical gradients. Last CTA: one fence.sys + flag barrier B.
//   tail_opt          (csrc/split_opt.cu) every rank applies the same update to the same values.
//
// Only single CTAs ever spin on a remote flag (the 1-CTA barrier kernel and the last CTA of a sweep), every spin has the
// common.cuh watchdog, sequence numbers only grow.
#include "common.cuh"

namespace dfm {

constexpr int kSaMaxRanks = 8;
constexpr int kSaThreads = 256;
constexpr int kSaFlagA = 0, kSaFlagB = kSaMaxRanks, kSaFlagC = 2 * kSaMaxRanks;

struct SaArgs {
  float* g[kSaMaxRanks];          // every rank's flat gradient buffer (peer-mapped)
  float* g_mc;                    // multicast mapping of the same buffer (NVLS) or null
  uint32_t* flags[kSaMaxRanks];   // every rank's flag words: [A x8 | B x8 | C x8]
  float* loss_parts[kSaMaxRanks]; // every rank's [kSaMaxRanks] inbox of per-rank log-loss sums
  uint32_t* bm_in[kSaMaxRanks];   // every rank's bitmap inbox: [kSaMaxRanks][nwords]
  uint32_t* bm_in_mc;             // multicast mapping of the inbox or null
  int rank, world;
  long long fmw_off, fmv_off;
  int N, nwords;
  long long d0_begin, d0_end, d1_begin, d1_end;
  const int* rows;                // ascending touched-row ids (union over ranks; identical on every rank)
  const int* row_count;
  uint32_t* bm_own;               // local live bitmap (bitmap_exchange reads and clears it)
  const float* loss_sum;          // this rank's log-loss sum
  unsigned int* done_counter;
  unsigned int* seq_ab;           // sequence number of barriers A/B (advanced by the all-reduce's last CTA)
  unsigned int* seq_c;            // sequence number of barrier C (advanced by bitmap_exchange's last CTA)
};

__device__ __forceinline__ void sa_st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 sa_ld_volatile_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float sa_ld_volatile_f1(const float* p) {
  float v;
  asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 sa_mm_ld_reduce_f4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float sa_mm_ld_reduce_f1(const float* p) {
  float v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void sa_mm_st_f4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void sa_mm_st_f1(float* p, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void sa_mm_st_u32(uint32_t* p, uint32_t v) {
  asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// flag exchange by threads 0..world-1 of the calling CTA: tell every peer we reached `seq`, wait until every peer did
__device__ __forceinline__ void sa_flag_barrier(const SaArgs& a, int base, uint32_t seq, int what) {
  if (threadIdx.x < a.world) {
    const int peer = threadIdx.x;
    sa_st_release_sys(pick8(a.flags, peer) + base + a.rank, seq);  // pick8: no dynamically indexed parameter table
    wait_flag_geq(pick8(a.flags, a.rank) + base + peer, seq, peer, what);
  }
}

// ---------------------------------------------------------------------------------------------- bitmap exchange
template <bool NVLS>
__global__ void __launch_bounds__(kSaThreads) bitmap_exchange_kernel(const SaArgs a) {
  __shared__ bool s_last;
  pdl_launch_dependents();
  pdl_wait(TR_BM_XCHG);
  const uint32_t seq = *a.seq_c + 1u;
  for (int wi = blockIdx.x * kSaThreads + threadIdx.x; wi < a.nwords; wi += gridDim.x * kSaThreads) {
    const uint32_t v = a.bm_own[wi];
    a.bm_own[wi] = 0u;
    const long long slot = (long long)a.rank * a.nwords + wi;
    if constexpr (NVLS) {
      sa_mm_st_u32(a.bm_in_mc + slot, v);
    } else {
#pragma unroll
      for (int r = 0; r < kSaMaxRanks; ++r)
        if (r < a.world) a.bm_in[r][slot] = v;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = (atomicAdd(a.done_counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence_system();  // every CTA's peer stores happen-before this fence through the done-counter chain
  __syncthreads();
  if (threadIdx.x == 0) {
    *a.done_counter = 0u;
    *a.seq_c = seq;
  }
  sa_flag_barrier(a, kSaFlagC, seq, 2);
  trace_end(TR_BM_XCHG);
}

// ---------------------------------------------------------------------------------------------- barrier A (+ loss deposit)
__global__ void sa_barrier_a_kernel(const SaArgs a) {
  pdl_launch_dependents();
  pdl_wait(TR_BAR_A);
  const uint32_t seq = *a.seq_ab + 1u;
  if (threadIdx.x < a.world) pick8(a.loss_parts, (int)threadIdx.x)[a.rank] = *a.loss_sum;  // ordered before the release below
  sa_flag_barrier(a, kSaFlagA, seq, 0);
  trace_end(TR_BAR_A);
}

// ---------------------------------------------------------------------------------------------- sparse all-reduce
template <bool NVLS, int G>
__global__ void __launch_bounds__(kSaThreads) sparse_allreduce_kernel(const SaArgs a) {
  __shared__ bool s_last;
  const int W = a.world;
  pdl_launch_dependents();
  pdl_wait(TR_ALLREDUCE);  // barrier A ran in the preceding 1-CTA kernel: every rank's gradients are complete and visible
  const uint32_t seq = *a.seq_ab + 1u;
  const int lane = threadIdx.x & 31;

  auto reduce4 = [&](long long e) -> float4 {  // e = element offset (multiple of 4)
    if constexpr (NVLS) {
      return sa_mm_ld_reduce_f4(a.g_mc + e);
    } else {
      float4 x[kSaMaxRanks];
#pragma unroll
      for (int r = 0; r < kSaMaxRanks; ++r)
        x[r] = r < W ? sa_ld_volatile_f4(a.g[r] + e) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 acc = x[0];  // fixed rank order: the sum does not depend on which rank computes it
#pragma unroll
      for (int r = 1; r < kSaMaxRanks; ++r) { acc.x += x[r].x; acc.y += x[r].y; acc.z += x[r].z; acc.w += x[r].w; }
      return acc;
    }
  };
  auto store4 = [&](long long e, float4 v) {
    if constexpr (NVLS) {
      sa_mm_st_f4(a.g_mc + e, v);
    } else {
#pragma unroll
      for (int r = 0; r < kSaMaxRanks; ++r)
        if (r < W) *reinterpret_cast<float4*>(a.g[r] + e) = v;
    }
  };

  // ---- dense part: the float4s of d0 | d1, split into W contiguous chunks (multiples of 32 float4 = 512 B)
  {
    const long long n0 = (a.d0_end - a.d0_begin) >> 2, n1 = (a.d1_end - a.d1_begin) >> 2, n = n0 + n1;
    const long long per = ((n + W - 1) / W + 31) / 32 * 32;
    const long long jb = (long long)a.rank * per, je = min(n, jb + per);
    const long long stride = (long long)gridDim.x * kSaThreads;
    constexpr int U = 2;
    long long j = jb + (long long)blockIdx.x * kSaThreads + threadIdx.x;
    auto elem = [&](long long jj) { return (jj < n0 ? (a.d0_begin >> 2) + jj : (a.d1_begin >> 2) + (jj - n0)) << 2; };
    for (; j + (U - 1) * stride < je; j += stride * U) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) v[u] = reduce4(elem(j + u * stride));
#pragma unroll
      for (int u = 0; u < U; ++u) store4(elem(j + u * stride), v[u]);
    }
    for (; j < je; j += stride) store4(elem(j), reduce4(elem(j)));
  }

  trace_max(TR_AR_DENSE, 2);
  // ---- touched rows: the row list (identical on every rank) split into W equal runs; one thread per (row, 16-byte
  // chunk), two items in flight per thread
  {
    const long long count = __ldcg(a.row_count);
    const long long per = (count + W - 1) / W;
    const long long lo = (long long)a.rank * per * G, hi = min(count, (long long)(a.rank + 1) * per) * G;
    const long long stride = (long long)gridDim.x * kSaThreads;
    auto one = [&](long long it, long long& e4, long long& e1) {
      const int r = __ldcg(a.rows + it / G);
      const int c = (int)(it % G);
      e4 = a.fmv_off + ((long long)r * G + c) * 4;
      e1 = c == 0 ? a.fmw_off + r : -1;  // the row's first-order gradient rides with chunk 0
    };
    auto scalar = [&](long long e) {
      if constexpr (NVLS) {
        sa_mm_st_f1(a.g_mc + e, sa_mm_ld_reduce_f1(a.g_mc + e));
      } else {
        float acc = 0.f;
#pragma unroll
        for (int r = 0; r < kSaMaxRanks; ++r)
          if (r < W) acc += sa_ld_volatile_f1(a.g[r] + e);
#pragma unroll
        for (int r = 0; r < kSaMaxRanks; ++r)
          if (r < W) a.g[r][e] = acc;
      }
    };
    long long it = lo + (long long)blockIdx.x * kSaThreads + threadIdx.x;
    for (; it + stride < hi; it += 2 * stride) {
      long long ea, sa_, eb, sb_;
      one(it, ea, sa_);
      one(it + stride, eb, sb_);
      const float4 va = reduce4(ea), vb = reduce4(eb);
      store4(ea, va);
      store4(eb, vb);
      if (sa_ >= 0) scalar(sa_);
      if (sb_ >= 0) scalar(sb_);
    }
    for (; it < hi; it += stride) {
      long long ea, sa_;
      one(it, ea, sa_);
      store4(ea, reduce4(ea));
      if (sa_ >= 0) scalar(sa_);
    }
  }
  trace_max(TR_AR_ROWS, 2);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = (atomicAdd(a.done_counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  // last CTA of this rank: ONE system fence (it waits for the NVLink write acknowledgements), then barrier B — when it
  // completes every rank's stores into this rank's gradient buffer are visible to the kernels that follow in the stream
  __threadfence_system();
  trace_max(TR_AR_FENCE, 2);
  __syncthreads();
  if (threadIdx.x == 0) {
    *a.done_counter = 0u;
    *a.seq_ab = seq;
  }
  sa_flag_barrier(a, kSaFlagB, seq, 1);
  trace_end(TR_ALLREDUCE);
}

static void fill_ptrs(SaArgs& a, void* const* g_ptrs, void* const* flag_ptrs, void* const* loss_ptrs,
                      void* const* bm_ptrs, int rank, int world) {
  for (int r = 0; r < kSaMaxRanks; ++r) {
    const int s = r < world ? r : 0;
    a.g[r] = g_ptrs ? (float*)g_ptrs[s] : nullptr;
    a.flags[r] = (uint32_t*)flag_ptrs[s];
    a.loss_parts[r] = loss_ptrs ? (float*)loss_ptrs[s] : nullptr;
    a.bm_in[r] = bm_ptrs ? (uint32_t*)bm_ptrs[s] : nullptr;
  }
  a.rank = rank; a.world = world;
}

template <bool NVLS, int G>
static int launch_sa(const SaArgs& a, int blocks, cudaStream_t st) {
  DFM_CHECK(launch_k(sparse_allreduce_kernel<NVLS, G>, dim3((unsigned)blocks), dim3(kSaThreads), 0, st, a));
  return DFM_OK;
}

}  // namespace dfm

using namespace dfm;

extern "C" {

int dfm_sparse_allreduce_flag_words(void) { return 3 * kSaMaxRanks; }

// Store this rank's touched-row bitmap (bm_own, nwords = ceil(N/32) words; cleared afterwards) into slot [rank] of every
// rank's inbox (bm_ptrs: host array of `world` device pointers to [8][nwords] inboxes in symmetric memory; bm_mc: multicast
// mapping of the inbox or NULL), then flag barrier C. seq_c: one zero-initialised device word (advanced by the kernel).
int dfm_bitmap_exchange(void* const* bm_ptrs, void* bm_mc, void* const* flag_ptrs, unsigned int* bm_own, int N, int rank,
                        int world, unsigned int* done_counter, unsigned int* seq_c, int use_nvls, void* stream) {
  wd_bind((cudaStream_t)stream);
  DFM_REQUIRE(world >= 1 && world <= kSaMaxRanks && rank >= 0 && rank < world, "bitmap_exchange: bad rank/world");
  DFM_REQUIRE(!use_nvls || bm_mc, "bitmap_exchange: NVLS requested without a multicast mapping");
  SaArgs a = {};
  fill_ptrs(a, nullptr, flag_ptrs, nullptr, bm_ptrs, rank, world);
  a.bm_in_mc = (uint32_t*)bm_mc; a.bm_own = bm_own; a.N = N; a.nwords = (N + 31) / 32;
  a.done_counter = done_counter; a.seq_c = seq_c;
  int blocks = ceil_div(a.nwords, kSaThreads);
  if (blocks > 32) blocks = 32;
  if (use_nvls)
    DFM_CHECK(launch_k(bitmap_exchange_kernel<true>, dim3(blocks), dim3(kSaThreads), 0, (cudaStream_t)stream, a));
  else
    DFM_CHECK(launch_k(bitmap_exchange_kernel<false>, dim3(blocks), dim3(kSaThreads), 0, (cudaStream_t)stream, a));
  return DFM_OK;
}

// Barrier A (deposits *loss_sum into loss_parts[rank] of every peer) + in-place all-reduce of the dense ranges d0 | d1 and
// of the fm_w / fm_v entries of the rows rows[0 .. *row_count) (dfm_union_rowlist), over the peer-mapped gradient buffers
// (g_ptrs; g_mc = NVLS multicast mapping or NULL) + barrier B. Every rank must pass the same ranges, N, K and hold the same
// row list. seq_ab: one zero-initialised device word (advanced by the kernel). done_counter: zeroed scratch word.
int dfm_sparse_allreduce(void* const* g_ptrs, void* g_mc, void* const* flag_ptrs, void* const* loss_ptrs, int rank,
                         int world, long long fmw_off, long long fmv_off, int N, int K, long long d0_begin,
                         long long d0_end, long long d1_begin, long long d1_end, const int* rows, const int* row_count,
                         const float* loss_sum, unsigned int* done_counter, unsigned int* seq_ab, int use_nvls,
                         int grid_blocks, void* stream) {
  wd_bind((cudaStream_t)stream);
  DFM_REQUIRE(world >= 1 && world <= kSaMaxRanks && rank >= 0 && rank < world, "sparse_allreduce: bad rank/world");
  DFM_REQUIRE(!use_nvls || g_mc, "sparse_allreduce: NVLS requested without a multicast mapping");
  DFM_REQUIRE(fmw_off % 4 == 0 && fmv_off % 4 == 0 && d0_begin % 4 == 0 && d0_end % 4 == 0 && d1_begin % 4 == 0 &&
                  d1_end % 4 == 0, "sparse_allreduce: offsets and ranges must be multiples of 4");
  DFM_REQUIRE(grid_blocks >= 1 && grid_blocks <= 148 * 8, "sparse_allreduce: grid_blocks %d not in [1,1184]", grid_blocks);
  SaArgs a = {};
  fill_ptrs(a, g_ptrs, flag_ptrs, loss_ptrs, nullptr, rank, world);
  a.g_mc = (float*)g_mc; a.fmw_off = fmw_off; a.fmv_off = fmv_off; a.N = N; a.nwords = (N + 31) / 32;
  a.d0_begin = d0_begin; a.d0_end = d0_end; a.d1_begin = d1_begin; a.d1_end = d1_end;
  a.rows = rows; a.row_count = row_count; a.loss_sum = loss_sum; a.done_counter = done_counter; a.seq_ab = seq_ab;
  cudaStream_t st = (cudaStream_t)stream;
  DFM_CHECK(launch_k(sa_barrier_a_kernel, dim3(1), dim3(32), 0, st, a));
  switch (K) {
#define SA_CASE(KK) \
    case KK: return use_nvls ? launch_sa<true, KK / 4>(a, grid_blocks, st) : launch_sa<false, KK / 4>(a, grid_blocks, st);
    SA_CASE(4) SA_CASE(8) SA_CASE(16) SA_CASE(32) SA_CASE(64) SA_CASE(128)
#undef SA_CASE
    default:
      dfm_set_error("sparse_allreduce: embedding_size %d unsupported (need 4,8,16,32,64,128)", K);
      return DFM_ERR_UNSUPPORTED;
  }
}

}  // extern "C"
