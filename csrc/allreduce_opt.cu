// dense_allreduce_opt — gradient all-reduce FUSED with the optimizer step, over NVLink 5 / NVSwitch peer memory.
//
// Replaces, in one kernel per step per rank (SURVEY K12 + X1; reference: Horovod's NCCL ring all-reduce of the
// 16-19 MB fusion buffer HVD:262 followed by 11 ApplyAdam launches PS:269):
//
//   1. cross-GPU barrier A   every rank's backward has finished writing its gradient buffer (and deposited its log-loss
//                            sum with every peer). Grids of at most one CTA per SM: CTA 0 of the sweep kernel runs it and
//                            releases the other CTAs through a local go word; larger grids: its own 1-CTA kernel
//   2. reduce-scatter        rank r sums ITS 1/W slice of the gradients across all W ranks
//                              P2P  : ld.volatile.global.v4.f32 from each peer's mapped buffer
//                              NVLS : multimem.ld_reduce.add.v4.f32 on the multicast address (reduced in the switch)
//   3. optimizer on slice    L2 decay of fm_w/fm_v + TF-1 Adam/Adagrad/Momentum/FTRL/GD on that slice only; optimizer
//                            slots are sharded W ways (the HBM sweep shrinks W-fold compared with a replicated step)
//   4. all-gather            updated fp32 weights and bf16 GEMM shadows stored straight into EVERY rank's replica
//                              P2P  : st.global.v4 to each peer pointer       NVLS : multimem.st
//   5. cross-GPU barrier B   by the last CTA to finish: all replicas are complete when any rank's kernel exits; it then
//                            publishes the GLOBAL loss sum and L2 term (scalars_out, pinned host ring)
//
// Gradients are NOT zeroed here (except an owner-routed range, which only its owner reads): peers may still be reading
// them; the step graph zeroes them on a forked branch that overlaps the next forward pass. Barriers use their own
// monotonically increasing sequence number (seq_ptr; never rewound, unlike the step counter), release/acquire at system
// scope, one flag per peer for A and one per peer for B. Only single CTAs ever spin on a peer's flag (the barrier kernel or
// CTA 0, and the last CTA of the sweep); CTAs waiting for CTA 0's local go word are all co-resident by construction.
// Peer-pointer tables arrive rotated by rank (entry 0 = this rank), so no kernel indexes a parameter array dynamically.
#include "common.cuh"
#include "optim_math.cuh"

#include <stdlib.h>

namespace dfm {

constexpr int kMaxRanks = 8;
constexpr int kArThreads = 256;
constexpr int kArMaxBlocks = 148 * 8;

struct ArOptArgs {
  // Peer tables are ROTATED by the host wrapper: entry j belongs to rank (rank + j) % world, so entry 0 is this rank and
  // the kernels index them with compile-time constants only (or pick8): no dynamically indexed parameter array, no stack
  // frame (see pick8 in common.cuh).
  float* g[kMaxRanks];             // every rank's flat gradient buffer (peer-mapped)
  float* p[kMaxRanks];             // every rank's flat parameter buffer
  __nv_bfloat16* sh[kMaxRanks];    // every rank's bf16 shadow buffer
  uint32_t* flags[kMaxRanks];      // every rank's flag array: [kMaxRanks] words for barrier A, then [kMaxRanks] for B
  float* reg_parts[kMaxRanks];     // every rank's [kMaxRanks] array of per-rank L2-loss partials
  float* g_mc;                     // multicast mappings (NVLS), may be null
  float* p_mc;
  __nv_bfloat16* sh_mc;
  float* s0;                       // local slot slices (index relative to slice_begin)
  float* s1;
  int rank, world;
  long long slice_begin, slice_end;  // this rank's element range (multiples of 4)
  OptHyper h;
  float grad_scale;
  long long decay_begin, decay_end;
  float l2;
  long long shadow_begin, shadow_end;
  unsigned long long* step_ptr;
  float* reg_local;                // scratch accumulator (1 float, zero between steps)
  unsigned int* done_counter;      // scratch (1 word, zero between steps)
  unsigned int* seq_ptr;           // barrier sequence number: monotonic, never rewound (unlike the step counter)
  // optional result publication (all may be null): barrier A deposits *loss_local in loss_parts[rank] of every peer; the
  // last CTA, after barrier B, writes the GLOBAL log-loss sum and the complete L2 term into scalars_out[0..1] and stores
  // them into record (step - 1) % pub_ring of a host-mapped ring (as dfm_flat_opt_step_pub does on one GPU)
  float* loss_parts[kMaxRanks];
  const float* loss_local;
  const float* reg_extra;          // optional: an L2 term this rank computed elsewhere (row-sharded tables), added to its part
  float* scalars_out;
  float* pub_dst;
  int pub_ring;
  long long local_begin, local_end;  // elements whose gradients arrive ALREADY reduced in this rank's own buffer (emb_bwd
                                     // routed them to their owner): read locally and zeroed here, no NVLink load
  int finalize;                    // 0: one bucket of several — no step advance, no L2 / loss publication (the last launch does)
  int barrier_inside;              // 1: barrier A runs in CTA 0 of the sweep kernel itself (grids of at most one CTA per SM)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_volatile_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_bf16x4(__nv_bfloat16* p, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// Cross-GPU barrier A as its own 1-CTA kernel: the sweep kernel that follows then needs no spin-wait at its start, so
// its grid is not limited to what is co-resident (it runs 8 CTAs/SM like the local optimizer).
__global__ void ar_barrier_kernel(const ArOptArgs a) {
  pdl_launch_dependents();  // let the sweep kernel's CTAs get resident; they block in pdl_wait() until this kernel exits
  trace_min(TR_BAR_A, 1);
  const uint32_t seq = *a.seq_ptr + 1u;
  if (threadIdx.x < a.world) {
    const int j = threadIdx.x;  // rotated slot; the peer behind it:
    const int peer = a.rank + j < a.world ? a.rank + j : a.rank + j - a.world;
    if (a.loss_local) pick8(a.loss_parts, j)[a.rank] = *a.loss_local;  // ordered before the release below (same thread)
    st_release_sys(pick8(a.flags, j) + a.rank, seq);
    wait_flag_geq(a.flags[0] + peer, seq, peer, 0);
  }
  trace_max(TR_BAR_A, 2);
}

template <bool NVLS>
__global__ void __launch_bounds__(kArThreads, 3) allreduce_opt_kernel(const ArOptArgs a) {
  __shared__ float s_lr;
  __shared__ float s_red[kArThreads / 32];
  __shared__ bool s_last;
  const int W = a.world;
  const unsigned long long t = *a.step_ptr + 1ull;
  const uint32_t seq = *a.seq_ptr + 1u;
  if (threadIdx.x == 0) s_lr = opt_step_lr(a.h, t);

  // barrier A: either the preceding 1-CTA kernel (ar_barrier_kernel), or — small grids — CTA 0 of this kernel while the
  // other CTAs wait for its local go flag. The separate kernel costs its launch plus ~6 us between its end and this
  // kernel's first instruction after griddepcontrol.wait (a kernel that wrote peer memory completes only when NVLink has
  // acknowledged those writes; scripts/step_trace.py at 8 GPUs: barrier end 70.1 us, sweep go 76.4 us).
  pdl_wait(TR_ALLREDUCE);
  if (a.barrier_inside) {
    uint32_t* go = a.flags[0] + 2 * kMaxRanks;
    if (blockIdx.x == 0) {
      if (threadIdx.x < W) {
        const int j = threadIdx.x;
        const int peer = a.rank + j < W ? a.rank + j : a.rank + j - W;
        if (a.loss_local) pick8(a.loss_parts, j)[a.rank] = *a.loss_local;
        st_release_sys(pick8(a.flags, j) + a.rank, seq);
        wait_flag_geq(a.flags[0] + peer, seq, peer, 0);
      }
      __syncthreads();
      if (threadIdx.x == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(go), "r"(seq) : "memory");
    } else if (threadIdx.x == 0) {
      uint32_t v;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(go) : "memory");
      } while ((int32_t)(v - seq) < 0);  // CTA 0 needs no other CTA to get here, and it has its own watchdog
    }
    trace_max(TR_BAR_A, 2);
  }
  __syncthreads();
  const float lr_t = s_lr;
  const int nslots = opt_num_slots(a.h.opt);
  float reg = 0.f;
  float* wl = a.p[0];
  const long long b4 = a.slice_begin >> 2, e4 = a.slice_end >> 2;
  // U float4 per thread per trip: every NVLink load of a trip (U x W peer loads, or U multimem loads) is issued before
  // the first use (branch-free main loop + scalar tail), so a trip costs ~one fabric round trip, not U x W of them.
  constexpr int U = 2;
  const long long stride = (long long)gridDim.x * kArThreads;
  auto reduce_one = [&](long long i) -> float4 {
    const long long e0 = i << 2;
    if (e0 >= a.local_begin && e0 < a.local_end) {  // owner-routed range: the sum is already here
      float4* lp = reinterpret_cast<float4*>(a.g[0]) + i;
      const float4 v = *lp;
      *lp = make_float4(0.f, 0.f, 0.f, 0.f);  // the owner zeroes its own slice (peers add to it again only after barrier B)
      return v;
    }
    if constexpr (NVLS) {
      return multimem_ld_reduce_f4(a.g_mc + (i << 2));
    } else {
      float4 x[kMaxRanks];
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r) {
        x[r] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < W) x[r] = ld_volatile_f4(a.g[r] + (i << 2));  // rotated table: peers are staggered across ranks
      }
      float4 acc = x[0];
#pragma unroll
      for (int r = 1; r < kMaxRanks; ++r) { acc.x += x[r].x; acc.y += x[r].y; acc.z += x[r].z; acc.w += x[r].w; }
      return acc;
    }
  };
  auto update_one = [&](long long i, float4 gg) {
    const long long e = i << 2;
    float4 w = reinterpret_cast<const float4*>(wl)[i];
    const long long li = i - b4;
    float4 s0 = nslots > 0 ? reinterpret_cast<float4*>(a.s0)[li] : make_float4(0, 0, 0, 0);
    float4 s1 = nslots > 1 ? reinterpret_cast<float4*>(a.s1)[li] : make_float4(0, 0, 0, 0);
    gg.x *= a.grad_scale; gg.y *= a.grad_scale; gg.z *= a.grad_scale; gg.w *= a.grad_scale;
    if (e >= a.decay_begin && e < a.decay_end) {
      reg += w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w;
      gg.x += a.l2 * w.x; gg.y += a.l2 * w.y; gg.z += a.l2 * w.z; gg.w += a.l2 * w.w;
    }
    opt_update(a.h, lr_t, gg.x, w.x, s0.x, s1.x);
    opt_update(a.h, lr_t, gg.y, w.y, s0.y, s1.y);
    opt_update(a.h, lr_t, gg.z, w.z, s0.z, s1.z);
    opt_update(a.h, lr_t, gg.w, w.w, s0.w, s1.w);
    if (nslots > 0) reinterpret_cast<float4*>(a.s0)[li] = s0;
    if (nslots > 1) reinterpret_cast<float4*>(a.s1)[li] = s1;
    const bool in_shadow = (e >= a.shadow_begin && e < a.shadow_end);
    uint2 pk = make_uint2(0u, 0u);
    if (in_shadow) {
      pk.x = pack_bf16x2(w.x, w.y);
      pk.y = pack_bf16x2(w.z, w.w);
    }
    if constexpr (NVLS) {
      multimem_st_f4(a.p_mc + e, w);
      if (in_shadow) multimem_st_bf16x4(a.sh_mc + (e - a.shadow_begin), pk);
    } else {
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r) {
        if (r < W) {
          reinterpret_cast<float4*>(a.p[r])[i] = w;
          if (in_shadow) *reinterpret_cast<uint2*>(a.sh[r] + (e - a.shadow_begin)) = pk;
        }
      }
    }
  };
  long long i0 = b4 + (long long)blockIdx.x * kArThreads + threadIdx.x;
  for (; i0 + (U - 1) * stride < e4; i0 += stride * U) {
    float4 g[U];
#pragma unroll
    for (int u = 0; u < U; ++u) g[u] = reduce_one(i0 + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) update_one(i0 + u * stride, g[u]);
  }
  for (; i0 < e4; i0 += stride) update_one(i0, reduce_one(i0));
  trace_max(TR_AR_DENSE, 2);  // last CTA done with its reduce + update + broadcast trips (stores still in flight)
  reg = warp_sum(reg);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = reg;
  __syncthreads();
  if (threadIdx.x == 0) {
    // gpu-scope fence per CTA only; the LAST CTA issues the one system-scope fence before barrier B (release
    // cumulativity: every CTA's peer stores happen-before it through the done-counter chain). A fence.sys waits for the
    // NVLink write acknowledgements (~5 us), so one per CTA wave or per thread is what made earlier versions slow.
    __threadfence();
    float tot = 0.f;
    for (int i = 0; i < kArThreads / 32; ++i) tot += s_red[i];
    if (tot != 0.f) atomicAdd(a.reg_local, tot);
    __threadfence();
    s_last = (atomicAdd(a.done_counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;

  // ---- last CTA of this rank: publish the L2-loss partial, advance the step, barrier B
  __threadfence();
  if (threadIdx.x < W && a.finalize) {
    const float part = 0.5f * a.l2 * atomicAdd(a.reg_local, 0.f) + (a.reg_extra ? __ldcg(a.reg_extra) : 0.f);
    pick8(a.reg_parts, (int)threadIdx.x)[a.rank] = part;  // slot [rank] on every peer
  }
  __threadfence_system();
  trace_max(TR_AR_FENCE, 2);  // NVLink has acknowledged every peer store of this rank
  __syncthreads();
  if (threadIdx.x == 0) {
    *a.reg_local = 0.f;
    *a.done_counter = 0u;
    if (a.finalize) *a.step_ptr = t;
    *a.seq_ptr = seq;
  }
  if (threadIdx.x < W) {
    const int j = threadIdx.x;
    const int peer = a.rank + j < W ? a.rank + j : a.rank + j - W;
    uint32_t* bflags_peer = pick8(a.flags, j) + kMaxRanks;
    const uint32_t* bflags_mine = a.flags[0] + kMaxRanks;
    st_release_sys(bflags_peer + a.rank, seq);
    wait_flag_geq(bflags_mine + peer, seq, peer, 1);
  }
  trace_end(TR_ALLREDUCE);
  if (a.scalars_out && a.finalize) {  // every peer's loss sum (barrier A) and L2 partial (before its barrier-B release) are visible now
    __syncthreads();
    if (threadIdx.x == 0) {
      float ls = 0.f, rs = 0.f;
      for (int r = 0; r < W; ++r) {  // fixed order: the same values on every rank
        ls += __ldcg(a.loss_parts[0] + r);
        rs += __ldcg(a.reg_parts[0] + r);
      }
      a.scalars_out[0] = ls;
      a.scalars_out[1] = rs;
      if (a.pub_dst) {
        float* rec = a.pub_dst + ((t - 1ull) % (unsigned long long)a.pub_ring) * 2ull;
        rec[0] = ls;
        rec[1] = rs;
      }
    }
  }
}

// bcast_state (reference: BroadcastGlobalVariablesHook(0), HVD:372 — an NCCL broadcast of ~36
// tensors at the start of every train() call): the source rank writes its buffer into every replica with one pass of
// multicast stores (NVLS) or per-peer P2P stores. Callers fence with a host barrier before and after (start/resume only).
template <bool NVLS>
__global__ void __launch_bounds__(256) bcast_kernel(const ArOptArgs a, long long n4) {
  const float4* src = reinterpret_cast<const float4*>(a.p[0]);
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 v = src[i];
    if constexpr (NVLS) {
      multimem_st_f4(a.p_mc + (i << 2), v);
    } else {
#pragma unroll
      for (int r = 1; r < kMaxRanks; ++r) {
        if (r < a.world) reinterpret_cast<float4*>(a.p[r])[i] = v;
      }
    }
  }
  __threadfence_system();
}

}  // namespace dfm

using namespace dfm;

extern "C" {

// Broadcast n fp32 elements (multiple of 4) from this rank's buffer into every peer's buffer at the same offset.
int dfm_bcast_f32(void* const* ptrs, void* mc, long long n, int rank, int world, int use_nvls, void* stream) {
  DFM_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, "bcast: bad rank/world");
  DFM_REQUIRE(n > 0 && n % 4 == 0, "bcast: n must be a positive multiple of 4");
  DFM_REQUIRE(!use_nvls || mc, "bcast: NVLS requested without a multicast mapping");
  ArOptArgs a = {};
  for (int r = 0; r < kMaxRanks; ++r) a.p[r] = (float*)ptrs[r < world ? (rank + r) % world : rank];  // rotated
  a.p_mc = (float*)mc; a.rank = rank; a.world = world;
  long long blocks = ceil_div_ll(n / 4, 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (use_nvls) bcast_kernel<true><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a, n / 4);
  else bcast_kernel<false><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a, n / 4);
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

int dfm_allreduce_opt_flag_words(void) { return 2 * kMaxRanks + 1; }  // barrier A, barrier B, local go flag
int dfm_allreduce_opt_max_blocks(void) { return kArMaxBlocks; }

// Pointer tables are host arrays of `world` device pointers (peer-mapped symmetric memory). mc pointers may be NULL
// (then the P2P path is used). s0/s1: this rank's slot slices. Ranges in elements, multiples of 4.
// loss_ptrs (host array of `world` device pointers to 8-float inboxes in symmetric memory), loss_local (this rank's
// log-loss sum), reg_extra (an L2 term of this rank computed elsewhere, added to its partial), [local_begin, local_end)
// (elements whose gradients dfm_emb_bwd_routed already reduced into their owner's buffer: read locally, then zeroed),
// finalize (0 = this launch is an early bucket of the step: own flags / seq / scratch, no step advance or publication), scalars_out (2 floats: global loss sum, L2 term) and pub_ring (pinned HOST ring of pub_ring_len records of 2
// floats) are optional: all NULL gives the plain step.
int dfm_allreduce_opt_step_pub(void* const* g_ptrs, void* const* p_ptrs, void* const* sh_ptrs, void* const* flag_ptrs,
                               void* const* reg_ptrs, void* g_mc, void* p_mc, void* sh_mc, float* s0, float* s1, int rank,
                               int world, long long slice_begin, long long slice_end, int opt, float lr, float beta1,
                               float beta2, float eps, float momentum, float grad_scale, long long decay_begin,
                               long long decay_end, float l2, long long shadow_begin, long long shadow_end,
                               unsigned long long* step_ptr, float* reg_local, unsigned int* done_counter,
                               unsigned int* seq_ptr, int use_nvls, int grid_blocks, void* const* loss_ptrs,
                               const float* loss_local, const float* reg_extra, float* scalars_out, float* pub_ring,
                               int pub_ring_len, long long local_begin, long long local_end, int finalize, void* stream);

int dfm_allreduce_opt_step(void* const* g_ptrs, void* const* p_ptrs, void* const* sh_ptrs, void* const* flag_ptrs,
                           void* const* reg_ptrs, void* g_mc, void* p_mc, void* sh_mc, float* s0, float* s1, int rank,
                           int world, long long slice_begin, long long slice_end, int opt, float lr, float beta1,
                           float beta2, float eps, float momentum, float grad_scale, long long decay_begin,
                           long long decay_end, float l2, long long shadow_begin, long long shadow_end,
                           unsigned long long* step_ptr, float* reg_local, unsigned int* done_counter,
                           unsigned int* seq_ptr, int use_nvls, int grid_blocks, void* stream) {
  return dfm_allreduce_opt_step_pub(g_ptrs, p_ptrs, sh_ptrs, flag_ptrs, reg_ptrs, g_mc, p_mc, sh_mc, s0, s1, rank, world,
                                    slice_begin, slice_end, opt, lr, beta1, beta2, eps, momentum, grad_scale, decay_begin,
                                    decay_end, l2, shadow_begin, shadow_end, step_ptr, reg_local, done_counter, seq_ptr,
                                    use_nvls, grid_blocks, nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0, 0, 1, stream);
}

int dfm_allreduce_opt_step_pub(void* const* g_ptrs, void* const* p_ptrs, void* const* sh_ptrs, void* const* flag_ptrs,
                               void* const* reg_ptrs, void* g_mc, void* p_mc, void* sh_mc, float* s0, float* s1, int rank,
                               int world, long long slice_begin, long long slice_end, int opt, float lr, float beta1,
                               float beta2, float eps, float momentum, float grad_scale, long long decay_begin,
                               long long decay_end, float l2, long long shadow_begin, long long shadow_end,
                               unsigned long long* step_ptr, float* reg_local, unsigned int* done_counter,
                               unsigned int* seq_ptr, int use_nvls, int grid_blocks, void* const* loss_ptrs,
                               const float* loss_local, const float* reg_extra, float* scalars_out, float* pub_ring,
                               int pub_ring_len, long long local_begin, long long local_end, int finalize, void* stream) {
  wd_bind((cudaStream_t)stream);
  DFM_REQUIRE(local_begin % 4 == 0 && local_end % 4 == 0 && local_end >= local_begin,
              "allreduce_opt: the owner-routed range must be ordered multiples of 4");
  DFM_REQUIRE((loss_ptrs == nullptr) == (loss_local == nullptr) && (loss_ptrs != nullptr || scalars_out == nullptr),
              "allreduce_opt: loss_ptrs, loss_local (and scalars_out) go together");
  DFM_REQUIRE(pub_ring == nullptr || (scalars_out != nullptr && pub_ring_len >= 1),
              "allreduce_opt: publication needs scalars_out and a ring length >= 1");
  DFM_REQUIRE(world >= 1 && world <= kMaxRanks, "allreduce_opt: world size %d not in [1,%d]", world, kMaxRanks);
  DFM_REQUIRE(rank >= 0 && rank < world, "allreduce_opt: bad rank");
  DFM_REQUIRE(slice_begin % 4 == 0 && slice_end % 4 == 0 && slice_end >= slice_begin, "allreduce_opt: bad slice");
  DFM_REQUIRE(!use_nvls || (g_mc && p_mc && sh_mc), "allreduce_opt: NVLS requested without multicast mappings");
  ArOptArgs a;
  for (int r = 0; r < kMaxRanks; ++r) {
    const int s = r < world ? (rank + r) % world : rank;  // rotated: entry 0 = this rank (see ArOptArgs)
    a.g[r] = (float*)g_ptrs[s];
    a.p[r] = (float*)p_ptrs[s];
    a.sh[r] = (__nv_bfloat16*)sh_ptrs[s];
    a.flags[r] = (uint32_t*)flag_ptrs[s];
    a.reg_parts[r] = (float*)reg_ptrs[s];
    a.loss_parts[r] = loss_ptrs ? (float*)loss_ptrs[s] : nullptr;
  }
  a.local_begin = local_begin; a.local_end = local_end; a.finalize = finalize;
  a.loss_local = loss_local; a.reg_extra = reg_extra; a.scalars_out = scalars_out; a.pub_dst = pub_ring; a.pub_ring = pub_ring_len;
  a.g_mc = (float*)g_mc; a.p_mc = (float*)p_mc; a.sh_mc = (__nv_bfloat16*)sh_mc;
  a.s0 = s0; a.s1 = s1; a.rank = rank; a.world = world;
  a.slice_begin = slice_begin; a.slice_end = slice_end;
  a.h.opt = opt; a.h.lr = lr; a.h.beta1 = beta1; a.h.beta2 = beta2; a.h.eps = eps; a.h.momentum = momentum;
  a.grad_scale = grad_scale; a.decay_begin = decay_begin; a.decay_end = decay_end; a.l2 = l2;
  a.shadow_begin = shadow_begin; a.shadow_end = shadow_end;
  a.step_ptr = step_ptr; a.reg_local = reg_local; a.done_counter = done_counter; a.seq_ptr = seq_ptr;
  // barrier A is the 1-CTA kernel below (flags[0..7]); barrier B runs in the sweep's last CTA (flags[8..15])
  DFM_REQUIRE(grid_blocks >= 1 && grid_blocks <= kArMaxBlocks, "allreduce_opt: grid_blocks %d not in [1,%d]",
              grid_blocks, kArMaxBlocks);
  const long long blocks = grid_blocks;
  // every CTA of a grid of at most one CTA per SM is resident at once, so CTAs spinning on CTA 0's go flag cannot keep it
  // from being scheduled; larger grids (the P2P flavour at 2 ranks) keep the separate 1-CTA barrier kernel
  static const bool inside_ok = !(getenv("DFM_AR_BARRIER_KERNEL") && getenv("DFM_AR_BARRIER_KERNEL")[0] == '1');
  a.barrier_inside = (inside_ok && grid_blocks <= 148) ? 1 : 0;
  if (!a.barrier_inside && (world > 1 || loss_local)) {  // world 1 with publication: the loss sum is still deposited
    ar_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a);
    DFM_CHECK(cudaGetLastError());
  }
  if (use_nvls)
    DFM_CHECK(launch_k(allreduce_opt_kernel<true>, dim3((unsigned)blocks), dim3(kArThreads), 0, (cudaStream_t)stream, a));
  else
    DFM_CHECK(launch_k(allreduce_opt_kernel<false>, dim3((unsigned)blocks), dim3(kArThreads), 0, (cudaStream_t)stream, a));
  return DFM_OK;
}

}  // extern "C"
