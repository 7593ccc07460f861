// Row-sharded embedding tables with the all-to-all FUSED into the embedding kernels (X5).
//
// fm_w / fm_v are split by row across the W GPUs of the box (row `id` lives on rank id % W at local row id / W); the
// MLP stays data parallel. Every rank maps every peer's table shard, gradient shard and touched-row bitmap through
// symmetric memory, so "all-to-all" is plain loads / stores / reductions on peer pointers issued from inside the
// compute kernels — there is no pack / exchange / unpack pass and no NCCL call:
//
//   emb_fm_fwd_a2a    dispatch + gather + scale + pool + combine in one kernel: each (sample, field) row is fetched
//                     with 128-bit loads straight from the OWNING rank's shard over NVSwitch, multiplied by the feature
//                     value, written as the bf16 deep input (and fp32 copy for backward), and reduced into the FM terms.
//   emb_bwd_a2a       sparse-gradient all-to-all + duplicate-id segment-sum in one kernel: lanes holding the same id are
//                     combined in-warp (match/shuffle), then pushed with red.global.add.v4.f32 INTO THE OWNER'S gradient
//                     shard over NVLink (the remote reduction is the cross-rank segment-sum); the owner's touched-row
//                     bitmap is set with a remote atomic OR.
//   sparse_row_opt    the owner walks its bitmap and applies the TF-1 update rule to touched rows only (lazy / "touched"
//                     L2 mode), zeroing the gradient rows and bitmap words it consumes.
//
// Ordering: emb_fm_fwd_a2a begins with a cross-GPU barrier ("every owner finished last step's row updates"),
// sparse_row_opt begins with one ("every rank finished pushing gradients"). A barrier is ONE flag exchange by CTA 0 with
// release/acquire at system scope on monotonically increasing sequence numbers (the other CTAs wait for CTA 0's local go
// word), so no grid size or CTA dispatch order is assumed.
#include "common.cuh"
#include "optim_math.cuh"

namespace dfm {

constexpr int kShMaxRanks = 8;
constexpr int kShThreads = 256;
constexpr int kShMaxBlocks = 1184;  // flag slots per barrier phase

struct ShardPtrs {
  const float* v[kShMaxRanks];   // table shards [rows_local, K]
  const float* w[kShMaxRanks];   // first-order shards [rows_local]
  float* gv[kShMaxRanks];        // gradient shards
  float* gw[kShMaxRanks];
  uint32_t* bitmap[kShMaxRanks]; // touched-row bitmaps [ceil(rows_local/32)]
  uint32_t* flags[kShMaxRanks];  // barrier flags [2 phases][kShMaxBlocks][kShMaxRanks]
  int rank, world;
};

__device__ __forceinline__ void sh_st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t sh_ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Barrier across ranks at kernel entry (common.cuh grid_entry_barrier): CTA 0 exchanges the flags of block slot 0 of this
// phase, the other CTAs wait for the local go word kept in block slot 1 (which no peer ever writes).
__device__ __forceinline__ void cross_rank_barrier(const ShardPtrs& sp, int phase, uint32_t seq) {
  const size_t slot = (size_t)phase * kShMaxBlocks * kShMaxRanks;
  grid_entry_barrier(sp.flags, slot, pick8(sp.flags, sp.rank) + slot + kShMaxRanks, sp.rank, sp.world, seq, 2 + phase);
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {  // peer data changes every step: never the nc path
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_peer_f1(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

constexpr int kShUnroll = 4;

// One warp per sample (lane = field-group * G + 4-column chunk), rows fetched from the owning rank's shard.
template <int G>
__global__ void __launch_bounds__(kShThreads)
emb_fm_fwd_a2a_kernel(const ShardPtrs sp, const int* __restrict__ ids, const float* __restrict__ vals,
                      __nv_bfloat16* __restrict__ x0, float* __restrict__ e32, int ldx, float* __restrict__ S,
                      float* __restrict__ yfm, int B, int F, long long N, const unsigned int* __restrict__ seq_ptr,
                      int do_barrier) {
  constexpr int K = 4 * G;
  constexpr int FG = 32 / G;
  // training: wait until every owner has finished last step's row updates. Evaluation / prediction on one rank only
  // (the reference evaluates on the chief, HVD:393) reads quiescent tables and must NOT wait for idle peers.
  if (do_barrier) cross_rank_barrier(sp, 0, *seq_ptr + 1u);
  const int lane = threadIdx.x & 31;
  const int W = sp.world;
  for (int sample = blockIdx.x * (kShThreads / 32) + (threadIdx.x >> 5); sample < B; sample += gridDim.x * (kShThreads / 32)) {
    const int g = lane % G, fg = lane / G;
    const int* sid = ids + (size_t)sample * F;
    const float* sval = vals + (size_t)sample * F;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f), ss = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int f0 = fg; f0 < F; f0 += FG * kShUnroll) {
      int id[kShUnroll];
      float val[kShUnroll];
      float4 row[kShUnroll];
#pragma unroll
      for (int u = 0; u < kShUnroll; ++u) {
        const int f = f0 + u * FG;
        id[u] = 0; val[u] = 0.f;
        if (f < F) {
          id[u] = __ldg(sid + f);
          val[u] = __ldg(sval + f);
          if ((unsigned)id[u] >= (unsigned long long)N) { id[u] = 0; val[u] = 0.f; }
        }
      }
#pragma unroll
      for (int u = 0; u < kShUnroll; ++u) {
        const int owner = id[u] % W;
        const size_t lrow = (size_t)(id[u] / W);
        row[u] = ld_peer_f4(pick8(sp.v, owner) + lrow * K + 4 * g);
      }
#pragma unroll
      for (int u = 0; u < kShUnroll; ++u) {
        const int f = f0 + u * FG;
        if (f < F) {
          const float4 e = make_float4(row[u].x * val[u], row[u].y * val[u], row[u].z * val[u], row[u].w * val[u]);
          s.x += e.x; s.y += e.y; s.z += e.z; s.w += e.w;
          ss.x += e.x * e.x; ss.y += e.y * e.y; ss.z += e.z * e.z; ss.w += e.w * e.w;
          uint2 pk;
          pk.x = pack_bf16x2(e.x, e.y);
          pk.y = pack_bf16x2(e.z, e.w);
          *reinterpret_cast<uint2*>(x0 + (size_t)sample * ldx + (size_t)f * K + 4 * g) = pk;
          *reinterpret_cast<float4*>(e32 + ((size_t)sample * F + f) * K + 4 * g) = e;  // e32 is dense [B, F*K]
        }
      }
    }
#pragma unroll
    for (int o = G; o < 32; o <<= 1) {
      s.x += __shfl_xor_sync(0xffffffffu, s.x, o); s.y += __shfl_xor_sync(0xffffffffu, s.y, o);
      s.z += __shfl_xor_sync(0xffffffffu, s.z, o); s.w += __shfl_xor_sync(0xffffffffu, s.w, o);
      ss.x += __shfl_xor_sync(0xffffffffu, ss.x, o); ss.y += __shfl_xor_sync(0xffffffffu, ss.y, o);
      ss.z += __shfl_xor_sync(0xffffffffu, ss.z, o); ss.w += __shfl_xor_sync(0xffffffffu, ss.w, o);
    }
    float acc = 0.f;
    if (fg == 0) {
      reinterpret_cast<float4*>(S + (size_t)sample * K)[g] = s;
      acc = 0.5f * ((s.x * s.x - ss.x) + (s.y * s.y - ss.y) + (s.z * s.z - ss.z) + (s.w * s.w - ss.w));
    }
    for (int f = lane; f < F; f += 32) {
      const int id = __ldg(sid + f);
      if ((unsigned)id < (unsigned long long)N) acc += ld_peer_f1(pick8(sp.w, id % W) + (size_t)(id / W)) * __ldg(sval + f);
    }
    acc = warp_sum(acc);
    if (lane == 0) yfm[sample] = acc;
  }
}

// One lane per (field, sample) pair in field-major order; duplicate ids combined in-warp, then pushed to the owner.
template <int G>
__global__ void __launch_bounds__(kShThreads)
emb_bwd_a2a_kernel(const ShardPtrs sp, const int* __restrict__ ids, const float* __restrict__ vals,
                   const float* __restrict__ e32, const float* __restrict__ S, const float* __restrict__ dy,
                   const __nv_bfloat16* __restrict__ dx0, int ldx, int B, int F, long long N) {
  constexpr int K = 4 * G;
  const int lane = threadIdx.x & 31;
  const int W = sp.world;
  const long long p = (long long)blockIdx.x * kShThreads + threadIdx.x;
  const bool active = p < (long long)B * F;
  const int f = active ? (int)(p / B) : 0;
  const int b = active ? (int)(p % B) : 0;
  int id = 0;
  float v = 0.f;
  if (active) {
    id = __ldg(ids + (size_t)b * F + f);
    v = __ldg(vals + (size_t)b * F + f);
  }
  const bool valid = active && v != 0.f && (unsigned)id < (unsigned long long)N;
  const unsigned key = valid ? (unsigned)id : (0x80000000u | (unsigned)lane);
  const unsigned mask = __match_any_sync(0xffffffffu, key);
  const int leader = __ffs(mask) - 1;
  const bool is_leader = valid && lane == leader;
  const float d = valid ? __ldg(dy + b) : 0.f;
  const int owner = valid ? id % W : 0;
  const size_t lrow = valid ? (size_t)(id / W) : 0;
  float* gdst = pick8(sp.gv, owner) + lrow * K;
  const float4* ep = reinterpret_cast<const float4*>(e32 + ((size_t)b * F + f) * K);  // dense [B, F*K]
  const float4* spp = reinterpret_cast<const float4*>(S + (size_t)b * K);
  const uint2* dxp = reinterpret_cast<const uint2*>(dx0 + (size_t)b * ldx + (size_t)f * K);
  // the row's G 16-byte chunks are split over gridDim.y CTAs (shorter serial chain per thread), as in emb_bwd
  const int g_per = (G + (int)gridDim.y - 1) / (int)gridDim.y;
  const int g_begin = (int)blockIdx.y * g_per, g_end = min(G, g_begin + g_per);
#pragma unroll 2
  for (int g = g_begin; g < g_end; ++g) {
    float4 gr = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) {
      const float4 e = __ldg(ep + g);
      const float4 s = __ldg(spp + g);
      const uint2 dx = __ldg(dxp + g);
      gr.x = v * (bf16_lo(dx.x) + d * (s.x - e.x));
      gr.y = v * (bf16_hi(dx.x) + d * (s.y - e.y));
      gr.z = v * (bf16_lo(dx.y) + d * (s.z - e.z));
      gr.w = v * (bf16_hi(dx.y) + d * (s.w - e.w));
    }
    if (mask == 0xffffffffu) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        gr.x += __shfl_xor_sync(0xffffffffu, gr.x, o); gr.y += __shfl_xor_sync(0xffffffffu, gr.y, o);
        gr.z += __shfl_xor_sync(0xffffffffu, gr.z, o); gr.w += __shfl_xor_sync(0xffffffffu, gr.w, o);
      }
    } else if (__popc(mask) > 1) {
      float4 acc = gr;
      unsigned rem = mask & ~(1u << leader);
      while (rem) {
        const int src = __ffs(rem) - 1;
        rem &= rem - 1;
        acc.x += __shfl_sync(mask, gr.x, src); acc.y += __shfl_sync(mask, gr.y, src);
        acc.z += __shfl_sync(mask, gr.z, src); acc.w += __shfl_sync(mask, gr.w, src);
      }
      gr = acc;
    }
    if (is_leader) red_add_f32x4(gdst + 4 * g, gr.x, gr.y, gr.z, gr.w);  // remote reduction over NVLink
  }
  if (blockIdx.y != 0) return;  // CTA-uniform: first-order gradient and touched bit once per pair
  float gw = valid ? d * v : 0.f;
  if (mask == 0xffffffffu) {
    gw = warp_sum(gw);
  } else if (__popc(mask) > 1) {
    float acc = gw;
    unsigned rem = mask & ~(1u << leader);
    while (rem) {
      const int src = __ffs(rem) - 1;
      rem &= rem - 1;
      acc += __shfl_sync(mask, gw, src);
    }
    gw = acc;
  }
  if (is_leader) {
    atomicAdd(pick8(sp.gw, owner) + lrow, gw);
    atomicOr(pick8(sp.bitmap, owner) + (lrow >> 5), 1u << (lrow & 31));  // mark the row touched on its owner
  }
}

struct SparseOptArgs {
  float* v; float* w; float* gv; float* gw;     // local shards
  float* v_s0; float* v_s1; float* w_s0; float* w_s1;
  uint32_t* bitmap;
  long long rows_local;
  OptHyper h;
  float l2;
  const unsigned long long* step_ptr;  // MLP optimizer's step counter (already advanced for this step or not: see host)
  int step_offset;
  unsigned int* seq_ptr;
  unsigned int* done_counter;
  float* reg_loss;
};

// Row update of one touched row by the G lanes of a sub-group: loads first (all independent), then math, then stores.
template <int G>
struct RowState {
  float4 gr, w, s0, s1;
  float gw1, w1, t0, t1;
};

template <int G>
__device__ __forceinline__ void row_load(const SparseOptArgs& a, long long row, int g, int nslots, RowState<G>& r) {
  constexpr int K = 4 * G;
  r.gr = reinterpret_cast<const float4*>(a.gv + row * K)[g];
  r.w = reinterpret_cast<const float4*>(a.v + row * K)[g];
  r.s0 = nslots > 0 ? reinterpret_cast<const float4*>(a.v_s0 + row * K)[g] : make_float4(0, 0, 0, 0);
  r.s1 = nslots > 1 ? reinterpret_cast<const float4*>(a.v_s1 + row * K)[g] : make_float4(0, 0, 0, 0);
  if (g == 0) {  // first-order weight of the same row
    r.gw1 = a.gw[row]; r.w1 = a.w[row];
    r.t0 = nslots > 0 ? a.w_s0[row] : 0.f;
    r.t1 = nslots > 1 ? a.w_s1[row] : 0.f;
  }
}

template <int G>
__device__ __forceinline__ float row_update_store(const SparseOptArgs& a, long long row, int g, int nslots, float lr_t,
                                                  RowState<G>& r) {
  constexpr int K = 4 * G;
  float reg = r.w.x * r.w.x + r.w.y * r.w.y + r.w.z * r.w.z + r.w.w * r.w.w;
  r.gr.x += a.l2 * r.w.x; r.gr.y += a.l2 * r.w.y; r.gr.z += a.l2 * r.w.z; r.gr.w += a.l2 * r.w.w;
  opt_update(a.h, lr_t, r.gr.x, r.w.x, r.s0.x, r.s1.x);
  opt_update(a.h, lr_t, r.gr.y, r.w.y, r.s0.y, r.s1.y);
  opt_update(a.h, lr_t, r.gr.z, r.w.z, r.s0.z, r.s1.z);
  opt_update(a.h, lr_t, r.gr.w, r.w.w, r.s0.w, r.s1.w);
  reinterpret_cast<float4*>(a.v + row * K)[g] = r.w;
  reinterpret_cast<float4*>(a.gv + row * K)[g] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (nslots > 0) reinterpret_cast<float4*>(a.v_s0 + row * K)[g] = r.s0;
  if (nslots > 1) reinterpret_cast<float4*>(a.v_s1 + row * K)[g] = r.s1;
  if (g == 0) {
    reg += r.w1 * r.w1;
    r.gw1 += a.l2 * r.w1;
    opt_update(a.h, lr_t, r.gw1, r.w1, r.t0, r.t1);
    a.w[row] = r.w1;
    a.gw[row] = 0.f;
    if (nslots > 0) a.w_s0[row] = r.t0;
    if (nslots > 1) a.w_s1[row] = r.t1;
  }
  return reg;
}

// Bitmap words are dealt to warps round-robin (word wd belongs to warp wd % nwarps): hot ids are small ids inside each
// field's range, so touched rows cluster, and a contiguous split left a few warps with hundreds of rows while 60 % of the
// SM cycles idled (ncu: 63 us, mostly stall_barrier). A warp reads up to 32 of its words with ONE load instruction,
// expands the set bits into a per-warp row list in shared memory (warp scan of the popcounts) and consumes the list
// 2 * (32 / G) rows at a time with every load of a trip in flight before the first update - instead of a chain of
// ~14 dependent word loads, each followed by its own dependent row loads.
template <int G>
__global__ void __launch_bounds__(kShThreads, 4)
sparse_row_opt_kernel(const ShardPtrs sp, const SparseOptArgs a) {
  constexpr int RPW = 32 / G;  // rows processed in parallel by one warp
  __shared__ float s_lr;
  __shared__ unsigned short s_rows[kShThreads / 32][32 * 32];  // (word slot in the trip) * 32 + bit
  const uint32_t seq = *a.seq_ptr + 1u;
  cross_rank_barrier(sp, 1, seq);
  const unsigned long long t = *a.step_ptr + (unsigned long long)a.step_offset;
  if (threadIdx.x == 0) s_lr = opt_step_lr(a.h, t);
  __syncthreads();
  const float lr_t = s_lr;
  const int nslots = opt_num_slots(a.h.opt);
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int g = lane % G, sub = lane / G;
  const long long nwords = (a.rows_local + 31) >> 5;
  const long long warp_global = (long long)blockIdx.x * (kShThreads / 32) + wib;
  const long long nwarps = (long long)gridDim.x * (kShThreads / 32);
  unsigned short* rows = s_rows[wib];
  float reg = 0.f;
  for (long long k0 = 0; k0 * nwarps + warp_global < nwords; k0 += 32) {  // warp-uniform
    const long long wd = (k0 + lane) * nwarps + warp_global;
    uint32_t bits = wd < nwords ? a.bitmap[wd] : 0u;
    if (bits) a.bitmap[wd] = 0u;
    // exclusive warp scan of the popcounts -> slot of this lane's first row in the list
    const int cnt = __popc(bits);
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += up;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == 0) continue;  // warp-uniform
    int slot = incl - cnt;
    while (bits) {
      rows[slot++] = (unsigned short)((lane << 5) + (__ffs(bits) - 1));
      bits &= bits - 1;
    }
    __syncwarp();
    for (int q = sub; q < total; q += 2 * RPW) {
      const int c0 = rows[q];
      const long long r0 = (((k0 + (c0 >> 5)) * nwarps + warp_global) << 5) + (c0 & 31);
      const bool two = q + RPW < total;
      const int c1 = two ? rows[q + RPW] : c0;
      const long long r1 = (((k0 + (c1 >> 5)) * nwarps + warp_global) << 5) + (c1 & 31);
      RowState<G> A, Bq;
      row_load<G>(a, r0, g, nslots, A);
      if (two) row_load<G>(a, r1, g, nslots, Bq);
      reg += row_update_store<G>(a, r0, g, nslots, lr_t, A);
      if (two) reg += row_update_store<G>(a, r1, g, nslots, lr_t, Bq);
    }
    __syncwarp();  // the list is rewritten by the next trip
  }
  reg = warp_sum(reg);
  if (lane == 0 && reg != 0.f && a.reg_loss) atomicAdd(a.reg_loss, 0.5f * a.l2 * reg);
  // Updated rows must be visible to peers before the next forward's barrier. One gpu-scope fence per CTA (thread 0, after
  // the CTA barrier: cumulative over the CTA's stores) chained through the done counter, and ONE system-scope fence in
  // the last CTA - a fence.sys per thread (303 K of them) was most of this kernel's 59 us (profiles/kernels_ncu.md).
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(a.done_counter, 1u) == gridDim.x - 1) {
      __threadfence_system();
      *a.done_counter = 0u;
      *a.seq_ptr = seq;
    }
  }
}

}  // namespace dfm

using namespace dfm;

static int fill_ptrs(ShardPtrs& sp, void* const* v, void* const* w, void* const* gv, void* const* gw, void* const* bm,
                     void* const* fl, int rank, int world) {
  DFM_REQUIRE(world >= 1 && world <= kShMaxRanks && rank >= 0 && rank < world, "sharded: bad rank/world %d/%d", rank, world);
  for (int r = 0; r < kShMaxRanks; ++r) {
    const int s = r < world ? r : 0;
    sp.v[r] = (const float*)v[s]; sp.w[r] = (const float*)w[s];
    sp.gv[r] = (float*)gv[s]; sp.gw[r] = (float*)gw[s];
    sp.bitmap[r] = (uint32_t*)bm[s]; sp.flags[r] = (uint32_t*)fl[s];
  }
  sp.rank = rank; sp.world = world;
  return DFM_OK;
}

#define DFM_SH_DISPATCH(K, CALL)                                                                            \
  switch (K) {                                                                                              \
    case 4: CALL(1); break; case 8: CALL(2); break; case 16: CALL(4); break; case 32: CALL(8); break;      \
    case 64: CALL(16); break; case 128: CALL(32); break;                                                    \
    default: dfm_set_error("embedding_size %d unsupported (need 4,8,16,32,64,128)", K); return DFM_ERR_UNSUPPORTED; \
  }

extern "C" {

int dfm_sharded_flag_words(void) { return 2 * kShMaxBlocks * kShMaxRanks; }

// grid_blocks must be identical on every rank (<= 1184). x0 bf16 has row stride ldx (padded activation rows); e32 f32 is
// dense [B, F*K].
int dfm_emb_fm_fwd_a2a(void* const* v, void* const* w, void* const* gv, void* const* gw, void* const* bm, void* const* fl,
                       int rank, int world, const int* ids, const float* vals, void* x0, float* e32, int ldx, float* S,
                       float* yfm, int B, int F, int K, long long N, const unsigned int* seq_ptr, int do_barrier,
                       int grid_blocks, void* stream) {
  wd_bind((cudaStream_t)stream);
  ShardPtrs sp;
  int rc = fill_ptrs(sp, v, w, gv, gw, bm, fl, rank, world);
  if (rc) return rc;
  DFM_REQUIRE(grid_blocks >= 1 && grid_blocks <= kShMaxBlocks, "fwd_a2a: grid_blocks out of range");
#define CALL(G) emb_fm_fwd_a2a_kernel<G><<<grid_blocks, kShThreads, 0, (cudaStream_t)stream>>>( \
      sp, ids, vals, (__nv_bfloat16*)x0, e32, ldx, S, yfm, B, F, N, seq_ptr, do_barrier)
  DFM_SH_DISPATCH(K, CALL)
#undef CALL
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

int dfm_emb_bwd_a2a(void* const* v, void* const* w, void* const* gv, void* const* gw, void* const* bm, void* const* fl,
                    int rank, int world, const int* ids, const float* vals, const float* e32, const float* S,
                    const float* dy, const void* dx0, int ldx, int B, int F, int K, long long N, void* stream) {
  wd_bind((cudaStream_t)stream);
  ShardPtrs sp;
  int rc = fill_ptrs(sp, v, w, gv, gw, bm, fl, rank, world);
  if (rc) return rc;
  const unsigned blocks = (unsigned)ceil_div_ll((long long)B * F, kShThreads);
  // split the K/4 chunks over up to 4 CTAs in y while the grid stays under ~8 CTAs per SM
  const int Gk = K / 4;
  unsigned ysplit = 1;
  while (ysplit < 4 && ysplit * 2 <= (unsigned)Gk && (unsigned long long)blocks * ysplit * 2 <= 148ull * 8) ysplit *= 2;
#define CALL(G) emb_bwd_a2a_kernel<G><<<dim3(blocks, ysplit), kShThreads, 0, (cudaStream_t)stream>>>( \
      sp, ids, vals, e32, S, dy, (const __nv_bfloat16*)dx0, ldx, B, F, N)
  DFM_SH_DISPATCH(K, CALL)
#undef CALL
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

int dfm_sparse_row_opt(void* const* v, void* const* w, void* const* gv, void* const* gw, void* const* bm, void* const* fl,
                       int rank, int world, float* v_s0, float* v_s1, float* w_s0, float* w_s1, long long rows_local,
                       int K, int opt, float lr, float beta1, float beta2, float eps, float momentum, float l2,
                       const unsigned long long* step_ptr, int step_offset, unsigned int* seq_ptr,
                       unsigned int* done_counter, float* reg_loss, int grid_blocks, void* stream) {
  wd_bind((cudaStream_t)stream);
  ShardPtrs sp;
  int rc = fill_ptrs(sp, v, w, gv, gw, bm, fl, rank, world);
  if (rc) return rc;
  DFM_REQUIRE(grid_blocks >= 1 && grid_blocks <= kShMaxBlocks, "sparse_row_opt: grid_blocks out of range");
  SparseOptArgs a;
  a.v = (float*)v[rank]; a.w = (float*)w[rank]; a.gv = (float*)gv[rank]; a.gw = (float*)gw[rank];
  a.v_s0 = v_s0; a.v_s1 = v_s1; a.w_s0 = w_s0; a.w_s1 = w_s1;
  a.bitmap = (uint32_t*)bm[rank]; a.rows_local = rows_local;
  a.h.opt = opt; a.h.lr = lr; a.h.beta1 = beta1; a.h.beta2 = beta2; a.h.eps = eps; a.h.momentum = momentum;
  a.l2 = l2; a.step_ptr = step_ptr; a.step_offset = step_offset; a.seq_ptr = seq_ptr; a.done_counter = done_counter;
  a.reg_loss = reg_loss;
#define CALL(G) sparse_row_opt_kernel<G><<<grid_blocks, kShThreads, 0, (cudaStream_t)stream>>>(sp, a)
  DFM_SH_DISPATCH(K, CALL)
#undef CALL
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

}  // extern "C"
