// Row-sharded embeddings, PUSH flavour of the fused all-to-all (large batches).
//
// csrc/sharded.cu pulls rows with random 256-byte peer LOADS; NVLink serves those at ~150 GB/s. Stores are
// fire-and-forget and stream at several times that, so for large batches the owner does the random access locally
// and pushes the result with P2P STORES into the requester's buffers:
//
//   a2a_dispatch    requester: (field, sample) lookups are bucketed by owning rank (warp-aggregated slot allocation)
//                   and written as 16-byte requests {local_row, dest, val} into the OWNER's inbox over NVLink
//   a2a_serve       owner:     [barrier] gathers each requested row from its local shard, scales it by the feature
//                   value and stores the fp32 row (and fm_w*val) straight into the REQUESTER's activation buffer
//   a2a_combine     requester: [barrier] pools the delivered rows: S = sum_f e, FM terms, bf16 deep input
//   a2a_grad_push   requester: gradient row of every lookup stored into the OWNER's gradient inbox (same slot as the request)
//   a2a_grad_apply  owner:     [barrier] segment-sums the inbox into its gradient shard with LOCAL red.add (duplicate ids
//                   across ranks and samples meet here), marks touched rows; sparse_row_opt (sharded.cu) then updates them
//
// One step = 3 cross-GPU barriers (+1 in sparse_row_opt); the forward needs no "owners finished updating" barrier of its
// own because a rank reaches the dispatch barrier only after its own row update of the previous step.
#include "common.cuh"

namespace dfm {

constexpr int kPsMaxRanks = 8;
constexpr int kPsThreads = 256;
constexpr int kPsMaxBlocks = 1184;

struct Req {
  int lrow;
  int dest;
  float val;
  int pad;
};

struct PushPtrs {
  Req* req[kPsMaxRanks];         // owner inbox: [src][cap] requests
  int* req_cnt[kPsMaxRanks];     // owner inbox: [src] number of requests
  float* rows_in[kPsMaxRanks];   // requester: [cap, K] delivered rows (e = v * val)
  float* wv_in[kPsMaxRanks];     // requester: [cap] delivered fm_w[id] * val
  float* grad_in[kPsMaxRanks];   // owner inbox: [src][cap][K] gradient rows
  float* gw_in[kPsMaxRanks];     // owner inbox: [src][cap] first-order gradients
  uint32_t* flags[kPsMaxRanks];  // barrier flags [3 phases][kPsMaxBlocks][kPsMaxRanks]
  int rank, world;
  long long cap;
};

// Barrier across ranks at kernel entry (common.cuh grid_entry_barrier: CTA 0 exchanges flags, the rest wait for its local
// go word in block slot 1 of the phase).
__device__ __forceinline__ void push_barrier(const PushPtrs& pp, int phase, uint32_t seq) {
  const size_t slot = (size_t)phase * kPsMaxBlocks * kPsMaxRanks;
  grid_entry_barrier(pp.flags, slot, pick8(pp.flags, pp.rank) + slot + kPsMaxRanks, pp.rank, pp.world, seq, 10 + phase);
}

// ------------------------------------------------------------------------------------------------ dispatch
__global__ void __launch_bounds__(kPsThreads)
a2a_dispatch_kernel(const PushPtrs pp, const int* __restrict__ ids, const float* __restrict__ vals, int B, int F,
                    long long N, int* __restrict__ cnt_local, int* __restrict__ pos_out, unsigned int* done_counter) {
  const int lane = threadIdx.x & 31;
  const int W = pp.world;
  const long long p = (long long)blockIdx.x * kPsThreads + threadIdx.x;
  const bool active = p < (long long)B * F;
  int id = 0;
  float v = 0.f;
  if (active) {
    const int f = (int)(p / B), b = (int)(p % B);  // field-major pair index, as in emb_bwd
    id = __ldg(ids + (size_t)b * F + f);
    v = __ldg(vals + (size_t)b * F + f);
  }
  const bool valid = active && v != 0.f && (unsigned)id < (unsigned long long)N;
  const int owner = valid ? id % W : 0;
  const unsigned key = valid ? (unsigned)owner : (100u + (unsigned)lane);
  const unsigned mask = __match_any_sync(0xffffffffu, key);
  const int leader = __ffs(mask) - 1;
  int base = 0;
  if (valid && lane == leader) base = atomicAdd(cnt_local + owner, __popc(mask));  // one atomic per owner per warp
  base = __shfl_sync(0xffffffffu, base, leader);
  const int pos = base + __popc(mask & ((1u << lane) - 1u));
  if (active) pos_out[p] = valid ? pos : -1;
  if (valid) {
    Req r;
    r.lrow = id / W; r.dest = (int)p; r.val = v; r.pad = 0;
    *reinterpret_cast<int4*>(pick8(pp.req, owner) + (size_t)pp.rank * pp.cap + pos) = *reinterpret_cast<const int4*>(&r);
  }
  // last CTA publishes the per-owner request counts into the owners' inboxes
  __shared__ bool s_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();  // one fence per CTA, cumulative over the CTA's request stores (ordered by the barrier above)
    s_last = (atomicAdd(done_counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    __threadfence();
    if (threadIdx.x < W) {
      const int n = atomicAdd(cnt_local + threadIdx.x, 0);
      pick8(pp.req_cnt, (int)threadIdx.x)[pp.rank] = n;
    }
    if (threadIdx.x == 0) *done_counter = 0u;
  }
}

// ------------------------------------------------------------------------------------------------ serve
template <int G>
__global__ void __launch_bounds__(kPsThreads)
a2a_serve_kernel(const PushPtrs pp, const float* __restrict__ V, const float* __restrict__ Wt,
                 const unsigned int* __restrict__ seq_ptr) {
  constexpr int K = 4 * G;
  push_barrier(pp, 0, *seq_ptr + 1u);  // every rank's requests and counts have arrived
  const int g = threadIdx.x % G;
  const long long group = ((long long)blockIdx.x * kPsThreads + threadIdx.x) / G;
  const long long ngroups = (long long)gridDim.x * kPsThreads / G;
  for (int src = 0; src < pp.world; ++src) {
    const int n = pick8(pp.req_cnt, pp.rank)[src];
    const Req* inbox = pick8(pp.req, pp.rank) + (size_t)src * pp.cap;
    float* rows_dst = pick8(pp.rows_in, src);
    float* wv_dst = pick8(pp.wv_in, src);
    for (long long e = group; e < n; e += ngroups) {
      const int4 raw = *reinterpret_cast<const int4*>(inbox + e);
      const int lrow = raw.x, dest = raw.y;
      const float val = __int_as_float(raw.z);
      const float4 row = __ldg(reinterpret_cast<const float4*>(V + (size_t)lrow * K) + g);
      *reinterpret_cast<float4*>(rows_dst + (size_t)dest * K + 4 * g) =
          make_float4(row.x * val, row.y * val, row.z * val, row.w * val);  // P2P store into the requester
      if (g == 0) wv_dst[dest] = __ldg(Wt + lrow) * val;
    }
  }
}

// ------------------------------------------------------------------------------------------------ combine
constexpr int kCombineUnroll = 4;

template <int G>
__global__ void __launch_bounds__(kPsThreads)
a2a_combine_kernel(const PushPtrs pp, const int* __restrict__ pos_of_pair, const float* __restrict__ rows,
                   const float* __restrict__ wv, __nv_bfloat16* __restrict__ x0, float* __restrict__ e32, int ldx,
                   float* __restrict__ S, float* __restrict__ yfm, int B, int F, const unsigned int* __restrict__ seq_ptr) {
  constexpr int K = 4 * G;
  constexpr int FG = 32 / G;
  push_barrier(pp, 1, *seq_ptr + 1u);  // every owner has delivered my rows
  const int lane = threadIdx.x & 31;
  const int g = lane % G, fg = lane / G;
  for (int sample = blockIdx.x * (kPsThreads / 32) + (threadIdx.x >> 5); sample < B; sample += gridDim.x * (kPsThreads / 32)) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f), ss = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int f0 = fg; f0 < F; f0 += FG * kCombineUnroll) {  // kCombineUnroll rows in flight per sub-group
      bool have[kCombineUnroll];
      float4 row[kCombineUnroll];
#pragma unroll
      for (int u = 0; u < kCombineUnroll; ++u) {
        const int f = f0 + u * FG;
        have[u] = f < F && pos_of_pair[(long long)f * B + sample] >= 0;
      }
#pragma unroll
      for (int u = 0; u < kCombineUnroll; ++u) {
        const long long p = (long long)(f0 + u * FG) * B + sample;
        row[u] = have[u] ? *reinterpret_cast<const float4*>(rows + (size_t)p * K + 4 * g) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < kCombineUnroll; ++u) {
        const int f = f0 + u * FG;
        if (f < F) {
          const float4 e = row[u];
          s.x += e.x; s.y += e.y; s.z += e.z; s.w += e.w;
          ss.x += e.x * e.x; ss.y += e.y * e.y; ss.z += e.z * e.z; ss.w += e.w * e.w;
          uint2 pk;
          pk.x = pack_bf16x2(e.x, e.y);
          pk.y = pack_bf16x2(e.z, e.w);
          *reinterpret_cast<uint2*>(x0 + (size_t)sample * ldx + (size_t)f * K + 4 * g) = pk;
          if (e32) *reinterpret_cast<float4*>(e32 + ((size_t)sample * F + f) * K + 4 * g) = e;  // dense [B, F*K]
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
      const long long p = (long long)f * B + sample;
      if (pos_of_pair[p] >= 0) acc += wv[p];
    }
    acc = warp_sum(acc);
    if (lane == 0) yfm[sample] = acc;
  }
}

// ------------------------------------------------------------------------------------------------ gradient push
template <int G>
__global__ void __launch_bounds__(kPsThreads)
a2a_grad_push_kernel(const PushPtrs pp, const int* __restrict__ ids, const float* __restrict__ vals,
                     const int* __restrict__ pos_of_pair, const float* __restrict__ rows, const float* __restrict__ S,
                     const float* __restrict__ dy, const __nv_bfloat16* __restrict__ dx0, int ldx, int B, int F) {
  constexpr int K = 4 * G;
  const int g = threadIdx.x % G;
  const long long p = ((long long)blockIdx.x * kPsThreads + threadIdx.x) / G;
  if (p >= (long long)B * F) return;
  const int pos = pos_of_pair[p];
  if (pos < 0) return;
  const int f = (int)(p / B), b = (int)(p % B);
  const int id = __ldg(ids + (size_t)b * F + f);
  const float v = __ldg(vals + (size_t)b * F + f);
  const int owner = id % pp.world;
  const float d = __ldg(dy + b);
  const float4 e = *reinterpret_cast<const float4*>(rows + (size_t)p * K + 4 * g);  // the delivered row, pair-major
  const float4 s = *reinterpret_cast<const float4*>(S + (size_t)b * K + 4 * g);
  const uint2 dx = *reinterpret_cast<const uint2*>(dx0 + (size_t)b * ldx + (size_t)f * K + 4 * g);
  const float4 gr = make_float4(v * (bf16_lo(dx.x) + d * (s.x - e.x)), v * (bf16_hi(dx.x) + d * (s.y - e.y)),
                                v * (bf16_lo(dx.y) + d * (s.z - e.z)), v * (bf16_hi(dx.y) + d * (s.w - e.w)));
  const size_t slot = (size_t)pp.rank * pp.cap + pos;
  *reinterpret_cast<float4*>(pick8(pp.grad_in, owner) + slot * K + 4 * g) = gr;  // P2P store into the owner's inbox
  if (g == 0) pick8(pp.gw_in, owner)[slot] = d * v;
}

// ------------------------------------------------------------------------------------------------ gradient apply
// A warp takes 32 consecutive inbox entries; its 32 / G sub-groups each walk G of them in order and MERGE RUNS of equal
// rows in registers before the red.add (dispatch numbers the requests of one requester warp consecutively and pairs are
// field-major, so a hot row - Criteo's numeric fields use one id for every sample - arrives as runs of up to 32 equal
// entries: same-address L2 atomics, which serialise at ~10 ns each, drop by the run length). Loads go 4 entries at a time
// in front of the (fire-and-forget) reductions. First-order gradients and bitmap bits are merged with match_any.
template <int G>
__global__ void __launch_bounds__(kPsThreads)
a2a_grad_apply_kernel(const PushPtrs pp, float* __restrict__ GV, float* __restrict__ GW, uint32_t* __restrict__ bitmap,
                      const unsigned int* __restrict__ seq_ptr) {
  constexpr int K = 4 * G;
  constexpr int U = G < 4 ? G : 4;
  push_barrier(pp, 2, *seq_ptr + 1u);  // every rank's gradient rows have arrived
  const int lane = threadIdx.x & 31;
  const int g = lane % G, sub = lane / G;
  const long long warp_global = (long long)blockIdx.x * (kPsThreads / 32) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (kPsThreads / 32);
  for (int src = 0; src < pp.world; ++src) {
    const int n = pick8(pp.req_cnt, pp.rank)[src];
    const Req* inbox = pick8(pp.req, pp.rank) + (size_t)src * pp.cap;
    const float* gin = pick8(pp.grad_in, pp.rank) + (size_t)src * pp.cap * K;
    const float* gwin = pick8(pp.gw_in, pp.rank) + (size_t)src * pp.cap;
    for (long long e0 = warp_global * 32; e0 < n; e0 += nwarps * 32) {
      const bool has = e0 + lane < n;
      const int my_row = has ? inbox[e0 + lane].lrow : -1;
      float my_gw = has ? gwin[e0 + lane] : 0.f;
      // ---- embedding rows: sub-group `sub` owns entries [sub * G, sub * G + G) of this trip
      const float* gsub = gin + (size_t)(e0 + sub * G) * K + 4 * g;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      int cur = -1;
#pragma unroll
      for (int i0 = 0; i0 < G; i0 += U) {
        int r[U];
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) r[u] = __shfl_sync(0xffffffffu, my_row, sub * G + i0 + u);
#pragma unroll
        for (int u = 0; u < U; ++u)
          v[u] = r[u] >= 0 ? *reinterpret_cast<const float4*>(gsub + (size_t)(i0 + u) * K) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (r[u] != cur) {
            if (cur >= 0) red_add_f32x4(GV + (size_t)cur * K + 4 * g, acc.x, acc.y, acc.z, acc.w);  // local segment-sum
            acc = make_float4(0.f, 0.f, 0.f, 0.f);
            cur = r[u];
          }
          acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w;
        }
      }
      if (cur >= 0) red_add_f32x4(GV + (size_t)cur * K + 4 * g, acc.x, acc.y, acc.z, acc.w);
      // ---- first-order gradient + touched bit: one atomic per distinct row of the trip
      const unsigned key = has ? (unsigned)my_row : (0x80000000u | (unsigned)lane);
      const unsigned mask = __match_any_sync(0xffffffffu, key);
      const int leader = __ffs(mask) - 1;
      if (__popc(mask) > 1) {
        float sum = my_gw;
        unsigned rem = mask & ~(1u << leader);
        while (rem) {
          const int srcl = __ffs(rem) - 1;
          rem &= rem - 1;
          sum += __shfl_sync(mask, my_gw, srcl);
        }
        my_gw = sum;
      }
      if (has && lane == leader) {
        atomicAdd(GW + my_row, my_gw);
        atomicOr(bitmap + (my_row >> 5), 1u << (my_row & 31));
      }
    }
  }
}

}  // namespace dfm

using namespace dfm;

static int fill_push(PushPtrs& pp, void* const* req, void* const* cnt, void* const* rows, void* const* wv,
                     void* const* gin, void* const* gwin, void* const* fl, int rank, int world, long long cap) {
  DFM_REQUIRE(world >= 1 && world <= kPsMaxRanks && rank >= 0 && rank < world, "a2a push: bad rank/world");
  for (int r = 0; r < kPsMaxRanks; ++r) {
    const int s = r < world ? r : 0;
    pp.req[r] = (Req*)req[s]; pp.req_cnt[r] = (int*)cnt[s]; pp.rows_in[r] = (float*)rows[s]; pp.wv_in[r] = (float*)wv[s];
    pp.grad_in[r] = (float*)gin[s]; pp.gw_in[r] = (float*)gwin[s]; pp.flags[r] = (uint32_t*)fl[s];
  }
  pp.rank = rank; pp.world = world; pp.cap = cap;
  return DFM_OK;
}

#define DFM_PS_DISPATCH(K, CALL)                                                                            \
  switch (K) {                                                                                              \
    case 4: CALL(1); break; case 8: CALL(2); break; case 16: CALL(4); break; case 32: CALL(8); break;      \
    case 64: CALL(16); break; case 128: CALL(32); break;                                                    \
    default: dfm_set_error("embedding_size %d unsupported (need 4,8,16,32,64,128)", K); return DFM_ERR_UNSUPPORTED; \
  }

extern "C" {

int dfm_a2a_push_flag_words(void) { return 3 * kPsMaxBlocks * kPsMaxRanks; }

// Forward, push flavour: dispatch -> serve -> combine. Pointer tables as in dfm_emb_fm_fwd_a2a (symmetric memory).
// cap = B*F (inbox capacity per source). cnt_local: W ints zeroed by the caller before each step. grid_blocks identical
// on every rank.
int dfm_a2a_push_forward(void* const* req, void* const* cnt, void* const* rows, void* const* wv, void* const* gin,
                         void* const* gwin, void* const* fl, int rank, int world, long long cap, const int* ids,
                         const float* vals, const float* V, const float* Wt, int* cnt_local, int* pos_of_pair,
                         unsigned int* done_counter, void* x0, float* e32, int ldx, float* S, float* yfm, int B, int F,
                         int K, long long N, const unsigned int* seq_ptr, int grid_blocks, void* stream) {
  wd_bind((cudaStream_t)stream);
  PushPtrs pp;
  int rc = fill_push(pp, req, cnt, rows, wv, gin, gwin, fl, rank, world, cap);
  if (rc) return rc;
  DFM_REQUIRE(grid_blocks >= 1 && grid_blocks <= kPsMaxBlocks, "a2a push: grid_blocks out of range");
  DFM_REQUIRE(cap >= (long long)B * F, "a2a push: inbox capacity smaller than B*F");
  cudaStream_t st = (cudaStream_t)stream;
  a2a_dispatch_kernel<<<(unsigned)ceil_div_ll((long long)B * F, kPsThreads), kPsThreads, 0, st>>>(
      pp, ids, vals, B, F, N, cnt_local, pos_of_pair, done_counter);
#define CALL(G)                                                                                                       \
  a2a_serve_kernel<G><<<grid_blocks, kPsThreads, 0, st>>>(pp, V, Wt, seq_ptr);                                        \
  a2a_combine_kernel<G><<<grid_blocks, kPsThreads, 0, st>>>(pp, pos_of_pair, (const float*)rows[rank],                \
                                                            (const float*)wv[rank], (__nv_bfloat16*)x0, e32, ldx, S, \
                                                            yfm, B, F, seq_ptr)
  DFM_PS_DISPATCH(K, CALL)
#undef CALL
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

// Backward, push flavour: grad_push -> grad_apply (sparse_row_opt follows, dfm_sparse_row_opt).
int dfm_a2a_push_backward(void* const* req, void* const* cnt, void* const* rows, void* const* wv, void* const* gin,
                          void* const* gwin, void* const* fl, int rank, int world, long long cap, const int* ids,
                          const float* vals, const int* pos_of_pair, const float* /*unused*/ e32, const float* S, const float* dy,
                          const void* dx0, int ldx, float* GV, float* GW, void* bitmap, int B, int F, int K,
                          const unsigned int* seq_ptr, int grid_blocks, void* stream) {
  wd_bind((cudaStream_t)stream);
  PushPtrs pp;
  int rc = fill_push(pp, req, cnt, rows, wv, gin, gwin, fl, rank, world, cap);
  if (rc) return rc;
  DFM_REQUIRE(grid_blocks >= 1 && grid_blocks <= kPsMaxBlocks, "a2a push: grid_blocks out of range");
  cudaStream_t st = (cudaStream_t)stream;
#define CALL(G)                                                                                                       \
  a2a_grad_push_kernel<G><<<(unsigned)ceil_div_ll((long long)B * F * G, kPsThreads), kPsThreads, 0, st>>>(            \
      pp, ids, vals, pos_of_pair, (const float*)rows[rank], S, dy, (const __nv_bfloat16*)dx0, ldx, B, F);             \
  a2a_grad_apply_kernel<G><<<grid_blocks, kPsThreads, 0, st>>>(pp, GV, GW, (uint32_t*)bitmap, seq_ptr)
  DFM_PS_DISPATCH(K, CALL)
#undef CALL
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

}  // extern "C"
