// Split dense optimizer for replicated embedding tables: the whole-table sweep leaves the step's critical path.
//
// The reference's loss regularises the WHOLE fm_w / fm_v tables (PS:245-246), so TF applies a dense optimizer update to
// every row every step — one ~150 MB sweep at the end of each step (flat_opt, csrc/optim.cu). But only the
// rows a batch looks up have a data gradient; for every other row the update needs nothing the backward pass produces
// (its gradient is exactly l2 * w). So the step is split three ways, with bit-identical results:
//
//   touched_bitmap   (side stream, at step start)  one bit per table row the batch looks up; lanes that share a word are
//                    combined in the warp first (Criteo's numeric fields put the same id in every sample)
//   sweep_untouched  (side stream, concurrent with the forward and backward passes)  L2 decay + TF-1 update of every row
//                    NOT in the bitmap: reads w, m, v only (their gradient buffer entries are zero and stay zero), so it
//                    also moves 25 % fewer bytes than the fused sweep; forward/backward only ever read touched rows, so
//                    the two never meet. Under data parallelism the bitmap is the union over ranks (inputs: one bitmap
//                    per rank, exchanged by bitmap_exchange in csrc/sparse_allreduce.cu) and every replica runs the same
//                    sweep on the same values.
//   tail_opt         (main stream, after emb_bwd)  touched rows + fm_bias + the MLP block: consumes and zeroes their
//                    gradients, refreshes the bf16 GEMM shadows, finishes the L2 loss value, advances the step counter and
//                    publishes the step's result into the host ring (what flat_opt's last CTA did).
#include "common.cuh"
#include "optim_math.cuh"

namespace dfm {

constexpr int kSplitMaxRanks = 8;
constexpr int kSplitThreads = 256;

struct SplitOptArgs {
  float* w;
  float* g;
  float* s0;
  float* s1;
  OptHyper h;
  float grad_scale;
  float l2;                      // 0 = no decay (the sweep still runs: dense Adam/Momentum move rows with g == 0)
  long long fmw_off, fmv_off;    // element offsets of fm_w / fm_v in the flat buffers (multiples of 4)
  int N, G;                      // table rows; float4 chunks per fm_v row (K = 4 G)
  int nwords;                    // ceil(N / 32)
  long long d0_begin, d0_end;    // dense ranges of the tail (elements, multiples of 4): fm_bias block, MLP block
  long long d1_begin, d1_end;
  long long shadow_begin, shadow_end;
  __nv_bfloat16* shadow;
  unsigned long long* step_ptr;
  float* reg_acc;                // running sum of w^2 over the decay range (pre-update); reset by the tail's last CTA
  float* reg_loss;               // out: 0.5 * l2 * reg_acc
  unsigned int* done_counter;
  const uint32_t* bm_in[kSplitMaxRanks];
  int world;
  int clear_inputs;              // sweep: zero bm_in[0] after reading it (single GPU: the input IS the live bitmap)
  uint32_t* bm_union;
  int* rows;                     // ascending ids of the rows set in bm_union (built by union_rowlist), capacity N
  int* row_count;                // number of entries in `rows`
  int* word_prefix;              // optional [nwords]: set bits in the words before each word (rank of a row = prefix + popc)
  // tail only
  float* loss_sum;               // this rank's log-loss sum (scalars[0]); overwritten with the global sum when world > 1
  const float* loss_parts;       // [world] per-rank sums deposited by the peers (world > 1)
  const float* pub_src;
  float* pub_dst;
  int pub_n, pub_ring;
  int do_dense, do_sparse, finalize;
  int dense_ctas;                // tail with both parts: CTAs [0, dense_ctas) take the dense ranges, the rest the rows
};

__device__ __forceinline__ void upd4(const SplitOptArgs& a, float lr_t, float4& w, float4 g, float4& s0, float4& s1,
                                     bool decay, float& reg) {
  g.x *= a.grad_scale; g.y *= a.grad_scale; g.z *= a.grad_scale; g.w *= a.grad_scale;
  if (decay) {
    reg += w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w;
    g.x += a.l2 * w.x; g.y += a.l2 * w.y; g.z += a.l2 * w.z; g.w += a.l2 * w.w;
  }
  opt_update(a.h, lr_t, g.x, w.x, s0.x, s1.x);
  opt_update(a.h, lr_t, g.y, w.y, s0.y, s1.y);
  opt_update(a.h, lr_t, g.z, w.z, s0.z, s1.z);
  opt_update(a.h, lr_t, g.w, w.w, s0.w, s1.w);
}

__device__ __forceinline__ float block_sum_to_thread0(float v, float* s_red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
  __syncthreads();
  float tot = 0.f;
  if (threadIdx.x == 0)
    for (int i = 0; i < kSplitThreads / 32; ++i) tot += s_red[i];
  return tot;
}

// ------------------------------------------------------------------------------------------------ touched bitmap
// One lane per (field, sample), field-major (a warp = 32 samples of one field, as in emb_bwd): lanes whose ids fall in
// the same bitmap word are merged before the atomic, and a word that already holds the bits is not written again.
__global__ void __launch_bounds__(256) touched_bitmap_kernel(const int* __restrict__ ids, int B, int F, int N,
                                                             uint32_t* __restrict__ bitmap) {
  pdl_launch_dependents();
  pdl_wait(TR_BITMAP);
  const long long p = (long long)blockIdx.x * 256 + threadIdx.x;
  const bool active = p < (long long)B * F;
  const int f = active ? (int)(p / B) : 0;
  const int b = active ? (int)(p % B) : 0;
  const int id = active ? __ldg(ids + (size_t)b * F + f) : -1;
  const bool valid = active && (unsigned)id < (unsigned)N;
  const unsigned lane = threadIdx.x & 31;
  const unsigned word = valid ? ((unsigned)id >> 5) : (0x80000000u | lane);
  const unsigned bit = valid ? (1u << (id & 31)) : 0u;
  const unsigned grp = __match_any_sync(0xffffffffu, word);
  const unsigned bits = __reduce_or_sync(grp, bit);
  if (valid && lane == (unsigned)(__ffs(grp) - 1)) {
    if ((__ldcg(bitmap + word) & bits) != bits) atomicOr(bitmap + word, bits);
  }
}

// ------------------------------------------------------------------------------------------------ union + row list
// ONE CTA: OR of the per-rank bitmaps -> bm_union, and the ascending list of touched rows (block prefix sum over the word
// popcounts) -> rows[0 .. *row_count). Deterministic, so every rank derives the same list from the same bitmaps and the
// exchange can split it by position. Off the critical path (side branch), latency-bound: ~3 us at N = 117581. The list is
// what lets the tail and the exchange give every touched row its own threads at once; walking the bitmap word by word made
// a warp chain 2-4 dependent DRAM round trips per word (measured: 14 us for 20 K rows).
constexpr int kRowlistThreads = 1024;
__global__ void __launch_bounds__(kRowlistThreads) union_rowlist_kernel(const SplitOptArgs a) {
  __shared__ int s_warp[kRowlistThreads / 32];
  pdl_launch_dependents();
  pdl_wait(TR_ROWLIST);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int wpt = (a.nwords + kRowlistThreads - 1) / kRowlistThreads;  // consecutive words per thread
  const int w_begin = t * wpt, w_end = min(a.nwords, w_begin + wpt);
  int cnt = 0;
  constexpr int kChunk = 8;  // words whose loads (x world) are in flight together
  for (int w0 = w_begin; w0 < w_end; w0 += kChunk) {
    uint32_t u[kChunk];
#pragma unroll
    for (int j = 0; j < kChunk; ++j) {
      u[j] = 0;
      if (w0 + j < w_end) {
#pragma unroll
        for (int r = 0; r < kSplitMaxRanks; ++r)
          if (r < a.world) u[j] |= __ldcg(a.bm_in[r] + w0 + j);
      }
    }
#pragma unroll
    for (int j = 0; j < kChunk; ++j) {
      const int wi = w0 + j;
      if (wi < w_end) {
        const int nrows = min(32, a.N - wi * 32);
        if (nrows < 32) u[j] &= (1u << nrows) - 1u;
        a.bm_union[wi] = u[j];
        if (a.clear_inputs) const_cast<uint32_t*>(a.bm_in[0])[wi] = 0u;
        cnt += __popc(u[j]);
      }
    }
  }
  // exclusive block scan of cnt
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int v = s_warp[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int x = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += x;
    }
    s_warp[lane] = v;  // inclusive over warps
  }
  __syncthreads();
  int pos = incl - cnt + (warp > 0 ? s_warp[warp - 1] : 0);
  for (int wi = w_begin; wi < w_end; ++wi) {
    uint32_t u = a.bm_union[wi];
    if (a.word_prefix) a.word_prefix[wi] = pos;
    while (u) {
      const int b = __ffs(u) - 1;
      u &= u - 1;
      a.rows[pos++] = wi * 32 + b;
    }
  }
  if (t == kRowlistThreads - 1) *a.row_count = pos;
  trace_end(TR_ROWLIST);
}

// ------------------------------------------------------------------------------------------------ untouched sweep
// One warp per word of the union bitmap = 32 consecutive table rows: the rows whose bit is clear.
template <int G>
__global__ void __launch_bounds__(kSplitThreads) sweep_untouched_kernel(const SplitOptArgs a) {
  __shared__ float s_lr;
  __shared__ float s_red[kSplitThreads / 32];
  pdl_launch_dependents();
  pdl_wait(TR_SWEEP);
  const unsigned long long t = *a.step_ptr + 1ull;
  if (threadIdx.x == 0) s_lr = opt_step_lr(a.h, t);
  __syncthreads();
  const float lr_t = s_lr;
  const int nslots = opt_num_slots(a.h.opt);
  const bool decay = a.l2 != 0.f;
  const int lane = threadIdx.x & 31;
  const int warps_per_cta = kSplitThreads / 32;
  float reg = 0.f;
  for (int wi = blockIdx.x * warps_per_cta + (threadIdx.x >> 5); wi < a.nwords; wi += gridDim.x * warps_per_cta) {
    const uint32_t u = __ldcg(a.bm_union + wi);
    const int row0 = wi * 32;
    const int nrows = min(32, a.N - row0);
    // first-order weights: lane j <-> row row0 + j
    if (lane < nrows && !((u >> lane) & 1u)) {
      const long long e = a.fmw_off + row0 + lane;
      float w = a.w[e];
      float s0 = nslots > 0 ? a.s0[e] : 0.f, s1 = nslots > 1 ? a.s1[e] : 0.f;
      float g = 0.f;
      if (decay) { reg += w * w; g = a.l2 * w; }
      opt_update(a.h, lr_t, g, w, s0, s1);
      a.w[e] = w;
      if (nslots > 0) a.s0[e] = s0;
      if (nslots > 1) a.s1[e] = s1;
    }
    // embedding rows: 32 rows x G float4, contiguous in memory; chunk index c = it * 32 + lane -> row c / G
    const long long base4 = (a.fmv_off >> 2) + (long long)row0 * G;
    float4* W4 = reinterpret_cast<float4*>(a.w) + base4;
    float4* S04 = reinterpret_cast<float4*>(a.s0) + base4;
    float4* S14 = reinterpret_cast<float4*>(a.s1) + base4;
    constexpr int U = (G >= 4) ? 4 : G;  // chunks per lane in flight
#pragma unroll 1
    for (int it0 = 0; it0 < G; it0 += U) {
      float4 w[U], s0[U], s1[U];
      bool on[U];
#pragma unroll
      for (int k = 0; k < U; ++k) {
        const int c = (it0 + k) * 32 + lane;
        const int r = c / G;
        on[k] = (it0 + k) < G && r < nrows && !((u >> r) & 1u);
        w[k] = s0[k] = s1[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (on[k]) {  // streaming (evict-first) accesses: every line is touched once per step
          w[k] = __ldcs(W4 + c);
          if (nslots > 0) s0[k] = __ldcs(S04 + c);
          if (nslots > 1) s1[k] = __ldcs(S14 + c);
        }
      }
#pragma unroll
      for (int k = 0; k < U; ++k) {
        if (on[k]) {
          const int c = (it0 + k) * 32 + lane;
          upd4(a, lr_t, w[k], make_float4(0.f, 0.f, 0.f, 0.f), s0[k], s1[k], decay, reg);
          __stcs(W4 + c, w[k]);
          if (nslots > 0) __stcs(S04 + c, s0[k]);
          if (nslots > 1) __stcs(S14 + c, s1[k]);
        }
      }
    }
  }
  const float tot = block_sum_to_thread0(reg, s_red);
  if (threadIdx.x == 0 && tot != 0.f) atomicAdd(a.reg_acc, tot);
  trace_end(TR_SWEEP);
}

// ------------------------------------------------------------------------------------------------ tail
template <int G>
__global__ void __launch_bounds__(kSplitThreads) tail_opt_kernel(const SplitOptArgs a) {
  __shared__ float s_lr;
  __shared__ float s_red[kSplitThreads / 32];
  pdl_launch_dependents();
  pdl_wait(TR_OPT);
  const unsigned long long t = *a.step_ptr + 1ull;
  if (threadIdx.x == 0) s_lr = opt_step_lr(a.h, t);
  __syncthreads();
  const float lr_t = s_lr;
  const int nslots = opt_num_slots(a.h.opt);
  const bool decay = a.l2 != 0.f;
  const int lane = threadIdx.x & 31;
  float reg = 0.f;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);

  // the first `dense_ctas` CTAs sweep the dense ranges, the others take one bitmap word (32 table rows) per warp: the two
  // parts run side by side (run one after the other in every CTA, the row part — a chain of dependent DRAM round trips per
  // word — added 13 us behind the dense part's 6)
  const int dense_ctas = a.do_dense ? (a.do_sparse ? a.dense_ctas : (int)gridDim.x) : 0;
  if (a.do_dense && (int)blockIdx.x < dense_ctas) {
    const long long n0 = (a.d0_end - a.d0_begin) >> 2, n1 = (a.d1_end - a.d1_begin) >> 2;
    const long long stride = (long long)dense_ctas * kSplitThreads, n = n0 + n1;
    auto index_of = [&](long long j) { return j < n0 ? (a.d0_begin >> 2) + j : (a.d1_begin >> 2) + (j - n0); };
    constexpr int UD = 2;  // float4 per thread in flight (x 4 arrays)
    for (long long j0 = (long long)blockIdx.x * kSplitThreads + threadIdx.x; j0 < n; j0 += stride * UD) {
      long long i[UD];
      float4 w[UD], g[UD], s0[UD], s1[UD];
#pragma unroll
      for (int x = 0; x < UD; ++x) {
        const long long j = j0 + x * stride;
        i[x] = j < n ? index_of(j) : -1;
        w[x] = g[x] = s0[x] = s1[x] = z4;
        if (i[x] >= 0) {
          w[x] = reinterpret_cast<float4*>(a.w)[i[x]];
          g[x] = reinterpret_cast<float4*>(a.g)[i[x]];
          if (nslots > 0) s0[x] = reinterpret_cast<float4*>(a.s0)[i[x]];
          if (nslots > 1) s1[x] = reinterpret_cast<float4*>(a.s1)[i[x]];
        }
      }
#pragma unroll
      for (int x = 0; x < UD; ++x) {
        if (i[x] < 0) continue;
        const long long e = i[x] << 2;
        float dummy = 0.f;
        upd4(a, lr_t, w[x], g[x], s0[x], s1[x], false, dummy);  // fm_bias and the MLP are not regularised (PS:244-246)
        reinterpret_cast<float4*>(a.w)[i[x]] = w[x];
        reinterpret_cast<float4*>(a.g)[i[x]] = z4;
        if (nslots > 0) reinterpret_cast<float4*>(a.s0)[i[x]] = s0[x];
        if (nslots > 1) reinterpret_cast<float4*>(a.s1)[i[x]] = s1[x];
        if (e >= a.shadow_begin && e < a.shadow_end) {
          uint2 pk;
          pk.x = pack_bf16x2(w[x].x, w[x].y);
          pk.y = pack_bf16x2(w[x].z, w[x].w);
          *reinterpret_cast<uint2*>(a.shadow + (e - a.shadow_begin)) = pk;
        }
      }
    }
  }

  if ((int)blockIdx.x < dense_ctas) trace_max(TR_OPT_DENSE, 2);
  if (a.do_sparse && (int)blockIdx.x >= dense_ctas) {
    // one thread per (touched row, 16-byte chunk), straight off the row list: every row of the step is in flight at once
    const long long items = (long long)__ldcg(a.row_count) * G;
    const long long stride = (long long)((int)gridDim.x - dense_ctas) * kSplitThreads;
    for (long long it = (long long)((int)blockIdx.x - dense_ctas) * kSplitThreads + threadIdx.x; it < items; it += stride) {
      const int r = __ldcg(a.rows + it / G);
      const int c = (int)(it % G);
      const long long i = (a.fmv_off >> 2) + (long long)r * G + c;
      float4 w = reinterpret_cast<float4*>(a.w)[i];
      const float4 g = reinterpret_cast<float4*>(a.g)[i];
      float4 s0 = nslots > 0 ? reinterpret_cast<float4*>(a.s0)[i] : z4;
      float4 s1 = nslots > 1 ? reinterpret_cast<float4*>(a.s1)[i] : z4;
      float w1 = 0.f, g1 = 0.f, s01 = 0.f, s11 = 0.f;
      const long long e1 = a.fmw_off + r;
      if (c == 0) {  // the row's first-order weight rides with chunk 0
        w1 = a.w[e1];
        g1 = a.g[e1] * a.grad_scale;
        if (nslots > 0) s01 = a.s0[e1];
        if (nslots > 1) s11 = a.s1[e1];
      }
      upd4(a, lr_t, w, g, s0, s1, decay, reg);
      reinterpret_cast<float4*>(a.w)[i] = w;
      reinterpret_cast<float4*>(a.g)[i] = z4;
      if (nslots > 0) reinterpret_cast<float4*>(a.s0)[i] = s0;
      if (nslots > 1) reinterpret_cast<float4*>(a.s1)[i] = s1;
      if (c == 0) {
        if (decay) { reg += w1 * w1; g1 += a.l2 * w1; }
        opt_update(a.h, lr_t, g1, w1, s01, s11);
        a.w[e1] = w1;
        a.g[e1] = 0.f;
        if (nslots > 0) a.s0[e1] = s01;
        if (nslots > 1) a.s1[e1] = s11;
      }
    }
  }

  if ((int)blockIdx.x >= dense_ctas) trace_max(TR_OPT_ROWS, 2);
  const float tot = block_sum_to_thread0(reg, s_red);
  if (threadIdx.x == 0) {
    if (tot != 0.f) atomicAdd(a.reg_acc, tot);
    trace_end(TR_OPT);
    if (!a.finalize) return;
    __threadfence();
    const unsigned int done = atomicAdd(a.done_counter, 1u);
    if (done == gridDim.x - 1) {  // last CTA: every CTA (and the sweep, joined before this kernel) has contributed
      __threadfence();
      const float reg_tot = atomicAdd(a.reg_acc, 0.f);
      *a.reg_loss = 0.5f * a.l2 * reg_tot;
      *a.reg_acc = 0.f;
      if (a.world > 1 && a.loss_parts) {  // global log-loss sum, same order on every rank
        float s = 0.f;
        for (int r = 0; r < a.world; ++r) s += __ldcg(a.loss_parts + r);
        *a.loss_sum = s;
      }
      *a.step_ptr = t;
      *a.done_counter = 0u;
      if (a.pub_dst) {
        __threadfence();
        float* rec = a.pub_dst + ((t - 1ull) % (unsigned long long)a.pub_ring) * (unsigned long long)a.pub_n;
        for (int i = 0; i < a.pub_n; ++i) rec[i] = __ldcg(a.pub_src + i);
      }
      trace_max(TR_OPT_FINAL, 2);
    }
  }
}

template <int G>
static int launch_sweep(const SplitOptArgs& a, int blocks, cudaStream_t st) {
  dfm_prefer_max_shared(reinterpret_cast<const void*>(sweep_untouched_kernel<G>));
  dfm_launch_no_window() = true;
  const cudaError_t e = launch_k(sweep_untouched_kernel<G>, dim3((unsigned)blocks), dim3(kSplitThreads), 0, st, a);
  dfm_launch_no_window() = false;
  DFM_CHECK(e);
  return DFM_OK;
}
template <int G>
static int launch_tail(const SplitOptArgs& a, int blocks, cudaStream_t st) {
  DFM_CHECK(launch_k(tail_opt_kernel<G>, dim3((unsigned)blocks), dim3(kSplitThreads), 0, st, a));
  return DFM_OK;
}

}  // namespace dfm

using namespace dfm;

#define DFM_SPLIT_DISPATCH(K, CALL)                                                               \
  switch (K) {                                                                                    \
    case 4: return CALL(1);                                                                       \
    case 8: return CALL(2);                                                                       \
    case 16: return CALL(4);                                                                      \
    case 32: return CALL(8);                                                                      \
    case 64: return CALL(16);                                                                     \
    case 128: return CALL(32);                                                                    \
    default:                                                                                      \
      dfm_set_error("split optimizer: embedding_size %d unsupported (need 4,8,16,32,64,128)", K); \
      return DFM_ERR_UNSUPPORTED;                                                                 \
  }

extern "C" {

int dfm_bitmap_words(int N) { return (N + 31) / 32; }

// bitmap[id >> 5] |= 1 << (id & 31) for every in-range id of the batch. bitmap: ceil(N/32) words.
int dfm_touched_bitmap(const int* ids, int B, int F, int N, unsigned int* bitmap, void* stream) {
  DFM_REQUIRE(B > 0 && F > 0 && N > 0 && ids && bitmap, "touched_bitmap: empty problem or null pointer");
  const long long threads = (long long)B * F;
  dfm_prefer_max_shared(reinterpret_cast<const void*>(touched_bitmap_kernel));
  DFM_CHECK(launch_k(touched_bitmap_kernel, dim3((unsigned)ceil_div_ll(threads, 256)), dim3(256), 0,
                     (cudaStream_t)stream, ids, B, F, N, bitmap));
  return DFM_OK;
}

static int fill_common(SplitOptArgs& a, float* w, float* g, float* s0, float* s1, int opt, float lr, float beta1,
                       float beta2, float eps, float momentum, float grad_scale, float l2, long long fmw_off,
                       long long fmv_off, int N, int K, unsigned long long* step_ptr, float* reg_acc) {
  DFM_REQUIRE(opt >= 0 && opt <= 4, "split optimizer: unknown optimizer id %d", opt);
  DFM_REQUIRE(fmw_off % 4 == 0 && fmv_off % 4 == 0 && N > 0, "split optimizer: table offsets must be multiples of 4");
  a = SplitOptArgs();
  a.w = w; a.g = g; a.s0 = s0; a.s1 = s1;
  a.h.opt = opt; a.h.lr = lr; a.h.beta1 = beta1; a.h.beta2 = beta2; a.h.eps = eps; a.h.momentum = momentum;
  a.grad_scale = grad_scale; a.l2 = l2; a.fmw_off = fmw_off; a.fmv_off = fmv_off; a.N = N; a.G = K / 4;
  a.nwords = (N + 31) / 32; a.step_ptr = step_ptr; a.reg_acc = reg_acc;
  return DFM_OK;
}

// bm_union = OR of `world` bitmaps (bm_in: host array of device pointers, each ceil(N/32) words; bm_in[0] is zeroed
// afterwards when clear_inputs != 0), rows[0 .. *row_count) = ascending ids of the rows set in it. One CTA.
int dfm_union_rowlist_prefix(const void* const* bm_in, int world, int clear_inputs, int N, unsigned int* bm_union, int* rows,
                             int* row_count, int* word_prefix, void* stream);

int dfm_union_rowlist(const void* const* bm_in, int world, int clear_inputs, int N, unsigned int* bm_union, int* rows,
                      int* row_count, void* stream) {
  return dfm_union_rowlist_prefix(bm_in, world, clear_inputs, N, bm_union, rows, row_count, nullptr, stream);
}

// Same, plus word_prefix[w] = number of listed rows below row 32 w, so the position of row r in the list is
// word_prefix[r >> 5] + popc(bm_union[r >> 5] & ((1 << (r & 31)) - 1)) (dfm_ps_remap_ids uses it).
int dfm_union_rowlist_prefix(const void* const* bm_in, int world, int clear_inputs, int N, unsigned int* bm_union, int* rows,
                             int* row_count, int* word_prefix, void* stream) {
  DFM_REQUIRE(world >= 1 && world <= kSplitMaxRanks, "union_rowlist: world %d not in [1,%d]", world, kSplitMaxRanks);
  DFM_REQUIRE(N > 0 && bm_union && rows && row_count, "union_rowlist: empty problem or null pointer");
  SplitOptArgs a = SplitOptArgs();
  for (int r = 0; r < kSplitMaxRanks; ++r) a.bm_in[r] = (const uint32_t*)bm_in[r < world ? r : 0];
  a.world = world; a.clear_inputs = clear_inputs; a.N = N; a.nwords = (N + 31) / 32;
  a.bm_union = bm_union; a.rows = rows; a.row_count = row_count; a.word_prefix = word_prefix;
  dfm_prefer_max_shared(reinterpret_cast<const void*>(union_rowlist_kernel));
  DFM_CHECK(launch_k(union_rowlist_kernel, dim3(1), dim3(kRowlistThreads), 0, (cudaStream_t)stream, a));
  return DFM_OK;
}

// Sweep of the rows whose bit is clear in bm_union; adds sum(w^2) of the swept rows to *reg_acc (when l2 != 0).
int dfm_sweep_untouched(float* w, float* s0, float* s1, int opt, float lr, float beta1, float beta2, float eps,
                        float momentum, float l2, long long fmw_off, long long fmv_off, int N, int K,
                        const unsigned int* bm_union, unsigned long long* step_ptr, float* reg_acc, int grid_blocks,
                        void* stream) {
  SplitOptArgs a;
  int rc = fill_common(a, w, nullptr, s0, s1, opt, lr, beta1, beta2, eps, momentum, 1.f, l2, fmw_off, fmv_off, N, K,
                       step_ptr, reg_acc);
  if (rc) return rc;
  a.bm_union = const_cast<uint32_t*>(bm_union);
  int blocks = grid_blocks > 0 ? grid_blocks : ceil_div(a.nwords, kSplitThreads / 32);
  if (blocks > 148 * 4) blocks = 148 * 4;
  cudaStream_t st = (cudaStream_t)stream;
#define CALL(G) launch_sweep<G>(a, blocks, st)
  DFM_SPLIT_DISPATCH(K, CALL)
#undef CALL
}

// Tail: dense ranges d0 / d1 (elements, multiples of 4; either may be empty) + the rows rows[0 .. *row_count); zeroes the
// gradients it consumes. finalize != 0: the last CTA writes *reg_loss = 0.5 * l2 * *reg_acc (and resets reg_acc), replaces
// *loss_sum by the sum of loss_parts[0..world) when world > 1, advances *step_ptr and stores pub_n floats from pub_src into
// record (step - 1) % pub_ring_len of pub_ring (pinned host memory; may be null).
int dfm_tail_opt(float* w, float* g, float* s0, float* s1, int opt, float lr, float beta1, float beta2, float eps,
                 float momentum, float grad_scale, float l2, long long fmw_off, long long fmv_off, int N, int K,
                 long long d0_begin, long long d0_end, long long d1_begin, long long d1_end, long long shadow_begin,
                 long long shadow_end, void* shadow, const int* rows, const int* row_count, unsigned long long* step_ptr,
                 float* reg_acc, float* reg_loss, unsigned int* done_counter, float* loss_sum, const float* loss_parts,
                 int world, const float* pub_src, int pub_n, float* pub_ring, int pub_ring_len, int do_dense,
                 int do_sparse, int finalize, void* stream) {
  DFM_REQUIRE(d0_begin % 4 == 0 && d0_end % 4 == 0 && d1_begin % 4 == 0 && d1_end % 4 == 0 && d0_end >= d0_begin &&
                  d1_end >= d1_begin, "tail_opt: dense ranges must be ordered multiples of 4");
  DFM_REQUIRE(shadow_begin % 4 == 0 && shadow_end % 4 == 0, "tail_opt: shadow range must be multiples of 4");
  DFM_REQUIRE((pub_ring == nullptr) || (pub_src != nullptr && pub_n >= 1 && pub_n <= 16 && pub_ring_len >= 1),
              "tail_opt: publication needs a source, 1..16 floats per record and a ring length >= 1");
  SplitOptArgs a;
  int rc = fill_common(a, w, g, s0, s1, opt, lr, beta1, beta2, eps, momentum, grad_scale, l2, fmw_off, fmv_off, N, K,
                       step_ptr, reg_acc);
  if (rc) return rc;
  a.d0_begin = d0_begin; a.d0_end = d0_end; a.d1_begin = d1_begin; a.d1_end = d1_end;
  a.shadow_begin = shadow_begin; a.shadow_end = shadow_end; a.shadow = (__nv_bfloat16*)shadow;
  a.rows = const_cast<int*>(rows); a.row_count = const_cast<int*>(row_count); a.reg_loss = reg_loss;
  a.done_counter = done_counter;
  a.loss_sum = loss_sum; a.loss_parts = loss_parts; a.world = world;
  a.pub_src = pub_src; a.pub_dst = pub_ring; a.pub_n = pub_n; a.pub_ring = pub_ring_len;
  a.do_dense = do_dense; a.do_sparse = do_sparse; a.finalize = finalize;
  // dense part: 2 CTAs per SM, two float4 x 4 arrays in flight per thread; row part: one warp per bitmap word
  long long dense_blocks = 0, row_blocks = 0;
  if (do_dense) {
    dense_blocks = ceil_div_ll(((d0_end - d0_begin) + (d1_end - d1_begin)) / 4, kSplitThreads * 2);
    if (dense_blocks > 148 * 2) dense_blocks = 148 * 2;
    if (dense_blocks < 1) dense_blocks = 1;
  }
  if (do_sparse) {  // the row count is only known on the device: enough CTAs for one (row, chunk) item per thread up to
    row_blocks = ceil_div_ll((long long)N * a.G, kSplitThreads);  // 4 CTAs per SM, grid-stride beyond that
    if (row_blocks > 148 * 4) row_blocks = 148 * 4;
  }
  a.dense_ctas = (int)dense_blocks;
  long long blocks = dense_blocks + row_blocks;
  if (blocks < 1) blocks = 1;
  cudaStream_t st = (cudaStream_t)stream;
#define CALL(G) launch_tail<G>(a, (int)blocks, st)
  DFM_SPLIT_DISPATCH(K, CALL)
#undef CALL
}

}  // extern "C"
