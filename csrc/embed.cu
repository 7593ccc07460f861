// Fused embedding / FM kernels (replicated tables).
//
// emb_fm_fwd  (reference ops K1+K2+K3: `embedding_lookup` x2, multiply, reduce_sum, square, subtract;
//              PS:177-187 / HVD:169-179, plus the reshape to the deep input PS:203):
//     one pass over ids/vals gathers fm_w and fm_v rows with 128-bit loads, scales by the feature value,
//     writes the deep input X0[b, f*K+k] as bf16, keeps S = sum_f e and sum_f e^2 in registers and emits
//         yfm[b] = sum_f fm_w[id]*val + 0.5 * sum_k (S_k^2 - sum_f e_k^2)
//     and S (fp32, needed by the backward pass).
// emb_bwd     (autograd of the above, reference op K11):
//     dE[b,f,:] = dX0[b,f,:] + dy[b] * (S[b,:] - e[b,f,:]);   g_v[id,:] += val * dE;   g_w[id] += dy[b] * val
//     accumulated with red.global.add.v4.f32 into the dense fp32 gradient buffer (the reference's embedding
//     gradient is dense as well because of the whole-table L2 term); duplicate ids inside a warp are
//     combined with match/shuffle first because same-address L2 atomics serialise.
//
// Lanes own 4 consecutive embedding columns each (K = 4*G, G | 32).
#include "common.cuh"

#include <stdlib.h>

namespace dfm {

constexpr int kEmbThreads = 256;
constexpr int kEmbUnroll = 4;

// Optional owner routing of the embedding gradients under ZeRO-1 data parallelism (parallel/dp.py): the flat gradient
// buffer is cut into `world` contiguous slices of `chunk` elements and element e is reduced by rank e / chunk. With a route,
// emb_bwd sends every row's gradient STRAIGHT to its owner's buffer (red.add on the peer-mapped pointer over NVLink) instead
// of accumulating locally, so the reduce-scatter of the 15 MB embedding gradient — 80 % of the all-reduce's first half —
// happens inside this kernel, overlapped with its own gathers, and only the ~20 K rows a batch touches cross the link
// instead of the dense buffer. world == 0: plain local accumulation.
struct EmbRoute {
  float* g[8];           // every rank's flat gradient buffer (peer-mapped)
  long long gw_off, gv_off;  // element offsets of g_w / g_v inside it
  long long chunk;       // slice length in elements (a multiple of the row length)
  int world;
};
// The route type is a template parameter of the kernel: the plain variant (NoRoute) must not carry the pointer table at
// all (a dynamically indexed kernel-parameter array is copied to a per-thread stack frame: 96 bytes of local memory per
// thread cost the single-GPU step 8 us), and the routed variant picks the peer pointer with a select chain instead of an
// indexed load.
struct NoRoute {
  static constexpr long long gw_off = 0, gv_off = 0;
};
__device__ __forceinline__ float* route_ptr(const NoRoute&, float* local, long long, long long e) { return local + e; }
__device__ __forceinline__ float* route_ptr(const EmbRoute& r, float*, long long off, long long e) {
  const long long ge = off + e;
  const int owner = (int)(ge / r.chunk);
  float* base = r.g[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) base = owner == i ? r.g[i] : base;
  return base + ge;
}

template <int G>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One warp per sample: lane = fg * G + g, g = 4-column chunk of the embedding row, fg = field group.
// Field group fg walks fields fg, fg+FG, ... so FG = 32/G row gathers of one sample are in flight per step.
template <int G>
__global__ void __launch_bounds__(kEmbThreads)
emb_fm_fwd_kernel(const int* __restrict__ ids, const float* __restrict__ vals, const float* __restrict__ fm_w,
                  const float* __restrict__ fm_v, __nv_bfloat16* __restrict__ x0, int ldx, float* __restrict__ S,
                  float* __restrict__ yfm, int B, int F, int N, float* __restrict__ zero_scalars, int n_zero) {
  constexpr int K = 4 * G;
  constexpr int FG = 32 / G;
  const int lane = threadIdx.x & 31;
  const int sample = blockIdx.x * (kEmbThreads / 32) + (threadIdx.x >> 5);
  const int g = lane % G, fg = lane / G;
  pdl_launch_dependents();
  pdl_wait(TR_EMB_FWD);
  // first kernel of a training step: also clears the step's loss accumulators (saves a memset node on the chain)
  if (blockIdx.x == 0 && threadIdx.x < n_zero) zero_scalars[threadIdx.x] = 0.f;
  if (sample >= B) return;  // warp-uniform
  const int* sid = ids + (size_t)sample * F;
  const float* sval = vals + (size_t)sample * F;

  float4 s = make_float4(0.f, 0.f, 0.f, 0.f), ss = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int f0 = fg; f0 < F; f0 += FG * kEmbUnroll) {
    int id[kEmbUnroll];
    float val[kEmbUnroll];
    float4 row[kEmbUnroll];
#pragma unroll
    for (int u = 0; u < kEmbUnroll; ++u) {
      const int f = f0 + u * FG;
      id[u] = 0; val[u] = 0.f;
      if (f < F) {
        id[u] = __ldg(sid + f);
        val[u] = __ldg(sval + f);
        if ((unsigned)id[u] >= (unsigned)N) { id[u] = 0; val[u] = 0.f; }  // out-of-range ids contribute nothing
      }
    }
#pragma unroll
    for (int u = 0; u < kEmbUnroll; ++u)
      row[u] = __ldg(reinterpret_cast<const float4*>(fm_v + (size_t)id[u] * K) + g);
#pragma unroll
    for (int u = 0; u < kEmbUnroll; ++u) {
      const int f = f0 + u * FG;
      if (f < F) {
        const float4 e = make_float4(row[u].x * val[u], row[u].y * val[u], row[u].z * val[u], row[u].w * val[u]);
        s.x += e.x; s.y += e.y; s.z += e.z; s.w += e.w;
        ss.x += e.x * e.x; ss.y += e.y * e.y; ss.z += e.z * e.z; ss.w += e.w * e.w;
        uint2 pk;
        pk.x = pack_bf16x2(e.x, e.y);
        pk.y = pack_bf16x2(e.z, e.w);
        *reinterpret_cast<uint2*>(x0 + (size_t)sample * ldx + (size_t)f * K + 4 * g) = pk;
      }
    }
  }
  // reduce over field groups (lanes with equal g): xor offsets G, 2G, ..., 16
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
  for (int f = lane; f < F; f += 32) {  // first-order term
    const int id = __ldg(sid + f);
    if ((unsigned)id < (unsigned)N) acc += __ldg(fm_w + id) * __ldg(sval + f);
  }
  acc = warp_sum(acc);
  if (lane == 0) yfm[sample] = acc;
  trace_end(TR_EMB_FWD);
}

// One LANE per (field, sample) pair, pairs in field-major order, so a warp holds 32 samples of the SAME field.
// Criteo-shaped data gives every sample the same id in the 13 numeric fields and zipf-hot ids in the categorical
// ones; same-address L2 atomics serialise (~10 ns each), so lanes that share an id are combined first:
//   all 32 lanes equal  -> butterfly reduce, one vector atomic per 16-byte chunk
//   partial duplicates  -> __match_any_sync groups, the group leader gathers its members with shuffles
//   unique id           -> direct vector atomic
template <int G, typename RouteT>
__global__ void __launch_bounds__(kEmbThreads)
emb_bwd_kernel(const int* __restrict__ ids, const float* __restrict__ vals, const float* __restrict__ fm_v,
               const float* __restrict__ S, const float* __restrict__ dy, const __nv_bfloat16* __restrict__ dx0,
               int ldx, float* __restrict__ g_w, float* __restrict__ g_v, int B, int F, int N, const RouteT route) {
  constexpr int K = 4 * G;
  const int lane = threadIdx.x & 31;
  const long long p = (long long)blockIdx.x * kEmbThreads + threadIdx.x;
  const bool active = p < (long long)B * F;
  const int f = active ? (int)(p / B) : 0;
  const int b = active ? (int)(p % B) : 0;
  int id = 0;
  float v = 0.f;
  pdl_wait(TR_EMB_BWD);
  if (active) {
    id = __ldg(ids + (size_t)b * F + f);
    v = __ldg(vals + (size_t)b * F + f);
  }
  const bool valid = active && v != 0.f && (unsigned)id < (unsigned)N;  // padding / out-of-range: zero gradient
  const unsigned key = valid ? (unsigned)id : (0x80000000u | (unsigned)lane);
  const unsigned mask = __match_any_sync(0xffffffffu, key);
  const int leader = __ffs(mask) - 1;
  const int cnt = __popc(mask);
  const float d = valid ? __ldg(dy + b) : 0.f;
  const float4* rowp = reinterpret_cast<const float4*>(fm_v + (size_t)(valid ? id : 0) * K);
  const float4* sp = reinterpret_cast<const float4*>(S + (size_t)b * K);
  const uint2* dxp = reinterpret_cast<const uint2*>(dx0 + (size_t)b * ldx + (size_t)f * K);

  // the row's G 16-byte chunks are split over gridDim.y CTAs (shorter serial chain per thread, 4x the threads)
  const int g_per = (G + (int)gridDim.y - 1) / (int)gridDim.y;
  const int g_begin = (int)blockIdx.y * g_per, g_end = min(G, g_begin + g_per);
#pragma unroll 2
  for (int g = g_begin; g < g_end; ++g) {
    float4 gr = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) {
      const float4 row = __ldg(rowp + g);
      const float4 s = __ldg(sp + g);
      const uint2 dx = __ldg(dxp + g);
      gr.x = v * (bf16_lo(dx.x) + d * (s.x - row.x * v));
      gr.y = v * (bf16_hi(dx.x) + d * (s.y - row.y * v));
      gr.z = v * (bf16_lo(dx.y) + d * (s.z - row.z * v));
      gr.w = v * (bf16_hi(dx.y) + d * (s.w - row.w * v));
    }
    if (mask == 0xffffffffu) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        gr.x += __shfl_xor_sync(0xffffffffu, gr.x, o); gr.y += __shfl_xor_sync(0xffffffffu, gr.y, o);
        gr.z += __shfl_xor_sync(0xffffffffu, gr.z, o); gr.w += __shfl_xor_sync(0xffffffffu, gr.w, o);
      }
      if (lane == 0 && valid) red_add_f32x4(route_ptr(route, g_v, route.gv_off, (long long)id * K + 4 * g), gr.x, gr.y, gr.z, gr.w);
    } else if (cnt == 1) {
      if (valid) red_add_f32x4(route_ptr(route, g_v, route.gv_off, (long long)id * K + 4 * g), gr.x, gr.y, gr.z, gr.w);
    } else {
      float4 acc = gr;
      unsigned rem = mask & ~(1u << leader);
      while (rem) {
        const int src = __ffs(rem) - 1;
        rem &= rem - 1;
        const float x = __shfl_sync(mask, gr.x, src), y = __shfl_sync(mask, gr.y, src);
        const float z = __shfl_sync(mask, gr.z, src), w = __shfl_sync(mask, gr.w, src);
        acc.x += x; acc.y += y; acc.z += z; acc.w += w;
      }
      if (lane == leader && valid)
        red_add_f32x4(route_ptr(route, g_v, route.gv_off, (long long)id * K + 4 * g), acc.x, acc.y, acc.z, acc.w);
    }
  }
  // first-order weights: same aggregation on the scalar dy*val (first chunk-slice only)
  pdl_launch_dependents();  // late: the optimizer's CTAs must not sit resident beside this kernel and the wgrad branch
  trace_end(TR_EMB_BWD);
  if (blockIdx.y != 0) return;  // CTA-uniform
  float gw = valid ? d * v : 0.f;
  if (mask == 0xffffffffu) {
    gw = warp_sum(gw);
    if (lane == 0 && valid) atomicAdd(route_ptr(route, g_w, route.gw_off, id), gw);
  } else if (cnt == 1) {
    if (valid) atomicAdd(route_ptr(route, g_w, route.gw_off, id), gw);
  } else {
    float acc = gw;
    unsigned rem = mask & ~(1u << leader);
    while (rem) {
      const int src = __ffs(rem) - 1;
      rem &= rem - 1;
      acc += __shfl_sync(mask, gw, src);
    }
    if (lane == leader && valid) atomicAdd(route_ptr(route, g_w, route.gw_off, id), acc);
  }
}

template <int G>
static int launch_fwd(const int* ids, const float* vals, const float* fm_w, const float* fm_v, __nv_bfloat16* x0,
                      int ldx, float* S, float* yfm, int B, int F, int N, float* zero_scalars, int n_zero,
                      cudaStream_t st) {
  DFM_CHECK(launch_k(emb_fm_fwd_kernel<G>, dim3((unsigned)ceil_div(B, kEmbThreads / 32)), dim3(kEmbThreads), 0, st, ids,
                     vals, fm_w, fm_v, x0, ldx, S, yfm, B, F, N, zero_scalars, n_zero));
  return DFM_OK;
}

template <int G>
static int launch_bwd(const int* ids, const float* vals, const float* fm_v, const float* S, const float* dy,
                      const __nv_bfloat16* dx0, int ldx, float* g_w, float* g_v, int B, int F, int N, cudaStream_t st,
                      const EmbRoute& route) {
  const long long threads = (long long)B * F;
  const unsigned xblocks = (unsigned)ceil_div_ll(threads, kEmbThreads);
  // split the K/4 chunks over up to 4 CTAs in y while the grid stays under ~8 CTAs per SM
  unsigned ysplit = 1;
  while (ysplit < 4 && ysplit * 2 <= (unsigned)G && (unsigned long long)xblocks * ysplit * 2 <= 148ull * 8) ysplit *= 2;
  // emb_bwd runs BESIDE the weight-gradient GEMMs (98 KB of shared memory per CTA). With the default carve-out its CTAs put
  // every SM they reach first into a small-shared-memory configuration, and a GEMM CTA cannot be placed there until the SM
  // has drained (scripts/step_trace.py: wgrad_0 entered 10 us late). DFM_EMB_BWD_CARVEOUT=0 restores the default.
  static const bool big_carveout = !(getenv("DFM_EMB_BWD_CARVEOUT") && getenv("DFM_EMB_BWD_CARVEOUT")[0] == '0');
  if (big_carveout) {
    dfm_prefer_max_shared(reinterpret_cast<const void*>(emb_bwd_kernel<G, EmbRoute>));
    dfm_prefer_max_shared(reinterpret_cast<const void*>(emb_bwd_kernel<G, NoRoute>));
  }
  if (route.world > 0) {
    DFM_CHECK(launch_k(emb_bwd_kernel<G, EmbRoute>, dim3(xblocks, ysplit), dim3(kEmbThreads), 0, st, ids, vals, fm_v, S, dy,
                       dx0, ldx, g_w, g_v, B, F, N, route));
  } else {
    DFM_CHECK(launch_k(emb_bwd_kernel<G, NoRoute>, dim3(xblocks, ysplit), dim3(kEmbThreads), 0, st, ids, vals, fm_v, S, dy,
                       dx0, ldx, g_w, g_v, B, F, N, NoRoute()));
  }
  return DFM_OK;
}

}  // namespace dfm

using namespace dfm;

#define DFM_DISPATCH_G(K, CALL)                                                         \
  switch (K) {                                                                          \
    case 4: return CALL(1);                                                             \
    case 8: return CALL(2);                                                             \
    case 16: return CALL(4);                                                            \
    case 32: return CALL(8);                                                            \
    case 64: return CALL(16);                                                           \
    case 128: return CALL(32);                                                          \
    default:                                                                            \
      dfm_set_error("embedding_size %d unsupported by the fused kernels (need 4,8,16,32,64,128)", K); \
      return DFM_ERR_UNSUPPORTED;                                                       \
  }

extern "C" {

// ids [B,F] int32, vals [B,F] f32, fm_w [N] f32, fm_v [N,K] f32 (16-byte aligned rows)
// -> x0 [B, ldx>=F*K] bf16, S [B,K] f32, yfm [B] f32
// zero_scalars/n_zero (optional): up to 32 floats cleared by the kernel (the step's loss accumulators).
int dfm_emb_fm_fwd(const int* ids, const float* vals, const float* fm_w, const float* fm_v, void* x0, int ldx, float* S,
                   float* yfm, int B, int F, int K, int N, float* zero_scalars, int n_zero, void* stream) {
  DFM_REQUIRE(n_zero >= 0 && n_zero <= 32, "emb_fm_fwd: n_zero must be in [0,32]");
  DFM_REQUIRE(B > 0 && F > 0 && N > 0, "emb_fm_fwd: empty problem");
  DFM_REQUIRE(ldx >= F * K && ldx % 4 == 0, "emb_fm_fwd: ldx must be >= F*K and a multiple of 4");
#define CALL(G) \
  launch_fwd<G>(ids, vals, fm_w, fm_v, (__nv_bfloat16*)x0, ldx, S, yfm, B, F, N, zero_scalars, n_zero, (cudaStream_t)stream)
  DFM_DISPATCH_G(K, CALL)
#undef CALL
}

// dy [B] f32, dx0 [B, ldx] bf16 -> g_w [N] f32 +=, g_v [N,K] f32 +=
int dfm_emb_bwd(const int* ids, const float* vals, const float* fm_v, const float* S, const float* dy, const void* dx0,
                int ldx, float* g_w, float* g_v, int B, int F, int K, int N, void* stream) {
  DFM_REQUIRE(B > 0 && F > 0 && N > 0, "emb_bwd: empty problem");
  EmbRoute route = {};
#define CALL(G) \
  launch_bwd<G>(ids, vals, fm_v, S, dy, (const __nv_bfloat16*)dx0, ldx, g_w, g_v, B, F, N, (cudaStream_t)stream, route)
  DFM_DISPATCH_G(K, CALL)
#undef CALL
}

// emb_bwd with owner routing: the row gradients go to the flat gradient buffer of the rank that reduces them (g_ptrs: host
// array of `world` device pointers to every rank's flat buffer; gw_off / gv_off: element offsets of fm_w / fm_v in it;
// element e belongs to rank e / chunk). No local accumulation: a rank's buffer ends up holding the SUM over ranks for the
// rows of its own slice and nothing for the others.
int dfm_emb_bwd_routed(const int* ids, const float* vals, const float* fm_v, const float* S, const float* dy,
                       const void* dx0, int ldx, void* const* g_ptrs, long long gw_off, long long gv_off, long long chunk,
                       int world, int B, int F, int K, int N, void* stream) {
  DFM_REQUIRE(B > 0 && F > 0 && N > 0, "emb_bwd_routed: empty problem");
  DFM_REQUIRE(world >= 1 && world <= 8 && chunk > 0 && chunk % K == 0 && gv_off % K == 0,
              "emb_bwd_routed: world in [1,8], chunk and gv_off multiples of the row length");
  EmbRoute route = {};
  for (int r = 0; r < 8; ++r) route.g[r] = (float*)g_ptrs[r < world ? r : 0];
  route.gw_off = gw_off; route.gv_off = gv_off; route.chunk = chunk; route.world = world;
#define CALL(G) \
  launch_bwd<G>(ids, vals, fm_v, S, dy, (const __nv_bfloat16*)dx0, ldx, nullptr, nullptr, B, F, N, (cudaStream_t)stream, route)
  DFM_DISPATCH_G(K, CALL)
#undef CALL
}

}  // extern "C"
