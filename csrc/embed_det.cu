// Deterministic embedding backward (SURVEY 7.3 hard part 8: "deterministic duplicate-id reduction (sorted segment-sum vs
// atomics)"; reference op K11: TF segment-sums duplicate ids before it applies them). Optional twin of emb_bwd
// (csrc/embed.cu), which accumulates with fp32 atomics in arrival order and is therefore not bit-reproducible run to run.
//
//   emb_sort_ids      (side branch, at step start — needs only the ids) keys = row id (invalid / zero-valued lookups get
//                     the sentinel N and sort to the end), values = lookup index b*F+f; cub::DeviceRadixSort (stable)
//   emb_bwd_det       one CTA per 128 consecutive SORTED lookups: every lookup's gradient chunk goes to shared memory, the
//                     first lookup of each run of equal ids sums its run front to back (fixed order), runs that lie
//                     strictly inside the CTA are written to the gradient buffer with plain stores, the (at most two) runs
//                     touching the CTA's ends go to a small record array
//   emb_bwd_det_fix   the records are in row order; the first record of each row sums its chain front to back and stores it
//
// Every touched row receives exactly ONE plain store: no atomics, no dependence on scheduling, and the gradient rows need no
// zeroing beforehand. Cost: a sort of B*F pairs off the critical path plus two kernels on it.
#include "common.cuh"

#include <cub/device/device_radix_sort.cuh>

namespace dfm {

constexpr int kDetThreads = 256;
constexpr int kDetT = 128;  // sorted lookups per CTA

struct DetWs {  // carved out of the caller's workspace
  int* keys_in;
  int* idx_in;
  int* keys_out;
  int* idx_out;
  int* rec_row;    // [2 * nCTA]
  float* rec_w;    // [2 * nCTA]
  float* rec_v;    // [2 * nCTA][K]
  void* cub_tmp;
  size_t cub_bytes;
};

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

static size_t cub_temp_bytes(int n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const int*)nullptr, (int*)nullptr, (const int*)nullptr, (int*)nullptr, n, 0,
                                  32, (cudaStream_t)0);
  return bytes;
}

static size_t carve(DetWs& w, void* base, int n, int K) {
  const int ncta = ceil_div(n, kDetT);
  uint8_t* p = (uint8_t*)base;
  size_t off = 0;
  auto take = [&](size_t bytes) { void* r = p ? p + off : nullptr; off += al256(bytes); return r; };
  w.keys_in = (int*)take((size_t)n * 4);
  w.idx_in = (int*)take((size_t)n * 4);
  w.keys_out = (int*)take((size_t)n * 4);
  w.idx_out = (int*)take((size_t)n * 4);
  w.rec_row = (int*)take((size_t)2 * ncta * 4);
  w.rec_w = (float*)take((size_t)2 * ncta * 4);
  w.rec_v = (float*)take((size_t)2 * ncta * K * 4);
  w.cub_bytes = cub_temp_bytes(n);
  w.cub_tmp = take(w.cub_bytes);
  return off;
}

__global__ void __launch_bounds__(256) det_keys_kernel(const int* __restrict__ ids, const float* __restrict__ vals, int n,
                                                       int N, int* __restrict__ keys, int* __restrict__ idx) {
  pdl_launch_dependents();
  pdl_wait();
  const int p = blockIdx.x * 256 + threadIdx.x;
  if (p >= n) return;
  const int id = ids[p];
  const bool valid = (unsigned)id < (unsigned)N && vals[p] != 0.f;  // padding / out-of-range: zero gradient (as emb_bwd)
  keys[p] = valid ? id : N;
  idx[p] = p;
}

template <int G>
__global__ void __launch_bounds__(kDetThreads)
emb_bwd_det_kernel(const int* __restrict__ keys, const int* __restrict__ idx, int n, const float* __restrict__ vals,
                   const float* __restrict__ fm_v, const float* __restrict__ S, const float* __restrict__ dy,
                   const __nv_bfloat16* __restrict__ dx0, int ldx, float* __restrict__ g_w, float* __restrict__ g_v, int F,
                   int N, int* __restrict__ rec_row, float* __restrict__ rec_w, float* __restrict__ rec_v) {
  constexpr int K = 4 * G;
  constexpr int LPP = kDetThreads / G;  // lookups per pass
  __shared__ int s_key[kDetT];
  __shared__ float s_w[kDetT];
  extern __shared__ float4 s_g[];  // [kDetT][G]
  const int base = blockIdx.x * kDetT;
  const int g = threadIdx.x % G, tl = threadIdx.x / G;
  pdl_wait(TR_EMB_BWD);
  // ---- per-lookup gradients -> shared memory
  for (int t = tl; t < kDetT; t += LPP) {
    const int i = base + t;
    int key = N;
    float4 gr = make_float4(0.f, 0.f, 0.f, 0.f);
    float gw = 0.f;
    if (i < n) {
      key = __ldg(keys + i);
      if (key < N) {
        const int p = __ldg(idx + i);
        const int b = p / F, f = p - b * F;
        const float v = __ldg(vals + p);
        const float d = __ldg(dy + b);
        const float4 row = __ldg(reinterpret_cast<const float4*>(fm_v + (size_t)key * K) + g);
        const float4 s = __ldg(reinterpret_cast<const float4*>(S + (size_t)b * K) + g);
        const uint2 dx = __ldg(reinterpret_cast<const uint2*>(dx0 + (size_t)b * ldx + (size_t)f * K) + g);
        gr.x = v * (bf16_lo(dx.x) + d * (s.x - row.x * v));
        gr.y = v * (bf16_hi(dx.x) + d * (s.y - row.y * v));
        gr.z = v * (bf16_lo(dx.y) + d * (s.z - row.z * v));
        gr.w = v * (bf16_hi(dx.y) + d * (s.w - row.w * v));
        gw = d * v;
      }
    }
    s_g[t * G + g] = gr;
    if (g == 0) { s_key[t] = key; s_w[t] = gw; }
  }
  __syncthreads();
  pdl_launch_dependents();
  if (threadIdx.x < 2) rec_row[2 * blockIdx.x + threadIdx.x] = -1;
  __syncthreads();
  // ---- the first lookup of every run of equal ids sums its run, front to back
  for (int t = tl; t < kDetT; t += LPP) {
    const int key = s_key[t];
    if (key >= N || (t > 0 && s_key[t - 1] == key)) continue;
    float4 acc = s_g[t * G + g];
    float aw = s_w[t];
    int j = t + 1;
    for (; j < kDetT && s_key[j] == key; ++j) {
      const float4 x = s_g[j * G + g];
      acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
      aw += s_w[j];
    }
    const bool at_front = t == 0, at_back = j == kDetT;
    if (!at_front && !at_back) {  // the run lies strictly inside this CTA: this is the row's whole gradient
      reinterpret_cast<float4*>(g_v + (size_t)key * K)[g] = acc;
      if (g == 0) g_w[key] = aw;
    } else {  // it may continue in a neighbouring CTA: hand it to the fix-up pass
      const int r = 2 * blockIdx.x + (at_front ? 0 : 1);
      reinterpret_cast<float4*>(rec_v + (size_t)r * K)[g] = acc;
      if (g == 0) { rec_row[r] = key; rec_w[r] = aw; }
    }
  }
  trace_end(TR_EMB_BWD);
}

template <int G>
__global__ void __launch_bounds__(kDetThreads)
emb_bwd_det_fix_kernel(const int* __restrict__ rec_row, const float* __restrict__ rec_w, const float* __restrict__ rec_v,
                       int nrec, float* __restrict__ g_w, float* __restrict__ g_v) {
  constexpr int K = 4 * G;
  const int g = threadIdx.x % G;
  pdl_launch_dependents();
  pdl_wait();
  for (int r = blockIdx.x * (kDetThreads / G) + threadIdx.x / G; r < nrec; r += gridDim.x * (kDetThreads / G)) {
    const int row = rec_row[r];
    if (row < 0) continue;
    int q = r - 1;  // previous non-empty record
    while (q >= 0 && rec_row[q] < 0) --q;
    if (q >= 0 && rec_row[q] == row) continue;  // not the first record of this row's chain
    float4 acc = reinterpret_cast<const float4*>(rec_v + (size_t)r * K)[g];
    float aw = rec_w[r];
    for (int j = r + 1; j < nrec; ++j) {
      const int rj = rec_row[j];
      if (rj < 0) continue;
      if (rj != row) break;
      const float4 x = reinterpret_cast<const float4*>(rec_v + (size_t)j * K)[g];
      acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
      aw += rec_w[j];
    }
    reinterpret_cast<float4*>(g_v + (size_t)row * K)[g] = acc;
    if (g == 0) g_w[row] = aw;
  }
}

}  // namespace dfm

using namespace dfm;

extern "C" {

// Bytes of device workspace dfm_emb_sort_ids / dfm_emb_bwd_det need for a batch of B x F lookups with embedding size K.
long long dfm_emb_bwd_det_workspace_bytes(int B, int F, int K) {
  DetWs w;
  return (long long)carve(w, nullptr, B * F, K);
}

// Sort the batch's lookups by row id into the workspace (stable radix sort). Needs only ids / vals: run it on a side
// stream at step start. No allocation: cub's temporary storage is part of the workspace.
int dfm_emb_sort_ids(const int* ids, const float* vals, int B, int F, int K, int N, void* workspace,
                     long long workspace_bytes, void* stream) {
  DFM_REQUIRE(B > 0 && F > 0 && N > 0 && ids && vals && workspace, "emb_sort_ids: empty problem or null pointer");
  const int n = B * F;
  DetWs w;
  const size_t need = carve(w, workspace, n, K);
  DFM_REQUIRE((long long)need <= workspace_bytes, "emb_sort_ids: workspace too small (%lld < %lld)", workspace_bytes,
              (long long)need);
  cudaStream_t st = (cudaStream_t)stream;
  DFM_CHECK(launch_k(det_keys_kernel, dim3(ceil_div(n, 256)), dim3(256), 0, st, ids, vals, n, N, w.keys_in, w.idx_in));
  int end_bit = 1;
  while (end_bit < 32 && (1ll << end_bit) <= (long long)N) ++end_bit;  // keys are in [0, N]
  size_t tmp = w.cub_bytes;
  DFM_CHECK(cub::DeviceRadixSort::SortPairs(w.cub_tmp, tmp, w.keys_in, w.keys_out, w.idx_in, w.idx_out, n, 0, end_bit, st));
  return DFM_OK;
}

// Deterministic twin of dfm_emb_bwd over the sorted lookups left in `workspace` by dfm_emb_sort_ids for the SAME batch.
// g_w [N] and g_v [N,K] receive one plain store per touched row (no accumulation into previous contents).
int dfm_emb_bwd_det(const float* vals, const float* fm_v, const float* S, const float* dy, const void* dx0, int ldx,
                    float* g_w, float* g_v, int B, int F, int K, int N, void* workspace, long long workspace_bytes,
                    void* stream) {
  DFM_REQUIRE(B > 0 && F > 0 && N > 0 && workspace, "emb_bwd_det: empty problem or null pointer");
  const int n = B * F;
  DetWs w;
  const size_t need = carve(w, workspace, n, K);
  DFM_REQUIRE((long long)need <= workspace_bytes, "emb_bwd_det: workspace too small");
  const int ncta = ceil_div(n, kDetT);
  cudaStream_t st = (cudaStream_t)stream;
#define DET_CASE(KK)                                                                                                    \
  case KK: {                                                                                                            \
    constexpr int G = KK / 4;                                                                                           \
    const size_t smem = (size_t)kDetT * G * sizeof(float4);                                                             \
    DFM_CHECK(launch_k(emb_bwd_det_kernel<G>, dim3(ncta), dim3(kDetThreads), smem, st, (const int*)w.keys_out,          \
                       (const int*)w.idx_out, n, vals, fm_v, S, dy, (const __nv_bfloat16*)dx0, ldx, g_w, g_v, F, N,      \
                       w.rec_row, w.rec_w, w.rec_v));                                                                   \
    const int fix_blocks = ceil_div(2 * ncta, kDetThreads / G);                                                         \
    DFM_CHECK(launch_k(emb_bwd_det_fix_kernel<G>, dim3(fix_blocks), dim3(kDetThreads), 0, st, (const int*)w.rec_row,    \
                       (const float*)w.rec_w, (const float*)w.rec_v, 2 * ncta, g_w, g_v));                               \
    return DFM_OK;                                                                                                      \
  }
  switch (K) {
    DET_CASE(4) DET_CASE(8) DET_CASE(16) DET_CASE(32) DET_CASE(64)
    default:
      dfm_set_error("emb_bwd_det: embedding_size %d unsupported (need 4,8,16,32,64)", K);
      return DFM_ERR_UNSUPPORTED;
  }
#undef DET_CASE
}

}  // extern "C"
