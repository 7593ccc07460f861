// Parameter-server pull as a zero-copy gather: the worker's GPU reads the rows its batch
// touches STRAIGHT out of the host-resident, page-locked parameter store over PCIe/C2C — no host-side gather, no
// staging buffer, no separate H2D copy. (Reference: TF's gRPC RecvTensor pulls of gathered rows + the whole fm_v.)
//   out[i, :] = table[ids[i], :]   table = device-mapped host pointer, ids = unique row ids of the batch (device)
#include "common.cuh"

namespace dfm {

template <int G>
__global__ void __launch_bounds__(256)
ps_gather_rows_kernel(const float* __restrict__ table, const long long* __restrict__ ids, int n,
                      float* __restrict__ out) {
  constexpr int K = 4 * G;
  const long long gtid = (long long)blockIdx.x * 256 + threadIdx.x;
  const long long row = gtid / G;
  const int g = (int)(gtid % G);
  if (row >= n) return;
  const long long id = ids[row];
  // host memory: bypass L1/L2 allocation, 128-bit loads
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(table + id * K + 4 * g));
  reinterpret_cast<float4*>(out + row * K)[g] = v;
}

__global__ void ps_gather_scalar_kernel(const float* __restrict__ table, const long long* __restrict__ ids, int n,
                                        float* __restrict__ out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) out[i] = table[ids[i]];
}

// ---- device-driven pull / push for the pipelined worker: the unique rows of a batch are a device-resident list with a
// device-resident count (touched_bitmap + union_rowlist), so the host never waits for a size (torch.unique would).

// inv[p] = position of ids[p] in the ascending row list; ids outside [0, N) map to `oob` (an id the compact table rejects)
__global__ void __launch_bounds__(256) ps_remap_ids_kernel(const int* __restrict__ ids, long long n, int N,
                                                           const uint32_t* __restrict__ bm,
                                                           const int* __restrict__ word_prefix, int oob,
                                                           int* __restrict__ inv) {
  const long long p = (long long)blockIdx.x * 256 + threadIdx.x;
  if (p >= n) return;
  const int id = ids[p];
  int r = oob;
  if ((unsigned)id < (unsigned)N) {
    const int w = id >> 5;
    r = __ldg(word_prefix + w) + __popc(__ldg(bm + w) & ((1u << (id & 31)) - 1u));
  }
  inv[p] = r;
}

// out[i, :] = table[rows[i], :] and out_w[i] = table_w[rows[i]] for i < *count (rows past *count are left alone)
template <int G>
__global__ void __launch_bounds__(256)
ps_pull_list_kernel(const float* __restrict__ table_v, const float* __restrict__ table_w, const int* __restrict__ rows,
                    const int* __restrict__ count, float* __restrict__ out_v, float* __restrict__ out_w) {
  constexpr int K = 4 * G;
  const long long items = (long long)__ldg(count) * G;
  for (long long it = (long long)blockIdx.x * 256 + threadIdx.x; it < items; it += (long long)gridDim.x * 256) {
    const long long i = it / G;
    const int g = (int)(it % G);
    const long long id = __ldg(rows + i);
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(table_v + id * K + 4 * g));
    reinterpret_cast<float4*>(out_v + i * K)[g] = v;
    if (g == 0) out_w[i] = table_w[id];
  }
}

// push: host_v[i, :] = g_v[i, :], host_w[i] = g_w[i], host_rows[i] = rows[i] for i < *count, host_count[0] = *count, and
// the device gradient rows are zeroed for the next step. host_* are page-locked host buffers (device-addressable).
template <int G>
__global__ void __launch_bounds__(256)
ps_push_list_kernel(float* __restrict__ g_v, float* __restrict__ g_w, const int* __restrict__ rows,
                    const int* __restrict__ count, float* __restrict__ host_v, float* __restrict__ host_w,
                    int* __restrict__ host_rows, int* __restrict__ host_count) {
  constexpr int K = 4 * G;
  const int n = __ldg(count);
  const long long items = (long long)n * G;
  if (blockIdx.x == 0 && threadIdx.x == 0) host_count[0] = n;
  for (long long it = (long long)blockIdx.x * 256 + threadIdx.x; it < items; it += (long long)gridDim.x * 256) {
    const long long i = it / G;
    const int g = (int)(it % G);
    float4* src = reinterpret_cast<float4*>(g_v + i * K) + g;
    reinterpret_cast<float4*>(host_v + i * K)[g] = *src;
    *src = make_float4(0.f, 0.f, 0.f, 0.f);
    if (g == 0) {
      host_w[i] = g_w[i];
      g_w[i] = 0.f;
      host_rows[i] = __ldg(rows + i);
    }
  }
}

}  // namespace dfm

using namespace dfm;

extern "C" {

// table_dev: device-visible address of the host store region (see dfm_host_register); K in {1,4,8,...,128}.
int dfm_ps_gather(const float* table_dev, const long long* ids, int n, int K, float* out, void* stream) {
  if (n <= 0) return DFM_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (K == 1) {
    ps_gather_scalar_kernel<<<ceil_div(n, 256), 256, 0, st>>>(table_dev, ids, n, out);
  } else {
#define CALL(G) ps_gather_rows_kernel<G><<<(unsigned)ceil_div_ll((long long)n * G, 256), 256, 0, st>>>(table_dev, ids, n, out)
    switch (K) {
      case 4: CALL(1); break; case 8: CALL(2); break; case 16: CALL(4); break; case 32: CALL(8); break;
      case 64: CALL(16); break; case 128: CALL(32); break;
      default: dfm_set_error("ps_gather: embedding_size %d unsupported", K); return DFM_ERR_UNSUPPORTED;
    }
#undef CALL
  }
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

int dfm_ps_remap_ids(const int* ids, long long n, int N, const unsigned int* bm_union, const int* word_prefix, int oob,
                     int* inv, void* stream) {
  DFM_REQUIRE(n > 0 && N > 0 && ids && bm_union && word_prefix && inv, "ps_remap_ids: empty problem or null pointer");
  ps_remap_ids_kernel<<<(unsigned)ceil_div_ll(n, 256), 256, 0, (cudaStream_t)stream>>>(ids, n, N, bm_union, word_prefix, oob,
                                                                                        inv);
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

#define DFM_PS_DISPATCH(K, CALL)                                                                   \
  switch (K) {                                                                                     \
    case 4: CALL(1); break; case 8: CALL(2); break; case 16: CALL(4); break; case 32: CALL(8); break; \
    case 64: CALL(16); break; case 128: CALL(32); break;                                            \
    default: dfm_set_error("ps: embedding_size %d unsupported", K); return DFM_ERR_UNSUPPORTED;     \
  }

// Pull of the listed rows (rows[0 .. *count), both device-resident) out of the mapped host store: fm_v rows and fm_w scalars.
int dfm_ps_pull_list(const float* table_v_dev, const float* table_w_dev, const int* rows, const int* count, int max_rows,
                     int K, float* out_v, float* out_w, void* stream) {
  DFM_REQUIRE(max_rows > 0 && rows && count, "ps_pull_list: empty problem or null pointer");
  long long blocks = ceil_div_ll((long long)max_rows * (K / 4), 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
#define CALL(G) ps_pull_list_kernel<G><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(table_v_dev, table_w_dev, rows, count, out_v, out_w)
  DFM_PS_DISPATCH(K, CALL)
#undef CALL
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

// Push of the listed gradient rows into page-locked host buffers (and zeroing of the device copies).
int dfm_ps_push_list(float* g_v, float* g_w, const int* rows, const int* count, int max_rows, int K, float* host_v,
                     float* host_w, int* host_rows, int* host_count, void* stream) {
  DFM_REQUIRE(max_rows > 0 && rows && count && host_v && host_w && host_rows && host_count,
              "ps_push_list: empty problem or null pointer");
  long long blocks = ceil_div_ll((long long)max_rows * (K / 4), 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
#define CALL(G) ps_push_list_kernel<G><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(g_v, g_w, rows, count, host_v, host_w, host_rows, host_count)
  DFM_PS_DISPATCH(K, CALL)
#undef CALL
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

}  // extern "C"
