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

}  // extern "C"
