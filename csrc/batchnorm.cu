// Batch normalisation AFTER ReLU for the optional `--batch_norm` path (reference K6: `batch_norm_layer`, PS:288-292 =
// `tf.contrib.layers.batch_norm(decay, center, scale, updates_collections=None)` applied after the fully-connected
// ReLU, PS:215-216, then dropout PS:218). Per replica, like the reference (no cross-rank statistics).
//
//   (z = relu(xW+b) is kept in fp32: BN multiplies rounding error of z by rstd, up to 31x at eps = 1e-3)
//   train:  mu, var = batch moments (biased) of z = relu(xW+b);  y = gamma * (z - mu) * rsqrt(var + eps) + beta
//           moving_mean/var <- decay * moving + (1 - decay) * batch        (in place: updates_collections=None)
//           out = dropout(y)  (same Philox stream as the GEMM epilogue)      eps = 0.001 (contrib default)
//   eval:   y = gamma * (z - moving_mean) * rsqrt(moving_var + eps) + beta
//   bwd:    dY (already multiplied by the dropout mask / keep by the producer) ->
//           dgamma = sum dY xhat, dbeta = sum dY,
//           dz_bn = gamma * rstd * (dY - mean(dY) - xhat * mean(dY xhat));   dZ = dz_bn * (z > 0)   (ReLU backward)
//
// A CLUSTER of 8 CTAs owns 32 columns: CTA r of the cluster reduces rows [r*B/8, (r+1)*B/8) (32 lanes x 8 row-lanes),
// the 8 partial column moments are exchanged through distributed shared memory (one cluster.sync), and every CTA then
// normalises its own rows, which are still in L1/L2. 13 x 8 = 104 CTAs at H = 400 instead of 13 (54 -> see profiles/).
#include <cooperative_groups.h>

#include "common.cuh"
#include "philox.cuh"

namespace cg = cooperative_groups;

namespace dfm {

constexpr int kBnThreads = 256;
constexpr int kBnCluster = 8;
constexpr float kBnEps = 1e-3f;

// Column sums of (a, b) over the CTA's row-lanes, then over the 8 CTAs of the cluster. Every thread of every CTA of the
// cluster must call it (two cluster barriers inside). Lane l of every warp returns the totals of column l.
__device__ __forceinline__ float2 bn_cluster_colsum2(float a, float b, float (*sa)[33], float (*sb)[33], float2* part) {
  cg::cluster_group cluster = cg::this_cluster();
  const int lane = threadIdx.x & 31, ty = threadIdx.x >> 5;
  sa[ty][lane] = a;
  sb[ty][lane] = b;
  __syncthreads();
  if (ty == 0) {
    float ra = 0.f, rb = 0.f;
#pragma unroll
    for (int k = 0; k < kBnThreads / 32; ++k) { ra += sa[k][lane]; rb += sb[k][lane]; }
    part[lane] = make_float2(ra, rb);
  }
  cluster.sync();  // partials of all 8 CTAs are visible cluster-wide
  float ra = 0.f, rb = 0.f;
#pragma unroll
  for (int r = 0; r < kBnCluster; ++r) {
    const float2 t = *cluster.map_shared_rank(part + lane, r);  // fixed order: identical totals in every CTA
    ra += t.x; rb += t.y;
  }
  cluster.sync();  // nobody leaves (and frees its shared memory) while a peer may still read it
  return make_float2(ra, rb);
}

__global__ void __cluster_dims__(1, kBnCluster, 1) __launch_bounds__(kBnThreads)
bn_fwd_kernel(const float* __restrict__ z, int ldz, __nv_bfloat16* __restrict__ out, int ldo, int B, int H,
              const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ mov_mean,
              float* __restrict__ mov_var, float* __restrict__ save_mean, float* __restrict__ save_rstd, float decay,
              int train, DropoutCfg drop, const unsigned long long* __restrict__ step_ptr) {
  __shared__ float sa[kBnThreads / 32][33], sb[kBnThreads / 32][33];
  __shared__ float2 part[32];
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + lane;
  const bool col_ok = j < H;
  const int chunk = ceil_div(B, kBnCluster);
  const int r_begin = blockIdx.y * chunk, r_end = min(B, r_begin + chunk);
  float mean, rstd;
  if (train) {  // uniform over the whole grid: every CTA of every cluster takes the same branch
    float s = 0.f, ss = 0.f;
    if (col_ok)
      for (int r = r_begin + ty; r < r_end; r += kBnThreads / 32) {
        const float v = z[(size_t)r * ldz + j];
        s += v; ss += v * v;
      }
    const float2 t = bn_cluster_colsum2(s, ss, sa, sb, part);
    mean = t.x / B;
    const float var = fmaxf(t.y / B - mean * mean, 0.f);
    rstd = rsqrtf(var + kBnEps);
    if (col_ok && ty == 0 && blockIdx.y == 0) {
      save_mean[j] = mean; save_rstd[j] = rstd;
      mov_mean[j] = decay * mov_mean[j] + (1.f - decay) * mean;
      mov_var[j] = decay * mov_var[j] + (1.f - decay) * var;
    }
  } else {
    mean = col_ok ? mov_mean[j] : 0.f;
    rstd = col_ok ? rsqrtf(mov_var[j] + kBnEps) : 0.f;
  }
  if (!col_ok) return;
  const float g = gamma[j], b = beta[j];
  unsigned long long step = 0;
  if (train && drop.enabled && step_ptr) step = *step_ptr;
  drop.step_lo = (uint32_t)step; drop.step_hi = (uint32_t)(step >> 32);
  for (int r = r_begin + ty; r < r_end; r += kBnThreads / 32) {
    float y = g * (z[(size_t)r * ldz + j] - mean) * rstd + b;
    if (train && drop.enabled) {
      const uint64_t elem = (uint64_t)(r + drop.row_offset) * (uint64_t)H + (uint64_t)j;
      const uint4 rnd = dropout_words(drop, elem >> 2);
      const uint32_t sel = (uint32_t)(elem & 3);
      const uint32_t u = sel == 0 ? rnd.x : sel == 1 ? rnd.y : sel == 2 ? rnd.z : rnd.w;
      y = (u < drop.threshold) ? y * drop.scale : 0.f;
    }
    out[(size_t)r * ldo + j] = __float2bfloat16(y);
  }
}

// dy: in = gradient w.r.t. the BN output (dropout already applied by the producer); out (in place) = gradient w.r.t.
// the pre-activation of the fully-connected layer (through BN and ReLU).
__global__ void __cluster_dims__(1, kBnCluster, 1) __launch_bounds__(kBnThreads)
bn_bwd_kernel(__nv_bfloat16* __restrict__ dy, int ldd, const float* __restrict__ z, int ldz, int B, int H,
              const float* __restrict__ gamma, const float* __restrict__ save_mean, const float* __restrict__ save_rstd,
              float* __restrict__ g_gamma, float* __restrict__ g_beta) {
  __shared__ float sa[kBnThreads / 32][33], sb[kBnThreads / 32][33];
  __shared__ float2 part[32];
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + lane;
  const bool col_ok = j < H;
  const int chunk = ceil_div(B, kBnCluster);
  const int r_begin = blockIdx.y * chunk, r_end = min(B, r_begin + chunk);
  const float mean = col_ok ? save_mean[j] : 0.f, rstd = col_ok ? save_rstd[j] : 0.f;
  float s = 0.f, sx = 0.f;
  if (col_ok)
    for (int r = r_begin + ty; r < r_end; r += kBnThreads / 32) {
      const float d = __bfloat162float(dy[(size_t)r * ldd + j]);
      const float xh = (z[(size_t)r * ldz + j] - mean) * rstd;
      s += d; sx += d * xh;
    }
  const float2 t = bn_cluster_colsum2(s, sx, sa, sb, part);
  if (!col_ok) return;
  if (ty == 0 && blockIdx.y == 0) {
    atomicAdd(g_beta + j, t.x);
    atomicAdd(g_gamma + j, t.y);
  }
  const float g = gamma[j], m1 = t.x / B, m2 = t.y / B;
  for (int r = r_begin + ty; r < r_end; r += kBnThreads / 32) {
    const float zz = z[(size_t)r * ldz + j];
    const float xh = (zz - mean) * rstd;
    const float d = __bfloat162float(dy[(size_t)r * ldd + j]);
    const float dz = zz > 0.f ? g * rstd * (d - m1 - xh * m2) : 0.f;
    dy[(size_t)r * ldd + j] = __float2bfloat16(dz);
  }
}

}  // namespace dfm

using namespace dfm;

extern "C" {

int dfm_bn_fwd(const void* z, int ldz, void* out, int ldo, int B, int H, const float* gamma, const float* beta,
               float* mov_mean, float* mov_var, float* save_mean, float* save_rstd, float decay, int train,
               float keep_prob, unsigned long long seed, int layer, int row_offset, const unsigned long long* step_ptr,
               void* stream) {
  DFM_REQUIRE(B > 0 && H > 0, "bn_fwd: empty problem");
  DropoutCfg d;
  d.enabled = (train && keep_prob < 1.0f) ? 1u : 0u;
  d.threshold = keep_prob >= 1.0f ? 0xFFFFFFFFu : (uint32_t)floor((double)keep_prob * 4294967296.0);
  d.scale = keep_prob >= 1.0f ? 1.0f : 1.0f / keep_prob;
  d.seed_lo = (uint32_t)seed; d.seed_hi = (uint32_t)(seed >> 32); d.step_lo = d.step_hi = 0;
  d.layer = (uint32_t)layer; d.row_offset = (uint32_t)row_offset;
  DFM_CHECK(launch_k(bn_fwd_kernel, dim3(ceil_div(H, 32), kBnCluster), dim3(kBnThreads), 0, (cudaStream_t)stream,
                     (const float*)z, ldz, (__nv_bfloat16*)out, ldo, B, H, gamma, beta, mov_mean, mov_var,
                     save_mean, save_rstd, decay, train, d, step_ptr));
  return DFM_OK;
}

int dfm_bn_bwd(void* dy_inout, int ldd, const void* z, int ldz, int B, int H, const float* gamma, const float* save_mean,
               const float* save_rstd, float* g_gamma, float* g_beta, void* stream) {
  DFM_REQUIRE(B > 0 && H > 0, "bn_bwd: empty problem");
  DFM_CHECK(launch_k(bn_bwd_kernel, dim3(ceil_div(H, 32), kBnCluster), dim3(kBnThreads), 0, (cudaStream_t)stream,
                     (__nv_bfloat16*)dy_inout, ldd, (const float*)z, ldz, B, H, gamma, save_mean, save_rstd,
                     g_gamma, g_beta));
  return DFM_OK;
}

}  // extern "C"
