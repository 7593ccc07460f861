// Fused output head: final-layer dot product, 4-way logit sum, sigmoid, numerically stable log-loss and its
// gradient, plus every reduction that hangs off dy (reference ops K7+K8: PS:221-232,244 / HVD:213-224,236).
//
//   y_d  = x[b,:] . w_out + b_out            y = fm_bias + yfm[b] + y_d          prob = sigmoid(y)
//   loss += max(y,0) - y*label + log1p(exp(-|y|))          (sum; divided by the global batch on the host side
//                                                           through inv_batch being folded into dy and loss_scale)
//   dy[b] = (prob - label) * inv_batch
//   dZ[b,j] = dy[b] * w_out[j] * (x[b,j] != 0) * mask_scale   <- gradient w.r.t. the last hidden pre-activation
//                                                              (x = dropout(relu(z)): its zero pattern IS the mask)
//   g_b_out += dy[b]    g_fm_bias += dy[b]
//   (g_w_out[j] = sum_b dy[b]*x[b,j] and g_b_last[j] = sum_b dZ[b,j] are weighted column sums: colsum_bf16)
#include "common.cuh"

namespace dfm {

constexpr int kLossThreads = 256;
constexpr int kLossWarps = kLossThreads / 32;

// One warp per sample, pure streaming: no column accumulators here (those are the two weighted column sums
// g_w_out = colsum(x, w=dy) and g_b_last = colsum(dz), run by colsum_bf16 on the wgrad branch of the step graph).
__global__ void __launch_bounds__(kLossThreads)
logit_loss_kernel(const __nv_bfloat16* __restrict__ x, int ldx, int H, const float* __restrict__ w_out,
                  const float* __restrict__ b_out, const float* __restrict__ fm_bias, const float* __restrict__ yfm,
                  const float* __restrict__ labels, int B, float inv_batch, float mask_scale, int train,
                  float* __restrict__ prob, float* __restrict__ logit, float* __restrict__ dy,
                  __nv_bfloat16* __restrict__ dz, int lddz, float* __restrict__ loss_sum,
                  float* __restrict__ g_b_out, float* __restrict__ g_fm_bias, float* __restrict__ g_b_top,
                  float* __restrict__ g_w_out) {
  __shared__ float s_scalar[2];  // loss, sum dy
  // optional column accumulators (train mode, both pointers given): s_col[0][j] = sum_b dZ[b,j] (bias gradient of the
  // last hidden layer), s_col[1][j] = sum_b dy[b] x[b,j] (gradient of deep_out.weight). One shared-memory atomic per
  // element here replaces two colsum_bf16 launches that re-read dZ and X from global memory.
  extern __shared__ float s_col[];  // [kLossWarps][2][H]: every warp owns a row pair, every lane its columns: no atomics
  const bool cols = train && g_w_out != nullptr;
  const bool has_gb = g_b_top != nullptr;  // the bias column sum is optional (the wgrad GEMM can produce it instead)
  if (threadIdx.x < 2) s_scalar[threadIdx.x] = 0.f;
  if (cols)
    for (int j = threadIdx.x; j < 2 * H * kLossWarps; j += kLossThreads) s_col[j] = 0.f;
  pdl_launch_dependents();
  pdl_wait(TR_LOSS);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* c0 = s_col + (size_t)warp * 2 * H;  // colsum(dZ) partial of this warp
  float* c1 = c0 + H;                        // X^T dy partial of this warp
  const float bias = __ldg(b_out) + __ldg(fm_bias);
  float loss_acc = 0.f, dy_acc = 0.f;
  for (int b = blockIdx.x * kLossWarps + warp; b < B; b += gridDim.x * kLossWarps) {
    const __nv_bfloat16* xr = x + (size_t)b * ldx;
    float dot = 0.f;
    for (int j = lane * 2; j < H; j += 64) {
      if (j + 1 < H) {
        const uint32_t u = *reinterpret_cast<const uint32_t*>(xr + j);
        dot += bf16_lo(u) * __ldg(w_out + j) + bf16_hi(u) * __ldg(w_out + j + 1);
      } else {
        dot += __bfloat162float(xr[j]) * __ldg(w_out + j);
      }
    }
    dot = warp_sum(dot);
    const float y = bias + yfm[b] + dot;
    const float p = 1.f / (1.f + __expf(-y));
    if (lane == 0) {
      if (prob) prob[b] = p;
      if (logit) logit[b] = y;
    }
    if (labels == nullptr) continue;  // predict mode
    const float lab = labels[b];
    const float l = fmaxf(y, 0.f) - y * lab + log1pf(__expf(-fabsf(y)));
    const float d = (p - lab) * inv_batch;
    if (lane == 0) { loss_acc += l; dy_acc += d; }
    if (!train) continue;
    if (lane == 0) dy[b] = d;
    __nv_bfloat16* dzr = dz + (size_t)b * lddz;
    for (int j = lane * 2; j < H; j += 64) {
      if (j + 1 < H) {
        const uint32_t u = *reinterpret_cast<const uint32_t*>(xr + j);
        const float z0 = (u & 0x7FFFu) ? d * __ldg(w_out + j) * mask_scale : 0.f;
        const float z1 = (u & 0x7FFF0000u) ? d * __ldg(w_out + j + 1) * mask_scale : 0.f;
        *reinterpret_cast<uint32_t*>(dzr + j) = pack_bf16x2(z0, z1);
        if (cols) {
          c0[j] += z0; c0[j + 1] += z1;
          c1[j] += d * bf16_lo(u); c1[j + 1] += d * bf16_hi(u);
        }
      } else {
        const float x0 = __bfloat162float(xr[j]);
        const float z0 = x0 != 0.f ? d * __ldg(w_out + j) * mask_scale : 0.f;
        dzr[j] = __float2bfloat16(z0);
        if (cols) { c0[j] += z0; c1[j] += d * x0; }
      }
    }
  }
  if (labels == nullptr) return;
  if (lane == 0) {
    atomicAdd(&s_scalar[0], loss_acc);
    atomicAdd(&s_scalar[1], dy_acc);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    atomicAdd(loss_sum, s_scalar[0]);
    if (train) {
      atomicAdd(g_b_out, s_scalar[1]);
      atomicAdd(g_fm_bias, s_scalar[1]);
    }
  }
  if (cols)
    for (int j = threadIdx.x; j < H; j += kLossThreads) {
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int w = 0; w < kLossWarps; ++w) { a0 += s_col[(size_t)w * 2 * H + j]; a1 += s_col[(size_t)w * 2 * H + H + j]; }
      if (has_gb && a0 != 0.f) atomicAdd(g_b_top + j, a0);
      if (a1 != 0.f) atomicAdd(g_w_out + j, a1);
    }
  trace_end(TR_LOSS);
}

// Weighted column sums of a bf16 [B,H] matrix: out[j] += sum_b a[b,j] * (w ? w[b] : 1).
// Used for every bias gradient (w = null) and for g_w_out = X_top^T dy (w = dy).
// Block = 32 lanes x 8 row-lanes over a 64-column x kColsumRows tile; grid = (ceil(H/64), ceil(B/kColsumRows)).
constexpr int kColsumRows = 64;
__global__ void __launch_bounds__(256) colsum_bf16_kernel(const __nv_bfloat16* __restrict__ a, int lda, int B, int H,
                                                          const float* __restrict__ w, float* __restrict__ out) {
  __shared__ float s_part[8][64];
  const int lane = threadIdx.x & 31, ty = threadIdx.x >> 5;
  pdl_launch_dependents();
  pdl_wait(TR_COLSUM);
  const int j = blockIdx.x * 64 + lane * 2;
  const int r0 = blockIdx.y * kColsumRows;
  float a0 = 0.f, a1 = 0.f;
  if (j < H) {
    const bool pair = (j + 1 < H) && ((lda & 1) == 0);
#pragma unroll
    for (int i = 0; i < kColsumRows / 8; ++i) {
      const int r = r0 + ty + i * 8;
      if (r < B) {
        const float wr = w ? __ldg(w + r) : 1.f;
        if (pair) {
          const uint32_t u = __ldg(reinterpret_cast<const uint32_t*>(a + (size_t)r * lda + j));
          a0 += bf16_lo(u) * wr;
          a1 += bf16_hi(u) * wr;
        } else {
          a0 += __bfloat162float(a[(size_t)r * lda + j]) * wr;
          if (j + 1 < H) a1 += __bfloat162float(a[(size_t)r * lda + j + 1]) * wr;
        }
      }
    }
  }
  s_part[ty][lane * 2] = a0;
  s_part[ty][lane * 2 + 1] = a1;
  __syncthreads();
  if (threadIdx.x < 64) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += s_part[k][threadIdx.x];
    const int jj = blockIdx.x * 64 + threadIdx.x;
    if (jj < H) atomicAdd(out + jj, t);
  }
}

// 200-threshold confusion histogram for streaming AUC (`tf.metrics.auc`, PS:249-251): hist[c][k] counts class-c
// samples whose prediction has exactly k thresholds strictly below it (thr sorted ascending, T entries).
// Thresholds are staged in shared memory once per CTA (the binary search is 8 dependent loads: ~5 us from global, cold);
// CTR predictions cluster in a few buckets, so equal (class, bucket) keys are counted per warp with match_any and added
// with one shared-memory atomic per distinct key.
__global__ void __launch_bounds__(256) auc_hist_kernel(const float* __restrict__ prob, const float* __restrict__ labels,
                                                       int B, const float* __restrict__ thr, int T,
                                                       float* __restrict__ hist) {
  extern __shared__ float sh[];  // [2*(T+1)] histogram, then [T] thresholds
  float* sthr = sh + 2 * (T + 1);
  for (int i = threadIdx.x; i < 2 * (T + 1); i += 256) sh[i] = 0.f;
  for (int i = threadIdx.x; i < T; i += 256) sthr[i] = thr[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  for (int b0 = blockIdx.x * 256 + (threadIdx.x & ~31); b0 < B; b0 += gridDim.x * 256) {  // warp-uniform trip count
    const int b = b0 + lane;
    unsigned key = 0x80000000u | (unsigned)lane;
    if (b < B) {
      const float p = prob[b];
      int lo = 0, hi = T;  // first index with thr[idx] >= p  == number of thresholds < p
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (sthr[mid] < p) lo = mid + 1; else hi = mid;
      }
      key = (unsigned)((labels[b] > 0.5f ? 1 : 0) * (T + 1) + lo);
    }
    const unsigned mask = __match_any_sync(0xffffffffu, key);
    if (b < B && lane == __ffs(mask) - 1) atomicAdd(&sh[key], (float)__popc(mask));
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * (T + 1); i += 256)
    if (sh[i] != 0.f) atomicAdd(hist + i, sh[i]);
}

}  // namespace dfm

using namespace dfm;

extern "C" {

// labels == NULL -> predict mode (only prob/logit written). train == 0 -> eval (loss only, no gradients).
// g_b_top / g_w_out (both or neither; train mode): additionally g_b_top[j] += sum_b dZ[b,j] and g_w_out[j] += sum_b dy[b]
// x[b,j], the two column sums that otherwise need dfm_colsum_bf16 launches.
int dfm_logit_loss_cols(const void* x, int ldx, int H, const float* w_out, const float* b_out, const float* fm_bias,
                        const float* yfm, const float* labels, int B, float inv_batch, float mask_scale, int train,
                        float* prob, float* logit, float* dy, void* dz, int lddz, float* loss_sum, float* g_b_out,
                        float* g_fm_bias, float* g_b_top, float* g_w_out, void* stream) {
  DFM_REQUIRE(B > 0 && H > 0, "logit_loss: empty problem");
  DFM_REQUIRE(ldx % 2 == 0 && (!train || lddz % 2 == 0), "logit_loss: leading dimensions must be even");
  DFM_REQUIRE(g_b_top == nullptr || g_w_out != nullptr, "logit_loss: g_b_top needs g_w_out as well");
  DFM_REQUIRE(g_w_out == nullptr || H <= 1024, "logit_loss: column sums need H <= 1024 (shared-memory accumulators)");
  int blocks = ceil_div(B, kLossWarps);
  if (blocks > 148 * 4) blocks = 148 * 4;
  const size_t smem = (g_w_out && train) ? 2 * (size_t)H * kLossWarps * sizeof(float) : 0;
  if (smem > 48 * 1024) {
    static bool big_smem_set = false;
    if (!big_smem_set) {
      DFM_CHECK(cudaFuncSetAttribute(logit_loss_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
      big_smem_set = true;
    }
  }
  DFM_CHECK(launch_k(logit_loss_kernel, dim3(blocks), dim3(kLossThreads), smem, (cudaStream_t)stream,
                     (const __nv_bfloat16*)x, ldx, H, w_out, b_out, fm_bias, yfm, labels, B, inv_batch, mask_scale, train,
                     prob, logit, dy, (__nv_bfloat16*)dz, lddz, loss_sum, g_b_out, g_fm_bias, g_b_top, g_w_out));
  return DFM_OK;
}

int dfm_logit_loss(const void* x, int ldx, int H, const float* w_out, const float* b_out, const float* fm_bias,
                   const float* yfm, const float* labels, int B, float inv_batch, float mask_scale, int train,
                   float* prob, float* logit, float* dy, void* dz, int lddz, float* loss_sum, float* g_b_out,
                   float* g_fm_bias, void* stream) {
  return dfm_logit_loss_cols(x, ldx, H, w_out, b_out, fm_bias, yfm, labels, B, inv_batch, mask_scale, train, prob, logit,
                             dy, dz, lddz, loss_sum, g_b_out, g_fm_bias, nullptr, nullptr, stream);
}

// out[j] += sum_b a[b,j] * (w ? w[b] : 1)
int dfm_colsum_bf16(const void* a, int lda, int B, int H, const float* w, float* out, void* stream) {
  DFM_REQUIRE(B > 0 && H > 0, "colsum: empty problem");
  dim3 grid(ceil_div(H, 64), ceil_div(B, kColsumRows));
  DFM_CHECK(launch_k(colsum_bf16_kernel, grid, dim3(256), 0, (cudaStream_t)stream, (const __nv_bfloat16*)a, lda, B, H,
                     w, out));
  return DFM_OK;
}

// hist [2, T+1] f32 +=
int dfm_auc_hist(const float* prob, const float* labels, int B, const float* thr, int T, float* hist, void* stream) {
  DFM_REQUIRE(B > 0 && T > 0 && T <= 4000, "auc_hist: need 0 < thresholds <= 4000 (shared-memory histogram)");
  int blocks = ceil_div(B, 256);
  if (blocks > 148) blocks = 148;
  auc_hist_kernel<<<blocks, 256, (3 * T + 2) * sizeof(float), (cudaStream_t)stream>>>(prob, labels, B, thr, T, hist);
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

}  // extern "C"
