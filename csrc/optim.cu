// Multi-tensor ("flat") dense optimizer step, single-GPU form (reference op K12: one ApplyAdam/ApplyAdagrad/
// ApplyMomentum/ApplyFtrl launch per variable, 11 launches -> here ONE launch over the flat parameter buffer).
//
// In the same sweep:
//   * the whole-table L2 term of the reference loss (`l2_reg * l2_loss(fm_w) + l2_reg * l2_loss(fm_v)`, PS:245-246):
//     g += l2 * w on the [decay_begin, decay_end) range, and the loss value 0.5*l2*|w|^2 is reduced into reg_loss
//     (reference op K9) — no separate full-table read;
//   * gradients are zeroed after they are consumed (the next step's atomics start from 0);
//   * bf16 shadow copies of the MLP weights (the tcgen05 GEMM operands) are refreshed;
//   * the device-resident step counter is advanced by the last CTA to finish (so CUDA graphs replay unchanged).
#include "common.cuh"
#include "optim_math.cuh"

namespace dfm {

struct FlatOptArgs {
  float* w;
  float* g;
  float* s0;
  float* s1;
  long long n;  // multiple of 4
  OptHyper h;
  float grad_scale;
  long long decay_begin, decay_end;  // multiples of 4
  float l2;
  long long shadow_begin, shadow_end;  // multiples of 4
  __nv_bfloat16* shadow;
  unsigned long long* step_ptr;
  float* reg_loss;
  unsigned int* done_counter;  // scratch: CTA completion counter, self-resetting
  // optional result publication (dfm_flat_opt_step_pub): the last CTA stores pub_n floats from pub_src into record
  // (t - 1) % pub_ring of a host-mapped ring, so the step's result reaches the host without a D2H copy in the stream
  const float* pub_src;
  float* pub_dst;
  int pub_n, pub_ring;
};

constexpr int kOptThreads = 256;

__global__ void __launch_bounds__(kOptThreads) flat_opt_kernel(const FlatOptArgs a) {
  __shared__ float s_lr;
  __shared__ float s_red[kOptThreads / 32];
  pdl_launch_dependents();
  pdl_wait(TR_OPT);
  const unsigned long long t = *a.step_ptr + 1ull;
  if (threadIdx.x == 0) s_lr = opt_step_lr(a.h, t);
  __syncthreads();
  const float lr_t = s_lr;
  const int nslots = opt_num_slots(a.h.opt);
  float reg = 0.f;
  const long long n4 = a.n >> 2;
  // Plain grid-stride float4 sweep at high occupancy (48 registers, 5 CTAs/SM): measured faster on B200 than a 4x
  // unrolled variant whose register count cut the occupancy (24.9 us vs 32.7 us for 4.7 M parameters).
  for (long long i = (long long)blockIdx.x * kOptThreads + threadIdx.x; i < n4; i += (long long)gridDim.x * kOptThreads) {
    const long long e = i << 2;
    float4 w = reinterpret_cast<float4*>(a.w)[i];
    float4 g = reinterpret_cast<float4*>(a.g)[i];
    float4 s0 = nslots > 0 ? reinterpret_cast<float4*>(a.s0)[i] : make_float4(0, 0, 0, 0);
    float4 s1 = nslots > 1 ? reinterpret_cast<float4*>(a.s1)[i] : make_float4(0, 0, 0, 0);
    g.x *= a.grad_scale; g.y *= a.grad_scale; g.z *= a.grad_scale; g.w *= a.grad_scale;
    if (e >= a.decay_begin && e < a.decay_end) {
      reg += w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w;
      g.x += a.l2 * w.x; g.y += a.l2 * w.y; g.z += a.l2 * w.z; g.w += a.l2 * w.w;
    }
    opt_update(a.h, lr_t, g.x, w.x, s0.x, s1.x);
    opt_update(a.h, lr_t, g.y, w.y, s0.y, s1.y);
    opt_update(a.h, lr_t, g.z, w.z, s0.z, s1.z);
    opt_update(a.h, lr_t, g.w, w.w, s0.w, s1.w);
    reinterpret_cast<float4*>(a.w)[i] = w;
    reinterpret_cast<float4*>(a.g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (nslots > 0) reinterpret_cast<float4*>(a.s0)[i] = s0;
    if (nslots > 1) reinterpret_cast<float4*>(a.s1)[i] = s1;
    if (e >= a.shadow_begin && e < a.shadow_end) {
      uint2 pk;
      pk.x = pack_bf16x2(w.x, w.y);
      pk.y = pack_bf16x2(w.z, w.w);
      *reinterpret_cast<uint2*>(a.shadow + (e - a.shadow_begin)) = pk;
    }
  }
  // block reduce of the L2 loss value
  reg = warp_sum(reg);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = reg;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
    for (int i = 0; i < kOptThreads / 32; ++i) tot += s_red[i];
    if (a.reg_loss && tot != 0.f) atomicAdd(a.reg_loss, 0.5f * a.l2 * tot);
    __threadfence();
    trace_end(TR_OPT);
    const unsigned int done = atomicAdd(a.done_counter, 1u);
    if (done == gridDim.x - 1) {  // every CTA has read the step by now
      *a.step_ptr = t;
      *a.done_counter = 0u;
      if (a.pub_dst) {  // every CTA's reg_loss atomic happened before its done-counter increment (fence above)
        __threadfence();
        float* rec = a.pub_dst + ((t - 1ull) % (unsigned long long)a.pub_ring) * (unsigned long long)a.pub_n;
        for (int i = 0; i < a.pub_n; ++i) rec[i] = __ldcg(a.pub_src + i);
        // no fence.sys here: it would hold the kernel (and the next step) for a PCIe round trip (~2.7 us measured);
        // the host reads the ring only after a stream / device synchronisation, which orders these stores
      }
    }
  }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = __float2bfloat16(in[i]);
}

}  // namespace dfm

using namespace dfm;

extern "C" {

// All ranges in elements and multiples of 4; buffers 16-byte aligned. `done_counter` = 4 bytes of zeroed scratch.
int dfm_flat_opt_step_pub(float* w, float* g, float* s0, float* s1, long long n, int opt, float lr, float beta1,
                          float beta2, float eps, float momentum, float grad_scale, long long decay_begin,
                          long long decay_end, float l2, long long shadow_begin, long long shadow_end, void* shadow,
                          unsigned long long* step_ptr, float* reg_loss, unsigned int* done_counter,
                          const float* pub_src, int pub_n, float* pub_ring, int pub_ring_len, void* stream) {
  DFM_REQUIRE((pub_ring == nullptr) || (pub_src != nullptr && pub_n >= 1 && pub_n <= 16 && pub_ring_len >= 1),
              "flat_opt_pub: publication needs a source, 1..16 floats per record and a ring length >= 1");
  DFM_REQUIRE(n > 0 && n % 4 == 0, "flat_opt: n must be a positive multiple of 4");
  DFM_REQUIRE(decay_begin % 4 == 0 && decay_end % 4 == 0 && shadow_begin % 4 == 0 && shadow_end % 4 == 0,
              "flat_opt: ranges must be multiples of 4");
  DFM_REQUIRE(opt >= 0 && opt <= 4, "flat_opt: unknown optimizer id %d", opt);
  FlatOptArgs a;
  a.w = w; a.g = g; a.s0 = s0; a.s1 = s1; a.n = n;
  a.h.opt = opt; a.h.lr = lr; a.h.beta1 = beta1; a.h.beta2 = beta2; a.h.eps = eps; a.h.momentum = momentum;
  a.grad_scale = grad_scale; a.decay_begin = decay_begin; a.decay_end = decay_end; a.l2 = l2;
  a.shadow_begin = shadow_begin; a.shadow_end = shadow_end; a.shadow = (__nv_bfloat16*)shadow;
  a.step_ptr = step_ptr; a.reg_loss = reg_loss; a.done_counter = done_counter;
  a.pub_src = pub_src; a.pub_dst = pub_ring; a.pub_n = pub_n; a.pub_ring = pub_ring_len;
  long long blocks = ceil_div_ll(n / 4, kOptThreads);
  if (blocks > 148 * 8) blocks = 148 * 8;
  DFM_CHECK(launch_k(flat_opt_kernel, dim3((unsigned)blocks), dim3(kOptThreads), 0, (cudaStream_t)stream, a));
  return DFM_OK;
}

int dfm_flat_opt_step(float* w, float* g, float* s0, float* s1, long long n, int opt, float lr, float beta1, float beta2,
                      float eps, float momentum, float grad_scale, long long decay_begin, long long decay_end, float l2,
                      long long shadow_begin, long long shadow_end, void* shadow, unsigned long long* step_ptr,
                      float* reg_loss, unsigned int* done_counter, void* stream) {
  return dfm_flat_opt_step_pub(w, g, s0, s1, n, opt, lr, beta1, beta2, eps, momentum, grad_scale, decay_begin, decay_end,
                               l2, shadow_begin, shadow_end, shadow, step_ptr, reg_loss, done_counter, nullptr, 0, nullptr,
                               0, stream);
}

int dfm_cast_f32_bf16(const float* in, void* out, long long n, void* stream) {
  DFM_REQUIRE(n > 0, "cast: empty");
  long long blocks = ceil_div_ll(n, 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  cast_f32_bf16_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(in, (__nv_bfloat16*)out, n);
  DFM_CHECK(cudaGetLastError());
  return DFM_OK;
}

}  // extern "C"
