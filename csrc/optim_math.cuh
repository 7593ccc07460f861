// Per-element optimizer math with TensorFlow-1 update rules (reference PS:260-267; SURVEY 2.2 optimizer table).
// Shared by the local flat optimizer (optim.cu), the fused all-reduce + optimizer (allreduce_opt.cu) and the
// sparse row optimizer (sharded.cu).
#pragma once
#include <cuda_runtime.h>

namespace dfm {

enum { OPT_ADAM = 0, OPT_ADAGRAD = 1, OPT_MOMENTUM = 2, OPT_FTRL = 3, OPT_GD = 4 };

struct OptHyper {
  int opt;
  float lr, beta1, beta2, eps, momentum;
};

// lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t) for Adam (t = 1-based step), lr otherwise.
__device__ __forceinline__ float opt_step_lr(const OptHyper& h, unsigned long long t) {
  if (h.opt != OPT_ADAM) return h.lr;
  // fp32 on purpose (TF keeps beta1_power / beta2_power in fp32 too). A double-precision pow() here ran serially in one
  // thread of EVERY CTA before the CTA could start: ~4 us per CTA wave, i.e. 8-25 us per optimizer launch on B200.
  const float tf = (float)t;
  const float b1t = exp2f(tf * log2f(h.beta1)), b2t = exp2f(tf * log2f(h.beta2));
  return h.lr * sqrtf(1.f - b2t) / (1.f - b1t);
}

// Updates (w, s0, s1) in registers for gradient g. lr_t from opt_step_lr.
__device__ __forceinline__ void opt_update(const OptHyper& h, float lr_t, float g, float& w, float& s0, float& s1) {
  switch (h.opt) {
    case OPT_ADAM: {
      s0 = h.beta1 * s0 + (1.f - h.beta1) * g;
      s1 = h.beta2 * s1 + (1.f - h.beta2) * g * g;
      w -= lr_t * s0 / (sqrtf(s1) + h.eps);
      break;
    }
    case OPT_ADAGRAD: {
      s0 += g * g;
      w -= h.lr * g * rsqrtf(s0);
      break;
    }
    case OPT_MOMENTUM: {
      s0 = h.momentum * s0 + g;
      w -= h.lr * s0;
      break;
    }
    case OPT_FTRL: {  // s0 = accumulator, s1 = linear; lr_power = -0.5, l1 = l2 = 0
      const float a_new = s0 + g * g;
      const float sq_new = sqrtf(a_new);
      const float sigma = (sq_new - sqrtf(s0)) / h.lr;
      s1 += g - sigma * w;
      w = -s1 * h.lr / sq_new;
      s0 = a_new;
      break;
    }
    default: {
      w -= h.lr * g;
      break;
    }
  }
}

__host__ __device__ inline int opt_num_slots(int opt) {
  return (opt == OPT_ADAM || opt == OPT_FTRL) ? 2 : (opt == OPT_GD ? 0 : 1);
}

}  // namespace dfm
