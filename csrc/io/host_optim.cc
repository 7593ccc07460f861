// Host-side optimizer application for the asynchronous parameter-server mode (in the reference the PS
// processes run TF's ApplyAdam / ApplyAdagrad / ... C++ kernels on the variables they own). The store here is one shared
// host mapping; a worker's applier thread calls these to fold a pushed gradient into it:
//
//   dfmio_sparse_opt_apply   touched rows of an embedding table: one fused pass per row
//                            (g += l2 * w; TF-1 rule; write w and the slot rows back) instead of ~15 tensor-op passes
//                            with gather / scatter temporaries; rows are split over `threads` std::threads
//   dfmio_dense_opt_apply    a whole dense variable (MLP weights / biases)
//
// Update rules = csrc/optim_math.cuh (TF-1 semantics: Adam epsilon outside the sqrt and lr_t = lr sqrt(1-b2^t)/(1-b1^t),
// Adagrad without epsilon, Momentum, FTRL lr_power -0.5, plain SGD). Deliberately UNLOCKED (Hogwild), like the
// reference's asynchronous PS (README-EN:14, PDF p35): concurrent appliers may interleave on a hot row.
#include <math.h>
#include <stdint.h>

#include <algorithm>
#include <thread>
#include <vector>

namespace {

enum { OPT_ADAM = 0, OPT_ADAGRAD = 1, OPT_MOMENTUM = 2, OPT_FTRL = 3, OPT_GD = 4 };

struct Hyper {
  int opt;
  float lr, beta1, beta2, eps, momentum, lr_t;
};

// One contiguous span (a table row or a whole dense variable). The optimizer switch sits OUTSIDE the element loops so
// each loop is straight-line code the compiler vectorises (vsqrtps / vdivps are IEEE-exact: results do not depend on the
// vector width); target_clones emits an AVX2 and a baseline SSE version and picks one at load time, so the library still
// runs on any x86-64 host it is copied to.
#if !defined(__SANITIZE_THREAD__)  // the ifunc resolver of a multiversioned function runs before TSan's runtime is up
__attribute__((target_clones("avx2", "default")))
#endif
void apply_span(const Hyper& h, float* __restrict__ w, float* __restrict__ s0, float* __restrict__ s1,
                const float* __restrict__ g, float l2, long long n) {
  switch (h.opt) {
    case OPT_ADAM: {
      const float b1 = h.beta1, b2 = h.beta2, c1 = 1.f - h.beta1, c2 = 1.f - h.beta2, lr_t = h.lr_t, eps = h.eps;
      for (long long i = 0; i < n; ++i) {
        const float gi = g[i] + l2 * w[i];
        const float m = b1 * s0[i] + c1 * gi;
        const float v = b2 * s1[i] + c2 * gi * gi;
        s0[i] = m;
        s1[i] = v;
        w[i] -= lr_t * m / (sqrtf(v) + eps);
      }
      break;
    }
    case OPT_ADAGRAD:
      for (long long i = 0; i < n; ++i) {
        const float gi = g[i] + l2 * w[i];
        const float a = s0[i] + gi * gi;
        s0[i] = a;
        w[i] -= h.lr * gi / sqrtf(a);
      }
      break;
    case OPT_MOMENTUM:
      for (long long i = 0; i < n; ++i) {
        const float gi = g[i] + l2 * w[i];
        const float acc = h.momentum * s0[i] + gi;
        s0[i] = acc;
        w[i] -= h.lr * acc;
      }
      break;
    case OPT_FTRL:
      for (long long i = 0; i < n; ++i) {
        const float gi = g[i] + l2 * w[i];
        const float a_new = s0[i] + gi * gi;
        const float sq_new = sqrtf(a_new);
        const float sigma = (sq_new - sqrtf(s0[i])) / h.lr;
        const float lin = s1[i] + (gi - sigma * w[i]);
        s1[i] = lin;
        w[i] = -lin * h.lr / sq_new;
        s0[i] = a_new;
      }
      break;
    default:
      for (long long i = 0; i < n; ++i) w[i] -= h.lr * (g[i] + l2 * w[i]);
      break;
  }
}

Hyper make_hyper(int opt, float lr, float beta1, float beta2, float eps, float momentum, long long step) {
  Hyper h{opt, lr, beta1, beta2, eps, momentum, lr};
  if (opt == OPT_ADAM) {
    const double t = (double)step;
    h.lr_t = (float)((double)lr * sqrt(1.0 - pow((double)beta2, t)) / (1.0 - pow((double)beta1, t)));
  }
  return h;
}

template <typename Fn>
void parallel_for(long long n, int threads, long long min_chunk, Fn fn) {
  int T = (int)std::max<long long>(1, std::min<long long>(threads, n / std::max<long long>(1, min_chunk)));
  if (T <= 1) {
    fn(0, n);
    return;
  }
  std::vector<std::thread> th;
  const long long per = (n + T - 1) / T;
  for (int t = 0; t < T; ++t) {
    const long long b = t * per, e = std::min(n, b + per);
    if (b >= e) break;
    th.emplace_back([=] { fn(b, e); });
  }
  for (auto& x : th) x.join();
}

}  // namespace

extern "C" {

// table [N, K] (and slot tables s0 / s1 of the same shape, NULL where the optimizer has fewer slots); rows [nrows] int64
// (unique); grads [nrows, K] dense. `step` is the 1-based update index (Adam bias correction). Returns 0, or -1 on bad
// arguments (unknown optimizer, missing slot table, row index outside [0, N)).
int dfmio_sparse_opt_apply(float* table, float* s0, float* s1, long long N, int K, const long long* rows, long long nrows,
                           const float* grads, int opt, float lr, float beta1, float beta2, float eps, float momentum,
                           float l2, long long step, int threads) {
  if (!table || !rows || !grads || N <= 0 || K <= 0 || nrows < 0 || opt < 0 || opt > 4) return -1;
  const int nslots = (opt == OPT_ADAM || opt == OPT_FTRL) ? 2 : (opt == OPT_GD ? 0 : 1);
  if ((nslots > 0 && !s0) || (nslots > 1 && !s1)) return -1;
  for (long long i = 0; i < nrows; ++i)
    if (rows[i] < 0 || rows[i] >= N) return -1;
  const Hyper h = make_hyper(opt, lr, beta1, beta2, eps, momentum, step);
  parallel_for(nrows, threads, 256, [=](long long b, long long e) {
    for (long long i = b; i < e; ++i) {
      const size_t base = (size_t)rows[i] * (size_t)K;
      apply_span(h, table + base, nslots > 0 ? s0 + base : nullptr, nslots > 1 ? s1 + base : nullptr,
                 grads + (size_t)i * (size_t)K, l2, K);
    }
  });
  return 0;
}

int dfmio_dense_opt_apply(float* w, float* s0, float* s1, const float* g, long long n, int opt, float lr, float beta1,
                          float beta2, float eps, float momentum, float l2, long long step, int threads) {
  if (!w || !g || n < 0 || opt < 0 || opt > 4) return -1;
  const int nslots = (opt == OPT_ADAM || opt == OPT_FTRL) ? 2 : (opt == OPT_GD ? 0 : 1);
  if ((nslots > 0 && !s0) || (nslots > 1 && !s1)) return -1;
  const Hyper h = make_hyper(opt, lr, beta1, beta2, eps, momentum, step);
  parallel_for(n, threads, 65536, [=](long long b, long long e) {
    apply_span(h, w + b, nslots > 0 ? s0 + b : nullptr, nslots > 1 ? s1 + b : nullptr, g + b, l2, e - b);
  });
  return 0;
}

}  // extern "C"
