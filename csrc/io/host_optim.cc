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
#include <condition_variable>
#include <functional>
#include <mutex>
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
// vector width). That needs -fno-math-errno (build.py): with errno semantics gcc keeps sqrtf scalar, and the whole pass
// ran at ~12 cycles per element (objdump showed sqrtss / divss only; 14.6 -> 4.7 ms per push once vectorised); target_clones emits an AVX2 and a baseline SSE version and picks one at load time, so the library still
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

// Software prefetch of a table row that will be UPDATED a few rows from now (write intent). Measured on the parameter-
// server push of a batch of 8192 (83,696 scattered rows of 128 bytes in three 15 MB tables): see scripts/ps_apply_bench.py.
constexpr long long kPrefetchAhead = 12;
inline void prefetch_row(const float* p, int K) {
  for (int b = 0; b < K * 4; b += 64) __builtin_prefetch(reinterpret_cast<const char*>(p) + b, 1);
}

Hyper make_hyper(int opt, float lr, float beta1, float beta2, float eps, float momentum, long long step) {
  Hyper h{opt, lr, beta1, beta2, eps, momentum, lr};
  if (opt == OPT_ADAM) {
    const double t = (double)step;
    h.lr_t = (float)((double)lr * sqrt(1.0 - pow((double)beta2, t)) / (1.0 - pow((double)beta1, t)));
  }
  return h;
}

// Persistent worker pool: the parameter-server applier calls into this file ~10 times per training step (two tables + the
// MLP variables); spawning and joining std::threads per call cost more than the arithmetic (measured: ~1 ms per step).
class Pool {
 public:
  static Pool& get() {
    static Pool p;
    return p;
  }
  // run fn(b, e) over [0, n) in `parts` contiguous chunks; the caller takes one chunk itself
  void run(long long n, int parts, const std::function<void(long long, long long)>& fn) {
    const long long per = (n + parts - 1) / parts;
    std::unique_lock<std::mutex> lk(mu_);
    ensure(parts - 1);
    fn_ = &fn;
    n_ = n;
    per_ = per;
    next_ = 1;  // chunk 0 is the caller's
    parts_ = parts;
    pending_ = parts - 1;
    ++gen_;
    cv_work_.notify_all();
    lk.unlock();
    fn(0, std::min(n, per));
    lk.lock();
    cv_done_.wait(lk, [&] { return pending_ == 0; });
    fn_ = nullptr;
  }

 private:
  Pool() = default;
  ~Pool() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
      cv_work_.notify_all();
    }
    for (auto& t : th_) t.join();
  }
  void ensure(int k) {  // mu_ held
    while ((int)th_.size() < k) th_.emplace_back([this] { loop(); });
  }
  void loop() {
    std::unique_lock<std::mutex> lk(mu_);
    unsigned long long seen = 0;
    for (;;) {
      cv_work_.wait(lk, [&] { return stop_ || (gen_ != seen && next_ < parts_); });
      if (stop_) return;
      while (next_ < parts_) {
        const int c = next_++;
        const long long b = c * per_, e = std::min(n_, b + per_);
        const auto* fn = fn_;
        lk.unlock();
        if (b < e) (*fn)(b, e);
        lk.lock();
        if (--pending_ == 0) cv_done_.notify_all();
      }
      seen = gen_;
    }
  }
  std::mutex mu_;
  std::condition_variable cv_work_, cv_done_;
  std::vector<std::thread> th_;
  const std::function<void(long long, long long)>* fn_ = nullptr;
  long long n_ = 0, per_ = 0;
  int next_ = 0, parts_ = 0, pending_ = 0;
  unsigned long long gen_ = 0;
  bool stop_ = false;
};

std::mutex g_call_mu;  // one parallel region at a time (callers are the applier threads of this process)

template <typename Fn>
void parallel_for(long long n, int threads, long long min_chunk, Fn fn) {
  int T = (int)std::max<long long>(1, std::min<long long>(threads, n / std::max<long long>(1, min_chunk)));
  if (T <= 1) {
    fn(0, n);
    return;
  }
  std::lock_guard<std::mutex> g(g_call_mu);
  const std::function<void(long long, long long)> f = fn;
  Pool::get().run(n, T, f);
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
      if (i + kPrefetchAhead < e) {
        const size_t pb = (size_t)rows[i + kPrefetchAhead] * (size_t)K;
        prefetch_row(table + pb, K);
        if (nslots > 0) prefetch_row(s0 + pb, K);
        if (nslots > 1) prefetch_row(s1 + pb, K);
      }
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

// One call per pushed training step of a parameter-server worker: the touched rows of both tables (rows int32 [nrows],
// unique; gv [nrows, K], gw [nrows]) with touched-row L2, then `nvars` dense variables (pointer arrays; no L2). The
// applier thread is a Python thread: one call instead of ~10 means one GIL hand-off per step instead of ~20 (measured: the
// per-variable calls spent 1.7 ms per step, almost all of it waiting for the interpreter lock held by the training thread).
int dfmio_ps_apply_push(float* tab_v, float* sv0, float* sv1, float* tab_w, float* sw0, float* sw1, long long N, int K,
                        const int* rows, long long nrows, const float* gv, const float* gw, int nvars, float* const* w,
                        float* const* s0, float* const* s1, const float* const* g, const long long* numel, int opt,
                        float lr, float beta1, float beta2, float eps, float momentum, float l2, long long step,
                        int threads) {
  if (!tab_v || !tab_w || !rows || !gv || !gw || N <= 0 || K <= 0 || nrows < 0 || opt < 0 || opt > 4 || nvars < 0) return -1;
  const int nslots = (opt == OPT_ADAM || opt == OPT_FTRL) ? 2 : (opt == OPT_GD ? 0 : 1);
  if ((nslots > 0 && (!sv0 || !sw0)) || (nslots > 1 && (!sv1 || !sw1))) return -1;
  for (long long i = 0; i < nrows; ++i)
    if (rows[i] < 0 || rows[i] >= N) return -1;
  const Hyper h = make_hyper(opt, lr, beta1, beta2, eps, momentum, step);
  parallel_for(nrows, threads, 256, [=](long long b, long long e) {
    for (long long i = b; i < e; ++i) {
      if (i + kPrefetchAhead < e) {  // the rows of a push are scattered over the tables: start their cache misses early
        const size_t pr = (size_t)rows[i + kPrefetchAhead], pb = pr * (size_t)K;
        prefetch_row(tab_v + pb, K);
        if (nslots > 0) prefetch_row(sv0 + pb, K);
        if (nslots > 1) prefetch_row(sv1 + pb, K);
        __builtin_prefetch(tab_w + pr, 1);
        if (nslots > 0) __builtin_prefetch(sw0 + pr, 1);
        if (nslots > 1) __builtin_prefetch(sw1 + pr, 1);
      }
      const size_t r = (size_t)rows[i], base = r * (size_t)K;
      apply_span(h, tab_v + base, nslots > 0 ? sv0 + base : nullptr, nslots > 1 ? sv1 + base : nullptr,
                 gv + (size_t)i * (size_t)K, l2, K);
      apply_span(h, tab_w + r, nslots > 0 ? sw0 + r : nullptr, nslots > 1 ? sw1 + r : nullptr, gw + i, l2, 1);
    }
  });
  for (int v = 0; v < nvars; ++v) {
    if (!w[v] || !g[v] || (nslots > 0 && !s0[v]) || (nslots > 1 && !s1[v])) return -1;
    float* wv = w[v];
    float* a0 = nslots > 0 ? s0[v] : nullptr;
    float* a1 = nslots > 1 ? s1[v] : nullptr;
    const float* gv_ = g[v];
    parallel_for(numel[v], threads, 32768, [=](long long b, long long e) {
      apply_span(h, wv + b, a0 ? a0 + b : nullptr, a1 ? a1 + b : nullptr, gv_ + b, 0.f, e - b);
    });
  }
  return 0;
}

}  // extern "C"
