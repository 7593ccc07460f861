"""Host-side push application of the async parameter server (no GPU needed): C++ fused pass (csrc/io/host_optim.cc) vs
the torch-op reference path, at BASELINE config 4's shape (117581 x 32 table, batch 8192 x 39 lookups, Adam)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deepfm_b200.data import native
from deepfm_b200.data.synthetic import SyntheticCTR
from deepfm_b200.optim import SLOT_INIT, TF1Optimizer

N, K, B, F = 117581, 32, 8192, 39
opt = TF1Optimizer("Adam", 5e-4)
ids, _, _ = SyntheticCTR(F, N, seed=0).batch(B, 0, ids_dtype=torch.int64)
rows = torch.unique(ids)
U = rows.numel()
g = torch.randn(U, K) * 0.01
dense_n = 1248 * 400 + 2 * 400 * 400 + 3 * 400 + 401
gd = torch.randn(dense_n) * 0.01


def fresh():
    return (torch.randn(N, K), [torch.full((N, K), v) for v in SLOT_INIT[opt.name]],
            torch.randn(dense_n), [torch.full((dense_n,), v) for v in SLOT_INIT[opt.name]])


def torch_apply(tab, slots, w, ws, t):
    wr = tab[rows]
    gg = g + 1e-4 * wr
    sl = [s[rows] for s in slots]
    opt.apply_dense(wr, gg, sl, t)
    tab[rows] = wr
    for s, v in zip(slots, sl):
        s[rows] = v
    opt.apply_dense(w, gd, ws, t)


def native_apply(tab, slots, w, ws, t, threads):
    native.sparse_opt_apply(tab, slots, rows, g, opt, 1e-4, t, threads)
    native.dense_opt_apply(w, ws, gd, opt, 0.0, t, threads)


def timeit(fn, reps=20):
    st = fresh()
    fn(*st, 1)
    t0 = time.perf_counter()
    for i in range(reps):
        fn(*st, 2 + i)
    return (time.perf_counter() - t0) / reps * 1e3


print("unique rows per push: %d of %d lookups; dense variables: %d floats" % (U, B * F, dense_n))
print("torch-op path: %.2f ms per push" % timeit(torch_apply))
for th in (1, 2, 4):
    print("native fused pass, %d thread(s): %.2f ms per push" % (th, timeit(lambda *a, th=th: native_apply(*a, th))))
