"""Fused cluster MLP forward (csrc/mlp_fused.cu) against the per-layer tcgen05 GEMM chain at the headline shape, with
in-kernel phase stamps. Finding (B200, B = 1024, 1248 -> 400 -> 400 -> 400): the fused kernel is bit-identical and SLOWER
(32 vs 16 us): broadcasting a 128 x 400 bf16 activation block to the 5 CTAs of a cluster through st.shared::cluster takes
4-7 us per layer (distributed shared memory moves ~17-21 B/cycle/SM), more than the kernel boundary it removes (~3 us).
The model therefore keeps the per-layer chain; the kernel stays as a tested experiment. 1 GPU: python scripts/fused_mlp_bench.py"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepfm_b200.ops.kernels import GemmPlan, MlpFwdPlan
M, K0, widths, keeps = 1024, 1248, [400, 400, 400], [0.5, 0.5, 0.5]
def buf(rows, width):
    ld = (width + 1 + 63) // 64 * 64
    b = torch.zeros((rows, ld), device="cuda", dtype=torch.bfloat16); return b, b[:, :width]
xb, x = buf(M, K0); x.copy_((torch.randn(M, K0, device="cuda") * 0.5).bfloat16())
Ws, bs, d = [], [], K0
for h in widths:
    Ws.append((torch.randn(h, d, device="cuda") / d ** 0.5).bfloat16()); bs.append(torch.zeros(h, device="cuda")); d = h
step = torch.tensor([1], device="cuda", dtype=torch.int64)
oa = [buf(M, h) for h in widths]; ob = [buf(M, h) for h in widths]
plans = []; cur = x
for l, h in enumerate(widths):
    plans.append(GemmPlan("fwd", cur, Ws[l], oa[l][1], M, h, cur.shape[1], bias=bs[l], relu=True, keep_prob=keeps[l], seed=1, layer=l, step_ptr=step)); cur = oa[l][1]
fused = MlpFwdPlan(x, Ws, [o[1] for o in ob], bs, M, relu=True, keep_probs=keeps, seed=1, step_ptr=step)
print(fused.info())
def timeit(fn, n=200):
    g = torch.cuda.CUDAGraph()
    fn(); torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(20): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n // 20): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / n
for L in (1, 2, 3):
    f = MlpFwdPlan(x, Ws[:L], [o[1] for o in ob[:L]], bs[:L], M, relu=True, keep_probs=keeps[:L], seed=1, step_ptr=step)
    print("L", L, "chain us %.2f" % timeit(lambda: [p.run() for p in plans[:L]]), "fused us %.2f" % timeit(lambda: f.run()))
f1 = MlpFwdPlan(oa[0][1], Ws[1:2], [ob[1][1]], bs[1:2], M, relu=True, keep_probs=[0.5], seed=1, step_ptr=step)
print("single 400x400 layer: chain us %.2f" % timeit(lambda: plans[1].run()), "fused-kernel us %.2f" % timeit(lambda: f1.run()))
print("equal", [bool(torch.equal(a[0], b[0])) for a, b in zip(oa, ob)])

import ctypes
from deepfm_b200.ops import lib as L_
dbg = torch.zeros(40 * 32, device="cuda", dtype=torch.int64)
L_.load().dfm_mlp_fwd_plan_set_debug(fused._h, ctypes.c_void_p(dbg.data_ptr()))
fused.run(); torch.cuda.synchronize(); dbg.zero_(); fused.run(); torch.cuda.synchronize()
d = dbg.view(40, 32).cpu()
t0 = int(d[:, 0][d[:, 0] > 0].min())
names = {0: "epi start", 1: "L0 acc done", 2: "L0 stored", 3: "L0 B1", 4: "L0 dsmem written", 5: "L0 fence", 6: "L0 B2", 7: "L1 acc done", 8: "L1 stored", 9: "L1 B1", 10: "L1 dsmem", 11: "L1 fence", 12: "L1 B2", 13: "L2 acc done"}
for k in sorted(names):
    col = d[:, k]; col = col[col > 0]
    if len(col): print("%-18s min %.2f max %.2f us" % (names[k], (int(col.min()) - t0) / 1e3, (int(col.max()) - t0) / 1e3))
