"""Warm CUDA-event timing of the tcgen05 GEMM variants (us per launch) to localise fixed overheads."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deepfm_b200.ops.kernels import GemmPlan
from deepfm_b200.ops import lib
lib.require_gpu()

def timeit(fn, n=200):
    for _ in range(10): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n

only = sys.argv[1] if len(sys.argv) > 1 else ""
M = 1024
step = torch.zeros(1, device="cuda", dtype=torch.int64)
for (N, K) in [(400, 1248), (400, 400), (128, 1248), (64, 128)]:
    A = torch.randn(M, K, device="cuda").bfloat16(); W = (torch.randn(N, K, device="cuda") * .1).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16); bias = torch.zeros(N, device="cuda")
    variants = {
        "fwd plain": dict(),
        "fwd bias+relu": dict(bias=bias, relu=True),
        "fwd bias+relu+drop": dict(bias=bias, relu=True, keep_prob=0.5, seed=1, step_ptr=step),
    }
    for name, kw in variants.items():
        if only and only not in name: continue
        p = GemmPlan("fwd", A, W, out, M, N, K, **kw)
        print("M%d N%d K%d %-20s %-28s %.2f us" % (M, N, K, name, p.info(), timeit(p.run)))
    if only: continue
    Wt = (torch.randn(K, N, device="cuda") * .1).bfloat16()
    mask = torch.relu(torch.randn(M, N, device="cuda")).bfloat16()
    p = GemmPlan("dgrad", A, Wt, out, M, N, K, mask=mask, mask_scale=2.0)
    print("M%d N%d K%d %-20s %-28s %.2f us" % (M, N, K, "dgrad mask", p.info(), timeit(p.run)))
    dZ = torch.randn(M, N, device="cuda").bfloat16(); X = torch.randn(M, K, device="cuda").bfloat16()
    g = torch.zeros(N, K, device="cuda")
    for sk in (1, 0):
        p = GemmPlan("wgrad", dZ, X, g, N, K, M, split_k=sk)
        print("M%d N%d K%d %-20s %-28s %.2f us" % (N, K, M, "wgrad split=%d" % sk, p.info(), timeit(p.run)))
# torch reference timing
A = torch.randn(M, 1248, device="cuda").bfloat16(); W = torch.randn(400, 1248, device="cuda").bfloat16()
print("torch matmul 1024x400x1248: %.2f us" % timeit(lambda: torch.matmul(A, W.t())))
x = torch.empty(1, device="cuda")
print("empty-ish kernel (fill 1 elem): %.2f us" % timeit(lambda: x.zero_()))
