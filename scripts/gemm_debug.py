"""GPU-side GEMM diagnostics: prints per-mode error maps so a descriptor/layout bug can be localised in one run."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deepfm_b200.ops.kernels import GemmPlan
from deepfm_b200.ops import lib
lib.require_gpu()
torch.manual_seed(0)

def report(name, out, ref, info):
    out, ref = out.float(), ref.float()
    err = (out - ref).abs()
    rel = float(err.max() / (ref.abs().max() + 1e-9))
    print("%s: rel_err=%.3e info=%s" % (name, rel, info))
    if rel > 1e-2:
        M, N = out.shape
        # per 32-row x 16-col block error summary
        bm, bn = 32, 16
        bad = []
        for i in range(0, min(M, 256), bm):
            row = ""
            for j in range(0, min(N, 256), bn):
                e = float(err[i:i+bm, j:j+bn].max() / (ref.abs().max() + 1e-9))
                row += "X" if e > 1e-2 else "."
            bad.append(row)
        print("\n".join(bad))
        print("out[0,:8]", out[0, :8].tolist()); print("ref[0,:8]", ref[0, :8].tolist())
        print("out[1,:8]", out[1, :8].tolist()); print("ref[1,:8]", ref[1, :8].tolist())

for (M, N, K) in [(128, 64, 64), (128, 64, 128), (256, 128, 256), (1024, 400, 1248)]:
    A = (torch.randn(M, K, device="cuda") * 0.5).bfloat16(); W = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    p = GemmPlan("fwd", A, W, out, M, N, K); p.run(); torch.cuda.synchronize()
    report("fwd  %s" % ((M, N, K),), out, A.float() @ W.float().t(), p.info())
    Wt = (torch.randn(K, N, device="cuda") * 0.1).bfloat16()
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    p = GemmPlan("dgrad", A, Wt, out, M, N, K); p.run(); torch.cuda.synchronize()
    report("dgrad%s" % ((M, N, K),), out, A.float() @ Wt.float(), p.info())
    At = (torch.randn(K, M, device="cuda") * 0.5).bfloat16()
    out = torch.zeros(M, N, device="cuda", dtype=torch.float32)
    p = GemmPlan("wgrad", At, Wt, out, M, N, K, split_k=1); p.run(); torch.cuda.synchronize()
    report("wgrad%s" % ((M, N, K),), out, At.float().t() @ Wt.float(), p.info())
print("gemm_debug done")
