import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deepfm_b200.models import DeepFMConfig
from deepfm_b200.models.fused import FusedDeepFM
from deepfm_b200.optim import TF1Optimizer
from deepfm_b200.data.synthetic import SyntheticCTR
from deepfm_b200.utils.checkpoint import save_checkpoint, load_checkpoint, latest_checkpoint
cfg = DeepFMConfig(field_size=39, feature_size=20000, embedding_size=32, deep_layers=[128, 64, 32], dropout=[0.9] * 3, l2_reg=1e-5)
data = SyntheticCTR(39, 20000, seed=0)
B = 256
m = FusedDeepFM(cfg, B, TF1Optimizer("Adam", 2e-3), seed=0)
for s in range(20):
    ids, vals, labels = data.batch(B, s); m.train_step(ids.cuda(), vals.cuda(), labels.cuda())
torch.cuda.synchronize()
sd = m.state_dict()
d = tempfile.mkdtemp()
prefix = save_checkpoint(d, sd, m.opt)
back = load_checkpoint(latest_checkpoint(d), list(m.named_parameters().keys()), "Adam")
for k in sd:
    if k not in back: print("MISSING", k); continue
    a, b = sd[k].cpu(), back[k].cpu()
    if a.shape != b.shape or not torch.equal(a.float(), b.float()): print("DIFF", k, a.shape, b.shape, float((a.float()-b.float().reshape(a.shape)).abs().max()) if a.numel()==b.numel() else "shape")
print("extra keys", [k for k in back if k not in sd])
P0 = m.P.clone(); S0 = m.S0.clone(); S1 = m.S1.clone()
m.load_state_dict(back, strict=False)
torch.cuda.synchronize()
print("P equal", torch.equal(P0, m.P), "S0 equal", torch.equal(S0, m.S0), "S1 equal", torch.equal(S1, m.S1), "step", m.global_step, m.opt.step)
if not torch.equal(P0, m.P):
    L = m.layout
    for n in L.names():
        a, b = L.view(P0, n), L.view(m.P, n)
        if not torch.equal(a, b): print("  P differs in", n, float((a-b).abs().max()))
if not torch.equal(S0, m.S0):
    L = m.layout
    for n in L.names(trainable_only=True):
        a, b = L.view(S0, n), L.view(m.S0, n)
        if not torch.equal(a, b): print("  S0 differs in", n, float((a-b).abs().max()))
