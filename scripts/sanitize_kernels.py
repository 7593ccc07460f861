"""Small-shape run of every single-GPU kernel for compute-sanitizer (memcheck / racecheck)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deepfm_b200.models import DeepFMConfig
from deepfm_b200.models.fused import FusedDeepFM
from deepfm_b200.optim import TF1Optimizer
for bn in (False, True):
    cfg = DeepFMConfig(field_size=39, feature_size=3000, embedding_size=32, deep_layers=[400, 128, 64], dropout=[0.5, 0.8, 1.0],
                       l2_reg=1e-4, batch_norm=bn)
    B = 200  # not a multiple of the 128-row tile
    m = FusedDeepFM(cfg, B, TF1Optimizer("Adam", 1e-3), seed=1, use_graph=False)
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(0, 3000, (B, 39), generator=g, dtype=torch.int32); vals = torch.rand(B, 39, generator=g)
    labels = (torch.rand(B, generator=g) < 0.3).float()
    for _ in range(2):
        m.train_step(ids.cuda(), vals.cuda(), labels.cuda())
    m.eval_step(ids[:77].cuda(), vals[:77].cuda(), labels[:77].cuda())
    m.predict(ids[:5].cuda(), vals[:5].cuda())
    torch.cuda.synchronize()
    print("bn=%s loss=%.4f step=%d" % (bn, float(m.last_loss()), m.global_step))
print("SANITIZE_RUN_OK")
