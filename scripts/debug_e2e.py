import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deepfm_b200.models import DeepFMConfig
from deepfm_b200.optim import TF1Optimizer
from deepfm_b200.runtime.trainer import Trainer
from deepfm_b200.data.synthetic import SyntheticCTR
from deepfm_b200.models.fused import pack_batch
cfg = DeepFMConfig(field_size=39, feature_size=20000, embedding_size=32, deep_layers=[128, 64, 32], dropout=[0.9] * 3, l2_reg=1e-5)
data = SyntheticCTR(39, 20000, seed=0)
B = 256
def run(mode, steps=120):
    tr = Trainer(cfg, B, TF1Optimizer("Adam", 2e-3), seed=0, backend="fused")
    out = []
    for s in range(steps):
        ids, vals, labels = data.batch(B, s)
        if mode == "device":
            tr.train_batch_device(ids.cuda(), vals.cuda(), labels.cuda())
            if s % 20 == 0: out.append(round(float(tr.last_loss()), 4))
        elif mode == "host":
            tr.train_batch(ids.pin_memory(), vals.pin_memory(), labels.pin_memory())
        elif mode == "packed":
            tr.train_packed(pack_batch(ids, vals, labels, pin=True))
    if mode != "device":
        l = tr.drain(); out = [round(x, 4) for x in l[::20]]
    ev = tr.evaluate([data.batch(B, 1000 + i) for i in range(8)])
    print(mode, out, "eval", round(ev["loss"], 4), round(ev["auc"], 4), flush=True)
for mode in ("device", "host", "packed"):
    run(mode)
os.environ["DFM_PDL"] = "0"
