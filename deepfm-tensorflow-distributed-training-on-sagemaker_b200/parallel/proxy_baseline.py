"""NCCL + cuBLAS + torch.optim PROXY of the reference's Horovod trainer (2-hvd-gpu) — the baseline, never the product.

The reference itself (TF 1.14/1.15 + Horovod + OpenMPI) cannot be installed offline here, so the
bar is this stand-in: the same model, built from stock PyTorch ops — `F.embedding`, `nn.Linear` (cuBLAS)
under bf16 autocast, stock dropout, `binary_cross_entropy_with_logits`, explicit whole-table L2 terms, autograd,
`torch.optim.Adam` — one process per GPU, gradients averaged with ONE flat NCCL all-reduce per step (the analogue of
Horovod's fusion buffer, HVD:262). Same Trainer interface as the fused engine so bench.py can time both.
"""
from __future__ import annotations

from collections import OrderedDict

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from ..models.config import DeepFMConfig
from ..models.params import init_params


class _Net(nn.Module):
    def __init__(self, cfg: DeepFMConfig, params):
        super().__init__()
        self.cfg = cfg
        self.fm_bias = nn.Parameter(params["fm_bias"].clone())
        self.fm_w = nn.Parameter(params["fm_w"].clone())
        self.fm_v = nn.Parameter(params["fm_v"].clone())
        self.mlp = nn.ModuleList()
        d = cfg.deep_input_dim
        for i, h in enumerate(cfg.deep_layers):
            lin = nn.Linear(d, h)
            lin.weight.data.copy_(params["mlp%d.weight" % i]); lin.bias.data.copy_(params["mlp%d.bias" % i])
            self.mlp.append(lin)
            d = h
        self.out = nn.Linear(d, 1)
        self.out.weight.data.copy_(params["deep_out.weight"].reshape(1, -1)); self.out.bias.data.copy_(params["deep_out.bias"])

    def forward(self, ids, vals, train: bool):
        cfg = self.cfg
        y_w = (F.embedding(ids, self.fm_w.unsqueeze(1)).squeeze(-1) * vals).sum(1)
        e = F.embedding(ids, self.fm_v) * vals.unsqueeze(-1)
        y_v = 0.5 * (e.sum(1).pow(2) - e.pow(2).sum(1)).sum(1)
        x = e.reshape(ids.shape[0], -1)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=x.is_cuda):
            for i, lin in enumerate(self.mlp):
                x = F.relu(lin(x))
                if train and cfg.dropout[i] < 1.0:
                    x = F.dropout(x, 1.0 - cfg.dropout[i], True)
            y_d = self.out(x).float().reshape(-1)
        return self.fm_bias + y_w + y_v + y_d


class ProxyDeepFM:
    def __init__(self, cfg: DeepFMConfig, batch_size: int, optimizer, seed: int, device, world_size: int = 1,
                 rank: int = 0):
        self.cfg, self.B, self.dev = cfg, batch_size, device
        self.world, self.rank = world_size, rank
        torch.manual_seed(seed + rank)
        self.net = _Net(cfg, init_params(cfg, seed)).to(device)
        name = optimizer.name
        ps = list(self.net.parameters())
        if name == "Adam":
            self.optim = torch.optim.Adam(ps, lr=optimizer.lr, eps=1e-8)
        elif name == "Adagrad":
            self.optim = torch.optim.Adagrad(ps, lr=optimizer.lr, initial_accumulator_value=1e-8, eps=0.0)
        elif name == "Momentum":
            self.optim = torch.optim.SGD(ps, lr=optimizer.lr, momentum=0.95)
        else:
            self.optim = torch.optim.SGD(ps, lr=optimizer.lr)
        self._loss = torch.zeros((), device=device)
        self.kernels_per_step = 0  # library kernels only; none of ours
        self._step = 0
        self.scalars = torch.zeros(2, device=device)

    def train_step(self, ids, vals, labels, slot_idx: int = 0):
        ids = ids.long()
        self.optim.zero_grad(set_to_none=True)
        y = self.net(ids, vals, True)
        data = F.binary_cross_entropy_with_logits(y, labels.reshape(-1))
        loss = data + 0.5 * self.cfg.l2_reg * (self.net.fm_w.pow(2).sum() + self.net.fm_v.pow(2).sum())
        loss.backward()
        if self.world > 1:  # Horovod-style: average every gradient through one fused buffer
            grads = [p.grad for p in self.net.parameters()]
            flat = torch._utils._flatten_dense_tensors(grads)
            dist.all_reduce(flat)
            flat.div_(self.world)
            for g, f in zip(grads, torch._utils._unflatten_dense_tensors(flat, grads)):
                g.copy_(f)
        self.optim.step()
        self._loss = loss.detach()
        self._step += 1
        return self._loss

    def last_loss(self):
        return self._loss

    @torch.no_grad()
    def predict(self, ids, vals):
        return torch.sigmoid(self.net(ids.long(), vals, False))

    @torch.no_grad()
    def eval_step(self, ids, vals, labels, update_auc=True):
        y = self.net(ids.long(), vals, False)
        return F.binary_cross_entropy_with_logits(y, labels.reshape(-1), reduction="sum").reshape(1), torch.sigmoid(y)

    def l2_penalty(self):
        return 0.5 * self.cfg.l2_reg * (self.net.fm_w.pow(2).sum() + self.net.fm_v.pow(2).sum()).detach()

    @property
    def global_step(self):
        return self._step

    def named_parameters(self):
        return OrderedDict(self.net.named_parameters())

    def state_dict(self):
        return self.net.state_dict()

    def load_state_dict(self, sd, strict=True):
        self.net.load_state_dict(sd, strict)
