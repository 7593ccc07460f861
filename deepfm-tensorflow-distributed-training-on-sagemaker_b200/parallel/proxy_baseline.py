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
                 rank: int = 0, use_graph: bool = False):
        """The best stock-PyTorch recipe we know for this model, so the ratio against it is not a launch-overhead ratio:

          * parameters and gradients are views of ONE flat buffer each (the DDP bucket idea): the gradient exchange is a
            single `dist.all_reduce(flat_grad, AVG)` with no flatten / unflatten copies;
          * the optimizer is `torch.optim.Adam(fused=True)` over the flat parameter: one multi-tensor kernel;
          * the whole-table L2 term is added to the flat gradient directly (`g += l2 * w`, one kernel) instead of being
            differentiated through `pow().sum()`;
          * `use_graph` (or DFM_PROXY_GRAPH=1; `bench.py --impl proxy --proxy_graph`): forward, backward, NCCL all-reduce
            and optimizer are captured in ONE CUDA graph (the PyTorch whole-network capture recipe), which is the fair
            stand-in for a TF-1 graph-mode `session.run(train_op)`. Adam / Momentum / GD (what torch can capture)."""
        import os
        self.cfg, self.B, self.dev = cfg, batch_size, device
        self._use_graph = bool(use_graph or os.environ.get("DFM_PROXY_GRAPH") == "1") and optimizer.name in ("Adam", "GD",
                                                                                                          "Momentum")
        self._graph = None
        self.world, self.rank = world_size, rank
        torch.manual_seed(seed + rank)
        self.net = _Net(cfg, init_params(cfg, seed)).to(device)
        name = optimizer.name
        ps = list(self.net.parameters())
        # ---- one flat parameter buffer + one flat gradient buffer; module parameters / .grad alias slices of them
        al = lambda n: (n + 63) // 64 * 64
        total = sum(al(p.numel()) for p in ps)
        self.flat = torch.zeros(total, device=device)
        self.flat_grad = torch.zeros(total, device=device)
        o = 0
        self._decay = []
        for p in ps:
            n = p.numel()
            self.flat[o:o + n].copy_(p.data.reshape(-1))
            p.data = self.flat[o:o + n].view_as(p)
            p.grad = self.flat_grad[o:o + n].view_as(p)
            if p is self.net.fm_w or p is self.net.fm_v:
                self._decay.append((o, o + n))
            o += al(n)
        self._decay = (min(a for a, _ in self._decay), max(b for _, b in self._decay))
        self.flat_param = nn.Parameter(self.flat)  # shares storage with every module parameter
        self.flat_param.grad = self.flat_grad
        cap = self._use_graph
        if name == "Adam":
            self.optim = torch.optim.Adam([self.flat_param], lr=optimizer.lr, eps=1e-8, fused=True, capturable=cap)
        elif name == "Adagrad":
            self.optim = torch.optim.Adagrad([self.flat_param], lr=optimizer.lr, initial_accumulator_value=1e-8, eps=0.0)
        elif name == "Momentum":
            self.optim = torch.optim.SGD([self.flat_param], lr=optimizer.lr, momentum=0.95)
        else:
            self.optim = torch.optim.SGD([self.flat_param], lr=optimizer.lr)
        self._loss = torch.zeros((), device=device)
        self.kernels_per_step = 0  # library kernels only; none of ours
        self._step = 0
        self.scalars = torch.zeros(2, device=device)

    def _capture(self):
        """Whole-network capture as in the PyTorch CUDA-graphs notes: warm up on a side stream (on a state snapshot that
        is restored afterwards), then record one step over static input buffers."""
        F_ = self.cfg.field_size
        self._sids = torch.zeros(self.B, F_, dtype=torch.long, device=self.dev)
        self._svals = torch.zeros(self.B, F_, device=self.dev)
        self._slabels = torch.zeros(self.B, device=self.dev)
        snap = self.flat.clone()
        side = torch.cuda.Stream(device=self.dev)
        side.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(side):
            for _ in range(3):
                self._eager_step(self._sids, self._svals, self._slabels)
        torch.cuda.current_stream(self.dev).wait_stream(side)
        torch.cuda.synchronize(self.dev)
        self.flat.copy_(snap)
        for st in self.optim.state.values():  # moments back to zero, step counters back to the start
            for k, v in st.items():
                if torch.is_tensor(v):
                    v.zero_()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._graph_loss = self._eager_step(self._sids, self._svals, self._slabels)
        self._graph = g

    def train_step(self, ids, vals, labels, slot_idx: int = 0):
        if self._use_graph:
            if self._graph is None:
                self._capture()
            self._sids.copy_(ids)
            self._svals.copy_(vals)
            self._slabels.copy_(labels.reshape(-1))
            self._graph.replay()
            self._loss = self._graph_loss
            self._step += 1
            return self._loss
        loss = self._eager_step(ids, vals, labels)
        self._step += 1
        return loss

    def _eager_step(self, ids, vals, labels):
        ids = ids.long()
        self.flat_grad.zero_()  # .grad tensors are views of it: autograd accumulates in place
        y = self.net(ids, vals, True)
        data = F.binary_cross_entropy_with_logits(y, labels.reshape(-1))
        data.backward()
        a, b = self._decay
        with torch.no_grad():
            reg = 0.5 * self.cfg.l2_reg * self.flat[a:b].pow(2).sum()
            if self.world > 1:  # Horovod-style: average every gradient through one fused buffer (HVD:262)
                dist.all_reduce(self.flat_grad, op=dist.ReduceOp.AVG)
            if self.cfg.l2_reg:
                self.flat_grad[a:b].add_(self.flat[a:b], alpha=self.cfg.l2_reg)
        self.optim.step()
        self._loss = data.detach() + reg
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
