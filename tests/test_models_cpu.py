"""Oracle, TF-1 optimizers, initialisers, AUC, Philox — CPU."""
import math

import numpy as np
import pytest
import torch

from deepfm_b200.models import DeepFMConfig, TorchDeepFM, init_params
from deepfm_b200.models.layout import FlatLayout
from deepfm_b200.models.oracle import oracle_forward, oracle_grads, oracle_loss
from deepfm_b200.optim import TF1Optimizer
from deepfm_b200.utils.auc import StreamingAUC, exact_auc
from deepfm_b200.utils.philox import dropout_keep_mask, philox4x32_10


def test_philox_known_answers_and_mask_rate():
    r = philox4x32_10(np.zeros(1, np.uint32), 0, 0, 0, 0, 0)
    assert [int(x[0]) for x in r] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    m = dropout_keep_mask(512, 400, 0.5, seed=1, step=3, layer=1)
    assert 0.49 < m.mean() < 0.51
    assert (dropout_keep_mask(16, 400, 0.5, 1, 3, 1, row_offset=100) == dropout_keep_mask(512, 400, 0.5, 1, 3, 1)[100:116]).all()
    assert dropout_keep_mask(4, 8, 1.0, 0, 0, 0).all()


def test_initialisers_match_tf1_statistics():
    cfg = DeepFMConfig(field_size=39, feature_size=117581, embedding_size=32, deep_layers=[400, 400, 400])
    p = init_params(cfg, seed=0)
    std_v = math.sqrt(1.0 / ((117581 + 32) / 2)) / 0.87962566103423978
    assert float(p["fm_v"].abs().max()) <= 2 * std_v * 1.0001  # truncated at 2 sigma
    assert abs(float(p["fm_v"].std()) - std_v * 0.8796) < 0.02 * std_v
    lim = math.sqrt(6.0 / (1248 + 400))
    assert float(p["mlp0.weight"].abs().max()) <= lim and float(p["mlp0.weight"].abs().max()) > 0.95 * lim
    assert float(p["fm_bias"]) == 0.0 and float(p["mlp0.bias"].abs().max()) == 0.0
    assert p["mlp0.weight"].shape == (400, 1248) and p["deep_out.weight"].shape == (400,)
    assert cfg.num_parameters() == 1 + 117581 * 33 + 1248 * 400 + 400 + 2 * (400 * 400 + 400) + 401


def test_oracle_forward_matches_manual_formula():
    cfg = DeepFMConfig(field_size=3, feature_size=10, embedding_size=2, deep_layers=[4], dropout=[1.0], l2_reg=0.1)
    p = init_params(cfg, seed=1)
    p["fm_bias"].fill_(0.3); p["mlp0.bias"].normal_(); p["deep_out.bias"].fill_(-0.2)
    ids = torch.tensor([[1, 2, 2], [0, 9, 5]]); vals = torch.tensor([[1.0, 0.5, 2.0], [0.0, 1.0, 1.0]])
    y = oracle_forward(p, cfg, ids, vals)
    for b in range(2):
        e = [p["fm_v"][ids[b, f]] * vals[b, f] for f in range(3)]
        yw = sum(float(p["fm_w"][ids[b, f]] * vals[b, f]) for f in range(3))
        yv = sum(0.5 * (float(sum(x[k] for x in e)) ** 2 - sum(float(x[k]) ** 2 for x in e)) for k in range(2))
        x = torch.cat(e)
        h = torch.relu(p["mlp0.weight"] @ x + p["mlp0.bias"])
        yd = float(h @ p["deep_out.weight"] + p["deep_out.bias"])
        assert float(y[b]) == pytest.approx(0.3 + yw + yv + yd, rel=1e-5, abs=1e-6)
    loss, data = oracle_loss(p, cfg, y, torch.tensor([1.0, 0.0]))
    reg = 0.05 * (float(p["fm_w"].pow(2).sum()) + float(p["fm_v"].pow(2).sum()))
    assert float(loss - data) == pytest.approx(reg, rel=1e-5)  # only fm_w / fm_v are regularised (PS:244-246)


def test_tf1_optimizer_rules():
    g = torch.tensor([0.5, -2.0])
    # Adam: epsilon outside the bias-corrected sqrt
    w = torch.tensor([1.0, 1.0]); o = TF1Optimizer("Adam", 0.1); sl = o.init_slots({"w": w})["w"]
    o.apply_dense(w, g, sl, 1)
    lr_t = 0.1 * math.sqrt(1 - 0.999) / (1 - 0.9)
    m, v = 0.1 * g, 0.001 * g * g
    assert torch.allclose(w, 1.0 - lr_t * m / (v.sqrt() + 1e-8))
    # Adagrad: accumulator starts at 1e-8, no epsilon
    w = torch.tensor([1.0, 1.0]); o = TF1Optimizer("Adagrad", 0.1); sl = o.init_slots({"w": w})["w"]
    o.apply_dense(w, g, sl, 1)
    assert torch.allclose(w, 1.0 - 0.1 * g / (g * g + 1e-8).sqrt())
    # Momentum 0.95
    w = torch.tensor([1.0, 1.0]); o = TF1Optimizer("Momentum", 0.1); sl = o.init_slots({"w": w})["w"]
    o.apply_dense(w, g, sl, 1); o.apply_dense(w, g, sl, 2)
    assert torch.allclose(w, 1.0 - 0.1 * g - 0.1 * (0.95 * g + g))
    # FTRL: lr_power -0.5, accumulator 0.1
    w = torch.tensor([1.0, 1.0]); o = TF1Optimizer("ftrl", 0.1); sl = o.init_slots({"w": w})["w"]
    o.apply_dense(w, g, sl, 1)
    a_new = 0.1 + g * g
    lin = g - (a_new.sqrt() - math.sqrt(0.1)) / 0.1 * 1.0
    assert torch.allclose(w, -lin / (a_new.sqrt() / 0.1))
    # GD works (missing in the reference)
    w = torch.tensor([1.0, 1.0]); o = TF1Optimizer("GD", 0.1); o.apply_dense(w, g, [], 1)
    assert torch.allclose(w, 1.0 - 0.1 * g)
    # sparse apply segment-sums duplicate ids first
    w = torch.zeros(4, 2); o = TF1Optimizer("Adagrad", 0.1); sl = o.init_slots({"w": w})["w"]
    o.apply_sparse(w, torch.tensor([1, 1, 3]), torch.ones(3, 2), sl, 1)
    assert torch.allclose(w[1], torch.full((2,), -0.1 * 2 / math.sqrt(4 + 1e-8))) and float(w[0].abs().sum()) == 0


def test_streaming_auc_close_to_exact():
    g = torch.Generator().manual_seed(0)
    p = torch.rand(20000, generator=g)
    y = (torch.rand(20000, generator=g) < p).float()
    auc = StreamingAUC()
    for s in range(0, 20000, 3000):
        auc.update(y[s:s + 3000], p[s:s + 3000])
    assert abs(auc.result() - exact_auc(y.numpy(), p.numpy())) < 2e-3  # 200-bucket trapezoid tolerance


def test_layout_ranges_and_views():
    cfg = DeepFMConfig(field_size=39, feature_size=117581, embedding_size=32, deep_layers=[400, 400, 400])
    L = FlatLayout(cfg)
    assert L.decay_range[0] == L.segments["fm_w"].offset and L.decay_range[1] == L.segments["mlp0.weight"].offset
    assert L.shadow_range == (L.segments["mlp0.weight"].offset, L.segments["mlp0.bias"].offset)
    assert all(s.offset % 64 == 0 for s in L.segments.values()) and L.trainable_numel % 4 == 0
    flat = torch.zeros(L.total_numel)
    L.view(flat, "fm_v").fill_(1)
    assert float(flat.sum()) == 117581 * 32


def test_l2_mode_touched_only_regularises_batch_rows():
    cfg = DeepFMConfig(field_size=2, feature_size=10, embedding_size=2, deep_layers=[4], dropout=[1.0], l2_reg=0.5,
                       l2_mode="touched")
    p = init_params(cfg, seed=2)
    ids = torch.tensor([[1, 2]]); vals = torch.zeros(1, 2)  # zero values: no data gradient on the tables
    r = oracle_grads(p, cfg, ids, vals, torch.tensor([1.0]))
    g = r["grads"]["fm_v"]
    assert float(g[[0, 3, 4, 5, 6, 7, 8, 9]].abs().sum()) == 0.0
    assert torch.allclose(g[1], 0.5 * p["fm_v"][1]) and torch.allclose(g[2], 0.5 * p["fm_v"][2])


def test_torch_backend_global_batch_scaling_matches_concat():
    """Two 'ranks' on halves of a batch, gradients summed == one rank on the whole batch."""
    cfg = DeepFMConfig(field_size=5, feature_size=50, embedding_size=4, deep_layers=[8], dropout=[0.5], l2_reg=1e-3)
    p = init_params(cfg, seed=0)
    g = torch.Generator().manual_seed(1)
    ids = torch.randint(0, 50, (8, 5), generator=g); vals = torch.rand(8, 5, generator=g)
    y = (torch.rand(8, generator=g) < 0.5).float()
    full = TorchDeepFM(cfg, 8, TF1Optimizer("GD", 0.1), seed=3, params=p)
    full.train_step(ids, vals, y)
    halves = [TorchDeepFM(cfg, 4, TF1Optimizer("GD", 0.1), seed=3, params=p, global_batch=8, row_offset=4 * r)
              for r in range(2)]
    rs = [h.compute_grads(ids[4 * r:4 * r + 4], vals[4 * r:4 * r + 4], y[4 * r:4 * r + 4]) for r, h in enumerate(halves)]
    for k in rs[0]["grads"]:
        tot = rs[0]["grads"][k] + rs[1]["grads"][k]
        if k in ("fm_w", "fm_v"):
            tot = tot + cfg.l2_reg * p[k]
        want = (p[k] - full.params[k]) / 0.1
        assert torch.allclose(tot, want, rtol=1e-4, atol=1e-6), k


def test_proxy_baseline_is_the_same_model():
    """The NCCL + cuBLAS proxy that stands in for the reference in bench.py must compute the reference's model: its logits
    equal the oracle's on the same parameters, and its first GD step equals the oracle-gradient step (exact L2 mode)."""
    from deepfm_b200.models import DeepFMConfig, init_params
    from deepfm_b200.models.oracle import oracle_forward, oracle_grads
    from deepfm_b200.optim import TF1Optimizer
    from deepfm_b200.parallel.proxy_baseline import ProxyDeepFM
    cfg = DeepFMConfig(field_size=6, feature_size=200, embedding_size=8, deep_layers=[16, 8], dropout=[1.0, 1.0], l2_reg=1e-3)
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(0, 200, (32, 6), generator=g)
    vals = torch.rand(32, 6, generator=g)
    labels = (torch.rand(32, generator=g) < 0.4).float()
    params = init_params(cfg, seed=0)
    m = ProxyDeepFM(cfg, 32, TF1Optimizer("GD", 0.1), 0, torch.device("cpu"))
    with torch.no_grad():
        y = m.net(ids, vals, False)
    assert torch.allclose(y, oracle_forward(params, cfg, ids, vals, False), rtol=1e-5, atol=1e-6)
    ref = oracle_grads({k: v.clone() for k, v in params.items()}, cfg, ids, vals, labels, True, 0, 0, 0)
    loss = m.train_step(ids, vals, labels)
    assert abs(float(loss) - float(ref["loss"])) < 1e-5 and m.global_step == 1
    assert torch.allclose(m.net.fm_v.detach(), params["fm_v"] - 0.1 * ref["grads"]["fm_v"], rtol=1e-4, atol=1e-6)
    assert torch.allclose(m.net.mlp[0].weight.detach(), params["mlp0.weight"] - 0.1 * ref["grads"]["mlp0.weight"],
                          rtol=1e-4, atol=1e-6)


def test_config_validation_and_fused_constraints():
    from deepfm_b200.models import DeepFMConfig
    ok = DeepFMConfig(field_size=39, feature_size=117581, embedding_size=32, deep_layers="400,400,400", dropout="0.5,0.5,0.5")
    assert ok.fused_constraints() == [] and ok.deep_layers == [400, 400, 400] and ok.num_parameters() == 4700975
    assert DeepFMConfig(deep_layers="128,64,32").fused_constraints() == []          # the notebooks' configuration
    for bad in (dict(dropout="0.5,0.0,0.5"), dict(dropout="1.5,1,1"), dict(deep_layers=""), dict(field_size=0),
                dict(l2_mode="lazy"), dict(deep_layers="8,8,8,8"), dict(batch_norm_decay=1.0)):
        with pytest.raises(ValueError):
            DeepFMConfig(**bad)
    odd = DeepFMConfig(field_size=13, feature_size=100, embedding_size=12, deep_layers=[30, 8], dropout=[1.0, 1.0])
    msgs = " ".join(odd.fused_constraints())
    assert "embedding_size" in msgs and "multiples of 8" in msgs and "field_size * embedding_size" in msgs
