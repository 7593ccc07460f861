"""sm_100a kernel numerics vs plain PyTorch fp32 references of the same op (SURVEY section 4 'Kernel' tests)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel_err(a, b):
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


GEMM_SHAPES = [
    # (M, N, K)
    (1024, 400, 1248),
    (1024, 400, 400),
    (256, 128, 1248),
    (100, 32, 64),
    (1024, 64, 128),
    (4096, 256, 512),
    (130, 80, 72),
]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_fwd(cuda_lib, M, N, K):
    from deepfm_b200.ops.kernels import GemmPlan
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    A = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16()
    W = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    out = torch.full((M, N), 7.0, device="cuda", dtype=torch.bfloat16)
    plan = GemmPlan("fwd", A, W, out, M, N, K, bias=bias, relu=True)
    plan.run()
    torch.cuda.synchronize()
    ref = torch.relu(A.float() @ W.float().t() + bias)
    err = _rel_err(out, ref)
    assert err < 1e-2, (err, plan.info())


def test_gemm_fwd_dropout_matches_philox(cuda_lib):
    from deepfm_b200.ops.kernels import GemmPlan
    from deepfm_b200.utils.philox import dropout_keep_mask
    M, N, K = 300, 400, 128
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    bias = torch.zeros(N, device="cuda")
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    step = torch.tensor([5], device="cuda", dtype=torch.int64)
    plan = GemmPlan("fwd", A, W, out, M, N, K, bias=bias, relu=False, keep_prob=0.5, seed=1234, layer=2,
                    row_offset=7, step_ptr=step)
    plan.run()
    torch.cuda.synchronize()
    keep = torch.from_numpy(dropout_keep_mask(M, N, 0.5, 1234, 5, 2, row_offset=7)).cuda()
    ref = torch.where(keep, (A.float() @ W.float().t()) * 2.0, torch.zeros((), device="cuda"))
    assert _rel_err(out, ref) < 1e-2
    assert 0.45 < keep.float().mean().item() < 0.55


@pytest.mark.parametrize("M,K0,widths,keeps", [
    (1024, 1248, [400, 400, 400], [0.5, 0.8, 1.0]),   # the headline shape: cluster of 5 x BN 80
    (300, 1248, [400, 400, 400], [0.5, 0.5, 0.5]),    # partial last row block
    (256, 312, [128, 64, 32], [0.9, 0.9, 0.9]),       # the reference's default widths: cluster of 2 x BN 64, idle CTAs
    (200, 72, [256], [1.0]),                          # one layer, K0 not a multiple of 64
    (130, 640, [448, 16, 400, 80], [0.7, 1.0, 0.6, 1.0]),  # four layers, widest allowed, narrowest allowed
])
def test_fused_mlp_forward_equals_the_per_layer_gemm_chain(cuda_lib, M, K0, widths, keeps):
    """csrc/mlp_fused.cu (one cluster kernel for all layers) against the chain of per-layer tcgen05 GEMM plans it replaces:
    same bf16 roundings at the same places and the same Philox streams, so every activation must be IDENTICAL; and both
    against the fp32 reference with the host-side Philox mask."""
    from deepfm_b200.ops.kernels import GemmPlan, MlpFwdPlan
    from deepfm_b200.utils.philox import dropout_keep_mask
    g = torch.Generator(device="cuda").manual_seed(M + K0 + len(widths))
    def buf(rows, width):  # padded rows with a ones column, like models/fused.py act_buf
        ld = (width + 1 + 63) // 64 * 64
        b = torch.full((rows, ld), 7.0, device="cuda", dtype=torch.bfloat16)
        return b, b[:, :width]
    xb, x = buf(M, K0)
    x.copy_((torch.randn(M, K0, device="cuda", generator=g) * 0.5).bfloat16())
    Ws, bs, d_in = [], [], K0
    for h in widths:
        Ws.append((torch.randn(h, d_in, device="cuda", generator=g) * (1.0 / d_in ** 0.5)).bfloat16())
        bs.append(torch.randn(h, device="cuda", generator=g) * 0.1)
        d_in = h
    step = torch.tensor([11], device="cuda", dtype=torch.int64)
    seed, row_offset = 4321, 5
    outs_a = [buf(M, h) for h in widths]
    outs_b = [buf(M, h) for h in widths]
    cur = x
    for l, h in enumerate(widths):
        GemmPlan("fwd", cur, Ws[l], outs_a[l][1], M, h, cur.shape[1], bias=bs[l], relu=True, keep_prob=keeps[l], seed=seed,
                 layer=l, row_offset=row_offset, step_ptr=step).run()
        cur = outs_a[l][1]
    plan = MlpFwdPlan(x, Ws, [o[1] for o in outs_b], bs, M, relu=True, keep_probs=keeps, seed=seed, row_offset=row_offset,
                      step_ptr=step)
    plan.run()
    torch.cuda.synchronize()
    ref = x.float()
    for l, h in enumerate(widths):
        ref = torch.relu(ref @ Ws[l].float().t() + bs[l])
        if keeps[l] < 1.0:
            keep = torch.from_numpy(dropout_keep_mask(M, h, keeps[l], seed, 11, l, row_offset=row_offset)).cuda()
            ref = torch.where(keep, ref / keeps[l], torch.zeros((), device="cuda"))
        got = outs_b[l][1].float()
        assert torch.equal(outs_b[l][0], outs_a[l][0]), (l, plan.info(), float((got - outs_a[l][1].float()).abs().max()))
        assert _rel_err(got, ref) < 2e-2, (l, _rel_err(got, ref), plan.info())
        assert bool((outs_b[l][0][:, h:] == 7.0).all())  # padding / ones column untouched
        ref = got  # next layer's reference starts from the rounded activations, as the kernels do


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_dgrad(cuda_lib, M, N, K):
    """out[M,N] = (dZ[M,K] @ W[K,N]) * (mask != 0) * scale — B operand is N-major (no transpose materialised)."""
    from deepfm_b200.ops.kernels import GemmPlan
    if N % 8 or K % 8:
        pytest.skip("TMA stride rule")
    g = torch.Generator(device="cuda").manual_seed(M * 3 + N + K)
    dZ = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16()
    W = (torch.randn(K, N, device="cuda", generator=g) * 0.1).bfloat16()
    mask = torch.relu(torch.randn(M, N, device="cuda", generator=g)).bfloat16()
    out = torch.full((M, N), 3.0, device="cuda", dtype=torch.bfloat16)
    plan = GemmPlan("dgrad", dZ, W, out, M, N, K, mask=mask, mask_scale=2.0)
    plan.run()
    torch.cuda.synchronize()
    ref = (dZ.float() @ W.float()) * (mask.float() != 0).float() * 2.0
    err = _rel_err(out, ref)
    assert err < 1e-2, (err, plan.info())
    out2 = torch.empty_like(out)
    GemmPlan("dgrad", dZ, W, out2, M, N, K).run()
    torch.cuda.synchronize()
    assert _rel_err(out2, dZ.float() @ W.float()) < 1e-2


@pytest.mark.parametrize("M,N,K", [(400, 1248, 1024), (400, 400, 1024), (128, 64, 256), (32, 1248, 100 * 8),
                                   (64, 128, 4096), (400, 1248, 16384)])
@pytest.mark.parametrize("split_k", [0, 1])
def test_gemm_wgrad(cuda_lib, M, N, K, split_k):
    """dW[M=out,N=in] += dZ[K=batch,M]^T @ X[K=batch,N] — both operands MN-major, fp32 atomic accumulate."""
    from deepfm_b200.ops.kernels import GemmPlan
    g = torch.Generator(device="cuda").manual_seed(M + N * 5 + K)
    dZ = (torch.randn(K, M, device="cuda", generator=g) * 0.1).bfloat16()
    X = (torch.randn(K, N, device="cuda", generator=g) * 0.5).bfloat16()
    out = torch.ones(M, N, device="cuda", dtype=torch.float32)
    plan = GemmPlan("wgrad", dZ, X, out, M, N, K, split_k=split_k)
    plan.run()
    torch.cuda.synchronize()
    ref = 1.0 + dZ.float().t() @ X.float()
    err = _rel_err(out, ref)
    assert err < 2e-3, (err, plan.info())


@pytest.mark.parametrize("B,F,K,N", [(1024, 39, 32, 117581), (77, 39, 64, 5000), (512, 13, 16, 1000), (33, 5, 128, 64),
                                     (64, 39, 8, 100), (64, 3, 4, 50)])
def test_emb_fm_fwd_bwd(cuda_lib, B, F, K, N):
    from deepfm_b200.ops import kernels as k
    g = torch.Generator(device="cuda").manual_seed(B + F + K)
    ids = torch.randint(0, N, (B, F), device="cuda", generator=g, dtype=torch.int64)
    ids[:, 0] = ids[0, 0]  # heavy duplicate id
    vals = torch.rand(B, F, device="cuda", generator=g)
    vals[:, -1] = 0.0  # padding-style zero values
    fm_w = torch.randn(N, device="cuda", generator=g) * 0.1
    fm_v = torch.randn(N, K, device="cuda", generator=g) * 0.1
    x0 = torch.zeros(B, F * K, device="cuda", dtype=torch.bfloat16)
    S = torch.zeros(B, K, device="cuda")
    yfm = torch.zeros(B, device="cuda")
    ids32 = ids.int()
    k.emb_fm_fwd(ids32, vals, fm_w, fm_v, x0, S, yfm)
    torch.cuda.synchronize()
    e = fm_v[ids] * vals.unsqueeze(-1)
    ref_y = (fm_w[ids] * vals).sum(1) + 0.5 * (e.sum(1).pow(2) - e.pow(2).sum(1)).sum(1)
    assert torch.allclose(yfm, ref_y, rtol=1e-4, atol=1e-5), (yfm - ref_y).abs().max()
    assert torch.allclose(S, e.sum(1), rtol=1e-5, atol=1e-6)
    assert torch.allclose(x0.float(), e.reshape(B, F * K), rtol=1e-2, atol=1e-4)
    # backward
    dy = torch.randn(B, device="cuda", generator=g) / B
    dx0 = (torch.randn(B, F * K, device="cuda", generator=g) * 0.01).bfloat16()
    g_w = torch.zeros(N, device="cuda")
    g_v = torch.zeros(N, K, device="cuda")
    k.emb_bwd(ids32, vals, fm_v, S, dy, dx0, g_w, g_v)
    torch.cuda.synchronize()
    fm_w_r = fm_w.clone().requires_grad_(True)
    fm_v_r = fm_v.clone().requires_grad_(True)
    e_r = fm_v_r[ids] * vals.unsqueeze(-1)
    y_r = (fm_w_r[ids] * vals).sum(1) + 0.5 * (e_r.sum(1).pow(2) - e_r.pow(2).sum(1)).sum(1)
    total = (y_r * dy).sum() + (e_r.reshape(B, F * K) * dx0.float()).sum()
    total.backward()
    assert torch.allclose(g_w, fm_w_r.grad, rtol=1e-4, atol=1e-6), (g_w - fm_w_r.grad).abs().max()
    assert torch.allclose(g_v, fm_v_r.grad, rtol=1e-3, atol=1e-6), (g_v - fm_v_r.grad).abs().max()


@pytest.mark.parametrize("B,H", [(1024, 400), (37, 32), (256, 64), (100, 33 * 2)])
def test_logit_loss(cuda_lib, B, H):
    from deepfm_b200.ops import kernels as k
    g = torch.Generator(device="cuda").manual_seed(B + H)
    x = torch.relu(torch.randn(B, H, device="cuda", generator=g)).bfloat16()
    w = torch.randn(H, device="cuda", generator=g) * 0.1
    b_out = torch.randn(1, device="cuda", generator=g)
    fm_bias = torch.randn(1, device="cuda", generator=g)
    yfm = torch.randn(B, device="cuda", generator=g)
    labels = (torch.rand(B, device="cuda", generator=g) > 0.7).float()
    prob = torch.empty(B, device="cuda"); logit = torch.empty(B, device="cuda"); dy = torch.empty(B, device="cuda")
    dz = torch.empty(B, H, device="cuda", dtype=torch.bfloat16)
    loss = torch.zeros(1, device="cuda"); gw = torch.zeros(H, device="cuda"); gb = torch.zeros(1, device="cuda")
    gfb = torch.zeros(1, device="cuda"); gbl = torch.zeros(H, device="cuda")
    k.logit_loss(x, w, b_out, fm_bias, yfm, labels, 1.0 / B, 2.0, True, prob, logit, dy, dz, loss, gb, gfb)
    k.colsum_bf16(x, gw, dy)
    k.colsum_bf16(dz, gbl)
    torch.cuda.synchronize()
    xr = x.float()
    y = fm_bias + b_out + yfm + xr @ w
    ref_loss = torch.nn.functional.binary_cross_entropy_with_logits(y, labels, reduction="sum")
    ref_dy = (torch.sigmoid(y) - labels) / B
    assert torch.allclose(logit, y, rtol=1e-4, atol=1e-4)
    assert torch.allclose(prob, torch.sigmoid(y), rtol=1e-3, atol=1e-5)
    assert torch.allclose(loss[0], ref_loss, rtol=1e-4)
    assert torch.allclose(dy, ref_dy, rtol=1e-3, atol=1e-7)
    ref_dz = ref_dy[:, None] * w[None, :] * (xr != 0).float() * 2.0
    assert _rel_err(dz, ref_dz) < 1e-2
    assert torch.allclose(gw, xr.t() @ ref_dy, rtol=1e-3, atol=1e-6)
    assert torch.allclose(gb[0], ref_dy.sum(), rtol=1e-3, atol=1e-6)
    assert torch.allclose(gfb[0], ref_dy.sum(), rtol=1e-3, atol=1e-6)
    assert torch.allclose(gbl, dz.float().sum(0), rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("opt", ["Adam", "Adagrad", "Momentum", "ftrl", "GD"])
def test_flat_opt_matches_tf1_reference(cuda_lib, opt):
    from deepfm_b200.ops import kernels as k
    from deepfm_b200.optim import TF1Optimizer, SLOT_INIT
    n = 4096 * 3 + 64
    g = torch.Generator(device="cuda").manual_seed(11)
    w = torch.randn(n, device="cuda", generator=g)
    w_ref = w.clone()
    ref = TF1Optimizer(opt, 1e-2)
    inits = SLOT_INIT[ref.name] + [0.0, 0.0]
    s0 = torch.full((n,), inits[0], device="cuda"); s1 = torch.full((n,), inits[1], device="cuda")
    slots_ref = [torch.full((n,), v, device="cuda") for v in SLOT_INIT[ref.name]]
    shadow = torch.zeros(1024, device="cuda", dtype=torch.bfloat16)
    step = torch.zeros(1, device="cuda", dtype=torch.int64)
    reg = torch.zeros(1, device="cuda"); done = torch.zeros(1, device="cuda", dtype=torch.int32)
    l2 = 1e-3
    for it in range(3):
        grad = torch.randn(n, device="cuda", generator=g)
        gbuf = grad.clone()
        reg.zero_()
        ref_reg = 0.5 * l2 * w_ref[64:2048].pow(2).sum()
        k.flat_opt_step(w, gbuf, s0, s1, ref.name, 1e-2, 0.9, 0.999, 1e-8, 0.95, 1.0, (64, 2048), l2, (2048, 3072),
                        shadow, step, reg, done)
        torch.cuda.synchronize()
        g_ref = grad.clone()
        g_ref[64:2048] += l2 * w_ref[64:2048]
        ref.apply_dense(w_ref, g_ref, slots_ref, it + 1)
        assert int(step.item()) == it + 1
        assert float(gbuf.abs().max()) == 0.0
        assert torch.allclose(w, w_ref, rtol=2e-4, atol=2e-6), (opt, it, (w - w_ref).abs().max())
        assert torch.allclose(reg[0], ref_reg, rtol=1e-4)
        assert torch.allclose(shadow.float(), w[2048:3072], rtol=1e-2, atol=1e-3)


def test_auc_hist_matches_streaming_auc(cuda_lib):
    from deepfm_b200.ops import kernels as k
    from deepfm_b200.utils.auc import StreamingAUC, thresholds
    B = 10000
    prob = torch.rand(B, device="cuda")
    prob[:5] = torch.tensor([0.0, 1.0, 0.5, 1.0 / 199, 198.0 / 199])
    labels = (torch.rand(B, device="cuda") < prob).float()
    thr = torch.from_numpy(thresholds()).cuda()
    hist = torch.zeros(2, 201, device="cuda")
    k.auc_hist(prob, labels, thr, hist)
    torch.cuda.synchronize()
    ref = StreamingAUC()
    ref.update(labels, prob)
    assert torch.equal(hist.cpu().double(), ref.hist)


def test_example_parse_gpu_matches_host_parser(cuda_lib, tmp_path):
    """GPU protobuf parser == C++/Python host parsers on real loader slots (packed + unpacked encodings, bad record)."""
    import struct
    import numpy as np
    from deepfm_b200.data import native
    from deepfm_b200.ops import kernels as k
    from deepfm_b200.utils import tfrecord as T
    F, B = 39, 256
    rng = np.random.default_rng(0)
    path = str(tmp_path / "tr.tfrecords")
    recs = []
    with native.NativeRecordWriter(path) as w:
        for i in range(B):
            ex = T.make_ctr_example(float(i % 2), rng.integers(0, 2 ** 31 - 1, F).tolist(), rng.random(F).astype(np.float32).tolist())
            recs.append(ex); w.write(ex)
    ld = native.NativeRawLoader([path], B, pin_memory=True)
    slot, rows, raw = ld.next()
    d = raw.cuda()
    off = ((B + 1) * 4 + 255) & ~255
    labels = torch.empty(B, device="cuda"); ids = torch.empty(B, F, device="cuda", dtype=torch.int32)
    vals = torch.empty(B, F, device="cuda"); err = torch.zeros(1, device="cuda", dtype=torch.int32)
    k.example_parse(d[off:], d[: (B + 1) * 4].view(torch.int32), B, F, labels, ids, vals, err)
    torch.cuda.synchronize()
    l, i, v = T.parse_ctr_batch(recs, F)
    assert int(err) == 0
    assert np.array_equal(labels.cpu().numpy(), l) and np.array_equal(ids.cpu().numpy(), i.astype(np.int32))
    assert np.array_equal(vals.cpu().numpy(), v)
    ld.release(slot); ld.close()
    # unpacked repeated fields + a record with the wrong number of ids
    unpacked = T._ld(1, T._ld(1, T._ld(1, b"values") + T._ld(2, T._ld(2, b"\x0d" + struct.pack("<f", 2.0) + b"\x0d" + struct.pack("<f", 3.0)))) +
                     T._ld(1, T._ld(1, b"label") + T._ld(2, T._ld(2, b"\x0d" + struct.pack("<f", 1.0)))) +
                     T._ld(1, T._ld(1, b"ids") + T._ld(2, T._ld(3, b"\x08\x05\x08\x07"))))
    bad = T.make_ctr_example(1.0, [1, 2, 3], [0.1, 0.2, 0.3])
    blob = torch.frombuffer(bytearray(unpacked + bad), dtype=torch.uint8).cuda()
    offs = torch.tensor([0, len(unpacked), len(unpacked) + len(bad)], dtype=torch.int32, device="cuda")
    labels = torch.empty(2, device="cuda"); ids = torch.empty(2, 2, device="cuda", dtype=torch.int32)
    vals = torch.empty(2, 2, device="cuda"); err = torch.zeros(1, device="cuda", dtype=torch.int32)
    k.example_parse(blob, offs, 2, 2, labels, ids, vals, err)
    torch.cuda.synchronize()
    assert ids[0].tolist() == [5, 7] and vals[0].tolist() == [2.0, 3.0] and float(labels[0]) == 1.0
    assert int(err) == 2 and ids[1].tolist() == [0, 0]


@pytest.mark.parametrize("B,F,K,N", [(1024, 39, 32, 117581), (200, 13, 16, 1000), (77, 5, 64, 37)])
def test_deterministic_emb_bwd_matches_atomic_version_and_is_bit_reproducible(cuda_lib, B, F, K, N):
    """csrc/embed_det.cu (sorted segment-sum, plain stores) vs csrc/embed.cu (fp32 atomics): same gradients up to summation
    order; two runs of the deterministic version are bit-identical; hot ids (runs spanning several CTAs), runs ending
    exactly at CTA boundaries, out-of-range ids and zero-valued (padding) lookups are handled."""
    import torch
    from deepfm_b200.ops import kernels as K_
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(B + F)
    ids = torch.randint(0, N, (B, F), generator=g, dtype=torch.int32)
    ids[:, 0] = 3                      # one row in every sample: a run of B lookups, many CTAs long
    ids[: B // 2, 1] = N - 1           # last table row, half the samples
    if F > 3:
        ids[::7, 2] = N + 5            # out of range
        ids[::5, 3] = -1
    vals = torch.rand(B, F, generator=g)
    vals[::11, F - 1] = 0.0            # padding lookups
    fm_v = torch.randn(N, K, generator=g) * 0.1
    S = torch.randn(B, K, generator=g)
    dy = torch.randn(B, generator=g) * 0.01
    dx0 = (torch.randn(B, F * K, generator=g) * 0.01).to(torch.bfloat16)
    ids, vals, fm_v, S, dy, dx0 = [t.to(dev) for t in (ids, vals, fm_v, S, dy, dx0)]
    gw_a, gv_a = torch.zeros(N, device=dev), torch.zeros(N, K, device=dev)
    K_.emb_bwd(ids, vals, fm_v, S, dy, dx0, gw_a, gv_a)
    det = K_.DetEmbBwd(B, F, K, N, dev)
    outs = []
    for _ in range(2):
        gw_d, gv_d = torch.full((N,), 7.0, device=dev), torch.full((N, K), 7.0, device=dev)  # stale contents: overwritten
        det.sort(ids, vals)
        det.run(vals, fm_v, S, dy, dx0, gw_d, gv_d)
        torch.cuda.synchronize()
        outs.append((gw_d, gv_d))
    valid = (ids >= 0) & (ids < N) & (vals != 0)
    touched = torch.zeros(N, dtype=torch.bool, device=dev)
    touched[ids[valid].long()] = True
    gw_d, gv_d = outs[0]
    assert torch.equal(gw_d[~touched], torch.full_like(gw_d[~touched], 7.0))  # untouched rows are not written
    scale = float(gv_a.abs().max()) + 1e-12
    assert float((gv_d[touched] - gv_a[touched]).abs().max()) / scale < 1e-5
    assert float((gw_d[touched] - gw_a[touched]).abs().max()) / (float(gw_a.abs().max()) + 1e-12) < 1e-5
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


def test_deterministic_mode_training_step_matches_default(cuda_lib, monkeypatch):
    """DFM_DETERMINISTIC=1 swaps the embedding backward inside the captured step; gradients agree with the atomic path."""
    import torch
    from deepfm_b200.models import DeepFMConfig, init_params
    from deepfm_b200.models.fused import FusedDeepFM
    cfg = DeepFMConfig(field_size=39, feature_size=5000, embedding_size=32, deep_layers=[400, 400, 400],
                       dropout=[0.5, 0.8, 1.0], l2_reg=0.0)
    B = 512
    params = init_params(cfg, seed=3)
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(0, cfg.feature_size, (B, cfg.field_size), generator=g, dtype=torch.int64)
    ids[:, 0] = 3
    vals = torch.rand(B, cfg.field_size, generator=g)
    labels = (torch.rand(B, generator=g) < 0.3).float()
    grads = []
    for det in ("0", "1", "1"):
        monkeypatch.setenv("DFM_DETERMINISTIC", det)
        m = FusedDeepFM(cfg, B, seed=77, params={k: v.clone() for k, v in params.items()}, use_graph=False)
        assert (m.det_emb is not None) == (det == "1")
        grads.append(m.compute_grads(ids.cuda(), vals.cuda(), labels.cuda()))
    for k in ("fm_v", "fm_w"):
        ref = grads[0][k]
        assert float((grads[1][k] - ref).abs().max()) / (float(ref.abs().max()) + 1e-12) < 1e-5, k
        assert torch.equal(grads[1][k], grads[2][k]), k  # bit-reproducible across engines / runs
    for k in grads[0]:
        if k.endswith("weight") and k.startswith("mlp"):
            assert torch.equal(grads[1][k], grads[2][k]), k  # store-based wgrads are deterministic too


def test_watchdog_record_is_readable_from_the_host(cuda_lib):
    """The flag-wait watchdog reports through 8 words of pinned host memory (csrc/error.cu) instead of a device printf:
    the record must exist, read as "nothing happened", and turn into the message once a kernel has filled it in (here the
    test plays the kernel and writes the words the way wait_flag_geq does)."""
    import ctypes
    from deepfm_b200.ops import lib as nlib
    lib = nlib.load()
    lib.dfm_watchdog_record.restype = ctypes.POINTER(ctypes.c_uint)
    rec = lib.dfm_watchdog_record()
    assert bool(rec)
    assert nlib.watchdog_message() is None
    rec[1], rec[2], rec[3], rec[4] = 5, 1, 4242, 4241
    rec[0] = 1
    try:
        msg = nlib.watchdog_message()
        assert msg is not None and "rank 5" in msg and "barrier 1" in msg and "4242" in msg and "4241" in msg
    finally:
        rec[0] = 0
    assert nlib.watchdog_message() is None
