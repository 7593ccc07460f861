"""Property-based shapes for the sm_100a kernels ('each kernel vs oracle over property-based shapes').

hypothesis draws the shapes (derandomised, so every run sees the same ones); each kernel is compared against the plain
PyTorch fp32 statement of the same op. Shapes stay inside the documented limits: bf16 leading
dimensions are multiples of 8 (TMA 16-byte stride rule), embedding_size in {4, 8, ..., 128}.
"""
import pytest
import torch
from hypothesis import HealthCheck, given, settings, strategies as st

pytestmark = pytest.mark.gpu

SETTINGS = dict(max_examples=10, deadline=None, derandomize=True,
                suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
mult8 = lambda lo, hi: st.integers(lo // 8, hi // 8).map(lambda v: v * 8)


def _rel_err(a, b):
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


@settings(**SETTINGS)
@given(M=st.integers(1, 2100), N=mult8(8, 520), K=mult8(8, 1600), relu=st.booleans())
def test_gemm_fwd_any_shape(cuda_lib, M, N, K, relu):
    from deepfm_b200.ops.kernels import GemmPlan
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K)
    A = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16()
    W = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    guard = torch.full((M + 2, N), 7.0, device="cuda", dtype=torch.bfloat16)  # rows M, M+1 must stay untouched
    out = guard[:M]
    GemmPlan("fwd", A, W, out, M, N, K, bias=bias, relu=relu).run()
    torch.cuda.synchronize()
    ref = A.float() @ W.float().t() + bias
    ref = torch.relu(ref) if relu else ref
    assert _rel_err(out, ref) < 1e-2, (M, N, K)
    assert bool((guard[M:] == 7.0).all()), "wrote past the last row"


@settings(**SETTINGS)
@given(M=st.integers(1, 2100), N=mult8(8, 1300), K=mult8(8, 520))
def test_gemm_dgrad_any_shape(cuda_lib, M, N, K):
    from deepfm_b200.ops.kernels import GemmPlan
    g = torch.Generator(device="cuda").manual_seed(M * 5 + N + K * 11)
    dZ = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16()
    W = (torch.randn(K, N, device="cuda", generator=g) * 0.1).bfloat16()
    mask = torch.relu(torch.randn(M, N, device="cuda", generator=g)).bfloat16()
    guard = torch.full((M + 2, N), 3.0, device="cuda", dtype=torch.bfloat16)
    out = guard[:M]
    GemmPlan("dgrad", dZ, W, out, M, N, K, mask=mask, mask_scale=2.0).run()
    torch.cuda.synchronize()
    ref = (dZ.float() @ W.float()) * (mask.float() != 0).float() * 2.0
    assert _rel_err(out, ref) < 1e-2, (M, N, K)
    assert bool((guard[M:] == 3.0).all()), "wrote past the last row"


@settings(**SETTINGS)
@given(M=mult8(8, 520), N=mult8(8, 1300), K=st.integers(1, 5000), split_k=st.sampled_from([0, 1, 3]))
def test_gemm_wgrad_any_shape(cuda_lib, M, N, K, split_k):
    from deepfm_b200.ops.kernels import GemmPlan
    g = torch.Generator(device="cuda").manual_seed(M + N * 13 + K)
    dZ = (torch.randn(K, M, device="cuda", generator=g) * 0.1).bfloat16()
    X = (torch.randn(K, N, device="cuda", generator=g) * 0.5).bfloat16()
    out = torch.ones(M, N, device="cuda", dtype=torch.float32)
    GemmPlan("wgrad", dZ, X, out, M, N, K, split_k=split_k).run()
    torch.cuda.synchronize()
    ref = 1.0 + dZ.float().t() @ X.float()
    assert _rel_err(out, ref) < 2e-3, (M, N, K, split_k)


@settings(**SETTINGS)
@given(B=st.integers(1, 1500), F=st.integers(1, 48), K=st.sampled_from([4, 8, 16, 32, 64, 128]),
       N=st.integers(2, 200000), dup=st.booleans())
def test_emb_fm_fwd_bwd_any_shape(cuda_lib, B, F, K, N, dup):
    from deepfm_b200.ops import kernels as k
    g = torch.Generator(device="cuda").manual_seed(B * 3 + F * 5 + K + N)
    ids = torch.randint(0, N, (B, F), device="cuda", generator=g, dtype=torch.int64)
    if dup:
        ids[:, 0] = ids[0, 0]
    vals = torch.rand(B, F, device="cuda", generator=g)
    fm_w = torch.randn(N, device="cuda", generator=g) * 0.1
    fm_v = torch.randn(N, K, device="cuda", generator=g) * 0.1
    x0 = torch.zeros(B, F * K, device="cuda", dtype=torch.bfloat16)
    S = torch.zeros(B, K, device="cuda")
    yfm = torch.zeros(B, device="cuda")
    ids32 = ids.int()
    k.emb_fm_fwd(ids32, vals, fm_w, fm_v, x0, S, yfm)
    e = fm_v[ids] * vals.unsqueeze(-1)
    ref_y = (fm_w[ids] * vals).sum(1) + 0.5 * (e.sum(1).pow(2) - e.pow(2).sum(1)).sum(1)
    assert torch.allclose(yfm, ref_y, rtol=1e-4, atol=1e-5)
    assert torch.allclose(S, e.sum(1), rtol=1e-5, atol=1e-6)
    assert torch.allclose(x0.float(), e.reshape(B, F * K), rtol=1e-2, atol=1e-4)
    dy = torch.randn(B, device="cuda", generator=g) / B
    dx0 = (torch.randn(B, F * K, device="cuda", generator=g) * 0.01).bfloat16()
    g_w = torch.zeros(N, device="cuda")
    g_v = torch.zeros(N, K, device="cuda")
    k.emb_bwd(ids32, vals, fm_v, S, dy, dx0, g_w, g_v)
    # closed form of the gradient: dE = dX0 + dy * (S - e); g_v[id] += dE * val; g_w[id] += dy * val
    dE = dx0.float().reshape(B, F, K) + dy[:, None, None] * (e.sum(1, keepdim=True) - e)
    ref_gv = torch.zeros(N, K, device="cuda").index_add_(0, ids.reshape(-1), (dE * vals.unsqueeze(-1)).reshape(-1, K))
    ref_gw = torch.zeros(N, device="cuda").index_add_(0, ids.reshape(-1), (dy[:, None] * vals).reshape(-1))
    assert torch.allclose(g_w, ref_gw, rtol=1e-3, atol=1e-6)
    assert torch.allclose(g_v, ref_gv, rtol=1e-3, atol=1e-6)


@settings(**SETTINGS)
@given(B=st.integers(1, 3000), H=mult8(8, 512))
def test_logit_loss_any_shape(cuda_lib, B, H):
    from deepfm_b200.ops import kernels as k
    g = torch.Generator(device="cuda").manual_seed(B + H)
    x = torch.relu(torch.randn(B, H, device="cuda", generator=g)).bfloat16()
    w_out = torch.randn(H, device="cuda", generator=g) * 0.1
    b_out = torch.randn(1, device="cuda", generator=g)
    fm_bias = torch.randn(1, device="cuda", generator=g)
    yfm = torch.randn(B, device="cuda", generator=g)
    labels = (torch.rand(B, device="cuda", generator=g) < 0.3).float()
    prob = torch.empty(B, device="cuda")
    logit = torch.empty(B, device="cuda")
    dy = torch.empty(B, device="cuda")
    dz = torch.empty(B, H, device="cuda", dtype=torch.bfloat16)
    loss_sum = torch.zeros(1, device="cuda")
    g_b = torch.zeros(1, device="cuda")
    g_f = torch.zeros(1, device="cuda")
    k.logit_loss(x, w_out, b_out, fm_bias, yfm, labels, 1.0 / B, 2.0, True, prob, logit, dy, dz, loss_sum, g_b, g_f)
    torch.cuda.synchronize()
    xr = x.float()
    y = xr @ w_out + b_out + fm_bias + yfm
    ref_dy = (torch.sigmoid(y) - labels) / B
    ref_loss = torch.nn.functional.binary_cross_entropy_with_logits(y, labels, reduction="sum")
    assert torch.allclose(logit, y, rtol=1e-4, atol=1e-4)
    assert torch.allclose(prob, torch.sigmoid(y), rtol=1e-3, atol=1e-5)
    assert torch.allclose(loss_sum[0], ref_loss, rtol=1e-4, atol=1e-5)
    assert torch.allclose(dy, ref_dy, rtol=1e-3, atol=1e-7)
    assert _rel_err(dz, ref_dy[:, None] * w_out[None, :] * (xr != 0).float() * 2.0) < 1e-2
    assert torch.allclose(g_b[0], ref_dy.sum(), rtol=1e-3, atol=1e-6)
    assert torch.allclose(g_f[0], ref_dy.sum(), rtol=1e-3, atol=1e-6)
