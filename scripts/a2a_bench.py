"""BASELINE.json config 5: embedding all-to-all + sparse-update microbenchmark, 1K-1M ids/GPU, fused P2P kernels vs
the NCCL `all_to_all_single` + torch-ops path (the baseline). torchrun; device-timed, max over ranks."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29561")
dist.init_process_group("nccl", device_id=dev, rank=rank, world_size=world)
from deepfm_b200.models import DeepFMConfig
from deepfm_b200.optim import TF1Optimizer
from deepfm_b200.parallel.sharded import ShardedDeepFM

N = int(os.environ.get("A2A_ROWS", 4_000_000)); K = int(os.environ.get("A2A_K", 64)); F = 39

def timeit(fn, n=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / n], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)

results = []
for n_ids in [int(x) for x in os.environ.get("A2A_SIZES", "1024,16384,131072,1048576").split(",")]:
    B = max(1, n_ids // F)
    cfg = DeepFMConfig(field_size=F, feature_size=N, embedding_size=K, deep_layers=[64], dropout=[1.0], l2_reg=0.0,
                       l2_mode="touched")
    m = ShardedDeepFM(cfg, B, TF1Optimizer("Adagrad", 1e-2), 1, dev, world, rank, use_graph=False,
                      a2a_mode=os.environ.get("A2A_MODE", "auto"))
    g = torch.Generator(device=dev).manual_seed(rank + 1)
    s = m.in_slots[0]
    s["ids"].copy_(torch.randint(0, N, (B, F), device=dev, generator=g, dtype=torch.int32))
    s["vals"].copy_(torch.rand(B, F, device=dev, generator=g) + 0.1)
    m.dy.normal_(generator=g); m.dx0.copy_((torch.randn(B, F * K, device=dev, generator=g) * 0.01).bfloat16())
    m._train_mode = True  # include the forward's cross-rank barrier, as inside a training step
    def fused():
        m._emb_forward(s)       # dispatch + gather + scale + pool + combine (P2P loads)
        m._emb_backward(s)      # grad push (remote red.add) + segment-sum + sparse row update on the owner
    t_fused = timeit(fused)
    t_fwd = timeit(lambda: m._emb_forward(s))

    # ---- baseline: NCCL all_to_all_single + torch ops (the "reference's NCCL path" stand-in)
    V = m.V; R = m.rows_local
    acc = torch.full((R, K), 1e-8, device=dev)
    def baseline():
        ids = s["ids"].reshape(-1).long(); vals = s["vals"].reshape(-1)
        owner = ids % world
        order = torch.argsort(owner, stable=True)
        send_ids = (ids[order] // world)
        counts = torch.bincount(owner, minlength=world)
        rcounts = torch.empty_like(counts)
        dist.all_to_all_single(rcounts, counts)
        sc, rc = counts.tolist(), rcounts.tolist()
        recv_ids = torch.empty(sum(rc), dtype=torch.long, device=dev)
        dist.all_to_all_single(recv_ids, send_ids, rc, sc)                 # dispatch ids
        rows = V.index_select(0, recv_ids)                                 # owner-side gather
        back = torch.empty(ids.numel(), K, device=dev)
        dist.all_to_all_single(back, rows, sc, rc)                         # combine rows
        e = torch.empty_like(back); e[order] = back
        e = e * vals[:, None]
        S = e.view(B, F, K).sum(1)                                         # pooling
        # backward: per-lookup grads -> owners
        dE = (m.dx0.float().view(B, F, K) + m.dy[:, None, None] * (S[:, None, :] - e.view(B, F, K))) * s["vals"][..., None]
        gsend = dE.reshape(-1, K)[order]
        grecv = torch.empty(sum(rc), K, device=dev)
        dist.all_to_all_single(grecv, gsend, rc, sc)                       # sparse-gradient all-to-all
        u, inv = torch.unique(recv_ids, return_inverse=True)
        gsum = torch.zeros(u.numel(), K, device=dev).index_add_(0, inv, grecv)   # segment-sum
        a = acc[u] + gsum * gsum                                           # sparse Adagrad on the owner
        acc[u] = a
        V[u] = V[u] - 1e-2 * gsum / a.sqrt()
    t_base = timeit(baseline, n=10)
    bytes_nvl = B * F * K * 4 * 2 * (world - 1) / world   # rows in + grads out, remote fraction
    if rank == 0:
        r = dict(ids_per_gpu=B * F, world=world, rows=N, K=K, mode="push" if m.push else "pull", fused_us=t_fused, fused_fwd_us=t_fwd, nccl_torch_us=t_base,
                 speedup=t_base / t_fused, nvlink_bytes_per_rank=bytes_nvl,
                 roofline_us_at_770GBs=bytes_nvl / 2 / 770e9 * 1e6, frac_of_roofline=(bytes_nvl / 2 / 770e9 * 1e6) / t_fused)
        results.append(r)
        print(json.dumps(r), flush=True)
    del m
    torch.cuda.synchronize(); dist.barrier()
dist.destroy_process_group()
