"""Multi-GPU check (torchrun): row-sharded embeddings with fused a2a kernels == the oracle with touched-row (lazy)
table updates on the concatenated batch; then device-timed us/step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29551")
dist.init_process_group("nccl", device_id=dev, rank=rank, world_size=world)
from deepfm_b200.models import DeepFMConfig, init_params
from deepfm_b200.models.oracle import oracle_grads
from deepfm_b200.optim import TF1Optimizer
from deepfm_b200.parallel.sharded import ShardedDeepFM

ok = True
def check(c, msg):
    global ok
    if not c:
        ok = False
        print("[rank %d] FAIL %s" % (rank, msg), flush=True)

def batch(cfg, B, seed):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, cfg.feature_size, (B, cfg.field_size), generator=g, dtype=torch.int64)
    ids[:, 0] = 3; ids[:, 1] = torch.randint(0, 7, (B,), generator=g)  # hot ids: cross-rank duplicates
    vals = torch.rand(B, cfg.field_size, generator=g) + 0.1
    labels = (torch.rand(B, generator=g) < 0.3).float()
    return ids, vals, labels

for K, opt, mode in [(32, "Adagrad", "pull"), (64, "Adam", "pull"), (32, "GD", "pull"), (64, "GD", "push"),
                     (32, "Adam", "push")]:
    cfg = DeepFMConfig(field_size=39, feature_size=5003, embedding_size=K, deep_layers=[400, 400, 400],
                       dropout=[0.5, 0.8, 1.0], l2_reg=1e-4, l2_mode="touched")
    Bl = 256; GB = Bl * world
    params = init_params(cfg, seed=3)
    m = ShardedDeepFM(cfg, Bl, TF1Optimizer(opt, 1e-2), 77, dev, world, rank, params=params, a2a_mode=mode)
    ref_p = {k: v.clone().to(dev) for k, v in params.items()}
    ref_opt = TF1Optimizer(opt, 1e-2)
    ref_slots = ref_opt.init_slots(ref_p)
    for step in range(3):
        ids, vals, labels = batch(cfg, GB, step)
        sl = slice(rank * Bl, (rank + 1) * Bl)
        m.train_step(ids[sl].to(dev), vals[sl].to(dev), labels[sl].to(dev))
        r = oracle_grads(ref_p, cfg, ids.to(dev), vals.to(dev), labels.to(dev), True, 77, step, 0, compute_dtype=torch.bfloat16)
        t = ref_opt.step + 1
        touched = torch.unique(ids.to(dev).reshape(-1))
        for k, g in r["grads"].items():
            if k in ("fm_w", "fm_v"):
                ref_opt.apply_sparse(ref_p[k], touched, g[touched], ref_slots[k], t)
            else:
                ref_opt.apply_dense(ref_p[k], g, ref_slots[k], t)
        ref_opt.step = t
    torch.cuda.synchronize(); dist.barrier()
    got = m.named_parameters()
    worst = 0.0
    for k in ref_p:
        a, b = got[k].float(), ref_p[k].float()
        scale = float(b.abs().max()) + 1e-8
        if opt == "GD":
            err = float((a - b).abs().max()) / scale; worst = max(worst, err)
            check(err < 3e-2, "K%d %s %s %s err %.3e (nan: got %d, oracle %d)" % (
                K, opt, mode, k, err, int(torch.isnan(a).sum()), int(torch.isnan(b).sum())))
        else:
            frac = float(((a - b).abs() > 3e-2 * scale).float().mean()); worst = max(worst, frac)
            check(frac < 0.20, "K%d %s %s %s frac_bad %.3f" % (K, opt, mode, k, frac))
    # untouched rows must be bit-identical to the initial table (lazy update)
    all_ids = torch.cat([batch(cfg, GB, s)[0].reshape(-1) for s in range(3)]).unique().to(dev)
    untouched = torch.ones(cfg.feature_size, dtype=torch.bool, device=dev); untouched[all_ids] = False
    check(bool(torch.equal(got["fm_v"][untouched], params["fm_v"].to(dev)[untouched])), "untouched rows changed")
    # timing
    ids, vals, labels = batch(cfg, GB, 50); sl = slice(rank * Bl, (rank + 1) * Bl)
    m.load_inputs(ids[sl].to(dev), vals[sl].to(dev), labels[sl].to(dev), 0)
    for _ in range(10): m.train_step_slot(0)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(100): m.train_step_slot(0)
    e1.record(); torch.cuda.synchronize()
    tt = torch.tensor([e0.elapsed_time(e1) * 10.0], device=dev); dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    if rank == 0:
        print("sharded K=%d opt=%-7s a2a=%s world=%d worst=%.3e us/step=%.1f loss=%.4f (%s)" % (K, opt, mode, world, worst, float(tt), float(m.last_loss()), m.comm_name), flush=True)
    del m
    torch.cuda.synchronize(); dist.barrier()
# ---- batch_norm + Adam: the start / resume broadcast (BN layouts have non-trainable statistics behind the trainable block;
# the slot broadcast used the padded length) and a few steps
bn_cfg = DeepFMConfig(field_size=39, feature_size=5003, embedding_size=32, deep_layers=[128, 64], dropout=[0.8, 1.0],
                      l2_reg=1e-4, l2_mode="touched", batch_norm=True)
for mode in ("pull", "push"):
    m = ShardedDeepFM(bn_cfg, 256, TF1Optimizer("Adam", 1e-3), 77, dev, world, rank, params=init_params(bn_cfg, seed=3),
                      a2a_mode=mode)
    m.broadcast_state(0)
    for step in range(3):
        ids, vals, labels = batch(bn_cfg, 256 * world, step)
        sl = slice(rank * 256, (rank + 1) * 256)
        m.train_step(ids[sl].to(dev), vals[sl].to(dev), labels[sl].to(dev))
    torch.cuda.synchronize()
    lossv = float(m.last_loss())
    check(lossv == lossv and 0 < lossv < 5, "bn %s: loss %r" % (mode, lossv))
    check(int(m.global_step) == 3, "bn %s: step counter %r" % (mode, m.global_step))
    if rank == 0:
        print("sharded batch_norm + Adam a2a=%s world=%d: broadcast_state + 3 steps ok, loss %.4f" % (mode, world, lossv), flush=True)
    del m
    torch.cuda.synchronize(); dist.barrier()

flag = torch.tensor([1 if ok else 0], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0: print("SHARDED_CHECK", "PASS" if int(flag) == 1 else "FAIL", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(flag) == 1 else 1)
