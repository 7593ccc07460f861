"""Multi-GPU check (run under torchrun): FusedDataParallel on W shards of a batch == the oracle on the whole batch,
for comm in {p2p, nvls (if offered), nccl}; prints device-timed us/step for each."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)

from deepfm_b200.models import DeepFMConfig, init_params
from deepfm_b200.models.oracle import oracle_grads
from deepfm_b200.optim import TF1Optimizer
from deepfm_b200.parallel.dp import FusedDataParallel

ok = True
def check(cond, msg):
    global ok
    if not cond:
        ok = False
        print("[rank %d] FAIL: %s" % (rank, msg), flush=True)

def batch(cfg, B, seed):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, cfg.feature_size, (B, cfg.field_size), generator=g, dtype=torch.int64)
    ids[:, 0] = 3
    vals = torch.rand(B, cfg.field_size, generator=g)
    labels = (torch.rand(B, generator=g) < 0.3).float()
    return ids, vals, labels

cfg = DeepFMConfig(field_size=39, feature_size=5000, embedding_size=32, deep_layers=[400, 400, 400],
                   dropout=[0.5, 0.8, 1.0], l2_reg=1e-4)
Bl = 256
GB = Bl * world
modes = ["p2p", "nvls", "nccl"]
for comm in modes:
    for opt in (["Momentum", "Adam"] if comm != "nccl" else ["Momentum"]):
        params = init_params(cfg, seed=3)
        try:
            m = FusedDataParallel(cfg, Bl, TF1Optimizer(opt, 1e-2), 77, dev, world, rank, comm=comm, params=params)
        except RuntimeError as e:
            if rank == 0:
                print("comm=%s unavailable: %s" % (comm, e), flush=True)
            break
        ref_p = {k: v.clone().to(dev) for k, v in params.items()}
        ref_opt = TF1Optimizer(opt, 1e-2)
        ref_slots = ref_opt.init_slots(ref_p)
        for step in range(3):
            ids, vals, labels = batch(cfg, GB, step)
            sl = slice(rank * Bl, (rank + 1) * Bl)
            m.train_step(ids[sl].to(dev), vals[sl].to(dev), labels[sl].to(dev))
            r = oracle_grads(ref_p, cfg, ids.to(dev), vals.to(dev), labels.to(dev), True, 77, step, 0,
                             compute_dtype=torch.bfloat16)
            ref_opt.apply_all(ref_p, r["grads"], ref_slots)
        torch.cuda.synchronize()
        check(m.global_step == 3, "step counter %d" % m.global_step)
        got = m.named_parameters()
        worst = 0.0
        for k in ref_p:
            a, b = got[k].float(), ref_p[k].float()
            scale = float(b.abs().max()) + 1e-8
            if opt == "Momentum":
                err = float((a - b).abs().max()) / scale
                worst = max(worst, err)
                check(err < 3e-2, "%s %s %s err %.3e" % (comm, opt, k, err))
            else:
                frac = float(((a - b).abs() > 3e-2 * scale).float().mean())
                worst = max(worst, frac)
                check(frac < 0.20, "%s %s %s frac_bad %.3f" % (comm, opt, k, frac))
        # replicas identical across ranks
        flat = m.P.clone()
        ref = flat.clone()
        dist.broadcast(ref, 0)
        check(bool(torch.equal(flat, ref)), "%s %s replicas differ across ranks" % (comm, opt))
        sh = m.shadow.clone(); shr = sh.clone(); dist.broadcast(shr, 0)
        check(bool(torch.equal(sh, shr)), "%s %s bf16 shadows differ across ranks" % (comm, opt))
        # chief-side slot assembly round-trips through a state dict
        if rank == 0:
            sd = m.state_dict()
            key = "fm_v/" + ("Momentum" if opt == "Momentum" else "Adam")
            check(key in sd and sd[key].shape == ref_p["fm_v"].shape, "slot %s missing" % key)
            if opt == "Momentum":
                e = float((sd[key] - ref_slots["fm_v"][0]).abs().max()) / (float(ref_slots["fm_v"][0].abs().max()) + 1e-8)
                check(e < 3e-2, "slot mismatch %.3e" % e)
        # timing
        ids, vals, labels = batch(cfg, GB, 100)
        sl = slice(rank * Bl, (rank + 1) * Bl)
        m.load_inputs(ids[sl].to(dev), vals[sl].to(dev), labels[sl].to(dev), 0)
        for _ in range(10):
            m.train_step_slot(0)
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(100):
            m.train_step_slot(0)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 10.0
        t = torch.tensor([us], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print("comm=%-5s opt=%-8s world=%d mode=%s worst=%.3e us/step=%.1f loss=%.4f" % (
                comm, opt, world, m.comm_name, worst, float(t), float(m.last_loss())), flush=True)
        del m
        torch.cuda.synchronize(); dist.barrier()

flag = torch.tensor([1 if ok else 0], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("DP_CHECK", "PASS" if int(flag) == 1 else "FAIL", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(flag) == 1 else 1)
