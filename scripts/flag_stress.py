"""Stress test of the cross-GPU flag / barrier protocol ("explicit memory-ordering tests for the P2P
flag/signal protocol"; run under torchrun at 2..8 ranks).

Thousands of data-parallel steps with the ranks deliberately skewed — every step each rank delays itself by a different,
changing amount both on the host (sleep before issuing) and on the device (a spin kernel in front of the step) — while
the barrier sequence numbers start just below 2^32 so they wrap during the run. A protocol bug shows up as a watchdog trap
(a rank waiting for a flag that was consumed or overwritten), a hang, or replicas that stop being bit-identical because a
rank read gradients or wrote weights outside its window. Checked: no trap, step counters agree, parameters and bf16
shadows bit-equal across ranks every `--check_every` steps and at the end, for the dense ZeRO-1 kernel and the sparse
touched-row exchange, P2P and NVLS.

    torchrun --nproc-per-node 2 scripts/flag_stress.py [--steps 10000]
"""
import argparse
import os
import random
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=10000)
ap.add_argument("--check_every", type=int, default=2500)
ap.add_argument("--modes", default="dense/p2p,dense/nvls,sparse/p2p,sparse/nvls")
a = ap.parse_args()

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)

from deepfm_b200.data.synthetic import SyntheticCTR
from deepfm_b200.models import DeepFMConfig
from deepfm_b200.optim import TF1Optimizer
from deepfm_b200.parallel.dp import FusedDataParallel
from deepfm_b200.parallel.dp_sparse import SparseDataParallel
from deepfm_b200.parallel.selfcheck import replicas_equal

cfg = DeepFMConfig(field_size=13, feature_size=4099, embedding_size=16, deep_layers=[64, 32], dropout=[0.9, 1.0], l2_reg=1e-4)
B = 128
ok = True
rng = random.Random(1234 + rank)
for mode in a.modes.split(","):
    exch, comm = mode.split("/")
    cls = SparseDataParallel if exch == "sparse" else FusedDataParallel
    try:
        m = cls(cfg, B, TF1Optimizer("Adam", 1e-3), 5, dev, world, rank, comm=comm)
    except RuntimeError as e:
        if rank == 0:
            print("%s unavailable: %s" % (mode, e), flush=True)
        continue
    # start every sequence number (and the flag words that mirror them) 40 below the 32-bit wrap
    start = -40
    for name in ("seq", "seq_ab", "seq_c"):
        if hasattr(m, name):
            getattr(m, name).fill_(start)
    m.arena.view("flags").fill_(start)
    torch.cuda.synchronize(); dist.barrier()
    gen = SyntheticCTR(13, 4099, seed=100 + rank)
    batches = [tuple(t.to(dev) for t in gen.batch(B, i)) for i in range(32)]
    t0 = time.time()
    for s in range(a.steps):
        # skew: a different rank lags each phase of the run, by up to ~300 us on the device and ~200 us on the host
        lag = (s // 97 + rank) % world == 0
        if lag or rng.random() < 0.05:
            torch.cuda._sleep(int(rng.uniform(1e4, 6e5)))
        if rng.random() < 0.02:
            time.sleep(rng.uniform(0, 2e-4))
        m.train_step(*batches[s % 32])
        if (s + 1) % a.check_every == 0 or s + 1 == a.steps:
            torch.cuda.synchronize()
            r = replicas_equal(m)
            steps_ok = m.global_step == s + 1
            if not (r["equal"] and steps_ok):
                ok = False
                print("[rank %d] FAIL %s at step %d: replicas %r, global_step %d" % (rank, mode, s + 1, r, m.global_step), flush=True)
    torch.cuda.synchronize()
    seqs = {n: int(getattr(m, n).item()) & 0xFFFFFFFF for n in ("seq", "seq_ab", "seq_c") if hasattr(m, n)}
    if rank == 0:
        print("%-11s world=%d steps=%d skewed, sequence numbers wrapped (now %s), %.1f s, loss %.4f: %s" % (
            mode, world, a.steps, seqs, time.time() - t0, float(m.last_loss()), "ok" if ok else "FAIL"), flush=True)
    del m
    torch.cuda.synchronize(); dist.barrier()
flag = torch.tensor([1 if ok else 0], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("FLAG_STRESS", "PASS" if int(flag) == 1 else "FAIL", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(flag) == 1 else 1)
