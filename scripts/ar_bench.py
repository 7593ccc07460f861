"""Times the fused all-reduce+optimizer kernel alone (torchrun, any world size) against NCCL all-reduce + local flat
optimizer on the same buffers. Device-timed, max over ranks."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29541")
dist.init_process_group("nccl", device_id=dev, rank=rank, world_size=world)
from deepfm_b200.models import DeepFMConfig
from deepfm_b200.optim import TF1Optimizer
from deepfm_b200.parallel.dp import FusedDataParallel

def timeit(fn, n=100):
    for _ in range(10): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / n], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)

cfg = DeepFMConfig(field_size=39, feature_size=117581, embedding_size=32, deep_layers=[400, 400, 400], dropout=[0.5] * 3)
for comm in ["p2p", "nvls", "nccl"]:
    try:
        m = FusedDataParallel(cfg, 1024, TF1Optimizer("Adam", 5e-4), 1, dev, world, rank, comm=comm, use_graph=False)
    except RuntimeError as e:
        if rank == 0: print("comm=%s unavailable (%s)" % (comm, e)); 
        continue
    n = m.layout.trainable_numel
    t_opt = timeit(m._optimizer_kernels)
    t_zero = timeit(lambda: m.G.zero_())
    t_step = timeit(lambda: m.train_step_slot(0))
    # sub-kernel probes on symmetric memory
    from deepfm_b200.ops import kernels as K
    s = m.in_slots[0]
    t_embb = timeit(lambda: K.emb_bwd(s["ids"], s["vals"], m._w("fm_v"), m.Ssum, m.dy, m.dx0, m._g("fm_w"), m._g("fm_v")))
    t_wg = timeit(lambda: m.wgrad[0].run())
    if rank == 0:
        print("world=%d comm=%-5s (%s) n=%.2fM  opt_kernel=%.1f us  G.zero=%.1f us  emb_bwd=%.1f us  wgrad0=%.1f us  full_step(no graph)=%.1f us  bytes/rank over NVLink ~%.1f MB"
              % (world, comm, m.comm_name, n / 1e6, t_opt, t_zero, t_embb, t_wg, t_step, 2 * 4 * n * (world - 1) / world / 1e6), flush=True)
    del m
    torch.cuda.synchronize(); dist.barrier()
# plain NCCL all-reduce of the same size for reference
buf = torch.zeros(4701120, device=dev)
t = timeit(lambda: dist.all_reduce(buf))
if rank == 0: print("NCCL all_reduce fp32 %.1f MB: %.1f us" % (buf.numel() * 4 / 1e6, t), flush=True)
# local (non-symmetric) memory control for the atomics
from deepfm_b200.models.fused import FusedDeepFM
from deepfm_b200.ops import kernels as K
m = FusedDeepFM(cfg, 1024, TF1Optimizer("Adam", 5e-4), 1, dev, use_graph=False)
s = m.in_slots[0]
t_embb = timeit(lambda: K.emb_bwd(s["ids"], s["vals"], m._w("fm_v"), m.Ssum, m.dy, m.dx0, m._g("fm_w"), m._g("fm_v")))
t_wg = timeit(lambda: m.wgrad[0].run()); t_opt = timeit(m._optimizer_kernels)
if rank == 0: print("local-memory control: emb_bwd=%.1f us wgrad0=%.1f us flat_opt=%.1f us" % (t_embb, t_wg, t_opt), flush=True)
dist.destroy_process_group()
