"""Grid-size sweep of the fused all-reduce+optimizer kernel (torchrun)."""
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
def timeit(fn, n=200):
    for _ in range(20): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / n], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
for feat in [int(x) for x in os.environ.get("AR_FEATS", "117581,2000").split(",")]:
    cfg = DeepFMConfig(field_size=39, feature_size=feat, embedding_size=32, deep_layers=[400, 400, 400], dropout=[0.5] * 3)
    for comm in os.environ.get("AR_COMMS", "p2p,nvls").split(","):
        for blocks in [int(x) for x in os.environ.get("AR_BLOCKS", "296,148,64,16").split(",")]:
            os.environ["DFM_AR_BLOCKS"] = str(blocks)
            try:
                m = FusedDataParallel(cfg, 1024, TF1Optimizer("Adam", 5e-4), 1, dev, world, rank, comm=comm, use_graph=False)
            except RuntimeError as e:
                break
            g = torch.cuda.CUDAGraph()
            m._optimizer_kernels(); torch.cuda.synchronize(); dist.barrier()
            with torch.cuda.graph(g):
                for _ in range(10): m._optimizer_kernels()
            t = timeit(g.replay, n=30) / 10
            if rank == 0:
                print("world=%d feat=%d n=%.2fM comm=%s blocks=%d: %.1f us/kernel (graph of 10)" % (world, feat, m.layout.trainable_numel / 1e6, comm, blocks, t), flush=True)
            del m, g
            torch.cuda.synchronize(); dist.barrier()
dist.destroy_process_group()
