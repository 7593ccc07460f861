"""Kernel timeline of ONE training step from %globaltimer stamps inside the kernels (trace build of the library).

    python -m deepfm_b200.build --trace
    DFM_KTRACE=1 python scripts/step_trace.py [--batch_size 1024] [--steps 5] [--out profiles/step_timeline.md]
    (multi-GPU: torchrun ... scripts/step_trace.py — every rank prints its own table, offsets relative to its own step)

ncu replays kernels one at a time, so it cannot show what overlaps what in a step that runs on three graph branches; this
can. Per kernel: `enter` = first CTA running, `go` = first CTA past griddepcontrol.wait (every prerequisite grid done),
`end` = last CTA at its final instruction, all in us from the step's first stamp. The measured step runs as a CUDA-graph
replay like any other; the stamps cost two atomics per CTA.
"""
import argparse
import ctypes
import os
import sys

os.environ.setdefault("DFM_KTRACE", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

NAMES = {0: "emb_fm_fwd", 1: "mlp_fwd0", 2: "mlp_fwd1", 3: "mlp_fwd2", 4: "logit_loss", 5: "dgrad0", 6: "dgrad1",
         7: "dgrad2", 8: "wgrad0", 9: "wgrad1", 10: "wgrad2", 11: "emb_bwd", 12: "tail_opt/flat_opt", 13: "touched_bitmap",
         14: "sweep_untouched", 15: "colsum (all)", 16: "barrier_A", 17: "allreduce", 18: "bitmap_exchange", 19: "union_rowlist",
         24: "  tail_opt: dense part done", 25: "  tail_opt: rows done", 26: "  tail_opt: finalised",
         20: "  allreduce: sweep done (last CTA)", 21: "  allreduce: rows done", 22: "  allreduce: fence.sys done"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch_size", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=5, help="traced steps (each printed)")
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--comm", default="auto")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    from deepfm_b200.data.synthetic import SyntheticCTR
    from deepfm_b200.models import DeepFMConfig
    from deepfm_b200.models.fused import pack_batch
    from deepfm_b200.ops import lib as nlib
    from deepfm_b200.optim import TF1Optimizer
    from deepfm_b200.runtime.trainer import Trainer
    lib = nlib.require_gpu()
    assert lib.dfm_trace_enabled() == 1, "not the trace build: python -m deepfm_b200.build --trace, DFM_KTRACE=1"
    cfg = DeepFMConfig(field_size=39, feature_size=117581, embedding_size=32, deep_layers="400,400,400",
                       dropout=[0.5, 0.5, 0.5], l2_reg=1e-4)
    B = a.batch_size
    tr = Trainer(cfg, B, TF1Optimizer("Adam", 5e-4 * world), seed=1234, device=dev, backend="fused", world_size=world,
                 rank=rank, comm=a.comm)
    gen = SyntheticCTR(39, 117581, seed=rank)
    pool = [pack_batch(*gen.batch(B, i)).to(dev) for i in range(64)]
    for s in range(a.warmup):
        tr.train_packed_device(pool[s % 64])
    torch.cuda.synchronize()
    nw = lib.dfm_trace_words()
    buf = (ctypes.c_ulonglong * nw)()
    lines = []
    for s in range(a.steps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        lib.dfm_trace_reset()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        tr.train_packed_device(pool[(a.warmup + s) % 64])
        e1.record()
        torch.cuda.synchronize()
        assert lib.dfm_trace_read(buf) == 0
        v = list(buf)
        mins = [x for i, x in enumerate(v) if (i & 3) < 2 and x != 0xFFFFFFFFFFFFFFFF]
        t0 = min(mins)
        lines.append("\n### rank %d, traced step %d: %.1f us by CUDA events (incl. the d2d input copy)\n" % (rank, s, e0.elapsed_time(e1) * 1e3))
        lines.append("| kernel | enter_us | go_us | end_us | go->end_us |")
        lines.append("|---|---|---|---|---|")
        rows = []
        for tid in range(nw // 4):
            en, go, end, ph = v[4 * tid:4 * tid + 4]
            if en == 0xFFFFFFFFFFFFFFFF and end == 0:
                continue
            f = lambda x: "" if x in (0, 0xFFFFFFFFFFFFFFFF) else "%.1f" % ((x - t0) / 1e3)
            dur = "" if (go == 0xFFFFFFFFFFFFFFFF or end == 0) else "%.1f" % ((end - go) / 1e3)
            key = go if go != 0xFFFFFFFFFFFFFFFF else end
            rows.append((key, "| %s | %s | %s | %s | %s |" % (NAMES.get(tid, "id%d" % tid), f(en), f(go), f(end), dur)))
        lines += [r for _, r in sorted(rows)]
    text = "\n".join(lines)
    for r in range(world):
        if r == rank:
            print(text, flush=True)
        if world > 1:
            dist.barrier()
    if a.out:
        mode = "w" if rank == 0 else "a"
        for r in range(world):
            if r == rank:
                with open(a.out, mode) as f:
                    if rank == 0:
                        f.write("# Step timeline from in-kernel %globaltimer stamps (scripts/step_trace.py, trace build)\n")
                        f.write("B=%d per GPU, world=%d, comm=%s, kernels/step=%d. enter = first CTA running, go = first CTA past "
                                "griddepcontrol.wait, end = last CTA done; us from the step's first stamp.\n" % (B, world, tr.comm_name, tr.kernels_per_step))
                    f.write(text + "\n")
            if world > 1:
                dist.barrier()
    tr.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
