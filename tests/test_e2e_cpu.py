"""End-to-end on CPU (torch backend): flags -> train -> checkpoint -> resume -> eval -> infer -> export -> serve,
plus the launcher with world_size=2 over gloo (SURVEY section 4 'End-to-end' and 'Distributed without a cluster')."""
import json
import os
import subprocess
import sys

import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
F, N = 13, 2000
COMMON = ["--field_size", str(F), "--feature_size", str(N), "--embedding_size", "8", "--deep_layers", "32,16",
          "--dropout", "0.9,0.9", "--batch_size", "64", "--learning_rate", "0.005", "--log_steps", "10",
          "--backend", "torch", "--loader_threads", "2"]


@pytest.fixture(scope="module")
def data(tmp_path_factory):
    from deepfm_b200.data.synthetic import write_synthetic_tfrecords
    d = tmp_path_factory.mktemp("data")
    write_synthetic_tfrecords(str(d / "train"), "tr", 2, 1280, F, N, seed=1)
    write_synthetic_tfrecords(str(d / "val"), "va", 1, 640, F, N, seed=2)
    write_synthetic_tfrecords(str(d / "val"), "te", 1, 100, F, N, seed=3)
    return d


def _run(args, env=None, timeout=600):
    e = dict(os.environ, PYTHONPATH=REPO, CUDA_VISIBLE_DEVICES="")
    e.update(env or {})
    r = subprocess.run([sys.executable, "-m", "deepfm_b200"] + args, capture_output=True, text=True, env=e,
                       timeout=timeout, cwd=REPO)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def test_flags_sagemaker_style():
    from deepfm_b200.runtime.flags import parse_flags
    f = parse_flags(["--field_size", "39", "--feature_size=117581", "--batch_norm", "True", "--enable_s3_shard", "False",
                     "--deep_layers", "128,64,32", "--model_dir", "s3://x", "--perform_shuffle", "False",
                     "--some_unknown", "1", "--clear_existing_model"], env={})
    assert f.field_size == 39 and f.feature_size == 117581 and f.batch_norm is True and f.enable_s3_shard is False
    assert f.checkpoint_dir == "s3://x" and f.clear_existing_model is True and f.hosts == ["localhost"]
    assert "--some_unknown" in f.unknown
    f = parse_flags([], env={"SM_HOSTS": '["algo-1","algo-2"]', "SM_CURRENT_HOST": "algo-2", "SM_NUM_GPUS": "4"})
    assert f.hosts == ["algo-1", "algo-2"] and f.current_host == "algo-2" and f.num_GPUs == 4


def test_shard_decision_table():
    from deepfm_b200.parallel.shard import hvd_shard, normalise, pipe_train_channel, ps_shard
    # README-EN:86-91 / HVD:92-120, world 8 = 2 hosts x 4 workers, this is global rank 5 (local 1)
    kw = dict(world_size=8, rank=5, local_rank=1, worker_per_host=4, num_hosts=2)
    assert hvd_shard(False, True, False, **kw) == (4, 1)
    assert hvd_shard(False, False, False, **kw) == (8, 5)
    assert hvd_shard(True, True, True, **kw) is None
    assert hvd_shard(True, False, True, **kw) == (2, 1)
    assert hvd_shard(True, False, True, world_size=4, rank=3, local_rank=3, worker_per_host=4, num_hosts=1) is None
    assert hvd_shard(True, True, False, **kw) == (4, 1)
    assert hvd_shard(True, False, False, **kw) == (8, 5)
    assert ps_shard(False, ["a", "b"], "b") == (2, 1) and ps_shard(True, ["a", "b"], "b") is None
    assert normalise(None) == (1, 0)
    assert pipe_train_channel(["evaluation", "training", "training-1"], 1) == "training-1"


def test_train_eval_infer_export_resume(data, tmp_path):
    ckpt, serv = str(tmp_path / "ckpt"), str(tmp_path / "model")
    base = COMMON + ["--training_data_dir", str(data / "train"), "--val_data_dir", str(data / "val"),
                     "--model_dir", ckpt]
    out = _run(base + ["--task_type", "train", "--num_epochs", "2", "--servable_model_dir", serv,
                       "--save_checkpoints_steps", "20"])
    assert "loss = " in out and "Saving checkpoints for 40" in out and "Saving checkpoints for 80" in out
    ev = json.loads([l for l in out.splitlines() if l.startswith('{"eval"')][-1])["eval"]
    assert ev["global_step"] == 80 and 0.0 < ev["loss"] < 0.8 and ev["auc"] > 0.55 and ev["examples"] == 640
    first_loss = float(out.split("loss = ")[1].split(",")[0])
    assert 0.3 < first_loss < 0.9  # step-10 loss, already below ln 2 (the reference logs 0.696 at step 1, PDF p34)
    assert os.path.exists(os.path.join(ckpt, "checkpoint")) and os.path.exists(os.path.join(ckpt, "metrics.jsonl"))
    # resume: a second train call continues from step 80
    out2 = _run(base + ["--task_type", "train", "--num_epochs", "1", "--perform_shuffle", "True"])  # shuffled epoch
    assert "Restoring parameters from" in out2 and "Saving checkpoints for 120" in out2
    # eval / infer / export as separate tasks
    out3 = _run(base + ["--task_type", "eval"])
    ev3 = json.loads([l for l in out3.splitlines() if l.startswith('{"eval"')][-1])["eval"]
    assert ev3["global_step"] == 120
    _run(base + ["--task_type", "infer"])
    preds = [float(x) for x in open(data / "val" / "pred.txt")]
    assert len(preds) == 100 and all(0.0 <= p <= 1.0 for p in preds)
    _run(base + ["--task_type", "export", "--servable_model_dir", serv])
    from deepfm_b200.serving import Predictor
    from deepfm_b200.data.pipeline import InputPipeline
    p = Predictor.load(serv, backend="torch-cpu")
    b = next(iter(InputPipeline([str(data / "val" / "te000.tfrecords")], F, 100, drop_remainder=False, pin_memory=False)))
    probs = p.predict({"feat_ids": b.ids.long(), "feat_vals": b.vals})["prob"]
    assert torch.allclose(probs, torch.tensor(preds), atol=2e-6)
    sig = json.load(open(os.path.join(serv, sorted(os.listdir(serv))[-1], "signature.json")))
    assert set(sig["serving_default"]["inputs"]) == {"feat_ids", "feat_vals"}


def test_launcher_dp2_gloo_matches_single_process(data, tmp_path):
    """Horovod-style DP over gloo with 2 processes == 1 process on the concatenated batches (sync invariant)."""
    from deepfm_b200.launch import LocalEstimator
    hp = dict(field_size=F, feature_size=N, embedding_size=8, deep_layers="32,16", dropout="1.0,1.0", batch_size=32,
              learning_rate=0.005, backend="torch", task_type="train", num_epochs=1, log_steps=1000,
              scale_lr_by_world=False, loader_threads=1)
    env = {"PYTHONPATH": REPO, "CUDA_VISIBLE_DEVICES": ""}
    d2 = str(tmp_path / "dp2")
    LocalEstimator(dict(hp, model_dir=d2), {"mpi": {"enabled": True, "processes_per_host": 2}}, env=env,
                   cwd=REPO).fit({"training": str(data / "train"), "evaluation": str(data / "val")}, timeout=600)
    d1 = str(tmp_path / "dp1")
    LocalEstimator(dict(hp, model_dir=d1, batch_size=64), env=env, cwd=REPO).fit(
        {"training": str(data / "train"), "evaluation": str(data / "val")}, timeout=600)
    from deepfm_b200.utils.checkpoint import latest_checkpoint
    from deepfm_b200.utils.tf_bundle import read_bundle
    a, b = read_bundle(latest_checkpoint(d1)), read_bundle(latest_checkpoint(d2))
    assert int(a["global_step"]) == int(b["global_step"]) == 40
    # rank r of 2 takes records r, r+2, ...; its batch k is global batch k's odd/even half -> same global batches
    for k in ("fm_w", "fm_v", "Deep-part/mlp0/weights", "Deep-part/deep_out/weights"):
        assert torch.allclose(a[k], b[k], rtol=1e-4, atol=1e-6), k


def test_pipe_mode_training(data, tmp_path):
    from deepfm_b200.launch import LocalEstimator
    hp = dict(field_size=F, feature_size=N, embedding_size=8, deep_layers="16", dropout="1.0", batch_size=64,
              backend="torch", task_type="train", num_epochs=2, log_steps=1000, model_dir=str(tmp_path / "pm"),
              loader_threads=1)
    LocalEstimator(hp, input_mode="Pipe", env={"PYTHONPATH": REPO, "CUDA_VISIBLE_DEVICES": ""}, cwd=REPO).fit(
        {"training": str(data / "train"), "evaluation": str(data / "val")}, timeout=600)
    from deepfm_b200.utils.checkpoint import latest_checkpoint
    assert latest_checkpoint(str(tmp_path / "pm")).endswith("model.ckpt-80")  # 2 epochs x 2560 rows / 64


def test_failure_detection_and_restart_from_checkpoint(data, tmp_path):
    """a worker that dies mid-epoch must fail the job promptly (no hang in a collective), and a restart
    of the same job resumes from the chief's latest checkpoint. Fault injection: DFM_FAULT_INJECT=<rank>:<step>."""
    import time
    from deepfm_b200.launch import LocalEstimator
    from deepfm_b200.utils.checkpoint import all_checkpoints, latest_checkpoint
    from deepfm_b200.utils.tf_bundle import read_bundle
    hp = dict(field_size=F, feature_size=N, embedding_size=8, deep_layers="16", dropout="1.0", batch_size=32,
              backend="torch", task_type="train", num_epochs=1, log_steps=1000, loader_threads=1,
              save_checkpoints_steps=10)
    inputs = {"training": str(data / "train"), "evaluation": str(data / "val")}
    env = {"PYTHONPATH": REPO, "CUDA_VISIBLE_DEVICES": ""}
    # (a) no restarts allowed: the job fails with the dead worker's exit code instead of hanging
    d0 = str(tmp_path / "f0")
    est = LocalEstimator(dict(hp, model_dir=d0), env=dict(env, DFM_FAULT_INJECT="0:13"), cwd=REPO)
    t0 = time.time()
    with pytest.raises(RuntimeError, match="17"):
        est.fit(inputs, timeout=300)
    assert time.time() - t0 < 200
    assert all_checkpoints(d0) == [10]  # the checkpoint before the fault survived
    # (b) two ranks, rank 1 dies after step 25 (once); the launcher tears the job down and restarts it
    d1 = str(tmp_path / "f1")
    marker = str(tmp_path / "fault.marker")
    est = LocalEstimator(dict(hp, model_dir=d1), {"mpi": {"enabled": True, "processes_per_host": 2}},
                         env=dict(env, DFM_FAULT_INJECT="1:25:" + marker), cwd=REPO, max_restarts=1, grace_secs=3.0)
    est.fit(inputs, timeout=600)
    assert est.restarts == 1 and os.path.exists(marker)
    # first attempt checkpointed at 10 and 20; the second resumed from 20 and ran its 40-step epoch
    assert int(read_bundle(latest_checkpoint(d1))["global_step"]) == 60


def test_example_notebooks_and_scripts_use_known_flags():
    """The launcher notebooks / example scripts (stand-ins for NB-PS / NB-HVD) must stay in sync with the flag table."""
    import ast
    import glob
    import json
    from deepfm_b200.runtime.flags import flag_definitions
    known = {n for n, _, _, _ in flag_definitions({})}
    sources = []
    for nb in sorted(glob.glob(os.path.join(REPO, "examples", "notebooks", "*.ipynb"))):
        doc = json.load(open(nb))
        assert doc["nbformat"] == 4
        sources += [("%s[%d]" % (os.path.basename(nb), i), "".join(c["source"]))
                    for i, c in enumerate(doc["cells"]) if c["cell_type"] == "code"]
    for py in sorted(glob.glob(os.path.join(REPO, "examples", "*.py"))):
        sources.append((os.path.basename(py), open(py).read()))
    assert len(sources) >= 9
    seen = set()
    for name, src in sources:
        tree = ast.parse(src, name)  # every cell compiles
        for node in ast.walk(tree):
            if isinstance(node, ast.Assign) and any(getattr(t, "id", "") == "hyperparameters" for t in node.targets):
                keys = {k.value for k in node.value.keys if isinstance(k, ast.Constant)}
                assert keys <= known, (name, keys - known)
                seen |= keys
    assert {"task_type", "field_size", "feature_size", "worker_per_host", "enable_data_multi_path"} <= seen


def test_uneven_shards_do_not_deadlock_the_job(data, tmp_path):
    """Horovod's failure mode with unequal per-rank data (PDF p22-23): 383 records over 2 ranks = 192 vs 191, i.e. 6 vs 5
    full batches of 32. The ranks agree on 5 steps before the epoch, so nobody waits in a collective for a 6th."""
    from deepfm_b200.data.synthetic import write_synthetic_tfrecords
    from deepfm_b200.launch import LocalEstimator
    from deepfm_b200.utils.checkpoint import latest_checkpoint
    from deepfm_b200.utils.tf_bundle import read_bundle
    write_synthetic_tfrecords(str(tmp_path / "train"), "tr", 1, 383, F, N, seed=4)
    hp = dict(field_size=F, feature_size=N, embedding_size=8, deep_layers="16", dropout="1.0", batch_size=32,
              backend="torch", task_type="train", num_epochs=2, log_steps=1000, loader_threads=1,
              model_dir=str(tmp_path / "ck"))
    LocalEstimator(hp, {"mpi": {"enabled": True, "processes_per_host": 2}},
                   env={"PYTHONPATH": REPO, "CUDA_VISIBLE_DEVICES": ""}, cwd=REPO).fit(
        {"training": str(tmp_path / "train"), "evaluation": str(data / "val")}, timeout=240)
    assert int(read_bundle(latest_checkpoint(str(tmp_path / "ck")))["global_step"]) == 10   # 2 epochs x min(6, 5)


def _agree_worker(rank, world, port, counts, q):
    """One rank of the pipe-mode agreement test: a stream of `counts[rank]` full batches and one short one."""
    import types
    import torch.distributed as dist
    from deepfm_b200.parallel.dist import DistEnv
    from deepfm_b200.runtime.estimator import Estimator
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    released = []

    class _Batch:
        def __init__(self, i, rows):
            self.i, self.rows = i, rows

        def release(self):
            released.append(self.i)

    B = 8
    stream = [_Batch(i, B) for i in range(counts[rank])]
    stream.insert(min(2, len(stream)), _Batch(-1, B - 3))  # a short batch in the middle is skipped, not a step
    me = types.SimpleNamespace(B=B, env=DistEnv(rank=rank, world_size=world, backend="gloo"))
    got = [b.i for b in Estimator._agreed_batches(me, iter(stream), block=3)]
    q.put((rank, got, sorted(released)))
    dist.barrier()
    dist.destroy_process_group()


def test_pipe_mode_ranks_with_uneven_streams_stop_together():
    """in pipe mode no rank knows its step count up front. Ranks whose FIFOs hold different numbers of
    records must still run the same number of synchronous steps (a lone extra step would hang in the collective, the
    Horovod failure of PDF p22-23): `_agreed_batches` agrees block by block over gloo, world size 2."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    counts = [10, 7]
    port = 29400 + os.getpid() % 500
    ps = [ctx.Process(target=_agree_worker, args=(r, 2, port, counts, q)) for r in range(2)]
    for p in ps:
        p.start()
    out = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, got0, rel0), (_, got1, rel1) = out
    # blocks of 3: both have 3, 3, then rank 1 holds only 1 more -> 7 steps on BOTH ranks, and they stop in the same block
    assert got0 == list(range(7)) and got1 == list(range(7))
    assert rel0 == [-1, 7, 8] and rel1 == [-1]  # fetched-but-unused batches go back to the loader; nothing leaks
