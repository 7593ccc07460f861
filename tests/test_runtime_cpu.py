"""Runtime plumbing that the reference gets from tf.estimator: hooks, checkpoint rotation, atomic
checkpoint publication, fault-injection spec parsing."""
import io
import json
import os

import pytest
import torch


def _state(step):
    return {"fm_bias": torch.tensor([0.5]), "fm_w": torch.arange(4.0), "global_step": torch.tensor(step)}


def test_checkpoint_rotation_state_file_and_atomic_publication(tmp_path):
    from deepfm_b200.optim import TF1Optimizer
    from deepfm_b200.utils.checkpoint import all_checkpoints, latest_checkpoint, save_checkpoint
    d = str(tmp_path / "ck")
    opt = TF1Optimizer("GD", 0.1)
    for step in (10, 20, 30, 40):
        save_checkpoint(d, _state(step), opt, keep=2)
    assert all_checkpoints(d) == [30, 40]                               # keep_checkpoint_max
    assert latest_checkpoint(d).endswith("model.ckpt-40")
    state = open(os.path.join(d, "checkpoint")).read()
    assert 'model_checkpoint_path: "model.ckpt-40"' in state and state.count("all_model_checkpoint_paths") == 2
    assert not [f for f in os.listdir(d) if ".tmp" in f]                # nothing half-written left behind
    # a crash that left only an index (no data file) and no state file must not be picked up as "latest"
    os.remove(os.path.join(d, "checkpoint"))
    open(os.path.join(d, "model.ckpt-50.index"), "wb").write(b"partial")
    assert latest_checkpoint(d).endswith("model.ckpt-40")
    # a state file that points at a vanished checkpoint falls back to the newest complete one
    open(os.path.join(d, "checkpoint"), "w").write('model_checkpoint_path: "model.ckpt-99"\n')
    assert latest_checkpoint(d).endswith("model.ckpt-40")
    assert latest_checkpoint(str(tmp_path / "nothing")) is None


def test_logging_hook_honours_log_steps_and_writes_jsonl(tmp_path):
    from deepfm_b200.runtime.hooks import LoggingHook
    out, mf = io.StringIO(), str(tmp_path / "m" / "metrics.jsonl")
    calls = []
    hook = LoggingHook(every_steps=5, global_batch=64, metrics_file=mf, out=out)
    hook.begin(0)
    for step in range(1, 13):
        hook.after_step(step, lambda: calls.append(1) or 0.25)
    assert len(calls) == 2                                              # the loss is read (device sync) only when due
    lines = [json.loads(x) for x in open(mf)]
    assert [x["step"] for x in lines] == [5, 10] and all(x["loss"] == 0.25 for x in lines)
    assert all(abs(x["examples_per_sec"] - 64 * x["steps_per_sec"]) < 1e-6 * x["examples_per_sec"] for x in lines)
    assert "loss = 0.25, step = 5" in out.getvalue() and "global_step/sec" in out.getvalue()
    hook.log_eval({"auc": 0.7, "loss": 0.4, "global_step": 10})
    assert json.loads(open(mf).read().splitlines()[-1])["eval"] is True
    silent = LoggingHook(1, 64, None, out=io.StringIO(), enabled=False)    # non-chief ranks log nothing
    silent.after_step(1, lambda: pytest.fail("must not read the loss"))


def test_checkpoint_saver_hook_periods(tmp_path):
    from deepfm_b200.optim import TF1Optimizer
    from deepfm_b200.runtime.hooks import CheckpointSaverHook

    class FakeTrainer:
        opt = TF1Optimizer("GD", 0.1)
        step = 0

        def state_dict(self):
            return _state(self.step)

    tr, d = FakeTrainer(), str(tmp_path / "ck")
    by_steps = CheckpointSaverHook(d, save_secs=600, save_steps=3, keep=5)
    saved = []
    for step in range(1, 10):
        tr.step = step
        if by_steps.due(step):
            saved.append(os.path.basename(by_steps.save(tr, out=io.StringIO())))
    assert saved == ["model.ckpt-3", "model.ckpt-6", "model.ckpt-9"]
    assert by_steps.save(tr, out=io.StringIO()) is None                 # same step twice: no duplicate write
    by_secs = CheckpointSaverHook(d, save_secs=600, save_steps=0)
    assert not by_secs.due(1)
    by_secs._last_t -= 601
    assert by_secs.due(1)
    assert not CheckpointSaverHook(d, enabled=False).due(3) and not CheckpointSaverHook("", save_steps=1).due(1)


def test_fault_injection_spec():
    from deepfm_b200.runtime.estimator import _parse_fault
    assert _parse_fault("", 0) is None
    assert _parse_fault("1:25", 0) is None and _parse_fault("1:25", 1) == (25, "")
    assert _parse_fault("0:7:/tmp/some:marker", 0) == (7, "/tmp/some:marker")
    assert _parse_fault("0:7:" + __file__, 0) is None                   # marker exists: the fault already fired
    with pytest.raises(SystemExit):
        _parse_fault("zero:seven", 0)


def test_inspect_checkpoint_lists_tf_names_for_model_dir_and_servable(tmp_path):
    from deepfm_b200.models import DeepFMConfig, init_params
    from deepfm_b200.optim import TF1Optimizer
    from deepfm_b200.serving.export import export_servable
    from deepfm_b200.utils.checkpoint import save_checkpoint
    from deepfm_b200.utils.inspect_checkpoint import main
    cfg = DeepFMConfig(field_size=4, feature_size=50, embedding_size=4, deep_layers=[8], dropout=[1.0])
    params = init_params(cfg, seed=1)
    sd = dict(params, global_step=torch.tensor(12))
    save_checkpoint(str(tmp_path / "ck"), sd, TF1Optimizer("GD", 0.1))
    out = io.StringIO()
    assert main([str(tmp_path / "ck"), "--stats"], out=out) == 0
    text = out.getvalue()
    for name in ("fm_bias", "fm_w", "fm_v", "Deep-part/mlp0/weights", "Deep-part/deep_out/biases", "global_step"):
        assert name in text
    assert "model.ckpt-12" in text and "= 12" in text and "l2" in text
    export_servable(str(tmp_path / "model"), cfg, params, global_step=12)
    out = io.StringIO()
    assert main([str(tmp_path / "model"), "--tensor", "fm_w"], out=out) == 0 and "(50,)" in out.getvalue()
    assert main([str(tmp_path / "nothing")], out=io.StringIO()) == 1


def test_bench_reference_arm_reports_unavailable():
    """bench.py contract: `--impl reference` prints one JSON line and exits 0 when the reference cannot be installed."""
    import subprocess
    import sys
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(repo, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "5",
                        "--warmup", "3"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and isinstance(line["unavailable"], str) and line["unavailable"]


def test_tensorboard_event_files_roundtrip_and_estimator_layout(tmp_path):
    """Event files are TFRecord-framed Event protos: version header first, then scalar summaries (train: model_dir,
    eval: model_dir/eval), readable back record by record with CRC checks."""
    import glob
    from deepfm_b200.runtime.hooks import LoggingHook
    from deepfm_b200.utils.tfevents import EventFileWriter, read_events
    w = EventFileWriter(str(tmp_path / "raw"))
    w.add_scalars({"loss": 0.693, "auc": 0.5}, step=1, wall_time=123.5)
    w.add_scalar("loss", 0.25, step=1 << 40)
    w.close()
    ev = list(read_events(w.path))
    assert ev[0][2] == "brain.Event:2" and ev[0][1] == 0 and ev[0][3] == []
    assert ev[1][0] == 123.5 and ev[1][1] == 1 and dict(ev[1][3]).keys() == {"loss", "auc"}
    assert abs(dict(ev[1][3])["loss"] - 0.693) < 1e-6 and ev[2][1] == 1 << 40 and ev[2][3] == [("loss", 0.25)]
    assert os.path.basename(w.path).startswith("events.out.tfevents.")
    # through the logging hook, the way the Estimator uses it
    d = str(tmp_path / "model")
    hook = LoggingHook(2, 64, None, out=io.StringIO(), event_dir=d)
    hook.begin(0)
    for step in range(1, 5):
        hook.after_step(step, lambda: 0.5)
    hook.log_eval({"auc": 0.71, "loss": 0.45, "global_step": 4})
    hook.close()
    (train_file,) = glob.glob(os.path.join(d, "events.out.tfevents.*"))
    (eval_file,) = glob.glob(os.path.join(d, "eval", "events.out.tfevents.*"))
    train = [e for e in read_events(train_file) if e[3]]
    assert [e[1] for e in train] == [2, 4] and {t for e in train for t, _ in e[3]} == {"loss", "global_step/sec", "examples/sec"}
    (e,) = [e for e in read_events(eval_file) if e[3]]
    assert e[1] == 4 and abs(dict(e[3])["auc"] - 0.71) < 1e-6


def test_inspect_tfrecords_reports_schema_problems(tmp_path):
    from deepfm_b200.data.synthetic import write_synthetic_tfrecords
    from deepfm_b200.utils.inspect_tfrecords import main
    paths = write_synthetic_tfrecords(str(tmp_path), "tr", 2, 300, 5, 1000, seed=0)
    out = io.StringIO()
    assert main(paths + ["--field_size", "5", "--feature_size", "1000"], out=out) == 0
    text = out.getvalue()
    assert "records            600" in text and "ids_out_of_range   0" in text
    out = io.StringIO()
    assert main(paths + ["--field_size", "5", "--feature_size", "3"], out=out) == 1 and "outside [0, 3)" in out.getvalue()
    out = io.StringIO()
    assert main(paths + ["--field_size", "7"], out=out) == 1 and "FAILED" in out.getvalue()      # wrong schema
    blob = bytearray(open(paths[0], "rb").read())
    blob[40] ^= 0xFF
    open(paths[0], "wb").write(bytes(blob))
    out = io.StringIO()
    assert main(paths + ["--field_size", "5"], out=out) == 1 and "FAILED" in out.getvalue()      # corrupted record
