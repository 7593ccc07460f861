"""Serving path on CPU: exported servable -> REST server (TF-Serving predict API) -> probabilities == the oracle."""
import json
import threading
import urllib.error
import urllib.request

import pytest
import torch

F, N = 6, 300


@pytest.fixture()
def served(tmp_path):
    from deepfm_b200.models import DeepFMConfig, init_params
    from deepfm_b200.serving.export import export_servable
    from deepfm_b200.serving.server import ModelServer
    cfg = DeepFMConfig(field_size=F, feature_size=N, embedding_size=8, deep_layers=[16, 8], dropout=[0.5, 0.5])
    params = init_params(cfg, seed=5)
    params["fm_bias"] = torch.tensor([0.3])
    export_servable(str(tmp_path / "model"), cfg, params, global_step=7)
    srv = ModelServer(str(tmp_path / "model"), port=0, backend="torch-cpu", max_batch=64, batch_timeout_ms=20.0).start()
    yield srv, cfg, params
    srv.stop()


def _call(port, path, payload=None):
    req = urllib.request.Request("http://127.0.0.1:%d%s" % (port, path),
                                 data=None if payload is None else json.dumps(payload).encode(),
                                 headers={"Content-Type": "application/json"})
    try:
        with urllib.request.urlopen(req, timeout=30) as r:
            return r.status, json.loads(r.read())
    except urllib.error.HTTPError as e:
        return e.code, json.loads(e.read())


def _oracle(cfg, params, ids, vals):
    from deepfm_b200.models.oracle import oracle_forward
    return torch.sigmoid(oracle_forward(params, cfg, ids, vals, train=False))


def test_rest_predict_row_and_columnar_formats_match_oracle(served):
    srv, cfg, params = served
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(0, N, (5, F), generator=g)
    vals = torch.rand(5, F, generator=g)
    ref = srv.predictor.predict({"feat_ids": ids, "feat_vals": vals})["prob"]
    assert torch.allclose(ref, _oracle(cfg, params, ids, vals).reshape(-1), atol=1e-5)
    code, out = _call(srv.port, "/v1/models/deepfm:predict",
                      {"instances": [{"feat_ids": i, "feat_vals": v} for i, v in zip(ids.tolist(), vals.tolist())]})
    assert code == 200 and len(out["predictions"]) == 5
    assert torch.allclose(torch.tensor([p["prob"] for p in out["predictions"]]), ref, atol=1e-6)
    code, out = _call(srv.port, "/invocations", {"inputs": {"feat_ids": ids.tolist(), "feat_vals": vals.tolist()}})
    assert code == 200 and torch.allclose(torch.tensor(out["outputs"]["prob"]), ref, atol=1e-6)


def test_rest_status_metadata_and_errors(served):
    srv, cfg, _ = served
    assert _call(srv.port, "/ping") == (200, {"status": "ok"})
    code, st = _call(srv.port, "/v1/models/deepfm")
    assert code == 200 and st["model_version_status"][0]["state"] == "AVAILABLE"
    assert st["model_version_status"][0]["version"] == srv.version and srv.version.isdigit()
    code, md = _call(srv.port, "/v1/models/deepfm/metadata")
    sig = md["metadata"]["signature_def"]["serving_default"]
    assert sig["inputs"]["feat_ids"] == {"dtype": "int64", "shape": [None, F]}  # the reference's receiver, PS:458-462
    assert sig["inputs"]["feat_vals"]["dtype"] == "float32" and "prob" in sig["outputs"]
    assert _call(srv.port, "/v1/models/other:predict", {"instances": []})[0] == 404
    code, err = _call(srv.port, "/v1/models/deepfm:predict", {"instances": [{"feat_ids": [1, 2], "feat_vals": [1.0, 1.0]}]})
    assert code == 400 and "shape" in err["error"]
    assert _call(srv.port, "/v1/models/deepfm:predict", {"foo": 1})[0] == 400


def test_concurrent_requests_are_micro_batched(served):
    srv, cfg, params = served
    g = torch.Generator().manual_seed(1)
    ids = torch.randint(0, N, (24, F), generator=g)
    vals = torch.rand(24, F, generator=g)
    ref = srv.predictor.predict({"feat_ids": ids, "feat_vals": vals})["prob"]
    before = srv.scorer.batches
    got = [None] * 24
    start = threading.Barrier(24)

    def one(i):
        start.wait()
        code, out = _call(srv.port, "/v1/models/deepfm:predict",
                          {"instances": [{"feat_ids": ids[i].tolist(), "feat_vals": vals[i].tolist()}]})
        got[i] = out["predictions"][0]["prob"] if code == 200 else None

    ts = [threading.Thread(target=one, args=(i,)) for i in range(24)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert all(x is not None for x in got)
    assert torch.allclose(torch.tensor(got), ref, atol=1e-6)           # every caller got ITS row back
    assert srv.scorer.batches - before < 24, "requests in flight together must share forward passes"
