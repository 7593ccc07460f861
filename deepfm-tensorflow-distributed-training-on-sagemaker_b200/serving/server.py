"""REST model server for an exported servable, speaking the TensorFlow-Serving predict API the reference's SavedModel
export (`export_savedmodel`, PS:451-467 / HVD:415-431) is meant for:

    python -m deepfm_b200.serving.server --model_dir /opt/ml/model --port 8501 [--model_name deepfm]

    POST /v1/models/deepfm:predict
        {"instances": [{"feat_ids": [..F ints..], "feat_vals": [..F floats..]}, ...]}         row format
        {"inputs": {"feat_ids": [[..], ..], "feat_vals": [[..], ..]}}                          columnar format
      -> {"predictions": [{"prob": p}, ...]}          (row)      {"outputs": {"prob": [p, ...]}}   (columnar)
    GET  /v1/models/deepfm            -> model version status       GET /v1/models/deepfm/metadata -> serving signature
    GET  /ping, POST /invocations     -> SageMaker inference-container conventions (same payloads as :predict)

Concurrent requests are MICRO-BATCHED: handler threads enqueue their rows, one scoring thread drains the queue into a
single `Predictor.predict` call of up to `max_batch` rows (waiting at most `batch_timeout_ms` for more rows), so a GPU
forward pass is shared by every request in flight instead of being launched per request.
Standard library only (http.server + threading): nothing to install on the serving box.
"""
from __future__ import annotations

import argparse
import json
import os
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import List, Optional, Tuple

import torch

from .export import latest_export, signature
from .predictor import Predictor


class _Pending:
    __slots__ = ("ids", "vals", "event", "probs", "error")

    def __init__(self, ids: torch.Tensor, vals: torch.Tensor):
        self.ids, self.vals = ids, vals
        self.event = threading.Event()
        self.probs: Optional[torch.Tensor] = None
        self.error: Optional[str] = None


class BatchingScorer:
    """One scoring thread; requests from any number of handler threads are fused into shared forward passes."""

    def __init__(self, predictor: Predictor, max_batch: int = 4096, batch_timeout_ms: float = 1.0):
        self.p, self.max_batch, self.timeout = predictor, int(max_batch), batch_timeout_ms / 1e3
        self._q: List[_Pending] = []
        self._cv = threading.Condition()
        self._stop = False
        self.batches = 0
        self.rows = 0
        self._t = threading.Thread(target=self._run, name="deepfm-scorer", daemon=True)
        self._t.start()

    def score(self, ids: torch.Tensor, vals: torch.Tensor, timeout: float = 30.0) -> torch.Tensor:
        item = _Pending(ids, vals)
        with self._cv:
            self._q.append(item)
            self._cv.notify()
        if not item.event.wait(timeout):
            raise TimeoutError("scoring timed out")
        if item.error:
            raise RuntimeError(item.error)
        return item.probs

    def _take(self) -> List[_Pending]:
        with self._cv:
            while not self._q and not self._stop:
                self._cv.wait()
            if self._stop:
                return []
            deadline = time.monotonic() + self.timeout
            while sum(x.ids.shape[0] for x in self._q) < self.max_batch:
                left = deadline - time.monotonic()
                if left <= 0:
                    break
                self._cv.wait(left)
            take, rows = [], 0
            while self._q and (not take or rows + self._q[0].ids.shape[0] <= self.max_batch):
                rows += self._q[0].ids.shape[0]
                take.append(self._q.pop(0))
            return take

    def _run(self) -> None:
        while not self._stop:
            take = self._take()
            if not take:
                continue
            try:
                ids = torch.cat([x.ids for x in take])
                vals = torch.cat([x.vals for x in take])
                probs = self.p.predict({"feat_ids": ids, "feat_vals": vals})["prob"]
                self.batches += 1
                self.rows += ids.shape[0]
                o = 0
                for x in take:
                    n = x.ids.shape[0]
                    x.probs = probs[o:o + n]
                    o += n
            except Exception as e:  # report to every waiting request, keep serving
                for x in take:
                    x.error = "%s: %s" % (type(e).__name__, e)
            for x in take:
                x.event.set()

    def close(self) -> None:
        with self._cv:
            self._stop = True
            self._cv.notify_all()
        self._t.join(timeout=5)


def parse_predict_request(body: dict, field_size: int) -> Tuple[torch.Tensor, torch.Tensor, bool]:
    """-> (ids int64 [n,F], vals float32 [n,F], row_format). Raises ValueError with a client-facing message."""
    if "instances" in body:
        inst = body["instances"]
        if not isinstance(inst, list) or not inst:
            raise ValueError("'instances' must be a non-empty list")
        try:
            ids = [r["feat_ids"] for r in inst]
            vals = [r["feat_vals"] for r in inst]
        except (KeyError, TypeError):
            raise ValueError("every instance needs 'feat_ids' and 'feat_vals'")
        row = True
    elif "inputs" in body:
        inp = body["inputs"]
        if not isinstance(inp, dict) or "feat_ids" not in inp or "feat_vals" not in inp:
            raise ValueError("'inputs' must hold 'feat_ids' and 'feat_vals'")
        ids, vals, row = inp["feat_ids"], inp["feat_vals"], False
    else:
        raise ValueError("request needs 'instances' (row format) or 'inputs' (columnar format)")
    try:
        ids_t = torch.tensor(ids, dtype=torch.int64)
        vals_t = torch.tensor(vals, dtype=torch.float32)
    except (ValueError, TypeError, RuntimeError):
        raise ValueError("feat_ids must be integers and feat_vals numbers, rectangular")
    if ids_t.dim() == 1:
        ids_t, vals_t = ids_t.reshape(1, -1), vals_t.reshape(1, -1)
    if ids_t.dim() != 2 or ids_t.shape[1] != field_size or vals_t.shape != ids_t.shape:
        raise ValueError("expected feat_ids and feat_vals of shape [n, %d], got %s and %s"
                         % (field_size, list(ids_t.shape), list(vals_t.shape)))
    return ids_t, vals_t, row


def make_handler(scorer: BatchingScorer, model_name: str, version: str, quiet: bool = True):
    cfg = scorer.p.cfg
    sig = signature(cfg)

    class Handler(BaseHTTPRequestHandler):
        protocol_version = "HTTP/1.1"

        def log_message(self, fmt, *args):  # noqa: N802
            if not quiet:
                BaseHTTPRequestHandler.log_message(self, fmt, *args)

        def _send(self, code: int, obj: dict) -> None:
            data = json.dumps(obj).encode()
            self.send_response(code)
            self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(data)))
            self.end_headers()
            self.wfile.write(data)

        def do_GET(self):  # noqa: N802
            path = self.path.rstrip("/")
            if path == "/ping":
                return self._send(200, {"status": "ok"})
            if path == "/v1/models/%s" % model_name:
                return self._send(200, {"model_version_status": [
                    {"version": version, "state": "AVAILABLE", "status": {"error_code": "OK", "error_message": ""}}]})
            if path == "/v1/models/%s/metadata" % model_name:
                return self._send(200, {"model_spec": {"name": model_name, "version": version},
                                        "metadata": {"signature_def": sig}})
            if path == "/stats":
                return self._send(200, {"batches": scorer.batches, "rows": scorer.rows})
            self._send(404, {"error": "unknown path %s" % self.path})

        def do_POST(self):  # noqa: N802
            path = self.path.rstrip("/")
            if path not in ("/v1/models/%s:predict" % model_name, "/invocations"):
                return self._send(404, {"error": "unknown path %s" % self.path})
            try:
                n = int(self.headers.get("Content-Length", "0"))
                body = json.loads(self.rfile.read(n) or b"{}")
                ids, vals, row = parse_predict_request(body, cfg.field_size)
            except ValueError as e:  # json.JSONDecodeError is a ValueError
                return self._send(400, {"error": str(e)})
            try:
                probs = scorer.score(ids, vals).tolist()
            except Exception as e:
                return self._send(500, {"error": str(e)})
            self._send(200, {"predictions": [{"prob": p} for p in probs]} if row else {"outputs": {"prob": probs}})

    return Handler


class _Httpd(ThreadingHTTPServer):
    request_queue_size = 256  # listen backlog: the default of 5 resets connections under a burst of clients
    daemon_threads = True


class ModelServer:
    """`ModelServer(model_dir).start()` serves in a background thread; `.port` is the bound port (0 = ephemeral)."""

    def __init__(self, model_dir: str, port: int = 8501, host: str = "127.0.0.1", model_name: str = "deepfm",
                 max_batch: int = 4096, batch_timeout_ms: float = 1.0, backend: str = "auto", quiet: bool = True):
        path = model_dir if os.path.exists(os.path.join(model_dir, "config.json")) else latest_export(model_dir)
        self.version = os.path.basename(os.path.normpath(path))
        self.predictor = Predictor.load(path, max_batch, backend)
        self.scorer = BatchingScorer(self.predictor, max_batch, batch_timeout_ms)
        self.httpd = _Httpd((host, port), make_handler(self.scorer, model_name, self.version, quiet))
        self.port = self.httpd.server_address[1]
        self._t: Optional[threading.Thread] = None

    def start(self) -> "ModelServer":
        self._t = threading.Thread(target=self.httpd.serve_forever, name="deepfm-http", daemon=True)
        self._t.start()
        return self

    def serve_forever(self) -> None:
        self.httpd.serve_forever()

    def stop(self) -> None:
        self.httpd.shutdown()
        self.httpd.server_close()
        self.scorer.close()


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="TF-Serving-style REST server for an exported DeepFM servable")
    ap.add_argument("--model_dir", required=True, help="servable_model_dir (newest timestamped export is served)")
    ap.add_argument("--port", type=int, default=8501)
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--model_name", default="deepfm")
    ap.add_argument("--max_batch", type=int, default=4096)
    ap.add_argument("--batch_timeout_ms", type=float, default=1.0)
    ap.add_argument("--backend", default="auto", choices=["auto", "fused", "torch", "torch-cpu"])
    a = ap.parse_args(argv)
    srv = ModelServer(a.model_dir, a.port, a.host, a.model_name, a.max_batch, a.batch_timeout_ms, a.backend, quiet=False)
    print("INFO:deepfm_b200.serving: model %s version %s on http://%s:%d/v1/models/%s:predict"
          % (a.model_name, srv.version, a.host, srv.port, a.model_name), flush=True)
    try:
        srv.serve_forever()
    except KeyboardInterrupt:
        pass
    finally:
        srv.stop()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
