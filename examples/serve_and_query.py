"""Serve an exported DeepFM servable over REST and query it (the role TensorFlow Serving plays for the reference's
SavedModel export, PS:451-467): start the server in-process on an ephemeral port, send one row-format and one columnar
request, print the probabilities.

    python examples/serve_and_query.py /tmp/deepfm_model        # <servable_model_dir> written by --task_type export
"""
import json
import sys
import urllib.request

import torch

from deepfm_b200.serving import ModelServer

model_dir = sys.argv[1]
srv = ModelServer(model_dir, port=0).start()
F = srv.predictor.cfg.field_size
N = srv.predictor.cfg.feature_size
ids = torch.randint(0, N, (4, F)).tolist()
vals = torch.rand(4, F).tolist()


def post(payload):
    req = urllib.request.Request("http://127.0.0.1:%d/v1/models/deepfm:predict" % srv.port, json.dumps(payload).encode(),
                                 {"Content-Type": "application/json"})
    return json.loads(urllib.request.urlopen(req, timeout=30).read())


print(post({"instances": [{"feat_ids": i, "feat_vals": v} for i, v in zip(ids, vals)]}))
print(post({"inputs": {"feat_ids": ids, "feat_vals": vals}}))
srv.stop()
