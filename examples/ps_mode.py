"""Stand-in for the reference notebook 1-ps-cpu/deepfm-sagemaker-ps-cpu.ipynb (NB-PS:56-143): asynchronous parameter
server. The reference starts 1 PS + 1 worker per c5.18xlarge; here N worker processes share one host-resident
parameter store (GPU workers run the fused kernels, CPU-only boxes fall back to the torch backend)."""
import sys
from deepfm_b200.launch import LocalEstimator

train_dir, val_dir = sys.argv[1], sys.argv[2]
workers = int(sys.argv[3]) if len(sys.argv) > 3 else 2
hyperparameters = {
    "task_type": "train", "field_size": 39, "feature_size": 117581, "embedding_size": 32, "deep_layers": "128,64,32",
    "dropout": "0.5,0.5,0.5", "batch_size": 1024, "num_epochs": 10, "log_steps": 10,
    "model_dir": "/tmp/deepfm_ps_ckpt", "servable_model_dir": "/tmp/deepfm_model", "enable_s3_shard": True,
    "ps_staleness": 4,
}
LocalEstimator(hyperparameters, distribution={"parameter_server": {"enabled": True, "workers_per_host": workers}},
               input_mode="File").fit({"training": train_dir, "evaluation": val_dir})
