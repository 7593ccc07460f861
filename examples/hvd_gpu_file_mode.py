"""Stand-in for the reference notebook 2-hvd-gpu/deepfm-sagemaker-hvd-gpu.ipynb, File-mode cell (NB-HVD:56-156):
one box, one process per GPU, Horovod-style synchronous data parallelism — here with the fused NVLS/P2P
all-reduce + optimizer kernel instead of NCCL. Same hyperparameters as the notebook (NB-HVD:94-107)."""
import sys
from deepfm_b200.launch import LocalEstimator

train_dir, val_dir = sys.argv[1], sys.argv[2]
gpus = int(sys.argv[3]) if len(sys.argv) > 3 else 8
hyperparameters = {
    "task_type": "train", "field_size": 39, "feature_size": 117581, "embedding_size": 32,
    "deep_layers": "128,64,32", "dropout": "0.5,0.5,0.5", "optimizer": "Adam", "learning_rate": 0.0005,
    "batch_size": 1024, "num_epochs": 10, "log_steps": 10, "worker_per_host": gpus,
    "enable_s3_shard": False, "enable_data_multi_path": True,   # as in the notebook (only matters in pipe mode)
    "checkpoint_dir": "/tmp/deepfm_ckpt", "servable_model_dir": "/tmp/deepfm_model", "clear_existing_model": True,
}
LocalEstimator(hyperparameters, distribution={"mpi": {"enabled": True, "processes_per_host": gpus}},
               input_mode="File").fit({"training": train_dir, "evaluation": val_dir})
