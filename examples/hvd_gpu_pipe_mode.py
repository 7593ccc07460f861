"""Stand-in for the Pipe-mode cell of the Horovod notebook (NB-HVD:172-312): one FIFO channel per epoch per channel,
streamed by the launcher; each worker shards the shared training stream (HVD:110-120)."""
import sys
from deepfm_b200.launch import LocalEstimator

train_dir, val_dir = sys.argv[1], sys.argv[2]
gpus = int(sys.argv[3]) if len(sys.argv) > 3 else 8
hyperparameters = {
    "task_type": "train", "field_size": 39, "feature_size": 117581, "embedding_size": 32, "deep_layers": "128,64,32",
    "dropout": "0.5,0.5,0.5", "batch_size": 1024, "num_epochs": 10, "log_steps": 10, "worker_per_host": gpus,
    "training_channel_name": "training", "evaluation_channel_name": "evaluation",
    "checkpoint_dir": "/tmp/deepfm_ckpt", "servable_model_dir": "/tmp/deepfm_model",
}
LocalEstimator(hyperparameters, distribution={"mpi": {"enabled": True, "processes_per_host": gpus}},
               input_mode="Pipe").fit({"training": train_dir, "evaluation": val_dir})
