"""deepfm_b200 — a Blackwell-native DeepFM CTR training / inference framework.

Capabilities mirror the two reference trainers (async parameter server on CPU,
Horovod data-parallel on GPU) of Chen188/deepfm-tensorflow-distributed-training-on-sagemaker,
re-designed for one 8xB200 box: PyTorch as the tensor/runtime layer, hand-written
sm_100a CUDA kernels behind a C ABI (``csrc/``, ``include/deepfm_b200.h``), a C++
host IO library, and NVLink/NVSwitch communication done from inside kernels.

Sub-packages
  models/    DeepFM definition, fp32 oracle, TF-1 initialisers
  ops/       ctypes bindings to the CUDA kernels + plain-torch reference ops
  optim/     Adam / Adagrad / Momentum / FTRL / GD with TF-1 semantics
  parallel/  DP (fused all-reduce+optimizer), row-sharded embeddings (fused a2a),
             async parameter server, shard decision table, rendezvous
  data/      TFRecord input pipeline (file + pipe mode), synthetic Criteo-shaped data
  runtime/   Estimator-like train / evaluate / predict / export driver and hooks
  serving/   servable artefact + inference entrypoint
  launch/    single-box launcher standing in for the SageMaker notebooks
  utils/     TFRecord / Example codec, libsvm converter, checkpoints, AUC, Philox
"""
__version__ = "0.1.0"
