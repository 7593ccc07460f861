from .config import DeepFMConfig  # noqa: F401
from .params import init_params, native_names, tf_name  # noqa: F401
from .torch_backend import TorchDeepFM  # noqa: F401


def build_model(cfg, batch_size, optimizer=None, seed=0, device=None, backend="auto", **kw):
    """backend: 'fused' (sm_100a kernels, requires a B200), 'torch' (plain PyTorch), 'auto' (fused on CUDA)."""
    import torch
    if backend == "auto":
        backend = "fused" if torch.cuda.is_available() else "torch"
    if backend == "fused":
        from .fused import FusedDeepFM
        return FusedDeepFM(cfg, batch_size, optimizer, seed, device, **kw)
    dev = device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu")
    kw.pop("use_graph", None); kw.pop("input_slots", None); kw.pop("grad_alloc", None)
    return TorchDeepFM(cfg, batch_size, optimizer, seed, dev, **kw)
