"""ctypes loader for the in-tree CUDA library (`ops/_lib/libdeepfm_b200.so`, ABI in include/deepfm_b200.h).

The library is sm_100a-only. On a GPU box the fused path is THE path: if the library is missing or the device
is not Blackwell, `require()` raises instead of silently falling back to eager PyTorch.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libdeepfm_b200.so")
if os.environ.get("DFM_KTRACE") == "1":  # timeline build (python -m deepfm_b200.build --trace; scripts/step_trace.py)
    LIB_PATH = os.path.join(_HERE, "_lib", "libdeepfm_b200_trace.so")
_lib: Optional[ctypes.CDLL] = None


class NativeError(RuntimeError):
    pass


def available() -> bool:
    return os.path.exists(LIB_PATH)


def load(build_if_missing: bool = True) -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if not build_if_missing:
            raise NativeError("native library not built: run `python -m deepfm_b200.build`")
        from ..build import build_cuda
        build_cuda(trace=LIB_PATH.endswith("_trace.so"))
    lib = ctypes.CDLL(LIB_PATH)
    lib.dfm_last_error.restype = ctypes.c_char_p
    lib.dfm_gemm_plan_destroy.restype = None
    if hasattr(lib, "dfm_mlp_fwd_plan_destroy"):
        lib.dfm_mlp_fwd_plan_destroy.restype = None
    _lib = lib
    return lib


def watchdog_message() -> Optional[str]:
    """The device-side flag-wait watchdog's report ("timed out waiting for rank r ..."), if one fired in this process. The
    record lives in pinned host memory, so this works after the trap has killed the CUDA context."""
    if _lib is None:
        return None
    buf = ctypes.create_string_buffer(256)
    return buf.value.decode() if _lib.dfm_watchdog_message(buf, 256) == 1 else None


def check(rc: int) -> None:
    if rc != 0:
        msg = load().dfm_last_error()
        wd = watchdog_message()
        raise NativeError("deepfm_b200 native call failed (%d): %s%s" % (rc, msg.decode() if msg else "?",
                                                                         "; " + wd if wd else ""))


def require_gpu() -> ctypes.CDLL:
    """Load the library and verify the current CUDA device is sm_100-class; raise loudly otherwise."""
    import torch
    if not torch.cuda.is_available():
        raise NativeError("CUDA device required for the fused sm_100a path")
    lib = load()
    r = lib.dfm_device_is_sm100()
    if r != 1:
        raise NativeError("libdeepfm_b200.so is built for sm_100a only; current device is not Blackwell (rc=%d)" % r)
    return lib
