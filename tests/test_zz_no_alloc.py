"""The allocation-free optimizer entry point, driven through the C ABI exactly as a foreign caller would.

`dfm_flat_opt_step_static` exists for callers that have no way to obtain device scratch: processes whose GPU
memory is already owned by another allocator or capped, and call sites that cannot tolerate the device-wide
serialisation of cudaMalloc/cudaFree. It must produce what `dfm_flat_opt_step` (caller-provided scratch) produces.
Both are called twice in a row on identical state, because the static scratch keeps the previous call's contents
and the second call is the one that would trip over anything left uncleared.

This file sorts last on purpose: a run with `-x` has already executed every other test before reaching it.
"""
import ctypes
from ctypes import c_float, c_int, c_longlong, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _p(t):
    return c_void_p(0 if t is None else t.data_ptr())


def _call(lib, fn, w, g, s0, s1, opt_id, shadow, step, reg, extra, stream):
    n = w.numel()
    args = [_p(w), _p(g), _p(s0), _p(s1), c_longlong(n), c_int(opt_id), c_float(1e-2), c_float(0.9), c_float(0.999),
            c_float(1e-8), c_float(0.95), c_float(1.0), c_longlong(64), c_longlong(8192), c_float(1e-3),
            c_longlong(8192), c_longlong(12288), _p(shadow), _p(step), _p(reg)] + extra + [c_void_p(stream)]
    rc = fn(*args)
    assert rc == 0, lib.dfm_last_error()


@pytest.mark.parametrize("opt_id,inits", [(0, (0.0, 0.0)), (1, (1e-8, 0.0)), (3, (0.1, 0.0))])
def test_static_scratch_entry_matches_caller_scratch_entry(cuda_lib, opt_id, inits):
    lib = cuda_lib
    lib.dfm_flat_opt_static_max_elems.restype = c_longlong
    lib.dfm_flat_opt_static_scratch_bytes.restype = c_longlong
    assert lib.dfm_flat_opt_static_max_elems() == 64 * 1024 * 1024
    assert lib.dfm_flat_opt_static_scratch_bytes() == 256 + 4 * 65536
    n = 1024 * 37 + 512  # not a multiple of the 1024-element CTA chunk
    gen = torch.Generator(device="cuda").manual_seed(5)
    w0 = torch.randn(n, device="cuda", generator=gen)
    state = {}
    for name in ("caller", "static"):
        state[name] = dict(w=w0.clone(), s0=torch.full((n,), inits[0], device="cuda"),
                           s1=torch.full((n,), inits[1], device="cuda"),
                           shadow=torch.zeros(4096, device="cuda", dtype=torch.bfloat16),
                           step=torch.zeros(1, device="cuda", dtype=torch.int64), reg=torch.zeros(1, device="cuda"))
    done = torch.zeros(1, device="cuda", dtype=torch.int32)
    stream = torch.cuda.current_stream().cuda_stream
    for it in range(2):  # two calls in a row: the second one sees the scratch the first one left behind
        grad = torch.randn(n, device="cuda", generator=gen)
        a, b = state["caller"], state["static"]
        ga, gb = grad.clone(), grad.clone()
        a["reg"].zero_(); b["reg"].zero_()
        _call(lib, lib.dfm_flat_opt_step, a["w"], ga, a["s0"], a["s1"], opt_id, a["shadow"], a["step"], a["reg"],
              [_p(done)], stream)
        _call(lib, lib.dfm_flat_opt_step_static, b["w"], gb, b["s0"], b["s1"], opt_id, b["shadow"], b["step"],
              b["reg"], [], stream)
        torch.cuda.synchronize()
        assert int(a["step"].item()) == int(b["step"].item()) == it + 1
        assert float(ga.abs().max()) == 0.0 and float(gb.abs().max()) == 0.0
        for key in ("w", "s0", "s1"):
            assert torch.allclose(b[key], a[key], rtol=2e-4, atol=2e-6), (key, it, (a[key] - b[key]).abs().max())
        assert torch.allclose(b["shadow"].float(), a["shadow"].float(), rtol=1e-2, atol=1e-3)
        assert torch.allclose(b["reg"], a["reg"], rtol=1e-4), (a["reg"], b["reg"])
        assert float(a["reg"]) > 0.0


def test_static_scratch_entry_rejects_sizes_beyond_its_scratch(cuda_lib):
    lib = cuda_lib
    lib.dfm_last_error.restype = ctypes.c_char_p
    lib.dfm_flat_opt_static_max_elems.restype = c_longlong
    too_big = lib.dfm_flat_opt_static_max_elems() + 4
    dummy = torch.zeros(8, device="cuda")
    step = torch.zeros(1, device="cuda", dtype=torch.int64)
    rc = lib.dfm_flat_opt_step_static(_p(dummy), _p(dummy), _p(dummy), _p(dummy), c_longlong(too_big), c_int(0),
                                      c_float(1e-2), c_float(0.9), c_float(0.999), c_float(1e-8), c_float(0.95),
                                      c_float(1.0), c_longlong(0), c_longlong(0), c_float(0.0), c_longlong(0),
                                      c_longlong(0), c_void_p(0), _p(step), c_void_p(0),
                                      c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 2  # DFM_ERR_ARG: refused before anything was launched
    assert b"exceeds the built-in scratch" in lib.dfm_last_error()
    torch.cuda.synchronize()
    assert int(step.item()) == 0
