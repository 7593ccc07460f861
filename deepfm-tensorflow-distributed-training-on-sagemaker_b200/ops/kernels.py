"""Thin typed wrappers over the C ABI (include/deepfm_b200.h): torch tensors in, kernels launched on the
current CUDA stream. No allocation, no synchronisation — callers own every buffer (CUDA-graph friendly)."""
from __future__ import annotations

import ctypes
from ctypes import c_float, c_int, c_longlong, c_uint64, c_void_p
from typing import Optional

import torch

from . import lib as _lib
from ..optim.tf1 import OPT_IDS


def _p(t: Optional[torch.Tensor]) -> c_void_p:
    return c_void_p(0 if t is None else t.data_ptr())


def _stream() -> c_void_p:
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _chk_dev(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("fused kernels need CUDA tensors")


def emb_fm_fwd(ids: torch.Tensor, vals: torch.Tensor, fm_w: torch.Tensor, fm_v: torch.Tensor, x0: torch.Tensor,
               S: torch.Tensor, yfm: torch.Tensor, zero_scalars: Optional[torch.Tensor] = None) -> None:
    """ids [B,F] int32, vals [B,F] f32, fm_w [N], fm_v [N,K] -> x0 [B, >=F*K] bf16, S [B,K] f32, yfm [B] f32."""
    _chk_dev(ids, vals, fm_w, fm_v, x0, S, yfm)
    assert ids.dtype == torch.int32 and vals.dtype == torch.float32 and x0.dtype == torch.bfloat16
    B, F = ids.shape
    N, K = fm_v.shape
    _lib.check(_lib.load().dfm_emb_fm_fwd(_p(ids), _p(vals), _p(fm_w), _p(fm_v), _p(x0), c_int(x0.stride(0)), _p(S),
                                          _p(yfm), c_int(B), c_int(F), c_int(K), c_int(N), _p(zero_scalars),
                                          c_int(zero_scalars.numel() if zero_scalars is not None else 0), _stream()))


def emb_bwd(ids, vals, fm_v, S, dy, dx0, g_w, g_v) -> None:
    _chk_dev(ids, vals, fm_v, S, dy, dx0, g_w, g_v)
    B, F = ids.shape
    N, K = fm_v.shape
    _lib.check(_lib.load().dfm_emb_bwd(_p(ids), _p(vals), _p(fm_v), _p(S), _p(dy), _p(dx0), c_int(dx0.stride(0)),
                                       _p(g_w), _p(g_v), c_int(B), c_int(F), c_int(K), c_int(N), _stream()))


class DetEmbBwd:
    """Deterministic embedding backward (csrc/embed_det.cu): `sort(ids, vals)` on a side stream at step start, `run(...)`
    in place of emb_bwd. Owns the workspace (sorted keys / indices, cub temporary storage, boundary records)."""

    def __init__(self, B: int, F: int, K: int, N: int, device):
        lib = _lib.load()
        lib.dfm_emb_bwd_det_workspace_bytes.restype = c_longlong
        self.B, self.F, self.K, self.N = B, F, K, N
        self.nbytes = int(lib.dfm_emb_bwd_det_workspace_bytes(c_int(B), c_int(F), c_int(K)))
        self.ws = torch.zeros(self.nbytes, dtype=torch.uint8, device=device)

    def sort(self, ids: torch.Tensor, vals: torch.Tensor) -> None:
        _lib.check(_lib.load().dfm_emb_sort_ids(_p(ids), _p(vals), c_int(self.B), c_int(self.F), c_int(self.K), c_int(self.N),
                                                _p(self.ws), c_longlong(self.nbytes), _stream()))

    def run(self, vals, fm_v, S, dy, dx0, g_w, g_v) -> None:
        _lib.check(_lib.load().dfm_emb_bwd_det(_p(vals), _p(fm_v), _p(S), _p(dy), _p(dx0), c_int(dx0.stride(0)), _p(g_w),
                                               _p(g_v), c_int(self.B), c_int(self.F), c_int(self.K), c_int(self.N),
                                               _p(self.ws), c_longlong(self.nbytes), _stream()))


class GemmPlan:
    """A tcgen05 GEMM bound to fixed buffers (owns the host-side TMA descriptors).

    mode 'fwd'  : out[M,N] bf16 = dropout(relu(A[M,K] @ B[N,K]^T + bias))
    mode 'dgrad': out[M,N] bf16 = (A[M,K] @ B[K,N]) * (mask != 0) * mask_scale
    mode 'wgrad': out[M,N] f32 += A[K,M]^T @ B[K,N]
    """
    MODES = {"fwd": 0, "dgrad": 1, "wgrad": 2}

    def __init__(self, mode: str, A: torch.Tensor, B: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int,
                 bias: Optional[torch.Tensor] = None, relu: bool = False, keep_prob: float = 1.0, seed: int = 0,
                 layer: int = 0, row_offset: int = 0, step_ptr: Optional[torch.Tensor] = None,
                 mask: Optional[torch.Tensor] = None, mask_scale: float = 1.0, split_k: int = 0):
        _chk_dev(A, B, out, bias, mask, step_ptr)
        assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16
        self._keep = (A, B, out, bias, step_ptr, mask)  # keep buffers alive as long as the descriptors exist
        self.mode, self.M, self.N, self.K = mode, M, N, K
        h = c_void_p()
        lib = _lib.load()
        _lib.check(lib.dfm_gemm_plan_create(
            ctypes.byref(h), c_int(self.MODES[mode]), _p(A), c_int(A.stride(0)), _p(B), c_int(B.stride(0)),
            c_int(M), c_int(N), c_int(K), _p(out), c_int(out.stride(0)), _p(bias), c_int(1 if relu else 0),
            c_float(keep_prob), c_uint64(seed), c_int(layer), c_int(row_offset), _p(step_ptr), _p(mask),
            c_int(mask.stride(0) if mask is not None else 0), c_float(mask_scale), c_int(split_k)))
        self._h = h

    def info(self):
        bn, gx, gy, gz = c_int(), c_int(), c_int(), c_int()
        _lib.load().dfm_gemm_plan_info(self._h, ctypes.byref(bn), ctypes.byref(gx), ctypes.byref(gy), ctypes.byref(gz))
        return {"bn": bn.value, "grid": (gx.value, gy.value, gz.value)}

    def set_overwrite(self, on: bool = True) -> "GemmPlan":
        """wgrad plans created with split_k=1: out = A^T B (plain stores) instead of out += (split-K atomics)."""
        _lib.check(_lib.load().dfm_gemm_plan_set_overwrite(self._h, c_int(1 if on else 0)))
        return self

    def set_exclusive(self, on: bool = True) -> "GemmPlan":
        """dgrad plans: no other GEMM runs beside this one, so it may take the whole SM for its TMA stages."""
        _lib.check(_lib.load().dfm_gemm_plan_set_exclusive(self._h, c_int(1 if on else 0)))
        return self

    def set_colsum(self, out: Optional[torch.Tensor]) -> "GemmPlan":
        """dgrad plans: out[n] += sum_m result[m,n] in the epilogue (the bias gradient of the layer below)."""
        self._keep = self._keep + (out,)
        _lib.check(_lib.load().dfm_gemm_plan_set_colsum(self._h, _p(out)))
        return self

    def set_trace_id(self, tid: int) -> "GemmPlan":
        """Timeline id of this plan's launches (only the trace build records anything; scripts/step_trace.py)."""
        _lib.check(_lib.load().dfm_gemm_plan_set_trace_id(self._h, c_int(tid)))
        return self

    def set_out_f32(self, on: bool = True) -> "GemmPlan":
        _lib.check(_lib.load().dfm_gemm_plan_set_out_f32(self._h, c_int(1 if on else 0)))
        return self

    def run(self) -> None:
        _lib.check(_lib.load().dfm_gemm_run(self._h, _stream()))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.load().dfm_gemm_plan_destroy(self._h)
                self._h = None
        except Exception:
            pass


def logit_loss(x, w_out, b_out, fm_bias, yfm, labels, inv_batch: float, mask_scale: float, train: bool,
               prob=None, logit=None, dy=None, dz=None, loss_sum=None, g_b_out=None, g_fm_bias=None, g_b_top=None,
               g_w_out=None) -> None:
    """g_b_top / g_w_out (train): also accumulate colsum(dZ) and X^T dy here instead of two colsum_bf16 launches."""
    B, H = x.shape
    _lib.check(_lib.load().dfm_logit_loss_cols(
        _p(x), c_int(x.stride(0)), c_int(H), _p(w_out), _p(b_out), _p(fm_bias), _p(yfm), _p(labels), c_int(B),
        c_float(inv_batch), c_float(mask_scale), c_int(1 if train else 0), _p(prob), _p(logit), _p(dy), _p(dz),
        c_int(dz.stride(0) if dz is not None else 0), _p(loss_sum), _p(g_b_out), _p(g_fm_bias), _p(g_b_top), _p(g_w_out),
        _stream()))


def colsum_bf16(a: torch.Tensor, out: torch.Tensor, w: Optional[torch.Tensor] = None) -> None:
    """out[j] += sum_b a[b,j] * (w[b] if w is given else 1)."""
    B, H = a.shape
    _lib.check(_lib.load().dfm_colsum_bf16(_p(a), c_int(a.stride(0)), c_int(B), c_int(H), _p(w), _p(out), _stream()))


def auc_hist(prob: torch.Tensor, labels: torch.Tensor, thr: torch.Tensor, hist: torch.Tensor) -> None:
    _lib.check(_lib.load().dfm_auc_hist(_p(prob), _p(labels), c_int(prob.numel()), _p(thr), c_int(thr.numel()),
                                        _p(hist), _stream()))


class MlpFwdPlan:
    """The whole MLP forward in one kernel (csrc/mlp_fused.cu): X_{l+1} = dropout(relu(X_l W_l^T + b_l)) for all layers, a
    thread-block cluster per 128 rows keeping the activations in distributed shared memory between layers. Same Philox
    streams as the per-layer `GemmPlan("fwd", ...)` chain, so both produce identical activations and masks.

    x: bf16 [M, K0] (view, stride = its padded buffer); weights[l]: bf16 [H_l, K_l]; outs[l]: bf16 [M, H_l] views."""

    def __init__(self, x: torch.Tensor, weights, outs, biases, M: int, relu: bool = True, keep_probs=None, seed: int = 0,
                 layer_ids=None, row_offset: int = 0, step_ptr: Optional[torch.Tensor] = None):
        L = len(weights)
        _chk_dev(x, step_ptr, *weights, *outs, *[b for b in biases if b is not None])
        self._keep = (x, tuple(weights), tuple(outs), tuple(biases), step_ptr)
        VP, IP, FP = c_void_p * L, c_int * L, c_float * L
        W = VP(*[w.data_ptr() for w in weights])
        O = VP(*[o.data_ptr() for o in outs])
        Bp = VP(*[(b.data_ptr() if b is not None else None) for b in biases])
        ldw = IP(*[w.stride(0) for w in weights])
        H = IP(*[w.shape[0] for w in weights])
        ldo = IP(*[o.stride(0) for o in outs])
        keep = FP(*[float(k) for k in (keep_probs or [1.0] * L)])
        ids = IP(*[int(i) for i in (layer_ids or range(L))])
        h = c_void_p()
        _lib.check(_lib.load().dfm_mlp_fwd_plan_create(
            ctypes.byref(h), _p(x), c_int(x.stride(0)), c_int(M), c_int(x.shape[1]), c_int(L), W, ldw, H, O, ldo, Bp,
            c_int(1 if relu else 0), keep, c_uint64(seed), ids, c_int(row_offset), _p(step_ptr)))
        self._h = h
        self.L = L

    @staticmethod
    def supported(widths, lds) -> bool:
        return 1 <= len(widths) <= 4 and all(h % 16 == 0 and 0 < h <= 448 for h in widths) and all(d % 8 == 0 for d in lds)

    def info(self):
        bn, cl, mt = c_int(), c_int(), c_int()
        _lib.load().dfm_mlp_fwd_plan_info(self._h, ctypes.byref(bn), ctypes.byref(cl), ctypes.byref(mt))
        return {"bn": bn.value, "cluster": cl.value, "m_tiles": mt.value}

    def set_trace_id(self, tid: int) -> "MlpFwdPlan":
        _lib.check(_lib.load().dfm_mlp_fwd_plan_set_trace_id(self._h, c_int(tid)))
        return self

    def run(self) -> None:
        _lib.check(_lib.load().dfm_mlp_fwd_run(self._h, _stream()))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.load().dfm_mlp_fwd_plan_destroy(self._h)
        except Exception:
            pass


def flat_opt_step(w, g, s0, s1, opt_name: str, lr: float, beta1: float, beta2: float, eps: float, momentum: float,
                  grad_scale: float, decay_range, l2: float, shadow_range, shadow, step_ptr, reg_loss,
                  done_counter, publish=None) -> None:
    """publish = (scalars [n] device f32, ring [R, n] PINNED HOST f32): the kernel's last CTA stores the scalars into
    ring[(step - 1) % R] directly (page-locked host memory is device-addressable under UVA)."""
    if publish is None:
        pub = (None, c_int(0), None, c_int(0))
    else:
        src, ring = publish
        assert src.is_cuda and ring.is_pinned() and ring.dtype == torch.float32 and ring.shape[1] == src.numel()
        pub = (_p(src), c_int(src.numel()), c_void_p(ring.data_ptr()), c_int(ring.shape[0]))
    _lib.check(_lib.load().dfm_flat_opt_step_pub(
        _p(w), _p(g), _p(s0), _p(s1), c_longlong(w.numel()), c_int(OPT_IDS[opt_name]), c_float(lr), c_float(beta1),
        c_float(beta2), c_float(eps), c_float(momentum), c_float(grad_scale), c_longlong(decay_range[0]),
        c_longlong(decay_range[1]), c_float(l2), c_longlong(shadow_range[0]), c_longlong(shadow_range[1]), _p(shadow),
        _p(step_ptr), _p(reg_loss), _p(done_counter), pub[0], pub[1], pub[2], pub[3], _stream()))


def cast_f32_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    _lib.check(_lib.load().dfm_cast_f32_bf16(_p(src), _p(dst), c_longlong(src.numel()), _stream()))


def bn_fwd(z, out, gamma, beta, mov_mean, mov_var, save_mean, save_rstd, decay: float, train: bool, keep_prob: float = 1.0,
           seed: int = 0, layer: int = 0, row_offset: int = 0, step_ptr: Optional[torch.Tensor] = None) -> None:
    """Batch norm after ReLU (+ inverted dropout in training), bf16 in/out, fp32 statistics (csrc/batchnorm.cu)."""
    B, H = z.shape
    _lib.check(_lib.load().dfm_bn_fwd(_p(z), c_int(z.stride(0)), _p(out), c_int(out.stride(0)), c_int(B), c_int(H),
                                      _p(gamma), _p(beta), _p(mov_mean), _p(mov_var), _p(save_mean), _p(save_rstd),
                                      c_float(decay), c_int(1 if train else 0), c_float(keep_prob), c_uint64(seed),
                                      c_int(layer), c_int(row_offset), _p(step_ptr), _stream()))


def bn_bwd(dy_inout, z, gamma, save_mean, save_rstd, g_gamma, g_beta) -> None:
    B, H = z.shape
    _lib.check(_lib.load().dfm_bn_bwd(_p(dy_inout), c_int(dy_inout.stride(0)), _p(z), c_int(z.stride(0)), c_int(B),
                                      c_int(H), _p(gamma), _p(save_mean), _p(save_rstd), _p(g_gamma), _p(g_beta),
                                      _stream()))


def example_parse(blob: torch.Tensor, offsets: torch.Tensor, n: int, field_size: int, labels: torch.Tensor,
                  ids: torch.Tensor, vals: torch.Tensor, err: torch.Tensor) -> None:
    """Decode n serialized tf.train.Example protos on the GPU into (labels, ids int32, vals); see csrc/parse_gpu.cu."""
    _chk_dev(blob, offsets, labels, ids, vals, err)
    _lib.check(_lib.load().dfm_example_parse(_p(blob), _p(offsets), c_int(n), c_int(field_size), _p(labels), _p(ids),
                                             _p(vals), _p(err), _stream()))


# ---------------------------------------------------------------------------------------------- split optimizer
def bitmap_words(N: int) -> int:
    return (N + 31) // 32


def touched_bitmap(ids: torch.Tensor, N: int, bitmap: torch.Tensor) -> None:
    """bitmap[id >> 5] |= 1 << (id & 31) for every in-range id (int32 [B,F]); bitmap int32 [ceil(N/32)]."""
    B, F = ids.shape
    _lib.check(_lib.load().dfm_touched_bitmap(_p(ids), c_int(B), c_int(F), c_int(N), _p(bitmap), _stream()))


class SplitOptimizer:
    """Bound arguments of the split dense optimizer (csrc/split_opt.cu) over one flat (w, g, s0, s1) buffer set:
    `sweep()` = every table row NOT in the touched bitmap (side stream, concurrent with forward/backward),
    `tail()`  = touched rows + fm_bias block + MLP block, then step counter / L2 loss value / result publication."""

    def __init__(self, layout, cfg, opt, w, g, s0, s1, shadow, step_ptr, reg_acc, reg_loss, done_counter, bm_union,
                 bm_in, clear_inputs: bool, loss_sum, loss_parts=None, world: int = 1, sweep_blocks: int = 0):
        seg_w, seg_v = layout.segments["fm_w"], layout.segments["fm_v"]
        self.o, self.cfg, self.L = opt, cfg, layout
        self.w, self.g, self.s0, self.s1, self.shadow = w, g, s0, s1, shadow
        self.step_ptr, self.reg_acc, self.reg_loss, self.done = step_ptr, reg_acc, reg_loss, done_counter
        self.bm_union, self.loss_sum, self.loss_parts, self.world = bm_union, loss_sum, loss_parts, world
        self._bm_in = bm_in  # keep alive
        self._bm_tab = (ctypes.c_void_p * len(bm_in))(*[t.data_ptr() if hasattr(t, "data_ptr") else int(t) for t in bm_in])
        self.clear_inputs = clear_inputs
        self.fmw_off, self.fmv_off = seg_w.offset, seg_v.offset
        self.N, self.K = cfg.feature_size, cfg.embedding_size
        self.d0 = (0, seg_w.offset)                          # fm_bias block (padded to the segment alignment)
        self.d1 = (layout.shadow_range[0], layout.trainable_numel)  # MLP weights, biases, head, BN beta/gamma
        self.sweep_blocks = sweep_blocks
        self.rows = torch.zeros(self.N, device=w.device, dtype=torch.int32)       # ascending touched-row ids
        self.row_count = torch.zeros(1, device=w.device, dtype=torch.int32)

    def _hyper(self):
        o = self.o
        return (c_int(OPT_IDS[o.name]), c_float(o.lr), c_float(o.beta1), c_float(o.beta2), c_float(o.eps),
                c_float(o.momentum))

    def union_rowlist(self) -> None:
        """One CTA: bm_union = OR of the input bitmaps; rows / row_count = ascending list of the rows set in it."""
        _lib.check(_lib.load().dfm_union_rowlist(self._bm_tab, c_int(len(self._bm_in)), c_int(1 if self.clear_inputs else 0),
                                                 c_int(self.N), _p(self.bm_union), _p(self.rows), _p(self.row_count),
                                                 _stream()))

    def sweep(self) -> None:
        _lib.check(_lib.load().dfm_sweep_untouched(
            _p(self.w), _p(self.s0), _p(self.s1), *self._hyper(), c_float(self.cfg.l2_reg), c_longlong(self.fmw_off),
            c_longlong(self.fmv_off), c_int(self.N), c_int(self.K), _p(self.bm_union), _p(self.step_ptr), _p(self.reg_acc),
            c_int(self.sweep_blocks), _stream()))

    def tail(self, publish=None, grad_scale: float = 1.0, dense: bool = True, sparse: bool = True,
             finalize: bool = True) -> None:
        if publish is None:
            pub = (None, c_int(0), None, c_int(0))
        else:
            src, ring = publish
            assert src.is_cuda and ring.is_pinned() and ring.dtype == torch.float32 and ring.shape[1] == src.numel()
            pub = (_p(src), c_int(src.numel()), c_void_p(ring.data_ptr()), c_int(ring.shape[0]))
        sh = self.L.shadow_range
        _lib.check(_lib.load().dfm_tail_opt(
            _p(self.w), _p(self.g), _p(self.s0), _p(self.s1), *self._hyper(), c_float(grad_scale),
            c_float(self.cfg.l2_reg), c_longlong(self.fmw_off), c_longlong(self.fmv_off), c_int(self.N), c_int(self.K),
            c_longlong(self.d0[0]), c_longlong(self.d0[1]), c_longlong(self.d1[0]), c_longlong(self.d1[1]),
            c_longlong(sh[0]), c_longlong(sh[1]), _p(self.shadow), _p(self.rows), _p(self.row_count), _p(self.step_ptr),
            _p(self.reg_acc),
            _p(self.reg_loss), _p(self.done), _p(self.loss_sum), _p(self.loss_parts), c_int(self.world), pub[0], pub[1],
            pub[2], pub[3], c_int(1 if dense else 0), c_int(1 if sparse else 0), c_int(1 if finalize else 0), _stream()))
