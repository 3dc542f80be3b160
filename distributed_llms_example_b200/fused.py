"""Fused lm_head + label-masked cross-entropy; on by default, ``B2_FUSED_LM_HEAD=0`` off.

Reference: T5 computes ``lm_logits = self.lm_head(sequence_output)`` and then
``CrossEntropyLoss(ignore_index=-100)(lm_logits.view(-1, V), labels.view(-1))``
(SP/transformers/models/t5/modeling_t5.py:1110-1117), reached from train-accelerator.py:220-221,
train-task.py:289-290 and train-torchrun.py:140. ``fused_lm_head_ce`` computes the same loss with
libb2ddp's tcgen05 kernel (csrc/lmhead_ce.cu): the [N, V] logits live in tensor memory one tile
at a time and never reach HBM in the forward pass; the backward recomputes them and writes
``dlogits`` once, in bf16, for the two library GEMMs that give ``d hidden`` and ``d weight``.

How it attaches to an unmodified model (``patch.patch_fused_lm_head``): T5's own ``forward``
still runs as written. While it runs with labels in training mode, ``self.lm_head`` answers with
a *lazy* logits tensor - a storage-less tensor subclass that remembers (hidden, weight) - and the
rebound ``CrossEntropyLoss`` recognises it and calls the fused kernel. Anything else that touches
the lazy tensor (``outputs.logits`` in an evaluation loop, ``argmax`` ...) materialises it with the
ordinary matmul first, so the model's outputs keep their meaning.
"""
from __future__ import annotations

import threading
from typing import Optional

import torch
import torch.nn.functional as F

from . import _lib

__all__ = ["fused_lm_head_ce", "LazyLogits", "lazy_lm_head_scope", "install_lazy_lm_head"]

_workspaces: dict[tuple[int, int], torch.Tensor] = {}


def _workspace(device: torch.device, stream: int, nbytes: int) -> torch.Tensor:
    key = (device.index if device.index is not None else torch.cuda.current_device(), stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 1 << 16), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


class _FusedLMHeadCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, hidden: torch.Tensor, weight: torch.Tensor, target: torch.Tensor, ignore_index: int,
                round_bf16: bool, keep_logits: bool):
        lib = _lib.load()
        n, d = hidden.shape
        v = weight.shape[0]
        hb = hidden.detach().to(torch.bfloat16).contiguous()
        wb = weight.detach().to(torch.bfloat16).contiguous()
        ld = (v + 7) // 8 * 8
        with torch.cuda.device(hidden.device):
            stream = torch.cuda.current_stream().cuda_stream
            row_lse = torch.empty(n, dtype=torch.float32, device=hidden.device)
            loss = torch.empty((), dtype=torch.float32, device=hidden.device)
            nvalid = torch.empty(1, dtype=torch.float32, device=hidden.device)
            logits = torch.empty((n, ld), dtype=torch.bfloat16, device=hidden.device) if keep_logits else None
            need = lib.b2_lmce_workspace_bytes(n, v)
            ws = _workspace(hidden.device, stream, need)
            _lib.check(lib.b2_lmce_forward(hb.data_ptr(), wb.data_ptr(), target.data_ptr(), n, v, d, ignore_index,
                                           1 if round_bf16 else 0, row_lse.data_ptr(), loss.data_ptr(),
                                           nvalid.data_ptr(), logits.data_ptr() if keep_logits else None, ld,
                                           ws.data_ptr(), ws.numel(), stream))
            _lib.count_launch(2)
        if keep_logits:
            ctx.save_for_backward(hb, wb, target, row_lse, nvalid, logits)
        else:
            ctx.save_for_backward(hb, wb, target, row_lse, nvalid)
        ctx.ignore_index, ctx.round_bf16, ctx.keep_logits = ignore_index, round_bf16, keep_logits
        ctx.hidden_dtype, ctx.weight_dtype = hidden.dtype, weight.dtype
        return loss

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor):
        hb, wb, target, row_lse, nvalid = ctx.saved_tensors[:5]
        lib = _lib.load()
        n, d = hb.shape
        v = wb.shape[0]
        ld = (v + 7) // 8 * 8
        with torch.cuda.device(hb.device):
            stream = torch.cuda.current_stream().cuda_stream
            g = grad_output.to(torch.float32).contiguous()
            dlogits = torch.empty((n, ld), dtype=torch.bfloat16, device=hb.device)
            if ctx.keep_logits:
                # the forward wrote the bf16 logits on their way out of tensor memory: the backward is
                # the HBM-bound masked-CE kernel on that buffer (one read, one write)
                logits = ctx.saved_tensors[5]
                _lib.check(lib.b2_ce_backward(logits.data_ptr(), target.data_ptr(), row_lse.data_ptr(), g.data_ptr(),
                                              nvalid.data_ptr(), n, v, ld, ctx.ignore_index, _lib.B2_BF16,
                                              dlogits.data_ptr(), stream))
            else:
                # nothing was kept: recompute the logits tile by tile on the tensor cores
                _lib.check(lib.b2_lmce_backward_dlogits(hb.data_ptr(), wb.data_ptr(), target.data_ptr(),
                                                        row_lse.data_ptr(), g.data_ptr(), nvalid.data_ptr(), n, v, d,
                                                        ctx.ignore_index, 1 if ctx.round_bf16 else 0,
                                                        dlogits.data_ptr(), ld, stream))
            _lib.count_launch()
            dl = dlogits[:, :v]
            d_hidden = d_weight = None
            if ctx.needs_input_grad[0]:
                d_hidden = (dl @ wb).to(ctx.hidden_dtype)          # [N, V] x [V, D], library GEMM
            if ctx.needs_input_grad[1]:
                d_weight = (dl.t() @ hb).to(ctx.weight_dtype)      # [V, N] x [N, D], library GEMM
        return d_hidden, d_weight, None, None, None, None


def fused_lm_head_ce(hidden: torch.Tensor, weight: torch.Tensor, target: torch.Tensor, ignore_index: int = -100,
                     round_bf16: bool = True, keep_logits: Optional[bool] = None) -> torch.Tensor:
    """``F.cross_entropy(F.linear(hidden, weight).view(-1, V), target.view(-1), ignore_index=...)`` with bf16
    tensor-core math and fp32 accumulation; ``hidden`` [..., D], ``weight`` [V, D], ``target`` [...].

    ``keep_logits=True`` (default, ``B2_LMCE_KEEP_LOGITS=0`` flips it): the forward also writes the bf16
    logits once (it never reads them) and the backward is the HBM-bound masked-CE kernel - the fastest
    pairing on B200. ``keep_logits=False``: no [N, V] tensor lives between forward and backward; the
    backward recomputes the logits on the tensor cores (N*V*2 bytes less memory, ~1.4x the time)."""
    if keep_logits is None:
        import os

        keep_logits = os.environ.get("B2_LMCE_KEEP_LOGITS", "1") != "0"
    if not hidden.is_cuda or not weight.is_cuda:
        raise RuntimeError("fused_lm_head_ce has no CPU path")
    if weight.dim() != 2 or hidden.shape[-1] != weight.shape[1]:
        raise ValueError(f"hidden {tuple(hidden.shape)} and weight {tuple(weight.shape)} do not match")
    if weight.shape[1] % 64 != 0:
        raise NotImplementedError(f"fused_lm_head_ce: d_model {weight.shape[1]} must be a multiple of 64")
    h2 = hidden.reshape(-1, hidden.shape[-1])
    t1 = target.reshape(-1)
    if t1.shape[0] != h2.shape[0] or t1.dtype != torch.int64:
        raise ValueError("target must be int64 with one entry per row of hidden")
    if keep_logits and not round_bf16:
        raise ValueError("keep_logits=True stores bf16 logits, so the loss must be taken on bf16 logits too "
                         "(round_bf16=True)")
    return _FusedLMHeadCE.apply(h2, weight, t1.contiguous(), int(ignore_index), bool(round_bf16), bool(keep_logits))


# ---- lazy logits: how the fused path reaches T5's unmodified forward ---------------------------
class LazyLogits(torch.Tensor):
    """A storage-less stand-in for ``F.linear(hidden, weight)``; see the module docstring."""

    @staticmethod
    def __new__(cls, hidden: torch.Tensor, weight: torch.Tensor):
        shape = tuple(hidden.shape[:-1]) + (weight.shape[0],)
        return torch.Tensor._make_wrapper_subclass(cls, shape, dtype=hidden.dtype, device=hidden.device,
                                                   requires_grad=False)

    def __init__(self, hidden: torch.Tensor, weight: torch.Tensor):
        self._b200_hidden, self._b200_weight = hidden, weight

    def materialize(self) -> torch.Tensor:
        return F.linear(self._b200_hidden, self._b200_weight.to(self._b200_hidden.dtype))

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        name = getattr(func, "__name__", "")
        self = args[0] if args and isinstance(args[0], LazyLogits) else None
        if self is not None and name in ("view", "reshape") and not kwargs:
            # T5: lm_logits.view(-1, lm_logits.size(-1)) - a reshape of the leading dims only stays lazy
            shape = tuple(args[1]) if (len(args) == 2 and isinstance(args[1], (tuple, list, torch.Size))) else args[1:]
            if len(shape) >= 1 and all(isinstance(d, int) for d in shape) and shape[-1] == self._b200_weight.shape[0]:
                hid = self._b200_hidden
                return LazyLogits(hid.reshape(*shape[:-1], hid.shape[-1]), self._b200_weight)
        if name in ("size", "dim", "__get__") and self is not None:
            with torch._C.DisableTorchFunctionSubclass():
                return func(*args, **kwargs)
        # anything else wants real logits
        args = tuple(a.materialize() if isinstance(a, LazyLogits) else a for a in args)
        kwargs = {k: (a.materialize() if isinstance(a, LazyLogits) else a) for k, a in kwargs.items()}
        return func(*args, **kwargs)

    @classmethod
    def __torch_dispatch__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        args = tuple(a.materialize() if isinstance(a, LazyLogits) else a for a in args)
        kwargs = {k: (a.materialize() if isinstance(a, LazyLogits) else a) for k, a in kwargs.items()}
        return func(*args, **kwargs)


_tls = threading.local()


class lazy_lm_head_scope:
    """While active on this thread, an lm_head installed by ``install_lazy_lm_head`` answers lazily."""

    def __init__(self, enabled: bool = True):
        self.enabled, self.prev = enabled, False

    def __enter__(self):
        self.prev = getattr(_tls, "active", False)
        _tls.active = self.enabled
        return self

    def __exit__(self, *exc):
        _tls.active = self.prev
        return False


def install_lazy_lm_head(linear: torch.nn.Linear) -> None:
    """Give ``linear`` (a bias-free lm_head) an instance-level forward that returns LazyLogits inside a
    ``lazy_lm_head_scope``; parameters, state_dict keys and weight tying are untouched."""
    if getattr(linear, "_b200_lazy", False):
        return
    if linear.bias is not None:
        raise NotImplementedError("fused lm_head + CE: an lm_head with a bias is not supported")
    cls_forward = type(linear).forward

    def forward(x: torch.Tensor) -> torch.Tensor:
        if getattr(_tls, "active", False) and x.is_cuda and linear.weight.shape[1] % 64 == 0:
            # only where the unfused matmul would run in bf16 anyway (bf16 autocast, or a bf16 model):
            # an fp32 model keeps its fp32 lm_head, this path never lowers precision by itself
            autocast_bf16 = torch.is_autocast_enabled("cuda") and torch.get_autocast_dtype("cuda") == torch.bfloat16
            if autocast_bf16 or (x.dtype == torch.bfloat16 and linear.weight.dtype == torch.bfloat16):
                return LazyLogits(x.to(torch.bfloat16), linear.weight)
        return cls_forward(linear, x)

    linear.forward = forward
    linear._b200_lazy = True


def lazy_parts(t: torch.Tensor) -> Optional[tuple]:
    """(hidden, weight) if ``t`` is a LazyLogits, else None (used by loss.CrossEntropyLoss)."""
    if isinstance(t, LazyLogits):
        return t._b200_hidden, t._b200_weight
    return None
