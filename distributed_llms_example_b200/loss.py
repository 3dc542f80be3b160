"""Drop-in ``torch.nn.CrossEntropyLoss`` for the seq2seq LM head loss.

Reference: T5 builds ``CrossEntropyLoss(ignore_index=-100)`` by its module-level name and
calls ``loss_fct(lm_logits.view(-1, V), labels.view(-1))``
(SP/transformers/models/t5/modeling_t5.py:21,1112-1117), reached from
train-accelerator.py:220-222, train-task.py:289-292 and train-torchrun.py:140. Rebinding
``transformers.models.t5.modeling_t5.CrossEntropyLoss`` to this class (see patch.py) is the
whole integration.

Contract kept from torch: mean over rows whose target != ignore_index, NaN when there is
none, fp32 0-d loss (also for bf16 logits under autocast), gradient in the logits dtype.
bf16 logits are consumed as they are: no fp32 up-cast copy, no [N, V] log-prob tensor.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from . import _lib

__all__ = ["CrossEntropyLoss", "cross_entropy"]

_DTYPES = {torch.float32: _lib.B2_F32, torch.bfloat16: _lib.B2_BF16}
_workspaces: dict[tuple[int, int], torch.Tensor] = {}


def _workspace(device: torch.device, stream: int, nbytes: int) -> torch.Tensor:
    """Zero-initialised scratch per (device, stream); the kernel leaves its counter at zero."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 1 << 16), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


class _MaskedCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits: torch.Tensor, target: torch.Tensor, ignore_index: int):
        lib = _lib.load()
        n, v = logits.shape
        with torch.cuda.device(logits.device):
            stream = torch.cuda.current_stream().cuda_stream
            row_lse = torch.empty(n, dtype=torch.float32, device=logits.device)
            loss = torch.empty((), dtype=torch.float32, device=logits.device)
            nvalid = torch.empty(1, dtype=torch.float32, device=logits.device)
            need = lib.b2_ce_workspace_bytes(n, v)
            ws = _workspace(logits.device, stream, need)
            _lib.check(lib.b2_ce_forward(
                logits.data_ptr(), target.data_ptr(), n, v, logits.stride(0), ignore_index,
                _DTYPES[logits.dtype], row_lse.data_ptr(), loss.data_ptr(),
                nvalid.data_ptr(), ws.data_ptr(), ws.numel(), stream))
            _lib.count_launch()
        ctx.save_for_backward(logits, target, row_lse, nvalid)
        ctx.ignore_index = ignore_index
        return loss

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor):
        logits, target, row_lse, nvalid = ctx.saved_tensors
        lib = _lib.load()
        n, v = logits.shape
        with torch.cuda.device(logits.device):
            g = grad_output.to(torch.float32).contiguous()
            dlogits = torch.empty_strided(logits.shape, logits.stride(), dtype=logits.dtype,
                                          device=logits.device)
            _lib.check(lib.b2_ce_backward(
                logits.data_ptr(), target.data_ptr(), row_lse.data_ptr(), g.data_ptr(),
                nvalid.data_ptr(), n, v, logits.stride(0), ctx.ignore_index,
                _DTYPES[logits.dtype], dlogits.data_ptr(),
                torch.cuda.current_stream().cuda_stream))
            _lib.count_launch()
        return dlogits, None, None


def cross_entropy(input: torch.Tensor, target: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """``F.cross_entropy(input, target, ignore_index=..., reduction="mean")`` for [N, V] CUDA logits."""
    if not input.is_cuda:
        raise RuntimeError("b200 CrossEntropyLoss has no CPU path: logits are on " + str(input.device))
    if input.dim() != 2 or target.dim() != 1 or target.shape[0] != input.shape[0]:
        raise ValueError(f"expected logits [N, V] and target [N], got {tuple(input.shape)} and {tuple(target.shape)}")
    if input.dtype not in _DTYPES:
        raise NotImplementedError(f"b200 CrossEntropyLoss: logits dtype {input.dtype} (fp32 and bf16 only)")
    if target.dtype != torch.int64:
        raise RuntimeError(f"expected int64 target, got {target.dtype}")
    if target.device != input.device:
        raise RuntimeError("logits and target are on different devices")
    if input.stride(1) != 1 or (input.shape[0] > 1 and input.stride(0) < input.shape[1]):
        input = input.contiguous()
    return _MaskedCE.apply(input, target.contiguous(), int(ignore_index))


class CrossEntropyLoss(nn.Module):
    """Same constructor as ``torch.nn.CrossEntropyLoss``; only the configuration the reference
    uses is implemented (no class weights, mean reduction, no label smoothing) and anything
    else raises instead of silently falling back."""

    def __init__(self, weight: Optional[torch.Tensor] = None, size_average=None, ignore_index: int = -100,
                 reduce=None, reduction: str = "mean", label_smoothing: float = 0.0):
        super().__init__()
        if weight is not None:
            raise NotImplementedError("b200 CrossEntropyLoss: class weights are not implemented")
        if size_average is not None or reduce is not None:
            raise NotImplementedError("b200 CrossEntropyLoss: legacy size_average/reduce are not implemented")
        if reduction != "mean":
            raise NotImplementedError(f"b200 CrossEntropyLoss: reduction={reduction!r} (mean only)")
        if label_smoothing != 0.0:
            raise NotImplementedError("b200 CrossEntropyLoss: label_smoothing is not implemented")
        self.ignore_index = ignore_index
        self.reduction = reduction
        self.label_smoothing = 0.0

    def forward(self, input: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        if type(input) is not torch.Tensor:
            # lazy logits from a patched lm_head (fused.py): the logits are never materialised,
            # the matmul and the loss run as one tensor-core kernel
            from .fused import fused_lm_head_ce, lazy_parts

            parts = lazy_parts(input)
            if parts is not None:
                return fused_lm_head_ce(parts[0], parts[1], target, self.ignore_index)
        return cross_entropy(input, target, self.ignore_index)

    def extra_repr(self) -> str:
        return f"ignore_index={self.ignore_index}"
