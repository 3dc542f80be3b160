"""Drop-in ``torch.optim.AdamW`` whose step is one multi-tensor CUDA launch.

Reference call sites: ``AdamW(optimizer_grouped_parameters, lr=5e-5)`` at
train-accelerator.py:174-187 / train-task.py:247-260 with ``optimizer.step()`` at
train-accelerator.py:224 / train-task.py:295, and HF Trainer's
``AdamW(groups, lr=, betas=, eps=, fused=True)`` (SP/transformers/trainer_optimizer.py:201-208)
stepped at SP/transformers/trainer.py:1760 for train-torchrun.py:140.

Semantics follow SP/torch/optim/adam.py:215-272 (``Adam.step``) and ``_single_tensor_adam``
(:347-550): params without ``.grad`` are skipped, state is created lazily at the first
step, ``group["lr"]`` is re-read every step (LambdaLR mutates it; it is 0 at step 0 with the
scripts' schedule), state_dict layout matches torch's (``step`` fp32 CPU scalar tensor,
``exp_avg``, ``exp_avg_sq``) so checkpoints move both ways. There is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Any, Callable, Iterable, Optional

import numpy as np
import torch
from torch.optim.optimizer import Optimizer

from . import _lib

__all__ = ["AdamW"]

_TABLE_DTYPE = np.dtype(
    [("p", "<u8"), ("g", "<u8"), ("m", "<u8"), ("v", "<u8"), ("numel", "<i8"),
     ("hyper", "<i4"), ("chunk_begin", "<i4")])
assert _TABLE_DTYPE.itemsize == C.sizeof(_lib.AdamWTensor) == 48


class AdamW(Optimizer):
    def __init__(self, params: Iterable, lr: float = 1e-3, betas: tuple[float, float] = (0.9, 0.999),
                 eps: float = 1e-8, weight_decay: float = 1e-2, amsgrad: bool = False, *,
                 maximize: bool = False, foreach: Optional[bool] = None, capturable: bool = False,
                 differentiable: bool = False, fused: Optional[bool] = None):
        if isinstance(lr, torch.Tensor):
            raise ValueError("b200 AdamW: tensor lr is not supported (host scalar lr only)")
        if not 0.0 <= lr:
            raise ValueError(f"Invalid learning rate: {lr}")
        if not 0.0 <= eps:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 0: {betas[0]}")
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 1: {betas[1]}")
        if not 0.0 <= weight_decay:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        for name, flag in (("amsgrad", amsgrad), ("maximize", maximize),
                           ("capturable", capturable), ("differentiable", differentiable)):
            if flag:
                raise NotImplementedError(f"b200 AdamW: {name}=True is not implemented")
        # foreach / fused only select torch's implementation; accepted and ignored so that
        # Trainer's fused=True (trainer_optimizer.py:201-208) constructs this class unchanged.
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False,
                        maximize=False, foreach=None, capturable=False, differentiable=False,
                        fused=None, decoupled_weight_decay=True)
        super().__init__(params, defaults)
        self._lib = _lib.load()
        self._table_key: Optional[tuple] = None
        self._table_dev: Optional[torch.Tensor] = None
        self._table_pinned: Optional[torch.Tensor] = None
        self._table_event: Optional[torch.cuda.Event] = None
        self._n_tensors = 0
        self._total_chunks = 0
        #: optional device fp32 scalar multiplied into every gradient inside the kernel
        #: (gradient-clipping fold-in); set by clip helpers, cleared after each step.
        self.grad_scale: Optional[torch.Tensor] = None
        #: when set to a list, step() appends a (start, end) CUDA event pair recorded on the
        #: launching stream tightly around the kernel launch (bench.py's roofline timing).
        self.kernel_events: Optional[list] = None

    def __setstate__(self, state: dict[str, Any]) -> None:
        super().__setstate__(state)
        for group in self.param_groups:
            group.setdefault("amsgrad", False)
            group.setdefault("maximize", False)
            group.setdefault("decoupled_weight_decay", True)

    # -- table ---------------------------------------------------------------------
    def _gather(self):
        """Params with grads, lazily initialised state, and one hyper set per (group, step)."""
        rows = []          # (p, g, m, v, numel, hyper_idx)
        hypers: list[tuple] = []
        hyper_index: dict[tuple, int] = {}
        device = None
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            lr = group["lr"]
            if isinstance(lr, torch.Tensor):
                lr = float(lr)
            for p in group["params"]:
                g = p.grad
                if g is None:
                    continue
                if g.is_sparse:
                    raise RuntimeError("AdamW does not support sparse gradients")
                if not p.is_cuda:
                    raise RuntimeError("b200 AdamW has no CPU path: parameter is on " + str(p.device))
                if p.dtype != torch.float32 or g.dtype != torch.float32:
                    raise NotImplementedError(
                        f"b200 AdamW: fp32 params and grads only (got {p.dtype} / {g.dtype}); "
                        "AMP keeps fp32 master params, which is what the reference scripts train")
                if not p.is_contiguous():
                    raise NotImplementedError("b200 AdamW: non-contiguous parameter")
                if device is None:
                    device = p.device
                elif p.device != device:
                    raise NotImplementedError("b200 AdamW: parameters on more than one device")
                if not g.is_contiguous():
                    g = g.contiguous()
                    p.grad = g
                state = self.state[p]
                if len(state) == 0:  # adam.py:162-190
                    state["step"] = torch.tensor(0.0, dtype=torch.float32)
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                state["step"] += 1
                step = int(state["step"].item()) if state["step"].device.type == "cpu" else int(state["step"])
                key = (lr, beta1, beta2, group["eps"], group["weight_decay"], step)
                idx = hyper_index.get(key)
                if idx is None:
                    idx = hyper_index[key] = len(hypers)
                    hypers.append(key)
                rows.append((p.data_ptr(), g.data_ptr(), state["exp_avg"].data_ptr(),
                             state["exp_avg_sq"].data_ptr(), p.numel(), idx))
        return rows, hypers, device

    def _upload_table(self, rows, device) -> None:
        n = len(rows)
        host = np.zeros(n, dtype=_TABLE_DTYPE)
        arr = np.asarray(rows, dtype=np.int64)
        host["p"], host["g"], host["m"], host["v"] = (arr[:, 0].astype(np.uint64), arr[:, 1].astype(np.uint64),
                                                      arr[:, 2].astype(np.uint64), arr[:, 3].astype(np.uint64))
        host["numel"], host["hyper"] = arr[:, 4], arr[:, 5].astype(np.int32)
        total = self._lib.b2_adamw_plan(host.ctypes.data_as(C.POINTER(_lib.AdamWTensor)), n)
        if total < 0:
            _lib.check(int(total))
        nbytes = host.nbytes
        if self._table_pinned is None or self._table_pinned.numel() < nbytes:
            self._table_pinned = torch.empty(max(nbytes, 4096), dtype=torch.uint8, pin_memory=True)
            self._table_dev = torch.empty(max(nbytes, 4096), dtype=torch.uint8, device=device)
            self._table_event = torch.cuda.Event()
        else:
            self._table_event.synchronize()  # previous async upload has consumed the pinned buffer
        self._table_pinned[:nbytes].copy_(torch.from_numpy(host.view(np.uint8)))
        self._table_dev[:nbytes].copy_(self._table_pinned[:nbytes], non_blocking=True)
        self._table_event.record()
        self._n_tensors, self._total_chunks = n, int(total)

    # -- step ----------------------------------------------------------------------
    @torch.no_grad()
    def step(self, closure: Optional[Callable[[], float]] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        rows, hypers, device = self._gather()
        if not rows:
            return loss
        if len(hypers) > _lib.B2_ADAMW_MAX_HYPERS:
            raise NotImplementedError(
                f"b200 AdamW: {len(hypers)} distinct (group, step) hyper sets in one step; "
                f"the kernel takes {_lib.B2_ADAMW_MAX_HYPERS}")
        key = tuple(rows)
        with torch.cuda.device(device):
            if key != self._table_key:
                self._upload_table(rows, device)
                self._table_key = key
            hp = (_lib.AdamWHyper * len(hypers))()
            for i, (lr, b1, b2, eps, wd, step) in enumerate(hypers):
                hp[i] = _lib.AdamWHyper(lr, b1, b2, eps, wd, step)
            gs = self.grad_scale
            if gs is not None and (gs.dtype != torch.float32 or gs.device != device or gs.numel() != 1):
                raise ValueError("grad_scale must be a 1-element fp32 tensor on the parameters' device")
            ev = None
            if self.kernel_events is not None:
                ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                ev[0].record()
            _lib.check(self._lib.b2_adamw_step(
                self._table_dev.data_ptr(), self._n_tensors, self._total_chunks, hp, len(hypers),
                gs.data_ptr() if gs is not None else None,
                torch.cuda.current_stream(device).cuda_stream))
            _lib.count_launch()
            if ev is not None:
                ev[1].record()
                self.kernel_events.append(ev)
        self.grad_scale = None
        return loss
