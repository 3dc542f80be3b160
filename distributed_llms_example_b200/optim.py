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
import sys
import weakref
from typing import Any, Callable, Iterable, Optional

import numpy as np
import torch
from torch.optim.optimizer import Optimizer

from . import _lib

__all__ = ["AdamW", "owner_of"]

_LIVE: "weakref.WeakSet[AdamW]" = weakref.WeakSet()


def owner_of(params: list) -> Optional["AdamW"]:
    """The live b200 AdamW whose parameters-with-gradients are exactly ``params``-with-gradients
    (used by patch.patch_clip_grad_norm to route torch.nn.utils.clip_grad_norm_)."""
    want = {id(p) for p in params if p.grad is not None}
    if not want:
        return None
    for opt in list(_LIVE):
        have = {id(p) for g in opt.param_groups for p in g["params"] if p.grad is not None}
        if have == want:
            return opt
    return None

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
        self._plan: Optional["_Plan"] = None
        self._all_params: Optional[list] = None
        self._table_dev: Optional[torch.Tensor] = None
        self._table_pinned: Optional[torch.Tensor] = None
        self._table_event: Optional[torch.cuda.Event] = None
        #: optional device fp32 scalar multiplied into every gradient inside the kernel
        #: (gradient-clipping fold-in); set by clip helpers, cleared after each step.
        self.grad_scale: Optional[torch.Tensor] = None
        #: when set to a list, step() appends a (start, end) CUDA event pair recorded on the
        #: launching stream tightly around the kernel launch (bench.py's roofline timing).
        self.kernel_events: Optional[list] = None
        self._norm_ws: Optional[torch.Tensor] = None
        #: where the last clip_grad_norm_ got its norm: "allreduce" (folded into the gradient
        #: all-reduce, no extra pass) or "kernel" (grad_sumsq_kernel, one read of the gradients)
        self.norm_source: Optional[str] = None
        _LIVE.add(self)

    _UNSUPPORTED = ("amsgrad", "maximize", "capturable", "differentiable")

    def _validate_groups(self) -> None:
        """The constructor's contract also holds for param_groups that arrive through
        load_state_dict / unpickling: a checkpoint written by torch.optim.AdamW(amsgrad=True) must not
        resume with different semantics in silence."""
        for i, group in enumerate(self.param_groups):
            for name in self._UNSUPPORTED:
                if group.get(name, False):
                    raise NotImplementedError(
                        f"b200 AdamW: param group {i} has {name}=True, which this optimizer does not implement "
                        "(the checkpoint was written by an optimizer with other semantics)")
            if isinstance(group.get("lr"), torch.Tensor):
                raise NotImplementedError("b200 AdamW: tensor lr is not supported (host scalar lr only)")

    def __setstate__(self, state: dict[str, Any]) -> None:
        super().__setstate__(state)
        self.__dict__.setdefault("_plan", None)
        self.__dict__.setdefault("_all_params", None)
        self.__dict__.setdefault("grad_scale", None)
        self.__dict__.setdefault("kernel_events", None)
        self.__dict__.setdefault("_norm_ws", None)
        self.__dict__.setdefault("norm_source", None)
        for group in self.param_groups:
            group.setdefault("amsgrad", False)
            group.setdefault("maximize", False)
            group.setdefault("decoupled_weight_decay", True)
        self._validate_groups()

    # -- plan: everything about a step that does not change from one step to the next -----
    def add_param_group(self, param_group: dict[str, Any]) -> None:
        super().add_param_group(param_group)
        self._plan = None
        self._all_params = None

    def load_state_dict(self, state_dict: dict[str, Any]) -> None:
        for i, group in enumerate(state_dict.get("param_groups", [])):  # refuse BEFORE mutating self
            for name in self._UNSUPPORTED:
                if group.get(name, False):
                    raise NotImplementedError(
                        f"b200 AdamW: the state dict's param group {i} has {name}=True, which this optimizer "
                        "does not implement")
        super().load_state_dict(state_dict)
        self._validate_groups()
        self._plan = None
        self._all_params = None

    @staticmethod
    def _check_grad(p: torch.Tensor, g: torch.Tensor) -> None:
        if g.is_sparse:
            raise RuntimeError("AdamW does not support sparse gradients")
        if g.dtype != torch.float32:
            raise NotImplementedError(f"b200 AdamW: fp32 gradients only (got {g.dtype})")
        if g.device != p.device:
            raise RuntimeError("b200 AdamW: gradient and parameter are on different devices")

    def _rebuild_plan(self, all_params, grads, pattern):
        """Slow path: validate every tensor, create missing state (adam.py:162-190), build the
        host table. Runs at the first step and whenever the set of params with grads changes."""
        live, group_of, device = [], [], None
        gi_of = {}
        for gi, group in enumerate(self.param_groups):
            for p in group["params"]:
                gi_of[id(p)] = gi
        for p, g in zip(all_params, grads):
            if g is None:
                continue
            if not p.is_cuda:
                raise RuntimeError("b200 AdamW has no CPU path: parameter is on " + str(p.device))
            if p.dtype != torch.float32:
                raise NotImplementedError(
                    f"b200 AdamW: fp32 parameters only (got {p.dtype}); AMP keeps fp32 master "
                    "params, which is what the reference scripts train")
            if not p.is_contiguous():
                raise NotImplementedError("b200 AdamW: non-contiguous parameter")
            if device is None:
                device = p.device
            elif p.device != device:
                raise NotImplementedError("b200 AdamW: parameters on more than one device")
            self._check_grad(p, g)
            state = self.state[p]
            if len(state) == 0:
                state["step"] = torch.tensor(0.0, dtype=torch.float32)
                state["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
            live.append(p)
            group_of.append(gi_of[id(p)])
        plan = _Plan()
        plan.pattern, plan.params, plan.device = pattern, live, device
        n = len(live)
        plan.group_of = np.asarray(group_of, dtype=np.int64)
        # torch keeps one 0-d CPU tensor per parameter for state["step"]; bumping 557 of them costs
        # ~0.9 ms per step even with _foreach_add_. Re-home them as views of ONE flat tensor so the
        # bump is a single vector add; state_dict() still sees a 0-d fp32 tensor per parameter.
        old = [float(self.state[p]["step"]) for p in live]
        plan.step_flat = torch.tensor(old, dtype=torch.float32)
        for i, p in enumerate(live):
            self.state[p]["step"] = plan.step_flat[i]
        plan.steps = np.asarray([int(v) for v in old], dtype=np.int64)
        host = np.zeros(n, dtype=_TABLE_DTYPE)
        host["p"] = [p.data_ptr() for p in live]
        host["m"] = [self.state[p]["exp_avg"].data_ptr() for p in live]
        host["v"] = [self.state[p]["exp_avg_sq"].data_ptr() for p in live]
        host["numel"] = [p.numel() for p in live]
        # one hyper class per distinct (group, step); all members advance together afterwards
        keys = plan.group_of * (1 << 40) + plan.steps
        uniq, inverse = np.unique(keys, return_inverse=True)
        if len(uniq) > _lib.B2_ADAMW_MAX_HYPERS:
            raise NotImplementedError(
                f"b200 AdamW: {len(uniq)} distinct (group, step) hyper sets in one step; "
                f"the kernel takes {_lib.B2_ADAMW_MAX_HYPERS}")
        host["hyper"] = inverse.astype(np.int32)
        plan.class_group = (uniq >> 40).astype(np.int64)
        plan.class_step = (uniq & ((1 << 40) - 1)).astype(np.int64)
        total = self._lib.b2_adamw_plan(host.ctypes.data_as(C.POINTER(_lib.AdamWTensor)), n) if n else 0
        if total < 0:
            _lib.check(int(total))
        plan.host, plan.total_chunks, plan.n = host, int(total), n
        plan.dirty = True
        return plan

    def _upload_table(self, plan) -> None:
        host = plan.host
        nbytes = host.nbytes
        if self._table_pinned is None or self._table_pinned.numel() < nbytes:
            self._table_pinned = torch.empty(max(nbytes, 4096), dtype=torch.uint8, pin_memory=True)
            self._table_dev = torch.empty(max(nbytes, 4096), dtype=torch.uint8, device=plan.device)
            self._table_event = torch.cuda.Event()
        else:
            self._table_event.synchronize()  # previous async upload has consumed the pinned buffer
        self._table_pinned[:nbytes].copy_(torch.from_numpy(host.view(np.uint8)))
        self._table_dev[:nbytes].copy_(self._table_pinned[:nbytes], non_blocking=True)
        self._table_event.record()
        plan.dirty = False

    def _refresh(self, check_params: bool = True) -> Optional["_Plan"]:
        """Current launch plan with gradient pointers re-read (cheap when nothing moved)."""
        all_params = self._all_params
        if all_params is None:
            all_params = self._all_params = [p for group in self.param_groups for p in group["params"]]
        grads = [p.grad for p in all_params]
        pattern = [g is not None for g in grads]
        plan = self._plan
        if plan is None or plan.pattern != pattern:
            plan = self._plan = self._rebuild_plan(all_params, grads, pattern)
        n = plan.n
        if n == 0:
            return None
        live = plan.params
        live_grads = grads if n == len(grads) else [g for g in grads if g is not None]
        gptrs = np.array(list(map(torch.Tensor.data_ptr, live_grads)), dtype=np.uint64)
        host = plan.host
        changed = np.nonzero(gptrs != host["g"])[0]
        if changed.size:
            for i in changed:  # new gradient tensors: re-validate just those
                g = live_grads[i]
                self._check_grad(live[i], g)
                if not g.is_contiguous():
                    g = live[i].grad = g.contiguous()
                    gptrs[i] = g.data_ptr()
            host["g"] = gptrs
            plan.dirty = True
        if check_params:
            pptrs = np.array(list(map(torch.Tensor.data_ptr, live)), dtype=np.uint64)
            if not np.array_equal(pptrs, host["p"]):  # p.data was re-homed (e.g. module.to(...))
                plan = self._plan = self._rebuild_plan(all_params, grads, pattern)
                plan.host["g"] = gptrs
        return plan

    # -- gradient clipping folded into the step ------------------------------------------
    @torch.no_grad()
    def clip_grad_norm_(self, max_norm: float) -> torch.Tensor:
        """``torch.nn.utils.clip_grad_norm_(params, max_norm)`` (norm_type 2) for this optimizer's
        parameters, as HF Trainer calls it before ``optimizer.step()`` (trainer.py:1753-1757).
        One read-only kernel computes the global norm and ``min(1, max_norm / (norm + 1e-6))`` on
        the device; the multiply happens inside the next ``step()``'s kernel, so ``p.grad`` itself
        is left unscaled. Returns the total norm as a 0-d device tensor (no host sync)."""
        plan = self._refresh(check_params=False)  # this kernel only reads gradients
        if plan is None:
            return torch.zeros((), dtype=torch.float32)
        device = plan.device
        # At world size > 1 the all-reduce has already summed the squares of every reduced gradient
        # element while it held them in registers: if its count covers exactly
        # this optimizer's gradients, the norm costs a one-thread kernel and no read of the gradients.
        comm_mod = sys.modules.get(__package__ + ".comm")
        if comm_mod is not None:
            total = int(plan.host["numel"].sum())
            for comm in comm_mod.live_comms():
                if comm.world < 2 or comm.device != device:
                    continue
                if not comm.sumsq_enabled:
                    comm.enable_sumsq(True)  # from the next step on (every rank gets here together)
                    continue
                if comm.sumsq_numel() == total:
                    out = torch.empty(2, dtype=torch.float32, device=device)  # [norm, coef]
                    with torch.cuda.device(device):
                        comm.sumsq_clip_(float(max_norm), out)
                    self.grad_scale = out[1:2]
                    self.norm_source = "allreduce"
                    return out[0]
                comm.enable_sumsq(True)  # count does not match these gradients: reset, use the kernel below
        self.norm_source = "kernel"
        with torch.cuda.device(device):
            if plan.dirty:
                self._upload_table(plan)
            if self._norm_ws is None or self._norm_ws.device != device:
                self._norm_ws = torch.zeros(self._lib.b2_grad_norm_workspace_bytes(), dtype=torch.uint8,
                                            device=device)
            out = torch.empty(2, dtype=torch.float32, device=device)  # [norm, coef]
            _lib.check(self._lib.b2_grad_norm_clip(
                self._table_dev.data_ptr(), plan.n, plan.total_chunks, float(max_norm),
                self._norm_ws.data_ptr(), self._norm_ws.numel(), out.data_ptr(), out.data_ptr() + 4,
                torch.cuda.current_stream(device).cuda_stream))
            _lib.count_launch()
        self.grad_scale = out[1:2]
        return out[0]

    def zero_grad(self, set_to_none: bool = True) -> None:
        """As torch, and drops a clip coefficient that was computed for these gradients but never
        consumed by a step (so it cannot leak into the next iteration)."""
        self.grad_scale = None
        super().zero_grad(set_to_none=set_to_none)

    # -- step ----------------------------------------------------------------------
    @torch.no_grad()
    def step(self, closure: Optional[Callable[[], float]] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        plan = self._refresh()
        if plan is None:
            return loss
        comm_mod = sys.modules.get(__package__ + ".comm")
        if comm_mod is not None:
            comm_mod.poll_all()  # never apply gradients a timed-out all-reduce filled with NaN
        n = plan.n
        # state["step"] += 1 for every live param, in one call; python mirror for the scalars
        plan.step_flat.add_(1)
        plan.steps += 1
        plan.class_step += 1
        n_h = len(plan.class_group)
        hp = (_lib.AdamWHyper * n_h)()
        for i in range(n_h):
            group = self.param_groups[int(plan.class_group[i])]
            lr = group["lr"]
            hp[i] = _lib.AdamWHyper(float(lr), group["betas"][0], group["betas"][1], group["eps"],
                                    group["weight_decay"], int(plan.class_step[i]))
        device = plan.device
        with torch.cuda.device(device):
            if plan.dirty:
                self._upload_table(plan)
            gs = self.grad_scale
            if gs is not None and (gs.dtype != torch.float32 or gs.device != device or gs.numel() != 1):
                raise ValueError("grad_scale must be a 1-element fp32 tensor on the parameters' device")
            ev = None
            if self.kernel_events is not None:
                ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                ev[0].record()
            _lib.check(self._lib.b2_adamw_step(
                self._table_dev.data_ptr(), n, plan.total_chunks, hp, n_h,
                gs.data_ptr() if gs is not None else None,
                torch.cuda.current_stream(device).cuda_stream))
            _lib.count_launch()
            if ev is not None:
                ev[1].record()
                self.kernel_events.append(ev)
        self.grad_scale = None
        return loss


class _Plan:
    """Cached per-optimizer launch plan (see AdamW._rebuild_plan)."""

    __slots__ = ("pattern", "params", "device", "group_of", "step_flat", "steps", "host", "class_group",
                 "class_step", "total_chunks", "n", "dirty", "grads_checked")
