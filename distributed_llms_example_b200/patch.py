"""Attaches the three plug-ins through public names, without editing the reference scripts.

Patch points
  * ``transformers.models.<family>.modeling_<family>.CrossEntropyLoss``: T5 (and BART, the
    default ``model_ckpt`` of valohai.yaml:8-20) build their loss from this module-level name
    (modeling_t5.py:21,1114), so rebinding it touches nothing else;
  * ``torch.optim.AdamW``: the scripts execute ``from torch.optim import AdamW`` at import
    (train-accelerator.py:14, train-task.py:17) and HF Trainer resolves it lazily
    (trainer_optimizer.py:203), so it must be rebound before the script module runs;
  * ``DistributedDataParallel.__init__``: registers the comm hook right after construction,
    because Accelerate's ``DistributedDataParallelKwargs(comm_hook=...)`` only carries built-in
    enum hooks;
  * ``average_gradients`` of train-task.py:65-69 is a function of the script itself; launch.py
    rebinds it in the script's namespace right after its definition.
"""
from __future__ import annotations

import importlib
import os
from typing import Any

import torch

_CE_MODULES = ("transformers.models.t5.modeling_t5", "transformers.models.bart.modeling_bart")
_saved: dict[str, Any] = {}


def patch_cross_entropy() -> None:
    from .loss import CrossEntropyLoss

    for name in _CE_MODULES:
        try:
            mod = importlib.import_module(name)
        except ImportError:
            continue
        if hasattr(mod, "CrossEntropyLoss"):
            _saved.setdefault(f"ce:{name}", mod.CrossEntropyLoss)
            mod.CrossEntropyLoss = CrossEntropyLoss


def patch_adamw() -> None:
    from .optim import AdamW

    _saved.setdefault("adamw", torch.optim.AdamW)
    torch.optim.AdamW = AdamW


_ddp_comm_kwargs: dict[str, Any] = {}


def _wants_hook(module, kwargs) -> bool:
    """A DDP module gets the peer-memory hook when its parameters live on a CUDA device and its
    process group has more than one rank (CPU / gloo modules and world size 1 are left alone)."""
    import torch.distributed as dist

    first = next(module.parameters(), None)
    return (first is not None and first.is_cuda and dist.is_initialized()
            and dist.get_world_size(kwargs.get("process_group")) > 1)


def patch_ddp(**comm_kwargs) -> None:
    """Every DistributedDataParallel built afterwards on a CUDA device with world size > 1 gets
    the peer-memory all-reduce hook (``comm_kwargs`` go to ``register_p2p_comm_hook``; calling
    this again only replaces them). Two constructor defaults change where the caller left them
    unset - none of the reference scripts sets either:

      * ``gradient_as_bucket_view=True``: ``p.grad`` aliases the Reducer's bucket, so the
        bucket -> ``p.grad`` copy-out after every all-reduce (one kernel per parameter, 2 x the
        gradient bytes) disappears and the optimizer's gradient addresses stay the same from
        step to step (``B2_DDP_BUCKET_VIEW=0`` keeps torch's default);
      * ``bucket_cap_mb`` = ``B2_DDP_BUCKET_MB`` (default 128): the peer-memory all-reduce has a
        fixed cost of ~25 us per bucket (gate, launch, final barrier) that torch's 25 MiB default
        cannot amortise (45-55 % of NVLink peak); at 128 MiB it runs at 65-80 %, the hook is
        called ~25 instead of ~95 times per step, and only the last bucket's ~0.3 ms is exposed
        after backward. ``B2_DDP_BUCKET_MB=0`` keeps torch's default.
    """
    from torch.nn.parallel import DistributedDataParallel as DDP

    _ddp_comm_kwargs.clear()
    _ddp_comm_kwargs.update(comm_kwargs)
    if "ddp_init" in _saved:
        return
    orig = DDP.__init__
    _saved["ddp_init"] = orig

    def __init__(self, module, *args, **kwargs):
        ours = _wants_hook(module, kwargs)
        if ours:
            cap = float(os.environ.get("B2_DDP_BUCKET_MB", "128") or 0)
            if cap > 0 and "bucket_cap_mb" not in kwargs:
                kwargs["bucket_cap_mb"] = cap
            if "gradient_as_bucket_view" not in kwargs and os.environ.get("B2_DDP_BUCKET_VIEW", "1") != "0":
                kwargs["gradient_as_bucket_view"] = True
        if ours and "cuda_graph" in _saved and next(module.parameters()).is_cuda:
            # DDP creates (and keeps alive) every parameter's AccumulateGrad node here, and a node remembers
            # the stream that was current at its creation. Created on the legacy default stream they make
            # any later CUDA-graph capture of the backward illegal ("would make the legacy stream depend on a
            # capturing stream"), so with graph replay on, DDP is constructed on a side stream - the
            # arrangement torch's own CUDA-graph + DDP recipe prescribes.
            side = torch.cuda.Stream(device=next(module.parameters()).device)
            side.wait_stream(torch.cuda.current_stream(side.device))
            with torch.cuda.stream(side):
                orig(self, module, *args, **kwargs)
            torch.cuda.current_stream(side.device).wait_stream(side)
        else:
            orig(self, module, *args, **kwargs)
        if ours:
            from . import comm

            self._b200_hook_state = comm.register_p2p_comm_hook(self, **_ddp_comm_kwargs)

    DDP.__init__ = __init__


def patch_clip_grad_norm() -> None:
    """``torch.nn.utils.clip_grad_norm_`` (what HF Trainer / Accelerate call, trainer.py:1753-1757)
    is routed to ``AdamW.clip_grad_norm_`` when the parameters are exactly those of a live b200
    AdamW and the call is the plain L2 clip. The scaling then happens inside the optimizer
    kernel: ``p.grad`` is NOT rewritten (nothing between clip and step reads it in the scripts)."""
    import torch.nn.utils as U

    from .optim import owner_of

    if "clip" in _saved:
        return
    orig = U.clip_grad_norm_
    _saved["clip"] = orig

    def clip_grad_norm_(parameters, max_norm, norm_type=2.0, error_if_nonfinite=False, foreach=None):
        params = [parameters] if isinstance(parameters, torch.Tensor) else list(parameters)
        if float(norm_type) == 2.0 and not error_if_nonfinite:
            opt = owner_of(params)
            if opt is not None:
                return opt.clip_grad_norm_(float(max_norm))
        return orig(params, max_norm, norm_type, error_if_nonfinite, foreach)

    U.clip_grad_norm_ = clip_grad_norm_


def patch_fused_lm_head() -> None:
    """On by default (``B2_FUSED_LM_HEAD=0`` or ``apply(fused_lm_head=False)`` switches it off): T5's ``lm_head`` + loss become
    one tcgen05 kernel (fused.py). ``T5ForConditionalGeneration.forward`` runs as written; it is only
    wrapped so that, when called with labels in training mode on a CUDA model, its ``lm_head`` answers
    with lazy logits that the rebound CrossEntropyLoss turns into the fused call. Needs
    ``patch_cross_entropy``."""
    from transformers.models.t5 import modeling_t5 as t5

    from .fused import install_lazy_lm_head, lazy_lm_head_scope

    if "t5_forward" in _saved:
        return
    orig = t5.T5ForConditionalGeneration.forward
    _saved["t5_forward"] = orig

    def forward(self, *args, **kwargs):
        lazy = (kwargs.get("labels") is not None and self.training and self.lm_head.weight.is_cuda
                and self.lm_head.bias is None and self.lm_head.weight.shape[1] % 64 == 0)
        if lazy:
            install_lazy_lm_head(self.lm_head)
        with lazy_lm_head_scope(lazy):
            return orig(self, *args, **kwargs)

    forward.__wrapped__ = orig
    t5.T5ForConditionalGeneration.forward = forward


def patch_cuda_graph() -> None:
    """On by default (``B2_CUDA_GRAPH=0`` or ``apply(cuda_graph=False)`` switches it off): the model's forward + backward are
    captured in CUDA graphs once the batch shapes repeat and replayed from then on (graphs.py). Applied
    after ``patch_fused_lm_head`` so that the captured forward already contains the fused kernel."""
    from . import graphs

    graphs.install()
    _saved.setdefault("cuda_graph", True)


def apply(cross_entropy: bool = True, adamw: bool = True, ddp: bool = True, fold_clip: bool = True,
          comm_kwargs: dict | None = None, fused_lm_head: bool | None = None,
          cuda_graph: bool | None = None) -> None:
    # Both extensions are on unless switched off: an unmodified script run through launch.py gets the
    # same configuration bench.py measures. Each falls back by itself where it does not apply (no bf16
    # autocast, another model family, changing batch shapes, extra forward arguments, eval mode).
    if fused_lm_head is None:
        fused_lm_head = os.environ.get("B2_FUSED_LM_HEAD", "1") != "0"
    if cuda_graph is None:
        cuda_graph = os.environ.get("B2_CUDA_GRAPH", "1") != "0"
    if cross_entropy:
        patch_cross_entropy()
    if fused_lm_head and cross_entropy:
        patch_fused_lm_head()
    if adamw:
        patch_adamw()
    if ddp:
        patch_ddp(**(comm_kwargs or {}))
    if fold_clip:
        patch_clip_grad_norm()
    if cuda_graph:
        patch_cuda_graph()


def remove() -> None:
    """Undo apply() (used by the benchmark to time the stock path in the same process)."""
    from torch.nn.parallel import DistributedDataParallel as DDP

    if _saved.pop("cuda_graph", None):  # outermost wrapper of T5's forward: comes off first
        from . import graphs

        graphs.uninstall()
    for key, val in list(_saved.items()):
        if key.startswith("ce:"):
            importlib.import_module(key[3:]).CrossEntropyLoss = val
        elif key == "adamw":
            torch.optim.AdamW = val
        elif key == "ddp_init":
            DDP.__init__ = val
        elif key == "t5_forward":
            from transformers.models.t5 import modeling_t5 as t5

            t5.T5ForConditionalGeneration.forward = val
        elif key == "clip":
            import torch.nn.utils as U

            U.clip_grad_norm_ = val
        del _saved[key]
