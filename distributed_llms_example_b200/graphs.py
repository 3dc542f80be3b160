"""CUDA-graph replay of the model's forward + backward (on by default; ``B2_CUDA_GRAPH=0`` or
``patch.apply(cuda_graph=False)`` switches it off).

Why: one flan-t5-large step issues ~9 400 eager kernels; the step is bound by the host issuing them
(profiles/r1_bench_launch_list_summary.txt), so the kernels of this package cannot move samples/s by
more than their few per cent of the step. The callers on either side of the hot path
(``outputs = model(**batch)`` at train-accelerator.py:220 / train-task.py:289 / train-torchrun.py:140 and
``loss.backward()`` at train-accelerator.py:222 / train-task.py:292) run the same launch sequence every
step as long as the batch shapes repeat - and the scripts pad every example to ``max_length``
(train-torchrun.py:102, train-accelerator.py:115-127, train-task.py:157-168) - so that sequence is
captured once and replayed.

How it attaches: ``T5ForConditionalGeneration.forward`` still runs as written, but only while it is
being captured. ``patch.patch_cuda_graph()`` wraps it; a training call with labels on a CUDA model first
runs eagerly ``B2_GRAPH_EAGER_STEPS`` (2) times per shape signature, then
``torch.cuda.make_graphed_callables`` captures forward and backward (this package's CE / fused lm_head
kernels included - they are ordinary stream-ordered launches), and later calls with the same signature
replay. The capture is made in two segments, encoder | decoder + loss (``B2_GRAPH_SEGMENTS``): the decoder's
gradients then reach DDP before the encoder's backward graph is launched, so their buckets are
all-reduced while it runs (2 GPUs: 99.8 -> 97.6 ms per step). The returned object is a ``Seq2SeqLMOutput`` holding the loss only; anything that needs logits,
another shape, eval mode or extra arguments takes the eager path. DDP, the comm hook, AdamW and the clip
are untouched: gradients leave the graph through autograd's AccumulateGrad like any others.

One library line is not capturable: ``transformers.masking_utils.eager_mask`` builds a scalar with
``torch.tensor(0.0, device=...)`` (a pageable host-to-device copy). It is replaced, while this patch is
applied, by an arithmetic equivalent with the same values.
"""
from __future__ import annotations

import os
from typing import Any, Optional

import torch

__all__ = ["GraphedForward", "capture_safe_eager_mask", "install", "uninstall"]

_KEYS = ("input_ids", "attention_mask", "labels", "decoder_input_ids")
_saved: dict[str, Any] = {}


def capture_safe_eager_mask(*args, dtype: torch.dtype = torch.float32, **kwargs):
    """``transformers.masking_utils.eager_mask`` without the host scalar: 0 where attention is
    allowed, ``finfo(dtype).min`` elsewhere (same values, two device-side ops)."""
    from transformers import masking_utils as mu

    kwargs.pop("allow_is_causal_skip", None)
    kwargs.pop("allow_torch_fix", None)
    mask = mu.sdpa_mask(*args, allow_is_causal_skip=False, allow_torch_fix=False, **kwargs)
    if mask is not None:
        mask = torch.zeros_like(mask, dtype=dtype).masked_fill_(~mask, torch.finfo(dtype).min)
    return mask


def _lib_count(n: int) -> None:
    from . import _lib

    _lib.count_launch(n)


class _ThreadLocalGraph(torch.cuda.graph):
    """torch.cuda.graph with cudaStreamCaptureModeThreadLocal: make_graphed_callables captures in
    "global" mode, in which a CUDA call from ANY thread invalidates the capture - and a process with an
    NCCL process group has a watchdog thread that polls events all the time (every DDP run)."""

    def __init__(self, cuda_graph, pool=None, stream=None, capture_error_mode: str = "thread_local"):
        super().__init__(cuda_graph, pool=pool, stream=stream, capture_error_mode="thread_local")


class _LossOnly(torch.nn.Module):
    """Positional view of the model for make_graphed_callables: tensors in, the loss out."""

    def __init__(self, model, orig_forward):
        super().__init__()
        self.model, self._orig = model, orig_forward

    def forward(self, input_ids, attention_mask, labels, decoder_input_ids):
        return self._orig(self.model, input_ids=input_ids, attention_mask=attention_mask, labels=labels,
                          decoder_input_ids=decoder_input_ids).loss


class _Encoder(torch.nn.Module):
    """First segment of the two-segment capture: the encoder stack, exactly as T5's forward calls it."""

    def __init__(self, model):
        super().__init__()
        self.encoder = model.encoder

    def forward(self, input_ids, attention_mask):
        return self.encoder(input_ids=input_ids, attention_mask=attention_mask)[0]


class _DecoderLoss(torch.nn.Module):
    """Second segment: T5's own forward, given the encoder's output (its ``encoder_outputs`` argument)."""

    def __init__(self, model, orig_forward):
        super().__init__()
        self.model, self._orig = model, orig_forward

    def forward(self, encoder_hidden, attention_mask, labels, decoder_input_ids):
        return self._orig(self.model, encoder_outputs=(encoder_hidden,), attention_mask=attention_mask, labels=labels,
                          decoder_input_ids=decoder_input_ids).loss


class GraphedForward:
    """Per-model cache {shape signature -> graphed callable}; see the module docstring."""

    def __init__(self, model, orig_forward):
        self.model, self.orig = model, orig_forward
        self.eager_steps = int(os.environ.get("B2_GRAPH_EAGER_STEPS", "2"))
        # every captured signature keeps its own activations alive (~25 GB for flan-t5-large at batch 8):
        # ragged batches (collator padding to the longest) would multiply that, so only the first few
        # signatures are captured and the rest stay eager
        self.max_graphs = int(os.environ.get("B2_GRAPH_MAX", "2"))
        # 2 segments (encoder | decoder + loss): the decoder's gradients leave the replayed backward before
        # the encoder's backward starts, so under DDP their buckets are all-reduced WHILE the encoder's
        # backward graph runs instead of after the whole backward. 1 = one graph for everything.
        self.segments = int(os.environ.get("B2_GRAPH_SEGMENTS", "2"))
        self.seen: dict[tuple, int] = {}
        self.graphs: dict[tuple, Any] = {}
        self.failed: Optional[str] = None
        self.replays = 0
        self.kernels_per_replay = 0

    def signature(self, kwargs) -> Optional[tuple]:
        if self.failed or not self.model.training or not torch.is_grad_enabled():
            return None
        extra = [k for k, v in kwargs.items() if k not in _KEYS and v is not None and v is not False]
        if extra or any(kwargs.get(k) is None for k in _KEYS):
            return None
        ts = [kwargs[k] for k in _KEYS]
        if not all(isinstance(t, torch.Tensor) and t.is_cuda for t in ts):
            return None
        ac = torch.is_autocast_enabled("cuda")
        return tuple((tuple(t.shape), t.dtype) for t in ts) + (ac, torch.get_autocast_dtype("cuda") if ac else None)

    def __call__(self, kwargs):
        from transformers.modeling_outputs import Seq2SeqLMOutput

        sig = self.signature(kwargs)
        if sig is None:
            return self.orig(self.model, **kwargs)
        g = self.graphs.get(sig)
        if g is None:
            n = self.seen.get(sig, 0)
            self.seen[sig] = n + 1
            if n < self.eager_steps or len(self.graphs) >= self.max_graphs:
                return self.orig(self.model, **kwargs)
            g = self._capture(sig, kwargs)
            if g is None:
                return self.orig(self.model, **kwargs)
        ac, acd = sig[-2], sig[-1]
        with torch.autocast("cuda", dtype=acd, enabled=bool(ac), cache_enabled=False):
            if isinstance(g, tuple):
                hidden = g[0](kwargs["input_ids"], kwargs["attention_mask"])
                loss = g[1](hidden, kwargs["attention_mask"], kwargs["labels"], kwargs["decoder_input_ids"])
            else:
                loss = g(*(kwargs[k] for k in _KEYS))
        self.replays += 1
        _lib_count(self.kernels_per_replay)
        return Seq2SeqLMOutput(loss=loss)

    def _capture(self, sig, kwargs):
        from . import _lib

        ac, acd = sig[-2], sig[-1]
        sample = tuple(kwargs[k].clone() for k in _KEYS)
        warmup = 3
        launches0 = _lib.launches
        real_graph = torch.cuda.graph
        torch.cuda.graph = _ThreadLocalGraph  # see _ThreadLocalGraph
        try:
            with torch.autocast("cuda", dtype=acd, enabled=bool(ac), cache_enabled=False):
                if self.segments >= 2:
                    enc = _Encoder(self.model)
                    with torch.no_grad():
                        probe = enc(sample[0], sample[1])  # shape / dtype of the segment boundary
                    hidden = torch.zeros_like(probe).requires_grad_(True)
                    g = torch.cuda.make_graphed_callables(
                        (enc, _DecoderLoss(self.model, self.orig)),
                        ((sample[0], sample[1]), (hidden, sample[1], sample[2], sample[3])),
                        num_warmup_iters=warmup, allow_unused_input=True)
                else:
                    g = torch.cuda.make_graphed_callables(_LossOnly(self.model, self.orig), sample,
                                                          num_warmup_iters=warmup, allow_unused_input=True)
        except Exception as exc:
            # A capture that dies half way leaves torch's CUDA generator registered with a dead graph;
            # there is no clean way back to the eager path, so say so instead of training on.
            self.failed = f"{type(exc).__name__}: {exc}"
            raise RuntimeError("b200: CUDA-graph capture of the model's forward/backward failed; rerun with "
                               f"B2_CUDA_GRAPH=0 (cause: {self.failed[:400]})") from exc
        finally:
            torch.cuda.graph = real_graph
        # this package's kernels inside one forward+backward (warm-up iterations + the captured one all
        # went through the same Python adapters): replayed with the graph, counted per replay
        self.kernels_per_replay = (_lib.launches - launches0) // (warmup + 1)
        self.graphs[sig] = g
        return g


def install() -> None:
    """Wrap T5ForConditionalGeneration.forward and make eager_mask capturable (idempotent)."""
    from transformers import masking_utils as mu
    from transformers.models.t5 import modeling_t5 as t5

    if "t5_forward_graph" in _saved:
        return
    orig = t5.T5ForConditionalGeneration.forward
    _saved["t5_forward_graph"] = orig
    _saved["eager_mask"] = mu.AttentionMaskInterface._global_mapping["eager"]
    mu.AttentionMaskInterface._global_mapping["eager"] = capture_safe_eager_mask

    def forward(self, *args, **kwargs):
        if args or not self.lm_head.weight.is_cuda:
            return orig(self, *args, **kwargs)
        gf = self.__dict__.get("_b200_graphed")
        if gf is None or gf.orig is not orig:
            gf = GraphedForward(self, orig)
            self.__dict__["_b200_graphed"] = gf
        return gf(kwargs)

    forward.__wrapped__ = orig
    t5.T5ForConditionalGeneration.forward = forward


def uninstall() -> None:
    from transformers import masking_utils as mu
    from transformers.models.t5 import modeling_t5 as t5

    if "t5_forward_graph" in _saved:
        t5.T5ForConditionalGeneration.forward = _saved.pop("t5_forward_graph")
        mu.AttentionMaskInterface._global_mapping["eager"] = _saved.pop("eager_mask")
