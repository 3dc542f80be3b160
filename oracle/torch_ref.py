"""The library ops the reference's hot path actually executes, wrapped once (CPU by default).

TEST INFRASTRUCTURE ONLY (see oracle/ref_numpy.py for the rule on who may import oracle/).

The reference scripts contain no arithmetic of their own: every number on the path is
produced by torch / transformers calls. These wrappers call exactly those library entry
points, so they are the closest thing to "the reference run here" that this image allows
(the scripts themselves cannot be imported: valohai, accelerate and evaluate are absent,
SURVEY.md section 0 F3-F5). oracle/make_golden.py uses them to generate tests/golden/.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def ce_ref(logits: torch.Tensor, labels: torch.Tensor, grad_out: float = 1.0, ignore_index: int = -100):
    """CrossEntropyLoss(ignore_index=-100) forward + backward as T5 calls it
    (SP/transformers/models/t5/modeling_t5.py:1112-1117). fp32 math on the given values
    (bf16 logits are up-cast first, which is what autocast does). Returns (loss, dlogits fp32)."""
    z = logits.detach().to(torch.float32).clone().requires_grad_(True)
    loss = F.cross_entropy(z, labels, ignore_index=ignore_index)
    (loss * grad_out).backward()
    return loss.detach(), z.grad.detach()


def adamw_ref(params, grads_per_step, *, lr_per_step, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
    """torch.optim.AdamW single-tensor path (foreach=False, fused=False: SP/torch/optim/adam.py
    :347-550), stepping through grads_per_step with lr_per_step[i] written to group['lr'] before
    step i, as LambdaLR does. Returns (params, exp_avg, exp_avg_sq) lists after the last step."""
    ps = [torch.nn.Parameter(p.detach().clone().to(torch.float32)) for p in params]
    opt = torch.optim.AdamW(ps, lr=lr_per_step[0], betas=betas, eps=eps, weight_decay=weight_decay,
                            foreach=False, fused=False)
    for lr, grads in zip(lr_per_step, grads_per_step):
        for g in opt.param_groups:
            g["lr"] = lr
        for p, g in zip(ps, grads):
            p.grad = g.detach().clone().to(torch.float32)
        opt.step()
    return ([p.detach() for p in ps], [opt.state[p]["exp_avg"] for p in ps],
            [opt.state[p]["exp_avg_sq"] for p in ps])


def allreduce_mean_ref(bufs):
    """default_hooks.allreduce_hook semantics: each rank divides by W, then SUM, accumulated in
    rank order in fp32 (SP/torch/distributed/algorithms/ddp_comm_hooks/default_hooks.py:18-54)."""
    w = len(bufs)
    acc = bufs[0].to(torch.float32) / w
    for b in bufs[1:]:
        acc = acc + b.to(torch.float32) / w
    return acc


def build_tokenizer(vocab_size: int = 32128):
    """In-memory word-level tokenizer with T5's special ids (pad 0, eos 1), right padding.
    Only used so that the real DataCollatorForSeq2Seq can be run offline."""
    from tokenizers import Tokenizer
    from tokenizers.models import WordLevel
    from transformers import PreTrainedTokenizerFast

    vocab = {"<pad>": 0, "</s>": 1, "<unk>": 2}
    tok = Tokenizer(WordLevel(vocab, unk_token="<unk>"))
    return PreTrainedTokenizerFast(tokenizer_object=tok, pad_token="<pad>", eos_token="</s>", unk_token="<unk>")


class _ShiftOnly:
    """Stands in for the model argument of DataCollatorForSeq2Seq: T5's own
    prepare_decoder_input_ids_from_labels without building a model."""

    def __init__(self, decoder_start_token_id=0, pad_token_id=0):
        from transformers import T5Config, T5ForConditionalGeneration

        cfg = T5Config(vocab_size=32, d_model=8, d_kv=4, d_ff=8, num_layers=1, num_heads=2,
                       pad_token_id=pad_token_id, decoder_start_token_id=decoder_start_token_id)
        self._m = T5ForConditionalGeneration(cfg)

    def prepare_decoder_input_ids_from_labels(self, labels):
        return self._m.prepare_decoder_input_ids_from_labels(labels)


def collate_ref(features, pad_to_multiple_of=None):
    """transformers.DataCollatorForSeq2Seq(tokenizer, model=model[, pad_to_multiple_of]) as at
    train-torchrun.py:113 / train-accelerator.py:155-159 / train-task.py:242."""
    from transformers import DataCollatorForSeq2Seq

    coll = DataCollatorForSeq2Seq(build_tokenizer(), model=_ShiftOnly(), pad_to_multiple_of=pad_to_multiple_of)
    return coll([dict(f) for f in features])
