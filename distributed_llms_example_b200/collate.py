"""Seq2seq collate with the output contract of ``DataCollatorForSeq2Seq(tokenizer, model=model)``.

Reference: train-torchrun.py:113, train-accelerator.py:155-159 (``pad_to_multiple_of=8`` only
under fp16), train-task.py:242 (``pad_to_multiple_of=8`` always) ->
SP/transformers/data/data_collator.py:529-615 plus T5's ``_shift_right``
(SP/transformers/models/t5/modeling_t5.py:595-613).

Integer host code, bit-exact by construction; it exists so that the step harness and the
benchmark can build batches without a tokenizer object (no SentencePiece model is available
offline). Right padding (T5's ``padding_side``), int64 outputs.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Sequence

import torch

__all__ = ["Seq2SeqCollator", "shift_right"]


def shift_right(labels: torch.Tensor, decoder_start_token_id: int = 0, pad_token_id: int = 0) -> torch.Tensor:
    """``T5ForConditionalGeneration.prepare_decoder_input_ids_from_labels`` (modeling_t5.py:595-613)."""
    shifted = labels.new_zeros(labels.shape)
    shifted[..., 1:] = labels[..., :-1].clone()
    shifted[..., 0] = decoder_start_token_id
    shifted.masked_fill_(shifted == -100, pad_token_id)
    return shifted


@dataclass
class Seq2SeqCollator:
    pad_token_id: int = 0
    label_pad_token_id: int = -100
    pad_to_multiple_of: Optional[int] = None
    decoder_start_token_id: int = 0
    pin_memory: bool = False

    def _round(self, n: int) -> int:
        m = self.pad_to_multiple_of
        return n if m is None else (n + m - 1) // m * m

    def __call__(self, features: Sequence[dict]) -> dict[str, torch.Tensor]:
        if not features:
            raise ValueError("cannot collate an empty batch")
        b = len(features)
        ls = self._round(max(len(f["input_ids"]) for f in features))
        input_ids = torch.full((b, ls), self.pad_token_id, dtype=torch.int64, pin_memory=self.pin_memory)
        attention = torch.zeros((b, ls), dtype=torch.int64, pin_memory=self.pin_memory)
        for i, f in enumerate(features):
            ids = torch.as_tensor(f["input_ids"], dtype=torch.int64)
            input_ids[i, : ids.numel()] = ids
            if "attention_mask" in f:
                attention[i, : ids.numel()] = torch.as_tensor(f["attention_mask"], dtype=torch.int64)
            else:
                attention[i, : ids.numel()] = 1
        batch = {"input_ids": input_ids, "attention_mask": attention}
        if "labels" in features[0] and features[0]["labels"] is not None:
            lt = self._round(max(len(f["labels"]) for f in features))
            labels = torch.full((b, lt), self.label_pad_token_id, dtype=torch.int64, pin_memory=self.pin_memory)
            for i, f in enumerate(features):
                lab = torch.as_tensor(f["labels"], dtype=torch.int64)
                labels[i, : lab.numel()] = lab
            batch["labels"] = labels
            batch["decoder_input_ids"] = shift_right(labels, self.decoder_start_token_id, self.pad_token_id)
        return batch
