"""Synthetic dialogue-summary pairs and T5-family configs.

No tokenizer files, datasets or checkpoints are reachable offline, so the benchmark and the
parity tests use token ids drawn uniformly from the vocabulary and random-init weights of
the published architectures. Shapes follow the reference: sources up to 512 tokens, targets
up to 128 (train-accelerator.py:115-127, train-task.py:157-168).

Label modes
  REF  : what the scripts really feed - targets tokenised with padding="max_length" and used
         as labels directly, i.e. right-padded with the pad id 0 to 128, no -100 anywhere
         (train-torchrun.py:102,107; train-accelerator.py:122-132; train-task.py:163-173).
  MASK : ragged targets padded by the collator with -100 (pad_to_multiple_of=8 as
         train-task.py:242), which exercises ignore_index.
"""
from __future__ import annotations

import torch

VOCAB = 32128
PAD, EOS = 0, 1
SRC_LEN, TGT_LEN = 512, 128

# Published hyper-parameters of the checkpoints the scripts are run with (no config.json offline).
T5_CONFIGS = {
    "t5-small": dict(d_model=512, d_ff=2048, num_layers=6, num_decoder_layers=6, num_heads=8,
                     feed_forward_proj="relu"),
    "flan-t5-base": dict(d_model=768, d_ff=2048, num_layers=12, num_decoder_layers=12, num_heads=12,
                         feed_forward_proj="gated-gelu"),
    "flan-t5-large": dict(d_model=1024, d_ff=2816, num_layers=24, num_decoder_layers=24, num_heads=16,
                          feed_forward_proj="gated-gelu"),
    "flan-t5-xl": dict(d_model=2048, d_ff=5120, num_layers=24, num_decoder_layers=24, num_heads=32,
                       feed_forward_proj="gated-gelu"),
}


def t5_config(name: str, dropout_rate: float = 0.1):
    from transformers import T5Config

    if name not in T5_CONFIGS:
        raise KeyError(f"unknown model {name!r}; known: {sorted(T5_CONFIGS)}")
    return T5Config(vocab_size=VOCAB, d_kv=64, pad_token_id=PAD, eos_token_id=EOS,
                    decoder_start_token_id=PAD, dropout_rate=dropout_rate, **T5_CONFIGS[name])


def build_model(name: str, dropout_rate: float = 0.1, seed: int = 0):
    """Random-init T5ForConditionalGeneration, same weights on every rank (seed 0)."""
    from transformers import T5ForConditionalGeneration

    torch.manual_seed(seed)
    return T5ForConditionalGeneration(t5_config(name, dropout_rate))


def make_features(n: int, rank: int = 0, label_mode: str = "REF", seed: int = 1234,
                  src_len: int = SRC_LEN, tgt_len: int = TGT_LEN) -> list[dict]:
    """n tokenised examples as dicts of int lists; generator seeded 1234 + rank (1234 is the
    reference's own partition seed, train-task.py:46)."""
    if label_mode not in ("REF", "MASK"):
        raise ValueError("label_mode must be REF or MASK")
    g = torch.Generator().manual_seed(seed + rank)
    feats = []
    for _ in range(n):
        ls = int(torch.randint(min(64, src_len - 1), src_len, (1,), generator=g))
        lt = int(torch.randint(min(8, tgt_len - 1), tgt_len, (1,), generator=g))
        src = torch.randint(2, 32100, (ls,), generator=g).tolist() + [EOS]
        tgt = torch.randint(2, 32100, (lt,), generator=g).tolist() + [EOS]
        # the scripts tokenise sources with padding="max_length" too
        attn = [1] * len(src) + [0] * (src_len - len(src))
        src = src + [PAD] * (src_len - len(src))
        if label_mode == "REF":
            tgt = tgt + [PAD] * (tgt_len - len(tgt))
        feats.append({"input_ids": src, "attention_mask": attn, "labels": tgt})
    return feats
