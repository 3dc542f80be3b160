"""Seeded inputs shared by oracle/make_golden.py and the tests (TEST INFRASTRUCTURE ONLY).

numpy's legacy RandomState stream is frozen by numpy's compatibility policy, so the same
seeds give the same bits on the build container and on the GPU box; only outputs are stored
in tests/golden/.
"""
from __future__ import annotations

import numpy as np


def ce_inputs(n: int, v: int, seed: int, frac_ignored: float):
    """logits ~ N(0,1)*4 fp32 [n, v]; labels uniform in [0, v) with -100 on a per-row basis
    (SURVEY.md section 8d microbench tensors)."""
    rs = np.random.RandomState(seed)
    logits = (rs.standard_normal((n, v)) * 4.0).astype(np.float32)
    labels = rs.randint(0, v, size=n).astype(np.int64)
    if frac_ignored >= 1.0:
        labels[:] = -100
    elif frac_ignored > 0:
        k = max(1, int(round(n * frac_ignored)))
        labels[rs.permutation(n)[:k]] = -100
    return logits, labels


def adamw_inputs(sizes, seed: int, steps: int):
    """params ~ N(0, 0.02), grads ~ N(0, 1e-2) per step, fp32."""
    rs = np.random.RandomState(seed)
    params = [(rs.standard_normal(s) * 0.02).astype(np.float32) for s in sizes]
    grads = [[(rs.standard_normal(s) * 1e-2).astype(np.float32) for s in sizes] for _ in range(steps)]
    return params, grads


def allreduce_inputs(world: int, n: int, seed: int):
    rs = np.random.RandomState(seed)
    return [(rs.standard_normal(n) * 1e-2).astype(np.float32) for _ in range(world)]


def collate_inputs(b: int, mode: str, seed: int = 41, src_len: int = 64, tgt_len: int = 24):
    """Tokenised examples like the scripts produce: sources padded to max_length with an
    attention mask; targets padded with pad id 0 (REF) or left ragged (MASK)."""
    rs = np.random.RandomState(seed + b)
    feats = []
    for _ in range(b):
        ls = int(rs.randint(8, src_len))
        lt = int(rs.randint(3, tgt_len))
        src = rs.randint(2, 32100, size=ls).tolist() + [1]
        tgt = rs.randint(2, 32100, size=lt).tolist() + [1]
        attn = [1] * len(src) + [0] * (src_len - len(src))
        src = src + [0] * (src_len - len(src))
        if mode == "REF":
            tgt = tgt + [0] * (tgt_len - len(tgt))
        feats.append({"input_ids": src, "attention_mask": attn, "labels": tgt})
    return feats


def lmce_inputs(n: int, v: int, d: int, seed: int, frac_ignored: float):
    """lm_head input [n, d] ~ N(0,1) and weight [v, d] ~ N(0, 4/sqrt(d)), both already rounded to
    bf16 values (stored as fp32), so logits ~ N(0, 4) like the CE microbench tensors; labels
    uniform in [0, v) with -100 on a per-row basis."""
    rs = np.random.RandomState(seed)

    def bf16(x):
        u = x.astype(np.float32).view(np.uint32).astype(np.uint64)
        return ((u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000).astype(np.uint32).view(np.float32)

    hidden = bf16(rs.standard_normal((n, d)))
    weight = bf16(rs.standard_normal((v, d)) * (4.0 / np.sqrt(d)))
    labels = rs.randint(0, v, size=n).astype(np.int64)
    if frac_ignored >= 1.0:
        labels[:] = -100
    elif frac_ignored > 0:
        k = max(1, int(round(n * frac_ignored)))
        labels[rs.permutation(n)[:k]] = -100
    return hidden, weight, labels
