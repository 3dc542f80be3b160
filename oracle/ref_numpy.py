"""CPU restatement (numpy) of the arithmetic on the reference's data-parallel hot path.

TEST INFRASTRUCTURE ONLY. Nothing under the product package imports this module; only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may.

Parity status: the reference (valohai/distributed-llms-example) ships no tests, fixtures or
golden vectors for this path ("parity unpinned" by the reference itself).
The algorithms live in its unpinned third-party dependencies; the versions on the image
are torch 2.11.0+cu128 and transformers 5.5.0. This restatement is therefore pinned against
outputs of THOSE libraries' CPU ops, generated in the build container by
oracle/make_golden.py and committed under tests/golden/ (tests/test_oracle.py checks it).

Each function cites the reference call site and the library lines it restates
(SP = site-packages of the image).
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def ulp_diff(a, b) -> np.ndarray:
    """Distance in units in the last place between two float32 arrays (test helper)."""
    a = np.asarray(a, F32).view(np.int32).astype(np.int64)
    b = np.asarray(b, F32).view(np.int32).astype(np.int64)
    a = np.where(a < 0, -(a & 0x7FFFFFFF), a)
    b = np.where(b < 0, -(b & 0x7FFFFFFF), b)
    return np.abs(a - b)


# --------------------------------------------------------------------------------------
# (a1/a2) label-masked cross-entropy
#   reference: train-accelerator.py:220-222, train-task.py:289-292, train-torchrun.py:140
#   library  : SP/transformers/models/t5/modeling_t5.py:1112-1117 ->
#              CrossEntropyLoss(ignore_index=-100) = log_softmax + nll_loss(mean over valid)
# --------------------------------------------------------------------------------------
def ce_forward(logits: np.ndarray, labels: np.ndarray, ignore_index: int = -100):
    """Returns (loss fp32 scalar, row_lse fp32 [N], n_valid). Math in float64, rounded once."""
    z = logits.astype(np.float64)
    m = z.max(axis=1, keepdims=True)
    lse = (m[:, 0] + np.log(np.exp(z - m).sum(axis=1)))
    valid = labels != ignore_index
    n_valid = int(valid.sum())
    rows = np.nonzero(valid)[0]
    picked = z[rows, labels[rows]]
    with np.errstate(invalid="ignore", divide="ignore"):
        loss = (lse[rows] - picked).sum() / n_valid if n_valid else np.float64("nan")
    return F32(loss), lse.astype(F32), n_valid


def ce_backward(logits: np.ndarray, labels: np.ndarray, grad_out: float = 1.0, ignore_index: int = -100):
    """d loss / d logits, float64 math: (softmax - onehot) * grad_out / n_valid, 0 on ignored rows."""
    z = logits.astype(np.float64)
    m = z.max(axis=1, keepdims=True)
    e = np.exp(z - m)
    p = e / e.sum(axis=1, keepdims=True)
    valid = labels != ignore_index
    n_valid = int(valid.sum())
    rows = np.nonzero(valid)[0]
    p[rows, labels[rows]] -= 1.0
    if n_valid:
        p[rows] *= np.float64(grad_out) / n_valid
    p[~valid] = 0.0  # torch writes exact zeros for ignored rows, also when every row is ignored
    return p


# --------------------------------------------------------------------------------------
# (a5) AdamW step
#   reference: train-accelerator.py:187,224; train-task.py:260,295; Trainer for train-torchrun.py:140
#   library  : SP/torch/optim/adam.py:436-536 (_single_tensor_adam, decoupled weight decay)
# --------------------------------------------------------------------------------------
def _fma32(a, b, c):
    # float32 fused multiply-add: the product of two float32 is exact in float64
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(F32)


def adamw_step(p, g, m, v, *, lr: float, beta1: float, beta2: float, eps: float,
               weight_decay: float, step: int, grad_scale: float | None = None):
    """One step on fp32 arrays, returns new (p, m, v). Same operation order and roundings as
    the single-tensor CPU path: mul_(1-lr*wd); lerp_(g, 1-b1) [an fma in ATen's vectorised
    kernel]; mul_(b2).addcmul_(g, g, 1-b2); sqrt()/bc2_sqrt + eps; addcdiv_(m, denom, -lr/bc1)."""
    p, g, m, v = (np.asarray(a, dtype=F32) for a in (p, g, m, v))
    if grad_scale is not None:
        g = g * F32(grad_scale)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    p = p * F32(1.0 - lr * weight_decay)
    m = _fma32(np.full_like(m, F32(1.0 - beta1)), g - m, m)
    v = v * F32(beta2) + (F32(1.0 - beta2) * g) * g
    denom = np.sqrt(v) / F32(bc2 ** 0.5) + F32(eps)
    p = p + (F32(-(lr / bc1)) * m) / denom
    return p.astype(F32), m.astype(F32), v.astype(F32)


def linear_schedule_factor(current_step: int, num_warmup_steps: int, num_training_steps: int) -> float:
    """get_scheduler('linear') LambdaLR factor (SP/transformers/optimization.py:101-104), used at
    train-accelerator.py:200-205, train-task.py:267-272; note factor(0) == 0 with 1 warm-up step."""
    if current_step < num_warmup_steps:
        return float(current_step) / float(max(1, num_warmup_steps))
    return max(0.0, float(num_training_steps - current_step) /
               float(max(1, num_training_steps - num_warmup_steps)))


# --------------------------------------------------------------------------------------
# (8f row 4) lm_head + cross-entropy as one op
#   reference: outputs = model(**batch); loss = outputs.loss at train-accelerator.py:220-221,
#              train-task.py:289-290, train-torchrun.py:140
#   library  : lm_logits = self.lm_head(sequence_output) then CrossEntropyLoss(ignore_index=-100)
#              (SP/transformers/models/t5/modeling_t5.py:1110-1117); under bf16 autocast the
#              matmul accumulates in fp32 and its output is rounded to bf16
# --------------------------------------------------------------------------------------
def lm_head_logits(hidden: np.ndarray, weight: np.ndarray, round_bf16: bool = True) -> np.ndarray:
    """hidden [N, D] @ weight[V, D]^T in float64 from bf16-valued inputs; with round_bf16 the result
    is rounded to bf16 (via fp32) like the library's bf16 GEMM output. Returned as float32."""
    z = (np.asarray(hidden, dtype=np.float64) @ np.asarray(weight, dtype=np.float64).T).astype(F32)
    return _round_bf16(z) if round_bf16 else z


def lm_head_ce_forward(hidden, weight, labels, ignore_index: int = -100, round_bf16: bool = True):
    """(loss, row_lse, n_valid) of ce_forward on lm_head_logits."""
    return ce_forward(lm_head_logits(hidden, weight, round_bf16), labels, ignore_index)


def lm_head_ce_backward(hidden, weight, labels, grad_out: float = 1.0, ignore_index: int = -100,
                        round_bf16: bool = True):
    """(dlogits, d hidden, d weight) in float64 math: dlogits from ce_backward, then the two
    contractions the chain rule gives (the bf16 rounding of the logits is treated as identity,
    as autograd treats the cast)."""
    dz = ce_backward(lm_head_logits(hidden, weight, round_bf16), labels, grad_out, ignore_index).astype(np.float64)
    return dz, dz @ np.asarray(weight, dtype=np.float64), dz.T @ np.asarray(hidden, dtype=np.float64)


# --------------------------------------------------------------------------------------
# (a3/a4) gradient mean all-reduce
#   reference: DDP Reducer via train-accelerator.py:189,222 / train-torchrun.py:127,140, and
#              train-task.py:65-69 (all_reduce SUM then "/= size")
#   library  : SP/torch/distributed/algorithms/ddp_comm_hooks/default_hooks.py:18-54
#              (tensor.div_(world) then all_reduce SUM)
# --------------------------------------------------------------------------------------
def allreduce_mean(bufs: list[np.ndarray], scale: float | None = None) -> np.ndarray:
    """sum_r (scale * b_r) accumulated in rank order 0..W-1 in float32 (scale = 1/W)."""
    w = len(bufs)
    s = F32(1.0 / w if scale is None else scale)
    acc = np.asarray(bufs[0], dtype=F32) * s
    for r in range(1, w):
        acc = acc + np.asarray(bufs[r], dtype=F32) * s
    return acc


def _round_bf16(x: np.ndarray) -> np.ndarray:
    """float32 -> nearest-even bfloat16 -> float32."""
    u = np.asarray(x, dtype=F32).view(np.uint32).astype(np.uint64)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000).astype(np.uint32)
    out = r.view(F32).copy()
    nan = np.isnan(x)
    out[nan] = np.nan
    return out


def allreduce_mean_bf16_wire(bufs: list[np.ndarray], scale: float | None = None) -> np.ndarray:
    """bf16 on the wire: bf16(sum in fp32, rank order, of bf16(scale * b_r))."""
    w = len(bufs)
    s = F32(1.0 / w if scale is None else scale)
    acc = _round_bf16(np.asarray(bufs[0], dtype=F32) * s)
    for r in range(1, w):
        acc = acc + _round_bf16(np.asarray(bufs[r], dtype=F32) * s)
    return _round_bf16(acc)


def grad_sumsq(x: np.ndarray) -> float:
    """sum(x*x) of float32 values, accumulated in float64: what the norm of
    torch.nn.utils.clip_grad_norm_ is the square root of (SP/transformers/trainer.py:1753-1757)."""
    x = np.asarray(x, dtype=np.float64).reshape(-1)
    return float(np.dot(x, x))


def clip_coef(total_norm: float, max_norm: float) -> float:
    """torch.nn.utils.clip_grad_norm_ coefficient: clamp(max_norm / (norm + 1e-6), max=1)
    (HF Trainer default max_grad_norm=1.0, SP/transformers/trainer.py:1753-1757)."""
    return min(1.0, max_norm / (total_norm + 1e-6))


# --------------------------------------------------------------------------------------
# (a6) seq2seq collate: integer work, bit-exact
#   reference: DataCollatorForSeq2Seq at train-torchrun.py:113, train-accelerator.py:155-159,
#              train-task.py:242
#   library  : SP/transformers/data/data_collator.py:529-615 and T5._shift_right
#              (SP/transformers/models/t5/modeling_t5.py:595-613)
# --------------------------------------------------------------------------------------
def shift_right(labels: np.ndarray, decoder_start_token_id: int = 0, pad_token_id: int = 0) -> np.ndarray:
    out = np.empty_like(labels)
    out[..., 1:] = labels[..., :-1]
    out[..., 0] = decoder_start_token_id
    out[out == -100] = pad_token_id
    return out


def collate_seq2seq(features: list[dict], pad_token_id: int = 0, label_pad_token_id: int = -100,
                    pad_to_multiple_of: int | None = None, decoder_start_token_id: int = 0) -> dict:
    """features: dicts with int lists 'input_ids', 'attention_mask', 'labels'. Right padding.
    input_ids/attention_mask are padded to the longest (rounded up to pad_to_multiple_of),
    labels to their own longest (rounded up likewise) with label_pad_token_id;
    decoder_input_ids = shift_right(labels)."""
    def rounded(n):
        if pad_to_multiple_of is None:
            return n
        return ((n + pad_to_multiple_of - 1) // pad_to_multiple_of) * pad_to_multiple_of

    b = len(features)
    ls = rounded(max(len(f["input_ids"]) for f in features))
    lt = rounded(max(len(f["labels"]) for f in features))
    input_ids = np.full((b, ls), pad_token_id, dtype=np.int64)
    attention = np.zeros((b, ls), dtype=np.int64)
    labels = np.full((b, lt), label_pad_token_id, dtype=np.int64)
    for i, f in enumerate(features):
        n = len(f["input_ids"])
        input_ids[i, :n] = f["input_ids"]
        attention[i, :n] = f["attention_mask"]
        labels[i, :len(f["labels"])] = f["labels"]
    return {"input_ids": input_ids, "attention_mask": attention, "labels": labels,
            "decoder_input_ids": shift_right(labels, decoder_start_token_id, pad_token_id)}
