"""Generates tests/golden/*.npz from the image's torch 2.11.0 / transformers 5.5.0 CPU ops.

Run in the build container:  python oracle/make_golden.py
The fixtures travel with the repo; the GPU box never needs /root/reference or this script.
Inputs are regenerated in the tests from numpy's legacy RandomState (bit-stable across
numpy versions) with the seeds stored next to the outputs, so only outputs are committed.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from oracle import torch_ref  # noqa: E402
from oracle.inputs import adamw_inputs, allreduce_inputs, ce_inputs, collate_inputs, lmce_inputs  # noqa: E402

GOLD = ROOT / "tests" / "golden"


def bf16_round(x: np.ndarray) -> np.ndarray:
    return torch.from_numpy(x).to(torch.bfloat16).to(torch.float32).numpy()


def gen_ce():
    out = {}
    for name, (n, v, seed, frac_ignored, dtype) in CE_CASES.items():
        logits, labels = ce_inputs(n, v, seed, frac_ignored)
        if dtype == "bf16":
            logits = bf16_round(logits)
        for g in (1.0, 0.0625):
            loss, dz = torch_ref.ce_ref(torch.from_numpy(logits), torch.from_numpy(labels), grad_out=g)
            out[f"{name}/g{g}/loss"] = loss.numpy()
            dz = dz.numpy()
            # whole gradient for small cases, a fixed sample + checksums for vocab-sized ones
            if dz.size <= 1 << 16:
                out[f"{name}/g{g}/dlogits"] = dz
            else:
                idx = np.random.RandomState(seed + 1).randint(0, dz.size, size=4096)
                out[f"{name}/g{g}/sample_idx"] = idx
                out[f"{name}/g{g}/sample_val"] = dz.reshape(-1)[idx]
                out[f"{name}/g{g}/row_abs_sum"] = np.abs(dz.astype(np.float64)).sum(axis=1)
        z = torch.from_numpy(logits)
        out[f"{name}/row_lse"] = torch.logsumexp(z, dim=1).numpy()
        out[f"{name}/meta"] = np.array([n, v, seed, int(frac_ignored * 100), dtype == "bf16"], dtype=np.int64)
    np.savez_compressed(GOLD / "ce.npz", **out)


CE_CASES = {
    # name: (rows, vocab, seed, fraction of -100 labels, logits dtype)
    "small_fp32": (16, 1000, 11, 0.25, "fp32"),
    "ragged_vocab_fp32": (7, 1003, 12, 0.3, "fp32"),   # vocab not a multiple of the vector width
    "t5vocab_fp32": (24, 32128, 13, 0.2, "fp32"),
    "t5vocab_bf16": (24, 32128, 14, 0.2, "bf16"),
    "no_ignored_fp32": (8, 32128, 15, 0.0, "fp32"),     # label mode REF: no -100 anywhere
    "all_ignored_fp32": (4, 512, 16, 1.0, "fp32"),      # NaN loss, zero gradient
    "single_row_bf16": (1, 32128, 17, 0.0, "bf16"),
}


LMCE_CASES = {
    # name: (rows, vocab, d_model, seed, fraction of -100 labels); shapes ragged against the
    # kernel's 128 x 256 x 64 tiles on purpose
    "small": (200, 1000, 128, 51, 0.25),
    "one_tile": (130, 300, 64, 52, 0.1),
    "all_ignored": (64, 512, 64, 53, 1.0),
}


def gen_lmce():
    """torch's own ops as the scripts reach them under bf16 autocast: a bf16 F.linear (fp32
    accumulation, bf16 output), then F.cross_entropy on the up-cast logits, then autograd."""
    import torch.nn.functional as F

    out = {}
    for name, (n, v, d, seed, frac) in LMCE_CASES.items():
        hidden, weight, labels = lmce_inputs(n, v, d, seed, frac)
        h = torch.from_numpy(hidden).to(torch.bfloat16).requires_grad_(True)
        w = torch.from_numpy(weight).to(torch.bfloat16).requires_grad_(True)
        logits = F.linear(h, w)
        assert logits.dtype == torch.bfloat16
        z32 = logits.float()
        z32.retain_grad()
        loss = F.cross_entropy(z32, torch.from_numpy(labels), ignore_index=-100)
        out[f"{name}/loss"] = loss.detach().numpy()
        out[f"{name}/row_lse"] = torch.logsumexp(z32.detach(), dim=1).numpy()
        if frac < 1.0:
            loss.backward()
            out[f"{name}/dlogits"] = z32.grad.numpy()
            out[f"{name}/dhidden"] = h.grad.float().numpy()   # bf16 GEMM outputs, as the library gives them
            out[f"{name}/dweight"] = w.grad.float().numpy()
        out[f"{name}/meta"] = np.array([n, v, d, seed, int(frac * 100)], dtype=np.int64)
    np.savez_compressed(GOLD / "lmce.npz", **out)


ADAMW_CASES = {
    # name: (tensor sizes, seed, weight_decay, steps, lr schedule)
    "accelerate_wd0": ([1024, 4096 + 3, 33, 1], 21, 0.0, 4, "linear1"),   # train-accelerator.py:174-205
    "trainer_wd001": ([8192 * 2 + 5, 512, 7], 22, 0.01, 4, "const"),       # train-torchrun.py:121
    "long_run": ([2048], 23, 0.01, 10, "linear1"),
}


def lr_schedule(kind: str, steps: int, base_lr: float = 5e-5):
    from oracle.ref_numpy import linear_schedule_factor

    if kind == "const":
        return [base_lr] * steps
    # LambdaLR: lr used at optimizer step i is base_lr * factor(i); factor(0) == 0
    return [base_lr * linear_schedule_factor(i, 1, 100) for i in range(steps)]


def gen_adamw():
    out = {}
    for name, (sizes, seed, wd, steps, sched) in ADAMW_CASES.items():
        params, grads = adamw_inputs(sizes, seed, steps)
        lrs = lr_schedule(sched, steps)
        p, m, v = torch_ref.adamw_ref([torch.from_numpy(a) for a in params],
                                      [[torch.from_numpy(g) for g in gs] for gs in grads],
                                      lr_per_step=lrs, weight_decay=wd)
        for i in range(len(sizes)):
            out[f"{name}/p{i}"] = p[i].numpy()
            out[f"{name}/m{i}"] = m[i].numpy()
            out[f"{name}/v{i}"] = v[i].numpy()
        out[f"{name}/lrs"] = np.array(lrs, dtype=np.float64)
    np.savez_compressed(GOLD / "adamw.npz", **out)


ALLREDUCE_CASES = {"w2": (2, 4099, 31), "w4": (4, 8192 + 7, 32), "w8": (8, 5000, 33)}


def gen_allreduce():
    out = {}
    for name, (w, n, seed) in ALLREDUCE_CASES.items():
        bufs = allreduce_inputs(w, n, seed)
        out[f"{name}/mean"] = torch_ref.allreduce_mean_ref([torch.from_numpy(b) for b in bufs]).numpy()
    np.savez_compressed(GOLD / "allreduce.npz", **out)


COLLATE_CASES = {"ref_b4": (4, "REF", None), "mask_b4_m8": (4, "MASK", 8), "mask_b3": (3, "MASK", None),
                 "mask_b1_m8": (1, "MASK", 8)}


def gen_collate():
    out = {}
    for name, (b, mode, mult) in COLLATE_CASES.items():
        feats = collate_inputs(b, mode)
        batch = torch_ref.collate_ref(feats, pad_to_multiple_of=mult)
        for k in ("input_ids", "attention_mask", "labels", "decoder_input_ids"):
            out[f"{name}/{k}"] = batch[k].numpy()
    np.savez_compressed(GOLD / "collate.npz", **out)


if __name__ == "__main__":
    GOLD.mkdir(parents=True, exist_ok=True)
    torch.set_num_threads(1)
    gen_ce()
    gen_adamw()
    gen_allreduce()
    gen_collate()
    gen_lmce()
    for f in sorted(GOLD.glob("*.npz")):
        print(f.name, f.stat().st_size, "bytes")
