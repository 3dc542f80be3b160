#!/usr/bin/env python
"""Host-side (CPU) cost per call of the plug-ins vs the stock torch calls they replace.
The B=8 training step is host-bound, so microseconds of Python on the path show up in samples/s.
Wall-clock around asynchronous calls (no device sync inside the timed loop), flan-t5-large shapes."""
import cProfile
import pstats
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from distributed_llms_example_b200.loss import cross_entropy
from distributed_llms_example_b200.optim import AdamW
from distributed_llms_example_b200.synthetic import build_model


def wall_us(fn, iters=30, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        torch.cuda.synchronize()
        t = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t) * 1e6)
    torch.cuda.synchronize()
    return float(np.median(ts))


with torch.device("meta"):
    shapes = [tuple(p.shape) for p in build_model("flan-t5-large").parameters()]
ps = [torch.nn.Parameter(torch.zeros(s, device="cuda")) for s in shapes]


def fresh_grads():
    for p in ps:
        p.grad = torch.zeros_like(p)


fresh_grads()
ours = AdamW(ps, lr=5e-5, weight_decay=0.01)
fused = torch.optim.AdamW(ps, lr=5e-5, weight_decay=0.01, fused=True)
print("ours  clip+step  us", wall_us(lambda: (ours.clip_grad_norm_(1.0), ours.step())))
print("ours  step only  us", wall_us(ours.step))
print("torch clip+step  us", wall_us(lambda: (torch.nn.utils.clip_grad_norm_(ps, 1.0), fused.step())))
print("torch step only  us", wall_us(fused.step))


def regrad_ours():
    fresh_grads()  # new gradient tensors every step, like zero_grad(set_to_none=True) + backward
    ours.clip_grad_norm_(1.0)
    ours.step()


def regrad_only():
    fresh_grads()


print("fresh grads only us", wall_us(regrad_only))
print("ours fresh grads + clip+step us", wall_us(regrad_ours))

z = (torch.randn(1024, 32128, device="cuda")).to(torch.bfloat16).requires_grad_(True)
y = torch.randint(0, 32128, (1024,), device="cuda")


def ce_ours():
    z.grad = None
    cross_entropy(z, y).backward()


def ce_torch():
    z.grad = None
    with torch.autocast("cuda", dtype=torch.bfloat16):
        torch.nn.functional.cross_entropy(z, y).backward()


print("ours  CE fwd+bwd us", wall_us(ce_ours))
print("torch CE fwd+bwd us", wall_us(ce_torch))

pr = cProfile.Profile()
pr.enable()
for _ in range(20):
    ours.clip_grad_norm_(1.0)
    ours.step()
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
