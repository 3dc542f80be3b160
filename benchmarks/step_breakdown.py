#!/usr/bin/env python
"""Host wall-clock per phase of one optimizer step (no device sync inside the loop): where the
host-bound step spends its time, for impl b200 vs stock. flan-t5-large, batch 8, N=1."""
import argparse
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from distributed_llms_example_b200.harness import StepConfig, StepHarness

ap = argparse.ArgumentParser()
ap.add_argument("--impl", default="b200")
ap.add_argument("--model", default="flan-t5-large")
ap.add_argument("--batch-size", type=int, default=8)
ap.add_argument("--semantics", default="trainer")
ap.add_argument("--steps", type=int, default=12)
args = ap.parse_args()

h = StepHarness(StepConfig(model=args.model, semantics=args.semantics, impl=args.impl, device="cuda:0",
                           batch_size=args.batch_size))
batches = [h.to_device(b) for b in h.host_batches(args.steps + 3)]
for b in batches[:3]:
    h.step([b])
torch.cuda.synchronize()
acc = {}


def tick(name, t0):
    t = time.perf_counter()
    acc[name] = acc.get(name, 0.0) + (t - t0)
    return t


t_all = time.perf_counter()
for b in batches[3:]:
    t = time.perf_counter()
    loss = h._forward(b)
    t = tick("forward", t)
    loss.backward()
    t = tick("backward", t)
    if args.semantics == "trainer":
        if args.impl == "b200":
            h.optimizer.clip_grad_norm_(1.0)
        else:
            torch.nn.utils.clip_grad_norm_(h.model.parameters(), 1.0)
        t = tick("clip", t)
    h.optimizer.step()
    t = tick("opt.step", t)
    h.scheduler.step()
    t = tick("sched.step", t)
    h.optimizer.zero_grad()
    t = tick("zero_grad", t)
host = time.perf_counter() - t_all
torch.cuda.synchronize()
total = time.perf_counter() - t_all
n = args.steps
print(args.impl, args.model, "B", args.batch_size, "host ms/step %.2f  wall ms/step %.2f" % (host / n * 1e3, total / n * 1e3),
      {k: round(v / n * 1e3, 3) for k, v in acc.items()})
