#!/usr/bin/env python
"""Feasibility probe: can T5's forward + backward be captured in CUDA graphs (torch.cuda.make_graphed_callables)
with this package's kernels attached, and what does a step cost then? One JSON line."""
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402


class Positional(torch.nn.Module):
    def __init__(self, model):
        super().__init__()
        self.model = model

    def forward(self, input_ids, attention_mask, labels, decoder_input_ids):
        return self.model(input_ids=input_ids, attention_mask=attention_mask, labels=labels,
                          decoder_input_ids=decoder_input_ids).loss


def main():
    from distributed_llms_example_b200 import patch
    from distributed_llms_example_b200.harness import StepConfig, StepHarness

    name = sys.argv[1] if len(sys.argv) > 1 else "flan-t5-large"
    bs = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    out = {"model": name, "batch": bs}
    h = StepHarness(StepConfig(model=name, semantics="trainer", impl="b200", device="cuda:0", batch_size=bs,
                               dropout=0.1, total_steps=1000))
    batches = [h.to_device(b) for b in h.host_batches(12)]
    keys = ("input_ids", "attention_mask", "labels", "decoder_input_ids")

    def timeit(fn, n=5):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / n * 1e3

    i = [0]

    def eager_step():
        b = batches[i[0] % len(batches)]
        i[0] += 1
        return h.step([b])

    out["eager_ms"] = timeit(eager_step)
    try:
        pos = Positional(h.model)
        sample = tuple(batches[0][k] for k in keys)
        with torch.autocast("cuda", dtype=torch.bfloat16, cache_enabled=False):
            graphed = torch.cuda.make_graphed_callables(pos, sample, num_warmup_iters=3)

        def graphed_step():
            b = batches[i[0] % len(batches)]
            i[0] += 1
            with torch.autocast("cuda", dtype=torch.bfloat16, cache_enabled=False):
                loss = graphed(*(b[k] for k in keys))
            loss.backward()
            torch.nn.utils.clip_grad_norm_(h.model.parameters(), 1.0)
            h.optimizer.step()
            h.scheduler.step()
            h.optimizer.zero_grad()
            return loss

        l0 = float(graphed_step())
        out["graphed_ms"] = timeit(graphed_step)
        out["graphed_loss_first"] = l0
        out["graphed_loss_last"] = float(graphed_step())
    except Exception as exc:
        import traceback

        out["graph_error"] = f"{type(exc).__name__}: {exc}"[:600]
        out["graph_trace"] = traceback.format_exc()[-1500:]
    patch.remove()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
