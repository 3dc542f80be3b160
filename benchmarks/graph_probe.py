#!/usr/bin/env python
"""A/B of the step harness with and without CUDA-graph replay of forward + backward (graphs.py), one GPU:
ms per optimizer step, and - with dropout 0 - the loss trajectories side by side. One JSON line.
    python benchmarks/graph_probe.py [model] [batch] [fused_lm_head 0/1]"""
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402


def run(name, bs, graph, fused, dropout, steps, timed):
    from distributed_llms_example_b200 import patch
    from distributed_llms_example_b200.harness import StepConfig, StepHarness

    h = StepHarness(StepConfig(model=name, semantics="trainer", impl="b200", device="cuda:0", batch_size=bs,
                               dropout=dropout, total_steps=1000, cuda_graph=graph, fused_lm_head=fused))
    batches = [h.to_device(b) for b in h.host_batches(steps + timed)]
    losses = [float(h.step([b])) for b in batches[:steps]]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for b in batches[steps:]:
        h.step([b])
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / max(timed, 1) * 1e3
    gf = h.model.__dict__.get("_b200_graphed")
    info = None if gf is None else {"replays": gf.replays, "graphs": len(gf.graphs), "failed": gf.failed}
    mem = torch.cuda.max_memory_allocated() / 2 ** 30
    h.close()
    patch.remove()
    del h
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
    return losses, ms, info, mem


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "flan-t5-large"
    bs = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    fused = bool(int(sys.argv[3])) if len(sys.argv) > 3 else False
    out = {"model": name, "batch": bs, "fused_lm_head": fused}
    try:
        le, ms_e, _, mem_e = run(name, bs, False, fused, 0.0, 8, 6)
        lg, ms_g, info, mem_g = run(name, bs, True, fused, 0.0, 8, 6)
        out.update(eager_losses=le, graph_losses=lg, eager_ms_dropout0=ms_e, graph_ms_dropout0=ms_g, graph_info=info,
                   max_rel_loss_diff=max(abs(a - b) / abs(a) for a, b in zip(le, lg)), eager_peak_gib=mem_e,
                   graph_peak_gib=mem_g)
        _, ms_e1, _, _ = run(name, bs, False, fused, 0.1, 6, 10)
        lg1, ms_g1, info1, _ = run(name, bs, True, fused, 0.1, 6, 10)
        out.update(eager_ms=ms_e1, graph_ms=ms_g1, graph_info_dropout=info1, graph_losses_dropout=lg1)
    except Exception as exc:
        import traceback

        out["error"] = f"{type(exc).__name__}: {exc}"[:500]
        out["trace"] = traceback.format_exc()[-2500:]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
