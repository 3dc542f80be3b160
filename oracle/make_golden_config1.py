"""Golden losses for BASELINE.json configs[0]: t5-small, synthetic 512/128-token pairs, world size 2 over
gloo on CPU, Trainer semantics (GA 2, clip 1.0, AdamW wd 0.01, linear schedule), 3 optimizer steps.

Runs the step harness with STOCK torch ops (DistributedDataParallel over gloo, CrossEntropyLoss,
torch.optim.AdamW) in the build container and stores the per-rank losses in
tests/golden/config1_t5small.npz:  python oracle/make_golden_config1.py"""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

STEPS, GA, WORLD = 3, 2, 2


def worker(rank, world, port, q, threads=2):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(threads)
        from distributed_llms_example_b200.harness import StepConfig, StepHarness

        h = StepHarness(StepConfig(model="t5-small", semantics="trainer", impl="stock", device="cpu", batch_size=1,
                                   grad_accum=GA, label_mode="MASK", bf16=False, dropout=0.0, total_steps=10))
        batches = h.host_batches(STEPS * GA)
        losses = [float(h.step(batches[i * GA:(i + 1) * GA])) for i in range(STEPS)]
        flat = torch.cat([p.detach().reshape(-1) for p in h.model.parameters()])
        parts = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(parts, flat)
        q.put((rank, losses, all(torch.equal(parts[0], p) for p in parts), float(flat.double().abs().sum()), None))
    except Exception:
        import traceback

        q.put((rank, [], False, 0.0, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def run(threads=2):
    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=worker, args=(r, WORLD, port, q, threads)) for r in range(WORLD)]
    for p in procs:
        p.start()
    outs = sorted(q.get(timeout=1800) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    return outs


if __name__ == "__main__":
    outs = run()
    for rank, losses, same, checksum, err in outs:
        assert err is None, err
        print(rank, losses, same, checksum)
    np.savez(ROOT / "tests" / "golden" / "config1_t5small.npz",
             **{f"rank{r}": np.array(l, dtype=np.float64) for r, l, *_ in outs},
             param_abs_sum=np.array([o[3] for o in outs]))
