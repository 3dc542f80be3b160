"""Golden losses of the reference's CPU path for the step-level parity test.

Runs the step harness with stock torch ops on CPU (fp32, dropout 0, t5-small, batch 2,
512/128-token synthetic pairs) and stores the per-step losses in tests/golden/step_t5small.npz.
Run in the build container:  python oracle/make_golden_step.py"""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_llms_example_b200.harness import StepConfig, StepHarness  # noqa: E402

CASES = {"accelerate_REF": ("accelerate", "REF", 1), "trainer_MASK_ga2": ("trainer", "MASK", 2)}
STEPS = 4

if __name__ == "__main__":
    torch.set_num_threads(8)
    out = {}
    for name, (sem, mode, ga) in CASES.items():
        h = StepHarness(StepConfig(model="t5-small", semantics=sem, impl="stock", device="cpu", batch_size=2,
                                   grad_accum=ga, label_mode=mode, bf16=False, dropout=0.0, total_steps=20))
        batches = h.host_batches(STEPS * ga)
        losses = [float(h.step(batches[i * ga:(i + 1) * ga])) for i in range(STEPS)]
        out[name] = np.array(losses, dtype=np.float64)
        print(name, losses)
    np.savez(ROOT / "tests" / "golden" / "step_t5small.npz", **out)
