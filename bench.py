#!/usr/bin/env python
"""Headline benchmark: T5 summarisation fine-tuning samples/s with the B200 hot path attached.

    python bench.py --gpus 1 --steps K --warmup W                  # this repo's kernels
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W     # N > 1, one rank per GPU
    python bench.py --impl reference ...   # the reference's CPU path (stock torch ops, gloo-era
                                           # semantics) on the box's host cores, rank 0 only
    python bench.py --impl stock ...       # stock GPU path (NCCL DDP + fused AdamW + ATen CE):
                                           # the optimisation target, not the driver's arm

A "step" is one optimizer step of the step harness (distributed_llms_example_b200/harness.py:
forward, masked CE, backward with bucketed gradient sync, clip, AdamW, schedule, zero_grad) on
synthetic 512-token-source / 128-token-target batches with random-init weights. One JSON line
is printed by rank 0:
  value     samples/s over all ranks with the step's batches already resident in HBM
  e2e       the same, through the same public API, with every step's batch copied from pinned
            host memory and the loss read back to the host inside the timed region
  roofline  the dominant kernel of the path at N=1 (the AdamW launch): algorithmic bytes
            (28 B per parameter) / CUDA-event duration on the launching stream, against the
            measured HBM copy peak of MEASURED_PEAKS.json
  cpu_baseline  the reference's CPU path timed on this box's host cores on a bounded sample
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

METRIC = "train_samples_per_second"
UNIT = "samples/s"
FALLBACK_HBM_GBS = 6650.0  # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "stock"])
    ap.add_argument("--model", default="flan-t5-large")
    ap.add_argument("--semantics", default="trainer", choices=["trainer", "accelerate", "task"])
    ap.add_argument("--batch-size", type=int, default=8, help="per device, per micro-batch")
    ap.add_argument("--grad-accum", type=int, default=1)
    ap.add_argument("--label-mode", default="REF", choices=["REF", "MASK"])
    ap.add_argument("--wire", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--bucket-cap-mb", type=float, default=None, help="DDP bucket size (default: torch's 25 MiB)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-steps", type=int, default=2, help="timed CPU steps for cpu_baseline")
    return ap.parse_args()


def hbm_peak():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            return float(json.loads(f.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows: list[list[str]] = []
        self.idx = str(gpu_index)
        self.proc = None

    def __enter__(self):
        if os.environ.get("B2_BENCH_NO_CLOCKS") == "1":  # e.g. under ncu: no child processes
            return self
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", self.idx], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 9:
                self.rows.append(parts)

    def __exit__(self, *exc):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self) -> dict:
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[5 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def timed_region(harness, steps_batches, *, read_loss: bool):
    """Barrier + synchronize on both sides, CUDA events on the launching stream, max over ranks."""
    dev = harness.device
    if dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    last = None
    for mbs in steps_batches:
        loss = harness.step(mbs)
        last = loss.item() if read_loss else loss
    e1.record()
    torch.cuda.synchronize(dev)
    if dist.is_initialized():
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if dist.is_initialized():
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()), (last if read_loss else float(last.item()))


def grad_sync_probe(harness, impl: str, world: int, bucket_mb: float = 25.0, iters: int = 20) -> dict:
    """Second half of BASELINE.json's metric: achieved gradient-sync bus GB/s against NVLink peak,
    measured live on a DDP-sized fp32 bucket (25 MiB), device-timed, max over ranks. For impl=b200
    the bucket is registered for peer access like the Reducer's buckets are from step 3 on."""
    dev = harness.device
    n = int(bucket_mb * (1 << 20)) // 4
    x = torch.randn(n, device=dev) * 1e-2
    path = "nccl div_ + all_reduce"
    if impl == "b200":
        comm = harness.hook_state.comm if harness.hook_state is not None else harness.comm
        off = None
        if comm.nvls and comm._pool is not None:
            # the Reducer's buckets live in the multicast-bound region: probe the same path
            with torch.cuda.use_mem_pool(comm._pool):
                xp = torch.empty(n, device=dev)
            off = comm.nvls_offset(xp)
            offs = [None] * world
            dist.all_gather_object(offs, off)
            if off is None or len(set(offs)) != 1:
                off = None
        if off is not None:
            xp.copy_(x)
            fn = lambda: comm.allreduce_nvls_inplace_(off, n, 1.0 / world)  # noqa: E731
            path = "zero-copy NVLS: in-switch reduction in place on buckets in multicast-bound memory"
        else:
            reg = comm.register(x)
            fn = lambda: comm.allreduce_registered_(reg, x, 1.0 / world)  # noqa: E731
            path = "zero-copy one-pass push over peer-mapped (IPC) buckets"
    else:
        def fn():
            x.div_(world)
            dist.all_reduce(x)
    for _ in range(5):
        fn()
    torch.cuda.synchronize(dev)
    dist.barrier()
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    b.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    bus = n * 4 * 2 * (world - 1) / world / ms / 1e6
    return {"bucket_mb": bucket_mb, "us_per_bucket": ms * 1e3, "bus_gbs": bus, "nvlink_peak_gbs": 900.0,
            "frac_of_nvlink_peak": bus / 900.0, "frac_of_measured_peer_copy_770": bus / 770.0,
            "impl": path}


def run_cpu(args, steps: int, warmup: int):
    """The reference's CPU path: the same harness with stock torch ops on the host cores, fp32,
    one sample per step (a bounded sample of the workload). Returns (samples/s, threads, desc)."""
    from distributed_llms_example_b200.harness import StepConfig, StepHarness

    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    cfg = StepConfig(model=args.model, semantics=args.semantics, impl="stock", device="cpu", batch_size=1,
                     grad_accum=1, label_mode=args.label_mode, bf16=False)
    h = StepHarness(cfg)
    batches = h.host_batches(steps + warmup)
    for b in batches[:warmup]:
        h.step([b])
    t0 = time.perf_counter()
    for b in batches[warmup:]:
        h.step([b]).item()
    dt = time.perf_counter() - t0
    sample = (f"{steps} optimizer steps of {args.model} ({args.semantics} semantics), batch 1 x 512/128 tokens, "
              f"fp32, stock torch CPU ops (CrossEntropyLoss + torch.optim.AdamW), {warmup} warm-up")
    return steps * 1.0 / dt, threads, sample, dt / steps * 1e3


def main():
    args = parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    config = {"workload": f"{args.model} summarization fine-tune step ({args.semantics} semantics), synthetic "
                          f"512-token sources / 128-token targets, label mode {args.label_mode}, random init",
              "per_device_batch": args.batch_size, "grad_accum": args.grad_accum,
              "global_batch": args.batch_size * args.grad_accum * max(world, 1),
              "src_len": 512, "tgt_len": 128, "parallelism": f"dp{world}",
              "l2": "inputs exceed L2 (weights + optimizer state >> 126 MB are re-read every step)"}

    if args.impl == "reference":
        if rank != 0:
            return  # rank 0 alone runs the CPU arm
        sps, threads, sample, ms = run_cpu(args, args.steps, args.warmup)
        config.update(per_device_batch=1, grad_accum=1, global_batch=1, parallelism="cpu")
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": sps, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": sps, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": sps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}), flush=True)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device for --impl b200/stock (there is no CPU fallback); "
                         "use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from distributed_llms_example_b200 import _lib
    from distributed_llms_example_b200.harness import StepConfig, StepHarness

    cfg = StepConfig(model=args.model, semantics=args.semantics, impl=args.impl, device=f"cuda:{local_rank}",
                     batch_size=args.batch_size, grad_accum=args.grad_accum, label_mode=args.label_mode,
                     bf16=True, wire=args.wire, bucket_cap_mb=args.bucket_cap_mb,
                     total_steps=max(1000, 4 * (args.steps + args.warmup)))
    h = StepHarness(cfg)
    n_params = sum(p.numel() for p in h.model.parameters())
    ga, K, W = args.grad_accum, args.steps, max(args.warmup, 0)
    host = h.host_batches((2 * K + W) * ga)
    groups = [host[i * ga:(i + 1) * ga] for i in range(2 * K + W)]
    warm, resident_src, e2e_src = groups[:W], groups[W:W + K], groups[W + K:]

    for mbs in warm:
        h.step(mbs)
    resident = [[h.to_device(mb) for mb in mbs] for mbs in resident_src]
    torch.cuda.synchronize()

    if args.impl == "b200":
        h.optimizer.kernel_events = []
    launches0 = _lib.launches
    cpp_fast0 = h.hook_state.cpp_handle.fast_calls if (h.hook_state is not None and h.hook_state.cpp_handle is not None) else 0
    with ClockSampler(local_rank) as clocks:
        # cudaProfilerStart/Stop bracket the timed region so that `ncu --profile-from-start off`
        # lists exactly these K steps' launches (no-ops when no profiler is attached)
        torch.cuda.profiler.start()
        ms_res, _ = timed_region(h, resident, read_loss=False)
        torch.cuda.profiler.stop()
        launches = _lib.launches - launches0
        if h.hook_state is not None and h.hook_state.cpp_handle is not None:
            launches += 2 * (h.hook_state.cpp_handle.fast_calls - cpp_fast0)  # gate + data kernel each
        opt_events = h.optimizer.kernel_events if args.impl == "b200" else None
        if args.impl == "b200":
            h.optimizer.kernel_events = None
        ms_e2e, last_loss = timed_region(h, e2e_src, read_loss=True)
    samples = args.batch_size * ga * world * K
    value = samples / (ms_res / 1e3)
    e2e = samples / (ms_e2e / 1e3)
    h2d = sum(h.batch_bytes(mb) for mb in e2e_src[0])

    roofline = None
    if opt_events:
        torch.cuda.synchronize()
        dur_ms = statistics.mean(a.elapsed_time(b) for a, b in opt_events)
        peak, how = hbm_peak()
        achieved = 28.0 * n_params / (dur_ms / 1e3) / 1e9
        traffic = None
        tf = ROOT / "profiles" / "adamw_traffic.json"
        if tf.exists():
            try:
                traffic = json.loads(tf.read_text()).get(args.model)
            except Exception:
                traffic = None
        roofline = {"kernel": "adamw_multi_kernel", "bound": "hbm", "achieved": achieved, "peak": peak,
                    "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": how,
                    "algorithmic_bytes_per_launch": 28 * n_params, "launch_ms": dur_ms,
                    "frac_of_8TBs_spec": achieved / 8000.0}

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
           "ms_per_step": ms_res / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "bf16", "data": "synthetic", "config": config, "impl": args.impl,
           "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                   "ms_per_step": ms_e2e / K},
           "gpu_launches": launches if args.impl == "b200" else 0,
           "clocks": clocks.summary(), "roofline": roofline, "final_loss": last_loss,
           "params": n_params, "lib": str(_lib.lib_path().name) if args.impl == "b200" else None}
    if world > 1:
        # optional extras must never cost the headline numbers above
        try:
            out["grad_sync"] = grad_sync_probe(h, args.impl, world)
            if h.hook_state is not None:  # which path the Reducer's buckets actually took
                regs = h.hook_state._registered.values()
                out["grad_sync"]["ddp_buckets"] = {
                    "reduced_per_step": h.hook_state.total_buckets_reduced // max(1, h.steps_done),
                    "cpp_hook": h.hook_state.cpp_handle is not None,
                    "nvls_in_place": sum(1 for r in regs if r[3] is not None),
                    "ipc_push_in_place": sum(1 for r in regs if r[3] is None),
                }
        except Exception as exc:  # collective: every rank takes the same branch or the run hangs
            out["grad_sync"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
    if world > 1:
        # the line goes out before any teardown (unmapping peers, destroying the process group)
        if rank == 0:
            print(json.dumps(out), flush=True)
        try:
            h.close()
        finally:
            dist.destroy_process_group()
        return
    h.close()
    if not args.no_cpu_baseline and args.impl == "b200":
        try:
            sps, threads, sample, _ = run_cpu(args, args.cpu_steps, 1)
            out["cpu_baseline"] = {"value": sps, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample}
        except Exception as exc:
            out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                   "sample": f"failed: {type(exc).__name__}: {exc}"[:300]}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
