#!/usr/bin/env python
"""Headline benchmark: T5 summarisation fine-tuning samples/s with the B200 hot path attached.

    python bench.py --gpus 1 --steps K --warmup W                  # this repo's kernels
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W     # N > 1, one rank per GPU
    python bench.py --impl reference ...   # the reference's CPU path (stock torch ops, gloo-era
                                           # semantics) on the box's host cores, rank 0 only
    python bench.py --impl stock ...       # stock GPU path (NCCL DDP + fused AdamW + ATen CE):
                                           # the optimisation target, not the driver's arm

The b200 arm runs with everything this package offers switched on: the three plug-ins, the fused
lm_head + CE kernel (--fused-lm-head 1) and CUDA-graph replay of the model's forward + backward
(--cuda-graph 1); `--cuda-graph 0 --fused-lm-head 0` is the plug-ins-only configuration of round 1.

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
  stock_gpu     the optimisation target of BASELINE.json's north_star, timed in the SAME process
                group right after the two regions above: the same harness with stock
                NCCL-DDP + torch.optim.AdamW(fused=True) + ATen cross-entropy; vs_baseline is
                value / stock_gpu.value (the reference publishes no number of its own)
  parity        N > 1 only, outside every timed region: each gradient-sync data path the run can
                take (staging one-/two-shot, bf16 wire, IPC push, NVLS staged / in place,
                average_gradients) all-reduces seeded 25 MiB and odd-sized buffers and is compared
                on the device with the rank-ordered fp32 sum built from a dist.all_gather, plus a
                4-step, dropout-0, same-seed loss comparison of this arm against the stock arm.
                Any failure makes the exit status non-zero (the JSON line is still printed).
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
    ap.add_argument("--cpu-steps", type=int, default=5, help="timed CPU steps for cpu_baseline")
    ap.add_argument("--cpu-omp1", action="store_true",
                    help="also time the CPU arm with 1 thread per rank (torchrun's default; minutes for flan-t5-large)")
    ap.add_argument("--cpu-world-max", type=int, default=2, help="--impl reference: at most this many gloo ranks")
    ap.add_argument("--no-stock", action="store_true", help="skip the stock-GPU leg (vs_baseline stays null)")
    ap.add_argument("--no-parity", action="store_true", help="skip the N>1 parity block and the loss comparison")
    ap.add_argument("--dropout", type=float, default=0.1)
    ap.add_argument("--cuda-graph", type=int, default=1, choices=[0, 1],
                    help="b200 arm: replay forward + backward from CUDA graphs (graphs.py); 0 = eager launches")
    ap.add_argument("--fused-lm-head", type=int, default=1, choices=[0, 1],
                    help="b200 arm: lm_head + CE as one tcgen05 kernel (fused.py); 0 = cuBLAS lm_head + CE kernels")
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


def _cpu_rank(rank: int, world: int, port: int, threads: int, model: str, semantics: str, label_mode: str,
              steps: int, warmup: int, q) -> None:
    """One rank of the reference's CPU path: gloo process group, DDP over gloo when world > 1
    (what `torchrun --nproc-per-node=W train-torchrun.py` gives on a machine without GPUs)."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS=str(threads))
    torch.set_num_threads(threads)
    if world > 1:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from distributed_llms_example_b200.harness import StepConfig, StepHarness

        cfg = StepConfig(model=model, semantics=semantics, impl="stock", device="cpu", batch_size=1,
                         grad_accum=1, label_mode=label_mode, bf16=False)
        h = StepHarness(cfg)
        batches = h.host_batches(steps + warmup)
        for b in batches[:warmup]:
            h.step([b])
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for b in batches[warmup:]:
            h.step([b]).item()
        if world > 1:
            dist.barrier()
        q.put((rank, time.perf_counter() - t0, torch.get_num_threads(), None))
    except Exception as exc:  # pragma: no cover
        q.put((rank, 0.0, 0, f"{type(exc).__name__}: {exc}"))
    finally:
        if world > 1:
            dist.destroy_process_group()


def run_cpu(args, steps: int, warmup: int, world: int = 1, threads: int | None = None):
    """The reference's CPU path: the same harness with stock torch ops on the host
    cores, fp32, one sample per rank per step (a bounded sample of the workload), `world` gloo
    ranks with `threads` OpenMP threads each (default: host cores // world). Returns
    (samples/s, threads per rank, description, ms per step)."""
    import socket

    import torch.multiprocessing as mp

    cores = os.cpu_count() or 1
    threads = threads or max(1, cores // world)
    sample = (f"{steps} optimizer steps of {args.model} ({args.semantics} semantics) after {warmup} warm-up, "
              f"batch 1 x 512/128 tokens per rank, fp32, stock torch CPU ops (CrossEntropyLoss + torch.optim.AdamW"
              f"{', DDP over gloo' if world > 1 else ''}), world {world} x {threads} threads on {cores} host cores")
    if world == 1:
        class _Q:
            def put(self, item):
                self.item = item

        q = _Q()
        _cpu_rank(0, 1, 0, threads, args.model, args.semantics, args.label_mode, steps, warmup, q)
        outs = [q.item]
    else:
        s_ = socket.socket()
        s_.bind(("127.0.0.1", 0))
        port = s_.getsockname()[1]
        s_.close()
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        procs = [ctx.Process(target=_cpu_rank, args=(r, world, port, threads, args.model, args.semantics,
                                                     args.label_mode, steps, warmup, q)) for r in range(world)]
        for p_ in procs:
            p_.start()
        outs = [q.get(timeout=3600) for _ in procs]
        for p_ in procs:
            p_.join(timeout=60)
    errs = [o[3] for o in outs if o[3]]
    if errs:
        raise RuntimeError(errs[0])
    dt = max(o[1] for o in outs)
    return steps * world * 1.0 / dt, threads, sample, dt / steps * 1e3


def _rank_ordered_mean(parts: list, scale: float, bf16_wire: bool = False) -> torch.Tensor:
    """The oracle's all-reduce (oracle/ref_numpy.py allreduce_mean) restated on the device from the
    all-gathered inputs: sum_r scale*b_r, separate fp32 multiply and add, ranks 0..W-1 in order."""
    def rnd(t):
        return t.to(torch.bfloat16).float() if bf16_wire else t

    acc = rnd(parts[0] * scale)
    for t in parts[1:]:
        acc = acc + rnd(t * scale)
    return rnd(acc)


def grad_sync_parity(comm, world: int, rank: int, dev) -> dict:
    """Every gradient-sync data path against the rank-ordered fp32 sum, on seeded buffers of a DDP
    bucket's size (25 MiB) and of an odd size (numel % 4 != 0). Bit-exact for the staging and IPC
    paths, <= 1e-6 of sum_r |scale*b_r| for the in-switch (NVLS) sums whose order is the
    hardware's; bit-identical across ranks for all; sum-of-squares fold within 2e-6."""
    from distributed_llms_example_b200.comm import average_gradients

    sizes = {"25MiB": (25 << 20) // 4, "odd": 1_000_003}
    scale = 1.0 / world
    out: dict = {}
    ok_all = True
    nvls = bool(comm.nvls) and comm._pool is not None
    comm.enable_sumsq(True)

    def check(name, x, parts, exact, bf16_wire=False, sumsq=True):
        nonlocal ok_all
        ref = _rank_ordered_mean(parts, scale, bf16_wire)
        mag = sum((t * scale).abs() for t in parts)
        err = (x - ref).abs()
        rel = float((err / mag.clamp_min(1e-30)).max())
        bit = bool(torch.equal(x, ref))
        outs = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(outs, x)
        same = all(torch.equal(outs[0], t) for t in outs)
        entry = {"max_rel": rel, "bit_exact": bit, "replicas_identical": same}
        ok = same and (bit if exact else rel <= 1e-6)
        if sumsq:
            got = torch.empty(2, device=dev)
            counted = comm.sumsq_numel()
            comm.sumsq_clip_(1.0, got)
            want = float(torch.linalg.vector_norm(x.double()))
            entry["norm_rel_err"] = abs(float(got[0]) - want) / max(want, 1e-30)
            ok = ok and counted == x.numel() and entry["norm_rel_err"] <= 2e-6
        entry["ok"] = bool(ok)
        out[name] = entry
        ok_all = ok_all and ok

    for sname, n in sizes.items():
        g = torch.Generator(device=dev).manual_seed(4242 + 17 * rank + n % 1000)
        src = torch.randn(n, device=dev, generator=g) * 1e-2
        parts = [torch.empty_like(src) for _ in range(world)]
        dist.all_gather(parts, src)
        for algo in ("one_shot", "two_shot"):
            x = src.clone()
            comm.allreduce_(x, scale, algo=algo)
            check(f"staging_{algo}/{sname}", x, parts, exact=True)
        x = src.clone()
        comm.allreduce_(x, scale, wire="bf16")
        check(f"staging_bf16_wire/{sname}", x, parts, exact=True, bf16_wire=True)
        x = src.clone()
        reg = comm.register(x)
        comm.allreduce_registered_(reg, x, scale)
        check(f"ipc_push/{sname}", x, parts, exact=True)
        comm.unregister(reg)
        if comm.nvls:
            x = src.clone()
            comm.allreduce_nvls_(x, scale)
            check(f"nvls_staged/{sname}", x, parts, exact=False)
        if nvls:
            with torch.cuda.use_mem_pool(comm._pool):
                xp = torch.empty(n, device=dev)
            off = comm.nvls_offset(xp)
            offs = [None] * world
            dist.all_gather_object(offs, off)
            if off is not None and len(set(offs)) == 1:
                xp.copy_(src)
                comm.allreduce_nvls_inplace_(off, n, scale)
                check(f"nvls_inplace/{sname}", xp, parts, exact=False)
            else:
                out[f"nvls_inplace/{sname}"] = {"ok": False, "error": f"offsets differ across ranks: {offs}"}
                ok_all = False
            del xp
    comm.enable_sumsq(False)
    # train-task.py:65-69: all_reduce(SUM) then "/= size", per parameter -> one flat exchange here
    torch.manual_seed(0)
    mlp = torch.nn.Sequential(torch.nn.Linear(1000, 2049), torch.nn.Linear(2049, 333, bias=False)).to(dev)
    g = torch.Generator(device=dev).manual_seed(9000 + rank)
    worst, bit, same = 0.0, True, True
    for p_ in mlp.parameters():
        p_.grad = torch.randn(p_.shape, device=dev, generator=g) * 1e-2
    want = []
    for p_ in mlp.parameters():
        parts = [torch.empty_like(p_.grad) for _ in range(world)]
        dist.all_gather(parts, p_.grad)
        want.append(_rank_ordered_mean(parts, scale))
    average_gradients(mlp, comm)
    for p_, w_ in zip(mlp.parameters(), want):
        bit = bit and bool(torch.equal(p_.grad, w_))
        worst = max(worst, float((p_.grad - w_).abs().max()))
        outs = [torch.empty_like(p_.grad) for _ in range(world)]
        dist.all_gather(outs, p_.grad)
        same = same and all(torch.equal(outs[0], t) for t in outs)
    out["average_gradients"] = {"max_abs": worst, "bit_exact": bit, "replicas_identical": same, "ok": bit and same}
    ok_all = ok_all and bit and same
    torch.cuda.synchronize(dev)
    comm.poll()
    out["ok"] = bool(ok_all)
    return out


def loss_parity(args, local_rank: int, steps: int = 4) -> dict:
    """This arm against the stock arm on the headline config with dropout 0, the same seeds and the
    same batches: `steps` optimizer steps each, loss per step compared. Two bars:
      * step 0 (identical parameters, identical bf16 logits in both arms) against an fp32
        cross-entropy of the stock model's own logits: <= 2e-5 relative. This isolates the loss
        kernel from torch's bf16-autocast cross-entropy, which carries ~1e-3 of its own error
        (measured here as `stock_ce_own_error`);
      * every step against the stock arm's reported loss: <= 1e-3 relative, the size of that error
        (the arms also differ in clip, AdamW and the all-reduce's summation order, all <= 1e-6)."""
    import torch.nn.functional as F

    from distributed_llms_example_b200.harness import StepConfig, StepHarness

    losses = {}
    ce32 = None
    for impl in ("b200", "stock"):
        cfg = StepConfig(model=args.model, semantics=args.semantics, impl=impl, device=f"cuda:{local_rank}",
                         batch_size=args.batch_size, grad_accum=args.grad_accum, label_mode=args.label_mode,
                         bf16=True, dropout=0.0, wire="fp32", bucket_cap_mb=args.bucket_cap_mb, total_steps=1000,
                         fused_lm_head=None if args.fused_lm_head is None else bool(args.fused_lm_head),
                         cuda_graph=None if args.cuda_graph is None else bool(args.cuda_graph))
        h = StepHarness(cfg)
        host = h.host_batches(steps * args.grad_accum, seed=4321)
        if impl == "stock":
            with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
                vals = []
                for mb in host[:args.grad_accum]:
                    out = h.model(**h.to_device(mb))
                    lab = mb["labels"].to(h.device).view(-1)
                    vals.append(float(F.cross_entropy(out.logits.float().view(-1, out.logits.size(-1)), lab,
                                                      ignore_index=-100)))
                    del out
                ce32 = sum(vals) / len(vals)
        ls = []
        for i in range(steps):
            ls.append(float(h.step(host[i * args.grad_accum:(i + 1) * args.grad_accum]).item()))
        losses[impl] = ls
        h.close()
        del h
        torch.cuda.empty_cache()
    rel = [abs(a - b) / max(abs(b), 1e-30) for a, b in zip(losses["b200"], losses["stock"])]
    step0 = abs(losses["b200"][0] - ce32) / abs(ce32)
    return {"steps": steps, "b200": losses["b200"], "stock": losses["stock"], "max_rel": max(rel),
            "tolerance": 1e-3, "step0_fp32_ce_of_stock_logits": ce32, "step0_rel_vs_fp32_ce": step0,
            "step0_tolerance": 2e-5, "stock_ce_own_error": abs(losses["stock"][0] - ce32) / abs(ce32),
            "ok": bool(max(rel) <= 1e-3 and step0 <= 2e-5), "dropout": 0.0}


def adamw_traffic(model: str):
    """ncu DRAM bytes per AdamW launch, valid only for the kernel source it was captured from."""
    import hashlib

    tf = ROOT / "profiles" / "adamw_traffic.json"
    if not tf.exists():
        return None, "no capture"
    try:
        d = json.loads(tf.read_text())
        src = ROOT / "distributed_llms_example_b200" / "csrc" / "adamw.cu"
        sha = hashlib.sha256(src.read_bytes()).hexdigest()
        if d.get("kernel_source_sha256") != sha:
            return None, "stale: adamw.cu changed since the ncu capture"
        return d.get("traffic", {}).get(model), f"ncu capture {d.get('capture', '')} of adamw.cu sha256 {sha[:12]}"
    except Exception as exc:
        return None, f"unreadable: {exc}"


def run_gpu_arm(args, impl: str, world: int, local_rank: int, K: int, W: int, with_clocks: bool):
    """W warm-up steps, K timed steps with resident batches, K timed steps end to end. Returns a dict
    of measurements plus the live harness (the caller closes it)."""
    from distributed_llms_example_b200 import _lib
    from distributed_llms_example_b200.harness import StepConfig, StepHarness

    cfg = StepConfig(model=args.model, semantics=args.semantics, impl=impl, device=f"cuda:{local_rank}",
                     batch_size=args.batch_size, grad_accum=args.grad_accum, label_mode=args.label_mode,
                     bf16=True, wire=args.wire, bucket_cap_mb=args.bucket_cap_mb, dropout=args.dropout,
                     total_steps=max(1000, 4 * (K + W)),
                     fused_lm_head=None if args.fused_lm_head is None else bool(args.fused_lm_head),
                     cuda_graph=None if args.cuda_graph is None else bool(args.cuda_graph))
    h = StepHarness(cfg)
    ga = args.grad_accum
    host = h.host_batches((2 * K + W) * ga)
    groups = [host[i * ga:(i + 1) * ga] for i in range(2 * K + W)]
    warm, resident_src, e2e_src = groups[:W], groups[W:W + K], groups[W + K:]
    for mbs in warm:
        h.step(mbs)
    # CUDA-graph replay captures on the third call with a given batch shape; with fewer warm-up
    # micro-batches than that, a few more untimed steps keep the one-off capture out of the timed region
    gf = h.model.__dict__.get("_b200_graphed")
    extra = 0
    while gf is not None and not gf.graphs and not gf.failed and extra < 4:
        h.step(h.host_batches(ga, seed=977 + extra))
        extra += 1
    resident = [[h.to_device(mb) for mb in mbs] for mbs in resident_src]
    torch.cuda.synchronize()
    if impl == "b200":
        h.optimizer.kernel_events = []
    launches0 = _lib.launches
    cpp = h.hook_state.cpp_handle if (h.hook_state is not None) else None
    cpp_fast0 = cpp.fast_calls if cpp is not None else 0

    class _NoClocks:
        def __enter__(self):
            return self

        def __exit__(self, *exc):
            return False

        def summary(self):
            return None

    with (ClockSampler(local_rank) if with_clocks else _NoClocks()) as clocks:
        # cudaProfilerStart/Stop bracket the timed region so that `ncu --profile-from-start off`
        # lists exactly these K steps' launches (no-ops when no profiler is attached)
        if impl == "b200":
            torch.cuda.profiler.start()
        ms_res, _ = timed_region(h, resident, read_loss=False)
        if impl == "b200":
            torch.cuda.profiler.stop()
        launches = _lib.launches - launches0
        if cpp is not None:
            launches += 2 * (cpp.fast_calls - cpp_fast0)  # gate + data kernel each
        opt_events = h.optimizer.kernel_events if impl == "b200" else None
        if impl == "b200":
            h.optimizer.kernel_events = None
        ms_e2e, last_loss = timed_region(h, e2e_src, read_loss=True)
    samples = args.batch_size * ga * world * K
    return {"h": h, "value": samples / (ms_res / 1e3), "e2e": samples / (ms_e2e / 1e3), "ms_res": ms_res,
            "ms_e2e": ms_e2e, "launches": launches, "opt_events": opt_events, "clocks": clocks.summary(),
            "final_loss": last_loss, "h2d": sum(h.batch_bytes(mb) for mb in e2e_src[0])}


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
        # the reference's gloo CPU path at the world size the driver asked for (valohai.yaml:7 runs
        # torchrun --nproc-per-node=4): one gloo rank per requested GPU, host cores // world threads each
        # (capped at --cpu-world-max ranks, default 2: every rank holds a full fp32 replica and each
        # step costs world samples of CPU work, and this arm has to end within minutes)
        cpu_world = max(1, min(args.gpus, args.cpu_world_max))
        sps, threads, sample, ms = run_cpu(args, args.steps, args.warmup, world=cpu_world)
        config.update(per_device_batch=1, grad_accum=1, global_batch=cpu_world, parallelism=f"cpu gloo dp{cpu_world}")
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": sps, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": sps, "unit": UNIT, "cores": threads * cpu_world, "kind": "port",
                             "sample": sample, "world": cpu_world, "threads_per_rank": threads,
                             "host_cores": os.cpu_count()},
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

    K, W = args.steps, max(args.warmup, 0)
    r = run_gpu_arm(args, args.impl, world, local_rank, K, W, with_clocks=True)
    h = r["h"]
    n_params = sum(p.numel() for p in h.model.parameters())

    roofline = None
    if r["opt_events"]:
        torch.cuda.synchronize()
        dur_ms = statistics.mean(a.elapsed_time(b) for a, b in r["opt_events"])
        peak, how = hbm_peak()
        achieved = 28.0 * n_params / (dur_ms / 1e3) / 1e9
        traffic, traffic_note = adamw_traffic(args.model)
        roofline = {"kernel": "adamw_multi_kernel", "bound": "hbm", "achieved": achieved, "peak": peak,
                    "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_note,
                    "peak_source": how, "algorithmic_bytes_per_launch": 28 * n_params, "launch_ms": dur_ms,
                    "frac_of_8TBs_spec": achieved / 8000.0}

    out = {"metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
           "ms_per_step": r["ms_res"] / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "bf16", "data": "synthetic", "config": config, "impl": args.impl,
           "e2e": {"value": r["e2e"], "unit": UNIT, "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": 4,
                   "ms_per_step": r["ms_e2e"] / K},
           "gpu_launches": r["launches"] if args.impl == "b200" else 0,
           "clocks": r["clocks"], "roofline": roofline, "final_loss": r["final_loss"],
           "params": n_params, "lib": str(_lib.lib_path().name) if args.impl == "b200" else None}
    if args.impl == "b200" and args.semantics == "trainer":
        out["clip_norm_source"] = h.optimizer.norm_source
    if args.impl == "b200":
        from distributed_llms_example_b200 import patch as _patch

        out["fused_lm_head_ce"] = "t5_forward" in _patch._saved
        gf = h.model.__dict__.get("_b200_graphed")
        out["cuda_graph"] = None if gf is None else {"replays": gf.replays, "graphs": len(gf.graphs), "failed": gf.failed}
    failed = False
    if world > 1:
        # optional extras must never cost the headline numbers above
        try:
            # the size the Reducer's buckets really have in this run (the median registered bucket),
            # and torch's 25 MiB default for reference
            mb = 25.0
            if h.hook_state is not None and h.hook_state._registered:
                mb = statistics.median(r_[1] for r_ in h.hook_state._registered.values()) * 4 / (1 << 20)
            elif args.bucket_cap_mb:
                mb = args.bucket_cap_mb
            out["grad_sync"] = grad_sync_probe(h, args.impl, world, bucket_mb=mb)
            if abs(mb - 25.0) > 1.0:
                out["grad_sync"]["at_25_mib"] = grad_sync_probe(h, args.impl, world, bucket_mb=25.0)
            if h.hook_state is not None:  # which path the Reducer's buckets actually took
                regs = h.hook_state._registered.values()
                out["grad_sync"]["ddp_buckets"] = {
                    "reduced_per_step": h.hook_state.total_buckets_reduced // max(1, h.steps_done),
                    "cpp_hook": h.hook_state.cpp_handle is not None,
                    "nvls_in_place": sum(1 for r_ in regs if r_[3] is not None),
                    "ipc_push_in_place": sum(1 for r_ in regs if r_[3] is None),
                    "gradient_as_bucket_view": bool(h.ddp.gradient_as_bucket_view),
                }
        except Exception as exc:  # collective: every rank takes the same branch or the run hangs
            out["grad_sync"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
        if args.impl == "b200" and not args.no_parity:
            try:
                comm = h.hook_state.comm if h.hook_state is not None else h.comm
                out["grad_sync"]["parity"] = grad_sync_parity(comm, world, rank, h.device)
                failed = failed or not out["grad_sync"]["parity"]["ok"]
            except Exception as exc:
                out["grad_sync"]["parity"] = {"ok": False, "error": f"{type(exc).__name__}: {exc}"[:400]}
                failed = True
    try:
        h.close()
    except Exception as exc:
        out["close_error"] = f"{type(exc).__name__}: {exc}"[:200]
    del h, r
    torch.cuda.empty_cache()

    if args.impl == "b200" and not args.no_stock:
        # the optimisation target, same process group, same box, same workload, right after
        try:
            s = run_gpu_arm(args, "stock", world, local_rank, K, W, with_clocks=False)
            out["stock_gpu"] = {"value": s["value"], "unit": UNIT, "ms_per_step": s["ms_res"] / K,
                                "e2e": s["e2e"], "e2e_ms_per_step": s["ms_e2e"] / K, "final_loss": s["final_loss"],
                                "what": "same harness, stock torch: NCCL DDP (default buckets) + "
                                        "torch.optim.AdamW(fused=True) + ATen cross-entropy + torch clip_grad_norm_"}
            if world > 1:
                try:
                    out["stock_gpu"]["grad_sync"] = grad_sync_probe(s["h"], "stock", world)
                except Exception as exc:
                    out["stock_gpu"]["grad_sync"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
            out["vs_baseline"] = out["value"] / s["value"]
            out["vs_baseline_e2e"] = out["e2e"]["value"] / s["e2e"]
            out["vs_baseline_is"] = ("value / stock_gpu.value measured in this run (the reference publishes no "
                                     "number; north_star's target is >= 1.2x this path)")
            s["h"].close()
            del s
            torch.cuda.empty_cache()
        except Exception as exc:
            out["stock_gpu"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
    if args.impl == "b200" and not args.no_parity and (world > 1 or os.environ.get("B2_BENCH_LOSS_PARITY") == "1"):
        try:
            out["loss_parity"] = loss_parity(args, local_rank)
            failed = failed or not out["loss_parity"]["ok"]
        except Exception as exc:
            out["loss_parity"] = {"ok": False, "error": f"{type(exc).__name__}: {exc}"[:400]}
            failed = True

    if world > 1:
        if rank == 0:
            print(json.dumps(out), flush=True)
        dist.destroy_process_group()
        if failed:
            raise SystemExit(3)
        return
    if not args.no_cpu_baseline and args.impl == "b200":
        try:
            sps, threads, sample, _ = run_cpu(args, args.cpu_steps, 1)
            out["cpu_baseline"] = {"value": sps, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                                   "world": 1, "threads_per_rank": threads, "host_cores": os.cpu_count()}
            if args.cpu_omp1:
                sps1, t1, sample1, _ = run_cpu(args, max(1, args.cpu_steps // 5), 1, threads=1)
                out["cpu_baseline"]["omp1"] = {"value": sps1, "unit": UNIT, "cores": t1, "sample": sample1}
        except Exception as exc:
            out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                   "sample": f"failed: {type(exc).__name__}: {exc}"[:300]}
    print(json.dumps(out), flush=True)
    if failed:
        raise SystemExit(3)


if __name__ == "__main__":
    main()
