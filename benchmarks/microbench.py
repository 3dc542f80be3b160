#!/usr/bin/env python
"""Kernel microbenchmarks of BASELINE config 4: AdamW over the real tensor-size lists, masked CE
over N x 32128, and the all-reduce bucket sweep, each next to the stock library kernel it replaces.

    python benchmarks/microbench.py adamw  [--models t5-small flan-t5-base flan-t5-large flan-t5-xl]
    python benchmarks/microbench.py ce     [--rows 128 1024 8192 32768]
    torchrun --nproc-per-node N benchmarks/microbench.py allreduce [--sizes-mb 1 2 4 ...]

Timing: CUDA events on the launching stream, warm-up then median of --iters; between timed
iterations a 512 MB buffer is written to flush the 126 MB L2 whenever the kernel's working set
could fit in it (CE at small N); AdamW and large buckets exceed L2 by themselves. One JSON line
per measurement. `--only ours` keeps the run short for ncu.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

PEAK = 6549.8
try:
    PEAK = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
except Exception:
    pass

_flush = None


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    _flush.zero_()


def time_ms(fn, iters: int, warmup: int, flush: bool):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush:
            flush_l2()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b))
    return statistics.median(ts), min(ts)


def emit(**kw):
    print(json.dumps(kw), flush=True)


# ------------------------------------------------------------------------------------------
def bench_adamw(args):
    from distributed_llms_example_b200.optim import AdamW
    from distributed_llms_example_b200.synthetic import build_model

    for name in args.models:
        with torch.device("meta"):
            shapes = [tuple(p.shape) for p in build_model(name).parameters()]
        n = sum(int(np.prod(s)) for s in shapes)
        torch.manual_seed(0)
        ps = [torch.nn.Parameter(torch.randn(s, device="cuda") * 0.02) for s in shapes]
        for p in ps:
            p.grad = torch.randn_like(p) * 1e-2
        impls = {"b200": lambda: AdamW(ps, lr=5e-5, weight_decay=0.01)}
        if args.only != "ours":
            impls["torch_fused"] = lambda: torch.optim.AdamW(ps, lr=5e-5, weight_decay=0.01, fused=True)
            impls["torch_foreach"] = lambda: torch.optim.AdamW(ps, lr=5e-5, weight_decay=0.01, foreach=True)
        for impl, make in impls.items():
            opt = make()
            if impl == "b200":
                # kernel-only time (events tight around the launch) and whole step() call
                opt.kernel_events = []
                med, best = time_ms(opt.step, args.iters, args.warmup, False)
                torch.cuda.synchronize()
                k = [a.elapsed_time(b) for a, b in opt.kernel_events[args.warmup:]]
                kmed = statistics.median(k)
                emit(bench="adamw", model=name, impl=impl, params=n, tensors=len(shapes), step_call_ms=med,
                     kernel_ms=kmed, gbs=28 * n / kmed / 1e6, frac_measured_peak=28 * n / kmed / 1e6 / PEAK,
                     frac_8tbs=28 * n / kmed / 1e6 / 8000, launches_per_step=1)
            else:
                med, best = time_ms(opt.step, args.iters, args.warmup, False)
                emit(bench="adamw", model=name, impl=impl, params=n, tensors=len(shapes), step_call_ms=med,
                     gbs=28 * n / med / 1e6, frac_measured_peak=28 * n / med / 1e6 / PEAK)
            del opt
        del ps
        torch.cuda.empty_cache()


# ------------------------------------------------------------------------------------------
def bench_ce(args):
    from distributed_llms_example_b200.loss import cross_entropy

    v = 32128
    for tdt, s in ((torch.bfloat16, 2), (torch.float32, 4)):
        for n in args.rows:
            if n * v * s > 6 << 30:
                continue
            torch.manual_seed(0)
            z = (torch.randn(n, v, device="cuda") * 4).to(tdt).requires_grad_(True)
            y = torch.randint(0, v, (n,), device="cuda")
            y[torch.rand(n, device="cuda") < 0.2] = -100
            flush = 3 * n * v * s < (256 << 20)
            frac_valid = float((y != -100).float().mean())

            def ours_fwd():
                return cross_entropy(z, y)

            loss = ours_fwd()

            def ours_bwd():
                z.grad = None
                loss.backward(retain_graph=True)

            fmed, _ = time_ms(ours_fwd, args.iters, args.warmup, flush)
            bmed, _ = time_ms(ours_bwd, args.iters, args.warmup, flush)
            sweep = {}
            if args.ce_sweep and tdt is torch.bfloat16:
                for sp in (1, 2, 4, 8):
                    for cap in (0, 4):
                        os.environ["B2_CE_SPLIT"], os.environ["B2_CE_FWD_CTAS_PER_SM"] = str(sp), str(cap)
                        sweep[f"split{sp}_cap{cap}"] = round(run(fwd) * 1e3, 1)
                os.environ.pop("B2_CE_SPLIT"); os.environ.pop("B2_CE_FWD_CTAS_PER_SM")
            fb, bb = n * v * s, n * v * s * (1 + frac_valid)  # backward reads only the valid rows
            emit(bench="ce", dtype=str(tdt).split(".")[-1], rows=n, impl="b200", l2_flush=flush, fwd_ms=fmed,
                 bwd_ms=bmed, fwd_gbs=fb / fmed / 1e6, bwd_gbs=bb / bmed / 1e6,
                 fwd_frac_measured_peak=fb / fmed / 1e6 / PEAK, bwd_frac_measured_peak=bb / bmed / 1e6 / PEAK,
                 fwd_bwd_ms=fmed + bmed)
            if args.only == "ours":
                continue
            z.grad = None

            def stock_fwd():
                with torch.autocast("cuda", dtype=torch.bfloat16, enabled=(tdt == torch.bfloat16)):
                    return torch.nn.functional.cross_entropy(z, y, ignore_index=-100)

            sloss = stock_fwd()

            def stock_bwd():
                z.grad = None
                sloss.backward(retain_graph=True)

            sf, _ = time_ms(stock_fwd, args.iters, args.warmup, flush)
            sb, _ = time_ms(stock_bwd, args.iters, args.warmup, flush)
            emit(bench="ce", dtype=str(tdt).split(".")[-1], rows=n, impl="torch", l2_flush=flush, fwd_ms=sf,
                 bwd_ms=sb, fwd_bwd_ms=sf + sb, speedup_fwd_bwd=(sf + sb) / (fmed + bmed))
            del z, y
            torch.cuda.empty_cache()


# ------------------------------------------------------------------------------------------
def bench_ce_kernel(args):
    """Kernel-only CE timing straight through the C ABI: `reps` back-to-back launches between two
    events, rotating over enough distinct input buffers that the rotation exceeds the 126 MB L2
    (so small batches are measured cold, like they are inside a training step)."""
    from distributed_llms_example_b200 import _lib as L

    lib = L.load()
    v = 32128
    stream = torch.cuda.current_stream().cuda_stream
    for tdt, s, code in ((torch.bfloat16, 2, L.B2_BF16), (torch.float32, 4, L.B2_F32)):
        for n in args.rows:
            nbuf = max(2, min(64, -(-(400 << 20) // (n * v * s))))
            if nbuf * n * v * s * 2 > 60 << 30:
                nbuf = 2
            torch.manual_seed(0)
            zs = [(torch.randn(n, v, device="cuda") * 4).to(tdt) for _ in range(nbuf)]
            ds = [torch.empty_like(z) for z in zs]
            y = torch.randint(0, v, (n,), device="cuda")
            y[torch.rand(n, device="cuda") < 0.2] = -100
            frac_valid = float((y != -100).float().mean())
            lse = torch.empty(n, device="cuda")
            loss = torch.empty(1, device="cuda")
            nvalid = torch.empty(1, device="cuda")
            g = torch.ones(1, device="cuda")
            nb = lib.b2_ce_workspace_bytes(n, v)
            ws = torch.zeros(nb, dtype=torch.uint8, device="cuda")
            reps = max(nbuf, 8)

            def fwd(i):
                z = zs[i % nbuf]
                L.check(lib.b2_ce_forward(z.data_ptr(), y.data_ptr(), n, v, v, -100, code, lse.data_ptr(),
                                          loss.data_ptr(), nvalid.data_ptr(), ws.data_ptr(), nb, stream))

            def bwd(i):
                z, d = zs[i % nbuf], ds[i % nbuf]
                L.check(lib.b2_ce_backward(z.data_ptr(), y.data_ptr(), lse.data_ptr(), g.data_ptr(),
                                           nvalid.data_ptr(), n, v, v, -100, code, d.data_ptr(), stream))

            def run(fn):
                for i in range(reps):
                    fn(i)
                torch.cuda.synchronize()
                ts = []
                for _ in range(args.iters):
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    for i in range(reps):
                        fn(i)
                    b.record()
                    b.synchronize()
                    ts.append(a.elapsed_time(b) / reps)
                return statistics.median(ts)

            fwd(0)
            f_ms, b_ms = run(fwd), run(bwd)
            poly_ms = None
            if tdt is torch.bfloat16:  # same process, same buffers: MUFU-only vs a quarter on the FMA pipe
                os.environ["B2_CE_POLY"] = "1" if os.environ.get("B2_CE_POLY") != "1" else "0"
                poly_ms = run(fwd)
                os.environ["B2_CE_POLY"] = "1" if os.environ.get("B2_CE_POLY") != "1" else "0"
            sweep = {}
            if args.ce_sweep and tdt is torch.bfloat16:
                for sp in (1, 2, 4, 8):
                    for cap in (0, 4):
                        os.environ["B2_CE_SPLIT"], os.environ["B2_CE_FWD_CTAS_PER_SM"] = str(sp), str(cap)
                        sweep[f"split{sp}_cap{cap}"] = round(run(fwd) * 1e3, 1)
                os.environ.pop("B2_CE_SPLIT"); os.environ.pop("B2_CE_FWD_CTAS_PER_SM")
            fb, bb = n * v * s, n * v * s * (1 + frac_valid)
            emit(bench="ce_kernel", dtype=str(tdt).split(".")[-1], rows=n, buffers=nbuf, reps=reps,
                 fwd_us=f_ms * 1e3, bwd_us=b_ms * 1e3, fwd_gbs=fb / f_ms / 1e6, bwd_gbs=bb / b_ms / 1e6,
                 fwd_frac_measured_peak=fb / f_ms / 1e6 / PEAK, bwd_frac_measured_peak=bb / b_ms / 1e6 / PEAK,
                 fwd_us_other_exp_mode=None if poly_ms is None else poly_ms * 1e3, fwd_us_sweep=sweep,
                 env={k: os.environ[k] for k in os.environ if k.startswith("B2_CE")})
            del zs, ds
            torch.cuda.empty_cache()


# ------------------------------------------------------------------------------------------
def bench_lmce(args):
    """Fused lm_head + CE (tcgen05) next to what it replaces: a bf16 library GEMM writing the
    [N, V] logits followed by this repo's CE kernels, and by ATen's. d_model 1024, vocab 32128.
    Tensor-pipe roofline: 2*N*V*D flop per pass against MEASURED_PEAKS.json bf16_tflops (burst)."""
    import torch.nn.functional as F

    from distributed_llms_example_b200 import _lib as L
    from distributed_llms_example_b200.loss import cross_entropy

    lib = L.load()
    v, d = 32128, args.d_model
    peak_tf = 1654.5
    try:
        peak_tf = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["bf16_tflops"])
    except Exception:
        pass
    stream = torch.cuda.current_stream().cuda_stream
    for n in args.rows:
        torch.manual_seed(0)
        h = torch.randn(n, d, device="cuda").to(torch.bfloat16)
        w = (torch.randn(v, d, device="cuda") * (4.0 / d ** 0.5)).to(torch.bfloat16)
        y = torch.randint(0, v, (n,), device="cuda")
        y[torch.rand(n, device="cuda") < 0.2] = -100
        lse = torch.empty(n, device="cuda")
        loss = torch.empty(1, device="cuda")
        nvalid = torch.empty(1, device="cuda")
        g = torch.ones(1, device="cuda")
        nb = lib.b2_lmce_workspace_bytes(n, v)
        ws = torch.zeros(nb, dtype=torch.uint8, device="cuda")
        ld = (v + 7) // 8 * 8
        dz = torch.empty(n, ld, dtype=torch.bfloat16, device="cuda")

        def fused_fwd():
            L.check(lib.b2_lmce_forward(h.data_ptr(), w.data_ptr(), y.data_ptr(), n, v, d, -100, 1, lse.data_ptr(),
                                        loss.data_ptr(), nvalid.data_ptr(), None, 0, ws.data_ptr(), nb, stream))

        zst = torch.empty(n, ld, dtype=torch.bfloat16, device="cuda")

        def fused_fwd_store():
            L.check(lib.b2_lmce_forward(h.data_ptr(), w.data_ptr(), y.data_ptr(), n, v, d, -100, 1, lse.data_ptr(),
                                        loss.data_ptr(), nvalid.data_ptr(), zst.data_ptr(), ld, ws.data_ptr(), nb,
                                        stream))

        def fused_fwd_store_plus_ce_bwd():
            fused_fwd_store()
            L.check(lib.b2_ce_backward(zst.data_ptr(), y.data_ptr(), lse.data_ptr(), g.data_ptr(), nvalid.data_ptr(), n,
                                       v, ld, -100, L.B2_BF16, dz.data_ptr(), stream))

        def fused_fwd_plus_recompute_bwd():
            fused_fwd()
            fused_bwd()

        def fused_bwd():
            L.check(lib.b2_lmce_backward_dlogits(h.data_ptr(), w.data_ptr(), y.data_ptr(), lse.data_ptr(), g.data_ptr(),
                                                 nvalid.data_ptr(), n, v, d, -100, 1, dz.data_ptr(), ld, stream))

        def gemm_only():
            return F.linear(h, w)

        def unfused_b200_fwd():
            return cross_entropy(F.linear(h, w), y)

        def unfused_aten_fwd():
            return F.cross_entropy(F.linear(h, w).float(), y, ignore_index=-100)

        hz = h.clone().requires_grad_(True)

        def unfused_b200_fwd_bwd_dlogits():
            z = F.linear(hz.detach(), w).requires_grad_(True)
            cross_entropy(z, y).backward()
            return z.grad

        flop = 2.0 * n * v * d
        flush = n * v * 2 < (200 << 20)
        rec = {"bench": "lmce", "rows": n, "vocab": v, "d_model": d, "flop_per_pass": flop}
        for name, fn in (("fused_fwd", fused_fwd), ("fused_fwd_store", fused_fwd_store),
                         ("fused_bwd_dlogits", fused_bwd),
                         ("fused_fwd_store_plus_ce_bwd", fused_fwd_store_plus_ce_bwd),
                         ("fused_fwd_plus_recompute_bwd", fused_fwd_plus_recompute_bwd),
                         ("cublas_gemm_only", gemm_only),
                         ("unfused_gemm_plus_b200_ce_fwd", unfused_b200_fwd),
                         ("unfused_gemm_plus_aten_ce_fwd", unfused_aten_fwd),
                         ("unfused_gemm_plus_b200_ce_fwd_bwd_dlogits", unfused_b200_fwd_bwd_dlogits)):
            if args.only == "ours" and not name.startswith("fused"):
                continue
            med, best = time_ms(fn, args.iters, args.warmup, flush)
            rec[name + "_us"] = med * 1e3
            if name in ("fused_fwd", "fused_fwd_store", "fused_bwd_dlogits", "cublas_gemm_only"):
                rec[name + "_tflops"] = flop / (med / 1e3) / 1e12
                rec[name + "_frac_of_measured_bf16_peak"] = flop / (med / 1e3) / 1e12 / peak_tf
        emit(**rec)
        del h, w, dz
        torch.cuda.empty_cache()


# ------------------------------------------------------------------------------------------
def bench_allreduce_emulated(args):
    """One GPU, W virtual ranks in ONE cooperative launch (b2_allreduce_emulated): the same device
    code as the multi-GPU kernels, runnable under ncu (which must never wrap a multi-rank command).
    Buffers are local, so timings say nothing about NVLink; the capture shows instruction mix,
    stalls and local traffic of the staging (1, 2), in-place pull (3) and push (4) kernels."""
    import ctypes

    from distributed_llms_example_b200 import _lib as L

    lib = L.load()
    w = 8
    for mb in args.sizes_mb:
        n = int(mb * (1 << 20)) // 4
        bufs = [torch.randn(n, device="cuda") * 1e-2 for _ in range(w)]
        scratch = torch.empty(lib.b2_allreduce_emulated_scratch_bytes(w, n), dtype=torch.uint8, device="cuda")
        ptrs = (ctypes.c_void_p * w)(*[b.data_ptr() for b in bufs])
        for algo in (2, 4):
            def fn():
                L.check(lib.b2_allreduce_emulated(ptrs, w, n, 0, 1.0 / w, algo, scratch.data_ptr(),
                                                  torch.cuda.current_stream().cuda_stream))
            med, _ = time_ms(fn, max(3, args.iters), 2, False)
            emit(bench="allreduce_emulated", world=w, size_mb=mb, algo={2: "two_shot_staging", 4: "push_in_place"}[algo],
                 ms=med, note="single GPU, local buffers: not an NVLink number")


# ------------------------------------------------------------------------------------------
def bench_allreduce(args):
    from distributed_llms_example_b200.comm import P2PComm

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm = P2PComm(staging_bytes=args.staging_mb << 20)
    nvls = world > 1 and comm.enable_nvls((int(sum(args.sizes_mb) * 1.1) + 256) << 20)
    pool = comm.nvls_pool() if nvls else None
    if rank == 0:
        print(json.dumps({"nvls_enabled": bool(nvls), "nvls_error": comm.nvls_error}), flush=True)

    def timed(fn, iters, warmup):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        b.synchronize()
        t = torch.tensor([a.elapsed_time(b) / iters], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for mb in args.sizes_mb:
        n = int(mb * (1 << 20)) // 4
        x = torch.randn(n, device="cuda") * 1e-2
        iters = max(5, min(args.iters, int(2000 / max(mb, 1))))
        rows = []
        for wire in args.wires:
            for algo in args.algos:
                if world == 1 and algo != "auto":
                    continue
                ms = timed(lambda: comm.allreduce_(x, 1.0 / world, wire=wire, algo=algo), iters, args.warmup)
                rows.append((f"b200_{wire}_{algo}", ms))
        if world > 1 and nvls:
            rows.append(("b200_fp32_nvls_staged", timed(lambda: comm.allreduce_nvls_(x, 1.0 / world), iters, args.warmup)))
            with torch.cuda.use_mem_pool(pool):
                xp = torch.empty(n, device="cuda")  # exactly one allocation inside the region
            xp.normal_().mul_(1e-2)
            off = comm.nvls_offset(xp)
            if off is not None:
                rows.append(("b200_fp32_nvls_zero_copy",
                             timed(lambda: comm.allreduce_nvls_inplace_(off, n, 1.0 / world), iters, args.warmup)))
            del xp
        if world > 1:
            reg = comm.register(x)
            rows.append(("b200_fp32_zero_copy_inplace",
                         timed(lambda: comm.allreduce_registered_(reg, x, 1.0 / world), iters, args.warmup)))
            comm.unregister(reg)
        if world > 1 and args.only != "ours":
            def nccl():
                x.div_(world)
                dist.all_reduce(x)
            rows.append(("nccl_div_allreduce", timed(nccl, iters, args.warmup)))
            rows.append(("nccl_allreduce_only", timed(lambda: dist.all_reduce(x), iters, args.warmup)))
        if rank == 0:
            for impl, ms in rows:
                bus = n * 4 * 2 * (world - 1) / world / ms / 1e6 if world > 1 else 0.0
                emit(bench="allreduce", world=world, size_mb=mb, impl=impl, ms=ms, bus_gbs=bus,
                     frac_of_900=bus / 900.0, alg_gbs=n * 4 / ms / 1e6)
        del x
    comm.check()
    comm.close()
    if world > 1:
        dist.destroy_process_group()


def bench_allreduce_tune(args):
    """Event-timed sweep of the zero-copy all-reduce kernels' launch shape (CTAs x packs per thread x
    loads in flight) at DDP-sized buckets: ncu must not wrap multi-rank commands, so this table is the
    tuning evidence for the NVLink-bound kernels. One JSON line per (path, size, shape)."""
    from distributed_llms_example_b200.comm import P2PComm

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm = P2PComm(staging_bytes=args.staging_mb << 20)
    nvls = comm.enable_nvls((int(sum(args.sizes_mb) * 1.1) + 256) << 20)
    pool = comm.nvls_pool() if nvls else None

    def timed(fn, iters, warmup=3):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        b.synchronize()
        t = torch.tensor([a.elapsed_time(b) / iters], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for sumsq in (False, True):
        comm.enable_sumsq(sumsq)
        for mb in args.sizes_mb:
            n = int(mb * (1 << 20)) // 4
            x = torch.randn(n, device="cuda") * 1e-2
            reg = comm.register(x)
            paths = [("ipc_push", lambda: comm.allreduce_registered_(reg, x, 1.0 / world))]
            if nvls:
                with torch.cuda.use_mem_pool(pool):
                    xp = torch.empty(n, device="cuda")
                xp.normal_().mul_(1e-2)
                off = comm.nvls_offset(xp)
                paths.append(("nvls_inplace", lambda: comm.allreduce_nvls_inplace_(off, n, 1.0 / world)))
            iters = max(5, min(args.iters, int(1500 / max(mb, 1))))
            shapes = [(b, p, u) for b in (48, 96, 128) for p in (2, 4, 8) for u in (0, 1, 2)]  # u: push 0/1/2 = tuned/more/fewer; NVLS (since the r2 sweep) 0|2 = 2, 1 = 8, 3 = 4 loads if not sumsq else [(-1, -1, -1)]
            for path, fn in paths:
                for (blocks, ppt, unroll) in shapes:
                    if blocks > 0:
                        comm.tune(blocks, ppt, unroll)
                    ms = timed(fn, iters)
                    if rank == 0:
                        bus = n * 4 * 2 * (world - 1) / world / ms / 1e6
                        emit(bench="allreduce_tune", world=world, size_mb=mb, path=path, max_blocks=blocks,
                             packs_per_thread=ppt, unroll=unroll, sumsq=sumsq, us=ms * 1e3, bus_gbs=bus,
                             frac_of_900=bus / 900.0)
            comm.tune(int(os.environ.get("B2_AR_BLOCKS", "128")), int(os.environ.get("B2_AR_PACKS_PER_THREAD", "4")),
                      int(os.environ.get("B2_AR_UNROLL", "0")))
            comm.unregister(reg)
            del x
        comm.sumsq_clip_(1.0, torch.empty(2, device="cuda")) if sumsq else None
    comm.enable_sumsq(False)
    comm.check()
    comm.close()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("bench", choices=["adamw", "ce", "ce_kernel", "lmce", "allreduce", "allreduce_emulated", "allreduce_tune"])
    ap.add_argument("--d-model", type=int, default=1024)
    ap.add_argument("--ce-sweep", action="store_true", help="ce_kernel: sweep row split x grid cap for the bf16 forward")
    ap.add_argument("--models", nargs="+", default=["t5-small", "flan-t5-base", "flan-t5-large", "flan-t5-xl"])
    ap.add_argument("--rows", nargs="+", type=int, default=[128, 1024, 8192, 32768])
    ap.add_argument("--sizes-mb", nargs="+", type=float, default=[1, 2, 4, 8, 16, 25, 32, 64, 128, 256, 512, 1024])
    ap.add_argument("--wires", nargs="+", default=["fp32", "bf16"])
    ap.add_argument("--algos", nargs="+", default=["auto", "one_shot", "two_shot"])
    ap.add_argument("--staging-mb", type=int, default=64)
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--only", default="all", choices=["all", "ours"])
    args = ap.parse_args()
    {"adamw": bench_adamw, "ce": bench_ce, "ce_kernel": bench_ce_kernel, "lmce": bench_lmce, "allreduce": bench_allreduce, "allreduce_tune": bench_allreduce_tune,
     "allreduce_emulated": bench_allreduce_emulated}[args.bench](args)


if __name__ == "__main__":
    main()
