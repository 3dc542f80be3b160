"""Host-side logic on CPU: collate parity, launcher rebinding, patch points, the comm-hook
signature DDP checks, and the N>1 control-plane path over gloo with world size 2."""
import inspect
import os
import socket
import sys
import textwrap

import numpy as np
import pytest
import torch
import torch.distributed as dist

from oracle import ref_numpy as O
from oracle.inputs import allreduce_inputs, collate_inputs
from oracle.make_golden import COLLATE_CASES


# ---- (a6) collate: product vs committed DataCollatorForSeq2Seq outputs, bit-exact -------------
@pytest.mark.parametrize("name", list(COLLATE_CASES))
def test_collator_bit_exact_vs_transformers_golden(golden, name):
    from distributed_llms_example_b200.collate import Seq2SeqCollator

    b, mode, mult = COLLATE_CASES[name]
    out = Seq2SeqCollator(pad_to_multiple_of=mult)(collate_inputs(b, mode))
    for k in ("input_ids", "attention_mask", "labels", "decoder_input_ids"):
        ref = golden["collate"][f"{name}/{k}"]
        assert out[k].dtype == torch.int64
        np.testing.assert_array_equal(out[k].numpy(), ref)


def test_collator_matches_live_transformers_collator_on_synthetic_batches():
    """Same comparison against the real DataCollatorForSeq2Seq + T5._shift_right, run live, on the
    benchmark's own synthetic features (both label modes, ragged and maximum lengths)."""
    from distributed_llms_example_b200.collate import Seq2SeqCollator
    from distributed_llms_example_b200.synthetic import make_features
    from oracle.torch_ref import collate_ref

    for mode, mult in (("REF", None), ("MASK", 8), ("MASK", None)):
        feats = make_features(5, rank=1, label_mode=mode, src_len=96, tgt_len=40)
        ours = Seq2SeqCollator(pad_to_multiple_of=mult)(feats)
        ref = collate_ref(feats, pad_to_multiple_of=mult)
        for k in ("input_ids", "attention_mask", "labels", "decoder_input_ids"):
            assert torch.equal(ours[k], ref[k]), (mode, mult, k)
        if mode == "REF":
            assert not bool((ours["labels"] == -100).any())  # the scripts produce no -100
        else:
            lab = ours["labels"]
            assert bool((lab == -100).any())
            # -100 only as a suffix of each row
            first = (lab == -100).float().argmax(dim=1)
            for i in range(lab.shape[0]):
                if (lab[i] == -100).any():
                    assert bool((lab[i, first[i]:] == -100).all())


def test_collator_edge_cases():
    from distributed_llms_example_b200.collate import Seq2SeqCollator, shift_right

    with pytest.raises(ValueError):
        Seq2SeqCollator()([])
    one = Seq2SeqCollator(pad_to_multiple_of=8)([{"input_ids": [5, 1], "attention_mask": [1, 1], "labels": [7, 1]}])
    assert one["input_ids"].shape == (1, 8) and one["labels"].tolist() == [[7, 1] + [-100] * 6]
    assert one["decoder_input_ids"].tolist() == [[0, 7, 1, 0, 0, 0, 0, 0]]
    np.testing.assert_array_equal(shift_right(one["labels"]).numpy(), O.shift_right(one["labels"].numpy()))


# ---- launcher / patch points ----------------------------------------------------------------
def test_launcher_rebinds_average_gradients_without_editing_the_script(tmp_path):
    from distributed_llms_example_b200 import launch

    script = tmp_path / "fake_train_task.py"
    script.write_text(textwrap.dedent("""
        import sys
        CALLS = []
        def average_gradients(model):
            CALLS.append(("script", model))
        def train(model):
            average_gradients(model)   # resolved through the module namespace, like train-task.py:293
        if __name__ == "__main__":
            train(sys.argv[1])
    """))
    seen = []
    globs = launch.run_script(str(script), ["the-model"], rebinds={"average_gradients": lambda m: seen.append(m)},
                              apply_patches=False)
    assert seen == ["the-model"] and globs["CALLS"] == []
    assert "def average_gradients" in script.read_text()  # file untouched
    assert sys.argv[0] != str(script)  # argv restored


def test_patch_points_rebind_and_restore():
    import transformers.models.t5.modeling_t5 as t5

    from distributed_llms_example_b200 import patch
    from distributed_llms_example_b200.loss import CrossEntropyLoss
    from distributed_llms_example_b200.optim import AdamW

    stock_ce, stock_adamw = t5.CrossEntropyLoss, torch.optim.AdamW
    patch.apply(ddp=False)
    try:
        assert t5.CrossEntropyLoss is CrossEntropyLoss and torch.optim.AdamW is AdamW
        assert isinstance(CrossEntropyLoss(ignore_index=-100), torch.nn.Module)
        assert issubclass(AdamW, torch.optim.Optimizer)  # LambdaLR requires an Optimizer instance
    finally:
        patch.remove()
    assert t5.CrossEntropyLoss is stock_ce and torch.optim.AdamW is stock_adamw


def _patch_ddp_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from torch.nn.parallel import DistributedDataParallel as DDP

        from distributed_llms_example_b200 import comm, patch

        calls = []
        real_register, real_wants = comm.register_p2p_comm_hook, patch._wants_hook
        comm.register_p2p_comm_hook = lambda ddp, **kw: calls.append((ddp, kw)) or "state"
        patch.apply(cross_entropy=False, adamw=False, fold_clip=False, comm_kwargs={"wire": "bf16"})
        try:
            # 1. a CPU module over gloo is left exactly as torch builds it
            plain = DDP(torch.nn.Linear(4, 4))
            ok_cpu = not calls and plain.gradient_as_bucket_view is False and not hasattr(plain, "_b200_hook_state")
            # 2. a module the predicate accepts (on the GPU box: CUDA parameters, world > 1) gets the hook
            #    exactly once, with the kwargs of patch.apply, and gradients as bucket views
            patch._wants_hook = lambda module, kwargs: True
            hooked = DDP(torch.nn.Linear(4, 4))
            ok_hook = (len(calls) == 1 and calls[0][0] is hooked and calls[0][1] == {"wire": "bf16"}
                       and hooked._b200_hook_state == "state" and hooked.gradient_as_bucket_view is True)
            # 3. the caller's own choices win over the plug-in's defaults
            os.environ["B2_DDP_BUCKET_MB"] = "50"
            own = DDP(torch.nn.Linear(4, 4), gradient_as_bucket_view=False, bucket_cap_mb=7)
            dflt = DDP(torch.nn.Linear(4, 4))
            ok_kwargs = (own.gradient_as_bucket_view is False and own.bucket_bytes_cap == 7 * 1024 * 1024
                         and dflt.bucket_bytes_cap == 50 * 1024 * 1024 and len(calls) == 3
                         and hooked.bucket_bytes_cap == 128 * 1024 * 1024     # the plug-in's own default
                         and plain.bucket_bytes_cap == 25 * 1024 * 1024)      # torch's, untouched
            # 4. apply() twice does not wrap twice; remove() restores torch's constructor
            patch.apply(cross_entropy=False, adamw=False, fold_clip=False)
            DDP(torch.nn.Linear(4, 4))
            ok_once = len(calls) == 4 and calls[-1][1] == {}
        finally:
            os.environ.pop("B2_DDP_BUCKET_MB", None)
            patch._wants_hook = real_wants
            comm.register_p2p_comm_hook = real_register
            patch.remove()
        after = DDP(torch.nn.Linear(4, 4))
        ok_removed = len(calls) == 4 and after.gradient_as_bucket_view is False
        q.put((rank, (ok_cpu, ok_hook, ok_kwargs, ok_once, ok_removed), None))
    except Exception:
        import traceback

        q.put((rank, (), traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_patch_ddp_registers_the_hook_once_on_gpu_modules_only():
    """patch.patch_ddp is how the comm hook reaches an unmodified script (accelerator.prepare and HF
    Trainer build DistributedDataParallel themselves): the wrapped constructor must attach the hook
    exactly once per DDP module, only where the predicate holds, and respect the caller's kwargs."""
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_patch_ddp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, oks, err in outs:
        assert err is None, f"rank {rank}:\n{err}"
        assert oks and all(oks), (rank, oks)


def test_wants_hook_predicate():
    from distributed_llms_example_b200 import patch

    assert patch._wants_hook(torch.nn.Linear(2, 2), {}) is False       # CPU parameters
    assert patch._wants_hook(torch.nn.ReLU(), {}) is False             # no parameters at all


def test_adamw_refuses_state_dicts_with_other_semantics():
    """A checkpoint written by torch.optim.AdamW(amsgrad=True) must not resume here with silently
    different optimizer semantics (the constructor rejects the flag; so must load_state_dict)."""
    from distributed_llms_example_b200.optim import AdamW

    p = torch.nn.Parameter(torch.zeros(4))
    ref = torch.optim.AdamW([p], lr=1e-3, amsgrad=True)
    p.grad = torch.ones(4)
    ref.step()
    ours = AdamW([p], lr=1e-3)
    with pytest.raises(NotImplementedError, match="amsgrad"):
        ours.load_state_dict(ref.state_dict())
    assert ours.param_groups[0]["amsgrad"] is False  # refused before anything was overwritten
    ok = torch.optim.AdamW([p], lr=1e-3)
    ok.step()
    ours.load_state_dict(ok.state_dict())  # a plain AdamW checkpoint still loads
    bad = ours.state_dict()
    bad["param_groups"][0]["maximize"] = True
    with pytest.raises(NotImplementedError, match="maximize"):
        ours.load_state_dict(bad)


def test_comm_hook_signature_is_what_ddp_checks():
    from distributed_llms_example_b200.comm import p2p_allreduce_hook

    sig = inspect.signature(p2p_allreduce_hook)
    assert sig.parameters["bucket"].annotation == dist.GradBucket
    assert sig.return_annotation == torch.futures.Future[torch.Tensor]


def test_adamw_constructor_contract_on_cpu():
    """Construction, param_groups, LambdaLR and state_dict work without a GPU; only step() needs one."""
    from transformers import get_scheduler

    from distributed_llms_example_b200.optim import AdamW

    ps = [torch.nn.Parameter(torch.zeros(3)), torch.nn.Parameter(torch.zeros(2))]
    opt = AdamW([{"params": ps[:1], "weight_decay": 0.01}, {"params": ps[1:], "weight_decay": 0.0}], lr=5e-5,
                betas=(0.9, 0.999), eps=1e-8, fused=True)  # Trainer passes fused=True
    sched = get_scheduler("linear", opt, num_warmup_steps=1, num_training_steps=10)
    assert opt.param_groups[0]["lr"] == 0.0  # lr is 0 before the first step
    assert opt.param_groups[0]["weight_decay"] == 0.01 and opt.param_groups[1]["weight_decay"] == 0.0
    opt.zero_grad()
    assert opt.step() is None  # no grads: nothing to do, no kernel
    sched.step()
    assert opt.param_groups[0]["lr"] == pytest.approx(5e-5)
    sd = opt.state_dict()
    assert len(sd["param_groups"]) == 2
    with pytest.raises(ValueError):
        AdamW(ps, lr=-1.0)


# ---- N > 1 over gloo -------------------------------------------------------------------------
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from distributed_llms_example_b200.comm import exchange_handles

        # 1. handle exchange: rank order, exact bytes
        mine = bytes([rank]) * 64
        got = exchange_handles(mine)
        ok_handles = got == [bytes([r]) * 64 for r in range(world)]
        try:
            exchange_handles(b"short")
            ok_len = False
        except ValueError:
            ok_len = True

        # 2. the hook contract DDP expects, with a torch-only stand-in hook defined HERE in the test:
        #    mean all-reduce in rank order, result identical on all ranks and equal to the oracle.
        def rank_ordered_mean_hook(state, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
            buf = bucket.buffer()
            parts = [torch.empty_like(buf) for _ in range(world)]
            dist.all_gather(parts, buf)
            acc = parts[0] / world
            for p in parts[1:]:
                acc = acc + p / world
            buf.copy_(acc)
            fut = torch.futures.Future()
            fut.set_result(buf)
            return fut

        def run(hook):
            torch.manual_seed(0)
            model = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 64))
            ddp = torch.nn.parallel.DistributedDataParallel(model)
            if hook is not None:
                ddp.register_comm_hook(None, hook)
            opt = torch.optim.AdamW(ddp.parameters(), lr=1e-3, weight_decay=0.01)
            g = torch.Generator().manual_seed(1234 + rank)
            for _ in range(3):
                ddp(torch.randn(8, 64, generator=g)).square().mean().backward()
                opt.step(); opt.zero_grad()
            return torch.cat([p.detach().reshape(-1) for p in model.parameters()])

        a, b = run(rank_ordered_mean_hook), run(None)
        hook_close = bool(torch.allclose(a, b, rtol=1e-6, atol=1e-8))
        parts = [torch.empty_like(a) for _ in range(world)]
        dist.all_gather(parts, a)
        replicas_equal = all(torch.equal(parts[0], p) for p in parts)

        # 3. train-task.py:65-69 per-tensor all_reduce + "/= size" equals the rank-ordered oracle at W=2
        bufs = allreduce_inputs(world, 1000, 77)
        t = torch.from_numpy(bufs[rank].copy())
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        t /= float(world)
        oracle_equal = bool(np.array_equal(t.numpy(), O.allreduce_mean(bufs)))
        q.put((rank, ok_handles, ok_len, hook_close, replicas_equal, oracle_equal, None))
    except Exception:
        import traceback

        q.put((rank, False, False, False, False, False, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_control_plane_and_hook_contract():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, ok_handles, ok_len, hook_close, replicas_equal, oracle_equal, err in outs:
        assert err is None, f"rank {rank}:\n{err}"
        assert ok_handles and ok_len
        assert hook_close and replicas_equal and oracle_equal


def test_product_refuses_cpu_tensors_instead_of_falling_back():
    from distributed_llms_example_b200.harness import StepConfig, StepHarness
    from distributed_llms_example_b200.loss import cross_entropy

    with pytest.raises(RuntimeError):
        cross_entropy(torch.zeros(2, 8), torch.zeros(2, dtype=torch.int64))
    with pytest.raises(RuntimeError):
        StepHarness(StepConfig(model="t5-small", impl="b200", device="cpu"))


# ---- zero-copy registration vote (no GPU: a stand-in communicator records the calls) -----------
class _FakeComm:
    """Duck-types the four members HookState.registration_for touches."""

    def __init__(self, rank, world, exportable=True):
        self.rank, self.world, self.group, self.wire = rank, world, None, "fp32"
        self.device = torch.device("cpu")
        self.exportable = exportable
        self.registered = []

    def export(self, buf):
        from distributed_llms_example_b200 import _lib

        if not self.exportable:
            raise _lib.B2Error(-2, "cannot export")
        return (bytes([self.rank]) * 64, 512 * self.rank)

    def register(self, buf, peers=None):
        self.registered.append(peers)
        return len(self.registered) - 1


class _Buf:
    def __init__(self, ptr, n):
        self._p, self._n = ptr, n

    def data_ptr(self):
        return self._p

    def numel(self):
        return self._n


def _vote_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from distributed_llms_example_b200.comm import HookState

        def state(exportable=True):
            st = HookState.__new__(HookState)  # skip the CUDA stream/event members: host logic only
            st.comm = _FakeComm(rank, world, exportable)
            st.zero_copy, st.zero_copy_min_bytes = True, 1 << 20
            st._last_seen, st._registered = {}, {}
            st._attempts, st._never, st.max_attempts = {}, set(), 3
            st.cpp_handle = None
            return st

        big = 1 << 20  # elements -> 4 MiB
        st = state()
        # step 0: first sighting -> nobody registers; step 1: same address everywhere -> register
        r0 = st.registration_for(3, _Buf(0x1000 + rank, big))
        r1 = st.registration_for(3, _Buf(0x1000 + rank, big))
        r2 = st.registration_for(3, _Buf(0x1000 + rank, big))  # cached, no collective
        peers = st.comm.registered[0]
        ok_basic = (r0 is None and r1 == (0x1000 + rank, big, 0, None) and r2 == r1 and
                    peers == [(bytes([r]) * 64, 512 * r) for r in range(world)])
        # a bucket below the size threshold never votes
        ok_small = st.registration_for(0, _Buf(0x2000, 16)) is None and 0 not in st._last_seen
        # one rank sees the bucket move between steps -> all ranks decline together, then agree later
        st2 = state()
        st2.registration_for(5, _Buf(0x3000, big))
        moved = st2.registration_for(5, _Buf(0x3000 + (0x100 if rank == 1 else 0), big))
        again = st2.registration_for(5, _Buf(0x3000 + (0x100 if rank == 1 else 0), big))
        ok_move = moved is None and again is not None and again[2] == 0
        # memory that cannot be exported on one rank (expandable segments) -> nobody registers
        st3 = state(exportable=(rank != 0))
        st3.registration_for(7, _Buf(0x4000, big))
        ok_noexport = st3.registration_for(7, _Buf(0x4000, big)) is None and not st3.comm.registered
        # ... and that verdict is permanent on EVERY rank after one collective: later calls return
        # without another all_gather_object (a rank that still voted would hang its peer here)
        ok_noexport = ok_noexport and 7 in st3._never
        for _ in range(3):
            ok_noexport = ok_noexport and st3.registration_for(7, _Buf(0x4000, big)) is None
        # a bucket that never settles is abandoned after max_attempts votes, on every rank alike
        st4 = state()
        for i in range(6):
            st4.registration_for(9, _Buf(0x5000 + 0x100 * i * (rank == 1), big))
        ok_noexport = ok_noexport and 9 in st4._never and not st4.comm.registered
        # a registered bucket that moves is an error, not a silent divergence
        try:
            st.registration_for(3, _Buf(0x9999, big))
            ok_raise = False
        except RuntimeError:
            ok_raise = True
        q.put((rank, ok_basic, ok_small, ok_move, ok_noexport, ok_raise, None))
    except Exception:
        import traceback

        q.put((rank, False, False, False, False, False, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_zero_copy_registration_is_a_collective_agreement():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_vote_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, *oks, err in outs:
        assert err is None, f"rank {rank}:\n{err}"
        assert all(oks), (rank, oks)


def test_bench_reference_arm_prints_one_json_line():
    """`bench.py --impl reference` (the CPU arm the driver runs next to ours) on the smallest model."""
    import json
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--model", "t5-small",
                          "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "train_samples_per_second" and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["higher_is_better"] is True


# ---- BASELINE config 1: t5-small, torchrun-style world size 2 over gloo on CPU -------------------
def _config1_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(2)
        from distributed_llms_example_b200.harness import StepConfig, StepHarness

        # small sequences keep the CPU test short; the step logic is what is under test
        import distributed_llms_example_b200.harness as H

        real = H.make_features
        H.make_features = lambda n, **kw: real(n, src_len=48, tgt_len=16, **kw)
        h = StepHarness(StepConfig(model="t5-small", semantics="trainer", impl="stock", device="cpu", batch_size=1,
                                   grad_accum=2, label_mode="MASK", bf16=False, dropout=0.0, total_steps=10))
        batches = h.host_batches(3 * 2)
        losses = [float(h.step(batches[i * 2:(i + 1) * 2])) for i in range(3)]
        flat = torch.cat([p.detach().reshape(-1) for p in h.model.parameters()])
        parts = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(parts, flat)
        q.put((rank, losses, all(torch.equal(parts[0], p) for p in parts), h.ddp is not None, None))
    except Exception:
        import traceback

        q.put((rank, [], False, False, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_config1_reference_cpu_path_world_2_gloo():
    """The reference's own CPU-runnable case (BASELINE.json configs[0]) through the step harness:
    DDP over gloo, gradient accumulation with no_sync, clip, AdamW, schedule. Ranks see different
    data (seed 1234 + rank), so losses differ, but the replicas must stay bit-identical."""
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_config1_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = sorted(q.get(timeout=600) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    for rank, losses, same, wrapped, err in outs:
        assert err is None, f"rank {rank}:\n{err}"
        assert wrapped and same and len(losses) == 3 and all(np.isfinite(losses))
    assert outs[0][1] != outs[1][1]  # different shards of synthetic data per rank


def test_trainer_semantics_decay_groups_match_hf_trainer():
    """The harness's Trainer semantics put weight_decay=0.01 on exactly the parameters HF Trainer
    decays (Trainer.get_decay_parameter_names; tied aliases of shared.weight aside)."""
    from transformers import Trainer

    from distributed_llms_example_b200.harness import _no_decay
    from distributed_llms_example_b200.synthetic import build_model

    m = build_model("t5-small")
    hf = set(Trainer.get_decay_parameter_names(Trainer.__new__(Trainer), m))
    own = {n for n, _ in m.named_parameters()}
    assert {n for n in own if not _no_decay(n)} == hf & own
    # exempt: the T5LayerNorm weights and the relative_attention_bias tables (name contains "bias")
    assert all("layer_norm" in n or "bias" in n for n in own - hf)


def test_synthetic_features_are_deterministic_per_rank_and_shaped_like_the_scripts():
    from distributed_llms_example_b200.synthetic import make_features

    a, b, c = make_features(3, rank=0), make_features(3, rank=0), make_features(3, rank=1)
    assert a == b and a != c  # seed 1234 + rank
    for f in a:
        assert len(f["input_ids"]) == 512 == len(f["attention_mask"]) and len(f["labels"]) == 128
        n = sum(f["attention_mask"])
        assert f["input_ids"][n - 1] == 1 and all(t == 0 for t in f["input_ids"][n:])  # eos, then pad 0
        assert -100 not in f["labels"]  # REF mode: what the scripts really feed
    m = make_features(3, rank=0, label_mode="MASK")
    assert all(8 < len(f["labels"]) <= 128 and f["labels"][-1] == 1 for f in m)  # ragged, eos-terminated


# ---- C++ DDP comm hook: attach + delegation to the Python hook, on CPU over gloo ----------------
def _cpp_hook_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from distributed_llms_example_b200.build import build_hook
        from distributed_llms_example_b200.comm import _load_cpp_hook

        build_hook()
        mod = _load_cpp_hook()
        calls = []

        def mean_hook(state, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
            calls.append(bucket.index())
            buf = bucket.buffer()
            buf.div_(world)
            return dist.all_reduce(buf, async_op=True).get_future().then(lambda f: f.value()[0])

        def run(cpp):
            torch.manual_seed(0)
            model = torch.nn.Sequential(torch.nn.Linear(32, 64), torch.nn.ReLU(), torch.nn.Linear(64, 8))
            ddp = torch.nn.parallel.DistributedDataParallel(model)
            handle = None
            if cpp:  # no bucket is ever entered in the table: every call takes the delegation path
                handle = mod.attach(ddp.reducer, mean_hook, None, 0, 0, 0, 1.0 / world, 0, 0, 0)
            opt = torch.optim.SGD(ddp.parameters(), lr=0.1)
            g = torch.Generator().manual_seed(1234 + rank)
            for _ in range(3):
                ddp(torch.randn(4, 32, generator=g)).square().mean().backward()
                opt.step(); opt.zero_grad()
            return torch.cat([p.detach().reshape(-1) for p in model.parameters()]), handle

        a, handle = run(True)
        b, _ = run(False)
        ok = bool(torch.allclose(a, b, rtol=1e-6, atol=1e-8)) and len(calls) >= 3
        ok = ok and handle.slow_calls == len(calls) and handle.fast_calls == 0
        handle.set_entry(5, 0x1000, 16, 1, 4096)  # table grows on demand
        # P2PComm.close() closes the handle: a later backward raises instead of using a freed communicator
        handle.close()
        try:
            torch.manual_seed(0)
            m2 = torch.nn.parallel.DistributedDataParallel(torch.nn.Linear(8, 8))
            h2 = mod.attach(m2.reducer, mean_hook, None, 0, 0, 0, 1.0 / world, 0, 0, 0)
            h2.close()
            m2(torch.randn(2, 8)).sum().backward()
            ok = False
        except RuntimeError as exc:
            ok = ok and "closed" in str(exc)
        q.put((rank, ok, None))
    except Exception:
        import traceback

        q.put((rank, False, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_cpp_comm_hook_attaches_and_delegates_to_python_hook():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_cpp_hook_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, ok, err in outs:
        assert err is None, f"rank {rank}:\n{err}"
        assert ok


def test_config1_full_sequence_lengths_match_committed_golden_losses():
    """BASELINE.json configs[0] at the lengths it names (512-token sources, 128-token targets): t5-small,
    world size 2 over gloo on CPU, Trainer semantics with gradient accumulation 2, 3 optimizer steps.
    The per-rank loss trajectories must equal the ones committed in tests/golden/config1_t5small.npz
    (oracle/make_golden_config1.py, stock torch ops in the build container) to 1e-4 relative - CPU
    reductions may be blocked differently on another host - and the replicas must stay bit-identical."""
    from oracle import make_golden_config1 as G

    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "config1_t5small.npz"))
    outs = G.run(threads=2)
    for rank, losses, same, checksum, err in outs:
        assert err is None, f"rank {rank}:\n{err}"
        assert same, "replicas diverged"
        np.testing.assert_allclose(losses, gold[f"rank{rank}"], rtol=1e-4)
    assert outs[0][1] != outs[1][1]  # each rank trains on its own shard (seed 1234 + rank)


# ---- round-2 host logic that needs no GPU: lazy logits and the graph-replay front door ---------------
def test_lazy_logits_stay_lazy_for_t5s_view_and_materialise_for_anything_else():
    """fused.LazyLogits is what a patched lm_head returns: T5's `lm_logits.view(-1, lm_logits.size(-1))`
    must keep it lazy (so the rebound CrossEntropyLoss can make the fused call), every other use must see
    the real `hidden @ weight.T`."""
    import torch.nn.functional as F

    from distributed_llms_example_b200.fused import LazyLogits, lazy_parts

    torch.manual_seed(0)
    h, w = torch.randn(2, 5, 64), torch.randn(100, 64)
    z = LazyLogits(h, w)
    assert z.shape == (2, 5, 100) and z.size(-1) == 100 and z.device == h.device and z.dtype == h.dtype
    v = z.view(-1, z.size(-1))
    assert isinstance(v, LazyLogits) and v.shape == (10, 100)
    hid, wt = lazy_parts(v)
    assert hid.shape == (10, 64) and wt is w and lazy_parts(torch.zeros(3)) is None
    real = h @ w.t()
    assert torch.equal(z.argmax(-1), real.argmax(-1))                      # materialises on demand
    assert torch.allclose(z.view(10, 100, 1).squeeze(-1), real.view(10, 100))  # a view that changes the last dim too
    y = torch.randint(0, 100, (10,))
    assert torch.allclose(F.cross_entropy(v, y), F.cross_entropy(real.view(10, 100), y))


def test_graph_replay_front_door_leaves_cpu_models_and_odd_calls_on_the_eager_path():
    """patch.apply(cuda_graph=True) wraps T5ForConditionalGeneration.forward. A CPU model, an eval-mode
    call, a call without labels or with extra arguments must all run T5's own forward unchanged (same
    loss as unpatched), nothing is captured, and remove() restores both the forward and eager_mask."""
    from transformers import masking_utils as mu
    from transformers.models.t5 import modeling_t5 as t5

    from distributed_llms_example_b200 import graphs, patch
    from distributed_llms_example_b200.synthetic import build_model

    fwd0, mask0 = t5.T5ForConditionalGeneration.forward, mu.AttentionMaskInterface._global_mapping["eager"]
    m = build_model("t5-small", 0.0).train()
    b = {"input_ids": torch.randint(2, 100, (2, 16)), "attention_mask": torch.ones(2, 16, dtype=torch.long),
         "labels": torch.randint(2, 100, (2, 8))}
    b["decoder_input_ids"] = m._shift_right(b["labels"])
    ref = float(m(**b).loss.detach())
    patch.apply(cross_entropy=False, adamw=False, ddp=False, fold_clip=False, cuda_graph=True, fused_lm_head=False)
    try:
        assert t5.T5ForConditionalGeneration.forward is not fwd0
        assert mu.AttentionMaskInterface._global_mapping["eager"] is graphs.capture_safe_eager_mask
        for _ in range(4):   # more calls than B2_GRAPH_EAGER_STEPS: a CUDA model would be captured by now
            assert float(m(**b).loss.detach()) == ref   # the capture-safe mask gives the same values
        assert m.__dict__.get("_b200_graphed") is None   # CPU model: the graph front door was never entered
        gf = graphs.GraphedForward(m, fwd0)
        assert gf.signature(b) is None                                     # CPU tensors
        assert gf.signature({**b, "output_attentions": True}) is None      # extra argument
        assert gf.signature({k: v for k, v in b.items() if k != "labels"}) is None
        m.eval()
        assert gf.signature(b) is None
    finally:
        patch.remove()
    assert t5.T5ForConditionalGeneration.forward is fwd0
    assert mu.AttentionMaskInterface._global_mapping["eager"] is mask0
