"""CUDA AdamW against the committed torch single-tensor outputs and the oracle (-m gpu).

Tolerance: <= 2 fp32 ulp per step on exp_avg / exp_avg_sq, 1e-6 relative on the
parameters after the run. torch's own three AdamW implementations differ from each other by
fma/ordering, so bit-exactness is reported, not required. Always >= 3 steps: with the scripts'
schedule the learning rate is 0 at the first step."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import ref_numpy as O
from oracle.inputs import adamw_inputs
from oracle.make_golden import ADAMW_CASES, lr_schedule

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ulp_diff = O.ulp_diff


def _L():
    from distributed_llms_example_b200 import _lib as L

    return L, L.load()


def c_abi_adamw(ps, gs, ms, vs, hyper, grad_scale=None):
    """One b2_adamw_step over torch CUDA tensors, table built by hand like a foreign binding would."""
    from distributed_llms_example_b200.optim import _TABLE_DTYPE

    L, lib = _L()
    n = len(ps)
    host = np.zeros(n, dtype=_TABLE_DTYPE)
    for i in range(n):
        host[i] = (ps[i].data_ptr(), gs[i].data_ptr(), ms[i].data_ptr(), vs[i].data_ptr(), ps[i].numel(), 0, 0)
    total = lib.b2_adamw_plan(host.ctypes.data_as(ctypes.POINTER(L.AdamWTensor)), n)
    assert total > 0
    table = torch.from_numpy(host.view(np.uint8)).to(DEV)
    hp = (L.AdamWHyper * 1)(L.AdamWHyper(*hyper))
    L.check(lib.b2_adamw_step(table.data_ptr(), n, total, hp, 1,
                              grad_scale.data_ptr() if grad_scale is not None else None,
                              torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()


@pytest.mark.parametrize("name", list(ADAMW_CASES))
def test_adamw_c_abi_matches_golden(golden, name):
    sizes, seed, wd, steps, sched = ADAMW_CASES[name]
    params, grads = adamw_inputs(sizes, seed, steps)
    lrs = lr_schedule(sched, steps)
    ps = [torch.from_numpy(p).to(DEV) for p in params]
    ms = [torch.zeros_like(p) for p in ps]
    vs = [torch.zeros_like(p) for p in ps]
    for s in range(steps):
        gs = [torch.from_numpy(g).to(DEV) for g in grads[s]]
        c_abi_adamw(ps, gs, ms, vs, (lrs[s], 0.9, 0.999, 1e-8, wd, s + 1))
    g = golden["adamw"]
    exact = total = 0
    for i in range(len(sizes)):
        m, v, p = ms[i].cpu().numpy(), vs[i].cpu().numpy(), ps[i].cpu().numpy()
        assert ulp_diff(m, g[f"{name}/m{i}"]).max() <= 2 * steps
        assert ulp_diff(v, g[f"{name}/v{i}"]).max() <= 2 * steps
        np.testing.assert_allclose(p, g[f"{name}/p{i}"], rtol=1e-6, atol=1e-9)
        exact += int((p == g[f"{name}/p{i}"]).sum())
        total += p.size
    print(f"{name}: {exact}/{total} parameters bit-identical to torch's single-tensor CPU path")


def test_adamw_optimizer_drop_in_matches_torch_and_oracle():
    """The reference's construction (train-accelerator.py:174-205: two groups, lr 5e-5, linear
    schedule with one warm-up step) driven through LambdaLR for 5 steps."""
    from transformers import get_scheduler

    from distributed_llms_example_b200.optim import AdamW

    torch.manual_seed(0)
    shapes = [(512, 384), (384,), (1000, 33), (7,), (1,)]

    def make():
        torch.manual_seed(0)
        return [torch.nn.Parameter(torch.randn(s, device=DEV) * 0.02) for s in shapes]

    ours, theirs = make(), make()
    groups = lambda ps, wd: [{"params": ps[:3], "weight_decay": wd}, {"params": ps[3:], "weight_decay": 0.0}]
    for wd in (0.0, 0.01):
        o = AdamW(groups(ours, wd), lr=5e-5)
        t = torch.optim.AdamW(groups(theirs, wd), lr=5e-5, foreach=False, fused=False)
        so = get_scheduler("linear", o, num_warmup_steps=1, num_training_steps=20)
        st = get_scheduler("linear", t, num_warmup_steps=1, num_training_steps=20)
        for step in range(5):
            torch.manual_seed(100 + step)
            for a, b in zip(ours, theirs):
                gr = torch.randn_like(a) * 1e-2
                a.grad, b.grad = gr.clone(), gr.clone()
            if step == 2:
                ours[1].grad = None  # params without grad are skipped (adam.py:151-158)
                theirs[1].grad = None
            o.step(); t.step(); so.step(); st.step()
            o.zero_grad(); t.zero_grad()
        for a, b in zip(ours, theirs):
            torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-9)
            assert float(o.state[a]["step"]) == float(t.state[b]["step"])
            assert ulp_diff(o.state[a]["exp_avg"].cpu().numpy(), t.state[b]["exp_avg"].cpu().numpy()).max() <= 10
            assert ulp_diff(o.state[a]["exp_avg_sq"].cpu().numpy(), t.state[b]["exp_avg_sq"].cpu().numpy()).max() <= 10
        # state_dict moves both ways (Trainer would call these when checkpointing)
        sd = o.state_dict()
        t2 = torch.optim.AdamW(groups(make(), wd), lr=5e-5)
        t2.load_state_dict(sd)
        o2 = AdamW(groups(make(), wd), lr=5e-5)
        o2.load_state_dict(t.state_dict())
        assert len(o2.state) == len(t.state)


def test_adamw_first_step_with_zero_lr_only_moves_the_moments():
    from distributed_llms_example_b200.optim import AdamW

    p = torch.nn.Parameter(torch.randn(5000, device=DEV))
    before = p.detach().clone()
    opt = AdamW([p], lr=0.0, weight_decay=0.01)
    p.grad = torch.randn_like(p)
    opt.step()
    assert torch.equal(p.detach(), before)
    torch.testing.assert_close(opt.state[p]["exp_avg"], 0.1 * p.grad, rtol=1e-6, atol=0)


def test_adamw_grad_scale_unaligned_and_flat():
    """Device-side gradient scale (clip fold-in), tensors that are not 16-byte aligned, and the
    flat-partition entry point, all against the oracle."""
    L, lib = _L()
    rs = np.random.RandomState(3)
    n = 8192 * 3 + 5
    base = [torch.from_numpy((rs.standard_normal(n + 1) * s).astype(np.float32)).to(DEV) for s in (0.02, 1e-2, 1e-3, 1e-6)]
    p, g, m, v = [b[1:] for b in base]  # data_ptr is 4 bytes past a 16-byte boundary
    v.abs_()
    hp = dict(lr=3e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=7)
    ref = O.adamw_step(p.cpu().numpy(), g.cpu().numpy(), m.cpu().numpy(), v.cpu().numpy(), grad_scale=0.25, **hp)
    scale = torch.tensor([0.25], device=DEV)
    c_abi_adamw([p], [g], [m], [v], tuple(hp.values()), grad_scale=scale)
    for got, want in zip((p, m, v), ref):
        assert ulp_diff(got.cpu().numpy(), want).max() <= 2
    # flat entry point on aligned storage
    q = [torch.from_numpy((rs.standard_normal(n) * s).astype(np.float32)).to(DEV) for s in (0.02, 1e-2, 1e-3, 1e-6)]
    q[3].abs_()
    ref = O.adamw_step(*[t.cpu().numpy() for t in q], **hp)
    h = L.AdamWHyper(*hp.values())
    L.check(lib.b2_adamw_step_flat(q[0].data_ptr(), q[1].data_ptr(), q[2].data_ptr(), q[3].data_ptr(), n,
                                   ctypes.byref(h), None, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    for got, want in zip((q[0], q[2], q[3]), ref):
        assert ulp_diff(got.cpu().numpy(), want).max() <= 2


@pytest.mark.parametrize("grad_std", [1e-4, 1e-1])  # below / above max_norm = 1.0
def test_clip_fold_in_matches_torch_clip_then_step(grad_std):
    """Trainer's sequence clip_grad_norm_(1.0) -> optimizer.step(), with the clip folded into the
    AdamW launch, against torch's clip + single-tensor AdamW; also through the patched
    torch.nn.utils.clip_grad_norm_ entry."""
    from distributed_llms_example_b200 import patch
    from distributed_llms_example_b200.optim import AdamW

    shapes = [(300, 257), (4099,), (64, 64, 3), (1,)]

    def make():
        torch.manual_seed(0)
        return [torch.nn.Parameter(torch.randn(s, device=DEV) * 0.02) for s in shapes]

    ours, theirs = make(), make()
    o = AdamW(ours, lr=1e-3, weight_decay=0.01)
    t = torch.optim.AdamW(theirs, lr=1e-3, weight_decay=0.01, foreach=False, fused=False)
    patch.patch_clip_grad_norm()
    try:
        for step in range(3):
            torch.manual_seed(50 + step)
            for a, b in zip(ours, theirs):
                g = torch.randn_like(a) * grad_std
                a.grad, b.grad = g.clone(), g.clone()
            grads_before = [a.grad.clone() for a in ours]
            if step == 0:
                n_o = o.clip_grad_norm_(1.0)
            else:
                n_o = torch.nn.utils.clip_grad_norm_(ours, 1.0)  # routed to the fold-in
            n_t = torch.nn.utils.clip_grad_norm_(theirs, 1.0)   # not ours: stock path
            torch.testing.assert_close(n_o, n_t, rtol=2e-6, atol=0)
            assert all(torch.equal(a.grad, g) for a, g in zip(ours, grads_before))  # grads not rewritten
            o.step(); t.step()
            assert o.grad_scale is None
        for a, b in zip(ours, theirs):
            torch.testing.assert_close(a.detach(), b.detach(), rtol=2e-6, atol=1e-9)
            assert ulp_diff(o.state[a]["exp_avg_sq"].cpu().numpy(), t.state[b]["exp_avg_sq"].cpu().numpy()).max() <= 64
    finally:
        patch.remove()


def test_adamw_rejects_what_it_does_not_implement():
    from distributed_llms_example_b200.optim import AdamW

    p = torch.nn.Parameter(torch.zeros(4, device=DEV))
    with pytest.raises(NotImplementedError):
        AdamW([p], amsgrad=True)
    cpu = torch.nn.Parameter(torch.zeros(4))
    opt = AdamW([cpu])
    cpu.grad = torch.zeros(4)
    with pytest.raises(RuntimeError):
        opt.step()  # no CPU fallback
    L, lib = _L()
    h = L.AdamWHyper(1e-3, 0.9, 0.999, 1e-8, 0.0, 0)  # step 0 is invalid
    assert lib.b2_adamw_step_flat(p.data_ptr(), p.data_ptr(), p.data_ptr(), p.data_ptr(), 4, ctypes.byref(h), None, None) == -1


def test_adamw_full_size_flan_t5_large_layout():
    """All 557 tensors of flan-t5-large (750 M parameters, the BASELINE config) in one launch:
    compared with torch's fused CUDA AdamW on the same inputs, 3 steps."""
    from distributed_llms_example_b200.optim import AdamW
    from distributed_llms_example_b200.synthetic import build_model

    with torch.device("meta"):
        shapes = [tuple(p.shape) for p in build_model("flan-t5-large").parameters()]
    assert len(shapes) == 557 and sum(int(np.prod(s)) for s in shapes) == 750_251_008
    torch.manual_seed(0)
    ours = [torch.nn.Parameter(torch.randn(s, device=DEV) * 0.02) for s in shapes]
    theirs = [torch.nn.Parameter(p.detach().clone()) for p in ours]
    o = AdamW(ours, lr=5e-5, weight_decay=0.01)
    t = torch.optim.AdamW(theirs, lr=5e-5, weight_decay=0.01, fused=True)
    for step in range(3):
        for a, b in zip(ours, theirs):
            a.grad = torch.randn_like(a) * 1e-2
            b.grad = a.grad
        o.step(); t.step()
    # |update| ~ lr = 5e-5 per step, so 1e-9 absolute is 2e-5 of one update
    for a, b in zip(ours, theirs):
        torch.testing.assert_close(a.detach(), b.detach(), rtol=1e-6, atol=1e-9)
