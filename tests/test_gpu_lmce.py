"""Fused lm_head + cross-entropy on the tensor cores (-m gpu), SURVEY.md 8f row 4.

Through the C ABI (b2_lmce_forward / b2_lmce_backward_dlogits) against
  * the numpy oracle (float64 matmul of the same bf16 inputs + ce_forward / ce_backward) and the
    committed torch golden (tests/golden/lmce.npz) on shapes ragged against the 128 x 256 x 64 tiles;
  * a plain PyTorch fp32 reference (fp32 matmul, F.cross_entropy, autograd) at the model's real
    shape: d_model 1024, vocab 32128, N in {1024, 8192}.
Tolerances (stated where used): with round_bf16 = 0 the kernel keeps its fp32 accumulator, so loss
and row_lse agree with the float64 oracle to 5e-5 absolute (fp32 accumulation over K) and dlogits
to one bf16 ulp (its storage type); with round_bf16 = 1 each logit is rounded to bf16 first, as the
unfused path does, and a logit whose sum sits on a rounding boundary may round the other way than in
the oracle - at most one bf16 ulp of one logit, on well under 0.1 % of the elements.
"""
import numpy as np
import pytest
import torch

from oracle import ref_numpy as O
from oracle.inputs import lmce_inputs
from oracle.make_golden import LMCE_CASES

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _L():
    from distributed_llms_example_b200 import _lib as L

    return L, L.load()


def run_forward(h, w, y, round_bf16, logits_out=None):
    L, lib = _L()
    n, d = h.shape
    v = w.shape[0]
    lo_ptr, lo_ld = (logits_out.data_ptr(), logits_out.stride(0)) if logits_out is not None else (None, 0)
    nbytes = lib.b2_lmce_workspace_bytes(n, v)
    ws = torch.zeros(nbytes, dtype=torch.uint8, device=DEV)
    lse = torch.full((n,), -1.0, device=DEV)
    loss = torch.full((1,), -1.0, device=DEV)
    nvalid = torch.full((1,), -1.0, device=DEV)
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(2):  # twice: the second call finds the workspace as the first one left it
        L.check(lib.b2_lmce_forward(h.data_ptr(), w.data_ptr(), y.data_ptr(), n, v, d, -100, int(round_bf16),
                                    lse.data_ptr(), loss.data_ptr(), nvalid.data_ptr(), lo_ptr, lo_ld, ws.data_ptr(),
                                    nbytes, s))
    torch.cuda.synchronize()
    return loss, lse, nvalid


def run_backward(h, w, y, lse, nvalid, round_bf16, grad_out=1.0):
    L, lib = _L()
    n, d = h.shape
    v = w.shape[0]
    ld = (v + 7) // 8 * 8
    dz = torch.full((n, ld), 7.0, dtype=torch.bfloat16, device=DEV)
    g = torch.tensor([grad_out], dtype=torch.float32, device=DEV)
    L.check(lib.b2_lmce_backward_dlogits(h.data_ptr(), w.data_ptr(), y.data_ptr(), lse.data_ptr(), g.data_ptr(),
                                         nvalid.data_ptr(), n, v, d, -100, int(round_bf16), dz.data_ptr(), ld,
                                         torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return dz[:, :v]


def to_dev(hidden, weight, labels):
    return (torch.from_numpy(hidden).to(DEV, torch.bfloat16), torch.from_numpy(weight).to(DEV, torch.bfloat16),
            torch.from_numpy(labels).to(DEV))


@pytest.mark.parametrize("name", list(LMCE_CASES))
def test_forward_and_dlogits_vs_oracle_exact_accumulator(name):
    n, v, d, seed, frac = LMCE_CASES[name]
    hidden, weight, labels = lmce_inputs(n, v, d, seed, frac)
    h, w, y = to_dev(hidden, weight, labels)
    loss, lse, nvalid = run_forward(h, w, y, round_bf16=False)
    o_loss, o_lse, o_nvalid = O.lm_head_ce_forward(hidden, weight, labels, round_bf16=False)
    assert float(nvalid) == o_nvalid
    np.testing.assert_allclose(lse.cpu().numpy(), o_lse, atol=5e-5)
    if frac >= 1.0:
        assert np.isnan(float(loss)) and np.isnan(o_loss)
    else:
        np.testing.assert_allclose(float(loss), o_loss, rtol=1e-5)
    for g in (1.0, 0.0625):
        dz = run_backward(h, w, y, lse, nvalid, False, g).float().cpu().numpy()
        if frac >= 1.0:
            assert not dz.any()
            continue
        ref = O.lm_head_ce_backward(hidden, weight, labels, grad_out=g, round_bf16=False)[0]
        assert np.all(np.abs(dz - ref) <= np.abs(ref) * 2.0 ** -8 + 1e-10)   # one bf16 ulp
        assert not dz[labels == -100].any()                                   # ignored rows: exact zeros


@pytest.mark.parametrize("name", list(LMCE_CASES))
def test_forward_and_dlogits_vs_torch_golden_bf16_logits(golden, name):
    """round_bf16 = 1: the numerics of the unfused path (bf16 logits tensor), against torch's outputs."""
    n, v, d, seed, frac = LMCE_CASES[name]
    hidden, weight, labels = lmce_inputs(n, v, d, seed, frac)
    h, w, y = to_dev(hidden, weight, labels)
    g = golden["lmce"]
    loss, lse, nvalid = run_forward(h, w, y, round_bf16=True)
    err = np.abs(lse.cpu().numpy() - g[f"{name}/row_lse"])
    assert err.max() <= 4e-2 and np.median(err) <= 2e-6
    if frac >= 1.0:
        assert np.isnan(float(loss))
        return
    np.testing.assert_allclose(float(loss), float(g[f"{name}/loss"]), rtol=2e-5)
    dz = run_backward(h, w, y, lse, nvalid, True).float().cpu().numpy()
    ref = g[f"{name}/dlogits"]
    bad = np.abs(dz - ref) > np.abs(ref) * 2.0 ** -8 + 1e-10
    assert bad.mean() < 1e-3                                            # rounding-boundary logits only
    assert np.all(np.abs(dz - ref) <= np.abs(ref) * 0.07 + 1e-9)        # and those by one ulp of the logit


@pytest.mark.parametrize("n,v,d", [(1, 32, 64), (127, 257, 64), (129, 8, 128), (300, 32128, 64), (256, 512, 2048)])
def test_ragged_shapes_against_oracle(n, v, d):
    hidden, weight, labels = lmce_inputs(n, v, d, 900 + n, 0.2)
    h, w, y = to_dev(hidden, weight, labels)
    # the store path on shapes ragged against its [32 x 64] boxes: exact bf16 rounding of the oracle's
    # logits up to rounding-boundary cases, nothing written outside [n, v]
    ld = (v + 7) // 8 * 8
    stored = torch.full((n + 3, ld), 7.0, dtype=torch.bfloat16, device=DEV)
    run_forward(h, w, y, round_bf16=True, logits_out=stored)
    want = O.lm_head_logits(hidden, weight, round_bf16=True)
    got = stored[:n, :v].float().cpu().numpy()
    # (absolute term: fp32 accumulation over K against the float64 oracle, visible only on logits near 0)
    assert (got != want).mean() < 2e-3 and np.all(np.abs(got - want) <= np.abs(want) * 2.0 ** -7 + 1e-4)
    # rows past n_rows are never touched; the < 8 padding columns up to the next multiple of 8 get zeros
    # (stores move whole 16-byte units), nothing beyond them is written
    assert bool(torch.all(stored[n:] == 7.0)) and bool(torch.all(stored[:n, v:] == 0.0))
    loss, lse, nvalid = run_forward(h, w, y, round_bf16=False)
    o_loss, o_lse, o_nvalid = O.lm_head_ce_forward(hidden, weight, labels, round_bf16=False)
    assert float(nvalid) == o_nvalid
    np.testing.assert_allclose(lse.cpu().numpy(), o_lse, atol=1e-4)
    np.testing.assert_allclose(float(loss), o_loss, rtol=2e-5)
    dz = run_backward(h, w, y, lse, nvalid, False).float().cpu().numpy()
    ref = O.lm_head_ce_backward(hidden, weight, labels, round_bf16=False)[0]
    assert np.all(np.abs(dz - ref) <= np.abs(ref) * 2.0 ** -8 + 1e-10)


def test_out_of_range_label_poisons_the_loss():
    hidden, weight, labels = lmce_inputs(64, 300, 64, 5, 0.0)
    labels[3] = 300
    loss, _, nvalid = run_forward(*to_dev(hidden, weight, labels), round_bf16=True)
    assert np.isnan(float(loss)) and float(nvalid) == 64


def test_rejects_bad_arguments():
    L, lib = _L()
    x = torch.zeros(1024, device=DEV)
    rc = lib.b2_lmce_forward(x.data_ptr(), x.data_ptr(), x.data_ptr(), 8, 8, 100, -100, 1, x.data_ptr(), x.data_ptr(),
                             x.data_ptr(), None, 0, x.data_ptr(), 1 << 20, None)
    assert rc == -1 and b"multiple of 64" in lib.b2_last_error()


@pytest.mark.parametrize("n", [1024, 8192])
def test_model_shape_vs_fp32_torch_reference(n):
    """d_model 1024, vocab 32128 (flan-t5-large), N = batch 8 / 64 x 128 target tokens. Reference: fp32
    matmul + F.cross_entropy + autograd in plain PyTorch on the same bf16-valued inputs; a sample of rows
    also against the float64 numpy oracle."""
    import torch.nn.functional as F

    v, d = 32128, 1024
    hidden, weight, labels = lmce_inputs(n, v, d, 77 + n, 0.2)
    h, w, y = to_dev(hidden, weight, labels)
    loss, lse, nvalid = run_forward(h, w, y, round_bf16=False)
    z = (h.float() @ w.float().t()).requires_grad_(True)
    ref_loss = F.cross_entropy(z, y, ignore_index=-100)
    ref_lse = torch.logsumexp(z.detach(), dim=1)
    assert float(nvalid) == float((y != -100).sum())
    np.testing.assert_allclose(float(loss), float(ref_loss.detach()), rtol=1e-5)
    np.testing.assert_allclose(lse.cpu().numpy(), ref_lse.cpu().numpy(), atol=1e-4)
    rows = np.random.RandomState(n).choice(n, size=48, replace=False)
    _, o_lse, _ = O.lm_head_ce_forward(hidden[rows], weight, labels[rows], round_bf16=False)
    np.testing.assert_allclose(lse.cpu().numpy()[rows], o_lse, atol=5e-5)
    ref_loss.backward()
    dz = run_backward(h, w, y, lse, nvalid, False).float()
    ref = z.grad
    assert bool(torch.all((dz - ref).abs() <= ref.abs() * 2.0 ** -8 + 1e-10))
    assert not bool(dz[y == -100].any())
    # the bf16-logits mode agrees with the exact mode to the unfused path's own rounding
    loss_r, lse_r, _ = run_forward(h, w, y, round_bf16=True)
    assert abs(float(loss_r) - float(loss)) <= 2e-4 * float(loss)
    assert float((lse_r - lse).abs().max()) <= 7e-2   # half a bf16 ulp of a logit in [16, 32) is 0.0625
    # forward + store: the logits written on the way out of tensor memory are the bf16 rounding of the
    # fp32 reference (a sum on a rounding boundary may go either way: one bf16 ulp, rarely), the padding
    # columns stay untouched, and loss / lse are those of the non-storing call, bit for bit
    ld = (v + 7) // 8 * 8 + 8   # one 16-byte unit of slack per row: must stay untouched
    stored = torch.full((n, ld), 7.0, dtype=torch.bfloat16, device=DEV)
    loss_s, lse_s, _ = run_forward(h, w, y, round_bf16=True, logits_out=stored)
    assert torch.equal(loss_s, loss_r) and torch.equal(lse_s, lse_r)
    want = z.detach().to(torch.bfloat16)
    diff = (stored[:, :v].float() - want.float()).abs()
    assert float((diff > 0).float().mean()) < 1e-3 and bool(torch.all(diff <= want.float().abs() * 2.0 ** -7 + 1e-4))
    assert bool(torch.all(stored[:, v:] == 7.0))


def test_autograd_function_matches_unfused_torch_path():
    """fused_lm_head_ce(hidden, weight, labels) against F.linear (bf16) + F.cross_entropy + autograd: the
    gradients are bf16 GEMM outputs in both paths (one bf16 ulp of the largest entry as the bar)."""
    import torch.nn.functional as F

    from distributed_llms_example_b200.fused import fused_lm_head_ce

    torch.manual_seed(0)
    n, v, d = 1024, 32128, 1024
    hid = (torch.randn(8, 128, d, device=DEV)).to(torch.bfloat16)
    wt = (torch.randn(v, d, device=DEV) * (4.0 / d ** 0.5))
    y = torch.randint(0, v, (8, 128), device=DEV)
    y[:, 100:] = -100
    h2, w2 = hid.clone().requires_grad_(True), wt.clone().requires_grad_(True)
    logits = F.linear(h2, w2.to(torch.bfloat16))
    ref = F.cross_entropy(logits.float().view(-1, v), y.view(-1), ignore_index=-100)
    (ref * 0.5).backward()
    for keep in (True, False):   # backward from the stored logits / by recomputing them
        h1, w1 = hid.clone().requires_grad_(True), wt.clone().requires_grad_(True)
        loss = fused_lm_head_ce(h1, w1, y, keep_logits=keep)
        (loss * 0.5).backward()
        assert loss.dtype == torch.float32 and h1.grad.dtype == torch.bfloat16 and w1.grad.dtype == torch.float32
        np.testing.assert_allclose(float(loss.detach()), float(ref.detach()), rtol=2e-5)
        for a, b in ((h1.grad.float(), h2.grad.float()), (w1.grad, w2.grad)):
            assert float((a - b).abs().max()) <= 2.0 ** -7 * float(b.abs().max()), keep


def test_lazy_lm_head_attaches_to_an_unmodified_t5():
    """patch.apply(fused_lm_head=True): T5ForConditionalGeneration.forward runs as written, its lm_head
    answers with lazy logits, the rebound CrossEntropyLoss makes the fused call; loss and gradients
    equal the unfused plug-in path's, and outputs.logits still materialises on demand."""
    from distributed_llms_example_b200 import patch
    from distributed_llms_example_b200.fused import LazyLogits
    from distributed_llms_example_b200.synthetic import build_model

    batch = {"input_ids": torch.randint(2, 32000, (2, 64), device=DEV),
             "attention_mask": torch.ones(2, 64, dtype=torch.long, device=DEV),
             "labels": torch.randint(2, 32000, (2, 24), device=DEV)}
    batch["labels"][:, 20:] = -100
    results = {}
    try:
        for fused in (False, True):
            patch.remove()
            patch.apply(ddp=False, fused_lm_head=fused)
            model = build_model("t5-small", dropout_rate=0.0).to(DEV).train()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = model(**batch)
            out.loss.backward()
            results[fused] = (float(out.loss), {k: p.grad.clone() for k, p in model.named_parameters()}, out.logits)
    finally:
        patch.remove()
    assert isinstance(results[True][2], LazyLogits) and type(results[False][2]) is torch.Tensor
    assert torch.equal(results[True][2].argmax(-1), results[False][2].argmax(-1))   # materialises on demand
    assert abs(results[True][0] - results[False][0]) <= 2e-5 * abs(results[False][0])
    for k, gref in results[False][1].items():
        g = results[True][1][k]
        assert float((g - gref).abs().max()) <= 2e-2 * float(gref.abs().max()) + 1e-8, k
