"""CUDA cross-entropy against the oracle and the committed torch outputs (run with -m gpu).

Tolerances: loss and row_lse 1e-5 relative against the fp32 oracle evaluated on
the same input values; dlogits 2e-5 relative for fp32 logits, one bf16 ulp (2^-8 relative)
for bf16 logits, exact zeros on ignored rows."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import ref_numpy as O
from oracle.inputs import ce_inputs
from oracle.make_golden import CE_CASES, bf16_round

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _lib():
    from distributed_llms_example_b200 import _lib as L

    return L, L.load()


def c_abi_ce(logits_t: torch.Tensor, labels_t: torch.Tensor, grad_out: float = 1.0, ignore_index: int = -100,
             do_backward: bool = True):
    """Straight through the C ABI (b2_ce_forward / b2_ce_backward) with raw device pointers."""
    L, lib = _lib()
    n, v = logits_t.shape
    dt = L.B2_BF16 if logits_t.dtype == torch.bfloat16 else L.B2_F32
    stream = torch.cuda.current_stream().cuda_stream
    lse = torch.empty(n, dtype=torch.float32, device=DEV)
    loss = torch.empty(1, dtype=torch.float32, device=DEV)
    nvalid = torch.empty(1, dtype=torch.float32, device=DEV)
    nbytes = lib.b2_ce_workspace_bytes(n, v)
    ws = torch.zeros(nbytes, dtype=torch.uint8, device=DEV)
    L.check(lib.b2_ce_forward(logits_t.data_ptr(), labels_t.data_ptr(), n, v, logits_t.stride(0) if n else v,
                              ignore_index, dt, lse.data_ptr(), loss.data_ptr(), nvalid.data_ptr(),
                              ws.data_ptr(), nbytes, stream))
    dz = None
    if do_backward:
        g = torch.tensor([grad_out], dtype=torch.float32, device=DEV)
        dz = torch.empty_strided(logits_t.shape, logits_t.stride(), dtype=logits_t.dtype, device=DEV)
        L.check(lib.b2_ce_backward(logits_t.data_ptr(), labels_t.data_ptr(), lse.data_ptr(), g.data_ptr(),
                                   nvalid.data_ptr(), n, v, logits_t.stride(0) if n else v, ignore_index, dt,
                                   dz.data_ptr(), stream))
    torch.cuda.synchronize()
    assert int(ws[:4].view(torch.int32).item()) == 0, "arrival counter must be left at zero"
    return loss.cpu().numpy()[0], lse.cpu().numpy(), nvalid.cpu().numpy()[0], dz


@pytest.mark.parametrize("name", list(CE_CASES))
def test_ce_matches_oracle_and_golden(golden, name):
    n, v, seed, frac, dtype = CE_CASES[name]
    logits, labels = ce_inputs(n, v, seed, frac)
    if dtype == "bf16":
        logits = bf16_round(logits)
    tdt = torch.bfloat16 if dtype == "bf16" else torch.float32
    zt = torch.from_numpy(logits).to(DEV, tdt)
    yt = torch.from_numpy(labels).to(DEV)
    g = golden["ce"]
    for go in (1.0, 0.0625):
        loss, lse, nvalid, dz = c_abi_ce(zt, yt, go)
        o_loss, o_lse, o_nvalid = O.ce_forward(logits, labels)
        assert nvalid == o_nvalid
        np.testing.assert_allclose(lse, o_lse, rtol=1e-5)
        np.testing.assert_allclose(lse, g[f"{name}/row_lse"], rtol=1e-5)
        if o_nvalid == 0:
            assert np.isnan(loss) and np.isnan(g[f"{name}/g1.0/loss"])
        else:
            np.testing.assert_allclose(loss, o_loss, rtol=1e-5)
            np.testing.assert_allclose(loss, g[f"{name}/g1.0/loss"], rtol=1e-5)
        dz_h = dz.float().cpu().numpy()
        o_dz = O.ce_backward(logits, labels, go)
        assert np.all(dz_h[labels == -100] == 0.0)
        if dtype == "bf16":
            tol = np.abs(o_dz) * 2.0 ** -8 + 1e-12
            assert np.all(np.abs(dz_h - o_dz) <= tol)
        else:
            np.testing.assert_allclose(dz_h, o_dz, rtol=2e-5, atol=1e-10)
        key = f"{name}/g{go}/dlogits"
        if key in g:
            ref = g[key]
            if dtype == "bf16":
                assert np.all(np.abs(dz_h - ref) <= np.abs(ref) * 2.0 ** -7 + 1e-12)
            else:
                np.testing.assert_allclose(dz_h, ref, rtol=2e-5, atol=1e-10)
        else:
            idx = g[f"{name}/g{go}/sample_idx"]
            ref = g[f"{name}/g{go}/sample_val"]
            rtol = 2.0 ** -7 if dtype == "bf16" else 2e-5
            np.testing.assert_allclose(dz_h.reshape(-1)[idx], ref, rtol=rtol, atol=1e-10)


def test_ce_module_is_a_drop_in_for_t5_loss():
    """Same call T5 makes (modeling_t5.py:1112-1117) under bf16 autocast: fp32 0-d loss, bf16 grad."""
    from distributed_llms_example_b200.loss import CrossEntropyLoss

    torch.manual_seed(0)
    b, t, v = 2, 128, 32128
    lm_logits = (torch.randn(b, t, v, device=DEV) * 4).to(torch.bfloat16).requires_grad_(True)
    labels = torch.randint(0, v, (b, t), device=DEV)
    labels[:, 100:] = -100
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss = CrossEntropyLoss(ignore_index=-100)(lm_logits.view(-1, lm_logits.size(-1)), labels.view(-1))
    assert loss.dtype == torch.float32 and loss.dim() == 0
    loss.backward()
    assert lm_logits.grad.dtype == torch.bfloat16 and lm_logits.grad.shape == lm_logits.shape
    # Oracle = the same bf16 values up-cast to fp32. torch's own autocast path
    # rounds log-probabilities to bf16 on this image and is ~1e-4 off the fp32 answer, so it
    # is only checked loosely.
    ref_in = lm_logits.detach().float().requires_grad_(True)
    ref = torch.nn.CrossEntropyLoss(ignore_index=-100)(ref_in.view(-1, v), labels.view(-1))
    ref.backward()
    torch.testing.assert_close(loss, ref, rtol=1e-5, atol=0)
    diff = (lm_logits.grad.float() - ref_in.grad).abs()
    assert bool((diff <= ref_in.grad.abs() * 2.0 ** -8 + 1e-12).all())
    with torch.autocast("cuda", dtype=torch.bfloat16):
        stock = torch.nn.CrossEntropyLoss(ignore_index=-100)(lm_logits.detach().view(-1, v), labels.view(-1))
    assert stock.dtype == torch.float32
    torch.testing.assert_close(loss, stock, rtol=2e-3, atol=0)
    assert bool((lm_logits.grad.view(-1, v)[labels.view(-1) == -100] == 0).all())


def test_ce_rejects_what_it_does_not_implement():
    from distributed_llms_example_b200.loss import CrossEntropyLoss, cross_entropy

    with pytest.raises(NotImplementedError):
        CrossEntropyLoss(label_smoothing=0.1)
    with pytest.raises(NotImplementedError):
        CrossEntropyLoss(reduction="sum")
    with pytest.raises(RuntimeError):
        cross_entropy(torch.zeros(2, 4), torch.zeros(2, dtype=torch.int64))  # CPU tensors: no fallback


def test_ce_edge_cases():
    L, lib = _lib()
    # empty batch: NaN loss like torch, nothing launched for backward
    z = torch.empty(0, 32128, device=DEV)
    y = torch.empty(0, dtype=torch.int64, device=DEV)
    loss, lse, nvalid, dz = c_abi_ce(z, y)
    assert np.isnan(loss) and nvalid == 0 and dz.numel() == 0
    # out-of-range label poisons the loss (torch would device-assert)
    z = torch.randn(4, 100, device=DEV)
    y = torch.tensor([1, 2, 100, -100], device=DEV)
    loss, *_ = c_abi_ce(z, y, do_backward=False)
    assert np.isnan(loss)
    # row pitch larger than the vocabulary (a column slice of a wider tensor), both dtypes
    for tdt in (torch.float32, torch.bfloat16):
        wide = (torch.randn(33, 1000 + 24, device=DEV) * 3).to(tdt)
        z = wide[:, :1000]
        y = torch.randint(0, 1000, (33,), device=DEV)
        y[::5] = -100
        loss, lse, nvalid, dz = c_abi_ce(z, y)
        o_loss, o_lse, o_n = O.ce_forward(z.float().cpu().numpy(), y.cpu().numpy())
        np.testing.assert_allclose(loss, o_loss, rtol=1e-5)
        np.testing.assert_allclose(lse, o_lse, rtol=1e-5)
        assert nvalid == o_n
    # -inf logits (masked vocabulary entries) do not produce NaN
    z = torch.randn(8, 4096, device=DEV)
    z[:, 2048:] = float("-inf")
    y = torch.randint(0, 2048, (8,), device=DEV)
    loss, lse, _, dz = c_abi_ce(z, y)
    ref = torch.nn.functional.cross_entropy(z, y)
    np.testing.assert_allclose(loss, ref.item(), rtol=1e-5)
    assert bool((dz[:, 2048:] == 0).all())


def test_ce_split_rows_agree_with_whole_rows_and_runs_are_deterministic():
    """Few rows -> a row is split over several CTAs; many rows -> one CTA per row. Same rows,
    same answers; and two runs give identical bits (fixed reduction order)."""
    torch.manual_seed(1)
    v = 32128
    big = (torch.randn(2048, v, device=DEV) * 4).to(torch.bfloat16)
    yb = torch.randint(0, v, (2048,), device=DEV)
    yb[::7] = -100
    loss_big, lse_big, _, dz_big = c_abi_ce(big, yb)
    small, ys = big[:16].contiguous(), yb[:16].contiguous()
    loss_s, lse_s, _, dz_s = c_abi_ce(small, ys)
    np.testing.assert_allclose(lse_s, lse_big[:16], rtol=1e-6)
    loss_big2, lse_big2, _, dz_big2 = c_abi_ce(big, yb)
    assert loss_big == loss_big2 and np.array_equal(lse_big, lse_big2) and torch.equal(dz_big, dz_big2)
    loss_s2, lse_s2, _, _ = c_abi_ce(small, ys)
    assert loss_s == loss_s2 and np.array_equal(lse_s, lse_s2)


@pytest.mark.parametrize("tdt", [torch.bfloat16, torch.float32])
def test_ce_full_size_properties(tdt):
    """BASELINE-sized batch (8192 x 32128 bf16; 4096 rows fp32): size-independent properties."""
    torch.manual_seed(2)
    n, v = (8192, 32128) if tdt == torch.bfloat16 else (4096, 32128)
    z = (torch.randn(n, v, device=DEV) * 4).to(tdt)
    y = torch.randint(0, v, (n,), device=DEV)
    y[torch.rand(n, device=DEV) < 0.2] = -100
    loss, lse, nvalid, dz = c_abi_ce(z, y, grad_out=1.0)
    valid = y != -100
    assert nvalid == int(valid.sum())
    # loss = mean over valid rows of (lse - z[y]) with torch's own logsumexp
    zf = z.float()
    ref_lse = torch.logsumexp(zf, dim=1)
    ref_loss = (ref_lse[valid] - zf[valid, y[valid]]).double().mean()
    np.testing.assert_allclose(lse, ref_lse.cpu().numpy(), rtol=1e-5)
    np.testing.assert_allclose(loss, ref_loss.item(), rtol=1e-5)
    # each valid row of (softmax - onehot)/n sums to 0; ignored rows are exact zeros
    row_sum = dz.double().sum(dim=1)
    bound = 1e-3 / nvalid if tdt == torch.bfloat16 else 1e-5 / nvalid
    assert float(row_sum.abs().max()) < bound * 10
    assert bool((dz[~valid] == 0).all())
    # linear in the upstream gradient (power-of-two scale is exact)
    _, _, _, dz2 = c_abi_ce(z, y, grad_out=0.5)
    assert torch.equal(dz2.float() * 2, dz.float())
    # the label column carries p - 1 < 0, every other entry is >= 0
    rows = torch.nonzero(valid).squeeze(1)
    assert bool((dz[rows, y[rows]].float() < 0).all())
