"""b2_ce_forward_noalloc: the cross-entropy forward for embedders that cannot allocate.

Who needs it: a host runtime that claims all device memory for its own pool when it starts
(any later cudaMalloc fails or is against its rules), or a thin binding that has no object in
which to keep a workspace alive from one call to the next. For those callers the scratch is
a fixed array inside libb2ddp.so, and this entry point takes inputs and outputs only.

What is checked, through the C ABI: on two calls in a row (different shapes, so the second
call finds the first call's leftovers in the shared array) the loss, the per-row
log-sum-exp and the valid-row count equal those of b2_ce_forward with a caller workspace,
and those of the CPU oracle (oracle/ref_numpy.py ce_forward) on the same values, within the
1e-5 relative tolerance the other cross-entropy tests use; a request above the stated row
limit is refused with B2_ERR_TOO_LARGE. The file is named so that pytest collects
it last; it also passes when run alone in a fresh process."""
import numpy as np
import pytest
import torch

from oracle import ref_numpy as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _run_both(L, lib, z, y):
    n, v = z.shape
    dt = L.B2_BF16 if z.dtype == torch.bfloat16 else L.B2_F32
    stream = torch.cuda.current_stream().cuda_stream
    # outputs are allocated by the CALLER (here: the test); the entry point itself allocates nothing
    ref = [torch.empty(n, device=DEV), torch.empty(1, device=DEV), torch.empty(1, device=DEV)]
    got = [torch.full((n,), -1.0, device=DEV), torch.full((1,), -1.0, device=DEV), torch.full((1,), -1.0, device=DEV)]
    nbytes = lib.b2_ce_workspace_bytes(n, v)
    ws = torch.zeros(nbytes, dtype=torch.uint8, device=DEV)
    L.check(lib.b2_ce_forward(z.data_ptr(), y.data_ptr(), n, v, z.stride(0), -100, dt, ref[0].data_ptr(),
                              ref[1].data_ptr(), ref[2].data_ptr(), ws.data_ptr(), nbytes, stream))
    L.check(lib.b2_ce_forward_noalloc(z.data_ptr(), y.data_ptr(), n, v, z.stride(0), -100, dt, got[0].data_ptr(),
                                      got[1].data_ptr(), got[2].data_ptr(), stream))
    torch.cuda.synchronize()
    return [t.cpu().numpy() for t in ref], [t.cpu().numpy() for t in got]


def test_noalloc_forward_matches_workspace_forward_twice_in_a_row():
    from distributed_llms_example_b200 import _lib as L

    lib = L.load()
    torch.manual_seed(7)
    # call 1: few rows (each row split over several CTAs, partials live in the scratch);
    # call 2: many rows, other dtype (one CTA per row, row losses live in the scratch)
    cases = [(96, 32128, torch.bfloat16), (1500, 32128, torch.float32)]
    for n, v, tdt in cases:
        z = (torch.randn(n, v, device=DEV) * 4).to(tdt)
        y = torch.randint(0, v, (n,), device=DEV)
        y[::4] = -100
        ref, got = _run_both(L, lib, z, y)
        np.testing.assert_allclose(got[0], ref[0], rtol=1e-5)   # row_lse
        np.testing.assert_allclose(got[1], ref[1], rtol=1e-5)   # loss
        assert got[2][0] == ref[2][0] == float((y != -100).sum())  # n_valid
        assert np.isfinite(got[1][0])
        # ... and it stands on its own against the oracle, not only against its sibling entry point
        o_loss, o_lse, o_nvalid = O.ce_forward(z.float().cpu().numpy(), y.cpu().numpy())
        np.testing.assert_allclose(got[0], o_lse, rtol=1e-5)
        np.testing.assert_allclose(got[1][0], o_loss, rtol=1e-5)
        assert got[2][0] == o_nvalid


def test_noalloc_rejects_more_rows_than_its_scratch_holds():
    from distributed_llms_example_b200 import _lib as L

    lib = L.load()
    n = L.B2_CE_NOALLOC_MAX_ROWS + 1
    out = torch.empty(3, device=DEV)
    dummy = torch.empty(16, device=DEV)
    rc = lib.b2_ce_forward_noalloc(dummy.data_ptr(), dummy.data_ptr(), n, 32128, 32128, -100, L.B2_F32,
                                   dummy.data_ptr(), out.data_ptr(), out.data_ptr() + 4, None)
    assert rc == -5  # B2_ERR_TOO_LARGE, nothing launched
    assert b"exceed" in lib.b2_last_error()
    torch.cuda.synchronize()
