"""The C-ABI boundary: libb2ddp.so loads without a GPU and exports every function that
include/b2ddp.h declares (no compute calls here)."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "b2ddp.h"


def declared_functions():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_three_plugins():
    names = declared_functions()
    for required in ("b2_ce_forward", "b2_ce_backward", "b2_adamw_step", "b2_allreduce", "b2_comm_create"):
        assert required in names


def test_library_exports_every_declared_symbol():
    from distributed_llms_example_b200 import _lib

    lib = _lib.load()
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} declared in b2ddp.h but not exported"
    assert lib.b2_abi_version() == _lib.ABI_VERSION
    assert lib.b2_last_error() == b""


def test_binding_covers_every_declared_symbol():
    from distributed_llms_example_b200 import _lib

    assert sorted(_lib.SIGNATURES) == declared_functions()


def test_struct_layouts_match_header():
    from distributed_llms_example_b200 import _lib

    assert ctypes.sizeof(_lib.AdamWTensor) == 48
    assert ctypes.sizeof(_lib.AdamWHyper) == 48
    assert _lib.AdamWTensor.chunk_begin.offset == 44


def test_host_side_plan_and_argument_errors():
    """Calls that never touch the GPU: the planner and argument validation."""
    import numpy as np

    from distributed_llms_example_b200 import _lib
    from distributed_llms_example_b200.optim import _TABLE_DTYPE

    lib = _lib.load()
    t = np.zeros(3, dtype=_TABLE_DTYPE)
    t["numel"] = [1, 8192, 8193]
    total = lib.b2_adamw_plan(t.ctypes.data_as(ctypes.POINTER(_lib.AdamWTensor)), 3)
    assert total == 4 and list(t["chunk_begin"]) == [0, 1, 2]
    t["numel"][1] = 0
    assert lib.b2_adamw_plan(t.ctypes.data_as(ctypes.POINTER(_lib.AdamWTensor)), 3) == -1
    assert b"numel" in lib.b2_last_error()
    with pytest.raises(_lib.B2Error):
        _lib.check(lib.b2_ce_forward(None, None, 4, 0, 0, -100, 0, None, None, None, None, 0, None))
    assert lib.b2_ce_workspace_bytes(128, 32128) >= 16 + 128 * 4
