"""ctypes binding of libb2ddp.so (include/b2ddp.h). There is no CPU or torch fallback:
if the library cannot be loaded, importing a product module raises."""
from __future__ import annotations

import ctypes as C
import threading
from pathlib import Path

from .build import LIB_PATH, build

B2_F32, B2_BF16 = 0, 1
B2_ALGO_AUTO, B2_ALGO_ONE_SHOT, B2_ALGO_TWO_SHOT, B2_ALGO_TWO_SHOT_INPLACE, B2_ALGO_PUSH_INPLACE = 0, 1, 2, 3, 4
B2_IPC_HANDLE_BYTES = 64
B2_MAX_WORLD = 8
B2_ADAMW_MAX_HYPERS = 16
B2_ADAMW_CHUNK = 8192
B2_CE_NOALLOC_MAX_ROWS = 8192
ABI_VERSION = 1


class B2Error(RuntimeError):
    """A libb2ddp call returned a non-zero status."""

    def __init__(self, status: int, message: str):
        super().__init__(f"libb2ddp error {status}: {message}")
        self.status = status


class AdamWHyper(C.Structure):
    _fields_ = [("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double),
                ("eps", C.c_double), ("weight_decay", C.c_double), ("step", C.c_int64)]


class AdamWTensor(C.Structure):
    _fields_ = [("p", C.c_void_p), ("g", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p),
                ("numel", C.c_int64), ("hyper", C.c_int32), ("chunk_begin", C.c_int32)]


_vp, _i64, _i32, _f32, _sz = C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_size_t

# name -> (restype, argtypes); exactly the declarations of include/b2ddp.h
SIGNATURES = {
    "b2_last_error": (C.c_char_p, []),
    "b2_abi_version": (C.c_int, []),
    "b2_ce_workspace_bytes": (_sz, [_i64, _i64]),
    "b2_ce_forward": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _vp, _sz, _vp]),
    "b2_ce_forward_noalloc": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _vp]),
    "b2_ce_backward": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i32, _vp, _vp]),
    "b2_lmce_workspace_bytes": (_sz, [_i64, _i64]),
    "b2_lmce_forward": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _vp, _i64, _vp, _sz,
                                  _vp]),
    "b2_lmce_backward_dlogits": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _i32, _vp, _i64,
                                           _vp]),
    "b2_adamw_plan": (_i64, [C.POINTER(AdamWTensor), C.c_int32]),
    "b2_adamw_step": (C.c_int, [_vp, C.c_int32, _i64, C.POINTER(AdamWHyper), C.c_int32, _vp, _vp]),
    "b2_adamw_step_flat": (C.c_int, [_vp, _vp, _vp, _vp, _i64, C.POINTER(AdamWHyper), _vp, _vp]),
    "b2_grad_norm_workspace_bytes": (_sz, []),
    "b2_grad_norm_clip": (C.c_int, [_vp, C.c_int32, _i64, _f32, _vp, _sz, _vp, _vp, _vp]),
    "b2_comm_create": (C.c_int, [_i32, _i32, _sz, C.POINTER(_vp)]),
    "b2_comm_get_handle": (C.c_int, [_vp, _vp]),
    "b2_comm_open_peers": (C.c_int, [_vp, _vp]),
    "b2_comm_max_numel": (_i64, [_vp, _i32]),
    "b2_allreduce": (C.c_int, [_vp, _vp, _i64, _i32, _f32, _i32, _vp]),
    "b2_comm_export": (C.c_int, [_vp, _vp, _vp, C.POINTER(C.c_uint64)]),
    "b2_comm_register": (C.c_int, [_vp, _vp, _sz, _vp, C.POINTER(C.c_uint64), C.POINTER(C.c_int)]),
    "b2_comm_unregister": (C.c_int, [_vp, _i32]),
    "b2_allreduce_registered": (C.c_int, [_vp, _i32, _i64, _f32, _vp]),
    "b2_nvls_supported": (C.c_int, [C.POINTER(C.c_int)]),
    "b2_nvls_create": (C.c_int, [_vp, _sz, C.POINTER(C.c_int), C.POINTER(_sz)]),
    "b2_nvls_import": (C.c_int, [_vp, C.c_int, _sz]),
    "b2_nvls_add_device": (C.c_int, [_vp]),
    "b2_nvls_bind": (C.c_int, [_vp]),
    "b2_allreduce_nvls": (C.c_int, [_vp, _vp, _i64, _f32, _vp]),
    "b2_nvls_region": (C.c_int, [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "b2_nvls_pool_attach": (C.c_int, [_vp]),
    "b2_pool_alloc": (_vp, [C.c_ssize_t, C.c_int, _vp]),
    "b2_pool_free": (None, [_vp, C.c_ssize_t, C.c_int, _vp]),
    "b2_allreduce_nvls_inplace": (C.c_int, [_vp, C.c_uint64, _i64, _f32, _vp]),
    "b2_comm_status": (C.c_int, [_vp]),
    "b2_comm_poll": (C.c_int, [_vp]),
    "b2_comm_tune": (C.c_int, [_vp, _i32, _i32, _i32]),
    "b2_comm_status_word": (_vp, [_vp]),
    "b2_comm_sumsq_enable": (C.c_int, [_vp, _i32]),
    "b2_comm_sumsq_numel": (_i64, [_vp]),
    "b2_comm_sumsq_clip": (C.c_int, [_vp, _f32, _vp, _vp, _vp]),
    "b2_comm_destroy": (C.c_int, [_vp]),
    "b2_allreduce_emulated_scratch_bytes": (_sz, [_i32, _i64]),
    "b2_allreduce_emulated": (C.c_int, [C.POINTER(_vp), _i32, _i64, _i32, _f32, _i32, _vp, _vp]),
    "b2_allreduce_emulated_sumsq": (C.c_int, [C.POINTER(_vp), _i32, _i64, _i32, _f32, _i32, _vp, _vp, _vp]),
}

_lock = threading.Lock()
_lib = None
launches = 0  # kernels launched through this binding (bench.py reports it as gpu_launches)


def lib_path() -> Path:
    return LIB_PATH


def load() -> C.CDLL:
    """Load (building first if sources changed and nvcc is present) and type the library."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        try:
            path = build()
        except Exception as exc:
            # Only a machine WITHOUT nvcc may run a prebuilt library whose sources it cannot check;
            # a compile or link error is never papered over with a stale .so.
            if "nvcc not found" not in str(exc) or not LIB_PATH.exists():
                raise ImportError(
                    f"libb2ddp.so could not be built ({exc}); this package has no CPU fallback — "
                    "run `python -m distributed_llms_example_b200.build --verbose`") from exc
            path = LIB_PATH
        lib = C.CDLL(str(path))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError here = header/library mismatch
            fn.restype, fn.argtypes = res, args
        if lib.b2_abi_version() != ABI_VERSION:
            raise ImportError(f"libb2ddp ABI {lib.b2_abi_version()} != binding {ABI_VERSION}; rebuild")
        _lib = lib
    return _lib


def check(status: int) -> None:
    if status != 0:
        raise B2Error(status, load().b2_last_error().decode(errors="replace"))


def count_launch(n: int = 1) -> None:
    global launches
    launches += n
