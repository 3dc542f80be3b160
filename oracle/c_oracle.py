"""ctypes wrapper of oracle/oracle.c (TEST INFRASTRUCTURE ONLY; see oracle/ref_numpy.py).

build() compiles the C restatement with plain gcc into oracle/_build/ (git-ignored, travels to
the GPU box with the tree); nothing under the product package imports this module."""
from __future__ import annotations

import ctypes as C
import shutil
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
SRC = HERE / "oracle.c"
LIB = HERE / "_build" / "liboracle_c.so"
_lib = None


def build(force: bool = False) -> Path:
    if LIB.exists() and not force and LIB.stat().st_mtime >= SRC.stat().st_mtime:
        return LIB
    gcc = shutil.which("gcc")
    if gcc is None:
        raise RuntimeError("gcc not found: cannot build the C oracle")
    LIB.parent.mkdir(exist_ok=True)
    subprocess.run([gcc, "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", str(LIB), str(SRC), "-lm"],
                   check=True)
    return LIB


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = C.CDLL(str(build()))
        _lib.o_ce.restype = C.c_int64
        _lib.o_grad_norm.restype = C.c_double
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def ce(logits, labels, grad_out=1.0, ignore_index=-100):
    """Returns (loss fp32, row_lse fp32 [N], n_valid, dlogits float64 [N, V])."""
    z, y = _f32(logits), np.ascontiguousarray(labels, dtype=np.int64)
    n, v = z.shape
    loss = np.zeros(1, np.float32)
    lse = np.zeros(n, np.float32)
    dz = np.zeros((n, v), np.float64)
    nv = lib().o_ce(_ptr(z, C.c_float), _ptr(y, C.c_int64), C.c_int64(n), C.c_int64(v), C.c_int64(ignore_index),
                    C.c_double(grad_out), _ptr(loss, C.c_float), _ptr(lse, C.c_float), _ptr(dz, C.c_double))
    return loss[0], lse, int(nv), dz


def adamw_step(p, g, m, v, *, lr, beta1, beta2, eps, weight_decay, step, grad_scale=None):
    p, g, m, v = (_f32(a).copy() for a in (p, g, m, v))
    lib().o_adamw_step(_ptr(p, C.c_float), _ptr(g, C.c_float), _ptr(m, C.c_float), _ptr(v, C.c_float),
                       C.c_int64(p.size), C.c_double(lr), C.c_double(beta1), C.c_double(beta2), C.c_double(eps),
                       C.c_double(weight_decay), C.c_int64(step), C.c_double(1.0 if grad_scale is None else grad_scale))
    return p, m, v


def allreduce_mean(bufs, scale=None):
    arrs = [_f32(b) for b in bufs]
    w, n = len(arrs), arrs[0].size
    ptrs = (C.POINTER(C.c_float) * w)(*[_ptr(a, C.c_float) for a in arrs])
    out = np.zeros(n, np.float32)
    lib().o_allreduce_mean(ptrs, C.c_int(w), C.c_int64(n), C.c_float(1.0 / w if scale is None else scale),
                           _ptr(out, C.c_float))
    return out


def grad_norm(grads):
    arrs = [_f32(g).reshape(-1) for g in grads]
    ptrs = (C.POINTER(C.c_float) * len(arrs))(*[_ptr(a, C.c_float) for a in arrs])
    numels = np.array([a.size for a in arrs], dtype=np.int64)
    return float(lib().o_grad_norm(ptrs, _ptr(numels, C.c_int64), C.c_int(len(arrs))))


def collate_seq2seq(features, pad_token_id=0, label_pad_token_id=-100, pad_to_multiple_of=None,
                    decoder_start_token_id=0):
    def rounded(k):
        return k if pad_to_multiple_of is None else (k + pad_to_multiple_of - 1) // pad_to_multiple_of * pad_to_multiple_of

    b = len(features)
    ls = rounded(max(len(f["input_ids"]) for f in features))
    lt = rounded(max(len(f["labels"]) for f in features))
    ids = np.array([t for f in features for t in f["input_ids"]], dtype=np.int64)
    mask = np.array([t for f in features for t in f["attention_mask"]], dtype=np.int64)
    lab = np.array([t for f in features for t in f["labels"]], dtype=np.int64)
    id_off = np.cumsum([0] + [len(f["input_ids"]) for f in features]).astype(np.int64)
    lab_off = np.cumsum([0] + [len(f["labels"]) for f in features]).astype(np.int64)
    out = {k: np.zeros((b, ls if k in ("input_ids", "attention_mask") else lt), np.int64)
           for k in ("input_ids", "attention_mask", "labels", "decoder_input_ids")}
    i64 = C.c_int64
    lib().o_collate(_ptr(ids, i64), _ptr(mask, i64), _ptr(id_off, i64), _ptr(lab, i64), _ptr(lab_off, i64),
                    i64(b), i64(ls), i64(lt), i64(pad_token_id), i64(label_pad_token_id), i64(decoder_start_token_id),
                    _ptr(out["input_ids"], i64), _ptr(out["attention_mask"], i64), _ptr(out["labels"], i64),
                    _ptr(out["decoder_input_ids"], i64))
    return out
