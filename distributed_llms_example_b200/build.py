"""Builds libb2ddp.so in-tree with nvcc for sm_100a (B200) only.

Plain ``nvcc -shared`` (no torch extension machinery): the library has a C ABI
(include/b2ddp.h) and links nothing from libtorch, so it can be bound from any host
language. The ``.so`` is git-ignored but travels with the tree to the GPU box.

    python -m distributed_llms_example_b200.build [--force] [--verbose]
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
INCLUDE = PKG_DIR.parent / "include"
LIB_PATH = PKG_DIR / "libb2ddp.so"
_STAMP = PKG_DIR / ".libb2ddp.stamp"

SOURCES = ["error.cu", "adamw.cu", "cross_entropy.cu", "lmhead_ce.cu", "gradnorm.cu", "allreduce.cu", "nvls.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-cudart", "static",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found: libb2ddp.so cannot be built (there is no CPU fallback)")


def _sources() -> list[Path]:
    return [CSRC / s for s in SOURCES if (CSRC / s).exists()]


def _fingerprint() -> str:
    h = hashlib.sha256()
    files = sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cuh")) + sorted(INCLUDE.glob("*.h"))
    for f in files:
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


class _BuildLock:
    """Inter-process lock (flock on a file next to the library): under torchrun every rank imports
    the package at once, and only one of them may run nvcc into build/ and replace the library."""

    def __init__(self, path: Path):
        self.path, self.fd = path, None

    def __enter__(self):
        import fcntl

        self.fd = os.open(str(self.path), os.O_CREAT | os.O_RDWR, 0o644)
        fcntl.flock(self.fd, fcntl.LOCK_EX)
        return self

    def __exit__(self, *exc):
        import fcntl

        fcntl.flock(self.fd, fcntl.LOCK_UN)
        os.close(self.fd)


def _up_to_date(fp: str) -> bool:
    return LIB_PATH.exists() and _STAMP.exists() and _STAMP.read_text() == fp


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every csrc/*.cu and link libb2ddp.so; no-op when sources are unchanged. Safe to call
    from several processes at once: one compiles (into a temporary name, renamed atomically), the
    others wait on the lock and find the finished library."""
    fp = _fingerprint()
    if not force and _up_to_date(fp):
        return LIB_PATH
    with _BuildLock(PKG_DIR / ".build.lock"):
        if not force and _up_to_date(fp):  # another process built it while this one waited
            return LIB_PATH
        return _build_locked(fp, verbose)


def _build_locked(fp: str, verbose: bool) -> Path:
    nvcc = _nvcc()
    obj_dir = PKG_DIR / "build"
    obj_dir.mkdir(exist_ok=True)
    objs = []
    procs = []
    for src in _sources():
        obj = obj_dir / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd[1:1] = ["-Xptxas", "-v"]
            print(" ".join(cmd), file=sys.stderr)
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for src, p in procs:
        out, _ = p.communicate()
        if verbose and out:
            print(out, file=sys.stderr)
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}:\n{out}")
    tmp = LIB_PATH.with_suffix(f".so.tmp{os.getpid()}")
    link = [nvcc, "-shared", "-cudart", "static", "-o", str(tmp), *map(str, objs)]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}")
    if _STAMP.exists():
        _STAMP.unlink()  # never a valid stamp next to a library it does not describe
    os.replace(tmp, LIB_PATH)  # atomic: a process that CDLLs concurrently sees the old or the new file, whole
    _STAMP.write_text(fp)
    return LIB_PATH


HOOK_SRC = PKG_DIR / "csrc_torch" / "ddp_hook.cpp"
HOOK_LIB = PKG_DIR / "_b2ddp_hook.so"


def build_hook(force: bool = False, verbose: bool = False) -> Path:
    """Compile the C++ DDP comm hook (the only source that includes torch headers) in-tree with
    torch.utils.cpp_extension; no-op when the source is unchanged."""
    stamp = PKG_DIR / ".b2ddp_hook.stamp"
    fp = hashlib.sha256(HOOK_SRC.read_bytes()).hexdigest()
    try:
        import torch

        fp += torch.__version__
    except Exception as exc:  # pragma: no cover
        raise RuntimeError(f"torch is required to build the DDP hook: {exc}")
    if not force and HOOK_LIB.exists() and stamp.exists() and stamp.read_text() == fp:
        return HOOK_LIB
    with _BuildLock(PKG_DIR / ".build.lock"):
        if not force and HOOK_LIB.exists() and stamp.exists() and stamp.read_text() == fp:
            return HOOK_LIB
        import torch.utils.cpp_extension as ce

        bdir = PKG_DIR / "build" / "hook"
        bdir.mkdir(parents=True, exist_ok=True)
        ce.load(name="_b2ddp_hook", sources=[str(HOOK_SRC)], extra_cflags=["-O2", "-std=c++17"], with_cuda=True,
                build_directory=str(bdir), verbose=verbose, is_python_module=False)
        tmp = HOOK_LIB.with_suffix(f".so.tmp{os.getpid()}")
        shutil.copy2(bdir / "_b2ddp_hook.so", tmp)
        os.replace(tmp, HOOK_LIB)
        stamp.write_text(fp)
    return HOOK_LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
    print(build_hook(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
