"""Gradient synchronisation over NVLink peer memory: communicator, DDP comm hook, and the
``average_gradients`` drop-in of train-task.py.

Reference paths replaced
  * DDP Reducer -> default all-reduce hook -> ncclAllReduce per ~25 MiB bucket
    (SP/torch/nn/parallel/distributed.py:1255-1284, default_hooks.py:18-54), driven by
    ``accelerator.prepare`` / ``accelerator.backward`` (train-accelerator.py:189,222) and HF
    Trainer (train-torchrun.py:127,140);
  * ``average_gradients(model)`` of train-task.py:65-69: one blocking ``dist.all_reduce`` plus
    one ``/= size`` kernel per parameter tensor.

The torch.distributed process group stays the control plane (rendezvous, handle exchange,
barriers); only gradient bytes move through libb2ddp's kernels. Single node, world in
{1, 2, 4, 8}. No NCCL/gloo fallback for the data path.
"""
import ctypes as C
import os
import weakref
from typing import Optional

import torch
import torch.distributed as dist

from . import _lib

__all__ = ["P2PComm", "HookState", "p2p_allreduce_hook", "register_p2p_comm_hook",
           "average_gradients", "exchange_handles", "default_comm", "poll_all", "live_comms"]

_LIVE: "weakref.WeakSet[P2PComm]" = weakref.WeakSet()


def live_comms() -> list:
    return [c for c in list(_LIVE) if c._ctx is not None]


def poll_all() -> None:
    """Raise B2Error(B2_ERR_TIMEOUT) if any live communicator of this process has seen a peer miss
    a barrier. One host load per communicator, no synchronisation: AdamW.step() calls this every
    step so that a poisoned (NaN-filled) gradient is never applied silently."""
    for c in live_comms():
        c.poll()

_WIRE = {"fp32": _lib.B2_F32, "bf16": _lib.B2_BF16}
_ALGO = {"auto": _lib.B2_ALGO_AUTO, "one_shot": _lib.B2_ALGO_ONE_SHOT, "two_shot": _lib.B2_ALGO_TWO_SHOT}


def exchange_handles(handle: bytes, group: Optional[dist.ProcessGroup] = None) -> list[bytes]:
    """All-gather one fixed-size opaque handle per rank, in rank order, over the control-plane
    process group (works on gloo and NCCL)."""
    world = dist.get_world_size(group)
    if len(handle) != _lib.B2_IPC_HANDLE_BYTES:
        raise ValueError(f"handle must be {_lib.B2_IPC_HANDLE_BYTES} bytes, got {len(handle)}")
    gathered: list = [None] * world
    dist.all_gather_object(gathered, handle, group=group)
    if any(not isinstance(h, (bytes, bytearray)) or len(h) != len(handle) for h in gathered):
        raise RuntimeError("handle exchange returned a malformed entry")
    return [bytes(h) for h in gathered]


class P2PComm:
    """This rank's end of the symmetric staging buffers. Collective: every rank of ``group``
    must construct it (and later call ``allreduce_``) in the same order."""

    def __init__(self, group: Optional[dist.ProcessGroup] = None, staging_bytes: Optional[int] = None,
                 wire: str = "fp32", algo: str = "auto"):
        if not torch.cuda.is_available():
            raise RuntimeError("P2PComm needs CUDA devices; there is no CPU path")
        if dist.is_available() and dist.is_initialized():
            self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        else:
            self.rank, self.world = 0, 1
        if self.world not in (1, 2, 4, 8):
            raise NotImplementedError(f"P2PComm: world size {self.world} (supported: 1, 2, 4, 8 on one node)")
        if wire not in _WIRE or algo not in _ALGO:
            raise ValueError(f"wire must be one of {list(_WIRE)}, algo one of {list(_ALGO)}")
        self.group, self.wire, self.algo = group, wire, algo
        self.nvls, self.nvls_error = False, ""
        self._pool = self._pool_allocator = self._region = None
        self.nvls_pool_bytes = 0
        self.device = torch.device("cuda", torch.cuda.current_device())
        if staging_bytes is None:
            staging_bytes = int(os.environ.get("B2_AR_STAGING_MB", "64")) << 20
        self._lib = _lib.load()
        self._ctx = C.c_void_p()
        self._hook_states: "weakref.WeakSet[HookState]" = weakref.WeakSet()
        self.sumsq_enabled = False
        _lib.check(self._lib.b2_comm_create(self.rank, self.world, staging_bytes, C.byref(self._ctx)))
        self._c_poll = self._lib.b2_comm_poll
        _LIVE.add(self)
        if self.world > 1:
            raw = C.create_string_buffer(_lib.B2_IPC_HANDLE_BYTES)
            _lib.check(self._lib.b2_comm_get_handle(self._ctx, raw))
            handles = exchange_handles(raw.raw, group)
            _lib.check(self._lib.b2_comm_open_peers(self._ctx, b"".join(handles)))
            dist.barrier(group)  # every rank has mapped every peer before the first kernel

    def allreduce_(self, tensor: torch.Tensor, scale: float = 1.0, *, wire: Optional[str] = None,
                   algo: Optional[str] = None, stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        """In place: tensor <- scale * sum over ranks (fixed rank order; identical on all ranks)."""
        if self._ctx is None:
            raise RuntimeError("P2PComm is closed")
        if tensor.device != self.device:
            raise RuntimeError(f"tensor on {tensor.device}, communicator on {self.device}")
        if tensor.dtype != torch.float32:
            raise NotImplementedError(f"P2PComm.allreduce_: fp32 buckets only (got {tensor.dtype})")
        if not tensor.is_contiguous():
            raise ValueError("P2PComm.allreduce_: tensor must be contiguous")
        s = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        _lib.check(self._lib.b2_allreduce(self._ctx, tensor.data_ptr(), tensor.numel(),
                                          _WIRE[wire or self.wire], float(scale),
                                          _ALGO[algo or self.algo], s))
        cap = self._lib.b2_comm_max_numel(self._ctx, _WIRE[wire or self.wire])
        _lib.count_launch(max(1, -(-tensor.numel() // max(cap, 1))))
        return tensor

    # -- zero-copy path: buffers that live across steps, mapped by every peer ----------------
    def register(self, tensor: torch.Tensor, peers: Optional[list] = None) -> int:
        """Collective (unless ``peers`` is given): export the allocation holding ``tensor``,
        all-gather (handle, offset) and map every peer's buffer. Returns a registration id for
        ``allreduce_registered_``. The tensor's memory must stay allocated until ``unregister``."""
        if tensor.device != self.device or tensor.dtype != torch.float32 or not tensor.is_contiguous():
            raise ValueError("register: contiguous fp32 tensor on the communicator's device required")
        mine = self.export(tensor)
        if peers is None:
            peers = [None] * self.world
            if self.world > 1:
                dist.all_gather_object(peers, mine, group=self.group)
            else:
                peers = [mine]
        handles = b"".join(p[0] for p in peers)
        offsets = (C.c_uint64 * self.world)(*[p[1] for p in peers])
        reg = C.c_int(-1)
        _lib.check(self._lib.b2_comm_register(self._ctx, tensor.data_ptr(), tensor.numel() * 4, handles, offsets,
                                              C.byref(reg)))
        return reg.value

    def export(self, tensor: torch.Tensor) -> tuple:
        """(64-byte IPC handle of the allocation containing ``tensor``, byte offset inside it)."""
        raw = C.create_string_buffer(_lib.B2_IPC_HANDLE_BYTES)
        off = C.c_uint64(0)
        _lib.check(self._lib.b2_comm_export(self._ctx, tensor.data_ptr(), raw, C.byref(off)))
        return (raw.raw, off.value)

    def unregister(self, reg_id: int) -> None:
        _lib.check(self._lib.b2_comm_unregister(self._ctx, reg_id))

    def allreduce_registered_(self, reg_id: int, tensor: torch.Tensor, scale: float = 1.0, *,
                              stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        """In-place two-shot directly on the registered buffers (no staging copy)."""
        s = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        _lib.check(self._lib.b2_allreduce_registered(self._ctx, reg_id, tensor.numel(), float(scale), s))
        _lib.count_launch(2 if self.world > 1 else 1)  # gate + data kernel
        return tensor

    # -- in-switch reduction (NVLS) -------------------------------------------------------------
    def enable_nvls(self, nbytes: Optional[int] = None) -> bool:
        """Collective. Builds the multicast object and binds one buffer per rank; returns False
        (on every rank) if any rank's device or driver cannot do it, so callers stay in step."""
        if self.nvls:
            return True
        if self.world < 2:
            return False
        import socket
        import tempfile

        nbytes = nbytes or (int(os.environ.get("B2_AR_NVLS_MB", "64")) << 20)
        lib, ctx, ok = self._lib, self._ctx, True
        sup = C.c_int(0)
        ok = lib.b2_nvls_supported(C.byref(sup)) == 0 and sup.value == 1

        def agree(flag: bool) -> bool:
            votes: list = [None] * self.world
            dist.all_gather_object(votes, bool(flag), group=self.group)
            return all(votes)

        if not agree(ok):
            return False
        # 1. rank 0 creates the multicast object; its POSIX fd travels over a Unix socket
        info = [None]
        srv = None
        fd = C.c_int(-1)
        rounded = C.c_size_t(0)
        if self.rank == 0:
            ok = lib.b2_nvls_create(ctx, nbytes, C.byref(fd), C.byref(rounded)) == 0
            if ok:
                path = os.path.join(tempfile.gettempdir(), f"b2ddp-nvls-{os.getpid()}-{id(self):x}.sock")
                srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                srv.bind(path)
                srv.listen(self.world)
                info = [(path, rounded.value)]
            else:
                self.nvls_error = lib.b2_last_error().decode(errors="replace")
        dist.broadcast_object_list(info, src=dist.get_global_rank(self.group, 0) if self.group else 0,
                                   group=self.group)
        if info[0] is None:
            return False
        path, rounded_bytes = info[0]
        if self.rank == 0:
            for _ in range(self.world - 1):
                conn, _addr = srv.accept()
                socket.send_fds(conn, [b"mc"], [fd.value])
                conn.close()
            srv.close()
            os.unlink(path)
            os.close(fd.value)
        else:
            cli = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            cli.connect(path)
            _msg, fds, _flags, _addr = socket.recv_fds(cli, 16, 1)
            cli.close()
            ok = len(fds) == 1 and lib.b2_nvls_import(ctx, fds[0], rounded_bytes) == 0
            if fds:
                os.close(fds[0])
        # 2. every device joins, then every rank binds its memory and maps both views
        ok = ok and lib.b2_nvls_add_device(ctx) == 0
        if not agree(ok):
            self.nvls_error = lib.b2_last_error().decode(errors="replace")
            return False
        ok = lib.b2_nvls_bind(ctx) == 0
        if not ok:
            self.nvls_error = lib.b2_last_error().decode(errors="replace")
        if not agree(ok):
            return False
        self.nvls = True
        return True

    def nvls_pool(self) -> "torch.cuda.MemPool":
        """A torch MemPool whose memory is the multicast-bound region: tensors allocated under
        ``torch.cuda.use_mem_pool(pool)`` can be all-reduced by the switch in place."""
        if not self.nvls:
            raise RuntimeError("nvls_pool: enable_nvls() has not succeeded")
        if self._pool is None:
            _lib.check(self._lib.b2_nvls_pool_attach(self._ctx))
            base, nbytes, pool_off = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
            _lib.check(self._lib.b2_nvls_region(self._ctx, C.byref(base), C.byref(nbytes), C.byref(pool_off)))
            self._region = (base.value, nbytes.value)
            self.nvls_pool_bytes = nbytes.value - pool_off.value
            self._pool_allocator = torch.cuda.memory.CUDAPluggableAllocator(
                str(_lib.lib_path()), "b2_pool_alloc", "b2_pool_free")
            self._pool = torch.cuda.MemPool(self._pool_allocator.allocator())
        return self._pool

    def nvls_offset(self, tensor: torch.Tensor) -> Optional[int]:
        """Byte offset of ``tensor`` inside the multicast-bound region, or None if it lives elsewhere."""
        if self._region is None:
            return None
        base, nbytes = self._region
        p = tensor.data_ptr()
        if base <= p and p + tensor.numel() * tensor.element_size() <= base + nbytes:
            return p - base
        return None

    def allreduce_nvls_inplace_(self, offset: int, numel: int, scale: float = 1.0, *,
                                stream: Optional[torch.cuda.Stream] = None) -> None:
        s = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        _lib.check(self._lib.b2_allreduce_nvls_inplace(self._ctx, offset, numel, float(scale), s))
        _lib.count_launch(2)

    def allreduce_nvls_(self, tensor: torch.Tensor, scale: float = 1.0, *,
                        stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        """In place, reduced inside the NVSwitch (identical on all ranks; fp32-rounding-equal to
        the rank-ordered sum)."""
        if not self.nvls:
            raise RuntimeError("allreduce_nvls_: enable_nvls() has not succeeded")
        if tensor.device != self.device or tensor.dtype != torch.float32 or not tensor.is_contiguous():
            raise ValueError("allreduce_nvls_: contiguous fp32 tensor on the communicator's device required")
        s = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        _lib.check(self._lib.b2_allreduce_nvls(self._ctx, tensor.data_ptr(), tensor.numel(), float(scale), s))
        _lib.count_launch()
        return tensor

    def tune(self, max_blocks: int = -1, packs_per_thread: int = -1, unroll: int = -1) -> None:
        """Launch shape of the zero-copy kernels (same call on every rank; see b2_comm_tune)."""
        _lib.check(self._lib.b2_comm_tune(self._ctx, max_blocks, packs_per_thread, unroll))

    def check(self) -> None:
        """Synchronising health check: raises if a kernel timed out waiting for a peer."""
        _lib.check(self._lib.b2_comm_status(self._ctx))

    def poll(self) -> None:
        """Non-synchronising health check (one host load of the mapped status word). A rank that
        waits longer than B2_AR_TIMEOUT_MS for a peer poisons the communicator on EVERY rank: the
        buckets reduced from then on are NaN-filled and this raises B2Error(B2_ERR_TIMEOUT)."""
        if self._ctx is not None:
            rc = self._c_poll(self._ctx)
            if rc != 0:
                _lib.check(rc)

    # -- sum of squares folded into the all-reduce (gradient-clipping fold-in) --
    def enable_sumsq(self, on: bool = True) -> None:
        """From the next all-reduce on, every bucket also accumulates sum(x*x) of its reduced values
        (identical on every rank). All ranks must switch at the same point of the program."""
        _lib.check(self._lib.b2_comm_sumsq_enable(self._ctx, 1 if on else 0))
        self.sumsq_enabled = bool(on)

    def sumsq_numel(self) -> int:
        return int(self._lib.b2_comm_sumsq_numel(self._ctx))

    def sumsq_clip_(self, max_norm: float, out: torch.Tensor, *, stream: Optional[torch.cuda.Stream] = None) -> None:
        """out[0] = sqrt(total), out[1] = min(1, max_norm / (out[0] + 1e-6)); resets the total. The
        stream must be ordered after the all-reduces (it is, after DDP's finalize_backward)."""
        s = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        _lib.check(self._lib.b2_comm_sumsq_clip(self._ctx, float(max_norm), out.data_ptr(), out.data_ptr() + 4, s))
        _lib.count_launch()

    def close(self) -> None:
        """Collective. Detaches every DDP hook that still points at this communicator first (a later
        backward on such a DDP module raises instead of touching freed memory), then unmaps."""
        if self._ctx is not None:
            torch.cuda.synchronize(self.device)
            if self.world > 1 and dist.is_initialized():
                dist.barrier(self.group)  # nobody unmaps while a peer may still read
            for st in list(self._hook_states):
                st.invalidate()
            for key in [k for k, p in _rebuild_pools.items() if p is self._pool]:
                _rebuild_pools.pop(key, None)
            self._lib.b2_comm_destroy(self._ctx)
            self._ctx = None
        global _default
        if _default is self:
            _default = None

    # No __del__: destroying without the barrier of close() could unmap memory a peer is still
    # reading, and a C++ hook may still hold the raw pointer. An unclosed communicator is reclaimed
    # at process exit.


_default: Optional["P2PComm"] = None


def default_comm(**kwargs) -> P2PComm:
    """Process-wide communicator on the default process group (created collectively on first use)."""
    global _default
    if _default is None:
        _default = P2PComm(**kwargs)
    return _default


class HookState:
    """State object handed to ``ddp.register_comm_hook``: the communicator plus the side stream
    the reductions run on, so they overlap the rest of backward like the Reducer's NCCL stream."""

    def __init__(self, comm: Optional[P2PComm] = None, zero_copy: Optional[bool] = None,
                 zero_copy_min_bytes: Optional[int] = None, **comm_kwargs):
        self.comm = comm or default_comm(**comm_kwargs)
        self.stream = torch.cuda.Stream(device=self.comm.device)
        self.scale = 1.0 / self.comm.world
        self.event = torch.cuda.Event()  # reused for every bucket: producer stream -> side stream
        self._c_allreduce_registered = self.comm._lib.b2_allreduce_registered
        self.buckets_reduced = 0
        # Zero-copy: once a bucket has shown up twice at the same address on EVERY rank (the
        # Reducer's buckets are stable after its one rebuild), it is registered for peer access
        # and reduced in place. Small buckets stay on the one-barrier staging path.
        if zero_copy is None:
            zero_copy = os.environ.get("B2_AR_ZERO_COPY", "1") != "0"
        if zero_copy_min_bytes is None:
            # measured crossover against the one-barrier staging path (profiles/README.md)
            dflt = {2: 16 << 20, 4: 8 << 20}.get(self.comm.world, 2 << 20)
            zero_copy_min_bytes = int(os.environ.get("B2_AR_ZERO_COPY_MIN_BYTES", str(dflt)))
        self.zero_copy = bool(zero_copy) and self.comm.world > 1 and self.comm.wire == "fp32"
        self.zero_copy_min_bytes = zero_copy_min_bytes
        self._last_seen: dict[int, tuple] = {}    # bucket index -> (data_ptr, numel) at its last call
        self._registered: dict[int, tuple] = {}   # bucket index -> (data_ptr, numel, reg_id | None, nvls offset | None)
        self._attempts: dict[int, int] = {}       # bucket index -> failed collective votes so far
        self._never: set[int] = set()             # buckets every rank agreed to leave on the staging path
        self.max_attempts = int(os.environ.get("B2_AR_REGISTER_ATTEMPTS", "3"))
        self._c_allreduce_nvls_inplace = self.comm._lib.b2_allreduce_nvls_inplace
        self.cpp_handle = None  # set when the C++ hook serves the registered buckets
        self.closed = False
        self.comm._hook_states.add(self)

    def invalidate(self) -> None:
        """Called by P2PComm.close(): the hook must not touch the communicator any more."""
        self.closed = True
        if self.cpp_handle is not None:
            self.cpp_handle.close()
        for reg in self._registered.values():
            if reg[2] is not None and self.comm._ctx is not None:
                try:
                    self.comm.unregister(reg[2])
                except _lib.B2Error:
                    pass
        self._registered.clear()

    @property
    def total_buckets_reduced(self) -> int:
        return self.buckets_reduced + (self.cpp_handle.fast_calls if self.cpp_handle is not None else 0)

    def registration_for(self, index: int, buf: torch.Tensor) -> Optional[tuple]:
        """Collective while the bucket is unregistered: every rank contributes (stable?, handle,
        offset, nvls offset); the bucket is registered only if it is stable everywhere. Returns
        (data_ptr, numel, ipc registration id | None, nvls offset | None) or None."""
        key = (buf.data_ptr(), buf.numel())
        reg = self._registered.get(index)
        if reg is not None:
            if reg[:2] != key:
                raise RuntimeError(
                    f"DDP bucket {index} moved after it was registered for peer access "
                    f"({reg[:2]} -> {key}); rebuild the HookState (or set B2_AR_ZERO_COPY=0)")
            return reg
        if not self.zero_copy or buf.numel() * 4 < self.zero_copy_min_bytes or index in self._never:
            return None
        stable = self._last_seen.get(index) == key
        self._last_seen[index] = key
        nvls_off = self.comm.nvls_offset(buf) if hasattr(self.comm, "nvls_offset") else None
        exportable = True
        try:
            mine = (stable,) + (self.comm.export(buf) if nvls_off is None else (b"", 0)) + (nvls_off,)
        except _lib.B2Error:
            exportable = False  # memory that cannot be exported (e.g. expandable segments)
            mine = (False, b"", 0, None)
        # Every rank contributes (stable, handle, offset, nvls offset, exportable, attempts so far)
        # and draws the SAME conclusion from the gathered votes, including "give up": a bucket that
        # cannot be exported anywhere, or that has failed max_attempts votes, gets a permanent
        # negative entry, so that a run on an allocator setting that defeats the zero-copy path
        # pays for a few collectives at start-up instead of one per bucket per step for ever.
        attempts = self._attempts.get(index, 0)
        votes: list = [None] * self.comm.world
        dist.all_gather_object(votes, mine + (exportable, attempts), group=self.comm.group)

        def give_up_or_retry():
            self._attempts[index] = attempts + 1
            if (not all(v[4] for v in votes)) or max(v[5] for v in votes) + 1 >= self.max_attempts:
                self._never.add(index)
            return None

        if not all(v[0] for v in votes):
            # "not stable yet" is normal for the first sighting; it only counts as a failed attempt
            # once the bucket has been seen before everywhere
            if attempts > 0 or not all(v[4] for v in votes):
                return give_up_or_retry()
            self._attempts[index] = 1
            return None
        offs = {v[3] for v in votes}
        if len(offs) == 1 and nvls_off is not None:
            # the bucket sits at the same offset of the multicast-bound region on every rank:
            # the switch reduces it in place, nothing to map
            reg = key + (None, nvls_off)
        elif None in offs and len(offs) == 1:
            # mapping can fail on one rank only (e.g. out of address space): vote again so that all
            # ranks either use the registration or fall back together
            try:
                reg_id = self.comm.register(buf, peers=[(v[1], v[2]) for v in votes])
            except _lib.B2Error:
                reg_id = None
            oks: list = [None] * self.comm.world
            dist.all_gather_object(oks, reg_id is not None, group=self.comm.group)
            if not all(oks):
                if reg_id is not None:
                    self.comm.unregister(reg_id)
                self._never.add(index)
                return None
            reg = key + (reg_id, None)
        else:
            return give_up_or_retry()  # mixed placement across ranks: stay on the staging path
        self._registered[index] = reg
        if self.cpp_handle is not None:  # from now on the C++ hook handles this bucket by itself
            self.cpp_handle.set_entry(index, reg[0], reg[1], 1 if reg[3] is not None else 0,
                                      reg[3] if reg[3] is not None else reg[2])
        return reg


def p2p_allreduce_hook(state: HookState, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
    """DDP communication hook: mean-all-reduce ``bucket.buffer()`` in place.

    Same result contract as ``default_hooks.allreduce_hook`` (divide by world size, then sum;
    SP/torch/distributed/algorithms/ddp_comm_hooks/default_hooks.py:18-54). Called by the
    Reducer on the autograd thread once per ready bucket (~100 times per step for
    flan-t5-large, and the step is host-bound), so the body is kept to a handful of C calls:
    one reusable event orders the side stream after the bucket's producer, the registered
    bucket goes straight to the C ABI, and the side stream is made current only around
    ``set_result``, which records the completion event that ``Reducer::finalize_backward``
    waits on - no host synchronisation anywhere.
    """
    comm = state.comm
    if state.closed or comm._ctx is None:
        raise RuntimeError("b200 DDP hook: its communicator has been closed; rebuild the DDP module's hook")
    comm.poll()  # a peer that missed a barrier earlier: raise here, on every rank, not train on NaN
    buf = bucket.buffer()
    side = state.stream
    producer = torch.cuda.current_stream(comm.device)
    state.event.record(producer)
    side.wait_event(state.event)
    reg = state._registered.get(bucket.index()) if state.zero_copy else None
    if reg is None or reg[0] != buf.data_ptr() or reg[1] != buf.numel():
        reg = state.registration_for(bucket.index(), buf) if state.zero_copy else None
    if reg is None:
        comm.allreduce_(buf, state.scale, stream=side)
    else:
        if reg[3] is not None:   # bucket lives in the multicast-bound region: in-switch, in place
            rc = state._c_allreduce_nvls_inplace(comm._ctx, reg[3], reg[1], state.scale, side.cuda_stream)
        else:                    # bucket mapped by the peers over IPC: one-pass reduce + push
            rc = state._c_allreduce_registered(comm._ctx, reg[2], reg[1], state.scale, side.cuda_stream)
        if rc != 0:
            _lib.check(rc)
        _lib.launches += 2  # gate + data kernel
    fut: torch.futures.Future[torch.Tensor] = torch.futures.Future(devices=[comm.device])
    torch.cuda.set_stream(side)
    try:
        fut.set_result(buf)
    finally:
        torch.cuda.set_stream(producer)
    state.buckets_reduced += 1
    return fut


_cpp_hook_module = None


def _load_cpp_hook():
    """The in-tree C++ hook extension (csrc_torch/ddp_hook.cpp), or None if it has not been built."""
    global _cpp_hook_module
    if _cpp_hook_module is None:
        import importlib.util

        from .build import HOOK_LIB

        if not HOOK_LIB.exists():
            return None
        spec = importlib.util.spec_from_file_location("_b2ddp_hook", str(HOOK_LIB))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _cpp_hook_module = mod
    return _cpp_hook_module


def register_p2p_comm_hook(ddp_model, state: Optional[HookState] = None, **comm_kwargs) -> HookState:
    """Attach the peer-memory all-reduce to a DDP module; collective on first use.

    By default (``B2_CPP_HOOK=auto``) the Reducer gets the C++ hook of csrc_torch/ddp_hook.cpp when
    that extension has been built and loads: it serves registered buckets without touching the
    interpreter and delegates everything else to ``p2p_allreduce_hook``. ``cpp_hook=False`` /
    ``B2_CPP_HOOK=0``, or an extension that is missing or fails to load, registers the Python hook
    through ``ddp_model.register_comm_hook`` (same kernels, ~15 us more host time per bucket);
    ``cpp_hook=True`` / ``B2_CPP_HOOK=1`` makes a missing extension an error."""
    nvls_buckets = comm_kwargs.pop("nvls_buckets", None)
    cpp_hook = comm_kwargs.pop("cpp_hook", None)
    state = state or HookState(**comm_kwargs)
    required = cpp_hook is True
    if cpp_hook is None:
        env = os.environ.get("B2_CPP_HOOK", "auto")
        cpp_hook, required = env != "0", env == "1"
    mod = None
    if cpp_hook:
        try:
            mod = _load_cpp_hook()
        except Exception as exc:  # e.g. built against another torch: fall back to the Python hook
            if required:
                raise
            import warnings

            warnings.warn(f"b200: C++ DDP hook could not be loaded ({exc}); using the Python hook")
        if mod is None and required:
            raise RuntimeError("the C++ DDP hook was requested but _b2ddp_hook.so is not built "
                               "(python -m distributed_llms_example_b200.build)")
    if mod is not None:
        lib = state.comm._lib
        fn_reg = C.cast(lib.b2_allreduce_registered, C.c_void_p).value
        fn_nvls = C.cast(lib.b2_allreduce_nvls_inplace, C.c_void_p).value
        side = state.stream
        ddp_model._check_comm_hook(p2p_allreduce_hook)  # the checks DDP.register_comm_hook makes
        status_word = lib.b2_comm_status_word(state.comm._ctx) or 0
        state.cpp_handle = mod.attach(ddp_model.reducer, p2p_allreduce_hook, state, state.comm._ctx.value, fn_reg,
                                      fn_nvls, state.scale, state.comm.device.index, side.stream_id, status_word)
        # ... and its bookkeeping (distributed.py:2064-2067): the hook list and the logger's hook name
        ddp_model.logger._set_comm_hook_name(p2p_allreduce_hook.__qualname__)
        ddp_model._comm_hooks.append((p2p_allreduce_hook, state))
    else:
        ddp_model.register_comm_hook(state, p2p_allreduce_hook)
    if nvls_buckets is None:
        # measured (profiles/README.md): the switch wins at world 8 (25 MiB bucket 90.7 us vs 100.7 us
        # for the IPC push, 146.6 us for NCCL); at world 2 it has no byte advantage and loses
        # (97 us vs 64 us), world 4 is unmeasured and stays on the IPC push
        env = os.environ.get("B2_AR_NVLS_BUCKETS", "auto")
        nvls_buckets = state.comm.world >= 8 if env == "auto" else env != "0"
    if nvls_buckets and state.zero_copy:
        _place_buckets_in_nvls_region(ddp_model, state)
    return state


_rebuild_pools: dict[int, "torch.cuda.MemPool"] = {}


def _place_buckets_in_nvls_region(ddp_model, state: HookState) -> bool:
    """Make the Reducer allocate its rebuilt buckets inside the communicator's multicast-bound
    region, so that the NVSwitch can reduce them in place (no copy, no peer mapping).

    The Reducer rebuilds its buckets exactly once, from ``DDP._pre_forward`` of the second
    iteration (SP/torch/nn/parallel/distributed.py:1578-1584), through the Python-visible
    ``reducer._rebuild_buckets()``. That one call is wrapped in ``torch.cuda.use_mem_pool`` with
    a pool whose allocator is libb2ddp's b2_pool_alloc; every rank makes the same allocations in
    the same order, so bucket i lands at the same offset everywhere (verified collectively by
    ``registration_for`` before the in-place path is used). Collective; returns False on every
    rank if NVLS is unavailable, leaving the IPC zero-copy path in charge."""
    comm = state.comm
    grad_bytes = sum(p.numel() * 4 for p in ddp_model.parameters() if p.requires_grad)
    # segments are 2 MiB-aligned and torch rounds each bucket up to a 2 MiB multiple
    need = int(grad_bytes * 1.25) + (128 << 20)
    if not comm.enable_nvls(need + (64 << 20)):  # + the staging area in front of the pool
        return False
    pool = comm.nvls_pool()
    if comm.nvls_pool_bytes < need:  # region was created earlier, too small: same answer on every rank
        return False
    _rebuild_pools[id(ddp_model.reducer)] = pool
    Reducer = type(ddp_model.reducer)
    if not getattr(Reducer, "_b200_rebuild_wrapped", False):
        orig = Reducer._rebuild_buckets

        def _rebuild_buckets(self):
            p = _rebuild_pools.get(id(self))
            if p is None:
                return orig(self)
            with torch.cuda.use_mem_pool(p):
                return orig(self)

        Reducer._rebuild_buckets = _rebuild_buckets
        Reducer._b200_rebuild_wrapped = True
    return True


@torch.no_grad()
def average_gradients(model: torch.nn.Module, comm: Optional[P2PComm] = None) -> None:
    """Drop-in for train-task.py:65-69 (same signature, same in-place effect on ``p.grad``).

    The reference issues one blocking NCCL all-reduce and one division kernel per parameter
    tensor (131-557 of each per step). Here all gradients go through ONE flat exchange:
    pack -> fused scale + all-reduce -> unpack, stream-ordered with the backward that produced
    the gradients and with the ``optimizer.step()`` that follows at train-task.py:295.
    """
    comm = comm or default_comm()
    comm.poll()  # raises if an earlier exchange was poisoned by a missing peer
    grads = [p.grad for p in model.parameters()]  # the reference dereferences every .grad too
    if any(g is None for g in grads):
        raise AttributeError("average_gradients: a parameter has no .grad (train-task.py:68 would fail too)")
    if not grads:
        return
    if any(g.dtype != torch.float32 for g in grads):
        raise NotImplementedError("average_gradients: fp32 gradients only")
    sizes = [g.numel() for g in grads]
    total = sum(sizes)
    # One persistent flat buffer per (communicator, size), registered once for peer access, so
    # every later step is: pack -> zero-copy all-reduce in place -> unpack.
    cache = getattr(comm, "_flat_cache", None)
    if cache is None:
        cache = comm._flat_cache = {}
    entry = cache.get(total)
    if entry is None:
        flat = torch.empty(total, dtype=torch.float32, device=comm.device)
        reg = comm.register(flat) if comm.world > 1 and comm.wire == "fp32" else None
        entry = cache[total] = (flat, reg)
    flat, reg = entry
    views = list(flat.split(sizes))
    torch._foreach_copy_(views, [g.reshape(-1) for g in grads])
    if reg is not None:
        comm.allreduce_registered_(reg, flat, 1.0 / comm.world)
    else:
        comm.allreduce_(flat, 1.0 / comm.world)
    torch._foreach_copy_(grads, [v.view_as(g) for v, g in zip(views, grads)])
