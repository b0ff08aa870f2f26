"""Sharded statevectors: one process per GPU, the state split by its top qubits.

Launch with ``python -m torch.distributed.run --nproc-per-node N ...`` (or any launcher that
sets RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT).  The only thing the ranks
need from the host side is the 128-byte NCCL unique id of rank 0; it travels over a one-shot
TCP socket on MASTER_PORT + 29 (no PyTorch on this path), or through a caller-supplied
``broadcast`` callable (e.g. ``torch.distributed.broadcast_object_list``).

Everything after the bootstrap goes through the same C ABI as the single-GPU path:
``qsv_apply_gates`` plans the global<->local qubit swaps itself (csrc/qsv_dist.cu),
``qsv_expval_pauli_batch`` all-reduces the per-rank partial sums, ``qsv_sample`` routes each
uniform to the rank that owns it.
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import time

import numpy as np

from . import _lib
from .exceptions import QiskitError
from .statevector import Statevector


class DistContext:
    """Rank / world of this process, read from the launcher's environment."""

    def __init__(self, rank=None, world=None, local_rank=None):
        self.rank = int(os.environ.get("RANK", "0")) if rank is None else int(rank)
        self.world = int(os.environ.get("WORLD_SIZE", "1")) if world is None else int(world)
        self.local_rank = int(os.environ.get("LOCAL_RANK", str(self.rank))) if local_rank is None else int(local_rank)
        self.addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
        self.port = int(os.environ.get("MASTER_PORT", "29500"))
        self._calls = 0

    def broadcast_bytes(self, payload: bytes | None, size: int) -> bytes:
        """Rank 0's ``payload`` to every rank over a throw-away TCP socket."""
        if self.world == 1:
            return payload
        port = self.port + 29 + (self._calls % 32)
        self._calls += 1
        if self.rank == 0:
            srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
            srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            srv.bind((self.addr, port))
            srv.listen(self.world)
            for _ in range(self.world - 1):
                conn, _ = srv.accept()
                conn.sendall(payload)
                conn.close()
            srv.close()
            return payload
        deadline = time.time() + 120
        while True:
            try:
                conn = socket.create_connection((self.addr, port), timeout=5)
                break
            except OSError:
                if time.time() > deadline:
                    raise
                time.sleep(0.05)
        buf = b""
        while len(buf) < size:
            chunk = conn.recv(size - len(buf))
            if not chunk:
                raise RuntimeError("unique-id broadcast: connection closed early")
            buf += chunk
        conn.close()
        return buf


def sharded_statevector(num_qubits: int, ctx: DistContext | None = None, broadcast=None) -> Statevector:
    """|0...0> on ``num_qubits`` qubits split over ``ctx.world`` GPUs (this rank's slab on
    GPU ``ctx.local_rank``).  ``broadcast(bytes_or_None) -> bytes`` overrides the socket bootstrap."""
    ctx = DistContext() if ctx is None else ctx
    lib = _lib.load()
    if ctx.world == 1:
        return Statevector.from_int(0, 1 << num_qubits, device=ctx.local_rank)
    uid = C.create_string_buffer(128)
    if ctx.rank == 0:
        _lib.check(lib.qsv_dist_unique_id(uid))
    raw = bytes(uid.raw) if ctx.rank == 0 else None
    raw = broadcast(raw) if broadcast is not None else ctx.broadcast_bytes(raw, 128)
    if len(raw) != 128:
        raise QiskitError("NCCL unique id must be 128 bytes")
    uid = C.create_string_buffer(raw, 128)
    h = C.c_void_p()
    device = ctx.local_rank % max(_lib.device_count(), 1)
    _lib.check(lib.qsv_create_dist(C.byref(h), int(num_qubits), device, ctx.rank, ctx.world, uid))
    sv = object.__new__(Statevector)
    sv._rng_generator = None
    sv._device = device
    sv._n = int(num_qubits)
    sv._h = h
    sv._sharded = (ctx.rank, ctx.world)
    return sv


def local_slab(sv: Statevector) -> np.ndarray:
    """This rank's 2^(n - log2 world) amplitudes in the identity layout (device -> host)."""
    lib = _lib.load()
    nl = C.c_int()
    _lib.check(lib.qsv_dist_info(sv._h, None, None, C.byref(nl)))
    out = np.empty(1 << nl.value, dtype=np.complex128)
    _lib.check(lib.qsv_download(sv._h, out.ctypes.data, 0, out.size))
    return out
