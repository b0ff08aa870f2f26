"""ctypes binding of libqsv_b200.so (the C ABI in include/qsv.h).

No CPU fallback: if the shared library is missing or no CUDA device is visible,
every compute entry point raises.  The library is built in-tree by
``python -m qiskit_b200.build`` (or ``__graft_entry__.build()``).
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import numpy as np

from .exceptions import QiskitError

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libqsv_b200.so")

QSV_OK = 0
ERR_INVALID, ERR_CUDA, ERR_NOMEM, ERR_NCCL, ERR_UNSUPPORTED, ERR_OVERFLOW = 1, 2, 3, 4, 5, 6
MAX_GATE_QUBITS = 12
GATE_MATRIX, GATE_DIAGONAL, GATE_PHASE = 0, 1, 2
FLAG_NO_FUSION = 1


class QsvGate(C.Structure):
    _fields_ = [
        ("kind", C.c_int32),
        ("k", C.c_int32),
        ("qargs", C.c_int32 * MAX_GATE_QUBITS),
        ("data", C.c_void_p),
        ("params", C.c_double * 4),
    ]


class QsvCounters(C.Structure):
    _fields_ = [
        ("kernel_launches", C.c_uint64),
        ("state_passes", C.c_uint64),
        ("read_passes", C.c_uint64),
        ("gates_applied", C.c_uint64),
        ("h2d_bytes", C.c_uint64),
        ("d2h_bytes", C.c_uint64),
        ("phases", C.c_uint64),
        ("swap_count", C.c_uint64),
        ("swap_bytes", C.c_uint64),
        ("write_passes", C.c_uint64),
        ("swap_ns", C.c_uint64),
        ("fused_swaps", C.c_uint64),
    ]


# every exported symbol of include/qsv.h with its prototype (tests check the list against the header)
_P = C.POINTER
_vp, _i, _u64, _dbl = C.c_void_p, C.c_int, C.c_uint64, C.c_double
PROTOTYPES = {
    "qsv_version": (C.c_int, []),
    "qsv_last_error": (C.c_char_p, []),
    "qsv_device_count": (_i, [_P(_i)]),
    "qsv_device_memory": (_i, [_i, _P(_u64), _P(_u64)]),
    "qsv_create": (_i, [_P(_vp), _i, _i]),
    "qsv_destroy": (_i, [_vp]),
    "qsv_clone": (_i, [_vp, _P(_vp)]),
    "qsv_copy_state": (_i, [_vp, _vp]),
    "qsv_num_qubits": (_i, [_vp, _P(_i)]),
    "qsv_init_basis": (_i, [_vp, _u64]),
    "qsv_upload": (_i, [_vp, _vp, _u64, _u64]),
    "qsv_download": (_i, [_vp, _vp, _u64, _u64]),
    "qsv_sync": (_i, [_vp]),
    "qsv_apply_matrix": (_i, [_vp, _vp, _i, _P(_i)]),
    "qsv_apply_diagonal": (_i, [_vp, _vp, _i, _P(_i)]),
    "qsv_apply_gates": (_i, [_vp, _P(QsvGate), _u64, _i]),
    "qsv_scale": (_i, [_vp, _dbl, _dbl]),
    "qsv_expval_pauli": (_i, [_vp, _u64, _u64, _dbl, _dbl, _i, _P(_dbl)]),
    "qsv_expval_pauli_batch": (_i, [_vp, _u64, _vp, _vp, _vp, _vp]),
    "qsv_norm2": (_i, [_vp, _P(_dbl)]),
    "qsv_inner": (_i, [_vp, _vp, _P(_dbl)]),
    "qsv_inner_static_scratch": (_i, [_i, _vp, _vp, _u64, _vp, _vp]),
    "qsv_probabilities": (_i, [_vp, _P(_i), _i, _vp]),
    "qsv_sample": (_i, [_vp, _P(_i), _i, _vp, _u64, _vp, _P(_dbl)]),
    "qsv_collapse": (_i, [_vp, _P(_i), _i, _u64, _dbl, _i]),
    "qsv_outcome_probability": (_i, [_vp, _P(_i), _i, _u64, _P(_dbl)]),
    "qsv_pack_samples": (_i, [_vp, _u64, _P(_i), _i, _vp]),
    "qsv_permute_qubits": (_i, [_vp, _vp, _P(_i)]),
    "qsv_tensor": (_i, [_vp, _vp, _vp, _P(_i), _P(_i), _i]),
    "qsv_axpy": (_i, [_vp, _vp, _dbl, _dbl]),
    "qsv_conjugate": (_i, [_vp]),
    "qsv_density_expval_pauli_batch": (_i, [_vp, _u64, _vp, _vp, _vp, _vp]),
    "qsv_density_probabilities": (_i, [_vp, _P(_i), _i, _vp]),
    "qsv_density_trace": (_i, [_vp, _P(_dbl)]),
    "qsv_density_purity": (_i, [_vp, _P(_dbl)]),
    "qsv_density_reset": (_i, [_vp, _P(_i), _i]),
    "qsv_sampled_expval": (_i, [_i, _vp, _vp, _u64, _u64, _vp, _vp, _vp, _vp]),
    "qsv_marginalize_samples": (_i, [_i, _vp, _u64, _P(_i), _i, _vp]),
    "qsv_sample_counts": (_i, [_i, _vp, _u64, _vp, _vp, _P(_u64)]),
    "qsv_get_counters": (_i, [_vp, _P(QsvCounters)]),
    "qsv_reset_counters": (_i, [_vp]),
    "qsv_process_counters": (_i, [_P(QsvCounters)]),
    "qsv_restore_defaults": (_i, [_vp]),
    "qsv_timer_begin": (_i, [_vp]),
    "qsv_timer_end": (_i, [_vp, _P(C.c_float)]),
    "qsv_set_option": (_i, [_vp, C.c_char_p, C.c_int64]),
    "qsv_flush_l2": (_i, [_vp]),
    "qsv_debug_plan": (_i, [_i, _P(QsvGate), _u64, _i, C.c_char_p]),
    "qsv_debug_fused_swap": (_i, [_i, _i, _i, _P(_i), _i, _P(QsvGate), _u64, _vp, _vp, _vp, _vp, _P(_i), _P(_i)]),
    "qsv_dist_unique_id": (_i, [_vp]),
    "qsv_create_dist": (_i, [_P(_vp), _i, _i, _i, _i, _vp]),
    "qsv_dist_info": (_i, [_vp, _P(_i), _P(_i), _P(_i)]),
    "qsv_dist_plan": (_i, [_i, _i, _P(_i), _u64, _vp, _vp, _vp, _i, _vp, _u64, _P(_u64), _P(_i)]),
    "qsv_dist_need_mask": (_i, [_P(QsvGate), _P(C.c_uint32)]),
}

_lib = None
_lock = threading.Lock()


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build it with `python -m qiskit_b200.build` "
                "(this engine has no CPU fallback)"
            )
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error() -> str:
    return load().qsv_last_error().decode("utf-8", "replace")


def check(status: int):
    """Map a C status to the exception the reference raises for the same fault."""
    if status == QSV_OK:
        return
    msg = last_error()
    if status == ERR_OVERFLOW:
        raise OverflowError(msg)       # pauli_exp_val.rs:44-48
    if status == ERR_INVALID:
        raise QiskitError(msg)
    if status == ERR_NOMEM:
        raise MemoryError(msg)
    if status == ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise RuntimeError(msg)


def device_count() -> int:
    n = C.c_int(0)
    check(load().qsv_device_count(C.byref(n)))
    return n.value


def require_device():
    if device_count() == 0:
        raise RuntimeError("qiskit_b200: no CUDA device visible and there is no CPU fallback")


def as_c128(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.complex128)


def int_array(vals):
    vals = [int(v) for v in vals]
    return (C.c_int * max(len(vals), 1))(*vals)


# --- handle pool: avoids cudaMalloc/cudaFree churn when primitives create one state per index ---
_pool: dict = {}
_pool_lock = threading.Lock()
_POOL_MAX_PER_KEY = 2


def acquire(n_qubits: int, device: int = 0):
    lib = load()
    with _pool_lock:
        free = _pool.get((n_qubits, device))
        if free:
            h = free.pop()
            lib.qsv_reset_counters(h)   # a recycled handle starts with clean counters, like a new one
            return h
    h = C.c_void_p()
    status = lib.qsv_create(C.byref(h), int(n_qubits), int(device))
    if status == ERR_NOMEM:  # pooled handles of other sizes may be holding the memory
        drain_pool()
        status = lib.qsv_create(C.byref(h), int(n_qubits), int(device))
    check(status)
    return h


def release(handle, n_qubits: int, device: int = 0):
    if handle is None or _lib is None:
        return
    with _pool_lock:
        free = _pool.setdefault((n_qubits, device), [])
        if len(free) < _POOL_MAX_PER_KEY and _lib.qsv_restore_defaults(handle) == QSV_OK:
            free.append(handle)   # tunables back to the defaults: the next owner starts clean
            return
    _lib.qsv_destroy(handle)


def process_counters() -> dict:
    """Counters summed over every handle of this process (qsv_process_counters)."""
    c = QsvCounters()
    check(load().qsv_process_counters(C.byref(c)))
    return {name: int(getattr(c, name)) for name, _ in c._fields_}


def drain_pool():
    with _pool_lock:
        for free in _pool.values():
            for h in free:
                _lib.qsv_destroy(h)
        _pool.clear()
