"""ctypes wrapper of oracle/pauli_expval.c (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Used as the multithreaded CPU baseline for the expectation-value kernels (the role
rayon plays in crates/accelerate/src/pauli_exp_val.rs) and cross-checked against the
NumPy restatement in tests/test_oracle_kats.py.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libpauli_expval_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "pauli_expval.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
    return _SO


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        lib = C.CDLL(_SO)
        lib.oracle_expval_pauli_no_x.restype = C.c_double
        lib.oracle_expval_pauli_no_x.argtypes = [C.c_void_p, C.c_uint, C.c_uint64]
        lib.oracle_expval_pauli_with_x.restype = C.c_double
        lib.oracle_expval_pauli_with_x.argtypes = [C.c_void_p, C.c_uint, C.c_uint64, C.c_uint64,
                                                   C.c_double, C.c_double, C.c_uint]
        lib.oracle_apply_1q.restype = None
        lib.oracle_apply_1q.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_void_p]
        _lib = lib
    return _lib


def expval_pauli_no_x(data: np.ndarray, num_qubits: int, z_mask: int) -> float:
    data = np.ascontiguousarray(data, dtype=np.complex128)
    return load().oracle_expval_pauli_no_x(data.ctypes.data, num_qubits, z_mask)


def expval_pauli_with_x(data, num_qubits, z_mask, x_mask, phase, x_max) -> float:
    data = np.ascontiguousarray(data, dtype=np.complex128)
    ph = complex(phase)
    return load().oracle_expval_pauli_with_x(data.ctypes.data, num_qubits, z_mask, x_mask,
                                             ph.real, ph.imag, x_max)


def expectation_value_sparse(psi, zs, xs, coeffs) -> complex:
    """statevector.py:533-583 with the C kernels."""
    n = int(round(np.log2(psi.size)))
    total = 0
    for z, x, coeff in zip(zs, xs, coeffs):
        qubits = np.arange(n)
        x_mask = int(np.dot(1 << qubits, x))
        z_mask = int(np.dot(1 << qubits, z))
        if x_mask == 0:
            v = expval_pauli_no_x(psi, n, z_mask)
        else:
            x_max = int(qubits[np.asarray(x, dtype=bool)][-1])
            v = expval_pauli_with_x(psi, n, z_mask, x_mask, (-1j) ** int(np.sum(x & z)), x_max)
        total = total + coeff * v
    return total
