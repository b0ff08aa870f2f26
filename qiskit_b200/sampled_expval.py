"""Counts post-processing on the device.

``sampled_expectation_value`` mirrors qiskit/result/sampled_expval.py:37-91 (same argument
meaning, validation order and error messages); the arithmetic of
crates/accelerate/src/sampled_exp_val.rs:46-67 runs in ``qsv_sampled_expval``.  Outcomes are
handled as integers (bit q = character ``len - 1 - q`` of the bit-string), so at most 64 bits.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .exceptions import QiskitError

OPERS = {"Z", "I", "0", "1"}


def _label_masks(label: str):
    """(z, zero, one) bit masks of a diagonal operator string (rightmost character = bit 0)."""
    z = zero = one = 0
    n = len(label)
    for pos, ch in enumerate(label):
        bit = 1 << (n - 1 - pos)
        if ch == "Z":
            z |= bit
        elif ch == "0":
            zero |= bit
        elif ch == "1":
            one |= bit
    return z, zero, one


def _device_expvals(keys: np.ndarray, weights, labels, device: int = 0) -> np.ndarray:
    """One value per label: sum_k w_k prod_q d_q(bit) / sum_k w_k on the device."""
    masks = np.array([_label_masks(s) for s in labels], dtype=np.uint64).reshape(-1, 3)
    z = np.ascontiguousarray(masks[:, 0])
    zero = np.ascontiguousarray(masks[:, 1])
    one = np.ascontiguousarray(masks[:, 2])
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float64)
    out = np.empty(len(labels), dtype=np.float64)
    _lib.check(_lib.load().qsv_sampled_expval(
        int(device), keys.ctypes.data, None if w is None else w.ctypes.data, keys.size, len(labels),
        z.ctypes.data, zero.ctypes.data, one.ctypes.data, out.ctypes.data))
    return out


def _oper_terms(oper):
    """(labels, coeffs) as qiskit/result/sampled_expval.py:64-76 builds them."""
    if isinstance(oper, str):
        return [oper.upper()], np.asarray([1.0])
    if hasattr(oper, "paulis") and hasattr(oper, "coeffs"):          # SparsePauliOp
        z = np.asarray(oper.paulis.z, dtype=bool)
        x = np.asarray(oper.paulis.x, dtype=bool)
        labels = ["".join("IZXY"[int(zb) + 2 * int(xb)] for zb, xb in zip(zr[::-1], xr[::-1]))
                  for zr, xr in zip(z, x)]
        return labels, np.asarray(oper.coeffs)
    if hasattr(oper, "z") and hasattr(oper, "x") and np.ndim(oper.z) == 1:   # Pauli
        label = "".join("IZXY"[int(zb) + 2 * int(xb)] for zb, xb in zip(oper.z[::-1], oper.x[::-1]))
        return [label], np.asarray([1.0])
    raise QiskitError("Invalid operator type")


def sampled_expectation_value(dist, oper, *, device: int = 0):
    """Expectation value of a diagonal operator from a sampled distribution
    (qiskit/result/sampled_expval.py:37-91).  ``dist``: dict of bit-string -> counts or
    (quasi-)probabilities; ``oper``: str, ``Pauli`` or ``SparsePauliOp`` over I, Z, 0, 1."""
    if hasattr(dist, "binary_probabilities"):
        dist = dist.binary_probabilities()
    if not isinstance(dist, dict):
        raise QiskitError("Invalid input distribution type")
    oper_strs, coeffs = _oper_terms(oper)
    bitstring_len = len(next(iter(dist)))
    if any(len(op) != bitstring_len for op in oper_strs):
        raise QiskitError(
            f"One or more operators not same length ({bitstring_len}) as input bitstrings"
        )
    for op in oper_strs:
        if set(op).difference(OPERS):
            raise QiskitError(f"Input operator {op} is not diagonal")
    if bitstring_len > 64:
        raise QiskitError("bit-strings longer than 64 bits are outside the device path")
    keys = np.fromiter((int(k, 2) for k in dist), dtype=np.uint64, count=len(dist))
    weights = np.fromiter((float(v) for v in dist.values()), dtype=np.float64, count=len(dist))
    vals = _device_expvals(keys, weights, oper_strs, device)
    if np.iscomplexobj(coeffs):
        return np.real_if_close(complex(np.sum(coeffs * vals)))
    return float(np.sum(coeffs * vals))


def marginalize_samples(samples: np.ndarray, src_bits, *, device: int = 0) -> np.ndarray:
    """out[i] bit j = samples[i] bit ``src_bits[j]`` (negative: 0): marginal_memory /
    ``BitArray.slice_bits`` on integer outcomes
    (crates/accelerate/src/results/marginalization.rs)."""
    samples = np.ascontiguousarray(samples, dtype=np.uint64)
    out = np.empty_like(samples)
    src = _lib.int_array(src_bits)
    _lib.check(_lib.load().qsv_marginalize_samples(int(device), samples.ctypes.data, samples.size, src,
                                                   len(src_bits), out.ctypes.data))
    return out


def sample_counts(samples: np.ndarray, *, device: int = 0):
    """(unique outcomes ascending, their counts): ``np.unique(samples, return_counts=True)`` of
    quantum_state.py:291-293 / ``BitArray.get_int_counts`` through a device hash table."""
    samples = np.ascontiguousarray(samples, dtype=np.uint64)
    keys = np.empty(samples.size, dtype=np.uint64)
    counts = np.empty(samples.size, dtype=np.uint64)
    n = C.c_uint64(0)
    _lib.check(_lib.load().qsv_sample_counts(int(device), samples.ctypes.data, samples.size, keys.ctypes.data,
                                             counts.ctypes.data, C.byref(n)))
    return keys[: n.value].copy(), counts[: n.value].copy()
