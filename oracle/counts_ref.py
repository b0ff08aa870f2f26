"""Pure-Python restatement of the reference's counts post-processing (TEST INFRASTRUCTURE,
SURVEY 8f N4).  See oracle/__init__.py for who may import this.  Paths relative to
/root/reference."""
from __future__ import annotations

import numpy as np

# crates/accelerate/src/sampled_exp_val.rs:26-33: diagonal of each operator character
OPERS = {"Z": (1.0, -1.0), "0": (1.0, 0.0), "1": (0.0, 1.0)}


def bitstring_expval(dist: dict, oper_str: str) -> float:
    """sampled_exp_val.rs:46-67: sum_bits val * prod_{non-I positions} diag[bit] / sum(val)."""
    inds = [i for i, ch in enumerate(oper_str) if ch != "I"]
    opers = [ch for ch in oper_str if ch != "I"]
    denom = float(np.sum(np.fromiter(dist.values(), dtype=float)))
    exp_val = 0.0
    for bits, val in dist.items():
        prod = 1.0
        for idx, oper in zip(inds, opers):
            prod *= OPERS[oper][int(bits[idx])]
        exp_val += val * prod
    return exp_val / denom


def sampled_expval(oper_strs, coeffs, dist: dict):
    """sampled_expval_float / sampled_expval_complex (sampled_exp_val.rs:101-130)."""
    return sum(c * bitstring_expval(dist, s) for s, c in zip(oper_strs, coeffs))


def counts_from_ints(samples, num_bits: int) -> dict:
    """np.unique(..., return_counts=True) -> bit-string keyed dict (quantum_state.py:291-293,
    bit_array.py get_counts)."""
    vals, cnt = np.unique(np.asarray(samples, dtype=np.uint64), return_counts=True)
    return {format(int(v), f"0{num_bits}b"): int(c) for v, c in zip(vals, cnt)}


def marginalize_ints(samples, src_bits) -> np.ndarray:
    """Bit j of the result = bit src_bits[j] of the sample (marginalization.rs map_memory /
    BitArray.slice_bits restated on integers)."""
    samples = np.asarray(samples, dtype=np.uint64)
    out = np.zeros_like(samples)
    for j, b in enumerate(src_bits):
        if b >= 0:
            out |= ((samples >> np.uint64(b)) & np.uint64(1)) << np.uint64(j)
    return out
