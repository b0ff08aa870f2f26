"""NumPy restatement of the reference statevector path (TEST INFRASTRUCTURE).

See oracle/__init__.py for who may import this.  Every function names the
reference lines it follows (paths relative to /root/reference).  States are
1-D C-contiguous complex128 arrays of length 2**n; qubit q is bit q of the
index (little-endian, statevector.py:937).

Gate lists used throughout the repo's tests: a list of ``(matrix, qargs)``
tuples, plus ``("phase", theta)`` entries for a definition's global phase
(statevector.py:1026-1028).
"""
from __future__ import annotations

import numpy as np

# --------------------------------------------------------------------------
# A1: gate application  (statevector.py:925-963)
# --------------------------------------------------------------------------


def evolve_operator(psi: np.ndarray, mat: np.ndarray, qargs=None) -> np.ndarray:
    """psi' = (mat on qargs) psi, out of place.

    Same three steps as ``Statevector._evolve_operator``: move the target axes
    to the front of a (2,)*n tensor (statevector.py:936-949), contract with a
    single ``np.dot`` on a (2^k, 2^(n-k)) view (:954), undo the transpose (:959).
    ``qargs=None`` is the full-system ``np.dot`` (:929-933).
    """
    psi = np.asarray(psi, dtype=np.complex128)
    mat = np.asarray(mat, dtype=np.complex128)
    if qargs is None:
        return np.dot(mat, psi)
    n = int(round(np.log2(psi.size)))
    k = len(qargs)
    front = [n - 1 - q for q in reversed(qargs)]
    order = front + [a for a in range(n) if a not in front]
    undo = np.argsort(order).tolist()
    t = np.transpose(psi.reshape((2,) * n), order)
    shape = t.shape
    t = np.dot(mat, t.reshape(2**k, 2 ** (n - k))).reshape(shape)
    return np.transpose(t, undo).reshape(2**n)


def apply_gate_list(psi: np.ndarray, gates) -> np.ndarray:
    """Run a flat gate list: the leaf calls ``_evolve_instruction`` makes
    (statevector.py:978-981 for matrices, :1026-1028 for global phase)."""
    for g in gates:
        if isinstance(g[0], str) and g[0] == "phase":
            psi = psi * np.exp(1j * float(g[1]))
        else:
            psi = evolve_operator(psi, g[0], list(g[1]))
    return psi


def basis_state(n: int, index: int = 0) -> np.ndarray:
    """``Statevector.from_int`` (statevector.py:790-816)."""
    psi = np.zeros(2**n, dtype=np.complex128)
    psi[index] = 1.0
    return psi


# --------------------------------------------------------------------------
# A3/A4: Pauli expectation kernels (crates/accelerate/src/pauli_exp_val.rs)
# --------------------------------------------------------------------------


def _parity(v: np.ndarray) -> np.ndarray:
    """popcount(v) & 1 for uint64 arrays."""
    v = v.copy()
    for s in (32, 16, 8, 4, 2, 1):
        v ^= v >> np.uint64(s)
    return (v & np.uint64(1)).astype(bool)


def expval_pauli_no_x(data: np.ndarray, num_qubits: int, z_mask: int) -> float:
    """pauli_exp_val.rs:39-65: sum_i (-1)^popcount(i & z_mask) * |a_i|^2."""
    if num_qubits >= 64:
        raise OverflowError(
            f"The value for num_qubits, {num_qubits}, is too large and would overflow"
        )
    val = data.real * data.real + data.imag * data.imag
    idx = np.arange(1 << num_qubits, dtype=np.uint64)
    neg = _parity(idx & np.uint64(z_mask))
    return float(np.sum(np.where(neg, -val, val)))


def expval_pauli_with_x(
    data: np.ndarray, num_qubits: int, z_mask: int, x_mask: int, phase: complex, x_max: int
) -> float:
    """pauli_exp_val.rs:70-123, formula kept literally.

    i0 = i with a zero inserted at bit x_max (:84-89), i1 = i0 ^ x_mask,
    v0 = Re(phase * a[i1] * conj(a[i0])), v1 = Re(phase * a[i0] * conj(a[i1])),
    term = s(i0) * v0 + s(i1) * v1 with s the Z-parity sign (:108-118).
    """
    if num_qubits > 64:
        raise OverflowError(
            f"The value for num_qubits, {num_qubits}, is too large and would overflow"
        )
    i = np.arange(1 << (num_qubits - 1), dtype=np.uint64)
    mask_l = np.uint64((1 << x_max) - 1)
    mask_u = np.uint64(((1 << 64) - 1) ^ ((1 << (x_max + 1)) - 1))
    i0 = ((i << np.uint64(1)) & mask_u) | (i & mask_l)
    i1 = i0 ^ np.uint64(x_mask)
    a0, a1 = data[i0], data[i1]
    ph = complex(phase)
    v0 = (ph * (a1 * np.conj(a0))).real
    v1 = (ph * (a0 * np.conj(a1))).real
    s0 = _parity(i0 & np.uint64(z_mask))
    s1 = _parity(i1 & np.uint64(z_mask))
    return float(np.sum(np.where(s0, -v0, v0) + np.where(s1, -v1, v1)))


# --------------------------------------------------------------------------
# A5/A14: Pauli input format and the expectation_value dispatcher
# --------------------------------------------------------------------------


def pauli_from_label(label: str):
    """(z, x, group_phase) of a Pauli label.

    Labels are big-endian (rightmost character = qubit 0); an optional prefix
    '', '-', 'i', '-i' (also '+', '1', '+1', 'j' variants) carries the group
    phase (-i)^phase (qiskit/quantum_info/operators/symplectic/pauli.py:280-284,
    :623-660).
    """
    phase = 0
    body = label
    for pref, ph in (("-i", 1), ("-j", 1), ("+i", 3), ("+j", 3), ("i", 3), ("j", 3),
                     ("-1", 2), ("+1", 0), ("-", 2), ("+", 0), ("1", 0)):
        if body.startswith(pref) and len(body) > len(pref) or body == pref:
            # only strip when what remains is a valid Pauli body (or empty)
            rest = body[len(pref):]
            if all(ch in "IXYZ" for ch in rest):
                phase, body = ph, rest
                break
    n = len(body)
    z = np.zeros(n, dtype=bool)
    x = np.zeros(n, dtype=bool)
    for pos, ch in enumerate(reversed(body)):
        if ch not in "IXYZ":
            raise ValueError(f"bad Pauli label {label!r}")
        x[pos] = ch in "XY"
        z[pos] = ch in "ZY"
    return z, x, phase


def expectation_value_pauli(psi, z, x, group_phase: int = 0, qargs=None) -> complex:
    """``Statevector._expectation_value_pauli`` (statevector.py:533-562)."""
    n = int(round(np.log2(psi.size)))
    z = np.asarray(z, dtype=bool)
    x = np.asarray(x, dtype=bool)
    qubits = np.arange(len(z)) if qargs is None else np.array(qargs)
    x_mask = int(np.dot(1 << qubits, x))
    z_mask = int(np.dot(1 << qubits, z))
    pauli_phase = (-1j) ** group_phase if group_phase else 1
    if x_mask == 0:
        return pauli_phase * expval_pauli_no_x(psi, n, z_mask)
    x_max = int(qubits[x][-1])
    y_phase = (-1j) ** int(np.sum(x & z))
    return pauli_phase * expval_pauli_with_x(psi, n, z_mask, x_mask, y_phase, x_max)


def expectation_value_sparse(psi, zs, xs, coeffs, qargs=None) -> complex:
    """SparsePauliOp branch of ``expectation_value`` (statevector.py:579-583):
    Python ``sum`` over terms in order, each term a phase-0 ``Pauli((z, x))``."""
    return sum(
        coeff * expectation_value_pauli(psi, z, x, 0, qargs)
        for z, x, coeff in zip(zs, xs, coeffs)
    )


def expectation_value_labels(psi, terms, qargs=None) -> complex:
    """Convenience: terms = [(label, coeff)].  A label's sign/phase prefix is
    folded into the coefficient the way ``SparsePauliOp`` stores it
    (sparse_pauli_op.py:166-182)."""
    total = 0
    for label, coeff in terms:
        z, x, ph = pauli_from_label(label)
        total = total + coeff * ((-1j) ** ph) * expectation_value_pauli(psi, z, x, 0, qargs)
    return total


def expectation_value_operator(psi, mat, qargs=None) -> complex:
    """Generic-operator branch (statevector.py:585-587)."""
    val = evolve_operator(psi, mat, qargs)
    return np.dot(np.conj(psi), val)


# --------------------------------------------------------------------------
# A6/A7: probabilities
# --------------------------------------------------------------------------


def subsystem_probabilities(probs: np.ndarray, n: int, qargs=None) -> np.ndarray:
    """``QuantumState._subsystem_probabilities`` for qubit dims
    (quantum_state.py:437-468): sum out the unmeasured axes, then permute so
    that output bit j is qubit qargs[j]."""
    if qargs is None:
        return probs
    t = probs.reshape((2,) * n)
    axes = [n - 1 - q for q in reversed(qargs)]
    summed = tuple(a for a in range(n) if a not in axes)
    if summed:
        t = np.sum(t, axis=summed)
        axes = np.argsort(np.argsort(axes))
    t = np.transpose(t, axes=axes)
    return t.reshape(t.size)


def probabilities(psi: np.ndarray, qargs=None, decimals=None) -> np.ndarray:
    """``Statevector.probabilities`` (statevector.py:663-673)."""
    n = int(round(np.log2(psi.size)))
    probs = subsystem_probabilities(np.abs(psi) ** 2, n, qargs)
    probs = np.clip(probs, a_min=0, a_max=1)
    if decimals is not None:
        probs = probs.round(decimals=decimals)
    return probs


# --------------------------------------------------------------------------
# A8/A9: sampling.  The arithmetic is NumPy's Generator.choice (called at
# quantum_state.py:265); its algorithm for p given, replace=True is
#   cdf = p.cumsum(); cdf /= cdf[-1]; u = rng.random(size);
#   idx = cdf.searchsorted(u, side="right")
# (numpy/random/_generator.pyx, Generator.choice).  ``choice_indices`` restates
# it so tests can feed explicit uniforms; ``tests/test_oracle_kats.py`` checks
# it against the real ``rng.choice`` call.
# --------------------------------------------------------------------------


def choice_indices(p: np.ndarray, uniforms: np.ndarray) -> np.ndarray:
    cdf = np.cumsum(p)
    cdf /= cdf[-1]
    return cdf.searchsorted(uniforms, side="right").astype(np.int64)


def index_to_ket_strings(inds: np.ndarray, m: int) -> np.ndarray:
    """``_index_to_ket_array(..., string_labels=True)`` for qubit dims
    (quantum_state.py:333-364): qubit 0 is the rightmost character."""
    inds = np.asarray(inds, dtype=np.int64)
    if m == 0:
        return np.array([""] * len(inds))
    out = None
    for q in range(m):
        col = ((inds >> q) & 1).astype(np.str_)
        out = col if out is None else np.char.add(col, out)
    return out


def sample_memory(psi, shots: int, qargs=None, rng=None, *, as_index: bool = False):
    """``QuantumState.sample_memory`` (quantum_state.py:236-265).

    With ``as_index=False`` it runs the literal reference recipe (string labels
    for all 2^m outcomes + ``rng.choice``), only feasible for small m."""
    rng = np.random.default_rng() if rng is None else rng
    probs = probabilities(psi, qargs)
    m = int(round(np.log2(probs.size)))
    if as_index:
        return choice_indices(probs, rng.random(shots))
    labels = index_to_ket_strings(np.arange(len(probs)), m)
    return rng.choice(labels, p=probs, size=shots)


def sample_counts(psi, shots, qargs=None, rng=None) -> dict:
    """quantum_state.py:267-294."""
    inds, counts = np.unique(sample_memory(psi, shots, qargs, rng), return_counts=True)
    return {str(k): int(v) for k, v in zip(inds, counts)}


# --------------------------------------------------------------------------
# A10: reset / measure
# --------------------------------------------------------------------------


def reset(psi, qargs=None, rng=None) -> np.ndarray:
    """``Statevector.reset`` (statevector.py:675-719)."""
    n = int(round(np.log2(psi.size)))
    if qargs is None:
        return basis_state(n, 0)
    rng = np.random.default_rng() if rng is None else rng
    probs = probabilities(psi, qargs)
    sample = rng.choice(len(probs), p=probs, size=1)
    proj = np.zeros(len(probs), dtype=complex)
    proj[sample] = 1 / np.sqrt(probs[sample])
    rot = np.eye(len(probs))
    rot[0, 0] = 0
    rot[sample, sample] = 0
    rot[0, sample] = 1
    rot = np.dot(rot, np.diag(proj))
    return evolve_operator(psi, rot, list(qargs))


def measure(psi, qargs=None, rng=None):
    """``QuantumState.measure`` (quantum_state.py:296-331) -> (outcome, state)."""
    n = int(round(np.log2(psi.size)))
    rng = np.random.default_rng() if rng is None else rng
    qa = list(range(n)) if qargs is None else list(qargs)
    probs = probabilities(psi, qargs)
    sample = rng.choice(len(probs), p=probs, size=1)
    outcome = index_to_ket_strings(sample, len(qa))[0]
    proj = np.zeros(len(probs), dtype=complex)
    proj[sample] = 1 / np.sqrt(probs[sample])
    return str(outcome), evolve_operator(psi, np.diag(proj), qa)


# --------------------------------------------------------------------------
# A11/A12: the primitives' loops
# --------------------------------------------------------------------------


def estimator_pub(n, gate_lists, observables, precision=0.0, seed=None) -> np.ndarray:
    """``StatevectorEstimator._run_pub`` (statevector_estimator.py:150-176) for
    already-bound circuits.  ``gate_lists`` and ``observables`` are object
    arrays / nested lists that broadcast against each other; an observable is
    a list of (label, coeff).  Returns ``evs``."""
    rng = np.random.default_rng(seed)
    gl = np.empty(np.shape(_obj_array(gate_lists)), dtype=object)
    gl[...] = _obj_array(gate_lists)
    ob = _obj_array(observables)
    bc, bo = np.broadcast_arrays(gl, ob)
    evs = np.zeros(bc.shape, dtype=np.float64)
    for index in np.ndindex(*bc.shape):
        psi = apply_gate_list(basis_state(n), bc[index])
        ev = np.real_if_close(expectation_value_labels(psi, bo[index]))
        if precision != 0:
            if not np.isreal(ev):
                raise ValueError("Given operator is not Hermitian and noise cannot be added.")
            ev = rng.normal(ev, precision)
        evs[index] = ev
    return evs


def _obj_array(nested):
    """Nested lists of gate lists / observables -> object ndarray whose leaves
    are the innermost lists of tuples."""
    def depth_is_leaf(v):
        return isinstance(v, list) and (len(v) == 0 or isinstance(v[0], tuple))

    if depth_is_leaf(nested):
        a = np.empty((), dtype=object)
        a[()] = nested
        return a
    subs = [_obj_array(v) for v in nested]
    out = np.empty((len(subs),) + subs[0].shape, dtype=object)
    for i, sub in enumerate(subs):
        out[i] = sub[()] if sub.ndim == 0 else sub
    return out


def samples_to_packed_array(sample_bits: np.ndarray, num_bits: int, indices) -> np.ndarray:
    """``_samples_to_packed_array`` (statevector_sampler.py:235-251).

    ``sample_bits`` is the uint8 matrix [shots, m] in ket-string order (column 0
    = highest measured qubit), ``indices[c]`` = position (in the sorted measured
    qubit list) feeding clbit c, or m for "unmeasured" (sentinel, :217-222)."""
    ary = np.pad(sample_bits[:, ::-1], ((0, 0), (0, 1)), constant_values=0)
    ary = ary[:, list(indices)[::-1]]
    pad = -num_bits % 8
    ary = np.pad(ary, ((0, 0), (pad, 0)), constant_values=0)
    return np.packbits(ary, axis=-1)


def sampler_pack_indices(sample_idx: np.ndarray, m: int, num_bits: int, indices) -> np.ndarray:
    """Index-array entry to the sampler's packing: builds the [shots, m] bit
    matrix the reference gets from its ket strings (statevector_sampler.py:193)
    and packs it."""
    sample_idx = np.asarray(sample_idx, dtype=np.int64)
    bits = ((sample_idx[:, None] >> np.arange(m - 1, -1, -1)[None, :]) & 1).astype(np.uint8)
    if m == 0:
        bits = np.zeros((len(sample_idx), 0), dtype=np.uint8)
    return samples_to_packed_array(bits, num_bits, indices)


# --------------------------------------------------------------------------
# Engine summation tree for sampling (NOT a reference function).
#
# NumPy's cumsum is one sequential fp64 chain; a GPU scan must use a tree.  The
# engine's contract (DESIGN.md "sampling") is stated here in NumPy so tests can
# demand bit-exact sample indices from the GPU against it, and count the (rare,
# +-1) differences against ``choice_indices`` (= NumPy's own recipe).
#
#   * p_i = fl(fl(re*re) + fl(im*im))  (two products, one add, no FMA);
#   * level 0 = the 2^m probabilities, level k+1 = the block totals of level k,
#     block = 1024 items = 32 segments x 32 items; the last level has one block;
#   * inside a block: w = inclusive prefix; items of a segment are summed
#     sequentially (s), the 32 segment totals are chained sequentially (E), and
#     w[seg, k] = fl(E[seg] + s[seg, k]); block total = w[31, 31];
#   * search for a uniform u descends from the top block with exclusive prefix
#     X = 0: pick the first item j of the block with fl(fl(X + w_j) / total) > u
#     (clamped to the block's last valid item), set X = fl(X + w_{j-1}) (X if
#     j = 0) and descend into item j's block; ``total`` is the top block's total.
# --------------------------------------------------------------------------

SEG = 32
LANES = 32
BLK = SEG * LANES


def tree_probabilities(psi: np.ndarray) -> np.ndarray:
    return psi.real * psi.real + psi.imag * psi.imag


def _inblock_prefix(v: np.ndarray):
    """v: 1-D, length multiple of 1024 -> (w same shape, block totals)."""
    b = v.reshape(-1, LANES, SEG)
    s = np.cumsum(b, axis=2)
    t = s[:, :, -1]
    e = np.zeros_like(t)
    e[:, 1:] = np.cumsum(t, axis=1)[:, :-1]
    w = e[:, :, None] + s
    return w.reshape(-1), w[:, -1, -1].copy()


def tree_levels(p: np.ndarray):
    """[(w, n_valid)] from level 0 upwards; the last entry has one block."""
    levels = []
    v = np.asarray(p, dtype=np.float64)
    while True:
        n_valid = v.size
        pad = (-n_valid) % BLK
        if pad:
            v = np.concatenate([v, np.zeros(pad)])
        w, totals = _inblock_prefix(v)
        levels.append((w, n_valid))
        if totals.size == 1:
            return levels
        v = totals


def tree_choice_indices(p: np.ndarray, uniforms: np.ndarray) -> np.ndarray:
    """Sample indices under the engine's tree (see the block comment above)."""
    levels = tree_levels(p)
    u = np.asarray(uniforms, dtype=np.float64)
    total = levels[-1][0][levels[-1][1] - 1]
    x = np.zeros(u.shape, dtype=np.float64)
    blk = np.zeros(u.shape, dtype=np.int64)
    for w, n_valid in reversed(levels):
        base = blk * BLK
        last = np.minimum(BLK, n_valid - base) - 1          # last valid item of the block
        lo = np.zeros(u.shape, dtype=np.int64)
        hi = last.copy()                                      # answer in [0, last]
        while np.any(lo < hi):
            mid = (lo + hi) >> 1
            hit = ((x + w[base + mid]) / total) > u
            hi = np.where(hit, mid, hi)
            lo = np.where(hit, lo, np.minimum(mid + 1, hi))
        j = lo
        prev = w[base + np.maximum(j - 1, 0)]
        x = np.where(j > 0, x + prev, x)
        blk = base + j
    return blk


def tree_choice_indices_sharded(p: np.ndarray, uniforms: np.ndarray, world: int) -> np.ndarray:
    """The engine's tree for a state split over ``world`` ranks (csrc/qsv_sample.cu): each rank
    builds the tree above over its own slab; the rank totals t_r are chained sequentially,
    chain[r+1] = fl(chain[r] + t_r); a uniform belongs to the first rank with
    fl(chain[r+1] / total) > u (clamped to the last rank) and descends that rank's tree with
    X = chain[r]."""
    p = np.asarray(p, dtype=np.float64)
    u = np.asarray(uniforms, dtype=np.float64)
    slabs = p.reshape(world, -1)
    levels = [tree_levels(s) for s in slabs]
    tot = [lv[-1][0][lv[-1][1] - 1] for lv in levels]
    chain = [0.0]
    for r in range(world):
        chain.append(tot[r] if r == 0 else chain[-1] + tot[r])
    total = chain[-1]
    owner = np.full(u.shape, world - 1, dtype=np.int64)
    for r in range(world - 2, -1, -1):
        owner = np.where((chain[r + 1] / total) > u, r, owner)
    out = np.zeros(u.shape, dtype=np.int64)
    for r in range(world):
        sel = np.nonzero(owner == r)[0]
        if sel.size == 0:
            continue
        x = np.full(sel.shape, chain[r], dtype=np.float64)
        blk = np.zeros(sel.shape, dtype=np.int64)
        uu = u[sel]
        for w, n_valid in reversed(levels[r]):
            base = blk * BLK
            last = np.minimum(BLK, n_valid - base) - 1
            lo = np.zeros(sel.shape, dtype=np.int64)
            hi = last.copy()
            while np.any(lo < hi):
                mid = (lo + hi) >> 1
                hit = ((x + w[base + mid]) / total) > uu
                hi = np.where(hit, mid, hi)
                lo = np.where(hit, lo, np.minimum(mid + 1, hi))
            j = lo
            prev = w[base + np.maximum(j - 1, 0)]
            x = np.where(j > 0, x + prev, x)
            blk = base + j
        out[sel] = blk + r * slabs.shape[1]
    return out
