"""Synthetic benchmark circuits and observables, Qiskit-free.

Each generator restates the *structure* of the reference's circuit-library entry
(cited per function) as a flat gate list ``[(matrix, qargs), ...]`` -- the format
both the engine (``Statevector.apply_gate_list``) and the oracle
(``oracle.statevector_ref.apply_gate_list``) consume -- or as a parameterised
``QuantumCircuit`` for the primitives.
"""
from __future__ import annotations

import numpy as np

from .circuit import ParameterVector, QuantumCircuit
from .gates import gate_matrix
from .pauli import SparsePauliOp


def named_to_matrices(named):
    """[(name, params, qargs)] -> [(matrix, qargs)]; ("unitary", matrix, qargs) passes through."""
    out = []
    for name, params, qargs in named:
        if name == "unitary":
            out.append((np.asarray(params, dtype=np.complex128), list(qargs)))
        else:
            out.append((gate_matrix(name, params), list(qargs)))
    return out


def haar_unitary(dim: int, rng: np.random.Generator) -> np.ndarray:
    """QR of a complex Ginibre matrix with the phases of R's diagonal fixed -- the
    algorithm of crates/circuit_library/src/quantum_volume.rs:37-85."""
    g = (rng.standard_normal((dim, dim)) + 1j * rng.standard_normal((dim, dim))) / np.sqrt(2)
    q, r = np.linalg.qr(g)
    d = np.diagonal(r)
    return q * (d / np.abs(d))


def quantum_volume(n: int, depth: int | None = None, seed: int = 0):
    """QuantumVolume-like model circuit (quantum_volume.rs:97-118): per layer a random
    permutation of the qubits, then a Haar SU(4) on each consecutive pair.  Same
    distribution as ``quantum_volume(n, depth, seed)``; not bit-identical (the reference
    draws from Rust's Pcg64Mcg)."""
    depth = n if depth is None else depth
    rng = np.random.default_rng(seed)
    gates = []
    for _ in range(depth):
        perm = rng.permutation(n)
        for w in range(n // 2):
            gates.append((haar_unitary(4, rng), [int(perm[2 * w]), int(perm[2 * w + 1])]))
    return gates


def random_two_qubit_stream(n: int, count: int, seed: int = 0, min_qubit: int = 0):
    """Unfusable stream for the roofline measurement: Haar SU(4) gates on random
    distinct qubit pairs; run with ``no_fusion=True`` so that 1 gate = 1 pass."""
    rng = np.random.default_rng(seed)
    gates = []
    for _ in range(count):
        a, b = rng.choice(np.arange(min_qubit, n), size=2, replace=False)
        gates.append((haar_unitary(4, rng), [int(a), int(b)]))
    return gates


def efficient_su2_named(n: int, reps: int, theta):
    """EfficientSU2(n, reps, entanglement='reverse_linear'): per repetition an RY layer,
    an RZ layer, then CX(q, q+1) for q = n-2 ... 0; a final RY, RZ layer
    (n_local/efficient_su2.py:31-45,116-134; entanglement.rs:29-42).  ``theta`` has
    2 n (reps + 1) entries in circuit order."""
    theta = np.asarray(theta, dtype=float)
    assert theta.size == 2 * n * (reps + 1)
    named, p = [], 0
    for r in range(reps + 1):
        for q in range(n):
            named.append(("ry", [theta[p]], [q]))
            p += 1
        for q in range(n):
            named.append(("rz", [theta[p]], [q]))
            p += 1
        if r < reps:
            for q in range(n - 2, -1, -1):
                named.append(("cx", [], [q, q + 1]))
    return named


def efficient_su2_circuit(n: int, reps: int) -> QuantumCircuit:
    """The same ansatz as a parameterised circuit (``ParameterVector('θ', 2n(reps+1))``);
    parameters sort numerically, so bare value arrays bind in circuit order."""
    theta = ParameterVector("θ", 2 * n * (reps + 1))
    qc = QuantumCircuit(n, name="EfficientSU2")
    p = 0
    for r in range(reps + 1):
        for q in range(n):
            qc.ry(theta[p], q)
            p += 1
        for q in range(n):
            qc.rz(theta[p], q)
            p += 1
        if r < reps:
            for q in range(n - 2, -1, -1):
                qc.cx(q, q + 1)
    return qc


def real_amplitudes_circuit(n: int, reps: int) -> QuantumCircuit:
    """real_amplitudes(n, reps) with reverse_linear entanglement
    (n_local/real_amplitudes.py): RY layer, then (CX chain, RY layer) * reps."""
    theta = ParameterVector("θ", n * (reps + 1))
    qc = QuantumCircuit(n, name="RealAmplitudes")
    p = 0
    for r in range(reps + 1):
        for q in range(n):
            qc.ry(theta[p], q)
            p += 1
        if r < reps:
            for q in range(n - 2, -1, -1):
                qc.cx(q, q + 1)
    return qc


def qft_named(n: int, do_swaps: bool = True, input_layer_seed: int | None = 3):
    """QFT per crates/synthesis/src/qft/qft_decompose_full.rs:79-110: for j = n-1..0:
    H(j), then CP(pi 2^(k-j)) on (j, k) for k = j-1..0; finally SWAP(i, n-1-i).  An RY
    input layer makes the state non-trivial."""
    named = []
    if input_layer_seed is not None:
        ang = np.random.default_rng(input_layer_seed).uniform(0, np.pi, n)
        for q in range(n):
            named.append(("ry", [ang[q]], [q]))
    for j in range(n - 1, -1, -1):
        named.append(("h", [], [j]))
        for k in range(j - 1, -1, -1):
            named.append(("cp", [np.pi * 2.0 ** (k - j)], [j, k]))
    if do_swaps:
        for i in range(n // 2):
            named.append(("swap", [], [i, n - 1 - i]))
    return named


def qft_circuit(n: int, measure: bool = True) -> QuantumCircuit:
    qc = QuantumCircuit(n, name="QFT")
    for name, params, qargs in qft_named(n):
        getattr(qc, name)(*params, *qargs)
    if measure:
        qc.measure_all()
    return qc


def tfim_named(n: int, steps: int = 10, dt: float = 0.1, j: float = 1.0, h: float = 1.0):
    """First-order Trotter steps of H = -J sum Z_i Z_{i+1} - h sum X_i: RY(pi/4) on all
    qubits, then per step RZZ(-2 J dt) on (i, i+1) and RX(-2 h dt) on every qubit
    (rotation angle 2 c t for a term c P, crates/synthesis/src/pauli_evolution.rs:83-160)."""
    named = [("ry", [np.pi / 4], [q]) for q in range(n)]
    for _ in range(steps):
        for i in range(n - 1):
            named.append(("rzz", [-2.0 * j * dt], [i, i + 1]))
        for q in range(n):
            named.append(("rx", [-2.0 * h * dt], [q]))
    return named


# ------------------------------------------------------------- observables --


def random_pauli_observable(n: int, terms: int, seed: int = 1, coeff_seed: int = 2) -> SparsePauliOp:
    """``random_pauli_list(n, terms, seed, phase=False)`` restated
    (symplectic/random.py:56-86: z bits then x bits from one generator) with
    standard-normal coefficients."""
    rng = np.random.default_rng(seed)
    z = rng.integers(2, size=(terms, n)).astype(bool)
    x = rng.integers(2, size=(terms, n)).astype(bool)
    coeffs = np.random.default_rng(coeff_seed).standard_normal(terms)
    from .pauli import PauliList

    return SparsePauliOp(PauliList(z, x), coeffs)


def heisenberg_observable(n: int, num_pairs: int = 324, field: float = 0.5) -> SparsePauliOp:
    """sum over the first ``num_pairs`` pairs (i<j, lexicographic, all-to-all) of
    XX + YY + ZZ, plus ``field`` * Z_q on every qubit: 3*num_pairs + n terms
    (1000 terms for n = 28, num_pairs = 324)."""
    terms = []
    pairs = [(i, j) for i in range(n) for j in range(i + 1, n)][:num_pairs]
    for i, j in pairs:
        for ch in "XYZ":
            terms.append((ch + ch, [i, j], 1.0))
    for q in range(n):
        terms.append(("Z", [q], field))
    return SparsePauliOp.from_sparse_list(terms, n)


def tfim_observable(n: int, j: float = 1.0, h: float = 1.0) -> SparsePauliOp:
    terms = [("ZZ", [i, i + 1], -j) for i in range(n - 1)] + [("X", [q], -h) for q in range(n)]
    return SparsePauliOp.from_sparse_list(terms, n)
