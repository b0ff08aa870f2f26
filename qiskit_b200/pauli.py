"""Pauli observables in the form the expectation-value path reads them.

The path touches only ``Pauli.z / .x / .phase``, ``SparsePauliOp.paulis.z / .x`` and
``.coeffs`` (statevector.py:543-583).  These light classes provide
that surface when real Qiskit is absent; real ``Pauli`` / ``SparsePauliOp`` objects
are accepted wherever these are (duck typing).
"""
from __future__ import annotations

import numpy as np

from .exceptions import QiskitError

_PREFIXES = (("-i", 1), ("-j", 1), ("+i", 3), ("+j", 3), ("i", 3), ("j", 3), ("-1", 2), ("+1", 0),
             ("-", 2), ("+", 0), ("1", 0))


def _parse_label(label: str):
    """label -> (z, x, group phase k with coefficient (-i)^k)
    (qiskit/quantum_info/operators/symplectic/pauli.py:623-660)."""
    body, phase = label, 0
    for pref, ph in _PREFIXES:
        if body.startswith(pref) and all(ch in "IXYZ" for ch in body[len(pref):]):
            body, phase = body[len(pref):], ph
            break
    if any(ch not in "IXYZ" for ch in body):
        raise QiskitError(f'Pauli string label "{label}" is not valid.')
    n = len(body)
    z = np.zeros(n, dtype=bool)
    x = np.zeros(n, dtype=bool)
    for q, ch in enumerate(reversed(body)):  # rightmost character = qubit 0
        x[q] = ch in "XY"
        z[q] = ch in "ZY"
    return z, x, phase


class Pauli:
    """n-qubit Pauli with group phase: ``(-i)^phase * P``."""

    def __init__(self, data):
        if isinstance(data, str):
            self.z, self.x, self.phase = _parse_label(data)
        elif isinstance(data, Pauli):
            self.z, self.x, self.phase = data.z.copy(), data.x.copy(), data.phase
        else:
            z, x, *rest = data
            self.z = np.asarray(z, dtype=bool).copy()
            self.x = np.asarray(x, dtype=bool).copy()
            self.phase = int(rest[0]) % 4 if rest else 0
            if self.z.shape != self.x.shape or self.z.ndim != 1:
                raise QiskitError("z and x must be 1-D arrays of equal length")

    def __len__(self):
        return len(self.z)

    @property
    def num_qubits(self):
        return len(self.z)

    def _count_y(self):
        return np.array([int(np.sum(self.x & self.z))])

    def to_label(self) -> str:
        chars = {(False, False): "I", (False, True): "X", (True, True): "Y", (True, False): "Z"}
        body = "".join(chars[(bool(z), bool(x))] for z, x in zip(self.z[::-1], self.x[::-1]))
        return ("", "-i", "-", "i")[self.phase] + body

    def to_matrix(self) -> np.ndarray:
        mats = {"I": np.eye(2), "X": np.array([[0, 1], [1, 0]]), "Y": np.array([[0, -1j], [1j, 0]]),
                "Z": np.diag([1.0, -1.0])}
        label = self.to_label().lstrip("-i")
        m = np.array([[1.0 + 0j]])
        for ch in label:
            m = np.kron(m, mats[ch])
        return ((-1j) ** self.phase) * m

    def __repr__(self):
        return f"Pauli('{self.to_label()}')"


class PauliList:
    """Stack of phase-free Paulis: boolean arrays ``z``, ``x`` of shape [T, n]."""

    def __init__(self, z, x):
        self.z = np.asarray(z, dtype=bool)
        self.x = np.asarray(x, dtype=bool)

    def __len__(self):
        return self.z.shape[0]

    @property
    def num_qubits(self):
        return self.z.shape[1]


class SparsePauliOp:
    """sum_t coeffs[t] * P_t.  A label's sign / phase prefix is folded into the
    coefficient, as the reference does (sparse_pauli_op.py:166-182)."""

    def __init__(self, data, coeffs=None):
        if isinstance(data, SparsePauliOp):
            self.paulis, self.coeffs = data.paulis, data.coeffs.copy()
            return
        if isinstance(data, PauliList):
            zs, xs, phases = data.z, data.x, np.zeros(len(data), dtype=int)
        else:
            if isinstance(data, (str, Pauli)):
                data = [data]
            ps = [d if isinstance(d, Pauli) else Pauli(d) for d in data]
            if not ps:
                raise QiskitError("SparsePauliOp needs at least one term")
            if len({len(p) for p in ps}) != 1:
                raise QiskitError("Pauli terms must act on the same number of qubits")
            zs = np.array([p.z for p in ps])
            xs = np.array([p.x for p in ps])
            phases = np.array([p.phase for p in ps])
        if coeffs is None:
            coeffs = np.ones(len(zs), dtype=complex)
        coeffs = np.asarray(coeffs, dtype=complex).copy()
        if coeffs.shape != (len(zs),):
            raise QiskitError("coeffs must have one entry per Pauli term")
        self.paulis = PauliList(zs, xs)
        self.coeffs = coeffs * ((-1j) ** phases)

    @classmethod
    def from_list(cls, terms, num_qubits=None):
        terms = list(terms)
        if not terms:
            if num_qubits is None:
                raise QiskitError("Input Pauli list is empty.")
            return cls(["I" * num_qubits], [0.0])
        labels, coeffs = zip(*terms)
        return cls(list(labels), list(coeffs))

    @classmethod
    def from_sparse_list(cls, terms, num_qubits):
        """[(pauli_chars, qubit_indices, coeff)] as in ``SparsePauliOp.from_sparse_list``."""
        labels, coeffs = [], []
        for chars, qubits, coeff in terms:
            lab = ["I"] * num_qubits
            for ch, q in zip(chars, qubits):
                lab[num_qubits - 1 - q] = ch
            labels.append("".join(lab))
            coeffs.append(coeff)
        return cls(labels, coeffs)

    def __len__(self):
        return len(self.coeffs)

    @property
    def num_qubits(self):
        return self.paulis.num_qubits

    def to_list(self):
        return [(Pauli((z, x)).to_label(), c) for z, x, c in zip(self.paulis.z, self.paulis.x, self.coeffs)]

    def to_matrix(self):
        return sum(c * Pauli((z, x)).to_matrix()
                   for z, x, c in zip(self.paulis.z, self.paulis.x, self.coeffs))


def pauli_masks(z: np.ndarray, x: np.ndarray, qubits=None):
    """Bit masks over state qubits for [T, n] boolean arrays
    (statevector.py:543-550: mask = dot(1 << qubits, bits))."""
    z = np.atleast_2d(np.asarray(z, dtype=bool))
    x = np.atleast_2d(np.asarray(x, dtype=bool))
    n = z.shape[1]
    qubits = np.arange(n) if qubits is None else np.asarray(qubits)
    if len(qubits) != n:
        raise QiskitError("qargs length does not match the Pauli's number of qubits")
    if n and int(qubits.max()) >= 63:
        raise OverflowError("qubit index too large for a 64-bit mask")
    weights = (np.uint64(1) << qubits.astype(np.uint64))
    z_mask = (z.astype(np.uint64) * weights).sum(axis=1, dtype=np.uint64)
    x_mask = (x.astype(np.uint64) * weights).sum(axis=1, dtype=np.uint64)
    n_y = (z & x).sum(axis=1)
    return z_mask, x_mask, n_y
