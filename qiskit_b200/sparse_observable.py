"""``SparseObservable`` inputs for the Estimator (SURVEY.md 8f / VERDICT r1 item 9).

A qubit-sparse observable: sum_t coeff_t * prod_q op_{t,q} with op in the alphabet of
crates/quantum_info/src/sparse_observable/mod.rs:2244-2275 -- the Paulis ``X Y Z`` and the six
single-qubit projectors ``+ - r l 0 1`` (eigenstates of X, Y, Z).  The reference's Estimator containers
coerce every observable to this type (containers/observables_array.py:247-303) and hand Pauli labels to
``Statevector.expectation_value`` (statevector_estimator.py:164-166; projector terms are a TODO there).

The device path evaluates Pauli terms, so the only arithmetic here is ``as_paulis`` -- every projector
is 1/2 (I +- P), the table of ``bit_term_as_pauli`` (mod.rs:207-219) -- followed by merging equal Pauli
strings; a term with k projectors becomes 2^k Pauli terms (the reference warns about the same growth,
mod.rs:3127-3130).  Storage is a plain list of (coeff, {qubit: letter}) terms: host-side glue, not a
port of the Rust CSR-like layout.

Duck-typed on the way in: anything with ``to_sparse_list()`` and ``num_qubits`` -- Qiskit's own
``SparseObservable`` included -- is accepted by ``from_any``.
"""
from __future__ import annotations

import numpy as np

from .exceptions import QiskitError
from .pauli import SparsePauliOp

PAULIS = "XYZ"
# projector -> (sign of the Pauli part, Pauli letter):  |+><+| = (I + X)/2, |-><-| = (I - X)/2, ...
PROJECTORS = {"+": (+1, "X"), "-": (-1, "X"), "r": (+1, "Y"), "l": (-1, "Y"), "0": (+1, "Z"), "1": (-1, "Z")}
ALPHABET = "I" + PAULIS + "".join(PROJECTORS)


class SparseObservable:
    def __init__(self, num_qubits: int, terms=()):
        self._n = int(num_qubits)
        self._terms: list[tuple[complex, dict]] = []
        for coeff, ops in terms:
            self._add(coeff, ops)

    def _add(self, coeff, ops: dict):
        clean = {}
        for q, ch in ops.items():
            q = int(q)
            if not 0 <= q < self._n:
                raise QiskitError(f"qubit index {q} out of range for {self._n} qubits")
            if ch not in ALPHABET:
                raise QiskitError(f"letter {ch!r} is not in the SparseObservable alphabet {ALPHABET!r}")
            if ch != "I":
                clean[q] = ch
        self._terms.append((complex(coeff), clean))

    # ---- constructors (same names and argument meaning as the reference's) ----
    @classmethod
    def from_label(cls, label: str):
        return cls.from_list([(label, 1.0)])

    @classmethod
    def from_list(cls, terms, num_qubits=None):
        """[(dense label, coeff)]; the label's RIGHT-most letter is qubit 0 (mod.rs:3010-3060)."""
        terms = list(terms)
        if not terms:
            if num_qubits is None:
                raise QiskitError("cannot construct an observable from an empty list without knowing `num_qubits`")
            return cls(num_qubits)
        n = len(terms[0][0]) if num_qubits is None else int(num_qubits)
        out = cls(n)
        for label, coeff in terms:
            if len(label) != n:
                raise QiskitError(f"label with length {len(label)} cannot be added to a {n}-qubit operator")
            out._add(coeff, {n - 1 - i: ch for i, ch in enumerate(label)})
        return out

    @classmethod
    def from_sparse_list(cls, terms, num_qubits: int):
        """[(letters, qubit indices, coeff)] (mod.rs:3076-3120); a qubit may appear once per term."""
        out = cls(num_qubits)
        for letters, qubits, coeff in terms:
            qubits = [int(q) for q in qubits]
            if len(letters) != len(qubits):
                raise QiskitError(f"label with length {len(letters)} does not match indices of length {len(qubits)}")
            if len(set(qubits)) != len(qubits):
                raise QiskitError("index set contains duplicates")
            out._add(coeff, dict(zip(qubits, letters)))
        return out

    @classmethod
    def from_sparse_pauli_op(cls, op):
        n = op.num_qubits
        out = cls(n)
        for z, x, c in zip(op.paulis.z, op.paulis.x, op.coeffs):
            ops = {}
            for q in range(n):
                if x[q] and z[q]:
                    ops[q] = "Y"
                elif x[q]:
                    ops[q] = "X"
                elif z[q]:
                    ops[q] = "Z"
            out._add(c, ops)
        return out

    @classmethod
    def from_any(cls, obj):
        """This class, or any object with the reference's ``to_sparse_list()`` / ``num_qubits``."""
        if isinstance(obj, cls):
            return obj
        return cls.from_sparse_list(obj.to_sparse_list(), obj.num_qubits)

    # ---- views ----
    @property
    def num_qubits(self):
        return self._n

    @property
    def coeffs(self):
        return np.array([c for c, _ in self._terms], dtype=complex)

    def __len__(self):
        return len(self._terms)

    def to_sparse_list(self):
        """[(letters, qubits ascending, coeff)] (mod.rs:3150-3190)."""
        return [("".join(ops[q] for q in sorted(ops)), sorted(ops), c) for c, ops in self._terms]

    def __repr__(self):
        body = " + ".join(f"({c})(" + " ".join(f"{ops[q]}_{q}" for q in sorted(ops, reverse=True)) + ")"
                          for c, ops in self._terms) or "0.0"
        return f"<SparseObservable with {len(self)} terms on {self._n} qubits: {body}>"

    # ---- the arithmetic the device path needs ----
    def as_paulis(self) -> "SparseObservable":
        """Every projector expanded into 1/2 (I +- P) (mod.rs:1073-1130, table :207-219)."""
        out = SparseObservable(self._n)
        for coeff, ops in self._terms:
            partial = [(coeff, {})]
            for q in sorted(ops):
                ch = ops[q]
                if ch in PAULIS:
                    for _, d in partial:
                        d[q] = ch
                else:
                    sign, pauli = PROJECTORS[ch]
                    nxt = []
                    for c, d in partial:
                        nxt.append((0.5 * c, dict(d)))
                        with_p = dict(d)
                        with_p[q] = pauli
                        nxt.append((0.5 * sign * c, with_p))
                    partial = nxt
            out._terms.extend(partial)
        return out

    def simplify(self, tol: float = 1e-8) -> "SparseObservable":
        """Merge terms with equal operator strings, drop |coeff| <= tol (mod.rs `canonicalize`)."""
        merged: dict = {}
        for coeff, ops in self._terms:
            key = tuple(sorted(ops.items()))
            merged[key] = merged.get(key, 0j) + coeff
        out = SparseObservable(self._n)
        for key in sorted(merged):
            if abs(merged[key]) > tol:
                out._terms.append((merged[key], dict(key)))
        return out

    def to_sparse_pauli_op(self, tol: float = 0.0) -> SparsePauliOp:
        """Pauli-only form as the operator type the engine's expectation-value entry takes; equal
        strings are merged so a projector-heavy observable does not evaluate the same term twice."""
        pauli_only = self.as_paulis().simplify(tol)
        if len(pauli_only) == 0:
            return SparsePauliOp(["I" * self._n], [0.0])
        return SparsePauliOp.from_sparse_list(pauli_only.to_sparse_list(), self._n)

    def to_matrix(self) -> np.ndarray:
        """Dense matrix (tests only: 4^n entries)."""
        one = {
            "I": np.eye(2), "X": np.array([[0, 1], [1, 0]]), "Y": np.array([[0, -1j], [1j, 0]]),
            "Z": np.diag([1, -1]), "0": np.diag([1, 0]), "1": np.diag([0, 1]),
            "+": np.array([[1, 1], [1, 1]]) / 2, "-": np.array([[1, -1], [-1, 1]]) / 2,
            "r": np.array([[1, -1j], [1j, 1]]) / 2, "l": np.array([[1, 1j], [-1j, 1]]) / 2,
        }
        total = np.zeros((1 << self._n, 1 << self._n), dtype=complex)
        for coeff, ops in self._terms:
            mat = np.array([[1.0 + 0j]])
            for q in range(self._n):                     # qubit 0 is the least significant index bit
                mat = np.kron(one[ops.get(q, "I")], mat)
            total += coeff * mat
        return total
