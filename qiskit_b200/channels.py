"""Quantum channels as the device ``DensityMatrix`` consumes them (SURVEY.md 8f, N2 remainder).

``DensityMatrix.evolve(channel)`` in the reference goes through ``SuperOp._evolve``
(qiskit/quantum_info/operators/channel/superop.py:234-292): the superoperator matrix acts on the
column-stacked density matrix, ``vec(rho)[row + d col] = rho[row, col]``.  On the device a density
matrix of n qubits is the 2n-qubit handle with index ``(row << n) | col``, so a k-qubit ``SuperOp`` on
``qargs`` is ONE dense 2k-qubit matrix on the handle qubits ``[q + n for q in qargs] + qargs`` (matrix
index bits 0..k-1 = row bits, k..2k-1 = column bits): no new kernel.

``SuperOp`` here is the minimal mirror (qubit channels only): from a 4^k x 4^k matrix, a list of Kraus
operators (``sum_i conj(K_i) (x) K_i``, superop.py / transformations.py ``_kraus_to_superop``) or a unitary.
"""
from __future__ import annotations

import numpy as np

from .exceptions import QiskitError


class SuperOp:
    """Superoperator matrix of a k-qubit channel in the column-stacking convention."""

    def __init__(self, data):
        if isinstance(data, SuperOp):
            self._data = data._data.copy()
        else:
            mat = np.asarray(getattr(data, "data", data), dtype=complex)
            if mat.ndim != 2 or mat.shape[0] != mat.shape[1]:
                raise QiskitError("Invalid shape for SuperOp matrix.")
            k4 = int(round(np.log2(mat.shape[0])))
            if (1 << k4) != mat.shape[0] or k4 % 2:
                raise QiskitError("SuperOp on the device needs a 4^k x 4^k matrix (qubit channels only)")
            self._data = mat
        self._k = int(round(np.log2(self._data.shape[0]))) // 2

    @classmethod
    def from_kraus(cls, kraus):
        """``sum_i conj(K_i) (x) K_i``: vec(K rho K^dagger) = (conj(K) (x) K) vec(rho), column stacking."""
        ops = [np.asarray(k, dtype=complex) for k in kraus]
        if not ops or any(o.shape != ops[0].shape or o.ndim != 2 or o.shape[0] != o.shape[1] for o in ops):
            raise QiskitError("Kraus operators must be square matrices of one size")
        return cls(sum(np.kron(np.conj(o), o) for o in ops))

    @classmethod
    def from_unitary(cls, mat):
        return cls.from_kraus([mat])

    @property
    def data(self):
        return self._data

    @property
    def num_qubits(self):
        return self._k

    def input_dims(self):
        return (2,) * self._k

    def output_dims(self):
        return (2,) * self._k

    def compose(self, other):
        """``other`` after ``self`` (superop.py:170-182: matrix product of the superoperators)."""
        other = other if isinstance(other, SuperOp) else SuperOp(other)
        if other._k != self._k:
            raise QiskitError("SuperOp dimensions do not match")
        return SuperOp(other._data @ self._data)

    def _evolve(self, state, qargs=None):
        from .densitymatrix import DensityMatrix

        if not isinstance(state, DensityMatrix):
            state = DensityMatrix(state)
        return state.evolve(self, qargs)


def as_superop_matrix(obj):
    """The superoperator matrix of a channel-like object, or None: this module's ``SuperOp``, Qiskit's
    ``SuperOp``, any other Qiskit ``QuantumChannel`` (converted through ``qiskit.quantum_info.SuperOp``), or an
    object with ``to_quantumchannel()`` (densitymatrix.py:332-333)."""
    if hasattr(obj, "to_quantumchannel"):
        obj = obj.to_quantumchannel()
    if isinstance(obj, SuperOp):
        return obj.data
    names = {c.__name__ for c in type(obj).__mro__}
    if "SuperOp" in names:
        return np.asarray(obj.data, dtype=complex)
    if "QuantumChannel" in names:
        from qiskit.quantum_info import SuperOp as _RefSuperOp  # noqa: PLC0415 - only with a real Qiskit

        return np.asarray(_RefSuperOp(obj).data, dtype=complex)
    return None
