"""Lower circuits / instructions to the engine's flat gate stream.

Host-side mirror of ``Statevector._evolve_instruction`` (statevector.py:965-1041)
and ``Operator._instruction_to_matrix`` (operators/operator.py:788-809): at every
nesting level try ``to_matrix()`` first, then the Reset / Barrier / Initialize
special cases, then recurse into ``definition`` mapping qubits and multiplying by
the definition's global phase.  Instead of touching the state per gate, matrix
leaves are appended to a ``GateStream`` that reaches the device in one
``qsv_apply_gates`` call; Reset / Initialize flush it first.

Duck-typed: reads ``obj.to_matrix()``, ``obj.definition`` (iterable of items with
``.operation/.qubits/.clbits``), ``definition.qubits``, ``definition.global_phase``,
``obj.name/.num_qubits/.params`` -- real Qiskit objects work unchanged.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .exceptions import QiskitError

# composite gates whose to_matrix() would build a huge dense matrix are unrolled through
# their definition instead (SURVEY.md 7.3 hard part 7)
MATRIX_MAX_QUBITS = 6


class GateStream:
    """Growable list of (kind, qargs, data) ready to be handed to qsv_apply_gates."""

    def __init__(self):
        self.kinds: list[int] = []
        self.qargs: list[tuple] = []
        self.data: list[np.ndarray | None] = []
        self.params: list[float] = []

    def __len__(self):
        return len(self.kinds)

    def matrix(self, mat, qargs):
        mat = _lib.as_c128(mat)
        k = len(qargs)
        if mat.shape != (1 << k, 1 << k):
            raise QiskitError(
                f"Operator input dimensions {mat.shape} are not equal to statevector subsystem "
                f"dimensions for qargs {list(qargs)}."
            )
        self.kinds.append(_lib.GATE_MATRIX)
        self.qargs.append(tuple(int(q) for q in qargs))
        self.data.append(mat)
        self.params.append(0.0)

    def diagonal(self, diag, qargs):
        diag = _lib.as_c128(diag).reshape(-1)
        if diag.shape != (1 << len(qargs),):
            raise QiskitError("diagonal length does not match the number of qargs")
        self.kinds.append(_lib.GATE_DIAGONAL)
        self.qargs.append(tuple(int(q) for q in qargs))
        self.data.append(diag)
        self.params.append(0.0)

    def phase(self, theta: float):
        self.kinds.append(_lib.GATE_PHASE)
        self.qargs.append(())
        self.data.append(None)
        self.params.append(float(theta))

    def to_c(self):
        n = len(self)
        arr = (_lib.QsvGate * max(n, 1))()
        for i in range(n):
            g = arr[i]
            g.kind = self.kinds[i]
            qa = self.qargs[i]
            g.k = len(qa)
            if len(qa) > _lib.MAX_GATE_QUBITS:
                raise QiskitError(
                    f"gate on {len(qa)} qubits exceeds the engine limit of {_lib.MAX_GATE_QUBITS}"
                )
            for j, q in enumerate(qa):
                g.qargs[j] = q
            d = self.data[i]
            g.data = d.ctypes.data if d is not None else None
            g.params[0] = self.params[i]
        return arr, n

    def clear(self):
        self.kinds.clear()
        self.qargs.clear()
        self.data.clear()
        self.params.clear()


def instruction_to_matrix(obj):
    """``Operator._instruction_to_matrix``: the matrix, or None when the object has
    none (QiskitError / CircuitError / TypeError swallowed, operator.py:797-809)."""
    to_matrix = getattr(obj, "to_matrix", None)
    if to_matrix is None:
        return None
    if getattr(obj, "num_qubits", 0) > MATRIX_MAX_QUBITS and getattr(obj, "definition", None) is not None:
        return None
    try:
        mat = to_matrix()
    except Exception:  # noqa: BLE001 - the reference swallows its own error types here
        return None
    if mat is None:
        return None
    return np.asarray(mat, dtype=np.complex128)


def _is_kind(obj, name: str) -> bool:
    """Reset / Barrier / Initialize are recognised by class name or instruction name, which
    covers both this package's classes and Qiskit's."""
    return type(obj).__name__.lower() == name or getattr(obj, "name", None) == name


def walk(state, obj, qargs, stream: GateStream):
    """Recursive walk; ``state`` is the device Statevector being updated (needed only by
    Reset / Initialize)."""
    mat = instruction_to_matrix(obj)
    if mat is not None:
        qa = list(range(obj.num_qubits)) if qargs is None else list(qargs)
        stream.matrix(mat, qa)
        return
    if _is_kind(obj, "reset"):
        state._flush(stream)
        state._reset_inplace(qargs)
        return
    if _is_kind(obj, "barrier"):
        return
    if _is_kind(obj, "initialize"):
        state._flush(stream)
        state._initialize_inplace(obj, qargs)
        return
    definition = getattr(obj, "definition", None)
    if definition is None:
        raise QiskitError(f"Cannot apply Instruction: {obj.name}")
    if not hasattr(definition, "qubits") or not hasattr(definition, "__iter__"):
        raise QiskitError(
            f"{obj.name} instruction definition is {type(definition)}; expected QuantumCircuit"
        )
    gphase = getattr(definition, "global_phase", 0)
    if gphase:
        stream.phase(float(gphase))
    index_of = {qubit: i for i, qubit in enumerate(definition.qubits)}
    for instruction in definition:
        if instruction.clbits:
            raise QiskitError(
                f"Cannot apply instruction with classical bits: {instruction.operation.name}"
            )
        if qargs is None:
            new_qargs = [index_of[q] for q in instruction.qubits]
        else:
            new_qargs = [qargs[index_of[q]] for q in instruction.qubits]
        walk(state, instruction.operation, new_qargs, stream)


def gate_list_to_stream(gates) -> GateStream:
    """Qiskit-free entry: ``gates`` is a list of ``(matrix, qargs)``, ``("phase", theta)``
    or ``("diag", diagonal, qargs)`` tuples (the format the oracle and benchmarks use)."""
    st = GateStream()
    for g in gates:
        if isinstance(g[0], str):
            if g[0] == "phase":
                st.phase(g[1])
            elif g[0] == "diag":
                st.diagonal(g[1], g[2])
            else:
                raise QiskitError(f"unknown gate-list entry {g[0]!r}")
        else:
            st.matrix(g[0], g[1])
    return st
