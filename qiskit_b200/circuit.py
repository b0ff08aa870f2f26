"""A small circuit container with the attribute surface the statevector path reads.

Qiskit's ``QuantumCircuit`` is the *input format* of the hot path, not part of it
(SURVEY.md 2.1): ``Statevector._evolve_instruction`` only iterates
``definition`` and calls ``to_matrix()`` (statevector.py:965-1041).  Real Qiskit
cannot be imported in the build image (no Rust toolchain for ``qiskit._accelerate``),
so this module supplies just enough of that surface -- ``Instruction`` /
``Gate`` with ``name, num_qubits, params, definition, to_matrix()``, and
``QuantumCircuit`` with ``data, qubits, cregs, parameters, global_phase,
assign_parameters, to_instruction, remove_final_measurements`` -- for the
primitives and the tests.  The walker in ``lowering.py`` is duck-typed and takes
real Qiskit objects equally.
"""
from __future__ import annotations

import copy
import re
from typing import Iterable, Sequence

import numpy as np

from .exceptions import QiskitError
from .gates import GATES, gate_matrix


class Parameter:
    """Symbolic circuit parameter (name only; no expressions)."""

    __slots__ = ("name", "_vector", "_index")

    def __init__(self, name: str, _vector=None, _index=None):
        self.name = name
        self._vector = _vector
        self._index = _index

    def __repr__(self):
        return f"Parameter({self.name})"

    def sort_key(self):
        # QuantumCircuit.parameters order: by name, ParameterVector elements numerically
        # (qiskit/circuit/quantumcircuit.py:4846-4851)
        if self._vector is not None:
            return (self._vector.name, self._index)
        m = re.fullmatch(r"(.*)\[(\d+)\]", self.name)
        if m:
            return (m.group(1), int(m.group(2)))
        return (self.name, -1)


class ParameterVector:
    def __init__(self, name: str, length: int):
        self.name = name
        self.params = [Parameter(f"{name}[{i}]", self, i) for i in range(length)]

    def __getitem__(self, i):
        return self.params[i]

    def __len__(self):
        return len(self.params)

    def __iter__(self):
        return iter(self.params)


class Instruction:
    """Mirror of ``qiskit.circuit.Instruction`` as read by the path."""

    def __init__(self, name, num_qubits, num_clbits=0, params=()):
        self.name = name
        self.num_qubits = int(num_qubits)
        self.num_clbits = int(num_clbits)
        self.params = list(params)
        self.definition = None

    def is_parameterized(self):
        return any(isinstance(p, Parameter) for p in self.params)

    def __copy__(self):
        # assign_parameters copies every operation of the circuit once per parameter set (host time on the
        # Estimator's critical path): plain attribute copy instead of copy.copy's __reduce_ex__ round trip
        new = object.__new__(type(self))
        new.__dict__.update(self.__dict__)
        return new

    def to_matrix(self):
        raise QiskitError(f"to_matrix not defined for {self.name}")

    def is_control_flow(self):
        return False

    def __repr__(self):
        return f"Instruction({self.name}, {self.num_qubits}, params={self.params})"


class Gate(Instruction):
    """Standard gate by name, or an explicit unitary (``UnitaryGate``)."""

    def __init__(self, name, num_qubits, params=(), matrix=None):
        super().__init__(name, num_qubits, 0, params)
        self._matrix = None if matrix is None else np.asarray(matrix, dtype=np.complex128)

    def to_matrix(self):
        if self._matrix is not None:
            return self._matrix
        if self.name in GATES:
            if self.is_parameterized():
                raise QiskitError(f"Cannot bind matrix of parameterised gate {self.name}")
            return gate_matrix(self.name, self.params)
        raise QiskitError(f"to_matrix not defined for {self.name}")


class Measure(Instruction):
    def __init__(self):
        super().__init__("measure", 1, 1)


class Reset(Instruction):
    def __init__(self):
        super().__init__("reset", 1, 0)


class Barrier(Instruction):
    def __init__(self, num_qubits):
        super().__init__("barrier", num_qubits, 0)


class Initialize(Instruction):
    """``qiskit.circuit.library.Initialize`` (statevector.py:989-1014): params are a
    label string per qubit, a single integer, or the full amplitude vector."""

    def __init__(self, params, num_qubits):
        super().__init__("initialize", num_qubits, 0, params)


class ControlFlowOp(Instruction):
    def is_control_flow(self):
        return True


class ClassicalRegister:
    def __init__(self, size: int, name: str = "c"):
        self.size, self.name = int(size), name

    def __len__(self):
        return self.size


class CircuitInstruction:
    __slots__ = ("operation", "qubits", "clbits")

    def __init__(self, operation, qubits=(), clbits=()):
        self.operation = operation
        self.qubits = tuple(qubits)
        self.clbits = tuple(clbits)

    @property
    def name(self):
        return self.operation.name

    def is_control_flow(self):
        return self.operation.is_control_flow()


class BitLocation:
    def __init__(self, index, registers):
        self.index, self.registers = index, registers


class QuantumCircuit:
    """Qubits and clbits are plain ints (their index)."""

    def __init__(self, num_qubits: int, *cregs, name: str = "circuit", global_phase: float = 0.0,
                 metadata=None):
        self.num_qubits = int(num_qubits)
        self.name = name
        self.global_phase = global_phase
        self.metadata = {} if metadata is None else metadata
        self.data: list[CircuitInstruction] = []
        self.cregs: list[ClassicalRegister] = []
        for c in cregs:
            if isinstance(c, int):
                if c > 0:
                    self.add_register(ClassicalRegister(c, "c"))
            else:
                self.add_register(c)

    # -- registers ----------------------------------------------------------
    @property
    def qubits(self):
        return list(range(self.num_qubits))

    @property
    def qregs(self):
        return [None]  # one flat register

    @property
    def num_clbits(self):
        return sum(c.size for c in self.cregs)

    @property
    def clbits(self):
        return list(range(self.num_clbits))

    def add_register(self, creg: ClassicalRegister):
        if any(c.name == creg.name for c in self.cregs):
            raise QiskitError(f"register name {creg.name!r} already exists")
        self.cregs.append(creg)

    def find_bit(self, bit: int) -> BitLocation:
        """Clbit index -> (index, [(register, offset)]) like ``QuantumCircuit.find_bit``."""
        off = 0
        for c in self.cregs:
            if off <= bit < off + c.size:
                return BitLocation(bit, [(c, bit - off)])
            off += c.size
        return BitLocation(bit, [])

    # -- iteration ----------------------------------------------------------
    def __iter__(self):
        return iter(self.data)

    def __len__(self):
        return len(self.data)

    def __getitem__(self, item):
        return self.data[item]

    # -- building -----------------------------------------------------------
    def append(self, operation, qargs: Sequence[int] = (), cargs: Sequence[int] = ()):
        if isinstance(operation, QuantumCircuit):
            operation = operation.to_instruction()
        qargs = [int(q) for q in qargs]
        if len(qargs) != operation.num_qubits:
            raise QiskitError(
                f"The amount of qubit arguments {len(qargs)} does not match the instruction "
                f"expectation ({operation.num_qubits})."
            )
        for q in qargs:
            if not 0 <= q < self.num_qubits:
                raise QiskitError(f"qubit index {q} out of range")
        if len(set(qargs)) != len(qargs):
            raise QiskitError("duplicate qubit arguments")
        self.data.append(CircuitInstruction(operation, qargs, [int(c) for c in cargs]))
        return self

    def _std(self, name, params, qubits):
        self.append(Gate(name, GATES[name].num_qubits, params), qubits)
        return self

    def unitary(self, matrix, qubits, label="unitary"):
        qubits = [qubits] if isinstance(qubits, int) else list(qubits)
        return self.append(Gate(label, len(qubits), [], matrix=matrix), qubits)

    def measure(self, qubit, clbit):
        qs = [qubit] if isinstance(qubit, int) else list(qubit)
        cs = [clbit] if isinstance(clbit, int) else list(clbit)
        for q, c in zip(qs, cs):
            self.append(Measure(), [q], [c])
        return self

    def measure_all(self):
        if not any(c.name == "meas" for c in self.cregs):
            self.add_register(ClassicalRegister(self.num_qubits, "meas"))
        off = 0
        for c in self.cregs:
            if c.name == "meas":
                break
            off += c.size
        self.barrier()
        for q in range(self.num_qubits):
            self.append(Measure(), [q], [off + q])
        return self

    def barrier(self, *qubits):
        qs = list(qubits) if qubits else list(range(self.num_qubits))
        return self.append(Barrier(len(qs)), qs)

    def reset(self, qubit):
        return self.append(Reset(), [qubit])

    def initialize(self, params, qubits=None):
        qs = list(range(self.num_qubits)) if qubits is None else (
            [qubits] if isinstance(qubits, int) else list(qubits))
        if isinstance(params, str):
            params = list(params)
        elif isinstance(params, (int, np.integer)):
            params = [complex(params)]
        return self.append(Initialize(list(params), len(qs)), qs)

    # -- parameters ---------------------------------------------------------
    def _walk_params(self, seen, out):
        for inst in self.data:
            for p in inst.operation.params:
                if isinstance(p, Parameter) and id(p) not in seen:
                    seen.add(id(p))
                    out.append(p)
            d = inst.operation.definition
            if isinstance(d, QuantumCircuit):
                d._walk_params(seen, out)

    @property
    def parameters(self):
        out: list[Parameter] = []
        self._walk_params(set(), out)
        return sorted(out, key=Parameter.sort_key)

    @property
    def num_parameters(self):
        return len(self.parameters)

    def assign_parameters(self, values, inplace=False):
        """Bind every parameter: a sequence (in ``parameters`` order) or a dict."""
        params = self.parameters
        if isinstance(values, dict):
            mapping = {}
            for key, val in values.items():
                if isinstance(key, ParameterVector):
                    for p, v in zip(key.params, val):
                        mapping[id(p)] = float(v)
                else:
                    mapping[id(key)] = float(val)
        else:
            values = list(np.asarray(values, dtype=float).ravel())
            if len(values) != len(params):
                raise ValueError(
                    f"Mismatching number of values and parameters. For partial binding please "
                    f"pass a dictionary of {{parameter: value}} pairs ({len(values)} != {len(params)})."
                )
            mapping = {id(p): v for p, v in zip(params, values)}
        target = self if inplace else self.copy_shallow()
        target._bind(mapping)
        return None if inplace else target

    def copy_shallow(self):
        new = copy.copy(self)
        new.data = [CircuitInstruction(i.operation.__copy__(), i.qubits, i.clbits) for i in self.data]
        for inst in new.data:
            inst.operation.params = list(inst.operation.params)
        new.cregs = list(self.cregs)
        return new

    def _bind(self, mapping):
        for inst in self.data:
            op = inst.operation
            op.params = [mapping.get(id(p), p) if isinstance(p, Parameter) else p for p in op.params]
            if isinstance(op.definition, QuantumCircuit):
                op.definition = op.definition.copy_shallow()
                op.definition._bind(mapping)

    def copy(self):
        return copy.deepcopy(self)

    # -- conversions ----------------------------------------------------------
    def to_instruction(self):
        inst = Instruction(self.name, self.num_qubits, self.num_clbits, [])
        inst.definition = self
        return inst

    def to_gate(self):
        g = Gate(self.name, self.num_qubits, [])
        g.definition = self
        return g

    def remove_final_measurements(self, inplace=False):
        """Strip measurements (and barriers) that no later operation follows."""
        circ = self if inplace else self.copy_shallow()
        busy: set[int] = set()
        keep = []
        for inst in reversed(circ.data):
            if inst.name in ("measure", "barrier") and not (set(inst.qubits) & busy):
                continue
            busy.update(inst.qubits)
            keep.append(inst)
        circ.data = keep[::-1]
        circ.cregs = []
        return None if inplace else circ


def _add_gate_methods():
    def make(name, gd):
        def method(self, *args):
            params = args[: gd.num_params]
            qubits = args[gd.num_params:]
            if len(qubits) == 1 and not isinstance(qubits[0], (int, np.integer)):
                # broadcast a 1-qubit gate over an iterable of qubits, e.g. qc.h(range(3))
                for q in qubits[0]:
                    self._std(name, params, [q])
                return self
            return self._std(name, params, list(qubits))

        method.__name__ = name
        method.__doc__ = f"Append a ``{name}`` gate: ({', '.join(['param'] * gd.num_params + ['qubit'] * gd.num_qubits)})."
        return method

    for name, gd in GATES.items():
        setattr(QuantumCircuit, name, make(name, gd))


_add_gate_methods()
