"""Stand-ins for the reference classes the adapter (qiskit_b200/adapter.py) derives from and reuses,
shaped like the real thing where the adapter can tell the difference:

* circuits hold ``Qubit`` / ``Clbit`` OBJECTS (not ints) in ``CircuitInstruction.qubits / .clbits``;
  ``find_bit`` maps them back; ``definition.qubits`` is a list of those objects
  (qiskit/circuit/quantumcircuit.py, statevector.py:1027-1040 builds its index map from it);
* ``circuit.parameters`` is sorted by name with ``ParameterVector`` elements in numeric order
  (quantumcircuit.py:4846-4851), bare value arrays bind in that order (estimator_pub.py:150-154);
* ``bound_circuit_to_instruction`` wraps the circuit as the ``definition`` of a matrix-less
  ``Instruction`` (primitives/utils.py:43-70);
* ``EstimatorPub.coerce`` yields ``observables`` as an object array of ``{label: coeff}`` dicts and
  ``parameter_values`` as a ``BindingsArray`` with ``bind_all`` (containers/estimator_pub.py:107-168,
  bindings_array.py:241-253, observables_array.py:103-125);
* ``PrimitiveJob`` runs the function on a worker thread (primitive_job.py:50-56).

TEST INFRASTRUCTURE: nothing under qiskit_b200/ imports this file.
"""
from __future__ import annotations

import copy
from concurrent.futures import ThreadPoolExecutor
from types import SimpleNamespace

import numpy as np

from qiskit_b200.circuit import Parameter, ParameterVector
from qiskit_b200.gates import GATES, gate_matrix


class QiskitError(Exception):
    pass


class Qubit:
    __slots__ = ()


class Clbit:
    __slots__ = ()


class Instruction:
    def __init__(self, name, num_qubits, num_clbits=0, params=()):
        self.name, self.num_qubits, self.num_clbits = name, num_qubits, num_clbits
        self.params = list(params)
        self.definition = None

    def is_control_flow(self):
        return False


class Gate(Instruction):
    def __init__(self, name, num_qubits, params=(), matrix=None):
        super().__init__(name, num_qubits, 0, params)
        self._matrix = matrix

    def to_matrix(self):
        if self._matrix is not None:
            return np.asarray(self._matrix, dtype=complex)
        if any(isinstance(p, Parameter) for p in self.params):
            raise TypeError("unbound parameter")
        return gate_matrix(self.name, [float(p) for p in self.params])


class CircuitInstruction:
    def __init__(self, operation, qubits=(), clbits=()):
        self.operation, self.qubits, self.clbits = operation, tuple(qubits), tuple(clbits)

    @property
    def name(self):
        return self.operation.name

    def is_control_flow(self):
        return self.operation.is_control_flow()


class ClassicalRegister:
    def __init__(self, size, name="c"):
        self.size, self.name = size, name
        self.bits = [Clbit() for _ in range(size)]

    def __len__(self):
        return self.size

    def __hash__(self):
        return hash((self.name, self.size))

    def __eq__(self, other):
        return isinstance(other, ClassicalRegister) and (self.name, self.size) == (other.name, other.size)


class QuantumCircuit:
    def __init__(self, num_qubits, *cregs, global_phase=0.0, metadata=None):
        self.qubits = [Qubit() for _ in range(num_qubits)]
        self.cregs = []
        self.clbits = []
        self.data = []
        self.global_phase = global_phase
        self.metadata = metadata
        self.name = "circuit"
        for c in cregs:
            self.add_register(c if isinstance(c, ClassicalRegister) else ClassicalRegister(int(c)))

    num_qubits = property(lambda self: len(self.qubits))
    num_clbits = property(lambda self: len(self.clbits))

    def add_register(self, creg):
        self.cregs.append(creg)
        self.clbits.extend(creg.bits)

    def find_bit(self, bit):
        if isinstance(bit, Qubit):
            return SimpleNamespace(index=next(i for i, q in enumerate(self.qubits) if q is bit), registers=[])
        idx = next(i for i, c in enumerate(self.clbits) if c is bit)
        regs = [(creg, creg.bits.index(bit)) for creg in self.cregs if any(b is bit for b in creg.bits)]
        return SimpleNamespace(index=idx, registers=regs)

    def __iter__(self):
        return iter(self.data)

    def __len__(self):
        return len(self.data)

    def __getitem__(self, item):
        return self.data[item]

    def append(self, operation, qargs=(), cargs=()):
        self.data.append(CircuitInstruction(operation, [self.qubits[int(q)] for q in qargs],
                                            [self.clbits[int(c)] for c in cargs]))
        return self

    def _std(self, name, params, qubits):
        return self.append(Gate(name, len(qubits), params), qubits)

    def h(self, q): return self._std("h", [], [q])
    def x(self, q): return self._std("x", [], [q])
    def ry(self, t, q): return self._std("ry", [t], [q])
    def rz(self, t, q): return self._std("rz", [t], [q])
    def rx(self, t, q): return self._std("rx", [t], [q])
    def cx(self, a, b): return self._std("cx", [], [a, b])
    def cp(self, t, a, b): return self._std("cp", [t], [a, b])

    def unitary(self, matrix, qubits):
        return self.append(Gate("unitary", len(qubits), [], matrix=np.asarray(matrix, dtype=complex)), qubits)

    def reset(self, q):
        return self.append(Instruction("reset", 1), [q])

    def measure(self, q, c):
        return self.append(Instruction("measure", 1, 1), [q], [c])

    def measure_all(self):
        creg = ClassicalRegister(self.num_qubits, "meas")
        self.add_register(creg)
        base = self.num_clbits - self.num_qubits
        for q in range(self.num_qubits):
            self.measure(q, base + q)
        return self

    @property
    def parameters(self):
        seen = {}
        for inst in self.data:
            for p in inst.operation.params:
                if isinstance(p, Parameter):
                    seen[id(p)] = p
        return sorted(seen.values(), key=lambda p: p.sort_key())

    def assign_parameters(self, values):
        params = self.parameters
        if len(values) != len(params):
            raise ValueError("Mismatching number of values and parameters.")
        mapping = {id(p): float(v) for p, v in zip(params, values)}
        out = copy.copy(self)
        out.data = []
        for inst in self.data:
            op = inst.operation
            if any(isinstance(p, Parameter) for p in op.params):
                op = copy.copy(op)
                op.params = [mapping[id(p)] if isinstance(p, Parameter) else p for p in op.params]
            out.data.append(CircuitInstruction(op, inst.qubits, inst.clbits))
        return out

    def to_instruction(self):
        return bound_circuit_to_instruction(self)

    def remove_final_measurements(self, inplace=False):
        out = copy.copy(self)
        data = list(self.data)
        while data and data[-1].operation.name in ("measure", "barrier"):
            data.pop()
        out.data = data
        return out


def bound_circuit_to_instruction(circuit):
    inst = Instruction(circuit.name, circuit.num_qubits, circuit.num_clbits, [])
    inst.definition = circuit
    return inst


# ------------------------------------------------------------------ quantum_info ------------
class OpShape:
    def __init__(self, n):
        self._n = n

    @classmethod
    def auto(cls, shape=None, dims_l=None, num_qubits_r=None):
        n = int(round(np.log2(shape[0])))
        if (1 << n) != shape[0]:
            raise QiskitError("not a qubit dimension")
        return cls(n)

    shape = property(lambda self: (1 << self._n,))
    num_qubits = property(lambda self: self._n)

    def dims_l(self, qargs=None):
        return (2,) * (self._n if qargs is None else len(qargs))


class Operator:
    def __init__(self, data, input_dims=None, output_dims=None):
        self.data = np.asarray(data, dtype=complex)
        self._dims = tuple(input_dims) if input_dims is not None else (2,) * int(round(np.log2(self.data.shape[-1])))

    dim = property(lambda self: (self.data.shape[1] if self.data.ndim == 2 else 1, self.data.shape[0]))

    def input_dims(self):
        return self._dims


class QuantumState:
    def __init__(self, op_shape=None):
        self._op_shape = op_shape
        self._rng_generator = None

    num_qubits = property(lambda self: self._op_shape.num_qubits)
    dim = property(lambda self: self._op_shape.shape[0])

    def dims(self, qargs=None):
        return self._op_shape.dims_l(qargs)

    @property
    def _rng(self):
        return np.random.default_rng() if self._rng_generator is None else self._rng_generator

    def seed(self, value=None):
        if value is None:
            self._rng_generator = None
        elif isinstance(value, np.random.Generator):
            self._rng_generator = value
        else:
            self._rng_generator = np.random.default_rng(value)

    def copy(self):
        return copy.deepcopy(self)

    def sample_counts(self, shots, qargs=None):
        inds, counts = np.unique(self.sample_memory(shots, qargs), return_counts=True)
        return dict(zip(inds.tolist(), counts.tolist()))


class Statevector(QuantumState):
    """Host statevector of the fake: holds ``_data`` like the reference; nothing else is needed here."""

    def __init__(self, data, dims=None):
        self._data = np.asarray(data, dtype=complex)
        super().__init__(op_shape=OpShape.auto(shape=self._data.shape, dims_l=dims, num_qubits_r=0))

    data = property(lambda self: self._data)

    def __eq__(self, other):
        return isinstance(other, Statevector) and np.allclose(self.data, other.data)


class SparsePauliOp:
    """(labels, coeffs) -> paulis.z / paulis.x / coeffs as the engine reads them."""

    def __init__(self, labels, coeffs=None):
        from qiskit_b200.pauli import SparsePauliOp as _Mirror

        m = _Mirror(list(labels), None if coeffs is None else list(coeffs))
        self.paulis, self.coeffs = m.paulis, m.coeffs


# ------------------------------------------------------------------ primitives --------------
class BaseEstimatorV2:
    pass


class BaseSamplerV2:
    pass


class BindingsArray:
    def __init__(self, values):
        self._values = np.asarray(values, dtype=float)

    shape = property(lambda self: self._values.shape[:-1])

    def as_array(self, parameters=None):
        return self._values

    def bind_all(self, circuit):
        out = np.empty(self.shape, dtype=object)
        for idx in np.ndindex(*self.shape):
            out[idx] = circuit.assign_parameters(self._values[idx]) if self._values.shape[-1] else circuit
        return out


def _coerce_values(circuit, values):
    n = len(circuit.parameters)
    arr = np.zeros((0,)) if values is None else np.asarray(values, dtype=float)
    if arr.ndim == 0:
        arr = arr.reshape(1)
    if n == 0 and arr.size == 0:
        arr = np.zeros((0,))
    if arr.shape[-1] != n:
        raise ValueError(f"The number of values ({arr.shape[-1]}) does not match the number of parameters ({n}).")
    return BindingsArray(arr)


def _obs_to_dict(obs):
    if isinstance(obs, str):
        return {obs: 1.0}
    if isinstance(obs, dict):
        return dict(obs)
    if hasattr(obs, "to_list"):
        return {label: complex(c).real for label, c in obs.to_list()}
    raise TypeError(f"unsupported observable {obs!r}")


def _obs_array(observables):
    if isinstance(observables, (str, dict)) or hasattr(observables, "to_list"):
        out = np.empty((), dtype=object)
        out[()] = _obs_to_dict(observables)
        return out
    subs = [_obs_array(o) for o in observables]
    out = np.empty((len(subs),) + subs[0].shape, dtype=object)
    for i, sub in enumerate(subs):
        out[i] = sub[()] if sub.ndim == 0 else sub
    return out


class EstimatorPub:
    def __init__(self, circuit, observables, parameter_values, precision):
        self.circuit, self.observables, self.parameter_values, self.precision = (
            circuit, observables, parameter_values, precision)
        self.shape = np.broadcast_shapes(observables.shape, parameter_values.shape)

    @classmethod
    def coerce(cls, pub, precision=None):
        if precision is not None and precision < 0:
            raise ValueError("precision must be non-negative")
        circuit, observables = pub[0], pub[1]
        values = pub[2] if len(pub) > 2 else None
        if len(pub) > 3 and pub[3] is not None:
            precision = pub[3]
        obs = _obs_array(observables)
        for d in obs.reshape(-1):
            for label in d:
                if len(label) != circuit.num_qubits:
                    raise ValueError("observable width does not match the circuit")
        return cls(circuit, obs, _coerce_values(circuit, values), precision)


class SamplerPub:
    def __init__(self, circuit, parameter_values, shots):
        self.circuit, self.parameter_values, self.shots = circuit, parameter_values, shots
        self.shape = parameter_values.shape

    @classmethod
    def coerce(cls, pub, shots=None):
        if isinstance(pub, QuantumCircuit):
            pub = (pub,)
        circuit = pub[0]
        values = pub[1] if len(pub) > 1 else None
        if len(pub) > 2 and pub[2] is not None:
            shots = pub[2]
        if shots is None or shots < 0:
            raise ValueError("shots must be a non-negative integer")
        return cls(circuit, _coerce_values(circuit, values), int(shots))


class DataBin:
    def __init__(self, *, shape=(), **fields):
        self.shape = shape
        self.__dict__.update(fields)


class BitArray:
    def __init__(self, array, num_bits):
        self.array, self.num_bits = np.asarray(array, dtype=np.uint8), num_bits

    num_shots = property(lambda self: self.array.shape[-2])

    def get_int_counts(self, loc=None):
        arr = self.array if loc is None else self.array[loc]
        vals = np.zeros(arr.shape[:-1], dtype=object)
        for b in range(arr.shape[-1]):
            vals = vals * 256 + arr[..., b].astype(object)
        keys, counts = np.unique(vals.reshape(-1).astype(np.int64), return_counts=True)
        return dict(zip(keys.tolist(), counts.tolist()))


class PubResult:
    def __init__(self, data, metadata=None):
        self.data, self.metadata = data, metadata or {}


class SamplerPubResult(PubResult):
    pass


class PrimitiveResult:
    def __init__(self, pub_results, metadata=None):
        self._results, self.metadata = list(pub_results), metadata or {}

    def __getitem__(self, i):
        return self._results[i]

    def __len__(self):
        return len(self._results)


class PrimitiveJob:
    _executor = ThreadPoolExecutor(max_workers=1)

    def __init__(self, function, *args):
        self._function, self._args, self._future = function, args, None

    def _submit(self):
        self._future = self._executor.submit(self._function, *self._args)

    def result(self):
        return self._future.result()


def _preprocess_circuit(circuit):
    """statevector_sampler.py:207-233 with this file's circuit class (final measurements only)."""
    from qiskit_b200.primitives import _MeasureInfo

    num_bits = {creg.name: creg.size for creg in circuit.cregs}
    mapping = {}
    active_q, active_c = set(range(circuit.num_qubits)), set(range(circuit.num_clbits))
    for item in circuit[::-1]:
        if item.name == "measure":
            loc = circuit.find_bit(item.clbits[0])
            q = circuit.find_bit(item.qubits[0]).index
            if loc.index in active_c and q in active_q:
                for creg in loc.registers:
                    mapping[creg] = q
                active_c.remove(loc.index)
        elif item.name not in ("barrier", "delay"):
            for qq in item.qubits:
                active_q.discard(circuit.find_bit(qq).index)
        if not active_c or not active_q:
            break
    qargs = sorted(set(mapping.values()))
    where = {v: k for k, v in enumerate(qargs)}
    stripped = circuit.remove_final_measurements(inplace=False)
    if any(i.name == "measure" for i in stripped):
        raise QiskitError("StatevectorSampler cannot handle mid-circuit measurements")
    sentinel = len(qargs)
    indices = {k: [sentinel] * v for k, v in num_bits.items()}
    for (creg, ind), q in mapping.items():
        indices[creg.name][ind] = where[q]
    info = [_MeasureInfo(creg_name=nm, num_bits=nb, num_bytes=(nb + 7) // 8, qreg_indices=indices[nm])
            for nm, nb in num_bits.items()]
    return stripped, qargs, info


def namespace():
    return SimpleNamespace(
        Statevector=Statevector, QuantumState=QuantumState, OpShape=OpShape, Operator=Operator,
        SparsePauliOp=SparsePauliOp, QuantumCircuit=QuantumCircuit, Instruction=Instruction,
        QiskitError=QiskitError, BaseEstimatorV2=BaseEstimatorV2, BaseSamplerV2=BaseSamplerV2,
        EstimatorPub=EstimatorPub, SamplerPub=SamplerPub, BitArray=BitArray, DataBin=DataBin,
        PubResult=PubResult, SamplerPubResult=SamplerPubResult, PrimitiveResult=PrimitiveResult,
        PrimitiveJob=PrimitiveJob, bound_circuit_to_instruction=bound_circuit_to_instruction,
        preprocess_circuit=_preprocess_circuit)
