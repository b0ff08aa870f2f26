"""Qiskit-side adapter: the device engine behind the reference's OWN classes.

``make_adapter(ns)`` builds three classes from the reference types it is given:

* ``DeviceStatevector(ns.Statevector)`` -- holds an opaque engine handle instead of ``_data`` and
  overrides exactly what sits on the hot path: ``evolve`` and the static ``_evolve_operator`` /
  ``_evolve_instruction`` (statevector.py:437-478, 925, 965), ``_expectation_value_pauli`` /
  ``expectation_value`` (:533-587), ``probabilities`` (:589-673), ``sample_memory`` / ``reset`` /
  ``measure`` (quantum_state.py:236-331, statevector.py:675-719), ``copy`` semantics, ``_multiply``,
  ``conjugate``, ``inner``, ``trace``, ``purity``; ``.data`` / ``__array__`` / ``_data`` download lazily
  and refuse above ``DATA_MAX_QUBITS``.
* ``DeviceEstimator(ns.BaseEstimatorV2)`` / ``DeviceSampler(ns.BaseSamplerV2)`` -- the reference's
  ``run`` signatures (base_estimator.py:78-93, base_sampler.py:59-74); ``run`` reuses the reference's
  ``EstimatorPub.coerce`` / ``SamplerPub.coerce`` / ``BindingsArray.bind_all`` / ``DataBin`` /
  ``BitArray`` / ``PrimitiveJob``; only ``_run_pub`` differs from statevector_estimator.py:150-176 and
  statevector_sampler.py:177-205: it builds a ``DeviceStatevector`` where the reference builds a
  ``Statevector`` and packs sample INDICES with ``qsv_pack_samples`` instead of looping over strings.

``ns`` is a namespace of reference classes.  ``qiskit_namespace()`` imports them from a real Qiskit
(``ImportError`` when Qiskit is not installed: this build image has no Rust toolchain for
``qiskit._accelerate``); ``tests/fake_qiskit.py`` provides stand-ins with the reference's shapes
(``Qubit`` objects in ``CircuitInstruction.qubits``, ``definition.qubits``, ``BindingsArray``,
``ObservablesArray`` of ``{label: coeff}`` dicts, ``PrimitiveJob``) so the adapter is exercised on CPU-less
CI and on the GPU box alike.  ``install()`` rebinds Qiskit's entry points to the device classes.
"""
from __future__ import annotations

import copy as _copy
import os
from types import SimpleNamespace

import numpy as np

from . import _lib
from .lowering import GateStream, walk
from .statevector import Statevector as _DevSV

DATA_MAX_QUBITS = 31   # .data / __array__ download 16 * 2^n bytes: refuse beyond 32 GiB


def qiskit_namespace():
    """The reference classes the adapter derives from / reuses (needs a working ``import qiskit``)."""
    import qiskit  # noqa: F401,PLC0415 - optional dependency
    from qiskit.circuit import Instruction, QuantumCircuit
    from qiskit.exceptions import QiskitError
    from qiskit.primitives.base import BaseEstimatorV2, BaseSamplerV2
    from qiskit.primitives.containers import (BitArray, DataBin, PrimitiveResult, PubResult,
                                              SamplerPubResult)
    from qiskit.primitives.containers.estimator_pub import EstimatorPub
    from qiskit.primitives.containers.sampler_pub import SamplerPub
    from qiskit.primitives.primitive_job import PrimitiveJob
    from qiskit.primitives.statevector_sampler import _preprocess_circuit
    from qiskit.primitives.utils import bound_circuit_to_instruction
    from qiskit.quantum_info import Operator, SparsePauliOp, Statevector
    from qiskit.quantum_info.operators.op_shape import OpShape
    from qiskit.quantum_info.states.quantum_state import QuantumState

    return SimpleNamespace(
        Statevector=Statevector, QuantumState=QuantumState, OpShape=OpShape, Operator=Operator,
        SparsePauliOp=SparsePauliOp, QuantumCircuit=QuantumCircuit, Instruction=Instruction,
        QiskitError=QiskitError, BaseEstimatorV2=BaseEstimatorV2, BaseSamplerV2=BaseSamplerV2,
        EstimatorPub=EstimatorPub, SamplerPub=SamplerPub, BitArray=BitArray, DataBin=DataBin,
        PubResult=PubResult, SamplerPubResult=SamplerPubResult, PrimitiveResult=PrimitiveResult,
        PrimitiveJob=PrimitiveJob, bound_circuit_to_instruction=bound_circuit_to_instruction,
        preprocess_circuit=_preprocess_circuit)


_PREFIX_CACHE = os.environ.get("QSV_PREFIX_CACHE", "1") != "0"


def _has_stochastic_op(circuit) -> bool:
    """reset / initialize draw from the state's rng (statevector.py:989-1014): no state reuse then."""
    for inst in circuit:
        op = inst.operation
        if getattr(op, "name", None) in ("reset", "initialize"):
            return True
        d = getattr(op, "definition", None)
        if d is not None and hasattr(d, "__iter__") and _has_stochastic_op(d):
            return True
    return False


def make_adapter(ns, *, device: int = 0):
    """Device classes derived from the reference classes in ``ns`` (see the module docstring)."""
    default_device = device

    class DeviceStatevector(ns.Statevector):
        """``Statevector`` whose amplitudes live in HBM behind an engine handle."""

        def __init__(self, data, dims=None, *, device=None):
            device = default_device if device is None else device
            try:
                if isinstance(data, DeviceStatevector):
                    dev = _DevSV(data._dev)                      # device-to-device clone
                elif isinstance(data, _DevSV):
                    dev = data                                   # adopt an engine state
                elif isinstance(data, ns.Statevector):
                    if dims is None:
                        dims = data.dims()
                    dev = _DevSV(np.asarray(data.data, dtype=complex), device=device)
                elif isinstance(data, (list, np.ndarray)):
                    dev = _DevSV(np.asarray(data, dtype=complex), device=device)
                elif isinstance(data, ns.Operator):
                    input_dim, _ = data.dim
                    if input_dim != 1:
                        raise ns.QiskitError("Input Operator is not a column-vector.")
                    dev = _DevSV(np.ravel(data.data), device=device)
                elif isinstance(data, (ns.QuantumCircuit, ns.Instruction)):
                    inst = data.to_instruction() if isinstance(data, ns.QuantumCircuit) else data
                    dev = _DevSV.from_int(0, 1 << inst.num_qubits, device=device)
                    dev._evolve_instruction_inplace(inst, None)
                else:
                    raise ns.QiskitError("Invalid input data format for Statevector")
            except ns.QiskitError:
                raise
            except Exception as exc:  # engine errors keep the reference's exception type
                if type(exc).__name__ == "QiskitError":
                    raise ns.QiskitError(str(exc)) from exc
                raise
            if dims is not None and not isinstance(dims, (int, np.integer)) and any(int(d) != 2 for d in dims):
                raise ns.QiskitError("The device engine holds qubit states only (dims = (2,) * n)")
            self._dev = dev
            ns.QuantumState.__init__(
                self, op_shape=ns.OpShape.auto(shape=(1 << dev.num_qubits,), dims_l=dims, num_qubits_r=0))

        # ---- host views (lazy, explicit) ----------------------------------------------------
        @property
        def _data(self):
            if self._dev.num_qubits > DATA_MAX_QUBITS:
                raise ns.QiskitError(
                    f"refusing to copy a {self._dev.num_qubits}-qubit state to the host "
                    f"({16 << self._dev.num_qubits} bytes); use probabilities / expectation_value / sample_*")
            return self._dev.data

        @_data.setter
        def _data(self, value):
            # reference code paths that are not overridden assign a host array (e.g. `_add`)
            self._dev = _DevSV(np.asarray(value, dtype=complex), device=self._dev._device)

        @property
        def data(self):
            return self._data

        def __array__(self, dtype=None, copy=None):
            arr = self._data
            return arr if dtype is None else arr.astype(dtype)

        def __len__(self):
            return 1 << self._dev.num_qubits

        def __getitem__(self, key):
            return self._dev[key]

        def __eq__(self, other):
            if isinstance(other, DeviceStatevector):
                return self._dev == other._dev
            return ns.Statevector.__eq__(self, other)

        __hash__ = None

        def __repr__(self):
            return f"DeviceStatevector(<device {self._dev._device}, {self._dev.num_qubits} qubits>, dims={self.dims()})"

        # ---- copy semantics (quantum_state.py:74-76) ---------------------------
        def _wrap(self, dev):
            ret = object.__new__(type(self))
            ret._dev = dev
            ret._op_shape = _copy.copy(self._op_shape)
            ret._rng_generator = self._rng_generator
            return ret

        def __copy__(self):
            return self._wrap(_DevSV(self._dev))

        def __deepcopy__(self, memo):
            ret = self._wrap(_DevSV(self._dev))
            ret._rng_generator = _copy.deepcopy(self._rng_generator, memo)
            return ret

        def _sync_rng(self):
            self._dev._rng_generator = self._rng_generator

        # ---- evolution -----------------------------------------------------------------------
        def evolve(self, other, qargs=None):
            """statevector.py:437-478 (never mutates ``self``)."""
            if qargs is None:
                qargs = getattr(other, "qargs", None)
            ret = _copy.copy(self)
            if isinstance(other, ns.QuantumCircuit):
                other = other.to_instruction()
            if isinstance(other, ns.Instruction):
                return DeviceStatevector._evolve_instruction(ret, other, qargs=qargs)
            if not isinstance(other, ns.Operator):
                dims = self.dims(qargs=qargs)
                other = ns.Operator(other, input_dims=dims, output_dims=dims)
            if self.dims(qargs) != other.input_dims():
                raise ns.QiskitError(
                    "Operator input dimensions are not equal to statevector subsystem dimensions.")
            return DeviceStatevector._evolve_operator(ret, other, qargs=qargs)

        @staticmethod
        def _evolve_operator(statevec, oper, qargs=None):
            """statevector.py:925-963: psi <- (U on qargs) psi, in place on ``statevec``."""
            mat = np.asarray(getattr(oper, "data", oper), dtype=np.complex128)
            try:
                statevec._dev._evolve_operator_inplace(mat, None if qargs is None else [int(q) for q in qargs])
            except Exception as exc:
                if type(exc).__name__ == "QiskitError":
                    raise ns.QiskitError(str(exc)) from exc
                raise
            return statevec

        @staticmethod
        def _evolve_instruction(statevec, obj, qargs=None):
            """statevector.py:965-1041: the same order of checks (matrix, reset / barrier / initialize,
            definition + global phase, errors for clbits / missing definitions) in lowering.walk; the
            matrix leaves reach the device as ONE gate stream."""
            statevec._sync_rng()
            stream = GateStream()
            try:
                walk(statevec._dev, obj, None if qargs is None else [int(q) for q in qargs], stream)
                statevec._dev._flush(stream)
            except Exception as exc:
                if type(exc).__name__ == "QiskitError":
                    raise ns.QiskitError(str(exc)) from exc
                raise
            return statevec

        # ---- expectation values ----------------------------------------------------------------
        def _expectation_value_pauli(self, pauli, qargs=None):
            return self._dev._expectation_value_pauli(pauli, qargs)

        def expectation_value(self, oper, qargs=None):
            """statevector.py:564-587; every term of a SparsePauliOp in one engine call."""
            if hasattr(oper, "paulis") and hasattr(oper, "coeffs"):
                return self._dev.expectation_value(oper, qargs)
            if hasattr(oper, "z") and hasattr(oper, "x") and np.ndim(oper.z) == 1:
                return self._dev._expectation_value_pauli(oper, qargs)
            val = self.evolve(oper, qargs=qargs)
            return self._dev.inner(val._dev)

        # ---- probabilities, sampling, reset, measure --------------------------------------------------
        def probabilities(self, qargs=None, decimals=None):
            return self._dev.probabilities(None if qargs is None else [int(q) for q in qargs], decimals)

        def sample_memory(self, shots, qargs=None):
            self._sync_rng()
            return self._dev.sample_memory(shots, None if qargs is None else [int(q) for q in qargs])

        def sample_indices(self, shots, qargs=None):
            """Outcome integers instead of ket strings (what DeviceSampler packs)."""
            self._sync_rng()
            return self._dev.sample_indices(shots, None if qargs is None else [int(q) for q in qargs])

        def reset(self, qargs=None):
            self._sync_rng()
            return self._wrap(self._dev.reset(None if qargs is None else [int(q) for q in qargs]))

        def measure(self, qargs=None):
            self._sync_rng()
            outcome, dev = self._dev.measure(None if qargs is None else [int(q) for q in qargs])
            return outcome, self._wrap(dev)

        # ---- algebra that would otherwise round-trip through the host ----------------------------------
        def _multiply(self, other):
            return self._wrap(self._dev._multiply(other))

        def conjugate(self):
            return self._wrap(self._dev.conjugate())

        def inner(self, other):
            if not isinstance(other, DeviceStatevector):
                other = DeviceStatevector(other, device=self._dev._device)
            return self._dev.inner(other._dev)

        def trace(self):
            return self._dev.trace()

        def purity(self):
            return self._dev.purity()

        def is_valid(self, atol=None, rtol=None):
            return self._dev.is_valid(atol, rtol)

        @classmethod
        def from_int(cls, i, dims):
            size = int(np.prod(dims))
            return cls(_DevSV.from_int(int(i), size, device=default_device),
                       dims=None if isinstance(dims, (int, np.integer)) else dims)

        @classmethod
        def from_instruction(cls, instruction):
            return cls(instruction)

        def counters(self):
            return self._dev.counters()

    class DeviceEstimator(ns.BaseEstimatorV2):
        """statevector_estimator.py:114-176 on the device engine."""

        def __init__(self, *, default_precision: float = 0.0, seed=None):
            self._default_precision = default_precision
            self._seed = seed

        @property
        def default_precision(self):
            return self._default_precision

        @property
        def seed(self):
            return self._seed

        def run(self, pubs, *, precision=None):
            if precision is None:
                precision = self._default_precision
            coerced_pubs = [ns.EstimatorPub.coerce(pub, precision) for pub in pubs]
            job = ns.PrimitiveJob(self._run, coerced_pubs)
            job._submit()
            return job

        def _run(self, pubs):
            return ns.PrimitiveResult([self._run_pub(pub) for pub in pubs], metadata={"version": 2})

        def _run_pub(self, pub):
            rng = np.random.default_rng(self._seed)
            circuit, precision = pub.circuit, pub.precision
            bound_circuits = pub.parameter_values.bind_all(circuit)
            bc_circuits, bc_obs = np.broadcast_arrays(bound_circuits, pub.observables)
            evs = np.zeros_like(bc_circuits, dtype=np.float64)
            stds = np.zeros_like(bc_circuits, dtype=np.float64)
            # the reference re-simulates every broadcast index; without reset / initialize the state only
            # depends on the bound circuit, so consecutive indices that share it share the state
            reuse = not _has_stochastic_op(circuit)
            self.last_prefix_stats = None
            raw = self._shared_prefix_evs(pub, bound_circuits, bc_circuits, bc_obs) if reuse else None
            state, state_of = None, None
            for index in np.ndindex(*bc_circuits.shape):
                bound_circuit = bc_circuits[index]
                if raw is not None:
                    expectation_value = raw[index]
                    if precision != 0:
                        if not np.isreal(expectation_value):
                            raise ValueError("Given operator is not Hermitian and noise cannot be added.")
                        expectation_value = rng.normal(expectation_value, precision)
                    evs[index] = expectation_value
                    continue
                if not (reuse and state is not None and state_of is bound_circuit):
                    state = None     # hand the previous buffer back to the pool first
                    state = DeviceStatevector.from_int(0, 2 ** bound_circuit.num_qubits)
                    state.seed(rng)
                    DeviceStatevector._evolve_instruction(state, ns.bound_circuit_to_instruction(bound_circuit))
                    state_of = bound_circuit
                paulis, coeffs = zip(*bc_obs[index].items())
                obs = ns.SparsePauliOp(paulis, coeffs)
                expectation_value = np.real_if_close(state.expectation_value(obs))
                if precision != 0:
                    if not np.isreal(expectation_value):
                        raise ValueError("Given operator is not Hermitian and noise cannot be added.")
                    expectation_value = rng.normal(expectation_value, precision)
                evs[index] = expectation_value
            data = ns.DataBin(evs=evs, stds=stds, shape=evs.shape)
            return ns.PubResult(data, metadata={"target_precision": precision,
                                                "circuit_metadata": pub.circuit.metadata})

        def _shared_prefix_evs(self, pub, bound_circuits, bc_circuits, bc_obs):
            """Shared-prefix caching across the parameter sets of the pub (prefix_cache.py):
            {broadcast index: expectation value}, or None when the sets share nothing (or the bindings
            container cannot give its values as one array) and the serial loop below should run."""
            from . import prefix_cache as pc

            if not _PREFIX_CACHE or bound_circuits.size < 2:
                return None
            circuit = pub.circuit
            try:
                values = np.asarray(pub.parameter_values.as_array(list(circuit.parameters)), dtype=float)
                values = values.reshape(bound_circuits.size, -1)
                first_inst, total = pc.first_use(circuit)
            except (AttributeError, TypeError, ValueError):
                return None
            order, share = pc.order_sets(values, first_inst, total)
            if share.size == 0 or float(share.max()) < pc.MIN_SHARE:
                return None
            flat = list(bound_circuits.reshape(-1))
            by_circuit = {}
            for index in np.ndindex(*bc_circuits.shape):
                by_circuit.setdefault(id(bc_circuits[index]), []).append(index)
            n = circuit.num_qubits
            raw = {}

            def lower(sid):
                stream = GateStream()
                walk(None, ns.bound_circuit_to_instruction(flat[sid]), None, stream)
                return stream

            def consume(sid, dev):
                state = DeviceStatevector(dev)               # adopts the engine state, no copy
                for index in by_circuit.get(id(flat[sid]), ()):
                    paulis, coeffs = zip(*bc_obs[index].items())
                    raw[index] = np.real_if_close(state.expectation_value(ns.SparsePauliOp(paulis, coeffs)))

            stats = pc.run_sets(order, lower, lambda: _DevSV.from_int(0, 1 << n, device=default_device), consume,
                                lanes=1, can_checkpoint=lambda: pc._device_can_hold_checkpoint(n, default_device))
            self.last_prefix_stats = stats.as_dict()
            return raw

    class DeviceSampler(ns.BaseSamplerV2):
        """statevector_sampler.py:136-205 on the device engine."""

        def __init__(self, *, default_shots: int = 1024, seed=None):
            self._default_shots = default_shots
            self._seed = seed

        @property
        def default_shots(self):
            return self._default_shots

        @property
        def seed(self):
            return self._seed

        def run(self, pubs, *, shots=None):
            import warnings

            if shots is None:
                shots = self._default_shots
            coerced_pubs = [ns.SamplerPub.coerce(pub, shots) for pub in pubs]
            if any(len(pub.circuit.cregs) == 0 for pub in coerced_pubs):
                warnings.warn(
                    "One of your circuits has no output classical registers and so the result "
                    "will be empty. Did you mean to add measurement instructions?", UserWarning)
            job = ns.PrimitiveJob(self._run, coerced_pubs)
            job._submit()
            return job

        def _run(self, pubs):
            return ns.PrimitiveResult([self._run_pub(pub) for pub in pubs], metadata={"version": 2})

        def _run_pub(self, pub):
            from .primitives import pack_samples

            circuit, qargs, meas_info = ns.preprocess_circuit(pub.circuit)
            bound_circuits = pub.parameter_values.bind_all(circuit)
            arrays = {item.creg_name: np.zeros(bound_circuits.shape + (pub.shots, item.num_bytes), dtype=np.uint8)
                      for item in meas_info}
            for index, bound_circuit in np.ndenumerate(bound_circuits):
                final_state = DeviceStatevector(ns.bound_circuit_to_instruction(bound_circuit))
                final_state.seed(self._seed)
                if qargs:
                    samples = final_state.sample_indices(pub.shots, qargs)
                else:
                    samples = np.zeros(pub.shots, dtype=np.uint64)
                for item in meas_info:
                    # qreg_indices[c] = position (within qargs) of the qubit read into clbit c, or the
                    # sentinel len(qargs) for clbits nobody measures (statevector_sampler.py:217-222)
                    arrays[item.creg_name][index] = pack_samples(samples, len(qargs), item.num_bits, item.qreg_indices)
            meas = {item.creg_name: ns.BitArray(arrays[item.creg_name], item.num_bits) for item in meas_info}
            return ns.SamplerPubResult(ns.DataBin(**meas, shape=pub.shape),
                                       metadata={"shots": pub.shots, "circuit_metadata": pub.circuit.metadata})

    return SimpleNamespace(DeviceStatevector=DeviceStatevector, DeviceEstimator=DeviceEstimator,
                           DeviceSampler=DeviceSampler, ns=ns)


_cached = None


def device_classes():
    """The adapter over the real Qiskit (ImportError when Qiskit is not importable)."""
    global _cached
    if _cached is None:
        _cached = make_adapter(qiskit_namespace())
    return _cached


def install():
    """Opt-in (SURVEY 8b B2 iii): rebind ``qiskit.quantum_info.Statevector`` and
    ``qiskit.primitives.StatevectorEstimator / StatevectorSampler`` to the device classes."""
    import qiskit.primitives as _prims          # noqa: PLC0415 - optional dependency
    import qiskit.quantum_info as _qi

    ad = device_classes()
    _lib.require_device()
    _prims.StatevectorEstimator = ad.DeviceEstimator
    _prims.StatevectorSampler = ad.DeviceSampler
    _qi.Statevector = ad.DeviceStatevector
    return ad
