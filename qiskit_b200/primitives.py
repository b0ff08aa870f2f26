"""``StatevectorEstimator`` / ``StatevectorSampler`` on the device engine.

Same ``run`` signatures, broadcasting, RNG use, result layout and metadata as
qiskit/primitives/statevector_estimator.py:114-176 and
statevector_sampler.py:136-251.  The containers below (``DataBin``, ``PubResult``,
``BitArray`` ...) are small stand-ins for ``qiskit.primitives.containers`` -- real
Qiskit is not importable in this image; INTEGRATION.md shows how the two classes
plug into Qiskit's own containers.
"""
from __future__ import annotations

import ctypes as C
import os
import warnings
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass

import numpy as np

from . import _lib
from .exceptions import QiskitError
from .lowering import GateStream, walk
from .pauli import Pauli, SparsePauliOp, pauli_masks
from .statevector import Statevector

# ------------------------------------------------------------- containers --


class DataBin:
    """Namespace of named arrays with a common leading ``shape``
    (primitives/containers/data_bin.py)."""

    def __init__(self, *, shape=(), **data):
        self._shape = tuple(shape)
        self._data = dict(data)
        for k, v in data.items():
            setattr(self, k, v)

    @property
    def shape(self):
        return self._shape

    def __len__(self):
        return len(self._data)

    def keys(self):
        return self._data.keys()

    def items(self):
        return self._data.items()

    def __getitem__(self, key):
        return self._data[key]

    def __contains__(self, key):
        return key in self._data

    def __repr__(self):
        return f"DataBin({', '.join(self._data)}, shape={self._shape})"


class PubResult:
    def __init__(self, data: DataBin, metadata=None):
        self.data = data
        self.metadata = {} if metadata is None else metadata


class SamplerPubResult(PubResult):
    def join_data(self, names=None):
        names = list(self.data.keys()) if names is None else list(names)
        arrays = [self.data[n] for n in names]
        out = arrays[0]
        for a in arrays[1:]:
            out = BitArray.concatenate_bits([out, a])
        return out


class PrimitiveResult:
    def __init__(self, pub_results, metadata=None):
        self._pub_results = list(pub_results)
        self.metadata = {} if metadata is None else metadata

    def __getitem__(self, i):
        return self._pub_results[i]

    def __len__(self):
        return len(self._pub_results)

    def __iter__(self):
        return iter(self._pub_results)


class PrimitiveJob:
    """One worker thread per job, as primitives/primitive_job.py:50-56."""

    def __init__(self, function, *args):
        self._function, self._args = function, args
        self._future = None

    def _submit(self):
        executor = ThreadPoolExecutor(max_workers=1)
        self._future = executor.submit(self._function, *self._args)
        executor.shutdown(wait=False)

    def result(self):
        return self._future.result()

    def done(self):
        return self._future.done()

    def status(self):
        if self._future.running():
            return "RUNNING"
        if self._future.cancelled():
            return "CANCELLED"
        if self._future.done():
            return "ERROR" if self._future.exception() is not None else "DONE"
        return "INITIALIZING"


class BitArray:
    """uint8 array [..., num_shots, num_bytes], big-endian bytes, with ``num_bits``
    valid bits (primitives/containers/bit_array.py:57-110)."""

    def __init__(self, array: np.ndarray, num_bits: int):
        array = np.asarray(array, dtype=np.uint8)
        if array.ndim < 2:
            raise ValueError("array must have shape (..., num_shots, num_bytes)")
        if array.shape[-1] != (num_bits + 7) // 8:
            raise ValueError("number of bytes does not match num_bits")
        self._array, self._num_bits = array, int(num_bits)

    @property
    def array(self):
        return self._array

    @property
    def num_bits(self):
        return self._num_bits

    @property
    def num_shots(self):
        return self._array.shape[-2]

    @property
    def shape(self):
        return self._array.shape[:-2]

    def _ints(self, loc=None) -> np.ndarray:
        arr = self._array if loc is None else self._array[loc]
        arr = arr.reshape(-1, arr.shape[-1])
        vals = np.zeros(arr.shape[0], dtype=object if self._num_bits > 63 else np.int64)
        for b in range(arr.shape[1]):
            vals = vals * 256 + arr[:, b].astype(vals.dtype)
        return vals

    def get_int_counts(self, loc=None) -> dict:
        vals, counts = np.unique(self._ints(loc), return_counts=True)
        return {int(v): int(c) for v, c in zip(vals, counts)}

    def get_counts(self, loc=None) -> dict:
        return {format(k, f"0{self._num_bits}b"): v for k, v in self.get_int_counts(loc).items()}

    def slice_bits(self, indices):
        """bit_array.py:515-548: keep the listed bit positions (bit 0 = least significant), in
        order; the gather runs on the device over the integer outcomes."""
        from .sampled_expval import marginalize_samples

        idx = [int(indices)] if isinstance(indices, (int, np.integer)) else [int(i) for i in indices]
        if any(i < 0 or i >= self._num_bits for i in idx):
            raise IndexError("index is out of bounds")
        if self._num_bits > 64:
            raise ValueError("slice_bits on more than 64 bits is outside the device path")
        vals = marginalize_samples(self._ints().astype(np.uint64), idx)
        nbytes = (len(idx) + 7) // 8
        out = np.zeros((vals.size, nbytes), dtype=np.uint8)
        for b in range(nbytes):
            out[:, nbytes - 1 - b] = (vals >> np.uint64(8 * b)) & np.uint64(255)
        return BitArray(out.reshape(self.shape + (self.num_shots, nbytes)), len(idx))

    def expectation_values(self, observables, *, device: int = 0) -> np.ndarray:
        """bit_array.py:613-656: expectation values of diagonal observables (I, Z, 0, 1),
        broadcast against this array's shape; real part, as ``sampled_expectation_value``.
        Every position's shots go to the device once with all the labels that need them."""
        from .sampled_expval import OPERS, _device_expvals, _oper_terms

        def is_leaf(o):
            return isinstance(o, (str, dict)) or hasattr(o, "paulis") or (hasattr(o, "z") and hasattr(o, "x"))

        def to_terms(o):
            if isinstance(o, dict):
                labels, coeffs = [k.upper() for k in o], np.asarray(list(o.values()))
            else:
                labels, coeffs = _oper_terms(o)
            for lab in labels:
                if len(lab) != self._num_bits:
                    raise ValueError(
                        f"The number of qubits of the observable ({len(lab)}) does not match "
                        f"the number of bits of the bit array ({self._num_bits})"
                    )
                if set(lab).difference(OPERS):
                    raise ValueError(f"Input operator {lab} is not diagonal")
            return labels, coeffs

        def build(o):  # nested lists -> object ndarray of (labels, coeffs), like ObservablesArray.coerce
            if is_leaf(o):
                out = np.empty((), dtype=object)
                out[()] = to_terms(o)
                return out
            subs = [build(x) for x in o]
            if not subs or len({x.shape for x in subs}) != 1:
                raise ValueError("ragged or empty observables array")
            out = np.empty((len(subs),) + subs[0].shape, dtype=object)
            for i, x in enumerate(subs):
                out[i] = x[()] if x.ndim == 0 else x
            return out

        obs = build(observables)
        if self._num_bits > 64:
            raise ValueError("expectation_values on more than 64 bits is outside the device path")
        arr_indices = np.fromiter(np.ndindex(self.shape), dtype=object, count=int(np.prod(self.shape, dtype=int)))
        arr_indices = arr_indices.reshape(self.shape)
        try:
            bc_indices, bc_obs = np.broadcast_arrays(arr_indices, obs)
        except ValueError as ex:
            raise ValueError("observables shape is not broadcastable with the bit array") from ex
        # one device call per position of the bit array: all labels any broadcast partner needs
        need: dict = {}
        for index in np.ndindex(bc_indices.shape):
            need.setdefault(bc_indices[index], {}).update(dict.fromkeys(bc_obs[index][0]))
        values: dict = {}
        for loc, labels in need.items():
            labels = list(labels)
            samples = self._ints(loc).astype(np.uint64)
            values[loc] = dict(zip(labels, _device_expvals(samples, None, labels, device)))
        out = np.zeros(bc_indices.shape, dtype=float)
        for index in np.ndindex(bc_indices.shape):
            labels, coeffs = bc_obs[index]
            vals = values[bc_indices[index]]
            out[index] = float(np.real(sum(c * vals[lab] for lab, c in zip(labels, coeffs))))
        return out

    def get_bitstrings(self, loc=None) -> list:
        return [format(int(v), f"0{self._num_bits}b") for v in self._ints(loc)]

    @staticmethod
    def concatenate_bits(arrays):
        total = sum(a.num_bits for a in arrays)
        shots = arrays[0].num_shots
        ints = np.zeros(arrays[0]._array[..., 0].size, dtype=object)
        for a in arrays:  # later registers are more significant
            pass
        shift = 0
        for a in arrays:
            ints = ints + (a._ints().astype(object) << shift)
            shift += a.num_bits
        nbytes = (total + 7) // 8
        out = np.zeros((ints.size, nbytes), dtype=np.uint8)
        for b in range(nbytes):
            out[:, nbytes - 1 - b] = [(int(v) >> (8 * b)) & 255 for v in ints]
        return BitArray(out.reshape(arrays[0].shape + (shots, nbytes)), total)


# ------------------------------------------------------------ pub coercion --


def _is_projector_label(label: str, num_qubits: int) -> bool:
    return (len(label) == num_qubits and any(ch in "+-rl01" for ch in label)
            and all(ch in "IXYZ+-rl01" for ch in label))


def _coerce_observable(obs, num_qubits: int) -> SparsePauliOp:
    """One observable -> SparsePauliOp (observables_array.py:247-303 accepts str, Pauli,
    SparsePauliOp and {label: coeff} mappings)."""
    if isinstance(obs, SparsePauliOp):
        op = obs
    elif hasattr(obs, "paulis") and hasattr(obs, "coeffs"):       # real qiskit SparsePauliOp
        op = SparsePauliOp.__new__(SparsePauliOp)
        op.paulis, op.coeffs = obs.paulis, np.asarray(obs.coeffs, dtype=complex)
    elif hasattr(obs, "to_sparse_list") and hasattr(obs, "num_qubits"):
        # SparseObservable (this package's or Qiskit's): projector letters are expanded into Paulis
        # (sparse_observable.py); the reference's estimator stops at them (statevector_estimator.py:165 TODO)
        from .sparse_observable import SparseObservable

        op = SparseObservable.from_any(obs).to_sparse_pauli_op()
    elif isinstance(obs, str) and _is_projector_label(obs, num_qubits):
        # a label over the extended alphabet (observables_array.py:265-266 -> SparseObservable.from_label):
        # with exactly one letter per qubit, "+" and "-" are the X-eigenstate projectors, not phase prefixes
        from .sparse_observable import SparseObservable

        op = SparseObservable.from_label(obs).to_sparse_pauli_op()
    elif isinstance(obs, (str, Pauli)):
        op = SparsePauliOp([obs])
    elif isinstance(obs, dict):
        if not obs:
            raise ValueError("Empty observable mapping")
        if any(isinstance(k, str) and _is_projector_label(k, num_qubits) for k in obs):
            from .sparse_observable import SparseObservable   # observables_array.py:267-277

            if not all(isinstance(k, str) and len(k) == num_qubits for k in obs):
                raise ValueError("projector labels cannot be mixed with Pauli objects or prefixed labels")
            op = SparseObservable.from_list(list(obs.items())).to_sparse_pauli_op()
        else:
            op = SparsePauliOp(list(obs.keys()), list(obs.values()))
    else:
        raise TypeError(f"Invalid observable type: {type(obs)}")
    if op.num_qubits != num_qubits:
        raise ValueError(
            f"The number of qubits of the circuit ({num_qubits}) does not match the number of "
            f"qubits of the observable ({op.num_qubits})."
        )
    return op


def _observables_array(observables, num_qubits: int) -> np.ndarray:
    """Nested lists of observables -> object ndarray (ObservablesArray.coerce)."""
    def is_leaf(o):
        return isinstance(o, (str, Pauli, SparsePauliOp, dict)) or hasattr(o, "paulis") or hasattr(o, "to_sparse_list")

    if is_leaf(observables):
        out = np.empty((), dtype=object)
        out[()] = _coerce_observable(observables, num_qubits)
        return out
    subs = [_observables_array(o, num_qubits) for o in observables]
    if not subs:
        return np.empty((0,), dtype=object)
    if len({s.shape for s in subs}) != 1:
        raise ValueError("ragged observables array")
    out = np.empty((len(subs),) + subs[0].shape, dtype=object)
    for i, s in enumerate(subs):
        out[i] = s[()] if s.ndim == 0 else s   # a 0-d object array would be stored as-is
    return out


def _bindings_array(parameter_values, circuit) -> np.ndarray:
    """Parameter values -> float array [..., num_parameters] (BindingsArray.coerce,
    containers/bindings_array.py): bare arrays bind in ``circuit.parameters`` order."""
    num_params = len(circuit.parameters)
    if parameter_values is None:
        vals = np.zeros((0,), dtype=float)
    elif isinstance(parameter_values, dict):
        params = list(circuit.parameters)
        cols = {}
        lead = None
        for key, v in parameter_values.items():
            keys = list(key) if isinstance(key, (tuple, list)) or hasattr(key, "params") else [key]
            v = np.asarray(v, dtype=float)
            if len(keys) == 1 and (v.ndim == 0 or v.shape[-1] != 1):
                v = v[..., None]
            lead = v.shape[:-1] if lead is None else np.broadcast_shapes(lead, v.shape[:-1])
            for j, p in enumerate(keys):
                cols[id(p)] = v[..., j]
        vals = np.zeros(lead + (num_params,), dtype=float)
        for j, p in enumerate(params):
            if id(p) not in cols:
                raise ValueError(f"no value given for parameter {p}")
            vals[..., j] = cols[id(p)]
    else:
        # bindings_array.py:345-375 (_infer_shape): the last axis runs over parameters, except
        # that for a 1-parameter circuit a missing trailing axis of length 1 is added
        vals = np.asarray(parameter_values, dtype=float)
        if num_params == 1 and not (vals.shape and vals.shape[-1] == 1):
            vals = vals[..., None]
        if vals.ndim == 0:
            raise ValueError("parameter values must have a trailing axis over parameters")
    if vals.shape[-1] != num_params:
        raise ValueError(
            f"The number of values ({vals.shape[-1]}) does not match the number of parameters "
            f"({num_params}) for the circuit."
        )
    return vals


@dataclass
class EstimatorPub:
    circuit: object
    observables: np.ndarray      # object array of SparsePauliOp
    parameter_values: np.ndarray  # float [..., num_parameters]
    precision: float

    @property
    def shape(self):
        return np.broadcast_shapes(self.observables.shape, self.parameter_values.shape[:-1])

    @classmethod
    def coerce(cls, pub, precision):
        """estimator_pub.py:107-199."""
        if isinstance(pub, EstimatorPub):
            return pub
        if precision is not None and precision < 0:
            raise ValueError("precision must be non-negative")
        if not isinstance(pub, (tuple, list)) or not 2 <= len(pub) <= 4:
            raise ValueError(f"The length of pub must be 2, 3 or 4, but got {len(pub) if hasattr(pub, '__len__') else pub}.")
        circuit = pub[0]
        obs = _observables_array(pub[1], circuit.num_qubits)
        vals = _bindings_array(pub[2] if len(pub) > 2 else None, circuit)
        if len(pub) > 3 and pub[3] is not None:
            precision = pub[3]
            if precision < 0:
                raise ValueError("precision must be non-negative")
        try:
            np.broadcast_shapes(obs.shape, vals.shape[:-1])
        except ValueError as ex:
            raise ValueError(
                f"The observables shape {obs.shape} and the parameter values shape "
                f"{vals.shape[:-1]} are not broadcastable."
            ) from ex
        return cls(circuit, obs, vals, precision)


@dataclass
class SamplerPub:
    circuit: object
    parameter_values: np.ndarray
    shots: int

    @property
    def shape(self):
        return self.parameter_values.shape[:-1]

    @classmethod
    def coerce(cls, pub, shots):
        """sampler_pub.py:88-168."""
        if isinstance(pub, SamplerPub):
            return pub
        if hasattr(pub, "num_qubits") and not isinstance(pub, (tuple, list)):
            pub = (pub,)
        if not 1 <= len(pub) <= 3:
            raise ValueError(f"The length of pub must be 1, 2 or 3, but got {len(pub)}.")
        circuit = pub[0]
        vals = _bindings_array(pub[1] if len(pub) > 1 else None, circuit)
        if len(pub) > 2 and pub[2] is not None:
            shots = pub[2]
        # sampler_pub.py:96-102, 149-153
        from numbers import Integral

        if shots is None or not isinstance(shots, Integral) or isinstance(shots, bool):
            raise TypeError("shots must be an integer")
        if shots <= 0:
            raise ValueError("shots must be positive")
        return cls(circuit, vals, int(shots))


def _has_reset(circuit) -> bool:
    for inst in circuit:
        op = inst.operation
        # `initialize` on a subset of qubits resets them first (statevector.py:989-1014): it draws
        # from the state's RNG exactly like `reset`, so it is just as stochastic
        if getattr(op, "name", None) in ("reset", "initialize"):
            return True
        d = getattr(op, "definition", None)
        if d is not None and hasattr(d, "__iter__") and _has_reset(d):
            return True
    return False


def _bound_instruction(circuit, values):
    bound = circuit.assign_parameters(values) if len(values) else circuit
    return bound.to_instruction()


# --------------------------------------------------------------- estimator --


class StatevectorEstimator:
    """statevector_estimator.py:114-176 on the device engine."""

    def __init__(self, *, default_precision: float = 0.0, seed=None, device: int = 0):
        self._default_precision = default_precision
        self._seed = seed
        self._device = device

    @property
    def default_precision(self):
        return self._default_precision

    @property
    def seed(self):
        return self._seed

    def run(self, pubs, *, precision=None):
        if precision is None:
            precision = self._default_precision
        coerced = [EstimatorPub.coerce(pub, precision) for pub in pubs]
        job = PrimitiveJob(self._run, coerced)
        job._submit()
        return job

    def _run(self, pubs):
        return PrimitiveResult([self._run_pub(pub) for pub in pubs], metadata={"version": 2})

    def _run_pub(self, pub: EstimatorPub) -> PubResult:
        rng = np.random.default_rng(self._seed)
        circuit, precision = pub.circuit, pub.precision
        shape = pub.shape
        pshape = pub.parameter_values.shape[:-1]
        vals_b = np.broadcast_to(pub.parameter_values, shape + pub.parameter_values.shape[-1:]) \
            if pshape != shape else pub.parameter_values
        obs_b = np.broadcast_to(pub.observables, shape)
        # which parameter set each broadcast index uses, so a state can be reused across
        # consecutive indices that share it -- legal only without `reset` (stochastic,
        # statevector_estimator.py:163 re-simulates every index)
        pidx = np.broadcast_to(np.arange(int(np.prod(pshape, dtype=int))).reshape(pshape), shape)
        reuse = not _has_reset(circuit)
        evs = np.zeros(shape, dtype=np.float64)
        stds = np.zeros(shape, dtype=np.float64)
        indices = list(np.ndindex(*shape))
        raw = None
        self.last_prefix_stats = None
        if reuse:
            raw = self._run_shared_prefix(circuit, vals_b, obs_b, pidx, indices)
            if raw is None:
                raw = self._run_lanes(circuit, vals_b, obs_b, pidx, indices)
        state, state_key = None, None
        for pos, index in enumerate(indices):
            if raw is not None:
                ev = raw[pos]
            else:
                key = int(pidx[index])
                if not (reuse and state is not None and key == state_key):
                    state = None  # release the previous buffer to the pool before simulating
                    state = Statevector.from_int(0, 1 << circuit.num_qubits, device=self._device)
                    state.seed(rng)
                    state._evolve_instruction_inplace(_bound_instruction(circuit, vals_b[index]), None)
                    state_key = key
                ev = np.real_if_close(state.expectation_value(obs_b[index]))
            if precision != 0:
                if not np.isreal(ev):
                    raise ValueError("Given operator is not Hermitian and noise cannot be added.")
                ev = rng.normal(ev, precision)
            evs[index] = ev
        data = DataBin(evs=evs, stds=stds, shape=evs.shape)
        return PubResult(
            data, metadata={"target_precision": precision,
                            "circuit_metadata": getattr(circuit, "metadata", None)}
        )


    # Shared-prefix caching across parameter sets (prefix_cache.py): when the ordered
    # parameter sets of the pub share leading gates (a sweep of late angles, parameter-shift pairs),
    # the state after the shared prefix is checkpointed on the device instead of being re-simulated.
    # Random parameter sets (the headline benchmark) share nothing and take `_run_lanes` unchanged.
    _PREFIX_CACHE = os.environ.get("QSV_PREFIX_CACHE", "1") != "0"

    def _run_shared_prefix(self, circuit, vals_b, obs_b, pidx, indices):
        from . import prefix_cache as pc

        if not self._PREFIX_CACHE:
            return None
        groups: dict = {}
        for pos, index in enumerate(indices):
            groups.setdefault(int(pidx[index]), []).append(pos)
        if len(groups) < 2:
            return None
        keys = list(groups)
        values = np.stack([np.asarray(vals_b[indices[groups[k][0]]], dtype=float) for k in keys])
        first_inst, total = pc.first_use(circuit)
        order, share = pc.order_sets(values, first_inst, total)
        if share.size == 0 or float(share.max()) < pc.MIN_SHARE:
            return None
        n = circuit.num_qubits
        out = [None] * len(indices)

        def lower(sid):
            stream = GateStream()
            walk(None, _bound_instruction(circuit, values[sid]), None, stream)
            return stream

        def make_state():
            return Statevector.from_int(0, 1 << n, device=self._device)

        def consume(sid, state):
            for pos in groups[keys[sid]]:
                out[pos] = np.real_if_close(state.expectation_value(obs_b[indices[pos]]))

        lanes = self._LANES if n <= self._LANE_MAX_QUBITS - 2 else 1
        stats = pc.run_sets(order, lower, make_state, consume, lanes=lanes,
                            can_checkpoint=lambda: pc._device_can_hold_checkpoint(n, self._device))
        self.last_prefix_stats = stats.as_dict()
        return out

    # Two host lanes, one engine handle (and CUDA stream) each.  The reference evaluates the broadcast
    # indices one after the other (statevector_estimator.py:160-171); without `reset` the values do not
    # depend on that order, so while one lane waits for the device (ctypes drops the GIL inside the
    # engine) the other binds, lowers and plans the next parameter set: the GPU never idles on host
    # work.  Values are identical to the serial loop (each evaluation is the same kernel sequence on
    # its own buffer); noise is drawn afterwards in index order from the same generator.
    _LANES = 2
    _LANE_MAX_QUBITS = 30  # 2 states + 2 basis-change scratch copies must fit (4 x 17 GB at 30 qubits)

    def _run_lanes(self, circuit, vals_b, obs_b, pidx, indices):
        groups: dict = {}
        for pos, index in enumerate(indices):
            groups.setdefault(int(pidx[index]), []).append(pos)
        if len(groups) < 2 or circuit.num_qubits > self._LANE_MAX_QUBITS or self._LANES < 2:
            return None
        out = [None] * len(indices)

        def work(positions):
            state = Statevector.from_int(0, 1 << circuit.num_qubits, device=self._device)
            state._evolve_instruction_inplace(_bound_instruction(circuit, vals_b[indices[positions[0]]]), None)
            for pos in positions:
                out[pos] = np.real_if_close(state.expectation_value(obs_b[indices[pos]]))

        from concurrent.futures import ThreadPoolExecutor

        with ThreadPoolExecutor(max_workers=self._LANES) as pool:
            for fut in [pool.submit(work, positions) for positions in groups.values()]:
                fut.result()
        return out


# ----------------------------------------------------------------- sampler --


@dataclass
class _MeasureInfo:
    creg_name: str
    num_bits: int
    num_bytes: int
    qreg_indices: list


def _final_measurement_mapping(circuit) -> dict:
    """statevector_sampler.py:254-286: (creg, offset) -> qubit for final measurements."""
    active_qubits = set(range(circuit.num_qubits))
    active_cbits = set(range(circuit.num_clbits))
    mapping = {}
    for item in circuit[::-1]:
        if item.name == "measure":
            loc = circuit.find_bit(item.clbits[0])
            cbit = loc.index
            qbit = circuit.find_bit(item.qubits[0]).index if not isinstance(item.qubits[0], int) \
                else item.qubits[0]
            if cbit in active_cbits and qbit in active_qubits:
                for creg in loc.registers:
                    mapping[creg] = qbit
                active_cbits.remove(cbit)
        elif item.name not in ["barrier", "delay"]:
            for qq in item.qubits:
                q = qq if isinstance(qq, int) else circuit.find_bit(qq).index
                active_qubits.discard(q)
        if not active_cbits or not active_qubits:
            break
    return mapping


def _preprocess_circuit(circuit):
    """statevector_sampler.py:207-232."""
    num_bits_dict = {creg.name: creg.size for creg in circuit.cregs}
    mapping = _final_measurement_mapping(circuit)
    qargs = sorted(set(mapping.values()))
    qargs_index = {v: k for k, v in enumerate(qargs)}
    stripped = circuit.remove_final_measurements(inplace=False)
    if any(inst.is_control_flow() for inst in stripped):
        raise QiskitError("StatevectorSampler cannot handle ControlFlowOp")
    if any(inst.name == "measure" for inst in stripped):
        raise QiskitError("StatevectorSampler cannot handle mid-circuit measurements")
    sentinel = len(qargs)
    indices = {key: [sentinel] * val for key, val in num_bits_dict.items()}
    for (creg, ind), qreg in mapping.items():
        indices[creg.name][ind] = qargs_index[qreg]
    meas_info = [
        _MeasureInfo(name, nbits, (nbits + 7) // 8, indices[name]) for name, nbits in num_bits_dict.items()
    ]
    return stripped, qargs, meas_info


def pack_samples(sample_idx: np.ndarray, num_measured: int, num_bits: int, indices) -> np.ndarray:
    """Index form of ``_samples_to_packed_array`` (statevector_sampler.py:235-251): clbit c
    takes bit ``indices[c]`` of the sample index (``num_measured`` = sentinel -> 0)."""
    sample_idx = np.ascontiguousarray(sample_idx, dtype=np.uint64)
    num_bytes = (num_bits + 7) // 8
    out = np.zeros((sample_idx.size, num_bytes), dtype=np.uint8)
    src = _lib.int_array([(-1 if i >= num_measured else i) for i in indices])
    _lib.check(_lib.load().qsv_pack_samples(sample_idx.ctypes.data, sample_idx.size, src, num_bits,
                                            out.ctypes.data))
    return out


class StatevectorSampler:
    """statevector_sampler.py:136-204 on the device engine."""

    def __init__(self, *, default_shots: int = 1024, seed=None, device: int = 0):
        self._default_shots = default_shots
        self._seed = seed
        self._device = device

    @property
    def default_shots(self):
        return self._default_shots

    @property
    def seed(self):
        return self._seed

    def run(self, pubs, *, shots=None):
        if shots is None:
            shots = self._default_shots
        coerced = [SamplerPub.coerce(pub, shots) for pub in pubs]
        if any(len(pub.circuit.cregs) == 0 for pub in coerced):
            warnings.warn(
                "One of your circuits has no output classical registers and so the result "
                "will be empty. Did you mean to add measurement instructions?",
                UserWarning,
            )
        job = PrimitiveJob(self._run, coerced)
        job._submit()
        return job

    def _run(self, pubs):
        return PrimitiveResult([self._run_pub(pub) for pub in pubs], metadata={"version": 2})

    _PREFIX_CACHE = StatevectorEstimator._PREFIX_CACHE

    def _run_shared_prefix(self, circuit, values_nd, indices, sample_into) -> bool:
        """Shared-prefix caching (prefix_cache.py) for the parameter sets of a Sampler pub: the seeded
        generator is per index (not shared across indices), so the visiting order does not change results."""
        from . import prefix_cache as pc

        if not self._PREFIX_CACHE or len(indices) < 2 or _has_reset(circuit) or circuit.num_qubits == 0:
            return False
        values = np.stack([np.asarray(values_nd[index], dtype=float) for index in indices])
        first_inst, total = pc.first_use(circuit)
        order, share = pc.order_sets(values, first_inst, total)
        if share.size == 0 or float(share.max()) < pc.MIN_SHARE:
            return False
        n = circuit.num_qubits

        def lower(sid):
            stream = GateStream()
            walk(None, _bound_instruction(circuit, values[sid]), None, stream)
            return stream

        stats = pc.run_sets(order, lower, lambda: Statevector.from_int(0, 1 << n, device=self._device),
                            lambda sid, state: sample_into(indices[sid], state), lanes=1,
                            can_checkpoint=lambda: pc._device_can_hold_checkpoint(n, self._device))
        self.last_prefix_stats = stats.as_dict()
        return True

    def _run_pub(self, pub: SamplerPub) -> SamplerPubResult:
        circuit, qargs, meas_info = _preprocess_circuit(pub.circuit)
        shape = pub.shape
        arrays = {
            item.creg_name: np.zeros(shape + (pub.shots, item.num_bytes), dtype=np.uint8)
            for item in meas_info
        }
        def sample_into(index, state):
            state.seed(self._seed)       # every index draws from its own generator (statevector_sampler.py:209)
            if qargs:
                samples = state.sample_indices(pub.shots, qargs)
            else:
                samples = np.zeros(pub.shots, dtype=np.uint64)
            for item in meas_info:
                arrays[item.creg_name][index] = pack_samples(samples, len(qargs), item.num_bits,
                                                             item.qreg_indices)

        indices = list(np.ndindex(*shape))
        self.last_prefix_stats = None
        if not self._run_shared_prefix(circuit, pub.parameter_values, indices, sample_into):
            for index in indices:
                state = Statevector(_bound_instruction(circuit, pub.parameter_values[index]),
                                    device=self._device)
                sample_into(index, state)
                state = None
        meas = {item.creg_name: BitArray(arrays[item.creg_name], item.num_bits) for item in meas_info}
        return SamplerPubResult(
            DataBin(**meas, shape=shape),
            metadata={"shots": pub.shots, "circuit_metadata": getattr(pub.circuit, "metadata", None)},
        )
