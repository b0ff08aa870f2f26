"""``BackendSamplerV2`` / ``BackendEstimatorV2`` over a device backend (SURVEY.md 8f N3 + N4).

Mirrors of qiskit/primitives/backend_sampler_v2.py and backend_estimator_v2.py for a backend
with the ``run(circuits, shots=, seed_simulator=, memory=)`` interface of ``BasicSimulator``:
shots are simulated on the device (qiskit_b200/basic_simulator.py), and the estimator's
counts -> expectation-value step (``_pauli_expval_with_variance``, backend_estimator_v2.py:87-112)
runs in ``qsv_sampled_expval`` (integer counts: exact, any summation order).
"""
from __future__ import annotations

import math

import numpy as np

from .circuit import ClassicalRegister
from .primitives import (BitArray, DataBin, EstimatorPub, PrimitiveJob, PrimitiveResult, PubResult,
                         SamplerPub, SamplerPubResult)
from .sampled_expval import _device_expvals


def _bound(circuit, values):
    return circuit.assign_parameters(values) if circuit.num_parameters else circuit


class BackendSamplerV2:
    """backend_sampler_v2.py:100-340: every parameter set of every pub is one backend circuit;
    the per-shot memory is split into one ``BitArray`` per classical register."""

    def __init__(self, *, backend, options: dict | None = None, default_shots: int = 1024, seed_simulator=None):
        options = dict(options or {})         # backend_sampler_v2.py:62-76: default_shots, seed_simulator
        self._backend = backend
        self._default_shots = int(options.get("default_shots", default_shots))
        self._seed = options.get("seed_simulator", seed_simulator)

    @property
    def backend(self):
        return self._backend

    def run(self, pubs, *, shots=None):
        if shots is None:
            shots = self._default_shots
        coerced = [SamplerPub.coerce(pub, shots) for pub in pubs]
        job = PrimitiveJob(self._run, coerced)
        job._submit()
        return job

    def _run(self, pubs):
        return PrimitiveResult([self._run_pub(pub) for pub in pubs], metadata={"version": 2})

    def _run_pub(self, pub: SamplerPub) -> SamplerPubResult:
        circuit, shape = pub.circuit, pub.shape
        cregs = [(c.name, c.size) for c in circuit.cregs]
        arrays = {name: np.zeros(shape + (pub.shots, (size + 7) // 8), dtype=np.uint8) for name, size in cregs}
        indices = list(np.ndindex(*shape))
        bound = [_bound(circuit, pub.parameter_values[index]) for index in indices]
        result = self._backend.run(bound, shots=pub.shots, seed_simulator=self._seed, memory=True).result()
        for k, index in enumerate(indices):
            mem = np.array([int(v, 16) for v in result.data(k)["memory"]], dtype=np.uint64)
            offset = 0
            for name, size in cregs:          # registers in order, first register = lowest bits
                vals = (mem >> np.uint64(offset)) & np.uint64((1 << size) - 1)
                nbytes = (size + 7) // 8
                for b in range(nbytes):
                    arrays[name][index][:, nbytes - 1 - b] = (vals >> np.uint64(8 * b)) & np.uint64(255)
                offset += size
        meas = {name: BitArray(arrays[name], size) for name, size in cregs}
        return SamplerPubResult(DataBin(**meas, shape=shape),
                                metadata={"shots": pub.shots, "circuit_metadata": getattr(circuit, "metadata", None)})


def _qwc_groups(paulis_zx, abelian: bool):
    """Group Pauli terms (z, x bool arrays) qubit-wise commuting, first fit; each group carries the
    basis Pauli that measures all members (backend_estimator_v2.py:440-511 uses
    ``PauliList.group_commuting(qubit_wise=True)``)."""
    groups = []  # [basis_z, basis_x, [member ids]]
    for t, (z, x) in enumerate(paulis_zx):
        placed = False
        if abelian:
            for g in groups:
                both = (g[0] | g[1]) & (z | x)          # qubits where both act
                if np.all((g[0][both] == z[both]) & (g[1][both] == x[both])):
                    new = (z | x) & ~(g[0] | g[1])
                    g[0] = g[0] | (z & new)
                    g[1] = g[1] | (x & new)
                    g[2].append(t)
                    placed = True
                    break
        if not placed:
            groups.append([z.copy(), x.copy(), [t]])
    return groups


class BackendEstimatorV2:
    """backend_estimator_v2.py:170-511: shots = ceil(1 / precision^2) per measurement circuit,
    terms grouped qubit-wise commuting, basis change (Sdg, H) + measure appended to the bound
    circuit, expectation values and standard errors from the counts."""

    def __init__(self, *, backend, options: dict | None = None, default_precision: float = 0.015625,
                 abelian_grouping: bool = True, seed_simulator=None):
        options = dict(options or {})         # backend_estimator_v2.py:45-63
        self._backend = backend
        self._default_precision = float(options.get("default_precision", default_precision))
        self._abelian = bool(options.get("abelian_grouping", abelian_grouping))
        self._seed = options.get("seed_simulator", seed_simulator)

    @property
    def backend(self):
        return self._backend

    def run(self, pubs, *, precision=None):
        if precision is None:
            precision = self._default_precision
        coerced = [EstimatorPub.coerce(pub, precision) for pub in pubs]
        for i, pub in enumerate(coerced):
            if pub.precision <= 0.0:
                raise ValueError(f"The {i}-th pub has precision less than or equal to 0 ({pub.precision}). "
                                 "But precision should be larger than 0.")
        job = PrimitiveJob(self._run, coerced)
        job._submit()
        return job

    def _run(self, pubs):
        return PrimitiveResult([self._run_pub(pub) for pub in pubs], metadata={"version": 2})

    def _run_pub(self, pub: EstimatorPub) -> PubResult:
        circuit, shape = pub.circuit, pub.shape
        shots = int(math.ceil(1.0 / pub.precision ** 2))
        pshape = pub.parameter_values.shape[:-1]
        pidx = np.broadcast_to(np.arange(int(np.prod(pshape, dtype=int))).reshape(pshape), shape)
        obs_b = np.broadcast_to(pub.observables, shape)
        flat_vals = pub.parameter_values.reshape(int(np.prod(pshape, dtype=int)), pub.parameter_values.shape[-1])
        # the Pauli terms every parameter set needs
        need: dict = {}
        for index in np.ndindex(*shape):
            op = obs_b[index]
            terms = need.setdefault(int(pidx[index]), {})
            for z, x in zip(np.asarray(op.paulis.z, bool), np.asarray(op.paulis.x, bool)):
                terms.setdefault((z.tobytes(), x.tobytes()), (z, x))
        circuits, meta = [], []
        for key, terms in need.items():
            bound = _bound(circuit, flat_vals[key])
            keys = list(terms)
            for gz, gx, members in _qwc_groups([terms[k] for k in keys], self._abelian):
                qubits = np.flatnonzero(gz | gx)
                if qubits.size == 0:
                    qubits = np.array([0])            # identity: measure one qubit (:514-519)
                meas = bound.copy_shallow()
                base = meas.num_clbits                # the new register sits above the circuit's own
                meas.add_register(ClassicalRegister(len(qubits), "__c"))
                for clbit, q in enumerate(qubits, start=base):
                    if gx[q]:
                        if gz[q]:
                            meas.sdg(int(q))
                        meas.h(int(q))
                    meas.measure(int(q), clbit)
                circuits.append(meas)
                meta.append((key, qubits, base, [keys[m] for m in members], [terms[keys[m]] for m in members]))
        result = self._backend.run(circuits, shots=shots, seed_simulator=self._seed, memory=False).result()
        expval_map = {}
        for k, (key, qubits, base, tkeys, tzx) in enumerate(meta):
            counts = result.data(k)["counts"]
            outcomes = np.array([int(h, 16) >> base for h in counts], dtype=np.uint64)
            weights = np.array(list(counts.values()), dtype=np.float64)
            pos = {int(q): c for c, q in enumerate(qubits)}
            labels = []
            for z, x in tzx:                           # every non-identity factor reads as Z (:87-112)
                lab = ["I"] * len(qubits)
                for q in np.flatnonzero(z | x):
                    lab[len(qubits) - 1 - pos[int(q)]] = "Z"
                labels.append("".join(lab))
            vals = _device_expvals(outcomes, weights, labels)
            for tk, v in zip(tkeys, vals):
                expval_map[key, tk] = (float(v), 1.0 - float(v) ** 2)
        evs = np.zeros(shape, dtype=float)
        variances = np.zeros(shape, dtype=float)
        for index in np.ndindex(*shape):
            op = obs_b[index]
            for z, x, coeff in zip(np.asarray(op.paulis.z, bool), np.asarray(op.paulis.x, bool), op.coeffs):
                ev, var = expval_map[int(pidx[index]), (z.tobytes(), x.tobytes())]
                evs[index] += np.real(ev * coeff)
                variances[index] += np.abs(coeff) * var ** 0.5
        stds = variances / np.sqrt(shots)
        return PubResult(DataBin(evs=evs, stds=stds, shape=evs.shape),
                         metadata={"target_precision": pub.precision, "shots": shots,
                                   "circuit_metadata": getattr(circuit, "metadata", None)})
