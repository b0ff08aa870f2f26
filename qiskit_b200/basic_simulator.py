"""``BasicSimulator`` on the device engine.

Mirror of qiskit/providers/basic_provider/basic_simulator.py for the run loop (:611-792): the
same walk over ``circuit.data``, the same use of ``np.random.default_rng(seed_simulator)``
(one ``random()`` per mid-circuit measure / reset, :250-268; one batch of uniforms for the
final-measurement sampling, :270-313), the same hex-keyed counts.  What changes is where the
statevector lives: gates are streamed to ``libqsv_b200.so`` (fused), measure / reset use the
device marginal + collapse kernels, final measurements use the device sampler.  The Clifford
shortcut (``use_clifford_optimization``) and the transpiler-facing ``Target`` are out of scope.
"""
from __future__ import annotations

import time
import uuid
import numpy as np

from . import _lib
from .exceptions import QiskitError
from .lowering import GateStream, instruction_to_matrix
from .statevector import Statevector


class BasicProviderError(QiskitError):
    """providers/basic_provider/exceptions.py"""


class Result:
    """The slice of ``qiskit.result.Result`` the simulator's callers use."""

    def __init__(self, backend_name, job_id, results):
        self.backend_name, self.job_id, self.results = backend_name, job_id, results
        self.success = all(r["success"] for r in results)

    def _exp(self, experiment=None):
        if experiment is None:
            if len(self.results) != 1:
                raise QiskitError("You have to select a circuit when there is more than one available")
            return self.results[0]
        if isinstance(experiment, int):
            return self.results[experiment]
        name = getattr(experiment, "name", experiment)
        for r in self.results:
            if r["name"] == name:
                return r
        raise QiskitError(f'Data for experiment "{name}" could not be found.')

    @staticmethod
    def _format(hexkey: str, header: dict) -> str:
        """result/postprocess.py format_counts_memory: zero-padded bit-string, one group per
        classical register, last register leftmost."""
        bits = format(int(hexkey, 16), f"0{header['memory_slots']}b")
        sizes = header.get("creg_sizes") or []
        if len(sizes) < 2:
            return bits
        parts, at = [], 0
        for _, size in reversed(sizes):
            parts.append(bits[at:at + size])
            at += size
        return " ".join(parts)

    def get_counts(self, experiment=None) -> dict:
        """result.py get_counts: hex keys -> formatted bit-strings."""
        exp = self._exp(experiment)
        return {self._format(k, exp["header"]): v for k, v in exp["data"]["counts"].items()}

    def get_memory(self, experiment=None) -> list:
        exp = self._exp(experiment)
        if "memory" not in exp["data"]:
            raise QiskitError("No memory for experiment: run with memory=True")
        return [self._format(k, exp["header"]) for k in exp["data"]["memory"]]

    def data(self, experiment=None) -> dict:
        return self._exp(experiment)["data"]


class BasicProviderJob:
    """Synchronous job (basic_provider_job.py): the result is ready when ``run`` returns."""

    def __init__(self, backend, job_id, result):
        self._backend, self._job_id, self._result = backend, job_id, result

    def result(self, timeout=None):
        return self._result

    def job_id(self):
        return self._job_id

    def status(self):
        return "DONE"

    def backend(self):
        return self._backend


class BasicSimulator:
    """``backend.run(circuits, shots=..., seed_simulator=..., memory=..., initial_statevector=...)``."""

    name = "basic_simulator"
    MAX_QUBITS_MEMORY = 33  # one B200 (basic_simulator.py:97 sizes this from host RAM)
    _DEFAULTS = {"shots": 1024, "memory": True, "initial_statevector": None, "seed_simulator": None}

    def __init__(self, *, device: int = 0, **fields):
        self._device = int(device)
        self.options = dict(self._DEFAULTS)
        self.set_options(**fields)

    def set_options(self, **fields):
        for k, v in fields.items():
            if k not in self._DEFAULTS:
                raise AttributeError(f"Options field {k} is not valid for this backend")
            self.options[k] = v

    # ------------------------------------------------------------ options ---
    def _set_run_options(self, run_options):
        """basic_simulator.py:366-400."""
        self._shots = self.options["shots"]
        self._memory = self.options["memory"]
        self._initial_statevector = self.options["initial_statevector"]
        self._seed_simulator = self.options["seed_simulator"]
        if run_options.get("initial_statevector") is not None:
            self._initial_statevector = np.array(run_options["initial_statevector"], dtype=complex)
        if self._initial_statevector is not None:
            norm = np.linalg.norm(self._initial_statevector)
            if round(norm, 12) != 1:
                raise BasicProviderError(f"Initial statevector is not normalized: norm {norm} != 1")
        if "shots" in run_options:
            self._shots = run_options["shots"]
        if "seed_simulator" in run_options:
            self._seed_simulator = run_options["seed_simulator"]
        elif self._seed_simulator is None:
            self._seed_simulator = int(np.random.randint(2147483647, dtype="int32"))
        if "memory" in run_options:
            self._memory = run_options["memory"]
        self._local_rng = np.random.default_rng(seed=self._seed_simulator)

    def _validate_measure_sampling(self, circuit):
        """basic_simulator.py:402-436: sample once iff shots > 1, no reset, measures only at the end."""
        self._sample_measure = False
        if self._shots <= 1:
            return
        measure_flag = False
        for instruction in circuit.data:
            if instruction.name == "reset":
                return
            if measure_flag:
                if instruction.name not in ("measure", "barrier", "id", "u0"):
                    return
            elif instruction.name == "measure":
                measure_flag = True
        self._sample_measure = measure_flag

    # --------------------------------------------------------- primitives ---
    def _measure_outcome(self, sv: Statevector, qubit: int):
        """basic_simulator.py:250-268: P(qubit) from the device marginal, one ``random()``."""
        probabilities = sv.probabilities([qubit])
        random_number = self._local_rng.random()
        if random_number < probabilities[0]:
            return 0, float(probabilities[0])
        return 1, float(probabilities[1])

    def _collapse(self, sv: Statevector, qubit: int, outcome: int, probability: float, to_zero: bool):
        """:315-352: project on the outcome, renormalise; reset also moves |1> to |0>."""
        _lib.check(_lib.load().qsv_collapse(sv._h, _lib.int_array([qubit]), 1, outcome,
                                            1.0 / np.sqrt(probability), 1 if to_zero else 0))

    def _sample_memory_ints(self, sv: Statevector, measure_params, num_samples: int) -> np.ndarray:
        """:270-313 with integer memory: sample the sorted measured qubits, scatter to the clbits."""
        measured_qubits = sorted({q for q, _ in measure_params})
        uniforms = self._local_rng.random(num_samples)   # what Generator.choice(p=...) draws
        samples = sv.sample_indices(num_samples, measured_qubits, uniforms=uniforms).astype(np.uint64)
        memory = np.zeros(num_samples, dtype=np.uint64)
        for qubit, cmembit in measure_params:            # later measures overwrite earlier ones
            pos = np.uint64(measured_qubits.index(qubit))
            bit = (samples >> pos) & np.uint64(1)
            membit = np.uint64(1 << cmembit)
            memory = (memory & ~membit) | (bit << np.uint64(cmembit))
        return memory

    # ------------------------------------------------------------ run loop --
    def _run_circuit(self, circuit) -> dict:
        """basic_simulator.py:611-792."""
        n, ncl = circuit.num_qubits, circuit.num_clbits
        if n > self.MAX_QUBITS_MEMORY:
            raise BasicProviderError(f"Number of qubits {n} is greater than maximum ({self.MAX_QUBITS_MEMORY})")
        if ncl > 64:
            raise BasicProviderError("more than 64 classical bits are outside the device path")
        if self._initial_statevector is not None and len(self._initial_statevector) != 1 << n:
            raise BasicProviderError(
                f"initial statevector is incorrect length: {len(self._initial_statevector)} != {1 << n}")
        start = time.time()
        self._validate_measure_sampling(circuit)
        ops = list(circuit.data)
        # unitary prefix shared by every shot: simulated once, cloned per shot
        split = next((i for i, op in enumerate(ops) if op.name in ("measure", "reset")), len(ops))
        if self._initial_statevector is None:
            base = Statevector.from_int(0, 1 << n, device=self._device)
        else:
            base = Statevector(self._initial_statevector, device=self._device)
        stream = GateStream()
        if circuit.global_phase:
            stream.phase(float(circuit.global_phase))
        self._stream_ops(ops[:split], stream)
        base._flush(stream)

        memory_ints = np.zeros(0, dtype=np.uint64)
        if self._sample_measure:
            measure_params = [(op.qubits[0], op.clbits[0]) for op in ops[split:] if op.name == "measure"]
            if ncl > 0:
                memory_ints = self._sample_memory_ints(base, measure_params, self._shots)
        else:
            shots_mem = []
            for _ in range(self._shots):
                sv = base.copy() if split < len(ops) else base
                classical_memory = 0
                for op in ops[split:]:
                    if op.name == "measure":
                        sv._flush(stream)
                        qubit, cmembit = op.qubits[0], op.clbits[0]
                        outcome, prob = self._measure_outcome(sv, qubit)
                        membit = 1 << cmembit
                        classical_memory = (classical_memory & ~membit) | (outcome << cmembit)
                        self._collapse(sv, qubit, outcome, prob, to_zero=False)
                    elif op.name == "reset":
                        sv._flush(stream)
                        outcome, prob = self._measure_outcome(sv, op.qubits[0])
                        self._collapse(sv, op.qubits[0], outcome, prob, to_zero=True)
                    else:
                        self._stream_ops([op], stream)
                sv._flush(stream)
                if ncl > 0:
                    shots_mem.append(classical_memory)
            memory_ints = np.array(shots_mem, dtype=np.uint64)

        vals, cnt = np.unique(memory_ints, return_counts=True)
        first = {}
        if len(vals):  # Counter(memory) keeps first-occurrence order
            _, first_idx = np.unique(memory_ints, return_index=True)
            order = np.argsort(first_idx)
            first = {hex(int(vals[i])): int(cnt[i]) for i in order}
        data = {"counts": first}
        if self._memory:
            data["memory"] = [hex(int(v)) for v in memory_ints]
        header = {
            "name": circuit.name, "n_qubits": n,
            "creg_sizes": [[c.name, c.size] for c in circuit.cregs],
            "memory_slots": ncl, "global_phase": circuit.global_phase,
            "metadata": circuit.metadata if circuit.metadata is not None else {},
        }
        return {"name": circuit.name, "seed_simulator": self._seed_simulator, "shots": self._shots,
                "data": data, "status": "DONE", "success": True, "header": header,
                "time_taken": time.time() - start}

    def _stream_ops(self, ops, stream: GateStream):
        for op in ops:
            if op.name in ("barrier", "id", "u0", "delay"):
                continue
            if op.name == "global_phase":
                stream.phase(float(op.operation.params[0]))
                continue
            mat = instruction_to_matrix(op.operation)
            if mat is None:
                raise BasicProviderError(f'{self.name} encountered unrecognized operation "{op.name}"')
            stream.matrix(mat, list(op.qubits))

    def run(self, run_input, **run_options) -> BasicProviderJob:
        """basic_simulator.py:438-490."""
        circuits = run_input if isinstance(run_input, (list, tuple)) else [run_input]
        for key in run_options:
            if key not in self._DEFAULTS:
                import warnings

                warnings.warn(f"Option {key} is not used by this backend", UserWarning, stacklevel=2)
        self._set_run_options({k: v for k, v in run_options.items() if k in self._DEFAULTS})
        _lib.require_device()
        job_id = str(uuid.uuid4())
        results = [self._run_circuit(c) for c in circuits]
        return BasicProviderJob(self, job_id, Result(self.name, job_id, results))
