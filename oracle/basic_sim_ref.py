"""NumPy restatement of BasicSimulator's run loop (TEST INFRASTRUCTURE).
Follows qiskit/providers/basic_provider/basic_simulator.py:250-352 and :611-760 line by line,
on the flat little-endian statevector of oracle/statevector_ref.py.  ``ops`` is a list of
("gate", matrix, qubits) | ("measure", qubit, clbit) | ("reset", qubit) | ("barrier",)."""
from __future__ import annotations

import math

import numpy as np

from . import statevector_ref as ref


def _measure_outcome(psi, n, qubit, rng):
    """:250-268."""
    probabilities = ref.probabilities(psi, [qubit])
    random_number = rng.random()
    if random_number < probabilities[0]:
        return 0, probabilities[0]
    return 1, probabilities[1]


def validate_measure_sampling(ops, shots) -> bool:
    """:402-436."""
    if shots <= 1:
        return False
    measure_flag = False
    for op in ops:
        if op[0] == "reset":
            return False
        if measure_flag:
            if op[0] not in ("measure", "barrier"):
                return False
        elif op[0] == "measure":
            measure_flag = True
    return measure_flag


def run(n, num_clbits, ops, shots, seed, initial=None, global_phase=0.0):
    """Returns the list of integer classical-memory values, one per shot (:648-760)."""
    rng = np.random.default_rng(seed)
    sample_measure = validate_measure_sampling(ops, shots)
    memory = []
    measure_sample_ops = []
    for _ in range(1 if sample_measure else shots):
        psi = ref.basis_state(n) if initial is None else np.array(initial, dtype=complex)
        psi = psi * np.exp(1j * global_phase)
        classical_memory = 0
        for op in ops:
            if op[0] == "gate":
                psi = ref.evolve_operator(psi, op[1], list(op[2]))
            elif op[0] == "measure":
                _, qubit, cmembit = op
                if sample_measure:
                    measure_sample_ops.append((qubit, cmembit))
                    continue
                outcome, probability = _measure_outcome(psi, n, qubit, rng)
                membit = 1 << cmembit
                classical_memory = (classical_memory & (~membit)) | (outcome << cmembit)
                upd = np.zeros((2, 2))
                upd[outcome, outcome] = 1 / math.sqrt(probability)
                psi = ref.evolve_operator(psi, upd, [qubit])
            elif op[0] == "reset":
                outcome, probability = _measure_outcome(psi, n, op[1], rng)
                upd = np.zeros((2, 2))
                upd[0, outcome] = 1 / math.sqrt(probability)
                psi = ref.evolve_operator(psi, upd, [op[1]])
        if num_clbits > 0:
            if sample_measure:
                # :270-313
                measured_qubits = sorted({q for q, _ in measure_sample_ops})
                probabilities = ref.probabilities(psi, measured_qubits)
                samples = rng.choice(range(2 ** len(measured_qubits)), shots, p=probabilities)
                for sample in samples:
                    cm = 0
                    for qubit, cmembit in measure_sample_ops:
                        pos = measured_qubits.index(qubit)
                        qubit_outcome = int((int(sample) & (1 << pos)) >> pos)
                        membit = 1 << cmembit
                        cm = (cm & (~membit)) | (qubit_outcome << cmembit)
                    memory.append(cm)
            else:
                memory.append(classical_memory)
    return memory
