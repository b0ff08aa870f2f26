"""BasicSimulator: the NumPy restatement against the reference's deterministic
goldens (test/python/providers/basic_provider/test_basic_simulator.py) on CPU; the device
backend against the restatement, shot by shot for the same seed, on GPU."""
import numpy as np
import pytest

from oracle import basic_sim_ref as bref
from oracle import gate_matrices as gm

X = gm.matrix("x", [])
H = gm.matrix("h", [])
CX = gm.matrix("cx", [])
SEED = 42


def g(mat, *qubits):
    return ("gate", mat, list(qubits))


REPEATED = (2, 4, [g(X, 1), ("measure", 0, 0), ("measure", 1, 1), ("measure", 1, 2), ("measure", 0, 3)])      # :57-74
PARTIAL = (5, 4, [g(X, 3), g(X, 1), ("barrier",), ("measure", 3, 1), ("barrier",), ("measure", 1, 0),
                  ("barrier",), ("measure", 0, 2), ("barrier",), ("measure", 3, 3)])                          # :95-132
MEMORY = (4, 4, [g(H, 0), g(CX, 0, 1), g(X, 3), ("measure", 0, 0), ("measure", 1, 1), ("measure", 2, 2),
                 ("measure", 3, 3)])                                                                          # :154-190
MIDCIRCUIT = (3, 3, [g(H, 0), g(CX, 0, 1), ("measure", 0, 0), ("reset", 0), g(H, 0), g(CX, 0, 2),
                     ("measure", 1, 1), g(H, 2), ("measure", 2, 2), ("measure", 0, 0)])


def test_oracle_reference_goldens():
    n, c, ops = REPEATED
    assert set(bref.run(n, c, ops, 100, SEED)) == {0b0110}
    n, c, ops = PARTIAL
    assert set(bref.run(n, c, ops, 100, SEED)) == {0b1011}
    n, c, ops = MEMORY
    mem = bref.run(n, c, ops, 50, SEED)
    assert len(mem) == 50 and set(mem) <= {0b1000, 0b1011}
    for q in range(5):  # :76-93
        assert set(bref.run(5, 1, [g(X, q), ("measure", q, 0)], 100, SEED)) == {1}
    # single shot and mid-circuit paths do not sample
    assert not bref.validate_measure_sampling(REPEATED[2], 1)
    assert not bref.validate_measure_sampling(MIDCIRCUIT[2], 100)
    assert bref.validate_measure_sampling(PARTIAL[2], 100)


def build_circuit(n, c, ops, cregs=None):
    from qiskit_b200 import ClassicalRegister, QuantumCircuit

    qc = QuantumCircuit(n, *(cregs or [ClassicalRegister(c, "cr")]))
    for op in ops:
        if op[0] == "gate":
            qc.unitary(op[1], op[2])
        elif op[0] == "measure":
            qc.measure(op[1], op[2])
        elif op[0] == "reset":
            qc.reset(op[1])
        else:
            qc.barrier()
    return qc


@pytest.mark.gpu
@pytest.mark.parametrize("case,shots", [(REPEATED, 100), (PARTIAL, 100), (MEMORY, 50), (MIDCIRCUIT, 200),
                                        (MIDCIRCUIT, 1), (MEMORY, 1)])
def test_device_matches_restatement_shot_by_shot(case, shots):
    from qiskit_b200.basic_simulator import BasicSimulator

    n, c, ops = case
    res = BasicSimulator().run(build_circuit(n, c, ops), shots=shots, seed_simulator=SEED, memory=True).result()
    want = bref.run(n, c, ops, shots, SEED)
    assert [int(m, 2) for m in res.get_memory()] == want
    counts = res.get_counts(0)
    assert sum(counts.values()) == shots
    assert {int(k, 2): v for k, v in counts.items()} == {k: want.count(k) for k in set(want)}


@pytest.mark.gpu
def test_device_reference_goldens_and_formatting():
    from qiskit_b200 import ClassicalRegister, QuantumCircuit
    from qiskit_b200.basic_simulator import BasicProviderError, BasicSimulator

    backend = BasicSimulator()
    assert backend.run(build_circuit(*REPEATED), shots=100, seed_simulator=SEED).result().get_counts(0) == {"0110": 100}
    assert backend.run(build_circuit(*PARTIAL), shots=100, seed_simulator=SEED).result().get_counts(0) == {"1011": 100}
    n, c, ops = MEMORY
    qc = build_circuit(n, c, ops, [ClassicalRegister(2, "cr0"), ClassicalRegister(2, "cr1")])
    memory = backend.run(qc, shots=50, seed_simulator=SEED, memory=True).result().get_memory()
    assert len(memory) == 50 and set(memory) <= {"10 00", "10 11"}       # :186-190
    # test_unitary (:192-215): X on every qubit through one n-qubit unitary
    xm = X
    for nq in range(1, 5):
        qc = QuantumCircuit(nq, nq)
        qc.unitary(xm, list(range(nq)))
        qc.measure(list(range(nq)), list(range(nq)))
        assert backend.run(qc, shots=1024).result().get_counts(0) == {"1" * nq: 1024}
        xm = np.kron(xm, X)
    # standard gates by name + Bell statistics (:134-152 style)
    qc = QuantumCircuit(2, 2)
    qc.h(0)
    qc.cx(0, 1)
    qc.measure([0, 1], [0, 1])
    counts = backend.run(qc, shots=4000, seed_simulator=7).result().get_counts()
    assert set(counts) == {"00", "11"} and abs(counts["00"] - 2000) < 200
    # initial statevector and its validation (:380-399, :366-380)
    qc = QuantumCircuit(2, 2)
    qc.measure([0, 1], [0, 1])
    init = np.array([0, 0, 0, 1], dtype=complex)
    assert backend.run(qc, shots=10, initial_statevector=init).result().get_counts() == {"11": 10}
    with pytest.raises(BasicProviderError, match="not normalized"):
        backend.run(qc, initial_statevector=np.array([1, 1, 0, 0]))
    with pytest.raises(BasicProviderError, match="incorrect length"):
        backend.run(qc, initial_statevector=np.array([1, 0]))
    with pytest.warns(UserWarning, match="not used by this backend"):
        backend.run(qc, shots=2, bogus=1)


@pytest.mark.gpu
def test_device_random_midcircuit_circuits():
    from helpers import random_unitary
    from qiskit_b200.basic_simulator import BasicSimulator

    rng = np.random.default_rng(3)
    n = 6
    ops = []
    for layer in range(6):
        for _ in range(4):
            k = int(rng.integers(1, 3))
            qs = [int(q) for q in rng.choice(n, size=k, replace=False)]
            ops.append(g(random_unitary(k, rng), *qs))
        q = int(rng.integers(n))
        ops.append(("measure", q, layer % 4))
        if layer % 2:
            ops.append(("reset", int(rng.integers(n))))
    shots = 40
    res = BasicSimulator().run(build_circuit(n, 4, ops), shots=shots, seed_simulator=11, memory=True).result()
    assert [int(m, 2) for m in res.get_memory()] == bref.run(n, 4, ops, shots, 11)
