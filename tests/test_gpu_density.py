"""GPU parity of the device DensityMatrix and the index-gather ops against the
NumPy oracle, through the C ABI.  Tolerance 1e-10 absolute (complex128), as north_star states."""
import types

import numpy as np
import pytest

from helpers import oracle_matrices, random_named_circuit, random_state, random_unitary
from oracle import density_ref as dref
from oracle import gate_matrices as ogm
from oracle import statevector_ref as ref
from qiskit_b200 import DensityMatrix, Pauli, QiskitError, SparsePauliOp, Statevector
from qiskit_b200 import circuit_library as lib
from test_density_oracle import dense_pauli, random_density

pytestmark = pytest.mark.gpu
TOL = 1e-10


def test_constructors_and_data_roundtrip():
    rho = random_density(3, 7)
    dm = DensityMatrix(rho)
    assert dm.num_qubits == 3 and dm.dims() == (2, 2, 2)
    np.testing.assert_array_equal(dm.data, rho)
    psi = random_state(4, 3)
    np.testing.assert_allclose(DensityMatrix(psi).data, dref.from_statevector(psi), atol=TOL)
    np.testing.assert_allclose(DensityMatrix(Statevector(psi)).data, dref.from_statevector(psi), atol=TOL)
    np.testing.assert_allclose(DensityMatrix(psi.reshape(-1, 1)).data, dref.from_statevector(psi), atol=TOL)
    np.testing.assert_array_equal(DensityMatrix.from_int(5, 8).data, dref.basis_state(3, 5))
    plus0 = DensityMatrix.from_label("+0").data
    np.testing.assert_allclose(plus0, dref.from_statevector(np.kron([1, 1], [1, 0]) / np.sqrt(2)), atol=TOL)
    with pytest.raises(QiskitError):
        DensityMatrix(np.zeros((3, 3)))
    with pytest.raises(QiskitError):
        DensityMatrix(np.zeros((2, 4)))


@pytest.mark.parametrize("qargs", [[0], [1], [2], [0, 2], [2, 0], [0, 1, 2], [2, 1, 0], None])
def test_evolve_operator_qargs(qargs):
    # test_densitymatrix.py:227-286
    rng = np.random.default_rng(17)
    rho = random_density(3, 5)
    u = random_unitary(3 if qargs is None else len(qargs), rng)
    dm = DensityMatrix(rho)
    out = dm.evolve(u, qargs)
    np.testing.assert_allclose(out.data, dref.evolve_operator(rho, u, qargs), atol=TOL)
    np.testing.assert_array_equal(dm.data, rho)  # receiver untouched


@pytest.mark.parametrize("n,count", [(4, 40), (6, 80), (7, 60)])
def test_random_circuits_mixed_state(n, count):
    named = random_named_circuit(n, count, seed=100 + n)
    rho = random_density(n, 9, rank=3)
    dm = DensityMatrix(rho)
    dm.apply_gate_list(lib.named_to_matrices(named))
    want = dref.apply_gate_list(rho, oracle_matrices(named))
    np.testing.assert_allclose(dm.data, want, atol=TOL)
    assert abs(dm.trace() - 1) < TOL
    assert abs(dm.purity() - dref.purity(want)) < TOL


def test_circuit_walk_matches_statevector_projector():
    n = 5
    named = random_named_circuit(n, 50, seed=3)
    sv = Statevector.from_int(0, 1 << n)
    sv.apply_gate_list(lib.named_to_matrices(named))
    dm = DensityMatrix.from_int(0, 1 << n)
    dm.apply_gate_list(lib.named_to_matrices(named))
    psi = sv.data
    np.testing.assert_allclose(dm.data, np.outer(psi, psi.conj()), atol=TOL)
    assert abs(dm.purity() - 1) < TOL
    np.testing.assert_allclose(np.abs(dm.to_statevector().inner(sv)), 1.0, atol=1e-8)


def test_expval_goldens_and_64_labels():
    psi = np.array([np.sqrt(2), 0, 0, 0, 0, 0, 0, 1 + 1j]) / 2
    dm = DensityMatrix(psi)
    for label, target in [("XXX", np.sqrt(2) / 2), ("YYY", -np.sqrt(2) / 2), ("ZZZ", 0), ("XYZ", 0), ("YIY", 0)]:
        assert abs(dm.expectation_value(Pauli(label)) - target) < TOL, label
    rho = random_density(2, 1020)
    dm = DensityMatrix(rho)
    for prefix, coeff in [("", 1), ("-", -1), ("i", 1j), ("-i", -1j)]:
        for a in "IXYZ":
            for b in "IXYZ":
                got = dm.expectation_value(Pauli(prefix + a + b))
                want = coeff * np.trace(dense_pauli(a + b) @ rho)
                assert abs(got - want) < TOL, prefix + a + b


@pytest.mark.parametrize("qubits", [[0, 1], [0, 2], [1, 0], [1, 2], [2, 0], [2, 1]])
def test_expval_qargs_and_operator(qubits):
    rho = random_density(3, 1020)
    dm = DensityMatrix(rho)
    z, x, ph = ref.pauli_from_label("XY")
    want = dref.expectation_value_pauli(rho, z, x, ph, qubits)
    assert abs(dm.expectation_value(Pauli("XY"), qubits) - want) < TOL
    assert abs(dm.expectation_value(dense_pauli("XY"), qubits) - want) < TOL
    op = random_unitary(2, np.random.default_rng(1))
    assert abs(dm.expectation_value(op, qubits) - dref.expectation_value_operator(rho, op, qubits)) < TOL


@pytest.mark.parametrize("n,terms", [(4, 30), (7, 200), (9, 64)])
def test_expval_sparse_batch(n, terms):
    rho = random_density(n, 21, rank=4)
    obs = lib.random_pauli_observable(n, terms)
    dm = DensityMatrix(rho)
    want = dref.expectation_value_sparse(rho, obs.paulis.z, obs.paulis.x, obs.coeffs)
    assert abs(dm.expectation_value(obs) - want) < TOL
    ident = SparsePauliOp(["I" * n, "Z" + "I" * (n - 1)], [2.0, 0.5])
    want = dref.expectation_value_sparse(rho, ident.paulis.z, ident.paulis.x, ident.coeffs)
    assert abs(dm.expectation_value(ident) - want) < TOL


def test_probabilities_goldens_and_marginals():
    ghz = np.zeros(8, complex)
    ghz[0] = ghz[7] = 1 / np.sqrt(2)
    dm = DensityMatrix(ghz)
    for qargs in [[0, 1, 2], [2, 1, 0], [1, 2, 0], [1, 0, 2]]:
        np.testing.assert_allclose(dm.probabilities(qargs), [0.5, 0, 0, 0, 0, 0, 0, 0.5], atol=TOL)
    np.testing.assert_allclose(dm.probabilities([2, 1]), [0.5, 0, 0, 0.5], atol=TOL)
    np.testing.assert_allclose(dm.probabilities([1]), [0.5, 0.5], atol=TOL)
    assert dm.probabilities_dict() == pytest.approx({"000": 0.5, "111": 0.5})
    rho = random_density(6, 2)
    dm = DensityMatrix(rho)
    for qargs in [None, [0], [5, 0], [3, 1, 4], [5, 4, 3, 2, 1, 0], [2, 0, 1, 5]]:
        np.testing.assert_allclose(dm.probabilities(qargs), dref.probabilities(rho, qargs), atol=TOL)
    np.testing.assert_allclose(dm.probabilities([1, 2], decimals=3), dref.probabilities(rho, [1, 2], 3), atol=TOL)


def test_sampling_follows_reference_rng_calls():
    # quantum_state.py:258-265: labels drawn with Generator.choice(p=probs); same seed -> same memory
    rho = random_density(4, 8)
    dm = DensityMatrix(rho)
    dm.seed(42)
    got = dm.sample_memory(500, [0, 2])
    probs = dref.probabilities(rho, [0, 2])
    labels = ref.index_to_ket_strings(np.arange(4), 2)
    want = np.random.default_rng(42).choice(labels, p=dm.probabilities([0, 2]), size=500)
    np.testing.assert_array_equal(got, want)
    np.testing.assert_allclose(dm.probabilities([0, 2]), probs, atol=TOL)
    dm.seed(7)
    counts = dm.sample_counts(2000)
    assert sum(counts.values()) == 2000 and all(len(k) == 4 for k in counts)


def test_reset_measure_initialize():
    state = DensityMatrix(np.diag([0.5, 0, 0, 0.5]))
    np.testing.assert_allclose(state.reset().data, np.diag([1, 0, 0, 0]), atol=TOL)
    np.testing.assert_allclose(state.reset([0, 1]).data, np.diag([1, 0, 0, 0]), atol=TOL)
    np.testing.assert_allclose(state.reset([0]).data, np.diag([0.5, 0, 0.5, 0]), atol=TOL)
    np.testing.assert_allclose(state.reset([1]).data, np.diag([0.5, 0.5, 0, 0]), atol=TOL)
    rho = random_density(5, 4)
    dm = DensityMatrix(rho)
    for qargs in [[0], [3], [4, 1], [2, 0, 3]]:
        np.testing.assert_allclose(dm.reset(qargs).data, dref.reset(rho, qargs), atol=TOL)
    # measure: same RNG call as the reference, post-state = P rho P / p
    dm.seed(11)
    outcome, post = dm.measure([1, 3])
    probs = dref.probabilities(rho, [1, 3])
    sample = int(np.random.default_rng(11).choice(4, p=dm.probabilities([1, 3]), size=1)[0])
    assert outcome == format(sample, "02b")
    proj = np.zeros(4, complex)
    proj[sample] = 1 / np.sqrt(probs[sample])
    np.testing.assert_allclose(post.data, dref.evolve_operator(rho, np.diag(proj), [1, 3]), atol=1e-9)
    assert abs(post.trace() - 1) < 1e-9
    # Initialize on a subset: reset + |psi><0|
    init = random_state(2, 5)
    inst = types.SimpleNamespace(name="initialize", num_qubits=2, params=list(init), definition=None)
    out = dm.evolve(inst, [3, 0])
    mat = np.zeros((4, 4), complex)
    mat[:, 0] = init
    np.testing.assert_allclose(out.data, dref.evolve_operator(dref.reset(rho, [3, 0]), mat, [3, 0]), atol=TOL)


def test_algebra_tensor_permutations():
    a, b = random_density(2, 1), random_density(3, 2)
    da, db = DensityMatrix(a), DensityMatrix(b)
    np.testing.assert_allclose(da.tensor(db).data, np.kron(a, b), atol=TOL)
    np.testing.assert_allclose(da.expand(db).data, np.kron(b, a), atol=TOL)
    np.testing.assert_allclose(db.reverse_qargs().data, dref.reverse_qargs(b), atol=0)
    np.testing.assert_allclose(db.partial_transpose([0, 2]).data, dref.partial_transpose(b, [0, 2]), atol=0)
    np.testing.assert_allclose(db.conjugate().data, np.conj(b), atol=0)
    b2 = random_density(3, 3)
    np.testing.assert_allclose((db + DensityMatrix(b2)).data, b + b2, atol=0)
    np.testing.assert_allclose((db - DensityMatrix(b2)).data, b - b2, atol=0)
    np.testing.assert_allclose((0.25j * db).data, 0.25j * b, atol=1e-15)
    assert abs(db.trace() - np.trace(b)) < TOL and abs(db.purity() - dref.purity(b)) < TOL
    assert db.is_valid() and not (2 * db).is_valid()


def test_statevector_device_ops():
    # tensor / expand / reverse_qargs / conjugate / add without a host round trip
    u, v = random_state(3, 1), random_state(5, 2)
    su, sv = Statevector(u), Statevector(v)
    np.testing.assert_allclose(su.tensor(sv).data, np.kron(u, v), atol=1e-15)
    np.testing.assert_allclose(su.expand(sv).data, np.kron(v, u), atol=1e-15)
    want = np.reshape(np.transpose(np.reshape(v, (2,) * 5), tuple(range(4, -1, -1))), -1)  # statevector.py:524-530
    np.testing.assert_array_equal(sv.reverse_qargs().data, want)
    np.testing.assert_array_equal(sv.conjugate().data, np.conj(v))
    v2 = random_state(5, 9)
    np.testing.assert_array_equal((sv + Statevector(v2)).data, v + v2)
    np.testing.assert_array_equal((sv - Statevector(v2)).data, v - v2)
    # from_circuit layout permutation (statevector.py:156-178)
    n = 5
    named = random_named_circuit(n, 30, seed=8)
    layout_idx = [3, 0, 4, 1, 2]

    class _Circ:
        num_qubits = n
        layout = types.SimpleNamespace(final_index_layout=lambda: layout_idx)

        def to_instruction(self):
            return self

    psi = ref.apply_gate_list(ref.basis_state(n), oracle_matrices(named))
    base = Statevector(psi)
    reversed_perm = [n - 1 - layout_idx[n - 1 - i] for i in range(n)]
    want = np.reshape(np.transpose(np.reshape(psi, (2,) * n), reversed_perm), -1)
    np.testing.assert_array_equal(base._permuted(layout_idx).data, want)


def test_larger_density_matrix_properties():
    # 12 qubits = a 24-qubit handle (268 MB): size-independent properties through the fused gate path
    n = 12
    gates = lib.quantum_volume(n, 4, seed=1)
    dm = DensityMatrix.from_int(0, 1 << n)
    dm.apply_gate_list(gates)
    sv = Statevector.from_int(0, 1 << n)
    sv.apply_gate_list(gates)
    assert abs(dm.trace() - 1) < TOL and abs(dm.purity() - 1) < TOL
    obs = lib.random_pauli_observable(n, 100)
    assert abs(dm.expectation_value(obs) - sv.expectation_value(obs)) < TOL
    np.testing.assert_allclose(dm.probabilities([0, 5, 11]), sv.probabilities([0, 5, 11]), atol=TOL)
    mixed = dm.reset([0, 1])
    assert abs(mixed.trace() - 1) < TOL and mixed.purity().real < 1 - 1e-3


def test_density_reset_and_measure_on_all_twelve_qubits():
    """reset / measure with more qubits than one kernel call or one diagonal takes:
    the channel and the projector factor per qubit, so the host applies them in chunks."""
    n = 12
    rng = np.random.default_rng(4)
    gates = [(ogm.ry(float(t)), [q]) for q, t in enumerate(rng.uniform(0.3, 2.8, n))] + \
            [(ogm.cx(), [q, q + 1]) for q in range(0, n - 1, 2)]
    rho = DensityMatrix.from_int(0, 1 << n)
    rho.apply_gate_list(gates)
    got = rho.reset(list(range(n))).data          # every qubit reset -> |0..0><0..0| (trace 1)
    assert abs(got[0, 0] - 1.0) < 1e-10 and np.count_nonzero(np.abs(got) > 1e-10) == 1
    rho.seed(9)
    outcome, post = rho.measure()
    idx = int(outcome, 2)
    d = post.data
    assert abs(d[idx, idx] - 1.0) < 1e-9 and np.count_nonzero(np.abs(d) > 1e-9) == 1


@pytest.mark.parametrize("n,qargs", [(1, None), (3, [0]), (3, [2]), (3, [0, 2]), (3, [2, 0]), (3, [2, 1, 0]), (6, [4, 1]),
                                     (7, [5, 0, 3]), (7, None)])
def test_evolve_by_a_channel(n, qargs):
    """DensityMatrix.evolve(QuantumChannel) (densitymatrix.py:329-336 -> SuperOp._evolve): the superoperator
    as one dense 2k-qubit matrix on the handle, against the oracle's restatement (pinned to
    test_superop.py:112-129 in tests/test_density_oracle.py)."""
    from qiskit_b200.channels import SuperOp

    rng = np.random.default_rng(n * 11 + (len(qargs) if qargs else 0))
    k = n if qargs is None else len(qargs)
    if k > 6:
        with pytest.raises(QiskitError):
            DensityMatrix(random_density(n, 3)).evolve(SuperOp(np.eye(1 << (2 * k))), qargs)
        return
    rho = random_density(n, 3)
    kraus = [rng.standard_normal((1 << k, 1 << k)) + 1j * rng.standard_normal((1 << k, 1 << k)) for _ in range(2)]
    chan = SuperOp.from_kraus(kraus)
    np.testing.assert_allclose(chan.data, dref.kraus_to_superop(kraus), atol=0)
    dm = DensityMatrix(rho)
    got = dm.evolve(chan, qargs).data
    np.testing.assert_allclose(got, dref.evolve_superop(rho, chan.data, qargs), atol=TOL * max(1.0, np.abs(got).max()))
    np.testing.assert_array_equal(dm.data, rho)                      # evolve never mutates
    dense = rng.standard_normal((1 << (2 * k),) * 2) + 1j * rng.standard_normal((1 << (2 * k),) * 2)   # non-CP map
    got = dm.evolve(SuperOp(dense), qargs).data
    np.testing.assert_allclose(got, dref.evolve_superop(rho, dense, qargs), atol=TOL * max(1.0, np.abs(got).max()))


def test_channel_goldens_on_device():
    """test_superop.py:112-129 on the device."""
    from qiskit_b200.channels import SuperOp

    rho = DensityMatrix(np.array([[0, 0], [0, 1]], dtype=complex))
    h = np.array([[1, 1], [1, -1]]) / np.sqrt(2)
    np.testing.assert_allclose(rho.evolve(SuperOp(np.eye(4))).data, [[0, 0], [0, 1]], atol=TOL)
    np.testing.assert_allclose(rho.evolve(SuperOp(np.kron(h.conj(), h))).data, np.array([[1, -1], [-1, 1]]) / 2, atol=TOL)
    depol1 = np.array([[1, 0, 0, 1], [0, 0, 0, 0], [0, 0, 0, 0], [1, 0, 0, 1]]) / 2
    np.testing.assert_allclose(rho.evolve(SuperOp(depol1)).data, np.eye(2) / 2, atol=TOL)
