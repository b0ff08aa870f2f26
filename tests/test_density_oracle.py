"""The DensityMatrix oracle against the reference's own golden values (CPU only).

Expected numbers are the constants asserted by the reference's
test/python/quantum_info/states/test_densitymatrix.py (line numbers in each test); the
64-label case uses the reference's own criterion (:1078-1163: Pauli kernel == dense matrix).
"""
import numpy as np
import pytest

from oracle import density_ref as dref
from oracle import statevector_ref as ref
from helpers import random_unitary

PAULI = {"I": np.eye(2), "X": np.array([[0, 1], [1, 0]]), "Y": np.array([[0, -1j], [1j, 0]]),
         "Z": np.diag([1.0, -1.0])}


def random_density(n, seed, rank=None):
    """Ginibre ensemble like quantum_info/states/random.py:82-120 (own seeds; not bit-identical)."""
    rng = np.random.default_rng(seed)
    dim = 1 << n
    rank = rank or dim
    g = rng.standard_normal((dim, rank)) + 1j * rng.standard_normal((dim, rank))
    rho = g @ g.conj().T
    return rho / np.trace(rho)


def dense_pauli(label):
    mat = np.eye(1)
    for ch in label:  # leftmost character = highest qubit
        mat = np.kron(mat, PAULI[ch])
    return mat


def test_expval_goldens():
    # test_densitymatrix.py:971-1000
    psi = np.array([1, 0, 0, 1]) / np.sqrt(2)
    rho = dref.from_statevector(psi)
    for label, target in [("II", 1), ("XX", 1), ("YY", -1), ("ZZ", 1), ("IX", 0), ("YZ", 0), ("ZX", 0), ("YI", 0)]:
        z, x, ph = ref.pauli_from_label(label)
        assert abs(dref.expectation_value_pauli(rho, z, x, ph) - target) < 1e-12, label
    psi = np.array([np.sqrt(2), 0, 0, 0, 0, 0, 0, 1 + 1j]) / 2
    rho = dref.from_statevector(psi)
    for label, target in [("XXX", np.sqrt(2) / 2), ("YYY", -np.sqrt(2) / 2), ("ZZZ", 0), ("XYZ", 0), ("YIY", 0)]:
        z, x, ph = ref.pauli_from_label(label)
        assert abs(dref.expectation_value_pauli(rho, z, x, ph) - target) < 1e-12, label


@pytest.mark.parametrize("prefix", ["", "-", "i", "-i"])
def test_expval_64_labels_vs_dense(prefix):
    # test_densitymatrix.py:1078-1163: all {,-,i,-i} x {I,X,Y,Z}^2 labels against Tr(P rho)
    rho = random_density(2, 1020)
    coeff = {"": 1, "-": -1, "i": 1j, "-i": -1j}[prefix]
    for a in "IXYZ":
        for b in "IXYZ":
            z, x, ph = ref.pauli_from_label(prefix + a + b)
            got = dref.expectation_value_pauli(rho, z, x, ph)
            want = coeff * np.trace(dense_pauli(a + b) @ rho)
            assert abs(got - want) < 1e-12, prefix + a + b


@pytest.mark.parametrize("qubits", [[0, 1], [0, 2], [1, 0], [1, 2], [2, 0], [2, 1]])
def test_expval_qargs(qubits):
    # test_densitymatrix.py:1165-1172
    rho = random_density(3, 1020)
    z, x, ph = ref.pauli_from_label("XY")
    got = dref.expectation_value_pauli(rho, z, x, ph, qubits)
    want = dref.expectation_value_operator(rho, dense_pauli("XY"), qubits)
    assert abs(got - want) < 1e-12


def test_probabilities_goldens():
    # GHZ (:462-487) and W (:489-518)
    ghz = np.zeros(8, complex)
    ghz[0] = ghz[7] = 1 / np.sqrt(2)
    rho = dref.from_statevector(ghz)
    for qargs in [[0, 1, 2], [2, 1, 0], [1, 2, 0], [1, 0, 2]]:
        np.testing.assert_allclose(dref.probabilities(rho, qargs), [0.5, 0, 0, 0, 0, 0, 0, 0.5], atol=1e-12)
    for qargs in [[0, 1], [2, 1], [1, 2]]:
        np.testing.assert_allclose(dref.probabilities(rho, qargs), [0.5, 0, 0, 0.5], atol=1e-12)
    w = np.zeros(8, complex)
    w[[1, 2, 4]] = 1 / np.sqrt(3)
    rho = dref.from_statevector(w)
    np.testing.assert_allclose(dref.probabilities(rho, [0, 1]), [1 / 3, 1 / 3, 1 / 3, 0], atol=1e-12)
    np.testing.assert_allclose(dref.probabilities(rho, [2]), [2 / 3, 1 / 3], atol=1e-12)
    # '+0' (:430-460): qubit 1 in |+>, qubit 0 in |0>
    plus0 = np.kron(np.array([1, 1]) / np.sqrt(2), np.array([1, 0]))
    rho = dref.from_statevector(plus0)
    np.testing.assert_allclose(dref.probabilities(rho, [0, 1]), [0.5, 0, 0.5, 0], atol=1e-12)
    np.testing.assert_allclose(dref.probabilities(rho, [1, 0]), [0.5, 0.5, 0, 0], atol=1e-12)


def test_reset_goldens():
    # test_densitymatrix.py:830-857
    state = np.diag([0.5, 0, 0, 0.5]).astype(complex)
    np.testing.assert_allclose(dref.reset(state), np.diag([1, 0, 0, 0]))
    np.testing.assert_allclose(dref.reset(state, [0, 1]), np.diag([1, 0, 0, 0]))
    np.testing.assert_allclose(dref.reset(state, [0]), np.diag([0.5, 0, 0.5, 0]))
    np.testing.assert_allclose(dref.reset(state, [1]), np.diag([0.5, 0.5, 0, 0]))


def test_evolve_matches_full_matrix_and_statevector():
    # :227-286: evolve on subsystems == Kronecker-expanded U rho U^dagger; pure states track the
    # statevector oracle
    rng = np.random.default_rng(5)
    rho = random_density(3, 11)
    for qargs in [[0], [1], [2], [0, 2], [2, 0], [0, 1, 2], [2, 1, 0]]:
        u = random_unitary(len(qargs), rng)
        full = np.empty((8, 8), complex)
        eye = np.eye(8, dtype=complex)
        for c in range(8):
            full[:, c] = ref.evolve_operator(eye[:, c].copy(), u, qargs)
        np.testing.assert_allclose(dref.evolve_operator(rho, u, qargs), full @ rho @ full.conj().T, atol=1e-12)
    psi = ref.basis_state(3)
    pure = dref.basis_state(3)
    for qargs in [[0], [1, 2], [2, 0], [1]]:
        u = random_unitary(len(qargs), rng)
        psi = ref.evolve_operator(psi, u, qargs)
        pure = dref.evolve_operator(pure, u, qargs)
    np.testing.assert_allclose(pure, np.outer(psi, psi.conj()), atol=1e-12)
    assert abs(dref.purity(pure) - 1) < 1e-12


def test_reverse_and_partial_transpose():
    a, b = random_density(1, 1), random_density(2, 2)
    rho = np.kron(a, b)  # a on qubit 2, b on qubits 1..0
    rev = dref.reverse_qargs(rho)
    b_rev = dref.reverse_qargs(b)
    np.testing.assert_allclose(rev, np.kron(b_rev, a), atol=1e-14)
    # partial transpose of a product state transposes that factor
    np.testing.assert_allclose(dref.partial_transpose(rho, [2]), np.kron(a.T, b), atol=1e-14)
    np.testing.assert_allclose(dref.partial_transpose(rho, [0, 1]), np.kron(a, b.T), atol=1e-14)


# ---- channels: SuperOp._evolve (operators/channel/superop.py:234-292) --------------------------------

def test_superop_evolve_reference_goldens():
    """test_superop.py:112-129: identity, Hadamard and completely depolarizing channel on |1><1|."""
    rho = np.array([[0, 0], [0, 1]], dtype=complex)
    np.testing.assert_allclose(dref.evolve_superop(rho, np.eye(4)), rho, atol=1e-15)
    h = np.array([[1, 1], [1, -1]]) / np.sqrt(2)
    np.testing.assert_allclose(dref.evolve_superop(rho, np.kron(h.conj(), h)), np.array([[1, -1], [-1, 1]]) / 2, atol=1e-15)
    depol1 = np.array([[1, 0, 0, 1], [0, 0, 0, 0], [0, 0, 0, 0], [1, 0, 0, 1]]) / 2      # channel_test_case.py:68-72, p = 1
    np.testing.assert_allclose(dref.evolve_superop(rho, depol1), np.eye(2) / 2, atol=1e-15)


@pytest.mark.parametrize("qargs", [[0], [1], [2], [0, 2], [2, 0], [0, 1, 2], [2, 1, 0]])
def test_superop_evolve_subsystem_equals_embedded_kraus(qargs):
    """test_superop.py:131-190 checks subsystems against tensored superoperators; here the same property
    through Kraus operators: evolving `qargs` = sum_i (K_i on qargs) rho (K_i on qargs)^dagger."""
    rng = np.random.default_rng(len(qargs) * 7 + qargs[0])
    rho = random_density(3, 5)
    k = len(qargs)
    kraus = [rng.standard_normal((1 << k, 1 << k)) + 1j * rng.standard_normal((1 << k, 1 << k)) for _ in range(3)]
    got = dref.evolve_superop(rho, dref.kraus_to_superop(kraus), qargs)
    want = np.zeros_like(rho)
    for op in kraus:
        cols = [ref.evolve_operator(rho[:, c].copy(), op, qargs) for c in range(rho.shape[1])]     # K rho
        left = np.stack(cols, axis=1)
        rows = [ref.evolve_operator(left[r, :].conj().copy(), op, qargs).conj() for r in range(rho.shape[0])]  # (K rho) K^dagger
        want += np.stack(rows, axis=0)
    np.testing.assert_allclose(got, want, atol=1e-12)
    if qargs == [0, 1, 2]:
        np.testing.assert_allclose(dref.evolve_superop(rho, dref.kraus_to_superop(kraus)), want, atol=1e-12)
