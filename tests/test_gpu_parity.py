"""GPU parity: the CUDA engine (through the C ABI) against the CPU oracle.

Tolerances: amplitudes and expectation values 1e-10 absolute (BASELINE.json
north_star); sample indices bit-exact against the oracle's statement of the
engine's summation tree, and counted (+-1 only) against NumPy's own recipe.
"""
import itertools
import json
import os

import numpy as np
import pytest

from oracle import gate_matrices as ogm
from oracle import statevector_ref as ref
from qiskit_b200 import (Pauli, PauliList, QiskitError, QuantumCircuit, SparsePauliOp, Statevector,
                         StatevectorEstimator, StatevectorSampler, ClassicalRegister, Parameter)
from qiskit_b200 import circuit_library as lib
from qiskit_b200.gates import gate_matrix

from helpers import oracle_matrices, random_named_circuit, random_state, random_unitary

pytestmark = pytest.mark.gpu
ATOL = 1e-10
KATS = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_kats.json")))


def engine_run(psi0, gates, **kw):
    sv = Statevector(psi0)
    sv.apply_gate_list(gates, **kw)
    return sv.data


# ----------------------------------------------------------- A1: gates -----

@pytest.mark.parametrize("n", range(1, 13))
def test_apply_matrix_all_ordered_qargs(n):
    """Every ordered qargs of length 1 and 2, random unitaries."""
    rng = np.random.default_rng(n)
    psi = random_state(n, 100 + n)
    sv0 = Statevector(psi)
    for k in (1, 2):
        if k > n:
            continue
        for qargs in itertools.permutations(range(n), k):
            u = random_unitary(k, rng)
            got = sv0.evolve(u, list(qargs)).data
            want = ref.evolve_operator(psi, u, list(qargs))
            assert np.max(np.abs(got - want)) < ATOL, (n, qargs)
    np.testing.assert_array_equal(sv0.data, psi)   # evolve never mutates (statevector.py:458)


@pytest.mark.parametrize("n,k", [(3, 3), (5, 3), (6, 4), (9, 5), (12, 6), (13, 3), (14, 7), (12, 12)])
def test_apply_wide_dense(n, k):
    rng = np.random.default_rng(n * 31 + k)
    psi = random_state(n, n + k)
    for _ in range(2):
        qargs = [int(q) for q in rng.choice(n, size=k, replace=False)]
        u = random_unitary(k, rng)
        got = Statevector(psi).evolve(u, qargs).data
        assert np.max(np.abs(got - ref.evolve_operator(psi, u, qargs))) < ATOL


def test_full_system_operator_and_nonunitary():
    psi = random_state(4, 3)
    rng = np.random.default_rng(0)
    m = rng.standard_normal((16, 16)) + 1j * rng.standard_normal((16, 16))
    got = Statevector(psi).evolve(m).data
    assert np.max(np.abs(got - m @ psi)) < 1e-9
    proj = np.diag([0, 1.5]).astype(complex)
    got = Statevector(psi).evolve(proj, [2]).data
    assert np.max(np.abs(got - ref.evolve_operator(psi, proj, [2]))) < ATOL


@pytest.mark.parametrize("n", [2, 3, 5, 8, 11, 13, 14, 16])
@pytest.mark.parametrize("fusion", [True, False])
def test_random_named_circuits(n, fusion):
    """Mixed standard gates (dense, diagonal, permutation, 3-qubit) through the fuser and
    the tile kernel, against the oracle's own gate table."""
    named = random_named_circuit(n, 60, seed=n * 7 + fusion)
    psi = random_state(n, n)
    got = engine_run(psi, lib.named_to_matrices(named), no_fusion=not fusion)
    want = ref.apply_gate_list(psi, oracle_matrices(named))
    assert np.max(np.abs(got - want)) < ATOL


@pytest.mark.parametrize("names", [["rz", "cp", "rzz", "cz", "z", "s", "t", "crz", "ccz", "p"],
                                   ["cx", "swap", "x", "y", "ccx", "cswap", "cy", "iswap"],
                                   ["h", "ry", "rx", "sx"], ["cx", "ry", "rz"]])
def test_gate_families(names):
    """Diagonal-only (phase tables), permutation-only, real-dense-only and ansatz-like streams."""
    for n in (4, 10, 15):
        named = random_named_circuit(n, 80, seed=n, names=names)
        psi = random_state(n, 5 * n)
        got = engine_run(psi, lib.named_to_matrices(named))
        want = ref.apply_gate_list(psi, oracle_matrices(named))
        assert np.max(np.abs(got - want)) < ATOL, names


def test_diagonal_entry_and_global_phase():
    psi = random_state(6, 1)
    d = np.exp(1j * np.random.default_rng(2).uniform(0, 6, 8))
    sv = Statevector(psi)
    sv.apply_gate_list([("diag", d, [4, 0, 2]), ("phase", 0.3)])
    want = ref.apply_gate_list(psi, [(np.diag(d), [4, 0, 2]), ("phase", 0.3)])
    assert np.max(np.abs(sv.data - want)) < ATOL
    sv = Statevector(psi)
    sv.apply_gate_list([("phase", 1.1)])
    assert np.max(np.abs(sv.data - psi * np.exp(1.1j))) < ATOL


@pytest.mark.parametrize("tile_bits,low_bits,threads", [(10, 3, 128), (11, 4, 256), (12, 4, 256),
                                                        (12, 2, 512), (13, 5, 512)])
def test_tile_configurations(tile_bits, low_bits, threads):
    n = 17
    named = random_named_circuit(n, 50, seed=tile_bits)
    psi = random_state(n, 9)
    sv = Statevector(psi)
    sv.set_option("tile_bits", tile_bits)
    sv.set_option("low_bits", low_bits)
    sv.set_option("tile_threads", threads)
    sv.apply_gate_list(lib.named_to_matrices(named))
    want = ref.apply_gate_list(psi, oracle_matrices(named))
    assert np.max(np.abs(sv.data - want)) < ATOL


@pytest.mark.parametrize("fused,direct_store", [(1, 1), (1, 0), (0, 0)])
def test_fused_pass_paths(fused, direct_store):
    """k_fused_pass (persistent, pipelined; with and without the direct register->global store of
    the last phase) and k_cluster_pass must agree with the oracle on deep mixed circuits."""
    for n, count, seed, names in [(12, 150, 1, None), (13, 200, 2, None), (17, 300, 3, None),
                                  (16, 300, 4, ["cx", "ry", "rz", "h", "cz", "swap"]),
                                  (15, 200, 5, ["cp", "h", "rzz", "rx", "ccx", "y", "cy"])]:
        named = random_named_circuit(n, count, seed=seed, names=names)
        psi = random_state(n, 11 * n + seed)
        sv = Statevector(psi)
        sv.set_option("use_fused", fused)
        sv.set_option("direct_store", direct_store)
        sv.apply_gate_list(lib.named_to_matrices(named))
        want = ref.apply_gate_list(psi, oracle_matrices(named))
        assert np.max(np.abs(sv.data - want)) < ATOL, (n, seed)
        assert sv.counters()["state_passes"] < count


@pytest.mark.parametrize("n", [1, 2, 5, 12, 13, 17])
def test_product_state_prologue_and_lazy_basis(n):
    """qsv_init_basis is lazy and qsv_apply_gates folds the leading one-qubit gates of every qubit
    into a product state written in one sweep: same amplitudes as gate-by-gate application."""
    rng = np.random.default_rng(n)
    names = ["ry", "rz", "h", "rx", "u", "x", "s"] + (["cx", "cz", "rzz"] if n > 1 else [])
    for start in (0, (1 << n) - 1, int(rng.integers(1 << n))):
        named = random_named_circuit(n, 8 * n + 5, seed=n + start, names=names)
        gates = lib.named_to_matrices(named)
        want = ref.apply_gate_list(ref.basis_state(n, start), oracle_matrices(named))
        for product_init, fold in ((1, 1), (1, 0), (0, 1)):
            sv = Statevector.from_int(start, 1 << n)
            sv.set_option("product_init", product_init)
            sv.set_option("product_fold", fold)   # first fused pass streams the template tile / separate sweep
            w0 = sv.counters()["write_passes"]      # pooled handles keep cumulative counters
            sv.apply_gate_list(gates)
            assert np.max(np.abs(sv.data - want)) < ATOL, (n, start, product_init)
            assert sv.counters()["write_passes"] == w0 + 1
    # every reader materialises a pending |index>
    sv = Statevector.from_int(3 % (1 << n), 1 << n)
    assert abs(sv._norm2() - 1) < 1e-15 and sv.data[3 % (1 << n)] == 1
    sv = Statevector.from_int(0, 1 << n)
    assert sv.probabilities()[0] == 1.0
    sv = Statevector.from_int(0, 1 << n)
    cp = sv.copy()
    assert cp.data[0] == 1 and sv.data[0] == 1
    sv = Statevector.from_int(0, 1 << n)
    assert (sv * 2j).data[0] == 2j


@pytest.mark.parametrize("n,depth", [(20, 20), (22, 4), (24, 2)])
def test_quantum_volume_spot_checks(n, depth):
    """configs[0] circuit in full (20 q, depth 20) and larger spot checks."""
    gates = lib.quantum_volume(n, depth, seed=0)
    sv = Statevector.from_int(0, 1 << n)
    sv.apply_gate_list(gates)
    want = ref.apply_gate_list(ref.basis_state(n), gates)
    assert np.max(np.abs(sv.data - want)) < ATOL
    c = sv.counters()
    assert c["state_passes"] < len(gates)           # fusion made fewer passes than gates
    sv2 = Statevector.from_int(0, 1 << n)
    sv2.apply_gate_list(gates[:10], no_fusion=True)
    assert sv2.counters()["state_passes"] == 10


def test_qft_and_tfim_and_su2_streams():
    n = 16
    for named in (lib.qft_named(n), lib.tfim_named(n, steps=3),
                  lib.efficient_su2_named(n, 2, np.random.default_rng(0).uniform(-3, 3, 2 * n * 3))):
        sv = Statevector.from_int(0, 1 << n)
        sv.apply_gate_list(lib.named_to_matrices(named))
        want = ref.apply_gate_list(ref.basis_state(n), oracle_matrices(named))
        assert np.max(np.abs(sv.data - want)) < ATOL


def test_evolve_copy_semantics_and_circuit_walk():
    """test_statevector.py:339-446 (evolve must not modify) and nested definitions."""
    inner = QuantumCircuit(2, name="inner", global_phase=0.3)
    inner.h(0); inner.cx(0, 1)
    outer = QuantumCircuit(3, global_phase=0.2)
    outer.append(inner.to_instruction(), [2, 0])
    outer.barrier()
    outer.rz(0.7, 1)
    psi = random_state(3, 4)
    sv = Statevector(psi)
    out = sv.evolve(outer)
    np.testing.assert_array_equal(sv.data, psi)
    want = np.exp(0.5j) * ref.apply_gate_list(psi, [(ogm.h(), [2]), (ogm.cx(), [2, 0]), (ogm.rz(0.7), [1])])
    assert np.max(np.abs(out.data - want)) < ATOL
    # evolve with qargs: the circuit acts on a subset
    sub = QuantumCircuit(2); sub.h(0); sub.cx(0, 1)
    psi5 = random_state(5, 6)
    out = Statevector(psi5).evolve(sub, [3, 1])
    want = ref.apply_gate_list(psi5, [(ogm.h(), [3]), (ogm.cx(), [3, 1])])
    assert np.max(np.abs(out.data - want)) < ATOL
    c = sv.copy()
    assert c == sv and c is not sv


def test_global_phase_kat():
    k = KATS["global_phase"]
    qc = QuantumCircuit(1, global_phase=k["gates"][0][1][0]); qc.p(0.5, 0)
    got = Statevector(qc).data
    np.testing.assert_allclose(got, np.array(k["expected_re"]) + 1j * np.array(k["expected_im"]), atol=1e-8)


def test_errors():
    sv = Statevector.from_int(0, 8)
    with pytest.raises(QiskitError):
        sv.evolve(np.eye(4), [0])
    with pytest.raises(QiskitError):
        sv.evolve(np.eye(2), [5])
    with pytest.raises(QiskitError):
        sv.evolve(np.eye(4), [1, 1])
    with pytest.raises(QiskitError):
        Statevector(np.ones(3))
    from qiskit_b200.circuit import Instruction
    with pytest.raises(QiskitError, match="Cannot apply Instruction"):
        sv.evolve(Instruction("mystery", 3))
    bad = QuantumCircuit(3, 1); bad.measure(0, 0)
    with pytest.raises(QiskitError, match="classical bits"):
        sv.evolve(bad)


# --------------------------------------------------- A3-A5: expectation -----

def test_expval_state_tables_and_sparse():
    for kat in KATS["expval_state"]:
        sv = Statevector(np.array(kat["state_re"]) + 1j * np.array(kat["state_im"]))
        for label, target in kat["labels"].items():
            assert abs(sv.expectation_value(Pauli(label)) - target) < 1e-7, label
    k = KATS["expval_sparse"]
    sv = Statevector(np.array(k["state_re"]) + 1j * np.array(k["state_im"]))
    op = SparsePauliOp(k["labels"], np.array(k["coeffs_re"]) + 1j * np.array(k["coeffs_im"]))
    assert abs(sv.expectation_value(op) - complex(k["expected_re"], k["expected_im"])) < 1e-7


@pytest.mark.parametrize("prefix", ["", "-", "i", "-i"])
def test_expval_pauli_64_labels(prefix):
    """test_statevector.py:1243-1316 against the oracle (which is pinned to the dense matrix)."""
    rng = np.random.default_rng(1020)
    vec = rng.standard_normal(4).astype(complex); vec += 1j * rng.standard_normal(4); vec /= np.linalg.norm(vec)
    sv = Statevector(vec)
    for a in "IXYZ":
        for b in "IXYZ":
            p = Pauli(prefix + a + b)
            want = ref.expectation_value_pauli(vec, p.z, p.x, p.phase)
            assert abs(sv.expectation_value(p) - want) < ATOL, prefix + a + b


@pytest.mark.parametrize("qubits", [[0, 1], [0, 2], [1, 0], [1, 2], [2, 0], [2, 1]])
def test_expval_pauli_qargs(qubits):
    psi = random_state(3, 1020)
    sv = Statevector(psi)
    for a in "IXYZ":
        for b in "IXYZ":
            p = Pauli(a + b)
            want = ref.expectation_value_pauli(psi, p.z, p.x, 0, qubits)
            assert abs(sv.expectation_value(p, qubits) - want) < ATOL


def test_expval_identity_non_normalised():
    """test_statevector.py:1328-1377: <I> is the squared norm, also off-normalisation."""
    psi = random_state(5, 2) * 1.7
    sv = Statevector(psi)
    assert abs(sv.expectation_value(Pauli("IIIII")) - 1.7**2) < ATOL
    assert abs(sv.expectation_value(SparsePauliOp(["IIIII"], [2.0])) - 2 * 1.7**2) < ATOL


@pytest.mark.parametrize("n,terms", [(1, 4), (2, 16), (6, 40), (10, 64), (13, 100), (20, 64)])
def test_expval_random_sparse_ops(n, terms):
    """configs[0] observable shape: random Pauli terms, all evaluated in one batch call."""
    psi = random_state(n, n)
    op = lib.random_pauli_observable(n, terms, seed=n, coeff_seed=n + 1)
    want = ref.expectation_value_sparse(psi, op.paulis.z, op.paulis.x, op.coeffs)
    got = Statevector(psi).expectation_value(op)
    assert abs(got - want) < ATOL


def test_expval_heisenberg_and_tfim_ops():
    n = 12
    psi = random_state(n, 8)
    sv = Statevector(psi)
    for op in (lib.heisenberg_observable(n, 40), lib.tfim_observable(n)):
        want = ref.expectation_value_sparse(psi, op.paulis.z, op.paulis.x, op.coeffs)
        assert abs(sv.expectation_value(op) - want) < ATOL


@pytest.mark.parametrize("n", [8, 9, 14])
def test_expval_many_z_terms_and_qwc_families(n):
    """> 512 Z-type terms (several WHT launches) and X/Y/Z families large enough to take the
    basis-change route, mixed with stray terms that stay on the direct route."""
    rng = np.random.default_rng(n)
    psi = random_state(n, 3 * n)
    sv = Statevector(psi)
    from qiskit_b200.pauli import PauliList
    z = rng.integers(2, size=(700, n)).astype(bool)
    op = SparsePauliOp(PauliList(z, np.zeros_like(z)), rng.standard_normal(700))
    want = ref.expectation_value_sparse(psi, op.paulis.z, op.paulis.x, op.coeffs)
    assert abs(sv.expectation_value(op) - want) < ATOL
    labels = []
    for basis in "XYZ":
        for _ in range(40):
            supp = rng.integers(2, size=n).astype(bool)
            labels.append("".join(basis if b else "I" for b in supp))
    for _ in range(30):      # mixed families: X on even qubits, Z on odd
        supp = rng.integers(2, size=n).astype(bool)
        labels.append("".join(("X" if q % 2 == 0 else "Z") if supp[q] else "I" for q in range(n)))
    labels += ["".join(rng.choice(list("IXYZ"), size=n)) for _ in range(20)]
    op = SparsePauliOp(labels, rng.standard_normal(len(labels)) + 1j * rng.standard_normal(len(labels)))
    want = ref.expectation_value_sparse(psi, op.paulis.z, op.paulis.x, op.coeffs)
    before = sv.counters()["read_passes"]
    assert abs(sv.expectation_value(op) - want) < ATOL
    assert sv.counters()["read_passes"] - before < len(labels) // 2     # batched, not per term
    np.testing.assert_array_equal(sv.data, psi)                          # the state is untouched


def test_expval_families_continue_and_restart():
    """Three rotated families on 20 qubits: X strings (2 passes from the state), the same with Y on
    qubit 0 (1 pass continuing the rotated copy in place, so the first family must leave amplitudes
    behind), all-Y strings (restart from the state, so the second family may leave |a|^2 only)."""
    n = 20
    rng = np.random.default_rng(20)
    psi = random_state(n, 21)
    sv = Statevector(psi)
    labels = []
    for first, other in (("X", "X"), ("Y", "X"), ("Y", "Y")):
        for _ in range(14):
            supp = rng.integers(2, size=n).astype(bool)
            supp[0] = True
            labels.append("".join((first if q == 0 else other) if supp[q] else "I" for q in reversed(range(n))))
    labels = list(dict.fromkeys(labels))
    op = SparsePauliOp(labels, rng.standard_normal(len(labels)))
    want = ref.expectation_value_sparse(psi, op.paulis.z, op.paulis.x, op.coeffs)
    before = sv.counters()
    assert abs(sv.expectation_value(op) - want) < ATOL
    after = sv.counters()
    assert after["state_passes"] - before["state_passes"] == 5     # 2 + 1 + 2 rotation passes
    assert after["read_passes"] - before["read_passes"] == 3
    np.testing.assert_array_equal(sv.data, psi)
    assert abs(sv.expectation_value(op) - want) < ATOL              # and again, from the reused scratch copy


def test_expval_generic_operator_and_inner():
    psi = random_state(4, 5)
    m = random_unitary(2, np.random.default_rng(1))
    got = Statevector(psi).expectation_value(m, [1, 3])
    assert abs(got - ref.expectation_value_operator(psi, m, [1, 3])) < ATOL
    phi = random_state(4, 6)
    assert abs(Statevector(psi).inner(Statevector(phi)) - np.vdot(psi, phi)) < ATOL
    assert Statevector(psi).is_valid() and not Statevector(2 * psi).is_valid()


def test_expval_overflow_and_range():
    sv = Statevector.from_int(0, 4)
    with pytest.raises(QiskitError):
        sv.expectation_value(Pauli("ZZZ"))


# ------------------------------------------------- A6-A7: probabilities -----

def test_probabilities_golden():
    named = {"+0": np.array([1, 0, 1, 0]) / np.sqrt(2)}
    v = np.zeros(8, complex); v[0] = v[7] = 1 / np.sqrt(2); named["ghz3"] = v
    v = np.zeros(8, complex); v[1] = v[2] = v[4] = 1 / np.sqrt(3); named["w3"] = v
    for kat in KATS["probabilities"]:
        sv = Statevector(named[kat["label"]])
        for qargs, target in kat["cases"]:
            np.testing.assert_allclose(sv.probabilities(qargs), target, atol=1e-12)


def test_probabilities_all_qargs_4q():
    """test_statevector.py:1379-1405's qargs enumeration, against the oracle."""
    psi = random_state(4, 12)
    sv = Statevector(psi)
    for r in range(1, 5):
        for qargs in itertools.permutations(range(4), r):
            np.testing.assert_allclose(sv.probabilities(list(qargs)), ref.probabilities(psi, list(qargs)), atol=ATOL)
    np.testing.assert_allclose(sv.probabilities(), ref.probabilities(psi), atol=ATOL)
    np.testing.assert_allclose(sv.probabilities(decimals=3), ref.probabilities(psi, decimals=3), atol=1e-12)


@pytest.mark.parametrize("n,qargs", [(14, [13, 0, 5]), (14, list(range(13, -1, -1))), (16, list(range(2, 15))),
                                     (18, [17]), (15, [3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14])])
def test_probabilities_marginals_larger(n, qargs):
    psi = random_state(n, n)
    np.testing.assert_allclose(Statevector(psi).probabilities(qargs), ref.probabilities(psi, qargs), atol=ATOL)


def test_probabilities_clip():
    """statevector.py:668 clips to [0, 1] (test_statevector.py:1584-1604)."""
    sv = Statevector(np.array([1.2, 0.0]))
    assert sv.probabilities().tolist() == [1.0, 0.0]


# --------------------------------------------------------- A8: sampling -----

@pytest.mark.parametrize("n", [1, 3, 10, 11, 16, 20, 21])
def test_sample_indices_bit_exact_vs_tree_oracle(n):
    psi = random_state(n, 40 + n)
    u = np.random.default_rng(123).random(100_000)
    got = Statevector(psi).sample_indices(len(u), uniforms=u).astype(np.int64)
    want = ref.tree_choice_indices(ref.tree_probabilities(psi), u)
    assert np.array_equal(got, want)
    numpy_idx = ref.choice_indices(ref.probabilities(psi), u)
    diff = got - numpy_idx
    assert np.max(np.abs(diff)) <= 1 and np.mean(diff != 0) < 1e-3


def test_sample_edge_uniforms_and_sparse_states():
    psi = np.zeros(1 << 12, complex); psi[5] = 0.6; psi[4000] = 0.8j
    u = np.array([0.0, 0.36 - 1e-12, 0.36, 0.36 + 1e-12, 0.999999999999, np.nextafter(1.0, 0)])
    got = Statevector(psi).sample_indices(len(u), uniforms=u).astype(np.int64)
    want = ref.tree_choice_indices(ref.tree_probabilities(psi), u)
    assert np.array_equal(got, want) and set(got.tolist()) <= {5, 4000}
    assert Statevector(psi).sample_indices(0, uniforms=np.zeros(0)).size == 0
    # the two level-0 search paths (stored segment ends for >= 4096 shots, whole-block rescan below)
    # are the same arithmetic: identical indices for identical uniforms, also on a sparse state
    for vec in (psi, random_state(15, 9)):
        sv = Statevector(vec)
        uu = np.random.default_rng(5).random(6000)
        whole = sv.sample_indices(len(uu), uniforms=uu)
        parts = np.concatenate([sv.sample_indices(3000, uniforms=uu[:3000]), sv.sample_indices(3000, uniforms=uu[3000:])])
        assert np.array_equal(whole, parts)


@pytest.mark.parametrize("n,qargs", [(6, [0, 3]), (12, [11, 2, 7]), (14, list(range(1, 14))), (13, [4])])
def test_sample_marginal_qargs(n, qargs):
    psi = random_state(n, 77)
    sv = Statevector(psi)
    u = np.random.default_rng(5).random(20_000)
    got = sv.sample_indices(len(u), qargs, uniforms=u).astype(np.int64)
    # the tree runs over the engine's own marginal probabilities
    want = ref.tree_choice_indices(sv.probabilities(qargs), u)
    numpy_idx = ref.choice_indices(ref.probabilities(psi, qargs), u)
    assert np.max(np.abs(got - numpy_idx)) <= 1
    assert np.mean(got != want) < 1e-3 and np.max(np.abs(got - want)) <= 1


def test_sample_memory_strings_counts_and_seed():
    """quantum_state.py:236-294; determinism with the same seed (test_statevector_sampler.py:127-136)."""
    qc = QuantumCircuit(3); qc.h(0); qc.cx(0, 1); qc.x(2)
    sv = Statevector(qc)
    sv.seed(100)
    mem = sv.sample_memory(50)
    assert set(mem.tolist()) <= {"100", "111"} and mem.dtype.kind == "U"
    sv.seed(100)
    assert np.array_equal(sv.sample_memory(50), mem)
    sv.seed(100)
    oracle_mem = ref.sample_memory(sv.data, 50, None, np.random.default_rng(100))
    assert np.array_equal(mem, oracle_mem)
    counts = sv.sample_counts(2000)
    assert set(counts) == {"100", "111"} and sum(counts.values()) == 2000
    assert sv.sample_memory(5, [2]).tolist() == ["1"] * 5
    with pytest.raises(ValueError, match="do not sum to 1"):
        Statevector(np.array([1.0, 1.0])).sample_memory(3)


def test_measure_and_reset():
    """statevector.py:675-719, quantum_state.py:296-331 against the oracle with the same RNG."""
    psi = random_state(5, 9)
    for qargs in ([1], [0, 3], [4, 2, 0]):
        sv = Statevector(psi); sv.seed(7)
        outcome, post = sv.measure(qargs)
        o_outcome, o_post = ref.measure(psi, qargs, np.random.default_rng(7))
        assert outcome == o_outcome and np.max(np.abs(post.data - o_post)) < ATOL
        sv = Statevector(psi); sv.seed(11)
        got = sv.reset(qargs).data
        assert np.max(np.abs(got - ref.reset(psi, qargs, np.random.default_rng(11)))) < ATOL
        np.testing.assert_array_equal(sv.data, psi)
    assert Statevector(psi).reset().data[0] == 1
    qc = QuantumCircuit(2); qc.h(0); qc.cx(0, 1); qc.reset(0); qc.h(1)
    sv = Statevector.from_int(0, 4); sv.seed(3)
    out = sv.evolve(qc).data
    h = ogm.h()
    mid = ref.apply_gate_list(ref.basis_state(2), [(h, [0]), (ogm.cx(), [0, 1])])
    want = ref.evolve_operator(ref.reset(mid, [0], np.random.default_rng(3)), h, [1])
    assert np.max(np.abs(out - want)) < ATOL


def test_initialize_instruction():
    qc = QuantumCircuit(3)
    qc.h(0)
    qc.initialize([0.6, 0.8j], [2])
    sv = Statevector.from_int(0, 8); sv.seed(1)
    got = sv.evolve(qc).data
    want = np.kron(np.array([0.6, 0.8j]), np.kron([1, 0], np.array([1, 1]) / np.sqrt(2)))
    assert np.max(np.abs(got - want)) < ATOL
    qc2 = QuantumCircuit(2); qc2.initialize("1+")
    assert np.max(np.abs(Statevector(qc2).data - np.kron([0, 1], np.array([1, 1]) / np.sqrt(2)))) < ATOL


def test_from_label_and_from_int():
    for label in ["0", "1", "+", "-", "r", "l", "01+-", "rl10"]:
        want = np.array([1.0 + 0j])
        one = {"0": [1, 0], "1": [0, 1], "+": [1, 1], "-": [1, -1], "r": [1, 1j], "l": [1, -1j]}
        for ch in label:
            v = np.array(one[ch], dtype=complex)
            want = np.kron(want, v / np.linalg.norm(v))
        assert np.max(np.abs(Statevector.from_label(label).data - want)) < ATOL, label
    assert Statevector.from_int(5, 8).data[5] == 1


# ------------------------------------------------------ A11: estimator ------

def _ra_circuit(reps):
    return lib.real_amplitudes_circuit(2, reps)


def test_estimator_reference_goldens():
    """test_statevector_estimator.py:66-123 through the public API."""
    psi1, psi2 = _ra_circuit(2), _ra_circuit(3)
    h1 = SparsePauliOp.from_list([("II", 1), ("IZ", 2), ("XI", 3)])
    h2 = SparsePauliOp.from_list([("IZ", 1)])
    h3 = SparsePauliOp.from_list([("ZI", 1), ("ZZ", 1)])
    t1, t2, t3 = [0, 1, 1, 2, 3, 5], [0, 1, 1, 2, 3, 5, 8, 13], [1, 2, 3, 4, 5, 6]
    est = StatevectorEstimator()
    res = est.run([(psi1, h1, [t1])]).result()
    np.testing.assert_allclose(res[0].data.evs, [1.5555572817900956], rtol=1e-12)
    assert res[0].data.evs.shape == (1,) and np.all(res[0].data.stds == 0)
    assert res.metadata == {"version": 2} and res[0].metadata["target_precision"] == 0.0
    np.testing.assert_allclose(est.run([(psi2, h1, t2)]).result()[0].data.evs, 2.97797666, rtol=1e-7)
    np.testing.assert_allclose(est.run([(psi1, [h2, h3], t1)]).result()[0].data.evs,
                               [-0.551653, 0.07535239], rtol=1e-6)
    r4 = est.run([(psi1, [h1, h3], [t1, t3]), (psi2, h2, t2)]).result()
    np.testing.assert_allclose(r4[0].data.evs, [1.55555728, -1.08766318], rtol=1e-7)
    np.testing.assert_allclose(r4[1].data.evs, [0.17849238], rtol=1e-7)
    hmol = SparsePauliOp.from_list([tuple(t) for t in KATS["estimator"][6]["obs"]])
    bound = psi1.assign_parameters(t1)
    np.testing.assert_allclose(est.run([(bound, hmol)]).result()[0].data.evs, -1.284366511861733, rtol=1e-12)


def test_estimator_endianness_noise_and_shapes():
    qc = QuantumCircuit(2); qc.x(0)
    est = StatevectorEstimator()
    evs = est.run([(qc, ["IZ", "ZI"])]).result()[0].data.evs
    np.testing.assert_allclose(evs, [-1.0, 1.0], atol=1e-12)
    k = KATS["estimator_noise"]
    psi1 = _ra_circuit(2)
    h1 = SparsePauliOp.from_list([("II", 1), ("IZ", 2), ("XI", 3)])
    noisy = StatevectorEstimator(default_precision=k["precision"], seed=k["seed"])
    got = noisy.run([(psi1, h1, [0, 1, 1, 2, 3, 5])]).result()[0].data.evs
    assert abs(float(got) - k["expected"]) < 1e-12
    r = est.run([(psi1, [["ZZ"], ["XX"], ["YY"]], np.zeros((2, 6)))]).result()[0]
    assert r.data.evs.shape == (3, 2)
    with pytest.raises(ValueError):
        est.run([(psi1, "ZZZ", [0] * 6)])


def test_estimator_matches_oracle_loop_random():
    n = 5
    qc = lib.efficient_su2_circuit(n, 2)
    vals = np.random.default_rng(3).uniform(-np.pi, np.pi, (3, 2 * n * 3))
    obs = lib.random_pauli_observable(n, 12, seed=4, coeff_seed=5)
    evs = StatevectorEstimator().run([(qc, obs, vals)]).result()[0].data.evs
    for i in range(3):
        gates = oracle_matrices(lib.efficient_su2_named(n, 2, vals[i]))
        psi = ref.apply_gate_list(ref.basis_state(n), gates)
        want = ref.expectation_value_sparse(psi, obs.paulis.z, obs.paulis.x, obs.coeffs)
        assert abs(evs[i] - want.real) < ATOL


def test_estimator_reset_trajectories_reproducible():
    """test_statevector_estimator.py:318-346: circuits with reset are re-simulated per index
    from one rng, reproducibly."""
    qc = QuantumCircuit(1); qc.h(0); qc.reset(0); qc.ry(Parameter("a"), 0)
    a = StatevectorEstimator(seed=12).run([(qc, ["Z", "X"], [[0.3]])]).result()[0].data.evs
    b = StatevectorEstimator(seed=12).run([(qc, ["Z", "X"], [[0.3]])]).result()[0].data.evs
    np.testing.assert_array_equal(a, b)
    np.testing.assert_allclose(a, [np.cos(0.3), np.sin(0.3)], atol=1e-12)


def test_estimator_initialize_on_entangled_subset_is_stochastic_and_seeded():
    """`initialize` on a subset resets those qubits first (statevector.py:989-1014): it draws from the
    state's rng like `reset`, so the Estimator must re-simulate every index from the pub's ONE seeded
    generator (statevector_estimator.py:151,160-171) -- the reset draw and the precision noise interleave."""
    s2 = np.sqrt(0.5)
    qc = QuantumCircuit(2); qc.h(0); qc.cx(0, 1); qc.initialize([s2, s2], [0]); qc.ry(Parameter("a"), 1)
    obs = ["ZI", "IX", "ZZ"]
    run = lambda seed: StatevectorEstimator(seed=seed, default_precision=0.1).run(  # noqa: E731
        [(qc, obs, [[0.4]])]).result()[0].data.evs
    a, b = run(21), run(21)
    np.testing.assert_array_equal(a, b)
    # the oracle's serial loop with the same generator: per index reset draw, then noise
    rng = np.random.default_rng(21)
    want = []
    hmat, cx = ogm.h(), ogm.cx()
    init = np.zeros((2, 2), complex); init[:, 0] = [s2, s2]
    for label in obs:
        psi = ref.evolve_operator(ref.evolve_operator(ref.basis_state(2), hmat, [0]), cx, [0, 1])
        psi = ref.reset(psi, [0], rng)
        psi = ref.evolve_operator(psi, init, [0])
        psi = ref.evolve_operator(psi, ogm.ry(0.4), [1])
        ev = np.real_if_close(ref.expectation_value_labels(psi, [(label, 1.0)]))
        want.append(rng.normal(ev, 0.1))
    np.testing.assert_allclose(a, want, atol=1e-10)


# -------------------------------------------------------- A12: sampler ------

def _counts_close(got, want, rtol=0.1):
    want = {int(k): v for k, v in want.items()}
    top = max(max(got), max(want))
    np.testing.assert_allclose([got.get(i, 0) for i in range(top + 1)],
                               [want.get(i, 0) for i in range(top + 1)], rtol=rtol)


def test_sampler_reference_counts_seed123():
    """test_statevector_sampler.py:37-79: rtol 0.1 in the reference; case :58 exactly."""
    sampler = StatevectorSampler(seed=123)
    bell = QuantumCircuit(2); bell.h(0); bell.cx(0, 1); bell.measure_all()
    res = sampler.run([bell], shots=10000).result()
    _counts_close(res[0].data.meas.get_int_counts(), {0: 5000, 3: 5000})
    assert res[0].metadata["shots"] == 10000 and res.metadata == {"version": 2}
    pqc = _ra_circuit(2); pqc.measure_all()
    cases = [([0] * 6, {0: 10000}), ([1] * 6, {0: 168, 1: 3389, 2: 470, 3: 5973}),
             ([0, 1, 1, 2, 3, 5], {0: 1339, 1: 3534, 2: 912, 3: 4215}),
             ([1, 2, 3, 4, 5, 6], {0: 634, 1: 291, 2: 6039, 3: 3036})]
    for theta, target in cases:
        got = sampler.run([(pqc, theta)], shots=10000).result()[0].data.meas.get_int_counts()
        _counts_close(got, target)
    got = sampler.run([(pqc, [1] * 6)], shots=10000).result()[0].data.meas.get_int_counts()
    assert got == {0: 168, 1: 3389, 2: 470, 3: 5973}
    pqc2 = _ra_circuit(3); pqc2.measure_all()
    got = sampler.run([(pqc2, list(range(8)))], shots=10000).result()[0].data.meas.get_int_counts()
    _counts_close(got, {0: 1898, 1: 6864, 2: 928, 3: 311})
    # batched parameter sets share the seed (statevector_sampler.py:188)
    r = sampler.run([(pqc, [[1] * 6, [1] * 6])], shots=500).result()[0].data.meas
    assert r.shape == (2,) and np.array_equal(r.array[0], r.array[1])


@pytest.mark.parametrize("kat", KATS["sampler_cregs"] + KATS["sampler_reverse"], ids=lambda k: k["src"])
def test_sampler_cregs_and_reverse_order(kat):
    cregs = [ClassicalRegister(size, name) for name, size in kat["cregs"]]
    qc = QuantumCircuit(kat["n"], *cregs)
    for name, params, qargs in kat["gates"]:
        getattr(qc, name)(*params, *qargs)
    for q, c in kat["measure"]:
        qc.measure(q, c)
    res = StatevectorSampler(seed=123).run([qc], shots=10000).result()[0]
    for name, want in kat["counts"].items():
        got = getattr(res.data, name).get_int_counts()
        assert set(got) == {int(k) for k in want}, (name, got)
        _counts_close(got, want)


def test_sampler_errors_and_no_cregs():
    qc = QuantumCircuit(2, 2); qc.measure(0, 0); qc.h(0); qc.measure(0, 1)
    with pytest.raises(QiskitError):
        StatevectorSampler().run([qc]).result()
    bare = QuantumCircuit(2); bare.h(0)
    with pytest.warns(UserWarning):
        res = StatevectorSampler().run([bare], shots=10).result()
    assert len(res[0].data) == 0


def test_edge_cases_small_and_empty():
    """Empty / minimal inputs behave like the reference: 0- and 1-qubit states, empty qargs, zero
    shots at the state level, empty operators, empty pub lists, invalid shots."""
    sv1 = Statevector.from_int(0, 2)
    np.testing.assert_array_equal(sv1.probabilities(), [1.0, 0.0])
    np.testing.assert_array_equal(sv1.probabilities([]), [1.0])          # everything marginalised
    assert sv1.sample_indices(0).shape == (0,) and sv1.sample_memory(0).shape == (0,)
    assert sv1.sample_counts(5) == {"0": 5}
    np.testing.assert_array_equal(Statevector(np.array([1.0 + 0j])).data, [1.0])   # zero qubits
    np.testing.assert_array_equal(sv1.apply_gate_list([]).data, [1.0, 0.0])
    empty = SparsePauliOp(PauliList(np.zeros((0, 1), bool), np.zeros((0, 1), bool)), np.zeros(0))
    assert sv1.expectation_value(empty) == 0
    assert sv1.expectation_value(Pauli("I")) == 1.0
    plus = Statevector.from_label("+")
    edge = np.array([0.0, 0.5, 1.0 - 2.0 ** -53])
    np.testing.assert_array_equal(plus.sample_indices(3, uniforms=edge),
                                  ref.tree_choice_indices(ref.tree_probabilities(plus.data), edge))
    with pytest.raises(ValueError, match="do not sum to 1"):              # Generator.choice's check
        Statevector(np.array([1.0, 1.0])).sample_indices(4)
    with pytest.raises(QiskitError):
        sv1.evolve(np.eye(2), [3])
    with pytest.raises(QiskitError):
        Statevector.from_int(0, 8).probabilities([1, 1])
    assert len(StatevectorEstimator().run([]).result()) == 0
    qc = QuantumCircuit(2)
    qc.h(0)
    qc.cx(0, 1)
    np.testing.assert_allclose(StatevectorEstimator().run([(qc, ["ZZ", "XX", "YY", "II"])]).result()[0].data.evs,
                               [1, 1, -1, 1], atol=ATOL)
    with pytest.raises(ValueError):
        StatevectorEstimator().run([(qc, "ZZZ")]).result()
    qm = QuantumCircuit(2)
    qm.h(0)
    qm.cx(0, 1)
    qm.measure_all()
    with pytest.raises(ValueError, match="positive"):                     # sampler_pub.py:152-153
        StatevectorSampler().run([qm], shots=0).result()
    with pytest.raises(TypeError):
        StatevectorSampler().run([qm], shots=1.5).result()
    assert sum(StatevectorSampler(seed=1).run([qm], shots=1).result()[0].data.meas.get_counts().values()) == 1


def test_kernel_matrix_with_bilinear_tables_opt_in():
    """FP_DIAGB (bilinear phase tables) is opt-in (measured slower, profiles/r2_bilinear_tables_ab.txt) and read
    from the environment once per process: the small-n kernel matrix (QFT / TFIM-like tables with register,
    thread and off-tile bits, all checked against the oracle) runs in a child process with QSV_BILINEAR=1."""
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, QSV_BILINEAR="1")
    res = subprocess.run([sys.executable, os.path.join(root, "scripts", "sanitize_matrix.py")], env=env,
                         capture_output=True, text=True, timeout=900, cwd=root)
    assert res.returncode == 0 and "sanitize matrix ok" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
