"""CPU-only tests: host logic, the C ABI surface, and product-vs-oracle gate tables."""
import ctypes
import os
import re

import numpy as np
import pytest

from oracle import gate_matrices as ogm
from oracle import statevector_ref as ref
from qiskit_b200 import _lib
from qiskit_b200.circuit import ParameterVector, QuantumCircuit, ClassicalRegister, Parameter
from qiskit_b200.gates import GATES, gate_matrix
from qiskit_b200.pauli import Pauli, SparsePauliOp, pauli_masks
from qiskit_b200 import primitives as prim
from qiskit_b200 import circuit_library as lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_abi_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "qsv.h")).read()
    declared = set(re.findall(r"\b(qsv_[a-z0-9_]+)\s*\(", header))
    declared -= {"qsv_state", "qsv_gate", "qsv_counters"}
    assert declared == set(_lib.PROTOTYPES), declared ^ set(_lib.PROTOTYPES)
    so = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(so, name), name
    assert _lib.load().qsv_version() >= 100


def test_struct_layout_matches_header():
    assert ctypes.sizeof(_lib.QsvGate) == 4 + 4 + 4 * 12 + 8 + 32
    assert ctypes.sizeof(_lib.QsvCounters) == 96   # 12 uint64 fields of qsv_counters


def test_no_device_fails_loudly():
    if _lib.device_count() > 0:
        pytest.skip("GPU present")
    from qiskit_b200 import Statevector
    with pytest.raises(RuntimeError, match="no CUDA device"):
        Statevector.from_int(0, 4)


@pytest.mark.parametrize("name", sorted(GATES))
def test_gate_table_matches_oracle_table(name):
    gd = GATES[name]
    if name not in ogm.BY_NAME:
        pytest.skip("not in oracle table")
    params = list(np.random.default_rng(len(name)).uniform(-3, 3, gd.num_params))
    m = gate_matrix(name, params)
    np.testing.assert_allclose(m, ogm.matrix(name, params), atol=1e-15)
    assert m.shape == (1 << gd.num_qubits,) * 2
    np.testing.assert_allclose(m @ m.conj().T, np.eye(m.shape[0]), atol=1e-14)
    if gd.diagonal:
        assert np.count_nonzero(m - np.diag(np.diagonal(m))) == 0


def test_pauli_labels_and_masks():
    p = Pauli("-iXYZ")
    assert p.phase == 1 and p.to_label() == "-iXYZ"
    assert p.x.tolist() == [False, True, True] and p.z.tolist() == [True, True, False]
    z, x, ph = ref.pauli_from_label("-iXYZ")
    assert (z == p.z).all() and (x == p.x).all() and ph == p.phase
    zm, xm, ny = pauli_masks(p.z, p.x)
    assert int(zm[0]) == 0b011 and int(xm[0]) == 0b110 and int(ny[0]) == 1
    zm, xm, _ = pauli_masks(p.z, p.x, [4, 0, 2])
    assert int(zm[0]) == (1 << 4) | 1 and int(xm[0]) == 1 | (1 << 2)
    np.testing.assert_allclose(Pauli("Y").to_matrix(), [[0, -1j], [1j, 0]])
    with pytest.raises(Exception):
        Pauli("XQ")


def test_sparse_pauli_op_folds_phase_into_coeffs():
    op = SparsePauliOp(["-XX", "iZI"], [2.0, 3.0])
    np.testing.assert_allclose(op.coeffs, [-2.0, 3.0j])
    dense = -2 * Pauli("XX").to_matrix() + 3j * Pauli("ZI").to_matrix()
    np.testing.assert_allclose(op.to_matrix(), dense)
    op2 = SparsePauliOp.from_sparse_list([("XZ", [0, 3], 1.5)], 4)
    assert op2.to_list()[0][0] == "ZIIX"


def test_circuit_parameters_sorted_and_bound():
    th = ParameterVector("t", 12)
    a = Parameter("a")
    qc = QuantumCircuit(2)
    qc.ry(th[10], 0); qc.ry(th[2], 1); qc.rz(a, 0); qc.cx(0, 1)
    assert [p.name for p in qc.parameters] == ["a", "t[2]", "t[10]"]
    bound = qc.assign_parameters([0.1, 0.2, 0.3])
    assert [inst.operation.params for inst in bound.data[:3]] == [[0.3], [0.2], [0.1]]
    assert qc.data[0].operation.params[0] is th[10]          # original untouched
    with pytest.raises(ValueError):
        qc.assign_parameters([0.1])


def test_final_measurement_mapping_and_preprocess():
    a, b, c = ClassicalRegister(1, "a"), ClassicalRegister(2, "b"), ClassicalRegister(3, "c")
    qc = QuantumCircuit(3, a, b, c)
    qc.h(range(3))
    qc.measure([0, 1, 2, 2], [0, 2, 4, 5])
    stripped, qargs, info = prim._preprocess_circuit(qc)
    assert qargs == [0, 1, 2] and len(stripped) == 3
    got = {i.creg_name: i.qreg_indices for i in info}
    assert got == {"a": [0], "b": [3, 1], "c": [3, 2, 2]}
    qc2 = QuantumCircuit(2, 2)
    qc2.measure(0, 0); qc2.h(0); qc2.measure(1, 1)
    from qiskit_b200 import QiskitError
    with pytest.raises(QiskitError, match="mid-circuit"):
        prim._preprocess_circuit(qc2)


@pytest.mark.parametrize("num_bits,indices,m", [(3, [2, 1, 0], 3), (5, [3, 1, 3, 0, 0], 3),
                                                 (9, [0, 1, 2, 3, 4, 5, 6, 7, 8], 9), (1, [4], 4),
                                                 (17, list(range(17)), 17)])
def test_pack_samples_matches_reference_packing(num_bits, indices, m):
    """The C packer (host-only entry of the ABI) against the oracle's restatement of
    _samples_to_packed_array."""
    rng = np.random.default_rng(num_bits)
    idx = rng.integers(0, 1 << m, size=257, dtype=np.uint64)
    want = ref.sampler_pack_indices(idx.astype(np.int64), m, num_bits, indices)
    got = prim.pack_samples(idx, m, num_bits, indices)
    assert got.dtype == np.uint8 and np.array_equal(got, want)


def test_pack_samples_empty():
    assert prim.pack_samples(np.zeros(0, dtype=np.uint64), 2, 2, [0, 1]).shape == (0, 1)


def test_estimator_pub_coercion_shapes_and_errors():
    qc = lib.real_amplitudes_circuit(2, 2)
    pub = prim.EstimatorPub.coerce((qc, "ZZ", [0, 1, 1, 2, 3, 5]), 0.0)
    assert pub.shape == ()
    pub = prim.EstimatorPub.coerce((qc, ["ZZ", "XX"], [[0, 1, 1, 2, 3, 5]] * 2), 0.0)
    assert pub.shape == (2,)
    pub = prim.EstimatorPub.coerce((qc, [["ZZ"], ["XX"], ["YY"]], [[0, 1, 1, 2, 3, 5]] * 2), 0.0)
    assert pub.shape == (3, 2)
    with pytest.raises(ValueError):
        prim.EstimatorPub.coerce((qc, "ZZZ", [0] * 6), 0.0)          # observable width
    with pytest.raises(ValueError):
        prim.EstimatorPub.coerce((qc, "ZZ", [0] * 5), 0.0)           # parameter count
    with pytest.raises(ValueError):
        prim.EstimatorPub.coerce((qc, "ZZ", [0] * 6), -1.0)          # negative precision
    with pytest.raises(ValueError):
        prim.EstimatorPub.coerce((qc, ["ZZ"] * 3, [[0] * 6] * 2), 0.0)   # not broadcastable
    one = QuantumCircuit(1); one.ry(Parameter("x"), 0)
    assert prim.EstimatorPub.coerce((one, "Z", [0.1, 0.2, 0.3]), 0.0).shape == (3,)
    assert prim.SamplerPub.coerce((one, [[0.1], [0.2]]), 10).shape == (2,)
    with pytest.raises(ValueError):
        prim.SamplerPub.coerce((one, [0.1], -5), None)


def test_bitarray_counts():
    arr = np.array([[0b00000001, 0b10000000], [0, 3], [0, 3]], dtype=np.uint8)
    ba = prim.BitArray(arr, 9)
    assert ba.num_shots == 3 and ba.shape == ()
    assert ba.get_int_counts() == {3: 2, 384: 1}
    assert ba.get_counts() == {"000000011": 2, "110000000": 1}


def test_library_generators_structure():
    assert len(lib.quantum_volume(20, 20, 0)) == 200
    named = lib.efficient_su2_named(28, 4, np.zeros(280))
    assert len(named) == 388 and sum(1 for g in named if g[0] == "cx") == 108
    assert len(lib.efficient_su2_circuit(28, 4).parameters) == 280
    q = lib.qft_named(30, input_layer_seed=None)
    assert sum(1 for g in q if g[0] == "h") == 30 and sum(1 for g in q if g[0] == "cp") == 435
    assert sum(1 for g in q if g[0] == "swap") == 15
    t = lib.tfim_named(36)
    assert sum(1 for g in t if g[0] == "rzz") == 350 and sum(1 for g in t if g[0] == "rx") == 360
    assert len(lib.heisenberg_observable(28)) == 1000
    assert len(lib.tfim_observable(36)) == 71
    for u, _ in lib.quantum_volume(4, 2, 1):
        np.testing.assert_allclose(u @ u.conj().T, np.eye(4), atol=1e-13)


def test_efficient_su2_circuit_equals_named_list():
    """The parameterised circuit, bound with a bare array, unrolls to the named gate list."""
    theta = np.random.default_rng(0).uniform(-np.pi, np.pi, 2 * 3 * 3)
    qc = lib.efficient_su2_circuit(3, 2).assign_parameters(theta)
    named = lib.efficient_su2_named(3, 2, theta)
    assert [(i.name, i.operation.params, list(i.qubits)) for i in qc.data] == \
           [(n, [float(p) for p in ps], q) for n, ps, q in named]


def test_lowering_walks_definitions_and_global_phase():
    """lowering.walk against the oracle on a nested definition (no device needed: a fake
    state records Reset calls)."""
    from qiskit_b200.lowering import GateStream, walk

    inner = QuantumCircuit(2, name="inner", global_phase=0.3)
    inner.h(0); inner.cx(0, 1)
    outer = QuantumCircuit(3, global_phase=0.2)
    outer.append(inner.to_instruction(), [2, 0])
    outer.barrier()
    outer.rz(0.7, 1)
    st = GateStream()
    walk(None, outer.to_instruction(), None, st)
    assert st.kinds == [2, 2, 0, 0, 0] and st.qargs[2:] == [(2,), (2, 0), (1,)]
    gates = []
    for kind, qa, d, p in zip(st.kinds, st.qargs, st.data, st.params):
        gates.append(("phase", p) if kind == 2 else (d, list(qa)))
    psi = ref.apply_gate_list(ref.basis_state(3), gates)
    want = np.exp(0.5j) * ref.apply_gate_list(
        ref.basis_state(3), [(ogm.h(), [2]), (ogm.cx(), [2, 0]), (ogm.rz(0.7), [1])])
    np.testing.assert_allclose(psi, want, atol=1e-15)
    bad = QuantumCircuit(1, 1); bad.measure(0, 0)
    from qiskit_b200 import QiskitError
    with pytest.raises(QiskitError, match="classical bits"):
        walk(None, bad.to_instruction(), None, GateStream())
