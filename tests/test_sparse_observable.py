"""SparseObservable inputs (qiskit_b200/sparse_observable.py): alphabet, constructors, `as_paulis`
against the reference's documented examples (crates/quantum_info/src/sparse_observable/mod.rs:3135-3150)
and against dense matrices, and the Estimator's observable coercion (observables_array.py:247-303)."""
import numpy as np
import pytest

from qiskit_b200 import QiskitError, SparseObservable, SparsePauliOp
from qiskit_b200.primitives import _coerce_observable, _observables_array


def test_reference_docstring_examples():
    # mod.rs:3139-3143: "+" -> 0.5 () + 0.5 (X_0)
    obs = SparseObservable.from_label("+").as_paulis()
    assert sorted(obs.to_sparse_list()) == [("", [], 0.5 + 0j), ("X", [0], 0.5 + 0j)]
    # mod.rs:3147-3149: 0.5 (|+><+| + |-><-|) == 0.5 (|r><r| + |l><l|) == I / 2
    left = SparseObservable.from_list([("+", 0.5), ("-", 0.5)]).as_paulis().simplify()
    right = SparseObservable.from_list([("r", 0.5), ("l", 0.5)]).as_paulis().simplify()
    assert left.to_sparse_list() == right.to_sparse_list() == [("", [], 0.5 + 0j)]


def test_label_order_and_sparse_list_round_trip():
    obs = SparseObservable.from_list([("IZ0+", 2.0), ("rIIX", -1.5j)])
    assert obs.num_qubits == 4 and len(obs) == 2
    assert obs.to_sparse_list() == [("+0Z", [0, 1, 2], 2.0 + 0j), ("Xr", [0, 3], -1.5j)]
    again = SparseObservable.from_sparse_list(obs.to_sparse_list(), 4)
    assert again.to_sparse_list() == obs.to_sparse_list()
    np.testing.assert_allclose(again.to_matrix(), obs.to_matrix())
    with pytest.raises(QiskitError):
        SparseObservable.from_sparse_list([("XX", [1, 1], 1.0)], 3)
    with pytest.raises(QiskitError):
        SparseObservable.from_list([("XQ", 1.0)])
    with pytest.raises(QiskitError):
        SparseObservable.from_sparse_list([("X", [5], 1.0)], 3)


def test_as_paulis_equals_dense_matrix_random():
    rng = np.random.default_rng(0)
    letters = "IXYZ+-rl01"
    n = 4
    terms = [("".join(rng.choice(list(letters), n)), complex(rng.normal(), rng.normal())) for _ in range(7)]
    obs = SparseObservable.from_list(terms)
    dense = obs.to_matrix()
    np.testing.assert_allclose(obs.as_paulis().to_matrix(), dense, atol=1e-14)
    spo = obs.to_sparse_pauli_op()
    assert isinstance(spo, SparsePauliOp)
    np.testing.assert_allclose(spo.to_matrix(), dense, atol=1e-14)
    labels = [lab for lab, _ in spo.to_list()]
    assert len(labels) == len(set(labels))                # equal strings merged
    # an observable that cancels to zero still coerces (to 0 * I)
    zero = SparseObservable.from_list([("0", 1.0), ("1", 1.0), ("I", -1.0)]).to_sparse_pauli_op()
    np.testing.assert_allclose(zero.to_matrix(), np.zeros((2, 2)), atol=1e-15)


def test_from_sparse_pauli_op_round_trip():
    op = SparsePauliOp.from_list([("XYZI", 1.0), ("IIZZ", -0.5), ("YYII", 0.25j)])
    obs = SparseObservable.from_sparse_pauli_op(op)
    np.testing.assert_allclose(obs.to_matrix(), op.to_matrix(), atol=1e-15)
    np.testing.assert_allclose(obs.to_sparse_pauli_op().to_matrix(), op.to_matrix(), atol=1e-15)


def test_estimator_coercion_accepts_projectors_and_duck_types():
    want = SparseObservable.from_list([("0+", 1.0)]).to_matrix()
    np.testing.assert_allclose(_coerce_observable("0+", 2).to_matrix(), want, atol=1e-15)
    d = _coerce_observable({"r1": 0.5, "ZZ": 1.0}, 2)
    np.testing.assert_allclose(d.to_matrix(), SparseObservable.from_list([("r1", 0.5), ("ZZ", 1.0)]).to_matrix(),
                               atol=1e-15)
    # "-Z" on ONE qubit is the prefixed Pauli label -Z; on TWO qubits it is |-><-| (x) Z (from_label semantics)
    np.testing.assert_allclose(_coerce_observable("-Z", 1).to_matrix(), -np.diag([1.0, -1.0]), atol=1e-15)
    np.testing.assert_allclose(_coerce_observable("-Z", 2).to_matrix(),
                               SparseObservable.from_label("-Z").to_matrix(), atol=1e-15)

    class Foreign:                                         # shaped like Qiskit's SparseObservable
        num_qubits = 3

        def to_sparse_list(self):
            return [("Z1", [0, 2], 2.0), ("", [], 0.5)]

    got = _coerce_observable(Foreign(), 3)
    np.testing.assert_allclose(got.to_matrix(),
                               SparseObservable.from_sparse_list(Foreign().to_sparse_list(), 3).to_matrix(), atol=1e-15)
    arr = _observables_array([Foreign(), "ZZZ", SparseObservable.from_label("01+")], 3)
    assert arr.shape == (3,) and all(isinstance(o, SparsePauliOp) for o in arr)
    with pytest.raises(ValueError):
        _coerce_observable(SparseObservable.from_label("0"), 2)
