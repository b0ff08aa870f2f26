"""BackendEstimatorV2 / BackendSamplerV2 over the device BasicSimulator (SURVEY 8f N3 + N4):
the reference's own cases (test/python/primitives/test_backend_estimator_v2.py:105-160,
test_backend_sampler_v2.py) at its tolerances, plus agreement with the exact StatevectorEstimator
within the reported standard errors."""
import numpy as np
import pytest

from qiskit_b200.backend_primitives import _qwc_groups


def _zx(label):
    z = np.array([c in "ZY" for c in reversed(label)])
    x = np.array([c in "XY" for c in reversed(label)])
    return z, x


def test_qubit_wise_commuting_groups():
    labels = ["II", "IZ", "ZI", "ZZ", "XX", "XI", "YZ"]
    groups = _qwc_groups([_zx(s) for s in labels], True)
    members = [g[2] for g in groups]
    assert sorted(sum(members, [])) == list(range(len(labels)))
    for gz, gx, mem in groups:           # every member agrees with the group basis where it acts
        for t in mem:
            z, x = _zx(labels[t])
            act = z | x
            assert np.all(gz[act] == z[act]) and np.all(gx[act] == x[act])
    assert members[0] == [0, 1, 2, 3]    # first fit: the diagonal terms share one circuit
    assert members[1:] == [[4, 5], [6]]  # XI joins XX; YZ fits neither basis
    single = _qwc_groups([_zx(s) for s in labels], False)
    assert [g[2] for g in single] == [[i] for i in range(len(labels))]


@pytest.mark.gpu
@pytest.mark.parametrize("abelian", [True, False])
def test_estimator_reference_cases(abelian):
    from qiskit_b200 import BackendEstimatorV2, BasicSimulator, SparsePauliOp, StatevectorEstimator
    from qiskit_b200 import circuit_library as lib

    psi1, psi2 = lib.real_amplitudes_circuit(2, 2), lib.real_amplitudes_circuit(2, 3)
    h1 = SparsePauliOp.from_list([("II", 1), ("IZ", 2), ("XI", 3)])
    h2 = SparsePauliOp.from_list([("IZ", 1)])
    h3 = SparsePauliOp.from_list([("ZI", 1), ("ZZ", 1)])
    t1, t2, t3 = [0, 1, 1, 2, 3, 5], [0, 1, 1, 2, 3, 5, 8, 13], [1, 2, 3, 4, 5, 6]
    est = BackendEstimatorV2(backend=BasicSimulator(), options={
        "default_precision": 5e-3, "seed_simulator": 12, "abelian_grouping": abelian})
    rtol = 3e-1                                                       # the reference's own tolerance
    res = est.run([(psi1, h1, [t1])]).result()
    np.testing.assert_allclose(res[0].data.evs, [1.5555572817900956], rtol=rtol)
    assert res[0].metadata["shots"] == 40000 and res[0].metadata["target_precision"] == 5e-3
    np.testing.assert_allclose(est.run([(psi2, h1, t2)]).result()[0].data.evs, [2.97797666], rtol=rtol)
    np.testing.assert_allclose(est.run([(psi1, [h2, h3], t1)]).result()[0].data.evs,
                               [-0.551653, 0.07535239], rtol=rtol)
    pubs = [(psi1, [h1, h3], [t1, t3]), (psi2, h2, t2)]
    r4 = est.run(pubs).result()
    np.testing.assert_allclose(r4[0].data.evs, [1.55555728, -1.08766318], rtol=rtol)
    np.testing.assert_allclose(r4[1].data.evs, [0.17849238], rtol=rtol)
    # against the exact values, within the reported standard errors
    exact = StatevectorEstimator().run(pubs).result()
    for got, want in zip(r4, exact):
        assert got.data.evs.shape == want.data.evs.shape
        assert np.all(np.abs(got.data.evs - want.data.evs) <= 5 * got.data.stds + 1e-12)
        assert np.all(got.data.stds > 0) and np.all(got.data.stds < 0.05)
    # same seed -> same counts -> same values
    again = est.run(pubs).result()
    np.testing.assert_array_equal(again[0].data.evs, r4[0].data.evs)


@pytest.mark.gpu
def test_estimator_y_basis_identity_and_precision():
    from qiskit_b200 import (BackendEstimatorV2, BasicSimulator, QuantumCircuit, SparsePauliOp,
                             StatevectorEstimator)
    from helpers import random_unitary

    rng = np.random.default_rng(5)
    qc = QuantumCircuit(4)
    for _ in range(6):
        qs = [int(q) for q in rng.choice(4, size=2, replace=False)]
        qc.unitary(random_unitary(2, rng), qs)
    obs = SparsePauliOp.from_list([("IIII", 0.5), ("XYZI", 1.0), ("YYII", -0.7), ("ZIZX", 0.3), ("IYIY", 1.1)])
    est = BackendEstimatorV2(backend=BasicSimulator(), seed_simulator=3)
    got = est.run([(qc, obs)], precision=0.01).result()[0]
    want = StatevectorEstimator().run([(qc, obs)]).result()[0]
    assert got.metadata["shots"] == 10000
    assert abs(got.data.evs - want.data.evs) <= 5 * got.data.stds
    only_id = est.run([(qc, SparsePauliOp.from_list([("IIII", 2.0)]))]).result()[0]
    assert only_id.data.evs == 2.0 and only_id.data.stds == 0.0
    with pytest.raises(ValueError, match="precision"):
        est.run([(qc, obs)], precision=0.0)


@pytest.mark.gpu
def test_sampler_reference_cases():
    from qiskit_b200 import (BackendSamplerV2, BasicSimulator, ClassicalRegister, Parameter, QuantumCircuit,
                             StatevectorSampler)

    shots = 10000
    sampler = BackendSamplerV2(backend=BasicSimulator(), options={"default_shots": shots, "seed_simulator": 123})
    bell = QuantumCircuit(2)
    bell.h(0)
    bell.cx(0, 1)
    bell.measure_all()
    res = sampler.run([bell]).result()
    counts = res[0].data.meas.get_int_counts()
    assert res[0].data.meas.num_shots == shots and set(counts) == {0, 3}
    assert abs(counts[0] - shots / 2) < 5 * np.sqrt(shots) / 2          # test_sampler_run
    # parameterised pub with a (3,) shape, shots given per run
    theta = Parameter("t")
    pqc = QuantumCircuit(1)
    pqc.ry(theta, 0)
    pqc.measure_all()
    r = sampler.run([(pqc, [[0.0], [np.pi / 2], [np.pi]])], shots=2000).result()[0]
    assert r.data.meas.shape == (3,) and r.data.meas.num_shots == 2000
    assert r.data.meas.get_int_counts(0) == {0: 2000} and r.data.meas.get_int_counts(2) == {1: 2000}
    assert abs(r.data.meas.get_int_counts(1).get(1, 0) - 1000) < 150
    # several classical registers, one of them wider than a byte (test_circuit_with_multiple_cregs)
    qc = QuantumCircuit(10, ClassicalRegister(1, "a"), ClassicalRegister(9, "b"))
    for q in (0, 2, 9):
        qc.x(q)
    qc.measure(0, 0)
    for q in range(1, 10):
        qc.measure(q, q)
    d = sampler.run([qc], shots=16).result()[0].data
    assert d.a.get_int_counts() == {1: 16}
    assert d.b.get_int_counts() == {(1 << 1) | (1 << 8): 16}
    assert d.b.array.shape == (16, 2)
    # same distribution as the exact sampler
    ref = StatevectorSampler(seed=1).run([bell], shots=shots).result()[0].data.meas.get_int_counts()
    assert set(ref) == set(counts)
    with pytest.raises(ValueError):
        sampler.run([bell], shots=0)
