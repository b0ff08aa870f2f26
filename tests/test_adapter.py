"""the adapter classes derive from the reference's classes and reuse its pub
containers.  Runs twice: against stand-ins with the reference's shapes (tests/fake_qiskit.py: Qubit
OBJECTS in CircuitInstruction.qubits, definition.qubits, sorted ParameterVector binding, BindingsArray,
ObservablesArray of dicts, PrimitiveJob on a worker thread) -- always -- and against the real Qiskit when
it is importable (skipped in this image: no Rust toolchain for qiskit._accelerate)."""
import numpy as np
import pytest

import fake_qiskit as fq
from oracle import gate_matrices as ogm
from oracle import statevector_ref as ref
from qiskit_b200 import adapter
from qiskit_b200.circuit import ParameterVector


def test_adapter_classes_derive_from_the_reference_classes():
    ad = adapter.make_adapter(fq.namespace())
    assert issubclass(ad.DeviceStatevector, fq.Statevector)
    assert issubclass(ad.DeviceEstimator, fq.BaseEstimatorV2)
    assert issubclass(ad.DeviceSampler, fq.BaseSamplerV2)
    # the static hooks the reference dispatches through (statevector.py:925, 965) are overridden
    assert ad.DeviceStatevector._evolve_operator is not fq.Statevector.__dict__.get("_evolve_operator")
    assert isinstance(ad.DeviceStatevector.__dict__["_evolve_instruction"], staticmethod)
    assert isinstance(ad.DeviceStatevector.__dict__["_evolve_operator"], staticmethod)
    with pytest.raises(ImportError):
        adapter.qiskit_namespace()            # this image has no qiskit; on a box that has it, see below


def _ansatz(n=3):
    """parameters inserted out of order: binding must follow the sorted order (t[2] < t[10])"""
    th = ParameterVector("t", 12)
    qc = fq.QuantumCircuit(n, metadata={"tag": "ansatz"})
    qc.ry(th[10], 0); qc.ry(th[2], 1); qc.h(2); qc.cx(0, 1); qc.rz(th[0], 2); qc.cx(1, 2)
    sub = fq.QuantumCircuit(2, global_phase=0.3)                 # nested definition with a global phase
    sub.h(0); sub.cp(0.7, 0, 1)
    qc.append(sub.to_instruction(), [2, 0])
    return qc


def _oracle_state(values):
    t0, t2, t10 = values                                        # sorted: t[0], t[2], t[10]
    gates = [(ogm.ry(t10), [0]), (ogm.ry(t2), [1]), (ogm.h(), [2]), (ogm.cx(), [0, 1]), (ogm.rz(t0), [2]),
             (ogm.cx(), [1, 2]), (ogm.h(), [2]), (ogm.cp(0.7), [2, 0])]
    return ref.apply_gate_list(ref.basis_state(3), gates) * np.exp(0.3j)


@pytest.mark.gpu
def test_device_statevector_on_reference_shaped_circuits():
    ad = adapter.make_adapter(fq.namespace())
    vals = [0.4, -1.1, 2.2]
    qc = _ansatz().assign_parameters(vals)
    sv = ad.DeviceStatevector(qc)
    want = _oracle_state(vals)
    assert isinstance(sv, fq.Statevector) and sv.num_qubits == 3 and sv.dims() == (2, 2, 2)
    np.testing.assert_allclose(sv.data, want, atol=1e-10)
    # evolve never mutates; operator path goes through the static _evolve_operator
    psi0 = ad.DeviceStatevector.from_int(0, 8)
    out = psi0.evolve(qc)
    np.testing.assert_allclose(out.data, want, atol=1e-10)
    np.testing.assert_allclose(psi0.data, ref.basis_state(3), atol=0)
    out2 = psi0.evolve(fq.Operator(ogm.cx()), qargs=[2, 0])
    np.testing.assert_allclose(out2.data, ref.evolve_operator(ref.basis_state(3), ogm.cx(), [2, 0]), atol=1e-12)
    # expectation value of a SparsePauliOp-like and deep copy with the rng
    op = fq.SparsePauliOp(["ZIX", "YYI", "IZZ"], [0.5, -1.5, 2.0])
    ev = sv.expectation_value(op)
    assert abs(ev - ref.expectation_value_sparse(want, op.paulis.z, op.paulis.x, op.coeffs)) < 1e-10
    sv.seed(7)
    cp = sv.copy()
    assert cp.sample_memory(5).tolist() == sv.sample_memory(5).tolist()
    np.testing.assert_allclose(sv.probabilities([2, 0]), ref.probabilities(want, [2, 0]), atol=1e-12)
    with pytest.raises(fq.QiskitError):
        ad.DeviceStatevector(np.ones(3))
    bad = fq.QuantumCircuit(1, 1); bad.measure(0, 0)
    with pytest.raises(fq.QiskitError, match="classical bits"):
        ad.DeviceStatevector.from_int(0, 2).evolve(bad)


@pytest.mark.gpu
def test_device_primitives_reuse_the_reference_pub_containers():
    ad = adapter.make_adapter(fq.namespace())
    qc = _ansatz()
    theta = np.array([[0.4, -1.1, 2.2], [1.0, 0.3, -0.5]])
    job = ad.DeviceEstimator(seed=3).run([(qc, ["ZIX", "IZZ"], theta[:, None, :])])
    assert isinstance(job, fq.PrimitiveJob)
    res = job.result()
    assert isinstance(res, fq.PrimitiveResult) and res.metadata == {"version": 2}
    evs = res[0].data.evs
    assert evs.shape == (2, 2) and res[0].metadata["circuit_metadata"] == {"tag": "ansatz"}
    for i in range(2):
        psi = _oracle_state(theta[i])
        for j, label in enumerate(["ZIX", "IZZ"]):
            assert abs(evs[i, j] - np.real(ref.expectation_value_labels(psi, [(label, 1.0)]))) < 1e-10
    # precision noise from the pub's single generator, in index order (statevector_estimator.py:151-171)
    noisy = ad.DeviceEstimator(seed=3).run([(qc, "ZIX", theta[0])], precision=0.1).result()[0].data.evs
    want = np.random.default_rng(3).normal(np.real(ref.expectation_value_labels(_oracle_state(theta[0]), [("ZIX", 1.0)])), 0.1)
    assert abs(noisy - want) < 1e-10
    # sampler: two cregs, reversed measure order, an unmeasured clbit (statevector_sampler.py:207-251)
    a, b = fq.ClassicalRegister(2, "a"), fq.ClassicalRegister(2, "b")
    mc = fq.QuantumCircuit(3, a, b)
    mc.x(0); mc.h(2)
    mc.measure(0, 1); mc.measure(2, 2)                             # a[1] <- q0, b[0] <- q2; a[0], b[1] unmeasured
    res = ad.DeviceSampler(seed=5).run([mc], shots=2000).result()[0]
    assert isinstance(res.data.a, fq.BitArray) and res.metadata["shots"] == 2000
    assert res.data.a.get_int_counts() == {2: 2000}
    cb = res.data.b.get_int_counts()
    assert set(cb) == {0, 1} and abs(cb[0] - 1000) < 150


@pytest.mark.gpu
def test_device_estimator_shares_prefixes_across_parameter_sets():
    """prefix_cache.py through the adapter: sets that differ only in the LAST parameter reuse the state
    after the shared gates; values equal the serial loop's (cache off) and the oracle's."""
    ad = adapter.make_adapter(fq.namespace())
    th = ParameterVector("t", 3)
    qc = fq.QuantumCircuit(3)
    qc.ry(th[0], 0); qc.ry(th[1], 1)
    for _ in range(4):
        qc.h(2); qc.cx(0, 1); qc.cx(1, 2); qc.h(0)
    qc.rz(th[2], 2); qc.h(2)
    theta = np.array([[0.4, -1.1, x] for x in (2.2, -0.3, 0.9, 1.7)])

    def oracle(v):
        gates = [(ogm.ry(v[0]), [0]), (ogm.ry(v[1]), [1])]
        for _ in range(4):
            gates += [(ogm.h(), [2]), (ogm.cx(), [0, 1]), (ogm.cx(), [1, 2]), (ogm.h(), [0])]
        gates += [(ogm.rz(v[2]), [2]), (ogm.h(), [2])]
        return ref.apply_gate_list(ref.basis_state(3), gates)

    est = ad.DeviceEstimator()
    evs = est.run([(qc, ["ZIX", "IZZ"], theta[:, None, :])]).result()[0].data.evs
    st = est.last_prefix_stats
    assert st is not None and st["sets"] == 4 and st["checkpoints"] == 1
    assert st["gates_applied"] == 20 + 3 * 2                     # one full stream, then two gates per set
    adapter._PREFIX_CACHE = False
    try:
        est2 = ad.DeviceEstimator()
        evs_off = est2.run([(qc, ["ZIX", "IZZ"], theta[:, None, :])]).result()[0].data.evs
        assert est2.last_prefix_stats is None
    finally:
        adapter._PREFIX_CACHE = True
    np.testing.assert_allclose(evs, evs_off, atol=1e-12)
    for i in range(4):
        psi = oracle(theta[i])
        for j, label in enumerate(["ZIX", "IZZ"]):
            assert abs(evs[i, j] - np.real(ref.expectation_value_labels(psi, [(label, 1.0)]))) < 1e-10


def test_real_qiskit_if_importable():
    try:
        ad = adapter.device_classes()
    except ImportError:
        pytest.skip("qiskit is not importable in this image (needs the Rust extension qiskit._accelerate)")
    from qiskit_b200 import _lib

    if _lib.device_count() == 0:
        pytest.skip("no CUDA device")
    from qiskit import QuantumCircuit
    from qiskit.circuit.library import efficient_su2
    from qiskit.quantum_info import SparsePauliOp, Statevector

    qc = efficient_su2(5, reps=2)
    vals = np.random.default_rng(0).uniform(-3, 3, qc.num_parameters)
    bound = qc.assign_parameters(vals)
    np.testing.assert_allclose(ad.DeviceStatevector(bound).data, Statevector(bound).data, atol=1e-10)
    op = SparsePauliOp.from_list([("ZZIII", 1.0), ("IXXII", 0.5), ("YIIIY", -0.3)])
    got = ad.DeviceEstimator().run([(qc, op, vals)]).result()[0].data.evs
    from qiskit.primitives import StatevectorEstimator

    want = StatevectorEstimator().run([(qc, op, vals)]).result()[0].data.evs
    assert abs(got - want) < 1e-10
    bell = QuantumCircuit(2); bell.h(0); bell.cx(0, 1); bell.measure_all()
    c = ad.DeviceSampler(seed=123).run([bell], shots=10000).result()[0].data.meas.get_int_counts()
    assert set(c) == {0, 3}
