"""Shared-prefix caching across parameter sets on the device (qiskit_b200/prefix_cache.py, SURVEY.md 8f
N1): StatevectorEstimator values with the cache on == cache off (1e-12) == the oracle's loop over bound
circuits (statevector_estimator.py:157-171; 1e-10), and the cache really skipped work."""
import numpy as np
import pytest

from oracle import statevector_ref as ref
from qiskit_b200 import StatevectorEstimator
from qiskit_b200 import circuit_library as lib
from qiskit_b200.circuit import Parameter, QuantumCircuit

from helpers import oracle_matrices

pytestmark = pytest.mark.gpu
ATOL = 1e-10


def _oracle_evs(n, reps, values, obs):
    out = []
    for v in values:
        psi = ref.apply_gate_list(ref.basis_state(n), oracle_matrices(lib.efficient_su2_named(n, reps, v)))
        out.append(ref.expectation_value_sparse(psi, obs.paulis.z, obs.paulis.x, obs.coeffs).real)
    return np.array(out)


def _run(circuit, obs, values, cache):
    est = StatevectorEstimator()
    est._PREFIX_CACHE = cache
    evs = est.run([(circuit, obs, values)]).result()[0].data.evs
    return evs, est.last_prefix_stats


@pytest.mark.parametrize("n", [6, 13])
def test_last_layer_sweep_matches_uncached_and_oracle(n):
    reps = 2
    qc = lib.efficient_su2_circuit(n, reps)
    p = len(qc.parameters)
    rng = np.random.default_rng(n)
    values = np.tile(rng.uniform(-np.pi, np.pi, p), (7, 1))
    values[:, -2 * n:] = rng.uniform(-np.pi, np.pi, (7, 2 * n))
    obs = lib.random_pauli_observable(n, 20, seed=1, coeff_seed=2)
    on, stats = _run(qc, obs, values, True)
    off, none = _run(qc, obs, values, False)
    assert none is None and stats is not None
    assert stats["sets"] == 7 and stats["checkpoints"] >= 1
    assert stats["gates_applied"] < 0.6 * stats["gates_total"]
    np.testing.assert_allclose(on, off, rtol=0, atol=1e-12)
    np.testing.assert_allclose(on, _oracle_evs(n, reps, values, obs), rtol=0, atol=ATOL)


def test_parameter_shift_pairs_and_broadcast_observables():
    """2P shifted parameter sets x 2 observables (shape (2P, 1) against (2,)): result positions must not
    depend on the order in which the cache visits the sets."""
    n, reps = 5, 1
    qc = lib.efficient_su2_circuit(n, reps)
    p = len(qc.parameters)
    base = np.random.default_rng(5).uniform(-np.pi, np.pi, p)
    rows = []
    for k in range(p):
        for sgn in (+1, -1):
            v = base.copy()
            v[k] += sgn * np.pi / 2
            rows.append(v)
    values = np.array(rows)[:, None, :]                                  # (2P, 1, P)
    obs_a = lib.random_pauli_observable(n, 9, seed=3, coeff_seed=4)
    obs_b = lib.random_pauli_observable(n, 5, seed=6, coeff_seed=7)
    est = StatevectorEstimator()
    evs = est.run([(qc, [obs_a, obs_b], values)]).result()[0].data.evs
    stats = est.last_prefix_stats
    assert evs.shape == (2 * p, 2)
    assert stats is not None and stats["gates_applied"] < stats["gates_total"]
    want = np.stack([_oracle_evs(n, reps, values[:, 0, :], o) for o in (obs_a, obs_b)], axis=1)
    np.testing.assert_allclose(evs, want, rtol=0, atol=ATOL)
    # gradient by the shift rule == finite-difference of the oracle (sanity of the use case itself)
    grad = 0.5 * (evs[0::2, 0] - evs[1::2, 0])
    eps = 1e-6
    fd = []
    for k in range(p):
        vp, vm = base.copy(), base.copy()
        vp[k] += eps
        vm[k] -= eps
        e = _oracle_evs(n, reps, [vp, vm], obs_a)
        fd.append((e[0] - e[1]) / (2 * eps))
    np.testing.assert_allclose(grad, fd, atol=1e-6)


def test_random_sets_keep_the_two_lane_path_and_reset_circuits_are_never_cached():
    n = 4
    qc = lib.efficient_su2_circuit(n, 1)
    values = np.random.default_rng(9).uniform(-3, 3, (4, len(qc.parameters)))
    obs = lib.random_pauli_observable(n, 6, seed=1, coeff_seed=2)
    evs, stats = _run(qc, obs, values, True)
    assert stats is None                                                 # nothing shared: old path
    np.testing.assert_allclose(evs, _oracle_evs(n, 1, values, obs), atol=ATOL)
    # a circuit with reset is stochastic per simulation: every index is re-simulated from the pub's one
    # generator (statevector_estimator.py:151,160-171) even when all parameter sets share a prefix
    a = Parameter("a")
    rc = QuantumCircuit(2)
    rc.h(0); rc.cx(0, 1); rc.reset(0); rc.ry(a, 1)                       # noqa: E702
    est = StatevectorEstimator(seed=3)
    r1 = est.run([(rc, ["ZI", "IZ"], [[0.1], [0.1]])]).result()[0].data.evs
    assert est.last_prefix_stats is None
    r2 = StatevectorEstimator(seed=3).run([(rc, ["ZI", "IZ"], [[0.1], [0.1]])]).result()[0].data.evs
    np.testing.assert_array_equal(r1, r2)


def test_sampler_parameter_sweep_is_identical_with_and_without_the_cache():
    """StatevectorSampler seeds every index from the same seed (statevector_sampler.py:209), so the order
    in which the cache visits the parameter sets cannot change any shot."""
    from qiskit_b200 import StatevectorSampler

    n, reps = 7, 2
    qc = lib.efficient_su2_circuit(n, reps)
    qc.measure_all()
    p = len(qc.parameters)
    rng = np.random.default_rng(2)
    values = np.tile(rng.uniform(-np.pi, np.pi, p), (5, 1))
    values[:, -2 * n:] = rng.uniform(-np.pi, np.pi, (5, 2 * n))       # last RY + RZ layer (RZ alone would not move any probability)
    values = values[[2, 0, 4, 3, 1]]
    out = {}
    for cache in (True, False):
        smp = StatevectorSampler(seed=11)
        smp._PREFIX_CACHE = cache
        res = smp.run([(qc, values)], shots=2000).result()[0]
        out[cache] = res.data.meas.array.copy()
        stats = smp.last_prefix_stats
        assert (stats is not None) == cache
        if cache:
            assert stats["sets"] == 5 and stats["gates_applied"] < 0.6 * stats["gates_total"]
    assert out[True].shape == (5, 2000, 1)
    np.testing.assert_array_equal(out[True], out[False])
    assert len({out[True][i].tobytes() for i in range(5)}) == 5          # the five sets really differ
