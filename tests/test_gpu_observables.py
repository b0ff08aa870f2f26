"""Estimator with SparseObservable inputs on the device: projector terms (X / Y / Z eigenstate projectors)
against <psi| M |psi> with the oracle's state and the dense matrix of the observable."""
import numpy as np
import pytest

from oracle import statevector_ref as ref
from qiskit_b200 import SparseObservable, StatevectorEstimator
from qiskit_b200 import circuit_library as lib

from helpers import oracle_matrices

pytestmark = pytest.mark.gpu
ATOL = 1e-10


def test_estimator_with_projector_observables():
    n, reps = 5, 1
    qc = lib.efficient_su2_circuit(n, reps)
    values = np.random.default_rng(0).uniform(-np.pi, np.pi, (3, len(qc.parameters)))
    rng = np.random.default_rng(1)
    terms = [("".join(rng.choice(list("IXYZ+-rl01"), n)), float(rng.normal())) for _ in range(6)]
    obs = SparseObservable.from_list(terms)
    label = "0+1rI"
    evs = StatevectorEstimator().run([(qc, [[obs], [label], [{"ZZ01l": 0.5, "XIIII": 2.0}]], values)]).result()[0].data.evs
    assert evs.shape == (3, 3)
    mats = [obs.to_matrix(), SparseObservable.from_label(label).to_matrix(),
            SparseObservable.from_list([("ZZ01l", 0.5), ("XIIII", 2.0)]).to_matrix()]
    for j, v in enumerate(values):
        psi = ref.apply_gate_list(ref.basis_state(n), oracle_matrices(lib.efficient_su2_named(n, reps, v)))
        for i, m in enumerate(mats):
            want = np.vdot(psi, m @ psi)
            assert abs(want.imag) < 1e-12
            assert abs(evs[i, j] - want.real) < ATOL, (i, j)
    # a projector onto a computational basis state is that outcome's probability
    from qiskit_b200 import Statevector

    sv = Statevector.from_int(0, 1 << n)
    sv.apply_gate_list(oracle_matrices(lib.efficient_su2_named(n, reps, values[0])))
    p = sv.probabilities()
    ev = StatevectorEstimator().run([(qc, "01101", values[0])]).result()[0].data.evs
    assert abs(float(ev) - p[0b01101]) < ATOL
