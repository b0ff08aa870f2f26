"""Small-n matrix of the engine's kernels for compute-sanitizer (memcheck / racecheck):
every kernel family once or twice at sizes where the tool's 10-100x slowdown does not matter.
Results are checked against the oracle too, so a silent corruption also fails.

    compute-sanitizer --tool memcheck  python scripts/sanitize_matrix.py
    compute-sanitizer --tool racecheck python scripts/sanitize_matrix.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np

from helpers import oracle_matrices, random_named_circuit, random_state
from oracle import statevector_ref as ref
from qiskit_b200 import DensityMatrix, Statevector
from qiskit_b200 import circuit_library as lib
from qiskit_b200.sampled_expval import sampled_expectation_value

ATOL = 1e-10
for n in (3, 9, 12, 13, 14):
    # fused passes (k_fused_pass at n >= 12, k_cluster_pass / k_tile_pass below), direct kernels
    named = random_named_circuit(n, 60, seed=n)
    gates = [g for g in oracle_matrices(named) if not isinstance(g[0], str)]
    for nf in (False, True):
        sv = Statevector.from_int(0, 1 << n)
        sv.apply_gate_list(gates, no_fusion=nf)
        want = ref.apply_gate_list(ref.basis_state(n), gates)
        assert np.max(np.abs(sv.data - want)) < ATOL, (n, nf)
    # QV (dense 2q phases), QFT (phase tables with off-tile bits), product-state prologue
    for gl in (lib.quantum_volume(n, 3, seed=n), lib.named_to_matrices(lib.qft_named(n)),
               lib.named_to_matrices(lib.efficient_su2_named(n, 2, np.linspace(-3, 3, 2 * n * 3)))):
        sv = Statevector.from_int(0, 1 << n)
        sv.apply_gate_list(gl)
        want = ref.apply_gate_list(ref.basis_state(n), gl)
        assert np.max(np.abs(sv.data - want)) < ATOL, n
    # expectation values: Z-only WHT path, X/Y families (basis change incl. |a|^2 store), direct pairs
    for obs in (lib.heisenberg_observable(n, max(1, n * (n - 1) // 4)), lib.random_pauli_observable(n, 40, seed=n),
                lib.tfim_observable(n) if n > 1 else None):
        if obs is None:
            continue
        ev = sv.expectation_value(obs)
        assert abs(ev - ref.expectation_value_sparse(want, obs.paulis.z, obs.paulis.x, obs.coeffs)) < ATOL, n
    # probabilities / marginals / sampling / collapse
    psi = random_state(n, n)
    sv = Statevector(psi)
    qa = [int(q) for q in np.random.default_rng(n).permutation(n)[: max(1, n // 2)]]
    assert np.max(np.abs(sv.probabilities(qa) - ref.probabilities(psi, qa))) < ATOL
    u = np.random.default_rng(1).random(5000)
    idx = sv.sample_indices(len(u), uniforms=u).astype(np.int64)
    assert np.array_equal(idx, ref.tree_choice_indices(ref.tree_probabilities(psi), u))
    sv.seed(3)
    got = sv.reset(qa[:2]).data
    assert np.max(np.abs(got - ref.reset(psi, qa[:2], np.random.default_rng(3)))) < 1e-9
    sv.seed(4)
    sv.measure()
# wide dense gate, density matrix, counts post-processing
psi = random_state(10, 5)
u7 = np.linalg.qr(np.random.default_rng(2).standard_normal((128, 128)) + 1j * np.random.default_rng(3).standard_normal((128, 128)))[0]
assert np.max(np.abs(Statevector(psi).evolve(u7, [9, 1, 3, 5, 7, 0, 2]).data - ref.evolve_operator(psi, u7, [9, 1, 3, 5, 7, 0, 2]))) < ATOL
rho = DensityMatrix.from_int(0, 1 << 6)
rho.apply_gate_list(lib.quantum_volume(6, 3, seed=1))
rho.reset([1, 4]); rho.probabilities([0, 5]); rho.purity()
rho.expectation_value(lib.random_pauli_observable(6, 12, seed=2))
shots = {format(int(k), "010b"): int(v) for k, v in zip(*np.unique(np.random.default_rng(0).integers(0, 1024, 5000), return_counts=True))}
sampled_expectation_value(shots, "ZZIIZ01IIZ")
print("sanitize matrix ok")
