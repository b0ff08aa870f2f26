"""Bounded differential fuzz of the gate / expectation-value path against the oracle (GPU box only).

    python scripts/fuzz_gpu.py [seconds=60] [first_seed=1000]

Every case: n in 12..17, a random stream over the whole standard-gate table plus Haar unitaries on 1-3
qubits, diagonals on up to 5 qubits and global phases; from |0..0> (product-state fold on) or from a random
state; fused and one-pass-per-gate; amplitudes and a random 12-term Pauli observable against the oracle
(1e-10).  Prints the failing seeds, exits 1 if there were any."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np

from helpers import oracle_matrices, random_named_circuit, random_state, random_unitary
from oracle import statevector_ref as ref
from qiskit_b200 import Statevector
from qiskit_b200 import circuit_library as lib

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
ATOL = 1e-10
t0 = time.time()
cases, bad = 0, []
while time.time() - t0 < budget:
    rng = np.random.default_rng(seed)
    n = int(rng.integers(12, 18))
    count = int(rng.integers(20, 120))
    gates = [g for g in oracle_matrices(random_named_circuit(n, count, seed)) if not isinstance(g[0], str)]
    # sprinkle Haar unitaries, wide diagonals and global phases
    for _ in range(int(rng.integers(0, 12))):
        k = int(rng.integers(1, 4))
        qa = [int(q) for q in rng.choice(n, size=k, replace=False)]
        gates.insert(int(rng.integers(0, len(gates) + 1)), (random_unitary(k, rng), qa))
    for _ in range(int(rng.integers(0, 8))):
        k = int(rng.integers(1, 6))
        qa = [int(q) for q in rng.choice(n, size=k, replace=False)]
        gates.insert(int(rng.integers(0, len(gates) + 1)), ("diag", np.exp(1j * rng.uniform(-3, 3, 1 << k)), qa))
    if rng.integers(2):
        gates.insert(int(rng.integers(0, len(gates) + 1)), ("phase", float(rng.uniform(-3, 3))))
    from_zero = bool(rng.integers(2))
    psi0 = ref.basis_state(n) if from_zero else random_state(n, seed)
    dense = []
    for g in gates:
        if isinstance(g[0], str) and g[0] == "diag":
            dense.append((np.diag(g[1]), g[2]))
        elif isinstance(g[0], str):
            dense.append(("phase", g[1]))
        else:
            dense.append(g)
    want = psi0
    for g in dense:
        want = want * np.exp(1j * g[1]) if isinstance(g[0], str) else ref.evolve_operator(want, g[0], g[1])
    obs = lib.random_pauli_observable(n, 12, seed=seed, coeff_seed=seed + 1)
    ev_want = ref.expectation_value_sparse(want, obs.paulis.z, obs.paulis.x, obs.coeffs)
    for nf in (() if os.environ.get("FUZZ_DRY") else (False, True)):
        sv = Statevector.from_int(0, 1 << n) if from_zero else Statevector(psi0)
        sv.apply_gate_list(gates, no_fusion=nf)
        err = float(np.max(np.abs(sv.data - want)))
        ev_err = abs(sv.expectation_value(obs) - ev_want)
        if not (err < ATOL and ev_err < ATOL):
            bad.append((seed, n, count, nf, from_zero, err, ev_err))
            print("FAIL", bad[-1], flush=True)
    cases += 1
    seed += 1
print(f"fuzz: {cases} cases (seeds {seed - cases}..{seed - 1}), {len(bad)} failures, {time.time() - t0:.1f} s")
sys.exit(1 if bad else 0)
