"""Shared test helpers: random states / circuits fed identically to oracle and engine."""
import numpy as np

from oracle import gate_matrices as ogm
from qiskit_b200.gates import GATES


def random_state(n, seed):
    rng = np.random.default_rng(seed)
    v = rng.standard_normal(1 << n) + 1j * rng.standard_normal(1 << n)
    return v / np.linalg.norm(v)


def random_unitary(k, rng):
    g = rng.standard_normal((1 << k, 1 << k)) + 1j * rng.standard_normal((1 << k, 1 << k))
    q, r = np.linalg.qr(g)
    return q * (np.diagonal(r) / np.abs(np.diagonal(r)))


def random_named_circuit(n, count, seed, names=None):
    """[(name, params, qargs)] over the standard-gate table (plus random unitaries)."""
    rng = np.random.default_rng(seed)
    names = names or [g for g in GATES if GATES[g].num_qubits <= n]
    out = []
    for _ in range(count):
        name = names[rng.integers(len(names))]
        gd = GATES[name]
        qargs = [int(q) for q in rng.choice(n, size=gd.num_qubits, replace=False)]
        params = [float(v) for v in rng.uniform(-np.pi, np.pi, gd.num_params)]
        out.append((name, params, qargs))
    return out


def oracle_matrices(named):
    """Matrices from the ORACLE's gate table (independent of the product's)."""
    out = []
    for name, params, qargs in named:
        if name == "unitary":
            out.append((np.asarray(params), list(qargs)))
        elif name == "phase":
            out.append(("phase", params[0]))
        else:
            out.append((ogm.matrix(name, params), list(qargs)))
    return out
