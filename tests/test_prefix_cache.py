"""Shared-prefix caching across parameter sets (qiskit_b200/prefix_cache.py): the host
logic on the CPU -- ordering, LCP on lowered gate streams, the checkpoint stack -- against the oracle's
gate-by-gate simulation, with a stand-in state that applies streams through oracle/statevector_ref.py and
counts what it was asked to do."""
import copy

import numpy as np

from oracle import statevector_ref as ref
from qiskit_b200 import _lib, prefix_cache as pc
from qiskit_b200 import circuit_library as lib
from qiskit_b200.circuit import Parameter, QuantumCircuit
from qiskit_b200.lowering import GateStream, walk


class OracleState:
    """`_flush(stream)` + `copy.copy`: what run_chunk needs from a state."""

    copies = 0
    gates = 0

    def __init__(self, n):
        self.n = n
        self.psi = ref.basis_state(n)

    def _flush(self, stream):
        for kind, qa, data, par in zip(stream.kinds, stream.qargs, stream.data, stream.params):
            OracleState.gates += 1
            if kind == _lib.GATE_MATRIX:
                self.psi = ref.evolve_operator(self.psi, data, list(qa))
            elif kind == _lib.GATE_PHASE:
                self.psi = self.psi * np.exp(1j * par)
            else:
                raise AssertionError(kind)

    def __copy__(self):
        OracleState.copies += 1
        new = type(self)(self.n)
        new.psi = self.psi.copy()
        return new


def _lower(circuit, values):
    stream = GateStream()
    walk(None, circuit.assign_parameters(values).to_instruction(), None, stream)
    return stream


def _run(circuit, values, order, lanes=1):
    n = circuit.num_qubits
    finals = {}
    OracleState.copies = OracleState.gates = 0
    stats = pc.run_sets(order, lambda s: _lower(circuit, values[s]), lambda: OracleState(n),
                        lambda s, st: finals.__setitem__(s, st.psi.copy()), lanes=lanes)
    stats.gates_seen_by_states = OracleState.gates
    for s in range(len(values)):
        direct = OracleState(n)
        direct._flush(_lower(circuit, values[s]))
        assert np.max(np.abs(finals[s] - direct.psi)) < 1e-12, s
    return stats


def test_first_use_and_ordering():
    a, b, c = Parameter("a"), Parameter("b"), Parameter("c")
    qc = QuantumCircuit(2)
    qc.ry(c, 0); qc.cx(0, 1); qc.rz(a, 1); qc.h(0); qc.rx(b, 0)        # noqa: E702
    assert [p.name for p in qc.parameters] == ["a", "b", "c"]
    first, total = pc.first_use(qc)
    assert (first, total) == ([2, 4, 0], 5)
    # rows differ only in b (last used) / in c (first used)
    vals = np.array([[0.1, 0.5, 0.3], [0.1, 0.2, 0.3], [0.1, 0.2, 0.9], [0.1, 0.7, 0.3]])
    order, share = pc.order_sets(vals, first, total)
    # sorted by (c, a, b): the three c = 0.3 rows are neighbours, ordered by b; the c = 0.9 row is last
    assert list(order) == [1, 0, 3, 2]
    np.testing.assert_allclose(share, [0.0, 0.8, 0.8, 0.0])
    # nested definitions are searched too
    outer = QuantumCircuit(2)
    outer.h(1)
    outer.append(qc.to_gate(), [0, 1])
    first2, total2 = pc.first_use(outer)
    assert total2 == 2 and set(first2) == {1}


def test_lcp_and_sub_stream():
    qc = lib.efficient_su2_circuit(3, 1)
    p = len(qc.parameters)
    v0 = np.linspace(0.1, 1.0, p)
    v1 = v0.copy(); v1[-1] += 0.5                                       # noqa: E702
    s0, s1 = _lower(qc, v0), _lower(qc, v1)
    assert pc.lcp(s0, s0) == len(s0)
    cut = pc.lcp(s0, s1)
    assert 0 < cut < len(s0)
    assert len(pc.sub_stream(s0, cut, len(s0))) == len(s0) - cut
    v2 = v0.copy(); v2[0] += 0.5                                        # noqa: E702
    assert pc.lcp(s0, _lower(qc, v2)) == 0


def test_sweep_of_the_last_layer_shares_one_checkpoint():
    n, reps = 4, 2
    qc = lib.efficient_su2_circuit(n, reps)
    p = len(qc.parameters)
    base = np.random.default_rng(0).uniform(-3, 3, p)
    values = np.tile(base, (6, 1))
    values[:, -1] = np.linspace(0, 1, 6)                               # only the very last angle moves
    first, total = pc.first_use(qc)
    order, share = pc.order_sets(values, first, total)
    assert share.max() > 0.8
    stats = _run(qc, values, order)
    d = stats.as_dict()
    assert d["sets"] == 6 and d["checkpoints"] == 1
    assert d["gates_applied"] < 0.4 * d["gates_total"]
    assert stats.gates_seen_by_states == d["gates_applied"]
    # two lanes: each lane keeps its own checkpoint, results identical
    stats2 = _run(qc, values, order, lanes=2)
    assert stats2.as_dict()["checkpoints"] == 2


def test_parameter_shift_pairs_reuse_growing_prefixes():
    n, reps = 3, 2
    qc = lib.real_amplitudes_circuit(n, reps)
    p = len(qc.parameters)
    base = np.random.default_rng(1).uniform(-3, 3, p)
    rows = []
    for k in range(p):
        for sgn in (+1, -1):
            v = base.copy(); v[k] += sgn * np.pi / 2                    # noqa: E702
            rows.append(v)
    values = np.array(rows)
    first, total = pc.first_use(qc)
    order, share = pc.order_sets(values, first, total)
    stats = _run(qc, values, order).as_dict()
    assert stats["sets"] == 2 * p
    assert stats["gates_applied"] < 0.8 * stats["gates_total"]
    assert stats["checkpoints"] >= 1 and stats["restores"] + stats["consumed"] >= stats["checkpoints"]


def test_random_sets_do_not_trigger_and_stay_correct_if_forced():
    qc = lib.efficient_su2_circuit(4, 1)
    values = np.random.default_rng(2).uniform(-3, 3, (5, len(qc.parameters)))
    first, total = pc.first_use(qc)
    order, share = pc.order_sets(values, first, total)
    assert share.max() < pc.MIN_SHARE                                   # the estimator keeps its old path
    stats = _run(qc, values, order).as_dict()                           # forced through the runner anyway
    assert stats["checkpoints"] == 0 and stats["gates_applied"] == stats["gates_total"]


def test_duplicate_sets_and_checkpoint_refusal():
    qc = lib.efficient_su2_circuit(3, 1)
    v = np.random.default_rng(3).uniform(-3, 3, len(qc.parameters))
    values = np.array([v, v, v])
    order, share = pc.order_sets(values, *pc.first_use(qc))
    assert list(share) == [0.0, 1.0, 1.0]
    stats = _run(qc, values, order).as_dict()
    assert stats["checkpoints"] == 1 and stats["gates_applied"] == stats["gates_total"] // 3
    # no memory for a checkpoint: everything is simulated from scratch, results still right
    n = qc.num_qubits
    finals = {}
    st = pc.run_chunk([0, 1, 2], lambda s: _lower(qc, values[s]), lambda: OracleState(n),
                      lambda s, state: finals.__setitem__(s, copy.copy(state).psi), can_checkpoint=lambda: False)
    assert st.checkpoints == 0 and st.gates_applied == st.gates_total
    assert np.max(np.abs(finals[0] - finals[2])) == 0


def test_estimator_integration_with_a_stand_in_state(monkeypatch):
    """StatevectorEstimator._run_shared_prefix end to end on the CPU: the device Statevector is replaced by
    the oracle-backed stand-in, so broadcasting, set ordering and result placement are what is tested."""
    from qiskit_b200 import primitives

    class FakeSV(OracleState):
        @classmethod
        def from_int(cls, i, dims, *, device=0):
            assert i == 0
            return cls(int(np.log2(dims)))

        def expectation_value(self, op):
            return ref.expectation_value_sparse(self.psi, op.paulis.z, op.paulis.x, op.coeffs)

    monkeypatch.setattr(primitives, "Statevector", FakeSV)
    monkeypatch.setattr(pc, "_device_can_hold_checkpoint", lambda n, device: True)
    n, reps = 4, 1
    qc = lib.efficient_su2_circuit(n, reps)
    p = len(qc.parameters)
    rng = np.random.default_rng(4)
    values = np.tile(rng.uniform(-3, 3, p), (5, 1))
    values[:, -3:] = rng.uniform(-3, 3, (5, 3))
    values = values[[3, 0, 4, 1, 2]][:, None, :]                         # (5, 1, P), deliberately unsorted
    obs = [lib.random_pauli_observable(n, 5, seed=s, coeff_seed=s + 1) for s in (1, 2, 3)]
    est = primitives.StatevectorEstimator()
    evs = est.run([(qc, obs, values)]).result()[0].data.evs
    assert evs.shape == (5, 3)
    st = est.last_prefix_stats
    assert st is not None and st["sets"] == 5 and st["gates_applied"] < st["gates_total"]
    for i in range(5):
        direct = OracleState(n)
        direct._flush(_lower(qc, values[i, 0]))
        for j, o in enumerate(obs):
            want = ref.expectation_value_sparse(direct.psi, o.paulis.z, o.paulis.x, o.coeffs).real
            assert abs(evs[i, j] - want) < 1e-12
    # switched off -> the old path is asked (which needs a device: here it is simply not reached)
    est2 = primitives.StatevectorEstimator()
    est2._PREFIX_CACHE = False
    assert est2._run_shared_prefix(qc, None, None, None, []) is None
