"""One rank of the multi-GPU parity test (launched by tests/test_gpu_dist.py through torchrun).
Checks the sharded engine against the oracle: amplitudes, expectation values, sample indices,
with both exchange transports (P2P kernel and NCCL send/recv)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import oracle_matrices, random_named_circuit
from oracle import statevector_ref as ref
from qiskit_b200 import _lib, circuit_library as lib
from qiskit_b200.sharded import DistContext, local_slab, sharded_statevector

ATOL = 1e-10


def main():
    ctx = DistContext()
    rank, world = ctx.rank, ctx.world
    g = int(np.log2(world))
    cases = []
    for n, seed in ((10, 1), (13, 2), (18, 3)):
        named = random_named_circuit(n, 80, seed=seed)
        cases.append((f"named{n}", n, [x for x in oracle_matrices(named) if not isinstance(x[0], str)]))
    cases.append(("qv16", 16, lib.quantum_volume(16, 6, seed=5)))
    cases.append(("tfim17", 17, lib.named_to_matrices(lib.tfim_named(17, steps=3))))
    cases.append(("qft15", 15, lib.named_to_matrices(lib.qft_named(15))))
    cases.append(("qv20", 20, lib.quantum_volume(20, 8, seed=9)))      # several swaps ride inside fused passes
    cases.append(("su2_19", 19, lib.named_to_matrices(lib.efficient_su2_named(
        19, 2, np.random.default_rng(4).uniform(-3, 3, 2 * 19 * 3)))))
    fused_seen = {}
    # transports: peer memory with the swap carried by a fused pass -- pushed by the last pass before it
    # ("scatter", the default where it applies) or pulled by the first pass after it ("gather") --, peer
    # memory with the standalone exchange kernel, and NCCL send/recv
    variants = (("p2p+scatter", 1, 3), ("p2p+gather", 1, 1), ("p2p", 1, 0), ("nccl", 0, 0))
    for label, p2p, fuse in variants:
        for name, n, gates in cases:
            sv = sharded_statevector(n, ctx)
            try:
                sv.set_option("dist_p2p", p2p)
                sv.set_option("dist_fuse_swap", fuse)
            except NotImplementedError:
                if rank == 0:
                    print("P2P transport unavailable (CUDA IPC); NCCL only", flush=True)
                continue
            want = ref.apply_gate_list(ref.basis_state(n), gates)
            sv.apply_gate_list(gates)
            c = sv.counters()
            # expectation values on the permuted layout (no canonicalisation needed)
            obs = lib.random_pauli_observable(n, 24, seed=n)
            # the engine evaluates an X / Y factor on a global qubit by making that qubit local first, so a
            # term needs all its X / Y qubits inside one slab: heavier terms (only possible on toy sizes:
            # 10 qubits over 8 ranks leave 7 local ones) must be refused loudly, not mis-evaluated
            heavy = np.asarray(obs.paulis.x).sum(axis=1) > n - g
            if heavy.any():
                from qiskit_b200 import SparsePauliOp
                from qiskit_b200.pauli import PauliList
                bad = SparsePauliOp(PauliList(obs.paulis.z[heavy], obs.paulis.x[heavy]), obs.coeffs[heavy])
                try:
                    sv.expectation_value(bad)
                    raise AssertionError("a term wider than a slab must raise")
                except NotImplementedError:
                    pass
                keep = ~heavy
                obs = SparsePauliOp(PauliList(obs.paulis.z[keep], obs.paulis.x[keep]), obs.coeffs[keep])
            ev = sv.expectation_value(obs)
            ev_ref = ref.expectation_value_sparse(want, obs.paulis.z, obs.paulis.x, obs.coeffs)
            assert abs(ev - ev_ref) < ATOL, (name, ev, ev_ref)
            tf = lib.tfim_observable(n)
            ev = sv.expectation_value(tf)
            ev_ref = ref.expectation_value_sparse(want, tf.paulis.z, tf.paulis.x, tf.coeffs)
            assert abs(ev - ev_ref) < ATOL, (name, "tfim", ev, ev_ref)
            assert abs(sv._norm2() - 1.0) < ATOL
            # sampling: bit-exact against the sharded summation tree
            u = np.random.default_rng(11).random(3000)
            idx = sv.sample_indices(len(u), uniforms=u).astype(np.int64)
            nl = n - g
            slab = local_slab(sv)
            lo = rank << nl
            err = float(np.max(np.abs(slab - want[lo:lo + (1 << nl)])))
            assert err < ATOL, (name, label, err)
            # bit-exact check on the engine's OWN amplitudes: every rank drops its slab in /tmp, the
            # norm all-reduce below doubles as the barrier, then each rank rebuilds the full state
            tag = f"/tmp/qsv_dist_{ctx.port}_{name}_{label}"
            np.save(f"{tag}_{rank}.npy", slab)
            sv._norm2()
            full = np.concatenate([np.load(f"{tag}_{r}.npy") for r in range(world)])
            idx_ref = ref.tree_choice_indices_sharded(ref.tree_probabilities(full), u, world)
            mism = int(np.sum(idx != idx_ref))
            assert mism == 0, (name, mism)
            if fuse:
                fused_seen[label] = fused_seen.get(label, 0) + c["fused_swaps"]
            else:
                assert c["fused_swaps"] == 0
            if rank == 0:
                print(f"[{name} transport={label}] amp err {err:.1e} swaps {c['swap_count']} (fused {c['fused_swaps']}) "
                      f"passes {c['state_passes']} sample mismatches {mism}", flush=True)
            del sv
    assert all(v > 0 for v in fused_seen.values()), f"no swap was carried by a fused pass: {fused_seen}"
    print(f"DIST_GPU_OK rank {rank} of {world}", flush=True)


if __name__ == "__main__":
    main()
