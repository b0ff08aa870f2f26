"""world_size-2/4 gloo worker: replays the engine's sharded schedule on NumPy slabs."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch
import torch.distributed as td

import dist_emulator as emu
from helpers import oracle_matrices, random_named_circuit
from oracle import statevector_ref as ref
from qiskit_b200 import circuit_library as lib


def main():
    td.init_process_group(backend="gloo")
    rank, world = td.get_rank(), td.get_world_size()

    def exchange(sends):
        out = {p: torch.empty(v.size * 2, dtype=torch.float64) for p, v in sends.items()}
        ops = []
        for p, v in sends.items():
            ops.append(td.P2POp(td.isend, torch.from_numpy(v.view(np.float64).reshape(-1).copy()), p))
            ops.append(td.P2POp(td.irecv, out[p], p))
        for r in td.batch_isend_irecv(ops):
            r.wait()
        return {p: t.numpy().view(np.complex128) for p, t in out.items()}

    cases = []
    for n, seed in ((7, 1), (9, 2), (10, 3)):
        named = random_named_circuit(n, 50, seed=seed)
        cases.append((n, [g for g in oracle_matrices(named) if not isinstance(g[0], str)]))
    cases.append((8, lib.quantum_volume(8, 6, seed=5)))
    cases.append((9, lib.named_to_matrices(lib.tfim_named(9, steps=3))))
    cases.append((8, lib.named_to_matrices(lib.qft_named(8))))
    for n, gates in cases:
        slab, perm = emu.run_rank(n, world, rank, gates, exchange)
        gathered = [torch.empty(slab.size * 2, dtype=torch.float64) for _ in range(world)]
        td.all_gather(gathered, torch.from_numpy(slab.view(np.float64).copy()))
        full = emu.assemble([t.numpy().view(np.complex128) for t in gathered], perm, n)
        want = ref.apply_gate_list(ref.basis_state(n), gates)
        err = float(np.max(np.abs(full - want)))
        assert err < 1e-12, (n, err)
    td.barrier()
    print(f"DIST_CPU_OK rank {rank} of {world}", flush=True)
    td.destroy_process_group()


if __name__ == "__main__":
    main()
