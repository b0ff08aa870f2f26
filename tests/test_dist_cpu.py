"""N > 1 host logic on CPU: the sharded schedule (qsv_dist_plan) replayed over gloo."""
import os
import subprocess
import sys

import numpy as np
import pytest

import dist_emulator as emu
from qiskit_b200 import circuit_library as lib
from oracle import gate_matrices as ogm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_need_mask_rules():
    assert emu.need_mask(ogm.cx()) == 0b10          # control (slot 0) may stay global
    assert emu.need_mask(ogm.cp(0.3)) == 0          # diagonal: nothing needs to be local
    assert emu.need_mask(ogm.rzz(0.2)) == 0
    assert emu.need_mask(ogm.swap()) == 0b11
    assert emu.need_mask(ogm.h()) == 0b1
    assert emu.need_mask(ogm.ccx()) == 0b100


def test_plan_tfim_one_swap_per_step_and_diagonals_free():
    n, world = 12, 4
    gates = lib.named_to_matrices(lib.tfim_named(n, steps=4))
    steps, perm = emu.plan(n, world, gates)
    swaps = [s for s in steps if s[0] == "swap"]
    assert 1 <= len(swaps) <= 5          # RY layer + 4 RX layers touch the 2 global qubits
    assert all(len(s[1]) == 2 for s in swaps)   # both global qubits come in together
    applied = sorted(g for s in steps if s[0] == "apply" for g in s[1])
    assert applied == list(range(len(gates)))
    only_diag = lib.named_to_matrices([("rzz", [0.1], [10, 11]), ("cp", [0.2], [11, 3]), ("rz", [0.3], [11])])
    steps, _ = emu.plan(n, world, only_diag)
    assert [s[0] for s in steps] == ["apply"]


def test_plan_rejects_nothing_and_keeps_order_per_qubit():
    n, world = 10, 2
    gates = lib.quantum_volume(n, 5, seed=3)
    steps, _ = emu.plan(n, world, gates)
    order = [g for s in steps if s[0] == "apply" for g in s[1]]
    assert sorted(order) == list(range(len(gates)))
    last = {}
    for gid in order:                      # gates sharing a qubit keep their relative order
        for q in gates[gid][1]:
            assert last.get(q, -1) < gid
            last[q] = gid


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_schedule_replay_gloo(world):
    port = 29650 + world
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "dist_cpu_worker.py")]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-3000:]
    assert res.stdout.count("DIST_CPU_OK") == world
