"""Dump the pass plan of a benchmark circuit (host only; works without a GPU).
    python scripts/plan_dump.py su2 28 | qft 30 | qv 30 [depth] | tfim 30 | hlayer 28"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from qiskit_b200 import _lib
from qiskit_b200 import circuit_library as lib
from qiskit_b200.lowering import gate_list_to_stream

mode = sys.argv[1]
n = int(sys.argv[2])
if mode == "su2":
    theta = np.random.default_rng(0).uniform(-np.pi, np.pi, 2 * n * 5)
    gates = lib.named_to_matrices(lib.efficient_su2_named(n, 4, theta))
elif mode == "qft":
    gates = lib.named_to_matrices(lib.qft_named(n))
elif mode == "qv":
    gates = lib.quantum_volume(n, int(sys.argv[3]) if len(sys.argv) > 3 else 4, seed=0)
elif mode == "tfim":
    gates = lib.named_to_matrices(lib.tfim_named(n, steps=int(sys.argv[3]) if len(sys.argv) > 3 else 2))
elif mode == "hlayer":
    gates = lib.named_to_matrices([("h", [], [q]) for q in range(n)])
st = gate_list_to_stream(gates)
arr, cnt = st.to_c()
path = "/tmp/plan_dump.txt"
_lib.check(_lib.load().qsv_debug_plan(n, arr, cnt, 0, path.encode()))
print(f"{cnt} gates")
print(open(path).read())
