"""Per-pass timing of a benchmark circuit (QSV_TRACE=1 makes the engine time every fused pass).
    QSV_TRACE=1 python scripts/trace_run.py su2|hlayer|qft|qv 28"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from qiskit_b200 import Statevector, _lib
from qiskit_b200 import circuit_library as lib
from qiskit_b200.lowering import gate_list_to_stream

mode, n = sys.argv[1], int(sys.argv[2])
if mode == "su2":
    theta = np.random.default_rng(0).uniform(-np.pi, np.pi, 2 * n * 5)
    gates = lib.named_to_matrices(lib.efficient_su2_named(n, 4, theta))
elif mode == "qft":
    gates = lib.named_to_matrices(lib.qft_named(n))
elif mode == "qv":
    gates = lib.quantum_volume(n, int(sys.argv[3]) if len(sys.argv) > 3 else 4, seed=0)
elif mode == "hlayer":
    gates = lib.named_to_matrices([("h", [], [q]) for q in range(n)])
st = gate_list_to_stream(gates)
arr, cnt = st.to_c()
sv = Statevector.from_int(0, 1 << n)
qsv = _lib.load()
for it in range(2):
    print(f"--- iteration {it}", file=sys.stderr)
    _lib.check(qsv.qsv_init_basis(sv._h, 0))
    _lib.check(qsv.qsv_apply_gates(sv._h, arr, cnt, 0))
    _lib.check(qsv.qsv_sync(sv._h))
print("norm2", sv._norm2())
