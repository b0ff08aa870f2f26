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
elif mode == "tfim":
    gates = lib.named_to_matrices(lib.tfim_named(n, steps=int(sys.argv[3]) if len(sys.argv) > 3 else 10))
elif mode == "u3":   # layers of Haar one-qubit gates + CZ ladder: generic complex 1q workload
    rng = np.random.default_rng(5)
    gates = []
    for layer in range(int(sys.argv[3]) if len(sys.argv) > 3 else 4):
        for q in range(n):
            g = rng.standard_normal((2, 2)) + 1j * rng.standard_normal((2, 2))
            u, r = np.linalg.qr(g)
            gates.append((u * (np.diagonal(r) / np.abs(np.diagonal(r))), [q]))
        for q in range(layer % 2, n - 1, 2):
            gates.append((np.diag([1, 1, 1, -1]).astype(complex), [q, q + 1]))
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
import ctypes as C
ms = C.c_float()
_lib.check(qsv.qsv_init_basis(sv._h, 0))
_lib.check(qsv.qsv_sync(sv._h))
_lib.check(qsv.qsv_timer_begin(sv._h))
_lib.check(qsv.qsv_apply_gates(sv._h, arr, cnt, 0))
_lib.check(qsv.qsv_timer_end(sv._h, C.byref(ms)))
print(f"total {ms.value:.3f} ms for {cnt} gates (includes per-pass trace syncs when QSV_TRACE is set)")
print("norm2", sv._norm2())
