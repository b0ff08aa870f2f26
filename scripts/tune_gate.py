"""GPU experiment: achieved HBM GB/s of single-gate passes for different kernels/tunables.
Run on the GPU box:  python scripts/tune_gate.py [n_qubits]"""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from qiskit_b200 import Statevector, _lib
from qiskit_b200 import circuit_library as lib
from qiskit_b200.lowering import gate_list_to_stream

n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
qsv = _lib.load()
sv = Statevector.from_int(0, 1 << n)
h = sv._h
rng = np.random.default_rng(0)
sv.apply_gate_list([(lib.haar_unitary(2, rng), [q]) for q in range(n)])   # dense non-trivial state

def time_gates(gates, flags=_lib.FLAG_NO_FUSION, reps=1):
    st = gate_list_to_stream(gates)
    arr, cnt = st.to_c()
    _lib.check(qsv.qsv_apply_gates(h, arr, min(cnt, 3), flags))
    _lib.check(qsv.qsv_sync(h)); _lib.check(qsv.qsv_reset_counters(h))
    ms = C.c_float()
    _lib.check(qsv.qsv_timer_begin(h))
    for _ in range(reps):
        _lib.check(qsv.qsv_apply_gates(h, arr, cnt, flags))
    _lib.check(qsv.qsv_timer_end(h, C.byref(ms)))
    passes = sv.counters()["state_passes"]
    return ms.value / passes, passes * 32.0 * (1 << n) / (ms.value / 1e3) / 1e9

def u(k): return lib.haar_unitary(1 << k, rng)
cases = {
    "2q both high (n-1,n-7)": [(u(2), [n - 1, n - 7])] * 8,
    "2q mid (12,17)": [(u(2), [12, 17])] * 8,
    "2q (5,6)": [(u(2), [5, 6])] * 8,
    "2q one low (0,n-3)": [(u(2), [0, n - 3])] * 8,
    "2q one low (2,9)": [(u(2), [2, 9])] * 8,
    "2q both low (0,1)": [(u(2), [0, 1])] * 8,
    "2q both low (1,3)": [(u(2), [3, 1])] * 8,
    "2q random pairs": lib.random_two_qubit_stream(n, 24, seed=7),
    "1q high": [(u(1), [n - 2])] * 8,
    "1q q=0": [(u(1), [0])] * 8,
    "1q real (ry) q=10": [(np.array([[0.8, -0.6], [0.6, 0.8]]), [10])] * 8,
}
configs = [("direct", dict(use_direct=1)),
           ("tile T12 L4 256", dict(use_direct=0, tile_bits=12, low_bits=4, tile_threads=256)),
           ("tile T11 L4 128", dict(use_direct=0, tile_bits=11, low_bits=4, tile_threads=128)),
           ("tile T11 L4 256", dict(use_direct=0, tile_bits=11, low_bits=4, tile_threads=256)),
           ("tile T12 L4 512", dict(use_direct=0, tile_bits=12, low_bits=4, tile_threads=512)),
           ("tile T13 L5 512", dict(use_direct=0, tile_bits=13, low_bits=5, tile_threads=512)),
           ("tile T10 L4 128", dict(use_direct=0, tile_bits=10, low_bits=4, tile_threads=128))]
print(f"n={n}  (GB/s algorithmic, 32*2^n bytes per pass; ms per pass)")
for cname, opts in configs:
    for k, v in opts.items():
        sv.set_option(k, v)
    row = []
    for name, gates in cases.items():
        ms, gbs = time_gates(gates)
        row.append(f"{name}: {gbs:6.0f} ({ms:.2f}ms)")
    print(f"[{cname}]\n   " + "\n   ".join(row))
# fused QV
for cname, opts in configs[1:]:
    for k, v in opts.items():
        sv.set_option(k, v)
    sv.set_option("use_direct", 1)
    qv = lib.quantum_volume(n, 3, seed=0)
    st = gate_list_to_stream(qv); arr, cnt = st.to_c()
    _lib.check(qsv.qsv_apply_gates(h, arr, cnt, 0)); _lib.check(qsv.qsv_sync(h)); _lib.check(qsv.qsv_reset_counters(h))
    ms = C.c_float(); _lib.check(qsv.qsv_timer_begin(h)); _lib.check(qsv.qsv_apply_gates(h, arr, cnt, 0)); _lib.check(qsv.qsv_timer_end(h, C.byref(ms)))
    p = sv.counters()["state_passes"]
    print(f"[fused QV depth3 {cname}] gates={cnt} passes={p} ms={ms.value:.1f} ms/pass={ms.value/p:.2f} gates/s={cnt/(ms.value/1e3):.0f}")
