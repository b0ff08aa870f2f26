"""Short workloads for ncu (launch list / --set full).  Run on the GPU box:

    python scripts/profile_run.py su2      # ONE estimator eval of BASELINE configs[1] (28 q)
    python scripts/profile_run.py stream   # 30 q, 6 unfusable 2-qubit gates (roofline kernel)
    python scripts/profile_run.py sampler  # 30 q QFT + 1e6 shots (configs[2])

Prints the CUDA-event time of the run so the plain run and the ncu launch list can be compared
(shares, not absolutes)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from qiskit_b200 import Statevector, _lib
from qiskit_b200 import circuit_library as lib
from qiskit_b200.lowering import gate_list_to_stream
from qiskit_b200.pauli import pauli_masks

mode = sys.argv[1] if len(sys.argv) > 1 else "su2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else {"su2": 28, "stream": 30, "sampler": 30, "qft": 30, "tfim": 32, "qv": 30}[mode]
qsv = _lib.load()
_lib.require_device()
sv = Statevector.from_int(0, 1 << n)
h = sv._h
ms = C.c_float()


def timed(fn, label):
    _lib.check(qsv.qsv_sync(h))
    _lib.check(qsv.qsv_reset_counters(h))
    _lib.check(qsv.qsv_timer_begin(h))
    out = fn()
    _lib.check(qsv.qsv_timer_end(h, C.byref(ms)))
    print(f"{label}: {ms.value:.3f} ms  counters={sv.counters()}")
    return out


if mode == "su2":
    reps = 4
    theta = np.random.default_rng(0).uniform(-np.pi, np.pi, 2 * n * (reps + 1))
    st = gate_list_to_stream(lib.named_to_matrices(lib.efficient_su2_named(n, reps, theta)))
    arr, cnt = st.to_c()
    obs = lib.heisenberg_observable(n, 324 if n >= 27 else max(1, (n * (n - 1)) // 4))
    z_mask, x_mask, n_y = pauli_masks(obs.paulis.z, obs.paulis.x)
    phases = np.ascontiguousarray((-1j) ** (n_y % 4), dtype=np.complex128)
    out = np.empty(len(z_mask))

    def gates():
        _lib.check(qsv.qsv_init_basis(h, 0))
        _lib.check(qsv.qsv_apply_gates(h, arr, cnt, 0))

    def expval():
        _lib.check(qsv.qsv_expval_pauli_batch(h, len(z_mask), z_mask.ctypes.data, x_mask.ctypes.data,
                                              phases.ctypes.data, out.ctypes.data))

    for it in range(2):
        timed(gates, f"[{it}] {cnt} gates")
        timed(expval, f"[{it}] {len(z_mask)} terms")
    print("ev =", float(np.real(np.dot(obs.coeffs, out))))
elif mode == "stream":
    gates = lib.random_two_qubit_stream(n, 6, seed=7)
    st = gate_list_to_stream(gates)
    arr, cnt = st.to_c()
    for it in range(2):
        timed(lambda: _lib.check(qsv.qsv_apply_gates(h, arr, cnt, _lib.FLAG_NO_FUSION)), f"[{it}] {cnt} unfused 2q gates")
    qv = gate_list_to_stream(lib.quantum_volume(n, 2, seed=0))
    arr2, cnt2 = qv.to_c()
    for it in range(2):
        timed(lambda: _lib.check(qsv.qsv_apply_gates(h, arr2, cnt2, 0)), f"[{it}] QV depth 2 fused ({cnt2} gates)")
elif mode == "sampler":
    st = gate_list_to_stream(lib.named_to_matrices(lib.qft_named(n)))
    arr, cnt = st.to_c()
    u = np.random.default_rng(123).random(1_000_000)
    idx = np.empty(u.size, dtype=np.uint64)
    total = C.c_double()
    for it in range(2):
        _lib.check(qsv.qsv_init_basis(h, 0))
        timed(lambda: _lib.check(qsv.qsv_apply_gates(h, arr, cnt, 0)), f"[{it}] QFT {cnt} gates")
        timed(lambda: _lib.check(qsv.qsv_sample(h, None, n, u.ctypes.data, u.size, idx.ctypes.data,
                                                 C.byref(total))), f"[{it}] sample 1e6 shots")
    print("total =", total.value, "idx[:4] =", idx[:4])
elif mode in ("qft", "tfim", "qv"):
    # one benchmark stream, timed twice (QSV_TRACE=1 prints the per-pass plan and times)
    if mode == "qft":
        gl = lib.named_to_matrices(lib.qft_named(n))
    elif mode == "tfim":
        gl = lib.named_to_matrices(lib.tfim_named(n, steps=10))
    else:
        gl = lib.quantum_volume(n, 4, seed=0)
    st = gate_list_to_stream(gl)
    arr, cnt = st.to_c()
    for it in range(2):
        _lib.check(qsv.qsv_init_basis(h, 0))
        timed(lambda: _lib.check(qsv.qsv_apply_gates(h, arr, cnt, 0)), f"[{it}] {mode} {n}q {cnt} gates")
    print("norm - 1 =", sv._norm2() - 1.0)
