"""Large configurations of BASELINE.json (configs[3], configs[4]) on 1..8 GPUs.

    python scripts/config_runs.py qv 33 33                 # 1 GPU, 137 GB state
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 \
        --master-port 29611 scripts/config_runs.py qv 33 33
    ... scripts/config_runs.py tfim 36 10                  # 8 GPUs

Prints one JSON line (rank 0): wall/device time of the gate stream, passes, swaps, achieved HBM
and NVLink bandwidth, norm (size-independent parity property), and for tfim the energy
<H> of the TFIM Hamiltonian before/after (expval all-reduce path).
torch.distributed is NOT used: ranks come from the launcher's environment, the NCCL id travels
over a TCP socket (qiskit_b200/sharded.py)."""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from qiskit_b200 import _lib
from qiskit_b200 import circuit_library as lib
from qiskit_b200.lowering import gate_list_to_stream
from qiskit_b200.sharded import DistContext, sharded_statevector

mode = sys.argv[1]
n = int(sys.argv[2])
depth = int(sys.argv[3]) if len(sys.argv) > 3 else (n if mode == "qv" else 10)
ctx = DistContext()
qsv = _lib.load()
g = int(np.log2(ctx.world))

if mode == "qv":
    gates = lib.quantum_volume(n, depth, seed=0)
    obs = None
elif mode == "tfim":
    gates = lib.named_to_matrices(lib.tfim_named(n, steps=depth))
    obs = lib.tfim_observable(n)
elif mode == "su2":
    theta = np.random.default_rng(0).uniform(-np.pi, np.pi, 2 * n * 5)
    gates = lib.named_to_matrices(lib.efficient_su2_named(n, 4, theta))
    obs = lib.heisenberg_observable(n, min(324, n * (n - 1) // 2))
else:
    raise SystemExit("mode: qv | tfim | su2")

t0 = time.perf_counter()
sv = sharded_statevector(n, ctx)
t_alloc = time.perf_counter() - t0
h = sv._h
st = gate_list_to_stream(gates)
arr, cnt = st.to_c()
out = {"mode": mode, "qubits": n, "depth": depth, "gpus": ctx.world, "gates": cnt,
       "state_bytes_per_gpu": 16 * (1 << (n - g)), "alloc_s": round(t_alloc, 3)}

ev0 = None
if obs is not None and mode == "tfim":
    ev0 = sv.expectation_value(obs).real    # on the RY(pi/4) product state? no: on |0..0>
_lib.check(qsv.qsv_sync(h))
_lib.check(qsv.qsv_reset_counters(h))
ms = C.c_float()
t0 = time.perf_counter()
_lib.check(qsv.qsv_timer_begin(h))
_lib.check(qsv.qsv_apply_gates(h, arr, cnt, 0))
_lib.check(qsv.qsv_timer_end(h, C.byref(ms)))
wall = time.perf_counter() - t0
c = sv.counters()
norm = sv._norm2()
local_bytes = 16.0 * (1 << (n - g))
dev_s = ms.value / 1e3
out.update({
    "gate_ms": round(ms.value, 2), "gate_wall_s": round(wall, 3),
    "gates_per_s": round(cnt / dev_s, 1),
    "state_passes": c["state_passes"], "write_passes": c["write_passes"], "phases": c["phases"],
    "swap_count": c["swap_count"], "swap_bytes_per_rank": c["swap_bytes"],
    "hbm_gbs_per_gpu": round((c["state_passes"] * 2 + c["write_passes"]) * local_bytes / dev_s / 1e9, 1),
    "norm_minus_1": norm - 1.0,
})
if obs is not None:
    t0 = time.perf_counter()
    ev = sv.expectation_value(obs)
    out["expval_s"] = round(time.perf_counter() - t0, 4)
    out["expval"] = ev.real
    out["expval_terms"] = len(obs.coeffs)
    if ev0 is not None:
        out["expval_initial"] = ev0
if ctx.rank == 0:
    print(json.dumps(out), flush=True)
del sv
