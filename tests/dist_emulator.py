"""CPU replay of the sharded gate schedule (TEST INFRASTRUCTURE).

The engine's scheduler (csrc/qsv_dist.cu: which gates run on the current layout, which
global<->local qubit swaps happen, which block of a gate a rank applies) is host logic.  This
module asks the library for the schedule through the host-only entry ``qsv_dist_plan`` and
executes it on NumPy slabs, one process per rank, with the exchange done by
``torch.distributed`` (gloo).  The result must equal the oracle's full-state simulation.
"""
import ctypes as C

import numpy as np

from oracle import statevector_ref as ref
from qiskit_b200 import _lib

MAXQ = _lib.MAX_GATE_QUBITS


def need_mask(mat) -> int:
    g = _lib.QsvGate()
    m = _lib.as_c128(mat)
    g.kind = _lib.GATE_MATRIX
    g.k = int(round(np.log2(m.shape[0])))
    g.data = m.ctypes.data
    out = C.c_uint32()
    _lib.check(_lib.load().qsv_dist_need_mask(C.byref(g), C.byref(out)))
    return out.value


def plan(n, world, gates, min_lpos=0, perm=None):
    """-> (steps, perm_out); steps = [("apply", [gate ids]) | ("swap", [(in_q, out_q), ...])]."""
    G = len(gates)
    ks = np.array([len(q) for _, q in gates], dtype=np.int32)
    qa = np.zeros((G, MAXQ), dtype=np.int32)
    need = np.zeros(G, dtype=np.uint32)
    for i, (m, q) in enumerate(gates):
        qa[i, :len(q)] = q
        need[i] = need_mask(m)
    cap = 16 * G + 64
    words = np.zeros(cap, dtype=np.int32)
    nw = C.c_uint64()
    perm_in = None if perm is None else (C.c_int * n)(*perm)
    perm_out = (C.c_int * n)()
    _lib.check(_lib.load().qsv_dist_plan(n, world, perm_in, G, ks.ctypes.data, qa.ctypes.data, need.ctypes.data,
                                         min_lpos, words.ctypes.data, cap, C.byref(nw), perm_out))
    steps, w = [], 0
    while w < nw.value:
        kind, cnt = int(words[w]), int(words[w + 1])
        w += 2
        if kind == 0:
            steps.append(("apply", [int(v) for v in words[w:w + cnt]]))
            w += cnt
        else:
            steps.append(("swap", [(int(words[w + 2 * i]), int(words[w + 2 * i + 1])) for i in range(cnt)]))
            w += 2 * cnt
    return steps, list(perm_out)


def restrict(mat, phys, n_local, rank):
    """Block of ``mat`` this rank applies: slots on global bits fixed to the rank's bits."""
    k = len(phys)
    keep = [j for j in range(k) if phys[j] < n_local]
    fixed = sum(((rank >> (phys[j] - n_local)) & 1) << j for j in range(k) if phys[j] >= n_local)
    idx = [fixed | sum(((s >> t) & 1) << keep[t] for t in range(len(keep))) for s in range(1 << len(keep))]
    return mat[np.ix_(idx, idx)], [phys[j] for j in keep]


def sub_slab_index(n_local, lpos, value):
    """Index tuple selecting the amplitudes whose bits ``lpos`` read ``value``."""
    idx = [slice(None)] * n_local
    for i, p in enumerate(lpos):
        idx[n_local - 1 - p] = (value >> i) & 1
    return tuple(idx)


def run_rank(n, world, rank, gates, exchange, min_lpos=0):
    """Execute the schedule for ``rank``.  ``exchange(sends: {peer: array}) -> {peer: array}``
    moves sub-slabs between ranks.  Returns (local slab, perm)."""
    g = int(np.log2(world))
    nl = n - g
    slab = np.zeros(1 << nl, dtype=np.complex128)
    if rank == 0:
        slab[0] = 1.0
    perm = list(range(n))
    steps, perm_out = plan(n, world, gates, min_lpos)
    for kind, body in steps:
        if kind == "apply":
            for gid in body:
                mat, qargs = gates[gid]
                phys = [perm[q] for q in qargs]
                sub, lq = restrict(np.asarray(mat, dtype=complex), phys, nl, rank)
                if not lq:
                    slab = slab * sub[0, 0]
                else:
                    slab = ref.evolve_operator(slab, sub, lq)
        else:
            gpos = [perm[a] for a, _ in body]
            lpos = [perm[b] for _, b in body]
            k = len(body)
            a_val = sum(((rank >> (gpos[i] - nl)) & 1) << i for i in range(k))
            t = slab.reshape((2,) * nl)
            sends = {}
            for j in range(1 << k):
                if j == a_val:
                    continue
                peer = rank
                for i in range(k):
                    bit = gpos[i] - nl
                    peer = (peer & ~(1 << bit)) | (((j >> i) & 1) << bit)
                sends[peer] = (j, np.ascontiguousarray(t[sub_slab_index(nl, lpos, j)]))
            recvd = exchange({p: v for p, (j, v) in sends.items()})
            t = t.copy()
            for peer, (j, v) in sends.items():
                t[sub_slab_index(nl, lpos, j)] = recvd[peer].reshape(v.shape)
            slab = t.reshape(-1)
            for (qa, qb) in body:
                perm[qa], perm[qb] = perm[qb], perm[qa]
    assert perm == perm_out
    return slab, perm


def assemble(slabs, perm, n):
    """Full logical state from per-rank slabs in layout ``perm`` (logical q on physical perm[q])."""
    phys = np.concatenate(slabs)
    idx = np.arange(1 << n)
    src = np.zeros_like(idx)
    for q in range(n):
        src |= ((idx >> q) & 1) << perm[q]
    return phys[src]
