"""CPU test of the fused swap pass's gather arithmetic (csrc/qsv_fused.cuh FpRemote, built by
csrc/qsv_gate.cu build_fp_remote) through the host-only emulator qsv_debug_fused_swap, which runs the
same __host__ __device__ helpers the kernel uses.

Checked for k = 1..3 swapped bits landing on tile rows, copy-in thread bits and outside the tile:
  1. every destination index of the slab is written exactly once;
  2. the source of destination x is the rank whose swapped global bits equal x's swapped local bits,
     at the address x with those bits replaced by the reader's own rank bits (the definition of the
     global<->local swap, csrc/qsv_dist.cu);
  3. the handshake is closed: the tile that owns a source address on the source rank carries the SAME
     tile number as the reading tile, the reader signals that rank, and every tile is read by exactly
     2^nin ranks (the increment of the arrival-counter target).
"""
import ctypes as C

import numpy as np
import pytest

from qiskit_b200 import _lib
from qiskit_b200 import circuit_library as lib
from qiskit_b200.lowering import gate_list_to_stream


def _emulate(n_local, lpos, a, arr, cnt, mode=1):
    qsv = _lib.load()
    size = 1 << n_local
    src_j = np.zeros(size, dtype=np.uint32)
    src_addr = np.zeros(size, dtype=np.uint64)
    tile_of = np.zeros(size, dtype=np.uint32)
    sig = np.zeros((size >> 12) * 16, dtype=np.uint32)
    nin, ok = C.c_int(), C.c_int()
    _lib.check(qsv.qsv_debug_fused_swap(mode, n_local, len(lpos), _lib.int_array(lpos), a, arr, cnt,
                                        src_j.ctypes.data, src_addr.ctypes.data, tile_of.ctypes.data,
                                        sig.ctypes.data, C.byref(nin), C.byref(ok)))
    return bool(ok.value), nin.value, src_j, src_addr, tile_of, sig.reshape(-1, 16)


def _gates_on(n_local, qubits, seed):
    """dense two-qubit gates on the given local qubits (they become tile bits of the first pass)"""
    rng = np.random.default_rng(seed)
    gates = []
    for i in range(0, len(qubits) - 1, 2):
        gates.append((lib.haar_unitary(4, rng), [qubits[i], qubits[i + 1]]))
    return gates


CASES = [
    # (n_local, lpos, gate qubits)                      where the swapped bits land
    (15, [14], [14, 13, 12, 11]),                       # k=1, tile row bit
    (15, [14], [5, 6, 7, 8]),                           # k=1, outside the tile
    (16, [15, 14], [15, 9, 14, 8, 13, 12, 11, 10]),     # k=2, both tile bits
    (16, [15, 14], [15, 9, 8, 7]),                      # k=2, one in, one out
    (16, [13, 15], [4, 5, 6, 7]),                       # k=2, both out, unordered lpos
    (17, [16, 15, 14], [16, 15, 14, 13]),               # k=3, all tile bits
    (17, [16, 9, 5], [16, 12, 9, 13, 11, 10]),          # k=3, row bit + thread bits (5 fills the low tile bits)
    (17, [14, 15, 16], [6, 7, 8, 10]),                  # k=3, all out
    (14, [13, 12], [13, 2, 12, 1]),                     # small slab: 4 tiles
    (24, [23, 22], [5, 6, 7, 8]),                       # 4096 tiles: the tile counter's bits are permuted
]


@pytest.mark.parametrize("mode", [1, 2, 3])
@pytest.mark.parametrize("n_local,lpos,gq", CASES)
def test_fused_swap_remote_map(n_local, lpos, gq, mode):
    """mode 1: gather (copy-in mapping of the first pass after the swap); modes 2 / 3: scatter through the
    copy-out / direct-store mapping of the last pass before it.  The swap is an involution, so the same
    three properties hold with "source" read as "destination"."""
    k = len(lpos)
    st = gate_list_to_stream(_gates_on(n_local, gq, seed=n_local + k))
    arr, cnt = st.to_c()
    size = 1 << n_local
    x = np.arange(size, dtype=np.uint64)
    lmask = sum(1 << p for p in lpos)
    per_rank = {}
    for a in range(1 << k):
        ok, nin, src_j, src_addr, tile_of, sig = _emulate(n_local, lpos, a, arr, cnt, mode)
        if mode == 3 and not ok:
            pytest.skip("the pass has no direct-store phase")
        assert ok, "the pass of the test stream must be a fused pass"
        # 1. coverage (the emulator fails on double coverage; here: nothing left out)
        assert int(np.sum(tile_of == 0xFFFFFFFF)) == 0
        # 2. swap semantics
        want_j = np.zeros(size, dtype=np.uint64)
        a_l = 0
        for i, p in enumerate(lpos):
            want_j |= ((x >> np.uint64(p)) & np.uint64(1)) << np.uint64(i)
            a_l |= ((a >> i) & 1) << p
        assert np.array_equal(src_j.astype(np.uint64), want_j)
        assert np.array_equal(src_addr, (x & np.uint64(~lmask & (size - 1))) | np.uint64(a_l))
        per_rank[a] = (nin, src_j, src_addr, tile_of, sig)
    # 3. handshake closure across the 2^k ranks of the group
    nin = per_rank[0][0]
    readers = {}   # (source rank, tile number on the source rank) -> set of reader ranks
    for a, (nin_a, src_j, src_addr, tile_of, sig) in per_rank.items():
        assert nin_a == nin
        for j in range(1 << k):
            sel = np.nonzero(src_j == j)[0]
            if sel.size == 0:
                continue
            tiles_here = tile_of[sel]                             # reading tiles on rank a
            tiles_there = per_rank[j][3][src_addr[sel].astype(np.int64)]   # owner tiles on rank j
            assert np.array_equal(tiles_here, tiles_there), (a, j)
            for t in np.unique(tiles_here):
                assert j in sig[t, : 1 << nin].tolist(), (a, j, int(t))
                readers.setdefault((j, int(t)), set()).add(a)
        # a tile signals 2^nin distinct ranks and nothing else
        assert all(len(set(row[: 1 << nin].tolist())) == (1 << nin) for row in sig)
        assert np.all(sig[:, 1 << nin:] == 0xFFFFFFFF)
    ntiles = size >> 12
    for j in range(1 << k):
        for t in range(ntiles):
            assert len(readers[(j, t)]) == (1 << nin), (j, t, readers.get((j, t)))
