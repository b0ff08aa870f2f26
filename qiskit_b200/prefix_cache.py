"""Shared-prefix caching across the parameter sets of one Estimator pub.

The reference simulates every bound circuit from |0...0> (statevector_estimator.py:157-163).  Two
parameter sets whose bound circuits start with the same gates -- a sweep over the last layer's angles,
the +-pi/2 pairs of a parameter-shift gradient, a scan of one angle -- share the state after that
common prefix.  Here the sets of a pub are ordered so that neighbours share long prefixes, every bound
circuit is lowered to its flat gate stream (lowering.py), the longest common prefix (LCP) of
neighbouring streams is measured on the streams themselves (same gate kind, qubits and matrix bytes),
and the state after the prefix the NEXT set can reuse is kept as a device-side checkpoint:

    set i:   start from the deepest checkpoint that is a prefix of stream i (or |0...0>),
             apply gates up to LCP(i, i+1), checkpoint (one device-to-device copy), apply the rest.

Checkpoints form a stack of increasing depth, all prefixes of the previous stream; those deeper than
LCP(i-1, i) are dropped before set i starts.  A checkpoint the next set cannot use is consumed in place
instead of copied.  Legal only for circuits without reset / initialize (their outcome is random per
simulation, SURVEY.md hard part 6); the caller checks that.

Values: every evaluation runs the same gates on the same engine; a cut only moves a pass boundary of
the fused-pass planner, so results agree with the uncached path to rounding (tests: 1e-12).
"""
from __future__ import annotations

import ctypes as C
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import _lib
from .lowering import GateStream

MIN_PREFIX_GATES = 8        # a shorter shared prefix is not worth a state copy
MIN_SHARE = 0.2             # trigger: some neighbouring pair shares >= 20 % of the circuit's instructions
MAX_CHECKPOINTS = 8         # per lane
FREE_STATES_NEEDED = 3      # device memory that must stay free after a checkpoint (expval scratch copies)


# ------------------------------------------------------------------ ordering --
def _parameters_of(p):
    """Parameter-like objects inside one `params` entry: a bare Parameter (this package's or Qiskit's)
    or a ParameterExpression (`.parameters`)."""
    sub = getattr(p, "parameters", None)
    if sub is not None and not callable(sub):
        try:
            return list(sub)
        except TypeError:
            return []
    if hasattr(p, "name") and not isinstance(p, (int, float, complex, np.number, np.ndarray, str)):
        return [p]
    return []


def first_use(circuit):
    """(first_inst, total): for every column of `circuit.parameters`, the index of the first top-level
    instruction whose operation (or nested definition) mentions that parameter; `total` = number of
    top-level instructions.  Parameters never found get `total`."""
    params = list(circuit.parameters)
    col = {}
    for j, p in enumerate(params):
        col[id(p)] = j
        try:
            col[("h", hash(p), getattr(p, "name", None))] = j
        except TypeError:
            pass
    data = list(circuit)
    first = [len(data)] * len(params)

    def visit(op, idx, depth=0):
        for p in getattr(op, "params", ()) or ():
            for q in _parameters_of(p):
                j = col.get(id(q))
                if j is None:
                    try:
                        j = col.get(("h", hash(q), getattr(q, "name", None)))
                    except TypeError:
                        j = None
                if j is not None and first[j] > idx:
                    first[j] = idx
        d = getattr(op, "definition", None)
        if d is not None and hasattr(d, "__iter__") and depth < 8:
            for inner in d:
                visit(inner.operation, idx, depth + 1)

    for idx, inst in enumerate(data):
        visit(inst.operation, idx)
    return first, len(data)


def order_sets(values: np.ndarray, first_inst, total: int):
    """values [S, P]: one row per parameter set.  Returns (order, share): `order` sorts the rows
    lexicographically with the columns taken in first-use order, so that neighbours agree on the longest
    run of early parameters; share[i] (i >= 1) estimates the fraction of the circuit's instructions that
    order[i-1] and order[i] have in common (position of the first instruction that uses a differing
    parameter), share[0] = 0."""
    values = np.asarray(values, dtype=float)
    n_sets, n_par = values.shape
    if n_sets == 0:
        return np.zeros(0, dtype=np.int64), np.zeros(0)
    if n_par == 0:
        return np.arange(n_sets), np.concatenate([[0.0], np.ones(n_sets - 1)])
    cols = sorted(range(n_par), key=lambda j: (first_inst[j], j))
    keyed = values[:, cols]
    order = np.lexsort(keyed.T[::-1]) if n_sets > 1 else np.zeros(1, dtype=np.int64)
    share = np.zeros(n_sets)
    firsts = np.asarray([first_inst[j] for j in cols], dtype=float)
    for i in range(1, n_sets):
        diff = np.nonzero(keyed[order[i - 1]] != keyed[order[i]])[0]
        share[i] = 1.0 if diff.size == 0 else (firsts[diff[0]] / total if total else 0.0)
    return order, share


# ------------------------------------------------------------------- streams --
def lcp(a: GateStream, b: GateStream) -> int:
    """Number of leading gates that are identical in both streams (kind, qubits, parameter, matrix bytes)."""
    n = min(len(a), len(b))
    for i in range(n):
        if a.kinds[i] != b.kinds[i] or a.qargs[i] != b.qargs[i] or a.params[i] != b.params[i]:
            return i
        da, db = a.data[i], b.data[i]
        if (da is None) != (db is None):
            return i
        if da is not None and (da.shape != db.shape or da.tobytes() != db.tobytes()):
            return i
    return n


def sub_stream(s: GateStream, lo: int, hi: int) -> GateStream:
    out = GateStream()
    out.kinds = s.kinds[lo:hi]
    out.qargs = s.qargs[lo:hi]
    out.data = s.data[lo:hi]
    out.params = s.params[lo:hi]
    return out


# -------------------------------------------------------------------- runner --
def _device_can_hold_checkpoint(num_qubits: int, device: int) -> bool:
    free = C.c_uint64()
    _lib.check(_lib.load().qsv_device_memory(device, C.byref(free), None))
    return free.value >= (FREE_STATES_NEEDED + 1) * (16 << num_qubits)


class PrefixStats:
    """What the cache did for one pub (summed over lanes)."""

    def __init__(self):
        self.sets = 0
        self.gates_total = 0        # gates the uncached path would have applied
        self.gates_applied = 0
        self.checkpoints = 0        # device-to-device copies made to keep a prefix
        self.restores = 0           # copies made to start a set from a kept prefix
        self.consumed = 0           # checkpoints used in place (no copy)

    def merge(self, other):
        for k, v in vars(other).items():
            setattr(self, k, getattr(self, k) + v)

    def as_dict(self):
        return dict(vars(self))


def run_chunk(set_ids, lower, make_state, consume, can_checkpoint=lambda: True) -> PrefixStats:
    """One lane: evaluate the parameter sets `set_ids` in this order.

    lower(set_id) -> GateStream of the bound circuit; make_state() -> |0...0> (needs `_flush(stream)` and
    `copy.copy`); consume(set_id, state) reads what the caller wants from the final state."""
    import copy as _copy

    stats = PrefixStats()
    stack: list = []            # (depth, state), increasing depth, all prefixes of `prev`
    prev = None
    nxt = lower(set_ids[0]) if len(set_ids) else None
    for pos, sid in enumerate(set_ids):
        cur = nxt
        nxt = lower(set_ids[pos + 1]) if pos + 1 < len(set_ids) else None
        l_prev = lcp(cur, prev) if prev is not None else 0
        while stack and stack[-1][0] > l_prev:
            stack.pop()
        l_next = lcp(cur, nxt) if nxt is not None else 0
        depth, state = 0, None
        if stack:
            depth = stack[-1][0]
            if depth > l_next:                 # nobody after this set can use it: take it over
                state = stack.pop()[1]
                stats.consumed += 1
            else:
                state = _copy.copy(stack[-1][1])
                stats.restores += 1
        if state is None:
            state = make_state()
        cut = l_next
        if (cut > depth and cut >= MIN_PREFIX_GATES and cut - depth >= MIN_PREFIX_GATES // 2
                and len(stack) < MAX_CHECKPOINTS and can_checkpoint()):
            state._flush(sub_stream(cur, depth, cut))
            stack.append((cut, _copy.copy(state)))
            stats.checkpoints += 1
            state._flush(sub_stream(cur, cut, len(cur)))
        else:
            state._flush(sub_stream(cur, depth, len(cur)))
        stats.sets += 1
        stats.gates_total += len(cur)
        stats.gates_applied += len(cur) - depth
        consume(sid, state)
        del state
        prev = cur
    stack.clear()
    return stats


def run_sets(order, lower, make_state, consume, *, lanes: int = 2, can_checkpoint=lambda: True) -> PrefixStats:
    """Split the ordered sets into `lanes` contiguous chunks (neighbours stay together) and run one
    chunk per host thread; ctypes drops the GIL inside the engine, so one lane lowers and plans while
    the other waits for the device."""
    order = [int(s) for s in order]
    lanes = max(1, min(int(lanes), len(order)))
    bounds = [round(i * len(order) / lanes) for i in range(lanes + 1)]
    chunks = [order[bounds[i]:bounds[i + 1]] for i in range(lanes)]
    total = PrefixStats()
    if lanes == 1:
        total.merge(run_chunk(chunks[0], lower, make_state, consume, can_checkpoint))
        return total
    with ThreadPoolExecutor(max_workers=lanes) as pool:
        for fut in [pool.submit(run_chunk, c, lower, make_state, consume, can_checkpoint) for c in chunks]:
            total.merge(fut.result())
    return total
