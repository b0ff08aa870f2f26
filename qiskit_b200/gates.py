"""Standard-gate matrices used when a gate stream is lowered for the engine.

Host-side mirror of ``Gate.to_matrix()`` (qiskit/circuit/gate.py:49-61); values as
in crates/circuit/src/gate_matrix.rs.  Index convention: ``qargs[0]`` is the least
significant bit of the row/column index.  Each entry records how many parameters
the gate takes, how many qubits, and whether the matrix is diagonal (the engine
applies diagonal gates as phase tables).
"""
from __future__ import annotations

import numpy as np

_SQ = np.sqrt(0.5)


class GateDef:
    __slots__ = ("name", "num_qubits", "num_params", "fn", "diagonal", "_const")

    def __init__(self, name, num_qubits, num_params, fn, diagonal=False):
        self.name, self.num_qubits, self.num_params = name, num_qubits, num_params
        self.fn, self.diagonal = fn, diagonal
        self._const = None

    def matrix(self, params=()):
        if len(params) != self.num_params:
            raise ValueError(f"gate {self.name} takes {self.num_params} parameters, got {len(params)}")
        if self.num_params == 0:
            # parameter-free gates (the CX chain of an ansatz is a third of its gates): built once, shared
            # read-only -- lowering a bound circuit is host time on the Estimator's critical path
            if self._const is None:
                m = np.ascontiguousarray(self.fn(), dtype=np.complex128)
                m.flags.writeable = False
                self._const = m
            return self._const
        return np.asarray(self.fn(*[float(p) for p in params]), dtype=np.complex128)


def _ctrl(u, n_ctrl=1):
    """Embed 2x2 ``u``: controls = low index bits, target = top bit."""
    u = np.asarray(u, dtype=np.complex128)
    dim = 2 << n_ctrl
    m = np.eye(dim, dtype=np.complex128)
    a, b = dim // 2 - 1, dim - 1
    m[np.ix_([a, b], [a, b])] = u
    return m


def _rot(axis, theta):
    c, s = np.cos(theta / 2), np.sin(theta / 2)
    if axis == "x":
        return [[c, -1j * s], [-1j * s, c]]
    if axis == "y":
        return [[c, -s], [s, c]]
    return [[np.exp(-0.5j * theta), 0], [0, np.exp(0.5j * theta)]]


def _u(theta, phi, lam):
    c, s = np.cos(theta / 2), np.sin(theta / 2)
    return [[c, -np.exp(1j * lam) * s], [np.exp(1j * phi) * s, np.exp(1j * (phi + lam)) * c]]


def _two_axis(theta, kind):
    c, s = np.cos(theta / 2), np.sin(theta / 2)
    m = np.zeros((4, 4), dtype=np.complex128)
    np.fill_diagonal(m, c)
    if kind == "xx":
        m[0, 3] = m[3, 0] = m[1, 2] = m[2, 1] = -1j * s
    elif kind == "yy":
        m[0, 3] = m[3, 0] = 1j * s
        m[1, 2] = m[2, 1] = -1j * s
    elif kind == "zx":
        m[0, 2] = m[2, 0] = -1j * s
        m[1, 3] = m[3, 1] = 1j * s
    return m


_X = [[0, 1], [1, 0]]
_Y = [[0, -1j], [1j, 0]]
_Z = [[1, 0], [0, -1]]
_H = [[_SQ, _SQ], [_SQ, -_SQ]]

_TABLE = [
    GateDef("id", 1, 0, lambda: np.eye(2), True),
    GateDef("h", 1, 0, lambda: _H),
    GateDef("x", 1, 0, lambda: _X),
    GateDef("y", 1, 0, lambda: _Y),
    GateDef("z", 1, 0, lambda: _Z, True),
    GateDef("s", 1, 0, lambda: [[1, 0], [0, 1j]], True),
    GateDef("sdg", 1, 0, lambda: [[1, 0], [0, -1j]], True),
    GateDef("t", 1, 0, lambda: [[1, 0], [0, np.exp(0.25j * np.pi)]], True),
    GateDef("tdg", 1, 0, lambda: [[1, 0], [0, np.exp(-0.25j * np.pi)]], True),
    GateDef("sx", 1, 0, lambda: [[0.5 + 0.5j, 0.5 - 0.5j], [0.5 - 0.5j, 0.5 + 0.5j]]),
    GateDef("sxdg", 1, 0, lambda: [[0.5 - 0.5j, 0.5 + 0.5j], [0.5 + 0.5j, 0.5 - 0.5j]]),
    GateDef("rx", 1, 1, lambda t: _rot("x", t)),
    GateDef("ry", 1, 1, lambda t: _rot("y", t)),
    GateDef("rz", 1, 1, lambda t: _rot("z", t), True),
    GateDef("p", 1, 1, lambda l: [[1, 0], [0, np.exp(1j * l)]], True),
    GateDef("u1", 1, 1, lambda l: [[1, 0], [0, np.exp(1j * l)]], True),
    GateDef("u", 1, 3, _u),
    GateDef("u3", 1, 3, _u),
    GateDef("u2", 1, 2, lambda phi, lam: _u(np.pi / 2, phi, lam)),
    GateDef("r", 1, 2, lambda t, phi: [
        [np.cos(t / 2), -1j * np.exp(-1j * phi) * np.sin(t / 2)],
        [-1j * np.exp(1j * phi) * np.sin(t / 2), np.cos(t / 2)]]),
    GateDef("cx", 2, 0, lambda: _ctrl(_X)),
    GateDef("cy", 2, 0, lambda: _ctrl(_Y)),
    GateDef("cz", 2, 0, lambda: _ctrl(_Z), True),
    GateDef("ch", 2, 0, lambda: _ctrl(_H)),
    GateDef("cp", 2, 1, lambda l: np.diag([1, 1, 1, np.exp(1j * l)]), True),
    GateDef("cu1", 2, 1, lambda l: np.diag([1, 1, 1, np.exp(1j * l)]), True),
    GateDef("crx", 2, 1, lambda t: _ctrl(_rot("x", t))),
    GateDef("cry", 2, 1, lambda t: _ctrl(_rot("y", t))),
    GateDef("crz", 2, 1, lambda t: _ctrl(_rot("z", t)), True),
    GateDef("swap", 2, 0, lambda: [[1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]]),
    GateDef("iswap", 2, 0, lambda: [[1, 0, 0, 0], [0, 0, 1j, 0], [0, 1j, 0, 0], [0, 0, 0, 1]]),
    GateDef("rxx", 2, 1, lambda t: _two_axis(t, "xx")),
    GateDef("ryy", 2, 1, lambda t: _two_axis(t, "yy")),
    GateDef("rzx", 2, 1, lambda t: _two_axis(t, "zx")),
    GateDef("rzz", 2, 1, lambda t: np.diag(np.exp(0.5j * t * np.array([-1, 1, 1, -1]))), True),
    GateDef("ccx", 3, 0, lambda: _ctrl(_X, 2)),
    GateDef("ccz", 3, 0, lambda: _ctrl(_Z, 2), True),
    GateDef("cswap", 3, 0, lambda: np.eye(8)[[0, 1, 2, 5, 4, 3, 6, 7]]),
]

GATES = {g.name: g for g in _TABLE}


def gate_matrix(name: str, params=()) -> np.ndarray:
    try:
        return GATES[name].matrix(params)
    except KeyError:
        raise KeyError(f"unknown standard gate {name!r}") from None
