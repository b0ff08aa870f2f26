"""Oracle gate-matrix table (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Independent restatement of the standard-gate matrices the statevector path
obtains through ``Gate.to_matrix()`` (reference: qiskit/circuit/gate.py:49-61;
values: crates/circuit/src/gate_matrix.rs, cited per entry).  Convention
(SURVEY.md A.1): the matrix row/column index has ``qargs[0]`` as its least
significant bit, so ``cx`` with qargs ``[ctrl, tgt]`` flips bit 1 when bit 0
is set.

The product has its own table (qiskit_b200/gates.py); tests/test_gates.py
cross-checks the two.
"""
from __future__ import annotations

import cmath
import math

import numpy as np

_S2 = 1.0 / math.sqrt(2.0)


def _c(rows):
    return np.array(rows, dtype=np.complex128)


def controlled(u: np.ndarray, n_ctrl: int = 1) -> np.ndarray:
    """Controls on the low ``n_ctrl`` index bits, 1-qubit target on the top bit.

    Follows ``make_n_controlled_gate`` (gate_matrix.rs:27-50): identity except
    on the two basis states whose control bits are all ones.
    """
    dim = 2 ** (n_ctrl + 1)
    m = np.eye(dim, dtype=np.complex128)
    lo, hi = dim // 2 - 1, dim - 1
    m[lo, lo], m[lo, hi], m[hi, lo], m[hi, hi] = u[0, 0], u[0, 1], u[1, 0], u[1, 1]
    return m


def h():  # gate_matrix.rs:53-56
    return _c([[_S2, _S2], [_S2, -_S2]])


def x():  # :58
    return _c([[0, 1], [1, 0]])


def y():  # :62
    return _c([[0, -1j], [1j, 0]])


def z():  # :60
    return _c([[1, 0], [0, -1]])


def s():  # :64
    return _c([[1, 0], [0, 1j]])


def sdg():  # :66
    return _c([[1, 0], [0, -1j]])


def t():  # :78
    return _c([[1, 0], [0, complex(_S2, _S2)]])


def tdg():  # :80-83
    return _c([[1, 0], [0, complex(_S2, -_S2)]])


def sx():  # :68-71
    return _c([[0.5 + 0.5j, 0.5 - 0.5j], [0.5 - 0.5j, 0.5 + 0.5j]])


def sxdg():  # :73-76
    return _c([[0.5 - 0.5j, 0.5 + 0.5j], [0.5 + 0.5j, 0.5 - 0.5j]])


def identity():
    return np.eye(2, dtype=np.complex128)


def rx(theta):  # :296-301
    c_, s_ = math.cos(theta / 2), math.sin(theta / 2)
    return _c([[c_, -1j * s_], [-1j * s_, c_]])


def ry(theta):  # :304-309
    c_, s_ = math.cos(theta / 2), math.sin(theta / 2)
    return _c([[c_, -s_], [s_, c_]])


def rz(theta):  # :312-315
    return _c([[cmath.exp(-0.5j * theta), 0], [0, cmath.exp(0.5j * theta)]])


def p(lam):  # :278-280
    return _c([[1, 0], [0, cmath.exp(1j * lam)]])


def u(theta, phi, lam):  # :318-325
    c_, s_ = math.cos(theta / 2), math.sin(theta / 2)
    return _c(
        [
            [c_, -cmath.exp(1j * lam) * s_],
            [cmath.exp(1j * phi) * s_, cmath.exp(1j * (phi + lam)) * c_],
        ]
    )


def u2(phi, lam):  # :334-346
    return _c(
        [
            [_S2, -cmath.exp(1j * lam) * _S2],
            [cmath.exp(1j * phi) * _S2, cmath.exp(1j * (phi + lam)) * _S2],
        ]
    )


def r(theta, phi):  # :283-293
    c_, s_ = math.cos(theta / 2), math.sin(theta / 2)
    return _c(
        [
            [c_, complex(-s_ * math.sin(phi), -s_ * math.cos(phi))],
            [complex(s_ * math.sin(phi), -s_ * math.cos(phi)), c_],
        ]
    )


def cx():  # :87
    return controlled(x())


def cy():  # :89
    return controlled(y())


def cz():  # :91
    return controlled(z())


def ch():  # :85
    return controlled(h())


def cp(lam):  # :357-364
    return np.diag([1, 1, 1, cmath.exp(1j * lam)]).astype(np.complex128)


def crx(theta):  # :367-370
    return controlled(rx(theta))


def cry(theta):  # :373-376
    return controlled(ry(theta))


def crz(theta):  # :379-382
    return controlled(rz(theta))


def swap():  # :127-132
    return _c([[1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]])


def iswap():  # :134-139
    return _c([[1, 0, 0, 0], [0, 0, 1j, 0], [0, 1j, 0, 0], [0, 0, 0, 1]])


def rxx(theta):  # :419-430
    c_, s_ = math.cos(theta / 2), -1j * math.sin(theta / 2)
    return _c([[c_, 0, 0, s_], [0, c_, s_, 0], [0, s_, c_, 0], [s_, 0, 0, c_]])


def ryy(theta):  # :433-444
    c_, s_ = math.cos(theta / 2), 1j * math.sin(theta / 2)
    return _c([[c_, 0, 0, s_], [0, c_, -s_, 0], [0, -s_, c_, 0], [s_, 0, 0, c_]])


def rzz(theta):  # :447-458
    em, ep = cmath.exp(-0.5j * theta), cmath.exp(0.5j * theta)
    return np.diag([em, ep, ep, em]).astype(np.complex128)


def rzx(theta):  # :461-472
    c_, s_ = math.cos(theta / 2), 1j * math.sin(theta / 2)
    return _c([[c_, 0, -s_, 0], [0, c_, 0, s_], [-s_, 0, c_, 0], [0, s_, 0, c_]])


def ccx():  # controls on bits 0,1, target bit 2 (make_n_controlled_gate!(X_GATE, 2))
    return controlled(x(), 2)


def ccz():
    return controlled(z(), 2)


def cswap():  # control bit 0, swaps bits 1 and 2
    m = np.eye(8, dtype=np.complex128)
    m[[3, 5]] = m[[5, 3]]
    return m


BY_NAME = {
    "id": identity, "h": h, "x": x, "y": y, "z": z, "s": s, "sdg": sdg, "t": t, "tdg": tdg,
    "sx": sx, "sxdg": sxdg, "rx": rx, "ry": ry, "rz": rz, "p": p, "u": u, "u3": u, "u2": u2, "r": r,
    "u1": p, "cx": cx, "cy": cy, "cz": cz, "ch": ch, "cp": cp, "cu1": cp, "crx": crx,
    "cry": cry, "crz": crz, "swap": swap, "iswap": iswap, "rxx": rxx, "ryy": ryy,
    "rzz": rzz, "rzx": rzx, "ccx": ccx, "ccz": ccz, "cswap": cswap,
}


def matrix(name: str, params=()) -> np.ndarray:
    """Matrix of standard gate ``name`` with float ``params``."""
    return BY_NAME[name](*[float(v) for v in params])
