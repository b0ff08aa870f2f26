"""NumPy restatement of the reference DensityMatrix path (TEST INFRASTRUCTURE).

See oracle/__init__.py for who may import this.  A density matrix is a C-contiguous
(2**n, 2**n) complex128 array; qubit q is bit q of the row and of the column index.
Every function names the reference lines it follows (paths relative to /root/reference).
"""
from __future__ import annotations

import numpy as np

from . import statevector_ref as svref


def from_statevector(psi: np.ndarray) -> np.ndarray:
    """densitymatrix.py:104: ``np.outer(v, conj(v))``."""
    psi = np.asarray(psi, dtype=np.complex128)
    return np.outer(psi, np.conj(psi))


def basis_state(n: int, index: int = 0) -> np.ndarray:
    """DensityMatrix.from_int (densitymatrix.py:576-602)."""
    rho = np.zeros((1 << n, 1 << n), dtype=np.complex128)
    rho[index, index] = 1.0
    return rho


def evolve_operator(rho: np.ndarray, mat: np.ndarray, qargs=None) -> np.ndarray:
    """rho' = U rho U^dagger with U = mat on qargs (densitymatrix.py:720-750).

    ``qargs=None``: ``np.dot(op, rho).dot(op.T.conj())`` (:731-733).  Otherwise the
    reference contracts ``mat`` with the row axes and ``mat.adjoint()`` with the column
    axes of the (2,)*2n tensor (``Operator._einsum_matmul``, :736-747); here the same two
    contractions are done with the statevector restatement on the rows, then on the
    columns of the matrix (rows: U a, columns: (U conj(a^T))^H).
    """
    rho = np.asarray(rho, dtype=np.complex128)
    mat = np.asarray(mat, dtype=np.complex128)
    if qargs is None:
        return np.dot(mat, rho).dot(mat.T.conj())
    dim = rho.shape[0]
    left = np.empty_like(rho)
    for c in range(dim):  # U applied to every column vector
        left[:, c] = svref.evolve_operator(np.ascontiguousarray(rho[:, c]), mat, qargs)
    out = np.empty_like(rho)
    for r in range(dim):  # row vector times U^dagger  ==  conj(U conj(row))
        out[r, :] = np.conj(svref.evolve_operator(np.conj(left[r, :]), mat, qargs))
    return out


def apply_gate_list(rho: np.ndarray, gates) -> np.ndarray:
    """Gate lists as in statevector_ref.apply_gate_list; a global phase cancels in U rho U^dagger."""
    for g in gates:
        if isinstance(g[0], str):
            if g[0] == "phase":
                continue
            if g[0] == "diag":
                rho = evolve_operator(rho, np.diag(np.asarray(g[1], dtype=np.complex128)), list(g[2]))
                continue
            raise ValueError(g[0])
        rho = evolve_operator(rho, g[0], list(g[1]))
    return rho


def density_expval_pauli_no_x(data_f: np.ndarray, num_qubits: int, z_mask: int) -> float:
    """crates/accelerate/src/pauli_exp_val.rs:128-154 on the F-order ravel ``data_f``."""
    rows = 1 << num_qubits
    i = np.arange(rows, dtype=np.uint64)
    val = data_f[i * np.uint64(rows + 1)].real.copy()
    val[svref._parity(i & np.uint64(z_mask)) == 1] *= -1.0
    return float(np.sum(val))


def density_expval_pauli_with_x(data_f, num_qubits: int, z_mask: int, x_mask: int, phase: complex, x_max: int) -> float:
    """crates/accelerate/src/pauli_exp_val.rs:159-194."""
    rows = 1 << num_qubits
    mask_u = ~((1 << (x_max + 1)) - 1) & ((1 << 63) - 1)
    mask_l = (1 << x_max) - 1
    i = np.arange(rows >> 1, dtype=np.uint64)
    index_vec = ((i << np.uint64(1)) & np.uint64(mask_u)) | (i & np.uint64(mask_l))
    index_mat = (index_vec ^ np.uint64(x_mask)) + np.uint64(rows) * index_vec
    val = 2.0 * (phase * data_f[index_mat]).real
    val[svref._parity(index_vec & np.uint64(z_mask)) == 1] *= -1.0
    return float(np.sum(val))


def expectation_value_pauli(rho, z, x, group_phase: int = 0, qargs=None) -> complex:
    """DensityMatrix._expectation_value_pauli (densitymatrix.py:366-397).  ``z``/``x`` are the
    Pauli's bool arrays, ``group_phase`` its ``Pauli.phase`` (power of -i)."""
    z = np.asarray(z, dtype=bool)
    x = np.asarray(x, dtype=bool)
    n = int(round(np.log2(rho.shape[0])))
    qubits = np.arange(len(z)) if qargs is None else np.array(qargs)
    x_mask = int(np.dot(1 << qubits, x))
    z_mask = int(np.dot(1 << qubits, z))
    pauli_phase = (-1j) ** group_phase if group_phase else 1
    if x_mask + z_mask == 0:
        return pauli_phase * np.trace(rho)
    data = np.ravel(rho, order="F")
    if x_mask == 0:
        return pauli_phase * density_expval_pauli_no_x(data, n, z_mask)
    x_max = int(qubits[x][-1])
    y_phase = (-1j) ** int(np.sum(x & z))
    return pauli_phase * density_expval_pauli_with_x(data, n, z_mask, x_mask, y_phase, x_max)


def expectation_value_sparse(rho, zs, xs, coeffs, qargs=None) -> complex:
    """densitymatrix.py:411-417: sum of coeff * <P> over the terms."""
    return sum(c * expectation_value_pauli(rho, z, x, 0, qargs) for z, x, c in zip(zs, xs, coeffs))


def expectation_value_operator(rho, mat, qargs=None) -> complex:
    """densitymatrix.py:418-420: ``trace(evolve(oper).data)``-like, i.e. Tr(O rho)."""
    n = int(round(np.log2(rho.shape[0])))
    mat = np.asarray(mat, dtype=np.complex128)
    if qargs is not None:
        full = np.empty_like(rho)
        eye = np.eye(rho.shape[0], dtype=np.complex128)
        for c in range(rho.shape[0]):
            full[:, c] = svref.evolve_operator(np.ascontiguousarray(eye[:, c]), mat, list(qargs))
        mat = full
    assert mat.shape == (1 << n, 1 << n)
    return complex(np.trace(np.dot(mat, rho)))


def probabilities(rho, qargs=None, decimals=None) -> np.ndarray:
    """densitymatrix.py:498-506."""
    n = int(round(np.log2(rho.shape[0])))
    probs = svref.subsystem_probabilities(np.abs(rho.diagonal()), n, qargs)
    probs = np.clip(probs, a_min=0, a_max=1)
    if decimals is not None:
        probs = probs.round(decimals=decimals)
    return probs


def purity(rho) -> complex:
    """densitymatrix.py:221-227."""
    return complex(np.trace(np.dot(rho, rho)))


def reset(rho, qargs=None) -> np.ndarray:
    """densitymatrix.py:508-537: all qubits -> |0><0|; a subset -> the reset SuperOp
    (row 0 of the superoperator = vec(identity)), i.e. partial trace over qargs (x) |0><0|."""
    n = int(round(np.log2(rho.shape[0])))
    if qargs is None:
        return basis_state(n, 0)
    qargs = list(qargs)
    m = len(qargs)
    out = np.zeros_like(rho)
    others = [q for q in range(n) if q not in qargs]
    for ro in range(1 << (n - m)):
        r0 = sum(((ro >> j) & 1) << others[j] for j in range(n - m))
        for co in range(1 << (n - m)):
            c0 = sum(((co >> j) & 1) << others[j] for j in range(n - m))
            acc = 0.0
            for k in range(1 << m):
                kb = sum(((k >> j) & 1) << qargs[j] for j in range(m))
                acc = acc + rho[r0 | kb, c0 | kb]
            out[r0, c0] = acc
    return out


def reverse_qargs(rho) -> np.ndarray:
    """densitymatrix.py:344-364."""
    n = int(round(np.log2(rho.shape[0])))
    axes = tuple(range(n - 1, -1, -1))
    axes = axes + tuple(len(axes) + i for i in axes)
    return np.reshape(np.transpose(np.reshape(rho, (2,) * (2 * n)), axes), rho.shape)


def partial_transpose(rho, qargs) -> np.ndarray:
    """densitymatrix.py:839-866: swap the row and column axes of the listed qubits."""
    n = int(round(np.log2(rho.shape[0])))
    arr = np.reshape(rho, (2,) * (2 * n))
    lst = list(range(2 * n))
    for i in n - 1 - np.array(qargs):  # qubit q is tensor axis n-1-q (:849)
        lst[i], lst[i + n] = lst[i + n], lst[i]
    return np.reshape(np.transpose(arr, lst), rho.shape)


def evolve_superop(rho: np.ndarray, sop: np.ndarray, qargs=None) -> np.ndarray:
    """``SuperOp._evolve`` (qiskit/quantum_info/operators/channel/superop.py:234-292) for qubit channels:
    full system = superoperator times the column-stacked (Fortran-order) density matrix; a subsystem =
    the reference's einsum contraction over the row and column tensor axes of ``qargs``."""
    n = int(round(np.log2(rho.shape[0])))
    if qargs is None:
        vec = np.ravel(rho, order="F")
        return np.reshape(np.dot(sop, vec), rho.shape, order="F")
    k = len(qargs)
    tensor = np.reshape(rho, (2,) * (2 * n))
    mat = np.reshape(sop, (2,) * (4 * k))
    indices = [n - 1 - q for q in qargs] + [2 * n - 1 - q for q in qargs]
    rank = tensor.ndim
    indices_tensor = list(range(rank))
    for j, index in enumerate(indices):
        indices_tensor[index] = rank + j
    mat_contract = list(reversed(range(rank, rank + len(indices))))
    mat_free = list(reversed(indices))
    out = np.einsum(tensor, indices_tensor, mat, mat_free + mat_contract)
    return np.reshape(out, rho.shape)


def kraus_to_superop(kraus) -> np.ndarray:
    """``_kraus_to_superop`` (operators/channel/transformations.py): sum_i conj(K_i) (x) K_i."""
    return sum(np.kron(np.conj(np.asarray(k, dtype=complex)), np.asarray(k, dtype=complex)) for k in kraus)
