"""Device-resident ``DensityMatrix``.

Same public behaviour as ``qiskit.quantum_info.DensityMatrix`` for qubit states
(qiskit/quantum_info/states/densitymatrix.py) on the engine of ``Statevector``: rho of n
qubits is the 2n-qubit handle with amplitude index ``(row << n) | col`` (C-order ravel).

* ``evolve`` (:305-342, ``_evolve_operator`` :720-750, ``_append_instruction`` :752-797):
  U on the row qubits, conj(U) on the column qubits -- one fused gate stream;
* ``expectation_value`` (:366-420): the strided kernels of
  crates/accelerate/src/pauli_exp_val.rs:128-194, all terms of a SparsePauliOp in one call;
* ``probabilities`` (:422-506), ``sample_memory / sample_counts / measure``
  (quantum_state.py:236-331: ``Generator.choice`` on the <= 2^n probabilities, the reference's
  own NumPy call), ``reset`` (:508-537), ``trace / purity / tensor / expand / reverse_qargs /
  partial_transpose / to_statevector``.
There is no CPU fallback: the matrix only crosses PCIe on explicit ``.data`` access.
"""
from __future__ import annotations

import ctypes as C
from numbers import Number

import numpy as np

from . import _lib
from .exceptions import QiskitError
from .lowering import GateStream, walk
from .pauli import pauli_masks
from .statevector import Statevector, _is_circuit, _is_instruction, _ket_strings

MAX_QUBITS = 16  # 2n-qubit handle: 16 qubits = 64 GiB


class _DensityStream(GateStream):
    """Gate stream of rho -> U rho U^dagger: every leaf matrix is appended twice, on the row
    qubits and (conjugated) on the column qubits; a global phase cancels."""

    def __init__(self, n: int):
        super().__init__()
        self._n = n

    def matrix(self, mat, qargs):
        mat = _lib.as_c128(mat)
        super().matrix(mat, [int(q) + self._n for q in qargs])
        super().matrix(np.conj(mat), qargs)

    def diagonal(self, diag, qargs):
        diag = _lib.as_c128(diag)
        super().diagonal(diag, [int(q) + self._n for q in qargs])
        super().diagonal(np.conj(diag), qargs)

    def phase(self, theta: float):
        return


class DensityMatrix:
    """DensityMatrix on the GPU.  ``DensityMatrix(data)`` accepts a 2^n x 2^n matrix, a vector
    of 2^n amplitudes (projector), a device ``Statevector``, a circuit / instruction (simulated
    from |0..0><0..0|) or another ``DensityMatrix``."""

    atol = 1e-8
    rtol = 1e-5

    def __init__(self, data, dims=None, *, device: int = 0):
        self._rng_generator = None
        self._device = int(device)
        self._h = None
        lib = _lib.load()
        if isinstance(data, DensityMatrix):
            self._n, self._device = data._n, data._device
            self._h = data._clone_handle()
            return
        if _is_circuit(data) or _is_instruction(data):
            inst = data.to_instruction() if _is_circuit(data) else data
            self._set_n(int(inst.num_qubits))
            self._h = _lib.acquire(2 * self._n, self._device)
            _lib.check(lib.qsv_init_basis(self._h, 0))
            self._evolve_instruction_inplace(inst, None)
            return
        if isinstance(data, Statevector):
            self._device = data._device
            self._from_device_vector(data)
            return
        if not isinstance(data, (list, np.ndarray)):
            if hasattr(data, "to_operator"):
                data = data.to_operator().data
            elif hasattr(data, "to_matrix"):
                data = data.to_matrix()
            else:
                raise QiskitError("Invalid input data format for DensityMatrix")
        arr = np.asarray(data, dtype=np.complex128)
        if arr.ndim == 2 and arr.shape[1] == 1 and arr.shape[0] != 1:
            arr = arr.reshape(-1)
        if arr.ndim == 1:
            self._from_device_vector(Statevector(arr, device=self._device))
            return
        if arr.ndim != 2 or arr.shape[0] != arr.shape[1]:
            raise QiskitError("Invalid DensityMatrix input: not a square matrix.")
        n = int(round(np.log2(arr.shape[0]))) if arr.shape[0] else -1
        if arr.shape[0] == 0 or (1 << n) != arr.shape[0]:
            raise QiskitError("Only qubit density matrices (2^n x 2^n) are supported on the device")
        if dims is not None:
            d = (dims,) if isinstance(dims, (int, np.integer)) else tuple(dims)
            if int(np.prod(d)) != arr.shape[0] or (len(d) != 1 and any(x != 2 for x in d)):
                raise QiskitError("Only qubit subsystem dimensions are supported on the device")
        self._set_n(n)
        self._h = _lib.acquire(2 * n, self._device)
        flat = _lib.as_c128(arr).reshape(-1)
        _lib.check(lib.qsv_upload(self._h, flat.ctypes.data, 0, flat.size))

    # ------------------------------------------------------------ plumbing --
    def _set_n(self, n: int):
        if n > MAX_QUBITS:
            raise QiskitError(f"DensityMatrix of {n} qubits exceeds the device limit of {MAX_QUBITS}")
        self._n = n

    def _from_device_vector(self, sv: Statevector):
        """rho = |v><v| (densitymatrix.py:104) as one Kronecker kernel on the device."""
        n = sv._n
        self._set_n(n)
        self._h = _lib.acquire(2 * n, self._device)
        rows = _lib.int_array(range(n, 2 * n))
        cols = _lib.int_array(range(n))
        _lib.check(_lib.load().qsv_tensor(self._h, sv._h, sv._h, rows, cols, 1))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None:
            try:
                _lib.release(h, 2 * self._n, self._device)
            except Exception:  # noqa: BLE001 - interpreter shutdown
                pass

    def _clone_handle(self):
        h = _lib.acquire(2 * self._n, self._device)
        _lib.check(_lib.load().qsv_copy_state(h, self._h))
        return h

    def _blank_like(self, n=None):
        ret = object.__new__(type(self))
        ret._rng_generator = self._rng_generator
        ret._device = self._device
        ret._n = self._n if n is None else n
        ret._h = None
        return ret

    def _new_like(self):
        ret = self._blank_like()
        ret._h = self._clone_handle()
        return ret

    def _flush(self, stream: GateStream, flags: int = 0):
        if len(stream):
            arr, n = stream.to_c()
            _lib.check(_lib.load().qsv_apply_gates(self._h, arr, n, flags))
            stream.clear()

    def _qargs(self, qargs):
        qa = list(range(self._n)) if qargs is None else [int(q) for q in qargs]
        if len(set(qa)) != len(qa) or any(q < 0 or q >= self._n for q in qa):
            raise QiskitError(f"invalid qargs {qa} for a {self._n}-qubit DensityMatrix")
        return qa

    # --------------------------------------------------------- basic props --
    @property
    def num_qubits(self):
        return self._n

    @property
    def dim(self):
        return 1 << self._n

    def dims(self, qargs=None):
        return (2,) * (self._n if qargs is None else len(qargs))

    @property
    def data(self) -> np.ndarray:
        """rho as a host array (device -> host copy of 16 * 4^n bytes)."""
        out = np.empty(1 << (2 * self._n), dtype=np.complex128)
        _lib.check(_lib.load().qsv_download(self._h, out.ctypes.data, 0, out.size))
        return out.reshape(1 << self._n, 1 << self._n)

    def __array__(self, dtype=None, copy=None):
        d = self.data
        return d if dtype is None else d.astype(dtype)

    def __eq__(self, other):
        return isinstance(other, DensityMatrix) and self._n == other._n and bool(
            np.allclose(self.data, other.data, rtol=self.rtol, atol=self.atol))

    def __repr__(self):
        return f"DensityMatrix(<device {self._device}, {self._n} qubits>, dims={self.dims()})"

    def copy(self):
        return self._new_like()

    def seed(self, value=None):
        """quantum_state.py:78-91."""
        if value is None:
            self._rng_generator = None
        elif isinstance(value, np.random.Generator):
            self._rng_generator = value
        else:
            self._rng_generator = np.random.default_rng(value)

    @property
    def _rng(self):
        return np.random.default_rng() if self._rng_generator is None else self._rng_generator

    # ------------------------------------------------------------ algebra ---
    def _scalar(self, fn) -> complex:
        out = (C.c_double * 2)()
        _lib.check(fn(self._h, out))
        return complex(out[0], out[1])

    def trace(self):
        """densitymatrix.py:217-219."""
        return self._scalar(_lib.load().qsv_density_trace)

    def purity(self):
        """densitymatrix.py:221-227: trace(rho . rho)."""
        return self._scalar(_lib.load().qsv_density_purity)

    def is_valid(self, atol=None, rtol=None):
        """densitymatrix.py:193-206: unit trace, Hermitian, positive semidefinite.  A validation
        helper, not part of the hot path: the eigenvalue test runs on a host copy like the
        reference's ``is_positive_semidefinite_matrix``."""
        atol = self.atol if atol is None else atol
        rtol = self.rtol if rtol is None else rtol
        if not np.allclose(self.trace(), 1, rtol=rtol, atol=atol):
            return False
        mat = self.data
        if not np.allclose(mat, mat.conj().T, rtol=rtol, atol=atol):
            return False
        vals = np.linalg.eigvalsh(mat)
        return bool(np.all(vals > -atol))

    def conjugate(self):
        """densitymatrix.py:213-215: element-wise conjugate."""
        ret = self._new_like()
        _lib.check(_lib.load().qsv_conjugate(ret._h))
        return ret

    def _multiply(self, other):
        """densitymatrix.py:287-303."""
        if not isinstance(other, Number):
            raise QiskitError("other is not a number")
        ret = self._new_like()
        c = complex(other)
        _lib.check(_lib.load().qsv_scale(ret._h, c.real, c.imag))
        return ret

    __mul__ = _multiply
    __rmul__ = _multiply

    def __truediv__(self, other):
        return self._multiply(1 / other)

    def __neg__(self):
        return self._multiply(-1)

    def _add(self, other):
        """densitymatrix.py:267-285."""
        if not isinstance(other, DensityMatrix):
            other = DensityMatrix(other, device=self._device)
        if other._n != self._n:
            raise QiskitError("DensityMatrix dimensions do not match")
        ret = self._new_like()
        _lib.check(_lib.load().qsv_axpy(ret._h, other._h, 1.0, 0.0))
        return ret

    __add__ = _add

    def __sub__(self, other):
        if not isinstance(other, DensityMatrix):
            other = DensityMatrix(other, device=self._device)
        if other._n != self._n:
            raise QiskitError("DensityMatrix dimensions do not match")
        ret = self._new_like()
        _lib.check(_lib.load().qsv_axpy(ret._h, other._h, -1.0, 0.0))
        return ret

    def _kron(self, hi: "DensityMatrix", lo: "DensityMatrix"):
        """np.kron(hi, lo): row = (r_hi, r_lo), col = (c_hi, c_lo)."""
        n_hi, n_lo = hi._n, lo._n
        n = n_hi + n_lo
        ret = self._blank_like(n)
        ret._set_n(n)
        ret._h = _lib.acquire(2 * n, self._device)
        # bit j of the factor's index (row << n_f | col) -> bit of the product index
        hi_pos = [n_lo + j for j in range(n_hi)] + [n + n_lo + j for j in range(n_hi)]
        lo_pos = list(range(n_lo)) + [n + j for j in range(n_lo)]
        _lib.check(_lib.load().qsv_tensor(ret._h, hi._h, lo._h, _lib.int_array(hi_pos),
                                          _lib.int_array(lo_pos), 0))
        return ret

    def tensor(self, other):
        """densitymatrix.py:229-246: self (x) other (``other`` takes the low qubits)."""
        if not isinstance(other, DensityMatrix):
            other = DensityMatrix(other, device=self._device)
        return self._kron(self, other)

    def expand(self, other):
        """densitymatrix.py:248-265: other (x) self."""
        if not isinstance(other, DensityMatrix):
            other = DensityMatrix(other, device=self._device)
        return self._kron(other, self)

    def _permuted(self, src_bit):
        ret = self._blank_like()
        ret._h = _lib.acquire(2 * self._n, self._device)
        _lib.check(_lib.load().qsv_permute_qubits(ret._h, self._h, _lib.int_array(src_bit)))
        return ret

    def reverse_qargs(self):
        """densitymatrix.py:344-364: qubit q <-> n-1-q on rows and columns."""
        n = self._n
        return self._permuted([n - 1 - k for k in range(n)] + [2 * n - 1 - k for k in range(n)])

    def partial_transpose(self, qargs):
        """densitymatrix.py:839-866: swap the row and column bit of every listed qubit."""
        n = self._n
        src = list(range(2 * n))
        for q in self._qargs(qargs):
            src[q], src[q + n] = q + n, q
        return self._permuted(src)

    def to_statevector(self, atol=None, rtol=None):
        """densitymatrix.py:807-837 (host eigen-decomposition of the 2^n x 2^n matrix, as the
        reference; the result is uploaded as a device Statevector)."""
        atol = self.atol if atol is None else atol
        rtol = self.rtol if rtol is None else rtol
        mat = self.data
        if not np.allclose(mat, mat.conj().T, rtol=rtol, atol=atol):
            raise QiskitError("Not a valid density matrix (non-hermitian).")
        evals, evecs = np.linalg.eig(mat)
        nonzero = np.logical_not(np.isclose(np.abs(evals), 0, atol=atol, rtol=rtol))
        if len(evals[nonzero]) != 1 or not np.isclose(evals[nonzero][0], 1, atol=atol, rtol=rtol):
            raise QiskitError("Density matrix is not a pure state")
        psi = evecs[:, np.argmax(evals)]
        return Statevector(psi, device=self._device)

    # ------------------------------------------------------------- evolve ---
    def evolve(self, other, qargs=None):
        """densitymatrix.py:305-342: returns a new state; ``self`` is untouched."""
        if qargs is None:
            qargs = getattr(other, "qargs", None)
        ret = self._new_like()
        if _is_circuit(other):
            other = other.to_instruction()
        if _is_instruction(other):
            ret._evolve_instruction_inplace(other, qargs)
            return ret
        from .channels import as_superop_matrix

        sop = as_superop_matrix(other)        # QuantumChannel branch (densitymatrix.py:329-336)
        if sop is not None:
            ret._evolve_superop_inplace(sop, qargs)
            return ret
        mat = np.asarray(getattr(other, "data", other), dtype=np.complex128)
        ret._evolve_operator_inplace(mat, qargs)
        return ret

    def _evolve_superop_inplace(self, sop, qargs=None):
        """``SuperOp._evolve`` (operators/channel/superop.py:234-292) on the device: the superoperator acts on
        the column-stacked density matrix (row bits low, column bits high), i.e. it is one dense 2k-qubit
        matrix on the handle qubits ``[q + n for q in qargs] + qargs`` (rows live on qubits q + n)."""
        qa = self._qargs(qargs)
        k = len(qa)
        sop = np.asarray(sop, dtype=np.complex128)
        if sop.shape != (1 << (2 * k), 1 << (2 * k)):
            raise QiskitError("Operator input dimensions are not equal to statevector subsystem dimensions.")
        if 2 * k > _lib.MAX_GATE_QUBITS:
            raise QiskitError(f"a SuperOp on {k} qubits exceeds the engine limit of {_lib.MAX_GATE_QUBITS // 2}")
        stream = GateStream()
        stream.matrix(sop, [q + self._n for q in qa] + qa)
        self._flush(stream)

    def _evolve_operator_inplace(self, mat, qargs=None):
        """Device counterpart of ``_evolve_operator`` (densitymatrix.py:720-750)."""
        qa = self._qargs(qargs)
        if mat.ndim != 2 or mat.shape[0] != mat.shape[1] or mat.shape[0] != (1 << len(qa)):
            raise QiskitError(
                "Operator input dimensions are not equal to density matrix subsystem dimensions."
            )
        stream = _DensityStream(self._n)
        stream.matrix(mat, qa)
        self._flush(stream)

    def _evolve_instruction_inplace(self, inst, qargs=None):
        """Device counterpart of ``_append_instruction`` (densitymatrix.py:752-797); the SuperOp
        branch (:771-776) is outside the device path."""
        if qargs is not None and len(qargs) != inst.num_qubits:
            raise QiskitError("qargs length does not match the instruction's number of qubits")
        if qargs is None and inst.num_qubits != self._n:
            raise QiskitError(
                f"Cannot apply a {inst.num_qubits}-qubit instruction to a {self._n}-qubit "
                "DensityMatrix without qargs"
            )
        stream = _DensityStream(self._n)
        walk(self, inst, None if qargs is None else [int(q) for q in qargs], stream)
        self._flush(stream)

    def apply_gate_list(self, gates):
        """In-place, Qiskit-free entry (``gates`` as in ``lowering.gate_list_to_stream``)."""
        stream = _DensityStream(self._n)
        for g in gates:
            if isinstance(g[0], str):
                if g[0] == "diag":
                    stream.diagonal(g[1], g[2])
                elif g[0] != "phase":
                    raise QiskitError(f"unknown gate-list entry {g[0]!r}")
            else:
                stream.matrix(g[0], g[1])
        self._flush(stream)
        return self

    def _reset_inplace(self, qargs):
        lib = _lib.load()
        if qargs is None:
            _lib.check(lib.qsv_init_basis(self._h, 0))
            return
        qa = self._qargs(qargs)
        # the reset channel factors per qubit: chunks of <= 10 qubits (the kernel's group size)
        for i in range(0, len(qa), 10):
            part = qa[i:i + 10]
            _lib.check(lib.qsv_density_reset(self._h, _lib.int_array(part), len(part)))

    def reset(self, qargs=None):
        """densitymatrix.py:508-537 (deterministic: the reset channel, no sampling)."""
        ret = self._new_like()
        ret._reset_inplace(qargs)
        return ret

    def _initialize_inplace(self, obj, qargs):
        """``Initialize`` = reset + state preparation (its definition, reached through
        densitymatrix.py:781-797): reset channel, then |psi><0| on both sides."""
        params = obj.params
        k = obj.num_qubits
        if all(isinstance(p, str) for p in params):
            init = Statevector.from_label("".join(params), device=self._device).data
        elif len(params) == 1:
            init = np.zeros(1 << k, dtype=complex)
            init[int(np.real(params[0]))] = 1
        else:
            init = np.asarray(params, dtype=complex)
        qa = self._qargs(qargs)
        self._reset_inplace(qa)
        mat = np.zeros((1 << len(qa), 1 << len(qa)), dtype=complex)
        mat[:, 0] = init
        self._evolve_operator_inplace(mat, qa)

    # ------------------------------------------------- expectation values ---
    def _expval_batch(self, z_mask, x_mask, n_y) -> np.ndarray:
        t = len(z_mask)
        z_mask = np.ascontiguousarray(z_mask, dtype=np.uint64)
        x_mask = np.ascontiguousarray(x_mask, dtype=np.uint64)
        if self._n < 64 and np.any((z_mask | x_mask) >> np.uint64(self._n)):
            raise QiskitError("Pauli acts on qubits outside the density matrix")
        phases = np.ascontiguousarray((-1j) ** (np.asarray(n_y) % 4), dtype=np.complex128)
        out = np.empty(t, dtype=np.float64)
        _lib.check(_lib.load().qsv_density_expval_pauli_batch(
            self._h, t, z_mask.ctypes.data, x_mask.ctypes.data, phases.ctypes.data, out.ctypes.data))
        return out

    def _expectation_value_pauli(self, pauli, qargs=None):
        """densitymatrix.py:366-397."""
        z_mask, x_mask, n_y = pauli_masks(pauli.z, pauli.x, qargs)
        phase = getattr(pauli, "phase", 0)
        pauli_phase = (-1j) ** int(phase) if phase else 1
        if int(x_mask[0]) + int(z_mask[0]) == 0:
            return pauli_phase * self.trace()
        return pauli_phase * float(self._expval_batch(z_mask, x_mask, n_y)[0])

    def expectation_value(self, oper, qargs=None):
        """densitymatrix.py:399-420.  All terms of a ``SparsePauliOp`` go to the device in one call."""
        if hasattr(oper, "paulis") and hasattr(oper, "coeffs"):      # SparsePauliOp
            z = np.asarray(oper.paulis.z, dtype=bool)
            x = np.asarray(oper.paulis.x, dtype=bool)
            coeffs = np.asarray(oper.coeffs, dtype=complex)
            z_mask, x_mask, n_y = pauli_masks(z, x, qargs)
            vals = self._expval_batch(z_mask, x_mask, n_y).astype(complex)
            ident = (z_mask | x_mask) == 0
            if np.any(ident):                                        # identity terms: trace (:384-385)
                vals[ident] = self.trace()
            return complex(np.dot(coeffs, vals))
        if hasattr(oper, "z") and hasattr(oper, "x") and np.ndim(oper.z) == 1:   # Pauli
            return self._expectation_value_pauli(oper, qargs)
        # trace(rho . O) (:418-420) = trace(O . rho): O on the row qubits of a copy, then the trace
        mat = np.asarray(getattr(oper, "data", oper), dtype=np.complex128)
        qa = self._qargs(qargs)
        if mat.ndim != 2 or mat.shape != (1 << len(qa), 1 << len(qa)):
            raise QiskitError("Operator dimensions do not match the density matrix subsystems")
        tmp = self._new_like()
        stream = GateStream()
        stream.matrix(mat, [q + self._n for q in qa])
        tmp._flush(stream)
        return tmp.trace()

    # ------------------------------------------- probabilities / sampling ---
    def probabilities(self, qargs=None, decimals=None) -> np.ndarray:
        """densitymatrix.py:422-506; output bit j is qubit ``qargs[j]``."""
        qa = None if qargs is None else self._qargs(qargs)
        m = self._n if qa is None else len(qa)
        out = np.empty(1 << m, dtype=np.float64)
        _lib.check(_lib.load().qsv_density_probabilities(
            self._h, None if qa is None else _lib.int_array(qa), m, out.ctypes.data))
        if decimals is not None:
            out = out.round(decimals=decimals)
        return out

    def probabilities_dict(self, qargs=None, decimals=None) -> dict:
        """quantum_state.py:210-234."""
        probs = self.probabilities(qargs, decimals)
        m = self._n if qargs is None else len(qargs)
        (nz,) = np.nonzero(probs)
        return dict(zip(_ket_strings(nz, m).tolist(), probs[nz].tolist()))

    def sample_memory(self, shots: int, qargs=None) -> np.ndarray:
        """quantum_state.py:236-265: ``rng.choice(labels, p=probs, size=shots)`` -- the
        reference's own NumPy call on the (device-computed) <= 2^n probabilities."""
        probs = self.probabilities(qargs)
        m = self._n if qargs is None else len(qargs)
        labels = _ket_strings(np.arange(len(probs)), m)
        return self._rng.choice(labels, p=probs, size=shots)

    def sample_counts(self, shots: int, qargs=None) -> dict:
        """quantum_state.py:267-294."""
        inds, counts = np.unique(self.sample_memory(shots, qargs), return_counts=True)
        return {str(k): int(v) for k, v in zip(inds, counts)}

    def measure(self, qargs=None):
        """quantum_state.py:296-331 -> (outcome string, collapsed state)."""
        qa = self._qargs(qargs)
        probs = self.probabilities(qa)
        sample = self._rng.choice(len(probs), p=probs, size=1)
        outcome = str(_ket_strings(sample, len(qa))[0])
        # the projector |s><s| / sqrt(p) is a product over qubits: diagonals of <= 8 qubits each
        # (the reference builds the full 2^m diagonal, quantum_state.py:321-329), the scale on the first
        idx = int(sample[0])
        ret = self._new_like()
        stream = _DensityStream(self._n)
        for i in range(0, len(qa), 8):
            part = qa[i:i + 8]
            proj = np.zeros(1 << len(part), dtype=complex)
            proj[(idx >> i) & ((1 << len(part)) - 1)] = 1 / np.sqrt(probs[idx]) if i == 0 else 1.0
            stream.diagonal(proj, part)
        ret._flush(stream)
        return outcome, ret

    # ------------------------------------------------------- constructors ---
    @classmethod
    def from_int(cls, i: int, dims, *, device: int = 0):
        """densitymatrix.py:576-602."""
        size = int(np.prod(dims))
        n = int(round(np.log2(size)))
        if (1 << n) != size or not 0 <= int(i) < size:
            raise QiskitError("from_int needs a power-of-two dimension and an index below it")
        ret = object.__new__(cls)
        ret._rng_generator = None
        ret._device = int(device)
        ret._h = None
        ret._set_n(n)
        ret._h = _lib.acquire(2 * n, ret._device)
        _lib.check(_lib.load().qsv_init_basis(ret._h, (int(i) << n) | int(i)))
        return ret

    @classmethod
    def from_label(cls, label: str, *, device: int = 0):
        """densitymatrix.py:540-573."""
        return cls(Statevector.from_label(label, device=device))

    @classmethod
    def from_instruction(cls, instruction, *, device: int = 0):
        """densitymatrix.py:604-632."""
        return cls(instruction, device=device)

    def counters(self) -> dict:
        c = _lib.QsvCounters()
        _lib.check(_lib.load().qsv_get_counters(self._h, C.byref(c)))
        return {name: int(getattr(c, name)) for name, _ in c._fields_}
