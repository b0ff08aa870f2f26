"""Device-resident ``Statevector``: the reference's state class on a B200.

Same public behaviour as ``qiskit.quantum_info.Statevector`` for the simulation
path (qiskit/quantum_info/states/statevector.py and quantum_state.py) -- ``evolve``
(never mutates its receiver, :437-478), ``expectation_value`` (:564-587),
``probabilities`` (:589-673), ``sample_memory / sample_counts / measure / reset /
seed`` (quantum_state.py:78-331) -- but the amplitude array lives in HBM behind an
opaque handle of libqsv_b200.so and only crosses PCIe on explicit ``.data`` access.
Qubit states only (dims = (2,)*n); there is no CPU fallback.
"""
from __future__ import annotations

import copy as _copy
import ctypes as C
from numbers import Number

import numpy as np

from . import _lib
from .exceptions import QiskitError
from .lowering import GateStream, walk
from .pauli import pauli_masks


def _is_circuit(obj) -> bool:
    return hasattr(obj, "to_instruction") and hasattr(obj, "num_qubits")


def _is_instruction(obj) -> bool:
    return hasattr(obj, "definition") and hasattr(obj, "num_qubits") and hasattr(obj, "params")


class Statevector:
    """Statevector on the GPU.  ``Statevector(data)`` accepts an array of 2^n
    amplitudes, a circuit / instruction (simulated from |0...0>), or another state."""

    atol = 1e-8   # operators/mixins/tolerances.py:26-28
    rtol = 1e-5

    def __init__(self, data, dims=None, *, device: int = 0):
        self._rng_generator = None
        self._device = int(device)
        self._h = None
        if isinstance(data, Statevector):
            self._n = data._n
            self._h = data._clone_handle()
            self._device = data._device
        elif _is_circuit(data) or _is_instruction(data):
            inst = data.to_instruction() if _is_circuit(data) else data
            self._n = int(inst.num_qubits)
            self._h = _lib.acquire(self._n, self._device)
            _lib.check(_lib.load().qsv_init_basis(self._h, 0))
            self._evolve_instruction_inplace(inst, None)
        else:
            vec = np.asarray(data, dtype=np.complex128)
            if vec.ndim == 2 and vec.shape[1] == 1:
                vec = vec.reshape(-1)
            if vec.ndim != 1:
                raise QiskitError("Invalid input: not a vector or column-vector.")
            n = int(round(np.log2(vec.size))) if vec.size else -1
            if vec.size == 0 or (1 << n) != vec.size:
                raise QiskitError("Only qubit statevectors (length 2^n) are supported on the device")
            if dims is not None:
                d = (dims,) if isinstance(dims, (int, np.integer)) else tuple(dims)
                if int(np.prod(d)) != vec.size or any(x != 2 for x in d) and len(d) != 1:
                    raise QiskitError("Only qubit subsystem dimensions are supported on the device")
            self._n = n
            self._h = _lib.acquire(n, self._device)
            self._upload(vec)

    # ------------------------------------------------------------ plumbing --
    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None:
            try:
                if getattr(self, "_sharded", None):      # sharded handles own an NCCL communicator
                    _lib.load().qsv_destroy(h)
                else:
                    _lib.release(h, self._n, self._device)
            except Exception:  # noqa: BLE001 - interpreter shutdown
                pass

    def _clone_handle(self):
        """Device-to-device copy into a (pooled) handle."""
        h = _lib.acquire(self._n, self._device)
        _lib.check(_lib.load().qsv_copy_state(h, self._h))
        return h

    def _upload(self, vec: np.ndarray):
        vec = _lib.as_c128(vec)
        _lib.check(_lib.load().qsv_upload(self._h, vec.ctypes.data, 0, vec.size))

    def _flush(self, stream: GateStream, flags: int = 0):
        if len(stream):
            arr, n = stream.to_c()
            _lib.check(_lib.load().qsv_apply_gates(self._h, arr, n, flags))
            stream.clear()

    def _new_like(self):
        ret = object.__new__(type(self))
        ret._rng_generator = self._rng_generator
        ret._device = self._device
        ret._n = self._n
        ret._h = self._clone_handle()
        return ret

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
        """Amplitudes as a host array (device -> host copy of 16 * 2^n bytes)."""
        out = np.empty(1 << self._n, dtype=np.complex128)
        _lib.check(_lib.load().qsv_download(self._h, out.ctypes.data, 0, out.size))
        return out

    def __array__(self, dtype=None, copy=None):
        arr = self.data
        return arr if dtype is None else arr.astype(dtype)

    def __len__(self):
        return 1 << self._n

    def __getitem__(self, key):
        if isinstance(key, str):
            key = int(key, 2)
        if isinstance(key, (int, np.integer)):
            if not 0 <= key < self.dim:
                raise QiskitError(f"Key {key} is greater than Statevector dimension {self.dim}.")
            out = np.empty(1, dtype=np.complex128)
            _lib.check(_lib.load().qsv_download(self._h, out.ctypes.data, int(key), 1))
            return out[0]
        raise QiskitError("Key must be int or a valid binary string.")

    def __repr__(self):
        return f"Statevector(<device {self._device}, {self._n} qubits>, dims={self.dims()})"

    def __eq__(self, other):
        if not isinstance(other, Statevector) or other._n != self._n:
            return False
        return bool(np.allclose(self.data, other.data, rtol=self.rtol, atol=self.atol))

    __hash__ = None

    def copy(self):
        """``QuantumState.copy`` is a deep copy, RNG included (quantum_state.py:74-76)."""
        return _copy.deepcopy(self)

    def __copy__(self):
        return self._new_like()

    def __deepcopy__(self, memo):
        ret = self._new_like()
        ret._rng_generator = _copy.deepcopy(self._rng_generator, memo)
        return ret

    def seed(self, value=None):
        """quantum_state.py:78-85."""
        if value is None:
            self._rng_generator = None
        elif isinstance(value, np.random.Generator):
            self._rng_generator = value
        else:
            self._rng_generator = np.random.default_rng(value)

    @property
    def _rng(self):
        if self._rng_generator is None:
            return np.random.default_rng()
        return self._rng_generator

    # ----------------------------------------------------------- factories --
    @classmethod
    def from_int(cls, i: int, dims, *, device: int = 0):
        """statevector.py:790-816 for qubit dims: ``dims`` is 2^n or a tuple of 2s."""
        size = int(np.prod(dims))
        n = int(round(np.log2(size)))
        if (1 << n) != size:
            raise QiskitError("Only qubit dimensions are supported on the device")
        ret = object.__new__(cls)
        ret._rng_generator = None
        ret._device = int(device)
        ret._n = n
        ret._h = _lib.acquire(n, device)
        _lib.check(_lib.load().qsv_init_basis(ret._h, int(i)))
        return ret

    @classmethod
    def from_label(cls, label: str, *, device: int = 0):
        """statevector.py:721-788: product of 0/1 and +,-,r,l single-qubit states."""
        if any(ch not in "01+-rl" for ch in label):
            raise QiskitError("Label contains invalid characters.")
        n = len(label)
        z_label = "".join("1" if ch == "1" else "0" for ch in label)
        ret = cls.from_int(int(z_label, 2) if n else 0, 1 << n, device=device)
        stream = GateStream()
        s2 = np.sqrt(0.5)
        x_mat = np.array([[1, 1], [1, -1]]) * s2
        y_mat = np.array([[1, 1], [1j, -1j]]) * s2
        for q, ch in enumerate(reversed(label)):
            if ch == "+":
                stream.matrix(x_mat, [q])
            elif ch == "-":
                stream.matrix(x_mat @ np.array([[0, 1], [1, 0]]), [q])
            elif ch == "r":
                stream.matrix(y_mat, [q])
            elif ch == "l":
                stream.matrix(y_mat @ np.array([[0, 1], [1, 0]]), [q])
        ret._flush(stream)
        return ret

    @classmethod
    def from_instruction(cls, instruction, *, device: int = 0):
        """statevector.py:818-843."""
        return cls(instruction, device=device)

    @classmethod
    def from_circuit(cls, circuit, ignore_set_layout: bool = False, *, device: int = 0):
        """statevector.py:115-180: simulate, then undo the transpiler's final layout: new qubit k
        takes old qubit ``final_index_layout()[k]`` (the reference's tensor transpose with
        ``reversed_perm``), as one bit-permutation kernel."""
        layout = None if ignore_set_layout else getattr(circuit, "layout", None)
        sv = cls.from_instruction(circuit, device=device)
        if layout is None:
            return sv
        idx = [int(v) for v in layout.final_index_layout()]
        return sv._permuted(idx)

    # ------------------------------------------------------------- algebra --
    def is_valid(self, atol=None, rtol=None):
        atol = self.atol if atol is None else atol
        rtol = self.rtol if rtol is None else rtol
        return bool(np.allclose(np.sqrt(self._norm2()), 1, rtol=rtol, atol=atol))

    def _norm2(self) -> float:
        out = C.c_double()
        _lib.check(_lib.load().qsv_norm2(self._h, C.byref(out)))
        return out.value

    def trace(self):
        return self._norm2()

    def purity(self):
        return self._norm2() ** 2

    def conjugate(self):
        """statevector.py:327-329, element-wise on the device."""
        ret = self._new_like()
        _lib.check(_lib.load().qsv_conjugate(ret._h))
        return ret

    def _kron(self, hi: "Statevector", lo: "Statevector"):
        """np.kron(hi, lo) on the device: ``lo`` takes the low qubits."""
        n = hi._n + lo._n
        ret = object.__new__(type(self))
        ret._rng_generator = self._rng_generator
        ret._device = self._device
        ret._n = n
        ret._h = _lib.acquire(n, self._device)
        _lib.check(_lib.load().qsv_tensor(ret._h, hi._h, lo._h, _lib.int_array(range(lo._n, n)),
                                          _lib.int_array(range(lo._n)), 0))
        return ret

    def tensor(self, other):
        """statevector.py:339-360: self (x) other."""
        if not isinstance(other, Statevector):
            other = Statevector(other, device=self._device)
        return self._kron(self, other)

    def expand(self, other):
        """statevector.py:378-397: other (x) self."""
        if not isinstance(other, Statevector):
            other = Statevector(other, device=self._device)
        return self._kron(other, self)

    def _permuted(self, src_bit):
        """New state with amplitude i taken from index j, bit src_bit[k] of j = bit k of i."""
        ret = object.__new__(type(self))
        ret._rng_generator = self._rng_generator
        ret._device = self._device
        ret._n = self._n
        ret._h = _lib.acquire(self._n, self._device)
        _lib.check(_lib.load().qsv_permute_qubits(ret._h, self._h, _lib.int_array(src_bit)))
        return ret

    def reverse_qargs(self):
        """statevector.py:512-531: qubit q <-> n-1-q."""
        return self._permuted([self._n - 1 - k for k in range(self._n)])

    def inner(self, other) -> complex:
        """<self|other> (statevector.py:399-413)."""
        if not isinstance(other, Statevector):
            other = Statevector(other, device=self._device)
        if other._n != self._n:
            raise QiskitError("Statevector dimensions do not match")
        out = (C.c_double * 2)()
        _lib.check(_lib.load().qsv_inner(self._h, other._h, out))
        return complex(out[0], out[1])

    def _multiply(self, other):
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
        if not isinstance(other, Statevector) or other._n != self._n:
            raise QiskitError("other is not a Statevector of the same size")
        ret = self._new_like()
        _lib.check(_lib.load().qsv_axpy(ret._h, other._h, 1.0, 0.0))
        return ret

    __add__ = _add

    def __sub__(self, other):
        if not isinstance(other, Statevector) or other._n != self._n:
            raise QiskitError("other is not a Statevector of the same size")
        ret = self._new_like()
        _lib.check(_lib.load().qsv_axpy(ret._h, other._h, -1.0, 0.0))
        return ret

    def equiv(self, other, rtol=None, atol=None) -> bool:
        """Equal up to a global phase (statevector.py:480-510)."""
        if not isinstance(other, Statevector):
            try:
                other = Statevector(other, device=self._device)
            except QiskitError:
                return False
        if other._n != self._n:
            return False
        rtol = self.rtol if rtol is None else rtol
        atol = self.atol if atol is None else atol
        a, b = self.data, other.data
        for v in a:
            if abs(v) > atol:
                a = a * np.exp(-1j * np.angle(v))
                break
        for v in b:
            if abs(v) > atol:
                b = b * np.exp(-1j * np.angle(v))
                break
        return bool(np.allclose(a, b, rtol=rtol, atol=atol))

    # ------------------------------------------------------------- evolve ---
    def evolve(self, other, qargs=None):
        """statevector.py:437-478: returns a new state; ``self`` is untouched."""
        if qargs is None:
            qargs = getattr(other, "qargs", None)
        ret = self._new_like()
        if _is_circuit(other):
            other = other.to_instruction()
        if _is_instruction(other):
            ret._evolve_instruction_inplace(other, qargs)
            return ret
        mat = getattr(other, "data", other)      # Operator-like or raw matrix
        mat = np.asarray(mat, dtype=np.complex128)
        ret._evolve_operator_inplace(mat, qargs)
        return ret

    def _evolve_operator_inplace(self, mat, qargs=None):
        """Device counterpart of ``_evolve_operator`` (statevector.py:925-963)."""
        k = self._n if qargs is None else len(qargs)
        if mat.ndim != 2 or mat.shape[0] != mat.shape[1] or mat.shape[0] != (1 << k):
            raise QiskitError(
                "Operator input dimensions are not equal to statevector subsystem dimensions."
            )
        qa = list(range(self._n)) if qargs is None else [int(q) for q in qargs]
        stream = GateStream()
        stream.matrix(mat, qa)
        self._flush(stream)

    def _evolve_instruction_inplace(self, inst, qargs=None):
        """Device counterpart of ``_evolve_instruction`` (statevector.py:965-1041)."""
        if qargs is not None and len(qargs) != inst.num_qubits:
            raise QiskitError("qargs length does not match the instruction's number of qubits")
        if qargs is None and inst.num_qubits != self._n:
            raise QiskitError(
                f"Cannot apply a {inst.num_qubits}-qubit instruction to a {self._n}-qubit "
                "Statevector without qargs"
            )
        stream = GateStream()
        walk(self, inst, None if qargs is None else [int(q) for q in qargs], stream)
        self._flush(stream)

    def apply_gate_list(self, gates, *, no_fusion: bool = False):
        """In-place, Qiskit-free entry used by benchmarks: ``gates`` as in
        ``lowering.gate_list_to_stream``."""
        from .lowering import gate_list_to_stream

        self._flush(gate_list_to_stream(gates), _lib.FLAG_NO_FUSION if no_fusion else 0)
        return self

    # ------------------------------------------------- expectation values ---
    def _expectation_value_pauli(self, pauli, qargs=None):
        """statevector.py:533-562."""
        z_mask, x_mask, n_y = pauli_masks(pauli.z, pauli.x, qargs)
        self._check_masks(z_mask | x_mask)
        phase = getattr(pauli, "phase", 0)
        pauli_phase = (-1j) ** int(phase) if phase else 1
        y_phase = (-1j) ** int(n_y[0])
        out = C.c_double()
        x_max = int(x_mask[0]).bit_length() - 1
        _lib.check(_lib.load().qsv_expval_pauli(self._h, int(z_mask[0]), int(x_mask[0]),
                                                y_phase.real, y_phase.imag, x_max, C.byref(out)))
        return pauli_phase * out.value

    def _check_masks(self, masks):
        if self._n < 64 and np.any(masks >> np.uint64(self._n)):
            raise QiskitError("Pauli acts on qubits outside the statevector")

    def expectation_value(self, oper, qargs=None):
        """statevector.py:564-587.  A ``SparsePauliOp`` is evaluated in ONE engine call
        (all terms batched) instead of one pass per term."""
        if hasattr(oper, "paulis") and hasattr(oper, "coeffs"):      # SparsePauliOp
            z = np.asarray(oper.paulis.z, dtype=bool)
            x = np.asarray(oper.paulis.x, dtype=bool)
            coeffs = np.asarray(oper.coeffs, dtype=complex)
            z_mask, x_mask, n_y = pauli_masks(z, x, qargs)
            self._check_masks(z_mask | x_mask)
            vals = self._expval_batch(z_mask, x_mask, n_y)
            return complex(np.dot(coeffs, vals))
        if hasattr(oper, "z") and hasattr(oper, "x") and np.ndim(oper.z) == 1:   # Pauli
            return self._expectation_value_pauli(oper, qargs)
        val = self.evolve(oper, qargs=qargs)
        return self.inner(val)

    def _expval_batch(self, z_mask, x_mask, n_y) -> np.ndarray:
        t = len(z_mask)
        z_mask = np.ascontiguousarray(z_mask, dtype=np.uint64)
        x_mask = np.ascontiguousarray(x_mask, dtype=np.uint64)
        phases = np.ascontiguousarray((-1j) ** (np.asarray(n_y) % 4), dtype=np.complex128)
        out = np.empty(t, dtype=np.float64)
        _lib.check(_lib.load().qsv_expval_pauli_batch(
            self._h, t, z_mask.ctypes.data, x_mask.ctypes.data, phases.ctypes.data, out.ctypes.data))
        return out

    # ------------------------------------------------------- probabilities --
    def probabilities(self, qargs=None, decimals=None) -> np.ndarray:
        """statevector.py:589-673; output bit j is qubit ``qargs[j]``."""
        m = self._n if qargs is None else len(qargs)
        out = np.empty(1 << m, dtype=np.float64)
        qa = None if qargs is None else _lib.int_array(qargs)
        _lib.check(_lib.load().qsv_probabilities(self._h, qa, m, out.ctypes.data))
        if decimals is not None:
            out = out.round(decimals=decimals)
        return out

    def probabilities_dict(self, qargs=None, decimals=None) -> dict:
        """quantum_state.py:210-234."""
        probs = self.probabilities(qargs, decimals)
        (inds,) = probs.nonzero()
        m = self._n if qargs is None else len(qargs)
        return dict(zip(_ket_strings(inds, m).tolist(), probs[inds].tolist()))

    def to_dict(self, decimals=None) -> dict:
        vec = self.data
        vals = vec if decimals is None else vec.round(decimals=decimals)
        (inds,) = vals.nonzero()
        return dict(zip(_ket_strings(inds, self._n).tolist(), vals[inds].tolist()))

    # ------------------------------------------------------------ sampling --
    def sample_indices(self, shots: int, qargs=None, uniforms=None) -> np.ndarray:
        """Outcome indices (bit j = qubit qargs[j]) for ``shots`` uniforms drawn from the
        state's RNG exactly as ``Generator.choice(p=...)`` draws them
        (quantum_state.py:258-265)."""
        if uniforms is None:
            uniforms = self._rng.random(int(shots))
        uniforms = np.ascontiguousarray(uniforms, dtype=np.float64)
        m = self._n if qargs is None else len(qargs)
        out = np.empty(uniforms.size, dtype=np.uint64)
        total = C.c_double()
        qa = None if qargs is None else _lib.int_array(qargs)
        _lib.check(_lib.load().qsv_sample(self._h, qa, m, uniforms.ctypes.data, uniforms.size,
                                          out.ctypes.data, C.byref(total)))
        if abs(total.value - 1.0) > np.sqrt(np.finfo(np.float64).eps):
            raise ValueError("probabilities do not sum to 1")  # Generator.choice's check
        return out

    def sample_memory(self, shots: int, qargs=None) -> np.ndarray:
        """quantum_state.py:236-265: array of ket strings, qubit 0 rightmost."""
        m = self._n if qargs is None else len(qargs)
        return _ket_strings(self.sample_indices(shots, qargs), m)

    def sample_counts(self, shots: int, qargs=None) -> dict:
        """quantum_state.py:267-294."""
        m = self._n if qargs is None else len(qargs)
        inds, counts = np.unique(self.sample_indices(shots, qargs), return_counts=True)
        return dict(zip(_ket_strings(inds, m).tolist(), counts.tolist()))

    def _sample_one(self, qargs):
        """One outcome and its (clipped) probability, consuming one uniform like
        ``rng.choice(len(probs), p=probs, size=1)`` (statevector.py:705)."""
        idx = int(self.sample_indices(1, qargs)[0])
        prob = C.c_double()
        _lib.check(_lib.load().qsv_outcome_probability(self._h, _lib.int_array(qargs), len(qargs), idx,
                                                       C.byref(prob)))
        return idx, prob.value

    def measure(self, qargs=None):
        """quantum_state.py:296-331 -> (outcome string, collapsed state)."""
        qa = list(range(self._n)) if qargs is None else [int(q) for q in qargs]
        idx, prob = self._sample_one(qa)
        ret = self._new_like()
        _lib.check(_lib.load().qsv_collapse(ret._h, _lib.int_array(qa), len(qa), idx,
                                            1.0 / np.sqrt(prob), 0))
        return str(_ket_strings(np.array([idx]), len(qa))[0]), ret

    def reset(self, qargs=None):
        """statevector.py:675-719."""
        ret = self._new_like()
        ret._reset_inplace(qargs)
        return ret

    def _reset_inplace(self, qargs):
        lib = _lib.load()
        if qargs is None:
            _lib.check(lib.qsv_init_basis(self._h, 0))
            return
        qa = [int(q) for q in qargs]
        idx, prob = self._sample_one(qa)
        _lib.check(lib.qsv_collapse(self._h, _lib.int_array(qa), len(qa), idx, 1.0 / np.sqrt(prob), 1))

    def _initialize_inplace(self, obj, qargs):
        """Initialize special case of ``_evolve_instruction`` (statevector.py:989-1014)."""
        params = obj.params
        k = obj.num_qubits
        if all(isinstance(p, str) for p in params):
            init = Statevector.from_label("".join(params), device=self._device).data
        elif len(params) == 1:
            init = np.zeros(1 << k, dtype=complex)
            init[int(np.real(params[0]))] = 1
        else:
            init = np.asarray(params, dtype=complex)
        if qargs is None:
            self._upload(init)
            return
        self._reset_inplace(qargs)
        mat = np.zeros((1 << len(qargs), 1 << len(qargs)), dtype=complex)
        mat[:, 0] = init
        self._evolve_operator_inplace(mat, qargs)

    # ------------------------------------------------------------ counters --
    def counters(self) -> dict:
        c = _lib.QsvCounters()
        _lib.check(_lib.load().qsv_get_counters(self._h, C.byref(c)))
        return {name: int(getattr(c, name)) for name, _ in c._fields_}

    def set_option(self, name: str, value: int):
        _lib.check(_lib.load().qsv_set_option(self._h, name.encode(), int(value)))


def _ket_strings(inds: np.ndarray, m: int) -> np.ndarray:
    """``_index_to_ket_array(..., string_labels=True)`` for qubits
    (quantum_state.py:333-364): qubit 0 is the rightmost character."""
    inds = np.asarray(inds, dtype=np.uint64)
    if m == 0:
        return np.array([""] * len(inds), dtype=np.str_)
    shifts = np.arange(m - 1, -1, -1, dtype=np.uint64)
    bits = ((inds[:, None] >> shifts[None, :]) & np.uint64(1)).astype(np.uint8) + ord("0")
    return np.ascontiguousarray(bits).view(f"S{m}").reshape(-1).astype(np.str_)
