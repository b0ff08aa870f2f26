"""qiskit_b200 -- a B200 (sm_100a) statevector engine behind Qiskit's simulation API.

One hot path of Qiskit, re-built for the GPU: ``Statevector.evolve /
expectation_value / probabilities / sample_memory``, ``StatevectorEstimator`` and
``StatevectorSampler`` with ``SparsePauliOp`` observables.  Host code is plain
Python calling ``libqsv_b200.so`` (hand-written CUDA, C ABI in include/qsv.h)
through ctypes; there is no PyTorch on the path and no CPU fallback.
"""
from .exceptions import QiskitError
from .circuit import (ClassicalRegister, Gate, Instruction, Parameter, ParameterVector,
                      QuantumCircuit)
from .pauli import Pauli, PauliList, SparsePauliOp
from .sparse_observable import SparseObservable
from .statevector import Statevector
from .densitymatrix import DensityMatrix
from .primitives import BitArray, StatevectorEstimator, StatevectorSampler
from .basic_simulator import BasicSimulator
from .sampled_expval import sampled_expectation_value
from .backend_primitives import BackendEstimatorV2, BackendSamplerV2

__all__ = [
    "QiskitError", "QuantumCircuit", "ClassicalRegister", "Parameter", "ParameterVector", "Gate",
    "Instruction", "Pauli", "PauliList", "SparsePauliOp", "SparseObservable", "Statevector", "DensityMatrix", "StatevectorEstimator",
    "StatevectorSampler", "BitArray", "BasicSimulator", "BackendEstimatorV2", "BackendSamplerV2", "sampled_expectation_value", "install",
]
__version__ = "0.1.0"


def install():
    """Opt-in: rebind Qiskit's statevector entry points to the device classes (SURVEY 8b B2 iii):
    ``qiskit.quantum_info.Statevector`` -> ``DeviceStatevector(Statevector)``,
    ``qiskit.primitives.StatevectorEstimator / StatevectorSampler`` -> ``DeviceEstimator(BaseEstimatorV2)`` /
    ``DeviceSampler(BaseSamplerV2)`` (qiskit_b200/adapter.py).  Needs a working ``import qiskit``;
    raises ImportError otherwise."""
    from . import adapter                      # noqa: PLC0415 - optional dependency inside

    return adapter.install()
