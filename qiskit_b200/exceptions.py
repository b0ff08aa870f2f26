"""Error types.  When real Qiskit is importable its ``QiskitError`` is re-used so
``except QiskitError`` in user code keeps working (qiskit/exceptions.py)."""
try:  # pragma: no cover - qiskit is absent from the build image
    from qiskit.exceptions import QiskitError  # type: ignore
except Exception:  # noqa: BLE001

    class QiskitError(Exception):
        """Stand-in for ``qiskit.exceptions.QiskitError``."""

        def __init__(self, *message):
            super().__init__(" ".join(str(m) for m in message))
            self.message = " ".join(str(m) for m in message)

        def __str__(self):
            return repr(self.message)
