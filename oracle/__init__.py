"""CPU oracle for the statevector hot path -- TEST INFRASTRUCTURE ONLY.

This package is a NumPy (plus one small C file) restatement of the reference's
algorithms for the path named in BASELINE.json (Qiskit 2.6.0.dev0,
``Statevector.evolve / expectation_value / probabilities / sample_memory`` and
the two statevector primitives).  It exists to *check* the CUDA engine in
``qiskit_b200``; nothing in the product imports it.

Who may import / execute it: ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``.  Nothing else.

Parity status: PINNED.  The restatement is checked against every golden value
the reference's own tests hold for this path (tests/golden/reference_kats.json,
tests/test_oracle_kats.py; sources: test/python/primitives/
test_statevector_estimator.py, test_statevector_sampler.py and
test/python/quantum_info/states/test_statevector.py).  The reference itself
cannot be imported or compiled in this image (``import qiskit`` needs the Rust
extension ``qiskit._accelerate``; there is no cargo/rustc), so there is no
``oracle/_ref``; NumPy -- the third-party code the reference calls for
``np.dot``, ``cumsum``, ``searchsorted`` and ``Generator`` -- is the same NumPy
the reference pins (``numpy>=2.0,<3``, requirements.txt:3).
"""
