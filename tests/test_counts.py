"""Counts post-processing: the oracle against the reference's golden values
(test/python/result/test_sampled_expval.py) on CPU, the device kernels against the oracle on GPU."""
import numpy as np
import pytest

from oracle import counts_ref as cref

PROBS = {  # test_sampled_expval.py:21-38
    "1000": 0.0022, "1001": 0.0045, "1110": 0.0081, "0001": 0.0036, "0010": 0.0319, "0101": 0.001,
    "1100": 0.0008, "1010": 0.0009, "1111": 0.3951, "0011": 0.0007, "0111": 0.01, "0000": 0.4666,
    "1101": 0.0355, "1011": 0.0211, "0110": 0.0081, "0100": 0.0099,
}
COUNTS = {"001": 67, "110": 113, "100": 83, "011": 205, "111": 4535, "101": 100, "010": 42, "000": 4855}
GOLDEN = [  # (dist, operator, expected, reference line)
    ({"00": 0.5, "11": 0.5}, "ZZ", 1.0, 50), ({"000": 0.5, "111": 0.5}, "ZZZ", 0.0, 52),
    ({"000": 0.5, "111": 0.5}, "III", 1.0, 54), ({"00": 0.5, "11": 0.5}, "IZ", 0.0, 56),
    ({"00": 0.5, "11": 0.5}, "ZI", 0.0, 57), (PROBS, "ZZZZ", 0.7554, 59), (COUNTS, "IZZ", 0.9356, 63),
    (PROBS, "0III", 0.5318, 101), (PROBS, "III0", 0.5285, 102), (PROBS, "1011", 0.0211, 103),
    (PROBS, "IZIZ", 0.8864, 109), (PROBS, "00ZI", 0.4376, 112),
]


@pytest.mark.parametrize("dist,oper,expected,line", GOLDEN, ids=[f"{g[1]}@{g[3]}" for g in GOLDEN])
def test_oracle_golden(dist, oper, expected, line):
    assert abs(cref.bitstring_expval(dist, oper) - expected) < 1e-7   # assertAlmostEqual (7 places)


def test_oracle_helpers():
    assert cref.counts_from_ints([3, 0, 3, 384], 9) == {"000000000": 1, "000000011": 2, "110000000": 1}
    np.testing.assert_array_equal(cref.marginalize_ints([0b1011, 0b0100], [3, 0, -1, 2]), [0b0011, 0b1000])
    assert abs(cref.sampled_expval(["ZZZZZ"], [1j], {"11111": 1}) - (-1j)) < 1e-15  # :114-123


# ------------------------------------------------------------------ device ---
@pytest.mark.gpu
@pytest.mark.parametrize("dist,oper,expected,line", GOLDEN, ids=[f"{g[1]}@{g[3]}" for g in GOLDEN])
def test_device_golden(dist, oper, expected, line):
    from qiskit_b200 import Pauli, SparsePauliOp
    from qiskit_b200.sampled_expval import sampled_expectation_value

    got = sampled_expectation_value(dist, oper)
    assert abs(got - expected) < 1e-7
    assert abs(got - cref.bitstring_expval(dist, oper)) < 1e-15
    if set(oper) <= {"I", "Z"}:
        assert abs(sampled_expectation_value(dist, Pauli(oper)) - got) < 1e-15
        assert abs(sampled_expectation_value(dist, SparsePauliOp([oper], [1])) - got) < 1e-15


@pytest.mark.gpu
def test_device_errors_and_complex():
    from qiskit_b200 import QiskitError, SparsePauliOp
    from qiskit_b200.sampled_expval import sampled_expectation_value

    with pytest.raises(QiskitError, match="not diagonal"):
        sampled_expectation_value(COUNTS, "IXZ")
    with pytest.raises(QiskitError, match="not same length"):
        sampled_expectation_value(COUNTS, "ZZ")
    with pytest.raises(QiskitError, match="Invalid input distribution"):
        sampled_expectation_value([1, 2], "ZZ")
    got = sampled_expectation_value({"11111": 1}, SparsePauliOp(["ZZZZZ"], [1j]))
    assert np.iscomplexobj(got) and abs(got - (-1j)) < 1e-15
    real = sampled_expectation_value({"11111": 1}, SparsePauliOp(["ZZZZZ"], [1]))  # np.real_if_close (:89)
    assert np.isrealobj(real) and real == -1.0
    assert isinstance(sampled_expectation_value({"11111": 1}, "ZZZZZ"), float)


@pytest.mark.gpu
@pytest.mark.parametrize("num_bits,shots,terms", [(9, 1000, 5), (30, 200_000, 37), (63, 50_000, 11)])
def test_device_random_samples_bit_exact(num_bits, shots, terms):
    """Integer counts: every partial sum is an integer below 2^53, so any summation order gives the
    reference's value exactly."""
    from qiskit_b200.sampled_expval import _device_expvals, marginalize_samples, sample_counts

    rng = np.random.default_rng(num_bits)
    pool = rng.integers(0, 1 << min(num_bits, 62), size=max(8, shots // 50), dtype=np.uint64)
    samples = pool[rng.integers(0, pool.size, size=shots)]
    labels = ["".join(rng.choice(list("IIZ01Z"), size=num_bits)) for _ in range(terms)]
    dist = cref.counts_from_ints(samples, num_bits)
    got = _device_expvals(samples, None, labels)
    want = np.array([cref.bitstring_expval(dist, s) for s in labels])
    np.testing.assert_array_equal(got, want)
    keys = np.array([int(k, 2) for k in dist], dtype=np.uint64)
    weights = np.array(list(dist.values()), dtype=float)
    np.testing.assert_array_equal(_device_expvals(keys, weights, labels), want)
    # histogram == np.unique
    k, c = sample_counts(samples)
    vals, cnt = np.unique(samples, return_counts=True)
    np.testing.assert_array_equal(k, vals)
    np.testing.assert_array_equal(c, cnt.astype(np.uint64))
    # marginalisation
    src = [int(b) for b in rng.permutation(num_bits)[: max(1, num_bits // 3)]] + [-1]
    np.testing.assert_array_equal(marginalize_samples(samples, src), cref.marginalize_ints(samples, src))


@pytest.mark.gpu
def test_bitarray_expectation_values_reference_case():
    # test_bit_array.py:693-738: incrementing bit-strings 0..59 in a (1, 2, 3) array of 10 shots
    from qiskit_b200 import BitArray, Pauli, SparsePauliOp

    data = np.frombuffer(np.arange(60, dtype=np.uint16).tobytes(), dtype=np.uint8)
    data = data.reshape(1, 2, 3, 10, 2)[..., ::-1]
    ba = BitArray(data, 9)
    assert ba.shape == (1, 2, 3)
    op, op2, op3 = "I" * 8 + "Z", "I" * 8 + "0", "I" * 8 + "1"
    np.testing.assert_allclose(ba.expectation_values(op), np.zeros(ba.shape))
    np.testing.assert_allclose(ba.expectation_values(op2), np.full(ba.shape, 0.5))
    np.testing.assert_allclose(ba.expectation_values(op3), np.full(ba.shape, 0.5))
    np.testing.assert_allclose(ba.slice_bits(6).expectation_values("Z"), np.ones(ba.shape))
    np.testing.assert_allclose(ba.slice_bits(5).expectation_values("0"), np.array([[[1, 1, 1], [0.2, 0, 0]]]))
    np.testing.assert_allclose(ba.expectation_values(Pauli(op)), np.zeros(ba.shape))
    np.testing.assert_allclose(ba.expectation_values(SparsePauliOp([op], [1])), np.zeros(ba.shape))
    # broadcasting: a (2, 1) list of observables against shape (1, 2, 3) -> (1, 2, 3)
    obs = [[op2], [op3]]
    got = ba.expectation_values(obs)
    assert got.shape == (1, 2, 3)
    with pytest.raises(ValueError):
        ba.expectation_values("Z")
    with pytest.raises(ValueError):
        ba.expectation_values("I" * 8 + "X")
    assert ba.slice_bits([0, 1]).num_bits == 2
