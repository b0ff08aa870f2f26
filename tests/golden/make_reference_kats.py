"""Writes tests/golden/reference_kats.json.

The reference (Qiskit 2.6.0.dev0) cannot be imported in this image (its
``qiskit._accelerate`` Rust extension cannot be built: no cargo/rustc), so the
golden values below are the constants the reference's OWN test-suite asserts for
this path, transcribed with their file:line (paths relative to /root/reference).
Circuits are given as flat gate lists ``[name, params, qargs]`` -- what
``Statevector._evolve_instruction`` reaches after unrolling the definitions.

Run:  python tests/golden/make_reference_kats.py
"""
import json
import os

EST = "test/python/primitives/test_statevector_estimator.py"
SMP = "test/python/primitives/test_statevector_sampler.py"
SV = "test/python/quantum_info/states/test_statevector.py"


def real_amplitudes_2q(reps, theta):
    """real_amplitudes(2, reps): RY layer then (CX(0,1), RY layer) * reps
    (qiskit/circuit/library/n_local/real_amplitudes.py; reverse_linear
    entanglement on 2 qubits is the single CX(0,1))."""
    g = [["ry", [theta[0]], [0]], ["ry", [theta[1]], [1]]]
    for r in range(reps):
        g.append(["cx", [], [0, 1]])
        g.append(["ry", [theta[2 * (r + 1)]], [0]])
        g.append(["ry", [theta[2 * (r + 1) + 1]], [1]])
    return g


H1 = [["II", 1], ["IZ", 2], ["XI", 3]]
H2 = [["IZ", 1]]
H3 = [["ZI", 1], ["ZZ", 1]]
TH1 = [0, 1, 1, 2, 3, 5]
TH2 = [0, 1, 1, 2, 3, 5, 8, 13]
TH3 = [1, 2, 3, 4, 5, 6]
H2MOL = [
    ["II", -1.052373245772859],
    ["IZ", 0.39793742484318045],
    ["ZI", -0.39793742484318045],
    ["ZZ", -0.01128010425623538],
    ["XX", 0.18093119978423156],
]

kats = {
    "estimator": [
        {"src": f"{EST}:77", "n": 2, "gates": real_amplitudes_2q(2, TH1), "obs": H1,
         "expected": 1.5555572817900956, "rtol": 1e-7},
        {"src": f"{EST}:85", "n": 2, "gates": real_amplitudes_2q(3, TH2), "obs": H1,
         "expected": 2.97797666, "rtol": 1e-7},
        {"src": f"{EST}:89", "n": 2, "gates": real_amplitudes_2q(2, TH1), "obs": H2,
         "expected": -0.551653, "rtol": 1e-6},
        {"src": f"{EST}:89", "n": 2, "gates": real_amplitudes_2q(2, TH1), "obs": H3,
         "expected": 0.07535239, "rtol": 1e-7},
        {"src": f"{EST}:97", "n": 2, "gates": real_amplitudes_2q(2, TH3), "obs": H3,
         "expected": -1.08766318, "rtol": 1e-7},
        {"src": f"{EST}:98", "n": 2, "gates": real_amplitudes_2q(3, TH2), "obs": H2,
         "expected": 0.17849238, "rtol": 1e-7},
        {"src": f"{EST}:123", "n": 2, "gates": real_amplitudes_2q(2, TH1), "obs": H2MOL,
         "expected": -1.284366511861733, "rtol": 1e-7},
        # endianness (EST:196-223): X on qubit 0 of 2
        {"src": f"{EST}:196-223", "n": 2, "gates": [["x", [], [0]]], "obs": [["IZ", 1]],
         "expected": -1.0, "rtol": 1e-12},
        {"src": f"{EST}:196-223", "n": 2, "gates": [["x", [], [0]]], "obs": [["ZI", 1]],
         "expected": 1.0, "rtol": 1e-12},
    ],
    # seed=1, precision=1.0 -> default_rng(1).normal(ev, 1.0)   (EST:273-286)
    "estimator_noise": {"src": f"{EST}:281", "seed": 1, "precision": 1.0, "n": 2,
                        "gates": real_amplitudes_2q(2, TH1), "obs": H1,
                        "expected": 1.901141473854881},
    "expval_state": [
        # Bell state table (SV:1207-1221)
        {"src": f"{SV}:1207-1221", "state_re": [0.7071067811865476, 0, 0, 0.7071067811865476],
         "state_im": [0, 0, 0, 0],
         "labels": {"II": 1, "XX": 1, "YY": -1, "ZZ": 1, "IX": 0, "YZ": 0, "ZX": 0, "YI": 0}},
        # psi = [sqrt2,0,0,0,0,0,0,1+1j]/2 (SV:1223-1234)
        {"src": f"{SV}:1223-1234",
         "state_re": [0.7071067811865476, 0, 0, 0, 0, 0, 0, 0.5],
         "state_im": [0, 0, 0, 0, 0, 0, 0, 0.5],
         "labels": {"XXX": 0.7071067811865476, "YYY": -0.7071067811865476, "ZZZ": 0, "XYZ": 0,
                    "YIY": 0}},
    ],
    # SparsePauliOp with complex coefficients (SV:1236-1241)
    "expval_sparse": {"src": f"{SV}:1236-1241",
                      "state_re": [0.7071067811865476, 0, 0, 0, 0, 0, 0, 0.5],
                      "state_im": [0, 0, 0, 0, 0, 0, 0, 0.5],
                      "labels": ["XXX", "IXI", "YYY", "III"],
                      "coeffs_re": [3.0, 5.5, 0.0, 23.0], "coeffs_im": [0.0, 0.0, -1.0, 0.0],
                      "expected_re": 25.121320343559642, "expected_im": 0.7071067811865476},
    # all 64 phase x Pauli labels vs the dense matrix, random_statevector(4, seed=1020)
    # (SV:1243-1316); the target is computed from the Kronecker matrix in the test.
    "expval_pauli_seed": 1020,
    "probabilities": [
        {"src": f"{SV}:675-704", "label": "+0", "cases": [
            [None, [0.5, 0, 0.5, 0]], [[0, 1], [0.5, 0, 0.5, 0]], [[1, 0], [0.5, 0.5, 0, 0]],
            [[0], [1, 0]], [[1], [0.5, 0.5]]]},
        {"src": f"{SV}:706-731", "label": "ghz3", "cases": [
            [[0, 1, 2], [0.5, 0, 0, 0, 0, 0, 0, 0.5]], [[2, 1, 0], [0.5, 0, 0, 0, 0, 0, 0, 0.5]],
            [[1, 2, 0], [0.5, 0, 0, 0, 0, 0, 0, 0.5]], [[1, 0, 2], [0.5, 0, 0, 0, 0, 0, 0, 0.5]],
            [[0, 1], [0.5, 0, 0, 0.5]], [[2, 1], [0.5, 0, 0, 0.5]], [[1, 2], [0.5, 0, 0, 0.5]],
            [[0], [0.5, 0.5]], [[1], [0.5, 0.5]], [[2], [0.5, 0.5]]]},
        {"src": f"{SV}:733-761", "label": "w3", "cases": [
            [[0, 1, 2], [0, 1 / 3, 1 / 3, 0, 1 / 3, 0, 0, 0]],
            [[2, 1, 0], [0, 1 / 3, 1 / 3, 0, 1 / 3, 0, 0, 0]],
            [[1, 2, 0], [0, 1 / 3, 1 / 3, 0, 1 / 3, 0, 0, 0]],
            [[1, 0, 2], [0, 1 / 3, 1 / 3, 0, 1 / 3, 0, 0, 0]],
            [[0, 1], [1 / 3, 1 / 3, 1 / 3, 0]], [[2, 1], [1 / 3, 1 / 3, 1 / 3, 0]],
            [[1, 2], [1 / 3, 1 / 3, 1 / 3, 0]],
            [[0], [2 / 3, 1 / 3]], [[1], [2 / 3, 1 / 3]], [[2], [2 / 3, 1 / 3]]]},
    ],
    # global phase: rz(0.5) expressed as p(0.5) with global phase 2pi-0.25 (SV:1407-1416)
    "global_phase": {"src": f"{SV}:1407-1416", "n": 1,
                     "gates": [["phase", [6.283185307179586 - 0.25], []], ["p", [0.5], [0]]],
                     "expected_re": [0.96891242, 0.0], "expected_im": [-0.24740396, 0.0],
                     "atol": 1e-8},
    # sampler, seed 123, 10 000 shots, reference asserts rtol=0.1 on the counts (SMP:37-79)
    "sampler": [
        {"src": f"{SMP}:50", "n": 2, "gates": [["h", [], [0]], ["cx", [], [0, 1]]],
         "counts": {"0": 5000, "3": 5000}},
        {"src": f"{SMP}:57", "n": 2, "gates": real_amplitudes_2q(2, [0] * 6),
         "counts": {"0": 10000}},
        {"src": f"{SMP}:58", "n": 2, "gates": real_amplitudes_2q(2, [1] * 6),
         "counts": {"0": 168, "1": 3389, "2": 470, "3": 5973}},
        {"src": f"{SMP}:59", "n": 2, "gates": real_amplitudes_2q(2, TH1),
         "counts": {"0": 1339, "1": 3534, "2": 912, "3": 4215}},
        {"src": f"{SMP}:60", "n": 2, "gates": real_amplitudes_2q(2, TH3),
         "counts": {"0": 634, "1": 291, "2": 6039, "3": 3036}},
        {"src": f"{SMP}:66", "n": 2, "gates": real_amplitudes_2q(3, [0, 1, 2, 3, 4, 5, 6, 7]),
         "counts": {"0": 1898, "1": 6864, "2": 928, "3": 311}},
    ],
    "sampler_seed": 123,
    "sampler_shots": 10000,
    # multiple classical registers (SMP:512-583): h on 3 qubits, measure q[0,1,2,2] -> clbits [0,2,4,5]
    # cregs a(1) b(2) c(3) laid out consecutively over clbits 0..5
    "sampler_cregs": [
        {"src": f"{SMP}:516-524", "n": 3, "gates": [["h", [], [0]], ["h", [], [1]], ["h", [], [2]]],
         "cregs": [["a", 1], ["b", 2], ["c", 3]], "measure": [[0, 0], [1, 2], [2, 4], [2, 5]],
         "counts": {"a": {"0": 5000, "1": 5000}, "b": {"0": 5000, "2": 5000},
                    "c": {"0": 5000, "6": 5000}}},
        {"src": f"{SMP}:527-539", "n": 3, "gates": [["h", [], [0]], ["h", [], [1]], ["h", [], [2]]],
         "cregs": [["a", 1], ["b", 5], ["c", 3]], "measure": [[0, 0], [1, 2], [2, 4], [2, 5]],
         "counts": {"a": {"0": 5000, "1": 5000},
                    "b": {"0": 2500, "2": 2500, "24": 2500, "26": 2500}, "c": {"0": 10000}}},
        {"src": f"{SMP}:542-550", "n": 3, "gates": [["h", [], [0]], ["h", [], [1]], ["h", [], [2]]],
         "cregs": [["a", 1], ["b", 2], ["c", 3]], "measure": [[1, 5]],
         "counts": {"a": {"0": 10000}, "b": {"0": 10000}, "c": {"0": 5000, "4": 5000}}},
        {"src": f"{SMP}:553-561", "n": 3, "gates": [["h", [], [0]], ["h", [], [1]], ["h", [], [2]]],
         "cregs": [["a", 1], ["b", 2], ["c", 3]], "measure": [[0, 5], [1, 5], [2, 5]],
         "counts": {"a": {"0": 10000}, "b": {"0": 10000}, "c": {"0": 5000, "4": 5000}}},
    ],
    # reversed measurement order (SMP:253-274): rx(x,0) rx(y,1) x(2); measure 0->2, 1->1, 2->0
    "sampler_reverse": [
        {"src": f"{SMP}:270", "n": 3,
         "gates": [["rx", [0.0], [0]], ["rx", [0.0], [1]], ["x", [], [2]]],
         "cregs": [["c", 3]], "measure": [[0, 2], [1, 1], [2, 0]], "counts": {"c": {"1": 10000}}},
        {"src": f"{SMP}:273", "n": 3,
         "gates": [["rx", [1.5707963267948966], [0]], ["rx", [0.0], [1]], ["x", [], [2]]],
         "cregs": [["c", 3]], "measure": [[0, 2], [1, 1], [2, 0]],
         "counts": {"c": {"1": 5000, "5": 5000}}},
    ],
}

if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_kats.json")
    with open(out, "w") as f:
        json.dump(kats, f, indent=1)
    print("wrote", out)
