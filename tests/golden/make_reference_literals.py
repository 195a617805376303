"""Writes tests/golden/reference_literals.json: the known-answer vectors that the reference's own unit tests hold for
this path, transcribed from /root/reference (file:line cited per case).  They are PyTorch-derived literals in the
reference, so they pin the oracle independently of our restatement.  Run from the repo root in the build container:

    python tests/golden/make_reference_literals.py
"""
import json
import os

cases = {
    # src/ops/conv.rs:708-765 test_conv (tolerance 1e-4: expect_eq_1e4, src/ops/mod.rs:1131-1136)
    "conv_3x3": {
        "cite": "src/ops/conv.rs:708-765",
        "tol": 1e-4,
        "input_shape": [1, 1, 3, 3],
        "input": [0.5946, 0.8249, 0.0448, 0.9552, 0.2041, 0.2501, 0.2693, 0.1007, 0.8862],
        "kernel_shape": [1, 1, 3, 3],
        "kernel": [0.3230, 0.7632, 0.4616, 0.8837, 0.5898, 0.3424, 0.2101, 0.7821, 0.6861],
        "same_padding": {"pads": [1, 1, 1, 1], "shape": [1, 1, 3, 3],
                         "expected": [1.5202, 1.5592, 0.9939, 1.7475, 2.6358, 1.3428, 1.0165, 1.1806, 0.8685]},
        "no_padding": {"pads": [0, 0, 0, 0], "shape": [1, 1, 1, 1], "expected": [2.6358]},
        "with_bias": {"pads": [0, 0, 0, 0], "bias": [1.0], "shape": [1, 1, 1, 1], "expected": [3.6358]},
    },
    # src/ops/conv.rs:922-956 test_conv_depthwise (default tolerance: rtol 1e-5, atol 1e-8)
    "conv_depthwise": {
        "cite": "src/ops/conv.rs:922-956",
        "tol": 1e-5,
        "input_shape": [1, 3, 2, 2],
        "input": [0.5946, 0.8249, 0.0448, 0.9552, 0.2041, 0.2501, 0.2693, 0.1007, 1.5202, 1.5592, 0.9939, 1.7475],
        "kernel_shape": [3, 1, 2, 2],
        "kernel": [-0.0862, -0.4111, 0.0813, 0.4993, -0.4641, 0.1715, -0.0532, -0.2429, -0.4325, 0.4273, 0.4180,
                   0.4338],
        "bias": [0.1, 0.2, 0.3],
        "groups": 3,
        "shape": [1, 3, 1, 1],
        "expected_minus_bias": [0.09020272, -0.09061745, 1.1822754],
    },
    # rten-vecmath/src/erf.rs:138 MAX_EXPECTED_DIFF vs libm; exp.rs:256,260 ULP bounds
    "vecmath": {"cite": "rten-vecmath/src/erf.rs:138; rten-vecmath/src/exp.rs:256-260",
                "erf_max_abs_diff": 6.631017e-7, "exp_max_ulps": 1.0, "sigmoid_max_ulps": 4.0},
    # rten-tensor/src/rng.rs:11-32 XorShiftRng(1234): first values, checked here against the algorithm's definition
    "xorshift": {"cite": "rten-tensor/src/rng.rs:11-32", "seed": 1234},
}

out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_literals.json")
with open(out, "w") as f:
    json.dump(cases, f, indent=1)
print("wrote", out)
