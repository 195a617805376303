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
    # src/ops/conv.rs:1155-1290 test_conv_transpose / _padding / _1d (values from torch.nn.functional.conv_transpose2d)
    "conv_transpose": {
        "cite": "src/ops/conv.rs:1155-1290",
        "tol": 1e-5,
        "input_shape": [1, 1, 2, 2], "input": [1.0, 2.0, 3.0, 4.0],
        "kernel_shape": [1, 1, 2, 2], "kernel": [0.1, 0.2, 0.3, 0.4],
        "strides": [2, 2],
        "no_padding": {"shape": [1, 1, 4, 4],
                       "expected": [0.1, 0.2, 0.2, 0.4, 0.3, 0.4, 0.6, 0.8, 0.3, 0.6, 0.4, 0.8, 0.9, 1.2, 1.2, 1.6]},
        "bias": 1.234,
        "fixed_padding_1": {"pads": [1, 1, 1, 1], "shape": [1, 1, 2, 2], "expected": [0.4, 0.6, 0.6, 0.4]},
        "same_shape": [1, 1, 4, 4],
        "one_d": {"input_shape": [1, 1, 2], "input": [1.0, 2.0], "kernel_shape": [1, 1, 2], "kernel": [0.1, 0.2],
                  "strides": [2], "shape": [1, 1, 4], "expected": [0.1, 0.2, 0.2, 0.4], "bias": 0.5},
        # src/ops/conv.rs:1293-1382 test_conv_transpose_output_size_and_padding: (input, kernel, padding, strides) ->
        # (output size, padding) or the error message; padding "same" | list
        "size_table": [
            {"in": [5, 5], "k": [3, 3], "pad": [0, 0, 0, 0], "s": [1, 1], "out": [7, 7], "pads": [0, 0, 0, 0]},
            {"in": [5, 5], "k": [3, 3], "pad": [0, 0, 0, 0], "s": [3, 3], "out": [15, 15], "pads": [0, 0, 0, 0]},
            {"in": [5, 5], "k": [3, 3], "pad": [1, 1, 1, 1], "s": [1, 1], "out": [5, 5], "pads": [1, 1, 1, 1]},
            {"in": [5, 5], "k": [3, 3], "pad": [2, 2, 2, 2], "s": [1, 1], "out": [3, 3], "pads": [2, 2, 2, 2]},
            {"in": [5, 5], "k": [3, 3], "pad": [1, 2, 1, 2], "s": [1, 1], "out": [5, 3], "pads": [1, 2, 1, 2]},
            {"in": [5, 5], "k": [3, 3], "pad": "same", "s": [1, 1], "out": [5, 5], "pads": [1, 1, 1, 1]},
            {"in": [5, 5], "k": [1, 1], "pad": "same", "s": [3, 3], "error": "Input is too small"},
            {"in": [5, 5], "k": [3, 3], "pad": [4, 4, 4, 4], "s": [1, 1], "error": "Input is too small"},
            {"in": [5, 5], "k": [3, 3], "pad": [0, 0, 0, 0], "s": [0, 0], "error": "Strides must be > 0"},
            {"in": [0, 0], "k": [3, 3], "pad": [0, 0, 0, 0], "s": [1, 1], "error": "Input width and height must be > 0"},
            {"in": [1, 1], "k": [3, 3], "pad": [0, 0], "s": [1, 1], "error": "Wrong number of pad values"},
        ],
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
    # src/ops/pooling.rs:416-470 test_average_pool / :568-600 test_max_pool (shared 4x4 input)
    "pool_4x4": {
        "cite": "src/ops/pooling.rs:416-470, 568-600",
        "tol": 1e-5,
        "input_shape": [1, 1, 4, 4],
        "input": [0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.1, 0.2, 0.3, 0.4, 0.6, 0.7, 0.8, 0.9],
        "average": [
            {"kernel": [2, 2], "strides": [2, 2], "shape": [1, 1, 2, 2], "expected": [0.35, 0.55, 0.4, 0.6]},
            {"kernel": [4, 4], "strides": [4, 4], "shape": [1, 1, 1, 1], "expected": [0.475]},
            {"kernel": [2, 4], "strides": [2, 4], "shape": [1, 1, 2, 1], "expected": [0.45, 0.5]},
            {"kernel": [2, 2], "strides": [1, 2], "shape": [1, 1, 3, 2], "expected": [0.35, 0.55, 0.35, 0.55, 0.4, 0.6]},
        ],
        "max": [
            {"kernel": [2, 2], "strides": [2, 2], "shape": [1, 1, 2, 2], "expected": [0.6, 0.8, 0.7, 0.9]},
            {"kernel": [4, 4], "strides": [4, 4], "shape": [1, 1, 1, 1], "expected": [0.9]},
        ],
    },
    # src/ops/pooling.rs:498-556 test_average_pool_padding (PyTorch avg_pool2d, padding=1), tolerance 1e-4
    "average_pool_padding": {
        "cite": "src/ops/pooling.rs:498-556",
        "tol": 1e-4,
        "channels": 5,
        "plane": [[0.0809, 0.5529, 0.1534, 0.7507], [0.4698, 0.7771, 0.9896, 0.4873],
                  [0.9750, 0.5160, 0.6419, 0.3670], [0.4101, 0.3762, 0.9689, 0.4389]],
        "kernel": [2, 2], "strides": [2, 2], "pads": [1, 1, 1, 1],
        "expected_exclude_pad": [[0.0809, 0.3531, 0.7507], [0.7224, 0.7312, 0.4271], [0.4101, 0.6725, 0.4389]],
        "expected_include_pad": [[0.0202, 0.1766, 0.1877], [0.3612, 0.7312, 0.2136], [0.1025, 0.3363, 0.1097]],
    },
    # src/ops/pooling.rs:558-566 test_global_average_pool
    "global_average_pool": {"cite": "src/ops/pooling.rs:558-566", "input_shape": [1, 2, 2, 2],
                            "input": [1.0, 2.0, 3.0, 4.0, 10.0, 20.0, 30.0, 40.0], "expected": [2.5, 25.0]},
    # src/ops/norm.rs:708-743 test_softmax, tolerance 1e-4
    "softmax": {
        "cite": "src/ops/norm.rs:708-743",
        "tol": 1e-4,
        "input": [0.1634, 0.8647, 0.6401, 0.8265, 0.0560, 0.2304],
        "expected_1d": [0.1172, 0.2362, 0.1887, 0.2274, 0.1052, 0.1253],
        "expected_2x3_axis1": [[0.2161, 0.4358, 0.3481], [0.4966, 0.2298, 0.2736]],
        "expected_2x3_axis0": [[0.3400, 0.6918, 0.6010], [0.6600, 0.3082, 0.3990]],
    },
    # src/ops/norm.rs:635-668 test_layer_normalization (axis -1, default epsilon 1e-5), tolerance 1e-4
    "layer_normalization": {
        "cite": "src/ops/norm.rs:635-668",
        "tol": 1e-4,
        "input": [[[0.9562, 0.0572], [0.4366, 0.5655], [0.2017, 0.0230], [0.7941, 0.1554], [0.3226, 0.120]]],
        "scale": [0.0751, 0.6952], "bias": [0.9993, 0.7632], "axis": -1, "epsilon": 1e-5,
        "expected": [[[1.0744, 0.0680], [0.9243, 1.4576], [1.0744, 0.0684], [1.0744, 0.0680], [1.0744, 0.0683]]],
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
