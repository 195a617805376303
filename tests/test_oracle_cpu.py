"""CPU tests: the oracle (CPU restatement of the reference) against the reference's own golden vectors
(tests/golden/reference_literals.json, transcribed from the reference's unit tests with file:line) and against
independent float64 arithmetic.  This is what pins the oracle; the GPU parity tests then compare against it."""
import json
import os

import numpy as np
import pytest

from conftest import expect_equal

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_literals.json")))


def test_xorshift_rng_definition(oracle):
    """rten-tensor/src/rng.rs:11-32: xorshift64 (13, 7, 17); next_f32 = (40 most significant bits as f32) * 2^-40."""
    state = GOLD["xorshift"]["seed"]
    want = []
    for _ in range(16):
        state ^= (state << 13) & 0xFFFFFFFFFFFFFFFF
        state ^= state >> 7
        state ^= (state << 17) & 0xFFFFFFFFFFFFFFFF
        want.append(np.float32(state >> 24) * (np.float32(1.0) / np.float32(1 << 40)))
    got = oracle.XorShiftRng(GOLD["xorshift"]["seed"]).fill((16,))
    assert got.tobytes() == np.array(want, np.float32).tobytes()


def test_conv_golden(oracle):
    g = GOLD["conv_3x3"]
    x = np.array(g["input"], np.float32).reshape(g["input_shape"])
    w = np.array(g["kernel"], np.float32).reshape(g["kernel_shape"])
    for case in ("same_padding", "no_padding", "with_bias"):
        c = g[case]
        bias = np.array(c["bias"], np.float32) if "bias" in c else None
        y = oracle.conv2d(x, w, bias, padding=tuple(c["pads"]))
        assert list(y.shape) == c["shape"]
        np.testing.assert_allclose(y.reshape(-1), c["expected"], rtol=0, atol=g["tol"])
    # "same" auto-pad gives the explicit (1,1,1,1) result for a 3x3 kernel (src/ops/conv.rs:744-753)
    y = oracle.conv2d(x, w, None, padding="same")
    np.testing.assert_allclose(y.reshape(-1), g["same_padding"]["expected"], rtol=0, atol=g["tol"])


def test_depthwise_golden(oracle):
    g = GOLD["conv_depthwise"]
    x = np.array(g["input"], np.float32).reshape(g["input_shape"])
    w = np.array(g["kernel"], np.float32).reshape(g["kernel_shape"])
    b = np.array(g["bias"], np.float32)
    y = oracle.conv2d(x, w, b, groups=g["groups"])
    assert list(y.shape) == g["shape"]
    expect_equal(y.reshape(-1), np.array(g["expected_minus_bias"], np.float32) + b, rtol=g["tol"], atol=1e-6)


def test_pool_golden(oracle):
    g = GOLD["pool_4x4"]
    x = np.array(g["input"], np.float32).reshape(g["input_shape"])
    for kind, key in (("avg", "average"), ("max", "max")):
        for c in g[key]:
            y = oracle.pool2d(kind, x, tuple(c["kernel"]), tuple(c["strides"]))
            assert list(y.shape) == c["shape"]
            np.testing.assert_allclose(y.reshape(-1), c["expected"], rtol=0, atol=g["tol"])
    g = GOLD["average_pool_padding"]
    plane = np.array(g["plane"], np.float32)
    x = np.broadcast_to(plane, (1, g["channels"], 4, 4)).copy()
    for include, key in ((False, "expected_exclude_pad"), (True, "expected_include_pad")):
        y = oracle.pool2d("avg", x, tuple(g["kernel"]), tuple(g["strides"]), tuple(g["pads"]), include)
        want = np.broadcast_to(np.array(g[key], np.float32), (1, g["channels"], 3, 3))
        np.testing.assert_allclose(y, want, rtol=0, atol=g["tol"])
    g = GOLD["global_average_pool"]
    y = oracle.global_average_pool(np.array(g["input"], np.float32).reshape(g["input_shape"]))
    assert y.reshape(-1).tolist() == g["expected"]


def test_softmax_and_layer_norm_golden(oracle):
    g = GOLD["softmax"]
    x = np.array(g["input"], np.float32)
    np.testing.assert_allclose(oracle.softmax(x, 0), g["expected_1d"], rtol=0, atol=g["tol"])
    np.testing.assert_allclose(oracle.softmax(x.reshape(2, 3), 1), g["expected_2x3_axis1"], rtol=0, atol=g["tol"])
    np.testing.assert_allclose(oracle.softmax(x.reshape(2, 3), 0), g["expected_2x3_axis0"], rtol=0, atol=g["tol"])
    g = GOLD["layer_normalization"]
    y = oracle.layer_norm(np.array(g["input"], np.float32), np.array(g["scale"], np.float32),
                          np.array(g["bias"], np.float32), g["axis"], g["epsilon"])
    np.testing.assert_allclose(y, g["expected"], rtol=0, atol=g["tol"])


def test_vecmath_error_bounds(oracle):
    """erf: max |diff| vs libm <= 6.631017e-7 (rten-vecmath/src/erf.rs:138); exp within 1 ULP, sigmoid within 4 ULP
    (rten-vecmath/src/exp.rs:256-260) over the ranges the reference's tests sweep."""
    import math

    g = GOLD["vecmath"]
    xs = np.linspace(-6.0, 6.0, 4001, dtype=np.float32)
    erf = np.array([oracle.vm_scalar("erf", float(v)) for v in xs], np.float32)
    want = np.array([math.erf(float(v)) for v in xs], np.float64)
    assert np.max(np.abs(erf - want)) <= g["erf_max_abs_diff"] * 1.05

    def ulps(a, b):
        a = np.asarray(a, np.float32).view(np.int32).astype(np.int64)
        b = np.asarray(b, np.float32).view(np.int32).astype(np.int64)
        return np.abs(a - b)

    xs = np.linspace(-80.0, 80.0, 4001, dtype=np.float32)
    got = np.array([oracle.vm_scalar("exp", float(v)) for v in xs], np.float32)
    want = np.exp(xs.astype(np.float64)).astype(np.float32)
    ok = np.isfinite(want) & (want > 1e-37)
    assert ulps(got[ok], want[ok]).max() <= g["exp_max_ulps"]
    xs = np.linspace(-20.0, 20.0, 4001, dtype=np.float32)
    got = np.array([oracle.vm_scalar("sigmoid", float(v)) for v in xs], np.float32)
    want = (1.0 / (1.0 + np.exp(-xs.astype(np.float64)))).astype(np.float32)
    assert ulps(got, want).max() <= g["sigmoid_max_ulps"]


@pytest.mark.parametrize("m,n,k", [(1, 1, 1), (6, 16, 8), (7, 33, 129), (64, 64, 300), (130, 70, 257)])
def test_gemm_against_float64(oracle, m, n, k):
    """GemmExecutor semantics out = alpha * a @ b + beta * out (src/gemm.rs:186-281), checked in float64 the way the
    reference's tests check against reference_gemm (src/gemm.rs:1080-1110)."""
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((m, k)), rng.fill((k, n))
    out0 = rng.fill((m, n))
    want = 0.5 * (a.astype(np.float64) @ b.astype(np.float64)) + 2.0 * out0
    got = oracle.gemm(a, b, out=out0.copy(), alpha=0.5, beta=2.0)
    expect_equal(got, want, rtol=1e-5, atol=1e-6)
    bias = rng.fill((m,))
    got = oracle.gemm(a, b, bias=bias)
    expect_equal(got, a.astype(np.float64) @ b.astype(np.float64) + bias[:, None], rtol=1e-5, atol=1e-6)


def test_gemm_edge_cases(oracle):
    """K == 0 scales the output by beta (src/gemm.rs:772-785); empty outputs are a no-op."""
    a, b = np.zeros((3, 0), np.float32), np.zeros((0, 4), np.float32)
    out = np.full((3, 4), 2.0, np.float32)
    got = oracle.gemm(a, b, out=out.copy(), beta=0.5)
    assert (got == 1.0).all()
    got = oracle.gemm(a, b, out=out.copy(), beta=0.0)
    assert (got == 0.0).all()
    assert oracle.gemm(np.zeros((0, 5), np.float32), np.zeros((5, 4), np.float32)).shape == (0, 4)


def test_conv_matches_direct_sum(oracle):
    """im2col + GEMM conv against a float64 direct convolution, incl. stride, dilation, groups, asymmetric pads."""
    rng = oracle.XorShiftRng(1234)
    x = rng.fill((2, 6, 9, 8))
    w = rng.fill((4, 3, 3, 2)) - 0.5
    bias = rng.fill((4,))
    pads, strides, dil, groups = (1, 0, 2, 1), (2, 1), (1, 2), 2
    y = oracle.conv2d(x, w, bias, padding=pads, groups=groups, strides=strides, dilations=dil)
    n, c, h, wd = x.shape
    o, cg, kh, kw = w.shape
    xp = np.zeros((n, c, h + pads[0] + pads[2], wd + pads[1] + pads[3]), np.float64)
    xp[:, :, pads[0]:pads[0] + h, pads[1]:pads[1] + wd] = x
    oh = (xp.shape[2] - dil[0] * (kh - 1) - 1) // strides[0] + 1
    ow = (xp.shape[3] - dil[1] * (kw - 1) - 1) // strides[1] + 1
    want = np.zeros((n, o, oh, ow))
    og = o // groups
    for oc in range(o):
        g0 = (oc // og) * cg
        for ky in range(kh):
            for kx in range(kw):
                patch = xp[:, g0:g0 + cg, ky * dil[0]:ky * dil[0] + oh * strides[0]:strides[0],
                           kx * dil[1]:kx * dil[1] + ow * strides[1]:strides[1]]
                want[:, oc] += np.einsum("nchw,c->nhw", patch, w[oc, :, ky, kx].astype(np.float64))
        want[:, oc] += bias[oc]
    assert y.shape == want.shape
    expect_equal(y, want, rtol=1e-5, atol=1e-6)


def test_conv_error_messages(oracle):
    """OpError variants and messages of conv() (src/ops/conv.rs:101-140)."""
    x = np.zeros((1, 4, 5, 5), np.float32)
    with pytest.raises(oracle.OracleError, match="Input channels"):
        oracle.conv2d(x, np.zeros((4, 3, 3, 3), np.float32))
    with pytest.raises(oracle.OracleError, match="Input too small"):
        oracle.conv2d(np.zeros((1, 1, 2, 2), np.float32), np.zeros((1, 1, 3, 3), np.float32))
    with pytest.raises(oracle.OracleError, match="divisible by group count"):
        oracle.conv2d(np.zeros((1, 4, 5, 5), np.float32), np.zeros((3, 2, 3, 3), np.float32), groups=2)


def test_graph_oracle_resnet_block(oracle):
    """The graph evaluator applies the same operators Model::run would: Conv -> Add -> Relu equals the sum by hand."""
    import graph_oracle
    from rten_fork_b200.rten_builder import ModelBuilder

    rng = oracle.XorShiftRng(7)
    b = ModelBuilder()
    x = b.add_value("x", [1, 4, 6, 6])
    b.inputs = [x]
    w = b.add_constant("w", rng.fill((4, 4, 3, 3)) - 0.5)
    y = b.op("Conv", [x, w], ModelBuilder.conv_attrs((1, 1, 1, 1)))
    y = b.op("Add", [y, x])
    b.outputs = [b.op("Relu", [y])]
    xin = rng.fill((1, 4, 6, 6)) - 0.5
    got = graph_oracle.run_graph(b, [xin])[0]
    want = np.maximum(oracle.conv2d(xin, b.nodes[w]["data"], None, padding=(1, 1, 1, 1)) + xin, 0)
    assert got.tobytes() == want.astype(np.float32).tobytes()


def test_conv_transpose_oracle_matches_reference_literals():
    """src/ops/conv.rs:1155-1382: conv_transpose2d literals (from torch), fixed / same padding, 1-D, and the
    output-size-and-padding table with its error messages."""
    import oracle as o

    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_literals.json")))["conv_transpose"]
    x = np.array(g["input"], np.float32).reshape(g["input_shape"])
    w = np.array(g["kernel"], np.float32).reshape(g["kernel_shape"])
    y = o.conv_transpose(x, w, None, (0, 0, 0, 0), tuple(g["strides"]))
    assert list(y.shape) == g["no_padding"]["shape"]
    np.testing.assert_allclose(y.reshape(-1), g["no_padding"]["expected"], rtol=g["tol"], atol=1e-8)
    yb = o.conv_transpose(x, w, np.array([g["bias"]], np.float32), (0, 0, 0, 0), tuple(g["strides"]))
    np.testing.assert_allclose(yb.reshape(-1), np.array(g["no_padding"]["expected"]) + g["bias"], rtol=g["tol"])
    yp = o.conv_transpose(x, w, None, tuple(g["fixed_padding_1"]["pads"]), tuple(g["strides"]))
    np.testing.assert_allclose(yp.reshape(-1), g["fixed_padding_1"]["expected"], rtol=g["tol"])
    assert list(o.conv_transpose(x, w, None, "same", tuple(g["strides"])).shape) == g["same_shape"]
    d = g["one_d"]
    y1 = o.conv_transpose(np.array(d["input"], np.float32).reshape(d["input_shape"]),
                          np.array(d["kernel"], np.float32).reshape(d["kernel_shape"]), None, (0, 0), tuple(d["strides"]))
    assert list(y1.shape) == d["shape"]
    np.testing.assert_allclose(y1.reshape(-1), d["expected"], rtol=g["tol"])
    for c in g["size_table"]:
        if "error" in c:
            with pytest.raises(o.OracleError, match=c["error"]):
                o.conv_transpose_output_size_and_padding(c["in"], c["k"], c["pad"], c["s"])
        else:
            out, pads = o.conv_transpose_output_size_and_padding(c["in"], c["k"], c["pad"], c["s"])
            assert list(out) == c["out"] and list(pads) == c["pads"]


def test_einsum_oracle_against_float64():
    """The Einsum restatement (permute -> MatMul -> permute, left-to-right folding) against numpy's float64 einsum on
    the equation forms the reference's tests use (src/ops/einsum.rs:740-960)."""
    import graph_oracle

    rng = np.random.default_rng(0)
    f = lambda *s: rng.standard_normal(s).astype(np.float32)  # noqa: E731
    cases = [("ij->ji", [f(3, 5)]), ("ij,jk->ik", [f(3, 5), f(5, 4)]), ("ij,jk", [f(3, 5), f(5, 4)]),
             ("ji,kj->ik", [f(5, 3), f(4, 5)]), ("ij,jk->ki", [f(3, 5), f(5, 4)]),
             ("bhwc,hkc->bhwk", [f(2, 3, 4, 6), f(3, 5, 6)]), ("i,i->", [f(7), f(7)]), ("ij,j->i", [f(3, 5), f(5)]),
             ("j,jk->k", [f(5), f(5, 4)]), ("ij,ij->", [f(3, 5), f(3, 5)]), ("ij,j->", [f(3, 5), f(5)]),
             ("i,j->ij", [f(3), f(4)]), ("bhqd,bhkd->bhqk", [f(2, 3, 4, 8), f(2, 3, 5, 8)]),
             ("ab,bc,cd,de->ea", [f(2, 3), f(3, 4), f(4, 5), f(5, 6)])]
    for eq, xs in cases:
        want = np.einsum(eq, *[x.astype(np.float64) for x in xs])
        got = graph_oracle.einsum(eq, xs)
        assert got.shape == want.shape, eq
        np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-5, err_msg=eq)
