"""Conv parity: CUDA path (C-ABI, host buffers) vs the oracle restatement and the reference's golden literals.

Cases restate src/ops/conv.rs:706-1153 (basic, SAME padding, uneven padding, pointwise incl. non-contiguous input /
batch / stride 2, depthwise, grouped, strided sweep, dilation sweep, 1-D, error strings).
"""
import json
import os

import numpy as np
import pytest

from conftest import expect_equal

pytestmark = pytest.mark.gpu

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_literals.json")))


def check_conv(rtg, oracle, tol, x, w, b, pads, groups, strides, dilations):
    """check_conv (src/ops/conv.rs:677-702): optimized result vs reference result, same inputs."""
    from rten_fork_b200 import ops

    got = ops.conv(x, w, b, pads, groups, strides, dilations)
    want = oracle.conv2d(x, w, b, padding=pads, groups=groups, strides=strides, dilations=dilations)
    expect_equal(got, want, rtol=tol)
    return got


def test_conv_golden_literals(rtg, ctx, oracle, math_mode):
    from rten_fork_b200 import ops

    g = GOLD["conv_3x3"]
    x = np.array(g["input"], np.float32).reshape(g["input_shape"])
    w = np.array(g["kernel"], np.float32).reshape(g["kernel_shape"])
    tol = max(g["tol"], math_mode["rtol"])
    for name in ("same_padding", "no_padding", "with_bias"):
        c = g[name]
        b = np.array(c["bias"], np.float32) if "bias" in c else None
        got = ops.conv(x, w, b, c["pads"], 1, (1, 1), (1, 1))
        want = np.array(c["expected"], np.float32).reshape(c["shape"])
        np.testing.assert_allclose(got, want, rtol=tol, atol=tol)
    d = GOLD["conv_depthwise"]
    x = np.array(d["input"], np.float32).reshape(d["input_shape"])
    w = np.array(d["kernel"], np.float32).reshape(d["kernel_shape"])
    b = np.array(d["bias"], np.float32)
    got = ops.conv(x, w, b, (0, 0, 0, 0), d["groups"], (1, 1), (1, 1))
    want = (np.array(d["expected_minus_bias"], np.float32) + b).reshape(d["shape"])
    expect_equal(got, want, rtol=1e-5, atol=1e-7)


def test_conv_same_padding(rtg, ctx, oracle, math_mode):  # src/ops/conv.rs:769-811
    from rten_fork_b200 import ops

    g = GOLD["conv_3x3"]
    x = np.array(g["input"], np.float32).reshape(g["input_shape"])
    w = np.array(g["kernel"], np.float32).reshape(g["kernel_shape"])
    got = ops.Conv(padding="same", groups=1, strides=(1, 1), dilations=(1, 1)).run([x, w])[0]
    expect_equal(got, oracle.conv2d(x, w, None, padding=(1, 1, 1, 1)), rtol=math_mode["rtol"])


def test_conv_uneven_padding(rtg, ctx, oracle, math_mode):  # :813-831
    rng = oracle.XorShiftRng(1234)
    w, x, b = rng.fill((10, 5, 3, 3)), rng.fill((1, 5, 10, 10)), rng.fill((10,))
    check_conv(rtg, oracle, math_mode["rtol"], x, w, b, (0, 0, 1, 1), 1, (1, 1), (1, 1))


def test_conv_depthwise_uneven_padding(rtg, ctx, oracle):  # :833-851
    rng = oracle.XorShiftRng(1234)
    w, x, b = rng.fill((10, 1, 3, 3)), rng.fill((1, 10, 10, 10)), rng.fill((10,))
    check_conv(rtg, oracle, 1e-5, x, w, b, (0, 0, 1, 1), 10, (1, 1), (1, 1))


def test_conv_pointwise(rtg, ctx, oracle, math_mode):  # :854-916
    tol = math_mode["rtol"]
    rng = oracle.XorShiftRng(1234)
    w, x, b = rng.fill((10, 5, 1, 1)), rng.fill((1, 5, 20, 20)), rng.fill((10,))
    assert check_conv(rtg, oracle, tol, x, w, b, (0, 0, 0, 0), 1, (1, 1), (1, 1)).shape == (1, 10, 20, 20)
    xt = x.transpose(0, 1, 3, 2)  # non-contiguous input
    assert not xt.flags.c_contiguous
    from rten_fork_b200 import ops

    got = ops.conv(xt, w, b, (0, 0, 0, 0), 1, (1, 1), (1, 1))
    expect_equal(got, oracle.conv2d(np.ascontiguousarray(xt), w, b), rtol=tol)
    x2 = rng.fill((2, 5, 20, 20))
    assert check_conv(rtg, oracle, tol, x2, w, b, (0, 0, 0, 0), 1, (1, 1), (1, 1)).shape == (2, 10, 20, 20)
    x3 = rng.fill((1, 5, 20, 20))
    assert check_conv(rtg, oracle, tol, x3, w, b, (0, 0, 0, 0), 1, (2, 2), (1, 1)).shape == (1, 10, 10, 10)


def test_conv_grouped(rtg, ctx, oracle, math_mode):  # :963-981 test_conv_not_depthwise_or_pointwise
    rng = oracle.XorShiftRng(1234)
    w, x, b = rng.fill((4, 2, 3, 3)), rng.fill((2, 4, 20, 20)), rng.fill((4,))
    check_conv(rtg, oracle, math_mode["rtol"], x, w, b, (1, 1, 1, 1), 2, (1, 1), (1, 1))
    # aligned group slices (C/g % 4 == 0) take the shared-NHWC path
    w, x, b = rng.fill((16, 8, 3, 3)), rng.fill((2, 32, 9, 11)), rng.fill((16,))
    check_conv(rtg, oracle, math_mode["rtol"], x, w, b, (1, 1, 1, 1), 4, (1, 1), (1, 1))


def test_conv_grouped_does_not_leak_non_finite_values_across_groups(rtg, ctx, oracle, math_mode):
    """The shared-NHWC grouped path reads 32-channel boxes; a group of 8 channels must not see its neighbour's
    channels (0-weight x Inf would give NaN where the reference's per-group im2col gives finite values)."""
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(77)
    w, x, b = rng.fill((16, 8, 3, 3)) - 0.5, rng.fill((2, 32, 9, 11)), rng.fill((16,))
    x[:, 8:16] = np.inf    # group 1 only
    x[0, 24:32, 3, 4] = np.nan  # group 3 only
    got = ops.conv(x, w, b, padding=(1, 1, 1, 1), groups=4, ctx=ctx)
    want = oracle.conv2d(np.where(np.isfinite(x), x, 0).astype(np.float32), w, b, (1, 1, 1, 1), 4, (1, 1), (1, 1))
    for grp in (0, 2):      # groups whose own inputs are finite
        sl = slice(4 * grp, 4 * grp + 4)
        assert np.isfinite(got[:, sl]).all(), "non-finite values leaked from a neighbouring group"
        np.testing.assert_allclose(got[:, sl], want[:, sl], rtol=math_mode["rtol"] * 4, atol=math_mode["rtol"] * 4 * np.abs(want).max())
    assert not np.isfinite(got[:, 4:8]).all()   # group 1 really is poisoned, as in the reference


@pytest.mark.parametrize("strides", [(2, 2), (3, 3), (1, 3)])
@pytest.mark.parametrize("pad", [0, 1])
def test_conv_strided(rtg, ctx, oracle, math_mode, strides, pad):  # :983-1005
    rng = oracle.XorShiftRng(1234)
    w = rng.fill((4, 3, 3, 3))
    for size in (3, 4, 5, 10, 20):
        x = rng.fill((2, 3, size, size))
        check_conv(rtg, oracle, math_mode["rtol"], x, w, None, (pad,) * 4, 1, strides, (1, 1))


@pytest.mark.parametrize("strides", [(2, 2), (3, 3), (1, 3)])
@pytest.mark.parametrize("pad", [0, 1])
def test_conv_strided_depthwise(rtg, ctx, oracle, strides, pad):  # :1007-1031
    rng = oracle.XorShiftRng(1234)
    w = rng.fill((3, 1, 3, 3))
    for size in (3, 4, 5, 10, 20):
        x = rng.fill((1, 3, size, size))
        check_conv(rtg, oracle, 1e-5, x, w, None, (pad,) * 4, 3, strides, (1, 1))


@pytest.mark.parametrize("dilations", [(2, 2), (3, 3), (1, 3)])
@pytest.mark.parametrize("pad", [0, 1])
def test_conv_dilated(rtg, ctx, oracle, math_mode, dilations, pad):  # :1081-1104
    rng = oracle.XorShiftRng(1234)
    w = rng.fill((4, 3, 3, 3))
    for size in (7, 10, 20):
        x = rng.fill((2, 3, size, size))
        check_conv(rtg, oracle, math_mode["rtol"], x, w, None, (pad,) * 4, 1, (1, 1), dilations)


@pytest.mark.parametrize("dilations", [(2, 2), (3, 3), (1, 3)])
@pytest.mark.parametrize("pad", [0, 1])
def test_conv_dilated_depthwise(rtg, ctx, oracle, dilations, pad):  # :1106-1130
    rng = oracle.XorShiftRng(1234)
    w = rng.fill((3, 1, 3, 3))
    for size in (7, 10, 20):
        x = rng.fill((2, 3, size, size))
        check_conv(rtg, oracle, 1e-5, x, w, None, (pad,) * 4, 3, (1, 1), dilations)


def test_conv_1d(rtg, ctx, oracle, math_mode):  # :1132-1153
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    x, w = rng.fill((1, 5, 20)), rng.fill((10, 5, 3))
    got = ops.conv(x, w, None, "same", 1, (1,), (1,))
    assert got.shape == (1, 10, 20)
    expect_equal(got, oracle.conv2d(x, w, None, padding="same", strides=(1,), dilations=(1,)), rtol=math_mode["rtol"])


def test_conv_errors(rtg, ctx, oracle):  # :1033-1079 and conv() checks :98-197
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    with pytest.raises(rtg.RtgError, match="Input too small for kernel size") as e:
        ops.conv(rng.fill((1, 1, 2, 2)), rng.fill((1, 1, 3, 3)), None, (0, 0, 0, 0), 1, (1, 1), (1, 1))
    assert e.value.kind == "InvalidValue"
    with pytest.raises(rtg.RtgError, match="Strides must be > 0"):
        ops.conv(rng.fill((1, 1, 2, 2)), rng.fill((1, 1, 2, 2)), None, (0, 0, 0, 0), 1, (0, 0), (1, 1))
    with pytest.raises(rtg.RtgError, match="Dilations must be > 0"):
        ops.conv(rng.fill((1, 1, 2, 2)), rng.fill((1, 1, 2, 2)), None, (0, 0, 0, 0), 1, (1, 1), (0, 1))
    with pytest.raises(rtg.RtgError, match=r"Input channels \(per group\) does not match kernel input channels") as e:
        ops.conv(rng.fill((1, 4, 5, 5)), rng.fill((4, 3, 3, 3)), None, (0, 0, 0, 0), 1, (1, 1), (1, 1))
    assert e.value.kind == "IncompatibleInputShapes"
    with pytest.raises(rtg.RtgError, match="must be divisible by group count"):
        ops.conv(rng.fill((1, 5, 5, 5)), rng.fill((4, 2, 3, 3)), None, (0, 0, 0, 0), 2, (1, 1), (1, 1))
    with pytest.raises(rtg.RtgError, match="expected 2 stride values"):
        ops.conv(rng.fill((1, 1, 4, 4)), rng.fill((1, 1, 3, 3)), None, (0, 0, 0, 0), 1, (1,), (1, 1))
    with pytest.raises(rtg.RtgError, match=r"kernel must have 4 dims \(OCHW\)"):
        ops.conv(rng.fill((1, 1, 4, 4)), rng.fill((1, 3, 3)), None, (0, 0, 0, 0), 1, (1, 1), (1, 1))
    with pytest.raises(rtg.RtgError, match="expected 2 pad values"):
        ops.conv(rng.fill((1, 1, 8)), rng.fill((1, 1, 3)), None, (0, 0, 0, 0), 1, (1,), (1,))


# ResNet-50 / MobileNetV2 layer shapes at reduced batch (BASELINE.json configs 2 and 4): every distinct
# (kernel, stride, channel) family the graphs contain.
RESNET_LAYERS = [
    # (C, O, H, k, stride, pad, groups)
    (3, 64, 64, 7, 2, 3, 1),      # conv1 (C=3: channel padding path)
    (64, 64, 28, 1, 1, 0, 1),     # pointwise
    (64, 64, 28, 3, 1, 1, 1),     # 3x3
    (64, 256, 28, 1, 1, 0, 1),
    (256, 128, 28, 1, 1, 0, 1),
    (128, 128, 28, 3, 2, 1, 1),   # 3x3 stride 2
    (256, 512, 28, 1, 2, 0, 1),   # 1x1 stride 2 downsample (im2col path in the reference)
    (512, 512, 7, 3, 1, 1, 1),    # K = 4608
    (512, 2048, 7, 1, 1, 0, 1),
    (32, 32, 28, 3, 1, 1, 32),    # MobileNetV2 depthwise
    (96, 96, 28, 3, 2, 1, 96),    # depthwise stride 2
    (144, 24, 14, 1, 1, 0, 1),    # MobileNetV2 projection (C % 32 != 0)
]


@pytest.mark.parametrize("C,O,H,k,s,p,g", RESNET_LAYERS)
def test_conv_model_layers(rtg, ctx, oracle, math_mode, C, O, H, k, s, p, g):
    rng = oracle.XorShiftRng(1234)
    x = rng.fill((2, C, H, H)) - 0.5
    x = np.abs(x) + 0.1  # keep sums away from cancellation so that the relative rule is meaningful
    w = rng.fill((O, C // g, k, k))
    b = rng.fill((O,))
    tol = 1e-5 if g == C else math_mode["rtol"]
    check_conv(rtg, oracle, tol, x.astype(np.float32), w, b, (p,) * 4, g, (s, s), (1, 1))


def test_conv_fused_epilogue_matches_unfused(rtg, ctx, oracle, math_mode):
    """act(conv + bias + residual) fused in the GEMM epilogue == Conv -> Add -> Relu run as separate operators."""
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    x, w, b = rng.fill((2, 64, 14, 14)), rng.fill((128, 64, 3, 3)), rng.fill((128,))
    res = rng.fill((2, 128, 14, 14))
    fused = ops.conv(x, w, b, (1, 1, 1, 1), 1, (1, 1), (1, 1), residual=res, activation="relu")
    want = oracle.unary("relu", oracle.binary("add", oracle.conv2d(x, w, b, padding=(1, 1, 1, 1)), res))
    expect_equal(fused, want, rtol=math_mode["rtol"])


def test_conv_nhwc_device_views(rtg, ctx, oracle, math_mode):
    """NHWC-strided device views are consumed and produced without layout copies (the graph executor's fast path)."""
    import torch
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    x, w, b = rng.fill((3, 32, 12, 10)), rng.fill((48, 32, 3, 3)), rng.fill((48,))
    xd = torch.from_numpy(x).cuda().permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)  # NCHW shape, NHWC memory
    out = ops.conv(xd, torch.from_numpy(w).cuda(), torch.from_numpy(b).cuda(), (1, 1, 1, 1), 1, (2, 2), (1, 1),
                   out_nhwc=True)
    ctx.sync()
    assert out.stride(1) == 1
    expect_equal(out.cpu().numpy(), oracle.conv2d(x, w, b, padding=(1, 1, 1, 1), strides=(2, 2)), rtol=math_mode["rtol"])


@pytest.mark.parametrize("C,O,H,k,s", [(256, 256, 14, 3, 1), (512, 512, 7, 3, 1), (1024, 256, 14, 1, 1), (128, 128, 28, 3, 2)])
def test_conv_split_k_nhwc(rtg, ctx, oracle, math_mode, C, O, H, k, s):
    """ResNet-50 layer3 / layer4 shapes have 50-200 output tiles for 148 SMs: the im2col launch splits K (tap and
    channel-block ranges per work unit) and the reduction pass applies bias + residual + Relu."""
    import torch
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    p = k // 2
    x, w, b = rng.fill((4, C, H, H)) + 0.1, rng.fill((O, C, k, k)), rng.fill((O,))
    oh = (H + 2 * p - k) // s + 1
    res = rng.fill((4, O, oh, oh))
    nhwc = lambda a: torch.from_numpy(a).cuda().contiguous(memory_format=torch.channels_last)  # noqa: E731
    out = ops.conv(nhwc(x), torch.from_numpy(w).cuda(), torch.from_numpy(b).cuda(), (p, p, p, p), 1, (s, s), (1, 1),
                   residual=nhwc(res), activation="relu", out_nhwc=True)
    ctx.sync()
    assert "splitK" in ctx.last_kernel, ctx.last_kernel
    want = oracle.unary("relu", oracle.binary("add", oracle.conv2d(x, w, b, padding=(p, p, p, p), strides=(s, s)), res))
    expect_equal(out.cpu().numpy(), want, rtol=math_mode["rtol"])


# ------------------------------------------------------------------------------------------------ ConvTranspose
def test_conv_transpose_golden_literals(rtg, ctx, oracle, math_mode):  # src/ops/conv.rs:1155-1290
    from rten_fork_b200 import ops

    g = GOLD["conv_transpose"]
    x = np.array(g["input"], np.float32).reshape(g["input_shape"])
    w = np.array(g["kernel"], np.float32).reshape(g["kernel_shape"])
    s = tuple(g["strides"])
    tol = max(g["tol"], math_mode["rtol"])
    np.testing.assert_allclose(ops.conv_transpose(x, w, None, (0, 0, 0, 0), s).reshape(-1), g["no_padding"]["expected"], rtol=tol)
    got = ops.conv_transpose(x, w, np.array([g["bias"]], np.float32), (0, 0, 0, 0), s)
    np.testing.assert_allclose(got.reshape(-1), np.array(g["no_padding"]["expected"]) + g["bias"], rtol=tol)
    got = ops.conv_transpose(x, w, None, tuple(g["fixed_padding_1"]["pads"]), s)
    assert list(got.shape) == g["fixed_padding_1"]["shape"]
    np.testing.assert_allclose(got.reshape(-1), g["fixed_padding_1"]["expected"], rtol=tol)
    assert list(ops.conv_transpose(x, w, None, "same", s).shape) == g["same_shape"]
    d = g["one_d"]
    got = ops.conv_transpose(np.array(d["input"], np.float32).reshape(d["input_shape"]),
                             np.array(d["kernel"], np.float32).reshape(d["kernel_shape"]),
                             np.array([d["bias"]], np.float32), (0, 0), tuple(d["strides"]))
    np.testing.assert_allclose(got.reshape(-1), np.array(d["expected"]) + d["bias"], rtol=tol)


@pytest.mark.parametrize("shape,wshape,pads,strides", [
    ((2, 8, 5, 7), (8, 6, 3, 3), (0, 0, 0, 0), (1, 1)),
    ((2, 8, 5, 7), (8, 6, 3, 3), (1, 1, 1, 1), (2, 2)),
    ((1, 32, 6, 6), (32, 20, 4, 4), (1, 1, 1, 1), (2, 2)),     # the usual 2x upsampling deconvolution
    ((3, 5, 4, 9), (5, 7, 2, 3), (0, 1, 1, 2), (3, 2)),        # uneven padding, C % 4 != 0
    ((2, 16, 8, 8), (16, 4, 3, 3), "same", (2, 2)),
    ((2, 16, 8, 8), (16, 4, 5, 3), "same", (1, 1)),            # the reference's "same" quirk: left pad = ceil(pad_h / 2)
    ((2, 6, 11), (6, 3, 4), (1, 2), (3,)),                     # 1-D
])
def test_conv_transpose_vs_oracle(rtg, ctx, oracle, math_mode, shape, wshape, pads, strides):
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    x, w, b = rng.fill(shape), rng.fill(wshape), rng.fill((wshape[1],))
    got = ops.conv_transpose(x, w, b, pads, strides)
    want = oracle.conv_transpose(x, w, b, pads, strides)
    expect_equal(got, want, rtol=math_mode["rtol"])


def test_conv_transpose_size_table_and_errors(rtg, ctx, oracle):  # src/ops/conv.rs:1293-1382, :441-505
    from rten_fork_b200 import ops

    for c in GOLD["conv_transpose"]["size_table"]:
        x = np.zeros((1, 2, *c["in"]), np.float32)
        w = np.zeros((2, 3, *c["k"]), np.float32)
        if "error" in c:
            with pytest.raises(rtg.RtgError, match=c["error"]):
                ops.conv_transpose_output_shape(x, w, None, c["pad"], c["s"])
        else:
            assert ops.conv_transpose_output_shape(x, w, None, c["pad"], c["s"]) == (1, 3, *c["out"])
    with pytest.raises(rtg.RtgError, match="Input channels does not match kernel input channels"):
        ops.conv_transpose(np.zeros((1, 4, 3, 3), np.float32), np.zeros((5, 2, 2, 2), np.float32))
    with pytest.raises(rtg.RtgError, match="expected 2 stride values"):
        ops.conv_transpose(np.zeros((1, 4, 3, 3), np.float32), np.zeros((4, 2, 2, 2), np.float32), strides=(1,))


def test_conv_transpose_in_a_model(rtg, ctx, oracle):
    """ConvTranspose through the OpRegistry / Graph::run path (attrs from the .rten file), after a fused Conv."""
    from rten_fork_b200.model import Model
    from rten_fork_b200.rten_builder import ModelBuilder
    import graph_oracle

    rng = oracle.XorShiftRng(1234)
    b = ModelBuilder()
    x = b.add_value("x", ["batch", 4, 6, 6])
    b.inputs = [x]
    y = b.op("Conv", [x, b.add_constant("w0", rng.fill((8, 4, 3, 3)) - 0.5), b.add_constant("b0", rng.fill((8,)))],
             ModelBuilder.conv_attrs((1, 1, 1, 1)))
    y = b.op("Relu", [y])
    y = b.op("ConvTranspose", [y, b.add_constant("w1", rng.fill((8, 5, 2, 2))), b.add_constant("b1", rng.fill((5,)))],
             ModelBuilder.conv_transpose_attrs((0, 0, 0, 0), (2, 2)))
    b.outputs = [y]
    xin = rng.fill((2, 4, 6, 6))
    got = Model.load(b.build(), ctx=ctx).run([xin])
    want = graph_oracle.run_graph(b, [xin])
    assert got[0].shape == (2, 5, 12, 12)
    expect_equal(got[0], want[0], rtol=1e-5, atol=1e-5)


DW_CASES = [
    # (N, C, H, W, stride, pads (t, l, b, r), activation, bias)
    (1, 4, 1, 1, 1, (1, 1, 1, 1), "none", True),
    (2, 20, 2, 5, 1, (1, 1, 1, 1), "relu", True),
    (1, 36, 7, 7, 1, (1, 1, 1, 1), "clip", True),
    (3, 100, 9, 33, 1, (0, 1, 2, 1), "clip", False),      # ragged: last strip partial, asymmetric padding
    (1, 144, 14, 14, 2, (1, 1, 1, 1), "clip", True),      # five 32-channel blocks, the last one half empty
    (2, 260, 30, 65, 1, (1, 1, 1, 1), "none", True),      # two column bands, 65 = 64 + 1
    (1, 32, 57, 130, 2, (0, 0, 1, 1), "relu", True),      # stride 2, odd sizes, no top / left padding
    (1, 96, 112, 112, 2, (1, 1, 1, 1), "clip", True),     # MobileNetV2's first stride-2 layer (several row bands)
    (2, 960, 7, 7, 1, (1, 1, 1, 1), "clip", True),        # 256-channel blocks, 8-column band
    (1, 8, 40, 3, 1, (2, 2, 2, 2), "none", False),        # padding wider than the kernel's reach
    (1, 64, 5, 300, 2, (1, 0, 0, 2), "clip", True),       # many column bands, stride 2
    (2, 144, 56, 56, 1, (1, 1, 1, 1), "clip", True),      # MobileNetV2's 144-channel layers: 4.5 channel blocks
    (1, 144, 56, 56, 2, (1, 1, 1, 1), "clip", True),
    (1, 180, 9, 120, 1, (1, 1, 1, 1), "relu", True),      # two column bands, the last channel block partly empty
]


@pytest.mark.parametrize("n,c,h,w,s,pads,act,with_bias", DW_CASES)
def test_depthwise3x3_tma_is_bit_exact(rtg, ctx, oracle, n, c, h, w, s, pads, act, with_bias):
    """The TMA-staged depthwise kernel (csrc/depthwise_tma.cuh) keeps the reference's summation order (bias, taps in
    (ky, kx) order, separate multiply and add, padding taps skipped: src/ops/conv/depthwise.rs:49-120): its results
    equal the oracle's bit for bit over band / block / row-band edges, asymmetric paddings and both strides."""
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(c * 1000 + h * 10 + w)
    x = rng.fill((n, c, h, w)) - 0.5
    wt = rng.fill((c, 1, 3, 3)) - 0.5
    b = rng.fill((c,)) - 0.5 if with_bias else None
    oh = (h + pads[0] + pads[2] - 3) // s + 1
    ow = (w + pads[1] + pads[3] - 3) // s + 1
    if oh <= 0 or ow <= 0:
        pytest.skip("empty output")
    got = ops.conv(x, wt, b, padding=pads, groups=c, strides=(s, s), activation=act, clip=(-0.25, 0.5), ctx=ctx)
    assert ctx.last_kernel.startswith("depthwise3x3_tma"), ctx.last_kernel
    want = oracle.conv2d(x, wt, b, pads, c, (s, s), (1, 1))
    if act == "relu":
        want = oracle.unary("relu", want)
    elif act == "clip":
        want = oracle.unary("clip", want, -0.25, 0.5)
    assert got.shape == want.shape == (n, c, oh, ow)
    np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("s", [1, 2])
def test_depthwise3x3_padding_taps_are_skipped_not_added(rtg, ctx, oracle, s):
    """The kernel may add the zero-filled padding instead of skipping the tap only where that cannot change a bit.  A
    channel with an infinite weight (inf x 0 would be NaN), one with a -0.0 bias over an all-zero input (-0.0 + 0.0 would
    be +0.0) and a Clip whose upper bound is -0.0 all take the predicated path: bit for bit the oracle's result
    (src/ops/conv/depthwise.rs:60-110 skips out-of-image taps)."""
    from rten_fork_b200 import ops

    n, c, h, w = 2, 40, 9, 21
    rng = oracle.XorShiftRng(77 + s)
    x = rng.fill((n, c, h, w)) - 0.5
    wt = rng.fill((c, 1, 3, 3)) - 0.5
    b = rng.fill((c,)) - 0.5
    wt[3, 0, 0, 0] = np.inf          # top-left tap: in the padding for the first row / column only
    wt[17, 0, 2, 2] = -np.inf
    b[5] = -0.0
    x[:, 5] = 0.0                    # every product of channel 5 is +-0.0
    wt[5] = -np.abs(wt[5])           # ... and -0.0 wherever the input is +0.0
    b[33] = -0.0
    for act, clip in (("none", (0.0, 0.0)), ("clip", (-0.25, 0.5)), ("clip", (-1.0, -0.0))):
        got = ops.conv(x, wt, b, padding=(1, 1, 1, 1), groups=c, strides=(s, s), activation=act, clip=clip, ctx=ctx)
        assert ctx.last_kernel.startswith("depthwise3x3_tma"), ctx.last_kernel
        want = oracle.conv2d(x, wt, b, (1, 1, 1, 1), c, (s, s), (1, 1))
        if act == "clip":
            want = oracle.unary("clip", want, clip[0], clip[1])
        assert got.tobytes() == want.tobytes(), (act, clip)

