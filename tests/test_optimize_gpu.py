"""GraphOptimizer parity (row a15): the reference's own optimizer tests (src/optimize.rs:546-728) restated on `.rten`
buffers written by the ModelBuilder mirror — constant propagation, fuse_transpose, fuse_silu, fuse_gelu,
fuse_layer_norm — plus Model::run addressed by node id and Model::partial_run (src/model.rs:580-646,
src/graph.rs:1147-1235, tests src/graph.rs:1965-2100)."""
import numpy as np
import pytest

from conftest import expect_equal

pytestmark = pytest.mark.gpu


def _load(ctx, b, optimize=True):
    from rten_fork_b200.model import Model

    return Model.load(b.build(), optimize=optimize, ctx=ctx)


def test_constant_propagation(rtg, ctx):
    """src/optimize.rs:546-600: add_1 = const_a + const_b is replaced by the constant [5, 7, 9]; add_2 consumes it."""
    from rten_fork_b200.rten_builder import ModelBuilder

    b = ModelBuilder()
    ca = b.add_constant("const_a", np.array([1, 2, 3], np.int32))
    cb = b.add_constant("const_b", np.array([4, 5, 6], np.int32))
    add_out = b.op("Add", [ca, cb], name="add_1")
    x = b.add_value("input", [3])
    add2_out = b.op("Add", [add_out, x], name="add_2")
    b.inputs, b.outputs = [x], [add_out, add2_out]
    m = _load(ctx, b)
    assert m.node_kind(m.find_node("add_1_out")) == "constant"       # the value became a constant node
    assert m.node_kind(m.find_node("add_2_out")) == "value"
    got = m.run([np.array([10, 20, 30], np.int32)])
    assert got[0].tolist() == [5, 7, 9] and got[1].tolist() == [15, 27, 39]
    unopt = _load(ctx, b, optimize=False)
    assert unopt.node_kind(unopt.find_node("add_1_out")) == "value"
    got2 = unopt.run([np.array([10, 20, 30], np.int32)])
    assert got2[0].tolist() == [5, 7, 9] and got2[1].tolist() == [15, 27, 39]


def test_constant_propagation_feeds_pattern_fusion(rtg, ctx, oracle):
    """A Gelu whose sqrt(2) is itself computed from constants (Sqrt(2.0)) only matches once constants have been
    propagated: the order of GraphOptimizer::optimize (src/optimize.rs:286-297)."""
    from rten_fork_b200.rten_builder import ModelBuilder

    b = ModelBuilder()
    x = b.add_value("x", ["n", 32])
    b.inputs = [x]
    two = b.add_constant("two", np.array(2.0, np.float32))
    sqrt2 = b.op("Sqrt", [two], name="sqrt2")
    d = b.op("Div", [x, sqrt2], name="div")
    e = b.op("Erf", [d], name="erf")
    a = b.op("Add", [e, b.add_constant("one", np.array(1.0, np.float32))], name="add")
    mu = b.op("Mul", [x, a], name="mul")
    y = b.op("Mul", [mu, b.add_constant("half", np.array(0.5, np.float32))], name="mul_half")
    b.outputs = [y]
    m = _load(ctx, b)
    assert m.source_op(m.output_ids[0]) == ("Gelu", "mul_half")
    xin = oracle.XorShiftRng(5).fill((4, 32)) * 4 - 2
    expect_equal(m.run([xin])[0], oracle.unary("gelu", xin), rtol=1e-6, atol=1e-7)


def test_fuse_transpose(rtg, ctx, oracle):
    """src/optimize.rs:602-621: Transpose -> MatMul becomes one fused operator that keeps the MatMul's node name."""
    from rten_fork_b200.rten_builder import ModelBuilder

    b = ModelBuilder()
    i1, i2 = b.add_value("a", [5, 3]), b.add_value("b", [5, 4])
    t = b.op("Transpose", [i1], ("TransposeAttrs", [(0, "vec_u32", None)]), name="transpose")   # perm: None
    y = b.op("MatMul", [t, i2], name="matmul")
    b.inputs, b.outputs = [i1, i2], [y]
    m = _load(ctx, b)
    op, node = m.source_op(m.output_ids[0])
    assert op == "FusedMatMul" and node == "matmul"   # the reference prints "FusedTranspose(MatMul)"
    rng = oracle.XorShiftRng(1234)
    a, bb = rng.fill((5, 3)), rng.fill((5, 4))
    expect_equal(m.run([a, bb])[0], oracle.gemm(np.ascontiguousarray(a.T), bb), rtol=1e-5)


def test_fuse_silu(rtg, ctx, oracle):
    """src/optimize.rs:623-640: x * Sigmoid(x) => Silu(x) on the Mul's node; bit-identical to the unfused pair."""
    from rten_fork_b200.rten_builder import ModelBuilder

    for order in (0, 1):
        b = ModelBuilder()
        x = b.add_value("x", ["n", 40])
        s = b.op("Sigmoid", [x], name="sigmoid")
        y = b.op("Mul", [x, s] if order == 0 else [s, x], name="mul")
        b.inputs, b.outputs = [x], [y]
        m = _load(ctx, b)
        assert m.source_op(m.output_ids[0]) == ("Silu", "mul")
        assert m.num_ops == 1
        xin = oracle.XorShiftRng(7).fill((6, 40)) * 8 - 4
        fused = m.run([xin])[0]
        unfused = _load(ctx, b, optimize=False).run([xin])[0]
        assert fused.tobytes() == unfused.tobytes()
        assert fused.tobytes() == oracle.unary("silu", xin).tobytes()


def test_fuse_silu_into_conv_epilogue(rtg, ctx, oracle, math_mode):
    """YOLO-style Conv -> x * Sigmoid(x): the Silu the optimiser creates folds into the convolution's epilogue."""
    import graph_oracle
    from rten_fork_b200.rten_builder import ModelBuilder

    rng = oracle.XorShiftRng(1234)
    b = ModelBuilder()
    x = b.add_value("x", ["n", 8, 10, 10])
    w = b.add_constant("w", rng.fill((16, 8, 3, 3)) - 0.5)
    bias = b.add_constant("b", rng.fill((16,)))
    c = b.op("Conv", [x, w, bias], ModelBuilder.conv_attrs((1, 1, 1, 1)), name="conv")
    s = b.op("Sigmoid", [c], name="sigmoid")
    y = b.op("Mul", [c, s], name="mul")
    b.inputs, b.outputs = [x], [y]
    m = _load(ctx, b)
    assert m.num_ops == 1 and m.source_op(m.output_ids[0])[0] == "FusedConv"
    xin = rng.fill((2, 8, 10, 10))
    want = graph_oracle.run_graph(b, [xin])[0]
    # mixed-sign conv outputs pass through x * sigmoid(x): compare against the scale of the tensor
    np.testing.assert_allclose(m.run([xin])[0], want, rtol=math_mode["rtol"], atol=math_mode["rtol"] * np.abs(want).max())


def test_fuse_gelu(rtg, ctx):
    """src/optimize.rs:642-664."""
    from rten_fork_b200.rten_builder import ModelBuilder

    b = ModelBuilder()
    x = b.add_value("x", ["n", 16])
    d = b.op("Div", [x, b.add_constant("sqrt2", np.array(np.sqrt(2.0), np.float32))], name="div")
    e = b.op("Erf", [d], name="erf")
    a = b.op("Add", [e, b.add_constant("one", np.array(1.0, np.float32))], name="add")
    mu = b.op("Mul", [x, a], name="mul")
    y = b.op("Mul", [mu, b.add_constant("half", np.array(0.5, np.float32))], name="mul_half")
    b.inputs, b.outputs = [x], [y]
    m = _load(ctx, b)
    assert m.source_op(m.output_ids[0]) == ("Gelu", "mul_half")


def _layer_norm_graph(keep_dims=True, eps_first=True):
    """src/optimize.rs:666-703 layer_norm_graph (the reference's graph says keep_dims=false, which its pattern never
    checks; the executable variant keeps the dims so that the primitives broadcast)."""
    from rten_fork_b200.rten_builder import ModelBuilder

    b = ModelBuilder()
    x = b.add_value("x", ["n", 5, 3])
    mean = b.op("ReduceMean", [x], ModelBuilder.reduce_mean_attrs([-1], keep_dims), name="mean")
    sub = b.op("Sub", [x, mean], name="sub")
    pw = b.op("Pow", [sub, b.add_constant("two", np.array(2.0, np.float32))], name="pow")
    var = b.op("ReduceMean", [pw], ModelBuilder.reduce_mean_attrs([-1], keep_dims), name="var_mean")
    eps = b.add_constant("eps", np.array(1e-6, np.float32))
    add_eps = b.op("Add", [eps, var] if eps_first else [var, eps], name="add_eps")
    sq = b.op("Sqrt", [add_eps], name="sqrt")
    div = b.op("Div", [sub, sq], name="div")
    bias = b.add_constant("bias", np.array([1.0, 2.0, 3.0], np.float32))
    scale = b.add_constant("scale", np.array([3.0, 4.0, 5.0], np.float32))
    mul = b.op("Mul", [div, scale], name="mul")
    y = b.op("Add", [mul, bias], name="final_add")
    b.inputs, b.outputs = [x], [y]
    return b


@pytest.mark.parametrize("eps_first", [True, False])
def test_fuse_layer_norm(rtg, ctx, oracle, eps_first):
    """src/optimize.rs:705-715: the nine primitives collapse into LayerNormalization{axis=-1, epsilon=1e-6} on the
    final Add's node; the fused operator's result agrees with the oracle's evaluation of the primitives."""
    import graph_oracle

    b = _layer_norm_graph(True, eps_first)
    m = _load(ctx, b)
    assert m.source_op(m.output_ids[0]) == ("LayerNormalization", "final_add")
    assert m.num_ops == 1
    assert _load(ctx, b, optimize=False).num_ops == 9
    xin = oracle.XorShiftRng(11).fill((4, 5, 3)) * 6 - 3
    want = graph_oracle.run_graph(b, [xin])[0]
    np.testing.assert_allclose(m.run([xin])[0], want, rtol=2e-5, atol=2e-5)
    np.testing.assert_allclose(_load(ctx, b, optimize=False).run([xin])[0], want, rtol=2e-5, atol=2e-5)
    # the reference's own (keep_dims = false) graph is matched by structure alone
    m2 = _load(ctx, _layer_norm_graph(False, eps_first))
    assert m2.source_op(m2.output_ids[0]) == ("LayerNormalization", "final_add")


def test_layer_norm_not_fused_over_other_axes(rtg, ctx):
    """Only means over axes == [-1] are fused (src/optimize.rs:455-478)."""
    from rten_fork_b200.rten_builder import ModelBuilder

    b = _layer_norm_graph()
    for nd in b.nodes:
        if nd["kind"] == "op" and nd["name"] == "var_mean":
            nd["attrs"] = ModelBuilder.reduce_mean_attrs([1], True)
    m = _load(ctx, b)
    assert m.source_op(m.output_ids[0])[0] == "Add"


def test_optimize_error(rtg, ctx):
    """src/optimize.rs:717-727: a graph whose output id names no node fails to load."""
    from rten_fork_b200.model import Model
    from rten_fork_b200.rten_builder import ModelBuilder

    b = ModelBuilder()
    b.add_value("x", [1])
    b.inputs, b.outputs = [123], [123]
    with pytest.raises(rtg.RtgError):
        Model.load(b.build(), ctx=ctx)


# ------------------------------------------------------------------------------------------- run by node id / partial run
def _two_branch_graph():
    """out = Relu(a + c0) * Relu(b + c1): each branch depends on one input (src/graph.rs test_partial_run shape)."""
    from rten_fork_b200.rten_builder import ModelBuilder

    m = ModelBuilder()
    a, b = m.add_value("a", [4]), m.add_value("b", [4])
    c0 = m.add_constant("c0", np.array([1, -2, 3, -4], np.float32))
    c1 = m.add_constant("c1", np.array([0.5, 0.5, -8, 2], np.float32))
    ra = m.op("Relu", [m.op("Add", [a, c0], name="add_a")], name="relu_a")
    rb = m.op("Relu", [m.op("Add", [b, c1], name="add_b")], name="relu_b")
    y = m.op("Mul", [ra, rb], name="mul")
    m.inputs, m.outputs = [a, b], [y]
    return m


def test_run_by_node_id_and_intermediate_outputs(rtg, ctx):
    """Model::run takes (node id, value) pairs and any value node as an output (src/model.rs:580-592)."""
    m = _load(ctx, _two_branch_graph(), optimize=False)
    a, b = np.array([1, 1, 1, 1], np.float32), np.array([2, 2, 2, 2], np.float32)
    ia, ib = m.find_node("a"), m.find_node("b")
    assert m.input_ids == [ia, ib]
    ra, y = m.run([(ib, b), (ia, a)], outputs=[m.find_node("relu_a_out"), m.output_ids[0]])   # any input order
    assert ra.tolist() == [2, 0, 4, 0] and y.tolist() == [5, 0, 0, 0]
    # an intermediate value can stand in for the inputs it depends on
    y2 = m.run([(m.find_node("relu_a_out"), np.array([1, 2, 3, 4], np.float32)), (ib, b)], outputs=m.output_ids)[0]
    assert y2.tolist() == [2.5, 5.0, 0.0, 16.0]
    with pytest.raises(rtg.RtgError, match="Missing input"):
        m.run([(ia, a)], outputs=m.output_ids)
    with pytest.raises(rtg.RtgError, match="not unique"):
        m.run([(ia, a), (ia, a)], outputs=m.output_ids)
    with pytest.raises(rtg.RtgError, match="Output node ID is invalid"):
        m.run([(ia, a), (ib, b)], outputs=[10_000])


def test_partial_run(rtg, ctx):
    """Model::partial_run (src/model.rs:639-646): with only `a`, branch a is evaluated and its leaf returned; feeding
    the leaf back with `b` gives the full result."""
    m = _load(ctx, _two_branch_graph(), optimize=False)
    a, b = np.array([1, 1, 1, 1], np.float32), np.array([2, 2, 2, 2], np.float32)
    ia, ib = m.find_node("a"), m.find_node("b")
    leaves = m.partial_run([(ia, a)], outputs=m.output_ids)
    assert [nid for nid, _ in leaves] == [m.find_node("relu_a_out")]
    assert leaves[0][1].tolist() == [2, 0, 4, 0]
    y = m.run(leaves + [(ib, b)], outputs=m.output_ids)[0]
    assert y.tolist() == m.run([a, b])[0].tolist() == [5, 0, 0, 0]
    # all inputs given: the leaf is the requested output itself; none given: nothing depends on constants only
    full = m.partial_run([(ia, a), (ib, b)], outputs=m.output_ids)
    assert [nid for nid, _ in full] == m.output_ids and full[0][1].tolist() == [5, 0, 0, 0]
    assert m.partial_run([], outputs=m.output_ids) == []


# ------------------------------------------------------------------------------------- epilogue fusions stay legal
@pytest.mark.parametrize("case", ["broadcast_residual", "depthwise_residual", "matmul_scalar_vector"])
def test_epilogue_fusion_falls_back_when_not_an_epilogue(rtg, ctx, oracle, case):
    """The load-time Conv -> Add / MatMul -> Add fusions only become epilogues when the operand really is one (same
    shape as the output; a bias of length N).  A broadcast operand, a depthwise layer or a shape-[1] constant run as the
    reference runs them: the operator, then Add with broadcasting, then the activation."""
    import graph_oracle
    from rten_fork_b200.rten_builder import ModelBuilder

    rng = oracle.XorShiftRng(1234)
    b = ModelBuilder()
    if case == "matmul_scalar_vector":
        x = b.add_value("x", ["n", 12])
        w = b.add_constant("w", rng.fill((12, 8)) - 0.5)
        y = b.op("MatMul", [x, w], name="matmul")
        y = b.op("Add", [y, b.add_constant("c", np.array([0.25], np.float32))], name="add")
        y = b.op("Relu", [y], name="relu")
        b.inputs, b.outputs = [x], [y]
        ins = [rng.fill((5, 12)) - 0.5]
    else:
        c = 8
        x = b.add_value("x", ["n", c, 9, 9])
        e = b.add_value("e", ["n", c, 1, 1] if case == "broadcast_residual" else ["n", c, 9, 9])
        groups = c if case == "depthwise_residual" else 1
        w = b.add_constant("w", rng.fill((c, c // groups, 3, 3)) - 0.5)
        bias = b.add_constant("b", rng.fill((c,)))
        y = b.op("Conv", [x, w, bias], ModelBuilder.conv_attrs((1, 1, 1, 1), groups), name="conv")
        y = b.op("Add", [y, e], name="add")
        y = b.op("Relu", [y], name="relu")
        b.inputs, b.outputs = [x, e], [y]
        ins = [rng.fill((2, c, 9, 9)) - 0.5, rng.fill((2, c, 1, 1) if case == "broadcast_residual" else (2, c, 9, 9)) - 0.5]
    m = _load(ctx, b)
    want = graph_oracle.run_graph(b, ins)[0]
    got = m.run(ins)[0]
    assert got.shape == want.shape
    expect_equal(got, want, rtol=1e-5, atol=2e-6)
    expect_equal(_load(ctx, b, optimize=False).run(ins)[0], want, rtol=1e-5, atol=2e-6)


@pytest.mark.parametrize("res_shape", ["same", "broadcast"])
@pytest.mark.parametrize("with_bias", [True, False])
def test_matmul_residual_rides_in_the_epilogue(rtg, ctx, res_shape, with_bias):
    """MatMul [-> Add(bias)] -> Add(activation) [-> Relu]: the residual connection of a transformer block is fused when it
    has the output's shape ((acc + bias) + residual in the epilogue, the order of the unfused chain, so optimised and
    unoptimised runs agree bit for bit in every math mode) and is added after the product when it only broadcasts."""
    from rten_fork_b200.rten_builder import ModelBuilder

    rng = np.random.default_rng(3)
    b = ModelBuilder()
    x = b.add_value("x", ["b", 24, 64])
    r = b.add_value("r", ["b", 24, 96] if res_shape == "same" else [1, 24, 96])
    b.inputs = [x, r]
    w_arr = rng.random((64, 96), dtype=np.float32) - 0.5
    bias_arr = rng.random(96, dtype=np.float32)
    w = b.add_constant("w", w_arr)
    y = b.op("MatMul", [x, w], name="dense")
    if with_bias:
        y = b.op("Add", [y, b.add_constant("bias", bias_arr)], name="bias_add")
    y = b.op("Add", [r, y], name="residual_add")
    y = b.op("Relu", [y], name="act")
    b.outputs = [y]
    xin = rng.random((3, 24, 64), dtype=np.float32) - 0.5
    rin = rng.random((3, 24, 96) if res_shape == "same" else (1, 24, 96), dtype=np.float32) - 0.5
    fused, plain = _load(ctx, b), _load(ctx, b, optimize=False)
    assert fused.source_op(fused.output_ids[0]) == ("FusedMatMul", "dense")      # the whole chain is one operator
    assert fused.num_ops == 1 and plain.num_ops == (4 if with_bias else 3)
    got, want = fused.run([xin, rin])[0], plain.run([xin, rin])[0]
    assert got.shape == (3, 24, 96)
    np.testing.assert_array_equal(got, want)
    ref = np.maximum(xin.astype(np.float64) @ w_arr.astype(np.float64) + (bias_arr if with_bias else 0) + rin, 0)
    np.testing.assert_allclose(got, ref, rtol=2e-3, atol=2e-3)


@pytest.mark.parametrize("case", ["mask_b11s", "mask_first", "full_shape", "vector", "general_broadcast", "ragged_len",
                                  "other_axis"])
def test_add_softmax_is_one_kernel_with_the_same_bits(rtg, ctx, case):
    """Add -> Softmax (BERT's attention mask): the add happens inside the Softmax kernel when the addend is [leading dims,
    1, ..., 1, len] and the axis is the last one; every other form (general broadcast, row length not a multiple of four,
    another axis) runs the two steps one after the other inside the same operator.  Optimised and unoptimised runs agree
    bit for bit in every case, and both follow the oracle's Softmax (rten-vecmath/src/softmax.rs)."""
    from rten_fork_b200.rten_builder import ModelBuilder
    import graph_oracle

    rng = np.random.default_rng(11)
    L = 10 if case == "ragged_len" else 16
    xs = (2, 3, 8, L)
    ms = {"mask_b11s": (2, 1, 1, L), "mask_first": (2, 1, 1, L), "full_shape": xs, "vector": (L,),
          "general_broadcast": (3, 1, L), "ragged_len": (2, 1, 1, L), "other_axis": (2, 1, 1, L)}[case]
    b = ModelBuilder()
    x = b.add_value("x", list(xs))
    m = b.add_value("m", list(ms))
    b.inputs = [x, m]
    s = b.op("Add", [m, x] if case == "mask_first" else [x, m], name="mask_add")
    y = b.op("Softmax", [s], ("SoftmaxAttrs", [(0, "i32", 1 if case == "other_axis" else -1)]), name="softmax")
    b.outputs = [y]
    xin = (rng.random(xs, dtype=np.float32) - 0.5) * 8
    min_ = np.where(rng.random(ms) < 0.3, np.float32(-10000.0), np.float32(0.0)).astype(np.float32)
    fused, plain = _load(ctx, b), _load(ctx, b, optimize=False)
    assert fused.num_ops == 1 and plain.num_ops == 2
    got, want = fused.run([xin, min_])[0], plain.run([xin, min_])[0]
    assert got.shape == xs
    np.testing.assert_array_equal(got, want)
    expect_equal(got, graph_oracle.run_graph(b, [xin, min_])[0], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("scale", [1.0, 12.0])
def test_gelu_in_the_matmul_epilogue_has_the_standalone_kernels_bits(rtg, ctx, scale):
    """MatMul -> Add(bias) -> Gelu: fused, the Gelu runs in the contraction kernel's epilogue on packed fp32 pairs
    (vm_gelu_x2: FFMA2 / FMUL2, branch-free reciprocal); unfused, in the elementwise kernel with the scalar formula.  Same
    IEEE operations lane by lane, so the two runs agree bit for bit — over pre-activations from tiny to far in both tails
    (|x| > 10.2, where exp(-x^2) underflows), and both follow rten-vecmath's gelu (rten-vecmath/src/erf.rs:82-91)."""
    from rten_fork_b200.rten_builder import ModelBuilder

    rng = np.random.default_rng(17)
    b = ModelBuilder()
    x = b.add_value("x", ["b", 40, 64])
    b.inputs = [x]
    w_arr = ((rng.random((64, 160), dtype=np.float32) - 0.5) * scale).astype(np.float32)
    bias_arr = ((rng.random(160, dtype=np.float32) - 0.5) * scale).astype(np.float32)
    y = b.op("MatMul", [x, b.add_constant("w", w_arr)], name="dense")
    y = b.op("Add", [y, b.add_constant("bias", bias_arr)], name="bias_add")
    y = b.op("Gelu", [y], name="act")
    b.outputs = [y]
    xin = (rng.random((3, 40, 64), dtype=np.float32) - 0.5) * 2
    fused, plain = _load(ctx, b), _load(ctx, b, optimize=False)
    assert fused.num_ops == 1 and plain.num_ops == 3
    got, want = fused.run([xin])[0], plain.run([xin])[0]
    pre = xin.astype(np.float64) @ w_arr.astype(np.float64) + bias_arr
    if scale > 1.0:
        assert np.abs(pre).max() > 10.2 and (pre < -10.2).any()      # both saturated tails are exercised
    np.testing.assert_array_equal(got, want)
    from math import erf
    ref = 0.5 * pre * (1.0 + np.vectorize(erf)(pre / np.sqrt(2.0)))
    np.testing.assert_allclose(got, ref, rtol=2e-3, atol=2e-3 * max(1.0, scale))

