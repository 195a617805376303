"""Model / Graph parity: a `.rten` file written by the ModelBuilder mirror is loaded by the C++ runtime and run on the
GPU; results are compared with the oracle's CPU evaluation of the same graph (reference: src/model.rs tests
:912-952 test_load_and_run_model builds models in V1 and V2 formats and runs them)."""
import numpy as np
import pytest

from conftest import expect_equal

pytestmark = pytest.mark.gpu


def _run_both(ctx, builder, inputs, optimize=True, version=2):
    import graph_oracle
    from rten_fork_b200.model import Model

    m = Model.load(builder.build(version=version), optimize=optimize, ctx=ctx)
    got = m.run(inputs)
    want = graph_oracle.run_graph(builder, inputs)
    return m, got, want


@pytest.mark.parametrize("version", [1, 2])
@pytest.mark.parametrize("optimize", [False, True])
def test_small_conv_model_formats(rtg, ctx, oracle, version, optimize):
    """V1 (headerless, inline constants) and V2 (header + external tensor data) containers, optimiser on / off."""
    from rten_fork_b200.rten_builder import ModelBuilder

    rng = oracle.XorShiftRng(1234)
    b = ModelBuilder()
    x = b.add_value("input", ["batch", 4, 9, 9])
    b.inputs = [x]
    w = b.add_constant("w", rng.fill((8, 4, 3, 3)) - 0.5)
    bias = b.add_constant("b", rng.fill((8,)))
    y = b.op("Conv", [x, w, bias], ModelBuilder.conv_attrs((1, 1, 1, 1)))
    y = b.op("Relu", [y])
    y = b.op("MaxPool", [y], ModelBuilder.pool_attrs("MaxPool", (2, 2), (2, 2)))
    b.outputs = [y]
    m, got, want = _run_both(ctx, b, [rng.fill((2, 4, 9, 9))], optimize, version)
    assert got[0].shape == (2, 8, 4, 4)
    expect_equal(got[0], want[0], rtol=2e-3 if ctx.math == rtg.MATH_TF32 else 1e-5, atol=1e-5)
    assert m.inputs[0]["name"] == "input" and m.inputs[0]["shape"] == [None, 4, 9, 9]
    assert m.num_ops == (2 if optimize else 3)  # Conv+Relu fused by the optimiser


@pytest.mark.parametrize("optimize", [False, True])
def test_resnet50_reduced(rtg, ctx, oracle, math_mode, optimize):
    """ResNet-50 topology (all 53 convs, bottlenecks, downsample branches) at width 16 / 64x64 input, batch 2."""
    from rten_fork_b200 import synth

    b = synth.resnet50(image=64, width=16, classes=100, seed=1)
    rng = oracle.XorShiftRng(1234)
    x = rng.fill((2, 3, 64, 64))
    m, got, want = _run_both(ctx, b, [x], optimize)
    assert got[0].shape == (2, 100)
    # logits mix signs, so the comparison is against the scale of the output; the reference's per-element rule is
    # applied to non-negative intermediate activations below and in test_conv_gpu.py
    tol = math_mode["rtol"] * 20
    np.testing.assert_allclose(got[0], want[0], rtol=tol, atol=tol * np.abs(want[0]).max())


def test_model_intermediate_activation_parity(rtg, ctx, oracle):
    """Post-ReLU activations of the first bottleneck block in 3xTF32 against the oracle."""
    import graph_oracle
    from rten_fork_b200 import synth
    from rten_fork_b200.model import Model

    b = synth.resnet50(image=32, width=16, blocks=(1, 1, 1, 1), classes=10, seed=2)
    b.outputs = [nd["outputs"][0] for nd in b.nodes if nd["kind"] == "op" and nd["name"] == "layer1.0.relu3"]
    assert len(b.outputs) == 1
    rng = oracle.XorShiftRng(1234)
    x = rng.fill((2, 3, 32, 32))
    prev = ctx.math
    ctx.set_math(rtg.MATH_3XTF32)
    try:
        got = Model.load(b.build(), optimize=True, ctx=ctx).run([x])
    finally:
        ctx.set_math(prev)
    want = graph_oracle.run_graph(b, [x])
    err = np.abs(got[0] - want[0])
    assert (err <= 1e-5 * np.abs(want[0]) + 2e-6 * np.abs(want[0]).max()).all(), err.max()


def test_mobilenet_v2_reduced(rtg, ctx, oracle, math_mode):
    from rten_fork_b200 import synth

    b = synth.mobilenet_v2(image=64, width_mult=0.25, classes=50, seed=3)
    rng = oracle.XorShiftRng(1234)
    x = rng.fill((2, 3, 64, 64))
    m, got, want = _run_both(ctx, b, [x], True)
    tol = math_mode["rtol"] * 20
    np.testing.assert_allclose(got[0], want[0], rtol=tol, atol=tol * np.abs(want[0]).max())


def test_model_errors(rtg, ctx):
    from rten_fork_b200.model import Model
    from rten_fork_b200.rten_builder import ModelBuilder

    with pytest.raises(rtg.RtgError) as e:
        Model.load(b"RTEN" + b"\x03\x00\x00\x00" + b"\x00" * 24, ctx=ctx)   # HeaderError::UnsupportedVersion
    assert e.value.kind == "Format"
    with pytest.raises(rtg.RtgError):
        Model.load(b"\x00" * 3, ctx=ctx)
    b = ModelBuilder()
    x = b.add_value("x", [1, 3])
    b.inputs = [x]
    b.outputs = [b.op("Cos", [x])]            # not in this registry => ReadOpError::OperatorUnavailable
    with pytest.raises(rtg.RtgError, match="operator Cos is not available"):
        Model.load(b.build(), ctx=ctx)
    # operator errors carry the node name (RunError::OperatorError, src/graph.rs:975-978)
    b = ModelBuilder()
    x = b.add_value("x", [1, 1, 2, 2])
    b.inputs = [x]
    w = b.add_constant("w", np.zeros((1, 1, 3, 3), np.float32))
    b.outputs = [b.op("Conv", [x, w], ModelBuilder.conv_attrs(), name="bad_conv")]
    m = Model.load(b.build(), ctx=ctx)
    with pytest.raises(rtg.RtgError, match="bad_conv.*Input too small for kernel size") as e:
        m.run([np.zeros((1, 1, 2, 2), np.float32)])
    assert e.value.kind == "InvalidValue"


def test_graph_in_place_and_pool_reuse(rtg, oracle):
    """Unary / binary operators run in place on owned temporaries and buffers come back through the pool
    (src/graph.rs:897-931, :1013-1037): a chain of elementwise ops allocates a bounded number of buffers."""
    from rten_fork_b200.model import Model
    from rten_fork_b200.rten_builder import ModelBuilder

    b = ModelBuilder()
    x = b.add_value("x", ["n", 1024])
    b.inputs = [x]
    y = x
    for i in range(8):
        y = b.op("Relu", [y])
        y = b.op("Add", [y, b.add_constant(f"c{i}", np.full((1024,), 0.5, np.float32))])
    b.outputs = [y]
    c = rtg.Context(0)
    m = Model.load(b.build(), ctx=c)
    xin = oracle.XorShiftRng(1234).fill((64, 1024)) - 0.5
    out = m.run([xin])[0]
    want = xin
    for i in range(8):
        want = np.maximum(want, 0) + np.float32(0.5)
    assert out.tobytes() == want.astype(np.float32).tobytes()
    s0 = c.pool_stats()
    m.run([xin])
    s1 = c.pool_stats()
    assert s1["alloc_count"] - s0["alloc_count"] <= 3   # the input upload (+ first op output); the rest is in place
    m.close()
    c.close()


def test_timing_report(rtg, ctx, oracle):
    from rten_fork_b200 import synth
    from rten_fork_b200.model import Model

    m = Model.load(synth.resnet50(image=32, width=8, blocks=(1, 1, 1, 1), classes=10).build(), ctx=ctx)
    m.set_timing(True)
    m.run([oracle.XorShiftRng(1).fill((1, 3, 32, 32))])
    rep = m.timing_report()
    assert "Conv" in rep and "Graph run of" in rep


@pytest.mark.parametrize("optimize", [False, True])
@pytest.mark.parametrize("divisor", [8.0, 3.0])
def test_fused_transpose_matmul_and_scale(rtg, ctx, oracle, math_mode, optimize, divisor):
    """Transpose -> MatMul runs on permuted views (GraphOptimizer::fuse_transpose, src/optimize.rs:333-377) and
    MatMul -> Div / Mul by a scalar folds into the epilogue; both must give what the unfused operators give."""
    from rten_fork_b200.rten_builder import ModelBuilder

    rng = oracle.XorShiftRng(99)
    b = ModelBuilder()
    q = b.add_value("q", [2, 20, 3, 8])   # [batch, seq, heads, d]
    k = b.add_value("k", [2, 20, 3, 8])
    b.inputs = [q, k]
    perm = lambda p: ("TransposeAttrs", [(0, "vec_u32", list(p))])
    qt = b.op("Transpose", [q], perm((0, 2, 1, 3)))
    kt = b.op("Transpose", [k], perm((0, 2, 3, 1)))
    s = b.op("MatMul", [qt, kt])
    s = b.op("Div", [s, b.add_constant("c", np.array(divisor, np.float32))])
    s = b.op("Mul", [b.add_constant("h", np.array(0.3, np.float32)), s])
    vt = b.op("Transpose", [k], perm((0, 2, 1, 3)))
    c = b.op("MatMul", [s, vt])
    c = b.op("Transpose", [c], perm((0, 2, 1, 3)))   # materialised: keeps the innermost dimension
    r = b.op("Transpose", [c], ("TransposeAttrs", []))  # no perm: full reversal, materialised
    b.outputs = [c, r]
    m, got, want = _run_both(ctx, b, [rng.fill((2, 20, 3, 8)), rng.fill((2, 20, 3, 8))], optimize)
    assert got[0].shape == (2, 20, 3, 8) and got[1].shape == (8, 3, 20, 2)
    for g, w in zip(got, want):
        expect_equal(g, w, rtol=math_mode["rtol"] * 2, atol=1e-6)
    assert got[1].tobytes() == np.ascontiguousarray(got[0].transpose()).tobytes()
    if optimize:
        assert m.num_ops == 5  # 2 MatMuls (+3 fused transposes, Div), the second Mul, 2 Transposes


def test_cli_runs_a_model_file(rtg, ctx, tmp_path, capsys):
    """`python -m rten_fork_b200.cli model.rten -n 2 -t -s dim0=3`: the rten-cli flow on the GPU path."""
    from rten_fork_b200 import cli, synth

    b = synth.resnet50(image=32, width=8, blocks=(1, 1, 1, 1), classes=10, seed=0)
    path = tmp_path / "tiny_resnet.rten"
    path.write_bytes(b.build())
    assert cli.main([str(path), "-n", "2", "-t", "-s", "dim0=3"]) == 0
    out = capsys.readouterr().out
    assert out.count("Model returned 1 outputs") == 2
    assert "shape: [3, 10]" in out and "FusedConv" in out


def test_shape_arithmetic_ops(rtg, ctx, oracle):
    """Concat / Slice / Shape / Expand as BERT-class exports use them (src/ops/concat.rs, slice.rs, layout.rs): the
    results are index-driven copies, compared bit-exactly with the oracle's evaluation of the same graph."""
    from rten_fork_b200.rten_builder import ModelBuilder

    rng = oracle.XorShiftRng(1234)
    b = ModelBuilder()
    x = b.add_value("x", ["batch", 6, 10])
    b.inputs = [x]
    i32 = lambda v: np.array(v, dtype=np.int32)  # noqa: E731
    # Slice with clamped / negative endpoints and a negative step on the last axis
    s1 = b.op("Slice", [x, b.add_constant("st", i32([1, -8])), b.add_constant("en", i32([1000, 9])),
                        b.add_constant("ax", i32([1, 2]))])
    s2 = b.op("Slice", [x, b.add_constant("st2", i32([-1])), b.add_constant("en2", i32([-100])),
                        b.add_constant("ax2", i32([2])), b.add_constant("sp2", i32([-3]))])
    # Shape -> Slice -> Concat: the usual "[batch, -1]" target of a Reshape, computed in the graph
    shp = b.op("Shape", [x])
    bdim = b.op("Slice", [shp, b.add_constant("z", i32([0])), b.add_constant("o", i32([1]))])
    target = b.op("Concat", [bdim, b.add_constant("m1", i32([-1]))], ModelBuilder.concat_attrs(0))
    flat = b.op("Reshape", [s1, target], ("ReshapeAttrs", []))
    cat = b.op("Concat", [x, x, b.op("Relu", [x])], ModelBuilder.concat_attrs(-1))
    row = b.add_constant("row", rng.fill((1, 1, 10)))
    ex = b.op("Expand", [row, shp])
    b.outputs = [s1, s2, shp, flat, cat, ex]
    xin = rng.fill((3, 6, 10)) - 0.5
    m, got, want = _run_both(ctx, b, [xin])
    assert got[0].shape == (3, 5, 7) and got[1].shape == (3, 6, 4) and got[3].shape == (3, 35)
    assert got[4].shape == (3, 6, 30) and got[5].shape == (3, 6, 10)
    for g, w in zip(got, want):
        assert g.dtype == w.dtype and g.shape == w.shape and g.tobytes() == w.tobytes()


def test_where_and_reduce_mean(rtg, ctx, oracle):
    """Where (broadcast select, bit-exact) and ReduceMean (src/ops/reduce.rs: sum / len over the given axes) as a
    mask-and-pool tail of a BERT-class export."""
    from rten_fork_b200.rten_builder import ModelBuilder

    rng = oracle.XorShiftRng(1234)
    b = ModelBuilder()
    x = b.add_value("x", ["batch", 5, 12])
    m = b.add_value("keep_mask", ["batch", 5, 1])
    b.inputs = [x, m]
    masked = b.op("Where", [m, x, b.add_constant("fill", np.array([-1.5], dtype=np.float32))])
    mean_last = b.op("ReduceMean", [masked], ModelBuilder.reduce_mean_attrs([-1], keep_dims=True))
    mean_mid = b.op("ReduceMean", [x], ModelBuilder.reduce_mean_attrs([1], keep_dims=False))
    mean_all = b.op("ReduceMean", [x], ModelBuilder.reduce_mean_attrs(None, keep_dims=False))
    b.outputs = [masked, mean_last, mean_mid, mean_all]
    xin = rng.fill((3, 5, 12)) - 0.5
    mask = (rng.fill((3, 5, 1)) > 0.4).astype(np.int32)
    _m, got, want = _run_both(ctx, b, [xin, mask])
    assert got[0].tobytes() == want[0].tobytes()
    assert got[1].shape == (3, 5, 1) and got[2].shape == (3, 12) and got[3].shape == ()
    for g, w in zip(got[1:], want[1:]):
        np.testing.assert_allclose(g, w, rtol=1e-5, atol=1e-6)


def test_run_async_pipelined_matches_run(rtg, ctx):
    """rtg_model_submit / rtg_ticket_wait: host inputs uploaded on a second stream into two staging slots, outputs
    written to the caller's host buffers; results must equal the synchronous Model.run bit for bit, whatever the number
    of submits in flight, and changing batch sizes must regrow the staging buffers."""
    import torch
    from rten_fork_b200 import synth

    b = synth.resnet50(image=32, width=8, blocks=(1, 1, 1, 1), classes=10, seed=0)
    model = rtg.Model.load(b.build(), ctx=ctx)
    gen = torch.Generator().manual_seed(7)
    xs = [torch.rand((n, 3, 32, 32), generator=gen).pin_memory() for n in (4, 4, 2, 8, 4, 4)]
    want = [model.run([x.numpy()])[0] for x in xs]
    outs = [torch.empty((8, 10)).pin_memory() for _ in xs]
    tickets = []
    got = []
    for x, o in zip(xs, outs):
        tickets.append(model.run_async([x], [o]))
        if len(tickets) == 2:  # two in flight
            got.append(tickets.pop(0).wait()[0])
    while tickets:
        got.append(tickets.pop(0).wait()[0])
    for g, w in zip(got, want):
        assert tuple(g.shape) == w.shape
        np.testing.assert_array_equal(g.numpy(), w)
    # numpy (pageable) buffers work too, and a too-small output buffer is an error, not an overrun
    o = np.empty((4, 10), np.float32)
    np.testing.assert_array_equal(model.run_async([xs[0].numpy()], [o]).wait()[0], want[0])
    with pytest.raises(rtg.RtgError, match="too small"):
        model.run_async([xs[3].numpy()], [np.empty((4, 10), np.float32)])
    np.testing.assert_array_equal(model.run_async([xs[1].numpy()], [o]).wait()[0], want[1])


EINSUM_CASES = [  # the equation forms of the reference's tests (src/ops/einsum.rs:740-960) plus an attention pair
    ("ij->ji", [(3, 5)]), ("ij,jk->ik", [(3, 5), (5, 4)]), ("ij,jk", [(3, 5), (5, 4)]), ("ji,kj->ik", [(5, 3), (4, 5)]),
    ("ij,jk->ki", [(3, 5), (5, 4)]), ("bhwc,hkc->bhwk", [(2, 3, 4, 6), (3, 5, 6)]), ("i,i->", [(7,), (7,)]),
    ("ij,j->i", [(3, 5), (5,)]), ("j,jk->k", [(5,), (5, 4)]), ("ij,ij->", [(3, 5), (3, 5)]), ("ij,j->", [(3, 5), (5,)]),
    ("i,j->ij", [(3,), (4,)]), ("bhqd,bhkd->bhqk", [(2, 12, 128, 64), (2, 12, 128, 64)]),
    ("bhqk,bhkd->bqhd", [(2, 4, 16, 16), (2, 4, 16, 32)]), ("ab,bc,cd,de->ea", [(2, 3), (3, 4), (4, 5), (5, 6)]),
    (" i j , j k -> i k ", [(3, 5), (5, 4)]),
]


@pytest.mark.parametrize("eq,shapes", EINSUM_CASES, ids=[c[0].strip().replace(" ", "") + f"#{i}" for i, c in enumerate(EINSUM_CASES)])
def test_einsum(rtg, ctx, oracle, math_mode, eq, shapes):
    """Einsum (src/ops/einsum.rs) lowered to permute -> batched MatMul on the contraction kernel -> permute, against
    the oracle's evaluation of the same lowering."""
    from rten_fork_b200.rten_builder import ModelBuilder

    rng = oracle.XorShiftRng(1234)
    b = ModelBuilder()
    ins = [b.add_value(f"x{i}", list(s)) for i, s in enumerate(shapes)]
    b.inputs = ins
    b.outputs = [b.op("Einsum", ins, ModelBuilder.einsum_attrs(eq))]
    xs = [rng.fill(s) - 0.5 for s in shapes]
    _m, got, want = _run_both(ctx, b, xs)
    assert got[0].shape == want[0].shape
    np.testing.assert_allclose(got[0], want[0], rtol=math_mode["rtol"] * 10, atol=math_mode["rtol"])


def test_einsum_errors(rtg, ctx):
    """OpError strings of EinsumExpr::parse / einsum() (src/ops/einsum.rs:27-93,151-203), and the forms this path
    does not build fail loudly."""
    from rten_fork_b200.rten_builder import ModelBuilder

    def run(eq, shapes):
        b = ModelBuilder()
        ins = [b.add_value(f"x{i}", list(s)) for i, s in enumerate(shapes)]
        b.inputs = ins
        b.outputs = [b.op("Einsum", ins, ModelBuilder.einsum_attrs(eq))]
        m = rtg.Model.load(b.build(), ctx=ctx)
        return m.run([np.zeros(s, np.float32) for s in shapes])

    for eq, shapes, msg in [("->ij", [(2, 2)], "Einsum equation must have at least one term"),
                            ("iJ,jk->ik", [(2, 2), (2, 2)], "Input term is invalid"),
                            ("ij,jk->ik", [(2, 2)], "Number of terms in Einsum equation does not match input tensor count"),
                            ("ijk,jk->ik", [(2, 2), (2, 2)], "Einsum term dimension count does not match input tensor"),
                            ("ij,jk->ik", [(2, 3), (4, 2)], "do not match"),
                            ("ii->i", [(2, 2)], "not supported on this path"),
                            ("...ij,...jk->...ik", [(2, 2, 2), (2, 2, 2)], "not supported on this path"),
                            ("ij->i", [(2, 2)], "not supported on this path")]:
        with pytest.raises(rtg.RtgError, match=msg):
            run(eq, shapes)


def test_cached_plan_replays_as_a_cuda_graph(rtg, oracle):
    """The second run of a (shapes, math mode) key is stream-captured and later runs replay one CUDA graph: results are
    bit-identical to the eager run, follow the input values, and a new batch size gets its own capture."""
    from rten_fork_b200 import synth
    from rten_fork_b200.model import Model

    c = rtg.Context(0)
    c.set_math(rtg.MATH_TF32_BF16X2)
    b = synth.resnet50(image=32, width=8, blocks=(1, 1, 1, 1), classes=10, seed=5)
    m = Model.load(b.build(), ctx=c)
    rng = oracle.XorShiftRng(3)
    x1, x2 = rng.fill((4, 3, 32, 32)), rng.fill((4, 3, 32, 32))
    eager = m.run([x1])[0]                    # first run of the key: eager
    assert m.graph_replays == 0
    l0 = c.launch_count
    captured = m.run([x1])[0]                 # second: captured, instantiated, launched
    assert m.graph_replays == 1
    per_run = c.launch_count - l0
    assert per_run > 10                       # the launch counter advances by the kernels inside the graph
    replay = m.run([x1])[0]
    assert m.graph_replays == 2 and c.launch_count - l0 == 2 * per_run
    assert eager.tobytes() == captured.tobytes() == replay.tobytes()
    other = m.run([x2])[0]                    # replays read the new input
    assert other.tobytes() != replay.tobytes()
    m.set_graph_capture(False)
    assert m.run([x2])[0].tobytes() == other.tobytes() and m.graph_replays == 3
    m.set_graph_capture(True)
    y8 = m.run([rng.fill((8, 3, 32, 32))])[0]  # another batch size: its own key, eager first
    assert y8.shape == (8, 10) and m.graph_replays == 3
    assert m.run([x1])[0].tobytes() == eager.tobytes() and m.graph_replays == 4
    # asynchronous submits replay too
    import torch

    xh, oh = torch.from_numpy(x1).pin_memory(), torch.empty((4, 10)).pin_memory()
    m.run_async([xh], [oh]).wait()
    assert oh.numpy().tobytes() == eager.tobytes() and m.graph_replays == 5
    m.close()
    c.close()


def test_graph_with_host_round_trip_stays_eager(rtg, oracle):
    """A graph that reads a device-computed shape back on the host (Shape -> Gather -> ... -> Reshape) cannot be
    captured: it keeps running eagerly and correctly."""
    from rten_fork_b200.model import Model
    from rten_fork_b200.rten_builder import ModelBuilder

    b = ModelBuilder()
    x = b.add_value("x", ["n", 6])
    b.inputs = [x]
    y = b.op("Relu", [x], name="relu")
    # the new shape arrives as a graph INPUT that lives on the device: Reshape has to read it back
    shp = b.add_value("shape", [2])
    b.inputs.append(shp)
    y = b.op("Reshape", [y, shp], ("ReshapeAttrs", []), name="reshape")
    b.outputs = [y]
    c = rtg.Context(0)
    m = Model.load(b.build(), ctx=c)
    import torch

    xin = oracle.XorShiftRng(9).fill((4, 6)) - 0.5
    sh = torch.tensor([3, 8], dtype=torch.int32, device="cuda")   # a device tensor: no host mirror
    xs = torch.from_numpy(xin).cuda()
    for _ in range(3):
        got = m.run([xs, sh])[0]
        assert got.shape == (3, 8) and got.tobytes() == np.maximum(xin, 0).reshape(3, 8).tobytes()
    assert m.graph_replays == 0
    m.close()
    c.close()


def test_loader_rejects_overflowing_constants_and_releases_the_pool(rtg, ctx):
    """Untrusted files: a shape whose element count leaves int64, a tensor-data offset that wraps u64, and an error
    after constants were uploaded (they must go back to the pool).  Reference errors: src/model.rs:468-522
    ("invalid tensor data offset", "length does not match shape")."""
    from rten_fork_b200.model import Model
    from rten_fork_b200.rten_builder import ModelBuilder

    def small(n_const=1):
        b = ModelBuilder()
        x = b.add_value("x", [1, 100])
        b.inputs = [x]
        ids = [b.add_constant(f"c{i}", np.ones((1, 100 + 7 * i), np.float32)) for i in range(n_const)]
        return b, x, ids

    b, x, (c,) = small()
    b.nodes[c]["shape_override"] = [0xFFFFFFFF, 0xFFFFFFFF, 4]      # 2^66 elements: n * 4 would wrap to a small count
    b.outputs = [b.op("Add", [x, c])]
    for version in (1, 2):
        with pytest.raises(rtg.RtgError, match="does not match shape|invalid tensor data offset") as e:
            Model.load(b.build(version=version), ctx=ctx)
        assert e.value.kind == "Format"
    b, x, (c,) = small()
    b.nodes[c]["offset_override"] = (1 << 64) - 8                    # tensor_data_offset + offset wraps around
    b.outputs = [b.op("Add", [x, c])]
    with pytest.raises(rtg.RtgError, match="invalid tensor data offset"):
        Model.load(b.build(version=2), ctx=ctx)
    # two constants are uploaded, then an operator this registry does not have fails the load: both buffers return
    b, x, ids = small(2)
    y = b.op("Add", [x, ids[0]])
    b.outputs = [b.op("Cos", [y])]
    data = b.build(version=2)
    with pytest.raises(rtg.RtgError, match="not available"):
        Model.load(data, ctx=ctx)
    s0 = ctx.pool_stats()
    with pytest.raises(rtg.RtgError, match="not available"):
        Model.load(data, ctx=ctx)
    s1 = ctx.pool_stats()
    assert s1["hit_count"] - s0["hit_count"] >= 2 and s1["len"] == s0["len"], "the failed load leaked its constants"
