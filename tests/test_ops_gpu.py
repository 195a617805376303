"""MatMul / Gemm / elementwise / pooling parity vs the oracle (C-ABI, host buffers).

Restates src/ops/matmul.rs:313-561 (gemm_op transposed / c broadcast / invalid, matmul broadcast, errors, zero-sized
dims), the vecmath error bounds (rten-vecmath) and the pooling tests.  Elementwise results must be BIT-EXACT: the
device code evaluates the same formulas with the same IEEE operations.
"""
import numpy as np
import pytest

from conftest import expect_equal

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("a_shape,b_shape", [
    ((3, 10), (10, 8)),
    ((2, 3, 10), (10, 8)),          # [A,M,K] x [K,N] collapse (src/ops/matmul.rs:155-169)
    ((3, 10), (2, 10, 8)),          # broadcast a
    ((2, 3, 10), (2, 10, 8)),       # batched
    ((2, 1, 3, 10), (1, 4, 10, 8)),  # broadcast both prefixes
    ((12, 128, 64), (12, 64, 128)),  # attention-shaped
    ((64, 768), (768, 256)),
])
def test_matmul_shapes(rtg, ctx, oracle, math_mode, a_shape, b_shape):
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill(a_shape), rng.fill(b_shape)
    expect_equal(ops.matmul(a, b), oracle.matmul(a, b), rtol=math_mode["rtol"])


def test_matmul_transposed_view(rtg, ctx, oracle, math_mode):
    """FusedTranspose(MatMul): the permuted view is consumed without materialising it (src/ops/fused.rs:46-78)."""
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    q, k = rng.fill((4, 16, 32)), rng.fill((4, 16, 32))
    kt = k.transpose(0, 2, 1)
    expect_equal(ops.matmul(q, kt), oracle.matmul(q, np.ascontiguousarray(kt)), rtol=math_mode["rtol"])


@pytest.mark.parametrize("a_shape,b_shape,kmajor_b", [
    ((4, 12, 128, 64), (4, 12, 64, 128), False),   # nested batch dims merge into one
    ((4, 1, 130, 64), (1, 3, 64, 36), False),      # two batch dims, each operand broadcast along one of them
    ((5, 200, 32), (5, 32, 20), False),            # partial M tile per entry
    ((2, 3, 4, 16, 8), (2, 3, 4, 8, 12), False),   # three nested dims -> one
    ((2, 3, 2, 16, 8), (2, 1, 2, 8, 12), False),   # not mergeable into two dims: per-entry launches
    ((6, 1, 64), (6, 64, 32), False),              # single-row entries
    ((3, 16, 8), (3, 8, 7), True),                 # N % 4 != 0: un-staged stores
    ((7, 300, 100), (7, 100, 260), True),          # K tail, several N tiles, K-major B
    ((2, 5, 64, 40), (5, 40, 48), False),          # B lacks the outer batch dim
])
def test_matmul_batched_single_launch(rtg, ctx, oracle, math_mode, a_shape, b_shape, kmajor_b):
    """Batched products run as one launch when TMA can walk the batch (src/ops/matmul.rs:171-238 loops per entry)."""
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(77)
    a = rng.fill(a_shape)
    if kmajor_b:
        bt = rng.fill(b_shape[:-2] + (b_shape[-1], b_shape[-2]))
        b = np.swapaxes(bt, -1, -2)
    else:
        b = rng.fill(b_shape)
    l0 = ctx.launch_count
    got = ops.matmul(a, b)
    launches = ctx.launch_count - l0
    expect_equal(got, oracle.matmul(a, np.ascontiguousarray(b)), rtol=math_mode["rtol"])
    if a_shape != (2, 3, 2, 16, 8):
        assert launches == 1 and "batched" in ctx.last_kernel, (launches, ctx.last_kernel)
    else:
        # per-entry launches; the TF32 + bf16 mode adds the B' pre-pass of each entry's (non-constant) B operand
        assert launches == (24 if math_mode["name"] == "tf32_bf16x2" else 12)


def test_matmul_errors_and_empty(rtg, ctx, oracle):  # src/ops/matmul.rs:484-561
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    with pytest.raises(rtg.RtgError, match="Columns of first matrix does not match rows of second matrix") as e:
        ops.matmul(rng.fill((3, 10)), rng.fill((9, 8)))
    assert e.value.kind == "IncompatibleInputShapes"
    with pytest.raises(rtg.RtgError, match="Cannot broadcast shapes"):
        ops.matmul(rng.fill((2, 3, 10)), rng.fill((3, 10, 8)))
    with pytest.raises(rtg.RtgError, match="Inputs must have >= 2 dimensions"):
        ops.matmul(rng.fill((10,)), rng.fill((10, 8)))
    for m, n, k in [(5, 10, 0), (5, 0, 10), (0, 5, 10)]:
        out = ops.matmul(np.zeros((m, k), np.float32), np.zeros((k, n), np.float32))
        assert out.shape == (m, n) and (out == 0).all()


@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
def test_gemm_op(rtg, ctx, oracle, math_mode, ta, tb):  # src/ops/matmul.rs:313-360
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    a = rng.fill((20, 30) if ta else (30, 20))
    b = rng.fill((10, 20) if tb else (20, 10))
    expect_equal(ops.gemm_op(a, b, None, 1.0, 1.0, ta, tb), oracle.gemm_op(a, b, None, 1.0, 1.0, ta, tb),
                 rtol=math_mode["rtol"])


@pytest.mark.parametrize("c_shape", [(10, 15), (15,), (1, 15), (10, 1), (1,)])
def test_gemm_op_c_broadcast(rtg, ctx, oracle, math_mode, c_shape):  # :362-400
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    a, b, c = rng.fill((10, 5)), rng.fill((5, 15)), rng.fill(c_shape)
    expect_equal(ops.gemm_op(a, b, c, 0.5, 2.0), oracle.gemm_op(a, b, c, 0.5, 2.0), rtol=math_mode["rtol"])


def test_gemm_op_invalid_c(rtg, ctx, oracle):  # :402-408
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    with pytest.raises(rtg.RtgError, match="Cannot broadcast c to output shape"):
        ops.gemm_op(rng.fill((10, 5)), rng.fill((5, 15)), rng.fill((10, 5)), 1.0, 1.0)


def test_resnet_fc_gemm(rtg, ctx, oracle, math_mode):
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    x, w, b = rng.fill((32, 2048)), rng.fill((1000, 2048)), rng.fill((1000,))
    expect_equal(ops.gemm_op(x, w, b, 1.0, 1.0, False, True), oracle.gemm_op(x, w, b, 1.0, 1.0, False, True),
                 rtol=math_mode["rtol"])


@pytest.mark.parametrize("kind", ["relu", "gelu", "sigmoid", "silu", "erf"])
def test_unary_bit_exact(rtg, ctx, oracle, kind):
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    x = (rng.fill((3, 1001)) - 0.5) * 20.0
    x[0, :8] = [0.0, -0.0, 1e-20, -1e-20, 104.5, -104.5, 88.0, -88.0]
    got = getattr(ops, kind)(x)
    want = oracle.unary(kind, x)
    assert got.tobytes() == want.tobytes(), f"{kind}: max diff {np.abs(got - want).max()}"
    y = x.copy()
    getattr(ops, kind)(y, in_place=True)
    assert y.tobytes() == want.tobytes()


@pytest.mark.parametrize("kind", ["erf", "gelu"])
def test_erf_gelu_bit_exact_dense_sweep(rtg, ctx, oracle, kind):
    """The device erf divides with a branch-free reciprocal (MUFU.RCP + one Newton step, the correctly rounded fast path);
    it has to give the bits of the oracle's IEEE division (rten-vecmath/src/erf.rs:28-58) for every operand, not only for
    a few random ones: every 5th float of both signs from 1e-3 to 20, plus zeros, huge values, infinities and NaN."""
    from rten_fork_b200 import ops

    lo, hi = np.float32(1e-3).view(np.uint32), np.float32(20.0).view(np.uint32)
    pos = np.arange(int(lo), int(hi), 5, dtype=np.uint32).view(np.float32)
    special = np.array([0.0, -0.0, 1e-30, -1e-30, 1e30, -1e30, 3.4e38, -3.4e38, np.inf, -np.inf, 10.19, 10.2, 10.21, 16.0,
                        np.nextafter(np.float32(16.0), np.float32(17.0))], dtype=np.float32)
    x = np.concatenate([pos, -pos, special])
    got = getattr(ops, kind)(x)
    want = oracle.unary(kind, x)
    # gelu(-inf) = -inf * 0 is NaN on both sides (sign / payload of a NaN are not part of the contract)
    bad = np.flatnonzero((got.view(np.uint32) != want.view(np.uint32)) & ~(np.isnan(got) & np.isnan(want)))
    assert bad.size == 0, f"{kind}: {bad.size} of {x.size} differ, first x = {x[bad[0]]!r}: {got[bad[0]]!r} vs {want[bad[0]]!r}"
    nan = np.array([np.nan], dtype=np.float32)
    assert np.isnan(getattr(ops, kind)(nan)).all()


def test_clip_bit_exact(rtg, ctx, oracle):
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    x = (rng.fill((7, 333)) - 0.5) * 20.0
    assert ops.clip(x, 0.0, 6.0).tobytes() == oracle.unary("clip", x, 0.0, 6.0).tobytes()
    assert ops.clip(x, None, 1.0).tobytes() == oracle.unary("clip", x, None, 1.0).tobytes()


def test_vecmath_error_bounds(rtg, ctx):
    """The reference's own accuracy claims (erf vs libm <= 6.631017e-7, rten-vecmath/src/erf.rs:138)."""
    import math
    from rten_fork_b200 import ops

    x = np.linspace(-6, 6, 20001, dtype=np.float32)
    got = ops.erf(x)
    want = np.array([math.erf(float(v)) for v in x])
    assert np.abs(got - want).max() <= 6.631017e-7 * 1.01


@pytest.mark.parametrize("kind", ["add", "sub", "mul", "div"])
@pytest.mark.parametrize("a_shape,b_shape", [
    ((2, 8, 5, 5), (2, 8, 5, 5)), ((2, 8, 5, 5), (8, 1, 1)), ((2, 8, 5, 5), (1,)), ((5,), (3, 1)),
    ((4, 1, 6), (1, 7, 1)), ((128, 768), (768,)),
    ((4, 3, 16, 16), (4, 1, 1, 16)),       # attention mask onto scores
    ((4, 3, 16, 16), ()), ((), (3, 40)),   # scalar operand on either side
    ((6, 20), (6, 1)), ((6, 1), (6, 20)),  # inner-broadcast operand
    ((2, 3, 4, 5, 8), (3, 1, 5, 1)),       # four outer dims after merging
    ((3, 5, 7), (5, 1)), ((70000,), (1,)),
    ((2, 3, 2, 3, 2, 3, 4), (3, 1, 3, 1, 3, 1)),  # too many outer dims for the row kernel
])
def test_binary_bit_exact(rtg, ctx, oracle, kind, a_shape, b_shape):
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill(a_shape) + 0.5, rng.fill(b_shape) + 0.5
    got, want = getattr(ops, kind)(a, b), oracle.binary(kind, a, b)
    assert got.shape == want.shape and got.tobytes() == want.tobytes()


def test_binary_broadcast_error(rtg, ctx, oracle):
    from rten_fork_b200 import ops

    with pytest.raises(rtg.RtgError, match="Cannot broadcast inputs to same shape"):
        ops.add(np.zeros((2, 3), np.float32), np.zeros((4,), np.float32))


@pytest.mark.parametrize("kind", ["max", "avg"])
@pytest.mark.parametrize("shape,k,s,pads", [
    ((2, 5, 9, 9), (3, 3), (2, 2), (1, 1, 1, 1)),      # ResNet stem MaxPool geometry
    ((1, 4, 8, 8), (2, 2), (2, 2), (0, 0, 0, 0)),
    ((2, 3, 7, 10), (3, 2), (1, 3), (0, 1, 1, 0)),
    ((1, 2, 5, 5), (2, 2), (1, 1), "same"),
])
def test_pool2d_bit_exact(rtg, ctx, oracle, kind, shape, k, s, pads):
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    x = rng.fill(shape) - 0.5
    if kind == "max":
        got, want = ops.max_pool(x, k, s, pads), oracle.pool2d("max", x, k, s, pads)
    else:
        got, want = ops.average_pool(x, k, s, pads), oracle.pool2d("avg", x, k, s, pads)
    assert got.shape == want.shape and got.tobytes() == want.tobytes()


@pytest.mark.parametrize("shape,k,s,pads", [
    ((2, 8, 9, 9), (3, 3), (2, 2), (1, 1, 1, 1)),      # odd output height: the last row pair has one row
    ((1, 64, 12, 12), (3, 3), (2, 2), (1, 1, 1, 1)),
    ((3, 4, 7, 5), (3, 3), (1, 1), (1, 1, 1, 1)),      # stride 1: windows of a pair share two rows
    ((1, 12, 6, 6), (3, 3), (2, 2), (0, 0, 0, 0)),
    ((2, 16, 8, 8), (2, 2), (2, 2), (0, 0, 0, 0)),
])
def test_max_pool_nhwc_kernels_bit_exact(rtg, ctx, oracle, shape, k, s, pads):
    """The NHWC MaxPool kernels (compile-time window; two output rows per thread for 3x3) against the oracle, on
    channels-last device tensors as Graph::run hands them over; -inf and NaN inputs included."""
    import torch
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(4321)
    x = rng.fill(shape) - 0.5
    x[0, 0, 0, 0] = -np.inf
    x[0, 1, 1, 1] = np.nan
    want = oracle.pool2d("max", x, k, s, pads)
    xd = torch.from_numpy(x).cuda().contiguous(memory_format=torch.channels_last)
    yd = torch.empty(want.shape, device="cuda").contiguous(memory_format=torch.channels_last)
    torch.cuda.synchronize()
    ops.max_pool(xd, k, s, pads, ctx=ctx, out=yd)
    ctx.sync()
    got = yd.cpu().numpy()
    assert ctx.last_kernel.startswith("maxpool_nhwc4"), ctx.last_kernel
    assert got.tobytes() == np.ascontiguousarray(want).tobytes()


def test_global_average_pool(rtg, ctx, oracle):
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(1234)
    x = rng.fill((3, 37, 7, 7))
    expect_equal(ops.global_average_pool(x), oracle.global_average_pool(x), rtol=1e-5)


def test_pool_matches_context_pool_stats(rtg, ctx):
    """TensorPool surface: alloc / add / alloc_count / hit_count (src/tensor_pool.rs:306-453)."""
    c = rtg.Context(0)
    s0 = c.pool_stats()
    p = c.pool_alloc(1 << 20)
    c.pool_free(p)
    q = c.pool_alloc(1 << 19)          # best fit: reuses the 1 MiB buffer
    s1 = c.pool_stats()
    assert q == p and s1["alloc_count"] == s0["alloc_count"] + 2 and s1["hit_count"] == s0["hit_count"] + 1
    c.pool_free(q)
    c.close()
