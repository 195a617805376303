"""GemmExecutor parity: CUDA path (through the C-ABI, host buffers) vs the oracle restatement.

Cases restate the reference's own GEMM tests (src/gemm.rs:1185-1777): size sweep, transposed inputs, alpha, beta
incl. K == 0, NaN-in-output with beta == 0, bias, gemv shapes, short output buffers.  Inputs come from the reference's
XorShiftRng(1234) (uniform [0,1)), tolerance is its expect_equal rule (rtol 1e-5) in 3xTF32 mode and the stated
2e-3 in TF32 mode.
"""
import numpy as np
import pytest

from conftest import expect_equal

pytestmark = pytest.mark.gpu


def _run(rtg, ctx, a, b, alpha=1.0, beta=0.0, bias=None, out=None):
    m, n = a.shape[0], b.shape[1]
    if out is None:
        out = np.zeros((m, n), dtype=np.float32)
    rtg.GemmExecutor(ctx).gemm_bias(out.reshape(-1), n, a, b, alpha, beta, bias)
    return out


# src/gemm.rs:1185-1238 test_gemm_with_kernel size lists
SIZE_CASES = (
    [(2, c, 20) for c in (2, 4, 5, 8, 1024, 1025)]          # vary columns
    + [(2, 2, d) for d in (2, 20, 256, 300)]                # vary depth
    + [(r, 2, 20) for r in (2, 8, 10, 16, 64, 80)]          # vary rows
    + [(s, s, s) for s in list(range(1, 20)) + [30, 64, 65]]  # squares
    + [(300, 300, 300), (128, 256, 512), (257, 129, 1000)]
)


@pytest.mark.parametrize("m,n,k", SIZE_CASES)
def test_gemm_sizes(rtg, ctx, oracle, math_mode, m, n, k):
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((m, k)), rng.fill((k, n))
    expect_equal(_run(rtg, ctx, a, b), oracle.gemm(a, b), rtol=math_mode["rtol"])


def test_gemm_transposed(rtg, ctx, oracle, math_mode):  # src/gemm.rs:1272-1293
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((20, 30)).T, rng.fill((10, 20)).T   # [30,20] x [20,10] as transposed views
    assert a.strides[0] == 4 and b.strides[0] == 4
    expect_equal(_run(rtg, ctx, a, b), oracle.gemm(a, b), rtol=math_mode["rtol"])
    a, b = rng.fill((300, 200)).T, rng.fill((150, 300)).T
    expect_equal(_run(rtg, ctx, a, b), oracle.gemm(a, b), rtol=math_mode["rtol"])


@pytest.mark.parametrize("alpha", [0.0, 0.5, 1.0, 2.0])
def test_gemm_alpha(rtg, ctx, oracle, math_mode, alpha):  # src/gemm.rs:1295-1315
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((10, 5)), rng.fill((5, 15))
    expect_equal(_run(rtg, ctx, a, b, alpha=alpha), oracle.gemm(a, b, alpha=alpha), rtol=math_mode["rtol"])


@pytest.mark.parametrize("m,n,k", [(10, 15, 5), (10, 15, 0), (100, 70, 300)])
@pytest.mark.parametrize("beta", [0.5, 1.0, 2.0])
def test_gemm_beta(rtg, ctx, oracle, math_mode, m, n, k, beta):  # src/gemm.rs:1317-1347
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((m, k)), rng.fill((k, n))
    init = rng.fill((m, n))
    got = _run(rtg, ctx, a, b, beta=beta, out=init.copy())
    want = oracle.gemm(a, b, out=init.copy().reshape(-1), beta=beta, out_row_stride=n).reshape(m, n)
    expect_equal(got, want, rtol=math_mode["rtol"])


@pytest.mark.parametrize("m,n,k", [(10, 15, 5), (10, 15, 0), (1, 64, 8)])
def test_gemm_beta_zero_ignores_nan(rtg, ctx, oracle, math_mode, m, n, k):  # src/gemm.rs:1349-1387
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((m, k)), rng.fill((k, n))
    got = _run(rtg, ctx, a, b, out=np.full((m, n), np.nan, dtype=np.float32))
    assert np.isfinite(got).all()
    expect_equal(got, oracle.gemm(a, b), rtol=math_mode["rtol"])


@pytest.mark.parametrize("m,n,k", [(10, 15, 5), (1, 40, 33), (130, 70, 64)])
def test_gemm_bias(rtg, ctx, oracle, math_mode, m, n, k):  # src/gemm.rs:1389-1408
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((m, k)), rng.fill((k, n))
    bias = np.arange(m, dtype=np.float32)
    expect_equal(_run(rtg, ctx, a, b, bias=bias), oracle.gemm(a, b, bias=bias), rtol=math_mode["rtol"])


@pytest.mark.parametrize("n,k", [(1, 1), (8, 8), (40, 20), (1024, 256), (100, 500)])
def test_gemv(rtg, ctx, oracle, math_mode, n, k):  # src/gemm.rs:1585-1777 (M == 1 shapes)
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((1, k)), rng.fill((k, n))
    expect_equal(_run(rtg, ctx, a, b), oracle.gemm(a, b), rtol=math_mode["rtol"])
    bt = rng.fill((n, k)).T  # transposed (column-major) B
    expect_equal(_run(rtg, ctx, a, bt), oracle.gemm(a, bt), rtol=math_mode["rtol"])
    expect_equal(_run(rtg, ctx, a, b, alpha=0.5, bias=np.array([3.0], np.float32)),
                 oracle.gemm(a, b, alpha=0.5, bias=np.array([3.0], np.float32)), rtol=math_mode["rtol"])


def test_gemm_strided_output(rtg, ctx, oracle):
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((7, 9)), rng.fill((9, 5))
    out = np.full(7 * 8, -1.0, dtype=np.float32)
    rtg.GemmExecutor(ctx).gemm(out, 8, a, b)
    got = out.reshape(7, 8)
    expect_equal(got[:, :5], oracle.gemm(a, b))
    assert (got[:, 5:] == -1.0).all()  # gaps between rows are not written


def test_gemm_panics_become_errors(rtg, ctx, oracle):  # src/gemm.rs:743-750,1167-1183
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((4, 5)), rng.fill((6, 3))
    with pytest.raises(rtg.RtgError, match="Columns of matrix `a` must match rows of matrix `b`"):
        rtg.GemmExecutor(ctx).gemm(np.zeros(12, np.float32), 3, a, b)
    b = rng.fill((5, 3))
    with pytest.raises(rtg.RtgError, match="Output buffer should be large enough"):
        rtg.GemmExecutor(ctx).gemm(np.zeros(11, np.float32), 3, a, b)
    with pytest.raises(rtg.RtgError, match="Bias vector length must match rows of matrix `a`"):
        rtg.GemmExecutor(ctx).gemm_bias(np.zeros(12, np.float32), 3, a, b, bias=np.zeros(3, np.float32))


def test_gemm_large_k_3xtf32_stays_within_tolerance(rtg, ctx, oracle):
    """K = 8192: the chunked fp32 accumulation keeps 3xTF32 inside the reference tolerance (checked against f64)."""
    ctx.set_math(rtg.MATH_3XTF32)
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((256, 8192)), rng.fill((8192, 192))
    got = _run(rtg, ctx, a, b)
    expect_equal(got, a.astype(np.float64) @ b.astype(np.float64), rtol=1e-5)


def test_gemm_device_resident(rtg, ctx, oracle, math_mode):
    import torch

    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((512, 384)), rng.fill((384, 640))
    ta, tb = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    out = torch.full((512, 640), float("nan"), device="cuda")
    torch.cuda.synchronize()
    rtg.GemmExecutor(ctx).gemm(out.view(-1), 640, ta, tb)
    ctx.sync()
    expect_equal(out.cpu().numpy(), oracle.gemm(a, b), rtol=math_mode["rtol"])
    # K-major ("prepacked") B: a transposed view of an [N, K] tensor
    tbt = tb.t().contiguous().t()
    rtg.GemmExecutor(ctx).gemm(out.view(-1), 640, ta, tbt)
    ctx.sync()
    expect_equal(out.cpu().numpy(), oracle.gemm(a, b), rtol=math_mode["rtol"])


@pytest.mark.parametrize("m,n,k", [(256, 256, 4096), (300, 200, 2000), (128, 128, 8224), (1024, 1024, 1024)])
def test_gemm_split_k(rtg, ctx, oracle, math_mode, m, n, k):
    """Few output tiles and a long K: the launch divides K over several CTAs per tile (partial sums in a workspace,
    fixed-order reduction with the epilogue).  Same tolerance as the unsplit kernel; repeated runs are bit-identical
    (Operator::is_deterministic, src/ops/mod.rs:850-861)."""
    rng = oracle.XorShiftRng(1234)
    a, b = rng.fill((m, k)), rng.fill((k, n))
    bias = rng.fill((m,))
    c0 = rng.fill((m, n))
    got = _run(rtg, ctx, a, b, alpha=0.5, beta=2.0, bias=bias, out=c0.copy())
    if k >= 2000:
        assert "splitK" in ctx.last_kernel, ctx.last_kernel
    expect_equal(got, oracle.gemm(a, b, alpha=0.5, beta=2.0, bias=bias, out=c0.copy()), rtol=math_mode["rtol"])
    again = _run(rtg, ctx, a, b, alpha=0.5, beta=2.0, bias=bias, out=c0.copy())
    assert again.tobytes() == got.tobytes()


@pytest.mark.parametrize("mode", ["3xtf32", "tf32_bf16x2", "tf32"])
def test_contraction_reruns_are_bit_identical(rtg, mode):
    """The kernel is deterministic (fixed accumulation order, split-K planes added in split order), so a result that
    differs between reruns is a pipeline race (a stage overwritten or read early).  Back-to-back launches with a small
    unrelated GEMM in between exercise the programmatic-dependent-launch overlap too (scripts/stress_race.py is the
    long version)."""
    import torch
    from rten_fork_b200 import ops

    # a context on torch's stream: the comparisons (torch kernels) and the launches under test are stream-ordered
    ctx = rtg.Context(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ctx.set_math({"3xtf32": rtg.MATH_3XTF32, "tf32": rtg.MATH_TF32, "tf32_bf16x2": rtg.MATH_TF32_BF16X2}[mode])
    try:
        torch.manual_seed(0)
        cl = torch.channels_last
        cases = []
        for (B, C, O, H, k) in [(4, 512, 512, 7, 3), (32, 256, 128, 56, 1), (32, 128, 128, 28, 3)]:
            x = (torch.rand((B, C, H, H), device="cuda") - 0.4).contiguous(memory_format=cl)
            w, b = torch.rand((O, C, k, k), device="cuda") - 0.5, torch.rand((O,), device="cuda")
            r = torch.rand((B, O, H, H), device="cuda").contiguous(memory_format=cl)
            p = k // 2
            cases.append(lambda x=x, w=w, b=b, p=p, r=r: ops.conv(x, w, b, padding=(p, p, p, p), groups=1, strides=(1, 1),
                                                                 residual=r, activation="relu", out_nhwc=True, ctx=ctx))
        a, bm = torch.rand((2048, 512), device="cuda") - 0.5, torch.rand((512, 2048), device="cuda") - 0.5
        cases.append(lambda: ops.matmul(a, bm, ctx=ctx))
        sa, sb = torch.rand((64, 64), device="cuda"), torch.rand((64, 64), device="cuda")
        for fn in cases:
            ref = fn().clone()
            for i in range(40):
                if i % 3 == 0:
                    ops.matmul(sa, sb, ctx=ctx)
                assert torch.equal(fn(), ref), f"rerun {i} differs ({ctx.last_kernel})"
    finally:
        torch.cuda.synchronize()
        ctx.close()
