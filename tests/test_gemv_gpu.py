"""The memory-bound path for M <= 8 (src/gemm.rs:651-704 gemv + src/gemm/kernels.rs:26-194; the reference takes it for
M == 1): both B layouts, the reference's bench_gemm gemv case 1 x 4096 x 512 (src/gemm.rs:1790-1838), the ResNet fc
shape at batch 1..8, alpha / beta / bias, against the oracle; reruns bit-identical."""
import numpy as np
import pytest

from conftest import expect_equal

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("m", [1, 2, 4, 5, 8])
@pytest.mark.parametrize("n,k", [(4096, 512), (1000, 2048), (64, 300), (128, 33), (4, 4096)])
@pytest.mark.parametrize("transposed_b", [False, True])
def test_gemv_paths(rtg, ctx, oracle, m, n, k, transposed_b):
    rng = oracle.XorShiftRng(1234)
    a = rng.fill((m, k)) - 0.5
    b = (rng.fill((n, k)) - 0.5).T if transposed_b else rng.fill((k, n)) - 0.5   # K-major view / row-major
    out = np.zeros((m, n), np.float32)
    rtg.GemmExecutor(ctx).gemm(out.reshape(-1), n, a, b)
    assert "gemv" in ctx.last_kernel, ctx.last_kernel
    want = a.astype(np.float64) @ b.astype(np.float64)
    mag = np.abs(a).astype(np.float64) @ np.abs(b).astype(np.float64)
    assert (np.abs(out - want) <= 1e-6 * mag + 1e-7).all()          # exact fp32 arithmetic: far inside rtol 1e-5
    ref = oracle.gemm(a, np.ascontiguousarray(b))
    assert (np.abs(out - ref) <= 2e-6 * mag + 1e-7).all()
    again = np.zeros((m, n), np.float32)
    rtg.GemmExecutor(ctx).gemm(again.reshape(-1), n, a, b)
    assert again.tobytes() == out.tobytes()


@pytest.mark.parametrize("transposed_b", [False, True])
def test_gemv_epilogue(rtg, ctx, oracle, transposed_b):
    """alpha, beta with an initialised output, the reference's scalar bias[0] for M == 1 (src/gemm.rs:785)."""
    rng = oracle.XorShiftRng(99)
    a = rng.fill((1, 640))
    b = rng.fill((256, 640)).T if transposed_b else rng.fill((640, 256))
    init = rng.fill((1, 256))
    bias = np.array([3.0], np.float32)
    got = init.copy()
    rtg.GemmExecutor(ctx).gemm_bias(got.reshape(-1), 256, a, b, 0.5, 2.0, bias)
    assert "gemv" in ctx.last_kernel
    want = oracle.gemm(a, np.ascontiguousarray(b), alpha=0.5, beta=2.0, bias=bias, out=init.copy().reshape(-1),
                       out_row_stride=256).reshape(1, 256)
    expect_equal(got, want, rtol=1e-5)


def test_fc_head_at_small_batch_uses_gemv(rtg, ctx, oracle):
    """ONNX Gemm with transB (the ResNet classifier) at batch 4 through the operator API."""
    from rten_fork_b200 import ops

    rng = oracle.XorShiftRng(7)
    x, w, c = rng.fill((4, 2048)) - 0.5, rng.fill((1000, 2048)) - 0.5, rng.fill((1000,))
    got = ops.gemm_op(x, w, c, 1.0, 1.0, False, True, ctx=ctx)
    assert "gemv" in ctx.last_kernel, ctx.last_kernel
    want = oracle.gemm_op(x, w, c, 1.0, 1.0, False, True)
    mag = np.abs(x) @ np.abs(w.T) + np.abs(c)
    assert (np.abs(got - want) <= 2e-6 * mag).all()
