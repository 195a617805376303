"""rtg_gemm_static_scratch: the split-K GEMM entry point that allocates nothing (scratch = the library's file-scope
device array, include/rten_gpu.h).  Collected last on purpose: called through the C
ABI on device tensors, checked against the existing entry point rtg_gemm on the same inputs under the tolerance of the
other GEMM tests, on two calls in a row that share the scratch array."""
import numpy as np
import pytest

from conftest import expect_equal

pytestmark = pytest.mark.gpu


def _both(rtg, ctx, a, b, alpha=1.0, beta=0.0, bias=None, splits=0, init=None):
    import torch

    m, n = a.shape[0], b.shape[1]
    ex = rtg.GemmExecutor(ctx)
    out_ref = (init.clone() if init is not None else torch.full((m, n), float("nan"), device="cuda"))
    out_new = out_ref.clone()
    torch.cuda.synchronize()
    ex.gemm_bias(out_ref.view(-1), n, a, b, alpha, beta, bias)              # existing entry point (pool scratch)
    ctx.sync()
    pool0, free0 = ctx.pool_stats(), torch.cuda.mem_get_info()[0]
    ex.gemm_static_scratch(out_new.view(-1), n, a, b, alpha, beta, bias, splits)
    ctx.sync()
    pool1, free1 = ctx.pool_stats(), torch.cuda.mem_get_info()[0]
    assert pool1["alloc_count"] == pool0["alloc_count"] and pool1["hit_count"] == pool0["hit_count"]  # pool untouched
    assert free1 == free0                                                   # no device memory taken from the driver
    assert "splitK" in ctx.last_kernel, ctx.last_kernel
    return out_ref.cpu().numpy(), out_new.cpu().numpy()


def test_static_scratch_gemm_matches_rtg_gemm_twice(rtg, ctx, oracle, math_mode):
    import torch

    rng = oracle.XorShiftRng(1234)
    # first call: alpha / beta / bias, the library's own split factor
    a = torch.from_numpy(rng.fill((256, 4096))).cuda()
    b = torch.from_numpy(rng.fill((4096, 256))).cuda()
    bias = torch.from_numpy(rng.fill((256,))).cuda()
    init = torch.from_numpy(rng.fill((256, 256))).cuda()
    ref, new = _both(rtg, ctx, a, b, alpha=0.5, beta=2.0, bias=bias, init=init)
    expect_equal(new, ref, rtol=math_mode["rtol"])
    expect_equal(new, oracle.gemm(a.cpu().numpy(), b.cpu().numpy(), alpha=0.5, beta=2.0, bias=bias.cpu().numpy(),
                                  out=init.cpu().numpy().copy()), rtol=math_mode["rtol"])
    # second call straight after it: other operands and shape, more splits; the array still holds the first call's
    # planes and nothing clears it in between
    a2 = torch.from_numpy(rng.fill((384, 2048)) - 0.25).cuda()
    b2 = torch.from_numpy(rng.fill((2048, 128))).cuda()
    ref2, new2 = _both(rtg, ctx, a2, b2, splits=4)
    tol = math_mode["rtol"]
    err = np.abs(new2 - ref2)
    scale = np.abs(a2.cpu().numpy()) @ np.abs(b2.cpu().numpy())   # mixed signs: relative to sum |a||b|
    assert (err <= tol * scale).all(), float((err / scale).max())
    assert np.isfinite(new2).all()


def test_static_scratch_gemm_rejects_what_does_not_fit(rtg, ctx):
    import torch

    ex = rtg.GemmExecutor(ctx)
    a = torch.zeros((4096, 512), device="cuda")
    b = torch.zeros((512, 4096), device="cuda")
    out = torch.zeros((4096, 4096), device="cuda")
    with pytest.raises(rtg.RtgError, match="RTG_STATIC_SCRATCH_BYTES") as e:   # 2 planes of 64 MiB > 16 MiB
        ex.gemm_static_scratch(out.view(-1), 4096, a, b, splits=2)
    assert e.value.kind == "UnsupportedValue"
    with pytest.raises(rtg.RtgError, match="multiples of 4"):                  # would need a gathered copy
        ex.gemm_static_scratch(torch.zeros(6 * 6, device="cuda"), 6, torch.zeros((6, 130), device="cuda"),
                               torch.zeros((130, 6), device="cuda"))
    with pytest.raises(rtg.RtgError, match="too short to split|at least two"):
        ex.gemm_static_scratch(torch.zeros(8 * 8, device="cuda"), 8, torch.zeros((8, 32), device="cuda"),
                               torch.zeros((32, 8), device="cuda"))
