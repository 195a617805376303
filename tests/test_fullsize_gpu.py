"""Parity at the sizes BASELINE.json's configs are benchmarked at: the exact bench graph
(`synth.resnet50(image=224, classes=1000, seed=0)` at batch 32), MobileNetV2 at batch 64, one BERT-base encoder layer
at batch 64 x seq 128, and the square GEMM sweep above 1024 (plus the K = 16384 case) — CUDA path through the C ABI
against the oracle (graphs) or float64 (single GEMMs), in all three math modes.

Tolerances.  Single contractions are held to the reference's own rule (rten-tensor/src/test_util.rs:44-61:
|x - y| <= 1e-8 + 1e-5 |y|; TF32 mode: the stated 2e-3).  Early post-ReLU activations of a graph are held to that
rule per element, once per contraction between the input and the value (relu1: one, layer1.0.relu3: four), plus an
absolute floor of 2e-6 of the tensor's scale per contraction (elements that are small sums of large terms: the oracle
itself sums in fp32 in another order).  Values 50 layers deep are
compared against the scale of the tensor at 20x the mode's rtol, like the reduced-size graph tests: two correct fp32
evaluations of a 53-convolution network differ by more than 1e-5 per element, and the reference has no test that says
otherwise (its model tests compare single operators, src/ops/conv.rs:677-702, src/gemm.rs:1185-1238).
"""
import numpy as np
import pytest

from conftest import expect_equal

pytestmark = pytest.mark.gpu

_cache = {}


def _node_output(b, name):
    outs = [nd["outputs"][0] for nd in b.nodes if nd["kind"] == "op" and nd["name"] == name]
    assert len(outs) == 1, name
    return outs[0]


def _case(key, make):
    if key not in _cache:
        _cache[key] = make()
    return _cache[key]


def _check_graph(got, want, names, depth, math_mode):
    """`depth`: name -> number of contractions between the graph input and that value, for the values held to the
    per-element rule: every contraction may contribute the mode's rtol once (and an absolute floor of 2e-6 of the
    tensor's scale for elements that are small sums of large terms); the rest are compared against the scale."""
    rtol = math_mode["rtol"]
    report = []
    for g, w, name in zip(got, want, names):
        assert g.shape == w.shape, (name, g.shape, w.shape)
        scale = float(np.abs(w).max())
        err = np.abs(g.astype(np.float64) - w.astype(np.float64))
        report.append(f"{name}: max err {err.max():.3e} (scale {scale:.3e}, rel-to-scale {err.max() / scale:.2e})")
        if name in depth:
            d = depth[name]
            bad = err > d * (rtol * np.abs(w) + 2e-6 * scale * (rtol / 1e-5))
            assert not bad.any(), f"{name}: {bad.sum()} of {bad.size} elements outside {d} x rtol {rtol}; " + "; ".join(report)
        else:
            assert err.max() <= 20 * rtol * scale, "; ".join(report)
    print("\n".join(report))


def _resnet50_b32(oracle):
    import graph_oracle
    from rten_fork_b200 import synth

    b = synth.resnet50(image=224, classes=1000, seed=0)   # bench.py's graph
    names = ["relu1", "layer1.0.relu3", "layer2.0.relu3", "layer3.0.relu3", "layer4.2.relu3", "fc"]
    b.outputs = [_node_output(b, n) for n in names]
    x = oracle.XorShiftRng(1234).fill((32, 3, 224, 224))
    return b, names, x, graph_oracle.run_graph(b, [x])


def test_resnet50_b32_full_size(rtg, ctx, oracle, math_mode):
    """BASELINE.json configs[1] exactly as bench.py runs it: M = 100 352-row launches, 784-tile grids, the split-K
    choices of layer3 / layer4 at batch 32 and ~420 MB of pool reuse."""
    from rten_fork_b200.model import Model

    b, names, x, want = _case("resnet", lambda: _resnet50_b32(oracle))
    m = Model.load(b.build(), ctx=ctx)
    try:
        got = m.run([x])
    finally:
        m.close()
    assert got[-1].shape == (32, 1000)
    _check_graph(got, want, names, {"relu1": 1, "layer1.0.relu3": 4}, math_mode)


def _mobilenet_b64(oracle):
    import graph_oracle
    from rten_fork_b200 import synth

    b = synth.mobilenet_v2(image=224, classes=1000, seed=0)
    names = ["features.0.relu6", "features.2.dw.relu6", "features.8.add", "classifier"]
    b.outputs = [_node_output(b, n) for n in names]
    x = oracle.XorShiftRng(1234).fill((64, 3, 224, 224))
    return b, names, x, graph_oracle.run_graph(b, [x])


def test_mobilenet_v2_b64_full_size(rtg, ctx, oracle, math_mode):
    """BASELINE.json configs[3]: every depthwise layer at its benchmarked size (C = 32 ... 960, 112 ... 7 pixels)."""
    from rten_fork_b200.model import Model

    b, names, x, want = _case("mbv2", lambda: _mobilenet_b64(oracle))
    m = Model.load(b.build(), ctx=ctx)
    try:
        got = m.run([x])
    finally:
        m.close()
    _check_graph(got, want, names, {"features.0.relu6": 1}, math_mode)


def _bert_layer(oracle):
    import graph_oracle
    from rten_fork_b200 import synth

    b = synth.bert_encoder(batch=64, seq=128, hidden=768, heads=12, ffn=3072, layers=1, vocab=1000, seed=0)
    names = ["encoder.layer.0.attention.LayerNorm", "encoder.layer.0.output.LayerNorm"]
    b.outputs = [_node_output(b, n) for n in names]
    rs = np.random.RandomState(0)
    ids = rs.randint(0, 1000, size=(64, 128)).astype(np.int32)
    mask = np.ones((64, 128), np.int32)
    mask[::3, 100:] = 0
    return b, names, [ids, mask], graph_oracle.run_graph(b, [ids, mask])


def test_bert_base_layer_b64_s128(rtg, ctx, oracle, math_mode):
    """One encoder layer of BASELINE.json configs[2] at full width: 8192 x 768 x {768, 3072} linears, 768 attention
    heads of 128 x 64 x 128 in one batched launch, fused Gelu / bias / scale epilogues."""
    from rten_fork_b200.model import Model

    b, names, ins, want = _case("bert", lambda: _bert_layer(oracle))
    m = Model.load(b.build(), ctx=ctx)
    try:
        got = m.run(ins)
    finally:
        m.close()
    _check_graph(got, want, names, {}, math_mode)


def _f64_case(m, n, k):
    import oracle as o

    rng = o.XorShiftRng(1234)
    a, b = rng.fill((m, k)), rng.fill((k, n))
    return a, b, a.astype(np.float64) @ b.astype(np.float64)


@pytest.mark.parametrize("m,n,k", [(2048, 2048, 2048), (4096, 4096, 4096), (16384, 256, 16384)])
def test_gemm_sweep_sizes_vs_float64(rtg, ctx, oracle, math_mode, m, n, k):
    """BASELINE.json configs[4] above the sizes the oracle comparison covers, and the K = 16384 depth of its largest
    entry, against a float64 product (inputs of the reference's own generator, src/gemm.rs:1854-1863)."""
    a, b, want = _case(("gemm", m, n, k), lambda: _f64_case(m, n, k))
    out = np.zeros((m, n), dtype=np.float32)
    rtg.GemmExecutor(ctx).gemm(out.reshape(-1), n, a, b)
    expect_equal(out, want, rtol=math_mode["rtol"])


# ---------------------------------------------------------------------------------------------- adversarial operands
# RTG_MATH_TF32_BF16X2 computes a*b as hi(a)*hi(b) [TF32] + bf16(lo(a))*bf16(hi(b)) + bf16(a)*bf16(lo(b)) [one bf16
# MMA], hi = round-to-nearest TF32.  include/rten_gpu.h states the per-product bound 2^-18 + 2^-19 (5.8e-6) relative
# to |a*b|; the fp32 accumulation (RN adds of 128-deep chunks) adds at most ~2e-6 of sum |a_k b_k| at K = 16384.
PRODUCT_BOUND = 2.0 ** -18 + 2.0 ** -19
ACCUM_SLACK = 2.0e-6


def _worst_pairs(n_pairs=64, n_samples=1 << 21):
    """Mantissa pairs that maximise the bf16 rounding of the correction operands (host model of the split)."""
    def rn_tf32(x):
        u = x.view(np.uint32).astype(np.uint64)
        return (((u + 0x1000) >> 13) << 13).astype(np.uint32).view(np.float32)

    def bf16(x):
        u = x.view(np.uint32).astype(np.uint64)
        return ((((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16) & 0xFFFFFFFF).astype(np.uint32).view(np.float32)

    rs = np.random.RandomState(7)
    ma = rs.randint(0, 1 << 16, n_samples).astype(np.uint32)   # mantissas near 1.0: the relative error is largest
    mb = rs.randint(0, 1 << 16, n_samples).astype(np.uint32)
    a = (ma | np.uint32(0x3F800000)).view(np.float32)
    b = (mb | np.uint32(0x3F800000)).view(np.float32)
    ha, hb = rn_tf32(a), rn_tf32(b)
    la, lb = a - ha, b - hb
    f = np.float64
    p = ha.astype(f) * hb.astype(f) + bf16(la).astype(f) * bf16(hb).astype(f) + bf16(a).astype(f) * bf16(lb).astype(f)
    rel = (p - a.astype(f) * b.astype(f)) / (a.astype(f) * b.astype(f))
    assert np.abs(rel).max() <= PRODUCT_BOUND, np.abs(rel).max()   # the model obeys the stated bound
    idx = np.argsort(rel)
    return a[idx[-n_pairs:]], b[idx[-n_pairs:]], a[idx[:n_pairs]], b[idx[:n_pairs]]


@pytest.mark.parametrize("k", [4096, 16384])
@pytest.mark.parametrize("mode", ["tf32_bf16x2", "3xtf32"])
@pytest.mark.parametrize("signs", ["coherent+", "coherent-", "mixed"])
def test_adversarial_operands(rtg, ctx, oracle, mode, k, signs):
    """Operands built so that every product carries (nearly) the largest rounding error of the correction operands,
    all with the same sign of error (coherent: nothing averages out over K), and with mixed operand signs.  Coherent
    rows must still meet the reference's expect_equal rule; every case must meet the stated per-product bound."""
    ap, bp, an, bn_ = _case("worst", _worst_pairs)
    pa, pb = (ap, bp) if signs != "coherent-" else (an, bn_)
    rs = np.random.RandomState(k)
    m, n = 128, 128
    # row i of A uses pair i % 64 everywhere, column j of B the matching b: product (i, j) with i % 64 == j % 64 is
    # fully coherent; power-of-two scales per k keep mantissas (hence rounding errors) while varying exponents
    scale = (2.0 ** rs.randint(-3, 4, size=k)).astype(np.float32)
    a = np.ascontiguousarray(np.repeat(pa[np.arange(m) % 64][:, None], k, axis=1) * scale[None, :])
    b = np.ascontiguousarray(np.repeat(pb[np.arange(n) % 64][None, :], k, axis=0) / scale[:, None])
    if signs == "mixed":
        sg = rs.choice(np.array([-1.0, 1.0], np.float32), size=k)
        a = a * sg[None, :]
        b = b * rs.choice(np.array([-1.0, 1.0], np.float32), size=k)[:, None]
    ctx.set_math({"3xtf32": rtg.MATH_3XTF32, "tf32_bf16x2": rtg.MATH_TF32_BF16X2}[mode])
    try:
        out = np.zeros((m, n), dtype=np.float32)
        rtg.GemmExecutor(ctx).gemm(out.reshape(-1), n, a, b)
    finally:
        ctx.set_math(rtg.MATH_3XTF32)
    want = a.astype(np.float64) @ b.astype(np.float64)
    mag = np.abs(a).astype(np.float64) @ np.abs(b).astype(np.float64)
    err = np.abs(out - want)
    bound = (PRODUCT_BOUND if mode == "tf32_bf16x2" else 2.0 ** -20) + ACCUM_SLACK
    assert (err <= bound * mag).all(), f"max err / sum|ab| = {(err / mag).max():.3e} > {bound:.3e}"
    if signs != "mixed":
        expect_equal(out, want, rtol=1e-5)
