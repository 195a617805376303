"""BERT-style encoder graph (BASELINE.json configs[2] topology at reduced size) through Model.run vs the oracle:
Gather embeddings, LayerNormalization, MatMul + bias Add (fused), head split / merge (Reshape, Transpose), QK^T with a
transposed-view operand, Softmax, the exported Gelu expression (fused by the optimiser like the reference's
fuse_gelu, src/optimize.rs:401-424), residual Adds, Cast / Unsqueeze mask arithmetic."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("optimize", [False, True])
def test_bert_encoder_reduced(rtg, ctx, oracle, math_mode, optimize):
    import graph_oracle
    from rten_fork_b200 import synth
    from rten_fork_b200.model import Model

    b = synth.bert_encoder(batch=2, seq=16, hidden=64, heads=4, ffn=128, layers=2, vocab=100, seed=4)
    rs = np.random.RandomState(0)
    ids = rs.randint(0, 100, size=(2, 16)).astype(np.int32)
    mask = np.ones((2, 16), np.int32)
    mask[1, 12:] = 0
    m = Model.load(b.build(), optimize=optimize, ctx=ctx)
    got = m.run([ids, mask])[0]
    want = graph_oracle.run_graph(b, [ids, mask])[0]
    assert got.shape == want.shape == (2, 16, 64)
    tol = math_mode["rtol"] * 20   # LayerNorm outputs mix signs: compare against the output scale
    np.testing.assert_allclose(got, want, rtol=tol, atol=tol * np.abs(want).max())
    if optimize:
        # per layer: 5 Gelu-expression ops collapse to 1 (fused into the FFN MatMul), 6 bias Adds fold into MatMuls
        unopt = Model.load(b.build(), optimize=False, ctx=ctx)
        assert m.num_ops <= unopt.num_ops - 2 * (5 + 6)
