"""Batch-sharded Model::run inside one process (rtg_sharded_model_*): one call over a batch,
chunks of dim 0 running concurrently on several contexts (several GPUs when the box has them; two contexts of GPU 0
otherwise, which exercises the same scatter / worker / gather logic), no collective."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _devices(n):
    import torch

    have = torch.cuda.device_count()
    return [i % have for i in range(n)]


@pytest.mark.parametrize("shards,batch", [(2, 8), (3, 7), (4, 3)])
def test_sharded_run_equals_the_chunks_run_alone(rtg, ctx, oracle, shards, batch):
    """Every shard's rows equal what Model.run gives for that chunk by itself (same kernels, same shapes: bitwise), and
    the whole result agrees with the oracle; ragged splits and more shards than items included."""
    import graph_oracle
    from rten_fork_b200 import synth
    from rten_fork_b200.model import Model, ShardedModel
    from rten_fork_b200.parallel import shard_range

    b = synth.resnet50(image=32, width=8, blocks=(1, 1, 1, 1), classes=10, seed=3)
    blob = b.build()
    x = oracle.XorShiftRng(21).fill((batch, 3, 32, 32))
    sm = ShardedModel(blob, _devices(shards), math=rtg.MATH_3XTF32)
    out = np.full((batch, 10), np.nan, np.float32)
    sm.run([x], [out])
    sm.run([x], [out])   # second call: every shard replays its captured plan
    ctx.set_math(rtg.MATH_3XTF32)
    m = Model.load(blob, ctx=ctx)
    for r in range(shards):
        lo, hi = shard_range(batch, shards, r)
        if hi > lo:
            assert out[lo:hi].tobytes() == m.run([x[lo:hi]])[0].tobytes(), f"shard {r}"
    want = graph_oracle.run_graph(b, [x])[0]
    np.testing.assert_allclose(out, want, rtol=2e-4, atol=2e-4 * np.abs(want).max())
    m.close()
    sm.close()


def test_sharded_load_refuses_graphs_that_mix_batch_items(rtg):
    from rten_fork_b200.model import ShardedModel
    from rten_fork_b200.rten_builder import ModelBuilder

    b = ModelBuilder()
    x = b.add_value("x", ["n", 4, 6])
    b.inputs = [x]
    y = b.op("Transpose", [x], ("TransposeAttrs", [(0, "vec_u32", [1, 0, 2])]), name="swap")   # batch leaves dim 0
    b.outputs = [y]
    with pytest.raises(rtg.RtgError, match="not separable by batch item.*swap") as e:
        ShardedModel(b.build(), _devices(2))
    assert e.value.kind == "UnsupportedValue"
    # a Reshape with a baked batch size (BERT-style head split) folds items together as far as the loader can tell
    b = ModelBuilder()
    x = b.add_value("x", ["n", 8])
    b.inputs = [x]
    shp = b.add_constant("shape", np.array([2, -1], np.int32))
    b.outputs = [b.op("Reshape", [x, shp], ("ReshapeAttrs", []), name="fold")]
    with pytest.raises(rtg.RtgError, match="folds the batch dimension"):
        ShardedModel(b.build(), _devices(2))


def test_sharded_run_checks_its_buffers(rtg, oracle):
    from rten_fork_b200 import synth
    from rten_fork_b200.model import ShardedModel

    sm = ShardedModel(synth.resnet50(image=32, width=8, blocks=(1, 1, 1, 1), classes=10).build(), _devices(2))
    x = oracle.XorShiftRng(1).fill((4, 3, 32, 32))
    with pytest.raises(rtg.RtgError, match="batch in dim 0"):
        sm.run([x], [np.zeros((3, 10), np.float32)])
    with pytest.raises(rtg.RtgError, match="wrong shape"):
        sm.run([x], [np.zeros((4, 11), np.float32)])
    sm.close()
