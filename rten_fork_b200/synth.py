"""Synthesises the `.rten` graphs BASELINE.json's configs name, with random-init weights (there is no network for
checkpoints and `rten-convert` needs the absent `onnx` module).  The graphs have the operator sequence an ONNX export
of the torchvision / HF models produces after BatchNorm folding: ResNet-50 v1.5 (53 Conv + Relu/Add/MaxPool/
GlobalAveragePool/Flatten/Gemm), MobileNetV2 (Conv/Clip/Add/...), BERT-base encoder (Gather/Add/LayerNormalization/
MatMul/Reshape/Transpose/Softmax/Erf-form Gelu ...).  ``width`` / ``layers`` scale the models down for parity tests.
"""
from __future__ import annotations

import numpy as np

from .rten_builder import ModelBuilder


class _Init:
    def __init__(self, seed):
        self.rng = np.random.default_rng(seed)

    def conv(self, o, c, kh, kw):
        bound = np.sqrt(6.0 / (c * kh * kw))  # He-uniform keeps activations O(1) through ReLU stacks
        return self.rng.uniform(-bound, bound, size=(o, c, kh, kw)).astype(np.float32)

    def linear(self, o, i, gain=1.0):
        bound = gain * np.sqrt(3.0 / i)
        return self.rng.uniform(-bound, bound, size=(o, i)).astype(np.float32)

    def bias(self, n):
        return (self.rng.uniform(0, 1, size=(n,)) * 0.1).astype(np.float32)


def _conv(b: ModelBuilder, init: _Init, x, c_in, c_out, k, stride=1, pad=0, groups=1, name="conv"):
    w = b.add_constant(name + ".weight", init.conv(c_out, c_in // groups, k, k))
    bias = b.add_constant(name + ".bias", init.bias(c_out))
    return b.op("Conv", [x, w, bias], ModelBuilder.conv_attrs((pad,) * 4, groups, (stride, stride)), name=name)


def resnet50(batch="batch", image=224, width=64, blocks=(3, 4, 6, 3), classes=1000, seed=0) -> ModelBuilder:
    """torchvision ``resnet50`` (v1.5: the stride sits on the 3x3 conv), BatchNorm folded into the convolutions."""
    b, init = ModelBuilder(), _Init(seed)
    x = b.add_value("input", [batch, 3, image, image])
    b.inputs = [x]
    y = _conv(b, init, x, 3, width, 7, 2, 3, name="conv1")
    y = b.op("Relu", [y], name="relu1")
    y = b.op("MaxPool", [y], ModelBuilder.pool_attrs("MaxPool", (3, 3), (2, 2), (1, 1, 1, 1)), name="maxpool")
    c_in = width
    for li, n_blocks in enumerate(blocks):
        planes = width * (2 ** li)
        for bi in range(n_blocks):
            stride = 2 if (bi == 0 and li > 0) else 1
            p = f"layer{li + 1}.{bi}"
            t = _conv(b, init, y, c_in, planes, 1, name=p + ".conv1")
            t = b.op("Relu", [t], name=p + ".relu1")
            t = _conv(b, init, t, planes, planes, 3, stride, 1, name=p + ".conv2")
            t = b.op("Relu", [t], name=p + ".relu2")
            t = _conv(b, init, t, planes, planes * 4, 1, name=p + ".conv3")
            sc = y
            if bi == 0:
                sc = _conv(b, init, y, c_in, planes * 4, 1, stride, name=p + ".downsample")
            t = b.op("Add", [t, sc], name=p + ".add")
            y = b.op("Relu", [t], name=p + ".relu3")
            c_in = planes * 4
    y = b.op("GlobalAveragePool", [y], name="avgpool")
    y = b.op("Flatten", [y], ("FlattenAttrs", [(0, "i32", 1)]), name="flatten")
    w = b.add_constant("fc.weight", init.linear(classes, c_in))
    bias = b.add_constant("fc.bias", init.bias(classes))
    y = b.op("Gemm", [y, w, bias], ModelBuilder.gemm_attrs(1.0, 1.0, False, True), name="fc",
             out_shape=[batch, classes])
    b.outputs = [y]
    return b


def mobilenet_v2(batch="batch", image=224, width_mult=1.0, classes=1000, seed=0) -> ModelBuilder:
    """torchvision ``mobilenet_v2``: ReLU6 is exported as Clip(min=0, max=6) with constant inputs."""
    b, init = ModelBuilder(), _Init(seed)
    lo = b.add_constant("clip_min", np.array(0.0, dtype=np.float32))
    hi = b.add_constant("clip_max", np.array(6.0, dtype=np.float32))

    def relu6(v, name):
        return b.op("Clip", [v, lo, hi], name=name)

    def ch(c):
        return max(8, int(c * width_mult + 4) // 8 * 8)

    x = b.add_value("input", [batch, 3, image, image])
    b.inputs = [x]
    c_in = ch(32)
    y = relu6(_conv(b, init, x, 3, c_in, 3, 2, 1, name="features.0"), "features.0.relu6")
    cfg = [(1, 16, 1, 1), (6, 24, 2, 2), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1)]
    idx = 1
    for t, c, n, s in cfg:
        c_out = ch(c)
        for i in range(n):
            stride = s if i == 0 else 1
            p = f"features.{idx}"
            hidden = c_in * t
            v = y
            if t != 1:
                v = relu6(_conv(b, init, v, c_in, hidden, 1, name=p + ".expand"), p + ".expand.relu6")
            v = relu6(_conv(b, init, v, hidden, hidden, 3, stride, 1, groups=hidden, name=p + ".dw"), p + ".dw.relu6")
            v = _conv(b, init, v, hidden, c_out, 1, name=p + ".project")
            if stride == 1 and c_in == c_out:
                v = b.op("Add", [y, v], name=p + ".add")
            y, c_in = v, c_out
            idx += 1
    last = ch(1280) if width_mult > 1.0 else max(ch(1280 * width_mult), 64) if width_mult < 1.0 else 1280
    y = relu6(_conv(b, init, y, c_in, last, 1, name="features.18"), "features.18.relu6")
    y = b.op("GlobalAveragePool", [y], name="avgpool")
    y = b.op("Flatten", [y], ("FlattenAttrs", [(0, "i32", 1)]), name="flatten")
    w = b.add_constant("classifier.weight", init.linear(classes, last))
    bias = b.add_constant("classifier.bias", init.bias(classes))
    y = b.op("Gemm", [y, w, bias], ModelBuilder.gemm_attrs(1.0, 1.0, False, True), name="classifier")
    b.outputs = [y]
    return b


def bert_encoder(batch=64, seq=128, hidden=768, heads=12, ffn=3072, layers=12, vocab=30522, seed=0) -> ModelBuilder:
    """BERT-base encoder as exported by HF + ONNX: embeddings (Gather + Add + LayerNormalization), ``layers`` x
    (Q/K/V MatMul+Add, Reshape/Transpose to heads, QK^T MatMul, scale Div, mask Add, Softmax, PV MatMul, output
    MatMul+Add, residual Add, LayerNormalization, FFN MatMul+Add, Gelu in its Erf form, MatMul+Add, Add, LN).
    Inputs: ``input_ids`` [batch, seq] i32 and ``attention_mask`` [batch, seq] i32 (rten-cli fills zeros / ones)."""
    b, init = ModelBuilder(), _Init(seed)
    rng = init.rng
    dh = hidden // heads
    ids = b.add_value("input_ids", [batch, seq])
    mask = b.add_value("attention_mask", [batch, seq])
    b.inputs = [ids, mask]

    def const(name, arr):
        return b.add_constant(name, np.asarray(arr))

    def f32(v):
        return np.array(v, dtype=np.float32)

    def i32(v):
        return np.array(v, dtype=np.int32)

    def layer_norm(x, name):
        g = const(name + ".weight", (1.0 + 0.1 * rng.uniform(-1, 1, hidden)).astype(np.float32))
        be = const(name + ".bias", (0.1 * rng.uniform(-1, 1, hidden)).astype(np.float32))
        return b.op("LayerNormalization", [x, g, be],
                    ("LayerNormalizationAttrs", [(0, "i32", -1), (1, "f32", 1e-12)]), name=name)

    def linear(x, n_in, n_out, name, gain=1.0):
        w = const(name + ".weight", np.ascontiguousarray(init.linear(n_out, n_in, gain).T))  # [K, N] as ONNX exports
        bias = const(name + ".bias", init.bias(n_out) - 0.05)
        y = b.op("MatMul", [x, w], name=name + ".matmul")
        return b.op("Add", [y, bias], name=name + ".add")

    word = const("embeddings.word", (0.05 * rng.standard_normal((vocab, hidden))).astype(np.float32))
    pos = const("embeddings.position", (0.05 * rng.standard_normal((1, seq, hidden))).astype(np.float32))
    tok = const("embeddings.token_type", (0.05 * rng.standard_normal((hidden,))).astype(np.float32))
    x = b.op("Gather", [word, ids], ("GatherAttrs", [(0, "i32", 0)]), name="embeddings.gather")
    x = b.op("Add", [x, pos], name="embeddings.add_pos")
    x = b.op("Add", [x, tok], name="embeddings.add_tok")
    x = layer_norm(x, "embeddings.LayerNorm")
    # extended attention mask: (1 - mask) * -10000, shaped [batch, 1, 1, seq]
    m = b.op("Cast", [mask], ("CastAttrs", [(0, "u8", 1)]), name="mask.cast")
    m = b.op("Unsqueeze", [m, const("mask.axes", i32([1, 2]))], name="mask.unsqueeze")
    m = b.op("Sub", [const("one", f32(1.0)), m], name="mask.sub")
    m = b.op("Mul", [m, const("neg", f32(-10000.0))], name="mask.mul")
    split = const("shape.split_heads", i32([batch, seq, heads, dh]))
    merge = const("shape.merge_heads", i32([batch, seq, hidden]))
    scale = const("attn.scale", f32(np.sqrt(dh)))
    for li in range(layers):
        p = f"encoder.layer.{li}"

        def heads_of(t, name, perm=(0, 2, 1, 3)):
            t = b.op("Reshape", [t, split], ("ReshapeAttrs", []), name=name + ".reshape")
            return b.op("Transpose", [t], ("TransposeAttrs", [(0, "vec_u32", list(perm))]), name=name + ".transpose")

        q = heads_of(linear(x, hidden, hidden, p + ".attention.query"), p + ".q")
        k = heads_of(linear(x, hidden, hidden, p + ".attention.key"), p + ".k", perm=(0, 2, 3, 1))  # K^T
        v = heads_of(linear(x, hidden, hidden, p + ".attention.value"), p + ".v")
        s = b.op("MatMul", [q, k], name=p + ".attention.scores")
        s = b.op("Div", [s, scale], name=p + ".attention.scale")
        s = b.op("Add", [s, m], name=p + ".attention.mask")
        s = b.op("Softmax", [s], ("SoftmaxAttrs", [(0, "i32", -1)]), name=p + ".attention.softmax")
        c = b.op("MatMul", [s, v], name=p + ".attention.context")
        c = b.op("Transpose", [c], ("TransposeAttrs", [(0, "vec_u32", [0, 2, 1, 3])]), name=p + ".context.transpose")
        c = b.op("Reshape", [c, merge], ("ReshapeAttrs", []), name=p + ".context.reshape")
        a = linear(c, hidden, hidden, p + ".attention.output")
        x = layer_norm(b.op("Add", [a, x], name=p + ".attention.residual"), p + ".attention.LayerNorm")
        h = linear(x, hidden, ffn, p + ".intermediate")
        # Gelu as exported: x * 0.5 * (1 + erf(x / sqrt(2)))
        e = b.op("Div", [h, const(p + ".sqrt2", f32(np.sqrt(2.0)))], name=p + ".gelu.div")
        e = b.op("Erf", [e], name=p + ".gelu.erf")
        e = b.op("Add", [e, const(p + ".gelu.one", f32(1.0))], name=p + ".gelu.add")
        g = b.op("Mul", [h, e], name=p + ".gelu.mul")
        g = b.op("Mul", [g, const(p + ".gelu.half", f32(0.5))], name=p + ".gelu.half")
        o = linear(g, ffn, hidden, p + ".output")
        x = layer_norm(b.op("Add", [o, x], name=p + ".output.residual"), p + ".output.LayerNorm")
    b.outputs = [x]
    return b
