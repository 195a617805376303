"""Writer for `.rten` model files: a hand-rolled FlatBuffers encoder plus a ``ModelBuilder`` with the same role as the
reference's test-only builder (src/model_builder.rs:1-1005).  There is no ``flatc`` / ``flatbuffers`` module in this
environment, so the wire format is produced directly from the schema (src/schema.fbs); field ids are the declaration
order of each table (a union occupies two ids: ``<name>_type`` then ``<name>``).

Container: V2 = 32-byte header ("RTEN", u32 version 2, u64 model offset, u64 model length, u64 tensor-data offset),
the FlatBuffers model, then 64-byte aligned raw tensor data (src/header.rs:60-110, docs/rten-file-format.md).
V1 = the bare FlatBuffers buffer with all constants inline.
"""
from __future__ import annotations

import struct

import numpy as np

OPERATOR_TYPES = """Add ArgMin ArgMax AveragePool BatchNormalization Cast Clip Concat ConstantOfShape Conv ConvTranspose
Cos CumSum Div Equal Erf Expand Flatten Gather Gemm GlobalAveragePool Greater GRU Identity LeakyRelu Less LessOrEqual
Log LogSoftmax LSTM MatMul MaxPool Mod Mul Pad Pow Range ReduceMean ReduceL2 Relu Reshape Resize Shape Sigmoid Sin
Slice Split Sqrt Squeeze Softmax Sub Tanh Transpose Unsqueeze Where ReduceProd ReduceSum ReduceMin ReduceMax NonZero
ScatterElements Tile Not Abs Max Mean Min Sum OneHot Round Floor Ceil Reciprocal TopK Neg Exp GreaterOrEqual Size Tan
Acos Asin Atan InstanceNormalization HardSigmoid HardSwish And Or Xor Trilu ScatterND NonMaxSuppression Sign
GatherElements LayerNormalization ReduceSumSquare RandomUniform Elu RandomUniformLike RandomNormal RandomNormalLike
Softplus GatherND Gelu Einsum If""".split()
OP_ID = {name: i for i, name in enumerate(OPERATOR_TYPES)}

ATTR_TYPES = """ArgMaxAttrs AveragePoolAttrs BatchNormalizationAttrs CastAttrs ConcatAttrs ConstantOfShapeAttrs ConvAttrs
ConvTransposeAttrs FlattenAttrs GatherAttrs GemmAttrs GRUAttrs LeakyReluAttrs LSTMAttrs MaxPoolAttrs ReduceMeanAttrs
ReshapeAttrs ResizeAttrs SplitAttrs SoftmaxAttrs TransposeAttrs ModAttrs ScatterElementsAttrs OneHotAttrs TopKAttrs
HardSigmoidAttrs TriluAttrs ScatterNDAttrs NonMaxSuppressionAttrs LayerNormalizationAttrs RandomUniformAttrs EluAttrs
RandomUniformLikeAttrs RandomNormalAttrs RandomNormalLikeAttrs GatherNDAttrs GeluAttrs EinsumAttrs IfAttrs""".split()
ATTR_ID = {name: i + 1 for i, name in enumerate(ATTR_TYPES)}  # union: 0 is NONE

_SCALAR = {"u8": ("<B", 1), "bool": ("<B", 1), "u16": ("<H", 2), "i32": ("<i", 4), "u32": ("<I", 4), "f32": ("<f", 4),
           "u64": ("<Q", 8)}


class FlatBufferBuilder:
    """Back-to-front FlatBuffers builder.  Object handles are distances from the END of the buffer."""

    def __init__(self):
        self.chunks = []  # reversed order
        self.size = 0
        self.max_align = 8

    def _prepend(self, b: bytes):
        self.chunks.append(b)
        self.size += len(b)

    def _align(self, alignment: int, extra: int):
        """Pad so that after ``extra`` more bytes are prepended the head is ``alignment``-aligned."""
        pad = (-(self.size + extra)) % alignment
        if pad:
            self._prepend(b"\0" * pad)
        self.max_align = max(self.max_align, alignment)

    def string(self, s: str) -> int:
        data = s.encode("utf-8")
        self._align(4, len(data) + 1 + 4)
        self._prepend(data + b"\0")
        self._prepend(struct.pack("<I", len(data)))
        return self.size

    def vector(self, kind: str, values) -> int:
        fmt, es = _SCALAR[kind]
        n = len(values)
        self._align(max(4, es), n * es)          # element block aligned to its element size
        if es < 4:
            self._align(4, n * es + 4)
        if isinstance(values, np.ndarray):
            self._prepend(values.astype({"i32": "<i4", "u32": "<u4", "f32": "<f4"}[kind]).tobytes())
        else:
            self._prepend(b"".join(struct.pack(fmt, v) for v in values))
        self._align(4, 4)
        self._prepend(struct.pack("<I", n))
        return self.size

    def offset_vector(self, handles) -> int:
        n = len(handles)
        self._align(4, 4 * n + 4)
        body = bytearray()
        # element i sits (size + 4*(n - i)) bytes from the end once everything is written
        for i, h in enumerate(handles):
            slot_pos = self.size + 4 * (n - i)
            body += struct.pack("<I", slot_pos - h)
        self._prepend(bytes(body))
        self._prepend(struct.pack("<I", n))
        return self.size

    def table(self, fields) -> int:
        """``fields``: list of (field_id, kind, value); kind is a scalar kind or "offset" (value = handle).
        Fields whose value is None are omitted (defaults / absent)."""
        fields = [(fid, kind, val) for fid, kind, val in fields if val is not None]
        # inline layout: soffset (4 bytes) then fields, largest alignment first
        def size_of(kind):
            return 4 if kind == "offset" else _SCALAR[kind][1]

        order = sorted(fields, key=lambda f: -size_of(f[1]))
        pos, layout = 4, {}
        for fid, kind, _ in order:
            sz = size_of(kind)
            pos = (pos + sz - 1) // sz * sz
            layout[fid] = pos
            pos += sz
        table_size = (pos + 3) // 4 * 4
        align = max([4] + [size_of(k) for _, k, _ in fields])
        self._align(align, table_size)
        table_end = self.size            # distance from buffer end to the END of the table
        table_start = table_end + table_size
        body = bytearray(table_size)
        n_slots = (max([f[0] for f in fields]) + 1) if fields else 0
        vt_size = 4 + 2 * n_slots
        struct.pack_into("<i", body, 0, vt_size)  # vtable sits immediately before the table
        for fid, kind, val in fields:
            off = layout[fid]
            if kind == "offset":
                slot_pos = table_start - off      # distance of this slot from the buffer end
                struct.pack_into("<I", body, off, slot_pos - val)
            else:
                struct.pack_into(_SCALAR[kind][0], body, off, val)
        vt = bytearray(vt_size)
        struct.pack_into("<HH", vt, 0, vt_size, table_size)
        for fid, _, _ in fields:
            struct.pack_into("<H", vt, 4 + 2 * fid, layout[fid])
        self._prepend(bytes(body))
        self._prepend(bytes(vt))
        return table_start

    def finish(self, root: int, identifier: bytes = b"RTEN") -> bytes:
        self._align(self.max_align, 8)
        self._prepend(identifier)
        self._prepend(struct.pack("<I", self.size + 4 - root))
        return b"".join(reversed(self.chunks))


class ModelBuilder:
    """Builds a graph node by node and serialises it (same node / id model as src/model_builder.rs)."""

    def __init__(self):
        self.nodes = []          # dicts
        self.inputs, self.outputs = [], []

    # -- nodes -----------------------------------------------------------------------------------------------------
    def add_value(self, name, shape=None) -> int:
        """ValueNode; ``shape`` items are ints or symbolic names (str)."""
        self.nodes.append({"kind": "value", "name": name, "shape": shape})
        return len(self.nodes) - 1

    def add_constant(self, name, array) -> int:
        array = np.ascontiguousarray(array)
        if array.dtype not in (np.float32, np.int32):
            raise TypeError("constants must be float32 or int32")
        self.nodes.append({"kind": "const", "name": name, "data": array})
        return len(self.nodes) - 1

    def add_operator(self, name, op_type, inputs, outputs, attrs=None) -> int:
        """``attrs``: None or (attrs_table_name, [(field_id, kind, value | list), ...])."""
        self.nodes.append({"kind": "op", "name": name, "type": op_type, "inputs": list(inputs), "outputs": list(outputs),
                           "attrs": attrs})
        return len(self.nodes) - 1

    def op(self, op_type, inputs, attrs=None, name=None, out_shape=None) -> int:
        """Add an operator with one fresh output value; returns the output's node id."""
        name = name or f"{op_type}_{len(self.nodes)}"
        out = self.add_value(name + "_out", out_shape)
        self.add_operator(name, op_type, [(-1 if i is None else i) for i in inputs], [out], attrs)
        return out

    # -- attribute helpers (field ids from src/schema.fbs) ---------------------------------------------------------------
    @staticmethod
    def conv_attrs(pads=(0, 0, 0, 0), groups=1, strides=(1, 1), dilations=(1, 1), same=False):
        return ("ConvAttrs", [(0, "u8", 0 if same else 1), (1, "vec_u32", None if same else list(pads)),
                              (2, "u32", groups), (3, "vec_u32", list(strides)), (4, "vec_u32", list(dilations))])

    @staticmethod
    def pool_attrs(kind, kernel, strides, pads=(0, 0, 0, 0), same=False, count_include_pad=False):
        f = [(0, "vec_u32", list(kernel)), (1, "u8", 0 if same else 1), (2, "vec_u32", None if same else list(pads)),
             (3, "vec_u32", list(strides))]
        if kind == "AveragePool":
            f.append((4, "bool", 1 if count_include_pad else None))
        return (kind + "Attrs", f)

    @staticmethod
    def conv_transpose_attrs(pads=(0, 0, 0, 0), strides=(1, 1), same=False):
        # ConvTransposeAttrs: strides:0 auto_pad:1 (default NotSet = 1) pads:2
        return ("ConvTransposeAttrs", [(0, "vec_u32", list(strides)), (1, "u8", 0 if same else None),
                                       (2, "vec_u32", None if same else list(pads))])

    @staticmethod
    def reduce_mean_attrs(axes=None, keep_dims=True):
        return ("ReduceMeanAttrs", [(0, "vec_i32", None if axes is None else list(axes)),
                                    (1, "bool", 1 if keep_dims else None)])

    @staticmethod
    def einsum_attrs(equation):
        return ("EinsumAttrs", [(0, "str", str(equation))])

    @staticmethod
    def concat_attrs(axis=0):
        return ("ConcatAttrs", [(0, "i32", int(axis))])

    @staticmethod
    def gemm_attrs(alpha=1.0, beta=1.0, transpose_a=False, transpose_b=False):
        return ("GemmAttrs", [(0, "f32", alpha), (1, "f32", beta), (2, "bool", 1 if transpose_a else None),
                              (3, "bool", 1 if transpose_b else None)])

    # -- serialisation ---------------------------------------------------------------------------------------------
    def build(self, version: int = 2, inline_limit: int = 64) -> bytes:
        """Serialise.  ``version`` 1 = headerless, all constants inline; 2 = header + external tensor data for
        constants with more than ``inline_limit`` elements."""
        fb = FlatBufferBuilder()
        tensor_data = bytearray()
        node_handles = []
        for nd in self.nodes:
            name_h = fb.string(nd["name"]) if nd["name"] is not None else None
            if nd["kind"] == "value":
                shape_h = None
                if nd["shape"] is not None:
                    dims = []
                    for d in nd["shape"]:
                        if isinstance(d, str):
                            dims.append(fb.table([(1, "offset", fb.string(d))]))
                        else:
                            dims.append(fb.table([(0, "u32", int(d))]))
                    shape_h = fb.offset_vector(dims)
                data_h, kind_id = fb.table([(0, "offset", shape_h)]), 3
            elif nd["kind"] == "const":
                arr = nd["data"]
                # "shape_override" / "offset_override": malformed files for the loader's own error tests
                shape_h = fb.vector("u32", [int(s) for s in nd.get("shape_override", arr.shape)])
                is_f = arr.dtype == np.float32
                if version == 2 and arr.size > inline_limit:
                    pad = (-len(tensor_data)) % 64
                    tensor_data += b"\0" * pad
                    off = len(tensor_data)
                    tensor_data += arr.tobytes()
                    data_h = fb.table([(0, "offset", shape_h), (3, "u16", 1 if is_f else 0),
                                       (4, "u64", nd.get("offset_override", off))])
                else:
                    payload = fb.table([(0, "offset", fb.vector("f32" if is_f else "i32", arr.reshape(-1)))])
                    data_h = fb.table([(0, "offset", shape_h), (1, "u8", 1 if is_f else 2), (2, "offset", payload),
                                       (3, "u16", 1 if is_f else 0)])
                kind_id = 2
            else:
                attrs_h, attrs_id = None, None
                if nd["attrs"] is not None:
                    tname, fields = nd["attrs"]
                    enc = []
                    for fid, kind, val in fields:
                        if val is None:
                            continue
                        if kind.startswith("vec_"):
                            enc.append((fid, "offset", fb.vector(kind[4:], val)))
                        elif kind == "str":
                            enc.append((fid, "offset", fb.string(val)))
                        else:
                            enc.append((fid, kind, val))
                    attrs_h, attrs_id = fb.table(enc), ATTR_ID[tname]
                ins_h = fb.vector("i32", nd["inputs"])
                outs_h = fb.vector("i32", nd["outputs"])
                t = OP_ID[nd["type"]]
                data_h = fb.table([(0, "u8", t if t else None), (1, "u8", attrs_id), (2, "offset", attrs_h),
                                   (3, "offset", ins_h), (4, "offset", outs_h)])
                kind_id = 1
            node_handles.append(fb.table([(0, "offset", name_h), (1, "u8", kind_id), (2, "offset", data_h)]))
        graph_h = fb.table([(0, "offset", fb.offset_vector(node_handles)), (1, "offset", fb.vector("u32", self.inputs)),
                            (2, "offset", fb.vector("u32", self.outputs))])
        model_h = fb.table([(0, "i32", 1), (1, "offset", graph_h)])
        model = fb.finish(model_h)
        if version == 1:
            return model
        model_off = 32
        tensor_off = (model_off + len(model) + 63) // 64 * 64
        header = b"RTEN" + struct.pack("<IQQQ", 2, model_off, len(model), tensor_off)
        return header + model + b"\0" * (tensor_off - model_off - len(model)) + bytes(tensor_data)
