"""ORACLE — TEST INFRASTRUCTURE ONLY.

Evaluates a graph (the node list of a ``ModelBuilder``) on the CPU with the oracle's restatement of each reference
operator, in the order ``Graph::run`` would (src/graph.rs:797-1073: nodes are stored topologically).  Used by
``tests/`` and ``bench.py --impl reference`` / ``cpu_baseline``; never imported by the product package.
"""
from __future__ import annotations

import numpy as np

import oracle as o


def _attrs(nd):
    if nd["attrs"] is None:
        return {}
    return {fid: val for fid, _kind, val in nd["attrs"][1]}


def _padding(a, auto_field, pads_field):
    if a.get(auto_field, 0) == 0:
        return "same"
    p = a.get(pads_field)
    return tuple(p) if p is not None else (0, 0, 0, 0)


def einsum(equation, ins):
    """src/ops/einsum.rs: terms folded left to right (einsum_path); each pairwise step permutes the operands to
    [batch.., M.., K..] x [batch.., K.., N..], multiplies them with the oracle's MatMul (reference GEMM kernel and
    accumulation order) and permutes the product into the requested label order (einsum_matmul)."""
    eq = "".join(equation.split())
    lhs, _, rhs = eq.partition("->")
    terms = lhs.split(",")
    if "->" not in eq:
        letters = "".join(terms)
        rhs = "".join(sorted(c for c in set(letters) if letters.count(c) == 1))
    sizes = {}
    for t, x in zip(terms, ins):
        for c, n in zip(t, x.shape):
            sizes[c] = n

    def view(term, x, order):
        x = np.asarray(x, np.float32)
        have = [c for c in order if c in term]
        x = np.transpose(x, [term.index(c) for c in have])
        shape = [sizes[c] if c in term else 1 for c in order]
        return np.broadcast_to(x.reshape(shape), [sizes[c] for c in order])

    def contract(ta, a, tb, b, out):
        batch = [c for c in out if c in ta and c in tb]
        ms = [c for c in out if c in ta and c not in tb]
        ns = [c for c in out if c in tb and c not in ta]
        ks = []
        for c in ta + tb:
            if c not in out and c not in ks:
                ks.append(c)
        prod = lambda ls: int(np.prod([sizes[c] for c in ls])) if ls else 1  # noqa: E731
        a3 = np.ascontiguousarray(view(ta, a, batch + ms + ks)).reshape(prod(batch), prod(ms), prod(ks))
        b3 = np.ascontiguousarray(view(tb, b, batch + ks + ns)).reshape(prod(batch), prod(ks), prod(ns))
        p = o.matmul(a3, b3).reshape([sizes[c] for c in batch + ms + ns])
        order = batch + ms + ns
        return np.array(np.transpose(p, [order.index(c) for c in out]), order="C")

    if len(terms) == 1:
        return np.array(view(terms[0], ins[0], list(rhs)), order="C")
    cur_t, cur = terms[0], ins[0]
    for i in range(1, len(terms)):
        if i + 1 < len(terms):
            later = set(rhs).union(*terms[i + 1:])
            step = []
            for c in cur_t + terms[i]:
                if c in later and c not in step:
                    step.append(c)
            step = "".join(step)
        else:
            step = rhs
        cur = contract(cur_t, cur, terms[i], ins[i], step)
        cur_t = step
    return cur


class _Pool:
    """TensorPool (src/tensor_pool.rs:103-199) for the timed CPU baseline: operator outputs come from buffers of dead
    values (best fit by size) instead of fresh allocations, which on the CPU cost a page fault per 4 KiB."""

    def __init__(self):
        self.free = []

    def alloc(self, shape):
        n = int(np.prod(shape))
        best = -1
        for i, b in enumerate(self.free):
            if b.size >= n and (best < 0 or b.size < self.free[best].size):
                best = i
        if best >= 0:
            return self.free.pop(best)[:n].reshape(shape)
        return np.empty(shape, dtype=np.float32)

    def add(self, arr):
        base = arr
        while isinstance(base.base, np.ndarray):
            base = base.base
        if base.dtype == np.float32 and base.flags.c_contiguous and base.flags.owndata:
            self.free.append(base.reshape(-1))


def run_graph(builder, inputs, outputs=None, reuse=False, pool=None):
    """``inputs``: arrays in graph-input order.  Returns the graph outputs (or the requested node ids).

    ``reuse``: execute like Graph::run_plan does (src/graph.rs:883-1051): values are reference counted, Relu / Clip /
    Add run in place on temporaries nobody else reads (src/graph.rs:897-931), dead temporaries go back to a pool that
    later convolutions allocate from.  Results are identical; only the timed CPU baseline uses it (pass the same
    ``pool`` to consecutive steps to keep its buffers)."""
    vals = {}
    for nid, x in zip(builder.inputs, inputs):
        vals[nid] = np.asarray(x)
    for i, nd in enumerate(builder.nodes):
        if nd["kind"] == "const":
            vals[i] = nd["data"]
    ids = outputs if outputs is not None else builder.outputs
    refs, temp = {}, set()
    if reuse:
        pool = pool if pool is not None else _Pool()
        for nd in builder.nodes:
            if nd["kind"] == "op":
                for i in nd["inputs"]:
                    if i is not None and i >= 0:
                        refs[i] = refs.get(i, 0) + 1
        for i in ids:
            refs[i] = refs.get(i, 0) + 1
    for nd in builder.nodes:
        if nd["kind"] != "op":
            continue
        ins = [vals[i] if i is not None and i >= 0 else None for i in nd["inputs"]]
        a = _attrs(nd)
        t = nd["type"]
        # in place: the first input is an owned temporary with no other reader
        inplace = reuse and nd["inputs"] and nd["inputs"][0] in temp and refs.get(nd["inputs"][0], 0) == 1
        if t == "Conv":
            y = o.conv2d(ins[0], ins[1], ins[2] if len(ins) > 2 else None, padding=_padding(a, 0, 1),
                         groups=a.get(2, 1), strides=tuple(a.get(3) or (1, 1)), dilations=tuple(a.get(4) or (1, 1)),
                         alloc=pool.alloc if reuse else None)
        elif reuse and t in ("Relu", "Gelu", "Sigmoid", "Erf"):
            y = o.unary(t.lower(), ins[0], in_place=inplace)
        elif reuse and t == "Add" and ins[0].dtype == np.float32 and ins[1].dtype == np.float32:
            # commutative: the executor may pick either operand as the in-place one (src/graph.rs:905-925)
            if inplace:
                y = o.binary("add", ins[0], ins[1], in_place=True)
            elif nd["inputs"][1] in temp and refs.get(nd["inputs"][1], 0) == 1 and ins[1].shape == ins[0].shape:
                y = o.binary("add", ins[1], ins[0], in_place=True)
            else:
                y = o.binary("add", ins[0], ins[1])
        elif t in ("Relu", "Gelu", "Sigmoid", "Erf"):
            y = o.unary(t.lower(), ins[0])
        elif t == "Sqrt":
            y = np.sqrt(ins[0].astype(np.float32))
        elif t == "Tanh":
            y = np.tanh(ins[0].astype(np.float64)).astype(np.float32)
        elif t == "Clip":
            lo = float(np.asarray(ins[1]).reshape(-1)[0]) if len(ins) > 1 and ins[1] is not None else None
            hi = float(np.asarray(ins[2]).reshape(-1)[0]) if len(ins) > 2 and ins[2] is not None else None
            y = o.unary("clip", ins[0], lo, hi)
        elif t == "Pow":
            y = np.power(ins[0].astype(np.float32), ins[1].astype(np.float32))
        elif t in ("Add", "Sub", "Mul", "Div"):
            y = o.binary(t.lower(), ins[0].astype(np.float32), ins[1].astype(np.float32))
        elif t in ("MaxPool", "AveragePool"):
            y = o.pool2d("max" if t == "MaxPool" else "avg", ins[0], tuple(a[0]), tuple(a.get(3) or (1, 1)),
                         _padding(a, 1, 2), bool(a.get(4, 0)))
        elif t == "GlobalAveragePool":
            y = o.global_average_pool(ins[0])
        elif t == "Flatten":
            ax = a.get(0, 0)
            x = ins[0]
            ax = ax + x.ndim if ax < 0 else ax
            y = np.ascontiguousarray(x).reshape(int(np.prod(x.shape[:ax])), -1)
        elif t == "Reshape":
            shape = [int(s) for s in ins[1]]
            shape = [ins[0].shape[k] if s == 0 and not a.get(0) else s for k, s in enumerate(shape)]
            y = np.ascontiguousarray(ins[0]).reshape(shape)
        elif t == "Gemm":
            y = o.gemm_op(ins[0], ins[1], ins[2] if len(ins) > 2 else None, a.get(0, 0.0), a.get(1, 0.0),
                          bool(a.get(2)), bool(a.get(3)))
        elif t == "Einsum":
            y = einsum(a[0], ins)
        elif t == "MatMul":
            y = o.matmul(np.ascontiguousarray(ins[0]), np.ascontiguousarray(ins[1]))
        elif t == "Transpose":
            y = np.ascontiguousarray(np.transpose(ins[0], a.get(0)))
        elif t == "Softmax":
            y = o.softmax(ins[0], a.get(0, 0))
        elif t == "LayerNormalization":
            y = o.layer_norm(ins[0], ins[1], ins[2] if len(ins) > 2 else None, a.get(0, 0), a.get(1, 0.0))
        elif t == "Gather":
            assert a.get(0, 0) == 0
            y = ins[0][ins[1]]
        elif t == "Cast":
            y = ins[0].astype(np.float32 if a.get(0, 0) == 1 else np.int32)
        elif t == "Unsqueeze":
            y = ins[0]
            for ax in sorted(int(v) for v in ins[1]):
                y = np.expand_dims(y, ax)
        elif t == "ConvTranspose":
            same = a.get(1, 1) == 0
            y = o.conv_transpose(ins[0], ins[1], ins[2] if len(ins) > 2 else None,
                                 "same" if same else tuple(a.get(2) or (0, 0, 0, 0)), tuple(a.get(0) or (1, 1)))
        elif t == "Where":
            y = np.ascontiguousarray(np.where(ins[0] != 0, ins[1], ins[2]))
        elif t == "ReduceMean":  # src/ops/reduce.rs:334-353: sum / len in f32 (summation order differs: tolerance)
            ax = a.get(0)
            ax = tuple(int(v) for v in ax) if ax else None
            x = ins[0].astype(np.float32)
            cnt = x.size if ax is None else int(np.prod([x.shape[v] for v in ax]))
            y = (x.sum(axis=ax, keepdims=bool(a.get(1)), dtype=np.float32) / np.float32(cnt)).astype(np.float32)
        elif t == "Concat":  # src/ops/concat.rs
            y = np.concatenate([x for x in ins], axis=a.get(0, 0))
        elif t == "Slice":  # src/ops/slice.rs: NumPy slicing semantics (clamped endpoints, negative steps)
            x = ins[0]
            starts, ends = [int(v) for v in ins[1]], [int(v) for v in ins[2]]
            axes = [int(v) for v in ins[3]] if len(ins) > 3 and ins[3] is not None else list(range(len(starts)))
            steps = [int(v) for v in ins[4]] if len(ins) > 4 and ins[4] is not None else [1] * len(starts)
            sl = [slice(None)] * x.ndim
            for st, en, ax, sp in zip(starts, ends, axes, steps):
                sl[ax] = slice(st, en, sp)
            y = np.ascontiguousarray(x[tuple(sl)])
        elif t == "Shape":
            y = np.array(ins[0].shape, dtype=np.int32)
        elif t == "Expand":
            y = np.ascontiguousarray(np.broadcast_to(ins[0], np.broadcast_shapes(ins[0].shape, tuple(int(v) for v in ins[1]))))
        elif t == "Identity":
            y = ins[0].copy()
        else:
            raise NotImplementedError(t)
        out_id = nd["outputs"][0]
        vals[out_id] = y
        if reuse:
            temp.add(out_id)
            for i in nd["inputs"]:
                if i is None or i < 0:
                    continue
                refs[i] -= 1
                if refs[i] == 0 and i in temp and not any(np.shares_memory(vals[i], vals[k]) for k in (out_id,)):
                    pool.add(vals[i])
                    del vals[i]
    return [vals[i] for i in ids]
