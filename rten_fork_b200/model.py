"""``Model`` — mirror of the reference's model API (src/model.rs:214-263 load / load_file, :580-646 run / run_one)
over the C-ABI ``rtg_model_*`` entry points.  The graph executes on the GPU; inputs may be numpy arrays (uploaded
inside ``run``) or torch CUDA tensors (consumed in place), outputs come back in the reference layout."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native as N
from .context import Context, as_tensor, default_context


class Ticket:
    """One Model.run_async in flight."""

    def __init__(self, handle, out, odescs, keep):
        self._h, self._out, self._odescs, self._keep = handle, out, odescs, keep

    def wait(self):
        if self._h is not None:
            h, self._h = self._h, None
            N.check(N.lib().rtg_ticket_wait(h))
        res = []
        for o, d in zip(self._out, self._odescs):
            shp = tuple(d.shape[k] for k in range(d.ndim))
            n = int(np.prod(shp)) if shp else 1
            res.append(o.reshape(-1)[:n].reshape(shp))
        return res

    def __del__(self):
        try:
            if self._h is not None:
                N.lib().rtg_ticket_wait(self._h)
        except Exception:
            pass


class Model:
    def __init__(self, handle, ctx: Context):
        self._h, self.ctx = handle, ctx

    # -- loading (ModelOptions::{with_all_ops, enable_optimization}, src/model.rs:155-207)
    @classmethod
    def load(cls, data: bytes, optimize: bool = True, ctx: Context | None = None) -> "Model":
        ctx = ctx or default_context()
        h = C.c_void_p()
        buf = (C.c_char * len(data)).from_buffer_copy(data)
        N.check(N.lib().rtg_model_load(ctx.handle, buf, C.c_size_t(len(data)), C.c_int32(bool(optimize)), C.byref(h)))
        return cls(h, ctx)

    @classmethod
    def load_file(cls, path: str, optimize: bool = True, ctx: Context | None = None) -> "Model":
        ctx = ctx or default_context()
        h = C.c_void_p()
        N.check(N.lib().rtg_model_load_file(ctx.handle, path.encode(), C.c_int32(bool(optimize)), C.byref(h)))
        return cls(h, ctx)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            N.lib().rtg_model_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- metadata (Model::input_ids / output_ids / node_info, src/model.rs:524-578)
    def _info(self, fn, i):
        name = C.c_char_p()
        shape = (C.c_int64 * N.RTG_MAX_DIMS)()
        nd, dt = C.c_int32(), C.c_int32()
        N.check(fn(self._h, C.c_int32(i), C.byref(name), shape, C.byref(nd), C.byref(dt)))
        dims = None if nd.value < 0 else [None if shape[k] < 0 else shape[k] for k in range(nd.value)]
        return {"name": (name.value or b"").decode(), "shape": dims}

    @property
    def inputs(self):
        n = C.c_int32()
        N.check(N.lib().rtg_model_num_inputs(self._h, C.byref(n)))
        return [self._info(N.lib().rtg_model_input_info, i) for i in range(n.value)]

    @property
    def outputs(self):
        n = C.c_int32()
        N.check(N.lib().rtg_model_num_outputs(self._h, C.byref(n)))
        return [self._info(N.lib().rtg_model_output_info, i) for i in range(n.value)]

    @property
    def num_ops(self) -> int:
        n = C.c_int32()
        N.check(N.lib().rtg_model_num_ops(self._h, C.byref(n)))
        return n.value

    # -- CUDA-graph replay of the cached plan (rtg_model_set_graph_capture / rtg_model_graph_replays)
    def set_graph_capture(self, enabled: bool):
        N.check(N.lib().rtg_model_set_graph_capture(self._h, C.c_int32(bool(enabled))))

    @property
    def graph_replays(self) -> int:
        n = C.c_uint64()
        N.check(N.lib().rtg_model_graph_replays(self._h, C.byref(n)))
        return n.value

    # -- RunOptions {timing} (src/graph.rs:467-483)
    def set_timing(self, enabled: bool):
        N.check(N.lib().rtg_model_set_timing(self._h, C.c_int32(bool(enabled))))

    def timing_report(self) -> str:
        N.lib().rtg_model_timing_report.restype = C.c_char_p
        return N.lib().rtg_model_timing_report(self._h).decode()

    # -- node lookup (Model::find_node / input_ids / output_ids, src/model.rs:525-553; Graph::get_source_node)
    def find_node(self, name: str) -> int:
        i = C.c_int32()
        N.check(N.lib().rtg_model_find_node(self._h, name.encode(), C.byref(i)))
        return i.value

    @property
    def input_ids(self):
        n, i = C.c_int32(), C.c_int32()
        N.check(N.lib().rtg_model_num_inputs(self._h, C.byref(n)))
        ids = []
        for k in range(n.value):
            N.check(N.lib().rtg_model_input_id(self._h, C.c_int32(k), C.byref(i)))
            ids.append(i.value)
        return ids

    @property
    def output_ids(self):
        n, i = C.c_int32(), C.c_int32()
        N.check(N.lib().rtg_model_num_outputs(self._h, C.byref(n)))
        ids = []
        for k in range(n.value):
            N.check(N.lib().rtg_model_output_id(self._h, C.c_int32(k), C.byref(i)))
            ids.append(i.value)
        return ids

    def node_kind(self, node_id: int) -> str:
        k = C.c_int32()
        N.check(N.lib().rtg_model_node_kind(self._h, C.c_int32(node_id), C.byref(k)))
        return ("operator", "constant", "value")[k.value]

    def source_op(self, value_id: int):
        """(operator name, node name) of the operator producing ``value_id`` after the load-time optimiser."""
        op, node = C.c_char_p(), C.c_char_p()
        N.check(N.lib().rtg_model_source_op(self._h, C.c_int32(value_id), C.byref(op), C.byref(node)))
        return (op.value or b"").decode(), (node.value or b"").decode()

    # -- Model::run: inputs in graph-input order (a list), or addressed by node id like the reference
    #    (``inputs`` a list of (node id, array) pairs / a dict, ``outputs`` a list of node ids); returns numpy arrays (or
    #    fills ``out`` buffers).  ``partial=True``: Model::partial_run, returns a list of (node id, array) leaves.
    def run(self, inputs, out=None, outputs=None, partial=False):
        by_id = isinstance(inputs, dict) or (len(inputs) > 0 and isinstance(inputs[0], tuple)) or outputs is not None \
            or partial
        ids = None
        if by_id:
            pairs = list(inputs.items()) if isinstance(inputs, dict) else list(inputs)
            ids = [int(k) for k, _ in pairs]
            inputs = [v for _, v in pairs]
        descs = (N.Tensor * max(1, len(inputs)))()
        keep = []
        for i, x in enumerate(inputs):
            if isinstance(x, np.ndarray) and not x.flags.c_contiguous:
                x = np.ascontiguousarray(x)
            keep.append(x)
            descs[i] = as_tensor(x)
        run = C.c_void_p()
        if by_id:
            outs = list(self.output_ids if outputs is None else outputs)
            id_arr = (C.c_int32 * max(1, len(ids)))(*ids)
            out_arr = (C.c_int32 * max(1, len(outs)))(*outs)
            fn = N.lib().rtg_model_partial_run if partial else N.lib().rtg_model_run_ids
            N.check(fn(self._h, id_arr, descs, C.c_int32(len(ids)), out_arr, C.c_int32(len(outs)), C.byref(run)))
        else:
            N.check(N.lib().rtg_model_run(self._h, descs, C.c_int32(len(inputs)), C.byref(run)))
        try:
            n = C.c_int32()
            N.check(N.lib().rtg_run_num_outputs(run, C.byref(n)))
            results = []
            for i in range(n.value):
                shape = (C.c_int64 * N.RTG_MAX_DIMS)()
                nd, dt = C.c_int32(), C.c_int32()
                N.check(N.lib().rtg_run_output_info(run, C.c_int32(i), shape, C.byref(nd), C.byref(dt)))
                shp = tuple(shape[k] for k in range(nd.value))
                if out is not None and out[i] is not None:
                    dst = out[i]
                else:
                    dst = np.empty(shp, dtype=np.float32 if dt.value == N.RTG_F32 else np.int32)
                if isinstance(dst, np.ndarray):
                    ptr, nbytes, dev = dst.ctypes.data, dst.nbytes, N.RTG_HOST
                else:  # torch tensor (pinned host or CUDA)
                    ptr, nbytes = dst.data_ptr(), dst.numel() * dst.element_size()
                    dev = dst.device.index if dst.is_cuda else N.RTG_HOST
                N.check(N.lib().rtg_run_output_copy(run, C.c_int32(i), C.c_void_p(ptr), C.c_size_t(nbytes), C.c_int32(dev)))
                if partial:
                    nid = C.c_int32()
                    N.check(N.lib().rtg_run_output_node(run, C.c_int32(i), C.byref(nid)))
                    results.append((nid.value, dst))
                else:
                    results.append(dst)
            return results
        finally:
            N.lib().rtg_run_free(run)

    def run_async(self, inputs, out):
        """rtg_model_submit: asynchronous run over host buffers.  ``inputs`` are contiguous host arrays (numpy, or
        pinned torch CPU tensors for real copy / compute overlap), ``out`` one host buffer per graph output with at
        least the output's size.  Returns a Ticket; ``ticket.wait()`` blocks until the outputs are in ``out`` and
        returns them shaped like the results.  Two submits may be in flight: the upload of the second overlaps the
        graph of the first."""
        keep = list(inputs) + list(out)
        descs = (N.Tensor * len(inputs))(*[as_tensor(x) for x in inputs])
        odescs = (N.Tensor * len(out))(*[as_tensor(o) for o in out])
        tk = C.c_void_p()
        N.check(N.lib().rtg_model_submit(self._h, descs, C.c_int32(len(inputs)), odescs, C.c_int32(len(out)), C.byref(tk)))
        return Ticket(tk, out, odescs, keep)

    def partial_run(self, inputs, outputs=None):
        """Model::partial_run (src/model.rs:639-646): list of (node id, value) leaves."""
        return self.run(inputs, outputs=outputs, partial=True)

    def run_one(self, x, out=None):
        """Model::run_one (src/model.rs:620-637): single input, single output."""
        return self.run([x], None if out is None else [out])[0]


class ShardedModel:
    """``rtg_sharded_model_*``: one ``run`` over a batch, split along dim 0 over several GPUs of this process (one
    context + worker thread + weight replica per entry of ``devices``; a device may appear twice).  No collective; the
    graph must be separable by batch item, which is checked at load (``RtgError`` UnsupportedValue otherwise)."""

    def __init__(self, data: bytes, devices, math: int = N.MATH_3XTF32, optimize: bool = True):
        devs = (C.c_int32 * len(devices))(*[int(d) for d in devices])
        h = C.c_void_p()
        buf = (C.c_char * len(data)).from_buffer_copy(data)
        N.check(N.lib().rtg_sharded_model_load(devs, C.c_int32(len(devices)), C.c_int32(math), buf, C.c_size_t(len(data)),
                                               C.c_int32(bool(optimize)), C.byref(h)))
        self._h, self.devices = h, list(devices)

    def run(self, inputs, out):
        """``inputs``: contiguous host arrays [batch, ...]; ``out``: one contiguous host buffer [batch, ...] per graph
        output (numpy, or pinned torch CPU tensors).  Returns ``out``."""
        keep = [np.ascontiguousarray(x) if isinstance(x, np.ndarray) else x for x in inputs]
        descs = (N.Tensor * len(keep))(*[as_tensor(x) for x in keep])
        odescs = (N.Tensor * len(out))(*[as_tensor(o) for o in out])
        N.check(N.lib().rtg_sharded_model_run(self._h, descs, C.c_int32(len(keep)), odescs, C.c_int32(len(out))))
        return out

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            N.lib().rtg_sharded_model_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
