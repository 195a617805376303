"""``python -m rten_fork_b200.cli model.rten`` — the rten-cli equivalent for the B200 path.

Mirrors the reference runner (rten-cli/src/main.rs:96-346): loads a ``.rten`` model, prints its inputs / outputs,
generates inputs from the shape metadata with the same heuristics (``*_mask`` -> ones i32, ``*_ids`` -> zeros i32,
anything else -> uniform [0, 1) f32; dynamic dimensions from ``-s name=size`` or 1), runs the graph ``-n`` times on
cuda:0 and reports the time of each run; ``-t`` prints the per-operator timing table (the RTEN_TIMING output,
docs/performance.md:25-45).  Everything executes through ``Model.run`` on the GPU: there is no CPU path behind it.
"""
from __future__ import annotations

import argparse
import sys
import time

import numpy as np


def _parse_dims(specs):
    """``-s dim=size`` or ``-s input.dim=size`` (DimSize::matches, rten-cli/src/main.rs:60-94)."""
    out = []
    for spec in specs or []:
        for part in spec.split(","):
            name, _, size = part.partition("=")
            if not size.isdigit():
                raise SystemExit(f'Invalid input shape specifier "{part}": expected "dim_name=size"')
            inp, _, dim = name.rpartition(".")
            out.append((inp or None, dim, int(size)))
    return out


def _resolve(name, shape, dims, quiet):
    res = []
    for d in shape:
        if isinstance(d, int):
            res.append(d)
            continue
        hit = [s for (inp, dn, s) in dims if dn == d and inp in (None, name)]
        if hit:
            res.append(hit[0])
        else:
            if not quiet:
                print(f'  Size not specified for dynamic dimension "{name}.{d}". Defaulting to 1.')
            res.append(1)
    return res


def generate_inputs(model, dims=(), quiet=False, seed=None):
    rng = np.random.default_rng(seed)
    arrays = []
    for info in model.inputs:
        name = info["name"] or "(unnamed input)"
        shape = _resolve(name, [d if d is not None else (info.get("dim_names") or {}).get(i, f"dim{i}")
                                for i, d in enumerate(info["shape"])], list(dims), quiet)
        if name.endswith("_mask"):
            arrays.append(np.ones(shape, dtype=np.int32))
        elif name.endswith("_ids"):
            arrays.append(np.zeros(shape, dtype=np.int32))
        else:
            arrays.append(rng.random(shape, dtype=np.float32))
        if not quiet:
            print(f'  Input "{name}" generated shape {list(shape)}')
    return arrays


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="rten_fork_b200.cli", description="Inspect and run RTen models on a B200.")
    ap.add_argument("model", help="path to a .rten model")
    ap.add_argument("-n", "--n_iters", type=int, default=1, help="number of times to run the model")
    ap.add_argument("-q", "--quiet", action="store_true", help="run model and don't produce other output")
    ap.add_argument("-t", "--timing", action="store_true", help="output timing info for each operator type")
    ap.add_argument("-s", "--shape", action="append", help='size of a dynamic dimension: "dimK=size" or "input.dimK=size", K the axis index '
                         "(the C ABI reports symbolic dimensions by position, not by name)")
    ap.add_argument("--no-optimize", action="store_true", help="skip the load-time graph optimiser (fusions)")
    ap.add_argument("--math", default="tf32_bf16x2", choices=["tf32_bf16x2", "3xtf32", "tf32"])
    args = ap.parse_args(argv)

    import rten_fork_b200 as R

    ctx = R.default_context(0)
    ctx.set_math({"tf32_bf16x2": R.MATH_TF32_BF16X2, "3xtf32": R.MATH_3XTF32, "tf32": R.MATH_TF32}[args.math])
    t0 = time.perf_counter()
    model = R.Model.load_file(args.model, optimize=not args.no_optimize, ctx=ctx)
    if not args.quiet:
        print(f"Model loaded in {(time.perf_counter() - t0) * 1e3:.2f}ms: {model.num_ops} operators")
        print("Inputs")
        for i in model.inputs:
            print(f'  {i["name"]}: {i["shape"]}')
        print("Outputs")
        for o in model.outputs:
            print(f'  {o["name"]}: {o["shape"]}')
        print()
    inputs = generate_inputs(model, _parse_dims(args.shape), args.quiet)
    if not args.quiet:
        print()
    if args.timing:
        model.set_timing(True)
    outs = None
    for _ in range(max(1, args.n_iters)):
        t = time.perf_counter()
        outs = model.run(inputs)
        if not args.quiet:
            print(f"  Model returned {len(outs)} outputs in {(time.perf_counter() - t) * 1e3:.2f}ms.")
        if args.timing:
            print(model.timing_report())
    if not args.quiet:
        print()
        for i, (o, info) in enumerate(zip(outs, model.outputs)):
            dtype = "f32" if o.dtype == np.float32 else "i32"
            print(f'  Output {i} "{info["name"]}" data type {dtype} shape: {list(o.shape)}')
    return 0


if __name__ == "__main__":
    sys.exit(main())
