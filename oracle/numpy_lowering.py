"""ORACLE — TEST INFRASTRUCTURE ONLY (parity unpinned, see oracle/README.md).

numpy restatement of what `--neptuneir-to-llvm` makes of the hot-path ops, i.e. of
lib/Passes/DataflowLowering.cpp in the reference:

  * field/temp -> identity-layout row-major memref of shape ub-lb      (:41-71)
  * wrap / unwrap / load forward their operand (the temp ALIASES the field)  (:131-159)
  * store = whole memref.copy, or subview+copy when `bounds` is given     (:165-220)
  * apply = alloc result, copy input 0 through, then for every p in [lb,ub)
        out[p - outLb] = body(in_k[p + off - inLb_k] ...)                 (:258-448)
    with the body's scalar ops evaluated one by one, in the body's own order, in IEEE-754
    round-to-nearest with no fusion (arith.mulf/addf -> llvm.fmul/fadd, no fast-math flags).
  * linear_opdef -> func.func, apply_linear -> func.call   (lib/Passes/StructureLowering.cpp:31-140)

The loop nest is evaluated as whole-array numpy slices: element-wise numpy add/sub/mul/div on
float64/float32 arrays are correctly-rounded IEEE operations and numpy never contracts a*b+c,
and the freshly allocated result never aliases an input, so the slice form is bit-identical to
the sequential loop nest.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import numpy as np

from . import mlir_subset as ms

_NP = {"f64": np.float64, "f32": np.float32}


def apply_numpy(ap: ms.Apply, inputs: Sequence[np.ndarray]) -> np.ndarray:
    """One `neptune_ir.apply` — DataflowLowering.cpp:258-448."""
    dt = _NP[ap.out_type.elem]
    out_shape = ap.out_type.bounds.shape
    if tuple(inputs[0].shape) != tuple(out_shape):
        raise ValueError("copy-through needs input 0 and the result to have equal extents (:282-287)")
    out = np.empty(out_shape, dtype=dt)          # memref.alloc   (:273-281)
    np.copyto(out, inputs[0])                    # memref.copy in0 -> out (:282-287)
    lb, ub = ap.bounds.lb, ap.bounds.ub
    if any(u <= l for l, u in zip(lb, ub)):
        return out
    env: Dict[str, object] = {}
    arg_index = {name: k for k, name in enumerate(ap.input_args)}
    for op in ap.body:
        if op.name == "access":                  # (:381-411) idx = iv + off - inLb
            k = arg_index[op.operands[0]]
            in_lb = ap.in_types[k].bounds.lb
            sl = tuple(slice(l + o - il, u + o - il) for l, u, o, il in zip(lb, ub, op.offsets, in_lb))
            for s, n in zip(sl, inputs[k].shape):
                if s.start < 0 or s.stop > n:
                    raise IndexError("access %s reads outside input %d (undefined behaviour in the reference)"
                                     % (op.result, k))
            env[op.result] = inputs[k][sl]
        elif op.name == "constant":
            env[op.result] = _NP[op.ty](op.value)
        elif op.name == "addf":
            env[op.result] = np.add(env[op.operands[0]], env[op.operands[1]], dtype=dt)
        elif op.name == "subf":
            env[op.result] = np.subtract(env[op.operands[0]], env[op.operands[1]], dtype=dt)
        elif op.name == "mulf":
            env[op.result] = np.multiply(env[op.operands[0]], env[op.operands[1]], dtype=dt)
        elif op.name == "divf":
            env[op.result] = np.divide(env[op.operands[0]], env[op.operands[1]], dtype=dt)
        elif op.name == "negf":
            env[op.result] = np.negative(env[op.operands[0]], dtype=dt)
        else:
            raise NotImplementedError(op.name)
    olb = ap.out_type.bounds.lb
    osl = tuple(slice(l - ol, u - ol) for l, u, ol in zip(lb, ub, olb))
    out[osl] = env[ap.yield_value]               # yield -> memref.store (:427-444)
    return out


def run_function(mod: ms.Module, name: str, args: Sequence[np.ndarray]) -> List[np.ndarray]:
    """Execute a func.func / opdef the way the lowered code would, on host numpy buffers.
    memref arguments are used in place (wrap forwards the buffer), so `store` mutates them."""
    fn = mod.functions[name]
    if len(args) != len(fn.args):
        raise ValueError("%s expects %d arguments" % (name, len(fn.args)))
    env: Dict[str, np.ndarray] = {}
    for (an, ty), a in zip(fn.args, args):
        env[an] = a
    for op in fn.ops:
        if op.name in ("wrap", "unwrap", "load"):
            src = env[op.operands[0]]
            ty = op.types[1] if op.name != "unwrap" else op.types[0]
            if isinstance(ty, ms.GridType):
                if tuple(src.shape) != tuple(ty.bounds.shape) or src.dtype != _NP[ty.elem]:
                    raise ValueError("%s: buffer %s/%s does not match %s" % (op.name, src.shape, src.dtype, ty))
            env[op.results[0]] = src             # forwarded, aliases (:131-159)
        elif op.name == "apply":
            env[op.results[0]] = apply_numpy(op.apply, [env[v] for v in op.apply.inputs])
        elif op.name == "apply_linear":
            env[op.results[0]] = run_function(mod, op.callee, [env[v] for v in op.operands])[0]
        elif op.name == "store":
            src, dst = env[op.operands[0]], env[op.operands[1]]
            if op.bounds is None:
                np.copyto(dst, src)              # memref.copy (:176-179)
            else:                                # subview x2 + copy (:181-218)
                slb, dlb = op.types[0].bounds.lb, op.types[1].bounds.lb
                ssl = tuple(slice(l - b, u - b) for l, u, b in zip(op.bounds.lb, op.bounds.ub, slb))
                dsl = tuple(slice(l - b, u - b) for l, u, b in zip(op.bounds.lb, op.bounds.ub, dlb))
                dst[dsl] = src[ssl]
        elif op.name == "reduce":
            # ReduceLowering (lib/Passes/DataflowLowering.cpp:589-698): only kind "sum"; acc starts at
            # zero and every element of the domain is added in row-major order, in the element type.
            # np.cumsum accumulates strictly left to right in the array's dtype (np.sum is pairwise).
            if op.callee != "sum":
                raise NotImplementedError('reduce kind %r (the reference lowers only "sum", :655-657)' % op.callee)
            src = env[op.operands[0]]
            ty = op.types[0]
            b = op.bounds if op.bounds is not None else ty.bounds
            sl = tuple(slice(l - o, u - o) for l, u, o in zip(b.lb, b.ub, ty.bounds.lb))
            flat = np.ascontiguousarray(src[sl]).reshape(-1)
            env[op.results[0]] = np.cumsum(flat, dtype=src.dtype)[-1] if flat.size else src.dtype.type(0)
        elif op.name == "constant":
            pass
        elif op.name == "return":
            return [env[v] for v in op.operands]
        else:
            raise NotImplementedError(op.name)
    return []


def run_steps(mod: ms.Module, u: np.ndarray, extra: Sequence[np.ndarray], nsteps: int,
              entry: str = "@entry") -> np.ndarray:
    """A time loop the way the reference's host drivers run one: call
    entry(out, in, extra...) once per step and swap the two buffers."""
    a = np.array(u, copy=True)
    b = np.empty_like(a)
    for _ in range(nsteps):
        run_function(mod, entry, [b, a] + list(extra))
        a, b = b, a
    return a
