"""ORACLE — TEST INFRASTRUCTURE ONLY (parity unpinned, see oracle/README.md).

Third restatement of the reference's lowering of `neptune_ir.apply`: the loop nest of
lib/Passes/DataflowLowering.cpp:258-448 hand-emitted as LLVM IR and JIT-compiled by LLVM itself (llvmlite, LLVM 20) —
the closest thing this image has to the reference's own back half, `mlir-translate --mlir-to-llvmir` + `llc`
(test/smoke_tests/smoke_apply.sh:22-29):

  * result = malloc'd buffer, memcpy of input 0 (memref.alloc + memref.copy, :273-287),
  * one header/body/latch loop per dimension with an i64 phi, constant bounds (scf.for -> cf),
  * memref.load/store on identity-layout row-major memrefs = linear index ((i0*n1)+i1)*n2+i2 with mul/add i64,
    getelementptr, scalar load/store (:381-411, :427-444),
  * arith.mulf/addf/subf/divf/negf -> fmul/fadd/fsub/fdiv/fneg with NO fast-math flags, so LLVM may neither contract
    nor reassociate; the module is compiled with the "generic" CPU the reference asks for (NeptuneCompiler.cpp:337),
    i.e. no FMA instructions exist to contract into.

It shares only the parser (oracle/mlir_subset.py) with the numpy restatement; the evaluation is LLVM's.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Sequence

import numpy as np

from . import mlir_subset as ms

_NP = {"f64": np.float64, "f32": np.float32}
_LL = {"f64": "double", "f32": "float"}


def _fconst(v: float, ty: str) -> str:
    """An exact LLVM float literal (hex form of the double that holds the value; f32 values are exactly representable)."""
    import struct
    if ty == "f32":
        v = float(np.float32(v))
    return "0x%016X" % struct.unpack("<Q", struct.pack("<d", v))[0]


def emit_apply_ir(ap: ms.Apply, name: str = "apply") -> str:
    """LLVM IR of `void name(ptr out, ptr in0, ptr in1, ...)` for one apply."""
    ty = ap.out_type.elem
    T = _LL[ty]
    S = 8 if ty == "f64" else 4
    rank = ap.bounds.rank
    oshape = ap.out_type.bounds.shape
    numel = int(np.prod(oshape))
    args = ", ".join(["ptr noalias %out"] + ["ptr noalias readonly %in" + str(k) for k in range(len(ap.inputs))])
    L: List[str] = []
    L.append("declare void @llvm.memcpy.p0.p0.i64(ptr, ptr, i64, i1)")
    L.append("define void @%s(%s) {" % (name, args))
    L.append("entry:")
    L.append("  call void @llvm.memcpy.p0.p0.i64(ptr %%out, ptr %%in0, i64 %d, i1 false)" % (numel * S))
    lb, ub = ap.bounds.lb, ap.bounds.ub
    empty = any(u <= l for l, u in zip(lb, ub))
    if empty:
        L.append("  ret void")
        L.append("}")
        return "\n".join(L)
    L.append("  br label %h0")
    arg_index = {n: k for k, n in enumerate(ap.input_args)}
    # loop headers
    for d in range(rank):
        prev = "entry" if d == 0 else "h%d" % (d - 1)
        # phi: from the enclosing block on entry, from the latch on repeat
        L.append("h%d:" % d)
        L.append("  %%i%d = phi i64 [ %d, %%%s ], [ %%i%d.next, %%latch%d ]" % (d, lb[d], ("entry" if d == 0 else "body%d" % (d - 1)), d, d))
        L.append("  %%c%d = icmp slt i64 %%i%d, %d" % (d, d, ub[d]))
        L.append("  br i1 %%c%d, label %%body%d, label %%exit%d" % (d, d, d))
        L.append("body%d:" % d)
        if d + 1 < rank:
            L.append("  br label %%h%d" % (d + 1))
    # innermost body
    tmp = [0]

    def fresh(p="t"):
        tmp[0] += 1
        return "%%%s%d" % (p, tmp[0])

    def linear(idx_terms, shape):
        """((i0*n1)+i1)*n2+i2 with mul/add i64; idx_terms are SSA names of the per-dim indices."""
        acc = idx_terms[0]
        for d in range(1, len(shape)):
            m = fresh("m")
            L.append("  %s = mul i64 %s, %d" % (m, acc, shape[d]))
            a = fresh("a")
            L.append("  %s = add i64 %s, %s" % (a, m, idx_terms[d]))
            acc = a
        return acc
    env: Dict[str, str] = {}
    for op in ap.body:
        if op.name == "access":
            k = arg_index[op.operands[0]]
            in_lb = ap.in_types[k].bounds.lb
            ishape = ap.in_types[k].bounds.shape
            terms = []
            for d in range(rank):
                x = fresh("x")
                L.append("  %s = add i64 %%i%d, %d" % (x, d, op.offsets[d] - in_lb[d]))   # iv + off - inLb (:381-411)
                terms.append(x)
            li = linear(terms, ishape)
            p = fresh("p")
            L.append("  %s = getelementptr inbounds %s, ptr %%in%d, i64 %s" % (p, T, k, li))
            v = fresh("v")
            L.append("  %s = load %s, ptr %s" % (v, T, p))
            env[op.result] = v
        elif op.name == "constant":
            env[op.result] = _fconst(op.value, ty)
        elif op.name in ("addf", "subf", "mulf", "divf"):
            v = fresh("v")
            L.append("  %s = f%s %s %s, %s" % (v, op.name[:3], T, env[op.operands[0]], env[op.operands[1]]))
            env[op.result] = v
        elif op.name == "negf":
            v = fresh("v")
            L.append("  %s = fneg %s %s" % (v, T, env[op.operands[0]]))
            env[op.result] = v
        else:
            raise ms.MlirError("operation %r is outside the oracle's apply-body subset" % op.name)
    olb = ap.out_type.bounds.lb
    terms = []
    for d in range(rank):
        x = fresh("x")
        L.append("  %s = sub i64 %%i%d, %d" % (x, d, olb[d]))                              # iv - outLb (:436-441)
        terms.append(x)
    li = linear(terms, oshape)
    p = fresh("p")
    L.append("  %s = getelementptr inbounds %s, ptr %%out, i64 %s" % (p, T, li))
    L.append("  store %s %s, ptr %s" % (T, env[ap.yield_value], p))
    # latches and exits, innermost first
    for d in reversed(range(rank)):
        L.append("  br label %%latch%d" % d)
        L.append("latch%d:" % d)
        L.append("  %%i%d.next = add i64 %%i%d, 1" % (d, d))
        L.append("  br label %%h%d" % d)
        L.append("exit%d:" % d)
        if d == 0:
            L.append("  ret void")
    L.append("}")
    return "\n".join(L)


class LlvmApply:
    """One apply JIT-compiled by LLVM for the generic x86-64 CPU (no FMA available, none requested)."""
    _engines: List[object] = []

    def __init__(self, ap: ms.Apply):
        import llvmlite.binding as llvm
        llvm.initialize_native_target()
        llvm.initialize_native_asmprinter()
        self.ap = ap
        self.ir = emit_apply_ir(ap)
        mod = llvm.parse_assembly(self.ir)
        mod.verify()
        target = llvm.Target.from_default_triple()
        tm = target.create_target_machine(cpu="generic", features="", opt=2)   # codegen -O2, no IR optimisation pipeline
        self.engine = llvm.create_mcjit_compiler(mod, tm)
        self.engine.finalize_object()
        self.asm = tm.emit_assembly(mod)
        addr = self.engine.get_function_address("apply")
        n_in = len(ap.inputs)
        self.fn = ctypes.CFUNCTYPE(None, *([ctypes.c_void_p] * (1 + n_in)))(addr)
        LlvmApply._engines.append(self.engine)   # keep the code alive

    def __call__(self, inputs: Sequence[np.ndarray]) -> np.ndarray:
        dt = _NP[self.ap.out_type.elem]
        out = np.empty(self.ap.out_type.bounds.shape, dtype=dt)
        ins = [np.ascontiguousarray(a, dtype=dt) for a in inputs]
        for a, t in zip(ins, self.ap.in_types):
            if tuple(a.shape) != tuple(t.bounds.shape):
                raise ValueError("input shape %s does not match %s" % (a.shape, t.bounds.shape))
        self.fn(out.ctypes.data, *[a.ctypes.data for a in ins])
        return out


def run_function(mod: ms.Module, name: str, args: Sequence[np.ndarray], _cache: Dict[int, LlvmApply] = {}) -> List[np.ndarray]:
    """As oracle.numpy_lowering.run_function, with every apply evaluated by LLVM-compiled code."""
    fn = mod.functions[name]
    if fn.rejected:
        raise ms.MlirError("%s is outside the oracle's subset: %s" % (name, fn.rejected))
    env: Dict[str, np.ndarray] = {an: a for (an, _), a in zip(fn.args, args)}
    for op in fn.ops:
        if op.name in ("wrap", "unwrap", "load"):
            env[op.results[0]] = env[op.operands[0]]
        elif op.name == "apply":
            key = id(op.apply)
            if key not in _cache:
                _cache[key] = LlvmApply(op.apply)
            env[op.results[0]] = _cache[key]([env[v] for v in op.apply.inputs])
        elif op.name == "apply_linear":
            env[op.results[0]] = run_function(mod, op.callee, [env[v] for v in op.operands])[0]
        elif op.name == "store":
            src, dst = env[op.operands[0]], env[op.operands[1]]
            if op.bounds is None:
                np.copyto(dst, src)
            else:
                slb, dlb = op.types[0].bounds.lb, op.types[1].bounds.lb
                ssl = tuple(slice(l - b, u - b) for l, u, b in zip(op.bounds.lb, op.bounds.ub, slb))
                dsl = tuple(slice(l - b, u - b) for l, u, b in zip(op.bounds.lb, op.bounds.ub, dlb))
                dst[dsl] = src[ssl]
        elif op.name == "return":
            return [env[v] for v in op.operands if v in env]
        elif op.name == "constant":
            pass
        else:
            raise ms.MlirError("host op %r is not evaluated by the LLVM restatement" % op.name)
    return []


def run_steps(mod: ms.Module, u: np.ndarray, extra: Sequence[np.ndarray], nsteps: int, entry: str = "@entry") -> np.ndarray:
    a = np.array(u, copy=True)
    b = np.empty_like(a)
    for _ in range(nsteps):
        run_function(mod, entry, [b, a] + list(extra))
        a, b = b, a
    return a
