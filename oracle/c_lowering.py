"""ORACLE — TEST INFRASTRUCTURE ONLY (parity unpinned, see oracle/README.md).

Restates the reference's CPU lowering as plain C, one C function per func.func / opdef, and
compiles it with gcc.  This is the closest thing to "neptune-opt --neptuneir-to-llvm | llc" that
can exist in a container without LLVM/MLIR: the C text below has the same
structure the reference emits, statement for statement —

  apply  (lib/Passes/DataflowLowering.cpp:258-448):
      out = malloc(numel*sizeof(T));                     memref.alloc        :273-281
      memcpy(out, in0, ...);                             memref.copy         :282-287
      for i0 in [lb0,ub0) ... for iR in [lbR,ubR):       scf.for nest        :289-308
          v = in_k[linear(iv + off - inLb_k)]            memref.load         :381-411
          scalar ops, one C statement per body op, in body order             :413-423
          out[linear(iv - outLb)] = yielded              memref.store        :427-444
  store  (:165-220)  memcpy / strided box copy
  wrap / unwrap / load forward the pointer (:131-159)

Build flags make gcc behave like the reference's code generator for this path: scalar IEEE
arithmetic, no contraction (`-ffp-contract=off`), no vectoriser, no fast-math; one thread, as
the reference has no parallel construct at all.  The reference never frees the
apply result; here it is freed at function exit so repeated timing runs stay bounded — the
per-sweep malloc + two full copies, which dominate the reference's extra traffic, are kept.

`emit_openmp_steps` is NOT the reference: it is a best-effort CPU context number (ping-pong
buffers, no redundant copies, OpenMP over the outer loop), reported next to the faithful one.
"""
from __future__ import annotations

import ctypes
import hashlib
import os
import subprocess
from typing import Dict, List, Sequence

import numpy as np

from . import mlir_subset as ms

_CT = {"f64": "double", "f32": "float"}
_HERE = os.path.dirname(os.path.abspath(__file__))
BUILD_DIR = os.path.join(_HERE, "_build")

FAITHFUL_FLAGS = ["-O2", "-fno-tree-vectorize", "-ffp-contract=off", "-fno-fast-math"]
OPENMP_FLAGS = ["-O3", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-march=native"]


def _lit(v: float, ty: str) -> str:
    if ty == "f32":
        return float(np.float32(v)).hex() + "f"
    return float(v).hex()


def _lin(idx: Sequence[str], shape: Sequence[int]) -> str:
    e = idx[0]
    for d in range(1, len(shape)):
        e = "(%s) * %dL + (%s)" % (e, shape[d], idx[d])
    return e


def _sym(v: str) -> str:
    return "v_" + v[1:].replace(".", "_").replace("-", "_").replace("$", "_")


def _emit_body(ap: ms.Apply, ins: Sequence[str], out: str, ind: str) -> List[str]:
    T = _CT[ap.out_type.elem]
    rank = ap.bounds.rank
    arg_index = {n: k for k, n in enumerate(ap.input_args)}
    L: List[str] = []
    for op in ap.body:
        r = _sym(op.result)
        if op.name == "access":
            k = arg_index[op.operands[0]]
            b = ap.in_types[k].bounds
            idx = ["i%d + (%d) - (%d)" % (d, op.offsets[d], b.lb[d]) for d in range(rank)]
            L.append("%sconst %s %s = %s[%s];" % (ind, T, r, ins[k], _lin(idx, b.shape)))
        elif op.name == "constant":
            L.append("%sconst %s %s = %s;" % (ind, _CT[op.ty], r, _lit(op.value, op.ty)))
        elif op.name in ("addf", "subf", "mulf", "divf"):
            c = {"addf": "+", "subf": "-", "mulf": "*", "divf": "/"}[op.name]
            L.append("%sconst %s %s = %s %s %s;" % (ind, T, r, _sym(op.operands[0]), c, _sym(op.operands[1])))
        elif op.name == "negf":
            L.append("%sconst %s %s = -%s;" % (ind, T, r, _sym(op.operands[0])))
        else:
            raise NotImplementedError(op.name)
    ob = ap.out_type.bounds
    oidx = ["i%d - (%d)" % (d, ob.lb[d]) for d in range(rank)]
    L.append("%s%s[%s] = %s;" % (ind, out, _lin(oidx, ob.shape), _sym(ap.yield_value)))
    return L


def _emit_apply(ap: ms.Apply, ins: Sequence[str], out: str) -> List[str]:
    T = _CT[ap.out_type.elem]
    numel = int(np.prod(ap.out_type.bounds.shape))
    rank = ap.bounds.rank
    L = ["  %s *%s = (%s *)malloc(%dL * sizeof(%s));" % (T, out, T, numel, T),
         "  memcpy(%s, %s, %dL * sizeof(%s));" % (out, ins[0], numel, T)]
    ind = "  "
    for d in range(rank):
        L.append("%sfor (long i%d = %d; i%d < %d; ++i%d) {" % (ind, d, ap.bounds.lb[d], d, ap.bounds.ub[d], d))
        ind += "  "
    L += _emit_body(ap, ins, out, ind)
    for d in range(rank):
        ind = ind[:-2]
        L.append(ind + "}")
    return L


def emit_module(mod: ms.Module) -> str:
    """C translation unit with one function per func.func/opdef; memrefs are bare pointers."""
    src = ["#include <stdlib.h>", "#include <string.h>", ""]
    protos, bodies = [], []
    for fn in mod.functions.values():
        cname = "nep_" + fn.name[1:]
        elem = None
        for _, ty in fn.args:
            elem = ty.elem
        T = _CT[elem]
        # a function whose return operand is a `reduce` result returns the scalar by value
        scalars = {op.results[0] for op in fn.ops if op.name == "reduce"}
        rets = [op for op in fn.ops if op.name == "return"]
        scalar_ret = bool(rets and rets[0].operands and rets[0].operands[0] in scalars)
        sig = "%s %s%s(%s)" % (T, "" if scalar_ret else "*", cname,
                               ", ".join("%s *%s" % (T, _sym(a)) for a, _ in fn.args))
        protos.append(sig + ";")
        L = [sig + " {"]
        frees: List[str] = []
        ret = "0"
        for op in fn.ops:
            if op.name in ("wrap", "unwrap", "load"):
                L.append("  %s *%s = %s;" % (T, _sym(op.results[0]), _sym(op.operands[0])))
            elif op.name == "apply":
                out = _sym(op.results[0])
                L += _emit_apply(op.apply, [_sym(v) for v in op.apply.inputs], out)
                frees.append(out)
            elif op.name == "apply_linear":
                L.append("  %s *%s = nep_%s(%s);" % (T, _sym(op.results[0]), op.callee[1:],
                                                      ", ".join(_sym(v) for v in op.operands)))
                frees.append(_sym(op.results[0]))
            elif op.name == "store":
                sb, db = op.types[0].bounds, op.types[1].bounds
                if op.bounds is None:
                    L.append("  memcpy(%s, %s, %dL * sizeof(%s));" % (
                        _sym(op.operands[1]), _sym(op.operands[0]), int(np.prod(sb.shape)), T))
                else:
                    rank = sb.rank
                    ind = "  "
                    for d in range(rank):
                        L.append("%sfor (long i%d = %d; i%d < %d; ++i%d) {" % (
                            ind, d, op.bounds.lb[d], d, op.bounds.ub[d], d))
                        ind += "  "
                    si = ["i%d - (%d)" % (d, sb.lb[d]) for d in range(rank)]
                    di = ["i%d - (%d)" % (d, db.lb[d]) for d in range(rank)]
                    L.append("%s%s[%s] = %s[%s];" % (ind, _sym(op.operands[1]), _lin(di, db.shape),
                                                      _sym(op.operands[0]), _lin(si, sb.shape)))
                    for d in range(rank):
                        ind = ind[:-2]
                        L.append(ind + "}")
            elif op.name == "reduce":
                # ReduceLowering / buildReduceLoops (DataflowLowering.cpp:589-698): acc = 0, then one
                # scf.for per dimension carrying acc, innermost `acc = acc + load` in the element type
                if op.callee != "sum":
                    raise NotImplementedError("reduce kind %r" % op.callee)
                sb = op.types[0].bounds
                dom = op.bounds if op.bounds is not None else sb
                acc = _sym(op.results[0])
                L.append("  %s %s = 0;" % (T, acc))
                ind = "  "
                for d in range(sb.rank):
                    L.append("%sfor (long i%d = %d; i%d < %d; ++i%d) {" % (ind, d, dom.lb[d], d, dom.ub[d], d))
                    ind += "  "
                si = ["i%d - (%d)" % (d, sb.lb[d]) for d in range(sb.rank)]
                L.append("%s%s = %s + %s[%s];" % (ind, acc, acc, _sym(op.operands[0]), _lin(si, sb.shape)))
                for d in range(sb.rank):
                    ind = ind[:-2]
                    L.append(ind + "}")
            elif op.name == "constant":
                pass
            elif op.name == "return":
                ret = _sym(op.operands[0]) if op.operands else "0"
        # opdefs hand their malloc'd result to the caller; a func.func returns a caller buffer
        for f in frees:
            if f != ret:
                L.append("  free(%s);" % f)
        L.append("  return %s;" % ret)
        L.append("}")
        bodies.append("\n".join(L))
    return "\n".join(src + protos + [""] + bodies) + "\n"


def emit_openmp_steps(mod: ms.Module, entry: str = "@entry") -> str:
    """Context-only CPU number (NOT the reference): ping-pong sweeps, OpenMP outer loop.
    Only for an entry of the shape wrap/load/apply/store/unwrap with a single apply."""
    fn = mod.functions[entry]
    aps = [o.apply for o in fn.ops if o.name == "apply"]
    if len(aps) != 1:
        raise ValueError("openmp context kernel needs exactly one apply")
    ap = aps[0]
    T = _CT[ap.out_type.elem]
    rank = ap.bounds.rank
    n_in = len(ap.inputs)
    args = ", ".join("const %s *restrict in%d" % (T, k) for k in range(n_in))
    L = ["#include <stdlib.h>", "#include <string.h>",
         "void nep_omp_sweep(%s *restrict out, %s) {" % (T, args),
         "  #pragma omp parallel for schedule(static)"]
    ind = "  "
    for d in range(rank):
        L.append("%sfor (long i%d = %d; i%d < %d; ++i%d) {" % (ind, d, ap.bounds.lb[d], d, ap.bounds.ub[d], d))
        ind += "  "
    L += _emit_body(ap, ["in%d" % k for k in range(n_in)], "out", ind)
    for d in range(rank):
        ind = ind[:-2]
        L.append(ind + "}")
    L.append("}")
    return "\n".join(L) + "\n"


def compile_c(src: str, flags: Sequence[str], tag: str) -> str:
    os.makedirs(BUILD_DIR, exist_ok=True)
    h = hashlib.sha256((src + " ".join(flags)).encode()).hexdigest()[:16]
    so = os.path.join(BUILD_DIR, "%s_%s.so" % (tag, h))
    if not os.path.exists(so):
        cfile = so[:-3] + ".c"
        with open(cfile, "w") as f:
            f.write(src)
        subprocess.check_call(["gcc", "-shared", "-fPIC", "-std=c11"] + list(flags) + [cfile, "-o", so + ".tmp"])
        os.replace(so + ".tmp", so)
    return so


class CompiledModule:
    """gcc-compiled faithful lowering of a module; call functions on numpy host buffers."""

    def __init__(self, text: str, tag: str = "oracle"):
        self.mod = ms.parse(text)
        self.lib = ctypes.CDLL(compile_c(emit_module(self.mod), FAITHFUL_FLAGS, tag))

    def call(self, name: str, args: Sequence[np.ndarray]):
        fn = getattr(self.lib, "nep_" + name[1:])
        fn.restype = ctypes.c_void_p
        fn.argtypes = [ctypes.c_void_p] * len(args)
        for a in args:
            assert a.flags["C_CONTIGUOUS"]
        return fn(*[a.ctypes.data for a in args])

    def call_scalar(self, name: str, args: Sequence[np.ndarray]) -> float:
        """Call a function that returns a `reduce` result by value."""
        fn = getattr(self.lib, "nep_" + name[1:])
        fn.restype = ctypes.c_double if args[0].dtype == np.float64 else ctypes.c_float
        fn.argtypes = [ctypes.c_void_p] * len(args)
        return fn(*[a.ctypes.data for a in args])

    def run_steps(self, u: np.ndarray, extra: Sequence[np.ndarray], nsteps: int, entry="@entry") -> np.ndarray:
        a = np.array(u, copy=True)
        b = np.empty_like(a)
        for _ in range(nsteps):
            self.call(entry, [b, a] + list(extra))
            a, b = b, a
        return a


class OpenMPSweep:
    def __init__(self, text: str, tag: str = "oracle_omp"):
        self.mod = ms.parse(text)
        self.lib = ctypes.CDLL(compile_c(emit_openmp_steps(self.mod), OPENMP_FLAGS, tag))

    def sweep(self, out: np.ndarray, ins: Sequence[np.ndarray]):
        fn = self.lib.nep_omp_sweep
        fn.restype = None
        fn.argtypes = [ctypes.c_void_p] * (1 + len(ins))
        fn(out.ctypes.data, *[a.ctypes.data for a in ins])
