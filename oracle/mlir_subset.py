"""ORACLE — TEST INFRASTRUCTURE ONLY (parity unpinned, see oracle/README.md).

Parser for the textual NeptuneIR subset that reaches the `neptune_ir.apply` hot path.
Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may import this;
the product path (neptune_pde_solver_b200/) has its own, independent C++ front-end.

Grammar follows the reference's op definitions:
  * types `!neptune_ir.field/temp<element = T, bounds = B, location = L>`
      include/Dialect/NeptuneIR/NeptuneIRTypes.td:12-59
  * attrs `#neptune_ir.bounds<lb = [..], ub = [..]>`, `#neptune_ir.location<"cell">`,
      `#neptune_ir.stencil_shape<[...]>`        include/Dialect/NeptuneIR/NeptuneIRAttrs.td:9-76
  * ops wrap/unwrap/load/yield/apply/access/store  include/Dialect/NeptuneIR/NeptuneIROps.td:20-259
  * apply_linear / linear_opdef / return (callers of the hot path, lowered to func.call /
      func.func by lib/Passes/StructureLowering.cpp:31-140)
  * arith.constant / addf / subf / mulf / divf / negf (the body whitelist of
      lib/Passes/VerifyAndAnnotate.cpp:167-204, float part)
and the region contract of `ApplyOp::verify` (lib/Dialect/NeptuneIR/NeptuneIRVerifier.cpp:141-171):
`rank` index block arguments, then one argument per input with the operand's type.
"""
from __future__ import annotations

import re

import numpy as np
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple


class MlirError(Exception):
    pass


@dataclass(frozen=True)
class Bounds:
    lb: Tuple[int, ...]
    ub: Tuple[int, ...]

    @property
    def rank(self):
        return len(self.lb)

    @property
    def shape(self):
        return tuple(u - l for l, u in zip(self.lb, self.ub))


@dataclass(frozen=True)
class GridType:          # !neptune_ir.field / !neptune_ir.temp
    kind: str            # "field" | "temp"
    elem: str            # "f64" | "f32"
    bounds: Bounds
    location: str


@dataclass(frozen=True)
class MemRefType:
    shape: Tuple[Optional[int], ...]
    elem: str


@dataclass
class BodyOp:
    name: str                      # access | constant | addf | subf | mulf | divf | negf
    result: str
    operands: Tuple[str, ...] = ()
    offsets: Tuple[int, ...] = ()  # access
    value: float = 0.0             # constant
    ty: str = "f64"


@dataclass
class Apply:
    result: str
    inputs: Tuple[str, ...]
    in_types: Tuple[GridType, ...]
    out_type: GridType
    bounds: Bounds
    shape_attr: Optional[Tuple[Tuple[int, ...], ...]]
    index_args: Tuple[str, ...]
    input_args: Tuple[str, ...]
    body: List[BodyOp]
    yield_value: str


@dataclass
class HostOp:
    name: str                      # wrap | unwrap | load | store | apply | apply_linear | return
    results: Tuple[str, ...] = ()
    operands: Tuple[str, ...] = ()
    types: Tuple[object, ...] = ()
    apply: Optional[Apply] = None
    callee: Optional[str] = None
    bounds: Optional[Bounds] = None


@dataclass
class Function:
    name: str
    kind: str                      # "func" | "linear_opdef" | "nonlinear_opdef"
    args: List[Tuple[str, object]]
    result_types: List[object]
    ops: List[HostOp] = field(default_factory=list)
    rejected: Optional[str] = None   # why the function is outside this restatement's subset (None: it is inside)


@dataclass
class Module:
    functions: Dict[str, Function]


_TOKEN = re.compile(r"""
    (?P<ws>\s+|//[^\n]*)
  | (?P<str>"[^"]*")
  | (?P<memref>memref<[^>]*>)
  | (?P<arrow>->)
  | (?P<id>[%@^\#!][A-Za-z0-9_.$\-]*[A-Za-z0-9_$])
  | (?P<num>[-+]?(?:\d+\.\d*(?:[eE][-+]?\d+)?|\d+[eE][-+]?\d+|\.\d+(?:[eE][-+]?\d+)?|0x[0-9a-fA-F]+|\d+|inf|nan))
  | (?P<word>[A-Za-z_][A-Za-z0-9_.]*)
  | (?P<punct>[{}()\[\]<>,:=?*x])
""", re.X)


def _tokenize(text: str) -> List[Tuple[str, str]]:
    out, pos = [], 0
    while pos < len(text):
        m = _TOKEN.match(text, pos)
        if not m:
            raise MlirError("unexpected character %r at offset %d" % (text[pos], pos))
        pos = m.end()
        k = m.lastgroup
        if k == "ws":
            continue
        out.append((k, m.group(k)))
    out.append(("eof", ""))
    return out


class _Parser:
    def __init__(self, text: str):
        # shaped types like memref<?x?xf64> / 512x512xf64 need a char-level scan; pre-extract
        self.toks = _tokenize(text)
        self.i = 0
        self.attr_alias: Dict[str, object] = {}
        self.type_alias: Dict[str, object] = {}

    # -- token helpers -------------------------------------------------------------------
    def peek(self, k=0):
        return self.toks[self.i + k]

    def next(self):
        t = self.toks[self.i]
        self.i += 1
        return t

    def accept(self, val):
        if self.peek()[1] == val:
            self.i += 1
            return True
        return False

    def expect(self, val):
        t = self.next()
        if t[1] != val:
            raise MlirError("expected %r, got %r (token %d)" % (val, t[1], self.i))
        return t

    # -- attributes / types --------------------------------------------------------------
    def int_list(self) -> Tuple[int, ...]:
        self.expect("[")
        vals = []
        if not self.accept("]"):
            while True:
                vals.append(int(self.next()[1]))
                if self.accept("]"):
                    break
                self.expect(",")
        return tuple(vals)

    def parse_bounds(self) -> Bounds:
        t = self.next()[1]
        if t in self.attr_alias:
            b = self.attr_alias[t]
            if not isinstance(b, Bounds):
                raise MlirError("%s is not a bounds attribute" % t)
            return b
        if t != "#neptune_ir.bounds":
            raise MlirError("expected bounds attribute, got %r" % t)
        self.expect("<")
        self.expect("lb"); self.expect("="); lb = self.int_list()
        self.expect(",")
        self.expect("ub"); self.expect("="); ub = self.int_list()
        self.expect(">")
        if len(lb) != len(ub):
            raise MlirError("bounds lb/ub rank mismatch")
        return Bounds(lb, ub)

    def parse_location(self) -> str:
        t = self.next()[1]
        if t in self.attr_alias:
            return self.attr_alias[t]
        if t != "#neptune_ir.location":
            raise MlirError("expected location attribute, got %r" % t)
        self.expect("<")
        s = self.next()[1].strip('"')
        self.expect(">")
        return s

    def parse_stencil_shape(self):
        t = self.next()[1]
        if t != "#neptune_ir.stencil_shape":
            raise MlirError("expected stencil_shape attribute")
        self.expect("<"); self.expect("[")
        offs = []
        while not self.accept("]"):
            if self.peek()[1] == "array":
                self.next(); self.expect("<"); self.next(); self.expect(":")  # array<i64: ...>
                v = []
                while not self.accept(">"):
                    v.append(int(self.next()[1])); self.accept(",")
                offs.append(tuple(v))
            else:
                offs.append(self.int_list())
            self.accept(",")
        self.expect(">")
        return tuple(offs)

    def parse_attr_value(self):
        t = self.peek()[1]
        if t == "#neptune_ir.bounds":
            return self.parse_bounds()
        if t == "#neptune_ir.location":
            return self.parse_location()
        if t == "#neptune_ir.stencil_shape":
            return self.parse_stencil_shape()
        if t in self.attr_alias:
            self.next()
            return self.attr_alias[t]
        raise MlirError("unsupported attribute %r" % t)

    def parse_type(self):
        t = self.next()[1]
        if t in self.type_alias:
            return self.type_alias[t]
        if t in ("f64", "f32", "index", "i32", "i64"):
            return t
        if t in ("!neptune_ir.temp", "!neptune_ir.field"):
            self.expect("<")
            self.expect("element"); self.expect("="); elem = self.parse_type()
            self.expect(",")
            self.expect("bounds"); self.expect("="); b = self.parse_bounds()
            loc = "cell"
            if self.accept(","):
                self.expect("location"); self.expect("="); loc = self.parse_location()
            self.expect(">")
            return GridType(t.split(".")[1], elem, b, loc)
        if t.startswith("memref<"):
            parts = t[len("memref<"):-1].replace(" ", "").split("x")
            if parts[-1] not in ("f64", "f32"):
                raise MlirError("bad memref element %r" % parts[-1])
            dims = tuple(None if d == "?" else int(d) for d in parts[:-1])
            return MemRefType(dims, parts[-1])
        raise MlirError("unsupported type %r" % t)

    # -- module --------------------------------------------------------------------------
    def parse_module(self) -> Module:
        funcs: Dict[str, Function] = {}
        while self.peek()[0] != "eof":
            t = self.peek()
            if t[0] == "id" and t[1][0] == "#" and self.peek(1)[1] == "=":
                name = self.next()[1]; self.expect("=")
                self.attr_alias[name] = self.parse_attr_value()
            elif t[0] == "id" and t[1][0] == "!" and self.peek(1)[1] == "=":
                name = self.next()[1]; self.expect("=")
                self.type_alias[name] = self.parse_type()
            elif t[1] == "module":
                self.next()
                self.expect("{")
                while not self.accept("}"):
                    f = self.parse_function()
                    funcs[f.name] = f
            else:
                f = self.parse_function()
                funcs[f.name] = f
        return Module(funcs)

    def parse_function(self) -> Function:
        t = self.next()[1]
        if t == "func.func":
            name = self.next()[1]
            self.expect("(")
            args = []
            if not self.accept(")"):
                while True:
                    a = self.next()[1]; self.expect(":"); ty = self.parse_type()
                    args.append((a, ty))
                    if self.accept(")"):
                        break
                    self.expect(",")
            res = []
            if self.accept("->"):
                if self.accept("("):
                    while not self.accept(")"):
                        res.append(self.parse_type()); self.accept(",")
                else:
                    res.append(self.parse_type())
            fn = Function(name, "func", args, res)
            self.expect("{")
            self.parse_host_block(fn)
            return fn
        if t in ("neptune_ir.linear_opdef", "neptune_ir.nonlinear_opdef"):
            name = self.next()[1]
            self.expect(":")
            self.expect("(")
            in_tys = []
            while not self.accept(")"):
                in_tys.append(self.parse_type()); self.accept(",")
            self.expect("->")
            res = [self.parse_type()]
            self.expect("{")
            lab = self.next()[1]
            if not lab.startswith("^"):
                raise MlirError("opdef region must start with a block label")
            self.expect("(")
            args = []
            while not self.accept(")"):
                a = self.next()[1]; self.expect(":"); ty = self.parse_type()
                args.append((a, ty)); self.accept(",")
            self.expect(":")
            if [ty for _, ty in args] != in_tys:
                raise MlirError("opdef block argument types differ from its signature")
            fn = Function(name, t.split(".")[1], args, res)
            self.parse_host_block(fn)
            return fn
        raise MlirError("unsupported top-level op %r" % t)

    def parse_host_block(self, fn: Function):
        """Ops up to the function's closing brace.  A function that holds an op outside this restatement's subset
        (neptune_ir.time_advance, solve_linear, ...) is kept as a REJECTED function carrying the error, and parsing
        resumes after its closing brace, so that the operator definitions of the reference's own fixtures
        (test/smoke_tests/smoke_time_advance.mlir) can be evaluated although their @entry cannot."""
        start = self.i
        try:
            while not self.accept("}"):
                fn.ops.append(self.parse_host_op())
        except MlirError as e:
            fn.ops = []
            fn.rejected = str(e)
            self.i = start
            depth = 1
            while depth > 0:
                t = self.next()
                if t[0] == "eof":
                    raise MlirError("%s (and the function's closing brace is missing)" % e)
                if t[1] == "{":
                    depth += 1
                elif t[1] == "}":
                    depth -= 1

    def parse_host_op(self) -> HostOp:
        results: Tuple[str, ...] = ()
        if self.peek()[0] == "id" and self.peek()[1][0] == "%" and self.peek(1)[1] == "=":
            results = (self.next()[1],)
            self.expect("=")
        op = self.next()[1]
        if op == "neptune_ir.wrap" or op == "neptune_ir.unwrap" or op == "neptune_ir.load":
            src = self.next()[1]
            self.expect(":"); t0 = self.parse_type(); self.expect("->"); t1 = self.parse_type()
            return HostOp(op.split(".")[1], results, (src,), (t0, t1))
        if op == "neptune_ir.store":
            v = self.next()[1]; self.expect("to"); f = self.next()[1]
            b = None
            if self.accept("{"):
                self.expect("bounds"); self.expect("="); b = self.parse_bounds(); self.expect("}")
            self.expect(":"); t0 = self.parse_type(); self.expect("to"); t1 = self.parse_type()
            return HostOp("store", (), (v, f), (t0, t1), bounds=b)
        if op == "neptune_ir.apply_linear" or op == "neptune_ir.apply_nonlinear":
            callee = self.next()[1]
            self.expect("(")
            ops = []
            while not self.accept(")"):
                ops.append(self.next()[1]); self.accept(",")
            self.expect(":"); self.expect("(")
            tys = []
            while not self.accept(")"):
                tys.append(self.parse_type()); self.accept(",")
            self.expect("->"); rt = self.parse_type()
            return HostOp("apply_linear", results, tuple(ops), tuple(tys) + (rt,), callee=callee)
        if op in ("func.return", "return", "neptune_ir.return"):
            ops = []
            if self.peek()[0] == "id" and self.peek()[1][0] == "%":
                while True:
                    ops.append(self.next()[1])
                    if not self.accept(","):
                        break
                self.expect(":")
                tys = [self.parse_type()]
                while self.accept(","):
                    tys.append(self.parse_type())
            return HostOp("return", (), tuple(ops))
        if op == "neptune_ir.reduce":        # include/Dialect/NeptuneIR/NeptuneIROps.td:272-299
            src = self.next()[1]
            b = None
            if self.accept("in"):
                b = self.parse_bounds()
            kind = None
            if self.accept("{"):
                self.expect("kind"); self.expect("="); kind = self.next()[1].strip('"'); self.expect("}")
            self.expect(":"); t0 = self.parse_type(); self.expect("->"); t1 = self.parse_type()
            return HostOp("reduce", results, (src,), (t0, t1), callee=kind, bounds=b)
        if op == "neptune_ir.apply":
            return HostOp("apply", results, apply=self.parse_apply(results[0]))
        if op == "arith.constant":          # host-level scalars (e.g. dt) never reach the sweep
            self.next(); self.expect(":"); self.parse_type()
            return HostOp("constant", results)
        raise MlirError("unsupported host-level op %r" % op)

    def parse_apply(self, result: str) -> Apply:
        self.expect("(")
        ins = []
        while not self.accept(")"):
            ins.append(self.next()[1]); self.accept(",")
        self.expect("attributes"); self.expect("{")
        bounds, shape = None, None
        while not self.accept("}"):
            k = self.next()[1]; self.expect("=")
            if k == "bounds":
                bounds = self.parse_bounds()
            elif k == "shape":
                shape = self.parse_stencil_shape()
            else:
                raise MlirError("unsupported apply attribute %r" % k)
            self.accept(",")
        if bounds is None:
            raise MlirError("apply is missing its required 'bounds' attribute")
        self.expect(":"); self.expect("(")
        in_tys = []
        while not self.accept(")"):
            in_tys.append(self.parse_type()); self.accept(",")
        self.expect("->"); out_ty = self.parse_type()
        self.expect("{")
        lab = self.next()[1]
        if not lab.startswith("^"):
            raise MlirError("apply region must start with a block label")
        self.expect("(")
        bargs = []
        while not self.accept(")"):
            a = self.next()[1]; self.expect(":"); ty = self.parse_type()
            bargs.append((a, ty)); self.accept(",")
        self.expect(":")
        rank = bounds.rank
        # ApplyOp::verify — NeptuneIRVerifier.cpp:141-171
        if rank == 0:
            raise MlirError("0-D apply not supported")
        if len(bargs) != rank + len(ins):
            raise MlirError("apply-like region block arg count must be (bounds rank + number of inputs) = %d, but got %d"
                            % (rank + len(ins), len(bargs)))
        for d in range(rank):
            if bargs[d][1] != "index":
                raise MlirError("region arg #%d must be index" % d)
        for k in range(len(ins)):
            if bargs[rank + k][1] != in_tys[k]:
                raise MlirError("region input arg #%d type mismatch" % (rank + k))
        body: List[BodyOp] = []
        yv = None
        while True:
            if self.peek()[1] == "neptune_ir.yield":
                self.next(); yv = self.next()[1]; self.expect(":"); self.parse_type()
                self.expect("}")
                break
            res = self.next()[1]; self.expect("=")
            op = self.next()[1]
            if op == "neptune_ir.access":
                src = self.next()[1]; offs = self.int_list()
                self.expect(":"); self.parse_type(); self.expect("->"); ty = self.parse_type()
                body.append(BodyOp("access", res, (src,), offsets=offs, ty=ty))
            elif op == "arith.constant":
                v = self.next()[1]; self.expect(":"); ty = self.parse_type()
                # f32 literals are rounded ONCE, decimal -> f32, as MLIR's APFloat does (not decimal -> f64 -> f32,
                # which can differ by one ulp on tie cases); the value is carried as the exactly representable double
                val = float(np.float32(v)) if ty == "f32" and not v.startswith("0x") else float(v)
                body.append(BodyOp("constant", res, value=val, ty=ty))
            elif op in ("arith.addf", "arith.subf", "arith.mulf", "arith.divf"):
                a = self.next()[1]; self.expect(","); b = self.next()[1]
                self.expect(":"); ty = self.parse_type()
                body.append(BodyOp(op.split(".")[1], res, (a, b), ty=ty))
            elif op == "arith.negf":
                a = self.next()[1]; self.expect(":"); ty = self.parse_type()
                body.append(BodyOp("negf", res, (a,), ty=ty))
            else:
                raise MlirError("operation %r is outside the oracle's apply-body subset" % op)
        return Apply(result, tuple(ins), tuple(in_tys), out_ty, bounds, shape,
                     tuple(a for a, _ in bargs[:rank]), tuple(a for a, _ in bargs[rank:]), body, yv)


def parse(text: str) -> Module:
    return _Parser(text).parse_module()
