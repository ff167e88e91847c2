"""Authoring helpers: NeptuneIR text for the structured-grid stencil programs of BASELINE.json.

The reference ships only 1-D `.mlir` inputs, so the 2-D/3-D programs the
north star names are written here, in the dialect's own assembly syntax:

* aliases / types / entry-function shape follow the reference's current-verifier form,
  `test/smoke_tests/smoke_time_advance.mlir:3-7` (aliases), `:13-29` (single-input apply with
  interior bounds), `:38-49` (two-input apply) and `:53-57,82-84` (wrap/load/store/unwrap);
* block arguments obey `ApplyOp::verify` (`lib/Dialect/NeptuneIR/NeptuneIRVerifier.cpp:141-171`):
  `rank` index arguments followed by one argument per input;
* the scalar bodies use the reference's own evaluation-order idioms:
  class S "k*centre first, fold neighbours around it, then scale" (`smoke_time_advance.mlir:21-27`),
  class A axpy epilogue (`:45-47`), class L left fold (`smoke_apply.mlir:6-13`).

Every function returns the text of a whole `module { ... }`. Nothing here touches a GPU.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple


def _f(v: float, ty: str) -> str:
    """MLIR float literal that round-trips exactly (repr is shortest round-trip for f64)."""
    if ty == "f32":
        import numpy as np
        v = float(np.float32(v))
    s = repr(float(v))
    if "e" not in s and "." not in s and "inf" not in s and "nan" not in s:
        s += ".0"
    return s


def _bounds(lb: Sequence[int], ub: Sequence[int]) -> str:
    return "#neptune_ir.bounds<lb = [%s], ub = [%s]>" % (
        ", ".join(str(int(x)) for x in lb), ", ".join(str(int(x)) for x in ub))


def _memref(rank: int, ty: str) -> str:
    return "memref<" + "x".join(["?"] * rank) + "x" + ty + ">"


def _header(shape: Sequence[int], ty: str, lb: Sequence[int] | None = None) -> Tuple[str, List[int], List[int]]:
    rank = len(shape)
    lb = [0] * rank if lb is None else [int(x) for x in lb]
    ub = [l + int(n) for l, n in zip(lb, shape)]
    txt = (
        '#loc = #neptune_ir.location<"cell">\n'
        "#b   = %s\n\n"
        "!temp  = !neptune_ir.temp<element = %s, bounds = #b, location = #loc>\n"
        "!field = !neptune_ir.field<element = %s, bounds = #b, location = #loc>\n\n"
    ) % (_bounds(lb, ub), ty, ty)
    return txt, lb, ub


def _off(o: Sequence[int]) -> str:
    return "[" + ", ".join(str(int(x)) for x in o) + "]"


def star_offsets(rank: int) -> List[Tuple[int, ...]]:
    """Neighbour order used by every star fixture: dimension-major, minus before plus."""
    out = []
    for d in range(rank):
        for s in (-1, 1):
            o = [0] * rank
            o[d] = s
            out.append(tuple(o))
    return out


def box_offsets(rank: int) -> List[Tuple[int, ...]]:
    """All 3^rank offsets in lexicographic order (-1,-1,-1) ... (1,1,1)."""
    import itertools
    return [tuple(o) for o in itertools.product((-1, 0, 1), repeat=rank)]


def _entry(rank: int, ty: str, n_extra_inputs: int, apply_text: str, in_names: Iterable[str]) -> str:
    mr = _memref(rank, ty)
    extra = "".join(", %%c%d: %s" % (k, mr) for k in range(n_extra_inputs))
    lines = ["  func.func @entry(%%out: %s, %%in: %s%s) -> %s {" % (mr, mr, extra, mr),
             "    %%fout = neptune_ir.wrap %%out : %s -> !field" % mr,
             "    %%fin  = neptune_ir.wrap %%in  : %s -> !field" % mr,
             "    %u0   = neptune_ir.load %fin : !field -> !temp"]
    for k in range(n_extra_inputs):
        lines.append("    %%fc%d = neptune_ir.wrap %%c%d : %s -> !field" % (k, k, mr))
        lines.append("    %%tc%d = neptune_ir.load %%fc%d : !field -> !temp" % (k, k))
    lines.append(apply_text)
    lines.append("    neptune_ir.store %r to %fout : !temp to !field")
    lines.append("    %%res = neptune_ir.unwrap %%fout : !field -> %s" % mr)
    lines.append("    func.return %%res : %s" % mr)
    lines.append("  }")
    return "\n".join(lines)


def _apply_open(rank: int, n_inputs: int, in_vals: Sequence[str], lb, ub, shape_attr: str = "") -> str:
    idx = ", ".join("%%i%d: index" % d for d in range(rank))
    args = ", ".join("%%a%d: !temp" % k for k in range(n_inputs))
    tys = ", ".join(["!temp"] * n_inputs)
    attrs = "bounds = %s%s" % (_bounds(lb, ub), shape_attr)
    return ("    %%r = neptune_ir.apply(%s) attributes {%s}\n"
            "      : (%s) -> !temp {\n"
            "      ^bb0(%s, %s):") % (", ".join(in_vals), attrs, tys, idx, args)


def _interior(lb, ub, radius=1):
    return [l + radius for l in lb], [u - radius for u in ub]


def _shape_attr(offsets) -> str:
    return ", shape = #neptune_ir.stencil_shape<[%s]>" % ", ".join(
        "array<i64: %s>" % ", ".join(str(x) for x in o) for o in offsets)


def heat_star(shape: Sequence[int], alpha: float, ty: str = "f64", lb=None,
              with_shape_attr: bool = False) -> str:
    """Explicit heat step on a star stencil (1-D 3-pt, 2-D 5-pt, 3-D 7-pt):

        u' = u + alpha * ((((n0 - k*u) + n1) + n2) + ...)      k = 2*rank

    Class S ("k*centre first, neighbours folded around it, then scale",
    reference idiom `smoke_time_advance.mlir:21-27`) followed by class A (axpy on the centre).
    Bounds are the interior shrunk by the radius, as in `smoke_time_advance.mlir:13`.
    """
    rank = len(shape)
    hdr, lb_, ub_ = _header(shape, ty, lb)
    ilb, iub = _interior(lb_, ub_)
    offs = star_offsets(rank)
    sa = _shape_attr([tuple([0] * rank)] + offs) if with_shape_attr else ""
    b = [_apply_open(rank, 1, ["%u0"], ilb, iub, sa)]
    b.append("        %%c = neptune_ir.access %%a0%s : !temp -> %s" % (_off([0] * rank), ty))
    for j, o in enumerate(offs):
        b.append("        %%n%d = neptune_ir.access %%a0%s : !temp -> %s" % (j, _off(o), ty))
    b.append("        %%k = arith.constant %s : %s" % (_f(2.0 * rank, ty), ty))
    b.append("        %%alpha = arith.constant %s : %s" % (_f(alpha, ty), ty))
    b.append("        %%t0 = arith.mulf %%k, %%c : %s" % ty)
    b.append("        %%t1 = arith.subf %%n0, %%t0 : %s" % ty)
    for j in range(1, len(offs)):
        b.append("        %%t%d = arith.addf %%t%d, %%n%d : %s" % (j + 1, j, j, ty))
    last = len(offs)
    b.append("        %%lap = arith.mulf %%alpha, %%t%d : %s" % (last, ty))
    b.append("        %%o = arith.addf %%c, %%lap : %s" % ty)
    b.append("        neptune_ir.yield %%o : %s" % ty)
    b.append("      }")
    return hdr + "module {\n" + _entry(rank, ty, 0, "\n".join(b), ["%u0"]) + "\n}\n"


def heat_box_2d9(shape: Sequence[int], alpha: float, ty: str = "f64", lb=None) -> str:
    """2-D 9-point explicit heat step with the isotropic 4/1 weights (BASELINE C2):

        lap = 4*(N+S+W+E) + (NW+NE+SW+SE) - 20*u ;  u' = u + (alpha/6) * lap

    written as a class-L left fold of `coef*tap` terms (reference idiom `smoke_apply.mlir:6-13`):
    edge taps carry the constant 4, corner taps are unit, the centre carries -20; the fold is
    then scaled (class S tail) and added to the centre (class A).
    """
    assert len(shape) == 2
    hdr, lb_, ub_ = _header(shape, ty, lb)
    ilb, iub = _interior(lb_, ub_)
    edges = [(-1, 0), (1, 0), (0, -1), (0, 1)]
    corners = [(-1, -1), (-1, 1), (1, -1), (1, 1)]
    b = [_apply_open(2, 1, ["%u0"], ilb, iub)]
    b.append("        %%c = neptune_ir.access %%a0[0, 0] : !temp -> %s" % ty)
    for j, o in enumerate(edges):
        b.append("        %%e%d = neptune_ir.access %%a0%s : !temp -> %s" % (j, _off(o), ty))
    for j, o in enumerate(corners):
        b.append("        %%d%d = neptune_ir.access %%a0%s : !temp -> %s" % (j, _off(o), ty))
    b.append("        %%four = arith.constant 4.0 : %s" % ty)
    b.append("        %%m20 = arith.constant -20.0 : %s" % ty)
    b.append("        %%s = arith.constant %s : %s" % (_f(alpha / 6.0, ty), ty))
    b.append("        %%t0 = arith.mulf %%m20, %%c : %s" % ty)
    prev = "%t0"
    n = 1
    for j in range(4):
        b.append("        %%p%d = arith.mulf %%four, %%e%d : %s" % (j, j, ty))
        b.append("        %%t%d = arith.addf %s, %%p%d : %s" % (n, prev, j, ty))
        prev = "%%t%d" % n
        n += 1
    for j in range(4):
        b.append("        %%t%d = arith.addf %s, %%d%d : %s" % (n, prev, j, ty))
        prev = "%%t%d" % n
        n += 1
    b.append("        %%lap = arith.mulf %%s, %s : %s" % (prev, ty))
    b.append("        %%o = arith.addf %%c, %%lap : %s" % ty)
    b.append("        neptune_ir.yield %%o : %s" % ty)
    b.append("      }")
    return hdr + "module {\n" + _entry(2, ty, 0, "\n".join(b), ["%u0"]) + "\n}\n"


def weighted_const(shape: Sequence[int], offsets: Sequence[Sequence[int]],
                   coefs: Sequence[float], ty: str = "f64", lb=None) -> str:
    """General constant-coefficient stencil as a class-L left fold:
    `t = c0*u[o0]; t = t + c1*u[o1]; ...` — a coefficient of exactly 1.0 is left un-multiplied,
    as in `smoke_apply.mlir:11-13`."""
    rank = len(shape)
    hdr, lb_, ub_ = _header(shape, ty, lb)
    rad = max(max(abs(int(x)) for x in o) for o in offsets)
    ilb, iub = _interior(lb_, ub_, rad)
    b = [_apply_open(rank, 1, ["%u0"], ilb, iub)]
    prev = None
    for j, (o, c) in enumerate(zip(offsets, coefs)):
        b.append("        %%v%d = neptune_ir.access %%a0%s : !temp -> %s" % (j, _off(o), ty))
        if float(c) == 1.0:
            term = "%%v%d" % j
        else:
            b.append("        %%k%d = arith.constant %s : %s" % (j, _f(c, ty), ty))
            b.append("        %%p%d = arith.mulf %%k%d, %%v%d : %s" % (j, j, j, ty))
            term = "%%p%d" % j
        if prev is None:
            prev = term
        else:
            b.append("        %%t%d = arith.addf %s, %s : %s" % (j, prev, term, ty))
            prev = "%%t%d" % j
    b.append("        neptune_ir.yield %s : %s" % (prev, ty))
    b.append("      }")
    return hdr + "module {\n" + _entry(rank, ty, 0, "\n".join(b), ["%u0"]) + "\n}\n"


def varcoef(shape: Sequence[int], offsets: Sequence[Sequence[int]], ty: str = "f64", lb=None) -> str:
    """Variable-coefficient stencil, one coefficient field per tap (BASELINE C4 with the 27
    box offsets): `t = c0[p]*u[p+o0]; t = t + c1[p]*u[p+o1]; ...` — every product is a `mulf`
    of two `access` results (not "linear" under `VerifyAndAnnotate.cpp:167-204`, hence it lives
    directly in `func.func`)."""
    rank = len(shape)
    K = len(offsets)
    hdr, lb_, ub_ = _header(shape, ty, lb)
    ilb, iub = _interior(lb_, ub_)
    ins = ["%u0"] + ["%%tc%d" % k for k in range(K)]
    b = [_apply_open(rank, 1 + K, ins, ilb, iub)]
    zero = _off([0] * rank)
    prev = None
    for j, o in enumerate(offsets):
        b.append("        %%v%d = neptune_ir.access %%a0%s : !temp -> %s" % (j, _off(o), ty))
        b.append("        %%k%d = neptune_ir.access %%a%d%s : !temp -> %s" % (j, j + 1, zero, ty))
        b.append("        %%p%d = arith.mulf %%k%d, %%v%d : %s" % (j, j, j, ty))
        if prev is None:
            prev = "%%p%d" % j
        else:
            b.append("        %%t%d = arith.addf %s, %%p%d : %s" % (j, prev, j, ty))
            prev = "%%t%d" % j
    b.append("        neptune_ir.yield %s : %s" % (prev, ty))
    b.append("      }")
    return hdr + "module {\n" + _entry(rank, ty, K, "\n".join(b), ins) + "\n}\n"


def jacobi_star(shape: Sequence[int], h2: float, ty: str = "f64", lb=None) -> str:
    """Poisson Jacobi sweep with a right-hand side (BASELINE C5):

        u' = (((((n0 + n1) + ... + n_{2r-1}) - h2*f) * (1/(2r))

    two inputs (u, f) as in the reference's two-input apply `smoke_time_advance.mlir:38-49`."""
    rank = len(shape)
    hdr, lb_, ub_ = _header(shape, ty, lb)
    ilb, iub = _interior(lb_, ub_)
    offs = star_offsets(rank)
    b = [_apply_open(rank, 2, ["%u0", "%tc0"], ilb, iub)]
    for j, o in enumerate(offs):
        b.append("        %%n%d = neptune_ir.access %%a0%s : !temp -> %s" % (j, _off(o), ty))
    b.append("        %%f = neptune_ir.access %%a1%s : !temp -> %s" % (_off([0] * rank), ty))
    b.append("        %%h2 = arith.constant %s : %s" % (_f(h2, ty), ty))
    b.append("        %%inv = arith.constant %s : %s" % (_f(1.0 / (2.0 * rank), ty), ty))
    b.append("        %%t1 = arith.addf %%n0, %%n1 : %s" % ty)
    for j in range(2, len(offs)):
        b.append("        %%t%d = arith.addf %%t%d, %%n%d : %s" % (j, j - 1, j, ty))
    b.append("        %%hf = arith.mulf %%h2, %%f : %s" % ty)
    b.append("        %%s = arith.subf %%t%d, %%hf : %s" % (len(offs) - 1, ty))
    b.append("        %%o = arith.mulf %%inv, %%s : %s" % ty)
    b.append("        neptune_ir.yield %%o : %s" % ty)
    b.append("      }")
    return hdr + "module {\n" + _entry(rank, ty, 1, "\n".join(b), ["%u0", "%tc0"]) + "\n}\n"


def laplacian_s_1d(n: int = 16, dxinv2: float = 100.0, ty: str = "f64") -> str:
    """The reference's own 1-D 3-point Laplacian body, verbatim in structure
    (`test/smoke_tests/smoke_time_advance.mlir:13-29`), inside a plain entry function."""
    hdr, lb_, ub_ = _header([n], ty)
    b = [_apply_open(1, 1, ["%u0"], [1], [n - 1])]
    b += ["        %%um1 = neptune_ir.access %%a0[-1] : !temp -> %s" % ty,
          "        %%uc  = neptune_ir.access %%a0[0]  : !temp -> %s" % ty,
          "        %%up1 = neptune_ir.access %%a0[1]  : !temp -> %s" % ty,
          "        %%two    = arith.constant 2.0 : %s" % ty,
          "        %%dxinv2 = arith.constant %s : %s" % (_f(dxinv2, ty), ty),
          "        %%t0 = arith.mulf %%two, %%uc : %s" % ty,
          "        %%t1 = arith.subf %%um1, %%t0 : %s" % ty,
          "        %%t2 = arith.addf %%t1, %%up1 : %s" % ty,
          "        %%lap_i = arith.mulf %%dxinv2, %%t2 : %s" % ty,
          "        neptune_ir.yield %%lap_i : %s" % ty,
          "      }"]
    return hdr + "module {\n" + _entry(1, ty, 0, "\n".join(b), ["%u0"]) + "\n}\n"


def field_sum(shape: Sequence[int], ty: str = "f64", lb=None, sub_lb=None, sub_ub=None) -> str:
    """`neptune_ir.reduce` kind "sum" of a whole field or of the sub-box [sub_lb, sub_ub)
    (`include/Dialect/NeptuneIR/NeptuneIROps.td:264-299`: `reduce %t (in #bounds)? {kind = "sum"}`)."""
    rank = len(shape)
    hdr, _, _ = _header(shape, ty, lb)
    mr = _memref(rank, ty)
    dom = "" if sub_lb is None else " in " + _bounds(sub_lb, sub_ub)
    lines = ["  func.func @entry(%%in: %s) -> %s {" % (mr, ty),
             "    %%fin = neptune_ir.wrap %%in : %s -> !field" % mr,
             "    %u0  = neptune_ir.load %fin : !field -> !temp",
             '    %%s   = neptune_ir.reduce %%u0%s {kind = "sum"} : !temp -> %s' % (dom, ty),
             "    func.return %%s : %s" % ty,
             "  }"]
    return hdr + "module {\n" + "\n".join(lines) + "\n}\n"


def heat_star_then_sum(shape: Sequence[int], alpha: float, ty: str = "f64") -> str:
    """One heat step stored to %out, then the sum of the new state over the interior: the
    "sweep, then a norm-like scalar" shape of a convergence check."""
    text = heat_star(shape, alpha, ty)
    rank = len(shape)
    mr = _memref(rank, ty)
    ilb, iub = [1] * rank, [int(n) - 1 for n in shape]
    text = text.replace("(%%out: %s, %%in: %s) -> %s {" % (mr, mr, mr), "(%%out: %s, %%in: %s) -> %s {" % (mr, mr, ty))
    tail_old = "    %%res = neptune_ir.unwrap %%fout : !field -> %s\n    func.return %%res : %s" % (mr, mr)
    tail_new = ('    %%s = neptune_ir.reduce %%r in %s {kind = "sum"} : !temp -> %s\n'
                "    func.return %%s : %s") % (_bounds(ilb, iub), ty, ty)
    assert tail_old in text
    return text.replace(tail_old, tail_new)


def euler_opdef(shape: Sequence[int], dxinv2: float, dt: float, ty: str = "f64") -> str:
    """Explicit Euler step written the way the reference composes it (`test/smoke_tests/
    smoke_time_advance.mlir:9-51`, `lib/Passes/HighLevelConvertion.cpp:77-120`): an operator definition
    `@L` (class-S Laplacian, the reference's own body generalised to `rank` dimensions), a call
    `apply_linear @L(%u)`, and an AXPY apply over (u, L(u)):

        l  = dxinv2 * ((((n0 - k*u) + n1) + n2) + ...)        // @L
        u' = u + dt * l                                        // apply(%u0, %l), both read at offset 0

    Unfused this is two sweeps and a temporary; the `neptuneir-to-b200` pipeline folds it into one
    (frontend/fusion.cpp)."""
    rank = len(shape)
    hdr, lb_, ub_ = _header(shape, ty)
    ilb, iub = _interior(lb_, ub_)
    offs = star_offsets(rank)
    mr = _memref(rank, ty)
    idx = ", ".join("%%i%d: index" % d for d in range(rank))
    zero = _off([0] * rank)
    L = ["  neptune_ir.linear_opdef @L : (!temp) -> !temp {",
         "  ^bb0(%x: !temp):",
         "    %%lap = neptune_ir.apply(%%x) attributes {bounds = %s} : (!temp) -> !temp {" % _bounds(ilb, iub),
         "      ^bb0(%s, %%a0: !temp):" % idx,
         "        %%c = neptune_ir.access %%a0%s : !temp -> %s" % (zero, ty)]
    for j, o in enumerate(offs):
        L.append("        %%n%d = neptune_ir.access %%a0%s : !temp -> %s" % (j, _off(o), ty))
    L.append("        %%k = arith.constant %s : %s" % (_f(2.0 * rank, ty), ty))
    L.append("        %%s = arith.constant %s : %s" % (_f(dxinv2, ty), ty))
    L.append("        %%t0 = arith.mulf %%k, %%c : %s" % ty)
    L.append("        %%t1 = arith.subf %%n0, %%t0 : %s" % ty)
    for j in range(1, len(offs)):
        L.append("        %%t%d = arith.addf %%t%d, %%n%d : %s" % (j + 1, j, j, ty))
    L.append("        %%l = arith.mulf %%s, %%t%d : %s" % (len(offs), ty))
    L.append("        neptune_ir.yield %%l : %s" % ty)
    L.append("    }")
    L.append("    neptune_ir.return %lap : !temp")
    L.append("  }")
    E = ["  func.func @entry(%%out: %s, %%in: %s) -> %s {" % (mr, mr, mr),
         "    %%fout = neptune_ir.wrap %%out : %s -> !field" % mr,
         "    %%fin  = neptune_ir.wrap %%in  : %s -> !field" % mr,
         "    %u0   = neptune_ir.load %fin : !field -> !temp",
         "    %lu   = neptune_ir.apply_linear @L(%u0) : (!temp) -> !temp",
         "    %%r = neptune_ir.apply(%%u0, %%lu) attributes {bounds = %s} : (!temp, !temp) -> !temp {" % _bounds(ilb, iub),
         "      ^bb0(%s, %%a0: !temp, %%a1: !temp):" % idx,
         "        %%x0 = neptune_ir.access %%a0%s : !temp -> %s" % (zero, ty),
         "        %%l0 = neptune_ir.access %%a1%s : !temp -> %s" % (zero, ty),
         "        %%dt = arith.constant %s : %s" % (_f(dt, ty), ty),
         "        %%t = arith.mulf %%dt, %%l0 : %s" % ty,
         "        %%o = arith.addf %%x0, %%t : %s" % ty,
         "        neptune_ir.yield %%o : %s" % ty,
         "    }",
         "    neptune_ir.store %r to %fout : !temp to !field",
         "    %%res = neptune_ir.unwrap %%fout : !field -> %s" % mr,
         "    func.return %%res : %s" % mr,
         "  }"]
    return hdr + "module {\n" + "\n".join(L) + "\n" + "\n".join(E) + "\n}\n"
