"""CPU tests of the host logic: the C++ NeptuneIR front end, the body recogniser, the pipeline
plan and the slab index algebra — all through the C ABI, no GPU needed."""
import numpy as np
import pytest

import helpers as H
from oracle import mlir_subset as ms

CASES = [(n, r, ty) for (n, r, tys) in H.CATALOGUE for ty in tys]
SMALL = {1: (64,), 2: (20, 28), 3: (9, 12, 14)}
OPNAME = {0: "access", 1: "constant", 2: "addf", 3: "subf", 4: "mulf", 5: "divf", 6: "negf"}


@pytest.mark.parametrize("name,rank,ty", CASES, ids=["%s-%dd-%s" % c for c in CASES])
def test_cxx_front_end_matches_oracle_parser(nb, name, rank, ty):
    """Two independent parsers (C++ product, Python oracle) must read the same apply."""
    text, n_extra = H.program(name, list(SMALL[rank]), ty, lb=None if name == "laplacian_s_1d" else [3] * rank)
    m = nb.Module(text)
    assert m.apply_count() == 1
    d = m.apply_desc(0)
    ap = [o for o in ms.parse(text).functions["@entry"].ops if o.name == "apply"][0].apply
    assert d.rank == rank and d.n_inputs == 1 + n_extra and d.dtype == (nb.F64 if ty == "f64" else nb.F32)
    assert tuple(d.lb) == ap.bounds.lb and tuple(d.ub) == ap.bounds.ub
    assert tuple(d.out_lb) == ap.out_type.bounds.lb and tuple(d.out_ub) == ap.out_type.bounds.ub
    assert len(d.ops) == len(ap.body)
    names = {o.result: i for i, o in enumerate(ap.body)}
    args = {n: k for k, n in enumerate(ap.input_args)}
    for o, p in zip(d.ops, ap.body):
        assert OPNAME[o["opcode"]] == p.name
        if p.name == "access":
            assert o["input"] == args[p.operands[0]] and tuple(o["offset"]) == p.offsets
        elif p.name == "constant":
            assert o["value"] == float(H.NPT[ty](p.value))
        else:
            assert [o["a"], o["b"]][:len(p.operands)] == [names[x] for x in p.operands]
    assert d.yield_value == names[ap.yield_value]
    assert nb.body_compiled(d.key()), "catalogue body has no compiled kernel"
    m.destroy()


def test_body_key_ignores_constants_and_extents(nb):
    a = nb.Module(H.mb.heat_star([8, 8, 8], 0.1)).apply_desc(0).key()
    b = nb.Module(H.mb.heat_star([64, 32, 16], 0.05, lb=[1, 2, 3])).apply_desc(0).key()
    c = nb.Module(H.mb.heat_star([8, 8, 8], 0.1, "f32")).apply_desc(0).key()
    d = nb.Module(H.mb.jacobi_star([8, 8, 8], 0.1)).apply_desc(0).key()
    assert a == b and a != c and a != d


def test_plan_fuses_store_and_skips_uploads(nb):
    m = nb.Module(H.mb.heat_star([8, 8, 8], 0.1))
    plan = m.plan("@entry")
    assert "store fused" in plan and "elided" in plan
    assert "arg 0 %out: no upload, download" in plan
    assert "arg 1 %in: upload, no download" in plan
    assert m.arg_count("@entry") == 2 and m.arg_numel("@entry", 0) == 512
    src = m.cuda_source(0)
    assert "A::mul(" in src and "tap<0, -1, 0, 0>" in src
    m.destroy()


BAD = {
    "index-dependent body": ("""%c = neptune_ir.access %a0[0] : !temp -> f64
        %x = arith.index_cast %i0 : index to i64""", "EUNSUPPORTED"),
    "control flow": ("""%c = neptune_ir.access %a0[0] : !temp -> f64
        %x = scf.if %c -> (f64) { scf.yield %c : f64 } else { scf.yield %c : f64 }""", "EUNSUPPORTED"),
    "undefined value": ("""%c = arith.addf %nope, %nope : f64""", "EINVAL"),
}


@pytest.mark.parametrize("what", sorted(BAD))
def test_rejected_bodies(nb, what):
    body, code = BAD[what]
    text = """
#b = #neptune_ir.bounds<lb = [0], ub = [16]>
!temp = !neptune_ir.temp<element = f64, bounds = #b, location = #neptune_ir.location<"cell">>
func.func @entry(%%u: !temp) -> !temp {
  %%r = neptune_ir.apply(%%u) attributes {bounds = #neptune_ir.bounds<lb = [1], ub = [15]>} : (!temp) -> !temp {
    ^bb0(%%i0: index, %%a0: !temp):
        %s
        neptune_ir.yield %%c : f64
  }
  func.return %%r : !temp
}""" % body
    with pytest.raises(nb.NeptuneError) as e:
        nb.Module(text)
    assert code in str(e.value)


def test_verifier_rules(nb):
    """ApplyOp::verify (lib/Dialect/NeptuneIR/NeptuneIRVerifier.cpp:141-171): block args = rank
    index args + one arg per input — the reference's stale one-index-arg form is rejected."""
    good = H.mb.heat_star([8, 8], 0.2)
    stale = good.replace("^bb0(%i0: index, %i1: index, %a0: !temp):", "^bb0(%i0: index, %a0: !temp):")
    with pytest.raises(nb.NeptuneError) as e:
        nb.Module(stale)
    assert "block arg count" in str(e.value)
    # taps that leave the input box over the apply bounds (UB in the reference) are diagnosed
    oob = good.replace("lb = [1, 1], ub = [7, 7]", "lb = [0, 1], ub = [7, 7]")
    with pytest.raises(nb.NeptuneError) as e:
        nb.Module(oob)
    assert "outside the input's bounds" in str(e.value)
    # the reference's CPU pipeline name is not served by this library
    with pytest.raises(nb.NeptuneError):
        nb.Module(good, "neptuneir-to-llvm")
    nb.Module(good, "neptuneir-to-b200{fma=true}").destroy()


def test_stencil_shape_attribute_is_cross_checked(nb):
    ok = H.mb.heat_star([8, 8, 8], 0.1, with_shape_attr=True)
    nb.Module(ok).destroy()
    bad = ok.replace("array<i64: 0, 0, 1>", "array<i64: 0, 0, 2>")
    with pytest.raises(nb.NeptuneError) as e:
        nb.Module(bad)
    assert "stencil_shape" in str(e.value)


def test_slab_partition_and_localize(nb):
    from neptune_pde_solver_b200.slab import Slab
    for n0, nr in [(16, 1), (16, 2), (17, 4), (512, 8), (5, 5)]:
        rows = [nb.slab_partition(n0, nr, r) for r in range(nr)]
        assert rows[0][0] == 0 and rows[-1][1] == n0
        assert all(rows[i][1] == rows[i + 1][0] for i in range(nr - 1))
        assert max(e - b for b, e in rows) - min(e - b for b, e in rows) <= 1
    g = nb.Module(H.mb.heat_star([16, 8, 8], 0.1, lb=[10, 0, 0])).apply_desc(0)
    covered = []
    for r in range(4):
        s = Slab(g, 4, r)
        l = s.local
        assert l.out_lb[0] == max(10, 10 + s.begin - 1) and l.out_ub[0] == min(26, 10 + s.end + 1)
        assert l.in_lb[0] == l.out_lb and l.in_ub[0] == l.out_ub
        assert l.lb[1:] == g.lb[1:] and l.ub[1:] == g.ub[1:]
        covered += list(range(l.lb[0], l.ub[0]))
        # the localized apply is the builder's own interior of the local box (halo == radius)
        again = nb.Module(H.mb.heat_star(list(s.local_shape((16, 8, 8))), 0.1, lb=[l.out_lb[0], 0, 0])).apply_desc(0)
        assert (again.lb, again.ub, again.out_lb, again.out_ub) == (l.lb, l.ub, l.out_lb, l.out_ub)
    assert covered == list(range(g.lb[0], g.ub[0]))      # every interior plane swept exactly once
    with pytest.raises(nb.NeptuneError):
        nb.slab_localize(g, 4, 0, 0)                       # halo narrower than the radius


def test_no_gpu_means_loud_failure(nb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(nb.NeptuneError) as e:
        nb.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_reduce_is_planned_and_other_kinds_rejected(nb):
    """`neptune_ir.reduce` (NeptuneIROps.td:264-299): kind "sum" becomes neptune_rt_field_reduce_sum;
    the kinds the reference does not lower (DataflowLowering.cpp:655-657) are diagnosed, not guessed."""
    m = nb.Module(H.mb.field_sum([8, 8, 8], "f32", lb=[1, 2, 3], sub_lb=[2, 2, 4], sub_ub=[5, 9, 10]))
    plan = m.plan("@entry")
    assert "neptune_rt_field_reduce_sum([2,2,4]..[5,9,10])" in plan and "arg 0 %in: upload, no download" in plan
    assert m.apply_count() == 0
    m.destroy()
    m = nb.Module(H.mb.heat_star_then_sum([8, 8, 8], 0.1))
    plan = m.plan("@entry")
    assert "neptune_rt_apply" in plan and "neptune_rt_field_reduce_sum([1,1,1]..[7,7,7])" in plan
    m.destroy()
    with pytest.raises(nb.NeptuneError) as e:
        nb.Module(H.mb.field_sum([8, 8], "f64").replace('"sum"', '"l2"'))
    assert "EUNSUPPORTED" in str(e.value) and "l2" in str(e.value)


def test_operator_call_plus_axpy_is_fused_into_one_sweep(nb):
    """`%l = apply_linear @L(%u); %r = apply(%u, %l){u + dt*l}` (the reference's explicit-step
    composition, test/smoke_tests/smoke_time_advance.mlir:33-51) becomes ONE apply whose body is @L's op
    list followed by the AXPY's; `fuse=false` keeps the two sweeps."""
    for shape in ([16], [12, 16], [8, 8, 8]):
        text = H.mb.euler_opdef(shape, 100.0, 1e-3)
        m = nb.Module(text)
        plan = m.plan("@entry")
        assert "fused into one sweep" in plan and "call @L" not in plan
        fused = [i for i in range(m.apply_count()) if m.apply_function(i) == "@entry"]
        assert len(fused) == 1
        d = m.apply_desc(fused[0])
        assert d.n_inputs == 1 and nb.body_compiled(d.key())
        src = m.cuda_source(fused[0])
        assert src.count("A::mul(") == 3 and "tap<1," not in src     # k*c, s*(...), dt*l — and no second input
        m.destroy()
        m = nb.Module(text, "neptuneir-to-b200{fuse=false}")
        plan = m.plan("@entry")
        assert "call @L" in plan and "fused into one sweep" not in plan
        m.destroy()
    with pytest.raises(nb.NeptuneError):
        nb.Module(H.mb.euler_opdef([8, 8], 1.0, 0.1), "neptuneir-to-b200{fuse=maybe}")


def test_fusion_is_refused_when_it_would_change_the_result(nb):
    """The temp must be read at offset 0 only, must not be input 0 (the copy-through source), and the
    consumer's bounds must lie inside the producer's."""
    text = H.mb.euler_opdef([12, 16], 100.0, 1e-3)
    # consumer reads the operator's result at a neighbour offset
    t1 = text.replace("%l0 = neptune_ir.access %a1[0, 0]", "%l0 = neptune_ir.access %a1[0, 1]")
    # consumer bounds wider than the operator's: there %lu holds copy-through values
    t2 = text.replace("neptune_ir.apply(%x) attributes {bounds = #neptune_ir.bounds<lb = [1, 1], ub = [11, 15]>}",
                      "neptune_ir.apply(%x) attributes {bounds = #neptune_ir.bounds<lb = [2, 2], ub = [10, 14]>}")
    assert t1 != text and t2 != text
    for t in (t1, t2):
        try:
            m = nb.Module(t)
        except nb.NeptuneError:
            continue            # unfused AXPY-at-an-offset body is not in the catalogue: loud, not wrong
        assert "fused into one sweep" not in m.plan("@entry")
        m.destroy()


def test_linear_opdef_rule_products_need_a_constant_factor(nb):
    """checkApplyLike(requireLinearBody) of the reference (VerifyAndAnnotate.cpp:183-204): inside a linear_opdef a mulf
    must have a constant operand and a divf a constant divisor.  The module is refused as a whole with the reference's
    wording (a failed VerifyAndAnnotate fails the pass pipeline there too: only well-formed IR that is merely off the hot
    path is set aside function by function); the same body in a func.func (variable coefficients) is legal."""
    import json
    import os
    pack = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_inputs.json")))
    text = pack["files"]["smoke_time_advance.mlir"]["text"]
    assert "%t  = arith.mulf %alpha, %lap0 : f64" in text
    bad = text.replace("%t  = arith.mulf %alpha, %lap0 : f64", "%t  = arith.mulf %x0, %lap0 : f64")
    with pytest.raises(nb.NeptuneError) as err:
        nb.Module(bad)
    assert "MulFOp in linear region must multiply by a constant" in str(err.value) and err.value.code == -1   # EINVAL
    m = nb.Module(text)                                          # the untouched fixture: operators lower, @entry is set aside
    assert m.diagnostic("@ac_lap") is None and m.diagnostic("@ac_A") is None and "time_advance" in m.diagnostic("@entry")
    m.destroy()
    m = nb.Module(H.program("varcoef_3d7", [8, 8, 8])[0])       # mulf(access, access) in a func.func: fine
    assert m.diagnostic("@entry") is None
    m.destroy()
