"""CPU tests of the oracle itself (parity unpinned: the reference holds no golden vectors for the
apply path, SURVEY.md §4/§8c — so the restatement is pinned three ways instead):
  1. two independent restatements agree bit for bit: numpy whole-array slices vs gcc-compiled
     scalar loop nest (`-O2 -fno-tree-vectorize -ffp-contract=off`);
  2. analytic known answers;
  3. committed golden vectors (tests/golden/, made by tests/golden/make_golden.py) guard against
     drift of the oracle between rounds.
"""
import os

import numpy as np
import pytest

import helpers as H
from oracle import c_lowering as oc
from oracle import mlir_subset as ms
from oracle import numpy_lowering as onp

SMALL = {1: (64,), 2: (20, 28), 3: (9, 12, 14)}
CASES = [(n, r, ty) for (n, r, tys) in H.CATALOGUE for ty in tys]


@pytest.mark.parametrize("name,rank,ty", CASES, ids=["%s-%dd-%s" % c for c in CASES])
def test_numpy_and_c_restatements_agree(name, rank, ty):
    shape = SMALL[rank]
    text, n_extra = H.program(name, list(shape), ty)
    u, extra = H.inputs(shape, n_extra, ty, seed=42)
    want = H.oracle_entry(text, u, extra)
    cm = oc.CompiledModule(text, tag="t_%s_%dd_%s" % (name, rank, ty))
    out = np.full_like(u, np.nan)
    cm.call("@entry", [out, u] + extra)
    assert H.bits_equal(out, want)
    # and over a short time loop
    assert H.bits_equal(cm.run_steps(u, extra, 5), onp.run_steps(ms.parse(text), u, extra, 5))


def test_reference_1d_laplacian_by_hand():
    """The reference's own body (test/smoke_tests/smoke_time_advance.mlir:13-29) at n=16:
    lap_i = dxinv2 * ((u[i-1] - 2*u[i]) + u[i+1]), ends copied through."""
    text = H.mb.laplacian_s_1d(16, 100.0)
    u = np.arange(16, dtype=np.float64) ** 2
    out = H.oracle_entry(text, u, [])
    want = u.copy()
    for i in range(1, 15):
        want[i] = 100.0 * ((u[i - 1] - 2.0 * u[i]) + u[i + 1])
    assert np.array_equal(out, want)
    assert np.all(out[1:15] == 200.0) and out[0] == 0.0 and out[15] == 225.0


def test_known_answers():
    shape = (10, 12, 14)
    text, _ = H.program("heat_star", list(shape))
    u = np.full(shape, 1.0)
    assert np.array_equal(H.oracle_entry(text, u, []), u)           # constant = fixed point
    lap = H.mb.weighted_const(list(shape), H.STAR7, [-6.0, 1, 1, 1, 1, 1, 1])
    z, y, x = np.meshgrid(*[np.arange(n, dtype=np.float64) for n in shape], indexing="ij")
    out = H.oracle_entry(lap, z * z + y * y + x * x, [])
    assert np.all(out[1:-1, 1:-1, 1:-1] == 6.0)                     # Laplacian of a quadratic
    # Jacobi fixed point: f = 0, u = const
    jt, _ = H.program("jacobi_star", [8, 8, 8])
    u = np.full((8, 8, 8), 3.0)
    assert np.allclose(H.oracle_entry(jt, u, [np.zeros_like(u)]), u, rtol=0, atol=1e-15)


def test_out_of_box_tap_is_an_error():
    text, _ = H.program("heat_star", [8, 8])
    bad = text.replace("lb = [1, 1], ub = [7, 7]", "lb = [0, 1], ub = [7, 7]")
    with pytest.raises(IndexError):
        H.oracle_entry(bad, np.zeros((8, 8)), [])


def test_golden_vectors():
    gdir = os.path.join(os.path.dirname(__file__), "golden")
    files = sorted(f for f in os.listdir(gdir) if f.endswith(".npz"))
    assert files, "no golden vectors committed"
    for f in files:
        g = np.load(os.path.join(gdir, f))
        text = str(g["mlir"])
        extra = [g["extra%d" % k] for k in range(int(g["n_extra"]))]
        got = onp.run_steps(ms.parse(text), g["u0"], extra, int(g["steps"]))
        assert H.bits_equal(got, g["result"]), f


@pytest.mark.parametrize("ty", ["f64", "f32"])
def test_reduce_sum_is_a_strict_left_fold(ty):
    """`neptune_ir.reduce` kind "sum" (lib/Passes/DataflowLowering.cpp:589-698): acc = 0, then
    acc = acc + x in row-major order, in the element type.  numpy (cumsum) and gcc restatements
    agree bit for bit, on the whole box and on a sub-box of a non-zero-based field."""
    dt = np.float64 if ty == "f64" else np.float32
    u = np.random.default_rng(5).uniform(-1, 1, size=(9, 12, 14)).astype(dt)
    for kw in ({}, {"lb": [2, 3, 4], "sub_lb": [3, 3, 5], "sub_ub": [9, 14, 17]}):
        text = H.mb.field_sum(list(u.shape), ty, **kw)
        got_np = onp.run_function(ms.parse(text), "@entry", [u])[0]
        got_c = oc.CompiledModule(text, tag="t_reduce_%s" % ty).call_scalar("@entry", [u])
        assert dt(got_c) == got_np and got_np.dtype == dt
        # by hand: python floats are f64, so only check the f64 case against a plain loop
        if ty == "f64":
            sub = u if not kw else u[1:7, 0:11, 1:13]
            acc = 0.0
            for v in sub.reshape(-1):
                acc = acc + float(v)
            assert acc == float(got_np)
    ones = np.ones((9, 12, 14), dtype=dt)
    assert onp.run_function(ms.parse(H.mb.field_sum([9, 12, 14], ty)), "@entry", [ones])[0] == 9 * 12 * 14
    with pytest.raises(NotImplementedError):
        onp.run_function(ms.parse(H.mb.field_sum([4, 4], ty).replace('"sum"', '"max"')), "@entry", [ones[0, :4, :4].copy()])


def test_sweep_then_reduce_program():
    text = H.mb.heat_star_then_sum([9, 12, 14], 0.1)
    u, _ = H.inputs((9, 12, 14), 0, "f64", seed=8)
    o1, o2 = np.zeros_like(u), np.zeros_like(u)
    s1 = onp.run_function(ms.parse(text), "@entry", [o1, u])[0]
    s2 = oc.CompiledModule(text, tag="t_sweep_reduce").call_scalar("@entry", [o2, u])
    assert H.bits_equal(o1, o2) and float(s1) == s2
    assert H.bits_equal(o1, H.oracle_entry(H.mb.heat_star([9, 12, 14], 0.1), u, []))


# ---- third restatement: the loop nest as LLVM IR, compiled by LLVM (oracle/llvm_lowering.py) ------------------
LLVM_CASES = [("heat_star", (12, 16, 20), "f64"), ("heat_star", (24, 32), "f64"), ("heat_star", (24, 32), "f32"),
              ("heat_star", (40,), "f64"), ("heat_box_2d9", (24, 32), "f64"), ("jacobi_star", (10, 12, 16), "f64"),
              ("jacobi_star", (10, 12, 16), "f32"), ("varcoef_3d27", (8, 10, 12), "f64"), ("varcoef_3d7", (8, 10, 12), "f64"),
              ("weighted_3d27", (8, 9, 10), "f64"), ("weighted_2d9", (16, 20), "f64"), ("euler_opdef", (10, 12, 16), "f64"),
              ("laplacian_s_1d", (16,), "f64")]


@pytest.mark.parametrize("name,shape,ty", LLVM_CASES, ids=["%s-%s-%s" % (n, "x".join(map(str, s)), t) for n, s, t in LLVM_CASES])
def test_llvm_restatement_equals_numpy_restatement(name, shape, ty):
    """fmul/fadd without fast-math flags on the generic x86-64 CPU, JIT-compiled by LLVM 20 (the closest analogue of the
    reference's mlir-translate + llc flow, smoke_apply.sh:22-29) == the numpy slices, bit for bit, over several steps."""
    from oracle import llvm_lowering as oll
    text, n_extra = H.program(name, list(shape), ty)
    u, extra = H.inputs(shape, n_extra, ty, seed=17)
    a = H.oracle_steps(text, u, extra, 4)
    b = oll.run_steps(ms.parse(text), u, extra, 4)
    assert H.bits_equal(a, b)


def test_llvm_restatement_emits_no_fused_multiply_add():
    from oracle import llvm_lowering as oll
    text, _ = H.program("heat_star", [12, 16, 20])
    ap = [op.apply for op in ms.parse(text).functions["@entry"].ops if op.name == "apply"][0]
    la = oll.LlvmApply(ap)
    assert "fmul double" in la.ir and "fadd double" in la.ir and "fast" not in la.ir and "contract" not in la.ir
    asm = la.asm.lower()
    assert "vfmadd" not in asm and "fma" not in asm and "mulsd" in asm and "addsd" in asm


def test_llvm_restatement_on_the_reference_fixtures():
    import json
    import os
    from oracle import llvm_lowering as oll
    pack = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_inputs.json")))
    for c in pack["cases"]:
        if c["text"] != "verbatim":
            continue
        mod = ms.parse(pack["files"][c["file"]]["text"])
        x = np.array([float.fromhex(v) for v in c["input"]])
        for want in c["outputs"]:
            y = oll.run_function(mod, c["function"], [x])[0]
            assert H.bits_equal(np.asarray(y), np.array([float.fromhex(v) for v in want])), c["function"]
            x = np.array(y, copy=True)
