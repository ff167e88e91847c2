"""GPU parity tests: the sm_100a path, called through the C ABI (ctypes over libneptune_rt_b200.so),
against the CPU oracle (oracle/numpy_lowering.py, a restatement of
lib/Passes/DataflowLowering.cpp:258-448) on the same seeded inputs.

Tolerance: the default pipeline keeps the body's op order and never contracts, so every comparison
here is BIT-EXACT (fp64 and fp32).  Only `neptuneir-to-b200{fma=true}` is compared with a
tolerance: 1e-12 relative max-norm (fp64), as BASELINE.json's north star states.
"""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

SHAPES = {  # rank -> (TMA-eligible extents, ragged extents that force the direct kernel)
    1: [(1000,), (37,)],
    2: [(96, 160), (45, 51)],
    3: [(40, 48, 64), (13, 21, 19)],
}

CASES = [(n, r, ty) for (n, r, tys) in H.CATALOGUE for ty in tys]


def _run_entry(nb, ctx, text, u, extra, variant=None):
    m = nb.Module(text)
    try:
        if variant is not None:
            m.stencil(ctx, 0).set("variant", variant)
        out = np.full_like(u, np.nan)
        m.run(ctx, "@entry", [out, u] + list(extra))
        used = m.stencil(ctx, 0).variant()
        return out, used
    finally:
        m.destroy()


@pytest.mark.parametrize("name,rank,ty", CASES, ids=["%s-%dd-%s" % c for c in CASES])
def test_entry_bit_exact_every_variant(nb, ctx, name, rank, ty):
    """One call of @entry(out, in, coef...) == the oracle, for every compiled kernel variant."""
    for si, shape in enumerate(SHAPES[rank]):
        text, n_extra = H.program(name, list(shape), ty)
        u, extra = H.inputs(shape, n_extra, ty, seed=1000 + 7 * rank + si)
        want = H.oracle_entry(text, u, extra)
        m = nb.Module(text)
        variants = m.stencil(ctx, 0).variants()
        m.destroy()
        assert "direct" in variants
        out, used = _run_entry(nb, ctx, text, u, extra)
        assert H.bits_equal(out, want), "auto variant %s differs on %s" % (used, shape)
        if si == 0 and rank >= 2:
            fast = ("rows", "tma") if rank == 2 else ("tma",)   # rank 2 marches down the rows where the body has that kernel
            assert used.startswith(fast), "aligned %d-D extents should take the staged kernel, took %r" % (rank, used)
        for v in variants:
            if v.startswith(("tma", "rows")) and si == 1:
                continue  # ragged rows are not TMA-eligible (checked in test_errors)
            out, used = _run_entry(nb, ctx, text, u, extra, variant=v)
            assert used == v
            assert H.bits_equal(out, want), "variant %s differs on %s" % (v, shape)


@pytest.mark.parametrize("name,shape,lb", [
    ("heat_star", (24, 32, 40), (5, -3, 100)),
    ("heat_box_2d9", (64, 72), (-7, 9)),
    ("varcoef_3d7", (12, 16, 24), (1, 2, 3)),
    ("heat_star", (50,), (11,)),
])
def test_nonzero_lower_bounds(nb, ctx, name, shape, lb):
    """Logical coordinates that do not start at 0: iv + off - inLb (DataflowLowering.cpp:400-404)."""
    text, n_extra = H.program(name, list(shape), "f64", list(lb))
    u, extra = H.inputs(shape, n_extra, "f64", seed=77)
    out, _ = _run_entry(nb, ctx, text, u, extra)
    assert H.bits_equal(out, H.oracle_entry(text, u, extra))


def test_copy_through_shell(nb, ctx):
    """Points of the result outside `bounds` carry input 0 (memref.copy, DataflowLowering.cpp:282-287)."""
    text, _ = H.program("heat_star", [20, 24, 32])
    u, _ = H.inputs((20, 24, 32), 0, "f64", seed=5)
    out, _ = _run_entry(nb, ctx, text, u, [])
    shell = np.ones(u.shape, bool)
    shell[1:-1, 1:-1, 1:-1] = False
    assert H.bits_equal(out[shell], u[shell])
    assert not np.array_equal(out[~shell], u[~shell])


@pytest.mark.parametrize("name,shape,ty,steps", [
    ("heat_star", (32, 40, 64), "f64", 25),
    ("heat_star", (64, 128), "f64", 40),
    ("heat_box_2d9", (72, 96), "f64", 30),
    ("jacobi_star", (24, 32, 48), "f64", 30),
    ("jacobi_star", (24, 32, 48), "f32", 30),
    ("weighted_3d27", (20, 24, 32), "f64", 10),
    ("varcoef_3d27", (16, 16, 32), "f64", 6),
    ("heat_star", (300,), "f64", 100),
])
def test_time_loop_bit_exact(nb, ctx, name, shape, ty, steps):
    """N steps of the host time loop (call entry, swap) stay bit-identical to the oracle."""
    text, n_extra = H.program(name, list(shape), ty)
    u, extra = H.inputs(shape, n_extra, ty, seed=31)
    want = H.oracle_steps(text, u, extra, steps)
    m = nb.Module(text)
    try:
        a, b = u.copy(), np.zeros_like(u)
        res = m.run_steps(ctx, "@entry", [b, a] + extra, steps)
        got = [b, a][res]
        assert res == (0 if steps % 2 else 1)
        assert H.bits_equal(got, want)
        # the same loop through neptune_rt_run_steps on device fields
        st = m.stencil(ctx, 0)
        fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
        fx = [nb.Field.from_numpy(ctx, e) for e in extra]
        r = st.run_steps(fa, fb, fx, steps)
        assert H.bits_equal(r.download(), want)
        for f in [fa, fb] + fx:
            f.free()
    finally:
        m.destroy()


@pytest.mark.parametrize("name,shape,ty", [
    ("heat_star", (300, 1040), "f64"),        # three 512-column strips, the last one ragged; several stages per item
    ("heat_box_2d9", (70, 2056), "f64"),      # corner taps; more strips than rows per chunk
    ("heat_box_2d9", (1030, 520), "f32"),     # VX = 4
    ("jacobi_star", (257, 1024), "f64"),      # pointwise right-hand side beside the marched input
    ("weighted_2d9", (64, 512), "f64"),
    ("heat_star", (3, 512), "f64"),           # a single interior row
])
def test_row_marching_2d_bit_exact(nb, ctx, name, shape, ty):
    """sweep_rows (rank-2 row marching over bulk-copied row segments) == the oracle, every strip width,
    whole sweeps, row-range partitions (the slab path) and a 6-step time loop."""
    text, n_extra = H.program(name, list(shape), ty)
    u, extra = H.inputs(shape, n_extra, ty, seed=77)
    want = H.oracle_entry(text, u, extra)
    want6 = H.oracle_steps(text, u, extra, 6)
    dt = np.float64 if ty == "f64" else np.float32
    m = nb.Module(text)
    try:
        st = m.stencil(ctx, 0)
        rows = [v for v in st.variants() if v.startswith("rows")]
        assert len(rows) >= 3, st.variants()
        for v in rows:
            for zc in ("0", "16"):
                st.set("variant", v)
                st.set("zchunk", zc)
                fin = nb.Field.from_numpy(ctx, u)
                fx = [nb.Field.from_numpy(ctx, e) for e in extra]
                fout = nb.Field(ctx, dt, [0, 0], list(shape))
                fout.fill(float("nan"))
                st.apply([fin] + fx, fout)
                assert st.variant() == v
                assert H.bits_equal(fout.download(), want), (v, zc)
                fout.fill(float("nan"))
                n = shape[0]
                cuts = sorted({0, 1, min(7, n), min(8, n), max(n - 1, 0), n})
                for p0, p1 in zip(cuts[:-1], cuts[1:]):
                    st.apply([fin] + fx, fout, planes=(p0, p1))
                assert H.bits_equal(fout.download(), want), (v, zc, "partition")
                fb = nb.Field.from_numpy(ctx, u)
                r = st.run_steps(fin, fb, fx, 6)
                assert H.bits_equal(r.download(), want6), (v, zc, "steps")
                for f in [fin, fout, fb] + fx:
                    f.free()
    finally:
        m.destroy()


def test_apply_planes_partition(nb, ctx):
    """Sweeping dim-0 plane ranges one by one writes exactly what one whole sweep writes."""
    shape = (30, 32, 64)
    text, _ = H.program("heat_star", list(shape))
    u, _ = H.inputs(shape, 0, "f64", seed=9)
    want = H.oracle_entry(text, u, [])
    m = nb.Module(text)
    try:
        st = m.stencil(ctx, 0)
        for variant in st.variants():
            st.set("variant", variant)
            fin = nb.Field.from_numpy(ctx, u)
            fout = nb.Field(ctx, np.float64, [0, 0, 0], list(shape))
            fout.fill(float("nan"))
            for p0, p1 in [(0, 1), (1, 7), (7, 8), (8, 29), (29, 30)]:
                st.apply([fin], fout, planes=(p0, p1))
            assert H.bits_equal(fout.download(), want), variant
            fin.free(); fout.free()
    finally:
        m.destroy()


@pytest.mark.parametrize("name,shape", [("heat_star", (30, 32, 64)), ("heat_box_2d9", (40, 1024)), ("heat_star", (33, 512))])
def test_apply_two_plane_ranges_in_one_launch(nb, ctx, name, shape):
    """neptune_rt_apply_planes2: the two boundary ranges of a slab as one launch (or two, where the variant
    cannot) + the interior == one whole sweep, and the gap is left untouched by the first call."""
    text, _ = H.program(name, list(shape))
    u, _ = H.inputs(shape, 0, "f64", seed=10)
    want = H.oracle_entry(text, u, [])
    n = shape[0]
    m = nb.Module(text)
    try:
        st = m.stencil(ctx, 0)
        for variant in st.variants():
            if variant.startswith("tb2"):
                continue   # two-step kernels take the two ranges in the slab time loop (test_gpu_multi.py)
            for w in (1, 2, 5):
                st.set("variant", variant)
                fin = nb.Field.from_numpy(ctx, u)
                fout = nb.Field(ctx, np.float64, [0] * len(shape), list(shape))
                fout.fill(-7.0)
                st.apply([fin], fout, planes=(0, w, n - w, n))
                got = fout.download()
                assert H.bits_equal(got[:w], want[:w]) and H.bits_equal(got[n - w:], want[n - w:]), (variant, w)
                assert np.all(got[w:n - w] == -7.0), (variant, w, "gap written")
                st.apply([fin], fout, planes=(w, n - w))
                assert H.bits_equal(fout.download(), want), (variant, w)
                fin.free(); fout.free()
        with pytest.raises(Exception):
            fin = nb.Field.from_numpy(ctx, u)
            fout = nb.Field(ctx, np.float64, [0] * len(shape), list(shape))
            st.apply([fin], fout, planes=(0, 2, n - 3, n))   # ranges of different widths
    finally:
        m.destroy()


OPDEF_PROGRAM = """
#loc = #neptune_ir.location<"cell">
#b   = #neptune_ir.bounds<lb = [0, 0], ub = [48, 64]>
#bi  = #neptune_ir.bounds<lb = [1, 1], ub = [47, 63]>
!temp  = !neptune_ir.temp<element = f64, bounds = #b, location = #loc>
!field = !neptune_ir.field<element = f64, bounds = #b, location = #loc>

module {
  // operator definition + call, as in test/smoke_tests/smoke_time_advance.mlir:9-31
  neptune_ir.linear_opdef @A : (!temp) -> !temp {
  ^bb0(%x: !temp):
    %r = neptune_ir.apply(%x) attributes {bounds = #bi} : (!temp) -> !temp {
      ^bb0(%i: index, %j: index, %a0: !temp):
        %c = neptune_ir.access %a0[0, 0] : !temp -> f64
        %n0 = neptune_ir.access %a0[-1, 0] : !temp -> f64
        %n1 = neptune_ir.access %a0[1, 0] : !temp -> f64
        %n2 = neptune_ir.access %a0[0, -1] : !temp -> f64
        %n3 = neptune_ir.access %a0[0, 1] : !temp -> f64
        %k = arith.constant 4.0 : f64
        %alpha = arith.constant 0.125 : f64
        %t0 = arith.mulf %k, %c : f64
        %t1 = arith.subf %n0, %t0 : f64
        %t2 = arith.addf %t1, %n1 : f64
        %t3 = arith.addf %t2, %n2 : f64
        %t4 = arith.addf %t3, %n3 : f64
        %lap = arith.mulf %alpha, %t4 : f64
        %o = arith.addf %c, %lap : f64
        neptune_ir.yield %o : f64
    }
    neptune_ir.return %r : !temp
  }
  func.func @entry(%out: memref<?x?xf64>, %in: memref<?x?xf64>) -> memref<?x?xf64> {
    %fout = neptune_ir.wrap %out : memref<?x?xf64> -> !field
    %fin  = neptune_ir.wrap %in  : memref<?x?xf64> -> !field
    %u0   = neptune_ir.load %fin : !field -> !temp
    %u1 = neptune_ir.apply_linear @A(%u0) : (!temp) -> !temp
    %u2 = neptune_ir.apply_linear @A(%u1) : (!temp) -> !temp
    %u3 = neptune_ir.apply_linear @A(%u2) : (!temp) -> !temp
    neptune_ir.store %u3 to %fout {bounds = #neptune_ir.bounds<lb = [4, 8], ub = [40, 56]>} : !temp to !field
    %res = neptune_ir.unwrap %fout : !field -> memref<?x?xf64>
    func.return %res : memref<?x?xf64>
  }
}
"""


def test_opdef_calls_and_bounded_store(nb, ctx):
    """linear_opdef + apply_linear chains (StructureLowering.cpp:31-140) and `store` with bounds
    (DataflowLowering.cpp:181-218): the destination keeps its old contents outside the box."""
    from oracle import mlir_subset as ms, numpy_lowering as onp
    u, _ = H.inputs((48, 64), 0, "f64", seed=3)
    out0 = np.random.default_rng(4).uniform(-1, 1, size=u.shape)
    want = out0.copy()
    onp.run_function(ms.parse(OPDEF_PROGRAM), "@entry", [want, u.copy()])
    m = nb.Module(OPDEF_PROGRAM)
    try:
        assert m.apply_count() == 1 and m.apply_function(0) == "@A"
        got = out0.copy()
        m.run(ctx, "@entry", [got, u.copy()])
        assert H.bits_equal(got, want)
    finally:
        m.destroy()


def test_fma_pipeline_within_tolerance(nb, ctx):
    """`neptuneir-to-b200{fma=true}` may contract a*b+c: <= 1e-12 relative max-norm after N steps."""
    shape, steps = (32, 40, 64), 50
    text, _ = H.program("heat_star", list(shape))
    u, _ = H.inputs(shape, 0, "f64", seed=12)
    want = H.oracle_steps(text, u, [], steps)
    m = nb.Module(text, "neptuneir-to-b200{fma=true}")
    try:
        a, b = u.copy(), np.zeros_like(u)
        res = m.run_steps(ctx, "@entry", [b, a], steps)
        got = [b, a][res]
        rel = np.max(np.abs(got - want)) / np.max(np.abs(want))
        assert rel <= 1e-12, rel
    finally:
        m.destroy()


def test_known_answers_on_device(nb, ctx):
    """Analytic checks that need no oracle: a constant field is a fixed point of the heat step;
    the 7-point Laplacian of x^2+y^2+z^2 on an integer grid is exactly 6."""
    shape = (16, 24, 32)
    text, _ = H.program("heat_star", list(shape))
    u = np.full(shape, 1.0)
    out, _ = _run_entry(nb, ctx, text, u, [])
    assert np.array_equal(out, u)
    lap = H.mb.weighted_const(list(shape), H.STAR7, [-6.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0])
    z, y, x = np.meshgrid(*[np.arange(n, dtype=np.float64) for n in shape], indexing="ij")
    q = z * z + y * y + x * x
    out, _ = _run_entry(nb, ctx, lap, q, [])
    assert np.array_equal(out[1:-1, 1:-1, 1:-1], np.full((14, 22, 30), 6.0))


def test_errors_are_loud(nb, ctx):
    import os
    # a legal linear body that is NOT in the prebuilt catalogue (radius-2 taps): the reference lowers any body
    # (DataflowLowering.cpp:413-423); here it is compiled at run time with NVRTC and runs, bit-exact
    text = H.mb.weighted_const([16, 16], [(0, 0), (0, 2), (2, 0)], [0.5, 0.25, 0.25])
    m = nb.Module(text)
    u, _ = H.inputs((16, 16), 0, "f64", seed=3)
    out = np.full_like(u, np.nan)
    m.run(ctx, "@entry", [out, u])
    assert H.bits_equal(out, H.oracle_entry(text, u, []))
    m.destroy()
    # with run-time compilation switched off the catalogue is closed: EUNSUPPORTED, never a CPU fallback
    os.environ["NEPTUNE_RT_JIT"] = "0"
    try:
        text = H.mb.weighted_const([16, 16], [(0, 0), (0, -2), (-2, 0)], [0.5, 0.25, 0.25])
        m = nb.Module(text)     # recognised (it is a legal linear body) ...
        with pytest.raises(nb.NeptuneError) as e:
            m.run(ctx, "@entry", [np.zeros((16, 16)), np.zeros((16, 16))])
        assert e.value.code == nb.runtime.EUNSUPPORTED and "catalogue" in str(e.value) and "no CPU fallback" in str(e.value)
        m.destroy()
    finally:
        del os.environ["NEPTUNE_RT_JIT"]
    # result aliasing an input
    text, _ = H.program("heat_star", [8, 8, 8])
    m = nb.Module(text)
    st = m.stencil(ctx, 0)
    f = nb.Field(ctx, np.float64, [0, 0, 0], [8, 8, 8])
    with pytest.raises(nb.NeptuneError) as e:
        st.apply([f], f)
    assert e.value.code == nb.runtime.EINVAL
    # a field whose bounds are not the apply's operand type
    g = nb.Field(ctx, np.float64, [0, 0, 0], [8, 8, 9])
    with pytest.raises(nb.NeptuneError):
        st.apply([g], f)
    # ragged rows cannot take a TMA variant when it is forced
    m2 = nb.Module(H.program("heat_star", [9, 9, 9])[0])
    st2 = m2.stencil(ctx, 0)
    tma = [v for v in st2.variants() if v.startswith("tma")][0]
    st2.set("variant", tma)
    with pytest.raises(nb.NeptuneError):
        m2.run(ctx, "@entry", [np.zeros((9, 9, 9)), np.zeros((9, 9, 9))])
    with pytest.raises(nb.NeptuneError):
        st2.set("variant", "no-such-kernel")
    f.free(); g.free(); m.destroy(); m2.destroy()


def test_empty_bounds_is_pure_copy(nb, ctx):
    text, _ = H.program("heat_star", [8, 8, 8])
    text = text.replace("lb = [1, 1, 1], ub = [7, 7, 7]", "lb = [3, 3, 3], ub = [3, 7, 7]")
    u, _ = H.inputs((8, 8, 8), 0, "f64", seed=2)
    out, _ = _run_entry(nb, ctx, text, u, [])
    assert H.bits_equal(out, u)
    assert H.bits_equal(out, H.oracle_entry(text, u, []))


def test_launch_accounting(nb, ctx):
    text, _ = H.program("heat_star", [16, 16, 32])
    m = nb.Module(text)
    u, _ = H.inputs((16, 16, 32), 0, "f64", seed=2)
    n0 = ctx.launch_count()
    a, b = u.copy(), np.zeros_like(u)
    m.run_steps(ctx, "@entry", [b, a], 7)
    # the entry's time loop is neptune_rt_run_steps: three fused pairs + one single sweep, and no copy-through
    # or store kernels (one kernel per launch group)
    assert ctx.launch_count() - n0 == 4
    assert H.bits_equal([b, a][0], H.oracle_steps(text, u, [], 7))
    m.stencil(ctx, 0).set("tblock", 1)
    n0 = ctx.launch_count()
    a, b = u.copy(), np.zeros_like(u)
    m.run_steps(ctx, "@entry", [b, a], 7)
    assert ctx.launch_count() - n0 == 7      # one kernel per sweep
    assert H.bits_equal(b, H.oracle_steps(text, u, [], 7))
    m.destroy()


@pytest.mark.parametrize("name,shape,ty,steps", [
    ("heat_star", (40, 48, 64), "f64", 8),
    ("heat_star", (40, 48, 64), "f64", 7),        # odd: fused pairs + one single sweep
    ("heat_star", (21, 37, 70), "f64", 6),        # extents that are not multiples of the tile
    ("heat_star", (150, 20, 136), "f64", 4),      # several z chunks
    ("heat_star", (24, 40, 72), "f32", 10),
    ("weighted_3d7", (20, 33, 66), "f64", 6),
    ("weighted_3d27", (18, 24, 64), "f64", 4),
])
def test_temporal_blocking_bit_exact(nb, ctx, name, shape, ty, steps):
    """tblock=2 (two time steps per launch, sweep_tb2.cuh) == the oracle's step-by-step loop, bitwise,
    for every compiled temporally blocked kernel; and it really halves the launches."""
    text, n_extra = H.program(name, list(shape), ty)
    assert n_extra == 0
    u, _ = H.inputs(shape, 0, ty, seed=41)
    want = H.oracle_steps(text, u, [], steps)
    m = nb.Module(text)
    try:
        st = m.stencil(ctx, 0)
        ran = []
        for tbv in ["tb2:64x16", "tb2m:128x14", "tb2q:128x14"]:
            try:
                st.set("variant", tbv)
            except nb.NeptuneError:
                assert tbv.startswith(("tb2m:", "tb2q:")) and name == "weighted_3d27"   # z-marching needs centre-column z taps
                continue
            ran.append(tbv)
            fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
            n0 = ctx.launch_count()
            r = st.run_steps(fa, fb, [], steps)
            assert ctx.launch_count() - n0 == steps // 2 + steps % 2
            got = r.download()
            assert H.bits_equal(got, want), "%s differs on %s after %d steps" % (tbv, shape, steps)
            if steps >= 2 and steps % 2 == 0:
                assert st.variant() == tbv
            fa.free(); fb.free()
        assert len(ran) >= (1 if name == "weighted_3d27" else 3)
        st.set("tblock", 1)
        fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
        n0 = ctx.launch_count()
        r = st.run_steps(fa, fb, [], steps)
        assert ctx.launch_count() - n0 == steps
        assert H.bits_equal(r.download(), want)
        fa.free(); fb.free()
    finally:
        m.destroy()


@pytest.mark.parametrize("name,shape,ty,steps", [
    ("heat_star", (300, 1040), "f64", 8),        # three 508-column strips, several stages
    ("heat_box_2d9", (70, 2056), "f64", 7),      # corner taps at both levels; odd step count
    ("heat_box_2d9", (1030, 520), "f32", 6),     # VX = 4: strips of 1016 columns
    ("weighted_2d9", (64, 512), "f64", 4),
    ("heat_star", (5, 512), "f64", 4),           # three interior rows: every stage is ragged
    ("heat_star", (48, 64), "f64", 6),
])
def test_temporal_blocking_2d_bit_exact(nb, ctx, name, shape, ty, steps):
    """sweep_rows2 (rank 2, two steps per launch, level 1 exchanged through shared memory) == the oracle's
    step-by-step loop, bitwise, for every compiled strip shape and several row-chunk sizes."""
    text, n_extra = H.program(name, list(shape), ty)
    assert n_extra == 0
    u, _ = H.inputs(shape, 0, ty, seed=47)
    want = H.oracle_steps(text, u, [], steps)
    m = nb.Module(text)
    try:
        st = m.stencil(ctx, 0)
        for tbv in ["auto", "tb2r:256x3"]:
            for zc in ("0", "16", "7"):
                if tbv != "auto":
                    st.set("variant", tbv)
                st.set("zchunk", zc)
                fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
                n0 = ctx.launch_count()
                r = st.run_steps(fa, fb, [], steps)
                assert ctx.launch_count() - n0 == steps // 2 + steps % 2
                assert H.bits_equal(r.download(), want), "%s zchunk %s differs on %s after %d steps" % (tbv, zc, shape, steps)
                if steps % 2 == 0:
                    assert st.variant().startswith("tb2r"), st.variant()
                fa.free(); fb.free()
    finally:
        m.destroy()


def test_temporal_blocking_falls_back_to_single_sweeps(nb, ctx):
    """Bodies/geometries the fused kernel does not take (extra inputs, 2-D, bounds touching the box)
    run step by step with the same result — still on the GPU, never on the CPU."""
    for name, shape in [("varcoef_3d7", (20, 24, 32)), ("jacobi_star", (48, 64))]:
        text, n_extra = H.program(name, list(shape))
        u, extra = H.inputs(shape, n_extra, "f64", seed=8)
        want = H.oracle_steps(text, u, extra, 6)
        m = nb.Module(text)
        st = m.stencil(ctx, 0).set("tblock", 2)
        fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
        fx = [nb.Field.from_numpy(ctx, e) for e in extra]
        n0 = ctx.launch_count()
        r = st.run_steps(fa, fb, fx, 6)
        assert ctx.launch_count() - n0 == 6
        assert H.bits_equal(r.download(), want)
        for f in [fa, fb] + fx:
            f.free()
        m.destroy()


@pytest.mark.parametrize("shape,ty,steps", [
    ((40, 48, 64), "f64", 8),
    ((21, 37, 70), "f64", 7),        # ragged extents, odd step count
    ((150, 20, 136), "f64", 4),      # several z chunks
    ((24, 40, 72), "f32", 10),
])
def test_temporal_blocking_with_a_pointwise_input(nb, ctx, shape, ty, steps):
    """Jacobi with a right-hand side (BASELINE configs[4]: "temporal blocking"): u and f are two inputs,
    f is read at offset 0 only, and a fused pair of steps is apply(apply(u, f), f).  The z-marching
    kernel streams f from global memory at both levels; the result equals the step-by-step oracle bitwise."""
    text, n_extra = H.program("jacobi_star", list(shape), ty)
    assert n_extra == 1
    u, extra = H.inputs(shape, n_extra, ty, seed=43)
    want = H.oracle_steps(text, u, extra, steps)
    m = nb.Module(text)
    try:
        st = m.stencil(ctx, 0)
        for tbv in ["auto", "tb2m:64x24", "tb2m:128x14"]:
            if tbv != "auto":
                st.set("variant", tbv)
            fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
            fx = [nb.Field.from_numpy(ctx, e) for e in extra]
            n0 = ctx.launch_count()
            r = st.run_steps(fa, fb, fx, steps)
            assert ctx.launch_count() - n0 == steps // 2 + steps % 2
            assert H.bits_equal(r.download(), want), "%s differs on %s after %d steps" % (tbv, shape, steps)
            if steps % 2 == 0 and tbv != "auto":
                assert st.variant() == tbv
            for f in [fa, fb] + fx:
                f.free()
    finally:
        m.destroy()


@pytest.mark.parametrize("name,shape,ty,chunk", [
    ("heat_star", (40, 48, 64), "f64", 3),
    ("heat_star", (37, 20, 66), "f32", 5),
    ("heat_box_2d9", (96, 128), "f64", 7),
    ("jacobi_star", (24, 32, 64), "f64", 1),
    ("varcoef_3d27", (12, 16, 32), "f64", 4),
    ("laplacian_s_1d", (16,), "f64", 2),
])
def test_pipelined_host_sweep_bit_exact(nb, ctx, name, shape, ty, chunk):
    """neptune_rt_apply_host (upload | sweep | download over plane chunks on three streams) == one
    oracle sweep, bitwise, including the copy-through shell and inputs given as host arrays."""
    text, n_extra = H.program(name, list(shape), ty)
    u, extra = H.inputs(shape, n_extra, ty, seed=17)
    want = H.oracle_entry(text, u, extra)
    m = nb.Module(text)
    try:
        st = m.stencil(ctx, 0)
        d = m.apply_desc(0)
        ins = [nb.Field(ctx, u.dtype, d.in_lb[k], d.in_ub[k]) for k in range(1 + n_extra)]
        out = nb.Field(ctx, u.dtype, d.out_lb, d.out_ub)
        got = np.full_like(u, np.nan)
        n0 = ctx.launch_count()
        st.apply_host(ins, [u] + extra, out, got, chunk)
        assert H.bits_equal(got, want)
        nch = -(-shape[0] // chunk)
        assert ctx.launch_count() - n0 == (nch if nch >= 3 else 1)   # one launch per chunk
        # second call with resident coefficient inputs (NULL host pointers) and a new u
        u2 = (u * 0.5).astype(u.dtype)
        want2 = H.oracle_entry(text, u2, extra)
        got2 = np.full_like(u, np.nan)
        st.apply_host(ins, [u2] + [None] * n_extra, out, got2, chunk)
        assert H.bits_equal(got2, want2)
        for f in ins + [out]:
            f.free()
    finally:
        m.destroy()


def test_entry_on_host_memrefs_pipelines_at_size(nb, ctx):
    """@entry(out, in) on host arrays large enough for the plane-chunk pipeline (8 chunks of 8 MiB here)."""
    shape = (64, 256, 512)
    text, _ = H.program("heat_star", list(shape))
    u, _ = H.inputs(shape, 0, "f64", seed=23)
    want = H.oracle_entry(text, u, [])
    m = nb.Module(text)
    try:
        assert "pipelined" in m.plan("@entry")
        out = np.full_like(u, np.nan)
        n0 = ctx.launch_count()
        m.run(ctx, "@entry", [out, u])
        assert ctx.launch_count() - n0 == 8
        assert H.bits_equal(out, want)
    finally:
        m.destroy()


@pytest.mark.parametrize("ty", ["f64", "f32"])
def test_reduce_sum_tolerance_tier(nb, ctx, ty):
    """`neptune_ir.reduce` kind "sum" (DataflowLowering.cpp:589-698).  The reference adds strictly left
    to right; the device sum is a fixed-partition fp64 tree, so parity is the tolerance tier, with the
    bound written here: |gpu - ref| <= n * eps(elem) * sum|x| (the classical bound on the REFERENCE's
    own rounding error; the fp64 tree is far inside it).  Deterministic run to run."""
    from oracle import mlir_subset as ms, numpy_lowering as onp
    dt = np.float64 if ty == "f64" else np.float32
    shape = (40, 33, 130)
    u = np.random.default_rng(21).uniform(-1, 1, size=shape).astype(dt)
    for kw in ({}, {"lb": [2, 3, 4], "sub_lb": [3, 3, 5], "sub_ub": [41, 30, 131]}):
        text = H.mb.field_sum(list(shape), ty, **kw)
        want = float(onp.run_function(ms.parse(text), "@entry", [u])[0])
        m = nb.Module(text)
        try:
            got = m.run(ctx, "@entry", [u])
            again = m.run(ctx, "@entry", [u])
        finally:
            m.destroy()
        assert len(got) == 1 and got == again
        sub = u if not kw else u[1:39, 0:27, 1:127]
        bound = sub.size * float(np.finfo(dt).eps) * float(np.sum(np.abs(sub), dtype=np.float64))
        assert abs(got[0] - want) <= bound, (got[0], want, bound)
        # and against an exactly rounded sum the fp64 tree is within a few ulps of the result type
        exact = float(np.sum(sub.astype(np.longdouble)))
        assert abs(got[0] - exact) <= 4 * float(np.finfo(dt).eps) * max(1.0, abs(exact)) + (1e-9 if ty == "f64" else 0.0)
    # API level: empty box sums to zero, out-of-box bounds are an error, ones sum to the point count
    f = nb.Field.from_numpy(ctx, np.ones(shape, dtype=dt))
    try:
        assert f.reduce_sum() == float(np.prod(shape))
        assert f.reduce_sum([5, 5, 5], [5, 9, 9]) == 0.0
        assert f.reduce_sum([1, 2, 3], [4, 5, 6]) == 27.0
        with pytest.raises(nb.NeptuneError):
            f.reduce_sum([0, 0, 0], [41, 33, 130])
    finally:
        f.free()


def test_sweep_then_reduce_program_on_device(nb, ctx):
    """apply -> store -> reduce of the new state: the field is bit-exact, the scalar within the bound."""
    from oracle import mlir_subset as ms, numpy_lowering as onp
    shape = (24, 40, 136)
    text = H.mb.heat_star_then_sum(list(shape), 0.1)
    u, _ = H.inputs(shape, 0, "f64", seed=31)
    want_out = np.zeros_like(u)
    want_s = float(onp.run_function(ms.parse(text), "@entry", [want_out, u.copy()])[0])
    m = nb.Module(text)
    try:
        out = np.zeros_like(u)
        got = m.run(ctx, "@entry", [out, u.copy()])
    finally:
        m.destroy()
    assert H.bits_equal(out, want_out)
    inner = want_out[1:-1, 1:-1, 1:-1]
    assert abs(got[0] - want_s) <= inner.size * 2.2e-16 * float(np.sum(np.abs(inner)))


@pytest.mark.parametrize("shape,ty,steps", [
    ((64,), "f64", 5),
    ((48, 512), "f64", 6),
    ((24, 40, 136), "f64", 6),
    ((20, 24, 64), "f32", 4),
])
def test_fused_operator_call_plus_axpy_bit_exact(nb, ctx, shape, ty, steps):
    """N3: the explicit Euler step composed as `apply_linear @L` + AXPY apply runs as ONE fused sweep on
    the device and equals the oracle, which executes the two applies one after the other (bit-exact: the
    fused body is the same ops in the same order per point); so does the unfused device plan."""
    if ty == "f32" and len(shape) != 3:
        pytest.skip("catalogue ships the f32 fused body for rank 3 only")
    text = H.mb.euler_opdef(list(shape), 100.0, 1e-3, ty)
    u, _ = H.inputs(shape, 0, ty, seed=91)
    want1 = H.oracle_entry(text, u, [])
    want = H.oracle_steps(text, u, [], steps)
    for pipeline, launches in (("neptuneir-to-b200", 1), ("neptuneir-to-b200{fuse=false}", 2)):
        m = nb.Module(text, pipeline)
        try:
            out = np.full_like(u, np.nan)
            n0 = ctx.launch_count()
            m.run(ctx, "@entry", [out, u.copy()])
            assert ctx.launch_count() - n0 == launches, pipeline
            assert H.bits_equal(out, want1), pipeline
            if launches == 1:
                a, b = u.copy(), np.zeros_like(u)
                res = m.run_steps(ctx, "@entry", [b, a], steps)
                assert H.bits_equal([b, a][res], want), "time loop"
        finally:
            m.destroy()


def test_committed_golden_fixtures_on_device(nb, ctx):
    """Every committed fixture of tests/golden/ (program text, seeded input, N-step result made by the
    oracle, tests/golden/make_golden.py) reproduced bitwise by the device time loop."""
    import os
    gdir = os.path.join(os.path.dirname(__file__), "golden")
    files = sorted(f for f in os.listdir(gdir) if f.endswith(".npz"))
    assert len(files) >= 10
    for f in files:
        g = np.load(os.path.join(gdir, f))
        text, steps = str(g["mlir"]), int(g["steps"])
        extra = [np.ascontiguousarray(g["extra%d" % k]) for k in range(int(g["n_extra"]))]
        a, b = np.array(g["u0"], copy=True), np.array(g["u0"], copy=True)
        m = nb.Module(text)
        try:
            res = m.run_steps(ctx, "@entry", [b, a] + extra, steps)
        finally:
            m.destroy()
        assert H.bits_equal([b, a][res], g["result"]), f


@pytest.mark.parametrize("name,shape,ty,steps", [
    ("jacobi_star", (768, 768, 768), "f64", 4),        # BASELINE configs[4] (C5) on one GPU, fp64
    ("jacobi_star", (768, 768, 768), "f32", 4),        # ... fp32
    ("jacobi_star", (96, 768, 768), "f64", 4),         # ... one GPU's share of 768^3 over 8 GPUs
    ("varcoef_3d27", (128, 1024, 1024), "f64", 2),     # BASELINE configs[3] (C4): one GPU's share of 1024^3 over 8, 27 coefficient fields
])
def test_full_size_properties_with_extra_inputs(nb, ctx, name, shape, ty, steps):
    """C4 and C5 at BASELINE.json's per-GPU sizes (the bodies with pointwise extra inputs), through properties that do
    not need the oracle on the whole box, all bitwise:
      * every kernel family agrees: the default path (fused pairs where the body has them) == single sweeps == the
        direct kernel;
      * linearity under an exact scaling of the state (and of the right-hand side): step(2u, 2f) == 2 step(u, f);
      * the shell is copied through;
      * a band of planes in the middle equals the oracle run on the band's dependency cone (extra inputs cut alike).
    Coefficient fields are made one at a time and only their cone rows stay on the host."""
    dt = H.NPT[ty]
    text, n_extra = H.program(name, list(shape), ty)
    rng = np.random.default_rng(20260407)
    u = rng.uniform(-1, 1, size=shape).astype(dt)
    mid = shape[0] // 2
    rows = slice(mid - 2 * steps, mid + 2 * steps)
    corner = (rows, slice(0, 96), slice(0, 160))
    scale_extra = name == "jacobi_star"                # the rhs enters additively: it scales with the state
    m = nb.Module(text)
    extra, extra_cone = [], []
    try:
        st = m.stencil(ctx, 0)
        for k in range(n_extra):
            c = (rng.uniform(-1, 1, size=shape) if scale_extra else rng.uniform(0.5, 1.5, size=shape) / n_extra).astype(dt)
            extra.append(nb.Field.from_numpy(ctx, c))
            extra_cone.append(np.ascontiguousarray(c[corner]))
            del c
        fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
        first = st.run_steps(fa, fb, extra, steps).download()
        fusable = name == "jacobi_star"
        assert st.variant().startswith("tb2") == fusable, st.variant()
        if fusable:
            st.set("tblock", 1)
            fa.upload(u); fb.upload(u)
            single = st.run_steps(fa, fb, extra, steps).download()
            assert not st.variant().startswith("tb2")
            assert H.bits_equal(first, single), "temporally blocked != single sweeps"
            del single
        st.set("variant", "direct")
        fa.upload(u); fb.upload(u)
        direct = st.run_steps(fa, fb, extra, steps).download()
        assert st.variant() == "direct" and H.bits_equal(direct, first), "staged kernel != direct kernel"
        del direct
        st.set("variant", "auto"); st.set("tblock", 0)
        two = dt(2.0)
        fa.upload(two * u); fb.upload(two * u)
        if scale_extra:
            for f in extra:
                f.upload(two * f.download())
        twice = st.run_steps(fa, fb, extra, steps).download()
        assert H.bits_equal(twice, two * first), "step(2u) != 2 step(u)"
        del twice
        for d in range(3):
            lo = [slice(None)] * 3; hi = [slice(None)] * 3
            lo[d] = 0; hi[d] = -1
            assert H.bits_equal(first[tuple(lo)], u[tuple(lo)]) and H.bits_equal(first[tuple(hi)], u[tuple(hi)])
        fa.free(); fb.free()
    finally:
        for f in extra:
            f.free()
        m.destroy()
    cone = np.ascontiguousarray(u[corner])
    ctext, _ = H.program(name, list(cone.shape), ty)
    if name == "jacobi_star":                          # the builder derives h^2 from the extents: keep the full box's value
        ctext = H.mb.jacobi_star(list(cone.shape), 1.0 / (max(shape) ** 2), ty)
    want = H.oracle_steps(ctext, cone, extra_cone, steps)
    got = first[corner]
    inner = (slice(steps, 3 * steps), slice(0, 96 - steps), slice(0, 160 - steps))
    assert H.bits_equal(np.ascontiguousarray(got[inner]), np.ascontiguousarray(want[inner]))


@pytest.mark.parametrize("name,shape", [
    ("heat_star", (512, 512, 512)),       # BASELINE configs[2] (C3), one GPU's share
    ("heat_box_2d9", (16384, 16384)),     # BASELINE configs[1] (C2)
    ("heat_star", (4096, 4096)),          # BASELINE configs[0] (C1)
])
def test_full_size_properties(nb, ctx, name, shape):
    """At BASELINE.json's full sizes the oracle takes minutes, so parity is checked through size-independent
    properties, all bitwise:
      * every kernel family agrees: 4 steps as two fused pairs == 4 single sweeps == 4 direct-kernel sweeps;
      * linearity under an exact scaling: step(2u) == 2 step(u) (a power of two commutes with every rounding);
      * the shell is copied through; a constant field is a fixed point;
      * a plane/row band cut out of the result equals the oracle run on that band's dependency cone."""
    text, _ = H.program(name, list(shape))
    rank = len(shape)
    rng = np.random.default_rng(20260403)
    u = rng.uniform(-1, 1, size=shape)
    m = nb.Module(text)
    try:
        st = m.stencil(ctx, 0)
        fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
        fused = st.run_steps(fa, fb, [], 4).download()
        assert st.variant().startswith("tb2"), st.variant()
        st.set("tblock", 1)
        fa.upload(u); fb.upload(u)
        single = st.run_steps(fa, fb, [], 4).download()
        assert not st.variant().startswith("tb2")
        assert H.bits_equal(fused, single), "temporally blocked != single sweeps"
        st.set("variant", "direct")
        fa.upload(u); fb.upload(u)
        direct = st.run_steps(fa, fb, [], 4).download()
        assert st.variant() == "direct" and H.bits_equal(direct, single), "staged kernel != direct kernel"
        del direct
        # linearity under scaling by 2
        st.set("variant", "auto"); st.set("tblock", 0)
        fa.upload(2.0 * u); fb.upload(2.0 * u)
        twice = st.run_steps(fa, fb, [], 4).download()
        assert H.bits_equal(twice, 2.0 * fused)
        del twice
        # shell copied through
        for d in range(rank):
            lo = [slice(None)] * rank; hi = [slice(None)] * rank
            lo[d] = 0; hi[d] = -1
            assert H.bits_equal(fused[tuple(lo)], u[tuple(lo)]) and H.bits_equal(fused[tuple(hi)], u[tuple(hi)])
        # constant field is a fixed point
        fa.fill(1.25); fb.fill(1.25)
        const = st.run_steps(fa, fb, [], 4).download()
        assert np.all(const == 1.25)
        del const
        fa.free(); fb.free()
    finally:
        m.destroy()
    # a band of 8 planes/rows in the middle, against the oracle on its dependency cone (4 steps -> 4 planes each side);
    # the cone's own cut faces are wrong after step 1 but cannot reach the band's centre in 4 steps
    mid = shape[0] // 2
    cone = u[mid - 8:mid + 8]
    if rank == 3:
        cone = cone[:, :96, :160]          # keep the oracle's share small: corner block that includes two real faces
        ref_shape = cone.shape
    else:
        cone = cone[:, :4096]
        ref_shape = cone.shape
    ctext, _ = H.program(name, list(ref_shape))
    want = H.oracle_steps(ctext, np.ascontiguousarray(cone), [], 4)
    if rank == 3:
        got = fused[mid - 8:mid + 8, :96, :160]
        assert H.bits_equal(got[4:12, :92, :156], want[4:12, :92, :156])
    else:
        got = fused[mid - 8:mid + 8, :4096]
        assert H.bits_equal(got[4:12, :4092], want[4:12, :4092])


def test_matrix_free_operator_on_caller_owned_device_vectors(nb, ctx):
    """`y = A(x)` on device memory the caller owns (torch CUDA tensors here), the hook a
    device-resident Krylov loop needs — no copies, same kernels, same bits as the oracle.  A few steps of a
    matrix-free Richardson iteration x <- x + w (b - A x) run entirely on the device with torch doing the
    vector algebra; every A(x) is checked against the oracle."""
    import torch
    shape = (24, 40, 136)
    lap = H.mb.weighted_const(list(shape), H.STAR7, [6.0, -1.0, -1.0, -1.0, -1.0, -1.0, -1.0])   # A = -Laplacian (class L body)
    m = nb.Module(lap)
    try:
        st = m.stencil(ctx, 0)
        dev = torch.device("cuda", 0)
        rng = np.random.default_rng(61)
        x_h = rng.uniform(-1, 1, size=shape)
        b_h = rng.uniform(-1, 1, size=shape)
        x = torch.from_numpy(x_h).to(dev)
        b = torch.from_numpy(b_h).to(dev)
        y = torch.empty_like(x)
        fx = nb.Field.wrap_device(ctx, x.data_ptr(), np.float64, shape)
        fy = nb.Field.wrap_device(ctx, y.data_ptr(), np.float64, shape)
        torch.cuda.synchronize()
        for it in range(3):
            st.apply([fx], fy)                    # y = A(x): interior = -Lap(x), shell = x (copy-through)
            ctx.sync()
            want = H.oracle_entry(lap, x.cpu().numpy(), [])
            assert H.bits_equal(y.cpu().numpy(), want), "A(x) differs at iteration %d" % it
            assert st.variant().startswith("tma")
            r = b - y
            r[0], r[-1] = 0, 0; r[:, 0], r[:, -1] = 0, 0; r[:, :, 0], r[:, :, -1] = 0, 0   # Dirichlet shell
            x += 0.08 * r                         # in place: fx still wraps x's memory
            torch.cuda.synchronize()
        fx.free(); fy.free()
        # the wrapped tensors are still alive and untouched by free()
        assert torch.isfinite(x).all() and torch.isfinite(y).all()
        # host memory is refused loudly
        with pytest.raises(nb.NeptuneError):
            nb.Field.wrap_device(ctx, x_h.ctypes.data, np.float64, shape)
    finally:
        m.destroy()


@pytest.mark.parametrize("name,shape,variants", [
    ("heat_star", (96, 112, 256), ["tb2:64x16", "tb2m:128x14", "tb2q:128x14"]),
    ("heat_box_2d9", (1536, 2032), ["tb2r:256x3"]),
    ("jacobi_star", (96, 96, 192), ["tb2m:64x24", "tb2m:128x14"]),
])
def test_temporally_blocked_kernels_repeat_bitwise(nb, ctx, name, shape, variants):
    """compute-sanitizer's racecheck is not available on the pool, so shared-memory hazards of the kernels that
    exchange level-1 values between threads are hunted the blunt way: sizes with many CTAs per SM slot and
    several items per CTA, 12 repetitions per kernel, every result compared bitwise with the single-step
    kernel's (which has no inter-thread exchange).  A missing barrier shows up as a flaky mismatch."""
    text, n_extra = H.program(name, list(shape))
    u, extra = H.inputs(shape, n_extra, "f64", seed=73)
    m = nb.Module(text)
    try:
        st = m.stencil(ctx, 0)
        fx = [nb.Field.from_numpy(ctx, e) for e in extra]
        st.set("tblock", 1)
        fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
        want = st.run_steps(fa, fb, fx, 6).download()
        st.set("tblock", 0)
        for v in variants:
            st.set("variant", v)
            for rep in range(12):
                st.set("zchunk", str([0, 16, 24, 33][rep % 4]))
                fa.upload(u); fb.upload(u)
                got = st.run_steps(fa, fb, fx, 6).download()
                assert st.variant() == v
                assert H.bits_equal(got, want), "%s repetition %d differs" % (v, rep)
        for f in [fa, fb] + fx:
            f.free()
    finally:
        m.destroy()


def test_jit_seven_point_with_permuted_taps_runs_on_every_kernel_family(nb, ctx):
    """A 7-point body whose taps come in another order than any catalogue file (so: another structure key).  The
    NVRTC path compiles the hand-written kernels for it: every single-step variant and the temporally blocked
    z-marchers must reproduce the oracle bit for bit — the body's own association order, not the catalogue's."""
    shape = (40, 48, 264)
    offs = [(0, 0, 1), (0, -1, 0), (1, 0, 0), (0, 0, 0), (-1, 0, 0), (0, 1, 0), (0, 0, -1)]
    text = H.mb.weighted_const(list(shape), offs, [0.11, 0.13, 0.07, 0.4, 0.09, 0.05, 0.15])
    m = nb.Module(text)
    assert not nb.runtime.body_compiled(m.apply_desc(0).key())
    u, _ = H.inputs(shape, 0, "f64", seed=11)
    want1 = H.oracle_entry(text, u, [])
    want6 = H.oracle_steps(text, u, [], 6)
    try:
        st = m.stencil(ctx, 0)          # first use: NVRTC (or the on-disk JIT cache)
        names = st.variants()
        assert "direct" in names and "tma:64x16" in names
        for v in names:
            st.set("variant", v)
            out = np.full_like(u, np.nan)
            m.run(ctx, "@entry", [out, u])
            assert st.variant() == v and H.bits_equal(out, want1), v
        st.set("variant", "auto")
        for tbv in ["tb2:64x16", "tb2m:128x14", "tb2q:128x14"]:
            st.set("variant", tbv)
            fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
            r = st.run_steps(fa, fb, [], 6)
            assert st.variant() == tbv and H.bits_equal(r.download(), want6), tbv
            fa.free(); fb.free()
    finally:
        m.destroy()


def test_jit_two_dimensional_body_takes_the_row_marchers(nb, ctx):
    shape = (96, 1040)
    offs = [(1, 0), (0, 0), (0, -1), (-1, 0), (0, 1)]
    text = H.mb.weighted_const(list(shape), offs, [0.2, 0.3, 0.1, 0.25, 0.15])
    m = nb.Module(text)
    assert not nb.runtime.body_compiled(m.apply_desc(0).key())
    u, _ = H.inputs(shape, 0, "f64", seed=12)
    try:
        out = np.full_like(u, np.nan)
        m.run(ctx, "@entry", [out, u])
        st = m.stencil(ctx, 0)
        assert st.variant().startswith("rows:"), st.variant()
        assert H.bits_equal(out, H.oracle_entry(text, u, []))
        a, b = u.copy(), np.zeros_like(u)
        res = m.run_steps(ctx, "@entry", [b, a], 8)
        assert st.variant().startswith("tb2r:"), st.variant()
        assert H.bits_equal([b, a][res], H.oracle_steps(text, u, [], 8))
    finally:
        m.destroy()


@pytest.mark.parametrize("ty", ["f64", "f32"])
def test_jacobi_1000_iterations_bit_exact(nb, ctx, ty):
    """BASELINE configs[4]'s iteration count at reduced size: 1000 Jacobi iterations with a right-hand side, fused in
    pairs by the entry's time loop, against the oracle's 1000-step loop, bitwise."""
    shape = (64, 64, 64)
    text, n_extra = H.program("jacobi_star", list(shape), ty)
    u, extra = H.inputs(shape, n_extra, ty, seed=1000)
    want = H.oracle_steps(text, u, extra, 1000)
    m = nb.Module(text)
    try:
        a, b = u.copy(), np.zeros_like(u)
        res = m.run_steps(ctx, "@entry", [b, a] + extra, 1000)
        assert m.stencil(ctx, 0).variant().startswith("tb2")
        assert H.bits_equal([b, a][res], want)
    finally:
        m.destroy()


def test_memref_descriptor_entry_points(nb, ctx):
    """neptune_rt_apply_memref_{1,2,3}d — what generated code calls — on MLIR memref descriptors with a non-zero
    offset: cached stencil + fields, pipelined upload | sweep | download, bit-exact; repeated calls reuse the cache."""
    import ctypes as C

    classes = {}

    def memref(rank, buf, offset, shape):
        if rank not in classes:
            class M(C.Structure):
                _fields_ = [("allocated", C.c_void_p), ("aligned", C.c_void_p), ("offset", C.c_int64),
                            ("sizes", C.c_int64 * rank), ("strides", C.c_int64 * rank)]
            classes[rank] = M
        m = classes[rank]()
        m.allocated = m.aligned = buf.ctypes.data
        m.offset = offset
        st = 1
        for d in reversed(range(rank)):
            m.sizes[d] = shape[d]
            m.strides[d] = st
            st *= shape[d]
        return m

    L = nb.runtime.lib()
    for name, shape in [("laplacian_s_1d", (64,)), ("heat_box_2d9", (40, 96)), ("heat_star", (24, 32, 64)), ("jacobi_star", (16, 24, 64))]:
        text, n_extra = H.program(name, list(shape))
        u, extra = H.inputs(shape, n_extra, "f64", seed=21)
        want = H.oracle_entry(text, u, extra)
        m = nb.Module(text)
        d = m.apply_desc(0)
        rank = len(shape)
        pad = 5                                           # elements before the data: descriptor offset
        bufs = []
        for a in [u] + extra:
            b = np.full(a.size + pad, np.nan)
            b[pad:] = a.ravel()
            bufs.append(b)
        outb = np.full(u.size + pad, np.nan)
        first = memref(rank, bufs[0], pad, shape)
        ins = (type(first) * len(bufs))(*[memref(rank, b, pad, shape) for b in bufs])
        om = memref(rank, outb, pad, shape)
        fn = getattr(L, "neptune_rt_apply_memref_%dd" % rank)
        for rep in range(2):
            outb[:] = np.nan
            fn(C.byref(d.c()), len(bufs), C.cast(ins, C.c_void_p), C.cast(C.byref(om), C.c_void_p))
            assert np.isnan(outb[:pad]).all()
            assert H.bits_equal(outb[pad:].reshape(shape), want), name
        # the call sequence the MLIR pass emits (mlir/lib/Passes/ApplyToB200.cpp): packed descriptor in a "global",
        # call_begin, one call_input per operand, call_run — through the _mlir_ciface_ names func-to-llvm produces
        dc = d.c()
        need = L.neptune_rt_desc_pack(C.byref(dc), None, 0)
        blob = C.create_string_buffer(need)
        assert L.neptune_rt_desc_pack(C.byref(dc), blob, need) == need
        blob2 = C.create_string_buffer(need)
        L.neptune_rt_desc_pack(C.byref(d.c()), blob2, need)
        assert blob.raw == blob2.raw                      # reproducible bytes (padding zeroed)
        for rep in range(2):
            outb[:] = np.nan
            call = L._mlir_ciface_neptune_rt_call_begin(blob, need)
            for k in range(len(bufs)):
                getattr(L, "_mlir_ciface_neptune_rt_call_input_%dd" % rank)(call, k, C.cast(C.byref(ins[k]), C.c_void_p))
            getattr(L, "_mlir_ciface_neptune_rt_call_run_%dd" % rank)(call, C.cast(C.byref(om), C.c_void_p))
            assert H.bits_equal(outb[pad:].reshape(shape), want), name
        m.destroy()
    L.neptune_rt_memref_cache_clear()
