"""The reference's own current-form fixtures as inputs (tests/golden/reference_inputs.json, packed verbatim by
tests/golden/make_reference_inputs.py from /root/reference/test/smoke_tests/smoke_time_advance{,_bs}.mlir).

CPU part: the packed text is the reference's file byte for byte (checked where /root/reference exists), the front end
lowers the operator definitions and rejects only @entry (neptune_ir.time_advance), the oracle reproduces the packed
outputs, and a scalar re-derivation written out by hand from the file's op list (the third independent statement of
the same arithmetic) agrees bit for bit.
GPU part: the device runs @ac_lap, @ac_A, the %ustar apply and @bs_A and matches the packed outputs bitwise; three of
the four bodies are outside the prebuilt catalogue and go through the NVRTC path."""
import hashlib
import json
import os

import numpy as np
import pytest

import helpers as H
from oracle import mlir_subset as ms, numpy_lowering as onp

PACK = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_inputs.json")))
REF = "/root/reference"


def _unhex(xs):
    return np.array([float.fromhex(x) for x in xs], dtype=np.float64)


def _text(case):
    t = PACK["files"][case["file"]]["text"]
    if case["text"] == "ustar_entry":
        from golden.make_reference_inputs import ustar_entry
        t = ustar_entry(t)
    return t


def test_packed_text_is_the_reference_file_verbatim():
    for name, f in PACK["files"].items():
        assert hashlib.sha256(f["text"].encode()).hexdigest() == f["sha256"]
        p = os.path.join(REF, f["path"])
        if os.path.exists(p):        # this container; the GPU box has no /root/reference
            assert open(p).read() == f["text"], name


def test_front_end_lowers_what_is_on_the_path_and_diagnoses_the_rest():
    import neptune_pde_solver_b200 as nb
    m = nb.Module(PACK["files"]["smoke_time_advance.mlir"]["text"])
    assert m.functions() == ["@ac_lap", "@ac_A", "@entry"]
    assert m.diagnostic("@ac_lap") is None and m.diagnostic("@ac_A") is None
    assert "time_advance" in m.diagnostic("@entry") and "no CPU fallback" in m.diagnostic("@entry")
    assert m.apply_count() == 2 and "neptune_rt_apply" in m.plan("@ac_A")
    m.destroy()
    m = nb.Module(PACK["files"]["smoke_time_advance_bs.mlir"]["text"])
    assert m.diagnostic("@bs_A") is None and "time_advance" in m.diagnostic("@entry")
    m.destroy()
    # a module none of whose functions lowers still fails, with the first diagnostic
    with pytest.raises(nb.NeptuneError) as e:
        nb.Module("func.func @f(%a: memref<?xf64>) -> memref<?xf64> {\n %x = neptune_ir.time_advance %a : memref<?xf64>\n}")
    assert "time_advance" in str(e.value)


@pytest.mark.parametrize("idx", range(len(PACK["cases"])))
def test_oracle_reproduces_the_packed_outputs(idx):
    c = PACK["cases"][idx]
    mod = ms.parse(_text(c))
    x = _unhex(c["input"])
    for want in c["outputs"]:
        if c["function"] == "@entry":
            y = np.full_like(x, np.nan)
            onp.run_function(mod, "@entry", [y, x])
        else:
            y = onp.run_function(mod, c["function"], [x])[0]
        assert H.bits_equal(y, _unhex(want))
        x = np.array(y, copy=True)


def _by_hand(function, u):
    """The file's op lists written out as scalar Python (IEEE double, one rounding per op, the file's own order):
    smoke_time_advance.mlir:17-28 (@ac_lap), :41-48 (@ac_A), :62-69 (%ustar); smoke_time_advance_bs.mlir:19-48 (@bs_A)."""
    n = len(u)
    out = [float(v) for v in u]               # copy-through of input 0 (DataflowLowering.cpp:282-287)

    def lap(x, i):
        t0 = 2.0 * x[i]
        t1 = x[i - 1] - t0
        t2 = t1 + x[i + 1]
        return 100.0 * t2
    if function == "@ac_lap":
        for i in range(1, n - 1):
            out[i] = lap(u, i)
    elif function == "@ac_A":
        l = [float(v) for v in u]
        for i in range(1, n - 1):
            l[i] = lap(u, i)
        for i in range(1, n - 1):
            t = 1.0e-4 * l[i]
            out[i] = u[i] - t
    elif function == "@entry":
        for i in range(1, n - 1):
            v = float(u[i])
            u2 = v * v
            u3 = u2 * v
            react = v - u3
            out[i] = v + 1.0e-2 * react
    elif function == "@bs_A":
        for i in range(1, n - 1):
            vm1, v0, vp1 = float(u[i - 1]), float(u[i]), float(u[i + 1])
            t0 = 2.0 * v0
            t1 = vm1 - t0
            t2 = t1 + vp1
            vxx = 100.0 * t2
            t3 = vp1 - vm1
            vx = 5.0 * t3
            t4 = 2.0e-2 * vxx
            t5 = 3.0e-2 * vx
            t6 = t4 + t5
            t7 = -5.0e-2 * v0
            L = t6 + t7
            dtL = 1.0e-2 * L
            out[i] = v0 - dtL
    return np.array(out)


@pytest.mark.parametrize("idx", range(len(PACK["cases"])))
def test_hand_derivation_agrees_with_the_packed_outputs(idx):
    c = PACK["cases"][idx]
    x = _unhex(c["input"])
    for want in c["outputs"]:
        y = _by_hand(c["function"], x)
        assert H.bits_equal(y, _unhex(want)), c["function"]
        x = y


@pytest.mark.gpu
@pytest.mark.parametrize("idx", range(len(PACK["cases"])))
def test_device_runs_the_reference_fixtures_bit_exact(idx):
    import neptune_pde_solver_b200 as nb
    c = PACK["cases"][idx]
    ctx = nb.Context(0)
    m = nb.Module(_text(c))
    try:
        x = _unhex(c["input"])
        for want in c["outputs"]:
            y = np.full_like(x, np.nan)
            if c["function"] == "@entry":
                m.run(ctx, "@entry", [y, x])
            else:
                m.run(ctx, c["function"], [x, y])     # operator definition: operands, then the result buffer
            assert H.bits_equal(y, _unhex(want)), c["function"]
            x = np.array(y, copy=True)
        if c["function"] == "@entry" and c["text"] == "verbatim":
            pytest.fail("unreachable")
    finally:
        m.destroy()
        ctx.close()


@pytest.mark.gpu
def test_reference_entry_with_time_advance_fails_loudly_on_device():
    import neptune_pde_solver_b200 as nb
    ctx = nb.Context(0)
    m = nb.Module(PACK["files"]["smoke_time_advance.mlir"]["text"])
    try:
        x = np.zeros(16)
        with pytest.raises(nb.NeptuneError) as e:
            m.run(ctx, "@entry", [x.copy(), x])
        assert e.value.code == nb.runtime.EUNSUPPORTED and "time_advance" in str(e.value)
    finally:
        m.destroy()
        ctx.close()
