"""Packs the reference's OWN current-form NeptuneIR fixtures as test inputs (tests/golden/reference_inputs.json).

The reference holds no numeric golden vectors for the apply path, but two of its smoke inputs are in
the current assembly form and contain hot-path operators:

    /root/reference/test/smoke_tests/smoke_time_advance.mlir     @ac_lap (:11-31), @ac_A (:34-51), the %ustar apply (:59-71)
    /root/reference/test/smoke_tests/smoke_time_advance_bs.mlir  @bs_A (:31-49, a composite linear body)

Their text is stored VERBATIM (with its sha256, so that tests/test_reference_inputs.py can check it against
/root/reference where that exists; the GPU box has no /root/reference, which is why the text travels as a fixture),
together with seeded inputs and the oracle's outputs as hex floats.  These are test INPUTS the builder did not author;
the expected outputs still come from this repo's oracle, so parity stays "unpinned" by the rules.

    python tests/golden/make_reference_inputs.py
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import mlir_subset as ms, numpy_lowering as onp  # noqa: E402

REF = "/root/reference/test/smoke_tests"
FILES = ["smoke_time_advance.mlir", "smoke_time_advance_bs.mlir"]


def hexlist(a):
    return [float(x).hex() for x in np.asarray(a, dtype=np.float64).ravel()]


def ustar_entry(text):
    """@entry of smoke_time_advance.mlir with the neptune_ir.time_advance statement (:73-80, outside the hot path) cut
    out and %ustar stored instead of %u1: the file's own pointwise apply (:59-71) as a runnable entry.  Derived
    mechanically from the verbatim text; everything else is untouched."""
    a = text.index("    // (I - alpha*Lap) u_{n+1} = u*")
    b = text.index("    neptune_ir.store %u1 to %fout")
    out = text[:a] + text[b:].replace("neptune_ir.store %u1 to %fout", "neptune_ir.store %ustar to %fout", 1)
    assert "time_advance" not in out
    return out


def main():
    pack = {"files": {}, "cases": []}
    texts = {}
    for f in FILES:
        t = open(os.path.join(REF, f)).read()
        texts[f] = t
        pack["files"][f] = {"path": "test/smoke_tests/" + f, "sha256": hashlib.sha256(t.encode()).hexdigest(), "text": t}
    rng = np.random.default_rng(20260501)

    def case(fname, text_kind, function, n, reps):
        text = texts[fname] if text_kind == "verbatim" else ustar_entry(texts[fname])
        mod = ms.parse(text)
        u = np.sin(2 * np.pi * np.arange(n) / n) + 0.25 * rng.uniform(-1, 1, n)
        outs = []
        x = u.copy()
        for _ in range(reps):
            if function == "@entry":
                y = np.full_like(x, np.nan)
                onp.run_function(mod, "@entry", [y, x])
            else:
                y = onp.run_function(mod, function, [x])[0]
            outs.append(hexlist(y))
            x = np.array(y, copy=True)
        pack["cases"].append({"file": fname, "text": text_kind, "function": function, "n": n, "input": hexlist(u), "outputs": outs})

    case("smoke_time_advance.mlir", "verbatim", "@ac_lap", 16, 3)
    case("smoke_time_advance.mlir", "verbatim", "@ac_A", 16, 3)        # apply_linear @ac_lap + two-input AXPY apply
    case("smoke_time_advance.mlir", "ustar_entry", "@entry", 16, 3)    # pointwise u + dt (u - u^3)
    case("smoke_time_advance_bs.mlir", "verbatim", "@bs_A", 32, 3)     # composite linear DAG
    with open(os.path.join(HERE, "reference_inputs.json"), "w") as f:
        json.dump(pack, f, indent=1)
    print("wrote reference_inputs.json:", [(c["function"], c["n"]) for c in pack["cases"]])


if __name__ == "__main__":
    main()
