"""Small runs of every kernel family for compute-sanitizer (memcheck / racecheck / synccheck):
python scripts/dev/sanitize_small.py   (under `compute-sanitizer --tool X`).  Bit-compares with the oracle too."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H  # noqa: E402
import neptune_pde_solver_b200 as nb  # noqa: E402

ctx = nb.Context(0)
CASES = [("heat_star", (36, 30, 264), ["tma:64x16", "tb2:64x16", "tb2m:128x14", "tb2q:128x14"]),
         ("heat_box_2d9", (40, 1040), ["rows:512", "tb2r:256x3"]),
         ("jacobi_star", (20, 30, 72), ["tma:64x16", "tb2m:64x24"]),
         ("varcoef_3d27", (8, 12, 64), ["tma:64x8"])]
for name, shape, variants in CASES:
    text, n_extra = H.program(name, list(shape))
    u, extra = H.inputs(shape, n_extra, "f64", seed=5)
    want = H.oracle_steps(text, u, extra, 4)
    m = nb.Module(text)
    st = m.stencil(ctx, 0)
    for v in variants:
        st.set("variant", v)
        if not v.startswith("tb2"):
            st.set("tblock", 1)
        fa, fb = nb.Field.from_numpy(ctx, u), nb.Field.from_numpy(ctx, u)
        fx = [nb.Field.from_numpy(ctx, e) for e in extra]
        r = st.run_steps(fa, fb, fx, 4)
        ok = H.bits_equal(r.download(), want)
        print(name, v, st.variant(), "bit-exact" if ok else "MISMATCH", flush=True)
        assert ok
        for f in [fa, fb] + fx:
            f.free()
        st.set("tblock", 0)
    m.destroy()
ctx.close()
print("done")
