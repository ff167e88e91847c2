"""Generates tests/golden/*.npz from the CPU oracle (numpy restatement of
lib/Passes/DataflowLowering.cpp:258-448; cross-checked against the gcc restatement in
tests/test_oracle.py).  The reference cannot be built or imported in this image (LLVM/MLIR
absent), so these vectors pin the oracle against drift, not against the reference.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H  # noqa: E402

CASES = [  # name, shape, dtype, steps, lb
    ("heat_star", (12, 16, 20), "f64", 8, None),       # C3 body
    ("heat_star", (24, 32), "f64", 10, None),          # C1 body
    ("heat_box_2d9", (24, 32), "f64", 10, None),       # C2 body
    ("varcoef_3d27", (8, 10, 12), "f64", 3, None),     # C4 body
    ("jacobi_star", (10, 12, 16), "f64", 12, None),    # C5 body f64
    ("jacobi_star", (10, 12, 16), "f32", 12, None),    # C5 body f32
    ("heat_star", (40,), "f64", 20, (7,)),
    ("weighted_3d27", (8, 9, 10), "f64", 2, (3, -2, 5)),
    ("euler_opdef", (10, 12, 16), "f64", 6, None),     # apply_linear @L + AXPY apply (N3)
    ("euler_opdef", (24, 32), "f64", 6, None),
]


def main():
    for name, shape, ty, steps, lb in CASES:
        text, n_extra = H.program(name, list(shape), ty, list(lb) if lb else None)
        u, extra = H.inputs(shape, n_extra, ty, seed=20260400 + len(shape))
        res = H.oracle_steps(text, u, extra, steps)
        d = dict(mlir=np.array(text), u0=u, steps=steps, n_extra=n_extra, result=res)
        for k, e in enumerate(extra):
            d["extra%d" % k] = e
        fn = "%s_%s_%s.npz" % (name, "x".join(map(str, shape)), ty)
        np.savez_compressed(os.path.join(HERE, fn), **d)
        print(fn, res.dtype, res.shape)


if __name__ == "__main__":
    main()
