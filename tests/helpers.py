"""Shared test helpers: the programs under test (NeptuneIR text) and seeded inputs.
The oracle (oracle/) is the checker here and nowhere in the product path."""
import itertools
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from neptune_pde_solver_b200 import mlir_builders as mb  # noqa: E402
from oracle import mlir_subset as ms  # noqa: E402
from oracle import numpy_lowering as onp  # noqa: E402

STAR5 = [(0, 0)] + mb.star_offsets(2)
STAR7 = [(0, 0, 0)] + mb.star_offsets(3)
BOX9 = mb.box_offsets(2)
BOX27 = mb.box_offsets(3)
NPT = {"f64": np.float64, "f32": np.float32}


def program(name, shape, ty="f64", lb=None):
    """NeptuneIR text of a catalogue body at the given extents. Returns (text, n_extra_inputs)."""
    if name == "heat_star":
        a = {1: 0.25, 2: 0.2, 3: 0.1}[len(shape)]
        return mb.heat_star(shape, a, ty, lb), 0
    if name == "heat_box_2d9":
        return mb.heat_box_2d9(shape, 0.1, ty, lb), 0
    if name == "jacobi_star":
        return mb.jacobi_star(shape, 1.0 / (max(shape) ** 2), ty, lb), 1
    if name == "euler_opdef":      # operator call + AXPY apply (two applies; fused on the device path)
        assert lb is None
        return mb.euler_opdef(shape, 100.0, 1e-3, ty), 0
    if name == "laplacian_s_1d":
        assert lb is None
        return mb.laplacian_s_1d(shape[0], 100.0, ty), 0
    if name == "weighted_2d5":
        return mb.weighted_const(shape, STAR5, [0.5, 0.125, 0.125, 0.125, 0.125], ty, lb), 0
    if name == "weighted_2d9":
        return mb.weighted_const(shape, BOX9, [0.05 + 0.01 * i for i in range(9)], ty, lb), 0
    if name == "weighted_3d7":
        return mb.weighted_const(shape, STAR7, [0.4, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1], ty, lb), 0
    if name == "weighted_3d27":
        return mb.weighted_const(shape, BOX27, [0.02 + 0.001 * i for i in range(27)], ty, lb), 0
    if name == "varcoef_2d5":
        return mb.varcoef(shape, STAR5, ty, lb), 5
    if name == "varcoef_3d7":
        return mb.varcoef(shape, STAR7, ty, lb), 7
    if name == "varcoef_3d27":
        return mb.varcoef(shape, BOX27, ty, lb), 27
    raise KeyError(name)


# (name, rank, dtypes) of every body in the compiled catalogue
CATALOGUE = [
    ("heat_star", 1, ("f64", "f32")), ("heat_star", 2, ("f64", "f32")), ("heat_star", 3, ("f64", "f32")),
    ("heat_box_2d9", 2, ("f64", "f32")), ("jacobi_star", 2, ("f64", "f32")), ("jacobi_star", 3, ("f64", "f32")),
    ("laplacian_s_1d", 1, ("f64", "f32")),
    ("weighted_2d5", 2, ("f64",)), ("weighted_2d9", 2, ("f64",)), ("weighted_3d7", 3, ("f64",)),
    ("weighted_3d27", 3, ("f64",)), ("varcoef_2d5", 2, ("f64",)), ("varcoef_3d7", 3, ("f64",)),
    ("varcoef_3d27", 3, ("f64",)),
]


def inputs(shape, n_extra, ty, seed):
    """Seeded O(1) inputs: a smooth mode plus noise for u, U(0.5,1.5)/K coefficient fields."""
    rng = np.random.default_rng(seed)
    dt = NPT[ty]
    grids = np.meshgrid(*[np.arange(n) / max(n, 1) for n in shape], indexing="ij")
    u = np.ones(shape)
    for g in grids:
        u = u * np.sin(2 * np.pi * g + 0.3)
    u = (u + 0.25 * rng.uniform(-1, 1, size=shape)).astype(dt)
    extra = [(rng.uniform(0.5, 1.5, size=shape) / max(n_extra, 1)).astype(dt) for _ in range(n_extra)]
    return u, extra


def oracle_entry(text, u, extra):
    """One call of @entry(out, in, extra...) through the numpy restatement; returns out."""
    mod = ms.parse(text)
    out = np.full_like(u, np.nan)
    onp.run_function(mod, "@entry", [out, u] + list(extra))
    return out


def oracle_steps(text, u, extra, nsteps):
    return onp.run_steps(ms.parse(text), u, extra, nsteps)


def bits_equal(a, b):
    return a.shape == b.shape and a.dtype == b.dtype and np.array_equal(
        a.view(np.uint64 if a.dtype == np.float64 else np.uint32),
        b.view(np.uint64 if b.dtype == np.float64 else np.uint32))
