"""The drop-in boundary exercised from plain C, the way a maintainer's generated code would use it: `neptune-opt-b200
--emit=c-desc` writes the stencil descriptor as C data, tests/c_driver/memref_driver.c (shaped like the driver the
reference's smoke script generates, test/smoke_tests/smoke_apply.sh:39-80) builds MLIR memref descriptors and calls
neptune_rt_apply_memref_{1,2,3}d; gcc compiles and links it against libneptune_rt_b200.so alone."""
import os
import subprocess

import numpy as np
import pytest

import helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "neptune_pde_solver_b200", "lib")
OPT = os.path.join(LIBDIR, "neptune-opt-b200")


def _build(tmp, text):
    mlir = os.path.join(tmp, "in.mlir")
    open(mlir, "w").write(text)
    r = subprocess.run([OPT, mlir, "--neptuneir-to-b200", "--emit=c-desc", "-o", os.path.join(tmp, "desc.h")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    exe = os.path.join(tmp, "memref_driver")
    r = subprocess.run(["gcc", "-std=c99", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-I", tmp,
                        os.path.join(ROOT, "tests", "c_driver", "memref_driver.c"), "-o", exe,
                        "-L", LIBDIR, "-lneptune_rt_b200", "-Wl,-rpath," + LIBDIR], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_neptune_opt_driver_surface(tmp_path):
    """The neptune-opt-shaped executable (src/neptuneOpt.cpp:38-47): pipeline flag, --pass-pipeline form, diagnostics."""
    text, _ = H.program("heat_star", [8, 8, 8])
    p = tmp_path / "a.mlir"
    p.write_text(text)
    r = subprocess.run([OPT, str(p), "--neptuneir-to-b200"], capture_output=True, text=True)
    assert r.returncode == 0 and "neptune_rt_apply(stencil #0" in r.stdout and "store fused" in r.stdout
    r = subprocess.run([OPT, str(p), "--pass-pipeline=builtin.module(neptuneir-to-b200{fma=true})", "--emit=keys"], capture_output=True, text=True)
    assert r.returncode == 0 and "prebuilt" in r.stdout
    r = subprocess.run([OPT, str(p), "--neptuneir-to-llvm"], capture_output=True, text=True)
    assert r.returncode == 2 and "no CPU fallback" in r.stderr
    bad = tmp_path / "b.mlir"
    bad.write_text(text.replace("^bb0(%i0: index, %i1: index, %i2: index,", "^bb0(%i0: index,"))
    r = subprocess.run([OPT, str(bad), "--neptuneir-to-b200"], capture_output=True, text=True)
    assert r.returncode == 1 and "block arg count" in r.stderr


def test_c_driver_compiles_and_links_without_a_gpu(tmp_path):
    text, _ = H.program("heat_star", [12, 16, 32])
    exe = _build(str(tmp_path), text)
    assert os.path.exists(exe)


@pytest.mark.gpu
@pytest.mark.parametrize("name,shape", [("laplacian_s_1d", (64,)), ("heat_box_2d9", (40, 96)), ("heat_star", (24, 32, 64)), ("jacobi_star", (16, 24, 64))])
def test_c_driver_matches_the_oracle(tmp_path, name, shape):
    text, n_extra = H.program(name, list(shape))
    u, extra = H.inputs(shape, n_extra, "f64", seed=31)
    exe = _build(str(tmp_path), text)
    files = []
    for k, a in enumerate([u] + extra):
        f = str(tmp_path / ("in%d.bin" % k))
        a.tofile(f)
        files.append(f)
    out = str(tmp_path / "out.bin")
    r = subprocess.run([exe, out, str(1 + n_extra)] + files, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr + r.stdout
    got = np.fromfile(out, dtype=np.float64).reshape(shape)
    assert H.bits_equal(got, H.oracle_entry(text, u, extra))


@pytest.mark.gpu
def test_c_driver_rejects_a_strided_memref_loudly(tmp_path):
    text, _ = H.program("heat_star", [12, 16, 32])
    u, _ = H.inputs((12, 16, 32), 0, "f64", seed=32)
    exe = _build(str(tmp_path), text)
    f = str(tmp_path / "in0.bin")
    u.tofile(f)
    env = dict(os.environ, NEPTUNE_DRIVER_BAD_STRIDE="1")
    r = subprocess.run([exe, str(tmp_path / "out.bin"), "1", f], capture_output=True, text=True, env=env, timeout=120)
    assert r.returncode != 0 and "identity layout" in r.stderr
