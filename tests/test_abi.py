"""The C-ABI library loads and exports every symbol include/*.h declares (no compute calls)."""
import ctypes
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    syms = []
    for h in ("neptune_rt.h", "neptune_b200.h"):
        text = open(os.path.join(ROOT, "include", h)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        syms += re.findall(r"\b(neptune_(?:rt|b200)_[a-z0-9_]+)\s*\(", text)
    return sorted(set(syms))


def test_every_declared_symbol_is_exported(nb):
    L = nb.lib()
    syms = declared_symbols()
    assert len(syms) >= 45
    for s in syms:
        assert hasattr(L, s), "header declares %s but the library does not export it" % s
    # and the Python binding covers the whole header
    assert set(syms) <= set(L._neptune_signatures), set(syms) - set(L._neptune_signatures)


def test_headers_are_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "neptune_b200.h"\nint main(void){neptune_rt_stencil_desc d; (void)d; return 0;}\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                           "-c", str(src), "-o", str(tmp_path / "t.o")])


def test_library_holds_sm100a_tma_kernels(nb):
    """The shipped .so holds sm_100a SASS whose staged kernels load through the Tensor Memory Accelerator (UTMALDG) and
    synchronise on mbarriers (SYNCS); the bit-exact tier contains NO fused multiply-add (DFMA == 0) while its FMA twin
    does; the default kernels read their windows with LDS, not generic loads (scripts/sass_census.py)."""
    import sys
    out = subprocess.run(["cuobjdump", "-lelf", nb.runtime.LIB_PATH], stdout=subprocess.PIPE, text=True).stdout
    assert "sm_100a" in out
    assert ctypes.sizeof(nb.runtime.BodyOp) == 48
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import sass_census as sc
    key = "16c0ec9675fa5599"          # 3-D 7-point fp64 heat step (BASELINE configs[2])
    syms = [s for s in sc.kernels_of() if key in s]
    by = {}
    for s in syms:
        name = sc.demangle(s)
        fam = name.split("neptune::")[1].split("<")[0]
        fma = name.split("(")[0].rstrip(">").endswith("true")
        by[(fam, fma)] = sc.census(s)
    for fam in ("sweep_tma", "sweep_tb2", "sweep_tb2m", "sweep_tb2q"):
        plain, fused = by[(fam, False)], by[(fam, True)]
        assert plain["UTMALDG"] >= 1 and plain["SYNCS"] >= 10, (fam, plain)
        assert plain["DFMA"] == 0 and plain["DADD"] > 0 and plain["DMUL"] > 0, (fam, plain)     # never contracted
        assert fused["DFMA"] > 0, (fam, fused)
        if fam != "sweep_tma":     # sweep_tma reads its ring through generic pointers for single-input bodies (measured faster, tiles.cuh)
            assert plain["LDS"] > 0, (fam, plain)
    assert by[("sweep_direct", False)]["DFMA"] == 0 and by[("sweep_direct", True)]["DFMA"] > 0
