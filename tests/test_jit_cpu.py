"""Host-only checks of the run-time compilation path and of the packed descriptor (no GPU needed: NVRTC is a compiler)."""
import ctypes as C
import os

import numpy as np
import pytest

import helpers as H


def test_nvrtc_compiles_an_unseen_body_to_an_sm100a_cubin(nb, tmp_path, monkeypatch):
    """A recognised body that is not in the prebuilt catalogue: NVRTC builds the hand-written kernels for its functor
    (csrc/runtime/jit.cpp), the cubin lands in the on-disk cache, the second request is served from it."""
    monkeypatch.setenv("NEPTUNE_B200_JIT_CACHE", str(tmp_path / "jit"))
    offs = [(0, 1, 0), (0, 0, 0), (0, 0, -1), (1, 0, 0), (0, -1, 0), (-1, 0, 0), (0, 0, 1)]
    text = H.mb.weighted_const([16, 16, 32], offs, [0.1, 0.4, 0.1, 0.1, 0.1, 0.1, 0.1])
    m = nb.Module(text)
    d = m.apply_desc(0)
    assert not nb.runtime.body_compiled(d.key())
    t_first = nb.runtime.body_jit_compile(d)
    files = sorted(os.listdir(tmp_path / "jit"))
    assert len(files) == 1 and files[0].endswith(".nbjit"), files           # one file: names and code cannot be mispaired
    path = tmp_path / "jit" / files[0]
    assert (os.stat(tmp_path / "jit").st_mode & 0o777) == 0o700 and (os.stat(path).st_mode & 0o077) == 0
    blob = open(path, "rb").read()
    assert blob[:8] == b"NBJIT001"
    _, names_bytes, cubin_bytes = np.frombuffer(blob[8:32], dtype=np.uint64)
    names = blob[32:32 + int(names_bytes)].decode().split()
    cubin = blob[32 + int(names_bytes):32 + int(names_bytes) + int(cubin_bytes)]
    assert len(blob) == 32 + int(names_bytes) + int(cubin_bytes) + 8
    assert cubin[:4] == b"\x7fELF" and len(cubin) > 100000
    assert any("sweep_tb2q" in n for n in names) and any("sweep_tma" in n for n in names) and any("sweep_direct" in n for n in names)
    assert t_first > 0.2            # a real compile ...
    # ... and a body the recogniser rejects is still rejected: this is not a fallback
    bad = text.replace("arith.mulf", "arith.maximumf", 1)
    with pytest.raises(nb.NeptuneError):
        nb.Module(bad)
    m.destroy()


def test_packed_descriptor_is_position_independent_and_reproducible(nb):
    text, _ = H.program("varcoef_3d7", [8, 10, 12])
    m = nb.Module(text)
    d = m.apply_desc(0)
    L = nb.runtime.lib()
    need = L.neptune_rt_desc_pack(C.byref(d.c()), None, 0)
    assert need == 8 * 4 + 12 * 8 + 2 * d.n_inputs * d.rank * 8 + len(d.ops) * C.sizeof(nb.runtime.BodyOp)
    a, b = C.create_string_buffer(need), C.create_string_buffer(need)
    assert L.neptune_rt_desc_pack(C.byref(d.c()), a, need) == need
    assert L.neptune_rt_desc_pack(C.byref(d.c()), b, need) == need      # a second, separately allocated copy of the descriptor
    assert a.raw == b.raw and a.raw[:4] == b"02BN"                       # little-endian 0x4e423230
    assert L.neptune_rt_desc_pack(C.byref(d.c()), a, need - 1) == need   # too small: only reports the size
    m.destroy()
