"""The MLIR-side sources under mlir/ cannot be compiled on this image (no MLIR).  What can be checked without a
compiler is checked here: the patch they need in the reference tree still applies, and the sources agree with each other
on what the patch exports and on the stages the shared patterns require."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATCH = os.path.join(ROOT, "mlir", "patches", "0001-export-dataflow-patterns.patch")
REFERENCE = "/root/reference"


def _read(*parts):
    return open(os.path.join(ROOT, *parts)).read()


def test_patch_only_adds_and_names_what_the_pass_uses():
    patch = open(PATCH).read()
    removed = [l for l in patch.splitlines() if l.startswith("-") and not l.startswith("---")]
    assert removed == [], removed                                      # nothing of the reference is moved or rewritten
    pass_src = _read("mlir", "lib", "Passes", "ApplyToB200.cpp")
    for name in ("createNeptuneIRTypeConverter", "populateNeptuneIRStoragePatterns"):
        assert name in patch and name in pass_src
    assert '#include "Passes/NeptuneIRDataflowPatterns.h"' in pass_src and "b/include/Passes/NeptuneIRDataflowPatterns.h" in patch
    assert "nextId" not in pass_src and "static unsigned id" not in pass_src   # no mutable static inside a pattern


def test_scf_is_loaded_legal_and_lowered_because_reduce_builds_loops():
    td = _read("mlir", "include", "Passes", "NeptuneIRB200Passes.td")
    pass_src = _read("mlir", "lib", "Passes", "ApplyToB200.cpp")
    pipe = _read("mlir", "lib", "Pipeline", "NeptuneIRToB200Pipeline.cpp")
    assert '"mlir::scf::SCFDialect"' in td
    assert re.search(r"addLegalDialect<[^;]*scf::SCFDialect", pass_src)
    stages = re.findall(r"pm\.addPass\(create(\w+)\(", pipe)
    assert "SCFToControlFlowPass" in stages
    assert stages.index("NeptuneIRApplyToB200Pass") < stages.index("SCFToControlFlowPass") < stages.index("ConvertControlFlowToLLVMPass")
    # the option the pass never implemented is gone from the record, the options struct and the pipeline
    assert '"fuse"' not in td and "opt.fuse" not in pipe and "Option<bool> fuse" not in _read("mlir", "include", "Passes", "NeptuneIRB200Passes.h")


@pytest.mark.skipif(not os.path.isdir(os.path.join(REFERENCE, "lib", "Passes")), reason="the reference tree is only present in the build container")
def test_patch_applies_to_the_reference_tree(tmp_path):
    work = tmp_path / "ref"
    os.makedirs(work / "lib" / "Passes")
    os.makedirs(work / "include" / "Passes")
    shutil.copy(os.path.join(REFERENCE, "lib", "Passes", "DataflowLowering.cpp"), work / "lib" / "Passes")
    subprocess.check_call(["git", "init", "-q"], cwd=work)
    subprocess.check_call(["git", "apply", "--check", PATCH], cwd=work)
    subprocess.check_call(["git", "apply", PATCH], cwd=work)
    src = open(work / "lib" / "Passes" / "DataflowLowering.cpp").read()
    hdr = open(work / "include" / "Passes" / "NeptuneIRDataflowPatterns.h").read()
    # every pattern the export lists is a struct the reference file defines
    listed = re.search(r"populateNeptuneIRStoragePatterns\([^{]*\{\s*patterns\.add<([^>]*)>", src, re.S).group(1)
    for name in [n.strip() for n in listed.split(",")]:
        assert re.search(r"struct %s\b" % name, src), name
    assert "createNeptuneIRTypeConverter" in hdr and "struct NeptuneIRTypeConverter" in src
