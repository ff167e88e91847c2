"""In-tree build of libneptune_rt_b200.so (no cmake needed; nvcc + g++ only).

Steps
  1. g++  : neptune-b200-gen  (front end + code generator, host only)
  2. gen  : catalogue/*.mlir -> build/gen/body_<key>.cu  (one point functor per body structure)
  3. nvcc : every generated unit + runtime/runtime.cu for sm_100a (-lineinfo), in parallel
  4. g++  : front end objects; link everything into lib/libneptune_rt_b200.so

`python -m neptune_pde_solver_b200.build` or `build()`; incremental (content-hashed stamps).
"""
from __future__ import annotations

import concurrent.futures as cf
import glob
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(HERE, "build")
GEN = os.path.join(BUILD, "gen")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libneptune_rt_b200.so")
CATALOGUE = os.path.join(HERE, "catalogue")
INCLUDE = os.path.join(ROOT, "include")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ARCH + ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default",
                     "--expt-relaxed-constexpr", "-I", CSRC, "-I", INCLUDE]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-I", CSRC, "-I", INCLUDE]

FRONTEND = ["frontend/parser.cpp", "frontend/recognizer.cpp", "frontend/codegen.cpp", "frontend/fusion.cpp"]
HOST_SRCS = FRONTEND + ["frontend/module.cpp", "runtime/registry.cpp", "runtime/jit.cpp"]
CUDA_SRCS = ["runtime/runtime.cu"]
KERNEL_HEADERS = sorted(glob.glob(os.path.join(CSRC, "kernels", "*"))) + sorted(glob.glob(os.path.join(CSRC, "frontend", "*.h"))) \
    + sorted(glob.glob(os.path.join(INCLUDE, "*.h")))


def _digest(paths, extra=""):
    h = hashlib.sha256(extra.encode())
    for p in paths:
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _run(cmd, verbose):
    if verbose:
        print("+", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed (%d): %s\n%s" % (r.returncode, " ".join(cmd), r.stdout))
    return r.stdout


def _compile(src, obj, cmd, deps, verbose):
    """Compile src -> obj unless the stamp (hash of src + deps + flags) is unchanged."""
    stamp = obj + ".stamp"
    want = _digest([src] + deps, " ".join(cmd))
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == want:
        return ""
    out = _run(cmd + ["-c", src, "-o", obj], verbose)
    with open(stamp, "w") as f:
        f.write(want)
    return out


def build(verbose: bool = False, jobs: int | None = None) -> str:
    os.makedirs(GEN, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    jobs = jobs or max(1, (os.cpu_count() or 4))
    # 1. generator
    gen_exe = os.path.join(BUILD, "neptune-b200-gen")
    gen_srcs = [os.path.join(CSRC, s) for s in FRONTEND + ["tools/gen_main.cpp"]]
    stamp = gen_exe + ".stamp"
    want = _digest(gen_srcs + KERNEL_HEADERS)
    if not (os.path.exists(gen_exe) and os.path.exists(stamp) and open(stamp).read() == want):
        _run([CXX] + CXX_FLAGS + gen_srcs + ["-o", gen_exe], verbose)
        open(stamp, "w").write(want)
    # 2. catalogue -> generated units
    mlirs = sorted(glob.glob(os.path.join(CATALOGUE, "*.mlir")))
    if not mlirs:
        raise RuntimeError("empty kernel catalogue: " + CATALOGUE)
    listing = _run([gen_exe, GEN] + mlirs, verbose)
    units = sorted(set(l.strip() for l in listing.splitlines() if l.strip().endswith(".cu")))
    for stale in set(glob.glob(os.path.join(GEN, "body_*.cu"))) - set(units):
        os.remove(stale)
    # 3./4. objects
    tasks = []
    for u in units:
        obj = os.path.join(BUILD, os.path.basename(u)[:-3] + ".o")
        tasks.append((u, obj, [NVCC] + NVCC_FLAGS, KERNEL_HEADERS))
    for s in CUDA_SRCS:
        obj = os.path.join(BUILD, os.path.basename(s)[:-3] + ".o")
        tasks.append((os.path.join(CSRC, s), obj, [NVCC] + NVCC_FLAGS, KERNEL_HEADERS))
    for s in HOST_SRCS:
        obj = os.path.join(BUILD, os.path.basename(s)[:-4] + ".o")
        tasks.append((os.path.join(CSRC, s), obj, [CXX] + CXX_FLAGS + ["-I", "/usr/local/cuda/include"], KERNEL_HEADERS))
    with cf.ThreadPoolExecutor(max_workers=jobs) as ex:
        futs = [ex.submit(_compile, *t, verbose) for t in tasks]
        outs = [f.result() for f in futs]
    if verbose:
        for o in outs:
            if o.strip():
                print(o)
    objs = [t[1] for t in tasks]
    for stale in set(glob.glob(os.path.join(BUILD, "body_*.o"))) - set(objs):
        os.remove(stale)
    link_stamp = LIB + ".stamp"
    want = _digest(objs)
    if not (os.path.exists(LIB) and os.path.exists(link_stamp) and open(link_stamp).read() == want):
        _run([NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-ldl", "-lpthread"], verbose)
        open(link_stamp, "w").write(want)
    # 5. neptune-opt-b200: the neptune-opt-shaped driver (csrc/tools/neptune_opt_b200.cpp), linked against the library
    opt_exe = os.path.join(LIBDIR, "neptune-opt-b200")
    opt_src = os.path.join(CSRC, "tools", "neptune_opt_b200.cpp")
    stamp = opt_exe + ".stamp"
    want = _digest([opt_src, LIB] + sorted(glob.glob(os.path.join(INCLUDE, "*.h"))))
    if not (os.path.exists(opt_exe) and os.path.exists(stamp) and open(stamp).read() == want):
        _run([CXX] + CXX_FLAGS + [opt_src, "-o", opt_exe, "-L", LIBDIR, "-lneptune_rt_b200", "-Wl,-rpath,$ORIGIN"], verbose)
        open(stamp, "w").write(want)
    return LIB


if __name__ == "__main__":
    lib = build(verbose="-v" in sys.argv)
    print(lib)
