"""neptune-b200: the `neptune_ir.apply` stencil sweep of levia-than/neptune-pde-solver on B200.

Only what the hot path needs lives here:
  csrc/       NeptuneIR front end, body recogniser, CUDA code generator, sm_100a sweep kernels,
              the neptune_rt_* / neptune_b200_* C ABI (include/*.h)
  catalogue/  the apply bodies the library ships compiled kernels for (.mlir)
  runtime.py  ctypes host side mirroring the reference's driver-level interface
  mlir_builders.py  authoring helpers for the 2-D/3-D NeptuneIR programs of BASELINE.json
  build.py    in-tree build of lib/libneptune_rt_b200.so
"""
from .runtime import (  # noqa: F401
    Context, Field, Module, NeptuneError, Stencil, StencilDesc, body_compiled, comm_unique_id, lib,
    neptune_opt, slab_localize, slab_partition, F32, F64,
)
