"""ctypes binding of libneptune_rt_b200.so — the host side of the `neptuneir-to-b200` path.

The reference's Python-side boundary is `ctypes.CDLL` over the object code its pipeline produced
(python_frontend/neptune/backend.py:75) driven by `neptune-opt`-style pipelines
(src/neptuneOpt.cpp:38-47, lib/Pipeline/NeptuneIRPassesPipeline.cpp:49-53).  This module is the
same boundary for the B200 path: `Module(text, "neptuneir-to-b200")` is the pipeline,
`Module.run(ctx, "@entry", [out, in, ...])` is the host driver's call of the compiled entry
(test/smoke_tests/smoke_apply.sh:39-80), and `Context`/`Field`/`Stencil` expose the
`neptune_rt_*` C ABI (include/neptune_rt.h) one to one.

There is no CPU fallback: if the CUDA library is missing this module raises at import of the
library handle, and without a GPU `Context()` raises `NeptuneError`.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import numpy as np

MAX_RANK = 3
F32, F64 = 0, 1
_NP = {F32: np.float32, F64: np.float64}
_DT = {np.dtype(np.float32): F32, np.dtype(np.float64): F64}

OK, EINVAL, ENODEV, ECUDA, EUNSUPPORTED, ENCCL, ENOMEM = 0, -1, -2, -3, -4, -5, -6
_CODE_NAMES = {EINVAL: "EINVAL", ENODEV: "ENODEV", ECUDA: "ECUDA", EUNSUPPORTED: "EUNSUPPORTED",
               ENCCL: "ENCCL", ENOMEM: "ENOMEM"}

OP_ACCESS, OP_CONST, OP_ADD, OP_SUB, OP_MUL, OP_DIV, OP_NEG = range(7)
FLAG_FMA = 1

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libneptune_rt_b200.so")


class NeptuneError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__("%s: %s" % (_CODE_NAMES.get(code, str(code)), message))
        self.code = code
        self.message = message


class BodyOp(C.Structure):
    _fields_ = [("opcode", C.c_int32), ("a", C.c_int32), ("b", C.c_int32), ("input", C.c_int32),
                ("offset", C.c_int64 * MAX_RANK), ("value", C.c_double)]


class StencilDescC(C.Structure):
    _fields_ = [("rank", C.c_int32), ("dtype", C.c_int32), ("n_inputs", C.c_int32), ("n_ops", C.c_int32),
                ("ops", C.POINTER(BodyOp)), ("yield_value", C.c_int32), ("flags", C.c_int32),
                ("lb", C.c_int64 * MAX_RANK), ("ub", C.c_int64 * MAX_RANK),
                ("out_lb", C.c_int64 * MAX_RANK), ("out_ub", C.c_int64 * MAX_RANK),
                ("in_lb", C.POINTER(C.c_int64)), ("in_ub", C.POINTER(C.c_int64))]


_lib = None


def lib() -> C.CDLL:
    """The CUDA library; fails loudly when it has not been built (`python -m neptune_pde_solver_b200.build`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libneptune_rt_b200.so is not built (%s); run `python -m neptune_pde_solver_b200.build` "
                          "— the B200 path has no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, i64p, i32, i64 = C.c_void_p, C.POINTER(C.c_int64), C.c_int, C.c_int64
    sig = {
        "neptune_rt_init": (i32, [i32, C.POINTER(vp)]),
        "neptune_rt_finalize": (i32, [vp]),
        "neptune_rt_ctx_alive": (i32, [vp]),
        "neptune_rt_sync": (i32, [vp]),
        "neptune_rt_last_error": (C.c_char_p, [vp]),
        "neptune_rt_launch_count": (i64, [vp]),
        "neptune_rt_stream": (vp, [vp]),
        "neptune_rt_timer_start": (i32, [vp]),
        "neptune_rt_timer_stop": (i32, [vp, C.POINTER(C.c_double)]),
        "neptune_rt_field_alloc": (i32, [vp, i32, i32, i64p, i64p, C.POINTER(vp)]),
        "neptune_rt_field_free": (i32, [vp]),
        "neptune_rt_field_upload": (i32, [vp, vp]),
        "neptune_rt_field_download": (i32, [vp, vp]),
        "neptune_rt_field_upload_planes": (i32, [vp, vp, i64, i64]),
        "neptune_rt_field_download_planes": (i32, [vp, vp, i64, i64]),
        "neptune_rt_field_fill": (i32, [vp, C.c_double]),
        "neptune_rt_field_wrap_device": (i32, [vp, i32, i32, i64p, i64p, vp, C.POINTER(vp)]),
        "neptune_rt_field_device_ptr": (vp, [vp]),
        "neptune_rt_field_bytes": (i64, [vp]),
        "neptune_rt_field_store": (i32, [vp, vp, i64p, i64p]),
        "neptune_rt_field_reduce_sum": (i32, [vp, i64p, i64p, i32, C.POINTER(C.c_double)]),
        "neptune_rt_reduce_sum_static": (i32, [vp, i32, i32, i64p, i64p, i64p, vp, vp]),
        "neptune_rt_body_key": (C.c_uint64, [C.POINTER(StencilDescC)]),
        "neptune_rt_body_compiled": (i32, [C.c_uint64]),
        "neptune_rt_body_jit_compile": (i32, [C.POINTER(StencilDescC), C.POINTER(C.c_double), C.c_char_p, C.c_size_t]),
        "neptune_rt_stencil_create": (i32, [vp, C.POINTER(StencilDescC), C.POINTER(vp)]),
        "neptune_rt_stencil_destroy": (i32, [vp]),
        "neptune_rt_stencil_set": (i32, [vp, C.c_char_p, C.c_char_p]),
        "neptune_rt_stencil_variant": (C.c_char_p, [vp]),
        "neptune_rt_stencil_variant_count": (i32, [vp]),
        "neptune_rt_stencil_variant_name": (C.c_char_p, [vp, i32]),
        "neptune_rt_apply": (i32, [vp, i32, C.POINTER(vp), vp]),
        "neptune_rt_apply_planes": (i32, [vp, i32, C.POINTER(vp), vp, i64, i64, vp]),
        "neptune_rt_apply_planes2": (i32, [vp, i32, C.POINTER(vp), vp, i64, i64, i64, i64, vp]),
        "neptune_rt_apply_host": (i32, [vp, i32, C.POINTER(vp), C.POINTER(vp), vp, vp, i64]),
        "neptune_rt_run_steps": (i32, [vp, vp, vp, i32, C.POINTER(vp), i64, C.POINTER(vp)]),
        "neptune_rt_slab_partition": (i32, [i64, i32, i32, i64p, i64p]),
        "neptune_rt_slab_localize": (i32, [C.POINTER(StencilDescC), i32, i32, i64, C.POINTER(StencilDescC), i64p, i64p]),
        "neptune_rt_stencil_set_slab": (i32, [vp, i64, i64, i64]),
        "neptune_rt_stencil_halo_mode": (C.c_char_p, [vp]),
        "neptune_rt_slab_schedule": (i64, [i64, i64, i64, i32, vp, i64]),
        "neptune_rt_comm_unique_id": (i32, [vp]),
        "neptune_rt_comm_init": (i32, [vp, vp, i32, i32]),
        "neptune_rt_halo_exchange": (i32, [vp, vp, i64, i64, i64, vp]),
        "neptune_rt_slab_trace": (i32, [vp, i32]),
        "neptune_rt_slab_trace_read": (i32, [vp, C.POINTER(C.c_double), i32, C.POINTER(i32)]),
        "neptune_rt_apply_memref": (None, [C.POINTER(StencilDescC), i32, C.POINTER(vp), vp]),
        "neptune_rt_apply_memref_1d": (None, [C.POINTER(StencilDescC), i32, vp, vp]),
        "neptune_rt_apply_memref_2d": (None, [C.POINTER(StencilDescC), i32, vp, vp]),
        "neptune_rt_apply_memref_3d": (None, [C.POINTER(StencilDescC), i32, vp, vp]),
        "neptune_rt_memref_cache_clear": (None, []),
        "neptune_rt_desc_pack": (i64, [C.POINTER(StencilDescC), vp, i64]),
        "neptune_rt_call_begin": (vp, [vp, i64]),
        "neptune_rt_call_input_1d": (None, [vp, i32, vp]),
        "neptune_rt_call_input_2d": (None, [vp, i32, vp]),
        "neptune_rt_call_input_3d": (None, [vp, i32, vp]),
        "neptune_rt_call_run_1d": (None, [vp, vp]),
        "neptune_rt_call_run_2d": (None, [vp, vp]),
        "neptune_rt_call_run_3d": (None, [vp, vp]),
        "_mlir_ciface_neptune_rt_call_begin": (vp, [vp, i64]),
        "_mlir_ciface_neptune_rt_call_input_1d": (None, [vp, i32, vp]),
        "_mlir_ciface_neptune_rt_call_input_2d": (None, [vp, i32, vp]),
        "_mlir_ciface_neptune_rt_call_input_3d": (None, [vp, i32, vp]),
        "_mlir_ciface_neptune_rt_call_run_1d": (None, [vp, vp]),
        "_mlir_ciface_neptune_rt_call_run_2d": (None, [vp, vp]),
        "_mlir_ciface_neptune_rt_call_run_3d": (None, [vp, vp]),
        "neptune_b200_compile": (i32, [C.c_char_p, C.c_char_p, C.POINTER(vp), C.c_char_p, C.c_size_t]),
        "neptune_b200_module_destroy": (i32, [vp]),
        "neptune_b200_apply_count": (i32, [vp]),
        "neptune_b200_apply_desc": (C.POINTER(StencilDescC), [vp, i32]),
        "neptune_b200_apply_function": (C.c_char_p, [vp, i32]),
        "neptune_b200_function_plan": (C.c_char_p, [vp, C.c_char_p]),
        "neptune_b200_function_arg_count": (i32, [vp, C.c_char_p]),
        "neptune_b200_function_arg_numel": (i64, [vp, C.c_char_p, i32]),
        "neptune_b200_last_error": (C.c_char_p, [vp]),
        "neptune_b200_function_count": (i32, [vp]),
        "neptune_b200_function_name": (C.c_char_p, [vp, i32]),
        "neptune_b200_function_diagnostic": (C.c_char_p, [vp, C.c_char_p]),
        "neptune_b200_apply_stencil": (vp, [vp, vp, i32]),
        "neptune_b200_apply_cuda_source": (C.c_char_p, [vp, i32]),
        "neptune_b200_run_host": (i32, [vp, vp, C.c_char_p, i32, C.POINTER(vp)]),
        "neptune_b200_last_scalars": (i32, [vp, C.POINTER(C.c_double), i32]),
        "neptune_b200_run_steps_host": (i32, [vp, vp, C.c_char_p, i32, C.POINTER(vp), i64, C.POINTER(i32)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)          # AttributeError here = header/library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    L._neptune_signatures = sig
    _lib = L
    return L


def _i64(v: Sequence[int]):
    return (C.c_int64 * len(v))(*[int(x) for x in v])


class StencilDesc:
    """Owning Python copy of a neptune_rt_stencil_desc."""

    def __init__(self, rank, dtype, n_inputs, ops, yield_value, lb, ub, out_lb, out_ub, in_lb, in_ub, flags=0):
        self.rank, self.dtype, self.n_inputs = int(rank), int(dtype), int(n_inputs)
        self.ops = list(ops)        # dicts: opcode, a, b, input, offset, value
        self.yield_value, self.flags = int(yield_value), int(flags)
        self.lb, self.ub = list(lb), list(ub)
        self.out_lb, self.out_ub = list(out_lb), list(out_ub)
        self.in_lb = [list(x) for x in in_lb]   # n_inputs x rank
        self.in_ub = [list(x) for x in in_ub]
        self._keep = None

    @classmethod
    def from_c(cls, d: StencilDescC) -> "StencilDesc":
        r, n = d.rank, d.n_inputs
        ops = []
        for i in range(d.n_ops):
            o = d.ops[i]
            ops.append(dict(opcode=o.opcode, a=o.a, b=o.b, input=o.input, offset=[o.offset[k] for k in range(r)],
                            value=o.value))
        return cls(r, d.dtype, n, ops, d.yield_value, [d.lb[k] for k in range(r)], [d.ub[k] for k in range(r)],
                   [d.out_lb[k] for k in range(r)], [d.out_ub[k] for k in range(r)],
                   [[d.in_lb[i * r + k] for k in range(r)] for i in range(n)],
                   [[d.in_ub[i * r + k] for k in range(r)] for i in range(n)], d.flags)

    def c(self) -> StencilDescC:
        d = StencilDescC()
        ops = (BodyOp * max(1, len(self.ops)))()
        for i, o in enumerate(self.ops):
            ops[i].opcode, ops[i].a, ops[i].b, ops[i].input = o["opcode"], o.get("a", -1), o.get("b", -1), o.get("input", -1)
            for k, v in enumerate(o.get("offset", [])):
                ops[i].offset[k] = int(v)
            ops[i].value = float(o.get("value", 0.0))
        flat_lb = _i64([x for row in self.in_lb for x in row] or [0])
        flat_ub = _i64([x for row in self.in_ub for x in row] or [0])
        d.rank, d.dtype, d.n_inputs, d.n_ops = self.rank, self.dtype, self.n_inputs, len(self.ops)
        d.ops = C.cast(ops, C.POINTER(BodyOp))
        d.yield_value, d.flags = self.yield_value, self.flags
        for k in range(self.rank):
            d.lb[k], d.ub[k], d.out_lb[k], d.out_ub[k] = self.lb[k], self.ub[k], self.out_lb[k], self.out_ub[k]
        d.in_lb = C.cast(flat_lb, C.POINTER(C.c_int64))
        d.in_ub = C.cast(flat_ub, C.POINTER(C.c_int64))
        self._keep = (ops, flat_lb, flat_ub, d)
        return d

    def key(self) -> int:
        return int(lib().neptune_rt_body_key(C.byref(self.c())))

    def radius0(self) -> int:
        return max([abs(o["offset"][0]) for o in self.ops if o["opcode"] == OP_ACCESS] or [0])


def body_compiled(key: int) -> bool:
    return bool(lib().neptune_rt_body_compiled(C.c_uint64(key)))


def body_jit_compile(desc: "StencilDesc") -> float:
    """NVRTC-compile the sweep kernels for a body outside the prebuilt catalogue (fills the on-disk cache; needs no
    GPU). Returns the seconds it took."""
    sec = C.c_double()
    err = C.create_string_buffer(4096)
    rc = lib().neptune_rt_body_jit_compile(C.byref(desc.c()), C.byref(sec), err, len(err))
    if rc:
        raise NeptuneError(rc, err.value.decode())
    return sec.value


def slab_partition(n0: int, nranks: int, rank: int):
    b, e = C.c_int64(), C.c_int64()
    rc = lib().neptune_rt_slab_partition(n0, nranks, rank, C.byref(b), C.byref(e))
    if rc:
        raise NeptuneError(rc, lib().neptune_rt_last_error(None).decode())
    return b.value, e.value


def slab_localize(desc: StencilDesc, nranks: int, rank: int, halo: int) -> StencilDesc:
    g = desc.c()
    l = StencilDescC()
    n = max(1, desc.n_inputs * desc.rank)
    lb, ub = (C.c_int64 * n)(), (C.c_int64 * n)()
    rc = lib().neptune_rt_slab_localize(C.byref(g), nranks, rank, halo, C.byref(l), lb, ub)
    if rc:
        raise NeptuneError(rc, lib().neptune_rt_last_error(None).decode())
    return StencilDesc.from_c(l)


class SlabLaunch(C.Structure):
    _fields_ = [("steps", C.c_int32), ("exchange", C.c_int32), ("ext", C.c_int64)]


def slab_schedule(nsteps: int, radius: int, ghost: int, fuse2: bool):
    """The launch plan of neptune_rt_run_steps in slab mode: [(steps, ext, exchange)] per launch."""
    n = lib().neptune_rt_slab_schedule(nsteps, radius, ghost, int(fuse2), None, 0)
    if n < 0:
        raise NeptuneError(EINVAL, "slab_schedule: impossible arguments")
    plan = (SlabLaunch * max(1, n))()
    lib().neptune_rt_slab_schedule(nsteps, radius, ghost, int(fuse2), plan, n)
    return [(plan[i].steps, plan[i].ext, bool(plan[i].exchange)) for i in range(n)]


def comm_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    rc = lib().neptune_rt_comm_unique_id(buf)
    if rc:
        raise NeptuneError(rc, lib().neptune_rt_last_error(None).decode())
    return buf.raw


class Context:
    """neptune_rt_ctx: one per process/GPU; owns the streams and the NCCL communicator."""

    def __init__(self, device: int = -1):
        self._L = lib()
        h = C.c_void_p()
        rc = self._L.neptune_rt_init(device, C.byref(h))
        if rc:
            raise NeptuneError(rc, self._L.neptune_rt_last_error(None).decode())
        self.h = h

    def _check(self, rc: int):
        if rc:
            raise NeptuneError(rc, self._L.neptune_rt_last_error(self.h).decode())

    def sync(self):
        self._check(self._L.neptune_rt_sync(self.h))

    def launch_count(self) -> int:
        return int(self._L.neptune_rt_launch_count(self.h))

    def stream(self) -> int:
        return int(self._L.neptune_rt_stream(self.h) or 0)

    def timer_start(self):
        self._check(self._L.neptune_rt_timer_start(self.h))

    def timer_stop(self) -> float:
        ms = C.c_double()
        self._check(self._L.neptune_rt_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def slab_trace(self, enable: bool = True):
        self._check(self._L.neptune_rt_slab_trace(self.h, int(enable)))

    def slab_trace_read(self):
        """[(kind, begin_ms, boundary_done_ms, exchange_done_ms, interior_done_ms)] of the last traced run_steps."""
        n = self._L.neptune_rt_slab_trace_read(self.h, None, 0, None)
        if n <= 0:
            return []
        buf, kinds = (C.c_double * (4 * n))(), (C.c_int * n)()
        self._L.neptune_rt_slab_trace_read(self.h, buf, n, kinds)
        return [(kinds[g],) + tuple(buf[4 * g + w] for w in range(4)) for g in range(n)]

    def comm_init(self, unique_id: bytes, nranks: int, rank: int):
        self._check(self._L.neptune_rt_comm_init(self.h, C.c_char_p(unique_id), nranks, rank))

    def close(self):
        if self.h:
            self._L.neptune_rt_finalize(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class Field:
    """neptune_rt_field: device storage of a !neptune_ir.field / !temp (row-major box ub-lb)."""

    def __init__(self, ctx: Context, dtype, lb: Sequence[int], ub: Sequence[int]):
        self.ctx = ctx
        self.dtype = dtype if isinstance(dtype, int) else _DT[np.dtype(dtype)]
        self.lb, self.ub = [int(x) for x in lb], [int(x) for x in ub]
        self.shape = tuple(u - l for l, u in zip(self.lb, self.ub))
        h = C.c_void_p()
        ctx._check(ctx._L.neptune_rt_field_alloc(ctx.h, self.dtype, len(self.lb), _i64(self.lb), _i64(self.ub), C.byref(h)))
        self.h = h

    @classmethod
    def from_numpy(cls, ctx: Context, a: np.ndarray, lb: Optional[Sequence[int]] = None) -> "Field":
        lb = [0] * a.ndim if lb is None else list(lb)
        f = cls(ctx, a.dtype, lb, [l + n for l, n in zip(lb, a.shape)])
        f.upload(a)
        return f

    @classmethod
    def wrap_device(cls, ctx: Context, device_ptr: int, dtype, shape: Sequence[int], lb: Optional[Sequence[int]] = None) -> "Field":
        """A field over device memory the caller owns (e.g. `tensor.data_ptr()` of a contiguous CUDA tensor)."""
        lb = [0] * len(shape) if lb is None else list(lb)
        ub = [l + int(n) for l, n in zip(lb, shape)]
        f = cls.__new__(cls)
        f.ctx = ctx
        f.dtype = dtype if isinstance(dtype, int) else _DT[np.dtype(dtype)]
        f.lb, f.ub = [int(x) for x in lb], [int(x) for x in ub]
        f.shape = tuple(int(n) for n in shape)
        f.h = C.c_void_p()
        ctx._check(ctx._L.neptune_rt_field_wrap_device(ctx.h, f.dtype, len(shape), _i64(f.lb), _i64(f.ub), C.c_void_p(device_ptr), C.byref(f.h)))
        return f

    def _host(self, a: np.ndarray) -> np.ndarray:
        if tuple(a.shape) != self.shape or a.dtype != _NP[self.dtype] or not a.flags["C_CONTIGUOUS"]:
            raise ValueError("host buffer %s/%s does not match field %s/%s" % (a.shape, a.dtype, self.shape, _NP[self.dtype]))
        return a

    def upload(self, a: np.ndarray):
        a = self._host(a)
        self.ctx._check(self.ctx._L.neptune_rt_field_upload(self.h, a.ctypes.data))
        self.ctx.sync()   # pageable host memory: the buffer may be reused by the caller

    def download(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty(self.shape, dtype=_NP[self.dtype])
        out = self._host(out)
        self.ctx._check(self.ctx._L.neptune_rt_field_download(self.h, out.ctypes.data))
        self.ctx.sync()
        return out

    def upload_planes(self, a: np.ndarray, p0: int, p1: int):
        a = self._host(a)
        self.ctx._check(self.ctx._L.neptune_rt_field_upload_planes(self.h, a.ctypes.data, p0, p1))

    def download_planes(self, a: np.ndarray, p0: int, p1: int):
        a = self._host(a)
        self.ctx._check(self.ctx._L.neptune_rt_field_download_planes(self.h, a.ctypes.data, p0, p1))

    def fill(self, value: float):
        self.ctx._check(self.ctx._L.neptune_rt_field_fill(self.h, float(value)))

    def store_from(self, src: "Field", lb=None, ub=None):
        """`neptune_ir.store src to self` (whole, or the sub-box [lb, ub))."""
        if lb is None:
            rc = self.ctx._L.neptune_rt_field_store(self.h, src.h, None, None)
        else:
            rc = self.ctx._L.neptune_rt_field_store(self.h, src.h, _i64(lb), _i64(ub))
        self.ctx._check(rc)

    def reduce_sum(self, lb=None, ub=None, allreduce: bool = False) -> float:
        """`neptune_ir.reduce` kind="sum" over [lb, ub) (None = whole box); tolerance tier (fp64 tree sum)."""
        r = C.c_double()
        self.ctx._check(self.ctx._L.neptune_rt_field_reduce_sum(
            self.h, None if lb is None else _i64(lb), None if ub is None else _i64(ub), int(allreduce), C.byref(r)))
        return r.value

    def device_ptr(self) -> int:
        return int(self.ctx._L.neptune_rt_field_device_ptr(self.h) or 0)

    def nbytes(self) -> int:
        return int(self.ctx._L.neptune_rt_field_bytes(self.h))

    def free(self):
        if self.h:
            if self.ctx.h:            # a closed ctx released its fields itself (neptune_rt_finalize)
                self.ctx._L.neptune_rt_field_free(self.h)
            self.h = None


class Stencil:
    """neptune_rt_stencil: one recognised apply body bound to compiled sm_100a kernels."""

    def __init__(self, ctx: Context, desc: Optional[StencilDesc] = None, _handle=None):
        self.ctx = ctx
        self.desc = desc
        self._owned = _handle is None
        if _handle is None:
            h = C.c_void_p()
            ctx._check(ctx._L.neptune_rt_stencil_create(ctx.h, C.byref(desc.c()), C.byref(h)))
            self.h = h
        else:
            self.h = C.c_void_p(_handle)

    def set(self, key: str, value) -> "Stencil":
        self.ctx._check(self.ctx._L.neptune_rt_stencil_set(self.h, key.encode(), str(value).encode()))
        return self

    def set_slab(self, ghost: int, global_lb0: int, global_ub0: int) -> "Stencil":
        """Slab mode: ghost planes per neighbour side and the global dim-0 apply bounds (lets
        run_steps fuse two steps per launch when ghost >= 2 * radius)."""
        self.ctx._check(self.ctx._L.neptune_rt_stencil_set_slab(self.h, ghost, global_lb0, global_ub0))
        return self

    def variant(self) -> str:
        return self.ctx._L.neptune_rt_stencil_variant(self.h).decode()

    def halo_mode(self) -> str:
        """'p2p' (copy engines + flags over NVLink), 'nccl', or '' before the first slab-mode run_steps."""
        return self.ctx._L.neptune_rt_stencil_halo_mode(self.h).decode()

    def variants(self) -> List[str]:
        n = self.ctx._L.neptune_rt_stencil_variant_count(self.h)
        return [self.ctx._L.neptune_rt_stencil_variant_name(self.h, i).decode() for i in range(n)]

    def apply(self, ins: Sequence[Field], out: Field, planes=None, stream: int = 0):
        arr = (C.c_void_p * len(ins))(*[f.h for f in ins])
        if planes is None:
            rc = self.ctx._L.neptune_rt_apply(self.h, len(ins), arr, out.h)
        elif len(planes) == 4:   # (p0, g0, g1, p1): planes [p0, g0) and [g1, p1) in one launch
            rc = self.ctx._L.neptune_rt_apply_planes2(self.h, len(ins), arr, out.h, planes[0], planes[1], planes[2], planes[3], stream or None)
        else:
            rc = self.ctx._L.neptune_rt_apply_planes(self.h, len(ins), arr, out.h, planes[0], planes[1], stream or None)
        self.ctx._check(rc)

    def apply_host(self, ins: Sequence[Field], ins_host: Sequence[Optional[np.ndarray]], out: Field,
                   out_host: np.ndarray, chunk_planes: int = 0):
        """One sweep on host arrays, upload | sweep | download pipelined over dim-0 plane chunks
        (neptune_rt_apply_host). ins_host[k] None = input k is already resident in ins[k]."""
        arr = (C.c_void_p * len(ins))(*[f.h for f in ins])
        keep = [None if h is None else np.ascontiguousarray(h) for h in ins_host]
        hp = (C.c_void_p * len(ins))(*[None if h is None else h.ctypes.data for h in keep])
        assert out_host.flags["C_CONTIGUOUS"]
        self.ctx._check(self.ctx._L.neptune_rt_apply_host(self.h, len(ins), arr, hp, out.h, out_host.ctypes.data, chunk_planes))

    def run_steps(self, a: Field, b: Field, extra: Sequence[Field], nsteps: int) -> Field:
        arr = (C.c_void_p * max(1, len(extra)))(*[f.h for f in extra])
        res = C.c_void_p()
        self.ctx._check(self.ctx._L.neptune_rt_run_steps(self.h, a.h, b.h, len(extra), arr, nsteps, C.byref(res)))
        return a if res.value == a.h.value else b

    def destroy(self):
        if self.h and self._owned and self.ctx.h:
            self.ctx._L.neptune_rt_stencil_destroy(self.h)
        self.h = None


class Module:
    """A NeptuneIR module taken through the `neptuneir-to-b200` pipeline."""

    def __init__(self, text: str, pipeline: str = "neptuneir-to-b200"):
        self._L = lib()
        h = C.c_void_p()
        err = C.create_string_buffer(2048)
        rc = self._L.neptune_b200_compile(text.encode(), pipeline.encode(), C.byref(h), err, len(err))
        if rc:
            raise NeptuneError(rc, err.value.decode())
        self.h = h

    def _check(self, rc: int):
        if rc:
            raise NeptuneError(rc, self._L.neptune_b200_last_error(self.h).decode())

    def apply_count(self) -> int:
        return self._L.neptune_b200_apply_count(self.h)

    def apply_desc(self, i: int) -> StencilDesc:
        p = self._L.neptune_b200_apply_desc(self.h, i)
        if not p:
            raise IndexError(i)
        return StencilDesc.from_c(p.contents)

    def apply_function(self, i: int) -> str:
        return self._L.neptune_b200_apply_function(self.h, i).decode()

    def cuda_source(self, i: int) -> str:
        return self._L.neptune_b200_apply_cuda_source(self.h, i).decode()

    def plan(self, function: str = "@entry") -> str:
        p = self._L.neptune_b200_function_plan(self.h, function.encode())
        if p is None:
            raise KeyError(function)
        return p.decode()

    def functions(self) -> List[str]:
        return [self._L.neptune_b200_function_name(self.h, i).decode() for i in range(self._L.neptune_b200_function_count(self.h))]

    def diagnostic(self, function: str) -> Optional[str]:
        """Why `function` was not lowered (None: it was)."""
        p = self._L.neptune_b200_function_diagnostic(self.h, function.encode())
        return None if p is None else p.decode()

    def arg_count(self, function: str = "@entry") -> int:
        return self._L.neptune_b200_function_arg_count(self.h, function.encode())

    def arg_numel(self, function: str, arg: int) -> int:
        return int(self._L.neptune_b200_function_arg_numel(self.h, function.encode(), arg))

    def stencil(self, ctx: Context, i: int = 0) -> Stencil:
        h = self._L.neptune_b200_apply_stencil(self.h, ctx.h, i)
        if not h:
            raise NeptuneError(EUNSUPPORTED, self._L.neptune_rt_last_error(ctx.h).decode())
        return Stencil(ctx, self.apply_desc(i), _handle=h)

    @staticmethod
    def _ptrs(args: Sequence[np.ndarray]):
        for a in args:
            if not isinstance(a, np.ndarray) or not a.flags["C_CONTIGUOUS"]:
                raise ValueError("entry arguments must be C-contiguous numpy arrays")
        return (C.c_void_p * len(args))(*[a.ctypes.data for a in args])

    def run(self, ctx: Context, function: str, args: Sequence[np.ndarray]):
        """One call of the compiled entry on host memrefs (in place, like the reference's driver)."""
        self._check(self._L.neptune_b200_run_host(self.h, ctx.h, function.encode(), len(args), self._ptrs(args)))
        n = self._L.neptune_b200_last_scalars(self.h, None, 0)
        if n <= 0:
            return []
        buf = (C.c_double * n)()
        self._L.neptune_b200_last_scalars(self.h, buf, n)
        return list(buf)

    def run_steps(self, ctx: Context, function: str, args: Sequence[np.ndarray], nsteps: int) -> int:
        """`for step: entry(out, in, extra...); swap(out, in)` resident on the device.
        Returns the index (0 or 1) of the argument that holds the final state."""
        res = C.c_int()
        self._check(self._L.neptune_b200_run_steps_host(self.h, ctx.h, function.encode(), len(args), self._ptrs(args),
                                                        nsteps, C.byref(res)))
        return res.value

    def destroy(self):
        if self.h:
            self._L.neptune_b200_module_destroy(self.h)
            self.h = None


def neptune_opt(text: str, pipeline: str = "neptuneir-to-b200") -> Module:
    """`neptune-opt in.mlir --neptuneir-to-b200` (the driver's pipeline flag, src/neptuneOpt.cpp:38-47)."""
    return Module(text, pipeline.lstrip("-"))
