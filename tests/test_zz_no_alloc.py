"""neptune_rt_reduce_sum_static: the box sum for callers that are not allowed to allocate.

Why the entry point exists (include/neptune_rt.h has the full contract): the ordinary
`neptune_rt_field_reduce_sum` cudaMallocs its partial-sum scratch the first time it runs.  A Krylov
solver that sized its own device pool at start-up, a framework whose caching allocator already holds
all of HBM, or compiler-generated code that owns every buffer cannot tolerate a library that goes to
cudaMalloc behind its back (it may simply fail, and it synchronises the whole device in the middle of
an overlapped time step).  The static variant keeps the partial sums in an array that is part of the
library image and takes plain device pointers: the caller brings the input and the 8-byte result slot.

This file is named test_zz_* so that pytest collects it LAST: `pytest -x` stops at the first failure and
a failure here must not hide the parity tests.  It has to pass both on its own in a fresh process
(where its first call is also the first CUDA work of the process on that GPU) and at the end of the whole suite.
"""
import ctypes as C

import numpy as np
import pytest

EINVAL = -1   # NEPTUNE_RT_EINVAL (include/neptune_rt.h)


def _i64(v):
    return (C.c_int64 * len(v))(*[int(x) for x in v])


def _static_sum(nb, ctx, field, lb, ub, result_field, slot):
    """One call through the C ABI: sum `field`[lb:ub) into element `slot` of `result_field` (both device memory the
    TEST allocated, standing in for the caller's pool), enqueued on the ctx's stream, nothing awaited."""
    L = nb.lib()
    rc = L.neptune_rt_reduce_sum_static(
        C.c_void_p(field.device_ptr()), field.dtype, len(field.shape), _i64(field.shape),
        None if lb is None else _i64(lb), None if ub is None else _i64(ub),
        C.c_void_p(result_field.device_ptr() + 8 * slot), C.c_void_p(ctx.stream()))
    assert rc == 0, L.neptune_rt_last_error(None).decode()


@pytest.mark.gpu
@pytest.mark.parametrize("ty", ["f64", "f32"])
def test_static_scratch_sum_matches_the_allocating_entry_point_twice_in_a_row(nb, ctx, ty):
    """Two calls back to back on one stream, no synchronisation between them, against
    neptune_rt_field_reduce_sum on the same boxes.  The second box has 6 rows, far fewer than the 592 CTAs,
    so almost every scratch slot still holds a partial sum of the FIRST call when the second starts: the
    result is right only if each call overwrites the whole array before reading it.  A third call sums an
    empty box into a slot that held a non-zero value.  Tolerance: the bound test_reduce_sum_tolerance_tier
    uses, n * eps(elem) * sum|x|; both entry points share partition and trees, so the values are in fact
    equal bit for bit, and that is asserted too."""
    dt = np.float64 if ty == "f64" else np.float32
    shape = (40, 33, 130)
    u = np.random.default_rng(77).uniform(-1, 1, size=shape).astype(dt)
    boxes = [(None, None), ([3, 5, 7], [5, 8, 120]), ([4, 4, 4], [4, 9, 9])]
    f = nb.Field.from_numpy(ctx, u)
    res = nb.Field.from_numpy(ctx, np.full((4,), 123.0))          # the caller's result slots, pre-filled with junk
    try:
        want = [f.reduce_sum(lb, ub) for lb, ub in boxes]         # existing entry point (allocates on first use)
        for rep in range(2):                                       # and the whole sequence twice
            for slot, (lb, ub) in enumerate(boxes):
                _static_sum(nb, ctx, f, lb, ub, res, slot)
            got = res.download()
            for slot, (lb, ub) in enumerate(boxes):
                sub = u if lb is None else u[tuple(slice(a, b) for a, b in zip(lb, ub))]
                bound = sub.size * float(np.finfo(dt).eps) * float(np.sum(np.abs(sub), dtype=np.float64))
                assert abs(got[slot] - want[slot]) <= bound, (rep, slot, got[slot], want[slot], bound)
                assert got[slot] == want[slot], (rep, slot, got[slot], want[slot])
            assert got[2] == 0.0 and got[3] == 123.0               # empty box; the slot nobody wrote is untouched
            assert abs(got[0] - float(np.sum(u.astype(np.longdouble)))) <= 4 * float(np.finfo(dt).eps) * max(1.0, abs(got[0])) + 1e-9
            res.upload(np.full((4,), 123.0))
    finally:
        res.free()
        f.free()


@pytest.mark.gpu
def test_static_scratch_sum_leaves_device_memory_alone_and_checks_its_pointers(nb, ctx):
    """After the first call on the device (which may make CUDA load the module), further calls change the
    amount of free device memory by nothing; host pointers are refused instead of being handed to a kernel."""
    import torch
    torch.cuda.init()                                   # torch's own start-up allocations happen before the measurement
    u = np.ones((64, 64, 64))
    f = nb.Field.from_numpy(ctx, u)
    res = nb.Field.from_numpy(ctx, np.zeros((1,)))
    try:
        _static_sum(nb, ctx, f, None, None, res, 0)
        ctx.sync()
        free0 = torch.cuda.mem_get_info(0)[0]
        for _ in range(50):
            _static_sum(nb, ctx, f, [1, 1, 1], [63, 63, 63], res, 0)
        ctx.sync()
        assert torch.cuda.mem_get_info(0)[0] == free0
        assert res.download()[0] == 62.0 ** 3
        L = nb.lib()
        host = np.zeros(1)
        rc = L.neptune_rt_reduce_sum_static(C.c_void_p(f.device_ptr()), f.dtype, 3, _i64(u.shape), None, None,
                                            C.c_void_p(host.ctypes.data), C.c_void_p(ctx.stream()))
        assert rc == EINVAL and b"current device" in L.neptune_rt_last_error(None)
    finally:
        res.free()
        f.free()


def test_static_scratch_sum_refuses_what_it_cannot_index(nb):
    """Argument checks run before any CUDA call, so they are testable without a GPU: an axis longer than
    NEPTUNE_RT_REDUCE_STATIC_MAX_EXTENT, a sub-box outside the array, a bad rank or dtype each return
    NEPTUNE_RT_EINVAL with a message that says what the limit is.  The pointers are never dereferenced."""
    L = nb.lib()
    fake = C.c_void_p(0x1000)
    limit = 1 << 30

    def call(dtype, rank, shape, lb=None, ub=None):
        return L.neptune_rt_reduce_sum_static(fake, dtype, rank, _i64(shape), None if lb is None else _i64(lb),
                                              None if ub is None else _i64(ub), fake, None)

    assert call(nb.F64, 1, [limit + 1]) == EINVAL
    msg = L.neptune_rt_last_error(None).decode()
    assert "shape[0]" in msg and str(limit) in msg and "NEPTUNE_RT_REDUCE_STATIC_MAX_EXTENT" in msg
    assert call(nb.F64, 3, [4, limit + 7, 4]) == EINVAL
    assert "shape[1]" in L.neptune_rt_last_error(None).decode()
    assert call(nb.F64, 2, [8, 8], [0, 0], [8, 9]) == EINVAL
    assert call(nb.F64, 4, [2, 2, 2, 2]) == EINVAL
    assert call(7, 1, [8]) == EINVAL
    assert L.neptune_rt_reduce_sum_static(None, nb.F64, 1, _i64([8]), None, None, fake, None) == EINVAL
