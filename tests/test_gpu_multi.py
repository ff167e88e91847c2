"""Multi-GPU parity (needs >= 2 B200s; skipped otherwise): the dim-0 slab decomposition with NCCL
halo exchange overlapped with the interior sweep (neptune_rt_run_steps) must reproduce the
single-domain oracle result BITWISE.  One process per GPU, NCCL id broadcast over a gloo group."""
import os
import socket

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, shape, steps, q, halo=None, tblock=0, ty="f64"):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import neptune_pde_solver_b200 as nb
        from neptune_pde_solver_b200.slab import Slab
        gtext, n_extra = H.program(name, list(shape), ty)
        u, extra = H.inputs(shape, n_extra, ty, seed=99)
        dt = H.NPT[ty]
        g = nb.Module(gtext).apply_desc(0)
        s = Slab(g, world, rank, halo)
        ctx = nb.Context(rank)
        uid = [nb.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], world, rank)
        st = s.stencil(ctx)
        st.set("tblock", tblock)
        lb, ub = s.local.out_lb, s.local.out_ub
        fa, fb = nb.Field(ctx, dt, lb, ub), nb.Field(ctx, dt, lb, ub)
        loc = s.scatter(u)
        fa.upload(loc)
        fb.upload(loc)
        fx = []
        for e in extra:
            f = nb.Field(ctx, dt, lb, ub)
            f.upload(s.scatter(e))
            fx.append(f)
        res = st.run_steps(fa, fb, fx, steps)
        got = s.owned(res.download()).copy()
        used = st.variant()
        parts = [None] * world
        dist.all_gather_object(parts, (s.begin, got))
        if rank == 0:
            full = np.concatenate([p[1] for p in sorted(parts, key=lambda p: p[0])], axis=0)
            want = H.oracle_steps(gtext, u, extra, steps)
            q.put((bool(H.bits_equal(full, want)), used))
        for f in [fa, fb] + fx:
            f.free()
        ctx.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name,shape,steps,halo,tblock", [
    ("heat_star", (64, 48, 64), 9, None, 0),
    ("heat_box_2d9", (96, 128), 7, None, 0),
    ("varcoef_3d7", (32, 16, 32), 4, None, 0),
    ("jacobi_star", (40, 32, 64), 6, None, 0),
    ("heat_star", (64, 48, 64), 9, 2, 0),      # two ghost planes: two steps per launch and per exchange, + one odd step
    ("heat_star", (96, 40, 136), 8, 2, 2),
    ("heat_star", (64, 48, 64), 5, 2, 1),      # two ghost planes, temporal blocking switched off
    ("weighted_3d27", (48, 24, 32), 6, 2, 0),
    ("heat_star", (48, 20, 256), 6, 2, 0),     # wide box: the z-marching kernel (tb2m) in slab mode
    ("heat_box_2d9", (96, 1040), 8, 2, 0),     # rank 2, two ghost rows: sweep_rows2 in slab mode
    ("jacobi_star", (64, 40, 72), 8, 2, 0),    # fused pairs with a pointwise right-hand side: f's ghost planes are exchanged once
    ("varcoef_3d27", (32, 16, 32), 3, None, 0),   # BASELINE configs[3]'s body: 27 coefficient fields, NCCL halos
    ("varcoef_3d27", (48, 16, 32), 4, 2, 0),      # ... computing ghost planes too (coefficient ghosts exchanged once)
    ("heat_star", (64, 48, 64), 10, 4, 0),        # deep halos: four ghost planes, one exchange per TWO fused pairs
    ("heat_star", (70, 48, 64), 9, 4, 0),         # ... uneven slabs (70 planes over the ranks) and an odd step count
    ("heat_star", (96, 40, 264), 12, 8, 0),       # eight ghost planes: one exchange per four fused pairs, wide box (tb2p / tb2q interior)
    ("jacobi_star", (64, 40, 72), 8, 4, 0),       # deep halos with a pointwise right-hand side
    ("heat_star", (64, 48, 64), 7, 3, 1),         # single steps with three ghost planes: one exchange per three steps
    ("heat_box_2d9", (96, 1040), 8, 4, 0),        # rank 2, four ghost rows
    ("heat_star", (24, 24, 64), 6, 4, 0),         # slabs thinner than two halos on four ranks: unsplit sweep, serial exchange (agreed by all ranks)
    ("varcoef_3d7", (48, 16, 32), 5, 2, 0),       # single-step kernel computing one ghost plane per launch
])
def test_slab_nccl_equals_single_domain(name, shape, steps, halo, tblock):
    import torch
    import torch.multiprocessing as mp
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs (ranks must not share a GPU: NCCL peers would wait on each other)")
    world = min(ngpu, 4)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, shape, steps, q, halo, tblock)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    ok, used = q.get(timeout=5)
    assert ok is True
    if halo == 2 and tblock != 1 and name in ("heat_star", "jacobi_star", "heat_box_2d9") and steps % 2 == 0:
        assert used.startswith("tb2"), used   # the temporally blocked kernel really ran in slab mode


@pytest.mark.parametrize("ty", ["f64", "f32"])
def test_jacobi_1000_iterations_slabs_bit_exact(ty):
    """BASELINE configs[4] at reduced size: 1000 Jacobi iterations with a right-hand side on 96^3 over the slabs of up
    to four GPUs, four ghost planes (two fused pairs per exchange), every iteration's rounding reproduced: the result
    equals the oracle's 1000-step loop bitwise."""
    import torch
    import torch.multiprocessing as mp
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(ngpu, 4)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, "jacobi_star", (96, 96, 96), 1000, q, 4, 0, ty)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
        assert p.exitcode == 0
    ok, used = q.get(timeout=5)
    assert ok is True and used.startswith("tb2"), used


def _worker_mode(rank, world, port, q):
    """Reports how the slab loop exchanged and its timeline (neptune_rt_stencil_halo_mode, neptune_rt_slab_trace)."""
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import neptune_pde_solver_b200 as nb
        from neptune_pde_solver_b200.slab import Slab
        shape = (64, 32, 256)
        gtext, _ = H.program("heat_star", list(shape))
        u, _ = H.inputs(shape, 0, "f64", seed=5)
        s = Slab(nb.Module(gtext).apply_desc(0), world, rank, 4)
        ctx = nb.Context(rank)
        uid = [nb.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], world, rank)
        st = s.stencil(ctx)
        lb, ub = s.local.out_lb, s.local.out_ub
        fa, fb = nb.Field(ctx, np.float64, lb, ub), nb.Field(ctx, np.float64, lb, ub)
        loc = s.scatter(u)
        fa.upload(loc)
        fb.upload(loc)
        assert st.halo_mode() == ""
        ctx.slab_trace(True)
        st.run_steps(fa, fb, [], 8)
        rows = ctx.slab_trace_read()
        ctx.slab_trace(False)
        plan = nb.runtime.slab_schedule(8, 1, 4, True)
        if rank == 0:
            q.put((st.halo_mode(), [r[0] for r in rows], [10 * p[0] + int(p[2]) for p in plan],
                   all(r[1] <= r[2] <= r[4] + 1e-3 and r[1] <= r[3] for r in rows)))
        fa.free(); fb.free()
        ctx.close()
    finally:
        dist.destroy_process_group()


def test_slab_loop_reports_its_exchange_mode_and_timeline():
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_mode, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    mode, kinds, planned, ordered = q.get(timeout=5)
    assert mode == ("nccl" if os.environ.get("NEPTUNE_RT_HALO") == "nccl" else "p2p")   # GPUs of one node: copy engines + flags
    assert kinds == planned          # the loop followed neptune_rt_slab_schedule: [20, 21, 20, 20] for 8 steps, 4 ghost planes
    assert ordered


def _worker_realloc(rank, world, port, q, cycles):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import neptune_pde_solver_b200 as nb
        from neptune_pde_solver_b200.slab import Slab
        shape, steps, halo = (64, 48, 64), 10, 4
        gtext, _ = H.program("heat_star", list(shape), "f64")
        g = nb.Module(gtext).apply_desc(0)
        s = Slab(g, world, rank, halo)
        ctx = nb.Context(rank)
        uid = [nb.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], world, rank)
        st = s.stencil(ctx)
        lb, ub = s.local.out_lb, s.local.out_ub
        free0 = 0
        ok, modes, ptrs = True, set(), set()
        for cyc in range(cycles):
            u, _ = H.inputs(shape, 0, "f64", seed=500 + cyc)
            fa, fb = nb.Field(ctx, np.float64, lb, ub), nb.Field(ctx, np.float64, lb, ub)
            ptrs.update((fa.device_ptr(), fb.device_ptr()))
            loc = s.scatter(u)
            fa.upload(loc)
            fb.upload(loc)
            if cyc % 5 == 4:
                st.set_slab(halo, g.lb[0], g.ub[0])       # same values: the agreement is dropped and made again
            res = st.run_steps(fa, fb, [], steps)
            modes.add(st.halo_mode())
            got = s.owned(res.download()).copy()
            parts = [None] * world
            dist.all_gather_object(parts, (s.begin, got))
            if rank == 0:
                full = np.concatenate([p[1] for p in sorted(parts, key=lambda p: p[0])], axis=0)
                ok = ok and bool(H.bits_equal(full, H.oracle_steps(gtext, u, [], steps)))
            # freed while the neighbours may still map them; rank 1 lags behind on purpose, so that its neighbours
            # free (and re-allocate) before it has unmapped their fields
            if rank == 1:
                import time
                time.sleep(0.02)
            fb.free()
            fa.free()
            if cyc == 1:
                # baseline after two whole cycles: by now NCCL's channel buffers and the CUDA modules are resident
                dist.barrier()
                ctx.sync()
                free0 = torch.cuda.mem_get_info(rank)[0]
        dist.barrier()
        ctx.sync()
        leaked = free0 - torch.cuda.mem_get_info(rank)[0]
        if rank == 0:
            q.put((ok, sorted(modes), len(ptrs), leaked))
        ctx.close()
    finally:
        dist.destroy_process_group()


def test_fields_freed_and_reallocated_between_slab_runs():
    """20 cycles of allocate -> run the slab loop (p2p: the neighbours IPC-map the fields) -> free, with one rank
    lagging: every cycle bit-equal to the single-domain oracle, although device addresses are reused by later
    fields (the agreement cache is keyed on field ids, not addresses) and the agreement is dropped by set_slab
    every fifth cycle.  Exported allocations are released only once the neighbours have unmapped them
    (ipc_release_field in runtime.cu), and in the end nothing stays allocated."""
    import torch
    import torch.multiprocessing as mp
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(ngpu, 4)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_realloc, args=(r, world, port, q, 20)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    ok, modes, distinct, leaked = q.get(timeout=5)
    assert ok is True
    assert modes in (["p2p"], ["nccl"])
    assert 2 <= distinct <= 40                 # (how many distinct addresses the 40 fields of rank 0 had: usually a handful)
    assert leaked <= 32 << 20, leaked          # 18 further cycles left nothing behind: deferred frees were released
