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


def _worker(rank, world, port, name, shape, steps, q, halo=None, tblock=0):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import neptune_pde_solver_b200 as nb
        from neptune_pde_solver_b200.slab import Slab
        gtext, n_extra = H.program(name, list(shape))
        u, extra = H.inputs(shape, n_extra, "f64", seed=99)
        g = nb.Module(gtext).apply_desc(0)
        s = Slab(g, world, rank, halo)
        ctx = nb.Context(rank)
        uid = [nb.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], world, rank)
        st = s.stencil(ctx)
        st.set("tblock", tblock)
        lb, ub = s.local.out_lb, s.local.out_ub
        fa, fb = nb.Field(ctx, np.float64, lb, ub), nb.Field(ctx, np.float64, lb, ub)
        loc = s.scatter(u)
        fa.upload(loc)
        fb.upload(loc)
        fx = []
        for e in extra:
            f = nb.Field(ctx, np.float64, lb, ub)
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
