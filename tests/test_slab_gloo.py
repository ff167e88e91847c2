"""world_size-2/3 CPU test (gloo) of the N>1 path's host logic: slab partition, localized
descriptors, ghost-plane exchange pattern and gather reproduce the single-domain result bit for
bit.  The sweep itself is stood in for by the oracle here (no GPU in this container); on the GPU
box the same Slab geometry drives neptune_rt_run_steps + NCCL (tests/test_gpu_multi.py)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as H


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, shape, steps, q, halo=None):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import neptune_pde_solver_b200 as nb
        from neptune_pde_solver_b200.slab import Slab
        from oracle import mlir_subset as ms, numpy_lowering as onp
        gtext, n_extra = H.program(name, list(shape))
        u, extra = H.inputs(shape, n_extra, "f64", seed=99)
        g = nb.Module(gtext).apply_desc(0)
        s = Slab(g, world, rank, halo)
        ltext, _ = H.program(name, list(s.local_shape(shape)), lb=[s.local.out_lb[0]] + [0] * (len(shape) - 1))
        l = nb.Module(ltext).apply_desc(0)
        if halo is None:
            assert (l.lb, l.ub, l.out_lb, l.out_ub) == (s.local.lb, s.local.ub, s.local.out_lb, s.local.out_ub)
        else:
            # wide halo (temporal blocking across the exchange): the local program's bounds are the
            # interior of the local box, i.e. they reach one plane into the ghosts, so `halo` sweeps
            # between two exchanges leave every owned plane exact (neptune_rt_run_steps, slab mode)
            assert (l.out_lb, l.out_ub) == (s.local.out_lb, s.local.out_ub)
            assert l.lb[0] <= s.local.lb[0] and l.ub[0] >= s.local.ub[0]
        mod = ms.parse(ltext)
        a = s.scatter(u)
        lex = [s.scatter(e) for e in extra]
        b = np.empty_like(a)
        per_exchange = 1 if halo is None else halo
        done = 0
        while done < steps:
            for _ in range(min(per_exchange, steps - done) - 1):
                onp.run_function(mod, "@entry", [b, a] + lex)
                a, b = b, a
                done += 1
            onp.run_function(mod, "@entry", [b, a] + lex)
            done += 1
            # ghost planes of the new state: owned boundary planes go to the neighbours
            reqs, bufs = [], []
            for peer, (s0, s1), (r0, r1) in s.send_recv_rows():
                snd = torch.from_numpy(np.ascontiguousarray(b[s0:s1]))
                rcv = torch.empty_like(snd)
                reqs += [dist.isend(snd, peer), dist.irecv(rcv, peer)]
                bufs.append((rcv, r0, r1))
            for r in reqs:
                r.wait()
            for rcv, r0, r1 in bufs:
                b[r0:r1] = rcv.numpy()
            a, b = b, a
        parts = [None] * world
        dist.all_gather_object(parts, (s.begin, s.end, s.owned(a).copy()))
        if rank == 0:
            full = np.concatenate([p[2] for p in sorted(parts, key=lambda p: p[0])], axis=0)
            want = H.oracle_steps(gtext, u, extra, steps)
            q.put(bool(H.bits_equal(full, want)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,name,shape,steps,halo", [
    (2, "heat_star", (16, 12, 10), 6, None),
    (3, "heat_box_2d9", (31, 24), 5, None),
    (2, "varcoef_3d7", (10, 8, 8), 3, None),
    (2, "heat_star", (16, 12, 10), 7, 2),     # two sweeps per exchange of two ghost planes (+ one odd step)
    (3, "heat_star", (27, 8, 10), 6, 2),
])
def test_slab_time_loop_equals_single_domain(world, name, shape, steps, halo):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, shape, steps, q, halo)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def _worker_plan(rank, world, port, name, shape, steps, q, halo, fuse2):
    """The launch plan neptune_rt_run_steps follows in slab mode (neptune_rt_slab_schedule: deep halos, fused pairs,
    ghost planes computed redundantly, boundary exchange only when the next launch would run short), executed on the CPU
    with the oracle standing in for the sweep kernels and gloo for the halo exchange."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import neptune_pde_solver_b200 as nb
        from neptune_pde_solver_b200.slab import Slab
        from oracle import mlir_subset as ms, numpy_lowering as onp
        gtext, n_extra = H.program(name, list(shape))
        u, extra = H.inputs(shape, n_extra, "f64", seed=77)
        g = nb.Module(gtext).apply_desc(0)
        r = g.radius0()
        s = Slab(g, world, rank, halo)
        llb = [s.local.out_lb[0]] + [0] * (len(shape) - 1)
        if name == "jacobi_star":     # the builder derives h^2 from the extents: keep the GLOBAL grid's constant
            ltext = H.mb.jacobi_star(list(s.local_shape(shape)), 1.0 / (max(shape) ** 2), "f64", llb)
        else:
            ltext, _ = H.program(name, list(s.local_shape(shape)), lb=llb)
        mod = ms.parse(ltext)
        ap = [op.apply for op in mod.functions["@entry"].ops if op.name == "apply"][0]
        rest_lb, rest_ub = tuple(g.lb[1:]), tuple(g.ub[1:])
        a = s.scatter(u)                      # ghost planes valid: the initial exchange
        lex = [s.scatter(e) for e in extra]   # coefficient / rhs ghosts valid too
        b = np.empty_like(a)
        plan = nb.runtime.slab_schedule(steps, r, s.halo, fuse2)
        assert sum(p[0] for p in plan) == steps
        valid = s.halo
        for (nst, ext, exchange) in plan:
            assert valid >= nst * r          # a launch never reads ghost planes that are out of date
            for k in range(1, nst + 1):
                # level k of the launch: the output range grown by the levels still to come (what the fused kernel
                # evaluates on its ring), in GLOBAL logical rows, clipped to the global apply bounds
                grow = ext + (nst - k) * r
                lo = max(g.lb[0], s.begin - (grow if s.up else 0))
                hi = min(g.ub[0], s.end + (grow if s.down else 0))
                ap.bounds = ms.Bounds((lo,) + rest_lb, (max(lo, hi),) + rest_ub)
                onp.run_function(mod, "@entry", [b, a] + lex)
                a, b = b, a
            valid -= nst * r
            if exchange:
                reqs, bufs = [], []
                for peer, (s0, s1), (r0, r1) in s.send_recv_rows():
                    snd = torch.from_numpy(np.ascontiguousarray(a[s0:s1]))
                    rcv = torch.empty_like(snd)
                    reqs += [dist.isend(snd, peer), dist.irecv(rcv, peer)]
                    bufs.append((rcv, r0, r1))
                for rq in reqs:
                    rq.wait()
                for rcv, r0, r1 in bufs:
                    a[r0:r1] = rcv.numpy()
                valid = s.halo
        parts = [None] * world
        dist.all_gather_object(parts, (s.begin, s.end, s.owned(a).copy()))
        if rank == 0:
            full = np.concatenate([p[2] for p in sorted(parts, key=lambda p: p[0])], axis=0)
            want = H.oracle_steps(gtext, u, extra, steps)
            n_ex = sum(1 for p in plan if p[2])
            q.put((bool(H.bits_equal(full, want)), n_ex))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,name,shape,steps,halo,fuse2,exchanges", [
    (2, "heat_star", (24, 10, 12), 9, 4, True, 2),      # four ghost planes: one exchange per two fused pairs, + one odd step
    (3, "heat_star", (30, 8, 10), 12, 8, True, 1),      # eight ghost planes: one exchange per four fused pairs
    (2, "heat_star", (20, 8, 10), 7, 3, False, 2),      # single steps, three ghost planes: one exchange per three steps
    (2, "jacobi_star", (24, 8, 12), 8, 4, True, 1),     # a pointwise right-hand side is read on computed ghost planes
    (2, "varcoef_3d7", (20, 8, 8), 5, 2, False, 2),     # coefficient fields, single steps over ghost planes
    (3, "heat_box_2d9", (36, 24), 8, 2, True, 3),       # the classic scheme: one exchange per fused pair
])
def test_slab_launch_plan_equals_single_domain(world, name, shape, steps, halo, fuse2, exchanges):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_plan, args=(r, world, port, name, shape, steps, q, halo, fuse2)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    ok, n_ex = q.get(timeout=5)
    assert ok is True
    assert n_ex == exchanges


def test_slab_schedule_invariants():
    import neptune_pde_solver_b200 as nb
    for nsteps in range(0, 40):
        for radius in (1, 2):
            for ghost_mult in (1, 2, 3, 4, 8):
                for fuse2 in (False, True):
                    ghost = ghost_mult * radius
                    if fuse2 and ghost < 2 * radius:
                        with pytest.raises(nb.NeptuneError):
                            nb.runtime.slab_schedule(nsteps, radius, ghost, fuse2)
                        continue
                    plan = nb.runtime.slab_schedule(nsteps, radius, ghost, fuse2)
                    assert sum(p[0] for p in plan) == nsteps
                    valid = ghost
                    for i, (nst, ext, exchange) in enumerate(plan):
                        assert nst in (1, 2) and (nst == 1 or fuse2)
                        assert valid >= nst * radius                     # enough valid ghost planes going in
                        assert ext == (0 if exchange or i == len(plan) - 1 else valid - nst * radius)
                        valid = ghost if exchange else valid - nst * radius
                    if plan:
                        assert not plan[-1][2]                           # nothing to exchange after the last launch
