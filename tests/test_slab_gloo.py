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
