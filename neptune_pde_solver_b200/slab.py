"""Dim-0 slab decomposition of an apply over the ranks of one node (one process per GPU).

The reference has no domain decomposition at all (no parallel construct, PETSc on
COMM_SELF); this is the multi-GPU form of the same sweep: rank g owns a contiguous range of dim-0
planes of the result box plus `halo` ghost planes on each side that has a neighbour.  A dim-0
plane is one contiguous block of the row-major memref (DataflowLowering.cpp:41-71), so ghost
planes travel straight out of / into field memory — NCCL send/recv on the device
(neptune_rt_halo_exchange), no pack kernels.  All index algebra is done by the C library
(neptune_rt_slab_partition / neptune_rt_slab_localize); this class only turns it into slices.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np

from . import runtime as rt


class Slab:
    def __init__(self, desc: rt.StencilDesc, nranks: int, rank: int, halo: int | None = None):
        self.g = desc
        self.nranks, self.rank = nranks, rank
        self.halo = desc.radius0() if halo is None else halo
        n0 = desc.out_ub[0] - desc.out_lb[0]
        self.begin, self.end = rt.slab_partition(n0, nranks, rank)       # owned planes, global storage rows
        self.local = rt.slab_localize(desc, nranks, rank, self.halo)     # local descriptor
        self.lo = self.local.out_lb[0] - desc.out_lb[0]                  # first stored plane (global row)
        self.hi = self.local.out_ub[0] - desc.out_lb[0]
        self.up = rank > 0
        self.down = rank + 1 < nranks

    def stencil(self, ctx: "rt.Context") -> "rt.Stencil":
        """The local stencil, told where its box sits in the global domain (ghost width and the
        global dim-0 apply bounds) so that run_steps may fuse two steps per launch."""
        st = rt.Stencil(ctx, self.local)
        st.set_slab(self.halo, self.g.lb[0], self.g.ub[0])
        return st

    # rows of the LOCAL storage
    @property
    def own_rows(self) -> Tuple[int, int]:
        return self.begin - self.lo, self.end - self.lo

    def local_shape(self, global_shape) -> Tuple[int, ...]:
        return (self.hi - self.lo,) + tuple(global_shape[1:])

    def scatter(self, global_array: np.ndarray) -> np.ndarray:
        """Local box (owned planes + ghosts) of a global host array (always a copy)."""
        return np.array(global_array[self.lo:self.hi], order="C", copy=True)

    def owned(self, local_array: np.ndarray) -> np.ndarray:
        b, e = self.own_rows
        return local_array[b:e]

    def send_recv_rows(self):
        """[(peer, send_rows, recv_rows)] in local storage rows, for a width-`halo` exchange."""
        b, e = self.own_rows
        h = self.halo
        out = []
        if self.up:
            out.append((self.rank - 1, (b, b + h), (b - h, b)))
        if self.down:
            out.append((self.rank + 1, (e - h, e), (e, e + h)))
        return out
