#!/usr/bin/env python
"""bench.py — throughput of the `neptune_ir.apply` sweep on B200 (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c3|c2|c1|c5]

One "step" is one sweep (one time step) of the workload's apply over its grid:
  c3 (default, the config the metric is quoted on): 3-D 7-point fp64 heat step, 512^3 PER GPU,
     dim-0 slab decomposition with NCCL halo exchange for N > 1 (weak scaling).
`value` is whole-job G point-updates/s with the fields resident in HBM; `e2e` is the same metric
through the reference-facing entry call `@entry(out, in)` on HOST memrefs (pinned), upload and
download inside the timed region every step.  `roofline` divides the algorithmic bytes of one
launch (16 B per point-update: one fp64 read + one fp64 write) by the CUDA-event
duration of that launch; `cpu_baseline` is the oracle's gcc-compiled restatement of the reference
lowering timed on this box's host cores (1 thread: the reference emits a single-threaded scf.for
nest).  `--impl reference` runs only that CPU arm.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "Gpoint-updates/s, 3D 7-pt fp64 heat step, 512^3 per GPU"
UNIT = "Gpoint-updates/s"

WORKLOADS = {
    # name: (builder, shape per GPU, dtype, alpha, bytes per update, description)
    "c3": ("heat_star", [512, 512, 512], "f64", 0.1, 16, "3D 7-pt fp64 heat step, 512^3 per GPU (BASELINE configs[2])"),
    "c2": ("heat_box_2d9", [16384, 16384], "f64", 0.1, 16, "2D 9-pt fp64 heat step, 16384^2 (BASELINE configs[1])"),
    "c1": ("heat_star", [4096, 4096], "f64", 0.2, 16, "2D 5-pt fp64 heat step, 4096^2 (BASELINE configs[0])"),
    "c5": ("jacobi_star", [768, 768, 768], "f64", None, 24, "3D 7-pt Jacobi with rhs, 768^3 f64 (BASELINE configs[4], 1 GPU)"),
    "c5f32": ("jacobi_star", [768, 768, 768], "f32", None, 12, "3D 7-pt Jacobi with rhs, 768^3 f32 (BASELINE configs[4], 1 GPU)"),
    "c5g": ("jacobi_star", [96, 768, 768], "f64", None, 24, "3D 7-pt Jacobi with rhs, fp64, 96x768x768 per GPU = 768^3 over 8 GPUs (BASELINE configs[4] as worded)"),
    "c5gf32": ("jacobi_star", [96, 768, 768], "f32", None, 12, "3D 7-pt Jacobi with rhs, fp32, 96x768x768 per GPU = 768^3 over 8 GPUs (BASELINE configs[4] as worded)"),
    "c4": ("varcoef27", [128, 1024, 1024], "f64", None, 232, "3D 27-pt variable-coefficient fp64 (27 coefficient fields), 128x1024x1024 per GPU = 1024^3 over 8 GPUs (BASELINE configs[3])"),
    "tiny": ("heat_star", [64, 64, 128], "f64", 0.1, 16, "3D 7-pt fp64 heat step, 64x64x128 (self-test)"),
}


def program_text(name, shape):
    from neptune_pde_solver_b200 import mlir_builders as mb
    kind, _, ty, alpha, _, _ = WORKLOADS[name]
    if kind == "heat_star":
        return mb.heat_star(shape, alpha, ty), 0
    if kind == "heat_box_2d9":
        return mb.heat_box_2d9(shape, alpha, ty), 0
    if kind == "jacobi_star":
        return mb.jacobi_star(shape, 1.0 / (max(shape) ** 2), ty), 1
    if kind == "varcoef27":
        return mb.varcoef(shape, mb.box_offsets(3), ty), 27
    raise KeyError(kind)


def program_for(kind, shape, ty, alpha, h2, lb=None):
    """The workload's NeptuneIR text at another extent with the SAME constants (band cones, local slab programs)."""
    from neptune_pde_solver_b200 import mlir_builders as mb
    if kind == "heat_star":
        return mb.heat_star(shape, alpha, ty, lb)
    if kind == "heat_box_2d9":
        return mb.heat_box_2d9(shape, alpha, ty, lb)
    if kind == "jacobi_star":
        return mb.jacobi_star(shape, h2, ty, lb)
    if kind == "varcoef27":
        return mb.varcoef(shape, mb.box_offsets(3), ty, lb)
    raise KeyError(kind)


PARITY_STEPS = 4        # N = 1; in slab mode parity_steps_for() raises it until the plan holds exchanging launches


def parity_steps_for(world, radius, ghost, fuse2):
    """Time steps of the in-run parity check. One GPU: 4. Slabs: the smallest even count whose launch plan
    (neptune_rt_slab_schedule — the function the time loop itself follows) contains at least TWO exchanging launches
    and at least one non-exchanging launch after the last exchange, so that the split boundary/interior launches, the
    halo push, the announce/join flags and a launch that consumes exchanged ghosts are all inside what is bit-compared."""
    if world == 1:
        return 4, None
    from neptune_pde_solver_b200.runtime import slab_schedule
    plan = None
    for S in range(4, 49, 2):
        plan = slab_schedule(S, radius, ghost, fuse2)
        ex = [i for i, (_, _, x) in enumerate(plan) if x]
        if len(ex) >= 2 and ex[-1] < len(plan) - 1:
            return S, plan
    return 48, plan


class BandRows:
    """The few dim-0 rows of a big host array that the band parity check reads (so that 27 coefficient fields of 1 GB
    each need not stay in host memory): rows [r0, r1) are kept for every range asked for at construction."""

    def __init__(self, a, ranges):
        self.shape, self.ndim, self.dtype = a.shape, a.ndim, a.dtype
        self.parts = {(r0, r1): np.array(a[r0:r1], copy=True) for (r0, r1) in ranges}

    def __getitem__(self, key):
        rows, rest = key[0], key[1:]
        return self.parts[(rows.start, rows.stop)][(slice(None),) + tuple(rest)]


def band_row_ranges(world, own_rows):
    ob, oe = own_rows
    S = PARITY_STEPS
    if world == 1:
        mid = (ob + oe) // 2
        return [(mid - 2 * S, mid + 2 * S)]
    return [(ob, ob + 2 * S), (oe - 2 * S, oe)]


def band_parity_check(kind, ty, alpha, h2, world, rank, own_rows, u_local, extra_local, run_fn, gather_fn):
    """Bitwise parity inside the benchmark run, on the ACTUAL configuration (kernel variant, slab decomposition,
    ghost depth, halo exchange): S = PARITY_STEPS steps from the benchmark's initial state (4 on one GPU; in slab mode
    as many as the launch plan needs to hold two exchanges, see parity_steps_for), then a 2S-plane band that straddles
    every slab boundary (N = 1: the middle of the domain) is compared with the oracle run on the band's dependency
    cone (4S planes x a corner block that includes two real faces; the cone's own cut faces cannot reach the band
    in S steps).  The oracle is the checker here, never the thing measured."""
    rank_nd = u_local.ndim
    cut = (slice(0, min(96, u_local.shape[1])), slice(0, min(160, u_local.shape[2]))) if rank_nd == 3 else (slice(0, min(4096, u_local.shape[1])),)
    ob, oe = own_rows
    S = PARITY_STEPS
    res_local = run_fn(S)                       # numpy array of the local box after S steps
    if world == 1:
        mid = (ob + oe) // 2
        packs = [{"init": [np.ascontiguousarray(a[(slice(mid - 2 * S, mid + 2 * S),) + cut]) for a in [u_local] + extra_local],
                  "res": np.ascontiguousarray(res_local[(slice(mid - S, mid + S),) + cut]), "where": "planes [%d, %d) of the single domain" % (mid - S, mid + S)}]
    else:
        mine = {"top_init": [np.ascontiguousarray(a[(slice(ob, ob + 2 * S),) + cut]) for a in [u_local] + extra_local],
                "bot_init": [np.ascontiguousarray(a[(slice(oe - 2 * S, oe),) + cut]) for a in [u_local] + extra_local],
                "top_res": np.ascontiguousarray(res_local[(slice(ob, ob + S),) + cut]),
                "bot_res": np.ascontiguousarray(res_local[(slice(oe - S, oe),) + cut])}
        parts = gather_fn(mine)
        if rank != 0:
            return None
        packs = []
        for r in range(world - 1):
            packs.append({"init": [np.concatenate([a, b], axis=0) for a, b in zip(parts[r]["bot_init"], parts[r + 1]["top_init"])],
                          "res": np.concatenate([parts[r]["bot_res"], parts[r + 1]["top_res"]], axis=0),
                          "where": "last %d planes of rank %d + first %d planes of rank %d" % (S, r, S, r + 1)})
    from oracle import mlir_subset as ms, numpy_lowering as onp
    ok, bands = True, []
    for pk in packs:
        cone = pk["init"][0]
        text = program_for(kind, list(cone.shape), ty, alpha, h2)
        want = onp.run_steps(ms.parse(text), cone, pk["init"][1:], S)
        inner = (slice(S, 3 * S),) + tuple(slice(0, n - S) for n in cone.shape[1:])
        trim = (slice(None),) + tuple(slice(0, n - S) for n in cone.shape[1:])
        a, b = pk["res"][trim], want[inner]
        view = np.uint64 if a.dtype == np.float64 else np.uint32
        eq = bool(a.shape == b.shape and np.array_equal(a.view(view), b.view(view)))
        ok &= eq
        bands.append({"band": pk["where"], "points": int(a.size), "bits_equal": eq})
    return {"bits_equal": ok, "steps": S, "bands": bands,
            "oracle": "oracle/numpy_lowering.py on each band's %d-plane dependency cone (restatement of DataflowLowering.cpp:258-448)" % (4 * S)}


def synth(shape, ty, seed):
    """u0 = product of sines + 0.25 U(-1,1), generated plane-block-wise."""
    dt = np.float64 if ty == "f64" else np.float32
    rng = np.random.default_rng(seed)
    u = np.empty(shape, dtype=dt)
    axes = [np.sin(2 * np.pi * np.arange(n) / n + 0.3) for n in shape]
    blk = max(1, (1 << 22) // int(np.prod(shape[1:])))
    for p in range(0, shape[0], blk):
        q = min(shape[0], p + blk)
        t = axes[0][p:q]
        for a in axes[1:]:
            t = np.multiply.outer(t, a)
        u[p:q] = (t + 0.25 * rng.uniform(-1, 1, size=t.shape)).astype(dt)
    return u


def kernel_digest():
    """Digest of the CUDA kernel sources: stamps ncu-derived numbers so that they go stale loudly."""
    import glob
    import hashlib
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "neptune_pde_solver_b200", "csrc", "kernels", "*.cuh"))):
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


def interior_points(shape):
    n = 1
    for s in shape:
        n *= (s - 2)
    return n


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        rows = [r for (t, r) in self.rows if (t0 is None or t >= t0) and (t1 is None or t <= t1 + 0.2)] or [r for _, r in self.rows]
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for k, nme in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle's gcc-compiled restatement of the reference lowering (kind "port")
# ------------------------------------------------------------------------------------------------
def cpu_reference_arm(workload, shape, budget_s, min_steps=2, max_steps=50, warmup=1, u=None):
    """Times `@entry(out, in)` of the faithful lowering (malloc + copy-through + scalar loop nest +
    store copy, 1 thread) on the same workload; returns (Gupd/s, ms/step, steps timed, cores)."""
    from oracle import c_lowering
    text, n_extra = program_text(workload, shape)
    ty = WORKLOADS[workload][2]
    cm = c_lowering.CompiledModule(text, tag="bench_" + workload)
    a = synth(shape, ty, 20260403) if u is None else np.array(u, copy=True)
    b = np.empty_like(a)
    extra = [synth(shape, ty, 20260500 + k) for k in range(n_extra)]
    for _ in range(max(0, warmup)):
        cm.call("@entry", [b, a] + extra)
        a, b = b, a
    t0 = time.perf_counter()
    cm.call("@entry", [b, a] + extra)
    a, b = b, a
    t_one = time.perf_counter() - t0
    n = int(max(min_steps, min(max_steps, budget_s / max(t_one, 1e-9))))
    times = [t_one]
    for _ in range(n - 1):
        t0 = time.perf_counter()
        cm.call("@entry", [b, a] + extra)
        a, b = b, a
        times.append(time.perf_counter() - t0)
    sec = float(np.mean(times))
    return interior_points(shape) / sec / 1e9, sec * 1e3, len(times), 1


def cpu_openmp_context(workload, shape, steps=3):
    """Context only (NOT the reference): ping-pong sweeps, OpenMP over every host core."""
    from oracle import c_lowering
    text, n_extra = program_text(workload, shape)
    ty = WORKLOADS[workload][2]
    sw = c_lowering.OpenMPSweep(text, tag="bench_omp_" + workload)
    a = synth(shape, ty, 20260403)
    b = a.copy()
    extra = [synth(shape, ty, 20260500 + k) for k in range(n_extra)]
    sw.sweep(b, [a] + extra)
    best = 1e30
    for _ in range(steps):
        t0 = time.perf_counter()
        sw.sweep(b, [a] + extra)
        best = min(best, time.perf_counter() - t0)
        a, b = b, a
    return interior_points(shape) / best / 1e9, len(os.sched_getaffinity(0))


def run_reference(args):
    """`--impl reference`: the reference's CPU lowering of the path (oracle port; the reference
    itself needs MLIR/LLVM 21 and cannot be built on this image), rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    shape = WORKLOADS[args.workload][1]
    # each "step" = one faithful sweep; K + W bounded so the run ends within a few minutes
    k = max(1, min(args.steps, 20))
    w = max(0, args.warmup)     # as asked (a 512^3 sweep on one core takes about 0.8 s, so even W = 10 stays within seconds)
    gups, ms, n, cores = cpu_reference_arm(args.workload, shape, budget_s=1e9, min_steps=k, max_steps=k, warmup=w)
    line = {
        "impl": "reference", "metric": METRIC, "value": gups, "unit": UNIT, "n_gpus": args.gpus,
        "steps": n, "warmup": w, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": WORKLOADS[args.workload][2], "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload][5], "grid_per_gpu": list(shape), "global_grid": list(shape),
                   "decomposition": "single domain on the host (the reference has no decomposition); at --gpus N > 1 this arm still "
                                    "sweeps ONE %s domain on one core, so the N-GPU ratio compares an N-GPU job with a 1-core job" % "x".join(map(str, shape)),
                   "points_per_step": interior_points(shape), "kernel_variant": "cpu scalar loop nest", "fma": False},
        "cpu_baseline": {"value": gups, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d full sweeps of the workload grid, gcc -O2 -ffp-contract=off scalar, "
                                   "incl. per-sweep malloc + copy-through + store copy "
                                   "(DataflowLowering.cpp:258-448,165-220); reference is single-threaded by construction" % n},
        "e2e": {"value": gups, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if args.omp_context:
        try:
            og, oc = cpu_openmp_context(args.workload, shape)
            line["cpu_context_openmp"] = {"value": og, "unit": UNIT, "cores": oc, "note": "NOT the reference: ping-pong, no copies, OpenMP"}
        except Exception as e:  # context number only
            line["cpu_context_openmp"] = {"error": str(e)[:200]}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def pinned_like(shape, dtype):
    """Pinned host buffer (torch is plumbing here: pinned memory + torch.distributed)."""
    import torch
    t = torch.empty(tuple(shape), dtype=torch.float64 if dtype == np.float64 else torch.float32, pin_memory=True)
    return t, t.numpy()


def run_b200(args):
    import torch
    import torch.distributed as dist
    import neptune_pde_solver_b200 as nb
    from neptune_pde_solver_b200.slab import Slab

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 with torch.distributed.run (one rank per GPU)")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    kind, shape1, ty, alpha, bytes_per_update, descr = WORKLOADS[args.workload]
    dt = np.float64 if ty == "f64" else np.float32
    # weak scaling: global grid = (N * n0) x n1 x n2, dim 0 sharded
    gshape = [shape1[0] * world] + list(shape1[1:])
    h2 = 1.0 / (max(gshape) ** 2)
    gtext, n_extra = program_text(args.workload, gshape)
    ctx = nb.Context(local)
    gdesc = nb.Module(gtext).apply_desc(0)
    if world > 1:
        # ghost planes per neighbour side: 2 * radius lets a launch fuse two steps (one exchange per fused pair);
        # deeper ghosts (--ghost) exchange once every ghost / (2 * radius) launches (neptune_rt_stencil_set_slab)
        fusable = kind in ("heat_star", "jacobi_star", "heat_box_2d9") and args.tblock != 1
        # default: four ghost planes where two steps fuse (r02, C3 at N = 2: 0.98 of linear with ghost 4 against 0.96 with
        # ghost 2 — half as many thin boundary launches for four redundant planes per cycle); --tblock d sets d * radius
        halo = args.ghost if args.ghost > 0 else ((args.tblock if args.tblock >= 2 else 4) * gdesc.radius0() if fusable else None)
        slab = Slab(gdesc, world, rank, halo=halo)
        lshape = list(slab.local_shape(gshape))
        from neptune_pde_solver_b200 import mlir_builders as mb  # local program: same body, local bounds
        ldesc = slab.local
        uid = [nb.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], world, rank)
        stencil = slab.stencil(ctx)
        lb = ldesc.out_lb
        ub = ldesc.out_ub
        own_points = 1
        for d in range(len(gshape)):
            own_points *= max(0, ldesc.ub[d] - ldesc.lb[d])
    else:
        lshape = gshape
        mod = nb.Module(gtext)
        stencil = mod.stencil(ctx, 0)
        lb, ub = gdesc.out_lb, gdesc.out_ub
        own_points = interior_points(gshape)
    if args.variant:
        stencil.set("variant", args.variant)
    if args.tblock:
        stencil.set("tblock", args.tblock)
    if args.zchunk and args.zchunk != "0":
        stencil.set("zchunk", args.zchunk)

    global PARITY_STEPS
    parity_plan = None
    if world > 1 and not args.no_parity:
        PARITY_STEPS, parity_plan = parity_steps_for(world, gdesc.radius0(), slab.halo, bool(fusable))
        if slab.own_rows[1] - slab.own_rows[0] < 2 * PARITY_STEPS:
            raise SystemExit("bench.py: slabs of %d planes are too thin for a %d-step parity band" % (slab.own_rows[1] - slab.own_rows[0], PARITY_STEPS))

    u = synth(lshape, ty, 20260403 + rank)
    fa = nb.Field(ctx, dt, lb, ub)
    fb = nb.Field(ctx, dt, lb, ub)
    fa.upload(u)
    fb.upload(u)   # both ping-pong buffers start with a valid boundary shell
    extra, extra_host = [], []
    for k in range(n_extra):
        f = nb.Field(ctx, dt, lb, ub)
        if kind == "varcoef27":
            # coefficients ~ U(0.5, 1.5)/27, seed 20260404 + k (+ 100 per rank): a stable weighted average
            c = (np.random.default_rng(20260404 + k + 100 * rank).uniform(0.5, 1.5, size=lshape) / 27.0).astype(dt)
        else:
            c = synth(lshape, ty, 20260500 + k + 100 * rank)
        f.upload(c)
        ctx.sync()
        extra.append(f)
        if not args.no_parity:
            extra_host.append(BandRows(c, band_row_ranges(world, slab.own_rows if world > 1 else (0, lshape[0]))))
        del c

    def barrier():
        ctx.sync()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- parity, in this run, on this configuration (before anything is timed) -------------------
    parity = None
    if not args.no_parity:
        if world > 1:
            ob, oe = slab.own_rows
            # the initial ghost planes of a rank hold its own synthetic data; the first exchange replaces them
            # with the neighbours' owned planes, so the global initial state is the concatenation of the owned parts
        else:
            ob, oe = 0, lshape[0]

        traced = []

        def run_fn(nsteps):
            if world > 1:
                ctx.slab_trace(True)
            r = stencil.run_steps(fa, fb, extra, nsteps)
            if world > 1:
                traced.extend(g[0] for g in ctx.slab_trace_read())    # 10 * steps advanced + (1 if the launch exchanged)
                ctx.slab_trace(False)
            return r.download()

        def gather_fn(mine):
            parts = [None] * world
            dist.all_gather_object(parts, mine)
            return parts

        parity = band_parity_check(kind, ty, alpha, h2, world, rank, (ob, oe), u, extra_host, run_fn, gather_fn)
        if world > 1:
            flag = [parity["bits_equal"] if rank == 0 else None]
            dist.broadcast_object_list(flag, src=0)
            ok = flag[0]
        else:
            ok = parity["bits_equal"]
        if world > 1:
            # the launches the checked run REALLY made (slab trace of this rank), beside the plan they were predicted from
            executed = [[k // 10, k % 10] for k in traced]
            n_ex = sum(1 for k in traced if k % 10)
            cnt = [None] * world
            dist.all_gather_object(cnt, n_ex)
            if rank == 0:
                parity["plan"] = {"columns": ["steps", "exchange", "ghost planes computed redundantly"],
                                  "predicted": [[st, int(x), ext] for (st, ext, x) in (parity_plan or [])],
                                  "executed_on_rank0": executed, "exchanging_launches_per_rank": cnt,
                                  "halo_mode": stencil.halo_mode()}
            if min(cnt) < 2:
                ok = False
                if rank == 0:
                    parity["bits_equal"] = False
                    parity["error"] = "the checked run made fewer than two exchanging launches: the exchange path is not covered"
        if rank == 0:
            parity["kernel_variant"] = stencil.variant()
        if not ok:
            if rank == 0:
                print(json.dumps({"error": "parity check failed", "parity_check": parity}), flush=True)
            raise SystemExit(3)
        fa.upload(u)
        fb.upload(u)
        del extra_host

    # ---- device-resident timing: W warm-up steps, then exactly K steps, REPS times; median ---------
    cur = stencil.run_steps(fa, fb, extra, args.warmup)
    other = fb if cur is fa else fa
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    barrier()
    t_wall0 = time.time()
    rep_ms, launches = [], 0
    for rep in range(max(1, args.reps)):
        barrier()
        n0 = ctx.launch_count()
        ctx.timer_start()
        cur = stencil.run_steps(cur, other, extra, args.steps)
        ms_rep = ctx.timer_stop()
        other = fb if cur is fa else fa
        barrier()
        launches = ctx.launch_count() - n0
        if world > 1:
            t = torch.tensor([ms_rep], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)      # a repetition takes as long as its slowest rank
            ms_rep = float(t[0])
        rep_ms.append(ms_rep)
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    ms = float(np.median(rep_ms))
    if world > 1:
        t = torch.tensor([float(own_points), float(launches)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ms_max, total_points, launches_all = ms, float(t[0]), int(t[1])
    else:
        ms_max, total_points, launches_all = ms, float(own_points), launches
    ms_per_step = ms_max / args.steps
    value = total_points / (ms_per_step * 1e-3) / 1e9
    variant = stencil.variant()

    # ---- sustained: the same loop for >= --sustained-seconds in one go, with its own clock samples. `value` above is a
    # burst figure (K steps last a few ms at K = 20: the GPU has not reached its power limit yet); under sw_power_cap the
    # clocks settle lower and this is the number a long run sees ------------------------------------------------------
    sustained = None
    if args.sustained_seconds > 0:
        n_s = max(args.steps, int(np.ceil(args.sustained_seconds * 1e3 / ms_per_step)))
        n_s += n_s % 2
        sampler2 = ClockSampler(local)
        barrier()
        if rank == 0:
            sampler2.start()
        t_s0 = time.time()
        ctx.timer_start()
        cur = stencil.run_steps(cur, other, extra, n_s)
        ms_s = ctx.timer_stop()
        other = fb if cur is fa else fa
        barrier()
        t_s1 = time.time()
        if world > 1:
            t = torch.tensor([ms_s], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_s = float(t[0])
        sustained = {"value": total_points / (ms_s / n_s * 1e-3) / 1e9, "unit": UNIT, "steps": n_s, "seconds": ms_s * 1e-3,
                     "ms_per_step": ms_s / n_s, "clocks": sampler2.stop(t_s0, t_s1) if rank == 0 else None,
                     "note": "one uninterrupted run of the timed loop, CUDA events on the sweep stream, max over ranks; "
                             "`value` is the burst figure of the K-step repetitions"}

    # ---- timeline of the slab time loop (one rank with two neighbours where there is one) ----------
    slab_trace = None
    if world > 1 and args.trace:
        ctx.slab_trace(True)
        cur = stencil.run_steps(cur, other, extra, args.steps)
        other = fb if cur is fa else fa
        rows = ctx.slab_trace_read()
        ctx.slab_trace(False)
        allrows = [None] * world
        dist.all_gather_object(allrows, rows)
        pick = 1 if world > 2 else 0
        slab_trace = {"rank": pick, "unit": "ms since the first group began",
                      "columns": ["kind (10*steps + exchanged)", "begin", "boundary_done", "exchange_done", "interior_done"],
                      "groups": [[g[0]] + [round(x, 4) for x in g[1:]] for g in allrows[pick][:12]]}

    # ---- roofline of the dominant kernel: whole-box sweep launches timed with CUDA events on the
    # launching stream (at N=1 the timed region above IS K such launches, one per step) ------------
    single = None
    if world == 1:
        kernel_ms = ms / max(1, launches)
        steps_per_launch = args.steps / max(1, launches)   # 2 when run_steps used the temporally blocked kernel
        if steps_per_launch > 1:
            # the single-step kernel beside it (one read + one write of the field per step is then
            # PHYSICAL traffic, so this fraction is a true DRAM-roofline fraction)
            stencil.set("tblock", 1)
            cur = stencil.run_steps(cur, other, extra, 3)
            other = fb if cur is fa else fa
            ctx.sync()
            ctx.timer_start()
            cur = stencil.run_steps(cur, other, extra, 20)
            other = fb if cur is fa else fa
            sms = ctx.timer_stop() / 20
            single = {"kernel": "sweep_%s" % stencil.variant().split(":")[0], "kernel_variant": stencil.variant(), "kernel_ms": sms,
                      "value": own_points / (sms * 1e-3) / 1e9, "unit": UNIT,
                      "achieved": own_points * bytes_per_update / (sms * 1e-3) / 1e9}
            stencil.set("tblock", args.tblock)
    else:
        steps_per_launch = 2.0 if variant.startswith("tb2") else 1.0
        kernel_ms = ms / (args.steps / steps_per_launch)    # per launch group: sweep launches and any exposed halo wait
    box_points = own_points
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_kind = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    achieved = box_points * bytes_per_update * steps_per_launch / (kernel_ms * 1e-3) / 1e9
    traffic, traffic_note = args.traffic, None
    if traffic is None and world == 1:
        try:   # per-launch DRAM bytes of this kernel from the committed ncu --set full capture, valid only for the kernel
            # sources it was captured from (entries are stamped with a digest of csrc/kernels/)
            ent = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload, {}).get(variant)
            if isinstance(ent, dict):
                if ent.get("digest") == kernel_digest():
                    traffic = ent.get("bytes")
                else:
                    traffic_note = "profiles/traffic.json entry for %s was captured from other kernel sources (digest %s, now %s): not reported" % (
                        variant, ent.get("digest"), kernel_digest())
            elif ent is not None:
                traffic_note = "profiles/traffic.json entry for %s carries no source digest: not reported" % variant
        except (OSError, ValueError):
            traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_kind": peak_kind, "frac_of_8TBps": achieved / 8000.0,
                "kernel": "sweep_%s<%s>" % (variant.split(":")[0], kind), "kernel_ms": kernel_ms,
                "algorithmic_bytes_per_launch": box_points * bytes_per_update * steps_per_launch,
                "steps_per_launch": steps_per_launch,
                "bytes_per_update": bytes_per_update}
    if traffic_note:
        roofline["traffic_note"] = traffic_note
    if steps_per_launch > 1:
        roofline["note"] = ("temporally blocked kernel: two time steps per launch, so algorithmic bytes (16 B per point-update) "
                            "exceed the DRAM bytes moved (see traffic); frac > 1 of the copy peak is possible by construction")
    if single:
        single["frac"] = single["achieved"] / peak
        roofline["single_step_kernel"] = single

    # ---- e2e: host buffers in, host buffers out, copies inside the timed region every step ---------
    # N = 1: the reference-facing entry call `@entry(out, in)` on pinned host memrefs (neptune_b200_run_host).
    # N > 1: every rank uploads its local box, runs ONE slab-mode step (halo exchange included) and downloads it;
    #        the single-GPU entry has no decomposition, so the multi-GPU public API (Slab + run_steps) is used.
    e2e, e_launches = None, 0
    field_bytes = int(np.prod(lshape)) * np.dtype(dt).itemsize
    if args.e2e_steps > 0:
        keep, hargs = [], []
        for k in range(2 + n_extra):
            t, a = pinned_like(lshape, dt)
            keep.append(t)
            hargs.append(a)
        hargs[1][...] = u
        for k in range(n_extra):
            hargs[2 + k][...] = synth(lshape, ty, 20260500 + k + 100 * rank)
        e_steps = max(1, min(args.steps, args.e2e_steps))
        if world == 1:
            emod = mod
            if args.variant:
                emod.stencil(ctx, 0).set("variant", args.variant)

            def e2e_step():
                emod.run(ctx, "@entry", hargs)          # upload inputs, sweep, download result, sync
                hargs[0], hargs[1] = hargs[1], hargs[0]  # the host time loop's swap
            api = ("neptune_b200_run_host(@entry(out, in)) on pinned host memrefs, one sweep per call; upload | sweep | download "
                   "pipelined over dim-0 plane chunks (neptune_rt_apply_host)")
        else:
            def e2e_step():
                fa.upload_planes(hargs[1], 0, lshape[0])
                for k in range(n_extra):
                    extra[k].upload_planes(hargs[2 + k], 0, lshape[0])
                r = stencil.run_steps(fa, fb, extra, 1)  # ghost exchange + boundary / interior sweep of this rank's slab
                r.download_planes(hargs[0], 0, lshape[0])
                ctx.sync()
                hargs[0], hargs[1] = hargs[1], hargs[0]
            api = ("per rank: neptune_rt_field_upload(local box, pinned) + neptune_rt_run_steps(1 step, slab mode, halo exchange included) + "
                   "neptune_rt_field_download; not pipelined; halo mode " + stencil.halo_mode())
        for _ in range(2):
            e2e_step()
        barrier()
        l0 = ctx.launch_count()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            e2e_step()
        barrier()
        e_sec = time.perf_counter() - t0
        e_launches = ctx.launch_count() - l0
        if world > 1:
            t = torch.tensor([e_sec], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e_sec = float(t[0])
        e_value = total_points / (e_sec / e_steps) / 1e9
        e2e = {"value": e_value, "unit": UNIT, "h2d_bytes_per_step": field_bytes * (1 + n_extra) * world,
               "d2h_bytes_per_step": field_bytes * world, "steps": e_steps, "ms_per_step": e_sec / e_steps * 1e3, "api": api}
        if args.link_probe:
            # what the host links give when every rank copies at once: bare cudaMemcpyAsync H2D and D2H of one field each,
            # concurrently on two streams (the ceiling of any host-buffer path on this box)
            import torch as _t
            dev = _t.empty(int(np.prod(lshape)), dtype=_t.float64 if dt == np.float64 else _t.float32, device="cuda")
            dev2 = _t.empty_like(dev)
            s1, s2 = _t.cuda.Stream(), _t.cuda.Stream()
            barrier()
            t0 = time.perf_counter()
            for _ in range(3):
                with _t.cuda.stream(s1):
                    dev.copy_(keep[0].view(-1), non_blocking=True)
                with _t.cuda.stream(s2):
                    keep[1].view(-1).copy_(dev2, non_blocking=True)
            s1.synchronize(); s2.synchronize()
            barrier()
            sec = (time.perf_counter() - t0) / 3
            if world > 1:
                t = _t.tensor([sec], dtype=_t.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                sec = float(t[0])
            e2e["host_link_probe"] = {"GBps_each_way_per_rank": field_bytes / sec / 1e9, "ranks_at_once": world,
                                      "note": "bare pinned cudaMemcpyAsync H2D + D2H of one field each, concurrently, all ranks at once"}
            del dev, dev2

    # ---- CPU baseline (rank 0, N=1 only): bounded sample of the same workload --------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cshape = list(shape1)
        g, cms, n, cores = cpu_reference_arm(args.workload, cshape, budget_s=args.cpu_seconds, min_steps=2, max_steps=40, warmup=1)
        cpu = {"value": g, "unit": UNIT, "cores": cores, "kind": "port", "ms_per_step": cms,
               "sample": "%d full sweeps of the %s grid through the oracle's gcc -O2 -ffp-contract=off scalar "
                         "restatement of DataflowLowering.cpp:258-448 + store copy :165-220 (1 thread: the "
                         "reference emits no parallel construct)" % (n, "x".join(map(str, cshape)))}
        if args.omp_context:
            try:
                og, oc = cpu_openmp_context(args.workload, cshape)
                cpu["context_openmp"] = {"value": og, "cores": oc, "note": "NOT the reference: ping-pong, no copies, OpenMP all cores"}
            except Exception as e:
                cpu["context_openmp"] = {"error": str(e)[:200]}

    if rank == 0:
        line = {
            "metric": METRIC if args.workload == "c3" else "Gpoint-updates/s, " + descr,
            "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": ty, "data": "synthetic",
            "config": {"workload": descr, "grid_per_gpu": list(shape1), "global_grid": gshape,
                       "decomposition": ("dim-0 slabs, %d ghost plane(s) per neighbour side, %d step(s) per launch, one exchange of %d plane(s) every %d step(s) "
                                         "(%s), overlapped with the interior sweep of the launch that produces them"
                                         % (slab.halo, 2 if variant.startswith("tb2") else 1, slab.halo, max(1, slab.halo // max(1, gdesc.radius0())),
                                            {"p2p": "pushed into the neighbour's field by a copy engine over NVLink, CUDA IPC + stream memory-op flags",
                                             "nccl": "ncclSend/ncclRecv"}.get(stencil.halo_mode(), "?"))) if world > 1 else "single domain",
                       "halo_mode": stencil.halo_mode() if world > 1 else None,
                       "points_per_step": int(total_points), "kernel_variant": variant,
                       "timing": "median of %d repetitions of exactly %d steps each, CUDA events on the sweep stream, max over ranks per repetition" % (len(rep_ms), args.steps),
                       "rep_ms": [round(x, 4) for x in rep_ms], "rep_spread": (max(rep_ms) - min(rep_ms)) / ms if ms > 0 else None,
                       "parity_check": parity,
                       "l2": "inputs larger than L2 (%.2f GB per field vs 126 MB L2); no flush needed" % (field_bytes / 1e9),
                       "fma": False},
            "sustained": sustained, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": int(launches_all), "e2e_gpu_launches": int(e_launches), "clocks": clocks,
            "achieved_GBps_step": total_points * bytes_per_update / (ms_per_step * 1e-3) / 1e9 / world,
        }
        if slab_trace:
            line["slab_trace"] = slab_trace
        print(json.dumps(line), flush=True)
    for f in [fa, fb] + extra:
        f.free()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--variant", default="")
    ap.add_argument("--zchunk", default="")
    ap.add_argument("--tblock", type=int, default=0, help="temporal block depth: 1 or 2 steps per launch (0 = library default); 4 / 8 = two steps per launch and, in slab mode, one exchange of tblock * radius ghost planes per tblock steps")
    ap.add_argument("--ghost", type=int, default=0, help="ghost planes per neighbour side in slab mode (0 = 4 * radius where two steps fuse [--tblock d >= 2: d * radius], else radius)")
    ap.add_argument("--reps", type=int, default=5, help="repetitions of the K-step timed region; the median is reported")
    ap.add_argument("--sustained-seconds", type=float, default=2.0, help="length of the additional uninterrupted run reported as `sustained` (0 = skip)")
    ap.add_argument("--no-parity", action="store_true", help="skip the in-run bitwise parity check against the oracle")
    ap.add_argument("--trace", action="store_true", help="N > 1: add per-launch-group timestamps of the slab time loop")
    ap.add_argument("--link-probe", action="store_true", help="add a bare pinned-memcpy H2D+D2H measurement (all ranks at once) to e2e")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--omp-context", dest="omp_context", action="store_true", default=True,
                    help="also report a best-effort all-core OpenMP CPU number beside the faithful 1-thread reference lowering (context, not the reference)")
    ap.add_argument("--no-omp-context", dest="omp_context", action="store_false")
    ap.add_argument("--traffic", type=float, default=None,
                    help="dram bytes per launch from the committed ncu --set full capture (profiles/)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
