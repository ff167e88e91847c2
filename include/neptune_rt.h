/* neptune_rt.h — C ABI of the B200 runtime behind the `neptuneir-to-b200` pipeline.
 *
 * This is the drop-in boundary for the one hot path this project accelerates: the
 * `neptune_ir.apply` stencil sweep.  In the reference that op is lowered by
 * `ApplyToSCFForLowering::matchAndRewrite` (lib/Passes/DataflowLowering.cpp:258-448) into a
 * malloc + memref.copy + scalar scf.for nest, and `neptune_ir.store` into a second full copy
 * (lib/Passes/DataflowLowering.cpp:165-220).  Here the same semantics are executed by
 * hand-written sm_100a kernels; a pass emits calls to the functions below the way
 * lib/Passes/RuntimeLowering.cpp:39-53,169-178 emits calls to `_neptune_rt_runtime_*`
 * (declared in include/Runtime/PETSc/NeptunePETScRuntime.h:79-105).
 *
 * Conventions (contrast with the reference shim):
 *   - plain C, opaque handles + POD out-params, never a struct returned by value;
 *   - every function returns 0 on success, a negative NEPTUNE_RT_E* code on failure, and
 *     leaves a message retrievable with neptune_rt_last_error();
 *   - every *_alloc / *_create has a matching *_free / *_destroy; the ctx owns streams,
 *     scratch memory and the NCCL communicator;
 *   - one host thread drives one ctx (one process per GPU); a ctx is not thread-safe;
 *   - there is NO CPU fallback: without a CUDA device neptune_rt_init fails.
 *
 * Normative semantics of one sweep (from DataflowLowering.cpp:258-448):
 *   out[q] = in0[q]                      for every q of the result box,
 *   out[p - out_lb] = body(in_k[p + off - in_lb_k] ...)   for every p in [lb, ub),
 * the body being the scalar op list of the apply region evaluated op by op, in the body's own
 * order, IEEE-754 round-to-nearest, no contraction (unless the stencil is created with
 * NEPTUNE_RT_FLAG_FMA, which moves the result to the 1e-12 tolerance tier).
 */
#ifndef NEPTUNE_RT_H
#define NEPTUNE_RT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NEPTUNE_RT_MAX_RANK 3
#define NEPTUNE_RT_MAX_INPUTS 32
#define NEPTUNE_RT_MAX_CONSTS 64

/* status codes */
#define NEPTUNE_RT_OK 0
#define NEPTUNE_RT_EINVAL (-1)      /* bad argument / malformed descriptor              */
#define NEPTUNE_RT_ENODEV (-2)      /* no CUDA device / driver (there is no CPU path)   */
#define NEPTUNE_RT_ECUDA (-3)       /* a CUDA call failed                               */
#define NEPTUNE_RT_EUNSUPPORTED (-4)/* outside the recognised set / cannot be compiled     */
#define NEPTUNE_RT_ENCCL (-5)       /* NCCL missing or a collective failed              */
#define NEPTUNE_RT_ENOMEM (-6)

/* element types: FieldType/TempType `element` (include/Dialect/NeptuneIR/NeptuneIRTypes.td:12-59) */
#define NEPTUNE_RT_F32 0
#define NEPTUNE_RT_F64 1

/* body op codes: neptune_ir.access (NeptuneIROps.td:205-228) plus the float part of the
 * linear-body whitelist of lib/Passes/VerifyAndAnnotate.cpp:167-204 */
#define NEPTUNE_RT_OP_ACCESS 0
#define NEPTUNE_RT_OP_CONST 1
#define NEPTUNE_RT_OP_ADD 2
#define NEPTUNE_RT_OP_SUB 3
#define NEPTUNE_RT_OP_MUL 4
#define NEPTUNE_RT_OP_DIV 5
#define NEPTUNE_RT_OP_NEG 6

/* One scalar op of the apply region.  The result of op i is value number i. */
typedef struct neptune_rt_body_op {
  int32_t opcode;                        /* NEPTUNE_RT_OP_*                                 */
  int32_t a, b;                          /* operand value numbers (ADD..NEG), else -1       */
  int32_t input;                         /* ACCESS: which apply input                       */
  int64_t offset[NEPTUNE_RT_MAX_RANK];   /* ACCESS: constant neighbour offset, logical dims */
  double value;                          /* CONST: literal (already rounded to the dtype)   */
} neptune_rt_body_op;

/* Everything `ApplyOp` (NeptuneIROps.td:164-197) carries that the lowering reads. */
typedef struct neptune_rt_stencil_desc {
  int32_t rank;                          /* 1..3                                            */
  int32_t dtype;                         /* NEPTUNE_RT_F32 / F64                            */
  int32_t n_inputs;                      /* variadic temp inputs                            */
  int32_t n_ops;
  const neptune_rt_body_op *ops;         /* body in program order                           */
  int32_t yield_value;                   /* value number passed to neptune_ir.yield         */
  int32_t flags;                         /* NEPTUNE_RT_FLAG_*                               */
  int64_t lb[NEPTUNE_RT_MAX_RANK], ub[NEPTUNE_RT_MAX_RANK];         /* apply `bounds`       */
  int64_t out_lb[NEPTUNE_RT_MAX_RANK], out_ub[NEPTUNE_RT_MAX_RANK]; /* result temp bounds   */
  const int64_t *in_lb;                  /* n_inputs x rank, row-major: input temp bounds   */
  const int64_t *in_ub;
} neptune_rt_stencil_desc;

#define NEPTUNE_RT_FLAG_FMA 1            /* allow a*b+c contraction (tolerance tier)        */

typedef struct neptune_rt_ctx neptune_rt_ctx;
typedef struct neptune_rt_field neptune_rt_field;
typedef struct neptune_rt_stencil neptune_rt_stencil;

/* ---- context ------------------------------------------------------------------------- */
/* Replaces the lazy `neptune_rt_init_noargs` (NeptunePETScRuntime.h:47). device < 0 picks
 * LOCAL_RANK (or 0). */
int neptune_rt_init(int device, neptune_rt_ctx **ctx);
/* Releases the ctx AND every field / stencil created on it that the caller has not freed (their handles become
 * invalid). neptune_rt_ctx_alive tells a handle owner that may outlive the ctx whether it still exists. */
int neptune_rt_finalize(neptune_rt_ctx *ctx);
int neptune_rt_ctx_alive(neptune_rt_ctx *ctx);
int neptune_rt_sync(neptune_rt_ctx *ctx);
/* Message of the last failure on this thread (ctx may be NULL for init failures). */
const char *neptune_rt_last_error(neptune_rt_ctx *ctx);
/* Number of kernels this ctx has launched so far (for accounting in benchmarks). */
int64_t neptune_rt_launch_count(neptune_rt_ctx *ctx);
/* The CUDA stream sweeps are launched on (a cudaStream_t, returned as void*). */
void *neptune_rt_stream(neptune_rt_ctx *ctx);
/* Device-side stopwatch on the sweep stream (CUDA events): start records an event, stop records a
 * second one, waits for it and returns the elapsed milliseconds between the two. */
int neptune_rt_timer_start(neptune_rt_ctx *ctx);
int neptune_rt_timer_stop(neptune_rt_ctx *ctx, double *ms);

/* ---- device fields --------------------------------------------------------------------- */
/* Storage of a !neptune_ir.field / !temp: identity-layout row-major box `ub - lb`
 * (NeptuneIRTypeConverter, lib/Passes/DataflowLowering.cpp:41-71), resident in HBM. */
int neptune_rt_field_alloc(neptune_rt_ctx *ctx, int dtype, int rank, const int64_t *lb,
                           const int64_t *ub, neptune_rt_field **field);
int neptune_rt_field_free(neptune_rt_field *field);
/* A field over device memory the CALLER owns (row-major box ub - lb, numel * sizeof(elem) bytes; 16-byte
 * aligned rows let the staged kernels run, anything else takes the direct kernel). Not freed by
 * neptune_rt_field_free. The matrix-free `y = A(x)` hook: the reference applies operators inside its Krylov
 * solves through a MatShell callback on host vectors (NeptunePETScRuntime.cpp:182-230, MatMultThunk); a
 * device-resident solver wraps its work vectors once and calls neptune_rt_apply. */
int neptune_rt_field_wrap_device(neptune_rt_ctx *ctx, int dtype, int rank, const int64_t *lb,
                                 const int64_t *ub, void *device_ptr, neptune_rt_field **field);
/* Whole-box copies between a contiguous row-major host buffer and the field (async on the ctx
 * stream; host memory should be pinned for full PCIe speed).  */
int neptune_rt_field_upload(neptune_rt_field *field, const void *host);
int neptune_rt_field_download(neptune_rt_field *field, void *host);
/* Planes [p0, p1) of dim 0 only (used by the pipelined host path and the halo tests). */
int neptune_rt_field_upload_planes(neptune_rt_field *field, const void *host, int64_t p0, int64_t p1);
int neptune_rt_field_download_planes(neptune_rt_field *field, void *host, int64_t p0, int64_t p1);
int neptune_rt_field_fill(neptune_rt_field *field, double value);
void *neptune_rt_field_device_ptr(neptune_rt_field *field);
int64_t neptune_rt_field_bytes(neptune_rt_field *field);
/* `neptune_ir.store` (DataflowLowering.cpp:165-220): whole copy when lb/ub are NULL, else the
 * sub-box [lb, ub) in logical coordinates of both operands. */
int neptune_rt_field_store(neptune_rt_field *dst, neptune_rt_field *src, const int64_t *lb,
                           const int64_t *ub);

/* `neptune_ir.reduce` kind = "sum" (NeptuneIROps.td:272-299, ReduceLowering in
 * lib/Passes/DataflowLowering.cpp:589-698): the sum of the field over [lb, ub) (logical coordinates;
 * NULL = the whole box). The reference adds element by element in row-major order; this is a
 * deterministic fixed-partition parallel sum accumulated in fp64 (rounded to f32 for f32 fields), so
 * parity is the tolerance tier: |gpu - ref| <= n * eps(elem) * sum|x|. With allreduce != 0 and a
 * communicator attached, the per-rank sums are added across ranks (ncclAllReduce): pass bounds clipped
 * to the rank's owned planes. Synchronous. */
int neptune_rt_field_reduce_sum(neptune_rt_field *field, const int64_t *lb, const int64_t *ub,
                                int allreduce, double *result);

/* The same box sum for callers that must not allocate. It has no counterpart in the reference: the
 * reference's reduce is a loop nest over a memref somebody else allocated, so the question never arises
 * there; it arises here because neptune_rt_field_reduce_sum cudaMallocs its partial-sum scratch (and a
 * pinned result word) on first use.
 *
 * Who needs this: a solver that owns the GPU's memory (PETSc-style pools sized at start-up, a framework's
 * caching allocator that has already reserved all of HBM, so a cudaMalloc from a library fails), compiler-
 * generated code whose buffers all come from its own bufferisation, and any inner loop that cannot afford
 * the device-wide synchronisation cudaMalloc/cudaFree imply (one first-use allocation inside an iteration
 * stalls the halo/interior overlap of the time step around it). This entry point takes no ctx and no field
 * handle, calls no cudaMalloc / cudaFree / cudaMallocHost nor any routine that does, and keeps its partial
 * sums in a static device array inside the library image: NEPTUNE_RT_REDUCE_STATIC_SCRATCH_BYTES bytes on
 * every GPU that loads the library, held until the process exits.
 *
 *   src         device pointer to a dense row-major array of `shape` (rank 1..3), dtype NEPTUNE_RT_F32/F64
 *   lb, ub      zero-based sub-box [lb, ub) of that array to sum; NULL = the whole array; an empty box sums to 0
 *   result_dev  device pointer, 8-byte aligned, caller-owned: receives the sum as a double (an f32 array's
 *               sum is rounded to f32 first, as neptune_rt_field_reduce_sum reports it)
 *   stream      a cudaStream_t (NULL = the default stream). The call only enqueues two kernels; it does not
 *               wait for them and copies nothing to the host.
 *
 * src and result_dev must belong to the CURRENT device (cudaSetDevice before the call); anything else is
 * NEPTUNE_RT_EINVAL. Same partition and same fixed trees as neptune_rt_field_reduce_sum, hence the same value
 * bit for bit, run to run.
 *
 * Limits: the scratch is one slot per CTA and does not grow with the box, so there is no element-count
 * limit; what is limited is what the kernel indexes in 32 bits: every shape[d] must be
 * <= NEPTUNE_RT_REDUCE_STATIC_MAX_EXTENT (2^30). A larger axis is refused with NEPTUNE_RT_EINVAL and a
 * message (neptune_rt_last_error(NULL)) naming the axis and the limit; fold the axis into two, or use
 * neptune_rt_field_reduce_sum. Device ordinals above 63 are NEPTUNE_RT_EUNSUPPORTED.
 *
 * Shared state: there is ONE scratch array per GPU for the whole process. Two calls that can run on the
 * same GPU at the same time (different streams, different threads) race on it and may both return garbage:
 * order them (same stream, or an event between them). The array keeps whatever the last call left in it;
 * that is harmless, because each call overwrites every slot before reading any, and nothing needs clearing.
 * The first call on each device looks the array's device address up (cudaGetSymbolAddress, which can make
 * the CUDA runtime load the library's module on that device); make that first call during set-up, before
 * any stream capture on that device. */
#define NEPTUNE_RT_REDUCE_STATIC_SCRATCH_BYTES 4736 /* 592 partial sums (4 CTAs per SM x 148 SMs) x 8 bytes */
#define NEPTUNE_RT_REDUCE_STATIC_MAX_EXTENT 1073741824LL
int neptune_rt_reduce_sum_static(const void *src, int dtype, int rank, const int64_t *shape,
                                 const int64_t *lb, const int64_t *ub, double *result_dev, void *stream);

/* ---- stencils -------------------------------------------------------------------------- */
/* 64-bit structure key of a body (constants' values, bounds and extents excluded): two applies
 * with the same key run the same compiled kernel. Pure host code. */
uint64_t neptune_rt_body_key(const neptune_rt_stencil_desc *desc);
/* 1 when the library holds compiled sm_100a kernels for this body key (prebuilt catalogue, or a body already
 * compiled at run time in this process). Pure host code. */
int neptune_rt_body_compiled(uint64_t key);
/* Bodies outside the prebuilt catalogue are compiled on first use by neptune_rt_stencil_create: NVRTC builds the same
 * hand-written sweep kernels for the body's generated point functor (sm_100a cubin, cached on disk under
 * $NEPTUNE_B200_JIT_CACHE or ~/.cache/neptune_b200_jit), so every body the recogniser accepts runs — the reference
 * lowers any body too (lib/Passes/DataflowLowering.cpp:413-423). This entry does only the compile step (no GPU
 * needed): it fills the cache ahead of time and reports the compile time. */
int neptune_rt_body_jit_compile(const neptune_rt_stencil_desc *desc, double *seconds, char *err, size_t err_len);
int neptune_rt_stencil_create(neptune_rt_ctx *ctx, const neptune_rt_stencil_desc *desc,
                              neptune_rt_stencil **stencil);
int neptune_rt_stencil_destroy(neptune_rt_stencil *stencil);
/* Tuning knobs, all optional: "variant" = "auto" | "direct" | "tma:<cfg>" | "tb2:<cfg>", "zchunk" =
 * planes per work item, "tblock" = time steps fused per launch by neptune_rt_run_steps (0 = auto:
 * two where the temporally blocked kernel applies, else one; 1; 2). Unknown keys return EINVAL. */
int neptune_rt_stencil_set(neptune_rt_stencil *stencil, const char *key, const char *value);
/* The kernel variants compiled for this stencil's body ("direct", "tma:64x16", ...). */
int neptune_rt_stencil_variant_count(neptune_rt_stencil *stencil);
const char *neptune_rt_stencil_variant_name(neptune_rt_stencil *stencil, int index);
/* Name of the kernel variant the last launch used ("" before the first launch). */
const char *neptune_rt_stencil_variant(neptune_rt_stencil *stencil);
/* One sweep, asynchronous on the ctx stream: out = apply(in[0..n_in)). `out` must not alias
 * an input (the reference allocates a fresh result, DataflowLowering.cpp:273-281). */
int neptune_rt_apply(neptune_rt_stencil *stencil, int n_in, neptune_rt_field *const *in,
                     neptune_rt_field *out);
/* Sweep only planes [p0, p1) of dim 0 of the result box (slab boundary / interior split). */
int neptune_rt_apply_planes(neptune_rt_stencil *stencil, int n_in, neptune_rt_field *const *in,
                            neptune_rt_field *out, int64_t p0, int64_t p1, void *stream);
/* Sweep planes [p0, g0) and [g1, p1) of dim 0 (equally wide, g0 < g1) and leave [g0, g1) alone: the two
 * boundary ranges of a slab that has a neighbour on both sides, as ONE launch where the kernel variant
 * takes two plane ranges (else as two). The slab time loop uses it before each halo exchange. */
int neptune_rt_apply_planes2(neptune_rt_stencil *stencil, int n_in, neptune_rt_field *const *in,
                             neptune_rt_field *out, int64_t p0, int64_t g0, int64_t g1, int64_t p1, void *stream);
/* One sweep on HOST buffers, pipelined over dim-0 plane chunks: upload of chunk i+1, sweep of chunk
 * i and download of chunk i-1 run concurrently (two copy engines + SMs), so the call costs about
 * max(upload, download) instead of their sum. in_host[k] == NULL means input k is already resident
 * in in[k]. chunk_planes = 0 picks about 16 chunks. Synchronous: returns when out_host is complete.
 * Host memory should be pinned. This is what `@entry(out, in)` of the reference's generated code
 * (one scf.for nest per call on host memrefs) becomes on the B200 path. */
int neptune_rt_apply_host(neptune_rt_stencil *stencil, int n_in, neptune_rt_field *const *in,
                          const void *const *in_host, neptune_rt_field *out, void *out_host,
                          int64_t chunk_planes);
/* nsteps sweeps ping-ponging a <-> b (in[0] = a on odd steps' input ...): the host time loop of
 * SURVEY.md §3.3 kept resident on the device. extra[] are inputs 1..n_in-1 (coefficients, rhs).
 * On return *result is a or b, whichever holds the last state. With a communicator attached
 * (neptune_rt_comm_init) every launch exchanges its ghost planes with the dim-0 neighbours,
 * overlapped with the interior sweep. Where the temporally blocked kernel applies (rank 3, one
 * input, radius 1, bounds inside the box) two steps are fused per launch; the result is bitwise
 * the same as nsteps single sweeps. */
int neptune_rt_run_steps(neptune_rt_stencil *stencil, neptune_rt_field *a, neptune_rt_field *b,
                         int n_extra, neptune_rt_field *const *extra, int64_t nsteps,
                         neptune_rt_field **result);

/* ---- slab decomposition + halo exchange (dim 0) ----------------------------------------- */
/* Planes [*begin, *end) of a dim-0 extent n0 owned by `rank` of `nranks` (balanced, contiguous).
 * Pure host code. */
int neptune_rt_slab_partition(int64_t n0, int nranks, int rank, int64_t *begin, int64_t *end);
/* Rewrites a global descriptor into the local one of `rank`: dim-0 bounds of the result and of
 * every input become [begin - halo, end + halo) clipped to the global box, the apply bounds are
 * intersected with the owned planes. out_* arrays must hold n_inputs*rank entries. Pure host. */
int neptune_rt_slab_localize(const neptune_rt_stencil_desc *global_desc, int nranks, int rank,
                             int64_t halo, neptune_rt_stencil_desc *local_desc,
                             int64_t *local_in_lb, int64_t *local_in_ub);
/* Tells a stencil created from a localized descriptor how its box sits in the global domain:
 * `ghost` planes on every dim-0 side that has a neighbour (the `halo` given to slab_localize) and
 * the GLOBAL apply bounds along dim 0. With ghost >= 2*radius neptune_rt_run_steps fuses two time
 * steps per launch; with G ghost planes it exchanges G planes once every G / (radius * steps per launch)
 * launches, the launches in between also computing the ghost planes that stay valid (deep halos: fewer
 * thin boundary launches and exchanges for a few redundant planes). Without this call it assumes
 * ghost = radius and runs one step per launch. Decisions that depend on rank-local facts (slab thickness,
 * alignment) are agreed by all ranks with one small ncclAllReduce per (fields, settings). */
int neptune_rt_stencil_set_slab(neptune_rt_stencil *stencil, int64_t ghost, int64_t global_lb0,
                                int64_t global_ub0);
/* The launch plan neptune_rt_run_steps follows in slab mode, as data (pure host code; the time loop itself calls the same
 * function): launch i advances plan[i].steps time steps (2 where two steps fuse) over the owned planes extended by
 * plan[i].ext ghost planes on every side that has a neighbour — computed redundantly with the neighbour, bit for bit
 * the same arithmetic — and when plan[i].exchange is set its `ghost` owned boundary planes are exchanged after it
 * (computed first, the exchange overlapping the interior). With G ghost planes that is one exchange per
 * G / (radius * steps per launch) launches. Returns the number of launches, or -1 for impossible arguments; writes at
 * most `capacity` entries. */
typedef struct neptune_rt_slab_launch { int32_t steps; int32_t exchange; int64_t ext; } neptune_rt_slab_launch;
int64_t neptune_rt_slab_schedule(int64_t nsteps, int64_t radius, int64_t ghost, int fuse2,
                                 neptune_rt_slab_launch *plan, int64_t capacity);
/* How the last slab-mode neptune_rt_run_steps of this stencil exchanged ghost planes ("" before the first):
 *   "p2p"  — the default between GPUs of one node: every rank PUSHES its boundary planes into the neighbour's field
 *            memory (CUDA IPC mapping) with a copy engine over NVLink, ordered by stream memory operations
 *            (cuStreamWriteValue32 / cuStreamWaitValue32) on four flag words per rank; no kernel, hence no SM, takes part,
 *            so the exchange really runs beside the persistent interior launch (which holds every SM's shared memory);
 *   "nccl" — ncclSend/ncclRecv groups on the high-priority stream: NEPTUNE_RT_HALO=nccl, caller-owned (wrapped) field
 *            memory, or when IPC / stream memory operations are unavailable. All ranks agree on one mode. */
const char *neptune_rt_stencil_halo_mode(neptune_rt_stencil *stencil);
/* 128-byte NCCL unique id (rank 0 creates it; the host harness broadcasts it). */
int neptune_rt_comm_unique_id(void *id128);
int neptune_rt_comm_init(neptune_rt_ctx *ctx, const void *id128, int nranks, int rank);
/* Exchange `width` planes with rank-1 / rank+1: owned boundary planes of `field` go to the
 * neighbours' ghost planes. own_begin/own_end are the owned planes in field storage rows. */
int neptune_rt_halo_exchange(neptune_rt_ctx *ctx, neptune_rt_field *field, int64_t own_begin,
                             int64_t own_end, int64_t width, void *stream);

/* Timeline of the slab time loop, for measuring what the exchange costs: after neptune_rt_slab_trace(ctx, 1) the
 * next neptune_rt_run_steps records four CUDA events per launch group; neptune_rt_slab_trace_read returns the
 * number of groups and writes, per group, milliseconds since the first group began: [0] group begins, [1] boundary
 * planes done, [2] halo exchange done, [3] interior done; kinds[g] = 10 * steps advanced + (1 if it exchanged). */
int neptune_rt_slab_trace(neptune_rt_ctx *ctx, int enable);
int neptune_rt_slab_trace_read(neptune_rt_ctx *ctx, double *ms4, int capacity_groups, int *kinds);

/* ---- memref-level entry points (what MLIR-generated code calls) -------------------------- */
/* MLIR's memref descriptor as the LLVM lowering lays it out, same field order as the reference's
 * NeptuneMemRef1D / NeptuneMemRef2D (include/Runtime/PETSc/NeptunePETScRuntime.h:22-42). */
typedef struct neptune_rt_memref_1d { void *allocated; void *aligned; int64_t offset; int64_t sizes[1]; int64_t strides[1]; } neptune_rt_memref_1d;
typedef struct neptune_rt_memref_2d { void *allocated; void *aligned; int64_t offset; int64_t sizes[2]; int64_t strides[2]; } neptune_rt_memref_2d;
typedef struct neptune_rt_memref_3d { void *allocated; void *aligned; int64_t offset; int64_t sizes[3]; int64_t strides[3]; } neptune_rt_memref_3d;
/* One sweep on host memrefs — the drop-in for the loop nest the reference emits for an apply
 * (lib/Passes/DataflowLowering.cpp:258-448): `in` points to n_in descriptors (generated code builds the array on
 * its stack), `out` to the result's. Element (i0, i1, ...) is aligned[offset + sum i_k * strides[k]]; sizes must
 * equal the apply's operand / result types and strides must be the identity (row-major contiguous) layout —
 * anything else aborts with a message naming the dimension (a silent wrong answer is worse). The stencil and the
 * device fields are created on first use and cached by the descriptor's content; each call is one pipelined
 * upload | sweep | download (neptune_rt_apply_host). Uses a lazily created default ctx (the analogue of
 * neptune_rt_init_noargs, NeptunePETScRuntime.cpp:91-98) and aborts on failure like the reference's shims
 * (NeptunePETScRuntime.cpp:15-30): generated code does not check statuses. */
void neptune_rt_apply_memref_1d(const neptune_rt_stencil_desc *desc, int n_in, const neptune_rt_memref_1d *in, const neptune_rt_memref_1d *out);
void neptune_rt_apply_memref_2d(const neptune_rt_stencil_desc *desc, int n_in, const neptune_rt_memref_2d *in, const neptune_rt_memref_2d *out);
void neptune_rt_apply_memref_3d(const neptune_rt_stencil_desc *desc, int n_in, const neptune_rt_memref_3d *in, const neptune_rt_memref_3d *out);
/* The same on bare host pointers to contiguous row-major boxes (offset 0, identity strides). */
void neptune_rt_apply_memref(const neptune_rt_stencil_desc *desc, int n_in,
                             const void *const *in_host, void *out_host);
/* Releases the stencils and device fields the memref entry points cached. */
void neptune_rt_memref_cache_clear(void);

/* ---- what the MLIR pass emits (mlir/lib/Passes/ApplyToB200.cpp) --------------------------- */
/* A descriptor as one position-independent byte string: the pass packs it at compile time and stores the bytes in
 * an llvm.mlir.global constant (a struct with pointers would need relocations in a constant global). Returns the
 * number of bytes the packed form takes; writes it when `capacity` suffices. Pure host code. */
int64_t neptune_rt_desc_pack(const neptune_rt_stencil_desc *desc, void *buf, int64_t capacity);
/* The call builder: an apply has a variable number of operands, a C function has not, so generated code makes
 *     %c = call @neptune_rt_call_begin(%desc_global, %bytes)
 *     call @neptune_rt_call_input_<rank>d(%c, k, %operand_k)          once per operand
 *     call @neptune_rt_call_run_<rank>d(%c, %result)                  result buffer allocated by the caller
 * with the functions declared `llvm.emit_c_interface`, so that func-to-llvm passes every memref as a pointer to its
 * descriptor; the `_mlir_ciface_` names that lowering produces are exported beside the plain ones. The call object
 * is cached per descriptor global; sweeps run through neptune_rt_apply_memref's cached, pipelined path. Aborts on
 * failure like the reference's shims (NeptunePETScRuntime.cpp:15-30). One host thread. */
void *neptune_rt_call_begin(const void *packed_desc, int64_t bytes);
void neptune_rt_call_input_1d(void *call, int32_t k, const neptune_rt_memref_1d *operand);
void neptune_rt_call_input_2d(void *call, int32_t k, const neptune_rt_memref_2d *operand);
void neptune_rt_call_input_3d(void *call, int32_t k, const neptune_rt_memref_3d *operand);
void neptune_rt_call_run_1d(void *call, const neptune_rt_memref_1d *result);
void neptune_rt_call_run_2d(void *call, const neptune_rt_memref_2d *result);
void neptune_rt_call_run_3d(void *call, const neptune_rt_memref_3d *result);
void *_mlir_ciface_neptune_rt_call_begin(const void *packed_desc, int64_t bytes);
void _mlir_ciface_neptune_rt_call_input_1d(void *call, int32_t k, const neptune_rt_memref_1d *operand);
void _mlir_ciface_neptune_rt_call_input_2d(void *call, int32_t k, const neptune_rt_memref_2d *operand);
void _mlir_ciface_neptune_rt_call_input_3d(void *call, int32_t k, const neptune_rt_memref_3d *operand);
void _mlir_ciface_neptune_rt_call_run_1d(void *call, const neptune_rt_memref_1d *result);
void _mlir_ciface_neptune_rt_call_run_2d(void *call, const neptune_rt_memref_2d *result);
void _mlir_ciface_neptune_rt_call_run_3d(void *call, const neptune_rt_memref_3d *result);

#ifdef __cplusplus
}
#endif
#endif /* NEPTUNE_RT_H */
