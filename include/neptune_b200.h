/* neptune_b200.h — C ABI of the `neptuneir-to-b200` pipeline front end.
 *
 * The reference drives its passes through `neptune-opt` (src/neptuneOpt.cpp:38-47) and
 * `NeptuneCompiler::runPipeline` (lib/Compiler/NeptuneCompiler.cpp:291-301) with the pipeline
 * registered in lib/Pipeline/NeptuneIRPassesPipeline.cpp:49-53 (`neptuneir-to-llvm`).  Those need
 * MLIR/LLVM 21, which this image does not have, so the same surface is offered MLIR-free:
 * textual NeptuneIR in, an executable module out.  The pipeline
 *
 *     parse -> ApplyOp::verify rules -> recognise apply bodies -> plan runtime calls
 *
 * replaces `verify-annotate -> structure-lowering -> dataflow-lowering{backend=cpu}`
 * (NeptuneIRPassesPipeline.cpp:10-26) for modules made of wrap/unwrap/load/store/apply/access/
 * yield (+ linear_opdef/apply_linear as callers).  Unrecognised bodies are rejected with a
 * diagnostic: there is no CPU fallback.  Everything in this header except *_run* is pure host
 * code and works without a GPU.
 */
#ifndef NEPTUNE_B200_H
#define NEPTUNE_B200_H

#include "neptune_rt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct neptune_b200_module neptune_b200_module;

/* Parse + verify + recognise. `pipeline` must be "neptuneir-to-b200" (options may follow in
 * braces: "neptuneir-to-b200{fma=true}").  Every function of the module is lowered on its own: one that
 * holds something outside the hot path (neptune_ir.time_advance, a body with control flow, ...) or breaks a
 * verifier rule is kept as a rejected function with its diagnostic (neptune_b200_function_diagnostic) and the
 * others — the operator definitions a time_advance refers to, for one — are still lowered, so the reference's
 * own fixtures (test/smoke_tests/smoke_time_advance.mlir) compile as far as they are on the path.  When NO
 * function lowers the call fails like a failed conversion: NEPTUNE_RT_EINVAL / NEPTUNE_RT_EUNSUPPORTED with the
 * first diagnostic in err (NUL-terminated, truncated). */
int neptune_b200_compile(const char *mlir_text, const char *pipeline, neptune_b200_module **module,
                         char *err, size_t err_len);
int neptune_b200_module_destroy(neptune_b200_module *module);

/* Introspection: the applies the module contains, in program order over all functions. */
int neptune_b200_apply_count(neptune_b200_module *module);
/* Borrowed pointer, valid until the module is destroyed. */
const neptune_rt_stencil_desc *neptune_b200_apply_desc(neptune_b200_module *module, int index);
/* Name of the function holding apply `index` ("@entry"). */
const char *neptune_b200_apply_function(neptune_b200_module *module, int index);
/* Text of the runtime-call plan of a function (one line per neptune_rt_* call), for tests and
 * for reading what the pipeline did — the analogue of the LLVM-dialect text the reference's
 * pipeline prints. Borrowed pointer. */
const char *neptune_b200_function_plan(neptune_b200_module *module, const char *function);
/* Number / element counts of a function's memref arguments (-1 if unknown function). */
int neptune_b200_function_arg_count(neptune_b200_module *module, const char *function);
int64_t neptune_b200_function_arg_numel(neptune_b200_module *module, const char *function, int arg);

/* Every function of the module, in program order, and why one was not lowered (NULL: it was). Borrowed pointers. */
int neptune_b200_function_count(neptune_b200_module *module);
const char *neptune_b200_function_name(neptune_b200_module *module, int index);
const char *neptune_b200_function_diagnostic(neptune_b200_module *module, const char *function);
/* Message of the last failed call on this module. Borrowed pointer. */
const char *neptune_b200_last_error(neptune_b200_module *module);
/* The runtime stencil apply `index` runs with on `ctx` (created on first use), for tuning through
 * neptune_rt_stencil_set and for driving neptune_rt_run_steps directly. Owned by the module. */
neptune_rt_stencil *neptune_b200_apply_stencil(neptune_b200_module *module, neptune_rt_ctx *ctx, int index);

/* CUDA source of the point functor generated for apply `index` (what the build compiles into the
 * catalogue). Borrowed pointer. */
const char *neptune_b200_apply_cuda_source(neptune_b200_module *module, int index);

/* Execute `function` the way the reference's compiled entry would be called by a host driver
 * (test/smoke_tests/smoke_apply.sh:39-80): args[i] are host pointers to the contiguous row-major
 * memref arguments; `wrap` aliases them, `store`/`unwrap` make results visible in them.
 * Uploads, runs every apply on the GPU, downloads what was written.
 * An operator definition (linear_opdef / nonlinear_opdef, which StructureLowering.cpp:31-140 turns into a
 * func.func) is called with its temp operands' buffers followed by ONE more buffer that receives the result. */
int neptune_b200_run_host(neptune_b200_module *module, neptune_rt_ctx *ctx, const char *function,
                          int n_args, void *const *args);
/* Scalar values (results of `neptune_ir.reduce`, NeptuneIROps.td:272-299) the function returned in
 * the last neptune_b200_run_host, in return-operand order. Writes at most `capacity` of them to
 * `out` and returns how many there are. */
int neptune_b200_last_scalars(neptune_b200_module *module, double *out, int capacity);
/* The host time loop "call entry(out, in, extra...) then swap out/in", nsteps times, kept
 * resident on the device: one upload, nsteps sweeps, one download. On return args[1] holds the
 * final state if nsteps is even, args[0] if odd (exactly what the swapped host loop leaves), and
 * *result_arg says which. */
int neptune_b200_run_steps_host(neptune_b200_module *module, neptune_rt_ctx *ctx,
                                const char *function, int n_args, void *const *args,
                                int64_t nsteps, int *result_arg);

#ifdef __cplusplus
}
#endif
#endif /* NEPTUNE_B200_H */
