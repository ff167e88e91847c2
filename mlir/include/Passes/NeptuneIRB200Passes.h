//===- NeptuneIRB200Passes.h - declarations of the B200 pass and pipeline -------*- C++ -*-===//
//
// Counterpart of include/Passes/NeptuneIRPasses.h:29-33 and include/Passes/NeptuneIRPassesPipeline.h of the
// reference for the one pass / one pipeline the B200 path adds.
//
//===----------------------------------------------------------------------===//
#ifndef NEPTUNEIR_B200_PASSES_H
#define NEPTUNEIR_B200_PASSES_H

#include "mlir/Pass/Pass.h"
#include "mlir/Pass/PassManager.h"
#include "mlir/Pass/PassRegistry.h"

#include "Dialect/NeptuneIR/NeptuneIRAttrs.h"
#include "Dialect/NeptuneIR/NeptuneIROps.h"

namespace mlir::Neptune::NeptuneIR {

#define GEN_PASS_DECL
#include "Passes/NeptuneIRB200Passes.h.inc"

#define GEN_PASS_REGISTRATION
#include "Passes/NeptuneIRB200Passes.h.inc"

/// verify-annotate -> structure-lowering -> apply-to-b200 -> scf-to-cf -> func/arith/cf/memref to LLVM.
/// Registered as `neptuneir-to-b200` by registerNeptuneB200Pipelines().
struct NeptuneToB200PipelineOptions : public PassPipelineOptions<NeptuneToB200PipelineOptions> {
  Option<bool> fma{*this, "fma", llvm::cl::desc("allow FMA contraction in the sweep kernels"), llvm::cl::init(false)};
  // (no `fuse` option here: operator-call + AXPY fusion is a planner step of the MLIR-free front end,
  //  neptune_b200_compile("neptuneir-to-b200{fuse=...}"); at the MLIR level every apply is lowered on its own)
};
void buildNeptuneToB200Pipeline(OpPassManager &pm, const NeptuneToB200PipelineOptions &options);
void registerNeptuneB200Pipelines();

} // namespace mlir::Neptune::NeptuneIR

#endif // NEPTUNEIR_B200_PASSES_H
