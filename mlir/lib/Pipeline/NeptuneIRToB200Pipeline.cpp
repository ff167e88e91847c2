//===- NeptuneIRToB200Pipeline.cpp - the `neptuneir-to-b200` pipeline ---------------------===//
//
// Shape: lib/Pipeline/NeptuneIRPassesPipeline.cpp:5-53 of the reference (buildNeptuneToLLVMPipeline +
// registerNeptunePipelines).  Differences: no runtime-lowering (PETSc) stage, and the dataflow stage is
// `neptune-ir-apply-to-b200`: an apply becomes runtime calls, not loops.  The SCF stage stays, because the storage
// patterns shared with the reference include ReduceLowering, which builds scf.for (DataflowLowering.cpp:608-637).
//
// Call registerNeptuneB200Pipelines() next to registerNeptunePipelines() in src/neptuneOpt.cpp:44 and in
// NeptuneCompiler's constructor; `neptune-opt in.mlir --neptuneir-to-b200` then works like `--neptuneir-to-llvm`.
//
//===----------------------------------------------------------------------===//
#include "Passes/NeptuneIRB200Passes.h"
#include "Passes/NeptuneIRPasses.h"

#include "mlir/Conversion/ArithToLLVM/ArithToLLVM.h"
#include "mlir/Conversion/ControlFlowToLLVM/ControlFlowToLLVM.h"
#include "mlir/Conversion/FuncToLLVM/ConvertFuncToLLVMPass.h"
#include "mlir/Conversion/MemRefToLLVM/MemRefToLLVM.h"
#include "mlir/Conversion/ReconcileUnrealizedCasts/ReconcileUnrealizedCasts.h"
#include "mlir/Conversion/SCFToControlFlow/SCFToControlFlow.h"
#include "mlir/Transforms/Passes.h"

using namespace mlir;

void mlir::Neptune::NeptuneIR::buildNeptuneToB200Pipeline(OpPassManager &pm, const NeptuneToB200PipelineOptions &options) {
  using namespace mlir::Neptune::NeptuneIR;

  // 0) the reference's own front passes: verifier rules + structure keys, opdef/apply_linear -> func/call
  pm.addPass(createNeptuneIRVerifyAnnotatePass());
  pm.addPass(createNeptuneIRStructureLoweringPass());

  // 1) the sweep: apply -> neptune_rt_* calls (wrap/unwrap/load/store/reduce keep the reference's lowerings)
  {
    NeptuneIRApplyToB200PassOptions opt;
    opt.fma = options.fma;
    pm.addPass(createNeptuneIRApplyToB200Pass(opt));
  }
  pm.addPass(createCanonicalizerPass());
  pm.addPass(createCSEPass());

  // 2) what is left is func / arith / memref / llvm.global, plus scf.for where the module held a neptune_ir.reduce:
  //    the tail of buildNeptuneToLLVMPipeline (:28-46), SCF stage included
  pm.addPass(createSCFToControlFlowPass());
  pm.addPass(createArithToLLVMConversionPass());
  pm.addPass(createConvertControlFlowToLLVMPass());
  pm.addPass(createFinalizeMemRefToLLVMConversionPass());
  pm.addPass(createConvertFuncToLLVMPass());   // honours llvm.emit_c_interface on the neptune_rt_call_* declarations
  pm.addPass(createReconcileUnrealizedCastsPass());
  pm.addPass(createCanonicalizerPass());
  pm.addPass(createSymbolDCEPass());
}

void mlir::Neptune::NeptuneIR::registerNeptuneB200Pipelines() {
  PassPipelineRegistration<NeptuneToB200PipelineOptions>(
      "neptuneir-to-b200", "Lower NeptuneIR stencil sweeps to calls into the B200 runtime (neptune_rt_*), the rest to LLVM.",
      mlir::Neptune::NeptuneIR::buildNeptuneToB200Pipeline);
}
