//===- ApplyToB200.cpp - neptune_ir.apply -> B200 runtime calls ----------------------------===//
//
// The MLIR half of the `neptuneir-to-b200` path.  It has the shape of the reference's dataflow lowering
// (lib/Passes/DataflowLowering.cpp): the same type converter and the same wrap / unwrap / load / store / reduce
// patterns, with ONE pattern exchanged — where ApplyToSCFForLowering (:258-448) builds an scf.for nest, the pattern
// below describes the apply to the runtime and calls it, the way RuntimeLowering.cpp:39-53,169-178 emits calls to
// `_neptune_rt_runtime_*`.
//
// Everything that knows about stencils lives MLIR-free in the runtime library (validation, structure keys, kernel
// selection, NVRTC for bodies outside the prebuilt catalogue); the only MLIR-dependent code is the walk that turns
// the region's operations into `neptune_rt_body_op`s (collectBody) and the IR that is emitted (emitRuntimeCalls).
//
// NOT COMPILED IN THIS REPOSITORY'S CI: the image has no MLIR.  Written against the MLIR API
// level of the reference (llvm-project release/21.x); mlir/CMakeLists.txt builds it where find_package(MLIR) succeeds.
//
// Two functions are expected from the reference tree. DataflowLowering.cpp keeps its type converter and patterns in an
// anonymous namespace today; mlir/patches/0001-export-dataflow-patterns.patch (pure additions, `git apply` it in the
// reference checkout; tests/test_mlir_sources.py checks that it still applies) exports them through a new header:
//     std::unique_ptr<TypeConverter> createNeptuneIRTypeConverter(MLIRContext *);    (the class at :51-71)
//     void populateNeptuneIRStoragePatterns(RewritePatternSet &, TypeConverter &);   (the list at :790-797)
//
//===----------------------------------------------------------------------===//
#include "Dialect/NeptuneIR/NeptuneIRDialect.h"
#include "Dialect/NeptuneIR/NeptuneIROps.h"
#include "Passes/NeptuneIRB200Passes.h"
#include "Passes/NeptuneIRDataflowPatterns.h"   // createNeptuneIRTypeConverter, populateNeptuneIRStoragePatterns

#include "mlir/Dialect/Arith/IR/Arith.h"
#include "mlir/Dialect/Func/IR/FuncOps.h"
#include "mlir/Dialect/Func/Transforms/FuncConversions.h"
#include "mlir/Dialect/LLVMIR/LLVMDialect.h"
#include "mlir/Dialect/MemRef/IR/MemRef.h"
#include "mlir/Dialect/SCF/IR/SCF.h"
#include "mlir/IR/BuiltinOps.h"
#include "mlir/Transforms/DialectConversion.h"

#include "neptune_rt.h"   // include/neptune_rt.h of the B200 runtime: descriptors, neptune_rt_desc_pack, neptune_rt_body_key

#include <string>
#include <vector>

using namespace mlir;
using namespace mlir::Neptune::NeptuneIR;

namespace mlir::Neptune::NeptuneIR {
#define GEN_PASS_DEF_NEPTUNEIRAPPLYTOB200PASS
#include "Passes/NeptuneIRB200Passes.h.inc"
} // namespace mlir::Neptune::NeptuneIR

namespace {

//===----------------------------------------------------------------------===//
// The adapter: region operations -> neptune_rt_body_op[]  (the only walk over MLIR operations)
//===----------------------------------------------------------------------===//
/// Fills `ops` with the body in program order; value number i is the result of ops[i].  Accepts exactly what the
/// MLIR-free recogniser accepts (csrc/frontend/recognizer.cpp): neptune_ir.access on a block argument with constant
/// offsets, float arith.constant, addf / subf / mulf / divf / negf.  Anything else (index arithmetic, integer math,
/// scf.if — classes P/X of SURVEY.md Appendix C) fails with a message naming the operation.
static LogicalResult collectBody(ApplyOp op, std::vector<neptune_rt_body_op> &ops, int32_t &yieldValue, std::string &why) {
  Block &body = op.getBody().front();
  const unsigned rank = op.getBounds().getLb().size();
  const unsigned nIn = op.getInputs().size();
  if (body.getNumArguments() != rank + nIn) {
    why = "apply-like region block arg count must be (bounds rank + number of inputs)";
    return failure();
  }
  llvm::DenseMap<Value, int32_t> number;
  auto push = [&](neptune_rt_body_op o, Value result) {
    number[result] = static_cast<int32_t>(ops.size());
    ops.push_back(o);
  };
  auto operand = [&](Value v, int32_t &out) -> bool {
    auto it = number.find(v);
    if (it == number.end()) return false;
    out = it->second;
    return true;
  };
  for (Operation &nested : body.without_terminator()) {
    neptune_rt_body_op o{};
    o.a = o.b = o.input = -1;
    if (auto acc = dyn_cast<AccessOp>(nested)) {
      auto arg = dyn_cast<BlockArgument>(acc.getInput());
      if (!arg || arg.getOwner() != &body || arg.getArgNumber() < rank) {
        why = "neptune_ir.access must read one of the apply's input block arguments";
        return failure();
      }
      o.opcode = NEPTUNE_RT_OP_ACCESS;
      o.input = static_cast<int32_t>(arg.getArgNumber() - rank);
      ArrayRef<int64_t> offs = acc.getOffsetsAttr().asArrayRef();
      if (offs.size() != rank) { why = "access offset rank differs from the apply's rank"; return failure(); }
      for (unsigned d = 0; d < rank; ++d) o.offset[d] = offs[d];
      push(o, acc.getResult());
      continue;
    }
    if (auto cst = dyn_cast<arith::ConstantOp>(nested)) {
      auto f = dyn_cast<FloatAttr>(cst.getValue());
      if (!f) { why = "integer/index constants in an apply body are outside the recognised stencil set"; return failure(); }
      o.opcode = NEPTUNE_RT_OP_CONST;
      o.value = f.getValueAsDouble();   // exact: f32 values are representable as doubles
      push(o, cst.getResult());
      continue;
    }
    auto binary = [&](int32_t opcode) -> LogicalResult {
      o.opcode = opcode;
      if (!operand(nested.getOperand(0), o.a) || !operand(nested.getOperand(1), o.b)) {
        why = "operand of '" + nested.getName().getStringRef().str() + "' is not a value of the body (index arguments cannot be used)";
        return failure();
      }
      push(o, nested.getResult(0));
      return success();
    };
    if (isa<arith::AddFOp>(nested)) { if (failed(binary(NEPTUNE_RT_OP_ADD))) return failure(); continue; }
    if (isa<arith::SubFOp>(nested)) { if (failed(binary(NEPTUNE_RT_OP_SUB))) return failure(); continue; }
    if (isa<arith::MulFOp>(nested)) { if (failed(binary(NEPTUNE_RT_OP_MUL))) return failure(); continue; }
    if (isa<arith::DivFOp>(nested)) { if (failed(binary(NEPTUNE_RT_OP_DIV))) return failure(); continue; }
    if (isa<arith::NegFOp>(nested)) {
      o.opcode = NEPTUNE_RT_OP_NEG;
      if (!operand(nested.getOperand(0), o.a)) { why = "operand of arith.negf is not a value of the body"; return failure(); }
      push(o, nested.getResult(0));
      continue;
    }
    why = "operation '" + nested.getName().getStringRef().str() +
          "' in an apply body is outside the recognised stencil set (straight-line float arithmetic over neptune_ir.access); "
          "there is no CPU fallback in neptuneir-to-b200";
    return failure();
  }
  auto y = cast<YieldOp>(body.getTerminator());
  if (y.getNumOperands() != 1 || !operand(y.getOperand(0), yieldValue)) {
    why = "neptune_ir.yield must return exactly one scalar computed by the body";   // DataflowLowering.cpp:427-429
    return failure();
  }
  return success();
}

//===----------------------------------------------------------------------===//
// Emission helpers (mechanism of RuntimeLowering.cpp:39-84)
//===----------------------------------------------------------------------===//
static func::FuncOp getOrCreateRuntimeFunc(Operation *anchor, StringRef name, ArrayRef<Type> inputs, ArrayRef<Type> results,
                                           PatternRewriter &rewriter) {
  ModuleOp m = anchor->getParentOfType<ModuleOp>();
  if (auto f = m.lookupSymbol<func::FuncOp>(name)) return f;
  OpBuilder::InsertionGuard g(rewriter);
  rewriter.setInsertionPointToStart(m.getBody());
  auto fn = rewriter.create<func::FuncOp>(m.getLoc(), name, FunctionType::get(m.getContext(), inputs, results));
  fn.setPrivate();
  // memref arguments reach the runtime as pointers to their descriptors (`_mlir_ciface_<name>`, exported by the runtime)
  fn->setAttr(LLVM::LLVMDialect::getEmitCWrapperAttrName(), UnitAttr::get(m.getContext()));
  return fn;
}

/// The packed descriptor as an internal constant byte array; returns its address.
static Value getOrCreateDescGlobal(Operation *anchor, PatternRewriter &rewriter, Location loc, StringRef name, StringRef bytes) {
  auto module = anchor->getParentOfType<ModuleOp>();
  MLIRContext *ctx = rewriter.getContext();
  auto arrTy = LLVM::LLVMArrayType::get(IntegerType::get(ctx, 8), bytes.size());
  if (!module.lookupSymbol<LLVM::GlobalOp>(name)) {
    OpBuilder::InsertionGuard g(rewriter);
    rewriter.setInsertionPointToStart(module.getBody());
    rewriter.create<LLVM::GlobalOp>(loc, arrTy, /*isConstant=*/true, LLVM::Linkage::Internal, name, rewriter.getStringAttr(bytes),
                                    /*alignment=*/8);
  }
  return rewriter.create<LLVM::AddressOfOp>(loc, LLVM::LLVMPointerType::get(ctx), name);
}

//===----------------------------------------------------------------------===//
// The pattern
//===----------------------------------------------------------------------===//
struct ApplyToB200Lowering : OpConversionPattern<ApplyOp> {
  ApplyToB200Lowering(TypeConverter &tc, MLIRContext *ctx, bool fma) : OpConversionPattern<ApplyOp>(tc, ctx), fma(fma) {}
  bool fma;

  LogicalResult matchAndRewrite(ApplyOp op, OpAdaptor adaptor, ConversionPatternRewriter &rewriter) const override {
    Location loc = op.getLoc();
    ArrayRef<int64_t> lbs = op.getBounds().getLb().asArrayRef();
    ArrayRef<int64_t> ubs = op.getBounds().getUb().asArrayRef();
    const unsigned rank = lbs.size();
    if (rank < 1 || rank > NEPTUNE_RT_MAX_RANK) return op.emitOpError("neptuneir-to-b200 lowers rank 1..3 applies");
    const unsigned nIn = adaptor.getOperands().size();
    if (nIn < 1 || nIn > NEPTUNE_RT_MAX_INPUTS) return op.emitOpError("neptuneir-to-b200 takes 1..32 apply inputs");
    auto outTemp = cast<TempType>(op.getResult().getType());
    Type elem = outTemp.getElementType();
    if (!elem.isF64() && !elem.isF32()) return op.emitOpError("neptuneir-to-b200 lowers f32 / f64 applies");
    auto resTy = dyn_cast<MemRefType>(getTypeConverter()->convertType(op.getResult().getType()));
    if (!resTy) return rewriter.notifyMatchFailure(op, "result not converted to memref");

    // 1. describe the apply
    std::vector<neptune_rt_body_op> ops;
    int32_t yieldValue = -1;
    std::string why;
    if (failed(collectBody(op, ops, yieldValue, why))) return op.emitOpError(why);
    std::vector<int64_t> inLb(nIn * rank), inUb(nIn * rank);
    for (unsigned k = 0; k < nIn; ++k) {
      auto t = cast<TempType>(op.getInputs()[k].getType());
      if (t.getElementType() != elem) return op.emitOpError("apply inputs must share the result's element type");
      for (unsigned d = 0; d < rank; ++d) {
        inLb[k * rank + d] = t.getBounds().getLb()[d];
        inUb[k * rank + d] = t.getBounds().getUb()[d];
      }
    }
    neptune_rt_stencil_desc desc{};
    desc.rank = rank;
    desc.dtype = elem.isF64() ? NEPTUNE_RT_F64 : NEPTUNE_RT_F32;
    desc.n_inputs = nIn;
    desc.n_ops = static_cast<int32_t>(ops.size());
    desc.ops = ops.data();
    desc.yield_value = yieldValue;
    desc.flags = fma ? NEPTUNE_RT_FLAG_FMA : 0;
    for (unsigned d = 0; d < rank; ++d) {
      desc.lb[d] = lbs[d];
      desc.ub[d] = ubs[d];
      desc.out_lb[d] = outTemp.getBounds().getLb()[d];
      desc.out_ub[d] = outTemp.getBounds().getUb()[d];
    }
    desc.in_lb = inLb.data();
    desc.in_ub = inUb.data();
    // taps must stay inside their input's box: undefined behaviour in the reference, an error here
    for (const neptune_rt_body_op &o : ops)
      if (o.opcode == NEPTUNE_RT_OP_ACCESS)
        for (unsigned d = 0; d < rank; ++d)
          if (lbs[d] < ubs[d] && (lbs[d] + o.offset[d] < inLb[o.input * rank + d] || ubs[d] + o.offset[d] > inUb[o.input * rank + d]))
            return op.emitOpError("a neptune_ir.access reads outside its input's bounds for some point of the apply's bounds");

    // 2. pack it into a constant global
    const int64_t need = neptune_rt_desc_pack(&desc, nullptr, 0);
    if (need <= 0) return op.emitOpError("the apply could not be described to the B200 runtime");
    std::string blob(static_cast<size_t>(need), '\0');
    neptune_rt_desc_pack(&desc, blob.data(), need);
    // named by content (structure key + FNV-1a of the packed bytes): no counter shared between pattern applications,
    // the same name whatever order the applies are visited in, and identical descriptors share one global
    uint64_t digest = 1469598103934665603ull;
    for (unsigned char ch : blob) { digest ^= ch; digest *= 1099511628211ull; }
    char name[80];
    std::snprintf(name, sizeof name, "__neptune_b200_desc_%016llx_%016llx", (unsigned long long)neptune_rt_body_key(&desc),
                  (unsigned long long)digest);
    Value descPtr = getOrCreateDescGlobal(op, rewriter, loc, name, blob);

    // 3. the calls
    MLIRContext *ctx = rewriter.getContext();
    Type ptrTy = LLVM::LLVMPointerType::get(ctx);
    Type i64 = rewriter.getI64Type(), i32 = rewriter.getI32Type();
    std::string sfx = "_" + std::to_string(rank) + "d";
    auto begin = getOrCreateRuntimeFunc(op, "neptune_rt_call_begin", {ptrTy, i64}, {ptrTy}, rewriter);
    Value bytes = rewriter.create<arith::ConstantIntOp>(loc, need, 64);
    Value call = rewriter.create<func::CallOp>(loc, begin, ValueRange{descPtr, bytes}).getResult(0);
    for (unsigned k = 0; k < nIn; ++k) {
      Value in = adaptor.getOperands()[k];
      auto inTy = cast<MemRefType>(in.getType());
      // one declaration per operand type: shapes are static, and func.func has no type polymorphism
      std::string fname = "neptune_rt_call_input" + sfx;
      auto fin = getOrCreateRuntimeFunc(op, fname, {ptrTy, i32, UnrankedOrRanked(inTy)}, {}, rewriter);
      Value kv = rewriter.create<arith::ConstantIntOp>(loc, k, 32);
      rewriter.create<func::CallOp>(loc, fin, ValueRange{call, kv, castForCall(rewriter, loc, in)});
    }
    auto out = rewriter.create<memref::AllocOp>(loc, resTy);   // as the reference: a fresh result (:273-281)
    auto frun = getOrCreateRuntimeFunc(op, "neptune_rt_call_run" + sfx, {ptrTy, UnrankedOrRanked(resTy)}, {}, rewriter);
    rewriter.create<func::CallOp>(loc, frun, ValueRange{call, castForCall(rewriter, loc, out)});
    rewriter.replaceOp(op, out.getResult());
    return success();
  }

private:
  /// The runtime takes fully dynamic memrefs of the apply's rank (memref<?x?x?xf64>): one C signature per rank.
  static MemRefType UnrankedOrRanked(MemRefType t) {
    SmallVector<int64_t> dyn(t.getRank(), ShapedType::kDynamic);
    return MemRefType::get(dyn, t.getElementType());
  }
  static Value castForCall(PatternRewriter &rewriter, Location loc, Value v) {
    auto t = cast<MemRefType>(v.getType());
    MemRefType dyn = UnrankedOrRanked(t);
    return t == dyn ? v : rewriter.create<memref::CastOp>(loc, dyn, v).getResult();
  }
};

//===----------------------------------------------------------------------===//
// The pass (shape of NeptuneIRDataflowLoweringPass::runOnOperation, DataflowLowering.cpp:739-819)
//===----------------------------------------------------------------------===//
struct NeptuneIRApplyToB200Pass final : public mlir::Neptune::NeptuneIR::impl::NeptuneIRApplyToB200PassBase<NeptuneIRApplyToB200Pass> {
  using Base::Base;

  void runOnOperation() override {
    ModuleOp module = getOperation();
    MLIRContext *ctx = module.getContext();
    std::unique_ptr<TypeConverter> converterOwner = createNeptuneIRTypeConverter(ctx);
    TypeConverter &typeConverter = *converterOwner;

    ConversionTarget target(*ctx);
    target.addLegalDialect<BuiltinDialect, arith::ArithDialect, memref::MemRefDialect, scf::SCFDialect, func::FuncDialect,
                           LLVM::LLVMDialect>();   // scf: what ReduceLowering emits; the pipeline lowers it afterwards
    target.addIllegalDialect<NeptuneIRDialect>();
    target.markUnknownOpDynamicallyLegal([&](Operation *op) { return typeConverter.isLegal(op); });
    target.addDynamicallyLegalOp<func::FuncOp>([&](func::FuncOp f) {
      return typeConverter.isSignatureLegal(f.getFunctionType()) && typeConverter.isLegal(&f.getBody());
    });
    target.addDynamicallyLegalOp<func::CallOp>([&](func::CallOp op) {
      return typeConverter.isLegal(op.getOperandTypes()) && typeConverter.isLegal(op.getResultTypes());
    });
    target.addDynamicallyLegalOp<func::ReturnOp>([&](func::ReturnOp r) { return typeConverter.isLegal(r.getOperandTypes()); });

    RewritePatternSet patterns(ctx);
    populateFunctionOpInterfaceTypeConversionPattern<func::FuncOp>(patterns, typeConverter);
    populateReturnOpTypeConversionPattern(patterns, typeConverter);
    populateCallOpTypeConversionPattern(patterns, typeConverter);
    // wrap / unwrap / load / store / reduce: the reference's own lowerings (DataflowLowering.cpp:131-220, :589-698)
    populateNeptuneIRStoragePatterns(patterns, typeConverter);
    // the sweep
    patterns.add<ApplyToB200Lowering>(typeConverter, ctx, fma);

    // (operator-call + AXPY fusion is a planner step of the MLIR-free front end, neptune_b200_compile; here every apply
    //  is lowered on its own, so the pass has no `fuse` option)
    if (failed(applyFullConversion(module, target, std::move(patterns)))) signalPassFailure();
  }
};

} // namespace
