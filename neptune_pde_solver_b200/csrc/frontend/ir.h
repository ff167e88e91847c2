// In-memory form of the NeptuneIR subset that reaches the `neptune_ir.apply` hot path.
// Mirrors the reference's ODS definitions (no MLIR dependency):
//   types  !neptune_ir.field / !neptune_ir.temp      include/Dialect/NeptuneIR/NeptuneIRTypes.td:12-59
//   attrs  #neptune_ir.bounds / location / stencil_shape   NeptuneIRAttrs.td:9-76
//   ops    wrap/unwrap/load/yield/apply/access/store        NeptuneIROps.td:20-259
#pragma once
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "neptune_rt.h"

namespace neptune {

struct Diagnostic : std::runtime_error {
  bool unsupported;  // true: well-formed IR outside the recognised set (no CPU fallback)
  Diagnostic(const std::string &m, bool unsup = false) : std::runtime_error(m), unsupported(unsup) {}
};

struct Bounds {
  std::vector<int64_t> lb, ub;
  int rank() const { return (int)lb.size(); }
  int64_t extent(int d) const { return ub[d] - lb[d]; }
  int64_t numel() const {
    int64_t n = 1;
    for (int d = 0; d < rank(); ++d) n *= extent(d);
    return n;
  }
  bool operator==(const Bounds &o) const { return lb == o.lb && ub == o.ub; }
};

struct Type {
  enum Kind { None, Float, Index, Int, Grid, MemRef } kind = None;
  int dtype = -1;               // NEPTUNE_RT_F32/F64 for Float, Grid, MemRef
  bool is_field = false;        // Grid
  Bounds bounds;                // Grid
  std::string location;         // Grid
  std::vector<int64_t> shape;   // MemRef, -1 = dynamic
  bool operator==(const Type &o) const {
    return kind == o.kind && dtype == o.dtype && is_field == o.is_field && bounds == o.bounds &&
           location == o.location && shape == o.shape;
  }
  bool operator!=(const Type &o) const { return !(*this == o); }
  std::string str() const;
};

struct Apply {
  std::string result;
  std::vector<std::string> inputs;
  std::vector<Type> in_types;
  Type out_type;
  Bounds bounds;
  bool has_shape = false;
  std::vector<std::vector<int64_t>> shape;   // stencil_shape offsets
  std::vector<neptune_rt_body_op> body;      // value-numbered, program order
  std::vector<int> body_dtype;               // element type of each body value
  int yield_value = -1;
  int yield_dtype = -1;
  int line = 0;
};

struct HostOp {
  enum Kind { Wrap, Unwrap, Load, Store, ApplyOp, ApplyLinear, Return, Constant, Reduce } kind;
  std::vector<std::string> results, operands;
  std::vector<Type> types;
  std::shared_ptr<Apply> apply;
  std::string callee;
  bool has_bounds = false;
  Bounds bounds;  // store, reduce (callee holds the reduce kind)
  int line = 0;
};

struct Function {
  std::string name;
  std::string kind;  // "func" | "linear_opdef" | "nonlinear_opdef"
  std::vector<std::pair<std::string, Type>> args;
  std::vector<Type> results;
  std::vector<HostOp> ops;
  // a function the pipeline could not lower (parser.cpp parse_function_body, module.cpp): kept with its diagnostic
  bool rejected = false, unsupported = false;
  std::string diagnostic;
};

struct Module {
  std::vector<Function> functions;  // program order
  const Function *find(const std::string &name) const {
    for (auto &f : functions)
      if (f.name == name) return &f;
    return nullptr;
  }
};

// parser.cpp
Module parse_module(const std::string &text);

// Owning stencil descriptor (the C struct points into the vectors).
struct StencilDesc {
  neptune_rt_stencil_desc c{};
  std::vector<neptune_rt_body_op> ops;
  std::vector<int64_t> in_lb, in_ub;
  StencilDesc() = default;
  StencilDesc(const StencilDesc &o) { *this = o; }
  StencilDesc &operator=(const StencilDesc &o) {
    c = o.c; ops = o.ops; in_lb = o.in_lb; in_ub = o.in_ub; fix();
    return *this;
  }
  void fix() { c.ops = ops.data(); c.n_ops = (int)ops.size(); c.in_lb = in_lb.data(); c.in_ub = in_ub.data(); }
  static StencilDesc from_c(const neptune_rt_stencil_desc &d);
};

// recognizer.cpp — apply -> descriptor; throws Diagnostic for anything outside the recognised set
StencilDesc recognize_apply(const Apply &ap, bool fma);
void validate_desc(const neptune_rt_stencil_desc &d);   // throws Diagnostic
uint64_t body_key(const neptune_rt_stencil_desc &d);

// Per-body facts the kernels are specialised on (canonical z,y,x dims; see codegen.cpp).
struct BodyInfo {
  int rank = 0, dtype = 0, n_inputs = 0, n_consts = 0;
  int radius[NEPTUNE_RT_MAX_INPUTS][3];   // per input, canonical dims (z,y,x)
  int staged = -1;                        // the one input with radius > 0 (-1: none, -2: several)
  std::vector<double> consts;             // values, in slot order
  std::vector<int> const_slot;            // per body op: slot of a CONST op, else -1
};
BodyInfo analyze_body(const neptune_rt_stencil_desc &d);
// canonical (z,y,x) offset of a logical offset vector: rank 3 (d0,d1,d2); rank 2 (0,d0,d1); rank 1 (0,0,d0)
void canonical_offset(int rank, const int64_t *logical, int out_zyx[3]);

// fusion.cpp — fold `%t = A(%x); %y = apply(%x, %t){.. %t[0] ..}` into one apply (bit-identical per point).
// accept(key): whether a kernel exists for the fused body (NULL: always fuse).  Returns the number of fusions.
int fuse_applies(Module &m, bool fma, const std::function<bool(uint64_t)> &accept);

// codegen.cpp — CUDA text of the point functor for a body
std::string functor_name(uint64_t key);
std::string generate_functor(const neptune_rt_stencil_desc &d);
std::string generate_registration_unit(const std::vector<const neptune_rt_stencil_desc *> &bodies);

}  // namespace neptune
