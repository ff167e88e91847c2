// Stencil recogniser of the `neptuneir-to-b200` pipeline: neptune_ir.apply -> neptune_rt_stencil_desc.
//
// The reference lowers ANY apply body by cloning its ops into an scf.for nest
// (lib/Passes/DataflowLowering.cpp:310-444).  The B200 path accepts the bodies whose value is a
// straight-line float expression over constant-offset neighbour reads — the linear whitelist of
// lib/Passes/VerifyAndAnnotate.cpp:167-204 extended with products of two accesses (variable
// coefficients) — and keeps the body's op list verbatim, so the kernels evaluate exactly the
// reference's expression DAG in the reference's order (bit-exact tier).
// Everything else (index-dependent bodies, control flow, integer bodies) is rejected with a
// diagnostic; there is no CPU fallback.
#include <algorithm>
#include <cstring>
#include <set>
#include <sstream>

#include "ir.h"

namespace neptune {

StencilDesc StencilDesc::from_c(const neptune_rt_stencil_desc &d) {
  StencilDesc s;
  s.c = d;
  s.ops.assign(d.ops, d.ops + d.n_ops);
  s.in_lb.assign(d.in_lb, d.in_lb + (size_t)d.n_inputs * d.rank);
  s.in_ub.assign(d.in_ub, d.in_ub + (size_t)d.n_inputs * d.rank);
  s.fix();
  return s;
}

void canonical_offset(int rank, const int64_t *logical, int out[3]) {
  out[0] = out[1] = out[2] = 0;
  for (int d = 0; d < rank; ++d) out[3 - rank + d] = (int)logical[d];
}

void validate_desc(const neptune_rt_stencil_desc &d) {
  if (d.rank < 1 || d.rank > NEPTUNE_RT_MAX_RANK)
    throw Diagnostic("apply rank " + std::to_string(d.rank) + " is outside the B200 path (rank 1..3)", true);
  if (d.dtype != NEPTUNE_RT_F32 && d.dtype != NEPTUNE_RT_F64) throw Diagnostic("element type must be f32 or f64", true);
  if (d.n_inputs < 1) throw Diagnostic("apply needs at least one input (input 0 is copied through, DataflowLowering.cpp:282-287)");
  if (d.n_inputs > NEPTUNE_RT_MAX_INPUTS)
    throw Diagnostic("apply has " + std::to_string(d.n_inputs) + " inputs; the B200 path supports at most " +
                         std::to_string(NEPTUNE_RT_MAX_INPUTS), true);
  if (d.n_ops < 1 || !d.ops) throw Diagnostic("empty apply body");
  if (d.yield_value < 0 || d.yield_value >= d.n_ops) throw Diagnostic("yield refers to an undefined value");
  if (!d.in_lb || !d.in_ub) throw Diagnostic("descriptor lacks input bounds");
  int nconst = 0;
  for (int i = 0; i < d.n_ops; ++i) {
    const neptune_rt_body_op &o = d.ops[i];
    switch (o.opcode) {
      case NEPTUNE_RT_OP_ACCESS:
        if (o.input < 0 || o.input >= d.n_inputs) throw Diagnostic("access refers to an undefined input");
        break;
      case NEPTUNE_RT_OP_CONST: ++nconst; break;
      case NEPTUNE_RT_OP_ADD: case NEPTUNE_RT_OP_SUB: case NEPTUNE_RT_OP_MUL: case NEPTUNE_RT_OP_DIV:
        if (o.a < 0 || o.a >= i || o.b < 0 || o.b >= i) throw Diagnostic("body op uses a value before its definition");
        break;
      case NEPTUNE_RT_OP_NEG:
        if (o.a < 0 || o.a >= i) throw Diagnostic("body op uses a value before its definition");
        break;
      default: throw Diagnostic("unknown body opcode " + std::to_string(o.opcode), true);
    }
  }
  if (nconst > NEPTUNE_RT_MAX_CONSTS)
    throw Diagnostic("apply body has " + std::to_string(nconst) + " constants; the B200 path supports at most " +
                         std::to_string(NEPTUNE_RT_MAX_CONSTS), true);
  for (int a = 0; a < d.rank; ++a) {
    if (d.out_ub[a] < d.out_lb[a]) throw Diagnostic("result bounds are inverted");
    // copy-through needs input 0 and the result to have equal extents (DataflowLowering.cpp:282-287)
    if (d.in_ub[a] - d.in_lb[a] != d.out_ub[a] - d.out_lb[a])
      throw Diagnostic("input 0 and the result must have equal extents: the lowering copies input 0 through "
                       "(DataflowLowering.cpp:282-287)");
    if (d.ub[a] > d.lb[a]) {
      if (d.lb[a] < d.out_lb[a] || d.ub[a] > d.out_ub[a])
        throw Diagnostic("apply bounds write outside the result temp (undefined behaviour in the reference)");
    }
  }
  bool empty = false;
  for (int a = 0; a < d.rank; ++a) empty |= d.ub[a] <= d.lb[a];
  if (!empty) {
    // the reference emits no bounds checks (DataflowLowering.cpp:380-411): an out-of-box tap is UB there;
    // here it is a diagnostic
    for (int i = 0; i < d.n_ops; ++i) {
      const neptune_rt_body_op &o = d.ops[i];
      if (o.opcode != NEPTUNE_RT_OP_ACCESS) continue;
      for (int a = 0; a < d.rank; ++a) {
        int64_t ilb = d.in_lb[(size_t)o.input * d.rank + a], iub = d.in_ub[(size_t)o.input * d.rank + a];
        if (d.lb[a] + o.offset[a] < ilb || d.ub[a] + o.offset[a] > iub) {
          std::ostringstream m;
          m << "access of input " << o.input << " at offset [";
          for (int b = 0; b < d.rank; ++b) m << (b ? ", " : "") << o.offset[b];
          m << "] reads outside the input's bounds over the apply bounds (undefined behaviour in the reference: "
               "shrink `bounds` by the stencil radius)";
          throw Diagnostic(m.str());
        }
      }
    }
  }
}

StencilDesc recognize_apply(const Apply &ap, bool fma) {
  StencilDesc s;
  neptune_rt_stencil_desc &c = s.c;
  std::memset(&c, 0, sizeof c);
  const int rank = ap.bounds.rank();
  auto where = "apply at line " + std::to_string(ap.line) + ": ";
  if (rank < 1 || rank > NEPTUNE_RT_MAX_RANK) throw Diagnostic(where + "rank " + std::to_string(rank) + " is outside the B200 path", true);
  if (ap.out_type.kind != Type::Grid) throw Diagnostic(where + "result must be a !neptune_ir.temp");
  if (ap.out_type.bounds.rank() != rank) throw Diagnostic(where + "result rank differs from the bounds rank");
  c.rank = rank;
  c.dtype = ap.out_type.dtype;
  c.n_inputs = (int)ap.inputs.size();
  c.flags = fma ? NEPTUNE_RT_FLAG_FMA : 0;
  for (int d = 0; d < rank; ++d) {
    c.lb[d] = ap.bounds.lb[d]; c.ub[d] = ap.bounds.ub[d];
    c.out_lb[d] = ap.out_type.bounds.lb[d]; c.out_ub[d] = ap.out_type.bounds.ub[d];
  }
  for (auto &t : ap.in_types) {
    if (t.kind != Type::Grid) throw Diagnostic(where + "inputs must be !neptune_ir.temp values");
    if (t.bounds.rank() != rank) throw Diagnostic(where + "input rank differs from the bounds rank");
    if (t.dtype != c.dtype) throw Diagnostic(where + "mixed element types are outside the recognised stencil set", true);
    for (int d = 0; d < rank; ++d) { s.in_lb.push_back(t.bounds.lb[d]); s.in_ub.push_back(t.bounds.ub[d]); }
  }
  for (size_t i = 0; i < ap.body.size(); ++i)
    if (ap.body_dtype[i] != c.dtype) throw Diagnostic(where + "body value type differs from the result element type", true);
  if (ap.yield_dtype != c.dtype) throw Diagnostic(where + "yield type differs from the result element type");
  s.ops = ap.body;
  c.yield_value = ap.yield_value;
  s.fix();
  try {
    validate_desc(c);
  } catch (const Diagnostic &d) {
    throw Diagnostic(where + d.what(), d.unsupported);
  }
  if (ap.has_shape) {
    // `shape` is read by nothing in the reference's lowering; here it is cross-checked
    std::set<std::vector<int64_t>> sh(ap.shape.begin(), ap.shape.end());
    for (auto &o : s.ops) {
      if (o.opcode != NEPTUNE_RT_OP_ACCESS) continue;
      std::vector<int64_t> v(o.offset, o.offset + rank);
      bool zero = std::all_of(v.begin(), v.end(), [](int64_t x) { return x == 0; });
      if (!zero && !sh.count(v)) throw Diagnostic(where + "access offset is not listed in the apply's stencil_shape attribute");
    }
  }
  return s;
}

uint64_t body_key(const neptune_rt_stencil_desc &d) {
  // FNV-1a over the structure: rank, dtype, n_inputs, per op (opcode, operands, input, offsets), yield.
  uint64_t h = 1469598103934665603ULL;
  auto mix = [&](int64_t v) {
    for (int i = 0; i < 8; ++i) { h ^= (uint64_t)((v >> (8 * i)) & 0xff); h *= 1099511628211ULL; }
  };
  mix(d.rank); mix(d.dtype); mix(d.n_inputs); mix(d.n_ops);
  for (int i = 0; i < d.n_ops; ++i) {
    const neptune_rt_body_op &o = d.ops[i];
    mix(o.opcode);
    switch (o.opcode) {
      case NEPTUNE_RT_OP_ACCESS:
        mix(o.input);
        for (int a = 0; a < d.rank; ++a) mix(o.offset[a]);
        break;
      case NEPTUNE_RT_OP_CONST: break;  // value is a run-time kernel parameter
      case NEPTUNE_RT_OP_NEG: mix(o.a); break;
      default: mix(o.a); mix(o.b);
    }
  }
  mix(d.yield_value);
  return h;
}

BodyInfo analyze_body(const neptune_rt_stencil_desc &d) {
  BodyInfo b;
  b.rank = d.rank; b.dtype = d.dtype; b.n_inputs = d.n_inputs;
  std::memset(b.radius, 0, sizeof b.radius);
  b.const_slot.assign(d.n_ops, -1);
  for (int i = 0; i < d.n_ops; ++i) {
    const neptune_rt_body_op &o = d.ops[i];
    if (o.opcode == NEPTUNE_RT_OP_CONST) {
      b.const_slot[i] = (int)b.consts.size();
      b.consts.push_back(o.value);
    } else if (o.opcode == NEPTUNE_RT_OP_ACCESS) {
      int zyx[3];
      canonical_offset(d.rank, o.offset, zyx);
      for (int a = 0; a < 3; ++a) b.radius[o.input][a] = std::max(b.radius[o.input][a], std::abs(zyx[a]));
    }
  }
  b.n_consts = (int)b.consts.size();
  for (int k = 0; k < d.n_inputs; ++k) {
    if (b.radius[k][0] || b.radius[k][1] || b.radius[k][2]) b.staged = (b.staged == -1) ? k : -2;
  }
  return b;
}

}  // namespace neptune
