// Producer -> consumer fusion of applies (the explicit time step
// `u + dt * A(u)` the reference writes as an operator call followed by an AXPY apply,
// test/smoke_tests/smoke_time_advance.mlir:33-51, lib/Passes/HighLevelConvertion.cpp:77-120).
//
//   %t = apply_linear @A(%x)            // or a plain neptune_ir.apply
//   %y = apply(%x, %t) { ... access %t_in[0] ... }
//
// becomes ONE apply whose body is A's op list followed by the consumer's, with the consumer's
// `access %t_in[0]` replaced by A's yielded value: one sweep, one read and one write of the field
// instead of two sweeps and a temporary.  Per point the arithmetic is the same ops in the same
// order, so the result is bit-identical to the unfused lowering.  Legal when
//   * the temp has exactly one use, as a consumer input other than input 0 (input 0 is the consumer's
//     copy-through source, lib/Passes/DataflowLowering.cpp:282-287), read at offset 0 only;
//   * the consumer's bounds lie inside the producer's (outside them %t holds the producer's
//     copy-through values, not A(x)) and both work on the same box;
//   * nothing between the two ops stores into a field (a `load`ed temp aliases its field's buffer).
// `accept(key)` lets the caller keep the unfused form when no kernel is compiled for the fused body.
#include <algorithm>
#include <functional>
#include <set>

#include "ir.h"

namespace neptune {

namespace {

// an operator definition that is exactly one apply over its block arguments
const Apply *single_apply_operator(const Function &f, std::string *ret) {
  const Apply *ap = nullptr;
  for (auto &op : f.ops) {
    if (op.kind == HostOp::ApplyOp) {
      if (ap) return nullptr;
      ap = op.apply.get();
    } else if (op.kind == HostOp::Return) {
      if (op.operands.size() != 1) return nullptr;
      *ret = op.operands[0];
    } else if (op.kind != HostOp::Constant) {
      return nullptr;
    }
  }
  if (!ap || *ret != ap->result) return nullptr;
  for (auto &in : ap->inputs) {
    bool is_arg = false;
    for (auto &a : f.args) is_arg |= a.first == in;
    if (!is_arg) return nullptr;
  }
  return ap;
}

bool inside(const Bounds &inner, const Bounds &outer) {
  if (inner.rank() != outer.rank()) return false;
  for (int d = 0; d < inner.rank(); ++d)
    if (inner.lb[d] < outer.lb[d] || inner.ub[d] > outer.ub[d]) return false;
  return true;
}

// consumer `c` with its input `k` replaced by producer `p` (whose inputs are already caller-level names)
std::shared_ptr<Apply> fuse_pair(const Apply &p, const Apply &c, int k) {
  auto f = std::make_shared<Apply>();
  f->result = c.result;
  f->out_type = c.out_type;
  f->bounds = c.bounds;
  f->line = c.line;
  std::vector<int> cmap(c.inputs.size(), -1), pmap(p.inputs.size(), -1);
  for (size_t i = 0; i < c.inputs.size(); ++i) {
    if ((int)i == k) continue;
    cmap[i] = (int)f->inputs.size();
    f->inputs.push_back(c.inputs[i]);
    f->in_types.push_back(c.in_types[i]);
  }
  for (size_t i = 0; i < p.inputs.size(); ++i) {
    auto it = std::find(f->inputs.begin(), f->inputs.end(), p.inputs[i]);
    if (it != f->inputs.end() && f->in_types[it - f->inputs.begin()] == p.in_types[i]) {
      pmap[i] = (int)(it - f->inputs.begin());
    } else {
      pmap[i] = (int)f->inputs.size();
      f->inputs.push_back(p.inputs[i]);
      f->in_types.push_back(p.in_types[i]);
    }
  }
  // producer ops first, then the consumer's, value numbers shifted
  std::vector<int> pv(p.body.size(), -1), cv(c.body.size(), -1);
  for (size_t i = 0; i < p.body.size(); ++i) {
    neptune_rt_body_op o = p.body[i];
    if (o.opcode == NEPTUNE_RT_OP_ACCESS) o.input = pmap[o.input];
    if (o.a >= 0) o.a = pv[o.a];
    if (o.b >= 0) o.b = pv[o.b];
    pv[i] = (int)f->body.size();
    f->body.push_back(o);
    f->body_dtype.push_back(p.body_dtype[i]);
  }
  for (size_t i = 0; i < c.body.size(); ++i) {
    neptune_rt_body_op o = c.body[i];
    if (o.opcode == NEPTUNE_RT_OP_ACCESS && o.input == k) {
      cv[i] = pv[p.yield_value];   // the temp at this point IS the producer's yield
      continue;
    }
    if (o.opcode == NEPTUNE_RT_OP_ACCESS) o.input = cmap[o.input];
    if (o.a >= 0) o.a = cv[o.a];
    if (o.b >= 0) o.b = cv[o.b];
    cv[i] = (int)f->body.size();
    f->body.push_back(o);
    f->body_dtype.push_back(c.body_dtype[i]);
  }
  f->yield_value = cv[c.yield_value];
  f->yield_dtype = c.yield_dtype;
  return f;
}

}  // namespace

int fuse_applies(Module &m, bool fma, const std::function<bool(uint64_t)> &accept) {
  int fused = 0;
  for (auto &fn : m.functions) {
    if (fn.rejected) continue;
    bool changed = true;
    while (changed) {
      changed = false;
      std::map<std::string, std::vector<int>> uses;
      for (size_t i = 0; i < fn.ops.size(); ++i)
        for (auto &o : fn.ops[i].operands) uses[o].push_back((int)i);
      for (size_t j = 0; j < fn.ops.size() && !changed; ++j) {
        HostOp &cop = fn.ops[j];
        if (cop.kind != HostOp::ApplyOp) continue;
        const Apply &c = *cop.apply;
        for (int k = 1; k < (int)c.inputs.size() && !changed; ++k) {
          const std::string &t = c.inputs[k];
          if (std::count(c.inputs.begin(), c.inputs.end(), t) != 1) continue;
          auto us = uses.find(t);
          if (us == uses.end() || us->second.size() != 1) continue;
          // the op that defines %t
          int i = -1;
          for (size_t q = 0; q < j; ++q)
            if (!fn.ops[q].results.empty() && fn.ops[q].results[0] == t) i = (int)q;
          if (i < 0) continue;
          const HostOp &pop = fn.ops[i];
          std::shared_ptr<Apply> prod;
          if (pop.kind == HostOp::ApplyOp) {
            prod = std::make_shared<Apply>(*pop.apply);
          } else if (pop.kind == HostOp::ApplyLinear) {
            const Function *callee = m.find(pop.callee);
            std::string ret;
            const Apply *ap = callee ? single_apply_operator(*callee, &ret) : nullptr;
            if (!ap || callee->args.size() != pop.operands.size()) continue;
            prod = std::make_shared<Apply>(*ap);
            for (auto &in : prod->inputs)   // operator block arguments -> call operands
              for (size_t a = 0; a < callee->args.size(); ++a)
                if (callee->args[a].first == in) { in = pop.operands[a]; break; }
          } else {
            continue;
          }
          // consumer reads %t at offset 0 only; same box; bounds nested; no store in between
          bool ok = prod->out_type.bounds == c.in_types[k].bounds && prod->out_type.dtype == c.in_types[k].dtype &&
                    inside(c.bounds, prod->bounds) && prod->yield_dtype == c.in_types[k].dtype;
          for (auto &o : c.body)
            if (o.opcode == NEPTUNE_RT_OP_ACCESS && o.input == k)
              for (int d = 0; d < c.bounds.rank(); ++d) ok &= o.offset[d] == 0;
          for (int q = i + 1; q < (int)j; ++q) ok &= fn.ops[q].kind != HostOp::Store;
          if (!ok) continue;
          std::shared_ptr<Apply> f = fuse_pair(*prod, c, k);
          if ((int)f->inputs.size() > NEPTUNE_RT_MAX_INPUTS) continue;
          try {
            StencilDesc d = recognize_apply(*f, fma);
            if (accept && !accept(body_key(d.c))) continue;
          } catch (const Diagnostic &) {
            continue;   // the fused body is outside the recognised set: keep the two sweeps
          }
          cop.apply = f;
          cop.operands = f->inputs;
          fn.ops.erase(fn.ops.begin() + i);
          ++fused;
          changed = true;
        }
      }
    }
  }
  return fused;
}

}  // namespace neptune
