// neptune_b200_* — the `neptuneir-to-b200` pipeline as a C ABI (include/neptune_b200.h).
//
// Compile time: parse -> ApplyOp::verify rules (parser.cpp) -> recognise every apply
// (recognizer.cpp) -> plan the runtime calls of each function.  This replaces
// verify-annotate -> structure-lowering -> dataflow-lowering{backend=cpu}
// (lib/Pipeline/NeptuneIRPassesPipeline.cpp:10-26) for modules on the apply hot path.
// Run time: the plan is executed against neptune_rt_* (runtime/runtime.cu); wrap/unwrap/load
// forward their operand exactly as the reference's lowerings do (DataflowLowering.cpp:131-159),
// `store` is a device copy (:165-220) unless it is the only use of an apply result, in which
// case the sweep writes straight into the destination field (the reference's alloc + copy-through
// + store-copy collapse into the one pass over the data).
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <sstream>

#include "ir.h"
#include "neptune_b200.h"

using namespace neptune;

namespace {

struct ApplyRec {
  std::string function;
  StencilDesc desc;
  std::string cuda_source;
  const Apply *ap = nullptr;
};

struct FnInfo {
  const Function *fn = nullptr;
  // memref argument -> the field type it is wrapped as (bounds, dtype); index = arg position
  std::vector<Type> arg_field;        // kind == None when the argument is never wrapped
  std::vector<bool> arg_needs_upload;
  std::vector<bool> arg_written;
  std::map<int, int> fused_store;     // op index of an apply -> op index of the store fused into it
  std::set<int> elided;               // store ops that became no-ops
  // entry of the form wrap* / load* / ONE apply on argument fields / fused whole store to an argument:
  // run from host buffers as an upload | sweep | download pipeline (neptune_rt_apply_host)
  int pipe_apply = -1;                // op index of that apply, -1 when the function has another shape
  std::vector<int> pipe_in_args;      // argument position of every apply input
  int pipe_out_arg = -1;
  std::string plan;
};

}  // namespace

struct neptune_b200_module {
  Module mod;
  bool fma = false;
  int n_fused = 0;   // producer -> consumer apply pairs folded into one sweep (fusion.cpp)
  std::vector<ApplyRec> applies;
  std::map<const Apply *, int> apply_index;
  std::map<std::string, FnInfo> fns;
  // run-time state, per ctx
  struct PerCtx {
    std::vector<neptune_rt_stencil *> stencils;
    std::map<std::string, std::vector<neptune_rt_field *>> arg_fields;  // function -> device fields of its memref args
    std::map<int, std::vector<neptune_rt_field *>> temps;               // apply index -> reusable result temps
    std::map<int, int> cursor;                                          // temps of that apply taken in this run
  };
  std::map<neptune_rt_ctx *, PerCtx> per_ctx;
  std::vector<double> last_scalars;   // scalar results returned by the last run_host
  std::string err;
};

namespace {

std::string bounds_str(const Bounds &b) {
  std::ostringstream o;
  o << "[";
  for (int d = 0; d < b.rank(); ++d) o << (d ? "," : "") << b.lb[d];
  o << "]..[";
  for (int d = 0; d < b.rank(); ++d) o << (d ? "," : "") << b.ub[d];
  o << "]";
  return o.str();
}

// Resolve a value name to the SSA name of the storage it aliases (wrap/unwrap/load forward).
std::string root_of(const std::map<std::string, std::string> &alias, std::string v) {
  for (;;) {
    auto it = alias.find(v);
    if (it == alias.end() || it->second == v) return v;
    v = it->second;
  }
}

void analyze_function(neptune_b200_module &m, const Function &fn, FnInfo &fi) {
  fi.fn = &fn;
  fi.arg_field.assign(fn.args.size(), Type());
  fi.arg_needs_upload.assign(fn.args.size(), false);
  fi.arg_written.assign(fn.args.size(), false);
  std::map<std::string, int> argpos;
  for (size_t i = 0; i < fn.args.size(); ++i) argpos[fn.args[i].first] = (int)i;
  std::map<std::string, std::string> alias;  // value -> value it forwards
  std::map<std::string, Type> vtype;
  for (auto &a : fn.args) vtype[a.first] = a.second;

  // uses of every apply result
  std::map<std::string, std::vector<int>> uses;
  for (size_t i = 0; i < fn.ops.size(); ++i)
    for (auto &o : fn.ops[i].operands) uses[o].push_back((int)i);

  std::set<std::string> touched;  // storage roots already read or written
  std::ostringstream plan;
  plan << fn.name << ":\n";
  if (m.n_fused) plan << "  (" << m.n_fused << " producer -> consumer apply pair(s) of this module fused into one sweep each)\n";
  for (size_t i = 0; i < fn.ops.size(); ++i) {
    const HostOp &op = fn.ops[i];
    switch (op.kind) {
      case HostOp::Wrap: {
        auto it = argpos.find(op.operands[0]);
        if (op.types[1].kind != Type::Grid) throw Diagnostic("line " + std::to_string(op.line) + ": wrap must produce a !neptune_ir.field");
        if (op.types[0].kind != Type::MemRef) throw Diagnostic("line " + std::to_string(op.line) + ": wrap operand must be a memref");
        if (it != argpos.end()) {
          Type &t = fi.arg_field[it->second];
          if (t.kind != Type::None && t != op.types[1])
            throw Diagnostic("line " + std::to_string(op.line) + ": memref argument wrapped with two different field types", true);
          t = op.types[1];
        }
        alias[op.results[0]] = op.operands[0];
        vtype[op.results[0]] = op.types[1];
        plan << "  " << op.results[0] << " = wrap " << op.operands[0] << "  -> neptune_rt_field_alloc(" << (op.types[1].dtype == NEPTUNE_RT_F64 ? "f64" : "f32")
             << ", " << bounds_str(op.types[1].bounds) << ")\n";
        break;
      }
      case HostOp::Unwrap:
      case HostOp::Load:
        alias[op.results[0]] = op.operands[0];
        vtype[op.results[0]] = op.types[1];
        plan << "  " << op.results[0] << " = " << (op.kind == HostOp::Load ? "load " : "unwrap ") << op.operands[0] << "  -> forwarded\n";
        break;
      case HostOp::ApplyOp: {
        const Apply &ap = *op.apply;
        const int idx = m.apply_index.at(&ap);
        std::set<std::string> in_roots;
        for (auto &v : ap.inputs) {
          std::string r = root_of(alias, v);
          in_roots.insert(r);
          if (argpos.count(r) && !touched.count(r)) fi.arg_needs_upload[argpos[r]] = true;
          touched.insert(r);
        }
        vtype[op.results[0]] = ap.out_type;
        // fuse `store %r to %f` when it is the only use of %r, copies the whole box, the field has the
        // temp's bounds and is not one of the apply's inputs
        auto &us = uses[op.results[0]];
        bool fused = false;
        if (us.size() == 1) {
          const HostOp &st = fn.ops[us[0]];
          if (st.kind == HostOp::Store && !st.has_bounds && st.operands[0] == op.results[0]) {
            std::string droot = root_of(alias, st.operands[1]);
            bool clobbered = false;  // nothing between the apply and the store may touch the destination
            for (int j = (int)i + 1; j < us[0]; ++j)
              for (auto &o : fn.ops[j].operands) clobbered |= root_of(alias, o) == droot;
            const Type &dt = vtype[st.operands[1]];
            if (!in_roots.count(droot) && !clobbered && dt.kind == Type::Grid && dt.bounds == ap.out_type.bounds && dt.dtype == ap.out_type.dtype) {
              fi.fused_store[(int)i] = us[0];
              fi.elided.insert(us[0]);
              fused = true;
              if (argpos.count(droot)) fi.arg_written[argpos[droot]] = true;
              touched.insert(droot);
            }
          }
        }
        char key[32];
        std::snprintf(key, sizeof key, "0x%016llx", (unsigned long long)body_key(m.applies[idx].desc.c));
        plan << "  " << op.results[0] << " = neptune_rt_apply(stencil #" << idx << " key " << key << ", in = [";
        for (size_t k = 0; k < ap.inputs.size(); ++k) plan << (k ? ", " : "") << ap.inputs[k];
        plan << "], bounds = " << bounds_str(ap.bounds) << ")";
        if (fused) plan << "  -> writes " << fn.ops[us[0]].operands[1] << " directly (store fused)";
        else plan << "  -> neptune_rt_field_alloc(temp)";
        plan << "\n";
        break;
      }
      case HostOp::ApplyLinear: {
        const Function *callee = m.mod.find(op.callee);
        if (!callee) throw Diagnostic("line " + std::to_string(op.line) + ": call to undefined operator " + op.callee);
        if (callee->args.size() != op.operands.size()) throw Diagnostic("line " + std::to_string(op.line) + ": operand count differs from " + op.callee);
        for (auto &v : op.operands) {
          std::string r = root_of(alias, v);
          if (argpos.count(r) && !touched.count(r)) fi.arg_needs_upload[argpos[r]] = true;
          touched.insert(r);
        }
        vtype[op.results[0]] = op.types.back();
        plan << "  " << op.results[0] << " = call " << op.callee << "(...)\n";
        break;
      }
      case HostOp::Store: {
        std::string sroot = root_of(alias, op.operands[0]), droot = root_of(alias, op.operands[1]);
        if (fi.elided.count((int)i)) {
          plan << "  store " << op.operands[0] << " to " << op.operands[1] << "  -> elided (fused into the apply)\n";
          break;
        }
        if (argpos.count(sroot) && !touched.count(sroot)) fi.arg_needs_upload[argpos[sroot]] = true;
        touched.insert(sroot);
        if (argpos.count(droot)) {
          // a partial store leaves the rest of the destination visible: it must be uploaded first
          if (op.has_bounds && !touched.count(droot)) fi.arg_needs_upload[argpos[droot]] = true;
          fi.arg_written[argpos[droot]] = true;
        }
        touched.insert(droot);
        plan << "  store " << op.operands[0] << " to " << op.operands[1] << "  -> neptune_rt_field_store("
             << (op.has_bounds ? bounds_str(op.bounds) : std::string("whole")) << ")\n";
        break;
      }
      case HostOp::Reduce: {
        if (op.types[0].kind != Type::Grid) throw Diagnostic("line " + std::to_string(op.line) + ": reduce input must be a !neptune_ir.temp");
        if (op.types[1].kind != Type::Float) throw Diagnostic("line " + std::to_string(op.line) + ": reduce of a non-float element type is not on the B200 path", true);
        std::string r = root_of(alias, op.operands[0]);
        if (argpos.count(r) && !touched.count(r)) fi.arg_needs_upload[argpos[r]] = true;
        touched.insert(r);
        vtype[op.results[0]] = op.types[1];
        plan << "  " << op.results[0] << " = reduce " << op.operands[0] << "  -> neptune_rt_field_reduce_sum("
             << (op.has_bounds ? bounds_str(op.bounds) : std::string("whole")) << ")\n";
        break;
      }
      case HostOp::Return:
        plan << "  return";
        for (auto &v : op.operands) plan << " " << v;
        plan << "\n";
        break;
      case HostOp::Constant: break;
    }
  }
  // pipelined host form?
  {
    int n_apply = 0, ai = -1;
    bool simple = true;
    for (size_t i = 0; i < fn.ops.size(); ++i) {
      const HostOp &op = fn.ops[i];
      if (op.kind == HostOp::ApplyOp) { ++n_apply; ai = (int)i; }
      else if (op.kind == HostOp::ApplyLinear || op.kind == HostOp::Reduce) simple = false;
      else if (op.kind == HostOp::Store && !fi.elided.count((int)i)) simple = false;
    }
    if (simple && n_apply == 1 && fi.fused_store.count(ai)) {
      const Apply &ap = *fn.ops[ai].apply;
      std::string droot = root_of(alias, fn.ops[fi.fused_store[ai]].operands[1]);
      std::vector<int> ins;
      for (auto &v : ap.inputs) {
        std::string r = root_of(alias, v);
        if (argpos.count(r)) ins.push_back(argpos[r]);
      }
      if (argpos.count(droot) && ins.size() == ap.inputs.size() && !fi.arg_needs_upload[argpos[droot]]) {
        fi.pipe_apply = ai;
        fi.pipe_in_args = ins;
        fi.pipe_out_arg = argpos[droot];
        plan << "  host call: upload | sweep | download pipelined over dim-0 plane chunks (neptune_rt_apply_host)\n";
      }
    }
  }
  for (size_t i = 0; i < fn.args.size(); ++i)
    if (fi.arg_field[i].kind == Type::Grid)
      plan << "  arg " << i << " " << fn.args[i].first << ": " << (fi.arg_needs_upload[i] ? "upload" : "no upload")
           << ", " << (fi.arg_written[i] ? "download" : "no download") << "\n";
  fi.plan = plan.str();
}

struct Exec {
  neptune_b200_module &m;
  neptune_rt_ctx *ctx;
  neptune_b200_module::PerCtx &pc;
  std::vector<neptune_rt_field *> scratch;  // temps of opdef calls, freed by the caller of run()
  std::map<std::string, double> scalars;    // results of reduce ops
  std::vector<double> returned;             // scalar values the top-level function returned

  int fail(int code, const std::string &msg) {
    m.err = msg;
    return code;
  }
  int rt_fail(int code) {
    m.err = neptune_rt_last_error(ctx);
    return code;
  }

  neptune_rt_stencil *stencil(int idx, int *rc) {
    if ((int)pc.stencils.size() < (int)m.applies.size()) pc.stencils.resize(m.applies.size(), nullptr);
    if (!pc.stencils[idx]) {
      *rc = neptune_rt_stencil_create(ctx, &m.applies[idx].desc.c, &pc.stencils[idx]);
      if (*rc) { rt_fail(*rc); return nullptr; }
    }
    return pc.stencils[idx];
  }

  // Execute `fn` with its arguments bound to device fields; *ret receives the returned value's field.
  int run(const Function &fn, const std::vector<neptune_rt_field *> &args, neptune_rt_field **ret) {
    const FnInfo &fi = m.fns.at(fn.name);
    std::map<std::string, neptune_rt_field *> env;
    for (size_t i = 0; i < fn.args.size(); ++i) env[fn.args[i].first] = args[i];
    for (size_t i = 0; i < fn.ops.size(); ++i) {
      const HostOp &op = fn.ops[i];
      switch (op.kind) {
        case HostOp::Wrap:
        case HostOp::Unwrap:
        case HostOp::Load:
          env[op.results[0]] = env[op.operands[0]];
          break;
        case HostOp::ApplyOp: {
          const Apply &ap = *op.apply;
          const int idx = m.apply_index.at(&ap);
          int rc = 0;
          neptune_rt_stencil *st = stencil(idx, &rc);
          if (!st) return rc;
          std::vector<neptune_rt_field *> ins;
          for (auto &v : ap.inputs) {
            if (!env.count(v) || !env[v]) return fail(NEPTUNE_RT_EINVAL, "apply input " + v + " has no storage");
            ins.push_back(env[v]);
          }
          neptune_rt_field *out = nullptr;
          auto fs = fi.fused_store.find((int)i);
          if (fs != fi.fused_store.end()) {
            out = env[fn.ops[fs->second].operands[1]];
          } else {
            // the n-th execution of this apply within one top-level run owns the n-th pooled temp, so
            // two live results of the same opdef never share storage
            auto &pool = pc.temps[idx];
            int &cur = pc.cursor[idx];
            if (cur >= (int)pool.size()) {
              const neptune_rt_stencil_desc &d = m.applies[idx].desc.c;
              neptune_rt_field *f = nullptr;
              if ((rc = neptune_rt_field_alloc(ctx, d.dtype, d.rank, d.out_lb, d.out_ub, &f))) return rt_fail(rc);
              pool.push_back(f);
            }
            out = pool[cur++];
          }
          if ((rc = neptune_rt_apply(st, (int)ins.size(), ins.data(), out))) return rt_fail(rc);
          env[op.results[0]] = out;
          break;
        }
        case HostOp::ApplyLinear: {
          const Function *callee = m.mod.find(op.callee);
          std::vector<neptune_rt_field *> cargs;
          for (auto &v : op.operands) cargs.push_back(env[v]);
          neptune_rt_field *r = nullptr;
          if (int rc = run(*callee, cargs, &r)) return rc;
          env[op.results[0]] = r;
          break;
        }
        case HostOp::Store: {
          if (fi.elided.count((int)i)) break;
          neptune_rt_field *src = env[op.operands[0]], *dst = env[op.operands[1]];
          if (!src || !dst) return fail(NEPTUNE_RT_EINVAL, "store operand has no storage");
          int rc;
          if (op.has_bounds) rc = neptune_rt_field_store(dst, src, op.bounds.lb.data(), op.bounds.ub.data());
          else rc = neptune_rt_field_store(dst, src, nullptr, nullptr);
          if (rc) return rt_fail(rc);
          break;
        }
        case HostOp::Reduce: {
          neptune_rt_field *src = env[op.operands[0]];
          if (!src) return fail(NEPTUNE_RT_EINVAL, "reduce operand has no storage");
          double v = 0;
          if (int rc = neptune_rt_field_reduce_sum(src, op.has_bounds ? op.bounds.lb.data() : nullptr,
                                                   op.has_bounds ? op.bounds.ub.data() : nullptr, 0, &v))
            return rt_fail(rc);
          scalars[op.results[0]] = v;
          break;
        }
        case HostOp::Return:
          returned.clear();
          for (auto &v : op.operands)
            if (scalars.count(v)) returned.push_back(scalars[v]);
          if (ret) *ret = (op.operands.empty() || scalars.count(op.operands[0])) ? nullptr : env[op.operands[0]];
          return NEPTUNE_RT_OK;
        case HostOp::Constant: break;
      }
    }
    if (ret) *ret = nullptr;
    return NEPTUNE_RT_OK;
  }
};

int bind_args(neptune_b200_module &m, neptune_rt_ctx *ctx, const Function &fn, std::vector<neptune_rt_field *> **out) {
  auto &pc = m.per_ctx[ctx];
  auto &v = pc.arg_fields[fn.name];
  const FnInfo &fi = m.fns.at(fn.name);
  if (v.empty()) {
    v.assign(fn.args.size(), nullptr);
    for (size_t i = 0; i < fn.args.size(); ++i) {
      // func.func: the field type a memref argument is wrapped as; operator definitions take temps directly
      const Type &t = fn.kind == "func" ? fi.arg_field[i] : fn.args[i].second;
      if (t.kind != Type::Grid) continue;
      if (int rc = neptune_rt_field_alloc(ctx, t.dtype, t.bounds.rank(), t.bounds.lb.data(), t.bounds.ub.data(), &v[i])) {
        m.err = neptune_rt_last_error(ctx);
        return rc;
      }
    }
  }
  *out = &v;
  return NEPTUNE_RT_OK;
}

}  // namespace

extern "C" int neptune_b200_compile(const char *text, const char *pipeline, neptune_b200_module **out, char *err, size_t err_len) {
  auto report = [&](const std::string &msg) {
    if (err && err_len) {
      std::snprintf(err, err_len, "%s", msg.c_str());
    }
  };
  if (out) *out = nullptr;
  if (!text || !pipeline || !out) { report("neptune_b200_compile: NULL argument"); return NEPTUNE_RT_EINVAL; }
  std::string pl = pipeline;
  bool fma = false;
  const std::string name = "neptuneir-to-b200";
  if (pl.compare(0, name.size(), name) != 0) {
    report("unknown pipeline '" + pl + "' (this library registers 'neptuneir-to-b200'; 'neptuneir-to-llvm' is the reference's CPU pipeline and is not provided here)");
    return NEPTUNE_RT_EINVAL;
  }
  std::string opts = pl.substr(name.size());
  bool fuse = true;
  if (!opts.empty()) {
    // {key=value key=value}: fma (allow contraction), fuse (producer -> consumer apply fusion, fusion.cpp)
    if (opts.size() < 2 || opts.front() != '{' || opts.back() != '}') { report("malformed pipeline options '" + opts + "'"); return NEPTUNE_RT_EINVAL; }
    std::istringstream is(opts.substr(1, opts.size() - 2));
    std::string kv;
    while (is >> kv) {
      if (!kv.empty() && kv.back() == ',') kv.pop_back();
      const bool on = kv.size() > 5 && (kv.substr(kv.find('=') + 1) == "true" || kv.substr(kv.find('=') + 1) == "1");
      const bool off = kv.size() > 5 && (kv.substr(kv.find('=') + 1) == "false" || kv.substr(kv.find('=') + 1) == "0");
      if (kv.compare(0, 4, "fma=") == 0 && (on || off)) fma = on;
      else if (kv.compare(0, 5, "fuse=") == 0 && (on || off)) fuse = on;
      else { report("unknown pipeline option '" + kv + "' (supported: fma=true|false, fuse=true|false)"); return NEPTUNE_RT_EINVAL; }
    }
  }
  auto *m = new neptune_b200_module();
  m->fma = fma;
  try {
    m->mod = parse_module(text);
    auto reject = [](Function &fn, const Diagnostic &d) {
      fn.ops.clear();
      fn.rejected = true;
      fn.unsupported = d.unsupported;
      fn.diagnostic = d.what();
    };
    // a function that calls a rejected operator cannot be lowered either (to a fixed point: calls nest)
    for (bool changed = true; changed;) {
      changed = false;
      for (auto &fn : m->mod.functions) {
        if (fn.rejected) continue;
        for (auto &op : fn.ops) {
          if (op.kind != HostOp::ApplyLinear) continue;
          const Function *callee = m->mod.find(op.callee);
          if (callee && callee->rejected) {
            reject(fn, Diagnostic("line " + std::to_string(op.line) + ": calls " + op.callee + ", which is not lowered: " + callee->diagnostic, callee->unsupported));
            changed = true;
            break;
          }
        }
      }
    }
    if (fuse) m->n_fused = fuse_applies(m->mod, fma, [](uint64_t key) { return neptune_rt_body_compiled(key) != 0; });
    for (auto &fn : m->mod.functions) {
      if (fn.rejected) continue;
      // recognise every apply of the function first: one body outside the recognised set rejects the function
      std::vector<ApplyRec> recs;
      try {
        for (auto &op : fn.ops)
          if (op.kind == HostOp::ApplyOp) {
            ApplyRec r;
            r.function = fn.name;
            r.ap = op.apply.get();
            r.desc = recognize_apply(*op.apply, fma);
            if (fn.kind == "linear_opdef") {
              // the linear-region rule of the reference's verifier (VerifyAndAnnotate.cpp:183-204): a product needs a
              // constant factor, a quotient a constant divisor
              const neptune_rt_stencil_desc &d = r.desc.c;
              for (int i = 0; i < d.n_ops; ++i) {
                const neptune_rt_body_op &o = d.ops[i];
                if (o.opcode == NEPTUNE_RT_OP_MUL && d.ops[o.a].opcode != NEPTUNE_RT_OP_CONST && d.ops[o.b].opcode != NEPTUNE_RT_OP_CONST)
                  throw Diagnostic("line " + std::to_string(op.line) + ": MulFOp in linear region must multiply by a constant (apply inside " +
                                   fn.name + ", a linear_opdef)");
                if (o.opcode == NEPTUNE_RT_OP_DIV && d.ops[o.b].opcode != NEPTUNE_RT_OP_CONST)
                  throw Diagnostic("line " + std::to_string(op.line) + ": DivFOp in linear region must divide by a constant (apply inside " +
                                   fn.name + ", a linear_opdef)");
              }
            }
            r.cuda_source = generate_functor(r.desc.c);
            recs.push_back(std::move(r));
          }
      } catch (const Diagnostic &d) {
        if (!d.unsupported) throw;   // invalid IR (e.g. the linear-region rule): the whole module fails, as in the reference
        reject(fn, d);
        continue;
      }
      for (auto &r : recs) {
        m->apply_index[r.ap] = (int)m->applies.size();
        m->applies.push_back(std::move(r));
      }
    }
    for (bool changed = true; changed;) {   // rejections found by the recogniser propagate to callers as well
      changed = false;
      for (auto &fn : m->mod.functions) {
        if (fn.rejected) continue;
        for (auto &op : fn.ops) {
          const Function *callee = op.kind == HostOp::ApplyLinear ? m->mod.find(op.callee) : nullptr;
          if (callee && callee->rejected) {
            reject(fn, Diagnostic("line " + std::to_string(op.line) + ": calls " + op.callee + ", which is not lowered: " + callee->diagnostic, callee->unsupported));
            changed = true;
            break;
          }
        }
      }
    }
    // (applies of functions rejected in that second round stay in the table; nothing runs them)
    int lowered = 0;
    for (auto &fn : m->mod.functions) {
      if (fn.rejected) continue;
      try {
        analyze_function(*m, fn, m->fns[fn.name]);
        ++lowered;
      } catch (const Diagnostic &d) {
        // a broken rule (the linear-region rule, a bounds mismatch ...) is an invalid module, not an unsupported one:
        // the reference's VerifyAndAnnotate fails the pass and with it the whole module
        if (!d.unsupported) throw;
        m->fns.erase(fn.name);
        reject(fn, d);
      }
    }
    if (lowered == 0) {
      // nothing of the module is on the B200 path: fail with the first diagnostic, as a failed conversion does
      for (auto &fn : m->mod.functions)
        if (fn.rejected) throw Diagnostic(fn.diagnostic, fn.unsupported);
      throw Diagnostic("the module holds no function");
    }
  } catch (const Diagnostic &d) {
    report(d.what());
    const int rc = d.unsupported ? NEPTUNE_RT_EUNSUPPORTED : NEPTUNE_RT_EINVAL;
    delete m;
    return rc;
  } catch (const std::exception &e) {
    report(std::string("internal error: ") + e.what());
    delete m;
    return NEPTUNE_RT_EINVAL;
  }
  *out = m;
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_b200_module_destroy(neptune_b200_module *m) {
  if (!m) return NEPTUNE_RT_OK;
  for (auto &pc : m->per_ctx) {
    // a ctx finalized before the module took its fields and stencils with it (neptune_rt_finalize): nothing left to free
    if (!neptune_rt_ctx_alive(pc.first)) continue;
    for (auto *s : pc.second.stencils) neptune_rt_stencil_destroy(s);
    for (auto &kv : pc.second.arg_fields)
      for (auto *f : kv.second) neptune_rt_field_free(f);
    for (auto &kv : pc.second.temps)
      for (auto *f : kv.second) neptune_rt_field_free(f);
  }
  delete m;
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_b200_apply_count(neptune_b200_module *m) { return m ? (int)m->applies.size() : 0; }
extern "C" const neptune_rt_stencil_desc *neptune_b200_apply_desc(neptune_b200_module *m, int i) {
  return (m && i >= 0 && i < (int)m->applies.size()) ? &m->applies[i].desc.c : nullptr;
}
extern "C" const char *neptune_b200_apply_function(neptune_b200_module *m, int i) {
  return (m && i >= 0 && i < (int)m->applies.size()) ? m->applies[i].function.c_str() : nullptr;
}
extern "C" const char *neptune_b200_apply_cuda_source(neptune_b200_module *m, int i) {
  return (m && i >= 0 && i < (int)m->applies.size()) ? m->applies[i].cuda_source.c_str() : nullptr;
}
extern "C" const char *neptune_b200_function_plan(neptune_b200_module *m, const char *fn) {
  if (!m || !fn) return nullptr;
  auto it = m->fns.find(fn);
  if (it == m->fns.end()) {
    const Function *f = m->mod.find(fn);
    return (f && f->rejected) ? f->diagnostic.c_str() : nullptr;   // the diagnostic stands where the plan would
  }
  return it->second.plan.c_str();
}
extern "C" int neptune_b200_function_arg_count(neptune_b200_module *m, const char *fn) {
  if (!m || !fn) return -1;
  const Function *f = m->mod.find(fn);
  return f ? (int)f->args.size() : -1;
}
extern "C" int64_t neptune_b200_function_arg_numel(neptune_b200_module *m, const char *fn, int arg) {
  if (!m || !fn) return -1;
  auto it = m->fns.find(fn);
  if (it == m->fns.end() || arg < 0 || arg >= (int)it->second.arg_field.size()) return -1;
  const Type &t = it->second.arg_field[arg];
  return t.kind == Type::Grid ? t.bounds.numel() : -1;
}
extern "C" const char *neptune_b200_last_error(neptune_b200_module *m) { return m ? m->err.c_str() : ""; }
extern "C" int neptune_b200_function_count(neptune_b200_module *m) { return m ? (int)m->mod.functions.size() : 0; }
extern "C" const char *neptune_b200_function_name(neptune_b200_module *m, int i) {
  return (m && i >= 0 && i < (int)m->mod.functions.size()) ? m->mod.functions[(size_t)i].name.c_str() : nullptr;
}
extern "C" const char *neptune_b200_function_diagnostic(neptune_b200_module *m, const char *fn) {
  if (!m || !fn) return nullptr;
  const Function *f = m->mod.find(fn);
  return (f && f->rejected) ? f->diagnostic.c_str() : nullptr;
}

extern "C" neptune_rt_stencil *neptune_b200_apply_stencil(neptune_b200_module *m, neptune_rt_ctx *ctx, int index) {
  if (!m || !ctx || index < 0 || index >= (int)m->applies.size()) return nullptr;
  Exec ex{*m, ctx, m->per_ctx[ctx], {}};
  int rc = 0;
  return ex.stencil(index, &rc);
}

extern "C" int neptune_b200_run_host(neptune_b200_module *m, neptune_rt_ctx *ctx, const char *function, int n_args, void *const *args) {
  if (!m || !ctx || !function || (n_args > 0 && !args)) return NEPTUNE_RT_EINVAL;
  const Function *fn = m->mod.find(function);
  if (!fn) { m->err = std::string("no function ") + function; return NEPTUNE_RT_EINVAL; }
  if (fn->rejected) {
    m->err = std::string(function) + " was not lowered by neptuneir-to-b200: " + fn->diagnostic;
    return fn->unsupported ? NEPTUNE_RT_EUNSUPPORTED : NEPTUNE_RT_EINVAL;
  }
  if (fn->kind != "func") {
    // an operator definition (linear_opdef / nonlinear_opdef), called the way StructureLowering.cpp:31-140 makes it
    // callable (func.func + func.call): args[0 .. n) are host buffers of its temp operands, args[n] receives the result
    if ((int)fn->args.size() + 1 != n_args) { m->err = "an operator definition takes its operands plus one result buffer"; return NEPTUNE_RT_EINVAL; }
    std::vector<neptune_rt_field *> *fields = nullptr;
    if (int rc = bind_args(*m, ctx, *fn, &fields)) return rc;
    for (size_t i = 0; i < fn->args.size(); ++i)
      if ((*fields)[i])
        if (int rc = neptune_rt_field_upload((*fields)[i], args[i])) { m->err = neptune_rt_last_error(ctx); return rc; }
    Exec ex{*m, ctx, m->per_ctx[ctx], {}};
    ex.pc.cursor.clear();
    neptune_rt_field *ret = nullptr;
    if (int rc = ex.run(*fn, *fields, &ret)) return rc;
    if (!ret) { m->err = "the operator returned no temp"; return NEPTUNE_RT_EINVAL; }
    if (int rc = neptune_rt_field_download(ret, args[n_args - 1])) { m->err = neptune_rt_last_error(ctx); return rc; }
    if (int rc = neptune_rt_sync(ctx)) { m->err = neptune_rt_last_error(ctx); return rc; }
    return NEPTUNE_RT_OK;
  }
  if ((int)fn->args.size() != n_args) { m->err = "argument count mismatch"; return NEPTUNE_RT_EINVAL; }
  const FnInfo &fi = m->fns.at(fn->name);
  std::vector<neptune_rt_field *> *fields = nullptr;
  if (int rc = bind_args(*m, ctx, *fn, &fields)) return rc;
  if (fi.pipe_apply >= 0) {
    Exec ex{*m, ctx, m->per_ctx[ctx], {}};
    int rc = 0;
    neptune_rt_stencil *st = ex.stencil(m->apply_index.at(fn->ops[fi.pipe_apply].apply.get()), &rc);
    if (!st) return rc;
    std::vector<neptune_rt_field *> ins;
    std::vector<const void *> hosts;
    for (int a : fi.pipe_in_args) { ins.push_back((*fields)[a]); hosts.push_back(args[a]); }
    rc = neptune_rt_apply_host(st, (int)ins.size(), ins.data(), hosts.data(), (*fields)[fi.pipe_out_arg], args[fi.pipe_out_arg], 0);
    if (rc) m->err = neptune_rt_last_error(ctx);
    return rc;
  }
  for (int i = 0; i < n_args; ++i)
    if ((*fields)[i] && fi.arg_needs_upload[i])
      if (int rc = neptune_rt_field_upload((*fields)[i], args[i])) { m->err = neptune_rt_last_error(ctx); return rc; }
  Exec ex{*m, ctx, m->per_ctx[ctx], {}};
  ex.pc.cursor.clear();
  if (int rc = ex.run(*fn, *fields, nullptr)) return rc;
  m->last_scalars = ex.returned;
  for (int i = 0; i < n_args; ++i)
    if ((*fields)[i] && fi.arg_written[i])
      if (int rc = neptune_rt_field_download((*fields)[i], args[i])) { m->err = neptune_rt_last_error(ctx); return rc; }
  if (int rc = neptune_rt_sync(ctx)) { m->err = neptune_rt_last_error(ctx); return rc; }
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_b200_last_scalars(neptune_b200_module *m, double *out, int capacity) {
  if (!m) return 0;
  const int n = (int)m->last_scalars.size();
  for (int i = 0; i < n && i < capacity && out; ++i) out[i] = m->last_scalars[i];
  return n;
}

extern "C" int neptune_b200_run_steps_host(neptune_b200_module *m, neptune_rt_ctx *ctx, const char *function, int n_args,
                                           void *const *args, int64_t nsteps, int *result_arg) {
  if (!m || !ctx || !function || n_args < 2 || !args || nsteps < 0) return NEPTUNE_RT_EINVAL;
  const Function *fn = m->mod.find(function);
  if (!fn) { m->err = std::string("no function ") + function; return NEPTUNE_RT_EINVAL; }
  if (fn->rejected) {
    m->err = std::string(function) + " was not lowered by neptuneir-to-b200: " + fn->diagnostic;
    return fn->unsupported ? NEPTUNE_RT_EUNSUPPORTED : NEPTUNE_RT_EINVAL;
  }
  if (fn->kind != "func" || (int)fn->args.size() != n_args) { m->err = "argument count mismatch"; return NEPTUNE_RT_EINVAL; }
  const FnInfo &fi = m->fns.at(fn->name);
  std::vector<neptune_rt_field *> *fields = nullptr;
  if (int rc = bind_args(*m, ctx, *fn, &fields)) return rc;
  if (!(*fields)[0] || !(*fields)[1] || fi.arg_field[0] != fi.arg_field[1]) {
    m->err = "run_steps needs entry(out, in, ...) with out and in wrapped as the same field type";
    return NEPTUNE_RT_EINVAL;
  }
  // the swapped host loop reads both buffers over time: upload every argument the entry reads,
  // and arg 0 too when the entry reads it before overwriting it
  for (int i = 0; i < n_args; ++i)
    if ((*fields)[i] && (fi.arg_needs_upload[i] || (i == 1 && fi.arg_needs_upload[0])))
      if (int rc = neptune_rt_field_upload((*fields)[i], args[i])) { m->err = neptune_rt_last_error(ctx); return rc; }
  Exec ex{*m, ctx, m->per_ctx[ctx], {}};
  // step t writes host argument (t even ? 0 : 1)
  const int res = nsteps == 0 ? 1 : ((nsteps - 1) % 2 == 0 ? 0 : 1);
  neptune_rt_field *last = (*fields)[res];
  if (fi.pipe_apply >= 0 && fi.pipe_out_arg == 0 && !fi.pipe_in_args.empty() && fi.pipe_in_args[0] == 1) {
    // entry(out, in, extra...) = ONE apply of `in` (+ pointwise/coefficient arguments) stored whole into `out`: the
    // swapped host loop is exactly neptune_rt_run_steps, which fuses pairs of steps into temporally blocked
    // launches where the body has such a kernel (bitwise the same as nsteps single sweeps)
    int rc = 0;
    neptune_rt_stencil *st = ex.stencil(m->apply_index.at(fn->ops[fi.pipe_apply].apply.get()), &rc);
    if (!st) return rc;
    std::vector<neptune_rt_field *> extra;
    for (size_t k = 1; k < fi.pipe_in_args.size(); ++k) extra.push_back((*fields)[fi.pipe_in_args[k]]);
    bool distinct = true;   // a time loop replaces input 0 only: the other inputs must be other arguments
    for (size_t k = 1; k < fi.pipe_in_args.size(); ++k) distinct &= fi.pipe_in_args[k] >= 2;
    if (distinct) {
      if ((rc = neptune_rt_run_steps(st, (*fields)[1], (*fields)[0], (int)extra.size(), extra.data(), nsteps, &last))) {
        m->err = neptune_rt_last_error(ctx);
        return rc;
      }
      if (nsteps > 0)
        if ((rc = neptune_rt_field_download(last, args[res]))) { m->err = neptune_rt_last_error(ctx); return rc; }
      if (result_arg) *result_arg = res;
      if ((rc = neptune_rt_sync(ctx))) { m->err = neptune_rt_last_error(ctx); return rc; }
      return NEPTUNE_RT_OK;
    }
  }
  std::vector<neptune_rt_field *> cur = *fields;
  for (int64_t t = 0; t < nsteps; ++t) {
    ex.pc.cursor.clear();
    if (int rc = ex.run(*fn, cur, nullptr)) return rc;
    std::swap(cur[0], cur[1]);
  }
  if (nsteps > 0)
    if (int rc = neptune_rt_field_download(last, args[res])) { m->err = neptune_rt_last_error(ctx); return rc; }
  if (result_arg) *result_arg = res;
  if (int rc = neptune_rt_sync(ctx)) { m->err = neptune_rt_last_error(ctx); return rc; }
  return NEPTUNE_RT_OK;
}
