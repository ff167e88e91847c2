// Parser for the textual NeptuneIR subset on the `neptune_ir.apply` hot path.
//
// Grammar = the reference's assembly formats:
//   wrap/unwrap/load     include/Dialect/NeptuneIR/NeptuneIROps.td:20-84
//   yield                NeptuneIROps.td:94-114
//   apply                NeptuneIROps.td:164-197   (+ region contract of ApplyOp::verify,
//                        lib/Dialect/NeptuneIR/NeptuneIRVerifier.cpp:141-171)
//   access               NeptuneIROps.td:205-228
//   store                NeptuneIROps.td:238-259
//   linear_opdef / nonlinear_opdef / apply_linear / return: callers of the hot path, which
//                        lib/Passes/StructureLowering.cpp:31-140 turns into func.func / func.call
//   arith.constant/addf/subf/mulf/divf/negf: float part of the body whitelist,
//                        lib/Passes/VerifyAndAnnotate.cpp:167-204
// Anything else is reported as unsupported: `neptuneir-to-b200` has no CPU fallback.
#include <cctype>
#include <cmath>
#include <cstring>
#include <sstream>

#include "ir.h"

namespace neptune {

std::string Type::str() const {
  std::ostringstream o;
  auto el = [&](int dt) { return dt == NEPTUNE_RT_F64 ? "f64" : dt == NEPTUNE_RT_F32 ? "f32" : "?"; };
  switch (kind) {
    case Float: o << el(dtype); break;
    case Index: o << "index"; break;
    case Int: o << "int"; break;
    case Grid: {
      o << "!neptune_ir." << (is_field ? "field" : "temp") << "<element = " << el(dtype) << ", bounds = [";
      for (int d = 0; d < bounds.rank(); ++d) o << (d ? "," : "") << bounds.lb[d];
      o << "]..[";
      for (int d = 0; d < bounds.rank(); ++d) o << (d ? "," : "") << bounds.ub[d];
      o << "]>";
      break;
    }
    case MemRef: {
      o << "memref<";
      for (auto s : shape) { if (s < 0) o << "?x"; else o << s << "x"; }
      o << el(dtype) << ">";
      break;
    }
    default: o << "<none>";
  }
  return o.str();
}

namespace {

struct Tok {
  enum Kind { Eof, Id, Num, Str, Word, Punct, Arrow } kind = Eof;
  std::string text;
  int line = 0;
};

class Lexer {
 public:
  explicit Lexer(const std::string &s) : s_(s) {}
  const Tok &peek() {
    if (!have_) { cur_ = lex(); have_ = true; }
    return cur_;
  }
  Tok next() {
    Tok t = peek();
    have_ = false;
    return t;
  }
  // raw scan up to the matching '>' (used for memref<...>), starting right after '<'
  std::string raw_until_gt() {
    have_ = false;  // caller guarantees the pending token was consumed
    size_t b = pos_;
    while (pos_ < s_.size() && s_[pos_] != '>') ++pos_;
    if (pos_ >= s_.size()) throw Diagnostic("unterminated memref type");
    std::string r = s_.substr(b, pos_ - b);
    ++pos_;
    return r;
  }
  int line() const { return line_; }

 private:
  static bool idch(char c) { return std::isalnum((unsigned char)c) || c == '_' || c == '.' || c == '$' || c == '-'; }
  void skip() {
    for (;;) {
      while (pos_ < s_.size() && std::isspace((unsigned char)s_[pos_])) {
        if (s_[pos_] == '\n') ++line_;
        ++pos_;
      }
      if (pos_ + 1 < s_.size() && s_[pos_] == '/' && s_[pos_ + 1] == '/') {
        while (pos_ < s_.size() && s_[pos_] != '\n') ++pos_;
        continue;
      }
      break;
    }
  }
  Tok lex() {
    skip();
    Tok t;
    t.line = line_;
    if (pos_ >= s_.size()) return t;
    char c = s_[pos_];
    if (c == '-' && pos_ + 1 < s_.size() && s_[pos_ + 1] == '>') {
      t.kind = Tok::Arrow; t.text = "->"; pos_ += 2; return t;
    }
    if (c == '"') {
      size_t e = s_.find('"', pos_ + 1);
      if (e == std::string::npos) throw Diagnostic("unterminated string literal at line " + std::to_string(line_));
      t.kind = Tok::Str; t.text = s_.substr(pos_ + 1, e - pos_ - 1); pos_ = e + 1; return t;
    }
    if (c == '%' || c == '@' || c == '^' || c == '#' || c == '!') {
      size_t b = pos_++;
      while (pos_ < s_.size() && idch(s_[pos_])) ++pos_;
      // an identifier never ends in '-' or '.' (e.g. "%a-" is "%a" then '-')
      while (pos_ > b + 1 && (s_[pos_ - 1] == '-' || s_[pos_ - 1] == '.')) --pos_;
      t.kind = Tok::Id; t.text = s_.substr(b, pos_ - b); return t;
    }
    bool neg = (c == '-' || c == '+');
    if (std::isdigit((unsigned char)c) || (c == '.' && pos_ + 1 < s_.size() && std::isdigit((unsigned char)s_[pos_ + 1])) ||
        (neg && pos_ + 1 < s_.size() && (std::isdigit((unsigned char)s_[pos_ + 1]) || s_[pos_ + 1] == '.'))) {
      size_t b = pos_;
      if (neg) ++pos_;
      if (pos_ + 1 < s_.size() && s_[pos_] == '0' && (s_[pos_ + 1] == 'x' || s_[pos_ + 1] == 'X')) {
        pos_ += 2;
        while (pos_ < s_.size() && std::isxdigit((unsigned char)s_[pos_])) ++pos_;
      } else {
        while (pos_ < s_.size() && std::isdigit((unsigned char)s_[pos_])) ++pos_;
        if (pos_ < s_.size() && s_[pos_] == '.') {
          ++pos_;
          while (pos_ < s_.size() && std::isdigit((unsigned char)s_[pos_])) ++pos_;
        }
        if (pos_ < s_.size() && (s_[pos_] == 'e' || s_[pos_] == 'E')) {
          size_t save = pos_;
          ++pos_;
          if (pos_ < s_.size() && (s_[pos_] == '+' || s_[pos_] == '-')) ++pos_;
          if (pos_ < s_.size() && std::isdigit((unsigned char)s_[pos_])) {
            while (pos_ < s_.size() && std::isdigit((unsigned char)s_[pos_])) ++pos_;
          } else {
            pos_ = save;
          }
        }
      }
      t.kind = Tok::Num; t.text = s_.substr(b, pos_ - b); return t;
    }
    if (std::isalpha((unsigned char)c) || c == '_') {
      size_t b = pos_;
      while (pos_ < s_.size() && (std::isalnum((unsigned char)s_[pos_]) || s_[pos_] == '_' || s_[pos_] == '.')) ++pos_;
      t.kind = Tok::Word; t.text = s_.substr(b, pos_ - b); return t;
    }
    t.kind = Tok::Punct; t.text = std::string(1, c); ++pos_;
    return t;
  }
  const std::string &s_;
  size_t pos_ = 0;
  int line_ = 1;
  bool have_ = false;
  Tok cur_;

 public:
  struct State { size_t pos; int line; bool have; Tok cur; };
  State save() const { return State{pos_, line_, have_, cur_}; }
  void restore(const State &st) { pos_ = st.pos; line_ = st.line; have_ = st.have; cur_ = st.cur; }
};

struct AttrValue {
  enum Kind { BoundsK, LocationK, ShapeK } kind;
  Bounds bounds;
  std::string location;
  std::vector<std::vector<int64_t>> shape;
};

class Parser {
 public:
  explicit Parser(const std::string &text) : lx_(text) {}

  Module parse() {
    Module m;
    for (;;) {
      const Tok &t = lx_.peek();
      if (t.kind == Tok::Eof) break;
      if (t.kind == Tok::Id && (t.text[0] == '#' || t.text[0] == '!') && t.text.find("neptune_ir.") != 1) {
        std::string name = lx_.next().text;
        expect("=");
        if (name[0] == '#') attr_alias_[name] = parse_attr_value();
        else type_alias_[name] = parse_type();
        continue;
      }
      if (t.kind == Tok::Word && t.text == "module") {
        lx_.next();
        if (accept_word("attributes")) skip_braces();
        expect("{");
        while (!accept("}")) m.functions.push_back(parse_function());
        continue;
      }
      m.functions.push_back(parse_function());
    }
    return m;
  }

 private:
  [[noreturn]] void fail(const std::string &msg, bool unsupported = false) {
    throw Diagnostic("line " + std::to_string(lx_.peek().line) + ": " + msg, unsupported);
  }
  bool accept(const char *p) {
    const Tok &t = lx_.peek();
    if ((t.kind == Tok::Punct || t.kind == Tok::Arrow) && t.text == p) { lx_.next(); return true; }
    return false;
  }
  bool accept_word(const char *w) {
    const Tok &t = lx_.peek();
    if (t.kind == Tok::Word && t.text == w) { lx_.next(); return true; }
    return false;
  }
  void expect(const char *p) {
    if (!accept(p)) fail(std::string("expected '") + p + "', got '" + lx_.peek().text + "'");
  }
  void expect_word(const char *w) {
    if (!accept_word(w)) fail(std::string("expected '") + w + "', got '" + lx_.peek().text + "'");
  }
  std::string expect_id(char prefix) {
    Tok t = lx_.next();
    if (t.kind != Tok::Id || t.text[0] != prefix)
      throw Diagnostic("line " + std::to_string(t.line) + ": expected " + std::string(1, prefix) + "-identifier, got '" + t.text + "'");
    return t.text;
  }
  int64_t parse_int() {
    Tok t = lx_.next();
    if (t.kind != Tok::Num) throw Diagnostic("line " + std::to_string(t.line) + ": expected integer, got '" + t.text + "'");
    return std::stoll(t.text, nullptr, 0);
  }
  std::vector<int64_t> parse_int_list() {
    std::vector<int64_t> v;
    expect("[");
    if (accept("]")) return v;
    for (;;) {
      v.push_back(parse_int());
      if (accept("]")) break;
      expect(",");
    }
    return v;
  }
  void skip_braces() {
    expect("{");
    int depth = 1;
    while (depth > 0) {
      Tok t = lx_.next();
      if (t.kind == Tok::Eof) fail("unterminated '{'");
      if (t.kind == Tok::Punct && t.text == "{") ++depth;
      if (t.kind == Tok::Punct && t.text == "}") --depth;
    }
  }

  // ---- attributes ---------------------------------------------------------------------------
  Bounds parse_bounds() {
    Tok t = lx_.next();
    if (t.kind == Tok::Id && attr_alias_.count(t.text)) {
      const AttrValue &a = attr_alias_[t.text];
      if (a.kind != AttrValue::BoundsK) fail(t.text + " is not a bounds attribute");
      return a.bounds;
    }
    if (t.text != "#neptune_ir.bounds") fail("expected #neptune_ir.bounds, got '" + t.text + "'");
    Bounds b;
    expect("<"); expect_word("lb"); expect("="); b.lb = parse_int_list();
    expect(","); expect_word("ub"); expect("="); b.ub = parse_int_list();
    expect(">");
    if (b.lb.size() != b.ub.size()) fail("bounds lb/ub rank mismatch");
    return b;
  }
  std::string parse_location() {
    Tok t = lx_.next();
    if (t.kind == Tok::Id && attr_alias_.count(t.text)) {
      const AttrValue &a = attr_alias_[t.text];
      if (a.kind != AttrValue::LocationK) fail(t.text + " is not a location attribute");
      return a.location;
    }
    if (t.text != "#neptune_ir.location") fail("expected #neptune_ir.location, got '" + t.text + "'");
    expect("<");
    Tok s = lx_.next();
    if (s.kind != Tok::Str) fail("expected location string");
    expect(">");
    return s.text;
  }
  std::vector<std::vector<int64_t>> parse_shape() {
    Tok t = lx_.next();
    if (t.kind == Tok::Id && attr_alias_.count(t.text)) {
      const AttrValue &a = attr_alias_[t.text];
      if (a.kind != AttrValue::ShapeK) fail(t.text + " is not a stencil_shape attribute");
      return a.shape;
    }
    if (t.text != "#neptune_ir.stencil_shape") fail("expected #neptune_ir.stencil_shape, got '" + t.text + "'");
    std::vector<std::vector<int64_t>> offs;
    expect("<"); expect("[");
    while (!accept("]")) {
      if (accept_word("array")) {           // array<i64: 1, 0>
        expect("<"); lx_.next(); expect(":");
        std::vector<int64_t> v;
        while (!accept(">")) { v.push_back(parse_int()); accept(","); }
        offs.push_back(v);
      } else {
        offs.push_back(parse_int_list());
      }
      accept(",");
    }
    expect(">");
    return offs;
  }
  AttrValue parse_attr_value() {
    AttrValue a;
    const std::string &t = lx_.peek().text;
    if (t == "#neptune_ir.bounds") { a.kind = AttrValue::BoundsK; a.bounds = parse_bounds(); }
    else if (t == "#neptune_ir.location") { a.kind = AttrValue::LocationK; a.location = parse_location(); }
    else if (t == "#neptune_ir.stencil_shape") { a.kind = AttrValue::ShapeK; a.shape = parse_shape(); }
    else if (attr_alias_.count(t)) { a = attr_alias_[t]; lx_.next(); }
    else fail("unsupported attribute '" + t + "'", true);
    return a;
  }

  // ---- types --------------------------------------------------------------------------------
  static int elem_of(const std::string &w) { return w == "f64" ? NEPTUNE_RT_F64 : w == "f32" ? NEPTUNE_RT_F32 : -1; }
  Type parse_type() {
    Tok t = lx_.next();
    Type ty;
    if (t.kind == Tok::Id && t.text[0] == '!') {
      if (type_alias_.count(t.text)) return type_alias_[t.text];
      if (t.text == "!neptune_ir.temp" || t.text == "!neptune_ir.field") {
        ty.kind = Type::Grid;
        ty.is_field = t.text == "!neptune_ir.field";
        expect("<"); expect_word("element"); expect("=");
        Type el = parse_type();
        if (el.kind != Type::Float) fail("field/temp element must be f32 or f64 on the B200 path", true);
        ty.dtype = el.dtype;
        expect(","); expect_word("bounds"); expect("="); ty.bounds = parse_bounds();
        ty.location = "cell";
        if (accept(",")) { expect_word("location"); expect("="); ty.location = parse_location(); }
        expect(">");
        return ty;
      }
      fail("unknown type '" + t.text + "'", true);
    }
    if (t.kind == Tok::Word) {
      if (elem_of(t.text) >= 0) { ty.kind = Type::Float; ty.dtype = elem_of(t.text); return ty; }
      if (t.text == "index") { ty.kind = Type::Index; return ty; }
      if (t.text == "i1" || t.text == "i8" || t.text == "i16" || t.text == "i32" || t.text == "i64") { ty.kind = Type::Int; return ty; }
      if (t.text == "memref") {
        // the lexer has not consumed beyond "memref": next raw char must be '<'
        const Tok &lt = lx_.peek();
        if (!(lt.kind == Tok::Punct && lt.text == "<")) fail("expected '<' after memref");
        lx_.next();
        std::string raw = lx_.raw_until_gt();
        std::string clean;
        for (char c : raw) if (!std::isspace((unsigned char)c)) clean += c;
        ty.kind = Type::MemRef;
        size_t p = 0;
        for (;;) {
          size_t x = clean.find('x', p);
          if (x == std::string::npos) break;
          std::string d = clean.substr(p, x - p);
          if (d == "?") ty.shape.push_back(-1);
          else if (!d.empty() && std::isdigit((unsigned char)d[0])) ty.shape.push_back(std::stoll(d));
          else break;
          p = x + 1;
        }
        ty.dtype = elem_of(clean.substr(p));
        if (ty.dtype < 0) fail("unsupported memref element type in 'memref<" + raw + ">'", true);
        return ty;
      }
    }
    fail("unsupported type '" + t.text + "'", true);
  }

  // ---- functions ----------------------------------------------------------------------------
  Function parse_function() {
    Tok t = lx_.next();
    Function fn;
    if (t.kind == Tok::Word && t.text == "func.func") {
      if (accept_word("private") || accept_word("public")) {}
      fn.kind = "func";
      fn.name = expect_id('@');
      expect("(");
      if (!accept(")")) {
        for (;;) {
          std::string a = expect_id('%');
          expect(":");
          fn.args.emplace_back(a, parse_type());
          if (accept(")")) break;
          expect(",");
        }
      }
      if (accept("->")) {
        if (accept("(")) {
          while (!accept(")")) { fn.results.push_back(parse_type()); accept(","); }
        } else {
          fn.results.push_back(parse_type());
        }
      }
      if (accept_word("attributes")) skip_braces();
      expect("{");
      parse_function_body(fn);
      return fn;
    }
    if (t.kind == Tok::Word && (t.text == "neptune_ir.linear_opdef" || t.text == "neptune_ir.nonlinear_opdef")) {
      fn.kind = t.text.substr(std::strlen("neptune_ir."));
      fn.name = expect_id('@');
      expect(":"); expect("(");
      std::vector<Type> in_tys;
      while (!accept(")")) { in_tys.push_back(parse_type()); accept(","); }
      expect("->");
      fn.results.push_back(parse_type());
      expect("{");
      expect_id('^');
      expect("(");
      while (!accept(")")) {
        std::string a = expect_id('%');
        expect(":");
        fn.args.emplace_back(a, parse_type());
        accept(",");
      }
      expect(":");
      if (fn.args.size() != in_tys.size()) fail("opdef block argument count differs from its signature");
      for (size_t i = 0; i < in_tys.size(); ++i)
        if (fn.args[i].second != in_tys[i]) fail("opdef block argument types differ from its signature");
      parse_function_body(fn);
      return fn;
    }
    throw Diagnostic("line " + std::to_string(t.line) + ": unsupported top-level operation '" + t.text + "'", true);
  }

  // The operations of a function up to its closing brace.  A function that holds something outside the hot path
  // (neptune_ir.time_advance, solve_linear, a body with control flow ...) is kept
  // as a REJECTED function carrying the diagnostic, and parsing resumes after its closing brace: the other
  // functions of the module — the operator definitions a time_advance refers to, for one — are still lowered.
  // (The reference's conversion fails the whole module; a module none of whose functions lowers still fails here.)
  void parse_function_body(Function &fn) {
    const Lexer::State at_open = lx_.save();
    try {
      while (!accept("}")) fn.ops.push_back(parse_host_op());
    } catch (const Diagnostic &d) {
      // only IR that is well formed but off the hot path is set aside; a syntax error or a broken verifier rule fails the
      // whole module, as the reference's verifier and applyFullConversion do (VerifyAndAnnotate.cpp, DataflowLowering.cpp:799-816)
      if (!d.unsupported) throw;
      fn.ops.clear();
      fn.rejected = true;
      fn.unsupported = d.unsupported;
      fn.diagnostic = d.what();
      lx_.restore(at_open);
      for (int depth = 1; depth > 0;) {
        Tok t = lx_.next();
        if (t.kind == Tok::Eof) throw Diagnostic(fn.diagnostic + " (and the function's closing brace is missing)", d.unsupported);
        if (t.kind == Tok::Punct && t.text == "{") ++depth;
        if (t.kind == Tok::Punct && t.text == "}") --depth;
      }
    }
  }

  HostOp parse_host_op() {
    HostOp op;
    op.line = lx_.peek().line;
    if (lx_.peek().kind == Tok::Id && lx_.peek().text[0] == '%') {
      op.results.push_back(lx_.next().text);
      expect("=");
    }
    Tok name = lx_.next();
    const std::string &n = name.text;
    if (n == "neptune_ir.wrap" || n == "neptune_ir.unwrap" || n == "neptune_ir.load") {
      op.kind = n == "neptune_ir.wrap" ? HostOp::Wrap : n == "neptune_ir.unwrap" ? HostOp::Unwrap : HostOp::Load;
      op.operands.push_back(expect_id('%'));
      expect(":");
      op.types.push_back(parse_type());
      expect("->");
      op.types.push_back(parse_type());
      return op;
    }
    if (n == "neptune_ir.store") {
      op.kind = HostOp::Store;
      op.operands.push_back(expect_id('%'));
      expect_word("to");
      op.operands.push_back(expect_id('%'));
      if (accept("{")) {
        expect_word("bounds"); expect("=");
        op.bounds = parse_bounds();
        op.has_bounds = true;
        expect("}");
      }
      expect(":");
      op.types.push_back(parse_type());
      expect_word("to");
      op.types.push_back(parse_type());
      return op;
    }
    if (n == "neptune_ir.apply_linear" || n == "neptune_ir.apply_nonlinear") {
      op.kind = HostOp::ApplyLinear;
      op.callee = expect_id('@');
      expect("(");
      while (!accept(")")) { op.operands.push_back(expect_id('%')); accept(","); }
      expect(":"); expect("(");
      while (!accept(")")) { op.types.push_back(parse_type()); accept(","); }
      expect("->");
      op.types.push_back(parse_type());
      return op;
    }
    if (n == "func.return" || n == "return" || n == "neptune_ir.return") {
      op.kind = HostOp::Return;
      if (lx_.peek().kind == Tok::Id && lx_.peek().text[0] == '%') {
        for (;;) {
          op.operands.push_back(expect_id('%'));
          if (!accept(",")) break;
        }
        expect(":");
        op.types.push_back(parse_type());
        while (accept(",")) op.types.push_back(parse_type());
      }
      return op;
    }
    if (n == "neptune_ir.reduce") {  // $input (`in` $bounds)? attr-dict : type -> type  (NeptuneIROps.td:272-299)
      op.kind = HostOp::Reduce;
      if (op.results.size() != 1) fail("neptune_ir.reduce must have exactly one result");
      op.operands.push_back(expect_id('%'));
      if (lx_.peek().kind == Tok::Word && lx_.peek().text == "in") {
        lx_.next();
        op.bounds = parse_bounds();
        op.has_bounds = true;
      }
      if (accept("{")) {
        expect_word("kind"); expect("=");
        Tok k = lx_.next();
        if (k.kind != Tok::Str) fail("reduce kind must be a string");
        op.callee = k.text;
        expect("}");
      }
      expect(":");
      op.types.push_back(parse_type());
      expect("->");
      op.types.push_back(parse_type());
      if (op.callee != "sum")
        throw Diagnostic("line " + std::to_string(name.line) + ": neptune_ir.reduce kind \"" + op.callee +
                             "\" is not lowered (the reference lowers only \"sum\", DataflowLowering.cpp:655-657)", true);
      return op;
    }
    if (n == "neptune_ir.apply") {
      op.kind = HostOp::ApplyOp;
      if (op.results.size() != 1) fail("neptune_ir.apply must have exactly one result");
      op.apply = std::make_shared<Apply>(parse_apply(op.results[0], name.line));
      op.operands = op.apply->inputs;   // so that use counts see the apply's inputs
      return op;
    }
    if (n == "arith.constant") {  // host-level scalar (e.g. a dt); never reaches the sweep
      op.kind = HostOp::Constant;
      lx_.next();
      expect(":");
      parse_type();
      return op;
    }
    throw Diagnostic("line " + std::to_string(name.line) + ": operation '" + n +
                         "' is not part of the neptuneir-to-b200 hot path (wrap/unwrap/load/store/apply/"
                         "access/yield); there is no CPU fallback",
                     true);
  }

  double parse_float_literal(int dtype) {
    Tok t = lx_.next();
    if (t.kind == Tok::Word) {
      if (t.text == "inf") return INFINITY;
      if (t.text == "nan") return NAN;
    }
    if (t.kind != Tok::Num) throw Diagnostic("line " + std::to_string(t.line) + ": expected float literal, got '" + t.text + "'");
    const std::string &s = t.text;
    double v;
    if (s.find("0x") != std::string::npos || s.find("0X") != std::string::npos) {
      // MLIR prints non-round-trippable floats as the hex bit pattern of the type
      unsigned long long bits = std::stoull(s, nullptr, 16);
      if (dtype == NEPTUNE_RT_F32) { uint32_t b = (uint32_t)bits; float f; std::memcpy(&f, &b, 4); v = f; }
      else std::memcpy(&v, &bits, 8);
    } else if (dtype == NEPTUNE_RT_F32) {
      v = (double)std::strtof(s.c_str(), nullptr);   // one rounding, decimal -> f32, as APFloat does (not decimal -> f64 -> f32)
    } else {
      v = std::strtod(s.c_str(), nullptr);
    }
    if (dtype == NEPTUNE_RT_F32) v = (double)(float)v;
    return v;
  }

  Apply parse_apply(const std::string &result, int line) {
    Apply ap;
    ap.result = result;
    ap.line = line;
    expect("(");
    while (!accept(")")) { ap.inputs.push_back(expect_id('%')); accept(","); }
    expect_word("attributes");
    expect("{");
    bool have_bounds = false;
    while (!accept("}")) {
      Tok k = lx_.next();
      expect("=");
      if (k.text == "bounds") { ap.bounds = parse_bounds(); have_bounds = true; }
      else if (k.text == "shape") { ap.shape = parse_shape(); ap.has_shape = true; }
      else fail("unsupported apply attribute '" + k.text + "'", true);
      accept(",");
    }
    if (!have_bounds) fail("neptune_ir.apply is missing its required 'bounds' attribute");
    expect(":"); expect("(");
    while (!accept(")")) { ap.in_types.push_back(parse_type()); accept(","); }
    expect("->");
    ap.out_type = parse_type();
    if (ap.in_types.size() != ap.inputs.size()) fail("apply operand/type count mismatch");
    expect("{");
    expect_id('^');
    expect("(");
    std::vector<std::pair<std::string, Type>> bargs;
    while (!accept(")")) {
      std::string a = expect_id('%');
      expect(":");
      bargs.emplace_back(a, parse_type());
      accept(",");
    }
    expect(":");
    // ApplyOp::verify, NeptuneIRVerifier.cpp:141-171
    const int rank = ap.bounds.rank();
    if (rank == 0) fail("0-D apply not supported");
    if ((int)bargs.size() != rank + (int)ap.inputs.size())
      fail("apply-like region block arg count must be (bounds rank + number of inputs) = " +
           std::to_string(rank + (int)ap.inputs.size()) + ", but got " + std::to_string(bargs.size()));
    for (int d = 0; d < rank; ++d)
      if (bargs[d].second.kind != Type::Index) fail("region arg #" + std::to_string(d) + " must be index");
    for (size_t k = 0; k < ap.inputs.size(); ++k)
      if (bargs[rank + k].second != ap.in_types[k]) fail("region input arg #" + std::to_string(rank + k) + " type mismatch");

    std::map<std::string, int> vn;       // body SSA name -> value number
    std::map<std::string, int> argidx;   // temp block arg -> input index
    for (size_t k = 0; k < ap.inputs.size(); ++k) argidx[bargs[rank + k].first] = (int)k;
    auto value = [&](const std::string &name) -> int {
      auto it = vn.find(name);
      if (it == vn.end()) {
        for (int d = 0; d < rank; ++d)
          if (bargs[d].first == name)
            throw Diagnostic("line " + std::to_string(lx_.peek().line) + ": body uses the index argument " + name +
                                 " as a value; index-dependent bodies are outside the recognised stencil set", true);
        fail("use of undefined value " + name + " in apply body (regions are IsolatedFromAbove)");
      }
      return it->second;
    };
    for (;;) {
      const Tok &t = lx_.peek();
      if (t.kind == Tok::Word && t.text == "neptune_ir.yield") {
        lx_.next();
        std::string v = expect_id('%');
        ap.yield_value = value(v);
        if (accept(",")) fail("neptune_ir.yield with several values is not supported (DataflowLowering.cpp:427-429)", true);
        expect(":");
        Type yt = parse_type();
        if (yt.kind != Type::Float) fail("yield of a non-float value", true);
        ap.yield_dtype = yt.dtype;
        expect("}");
        break;
      }
      std::string res = expect_id('%');
      expect("=");
      Tok opn = lx_.next();
      neptune_rt_body_op bo;
      std::memset(&bo, 0, sizeof bo);
      bo.a = bo.b = bo.input = -1;
      int dt = -1;
      const std::string &o = opn.text;
      if (o == "neptune_ir.access") {
        std::string src = expect_id('%');
        auto it = argidx.find(src);
        if (it == argidx.end()) fail("neptune_ir.access operand " + src + " is not a temp block argument of the apply");
        std::vector<int64_t> offs = parse_int_list();
        expect(":");
        Type st = parse_type();
        expect("->");
        Type rt = parse_type();
        if (st != ap.in_types[it->second]) fail("access source type differs from the block argument type");
        if ((int)offs.size() != rank) fail("access offset rank " + std::to_string(offs.size()) + " differs from the apply rank");
        if (rt.kind != Type::Float || rt.dtype != st.dtype) fail("access result type must be the temp's element type");
        bo.opcode = NEPTUNE_RT_OP_ACCESS;
        bo.input = it->second;
        for (int d = 0; d < rank; ++d) bo.offset[d] = offs[d];
        dt = rt.dtype;
      } else if (o == "arith.constant") {
        // literal is rounded to the constant's type after the type is known: read it raw first
        Tok lit = lx_.peek();
        // parse with f64 precision now, round below
        double v = parse_float_literal(NEPTUNE_RT_F64);
        expect(":");
        Type ct = parse_type();
        if (ct.kind != Type::Float) fail("integer/index constants in an apply body are outside the recognised stencil set", true);
        if (ct.dtype == NEPTUNE_RT_F32) {
          if (lit.text.find("0x") != std::string::npos) {
            uint32_t b = (uint32_t)std::stoull(lit.text, nullptr, 16); float f; std::memcpy(&f, &b, 4); v = f;
          } else {
            v = (double)std::strtof(lit.text.c_str(), nullptr);
          }
        }
        bo.opcode = NEPTUNE_RT_OP_CONST;
        bo.value = v;
        dt = ct.dtype;
      } else if (o == "arith.addf" || o == "arith.subf" || o == "arith.mulf" || o == "arith.divf") {
        std::string a = expect_id('%');
        expect(",");
        std::string b = expect_id('%');
        if (accept_word("fastmath")) fail("fastmath flags are not supported on the bit-exact path", true);
        expect(":");
        Type ct = parse_type();
        if (ct.kind != Type::Float) fail("non-float arithmetic type", true);
        bo.opcode = o == "arith.addf" ? NEPTUNE_RT_OP_ADD : o == "arith.subf" ? NEPTUNE_RT_OP_SUB
                  : o == "arith.mulf" ? NEPTUNE_RT_OP_MUL : NEPTUNE_RT_OP_DIV;
        bo.a = value(a);
        bo.b = value(b);
        dt = ct.dtype;
      } else if (o == "arith.negf") {
        std::string a = expect_id('%');
        expect(":");
        Type ct = parse_type();
        if (ct.kind != Type::Float) fail("non-float arithmetic type", true);
        bo.opcode = NEPTUNE_RT_OP_NEG;
        bo.a = value(a);
        dt = ct.dtype;
      } else {
        throw Diagnostic("line " + std::to_string(opn.line) + ": operation '" + o +
                             "' in an apply body is outside the recognised stencil set (access, arith.constant/"
                             "addf/subf/mulf/divf/negf); neptuneir-to-b200 has no CPU fallback",
                         true);
      }
      if (vn.count(res)) fail("redefinition of " + res);
      vn[res] = (int)ap.body.size();
      ap.body.push_back(bo);
      ap.body_dtype.push_back(dt);
    }
    return ap;
  }

  Lexer lx_;
  std::map<std::string, AttrValue> attr_alias_;
  std::map<std::string, Type> type_alias_;
};

}  // namespace

Module parse_module(const std::string &text) { return Parser(text).parse(); }

}  // namespace neptune
