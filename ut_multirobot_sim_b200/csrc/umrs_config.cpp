// Config ingestion without a Lua interpreter.
//
// The reference reads its parameters through config_reader (a Lua state loaded
// with a list of files, globals looked up by dotted name: lua_script.h:82-119,
// CONFIG_FLOAT/INT/BOOL/STRING/STRINGLIST/VECTOR3FLIST macros, e.g.
// simulator.cpp:65-97, ackermann_model.cpp:18-23, diff_drive_model.cpp:43-55,
// omnidirectional_model.cpp:21-31, human_object.cpp:30-43).  liblua is not
// available here, and the files under config/*.lua only use a small, flat subset
// of the language; this evaluator covers that subset:
//   * `name = expr`, `a.b = expr`, `local name = expr`, optional `;`
//   * `function Name(args) return expr end`  (single-return helpers: Vector2,
//     Vector3, DegToRad -- sim_config.lua:1-11)
//   * `require("config.sim_config")` (ut_automata_config.lua:1): the dotted module
//     path is resolved against the root directory like Lua's package.path "./?.lua"
//   * expressions: numbers, strings, true/false/nil, table constructors
//     ({a, b}, {k = v}, {["k"] = v}), calls, a.b / a[i], unary - and not,
//     + - * / % ^ .. with Lua's precedence, parentheses, math.pi and a few math.*
// Numbers are doubles as in Lua 5.1; a float key is the double narrowed to float
// (config_reader's CONFIG_FLOAT does the same cast).  Files are loaded in order
// into one global table, later files overriding earlier ones, like
// ConfigReader({env_config, robot_config}) (simulator.cpp:106).
// Anything outside the subset is a load error (UMRS_ERR_UNSUPPORTED), never a
// silent default.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/umrs.h"

namespace {

struct Value;
typedef std::shared_ptr<Value> ValuePtr;
struct Node;
typedef std::shared_ptr<Node> NodePtr;

enum Kind { kNil, kBool, kNumber, kString, kTable, kFunction, kBuiltin };

struct Value {
  Kind kind = kNil;
  bool b = false;
  double num = 0.0;
  std::string str;
  // table: positional part (1-based in Lua) and string-keyed part
  std::vector<ValuePtr> arr;
  std::map<std::string, ValuePtr> fields;
  // function: parameter names + the single returned expression
  std::vector<std::string> params;
  NodePtr body;
  int builtin = 0;
};

ValuePtr MakeNil() { return std::make_shared<Value>(); }
ValuePtr MakeNum(double d) { auto v = std::make_shared<Value>(); v->kind = kNumber; v->num = d; return v; }
ValuePtr MakeBool(bool b) { auto v = std::make_shared<Value>(); v->kind = kBool; v->b = b; return v; }
ValuePtr MakeStr(const std::string& s) { auto v = std::make_shared<Value>(); v->kind = kString; v->str = s; return v; }
ValuePtr MakeTable() { auto v = std::make_shared<Value>(); v->kind = kTable; return v; }

// ---- lexer ----------------------------------------------------------------------
enum Tok { tEof, tName, tNumber, tString, tOp };

struct Token {
  Tok t = tEof;
  std::string s;
  double num = 0.0;
  int line = 0;
};

struct ParseError {
  std::string msg;
};

class Lexer {
 public:
  Lexer(const std::string& src, const std::string& file) : s_(src), file_(file) {}

  std::vector<Token> Run() {
    std::vector<Token> out;
    for (;;) {
      Token t = Next();
      out.push_back(t);
      if (t.t == tEof) break;
    }
    return out;
  }

 private:
  [[noreturn]] void Fail(const std::string& m) {
    throw ParseError{file_ + ":" + std::to_string(line_) + ": " + m};
  }
  bool LongBracket(std::string* body) {
    // at '[' : [[ ... ]] or [=[ ... ]=]
    size_t p = i_ + 1, eq = 0;
    while (p < s_.size() && s_[p] == '=') { ++p; ++eq; }
    if (p >= s_.size() || s_[p] != '[') return false;
    const std::string close = "]" + std::string(eq, '=') + "]";
    const size_t end = s_.find(close, p + 1);
    if (end == std::string::npos) Fail("unterminated long bracket");
    *body = s_.substr(p + 1, end - p - 1);
    for (char c : *body) if (c == '\n') ++line_;
    i_ = end + close.size();
    return true;
  }
  Token Next() {
    for (;;) {
      while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\t' || s_[i_] == '\r' || s_[i_] == '\n')) {
        if (s_[i_] == '\n') ++line_;
        ++i_;
      }
      if (i_ + 1 < s_.size() && s_[i_] == '-' && s_[i_ + 1] == '-') {
        i_ += 2;
        std::string dummy;
        if (i_ < s_.size() && s_[i_] == '[' && LongBracket(&dummy)) continue;
        while (i_ < s_.size() && s_[i_] != '\n') ++i_;
        continue;
      }
      break;
    }
    Token t;
    t.line = line_;
    if (i_ >= s_.size()) return t;
    const char c = s_[i_];
    if (isalpha((unsigned char)c) || c == '_') {
      size_t j = i_;
      while (j < s_.size() && (isalnum((unsigned char)s_[j]) || s_[j] == '_')) ++j;
      t.t = tName; t.s = s_.substr(i_, j - i_); i_ = j;
      return t;
    }
    if (isdigit((unsigned char)c) || (c == '.' && i_ + 1 < s_.size() && isdigit((unsigned char)s_[i_ + 1]))) {
      // Lua numerals: decimal with optional fraction/exponent ("39.", "1e-3"), hex
      const char* b = s_.c_str() + i_;
      char* e = nullptr;
      if (c == '0' && i_ + 1 < s_.size() && (s_[i_ + 1] == 'x' || s_[i_ + 1] == 'X')) {
        t.num = (double)strtoull(b, &e, 16);
      } else {
        t.num = strtod(b, &e);
      }
      if (e == b) Fail("malformed number");
      t.t = tNumber; i_ += (size_t)(e - b);
      return t;
    }
    if (c == '"' || c == '\'') {
      std::string v;
      size_t j = i_ + 1;
      for (;; ++j) {
        if (j >= s_.size() || s_[j] == '\n') Fail("unterminated string");
        if (s_[j] == c) break;
        if (s_[j] == '\\' && j + 1 < s_.size()) {
          ++j;
          switch (s_[j]) {
            case 'n': v += '\n'; break;
            case 't': v += '\t'; break;
            case 'r': v += '\r'; break;
            case '\\': v += '\\'; break;
            case '"': v += '"'; break;
            case '\'': v += '\''; break;
            default: Fail("unsupported escape in string");
          }
        } else {
          v += s_[j];
        }
      }
      t.t = tString; t.s = v; i_ = j + 1;
      return t;
    }
    if (c == '[') {
      std::string body;
      if (LongBracket(&body)) { t.t = tString; t.s = body; return t; }
    }
    static const char* two[] = {"..", "==", "~=", "<=", ">="};
    for (const char* op : two) {
      if (s_.compare(i_, 2, op) == 0) {
        if (s_.compare(i_, 3, "...") == 0) Fail("varargs are outside the supported subset");
        t.t = tOp; t.s = op; i_ += 2;
        return t;
      }
    }
    if (strchr("+-*/%^=,;.(){}[]<>#", c)) {
      t.t = tOp; t.s = std::string(1, c); ++i_;
      return t;
    }
    Fail(std::string("unexpected character '") + c + "'");
  }

  const std::string& s_;
  std::string file_;
  size_t i_ = 0;
  int line_ = 1;
};

}  // namespace

// ---- AST + parser ---------------------------------------------------------------
namespace {

enum NodeKind { nConst, nName, nIndex, nCall, nTable, nUnary, nBinary };

struct Node {
  NodeKind kind = nConst;
  ValuePtr value;                 // nConst
  std::string name;               // nName; operator for nUnary / nBinary
  NodePtr a, b;                   // nIndex: a[b]; nUnary: a; nBinary: a op b; nCall: callee a
  std::vector<NodePtr> args;      // nCall arguments; nTable positional items
  std::vector<std::pair<NodePtr, NodePtr>> keyed;  // nTable keyed items
  int line = 0;
};

enum StmtKind { sAssign, sLocal, sFunction, sRequire, sCall };

struct Stmt {
  StmtKind kind = sAssign;
  std::vector<std::string> path;  // a.b.c target (sAssign / sLocal / sFunction name)
  NodePtr expr;                   // value / call expression
  std::vector<std::string> params;
  std::string module;             // sRequire
  int line = 0;
};

class Parser {
 public:
  Parser(std::vector<Token> toks, const std::string& file) : t_(std::move(toks)), file_(file) {}

  std::vector<Stmt> Chunk() {
    std::vector<Stmt> out;
    while (Peek().t != tEof) {
      if (IsOp(";")) { ++p_; continue; }
      out.push_back(Statement());
    }
    return out;
  }

 private:
  [[noreturn]] void Fail(const std::string& m) {
    throw ParseError{file_ + ":" + std::to_string(Peek().line) + ": " + m};
  }
  const Token& Peek(int k = 0) const { return t_[std::min(p_ + (size_t)k, t_.size() - 1)]; }
  bool IsOp(const char* s, int k = 0) const { return Peek(k).t == tOp && Peek(k).s == s; }
  bool IsName(const char* s, int k = 0) const { return Peek(k).t == tName && Peek(k).s == s; }
  void ExpectOp(const char* s) {
    if (!IsOp(s)) Fail(std::string("expected '") + s + "'");
    ++p_;
  }
  std::string ExpectName() {
    if (Peek().t != tName) Fail("expected a name");
    return t_[p_++].s;
  }
  static bool Reserved(const std::string& s) {
    static const char* kw[] = {"and", "break", "do", "else", "elseif", "end", "false", "for",
                               "function", "if", "in", "local", "nil", "not", "or", "repeat",
                               "return", "then", "true", "until", "while"};
    for (const char* k : kw) if (s == k) return true;
    return false;
  }

  Stmt Statement() {
    Stmt st;
    st.line = Peek().line;
    bool local = false;
    if (IsName("local")) { local = true; ++p_; }
    if (IsName("function")) {
      ++p_;
      st.kind = sFunction;
      st.path.push_back(ExpectName());
      while (IsOp(".")) { ++p_; st.path.push_back(ExpectName()); }
      ExpectOp("(");
      if (!IsOp(")")) {
        st.params.push_back(ExpectName());
        while (IsOp(",")) { ++p_; st.params.push_back(ExpectName()); }
      }
      ExpectOp(")");
      if (!IsName("return")) Fail("only single-return helper functions are supported");
      ++p_;
      st.expr = Expr(0);
      if (IsOp(";")) ++p_;
      if (!IsName("end")) Fail("expected 'end' (function bodies are one return statement)");
      ++p_;
      if (local) st.kind = sFunction;  // scope is not modelled: locals live in the global table too
      return st;
    }
    if (Peek().t != tName || Reserved(Peek().s))
      Fail("unsupported statement '" + Peek().s + "' (control flow is outside the supported subset)");
    // require("module") / require "module"
    if (IsName("require") && !local) {
      ++p_;
      st.kind = sRequire;
      const bool paren = IsOp("(");
      if (paren) ++p_;
      if (Peek().t != tString) Fail("require needs a string literal");
      st.module = t_[p_++].s;
      if (paren) ExpectOp(")");
      return st;
    }
    const size_t save = p_;
    st.path.push_back(ExpectName());
    while (IsOp(".")) { ++p_; st.path.push_back(ExpectName()); }
    if (IsOp("=")) {
      ++p_;
      st.kind = local ? sLocal : sAssign;
      st.expr = Expr(0);
      return st;
    }
    if (local) {  // `local x` without a value
      st.kind = sLocal;
      st.expr = std::make_shared<Node>();
      st.expr->value = MakeNil();
      return st;
    }
    // a bare call statement: evaluated for errors, result dropped
    p_ = save;
    st.kind = sCall;
    st.path.clear();
    st.expr = Suffixed();
    if (st.expr->kind != nCall) Fail("expected '=' or a call");
    return st;
  }

  // precedence climbing, Lua 5.1 table: or < and < comparison < .. < +- < */% < unary < ^
  static int LeftPrec(const std::string& op, bool* right_assoc) {
    *right_assoc = false;
    if (op == "or") return 1;
    if (op == "and") return 2;
    if (op == "<" || op == ">" || op == "<=" || op == ">=" || op == "~=" || op == "==") return 3;
    if (op == "..") { *right_assoc = true; return 4; }
    if (op == "+" || op == "-") return 5;
    if (op == "*" || op == "/" || op == "%") return 6;
    if (op == "^") { *right_assoc = true; return 8; }
    return -1;
  }

  NodePtr Expr(int min_prec) {
    NodePtr lhs;
    if (IsOp("-") || IsName("not") || IsOp("#")) {
      auto n = std::make_shared<Node>();
      n->kind = nUnary; n->name = t_[p_].s; n->line = t_[p_].line;
      ++p_;
      n->a = Expr(7);   // unary binds tighter than * but looser than ^
      lhs = n;
    } else {
      lhs = Suffixed();
    }
    for (;;) {
      const Token& t = Peek();
      if (!(t.t == tOp || (t.t == tName && (t.s == "and" || t.s == "or")))) break;
      bool right = false;
      const int prec = LeftPrec(t.s, &right);
      if (prec < 0 || prec < min_prec) break;
      auto n = std::make_shared<Node>();
      n->kind = nBinary; n->name = t.s; n->line = t.line;
      ++p_;
      n->a = lhs;
      n->b = Expr(right ? prec : prec + 1);
      lhs = n;
    }
    return lhs;
  }

  NodePtr Primary() {
    auto n = std::make_shared<Node>();
    n->line = Peek().line;
    const Token& t = Peek();
    if (t.t == tNumber) { n->value = MakeNum(t.num); ++p_; return n; }
    if (t.t == tString) { n->value = MakeStr(t.s); ++p_; return n; }
    if (t.t == tName) {
      if (t.s == "true" || t.s == "false") { n->value = MakeBool(t.s == "true"); ++p_; return n; }
      if (t.s == "nil") { n->value = MakeNil(); ++p_; return n; }
      if (t.s == "function") Fail("anonymous functions are outside the supported subset");
      if (Reserved(t.s)) Fail("unexpected keyword '" + t.s + "'");
      n->kind = nName; n->name = t.s; ++p_;
      return n;
    }
    if (IsOp("(")) {
      ++p_;
      NodePtr e = Expr(0);
      ExpectOp(")");
      return e;
    }
    if (IsOp("{")) return TableCtor();
    Fail("unexpected token '" + t.s + "' in expression");
  }

  NodePtr Suffixed() {
    NodePtr e = Primary();
    for (;;) {
      if (IsOp(".")) {
        ++p_;
        auto n = std::make_shared<Node>();
        n->kind = nIndex; n->line = Peek().line; n->a = e;
        n->b = std::make_shared<Node>();
        n->b->value = MakeStr(ExpectName());
        e = n;
      } else if (IsOp("[")) {
        ++p_;
        auto n = std::make_shared<Node>();
        n->kind = nIndex; n->line = Peek().line; n->a = e;
        n->b = Expr(0);
        ExpectOp("]");
        e = n;
      } else if (IsOp("(")) {
        ++p_;
        auto n = std::make_shared<Node>();
        n->kind = nCall; n->line = Peek().line; n->a = e;
        if (!IsOp(")")) {
          n->args.push_back(Expr(0));
          while (IsOp(",")) { ++p_; n->args.push_back(Expr(0)); }
        }
        ExpectOp(")");
        e = n;
      } else if (Peek().t == tString || IsOp("{")) {   // f "str" / f {table}
        auto n = std::make_shared<Node>();
        n->kind = nCall; n->line = Peek().line; n->a = e;
        n->args.push_back(Primary());
        e = n;
      } else {
        break;
      }
    }
    return e;
  }

  NodePtr TableCtor() {
    auto n = std::make_shared<Node>();
    n->kind = nTable; n->line = Peek().line;
    ExpectOp("{");
    while (!IsOp("}")) {
      if (IsOp("[")) {
        ++p_;
        NodePtr k = Expr(0);
        ExpectOp("]");
        ExpectOp("=");
        n->keyed.emplace_back(k, Expr(0));
      } else if (Peek().t == tName && IsOp("=", 1)) {
        auto k = std::make_shared<Node>();
        k->value = MakeStr(ExpectName());
        ExpectOp("=");
        n->keyed.emplace_back(k, Expr(0));
      } else {
        n->args.push_back(Expr(0));
      }
      if (IsOp(",") || IsOp(";")) ++p_;
      else if (!IsOp("}")) Fail("expected ',' or '}' in table constructor");
    }
    ExpectOp("}");
    return n;
  }

  std::vector<Token> t_;
  std::string file_;
  size_t p_ = 0;
};

}  // namespace

// ---- evaluator ------------------------------------------------------------------
namespace {

enum Builtin { bSin = 1, bCos, bTan, bSqrt, bAbs, bFloor, bCeil, bMax, bMin, bRad, bDeg, bAtan2,
               bToNumber, bToString };

struct Interp {
  ValuePtr globals = MakeTable();
  std::string root;                       // directory `require` resolves against
  std::vector<std::string> loading;       // require cycle guard
  std::string error;

  Interp() {
    auto math = MakeTable();
    math->fields["pi"] = MakeNum(3.14159265358979323846);   // math.pi (lmathlib.c PI)
    math->fields["huge"] = MakeNum(HUGE_VAL);
    const struct { const char* n; int b; } fn[] = {
        {"sin", bSin}, {"cos", bCos}, {"tan", bTan}, {"sqrt", bSqrt}, {"abs", bAbs},
        {"floor", bFloor}, {"ceil", bCeil}, {"max", bMax}, {"min", bMin}, {"rad", bRad},
        {"deg", bDeg}, {"atan2", bAtan2}};
    for (const auto& f : fn) {
      auto v = std::make_shared<Value>();
      v->kind = kBuiltin; v->builtin = f.b;
      math->fields[f.n] = v;
    }
    globals->fields["math"] = math;
    auto tn = std::make_shared<Value>(); tn->kind = kBuiltin; tn->builtin = bToNumber;
    auto ts = std::make_shared<Value>(); ts->kind = kBuiltin; ts->builtin = bToString;
    globals->fields["tonumber"] = tn;
    globals->fields["tostring"] = ts;
  }

  [[noreturn]] void Fail(const Node& n, const std::string& file, const std::string& m) {
    throw ParseError{file + ":" + std::to_string(n.line) + ": " + m};
  }

  static bool Truthy(const ValuePtr& v) { return !(v->kind == kNil || (v->kind == kBool && !v->b)); }

  static std::string NumToString(double d) {   // lua_Number2str: "%.14g"
    char buf[64];
    snprintf(buf, sizeof(buf), "%.14g", d);
    return buf;
  }

  bool ToNumber(const ValuePtr& v, double* out) {
    if (v->kind == kNumber) { *out = v->num; return true; }
    if (v->kind == kString) {   // Lua coerces numeric strings in arithmetic
      char* e = nullptr;
      const double d = strtod(v->str.c_str(), &e);
      if (e != v->str.c_str() && *e == '\0') { *out = d; return true; }
    }
    return false;
  }

  typedef std::map<std::string, ValuePtr> Scope;

  ValuePtr Index(const ValuePtr& t, const ValuePtr& k, const Node& n, const std::string& file) {
    if (t->kind != kTable) Fail(n, file, "attempt to index a non-table value");
    if (k->kind == kString) {
      auto it = t->fields.find(k->str);
      return it == t->fields.end() ? MakeNil() : it->second;
    }
    if (k->kind == kNumber) {
      const double d = k->num;
      if (d >= 1.0 && d <= (double)t->arr.size() && d == floor(d)) return t->arr[(size_t)d - 1];
      return MakeNil();
    }
    Fail(n, file, "unsupported table key type");
  }

  ValuePtr Eval(const NodePtr& np, const Scope* scope, const std::string& file) {
    const Node& n = *np;
    switch (n.kind) {
      case nConst:
        return n.value ? n.value : MakeNil();
      case nName: {
        if (scope) {
          auto it = scope->find(n.name);
          if (it != scope->end()) return it->second;
        }
        auto it = globals->fields.find(n.name);
        return it == globals->fields.end() ? MakeNil() : it->second;
      }
      case nIndex:
        return Index(Eval(n.a, scope, file), Eval(n.b, scope, file), n, file);
      case nTable: {
        auto t = MakeTable();
        for (const auto& a : n.args) t->arr.push_back(Eval(a, scope, file));
        for (const auto& kv : n.keyed) {
          ValuePtr k = Eval(kv.first, scope, file);
          ValuePtr v = Eval(kv.second, scope, file);
          if (k->kind == kString) {
            t->fields[k->str] = v;
          } else if (k->kind == kNumber && k->num == floor(k->num) && k->num >= 1 &&
                     k->num <= (double)t->arr.size() + 1) {
            if ((size_t)k->num == t->arr.size() + 1) t->arr.push_back(v);
            else t->arr[(size_t)k->num - 1] = v;
          } else {
            Fail(n, file, "unsupported table key");
          }
        }
        return t;
      }
      case nUnary: {
        ValuePtr a = Eval(n.a, scope, file);
        if (n.name == "not") return MakeBool(!Truthy(a));
        if (n.name == "#") {
          if (a->kind == kString) return MakeNum((double)a->str.size());
          if (a->kind == kTable) return MakeNum((double)a->arr.size());
          Fail(n, file, "attempt to get length of a non-table value");
        }
        double d;
        if (!ToNumber(a, &d)) Fail(n, file, "attempt to negate a non-number");
        return MakeNum(-d);
      }
      case nBinary: {
        if (n.name == "and") {
          ValuePtr a = Eval(n.a, scope, file);
          return Truthy(a) ? Eval(n.b, scope, file) : a;
        }
        if (n.name == "or") {
          ValuePtr a = Eval(n.a, scope, file);
          return Truthy(a) ? a : Eval(n.b, scope, file);
        }
        ValuePtr a = Eval(n.a, scope, file), b = Eval(n.b, scope, file);
        if (n.name == "..") {
          auto str = [&](const ValuePtr& v) -> std::string {
            if (v->kind == kString) return v->str;
            if (v->kind == kNumber) return NumToString(v->num);
            Fail(n, file, "attempt to concatenate a non-string value");
          };
          return MakeStr(str(a) + str(b));
        }
        if (n.name == "==" || n.name == "~=") {
          bool eq = a->kind == b->kind;
          if (eq) {
            switch (a->kind) {
              case kNil: break;
              case kBool: eq = a->b == b->b; break;
              case kNumber: eq = a->num == b->num; break;
              case kString: eq = a->str == b->str; break;
              default: eq = a.get() == b.get();
            }
          }
          return MakeBool(n.name == "==" ? eq : !eq);
        }
        double x, y;
        if (!ToNumber(a, &x) || !ToNumber(b, &y))
          Fail(n, file, "arithmetic on a non-number ('" + n.name + "')");
        if (n.name == "+") return MakeNum(x + y);
        if (n.name == "-") return MakeNum(x - y);
        if (n.name == "*") return MakeNum(x * y);
        if (n.name == "/") return MakeNum(x / y);
        if (n.name == "%") return MakeNum(x - floor(x / y) * y);   // luai_nummod
        if (n.name == "^") return MakeNum(pow(x, y));
        if (n.name == "<") return MakeBool(x < y);
        if (n.name == "<=") return MakeBool(x <= y);
        if (n.name == ">") return MakeBool(x > y);
        if (n.name == ">=") return MakeBool(x >= y);
        Fail(n, file, "unsupported operator '" + n.name + "'");
      }
      case nCall: {
        ValuePtr f = Eval(n.a, scope, file);
        std::vector<ValuePtr> args;
        for (const auto& a : n.args) args.push_back(Eval(a, scope, file));
        if (f->kind == kFunction) {
          Scope local;
          for (size_t i = 0; i < f->params.size(); ++i)
            local[f->params[i]] = i < args.size() ? args[i] : MakeNil();
          return Eval(f->body, &local, file);
        }
        if (f->kind == kBuiltin) return CallBuiltin(f->builtin, args, n, file);
        Fail(n, file, "attempt to call a non-function value");
      }
    }
    return MakeNil();
  }

  ValuePtr CallBuiltin(int b, const std::vector<ValuePtr>& args, const Node& n, const std::string& file) {
    if (b == bToString) {
      if (args.empty()) return MakeStr("nil");
      const ValuePtr& v = args[0];
      if (v->kind == kString) return v;
      if (v->kind == kNumber) return MakeStr(NumToString(v->num));
      if (v->kind == kBool) return MakeStr(v->b ? "true" : "false");
      return MakeStr("nil");
    }
    if (b == bToNumber) {
      double d;
      if (!args.empty() && ToNumber(args[0], &d)) return MakeNum(d);
      return MakeNil();
    }
    std::vector<double> x;
    for (const auto& a : args) {
      double d;
      if (!ToNumber(a, &d)) Fail(n, file, "math function called with a non-number");
      x.push_back(d);
    }
    if (x.empty()) Fail(n, file, "math function called without arguments");
    switch (b) {
      case bSin: return MakeNum(sin(x[0]));
      case bCos: return MakeNum(cos(x[0]));
      case bTan: return MakeNum(tan(x[0]));
      case bSqrt: return MakeNum(sqrt(x[0]));
      case bAbs: return MakeNum(fabs(x[0]));
      case bFloor: return MakeNum(floor(x[0]));
      case bCeil: return MakeNum(ceil(x[0]));
      case bRad: return MakeNum(x[0] * (3.14159265358979323846 / 180.0));   // lmathlib RADIANS_PER_DEGREE
      case bDeg: return MakeNum(x[0] / (3.14159265358979323846 / 180.0));
      case bAtan2: if (x.size() < 2) Fail(n, file, "math.atan2 needs two arguments"); return MakeNum(atan2(x[0], x[1]));
      case bMax: { double m = x[0]; for (double d : x) m = d > m ? d : m; return MakeNum(m); }
      case bMin: { double m = x[0]; for (double d : x) m = d < m ? d : m; return MakeNum(m); }
    }
    Fail(n, file, "unknown builtin");
  }

  void Assign(const std::vector<std::string>& path, const ValuePtr& v, const Stmt& st,
              const std::string& file) {
    ValuePtr t = globals;
    for (size_t i = 0; i + 1 < path.size(); ++i) {
      auto it = t->fields.find(path[i]);
      if (it == t->fields.end() || it->second->kind != kTable)
        throw ParseError{file + ":" + std::to_string(st.line) + ": attempt to index '" + path[i] +
                         "' (not a table)"};
      t = it->second;
    }
    if (v->kind == kNil) t->fields.erase(path.back());
    else t->fields[path.back()] = v;
  }

  static bool ReadFile(const std::string& path, std::string* out) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    char buf[4096];
    size_t n;
    out->clear();
    while ((n = fread(buf, 1, sizeof(buf), f)) > 0) out->append(buf, n);
    fclose(f);
    return true;
  }

  // UMRS_OK / UMRS_ERR_IO / UMRS_ERR_UNSUPPORTED
  int LoadFile(const std::string& path) {
    std::string src;
    if (!ReadFile(path, &src)) {
      error = "cannot read " + path;
      return UMRS_ERR_IO;
    }
    for (const auto& l : loading)
      if (l == path) return UMRS_OK;   // already being loaded (package.loaded semantics)
    loading.push_back(path);
    int rc = UMRS_OK;
    try {
      Lexer lx(src, path);
      Parser ps(lx.Run(), path);
      for (const Stmt& st : ps.Chunk()) {
        switch (st.kind) {
          case sAssign:
          case sLocal:
            Assign(st.path, Eval(st.expr, nullptr, path), st, path);
            break;
          case sFunction: {
            auto f = std::make_shared<Value>();
            f->kind = kFunction; f->params = st.params; f->body = st.expr;
            Assign(st.path, f, st, path);
            break;
          }
          case sCall:
            Eval(st.expr, nullptr, path);
            break;
          case sRequire: {
            std::string rel = st.module;
            for (char& c : rel) if (c == '.') c = '/';
            rel += ".lua";
            const std::string full = root.empty() ? rel : root + "/" + rel;
            rc = LoadFile(full);
            if (rc != UMRS_OK) {
              error = path + ":" + std::to_string(st.line) + ": require(\"" + st.module + "\"): " + error;
              loading.pop_back();
              return rc;
            }
            break;
          }
        }
      }
    } catch (const ParseError& e) {
      error = e.msg;
      rc = UMRS_ERR_UNSUPPORTED;
    }
    loading.pop_back();
    return rc;
  }

  // dotted lookup as lua_script.h:82-119
  ValuePtr Lookup(const char* key) const {
    ValuePtr t = globals;
    std::string var;
    for (const char* p = key;; ++p) {
      if (*p == '.' || *p == '\0') {
        if (t->kind != kTable) return nullptr;
        auto it = t->fields.find(var);
        if (it == t->fields.end()) return nullptr;
        t = it->second;
        var.clear();
        if (*p == '\0') break;
      } else {
        var += *p;
      }
    }
    return t;
  }
};

}  // namespace

struct umrs_config {
  Interp in;
};

// ---- C-ABI ----------------------------------------------------------------------
namespace {
thread_local std::string g_config_error;

// CONFIG_FLOAT semantics: missing / non-numeric key -> 0 and one more "missing"
// (the reference prints an error and keeps the type's default, lua_script.h:72-75).
float GetF(const umrs_config* c, const char* key, int* missing) {
  ValuePtr v = c->in.Lookup(key);
  if (!v || v->kind != kNumber) { ++*missing; return 0.f; }
  return (float)v->num;
}
int GetB(const umrs_config* c, const char* key, int* missing) {
  ValuePtr v = c->in.Lookup(key);
  if (!v || v->kind != kBool) { ++*missing; return 0; }
  return v->b ? 1 : 0;
}
}  // namespace

extern "C" {

const char* umrs_config_last_error(void) { return g_config_error.c_str(); }

int umrs_config_load(const char* const* files, int32_t n_files, const char* root_dir,
                     umrs_config** out) {
  if (out == NULL) return UMRS_ERR_INVALID;
  *out = NULL;
  if (n_files < 0 || (n_files > 0 && files == NULL)) return UMRS_ERR_INVALID;
  umrs_config* c = new (std::nothrow) umrs_config();
  if (!c) return UMRS_ERR_NOMEM;
  c->in.root = root_dir ? root_dir : "";
  for (int32_t i = 0; i < n_files; ++i) {
    if (files[i] == NULL) { delete c; return UMRS_ERR_INVALID; }
    std::string path = files[i];
    // a relative path is tried as given, then against the root (the reference is
    // started from its repository root, README.md:40-54)
    FILE* f = fopen(path.c_str(), "rb");
    if (f) fclose(f);
    else if (!c->in.root.empty() && path[0] != '/') path = c->in.root + "/" + path;
    const int rc = c->in.LoadFile(path);
    if (rc != UMRS_OK) {
      g_config_error = c->in.error;
      delete c;
      return rc;
    }
  }
  g_config_error.clear();
  *out = c;
  return UMRS_OK;
}

int umrs_config_destroy(umrs_config* c) {
  delete c;
  return UMRS_OK;
}

int umrs_config_has(const umrs_config* c, const char* key) {
  if (c == NULL || key == NULL) return 0;
  return c->in.Lookup(key) ? 1 : 0;
}

int umrs_config_get_double(const umrs_config* c, const char* key, double* out) {
  if (c == NULL || key == NULL || out == NULL) return UMRS_ERR_INVALID;
  ValuePtr v = c->in.Lookup(key);
  if (!v || v->kind != kNumber) return UMRS_ERR_INVALID;
  *out = v->num;
  return UMRS_OK;
}

int umrs_config_get_float(const umrs_config* c, const char* key, float* out) {
  double d;
  if (out == NULL) return UMRS_ERR_INVALID;
  const int rc = umrs_config_get_double(c, key, &d);
  if (rc == UMRS_OK) *out = (float)d;
  return rc;
}

int umrs_config_get_int(const umrs_config* c, const char* key, int32_t* out) {
  double d;
  if (out == NULL) return UMRS_ERR_INVALID;
  const int rc = umrs_config_get_double(c, key, &d);
  if (rc == UMRS_OK) *out = (int32_t)d;   // lua_tointeger truncation
  return rc;
}

int umrs_config_get_bool(const umrs_config* c, const char* key, int32_t* out) {
  if (c == NULL || key == NULL || out == NULL) return UMRS_ERR_INVALID;
  ValuePtr v = c->in.Lookup(key);
  if (!v || v->kind != kBool) return UMRS_ERR_INVALID;
  *out = v->b ? 1 : 0;
  return UMRS_OK;
}

int umrs_config_get_string(const umrs_config* c, const char* key, char* buf, int32_t cap) {
  if (c == NULL || key == NULL || buf == NULL || cap <= 0) return UMRS_ERR_INVALID;
  ValuePtr v = c->in.Lookup(key);
  if (!v || v->kind != kString || (int64_t)v->str.size() + 1 > cap) return UMRS_ERR_INVALID;
  memcpy(buf, v->str.c_str(), v->str.size() + 1);
  return UMRS_OK;
}

int umrs_config_list_length(const umrs_config* c, const char* key, int32_t* n) {
  if (c == NULL || key == NULL || n == NULL) return UMRS_ERR_INVALID;
  ValuePtr v = c->in.Lookup(key);
  if (!v || v->kind != kTable) return UMRS_ERR_INVALID;
  *n = (int32_t)v->arr.size();
  return UMRS_OK;
}

int umrs_config_get_list_string(const umrs_config* c, const char* key, int32_t index, char* buf,
                                int32_t cap) {
  if (c == NULL || key == NULL || buf == NULL || cap <= 0) return UMRS_ERR_INVALID;
  ValuePtr v = c->in.Lookup(key);
  if (!v || v->kind != kTable || index < 0 || index >= (int32_t)v->arr.size()) return UMRS_ERR_INVALID;
  const ValuePtr& e = v->arr[(size_t)index];
  if (e->kind != kString || (int64_t)e->str.size() + 1 > cap) return UMRS_ERR_INVALID;
  memcpy(buf, e->str.c_str(), e->str.size() + 1);
  return UMRS_OK;
}

int umrs_config_get_list_vector3(const umrs_config* c, const char* key, int32_t index, float out3[3]) {
  if (c == NULL || key == NULL || out3 == NULL) return UMRS_ERR_INVALID;
  ValuePtr v = c->in.Lookup(key);
  if (!v || v->kind != kTable || index < 0 || index >= (int32_t)v->arr.size()) return UMRS_ERR_INVALID;
  const ValuePtr& e = v->arr[(size_t)index];
  if (e->kind != kTable) return UMRS_ERR_INVALID;
  // {x, y, z} positional (default_init_config.lua:2) or Vector3's {x=, y=, z=}
  const char* names[3] = {"x", "y", "z"};
  for (int k = 0; k < 3; ++k) {
    ValuePtr f;
    if ((size_t)k < e->arr.size()) f = e->arr[(size_t)k];
    else {
      auto it = e->fields.find(names[k]);
      if (it != e->fields.end()) f = it->second;
    }
    if (!f || f->kind != kNumber) return UMRS_ERR_INVALID;
    out3[k] = (float)f->num;
  }
  return UMRS_OK;
}

int umrs_config_laser_params(const umrs_config* c, umrs_laser_params* o, int32_t* n_missing) {
  if (c == NULL || o == NULL) return UMRS_ERR_INVALID;
  int miss = 0;
  o->loc_x = GetF(c, "laser_loc.x", &miss);                  // simulator.cpp:70-71
  o->loc_y = GetF(c, "laser_loc.y", &miss);
  o->angle_min = GetF(c, "laser_angle_min", &miss);          // :87-91
  o->angle_max = GetF(c, "laser_angle_max", &miss);
  o->angle_increment = GetF(c, "laser_angle_increment", &miss);
  o->range_min = GetF(c, "laser_min_range", &miss);
  o->range_max = GetF(c, "laser_max_range", &miss);
  o->noise_stddev = GetF(c, "laser_noise_stddev", &miss);    // :75
  if (n_missing) *n_missing = miss;
  return UMRS_OK;
}

int umrs_config_robot_params(const umrs_config* c, const char* robot_type, umrs_robot_params* o,
                             int32_t* n_missing) {
  if (c == NULL || o == NULL) return UMRS_ERR_INVALID;
  const int type = umrs_model_type_from_string(robot_type);
  if (type < 0) return UMRS_ERR_INVALID;   // the reference exits: simulator.cpp:139-141
  memset(o, 0, sizeof(*o));
  o->model_type = type;
  int miss = 0;
  if (type == UMRS_ACKERMANN_DRIVE) {            // ackermann_model.cpp:18-22
    o->ak_min_turn_radius = GetF(c, "ak_min_turn_radius", &miss);
    o->ak_max_speed = GetF(c, "ak_max_speed", &miss);
    o->ak_max_accel = GetF(c, "ak_max_accel", &miss);
    o->ak_angular_error_bias = GetF(c, "ak_angular_error_bias", &miss);
    o->ak_angular_error_rate = GetF(c, "ak_angular_error_rate", &miss);
  } else if (type == UMRS_DIFF_DRIVE) {          // diff_drive_model.cpp:43-50
    o->invert_linear_vel_cmds = GetB(c, "invert_linear_vel_cmds", &miss);
    o->invert_angular_vel_cmds = GetB(c, "invert_angular_vel_cmds", &miss);
    o->linear_pos_accel_limit = GetF(c, "linear_pos_accel_limit", &miss);
    o->linear_neg_accel_limit = GetF(c, "linear_neg_accel_limit", &miss);
    o->angular_pos_accel_limit = GetF(c, "angular_pos_accel_limit", &miss);
    o->angular_neg_accel_limit = GetF(c, "angular_neg_accel_limit", &miss);
    o->max_angular = GetF(c, "max_angular", &miss);
    o->max_linear_vel = GetF(c, "max_linear_vel", &miss);
  } else {                                       // omnidirectional_model.cpp:21-24
    o->co_max_accel = GetF(c, "co_max_accel", &miss);
    o->co_max_angle_accel = GetF(c, "co_max_angle_accel", &miss);
    o->co_max_speed = GetF(c, "co_max_speed", &miss);
    o->co_max_angle_vel = GetF(c, "co_max_angle_vel", &miss);
  }
  o->car_length = GetF(c, "car_length", &miss);              // simulator.cpp:65-68
  o->car_width = GetF(c, "car_width", &miss);
  o->rear_axle_offset = GetF(c, "rear_axle_offset", &miss);
  // robot bodies are not part of the reference's scans; the caller opts in
  o->footprint_enabled = 0;
  int lm = 0;
  umrs_config_laser_params(c, &o->laser, &lm);
  if (n_missing) *n_missing = miss + lm;
  return UMRS_OK;
}

int umrs_config_human_params(const umrs_config* c, umrs_human_params* o, int32_t* n_missing) {
  if (c == NULL || o == NULL) return UMRS_ERR_INVALID;
  int miss = 0;
  memset(o, 0, sizeof(*o));
  o->radius = GetF(c, "hu_radius", &miss);                   // human_object.cpp:30-43
  o->num_segments = (int32_t)GetF(c, "hu_num_segments", &miss);  // CONFIG_FLOAT, used as int (:31,:101)
  o->start_x = GetF(c, "hu_start_x", &miss);
  o->start_y = GetF(c, "hu_start_y", &miss);
  o->start_theta = GetF(c, "hu_start_theta", &miss);
  o->goal_x = GetF(c, "hu_goal_x", &miss);
  o->goal_y = GetF(c, "hu_goal_y", &miss);
  o->goal_theta = GetF(c, "hu_goal_theta", &miss);
  o->max_speed = GetF(c, "hu_max_speed", &miss);
  o->avg_speed = GetF(c, "hu_avg_speed", &miss);
  o->max_omega = GetF(c, "hu_max_omega", &miss);
  o->avg_omega = GetF(c, "hu_avg_omega", &miss);
  o->reach_goal_threshold = GetF(c, "hu_reach_goal_threshold", &miss);
  o->mode = (int32_t)GetF(c, "hu_mode", &miss);              // CONFIG_INT (:43)
  if (n_missing) *n_missing = miss;
  return UMRS_OK;
}

}  // extern "C"
