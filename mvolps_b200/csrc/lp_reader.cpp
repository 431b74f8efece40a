// lp_reader.cpp -- CPLEX-LP and MPS readers behind glp_read_lp / glp_read_mps
// (util.cpp:283-293; file type chosen by extension in 2test.cpp:65-81).
// Written from the public format descriptions (CPLEX LP file format; MPS
// fixed/free format incl. OBJSENSE, RANGES, MARKER INTORG/INTEND).
#include "lp_model.hpp"

#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <fstream>
#include <map>
#include <sstream>

namespace {

std::string lower(std::string s) {
  for (auto &c : s) c = (char)std::tolower((unsigned char)c);
  return s;
}

// The readers fill an array-of-structs image of the problem (rows and columns are touched again
// and again while sections are parsed) and commit it to the glp_prob's copy-on-write arrays at the end.
struct MvRow {
  int type = GLP_FR;
  double lb = 0.0, ub = 0.0;
  std::vector<int> ind;
  std::vector<double> val;
  int stat = GLP_BS;
  MvName name;
};
struct MvCol {
  int type = GLP_FX;
  double lb = 0.0, ub = 0.0;
  int kind = GLP_CV;
  double coef = 0.0;
  int stat = GLP_NS;
  MvName name;
};
struct BuildProb {
  int dir = GLP_MIN;
  double c0 = 0.0;
  std::string name;
  std::vector<MvRow> rows;
  std::vector<MvCol> cols;
};
void commit(const BuildProb &b, glp_prob *P) {
  P->dir = b.dir;
  P->c0 = b.c0;
  P->name = b.name;
  mv_add_cols(P, (int)b.cols.size());
  mv_add_rows(P, (int)b.rows.size());
  auto cm = P->colm.mut();
  auto cb = P->cbnd.mut();
  auto cs = P->cstat.mut();
  for (size_t j = 0; j < b.cols.size(); j++) {
    const MvCol &c = b.cols[j];
    cm[j].kind = c.kind; cm[j].coef = c.coef; cm[j].name = c.name;
    cb[j] = MvBnd{c.type, c.lb, c.ub};
    cs[j] = c.stat;
  }
  auto rm = P->rowm.mut();
  auto rb = P->rbnd.mut();
  auto rs = P->rstat.mut();
  for (size_t i = 0; i < b.rows.size(); i++) {
    const MvRow &r = b.rows[i];
    rm[i].ind.assign(r.ind.data(), r.ind.data() + r.ind.size());
    rm[i].val.assign(r.val.data(), r.val.data() + r.val.size());
    rm[i].name = r.name;
    rb[i] = MvBnd{r.type, r.lb, r.ub};
    rs[i] = r.stat;
  }
}

struct Builder {
  BuildProb *P;
  std::map<std::string, int> colidx, rowidx;
  explicit Builder(BuildProb *p) : P(p) {}
  int col(const std::string &name) {
    auto it = colidx.find(name);
    if (it != colidx.end()) return it->second;
    MvCol c;
    c.type = GLP_LO;  // default bounds 0 <= x < +inf
    c.stat = GLP_NL;
    c.name = name;
    P->cols.push_back(c);
    int j = (int)P->cols.size();
    colidx[name] = j;
    return j;
  }
  void set_row(int i, std::map<int, double> &coefs) {
    MvRow &r = P->rows[i - 1];
    r.ind.clear();
    r.val.clear();
    for (auto &kv : coefs)
      if (kv.second != 0.0) { r.ind.push_back(kv.first); r.val.push_back(kv.second); }
  }
};

void set_typed(int &type, double &lb, double &ub, int &stat, bool hasl, double l,
               bool hasu, double u) {
  if (hasl && hasu) {
    if (l == u) { type = GLP_FX; lb = ub = l; stat = GLP_NS; }
    else { type = GLP_DB; lb = l; ub = u; stat = GLP_NL; }
  } else if (hasl) { type = GLP_LO; lb = l; ub = 0; stat = GLP_NL; }
  else if (hasu) { type = GLP_UP; lb = 0; ub = u; stat = GLP_NU; }
  else { type = GLP_FR; lb = ub = 0; stat = GLP_NF; }
}

// ------------------------------------------------------------------ LP format
enum TokKind { T_EOF, T_NUM, T_NAME, T_OP, T_COLON, T_PLUS, T_MINUS };
struct Tok {
  TokKind kind = T_EOF;
  std::string text;
  double num = 0;
};

bool name_char(char c) {
  return std::isalnum((unsigned char)c) || std::strchr("!\"#$%&()/,.;?@_`'{}|~", c);
}

std::vector<Tok> tokenize(const std::string &src) {
  std::vector<Tok> out;
  size_t i = 0, n = src.size();
  while (i < n) {
    char c = src[i];
    if (c == '\\') { while (i < n && src[i] != '\n') i++; continue; }
    if (std::isspace((unsigned char)c)) { i++; continue; }
    Tok t;
    if (std::isdigit((unsigned char)c) || (c == '.' && i + 1 < n && std::isdigit((unsigned char)src[i + 1]))) {
      char *end = nullptr;
      t.kind = T_NUM;
      t.num = std::strtod(src.c_str() + i, &end);
      size_t len = end - (src.c_str() + i);
      t.text = src.substr(i, len);
      i += len;
    } else if (c == '+') { t.kind = T_PLUS; t.text = "+"; i++; }
    else if (c == '-') { t.kind = T_MINUS; t.text = "-"; i++; }
    else if (c == ':') { t.kind = T_COLON; t.text = ":"; i++; }
    else if (c == '<' || c == '>' || c == '=') {
      t.kind = T_OP;
      char a = c;
      i++;
      char b = (i < n && (src[i] == '=' || src[i] == '<' || src[i] == '>')) ? src[i] : 0;
      if (b) i++;
      if (a == '<' || b == '<') t.text = "<=";
      else if (a == '>' || b == '>') t.text = ">=";
      else t.text = "=";
    } else if (name_char(c)) {
      size_t s = i;
      while (i < n && name_char(src[i])) i++;
      t.kind = T_NAME;
      t.text = src.substr(s, i - s);
    } else {
      t.kind = T_NAME;  // unknown char: surfaces as a parse error upstream
      t.text = std::string(1, c);
      i++;
    }
    out.push_back(t);
  }
  out.push_back(Tok{});
  return out;
}

enum Section { S_NONE, S_OBJ, S_ST, S_BOUNDS, S_GENERAL, S_BINARY, S_END };

struct LpParser {
  std::vector<Tok> tk;
  size_t pos = 0;
  Builder b;
  std::string err;
  explicit LpParser(BuildProb *P) : b(P) {}

  const Tok &peek(size_t k = 0) const { return tk[std::min(pos + k, tk.size() - 1)]; }
  bool is_inf_word(const Tok &t) const {
    if (t.kind != T_NAME) return false;
    std::string w = lower(t.text);
    return w == "inf" || w == "infinity";
  }
  // section keyword at the current position? returns section and token count
  Section keyword(size_t &len) const {
    const Tok &t = peek();
    len = 1;
    if (t.kind != T_NAME) return S_NONE;
    if (peek(1).kind == T_COLON) return S_NONE;  // "max:" style names stay names
    std::string w = lower(t.text);
    if (w == "maximize" || w == "maximise" || w == "maximum" || w == "max") return S_OBJ;
    if (w == "minimize" || w == "minimise" || w == "minimum" || w == "min") return S_OBJ;
    if (w == "st" || w == "s.t." || w == "st.") return S_ST;
    if ((w == "subject" || w == "such") && peek(1).kind == T_NAME) {
      std::string w2 = lower(peek(1).text);
      if ((w == "subject" && w2 == "to") || (w == "such" && w2 == "that")) { len = 2; return S_ST; }
    }
    if (w == "bounds" || w == "bound") return S_BOUNDS;
    if (w == "general" || w == "generals" || w == "gen" || w == "integer" ||
        w == "integers" || w == "int") return S_GENERAL;
    if (w == "binary" || w == "binaries" || w == "bin") return S_BINARY;
    if (w == "end") return S_END;
    return S_NONE;
  }

  // [+|-] [number] name | [+|-] number ; accumulates into coefs / constant
  bool parse_expr(std::map<int, double> &coefs, double &constant) {
    bool any = false;
    while (true) {
      size_t klen;
      if (keyword(klen) != S_NONE) break;
      const Tok &t = peek();
      if (t.kind == T_OP || t.kind == T_EOF) break;
      double sign = 1.0;
      bool had_sign = false;
      while (peek().kind == T_PLUS || peek().kind == T_MINUS) {
        if (peek().kind == T_MINUS) sign = -sign;
        pos++;
        had_sign = true;
      }
      double coef = 1.0;
      bool had_num = false;
      if (peek().kind == T_NUM) { coef = peek().num; pos++; had_num = true; }
      size_t k2;
      if (peek().kind == T_NAME && keyword(k2) == S_NONE && !is_inf_word(peek())) {
        if (peek(1).kind == T_COLON) {
          if (had_sign || had_num) { err = "unexpected label inside expression"; return false; }
          break;  // a constraint label: caller handles it
        }
        int j = b.col(peek().text);
        pos++;
        coefs[j] += sign * coef;
      } else if (had_num) {
        constant += sign * coef;
      } else {
        if (had_sign) { err = "dangling sign in expression"; return false; }
        break;
      }
      any = true;
    }
    (void)any;
    return true;
  }

  bool parse_number(double &v) {
    double sign = 1.0;
    while (peek().kind == T_PLUS || peek().kind == T_MINUS) {
      if (peek().kind == T_MINUS) sign = -sign;
      pos++;
    }
    if (peek().kind == T_NUM) { v = sign * peek().num; pos++; return true; }
    if (is_inf_word(peek())) { v = sign * HUGE_VAL; pos++; return true; }
    err = "number expected near '" + peek().text + "'";
    return false;
  }

  bool parse_objective() {
    std::string w = lower(peek().text);
    b.P->dir = (w[1] == 'a') ? GLP_MAX : GLP_MIN;  // mAx... / mIn...
    pos++;
    if (peek().kind == T_NAME && peek(1).kind == T_COLON) pos += 2;  // objective name
    std::map<int, double> coefs;
    double c0 = 0.0;
    if (!parse_expr(coefs, c0)) return false;
    for (auto &kv : coefs) b.P->cols[kv.first - 1].coef = kv.second;
    b.P->c0 = c0;
    return true;
  }

  bool parse_constraints() {
    while (true) {
      size_t klen;
      if (keyword(klen) != S_NONE || peek().kind == T_EOF) return true;
      std::string rname;
      if (peek().kind == T_NAME && peek(1).kind == T_COLON) { rname = peek().text; pos += 2; }
      std::map<int, double> coefs;
      double lhs_const = 0.0;
      if (!parse_expr(coefs, lhs_const)) return false;
      if (peek().kind != T_OP) { err = "constraint sense expected near '" + peek().text + "'"; return false; }
      std::string op = peek().text;
      pos++;
      bool ranged = false;
      double rlo = 0.0, rhi = 0.0, rhs = 0.0;
      if (coefs.empty() && op != "=") {
        // double-sided row: "lo <= expr <= hi" (or "hi >= expr >= lo")
        const double left = lhs_const;
        double mid_const = 0.0;
        if (!parse_expr(coefs, mid_const)) return false;
        if (coefs.empty() || peek().kind != T_OP || peek().text != op) {
          err = "double-sided constraint needs the same sense twice";
          return false;
        }
        pos++;
        double right;
        if (!parse_number(right)) return false;
        rlo = (op == "<=" ? left : right) - mid_const;
        rhi = (op == "<=" ? right : left) - mid_const;
        ranged = true;
      } else {
        if (!parse_number(rhs)) return false;
        rhs -= lhs_const;
      }
      b.P->rows.emplace_back();
      int i = (int)b.P->rows.size();
      MvRow &r = b.P->rows[i - 1];
      r.name = rname.empty() ? "r." + std::to_string(i) : rname;
      if (ranged) set_typed(r.type, r.lb, r.ub, r.stat, rlo > -HUGE_VAL, rlo, rhi < HUGE_VAL, rhi);
      else if (op == "<=") set_typed(r.type, r.lb, r.ub, r.stat, false, 0, true, rhs);
      else if (op == ">=") set_typed(r.type, r.lb, r.ub, r.stat, true, rhs, false, 0);
      else set_typed(r.type, r.lb, r.ub, r.stat, true, rhs, true, rhs);
      r.stat = GLP_BS;
      b.set_row(i, coefs);
    }
  }

  void apply_bounds(int j, bool setl, double l, bool setu, double u) {
    MvCol &c = b.P->cols[j - 1];
    double cl = mv_eff_lb(c.type, c.lb), cu = mv_eff_ub(c.type, c.ub);
    if (setl) cl = l;
    if (setu) cu = u;
    set_typed(c.type, c.lb, c.ub, c.stat, cl > -HUGE_VAL, cl, cu < HUGE_VAL, cu);
  }

  bool parse_bounds() {
    while (true) {
      size_t klen;
      if (keyword(klen) != S_NONE || peek().kind == T_EOF) return true;
      size_t k2;
      if (peek().kind == T_NAME && !is_inf_word(peek())) {
        int j = b.col(peek().text);
        pos++;
        if (peek().kind == T_NAME && lower(peek().text) == "free" && keyword(k2) == S_NONE) {
          pos++;
          apply_bounds(j, true, -HUGE_VAL, true, HUGE_VAL);
          continue;
        }
        if (peek().kind != T_OP) { err = "bound sense expected after '" + b.P->cols[j - 1].name.str() + "'"; return false; }
        std::string op = peek().text;
        pos++;
        double v;
        if (!parse_number(v)) return false;
        if (op == "<=") apply_bounds(j, false, 0, true, v);
        else if (op == ">=") apply_bounds(j, true, v, false, 0);
        else apply_bounds(j, true, v, true, v);
      } else {
        double l;
        if (!parse_number(l)) return false;
        if (peek().kind != T_OP) { err = "bound sense expected"; return false; }
        std::string op = peek().text;
        pos++;
        if (peek().kind != T_NAME) { err = "variable name expected in bound"; return false; }
        int j = b.col(peek().text);
        pos++;
        if (op == "<=") apply_bounds(j, true, l, false, 0);
        else if (op == ">=") apply_bounds(j, false, 0, true, l);
        else apply_bounds(j, true, l, true, l);
        if (peek().kind == T_OP && op == "<=" && peek().text == "<=") {
          pos++;
          double u;
          if (!parse_number(u)) return false;
          apply_bounds(j, false, 0, true, u);
        }
      }
    }
  }

  bool parse_kinds(bool binary) {
    while (true) {
      size_t klen;
      if (keyword(klen) != S_NONE || peek().kind == T_EOF) return true;
      if (peek().kind != T_NAME) { err = "variable name expected near '" + peek().text + "'"; return false; }
      int j = b.col(peek().text);
      pos++;
      MvCol &c = b.P->cols[j - 1];
      c.kind = GLP_IV;
      if (binary) set_typed(c.type, c.lb, c.ub, c.stat, true, 0.0, true, 1.0);
    }
  }

  bool run(const std::string &src) {
    tk = tokenize(src);
    size_t klen;
    if (keyword(klen) != S_OBJ) { err = "'maximize' or 'minimize' expected"; return false; }
    if (!parse_objective()) return false;
    bool ended = false;
    while (!ended) {
      Section s = keyword(klen);
      if (s == S_NONE) {
        err = peek().kind == T_EOF ? "missing 'end'" : "unexpected '" + peek().text + "'";
        if (peek().kind == T_EOF) { err.clear(); break; }  // tolerate a missing End
        return false;
      }
      pos += klen;
      switch (s) {
        case S_ST: if (!parse_constraints()) return false; break;
        case S_BOUNDS: if (!parse_bounds()) return false; break;
        case S_GENERAL: if (!parse_kinds(false)) return false; break;
        case S_BINARY: if (!parse_kinds(true)) return false; break;
        case S_END: ended = true; break;
        default: err = "objective given twice"; return false;
      }
    }
    if (b.P->cols.empty()) { err = "problem has no columns"; return false; }
    return true;
  }
};

}  // namespace

int mv_read_lp(glp_prob *P, const char *fname, std::string &err) {
  std::ifstream f(fname);
  if (!f) { err = "unable to open file"; return 1; }
  std::stringstream ss;
  ss << f.rdbuf();
  BuildProb image;
  LpParser p(&image);
  if (!p.run(ss.str())) { err = p.err; return 1; }
  commit(image, P);
  return 0;
}

// ------------------------------------------------------------------ MPS format
// Free-style field splitting (also accepts fixed-format files whose names have
// no embedded blanks).  Sections: NAME, OBJSENSE, ROWS, COLUMNS (with MARKER
// INTORG/INTEND), RHS (objective entry = minus the constant term), RANGES,
// BOUNDS (UP LO FX FR MI PL BV LI UI), ENDATA.  The first N row is the
// objective; further N rows are dropped.
static int read_mps_image(BuildProb *P, const char *fname, std::string &err);
int mv_read_mps(glp_prob *P, const char *fname, std::string &err) {
  BuildProb image;
  if (read_mps_image(&image, fname, err)) return 1;
  commit(image, P);
  return 0;
}
static int read_mps_image(BuildProb *P, const char *fname, std::string &err) {
  std::ifstream f(fname);
  if (!f) { err = "unable to open file"; return 1; }
  Builder b(P);
  P->dir = GLP_MIN;
  std::string line, section, objrow;
  std::map<std::string, char> rowtype;
  std::vector<std::map<int, double>> rowcoef;  // by row index-1
  std::vector<bool> col_bounded;
  bool integer_mode = false;
  int lineno = 0;
  auto fail = [&](const std::string &m) {
    err = "line " + std::to_string(lineno) + ": " + m;
    return 1;
  };
  while (std::getline(f, line)) {
    lineno++;
    if (line.empty() || line[0] == '*') continue;
    std::istringstream ls(line);
    std::vector<std::string> w;
    for (std::string t; ls >> t;) w.push_back(t);
    if (w.empty()) continue;
    if (!std::isspace((unsigned char)line[0])) {  // section header
      section = w[0];
      for (auto &c : section) c = (char)std::toupper((unsigned char)c);
      if (section == "NAME") { if (w.size() > 1) P->name = w[1]; }
      else if (section == "OBJSENSE" && w.size() > 1) {
        std::string s = lower(w[1]);
        P->dir = (s == "max" || s == "maximize") ? GLP_MAX : GLP_MIN;
      } else if (section == "ENDATA") break;
      continue;
    }
    if (section == "OBJSENSE") {
      std::string s = lower(w[0]);
      P->dir = (s == "max" || s == "maximize") ? GLP_MAX : GLP_MIN;
    } else if (section == "ROWS") {
      if (w.size() < 2) return fail("row type and name expected");
      char t = (char)std::toupper((unsigned char)w[0][0]);
      if (!std::strchr("NLGE", t)) return fail("unknown row type");
      if (t == 'N') {
        if (objrow.empty()) objrow = w[1];
        rowtype[w[1]] = 'N';
        continue;
      }
      P->rows.emplace_back();
      MvRow &r = P->rows.back();
      r.name = w[1];
      if (t == 'L') set_typed(r.type, r.lb, r.ub, r.stat, false, 0, true, 0.0);
      if (t == 'G') set_typed(r.type, r.lb, r.ub, r.stat, true, 0.0, false, 0);
      if (t == 'E') set_typed(r.type, r.lb, r.ub, r.stat, true, 0.0, true, 0.0);
      r.stat = GLP_BS;
      rowtype[w[1]] = t;
      b.rowidx[w[1]] = (int)P->rows.size();
      rowcoef.emplace_back();
    } else if (section == "COLUMNS") {
      if (w.size() >= 3 && w[1] == "'MARKER'") {
        if (w[2] == "'INTORG'") integer_mode = true;
        else if (w[2] == "'INTEND'") integer_mode = false;
        continue;
      }
      if (w.size() < 3 || w.size() % 2 == 0) return fail("column entry expected");
      bool fresh = !b.colidx.count(w[0]);
      int j = b.col(w[0]);
      if (fresh) {
        col_bounded.push_back(false);
        if (integer_mode) P->cols[j - 1].kind = GLP_IV;
      }
      for (size_t t = 1; t + 1 < w.size(); t += 2) {
        double v = std::strtod(w[t + 1].c_str(), nullptr);
        auto rt = rowtype.find(w[t]);
        if (rt == rowtype.end()) return fail("unknown row '" + w[t] + "'");
        if (rt->second == 'N') { if (w[t] == objrow) P->cols[j - 1].coef = v; }
        else rowcoef[b.rowidx[w[t]] - 1][j] += v;
      }
    } else if (section == "RHS" || section == "RANGES") {
      size_t s = (w.size() % 2 == 1) ? 1 : 0;  // optional set name
      for (size_t t = s; t + 1 < w.size(); t += 2) {
        double v = std::strtod(w[t + 1].c_str(), nullptr);
        auto rt = rowtype.find(w[t]);
        if (rt == rowtype.end()) return fail("unknown row '" + w[t] + "'");
        if (rt->second == 'N') { if (section == "RHS" && w[t] == objrow) P->c0 = -v; continue; }
        MvRow &r = P->rows[b.rowidx[w[t]] - 1];
        if (section == "RHS") {
          if (rt->second == 'L') r.ub = v;
          else if (rt->second == 'G') r.lb = v;
          else r.lb = r.ub = v;
        } else {
          double l, u;
          if (rt->second == 'L') { u = r.ub; l = u - std::fabs(v); }
          else if (rt->second == 'G') { l = r.lb; u = l + std::fabs(v); }
          else if (v >= 0) { l = r.lb; u = l + v; }
          else { u = r.ub; l = u + v; }
          int st = r.stat;
          set_typed(r.type, r.lb, r.ub, st, true, l, true, u);
        }
      }
    } else if (section == "BOUNDS") {
      if (w.size() < 3) return fail("bound entry expected");
      std::string bt = w[0];
      for (auto &c : bt) c = (char)std::toupper((unsigned char)c);
      bool valued = !(bt == "FR" || bt == "MI" || bt == "PL" || bt == "BV");
      // fields: type [set] column [value]
      size_t ci = (w.size() == (valued ? 4u : 3u)) ? 2 : 1;
      if (ci >= w.size()) return fail("bound entry expected");
      auto it = b.colidx.find(w[ci]);
      if (it == b.colidx.end()) return fail("unknown column '" + w[ci] + "'");
      MvCol &c = P->cols[it->second - 1];
      double v = valued && ci + 1 < w.size() ? std::strtod(w[ci + 1].c_str(), nullptr) : 0.0;
      double cl = mv_eff_lb(c.type, c.lb), cu = mv_eff_ub(c.type, c.ub);
      if (bt == "UP") { cu = v; if (v < 0 && cl == 0.0 && !col_bounded[it->second - 1]) cl = -HUGE_VAL; }
      else if (bt == "LO") cl = v;
      else if (bt == "FX") cl = cu = v;
      else if (bt == "FR") { cl = -HUGE_VAL; cu = HUGE_VAL; }
      else if (bt == "MI") cl = -HUGE_VAL;
      else if (bt == "PL") cu = HUGE_VAL;
      else if (bt == "BV") { cl = 0; cu = 1; c.kind = GLP_IV; }
      else if (bt == "LI") { cl = v; c.kind = GLP_IV; }
      else if (bt == "UI") { cu = v; c.kind = GLP_IV; }
      else return fail("unknown bound type '" + bt + "'");
      col_bounded[it->second - 1] = true;
      set_typed(c.type, c.lb, c.ub, c.stat, cl > -HUGE_VAL, cl, cu < HUGE_VAL, cu);
    }
  }
  // integer columns between markers without explicit bounds are binary-ranged
  // in classic MPS; GLPK keeps [0,+inf), and so do we.
  for (size_t i = 0; i < rowcoef.size(); i++) b.set_row((int)i + 1, rowcoef[i]);
  if (P->cols.empty()) { err = "problem has no columns"; return 1; }
  return 0;
}
