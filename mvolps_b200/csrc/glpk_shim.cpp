// glpk_shim.cpp -- the 34-function GLPK subset MVOLPS links against
// (SURVEY.md section 8b), implemented over the engine C ABI
// (include/mvolps_engine.h).  Host-only C++; every LP solve and tableau-row
// query is forwarded to the engine (CUDA in the product library).
//
// Semantics follow the public GLPK API manual; the call sites served are
// cited in include/glpk.h.
#include "lp_model.hpp"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

static int g_term_out = GLP_ON;

static void shim_abort(const char *what) {
  std::fprintf(stderr, "glpk-shim: %s\n", what);
  std::abort();
}

double mv_eff_lb(int type, double lb) {
  return (type == GLP_LO || type == GLP_DB || type == GLP_FX) ? lb : -HUGE_VAL;
}
double mv_eff_ub(int type, double ub) {
  return (type == GLP_UP || type == GLP_DB || type == GLP_FX) ? ub : HUGE_VAL;
}
/* Integer snap of the values the shim reports (glp_get_col_prim / glp_get_row_prim / the
 * objective computed from them).  The reference tests integrality EXACTLY (util.cpp:435-451,
 * trunc(x) != x) on what GLPK returns; a dense tableau that is never refactorised returns e.g.
 * 0.3/0.1 = 2.9999999999999996 for a vertex at 3, which that test would branch on (and the up
 * child x >= 3 is then infeasible by one ulp: the optimum would be lost).  Values within
 * g_snap_tol of an integer are therefore reported as that integer.  MVOLPS_SNAP_TOL=0 (or
 * mvb_glp_snap_tol(0)) turns it off; DESIGN.md section 5, tests/test_shim_and_bnb.py. */
static double g_snap_tol = [] {
  const char *e = std::getenv("MVOLPS_SNAP_TOL");
  return e ? std::atof(e) : MV_SNAP_TOL;
}();
extern "C" double mvb_glp_snap_tol(double tol) {
  const double old = g_snap_tol;
  if (tol >= 0.0) g_snap_tol = tol;
  return old;
}
double mv_snap(double x) {
  /* r = rint(x) without the libm call: below 2^51 adding and subtracting 2^52+2^51
   * rounds to nearest-even exactly as rint does; copysign keeps rint's signed zero */
  const double ax = std::fabs(x);
  if (!(ax < 2251799813685248.0)) return x;  // already integral (or inf/nan)
  const double r = std::copysign((ax + 6755399441055744.0) - 6755399441055744.0, x);
  return std::fabs(x - r) <= g_snap_tol ? r : x;
}

static void drop_engine(glp_prob *P) {
  if (P->eng) mvb_lp_free(P->eng);
  P->eng = nullptr;
  P->ghost = false;
  P->owner = 0;
  P->eng_rows = 0;
  P->eng_stale = true;
  P->eng_changed = true;
  P->dirty.clear();
  P->head.clear();
  P->nb.clear();
  P->at.clear();
  P->basis_lazy = false;
}

// 0-based index of row i / column j after the range check GLPK makes
static size_t row_at(glp_prob *P, int i, const char *fn) {
  if (i < 1 || i > P->nrows()) shim_abort(fn);
  return (size_t)i - 1;
}
static size_t col_at(glp_prob *P, int j, const char *fn) {
  if (j < 1 || j > P->ncols()) shim_abort(fn);
  return (size_t)j - 1;
}

void mv_add_rows(glp_prob *P, int count) {
  const size_t m = P->rbnd.size() + (size_t)count;
  P->rowm.resize(m);
  P->rbnd.resize(m, MV_ROW_BND0);
  P->rstat.resize(m, GLP_BS);
  P->rprim.resize(m, 0.0);
}
void mv_add_cols(glp_prob *P, int count) {
  const size_t n = P->cbnd.size() + (size_t)count;
  P->colm.resize(n);
  P->cbnd.resize(n, MV_COL_BND0);
  P->cstat.resize(n, GLP_NS);
  P->cprim.resize(n, 0.0);
}

// the setters note a variable's bounds as the engine holds them the first time they touch it
static void note_dirty(glp_prob *P, int id, const MvBnd &old) {
  for (const MvDirty &d : P->dirty)
    if (d.id == id) return;
  P->dirty.push_back(MvDirty{id, mv_eff_lb(old.type, old.lb), mv_eff_ub(old.type, old.ub)});
}

extern "C" {

glp_prob *glp_create_prob(void) { return new glp_prob(); }

void glp_erase_prob(glp_prob *P) {
  drop_engine(P);
  mvb_lp *none = nullptr;
  (void)none;
  glp_prob fresh;
  P->dir = fresh.dir;
  P->c0 = 0.0;
  P->rowm.clear(); P->colm.clear();
  P->rbnd.clear(); P->cbnd.clear();
  P->rstat.clear(); P->cstat.clear();
  P->rprim.clear(); P->cprim.clear();
  P->status = GLP_UNDEF;
  P->obj_val = 0.0;
  P->it_cnt = 0;
  P->name.clear();
  P->last = mvb_result{};
}

void glp_delete_prob(glp_prob *P) {
  if (!P) return;
  drop_engine(P);
  delete P;
}

void glp_copy_prob(glp_prob *dest, glp_prob *src, int names) {
  if (dest == src) shim_abort("glp_copy_prob: copying problem object to itself");
  glp_erase_prob(dest);
  dest->dir = src->dir;
  dest->c0 = src->c0;
  dest->rowm = src->rowm; dest->colm = src->colm;
  dest->rbnd = src->rbnd; dest->cbnd = src->cbnd;
  dest->rstat = src->rstat; dest->cstat = src->cstat;
  dest->rprim = src->rprim; dest->cprim = src->cprim;
  dest->status = src->status;
  dest->obj_val = src->obj_val;
  dest->it_cnt = src->it_cnt;
  if (names) dest->name = src->name;
  else {
    for (auto &r : dest->rowm.mut()) r.name.clear();
    for (auto &c : dest->colm.mut()) c.name.clear();
  }
  if ((src->eng || src->ghost) && !src->eng_stale) {
    if (!src->ghost) {
      dest->eng = mvb_lp_clone(src->eng);
      if (!dest->eng) shim_abort(mvb_last_error());
    }
    dest->ghost = src->ghost;
    dest->owner = src->owner;
    dest->eng_rows = src->eng_rows;
    dest->eng_stale = false;
    dest->eng_changed = src->eng_changed;
    dest->dirty = src->dirty;
    dest->head = src->head;
    dest->nb = src->nb;
    dest->at = src->at;
    dest->basis_lazy = src->basis_lazy;
    dest->last = src->last;
  }
}

int glp_read_lp(glp_prob *P, const void *parm, const char *fname) {
  (void)parm;
  glp_erase_prob(P);
  std::string err;
  int rc = mv_read_lp(P, fname, err);
  if (rc) {
    std::fprintf(stderr, "glp_read_lp: %s: %s\n", fname, err.c_str());
    glp_erase_prob(P);
  }
  return rc;
}

int glp_read_mps(glp_prob *P, int fmt, const void *parm, const char *fname) {
  (void)parm;
  (void)fmt;
  glp_erase_prob(P);
  std::string err;
  int rc = mv_read_mps(P, fname, err);
  if (rc) {
    std::fprintf(stderr, "glp_read_mps: %s: %s\n", fname, err.c_str());
    glp_erase_prob(P);
  }
  return rc;
}

const char *glp_version(void) {
  static char buf[64];
  std::snprintf(buf, sizeof buf, "%d.%d (mvolps-b200 %s)", GLP_MAJOR_VERSION,
                GLP_MINOR_VERSION, mvb_backend());
  return buf;
}

int glp_term_out(int flag) {
  int old = g_term_out;
  if (flag != GLP_ON && flag != GLP_OFF) shim_abort("glp_term_out: bad flag");
  g_term_out = flag;
  return old;
}

int glp_get_num_rows(glp_prob *P) { return P->nrows(); }
int glp_get_num_cols(glp_prob *P) { return P->ncols(); }
int glp_get_num_int(glp_prob *P) {
  int c = 0;
  for (const MvColM &col : P->colm) c += col.kind == GLP_IV;
  return c;
}

int glp_get_obj_dir(glp_prob *P) { return P->dir; }
void glp_set_obj_dir(glp_prob *P, int dir) {
  if (dir != GLP_MIN && dir != GLP_MAX) shim_abort("glp_set_obj_dir: bad dir");
  if (P->dir != dir) P->eng_stale = true;
  P->dir = dir;
}
double glp_get_obj_coef(glp_prob *P, int j) {
  if (j == 0) return P->c0;
  return P->colm[col_at(P, j, "glp_get_obj_coef: column out of range")].coef;
}
void glp_set_obj_coef(glp_prob *P, int j, double coef) {
  if (j == 0) {
    // the constant term does not touch the LP: a held solution keeps its point, its value moves
    if (coef != P->c0 && P->status != GLP_UNDEF) {
      double z = coef;
      for (int k = 0; k < P->ncols(); k++) z = std::fma(P->colm[k].coef, P->cprim[k], z);
      P->obj_val = z;
    }
    P->c0 = coef;
    return;
  }
  const size_t k = col_at(P, j, "glp_set_obj_coef: column out of range");
  if (P->colm[k].coef != coef) {
    P->eng_stale = true;
    P->colm.w(k).coef = coef;
  }
}
double glp_get_obj_val(glp_prob *P) { return P->obj_val; }

int glp_add_rows(glp_prob *P, int nrs) {
  if (nrs < 1) shim_abort("glp_add_rows: invalid number of rows");
  int first = P->nrows() + 1;
  mv_add_rows(P, nrs);
  return first;
}

int glp_add_cols(glp_prob *P, int ncs) {
  if (ncs < 1) shim_abort("glp_add_cols: invalid number of columns");
  int first = P->ncols() + 1;
  mv_add_cols(P, ncs);
  P->eng_stale = true;
  return first;
}

int glp_get_row_type(glp_prob *P, int i) { return P->rbnd[row_at(P, i, "glp_get_row_type")].type; }
double glp_get_row_lb(glp_prob *P, int i) {
  const MvBnd &r = P->rbnd[row_at(P, i, "glp_get_row_lb")];
  return (r.type == GLP_FR || r.type == GLP_UP) ? -DBL_MAX : r.lb;
}
double glp_get_row_ub(glp_prob *P, int i) {
  const MvBnd &r = P->rbnd[row_at(P, i, "glp_get_row_ub")];
  return (r.type == GLP_FR || r.type == GLP_LO) ? +DBL_MAX : r.ub;
}

static void store_bnds(int type, double lb, double ub, int &t, double &l,
                       double &u, int &stat) {
  switch (type) {
    case GLP_FR: l = u = 0.0; if (stat != GLP_BS) stat = GLP_NF; break;
    case GLP_LO: l = lb; u = 0.0; if (stat != GLP_BS) stat = GLP_NL; break;
    case GLP_UP: l = 0.0; u = ub; if (stat != GLP_BS) stat = GLP_NU; break;
    case GLP_DB:
      l = lb; u = ub;
      if (!(stat == GLP_BS || stat == GLP_NL || stat == GLP_NU))
        stat = std::fabs(lb) <= std::fabs(ub) ? GLP_NL : GLP_NU;
      break;
    case GLP_FX: l = u = lb; if (stat != GLP_BS) stat = GLP_NS; break;
    default: shim_abort("set bounds: invalid type");
  }
  t = type;
}

void glp_set_row_bnds(glp_prob *P, int i, int type, double lb, double ub) {
  const size_t k = row_at(P, i, "glp_set_row_bnds: row out of range");
  if ((P->eng || P->ghost) && i <= P->eng_rows) note_dirty(P, P->ncols() + (int)k, P->rbnd[k]);
  MvBnd b = P->rbnd[k];
  int stat = P->rstat[k];
  store_bnds(type, lb, ub, b.type, b.lb, b.ub, stat);
  P->rbnd.w(k) = b;
  if (stat != P->rstat[k]) P->rstat.w(k) = stat;
}

int glp_get_mat_row(glp_prob *P, int i, int ind[], double val[]) {
  const MvRowM &r = P->rowm[row_at(P, i, "glp_get_mat_row: row out of range")];
  int len = (int)r.ind.size();
  for (int t = 0; t < len; t++) {
    if (ind) ind[t + 1] = r.ind[t];
    if (val) val[t + 1] = r.val[t];
  }
  return len;
}

void glp_set_mat_row(glp_prob *P, int i, int len, const int ind[],
                     const double val[]) {
  const size_t k = row_at(P, i, "glp_set_mat_row: row out of range");
  int n = P->ncols();
  if (len < 0 || len > n) shim_abort("glp_set_mat_row: invalid row length");
  std::vector<std::pair<int, double>> e;
  e.reserve(len);
  bool ascending = true;
  for (int t = 1; t <= len; t++) {
    if (ind[t] < 1 || ind[t] > n) shim_abort("glp_set_mat_row: column index out of range");
    if (val[t] != 0.0) {
      if (!e.empty() && ind[t] < e.back().first) ascending = false;
      e.emplace_back(ind[t], val[t]);
    }
  }
  if (!ascending) std::sort(e.begin(), e.end(), [](auto &a, auto &b) { return a.first < b.first; });
  for (size_t t = 1; t < e.size(); t++)
    if (e[t].first == e[t - 1].first) shim_abort("glp_set_mat_row: duplicate column indices");
  MvRowM &r = P->rowm.w(k);
  int *ri = r.ind.fresh(e.size());
  double *rv = r.val.fresh(e.size());
  for (size_t t = 0; t < e.size(); t++) { ri[t] = e[t].first; rv[t] = e[t].second; }
  if (i <= P->eng_rows) P->eng_stale = true;  // a row the tableau already holds
}

int glp_get_row_stat(glp_prob *P, int i) { return P->rstat[row_at(P, i, "glp_get_row_stat")]; }
double glp_get_row_prim(glp_prob *P, int i) { return P->rprim[row_at(P, i, "glp_get_row_prim")]; }

int glp_get_col_type(glp_prob *P, int j) { return P->cbnd[col_at(P, j, "glp_get_col_type")].type; }
double glp_get_col_lb(glp_prob *P, int j) {
  const MvBnd &c = P->cbnd[col_at(P, j, "glp_get_col_lb")];
  return (c.type == GLP_FR || c.type == GLP_UP) ? -DBL_MAX : c.lb;
}
double glp_get_col_ub(glp_prob *P, int j) {
  const MvBnd &c = P->cbnd[col_at(P, j, "glp_get_col_ub")];
  return (c.type == GLP_FR || c.type == GLP_LO) ? +DBL_MAX : c.ub;
}
static void set_col_bnds(glp_prob *P, size_t k, int type, double lb, double ub) {
  if (P->eng || P->ghost) note_dirty(P, (int)k, P->cbnd[k]);
  MvBnd b = P->cbnd[k];
  int stat = P->cstat[k];
  store_bnds(type, lb, ub, b.type, b.lb, b.ub, stat);
  P->cbnd.w(k) = b;
  if (stat != P->cstat[k]) P->cstat.w(k) = stat;
}
void glp_set_col_bnds(glp_prob *P, int j, int type, double lb, double ub) {
  set_col_bnds(P, col_at(P, j, "glp_set_col_bnds: column out of range"), type, lb, ub);
}
int glp_get_col_kind(glp_prob *P, int j) {
  const size_t k = col_at(P, j, "glp_get_col_kind");
  const MvBnd &c = P->cbnd[k];
  const int kind = P->colm[k].kind;
  if (kind == GLP_IV && c.type == GLP_DB && c.lb == 0.0 && c.ub == 1.0) return GLP_BV;
  return kind;
}
void glp_set_col_kind(glp_prob *P, int j, int kind) {
  const size_t k = col_at(P, j, "glp_set_col_kind");
  if (kind == GLP_BV) {
    P->colm.w(k).kind = GLP_IV;
    set_col_bnds(P, k, GLP_DB, 0.0, 1.0);
  } else if (kind == GLP_CV || kind == GLP_IV) {
    if (P->colm[k].kind != kind) P->colm.w(k).kind = kind;
  } else shim_abort("glp_set_col_kind: invalid kind");
}
double glp_get_col_prim(glp_prob *P, int j) { return P->cprim[col_at(P, j, "glp_get_col_prim")]; }
int glp_get_col_stat(glp_prob *P, int j) { return P->cstat[col_at(P, j, "glp_get_col_stat")]; }
int glp_get_status(glp_prob *P) { return P->status; }
int glp_get_it_cnt(glp_prob *P) { return P->it_cnt; }

}  // extern "C"

/* bring the engine's LP in line with the problem object: cold build when the
 * matrix/objective changed, otherwise append the new rows (cut.cpp:23-43) and
 * forward changed bounds (bs.cpp:274,282) */
static int sync_engine_delta(glp_prob *P, bool *changed);
static int sync_engine(glp_prob *P, bool *changed) {
  bool now = false;
  const int rc = sync_engine_delta(P, &now);
  if (now) P->eng_changed = true;
  *changed = P->eng_changed;
  return rc;
}
static int sync_engine_delta(glp_prob *P, bool *changed) {
  *changed = false;
  int m = P->nrows(), n = P->ncols();
  const bool ghost = P->ghost;  // mirror only: the owner's rank makes the engine calls
  if ((P->eng || ghost) && !P->eng_stale && P->eng_rows > m) P->eng_stale = true;
  if ((!P->eng && !ghost) || P->eng_stale) {
    *changed = true;
    if (!ghost) {
      std::vector<double> lb(n + m), ub(n + m);
      for (int j = 0; j < n; j++) {
        lb[j] = mv_eff_lb(P->cbnd[j].type, P->cbnd[j].lb);
        ub[j] = mv_eff_ub(P->cbnd[j].type, P->cbnd[j].ub);
      }
      for (int i = 0; i < m; i++) {
        lb[n + i] = mv_eff_lb(P->rbnd[i].type, P->rbnd[i].lb);
        ub[n + i] = mv_eff_ub(P->rbnd[i].type, P->rbnd[i].ub);
      }
      if (P->eng) mvb_lp_free(P->eng);
      P->eng = mvb_lp_create(m, n, m + 32 + m / 8);
      if (!P->eng) return -1;
      // dense A built straight into the engine's (pinned) staging buffer: the upload is one DMA
      std::vector<double> own, cost(n);
      double *A = (double *)mvb_host_staging((size_t)m * n * sizeof(double));
      if (!A) { own.resize((size_t)m * n); A = own.data(); }
      for (int i = 0; i < m; i++) {
        const MvRowM &r = P->rowm[i];
        double *row = A + (size_t)i * n;
        const int len = (int)r.ind.size();
        const int *ind = r.ind.data();
        const double *val = r.val.data();
        if (len == n) {  // a dense row lists every column, ascending: a straight copy
          std::memcpy(row, val, (size_t)n * sizeof(double));
        } else {
          std::memset(row, 0, (size_t)n * sizeof(double));
          for (int t = 0; t < len; t++) row[ind[t] - 1] = val[t];
        }
      }
      for (int j = 0; j < n; j++)
        cost[j] = P->dir == GLP_MAX ? -P->colm[j].coef : P->colm[j].coef;
      if (mvb_lp_load(P->eng, A, cost.data(), lb.data(), ub.data())) return -1;
    }
    P->eng_rows = m;
    P->eng_stale = false;
    P->dirty.clear();
    return 0;
  }
  const int old_rows = P->eng_rows;
  if (P->eng_rows < m) {
    std::vector<double> a(ghost ? 0 : n);
    for (int i = P->eng_rows; i < m; i++) {
      if (!ghost) {
        std::fill(a.begin(), a.end(), 0.0);
        const MvRowM &r = P->rowm[i];
        for (size_t t = 0; t < r.ind.size(); t++) a[r.ind[t] - 1] = r.val[t];
        const double l = mv_eff_lb(P->rbnd[i].type, P->rbnd[i].lb);
        const double u = mv_eff_ub(P->rbnd[i].type, P->rbnd[i].ub);
        if (mvb_lp_append_row(P->eng, a.data(), l, u)) return -1;
      }
      *changed = true;
    }
    P->eng_rows = m;
  }
  /* bounds set since the last sync (the setters log the engine id and what the engine holds):
   * forward the ones that really changed, in ascending id order as a full scan would */
  if (!P->dirty.empty()) {
    if (P->dirty.size() > 1)
      std::sort(P->dirty.begin(), P->dirty.end(), [](const MvDirty &a, const MvDirty &b) { return a.id < b.id; });
    for (const MvDirty &d : P->dirty) {
      const int k = d.id;
      if (k >= n + old_rows) continue;  // a row appended above
      double l, u;
      if (k < n) {
        l = mv_eff_lb(P->cbnd[k].type, P->cbnd[k].lb);
        u = mv_eff_ub(P->cbnd[k].type, P->cbnd[k].ub);
      } else {
        l = mv_eff_lb(P->rbnd[k - n].type, P->rbnd[k - n].lb);
        u = mv_eff_ub(P->rbnd[k - n].type, P->rbnd[k - n].ub);
      }
      if (l != d.lb || u != d.ub) {
        if (!ghost && mvb_lp_set_bounds(P->eng, k, l, u)) return -1;
        *changed = true;
      }
    }
    P->dirty.clear();
  }
  return 0;
}

static int stat_of(int at) {
  switch (at) {
    case MVB_AT_LB: return GLP_NL;
    case MVB_AT_UB: return GLP_NU;
    case MVB_AT_FR: return GLP_NF;
    default: return GLP_NS;
  }
}

/* copy status, primal values and basis back from the engine */
int mv_pull_solution(glp_prob *P, const mvb_result &res) {
  const double *x;
  const int *head, *nb, *at;
  if (mvb_lp_solution_view(P->eng, &x, &head, &nb, &at)) return -1;
  mv_install_solution(P, res, x, head, nb, at, /*keep_basis=*/false);
  return 0;
}

int mv_need_basis(glp_prob *P) {
  if (!P->basis_lazy) return 0;
  if (!P->eng) return -1;
  const int m = P->nrows(), n = P->ncols();
  if (mvb_lp_get_basis(P->eng, P->head.fresh(m), P->nb.fresh(n), P->at.fresh(n))) return -1;
  P->basis_lazy = false;
  return 0;
}

void mv_install_solution(glp_prob *P, const mvb_result &res, const double *x, const int *head,
                         const int *nb, const int *at, bool keep_basis) {
  int m = P->nrows(), n = P->ncols();
  if (keep_basis) {
    P->head.assign(head, head + m);
    P->nb.assign(nb, nb + n);
    P->at.assign(at, at + n);
    P->basis_lazy = false;
  } else {
    P->basis_lazy = true;  // they stay in the engine's mirror until somebody asks (mv_need_basis)
  }
  // fresh arrays: a copy that still shares its parent's solution does not pay for a copy of it
  int *cstat = P->cstat.fresh(n), *rstat = P->rstat.fresh(m);
  double *cprim = P->cprim.fresh(n), *rprim = P->rprim.fresh(m);
  for (int i = 0; i < m; i++) {
    int k = head[i];
    if (k < n) cstat[k] = GLP_BS; else rstat[k - n] = GLP_BS;
  }
  for (int j = 0; j < n; j++) {
    int k = nb[j], st = stat_of(at[j]);
    if (k < n) cstat[k] = st; else rstat[k - n] = st;
  }
  const MvColM *cm = P->colm.data();
  double z = P->c0;
  for (int j = 0; j < n; j++) {
    const double v = mv_snap(x[j]);
    cprim[j] = v;
    z = std::fma(cm[j].coef, v, z);
  }
  for (int i = 0; i < m; i++) rprim[i] = mv_snap(x[n + i]);
  P->obj_val = z;
  P->it_cnt = res.iters;
  P->last = res;
  P->eng_changed = false;
  switch (res.status) {
    case MVB_OPT: P->status = GLP_OPT; break;
    case MVB_NOFEAS: P->status = GLP_NOFEAS; break;
    case MVB_UNBND: P->status = GLP_UNBND; break;
    default: P->status = res.primal_feasible ? GLP_FEAS : GLP_INFEAS; break;
  }
}

void mv_make_ghost(glp_prob *P, int owner) {
  mv_need_basis(P);  // the mirror must be complete before the engine object goes
  if (P->eng) mvb_lp_free(P->eng);
  P->eng = nullptr;
  P->ghost = true;
  P->owner = owner;
}

bool mv_answer_cached(glp_prob *P, bool changed) {
  if (changed || P->status != GLP_OPT || P->last.status != MVB_OPT) return false;
  P->it_cnt = 0;
  P->last.iters = P->last.pivots = P->last.flips = 0;
  return true;
}

int mv_sync_engine(glp_prob *P, bool *changed) {
  bool c = false;
  int rc = sync_engine(P, &c);
  if (changed) *changed = c;
  return rc;
}

#ifdef MVB_WITH_SHARDS
// ---- the column-sharded path behind glp_simplex (product library only; include/mvolps_dist.h) ----
#include "../../include/mvolps_dist.h"
static const mvb_comm *g_shard_comm = nullptr;
static int g_shard_local = 0;
static long long g_shard_min_bytes = 1LL << 30;

extern "C" int mvb_glp_use_shards(const mvb_comm *comm, int local_world, long long min_bytes) {
  g_shard_comm = comm;
  g_shard_local = comm ? 0 : local_world;
  g_shard_min_bytes = min_bytes > 0 ? min_bytes : (1LL << 30);
  return 0;
}

// 1: solved on the shard group (solution installed), 0: not eligible, -1: error
static int simplex_on_shards(glp_prob *P) {
  const int world = g_shard_comm ? g_shard_comm->world : g_shard_local;
  if (world < 2) return 0;
  const int m = P->nrows(), n = P->ncols();
  if ((P->eng && !P->eng_stale) || P->ghost) return 0;                // a tableau exists: warm start on one GPU
  if ((long long)(m + 1) * n * 8 < g_shard_min_bytes) return 0;
  std::vector<double> lb(n + m), ub(n + m);
  for (int j = 0; j < n; j++) {
    lb[j] = mv_eff_lb(P->cbnd[j].type, P->cbnd[j].lb);
    ub[j] = mv_eff_ub(P->cbnd[j].type, P->cbnd[j].ub);
    const bool zero_ok = lb[j] == 0.0 || (!(lb[j] > -HUGE_VAL) && (ub[j] == 0.0 || !(ub[j] < HUGE_VAL)));
    if (!zero_ok) return 0;  // the sharded path starts with every non-basic at value 0
  }
  for (int i = 0; i < m; i++) {
    lb[n + i] = mv_eff_lb(P->rbnd[i].type, P->rbnd[i].lb);
    ub[n + i] = mv_eff_ub(P->rbnd[i].type, P->rbnd[i].ub);
  }
  {
    // primal phase 2 or the dual simplex from the slack basis; the composite phase 1 is not sharded
    std::vector<double> c(n);
    for (int j = 0; j < n; j++) c[j] = P->dir == GLP_MAX ? -P->colm[j].coef : P->colm[j].coef;
    if (!mvb_shard_start_mode(m, n, c.data(), lb.data(), ub.data())) return 0;
  }
  const bool local = g_shard_comm == nullptr;
  std::vector<mvb_comm *> own;
  std::vector<mvb_shard *> shards;
  auto cleanup = [&] {
    for (mvb_shard *s : shards) mvb_shard_free(s);
    for (mvb_comm *c : own) mvb_comm_free(c);
  };
  for (int r = local ? 0 : g_shard_comm->rank; r < (local ? world : g_shard_comm->rank + 1); r++) {
    int c0 = 0, nc = 0;
    mvb_shard_partition(n, world, r, &c0, &nc);
    const mvb_comm *cm = g_shard_comm;
    if (local) { own.push_back(mvb_comm_local(r, world)); cm = own.back(); }
    mvb_shard *s = mvb_shard_create(m, n, c0, nc, cm);
    if (!s) { cleanup(); return -1; }
    shards.push_back(s);
    std::vector<double> A((size_t)m * std::max(nc, 1), 0.0), cost(std::max(nc, 1));
    for (int i = 0; i < m; i++) {
      const MvRowM &row = P->rowm[i];
      for (size_t t = 0; t < row.ind.size(); t++) {
        const int j = row.ind[t] - 1;
        if (j >= c0 && j < c0 + nc) A[(size_t)i * nc + (j - c0)] = row.val[t];
      }
    }
    for (int j = 0; j < nc; j++) cost[j] = P->dir == GLP_MAX ? -P->colm[c0 + j].coef : P->colm[c0 + j].coef;
    if (mvb_shard_load(s, A.data(), cost.data(), lb.data(), ub.data())) { cleanup(); return -1; }
  }
  mvb_result res;
  mvb_opts opts{};
  if (mvb_shard_solve(shards.data(), (int)shards.size(), &opts, &res)) { cleanup(); return -1; }
  std::vector<double> x(n + m);
  std::vector<int> head(m), nb(n), at(n);
  if (mvb_shard_get_primal(shards[0], x.data()) || mvb_shard_get_basis(shards[0], head.data(), nb.data(), at.data())) {
    cleanup();
    return -1;
  }
  mv_install_solution(P, res, x.data(), head.data(), nb.data(), at.data());
  // the tableau stays with the shards: this object holds the solution, not a factorisation
  P->eng_stale = true;
  P->eng_changed = true;
  cleanup();
  return 1;
}
#endif

extern "C" {

int glp_simplex(glp_prob *P, const glp_smcp *parm) {
  (void)parm;
  int m = P->nrows(), n = P->ncols();
  if (n == 0) {
    P->status = GLP_UNDEF;
    return GLP_EFAIL;
  }
  if (m == 0) {
    /* no rows: every column sits at the bound its cost prefers (GLPK solves such problems
     * too); nothing to send to the engine.  MVOLPS never gets here. */
    double z = P->c0;
    P->status = GLP_OPT;
    int *cstat = P->cstat.fresh(n);
    double *cprim = P->cprim.fresh(n);
    for (int j = 0; j < n; j++) {
      const MvBnd &c = P->cbnd[j];
      const double coef = P->colm[j].coef;
      const double cost = P->dir == GLP_MAX ? -coef : coef;
      const double l = mv_eff_lb(c.type, c.lb), u = mv_eff_ub(c.type, c.ub);
      double v;
      if (c.type == GLP_FX) { v = c.lb; cstat[j] = GLP_NS; }
      else if (cost > 0.0 || (cost == 0.0 && l > -HUGE_VAL)) { v = l; cstat[j] = GLP_NL; }
      else if (cost < 0.0 || u < HUGE_VAL) { v = u; cstat[j] = GLP_NU; }
      else { v = 0.0; cstat[j] = GLP_NF; }
      if (!(v > -HUGE_VAL && v < HUGE_VAL)) { P->status = GLP_UNBND; v = 0.0; }
      cprim[j] = v;
      z = std::fma(coef, v, z);
    }
    P->obj_val = z;
    P->it_cnt = 0;
    return 0;
  }
  if (P->ghost) shim_abort("glp_simplex: the object is a ghost of a multi-GPU tree (no device tableau on this rank)");
#ifdef MVB_WITH_SHARDS
  {
    const int rc = simplex_on_shards(P);
    if (rc < 0) {
      std::fprintf(stderr, "glp_simplex: sharded engine error: %s\n", mvb_last_error());
      return GLP_EFAIL;
    }
    if (rc > 0) return P->last.status == MVB_ITLIM ? GLP_EITLIM : 0;
  }
#endif
  bool changed = false;
  if (sync_engine(P, &changed)) {
    std::fprintf(stderr, "glp_simplex: engine error: %s\n", mvb_last_error());
    return GLP_EFAIL;
  }
  /* bs.cpp:117 re-solves a copy of an LP that was solved to optimality as a child
   * (bs.cpp:279,287): nothing changed, the answer is the one already held */
  if (mv_answer_cached(P, changed)) return 0;
  mvb_result res;
  mvb_opts opts{};
  if (mvb_solve_batch(&P->eng, 1, &opts, &res)) {
    std::fprintf(stderr, "glp_simplex: engine error: %s\n", mvb_last_error());
    return GLP_EFAIL;
  }
  if (mv_pull_solution(P, res)) return GLP_EFAIL;
  if (g_term_out == GLP_ON)
    std::printf("mvolps-b200[%s]: %d rows, %d cols: status %d, %d iterations "
                "(%d pivots, %d flips), obj = %.12g\n",
                mvb_backend(), m, n, res.status, res.iters, res.pivots,
                res.flips, P->obj_val);
  return res.status == MVB_ITLIM ? GLP_EITLIM : 0;
}

int glp_eval_tab_row(glp_prob *P, int k, int ind[], double val[]) {
  int m = P->nrows(), n = P->ncols();
  if (!P->eng || P->eng_stale || P->eng_rows != m || mv_need_basis(P) || (int)P->head.size() != m)
    shim_abort("glp_eval_tab_row: basis factorization does not exist");
  if (k < 1 || k > m + n) shim_abort("glp_eval_tab_row: variable number out of range");
  int id = k <= m ? n + (k - 1) : k - m - 1;
  int row = -1;
  for (int i = 0; i < m; i++)
    if (P->head[i] == id) { row = i + 1; break; }
  if (row < 0) shim_abort("glp_eval_tab_row: variable must be basic");
  std::vector<double> t(n);
  if (mvb_lp_get_tab_rows(P->eng, &row, 1, t.data())) shim_abort(mvb_last_error());
  std::vector<std::pair<int, double>> e;
  for (int p = 0; p < n; p++) {
    if (t[p] == 0.0) continue;
    int v = P->nb[p];
    e.emplace_back(v < n ? m + v + 1 : v - n + 1, t[p]);
  }
  std::sort(e.begin(), e.end(), [](auto &a, auto &b) { return a.first < b.first; });
  for (size_t s = 0; s < e.size(); s++) {
    ind[s + 1] = e[s].first;
    val[s + 1] = e[s].second;
  }
  return (int)e.size();
}

}  // extern "C"
