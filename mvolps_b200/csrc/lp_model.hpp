// lp_model.hpp -- host-side problem object behind the glpk-subset C ABI.
//
// The reference keeps one glp_prob per open B&B node (util.cpp:25-42) and
// mutates it through the GLPK API; this struct is what those calls act on.
// The dense simplex tableau itself lives in the engine (mvb_lp, on the GPU in
// the product build); this object owns the sparse rows, typed bounds, kinds,
// basis statuses and the last solution, and remembers what the engine has
// already seen so that a re-solve ships only bound deltas and new cut rows.
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "../../include/glpk.h"
#include "../../include/mvolps_engine.h"

// Copy-on-write vector: every open B&B node holds a copy of the problem
// (util.cpp:25-42) but only cut rows ever differ, so copies share the
// coefficient storage and glp_copy_prob costs a reference count per row.
template <class T>
class CowVec {
  std::shared_ptr<std::vector<T>> p_;
  std::vector<T> &mut() {
    if (!p_) p_ = std::make_shared<std::vector<T>>();
    else if (p_.use_count() > 1) p_ = std::make_shared<std::vector<T>>(*p_);
    return *p_;
  }

 public:
  size_t size() const { return p_ ? p_->size() : 0; }
  const T &operator[](size_t i) const { return (*p_)[i]; }
  void clear() { p_.reset(); }
  void push_back(const T &v) { mut().push_back(v); }
};

// Row/column names are never read through the 34-function boundary, only by the
// readers' messages: shared, so a problem copy does not duplicate the strings.
struct MvName {
  std::shared_ptr<const std::string> p;
  MvName &operator=(const std::string &s) {
    p = s.empty() ? nullptr : std::make_shared<const std::string>(s);
    return *this;
  }
  void clear() { p.reset(); }
  const std::string &str() const {
    static const std::string none;
    return p ? *p : none;
  }
};

struct MvRow {
  int type = GLP_FR;
  double lb = 0.0, ub = 0.0;
  CowVec<int> ind;     // 1-based column numbers, ascending
  CowVec<double> val;
  int stat = GLP_BS;
  double prim = 0.0;
  MvName name;
};

struct MvCol {
  int type = GLP_FX;
  double lb = 0.0, ub = 0.0;
  int kind = GLP_CV;
  double coef = 0.0;
  int stat = GLP_NS;
  double prim = 0.0;
  MvName name;
};

struct glp_prob {
  int dir = GLP_MIN;
  double c0 = 0.0;
  std::vector<MvRow> rows;  // rows[i-1]
  std::vector<MvCol> cols;  // cols[j-1]
  int status = GLP_UNDEF;
  double obj_val = 0.0;
  int it_cnt = 0;
  std::string name;

  // ---- engine mirror ----
  mvb_lp *eng = nullptr;
  int eng_rows = 0;            // rows the engine tableau already contains
  bool eng_stale = true;       // matrix/objective changed: rebuild from scratch
  bool eng_changed = true;     // the engine's LP differs from the one the held solution belongs to
                               // (sticky across syncs and copies until the next solve is pulled)
  // Round-synchronous multi-GPU tree (MVB_DIST_SYNC): every rank holds every node's problem object
  // and takes the same decisions, but only the node's OWNER holds its device tableau.  On the
  // other ranks the object is a ghost: the mirror below is kept exactly as on the owner (so
  // `changed`, cached answers and copies behave identically) and solutions arrive by exchange.
  bool ghost = false;
  int owner = 0;
  std::vector<double> eng_lb, eng_ub;  // bounds last sent, by engine id
  std::vector<int> dirty;              // engine ids whose bounds were set since the last sync
  std::vector<int> head, nb, at;       // basis after the last solve
  mvb_result last{};
};

// bound of a typed variable as the engine sees it (+-HUGE_VAL when absent)
double mv_eff_lb(int type, double lb);
double mv_eff_ub(int type, double ub);

// LP-format / MPS readers (lp_reader.cpp). 0 = ok, non-zero = error
int mv_read_lp(glp_prob *P, const char *fname, std::string &err);
int mv_read_mps(glp_prob *P, const char *fname, std::string &err);

// values closer than this to an integer are reported as that integer
constexpr double MV_SNAP_TOL = 1e-9;
double mv_snap(double x);

// glpk_shim.cpp helpers shared with the batched B&B driver
// push deltas to the engine; *changed: the engine's LP differs from the one last solved
int mv_sync_engine(glp_prob *P, bool *changed = nullptr);
// true when the LP is exactly the one whose optimal solution the object already holds (a
// solve would start from the optimal tableau, do zero iterations and return the same
// values); the iteration counters are then set to that zero-iteration solve's
bool mv_answer_cached(glp_prob *P, bool changed);
int mv_pull_solution(glp_prob *P, const mvb_result &res);   // read result back
// the same from caller-held arrays (x: n+m by engine id; head: m; nb, at: n): how a ghost
// receives the solution its owner computed
void mv_install_solution(glp_prob *P, const mvb_result &res, const double *x, const int *head,
                         const int *nb, const int *at);
// give up the device tableau but keep the mirror: the object becomes a ghost owned by `owner`
void mv_make_ghost(glp_prob *P, int owner);
