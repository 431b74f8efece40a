// lp_model.hpp -- host-side problem object behind the glpk-subset C ABI.
//
// The reference keeps one glp_prob per open B&B node (util.cpp:25-42) and
// mutates it through the GLPK API; this struct is what those calls act on.
// The dense simplex tableau itself lives in the engine (mvb_lp, on the GPU in
// the product build); this object owns the sparse rows, typed bounds, kinds,
// basis statuses and the last solution, and remembers what the engine has
// already seen so that a re-solve ships only bound deltas and new cut rows.
//
// Layout: structure of arrays, every array copy-on-write.  A B&B run makes
// two copies of the problem per branched node (util.cpp:34) that differ from
// their parent in ONE column bound (bs.cpp:274,282), perhaps one cut row
// (cut.cpp:23-43), and -- once solved -- the solution.  So glp_copy_prob is a
// handful of reference counts; the model (coefficients, kinds, objective,
// names) is shared by every node of the tree; a child detaches the bound array
// it writes; a solve installs fresh solution arrays without copying the old.
#pragma once

#include <algorithm>
#include <memory>
#include <string>
#include <vector>  // (the driver and the readers)

#include "../../include/glpk.h"
#include "../../include/mvolps_engine.h"
#include "cow_vec.hpp"

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

// ---- model part: the same for every node of a B&B tree ----
struct MvRowM {
  CowVec<int> ind;     // 1-based column numbers, ascending
  CowVec<double> val;
  MvName name;
};
struct MvColM {
  int kind = GLP_CV;
  double coef = 0.0;
  MvName name;
};
// ---- typed bounds: what branching (columns) and cut rows (rows) change ----
struct MvBnd {
  int type;
  double lb, ub;
};
constexpr MvBnd MV_ROW_BND0{GLP_FR, 0.0, 0.0};  // a new row is free, basic
constexpr MvBnd MV_COL_BND0{GLP_FX, 0.0, 0.0};  // a new column is fixed at zero, non-basic

// a bound the setters changed since the last sync: engine id and the effective bounds the
// engine holds for it (captured the first time the variable is touched)
struct MvDirty {
  int id;
  double lb, ub;
};

struct glp_prob {
  int dir = GLP_MIN;
  double c0 = 0.0;
  CowVec<MvRowM> rowm;  // rowm[i-1]
  CowVec<MvColM> colm;  // colm[j-1]
  CowVec<MvBnd> rbnd, cbnd;
  CowVec<int> rstat, cstat;      // GLP_BS / GLP_N*
  CowVec<double> rprim, cprim;   // last primal values (snapped, see mv_snap)
  int status = GLP_UNDEF;
  double obj_val = 0.0;
  int it_cnt = 0;
  std::string name;

  int nrows() const { return (int)rbnd.size(); }
  int ncols() const { return (int)cbnd.size(); }

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
  std::vector<MvDirty> dirty;          // bounds set since the last sync
  CowVec<int> head, nb, at;            // basis after the last solve (see basis_lazy)
  bool basis_lazy = false;             // the basis arrays are still in the engine only: mv_need_basis fetches them
                                       // (only the cut step and glp_eval_tab_row ever read them)
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
                         const int *nb, const int *at, bool keep_basis = true);
// make P->head / nb / at valid (fetches them from the engine when the last pull left them there)
int mv_need_basis(glp_prob *P);
// give up the device tableau but keep the mirror: the object becomes a ghost owned by `owner`
void mv_make_ghost(glp_prob *P, int owner);
// new rows / columns with the defaults GLPK gives them
void mv_add_rows(glp_prob *P, int count);
void mv_add_cols(glp_prob *P, int count);
