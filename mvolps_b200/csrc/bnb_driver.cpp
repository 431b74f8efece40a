// bnb_driver.cpp -- batched branch-and-bound over the engine C ABI.
//
// Restates bs.cpp:54-348 (loop), util.cpp:154-230 (node / variable choice),
// util.cpp:414-473 (status + violated list), util.cpp:11-23 (getFract),
// gmi.cpp:11-117 and cut.cpp:6-46 (cuts), util.cpp:25-42 (object ids).  The
// node order, ids, prune tags and incumbent updates are those of the serial
// reference; only the LP solves are grouped into engine launches (see
// include/mvolps_bnb.h).  Host-only C++; no LP arithmetic happens here.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <limits>
#include <map>
#include <memory>
#include <charconv>
#include <atomic>
#include <future>
#include <sstream>
#include <type_traits>
#include <string>
#include <vector>

#include "../../include/mvolps_bnb.h"
#include "../../include/mvolps_spec.h"
#include "lp_model.hpp"
#include "par_for.hpp"

namespace {

// Text formatting identical to util.h:124-128 (operator<< into an ostringstream with the
// default flags: integers in decimal, doubles as %g with precision 6) without building
// a stream per call -- every node logs several events.
inline void put(std::string &s, const char *v) { s += v; }
inline void put(std::string &s, const std::string &v) { s += v; }
inline void put(std::string &s, double v) {
  char b[48];
  auto r = std::to_chars(b, b + sizeof b, v, std::chars_format::general, 6);
  s.append(b, r.ptr);
}
template <typename I, typename = std::enable_if_t<std::is_integral<I>::value>>
inline void put(std::string &s, I v) {
  char b[24];
  auto r = std::to_chars(b, b + sizeof b, v);
  s.append(b, r.ptr);
}
template <typename... Args>
std::string sstr(Args &&...args) {
  std::string s;
  (put(s, args), ...);
  return s;
}

double get_fract(double x) {  // util.cpp:11-23
  double ip;
  double f = std::modf(x, &ip);
  if (f < 0.0) f += 1;
  return f;
}

enum Prune { P_INTG, P_FEAS, P_BNDS, P_NONE };  // util.h:27

// Best-first speculation: the two children of an open node, built and solved ahead of the
// node's turn (they are pure functions of the node's own LP solution when no cuts are added)
struct Spec {
  glp_prob *down = nullptr, *up = nullptr;  // S2 (bs.cpp:274) and S3 (bs.cpp:282)
  int pick = 0;
  ~Spec() {
    if (down) glp_delete_prob(down);
    if (up) glp_delete_prob(up);
  }
};

struct Node {  // NodeData, util.h:35-54
  glp_prob *prob = nullptr;
  double upperBound = std::numeric_limits<double>::infinity();
  bool initial = false;
  int oid = 0;
  bool spec_tried = false;      // looked at by a speculation pass (with or without children built)
  std::unique_ptr<Spec> spec;
  ~Node() { if (prob) glp_delete_prob(prob); }
};

struct TreeRec {
  int oid = 0, pid = 0, depth = 0;  // oid 0: no such node
  Prune prune = P_NONE;
  std::vector<int> kids;
};

// the tree, indexed by object id (ids are dense: util.cpp:25-37 hands them out 1, 2, 3, ...); a deque
// so that references stay valid while children are added
struct TreeTab {
  std::deque<TreeRec> v;
  TreeRec &operator[](int oid) {
    if ((size_t)oid >= v.size()) v.resize((size_t)oid + 1);
    return v[(size_t)oid];
  }
};

struct Cut {  // CutContainer, cut.h:7-13
  std::vector<int> inds;
  std::vector<double> vals;
  double lb = 0.0;
};

}  // namespace

struct mvb_bnb {
  mvb_bnb_stats st{};
  std::string tree, solution, events;
  std::vector<double> incumbent;  // 1-based
};

namespace {

struct Driver {
  const mvb_bnb_params &pr;
  glp_prob *root_prob;
  mvb_bnb *out;
  int next_id = 1;
  TreeTab tree;
  // CutPool (cut.h:15-23).  The reference keeps every cut it ever generated and only ever reads
  // the last one (cut.cpp:16-21); here the cuts of finished nodes are dropped except the last, and
  // pool_base counts them so that the pool "size" (cut ids) stays the reference's
  std::vector<Cut> pool;
  size_t pool_base = 0;
  void pool_trim() {
    if (pool.size() <= 1) return;
    pool_base += pool.size() - 1;
    Cut last = std::move(pool.back());
    pool.clear();
    pool.push_back(std::move(last));
  }
  // gmi.cpp never sets CutContainer::oid on success and bs.cpp:252 tests it
  // (`result.oid != -1`): the test reads whatever the caller's stack slot last
  // held.  As compiled by g++ -O2 that slot is stable across loop iterations,
  // so once any column was rejected (oid = -1, gmi.cpp:20,25) every later cut
  // is dropped.  MVB_CUT_REF models exactly that; MVB_CUT_GMI pools every cut.
  int ref_oid_slot = 0;
  unsigned long long ehash = MVB_FNV_OFFSET;

  const mvb_comm *comm = nullptr;  // node-sharded run (mvb_bnb_run_dist)

  Driver(const mvb_bnb_params &p, glp_prob *r, mvb_bnb *o) : pr(p), root_prob(r), out(o) {}

  // max-reduction over the ranks, in chunks the communicator accepts
  bool reduce_max(double *v, int count) {
    for (int o = 0; o < count; o += 32)
      if (comm->allreduce_max(comm->ctx, v + o, std::min(32, count - o))) return false;
    return true;
  }

  void hash_mix(unsigned long long v) { ehash = (ehash ^ v) * MVB_FNV_PRIME; }

  static const char *dir_of(int oid) {  // bs.cpp:43-52
    if (oid <= 1) return "M";
    return oid % 2 == 0 ? "R" : "L";
  }
  int pid_of(int oid) { return oid == 1 ? 0 : tree[oid].pid; }  // bs.cpp:27-34

  std::unique_ptr<Node> make_node(glp_prob *take) {  // NodeData ctor, util.cpp:25-37
    auto n = std::make_unique<Node>();
    n->oid = next_id++;
    n->prob = take;
    return n;
  }

  // ---- node hand-over between ranks (work stealing) ----
  // A node is the root problem + its column bounds + the cut rows appended on
  // the way down (bs.cpp:274,282; cut.cpp:38-45), so only those travel:
  // [upperBound, depth, extra rows, per column (type, lb, ub), per extra row
  // (type, lb, ub, len, (ind, val) x len)], all as doubles.
  void pack_node(const Node &nd, std::vector<double> &buf) {
    const glp_prob *a = nd.prob;
    const size_t m0 = (size_t)root_prob->nrows();
    buf.push_back(nd.upperBound);
    buf.push_back((double)tree[nd.oid].depth);
    buf.push_back((double)((size_t)a->nrows() - m0));
    for (const MvBnd &c : a->cbnd) { buf.push_back(c.type); buf.push_back(c.lb); buf.push_back(c.ub); }
    for (size_t i = m0; i < (size_t)a->nrows(); i++) {
      const MvRowM &r = a->rowm[i];
      const MvBnd &b = a->rbnd[i];
      buf.push_back(b.type); buf.push_back(b.lb); buf.push_back(b.ub);
      buf.push_back((double)r.ind.size());
      for (size_t t = 0; t < r.ind.size(); t++) { buf.push_back(r.ind[t]); buf.push_back(r.val[t]); }
    }
  }
  // rebuilds the node through the glpk boundary (the calls a host program would make), so the
  // engine mirror sees a fresh LP: its first solve starts from the slack basis
  std::unique_ptr<Node> unpack_node(const double *&p) {
    glp_prob *c = glp_create_prob();
    glp_copy_prob(c, root_prob, GLP_ON);
    const double ub = *p++;
    const int depth = (int)*p++;
    const int extra = (int)*p++;
    const int n = glp_get_num_cols(c);
    for (int j = 1; j <= n; j++, p += 3) {
      const MvBnd &r = root_prob->cbnd[j - 1];
      if ((int)p[0] != r.type || p[1] != r.lb || p[2] != r.ub) glp_set_col_bnds(c, j, (int)p[0], p[1], p[2]);
    }
    std::vector<int> ind;
    std::vector<double> val;
    for (int k = 0; k < extra; k++) {
      const int type = (int)p[0], len = (int)p[3];
      const int i = glp_add_rows(c, 1);
      glp_set_row_bnds(c, i, type, p[1], p[2]);
      ind.assign(1, 0); val.assign(1, 0.0);
      for (int t = 0; t < len; t++) { ind.push_back((int)p[4 + 2 * t]); val.push_back(p[5 + 2 * t]); }
      glp_set_mat_row(c, i, len, ind.data(), val.data());
      p += 4 + 2 * (size_t)len;
    }
    auto nd = make_node(c);
    nd->upperBound = ub;
    tree[nd->oid] = TreeRec{nd->oid, 1, depth, P_NONE, {}};
    tree[1].kids.push_back(nd->oid);
    return nd;
  }
  struct StealPair { int donor, thief, k; };
  // the matching every rank derives from the same open-node counts (cnt < 0: retired rank):
  // poorest unused rank below MVB_STEAL_LOW with the richest unused rank, half of the
  // difference, ties to the lowest rank; a rank takes part in one pair per round
  static std::vector<StealPair> steal_matching(const std::vector<double> &cnt) {
    std::vector<StealPair> pairs;
    std::vector<char> used(cnt.size(), 0);
    while (true) {
      int lo = -1, hi = -1;
      for (int r = 0; r < (int)cnt.size(); r++) {
        if (used[r] || cnt[r] < 0) continue;
        if (lo < 0 || cnt[r] < cnt[lo]) lo = r;
        if (hi < 0 || cnt[r] > cnt[hi]) hi = r;
      }
      if (lo < 0 || lo == hi || cnt[lo] >= MVB_STEAL_LOW || cnt[hi] - cnt[lo] < 2) break;
      const int k = std::min<int>((int)((cnt[hi] - cnt[lo]) / 2), MVB_STEAL_MAX);
      pairs.push_back(StealPair{hi, lo, k});
      used[lo] = used[hi] = 1;
    }
    return pairs;
  }

  double t_sync = 0, t_solve = 0, t_pull = 0, t_copy = 0;  // MVOLPS_BNB_TIMING=1 prints them
  double t_consume = 0, t_final = 0, t_spec = 0, t_pick = 0, t_rows = 0, t_gen = 0;
  static double now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
  }

  // ---- round-synchronous mode (MVB_DIST_SYNC): every rank runs this very driver
  // and takes the same decisions; after the replicated ramp-up every open node gets an OWNER
  // rank that alone keeps its device tableau (children inherit it); the LPs of a round are solved
  // by their owners and the solutions exchanged (one all-gather per launch), then consumed in the
  // serial order: oids, prune tags, events and tree text are the serial run's on every rank.
  bool sync_mode = false, dealt = false;
  static size_t rec_bytes(const glp_prob *p) {
    const size_t m = (size_t)p->nrows(), n = (size_t)p->ncols();
    return (sizeof(mvb_result) + 8 * (n + m) + 4 * (m + 2 * n) + 7) / 8 * 8;
  }
  bool allgather_segments(const std::vector<size_t> &seg, const std::vector<char> &mine, std::vector<char> &all,
                          size_t &stride) {
    stride = 8;
    for (size_t b : seg) stride = std::max(stride, b);
    std::vector<char> send(stride, 0);
    std::copy(mine.begin(), mine.end(), send.begin());
    all.resize(stride * seg.size());
    return comm->allgather_host(comm->ctx, send.data(), all.data(), stride) == 0;
  }

  // the solutions of one launch into their problem objects: every LP on its own (statuses, primal
  // values with the integer snap, objective), so the loop runs on the host worker threads
  bool pull_many(const std::vector<glp_prob *> &ps, const std::vector<mvb_result> &res, bool counted) {
    std::atomic<bool> bad{false};
    par_for(ps.size(), [&](size_t a, size_t b) {
      for (size_t k = a; k < b; k++)
        if (mv_pull_solution(ps[k], res[k])) bad = true;
    });
    if (bad) return false;
    if (counted)
      for (size_t k = 0; k < ps.size(); k++) out->st.simplex_iters += res[k].iters;
    return true;
  }

  // counted = false: speculative solves; they enter lp_solves / simplex_iters when (and if) the
  // serial order adopts them, so the statistics are those of the unspeculated run
  bool solve_many(const std::vector<glp_prob *> &all, bool counted = true) {
    if (all.empty()) return true;
    std::vector<glp_prob *> ps;
    std::vector<mvb_lp *> engs;
    double t0 = now();
    for (auto *p : all) {
      bool changed = false;
      if (mv_sync_engine(p, &changed)) return false;
      if (mv_answer_cached(p, changed)) { if (counted) out->st.cached_solves++; continue; }
      ps.push_back(p);
      if (!p->ghost) engs.push_back(p->eng);
    }
    if (counted) out->st.lp_solves += (int)all.size();
    double t1 = now();
    if (ps.empty()) { t_sync += t1 - t0; return true; }
    std::vector<mvb_result> res(std::max<size_t>(engs.size(), 1));
    mvb_opts o{};
    if (!fly_wait()) return false;  // (the engine solves one batch at a time)
    if (!engs.empty() && mvb_solve_batch(engs.data(), (int)engs.size(), &o, res.data())) return false;
    out->st.solved_here += (int)engs.size();
    double t2 = now();
    if (!(sync_mode && dealt)) {
      if (!pull_many(ps, res, counted)) return false;
    } else {
      // records of my LPs, in batch order: result | x | head | nb | at
      std::vector<size_t> seg(comm->world, 0);
      for (auto *p : ps) seg[p->owner] += rec_bytes(p);
      std::vector<char> mine(seg[comm->rank]), allrec;
      size_t off = 0, e = 0;
      for (auto *p : ps) {
        if (p->ghost) continue;
        const size_t m = (size_t)p->nrows(), n = (size_t)p->ncols();
        char *r = mine.data() + off;
        std::memcpy(r, &res[e++], sizeof(mvb_result));
        double *x = (double *)(r + sizeof(mvb_result));
        int *hd = (int *)(x + n + m);
        if (mvb_lp_get_primal(p->eng, x) || mvb_lp_get_basis(p->eng, hd, hd + m, hd + m + n)) return false;
        off += rec_bytes(p);
      }
      size_t stride = 0;
      if (!allgather_segments(seg, mine, allrec, stride)) return false;
      std::vector<size_t> cur(comm->world, 0);
      for (auto *p : ps) {
        const size_t m = (size_t)p->nrows(), n = (size_t)p->ncols();
        const char *r = allrec.data() + (size_t)p->owner * stride + cur[p->owner];
        cur[p->owner] += rec_bytes(p);
        mvb_result rr;
        std::memcpy(&rr, r, sizeof rr);
        const double *x = (const double *)(r + sizeof(mvb_result));
        const int *hd = (const int *)(x + n + m);
        mv_install_solution(p, rr, x, hd, hd + m, hd + m + n);
        if (counted) out->st.simplex_iters += rr.iters;
      }
    }
    t_sync += t1 - t0; t_solve += t2 - t1; t_pull += now() - t2;
    out->st.launches++;
    return true;
  }

  // solve_many in two halves (mvb_solve_batch_begin / _end): the launch of one round's children
  // stays in flight while the host consumes the next round's nodes (FIFO without cuts: those were
  // created -- and solved -- rounds ago; Driver::run keeps one batch in flight)
  std::vector<glp_prob *> fly_ps;
  bool fly_counted = true;
  bool solve_many_begin(const std::vector<glp_prob *> &all, bool counted = true) {
    fly_ps.clear();
    fly_counted = counted;
    if (all.empty()) return true;
    std::vector<mvb_lp *> engs;
    const double t0 = now();
    for (auto *p : all) {
      bool changed = false;
      if (mv_sync_engine(p, &changed)) return false;
      if (mv_answer_cached(p, changed)) { if (counted) out->st.cached_solves++; continue; }
      fly_ps.push_back(p);
      engs.push_back(p->eng);
    }
    if (counted) out->st.lp_solves += (int)all.size();
    const double t1 = now();
    t_sync += t1 - t0;
    if (fly_ps.empty()) return true;
    mvb_opts o{};
    if (mvb_solve_batch_begin(engs.data(), (int)engs.size(), &o)) return false;
    out->st.solved_here += (int)engs.size();
    t_solve += now() - t1;
    return true;
  }
  bool solve_many_end() {
    if (fly_ps.empty()) return true;
    std::vector<mvb_result> res(fly_ps.size());
    const double t0 = now();
    if (mvb_solve_batch_end(res.data())) return false;
    const double t1 = now();
    if (!pull_many(fly_ps, res, fly_counted)) return false;
    t_solve += t1 - t0;
    t_pull += now() - t1;
    out->st.launches++;
    fly_ps.clear();
    return true;
  }

  // The read-back of the launch in flight (wait for the device, host mirrors, solutions into the problem
  // objects) runs on a thread of its own while this one consumes the next round: it touches the
  // launch's LPs and problem objects only, which nobody else looks at before fly_wait().  Whoever
  // needs the engine for another solve, or the children themselves, calls fly_wait() first.
  std::future<bool> fly_done;
  bool fly_async = false, fly_ok = true;
  std::string fly_err;
  void fly_start() {
    fly_ok = true;
    if (fly_ps.empty() || getenv("MVOLPS_BNB_NO_ASYNC_PULL")) return;
    // (a thread per launch: worth it from a few tens of thousands of solution values on -- on the closing
    // instance, 38 values per LP, it cost 5-10 %)
    const glp_prob *p0 = fly_ps[0];
    if (fly_ps.size() * (size_t)(p0->nrows() + p0->ncols()) < 24576) return;
    fly_async = true;
    fly_done = std::async(std::launch::async, [this] {
      const bool ok = solve_many_end();
      if (!ok) fly_err = mvb_last_error();
      return ok;
    });
  }
  bool fly_wait() {  // idempotent; false: the read-back failed
    if (fly_async) {
      fly_async = false;
      fly_ok = fly_done.get();
      if (!fly_ok) std::fprintf(stderr, "mvb_bnb_run: read-back of a launch failed: %s\n", fly_err.c_str());
    }
    return fly_ok;
  }

  // sync mode, end of the ramp-up: owners are dealt round-robin over the open nodes; a rank gives
  // up the device tableau of every node it does not own
  void deal_owners(std::deque<std::unique_ptr<Node>> &open) {
    for (size_t k = 0; k < open.size(); k++) {
      const int owner = (int)(k % (size_t)comm->world);
      Node *nd = open[k].get();
      std::vector<glp_prob *> objs{nd->prob};
      if (nd->spec) { objs.push_back(nd->spec->down); objs.push_back(nd->spec->up); }
      for (glp_prob *p : objs) {
        if (!p) continue;
        if (owner != comm->rank) mv_make_ghost(p, owner);
        else p->owner = owner;
      }
    }
    dealt = true;
  }

  // printInfo, util.cpp:414-473: -1 infeasible, 1 integral, 0 otherwise
  int classify(glp_prob *a, std::vector<int> &violated) {
    violated.clear();
    int status = glp_get_status(a);
    if (status == GLP_NOFEAS || status == GLP_INFEAS || status == GLP_UNBND) return -1;
    // same test as through glp_get_col_prim / glp_get_obj_coef / glp_get_col_kind, on the
    // problem object directly (this runs for every node over every column)
    const int cols = a->ncols();
    const MvColM *c = a->colm.data();
    const double *prim = a->cprim.data();
    for (int i = 1; i <= cols; i++) {
      const double x = prim[i - 1];
      if (x != 0 && c[i - 1].coef != 0)
        if (std::trunc(x) != x && c[i - 1].kind != GLP_CV) violated.push_back(i);
    }
    return violated.empty() ? 1 : 0;
  }

  int pick_var(glp_prob *a, const std::vector<int> &vars) {  // util.cpp:190-230
    if (pr.var_strategy == MVB_VAR_VFP_NODE) {  // the same loop on the node's own solution
      double best = std::abs(get_fract(glp_get_col_prim(a, vars.front())) - 0.5);
      int index = vars.front();
      for (int i : vars) {
        double cur = std::abs(get_fract(glp_get_col_prim(a, i)) - 0.5);
        if (cur < best) { best = cur; index = i; }
      }
      return index;
    }
    if (pr.var_strategy == MVB_VAR_VFP) {
      // evaluated on the never-solved root object, as the reference does
      double best = std::abs(get_fract(glp_get_col_prim(root_prob, vars.front())) - 0.5);
      int index = vars.front();
      for (int i : vars) {
        double cur = std::abs(get_fract(glp_get_col_prim(root_prob, i)) - 0.5);
        if (cur < best) { best = cur; index = i; }
      }
      return index;
    }
    if (pr.var_strategy == MVB_VAR_VGO) {
      double bestCoef = 0.0;
      int index = vars.front();  // the reference leaves this uninitialised
      for (int i : vars) {
        double cur = glp_get_obj_coef(root_prob, i);
        if (cur > bestCoef) { bestCoef = cur; index = i; }
      }
      return index;
    }
    return vars.front();
  }

  // best-first only: upperBound / spec_tried of open[k], mirrored in contiguous arrays so that
  // the per-pop scan (util.cpp:170-187 walks the whole container) reads doubles, not nodes
  std::vector<double> open_ub;
  std::vector<char> open_tried;

  size_t pick_node(const std::deque<std::unique_ptr<Node>> &q) {  // util.cpp:154-188
    if (pr.node_strategy == MVB_NODE_BEST) {
      size_t best = 0;
      const double *ub = open_ub.data();
      for (size_t k = 1; k < q.size(); k++)
        if (ub[best] < ub[k]) best = k;
      return best;
    }
    return 0;
  }

  // Best-first speculation.  With "-bs 1" the reference pops ONE node per
  // iteration, so the device would see two child LPs per launch.  Without cuts the children of
  // an open node depend on nothing but that node's own LP solution (violated list, branching
  // variable, floor/ceil of its value: util.cpp:190-230, bs.cpp:269-288), so they can be built
  // and solved before the node's turn: when the best node has not been looked at yet, its
  // children and those of the next `width - 1` unvisited nodes in pop order are solved as one
  // launch.  The serial loop below then adopts them (ids, events, incumbent, prune tags in the
  // reference's order) and a speculated node that gets fathomed first simply drops them.
  bool speculate(std::deque<std::unique_ptr<Node>> &open, size_t best, double bestLower, int width) {
    std::vector<size_t> cand;
    for (size_t k = 0; k < open.size(); k++)
      if (k != best && !open_tried[k]) cand.push_back(k);
    const size_t take = std::min<size_t>((size_t)std::max(width - 1, 0), cand.size());
    std::partial_sort(cand.begin(), cand.begin() + take, cand.end(), [&](size_t a, size_t b) {
      const double ua = open_ub[a], ub = open_ub[b];
      return ua > ub || (ua == ub && a < b);  // pick_node's order: first maximum
    });
    cand.resize(take);
    cand.insert(cand.begin(), best);
    std::vector<glp_prob *> ps;
    std::vector<int> vars;
    for (size_t k : cand) {
      Node *nd = open[k].get();
      nd->spec_tried = true;
      open_tried[k] = 1;
      glp_prob *a = nd->prob;
      bool changed = false;
      if (mv_sync_engine(a, &changed)) return false;
      if (!mv_answer_cached(a, changed)) continue;  // not solved yet (root): the serial path does it
      if (classify(a, vars) != 0) continue;          // integral or infeasible: no children
      if (!(glp_get_obj_val(a) > bestLower)) continue;  // fathomed already
      auto sp = std::make_unique<Spec>();
      sp->pick = pick_var(a, vars);
      const double bound = glp_get_col_prim(a, sp->pick);
      sp->down = glp_create_prob();
      glp_copy_prob(sp->down, a, GLP_ON);
      sp->up = glp_create_prob();
      glp_copy_prob(sp->up, a, GLP_ON);
      glp_set_col_bnds(sp->down, sp->pick, GLP_UP, 0, std::floor(bound));  // bs.cpp:274
      glp_set_col_bnds(sp->up, sp->pick, GLP_LO, std::ceil(bound), 0);     // bs.cpp:282
      ps.push_back(sp->down);
      ps.push_back(sp->up);
      nd->spec = std::move(sp);
      out->st.spec_nodes++;
    }
    return solve_many(ps, /*counted=*/false);
  }

  bool cuts_wanted(int oid) const;
  std::vector<int> row_of;  // scratch of cut_rows_of
  // what cut_rows_of found out when only ONE row has to cross PCIe (textbook GMI, the last pooled cut
  // becomes the row, no free non-basic variable in the node: whether a column produces a cut is then
  // decided by the fraction of its value alone): lean_k = position among the accepted columns of the
  // last one that produces (-1: none), lean_pooled = how many produce; lean = false: every row
  bool lean = false;
  int lean_k = -1, lean_j = 0, lean_pooled = 0;
  void cut_rows_of(glp_prob *a, std::vector<int> &js, std::vector<int> &rows);
  // what the lean path knows about a node of the batch before the serial loop reaches it
  struct PreCut {
    bool lean = false, have = false;
    int j = 0, pooled = 0;
    Cut cut;
  };
  std::vector<PreCut> pre_cuts;  // by position in the batch (prefetch_cut_rows)
  void gen_cuts(glp_prob *a, const double *trows, PreCut *pre = nullptr);
  void gen_cut_ref(glp_prob *a, int j, const double *trow);
  void gen_cut_gmi(glp_prob *a, int j, const double *trow);
  static bool build_cut_gmi(glp_prob *a, int j, const double *trow, Cut &c);
  bool gmi_produces(glp_prob *a, int j, const double *trow);
  int add_cut_constraint(glp_prob *a, int cID);
  int deepest_cut(glp_prob *a, size_t first);

  // K7 for a whole round: the tableau rows the cut step (bs.cpp:249-258, gmi.cpp:36) will read,
  // for every node of the batch that can still branch (a superset: the incumbent only improves
  // while the batch is consumed), gathered by ONE engine call.  off[k]: offset of node k's rows
  // in buf, or npos when it has none.
  static constexpr size_t npos = (size_t)-1;
  bool prefetch_cut_rows(const std::vector<std::unique_ptr<Node>> &batch, double bestLower,
                         std::vector<double> &buf, std::vector<size_t> &off) {
    off.assign(batch.size(), npos);
    pre_cuts.assign(batch.size(), PreCut{});
    std::vector<mvb_lp *> lps;
    std::vector<int> rows, nrows, js, r1, vars;
    size_t tot = 0;
    for (size_t k = 0; k < batch.size(); k++) {
      Node *nd = batch[k].get();
      glp_prob *a = nd->prob;
      if (!cuts_wanted(nd->oid) || classify(a, vars) != 0) continue;
      if (glp_get_obj_val(a) <= bestLower) continue;  // fathomed whatever happens (bs.cpp:210)
      cut_rows_of(a, js, r1);
      pre_cuts[k].lean = lean;
      pre_cuts[k].j = lean_j;
      pre_cuts[k].pooled = lean ? lean_pooled : 0;
      off[k] = tot;
      tot += r1.size() * (size_t)a->ncols();
      if (r1.empty()) continue;
      lps.push_back(a->eng);
      nrows.push_back((int)r1.size());
      rows.insert(rows.end(), r1.begin(), r1.end());
    }
    buf.resize(tot);
    if (!(sync_mode && dealt)) {
      if (lps.empty()) return true;
      const double t0 = now();
      const int rc = mvb_lp_get_tab_rows_batch(lps.data(), (int)lps.size(), rows.data(), nrows.data(), buf.data());
      t_rows += now() - t0;
      out->st.launches_aux++;
      if (rc == 0) precompute_cuts(batch, buf, off);
      return rc == 0;
    }
    // sync mode: every rank gathers the rows of the nodes it owns, one all-gather shares them
    if (tot == 0) return true;
    std::vector<size_t> seg(comm->world, 0);
    std::vector<mvb_lp *> my_lps;
    std::vector<int> my_rows, my_nrows;
    {
      size_t t = 0;
      for (size_t k = 0; k < batch.size(); k++) {
        if (off[k] == npos) continue;
        glp_prob *a = batch[k]->prob;
        const size_t next = [&] { for (size_t j = k + 1; j < batch.size(); j++) if (off[j] != npos) return off[j]; return tot; }();
        const size_t cnt = (next - off[k]) / (size_t)a->ncols();
        seg[a->owner] += (next - off[k]) * sizeof(double);
        if (cnt && !a->ghost) {
          my_lps.push_back(a->eng);
          my_nrows.push_back((int)cnt);
          my_rows.insert(my_rows.end(), rows.begin() + t, rows.begin() + t + cnt);
        }
        t += cnt;
      }
    }
    std::vector<char> mine(seg[comm->rank]), allrows;
    if (!my_lps.empty() &&
        mvb_lp_get_tab_rows_batch(my_lps.data(), (int)my_lps.size(), my_rows.data(), my_nrows.data(), (double *)mine.data()))
      return false;
    out->st.launches_aux++;
    size_t stride = 0;
    if (!allgather_segments(seg, mine, allrows, stride)) return false;
    std::vector<size_t> cur(comm->world, 0);
    for (size_t k = 0; k < batch.size(); k++) {
      if (off[k] == npos) continue;
      glp_prob *a = batch[k]->prob;
      const size_t next = [&] { for (size_t j = k + 1; j < batch.size(); j++) if (off[j] != npos) return off[j]; return tot; }();
      const size_t bytes = (next - off[k]) * sizeof(double);
      std::memcpy(buf.data() + off[k], allrows.data() + (size_t)a->owner * stride + cur[a->owner], bytes);
      cur[a->owner] += bytes;
    }
    precompute_cuts(batch, buf, off);
    return true;
  }

  // lean path: the one cut a node can contribute depends on that node alone, so the cuts of the whole
  // batch are built on the host worker threads before the serial loop consumes the nodes in order
  void precompute_cuts(const std::vector<std::unique_ptr<Node>> &batch, const std::vector<double> &buf,
                       const std::vector<size_t> &off) {
    const double t0 = now();
    par_for(batch.size(), [&](size_t k0, size_t k1) {
      for (size_t k = k0; k < k1; k++) {
        PreCut &pc = pre_cuts[k];
        if (off[k] == npos || !pc.lean || pc.pooled == 0) continue;
        pc.have = build_cut_gmi(batch[k]->prob, pc.j, buf.data() + off[k], pc.cut);
      }
    }, 4);
    t_gen += now() - t0;
  }

  // root_heuristic: the root LP solution with its integer columns rounded down; feasible -> objective
  bool floor_point(glp_prob *a, std::vector<double> &x, double &obj) {
    const int n = a->ncols(), m = a->nrows();
    x.assign(n + 1, 0.0);
    obj = a->c0;
    for (int j = 1; j <= n; j++) {
      const MvBnd &c = a->cbnd[j - 1];
      double v = a->cprim[j - 1];
      if (a->colm[j - 1].kind != GLP_CV) v = std::floor(v);
      if (v < mv_eff_lb(c.type, c.lb) || v > mv_eff_ub(c.type, c.ub)) return false;
      x[j] = v;
      obj = std::fma(a->colm[j - 1].coef, v, obj);
    }
    for (int i = 0; i < m; i++) {
      const MvRowM &r = a->rowm[i];
      const MvBnd &b = a->rbnd[i];
      double act = 0.0;
      for (size_t t = 0; t < r.ind.size(); t++) act += r.val[t] * x[r.ind[t]];
      const double tol = 1e-9 * (1.0 + std::fabs(act));
      if (act < mv_eff_lb(b.type, b.lb) - tol || act > mv_eff_ub(b.type, b.ub) + tol) return false;
    }
    return true;
  }

  void print_tree(int oid, std::string &os) {  // tree_print.h:12-22
    TreeRec &t = tree[oid];
    char buf[16];
    os.append((size_t)t.depth, ' ');
    os += '-';
    os.append(buf, std::to_chars(buf, buf + sizeof buf, oid).ptr);
    if (t.prune == P_INTG) os += " I";
    else if (t.prune == P_FEAS) os += " F";
    else if (t.prune == P_BNDS) os += " B";
    os += '\n';
    for (int k : t.kids) print_tree(k, os);
  }

  bool run();
};

}  // namespace

namespace {

// CutPool::addCutConstraint, cut.cpp:11-46: pooled cut cID (the reference always
// passes -1 = the LAST pooled cut) becomes a new GLP_LO row of a (which is not
// re-solved before its children are cloned)
int Driver::add_cut_constraint(glp_prob *a, int cID) {
  if (pool.empty()) return -1;
  if (cID < 0) cID = (int)pool.size() - 1;
  const Cut &c = pool[cID];
  int index = glp_add_rows(a, 1);
  glp_set_mat_row(a, index, (int)c.inds.size() - 1, c.inds.data(), c.vals.data());
  glp_set_row_bnds(a, index, GLP_LO, c.lb, 0);
  return (int)pool_base + cID;
}

// cut_select = 1 (the "heuristic to pick the best cut" cut.cpp:16-19 leaves as a TODO): among
// the cuts pool[first..) the one the node's LP solution violates most, (lb - a.x)/|a|, first
// on ties; -1 when none of them is violated
int Driver::deepest_cut(glp_prob *a, size_t first) {
  int best = -1;
  double bestv = 0.0;
  for (size_t k = first; k < pool.size(); k++) {
    const Cut &c = pool[k];
    double act = 0.0, nrm = 0.0;
    for (size_t t = 1; t < c.inds.size(); t++) {
      act += c.vals[t] * a->cprim[c.inds[t] - 1];
      nrm += c.vals[t] * c.vals[t];
    }
    if (!(nrm > 0.0)) continue;
    const double v = (c.lb - act) / std::sqrt(nrm);
    if (v > bestv) { bestv = v; best = (int)k; }
  }
  return best;
}

// literal restatement of generateCut3, gmi.cpp:11-117 (see SURVEY.md 3.4 for
// its quirks; `temp` is zero-initialised here where the reference reads an
// indeterminate value for GLP_BV columns)
void Driver::gen_cut_ref(glp_prob *a, int j, const double *trow) {
  int m = glp_get_num_rows(a), n = glp_get_num_cols(a);
  std::vector<double> work(m + n + 1, 0.0);
  // glp_eval_tab_row ordering: ascending variable number, zeros dropped
  std::vector<std::pair<int, double>> e;
  for (int p = 0; p < n; p++) {
    if (trow[p] == 0.0) continue;
    int v = a->nb[p];
    e.emplace_back(v < n ? m + v + 1 : v - n + 1, trow[p]);
  }
  std::sort(e.begin(), e.end(), [](auto &x, auto &y) { return x.first < y.first; });
  double rhs = glp_get_col_prim(a, j);
  double temp = 0.0;
  for (auto &kv : e) {
    double val = kv.second;
    int kind;
    double ub;
    if (kv.first <= m) {
      kind = GLP_CV;
      ub = glp_get_row_ub(a, kv.first);
    } else {
      kind = glp_get_col_kind(a, kv.first - m);
      ub = glp_get_col_ub(a, kv.first - m);
    }
    double fRhs = get_fract(rhs), fVal = get_fract(val);
    if (kind == GLP_IV) temp = fRhs >= fVal ? fVal : (fRhs / (1.0 - fRhs)) * (1.0 - fVal);
    if (kind == GLP_CV) temp = val >= 0.0 ? val : (fRhs / (1.0 - fRhs)) * (-1.0 * val);
    work[kv.first] = -1.0 * temp;
    rhs -= temp * ub;
  }
  for (int i = 1; i <= m; i++) {
    const MvRowM &r = a->rowm[i - 1];
    for (size_t k = 1; k <= r.val.size(); k++)  // indexes by position, as gmi.cpp:86-88
      work[m + k] += work[i] * r.val[k - 1];
  }
  Cut c;
  c.inds.push_back(0);
  c.vals.push_back(rhs);
  for (int i = 1; i <= n; i++) { c.inds.push_back(i); c.vals.push_back(work[m + i]); }
  c.lb = rhs;
  pool.push_back(std::move(c));
}

// whether gen_cut_gmi(a, j, trow) pools a cut (its two early exits, nothing else)
bool Driver::gmi_produces(glp_prob *a, int j, const double *trow) {
  const double f0 = get_fract(glp_get_col_prim(a, j));
  if (f0 < 1e-6 || f0 > 1.0 - 1e-6) return false;
  const int n = a->ncols();
  const int *at = a->at.data();
  for (int p = 0; p < n; p++)
    if (trow[p] != 0.0 && at[p] == MVB_AT_FR) return false;  // free non-basic in the row: no cut
  return true;
}

// textbook Gomory mixed-integer cut of the row of basic integer column j:
// with non-basics shifted to their active bounds (y >= 0),
//   x_j + sum a_k y_k = beta,  f0 = frac(beta),
//   integer y_k   : g = f_k/f0 if f_k <= f0 else (1-f_k)/(1-f0), f_k = frac(a_k)
//   continuous y_k: g = a_k/f0 if a_k >= 0 else -a_k/(1-f0)
//   sum g_k y_k >= 1, then y_k substituted back to structurals.
void Driver::gen_cut_gmi(glp_prob *a, int j, const double *trow) {
  Cut c;
  if (build_cut_gmi(a, j, trow, c)) pool.push_back(std::move(c));
}

// the cut itself: reads the node, touches nothing else (the batch's cuts are built on the host
// worker threads ahead of the serial loop, precompute_cuts)
bool Driver::build_cut_gmi(glp_prob *a, int j, const double *trow, Cut &c) {
  int n = glp_get_num_cols(a);
  double beta = glp_get_col_prim(a, j);
  double f0 = get_fract(beta);
  if (f0 < 1e-6 || f0 > 1.0 - 1e-6) return false;
  std::vector<double> g(n + 1, 0.0);  // cut over structurals, 1-based
  double rhs = 1.0;
  for (int p = 0; p < n; p++) {
    double alpha = trow[p];
    if (alpha == 0.0) continue;
    int v = a->nb[p], at = a->at[p];
    if (at == MVB_AT_FX) continue;
    if (at == MVB_AT_FR) return false;  // free non-basic in the row: no cut
    bool is_int = false;
    double bound;
    if (v < n) {
      const MvBnd &c = a->cbnd[v];
      bound = at == MVB_AT_LB ? c.lb : c.ub;
      is_int = a->colm[v].kind == GLP_IV && std::trunc(bound) == bound;
    } else {
      const MvBnd &r = a->rbnd[v - n];
      bound = at == MVB_AT_LB ? r.lb : r.ub;
    }
    // x_j = beta + sum alpha*(x_k - v_k): y = x_k - l (at LB) or u - x_k (at UB)
    double ak = at == MVB_AT_LB ? -alpha : alpha;  // coefficient of y_k on the lhs
    double gk;
    if (is_int) {
      double fk = get_fract(ak);
      gk = fk <= f0 ? fk / f0 : (1.0 - fk) / (1.0 - f0);
    } else {
      gk = ak >= 0.0 ? ak / f0 : -ak / (1.0 - f0);
    }
    if (gk == 0.0) continue;
    // gk*y_k with y_k = s*(x_k - bound), s = +1 at LB, -1 at UB
    double s = at == MVB_AT_LB ? 1.0 : -1.0;
    rhs += gk * s * bound;
    if (v < n) g[v + 1] += gk * s;
    else {
      // a non-basic auxiliary: its row in structural space (this loop is the cut step's cost: up to
      // m rows of n entries per cut)
      const MvRowM &r = a->rowm[v - n];
      const size_t len = r.ind.size();
      const double c = gk * s;
      if ((int)len == n && r.ind[0] == 1 && r.ind[len - 1] == n) {
        // a dense row (indices 1..n in order, glp_set_mat_row sorts them): same products and sums,
        // element by element, without the index loads -- the compiler vectorises it
        const double *__restrict__ rv = r.val.data();
        double *__restrict__ gp = g.data() + 1;
        for (int t = 0; t < n; t++) gp[t] += c * rv[t];
      } else {
        for (size_t t = 0; t < len; t++) g[r.ind[t]] += c * r.val[t];
      }
    }
  }
  c.inds.clear();
  c.vals.clear();
  c.inds.reserve(n + 1);
  c.vals.reserve(n + 1);
  c.inds.push_back(0);
  c.vals.push_back(rhs);
  for (int i = 1; i <= n; i++) { c.inds.push_back(i); c.vals.push_back(g[i]); }
  c.lb = rhs;
  return true;
}

// "-cf" (2test.cpp:135-149): whether the cut step of bs.cpp:249-258 runs at the node with this
// oid; a fixed hash of the oid, so a run is reproducible and identical on every rank
bool Driver::cuts_wanted(int oid) const {
  if (!(pr.cut_skip > 0.0)) return true;
  unsigned long long z = (unsigned long long)oid * 0x9E3779B97F4A7C15ULL;  // splitmix64 finaliser
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  z ^= z >> 31;
  return (double)(z >> 11) * (1.0 / 9007199254740992.0) >= pr.cut_skip;
}

// the tableau rows gen_cuts needs from node a: one per column that generates (basic, of an
// admitted kind), in column order; js[k] < 0 marks a rejected column (no row)
void Driver::cut_rows_of(glp_prob *a, std::vector<int> &js, std::vector<int> &rows) {
  int m = glp_get_num_rows(a), n = glp_get_num_cols(a);
  js.clear();
  rows.clear();
  mv_need_basis(a);
  row_of.assign(n, 0);
  for (int i = 0; i < m; i++)
    if (a->head[i] < n) row_of[a->head[i]] = i + 1;
  js.reserve(n);
  {
    // (glp_get_col_kind / glp_get_col_stat of every column, read from the arrays they read)
    const MvColM *cm = a->colm.data();
    const MvBnd *cb = a->cbnd.data();
    const int *cs = a->cstat.data();
    for (int j = 1; j <= n; j++) {
      int kind = cm[j - 1].kind;
      if (kind == GLP_IV && cb[j - 1].type == GLP_DB && cb[j - 1].lb == 0.0 && cb[j - 1].ub == 1.0) kind = GLP_BV;
      bool ok = pr.cut_mode == MVB_CUT_REF ? kind == GLP_IV : kind != GLP_CV;
      if (!ok || cs[j - 1] != GLP_BS || !row_of[j - 1]) { js.push_back(-j); continue; }
      js.push_back(j);
      rows.push_back(row_of[j - 1]);
    }
  }
  lean = false;
  if (pr.cut_mode == MVB_CUT_GMI && pr.cut_select == 0) {
    bool any_free = false;
    for (int p = 0; p < n; p++) any_free |= a->at[p] == MVB_AT_FR;
    if (!any_free) {
      // gmi_produces without the row: with no free non-basic variable it is the fraction test alone
      lean = true;
      lean_k = -1;
      lean_pooled = 0;
      int k = 0, last_row = 0;
      for (int j : js) {
        if (j < 0) continue;
        const double f0 = get_fract(glp_get_col_prim(a, j));
        if (!(f0 < 1e-6 || f0 > 1.0 - 1e-6)) { lean_k = k; lean_j = j; lean_pooled++; last_row = rows[k]; }
        k++;
      }
      rows.clear();
      if (lean_k >= 0) rows.push_back(last_row);
    }
  }
}

// bs.cpp:249-258: every basic GLP_IV column generates; the rows were gathered from the engine
// tableaus of all branching candidates of the round in one call (prefetch_cut_rows), so only
// those rows cross PCIe, once per round
void Driver::gen_cuts(glp_prob *a, const double *t, PreCut *pre) {
  if (pre && pre->lean) {  // everything was worked out ahead of the loop (prefetch_cut_rows / precompute_cuts)
    pool_trim();
    const size_t pool0 = pool.size();
    if (pre->pooled) {
      pool_base += pre->pooled - 1;
      if (pre->have) pool.push_back(std::move(pre->cut));
    }
    if (pool.size() > pool0) add_cut_constraint(a, -1);
    return;
  }
  int n = glp_get_num_cols(a);
  std::vector<int> js, rows;
  cut_rows_of(a, js, rows);
  pool_trim();
  const size_t pool0 = pool.size();
  if (lean) {
    // (cut_rows_of has walked the columns already: `t` holds the one row of the last column that
    // produces a cut; the others are counted, as below)
    if (lean_pooled) {
      pool_base += lean_pooled - 1;
      gen_cut_gmi(a, lean_j, t);
    }
  } else if (pr.cut_select == 0) {
    // Only the LAST cut the loop of bs.cpp:250-255 pools is ever read (cut.cpp:16-21), so only that
    // one is built: walk the columns in order to find which generator call is the last to pool a
    // cut (same accept / reject rules, same ref_oid_slot bookkeeping), count the others.
    int last_j = 0;
    size_t last_k = 0, k = 0, pooled = 0;
    for (int j : js) {
      if (j < 0) { ref_oid_slot = -1; continue; }
      const bool pools = pr.cut_mode == MVB_CUT_GMI ? gmi_produces(a, j, t + k * n) : ref_oid_slot != -1;
      if (pools) { last_j = j; last_k = k; pooled++; }
      k++;
    }
    if (pooled) {
      pool_base += pooled - 1;
      if (pr.cut_mode == MVB_CUT_GMI) gen_cut_gmi(a, last_j, t + last_k * n);
      else gen_cut_ref(a, last_j, t + last_k * n);
    }
  } else {
    size_t k = 0;
    for (int j : js) {
      if (j < 0) { ref_oid_slot = -1; continue; }  // rejected column
      if (pr.cut_mode == MVB_CUT_GMI) gen_cut_gmi(a, j, t + k * n);
      else if (ref_oid_slot != -1) gen_cut_ref(a, j, t + k * n);
      k++;
    }
  }
  if (pr.cut_select == 1) {
    const int pick = deepest_cut(a, pool0);
    if (pick >= 0) add_cut_constraint(a, pick);
  } else if (pr.cut_mode == MVB_CUT_GMI) {
    // a textbook GMI cut uses the node's own bounds: it holds in this node's subtree only, so
    // the row must be one this node produced (none produced: none added)
    if (pool.size() > pool0) add_cut_constraint(a, -1);
  } else {
    add_cut_constraint(a, -1);  // cut.cpp:16-21: the last pooled cut, whichever node made it
  }
}

}  // namespace

namespace {

bool Driver::run() {
  const double t_run0 = now();
  {
    // host worker threads of the per-LP loops: the ranks of a multi-GPU run share the cores
    const int hw = (int)std::thread::hardware_concurrency(), w = comm ? std::max(1, comm->world) : 1;
    // (half the cores of a rank's share: at 8 ranks on 32 cores 2 threads per rank measured 720 k nodes/s,
    // 4 threads 678 k -- profiles/r2_runs/r2z_n8_bench.json against r3f_n8_bench.json)
    ParPool::get().set_threads(std::max(1, std::min(4, hw / (2 * w))));
  }
  const double inf = std::numeric_limits<double>::infinity();
  std::deque<std::unique_ptr<Node>> open;  // leafContainer, bs.cpp:79
  std::vector<std::string> evs;            // event lines (candidates filled late)
  std::vector<double> ev_ms;               // timed_events: BaseMessagePOD::timeSpan (message.h:142-146)
  // no_event_text: nothing is formatted or kept; the hash takes what it would have taken from the
  // line -- the characters of "type oid pid", up to the third blank -- as the event happens
  const bool keep_text = !pr.no_event_text;
  size_t nev = 0;
  auto ev_line = [&](const char *type, int oid, const std::string &fields) {
    if (!keep_text) {
      char buf[32];
      for (const char *c = type; *c; c++) hash_mix((unsigned char)*c);
      hash_mix((unsigned char)' ');
      for (const char *c = buf, *e = std::to_chars(buf, buf + sizeof buf, oid).ptr; c < e; c++) hash_mix((unsigned char)*c);
      hash_mix((unsigned char)' ');
      for (const char *c = buf, *e = std::to_chars(buf, buf + sizeof buf, pid_of(oid)).ptr; c < e; c++) hash_mix((unsigned char)*c);
      return nev++;
    }
    if (pr.timed_events) ev_ms.push_back(1e3 * (now() - t_run0));
    evs.push_back(sstr(type, " ", oid, " ", pid_of(oid), " ", dir_of(oid), fields));
    return evs.size() - 1;
  };
  // the fields of an event line (formatted only when the text is kept)
  auto fld = [&](auto &&...a) { return keep_text ? sstr(a...) : std::string(); };
  // One launch in flight (FIFO without cuts, not the round-synchronous mode): the children of round t
  // are solved while the host consumes round t + 1, whose nodes were created -- and solved -- long
  // before; the children's bounds (and the tail of their "candidate" events) are filled in when the
  // launch is read back, which happens before any of them is popped, handed to another rank, or the
  // next launch begins.  Decisions, ids, events and statistics are those of the unpipelined loop.
  struct {
    std::vector<Node *> nodes;
    std::vector<size_t> slots;
    int min_oid = std::numeric_limits<int>::max();
    bool active = false;
  } fly;
  auto finish_fly = [&]() -> bool {
    if (!fly.active) return true;
    fly.active = false;
    if (fly_async || !fly_ok) { if (!fly_wait()) return false; }
    else if (!solve_many_end()) return false;  // (also a no-op after fly_wait: its list is empty then)
    for (size_t i = 0; i < fly.nodes.size(); i++) {
      Node *k = fly.nodes[i];
      k->upperBound = glp_get_obj_val(k->prob);
      if (keep_text) evs[fly.slots[i]] += sstr(" ", k->upperBound);
    }
    fly.nodes.clear();
    fly.slots.clear();
    fly.min_oid = std::numeric_limits<int>::max();
    return true;
  };

  {  // S1, bs.cpp:80-87
    glp_prob *c = glp_create_prob();
    glp_copy_prob(c, root_prob, GLP_ON);
    auto s1 = make_node(c);
    s1->initial = true;
    tree[s1->oid] = TreeRec{s1->oid, 0, 0, P_NONE, {}};
    open_ub.push_back(s1->upperBound);
    open_tried.push_back(0);
    open.push_back(std::move(s1));
  }
  double bestLower = -inf;
  std::string solution;
  int count = 0;
  bool stop = false;
  int width = pr.batch_width > 0 ? pr.batch_width : 256;
  int limit = pr.max_iters;  // node budget of this rank (total_budget: re-derived when the queue is dealt out)

  struct Pending {  // children awaiting their LP solve (bs.cpp:279,287)
    std::unique_ptr<Node> node;
    size_t ev_slot;
    bool solved;  // adopted from a speculation pass
  };
  const bool spec_on = pr.node_strategy == MVB_NODE_BEST && pr.cut_mode == MVB_CUT_NONE && width > 1 &&
                       !pr.no_speculation;

  sync_mode = comm != nullptr && comm->world > 1 && pr.dist_mode == MVB_DIST_SYNC;
  const bool dist = comm != nullptr && comm->world > 1 && !sync_mode;
  const bool pipe = pr.node_strategy != MVB_NODE_BEST && pr.cut_mode == MVB_CUT_NONE && !sync_mode &&
                    !getenv("MVOLPS_BNB_NO_PIPELINE");
  const bool steal = dist && !pr.no_steal && comm->sendrecv_host != nullptr;
  bool split = false;        // dist: the queue has been dealt out, ranks now work on their own share
  bool retired = false;      // dist: this rank hit its budget and only takes part in the exchanges
  int inc_rank = 0;
  out->st.split_at = -1;
  while (true) {
    if (dist && split) {     // termination is global: somebody still has open nodes
      // ONE reduction per round: the open-node counts and, in the last slot, the incumbent -- the bound
      // test (bs.cpp:210) sees the best incumbent of all ranks
      std::vector<double> cnt(comm->world + 1, -inf);
      cnt[comm->rank] = retired ? -1.0 : (double)open.size();
      cnt[comm->world] = bestLower;
      if (!reduce_max(cnt.data(), comm->world + 1)) return false;
      if (cnt[comm->world] > bestLower) { bestLower = cnt[comm->world]; solution.clear(); out->st.has_incumbent = 0; }
      cnt.pop_back();
      if (*std::max_element(cnt.begin(), cnt.end()) <= 0.0) break;
      if (steal) {           // work stealing: poor ranks take over open nodes of rich ones
        const auto pairs = steal_matching(cnt);
        if (!pairs.empty()) out->st.steal_rounds++;
        for (const StealPair &sp : pairs) {
          if (sp.donor == comm->rank) {
            if (!finish_fly()) return false;  // the newest nodes go: they may be the ones in flight
            std::vector<double> buf(1, 0.0);
            for (int k = 0; k < sp.k; k++) {  // newest first
              pack_node(*open.back(), buf);
              open.pop_back();
            }
            buf[0] = (double)buf.size();
            if (comm->sendrecv_host(comm->ctx, buf.data(), sizeof(double), sp.thief, nullptr, 0, -1)) return false;
            if (comm->sendrecv_host(comm->ctx, buf.data(), buf.size() * sizeof(double), sp.thief, nullptr, 0, -1))
              return false;
            out->st.nodes_sent += sp.k;
          } else if (sp.thief == comm->rank) {
            double len = 0;
            if (comm->sendrecv_host(comm->ctx, nullptr, 0, -1, &len, sizeof(double), sp.donor)) return false;
            std::vector<double> buf((size_t)len);
            if (comm->sendrecv_host(comm->ctx, nullptr, 0, -1, buf.data(), buf.size() * sizeof(double), sp.donor))
              return false;
            const double *p = buf.data() + 1;
            for (int k = 0; k < sp.k; k++) {
              auto nd = unpack_node(p);
              ev_line("candidate", nd->oid, fld(" ", nd->upperBound));
              open.push_back(std::move(nd));
            }
            out->st.nodes_received += sp.k;
          }
        }
      }
    } else if (open.empty() || stop) {
      break;
    }
    // ---- the nodes this round pops, in pop order ----
    std::vector<std::unique_ptr<Node>> batch;
    if (pr.node_strategy == MVB_NODE_BEST) {
      const double tp0 = now();
      size_t idx = pick_node(open);
      t_pick += now() - tp0;
      if (spec_on && !open[idx]->spec_tried && !open[idx]->initial) {
        const double ts0 = now();
        if (!speculate(open, idx, bestLower, width)) return false;
        t_spec += now() - ts0;
      }
      const double tp1 = now();
      batch.push_back(std::move(open[idx]));
      open.erase(open.begin() + idx);
      open_ub.erase(open_ub.begin() + idx);
      open_tried.erase(open_tried.begin() + idx);
      t_pick += now() - tp1;
    } else {
      size_t w = std::min<size_t>(width, open.size());
      if (limit > 0) w = std::min<size_t>(w, (size_t)std::max(1, limit - count));
      for (size_t k = 0; k < w; k++) { batch.push_back(std::move(open.front())); open.pop_front(); }
      for (auto &nd : batch)
        if (nd->oid >= fly.min_oid) { if (!finish_fly()) return false; break; }  // (short queue: ramp-up, the end)
    }
    // ---- LP solve #1 of every popped node (bs.cpp:114-117), one launch ----
    {
      std::vector<glp_prob *> ps;
      for (auto &nd : batch) ps.push_back(nd->prob);
      if (!solve_many(ps)) return false;
    }
    std::vector<double> cut_buf;
    std::vector<size_t> cut_off;
    if (pr.cut_mode != MVB_CUT_NONE && !prefetch_cut_rows(batch, bestLower, cut_buf, cut_off)) return false;
    std::vector<Pending> kids;
    size_t consumed = 0;
    const double tcons0 = now();
    for (auto &nd : batch) {
      consumed++;
      Node *node = nd.get();
      glp_prob *a = node->prob;
      TreeRec &rec = tree[node->oid];
      ev_line("pregnant", node->oid, fld(" ", glp_get_obj_val(a), " 1 2"));
      std::vector<int> vars;
      int status = classify(a, vars);
      if (node->initial) {  // bs.cpp:134-149
        if (status == -1) { rec.prune = P_FEAS; stop = true; break; }
        if (status == 1) { node->upperBound = glp_get_obj_val(a); rec.prune = P_INTG; stop = true; break; }
      }
      if (node->initial && pr.root_heuristic && status == 0) {
        std::vector<double> hx;
        double hobj;
        if (floor_point(a, hx, hobj) && hobj > bestLower) {
          bestLower = hobj;
          ev_line("heuristic", node->oid, fld(" ", hobj));
          solution = sstr("[", node->oid, "] Heuristic solution (root LP rounded down) = ", hobj, "\n");
          out->incumbent = hx;
          out->st.has_incumbent = 1;
          out->st.incumbent_oid = node->oid;
          out->st.incumbent = bestLower;
        }
      }
      node->upperBound = glp_get_obj_val(a);
      if (status == 1) {  // bs.cpp:158-193
        rec.prune = P_INTG;
        ev_line("integer", node->oid, fld(" ", node->upperBound));
        if (node->upperBound > bestLower) {
          bestLower = node->upperBound;
          solution = sstr("[", node->oid, "] Solution is: ");
          int n = glp_get_num_cols(a);
          out->incumbent.assign(n + 1, 0.0);
          for (int i = 1; i <= n; i++) {
            out->incumbent[i] = glp_get_col_prim(a, i);
            if (glp_get_col_prim(a, i) != 0 && glp_get_obj_coef(a, i) != 0)
              solution += sstr(glp_get_obj_coef(a, i), "*(x[", i, "] = ", glp_get_col_prim(a, i), ") + ");
          }
          solution += sstr(glp_get_obj_coef(a, 0), " = ", glp_get_obj_val(a), "\n");
          out->st.has_incumbent = 1;
          out->st.incumbent_oid = node->oid;
          out->st.incumbent = bestLower;
        }
      } else if (status == -1) {  // bs.cpp:194-209
        rec.prune = P_FEAS;
        ev_line("infeasible", node->oid, " 1 2");
      } else if (glp_get_obj_val(a) <= bestLower) {  // bs.cpp:210-223
        rec.prune = P_BNDS;
        ev_line("fathomed", node->oid, "");
      } else {  // branch, bs.cpp:224-324
        double acc = 0;
        for (int i : vars) acc += get_fract(glp_get_col_prim(a, i));
        ev_line("branched", node->oid, fld(" ", node->upperBound, " ", acc, " ", vars.size(), " 1 2"));
        if (pr.cut_mode != MVB_CUT_NONE && cuts_wanted(node->oid)) {
          const size_t bk = consumed - 1;
          std::vector<double> one;
          const double *trows = cut_off[bk] != npos ? cut_buf.data() + cut_off[bk] : nullptr;
          if (!trows) {  // not prefetched (cannot happen with the superset rule; kept as a safe path)
            std::vector<int> js, r1;
            cut_rows_of(a, js, r1);
            one.resize(r1.size() * (size_t)a->ncols());
            if (!r1.empty() && mvb_lp_get_tab_rows(a->eng, r1.data(), (int)r1.size(), one.data())) return false;
            trows = one.data();
          }
          const size_t rows0 = (size_t)a->nrows();
          const double tg0 = now();
          gen_cuts(a, trows, cut_off[bk] != npos && bk < pre_cuts.size() ? &pre_cuts[bk] : nullptr);
          t_gen += now() - tg0;
          // the new row goes to a's engine now (queued; all rows of the round are formed by one
          // kernel), so the children clone a tableau that already holds it
          if ((size_t)a->nrows() != rows0 && mv_sync_engine(a)) return false;
        }
        int pick = pick_var(a, vars);
        double bound = glp_get_col_prim(a, pick);
        std::unique_ptr<Node> s2, s3;
        const bool adopted = node->spec && node->spec->pick == pick;
        if (adopted) {
          // the children were built and solved ahead of time (speculate): they enter the
          // statistics now, as the two solves of bs.cpp:279,287
          s2 = make_node(node->spec->down);
          s3 = make_node(node->spec->up);
          node->spec->down = node->spec->up = nullptr;
          out->st.lp_solves += 2;
          out->st.simplex_iters += s2->prob->last.iters + s3->prob->last.iters;
          out->st.spec_adopted++;
        } else {
          // S2/S3 are copies of a (util.cpp:34); a itself is dead afterwards, so
          // S3 takes it over instead of copying a third time
          const double tc0 = now();
          glp_prob *c2 = glp_create_prob();
          glp_copy_prob(c2, a, GLP_ON);
          t_copy += now() - tc0;
          s2 = make_node(c2);
          node->prob = nullptr;
          s3 = make_node(a);
          glp_set_col_bnds(s2->prob, pick, GLP_UP, 0, std::floor(bound));  // bs.cpp:274
          glp_set_col_bnds(s3->prob, pick, GLP_LO, std::ceil(bound), 0);   // bs.cpp:282
        }
        for (Node *s : {s2.get(), s3.get()}) {
          tree[s->oid] = TreeRec{s->oid, node->oid, tree[node->oid].depth + 1, P_NONE, {}};
          tree[node->oid].kids.push_back(s->oid);
        }
        size_t e2 = ev_line("candidate", s2->oid, "");
        size_t e3 = ev_line("candidate", s3->oid, "");
        kids.push_back(Pending{std::move(s2), e2, adopted});
        kids.push_back(Pending{std::move(s3), e3, adopted});
        if (count > 200000) { out->st.hit_limit = 1; stop = true; }  // bs.cpp:320-323 (exit(-1) there)
      }
      count++;
      if (limit > 0 && count >= limit) { out->st.hit_limit = 1; stop = true; }
      if (stop) break;
    }
    t_consume += now() - tcons0;
    // nodes of the batch not consumed because the loop stopped go back unopened
    for (size_t k = batch.size(); k-- > consumed;)
      if (batch[k] && batch[k]->prob) open.push_front(std::move(batch[k]));
    // ---- LP solves #2/#3 of the new children (bs.cpp:279,287), one launch ----
    if (pipe) {
      if (!finish_fly()) return false;  // one launch in flight at a time: the previous round's comes home now
      std::vector<glp_prob *> ps;
      for (auto &k : kids) ps.push_back(k.node->prob);
      if (!solve_many_begin(ps)) return false;
      fly.active = true;
      fly_start();
      for (auto &k : kids) {
        fly.nodes.push_back(k.node.get());
        fly.slots.push_back(k.ev_slot);
        fly.min_oid = std::min(fly.min_oid, k.node->oid);
        open.push_back(std::move(k.node));
      }
    } else {
      std::vector<glp_prob *> ps;
      for (auto &k : kids)
        if (!k.solved) ps.push_back(k.node->prob);
      if (!solve_many(ps)) return false;
      for (auto &k : kids) {
        k.node->upperBound = glp_get_obj_val(k.node->prob);
        if (keep_text) evs[k.ev_slot] += sstr(" ", k.node->upperBound);
        if (pr.node_strategy == MVB_NODE_BEST) {
          open_ub.push_back(k.node->upperBound);
          open_tried.push_back(0);
        }
        open.push_back(std::move(k.node));
      }
    }
    if (sync_mode && !dealt && !stop && open.size() >= (size_t)2 * comm->world) {
      deal_owners(open);
      out->st.split_at = count;
    }
    if (dist) {
      if (!split && !stop && pr.node_strategy != MVB_NODE_BEST && open.size() >= (size_t)4 * comm->world) {
        // deal the FIFO queue round-robin; from here on every rank owns its share
        if (!finish_fly()) return false;  // (the nodes that go to other ranks are dropped here)
        std::deque<std::unique_ptr<Node>> mine;
        for (size_t k = 0; k < open.size(); k++)
          if ((int)(k % comm->world) == comm->rank) mine.push_back(std::move(open[k]));
        open.swap(mine);
        split = true;
        out->st.split_at = count;
        if (pr.total_budget && limit > 0)  // every rank gets an equal share of what is left of the budget
          limit = count + std::max(1, (limit - count + comm->world - 1) / comm->world);
      }
      if (split) {
        if (stop) {
          if (!finish_fly()) return false;
          retired = true; stop = false; open.clear();
        }
        // (the incumbent travels with the open-node counts at the top of the next round)
      }
    }
  }
  if (!finish_fly()) return false;
  if (dist) {
    // final exchange: who holds the optimum (lowest rank on ties), its x, and the node counts
    double best = out->st.has_incumbent ? bestLower : -inf;
    double g = best;
    if (!reduce_max(&g, 1)) return false;
    double who = (out->st.has_incumbent && best == g) ? -(double)comm->rank : -inf;
    if (!reduce_max(&who, 1)) return false;
    if (g > -inf) {
      inc_rank = (int)(-who);
      int n = glp_get_num_cols(root_prob);
      std::vector<double> x(n + 1, -inf);
      if (comm->rank == inc_rank) x = out->incumbent;
      x.resize(n + 1, -inf);
      if (!reduce_max(x.data(), n + 1)) return false;
      out->incumbent = x;
      out->st.has_incumbent = 1;
      out->st.incumbent = g;
      bestLower = g;
    }
    std::vector<double> cnt(comm->world, -inf);
    cnt[comm->rank] = (double)count;
    if (!reduce_max(cnt.data(), comm->world)) return false;
    // the lock-step ramp-up (and a run that never dealt its queue out) is replicated work: its
    // nodes count once, not once per rank
    const int sa = out->st.split_at;
    if (sa < 0) out->st.total_nodes = count;
    else {
      out->st.total_nodes = sa;
      for (double c : cnt) out->st.total_nodes += (long long)c - sa;
    }
  } else {
    out->st.total_nodes = count;
  }
  out->st.incumbent_rank = inc_rank;

  const double tfin0 = now();
  out->tree.clear();
  print_tree(1, out->tree);
  out->solution = solution;
  for (size_t e = 0; e < evs.size(); e++) {
    const std::string &l = evs[e];
    // timed: "timeSpan type oid pid direction fields", the field order of message.h:141-226; events
    // of one device batch carry the time at which the host consumed them
    if (pr.timed_events) out->events += sstr(ev_ms[e], " ");
    out->events += l + "\n";
    // hash type, oid, pid (the leading three fields), not the fp fields
    int sp = 0;
    for (char ch : l) {
      if (ch == ' ' && ++sp == 3) break;
      hash_mix((unsigned char)ch);
    }
  }
  for (const TreeRec &t : tree.v)
    if (t.oid) { hash_mix((unsigned)t.oid); hash_mix((unsigned)t.prune); }
  t_final = now() - tfin0;
  if (getenv("MVOLPS_BNB_TIMING"))
    std::fprintf(stderr, "bnb timing: sync %.1f ms, solve %.1f ms, pull %.1f ms, copy_prob %.1f ms, consume (incl. copy) "
                 "%.1f ms, final text %.1f ms, pick_node %.1f ms, speculate (incl. its solves) %.1f ms, cut rows %.1f ms, cut arithmetic + new row %.1f ms, run %.1f ms\n",
                 1e3 * t_sync, 1e3 * t_solve, 1e3 * t_pull, 1e3 * t_copy, 1e3 * t_consume, 1e3 * t_final,
                 1e3 * t_pick, 1e3 * t_spec, 1e3 * t_rows, 1e3 * t_gen,
                 1e3 * (now() - t_run0));
  out->st.iterations = count;
  out->st.oids = next_id - 1;
  out->st.event_hash = ehash;
  // (the open queue is destroyed by the caller's thread: freeing its 10^5 problem objects on the
  // worker threads measured SLOWER -- 106 k against 124 k nodes/s -- every block goes back to the
  // allocating thread's arena, under its lock)
  return true;
}

}  // namespace

extern "C" {

static mvb_bnb *bnb_run(glp_prob *prob, const mvb_bnb_params *params, const mvb_comm *comm) {
  mvb_bnb_params p{};
  if (params) p = *params;
  auto *h = new mvb_bnb();
  Driver d(p, prob, h);
  d.comm = comm;
  if (comm && comm->world > 1 && !comm->allreduce_max) {
    std::fprintf(stderr, "mvb_bnb_run_dist: communicator has no allreduce_max\n");
    delete h;
    return nullptr;
  }
  if (comm && comm->world > 1 && p.dist_mode == MVB_DIST_SYNC && !comm->allgather_host) {
    std::fprintf(stderr, "mvb_bnb_run_dist: MVB_DIST_SYNC needs a communicator with allgather_host\n");
    delete h;
    return nullptr;
  }
  if (!d.run()) {
    std::fprintf(stderr, "mvb_bnb_run: engine error: %s\n", mvb_last_error());
    delete h;
    return nullptr;
  }
  if (!p.quiet) {  // bs.cpp:329-343
    std::printf("[I = Integral node, F = Infeasible node, B = Worse bound node]\n%s\n%s\n",
                h->tree.c_str(), h->solution.c_str());
    std::printf("Solution found after %d iterations\n", h->st.iterations);
  }
  return h;
}

mvb_bnb *mvb_bnb_run(glp_prob *prob, const mvb_bnb_params *params) { return bnb_run(prob, params, nullptr); }
mvb_bnb *mvb_bnb_run_dist(glp_prob *prob, const mvb_bnb_params *params, const mvb_comm *comm) {
  return bnb_run(prob, params, comm);
}

int mvb_bnb_steal_plan(const double *cnt, int world, int *donor, int *thief, int *nodes, int cap) {
  const auto pairs = Driver::steal_matching(std::vector<double>(cnt, cnt + world));
  int n = 0;
  for (const auto &p : pairs) {
    if (n >= cap) break;
    donor[n] = p.donor; thief[n] = p.thief; nodes[n] = p.k;
    n++;
  }
  return n;
}

void mvb_bnb_free(mvb_bnb *h) { delete h; }
const mvb_bnb_stats *mvb_bnb_get_stats(const mvb_bnb *h) { return &h->st; }
const char *mvb_bnb_tree_text(const mvb_bnb *h) { return h->tree.c_str(); }
const char *mvb_bnb_solution_text(const mvb_bnb *h) { return h->solution.c_str(); }
const char *mvb_bnb_event_text(const mvb_bnb *h) { return h->events.c_str(); }
int mvb_bnb_get_incumbent(const mvb_bnb *h, double *x, int cap) {
  if (!h->st.has_incumbent) return 0;
  int n = (int)h->incumbent.size() - 1;
  for (int i = 0; i <= n && i < cap; i++) x[i] = h->incumbent[i];
  return n;
}

}  // extern "C"
