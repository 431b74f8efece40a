/* simplex_oracle.c -- CPU ORACLE of the LP-relaxation engine.  TEST
 * INFRASTRUCTURE ONLY: it may be imported, linked or executed only by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs.  The product (mvolps_b200/) never links or loads it.
 *
 * PARITY UNPINNED at the reference level: the reference solves every node LP
 * with GLPK's glp_simplex (bs.cpp:117, bs.cpp:279, bs.cpp:287) and reads
 * tableau rows with glp_eval_tab_row (gmi.cpp:36).  GLPK is linked with
 * -lglpk (Makefile:2), is not vendored, not a submodule (.gitmodules lists
 * only spdlog) and not version-pinned (only glp_version() is logged,
 * util.cpp:278); the reference holds no tests, fixtures or golden vectors.
 * GLPK's pivot sequence is therefore not observable, and this file restates
 * instead the published textbook algorithm the north_star names -- a dense
 * bounded-variable tableau simplex (Dantzig pricing, lowest-index ties,
 * textbook ratio tests; Chvatal, "Linear Programming", ch. 8; Bland 1977 for
 * the anti-cycling rule) -- in plain sequential C, as the executable
 * specification the CUDA kernels must match bit for bit (pivot sequence,
 * tableau, objective).  Solver-independent pinning of the *results* (LP and
 * ILP objectives, integer optimum) is done against SciPy/HiGHS in tests/.
 *
 * Every fp64 step is written with explicit fma()/IEEE divide; build with
 * -ffp-contract=off so the compiler adds no contraction of its own.
 */
#include "../include/mvolps_engine.h"
#include "../include/mvolps_spec.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

struct mvb_lp {
  int m, n, ld, mcap;
  double *T;    /* (mcap+1) x ld, row 0 = reduced costs */
  double *beta; /* mcap+1, 1-based */
  double *lb, *ub; /* n+mcap by variable id */
  double *cost;    /* n */
  int *head;       /* mcap+1, 1-based: id basic in row i */
  int *nb, *at;    /* n */
  double *work;    /* scratch: ld + mcap+1 + n*chunks */
  int *pivlog;
  int pivlog_cap, pivlog_n;
  mvb_result res;
};

static __thread char g_err[256];
static void set_err(const char *s) { snprintf(g_err, sizeof g_err, "%s", s); }

const char *mvb_backend(void) { return "oracle-c"; }
const char *mvb_last_error(void) { return g_err; }
int mvb_device_init(int device) { (void)device; return 0; }
unsigned long long mvb_kernel_launches(void) { return 0ULL; }

#include <time.h>
static double g_slots[16];
int mvb_timer_record(int slot) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  g_slots[slot & 15] = ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
  return 0;
}
double mvb_timer_elapsed_ms(int a, int b) { return g_slots[b & 15] - g_slots[a & 15]; }
int mvb_flush_l2(void) { return 0; }
double mvb_last_kernel_ms(void) { return 0.0; }
int mvb_last_kernel_mode(void) { return -1; }
int mvb_engine_option(const char *name, int value) { (void)name; (void)value; return -1; }
int mvb_last_launch(int *c, int *g, int *t) { if (c) *c = 0; if (g) *g = 0; if (t) *t = 0; return 0; }

static int round_up(int x, int a) { return (x + a - 1) / a * a; }

static void lp_release(mvb_lp *lp) {
  free(lp->T); free(lp->beta); free(lp->lb); free(lp->ub); free(lp->cost);
  free(lp->head); free(lp->nb); free(lp->at); free(lp->work); free(lp->pivlog);
}

static int lp_alloc(mvb_lp *lp, int m, int n, int mcap) {
  if (mcap < m) mcap = m;
  memset(lp, 0, sizeof *lp);
  lp->m = m; lp->n = n; lp->mcap = mcap;
  lp->ld = round_up(n > 0 ? n : 1, MVB_LD_ALIGN);
  size_t tsz = (size_t)(mcap + 1) * lp->ld;
  lp->T = (double *)calloc(tsz, sizeof(double));
  lp->beta = (double *)calloc(mcap + 1, sizeof(double));
  lp->lb = (double *)calloc(n + mcap, sizeof(double));
  lp->ub = (double *)calloc(n + mcap, sizeof(double));
  lp->cost = (double *)calloc(n > 0 ? n : 1, sizeof(double));
  lp->head = (int *)calloc(mcap + 1, sizeof(int));
  lp->nb = (int *)calloc(n > 0 ? n : 1, sizeof(int));
  lp->at = (int *)calloc(n > 0 ? n : 1, sizeof(int));
  lp->work = (double *)calloc((size_t)lp->ld * 2 + mcap + 2, sizeof(double));
  if (!lp->T || !lp->beta || !lp->lb || !lp->ub || !lp->cost || !lp->head ||
      !lp->nb || !lp->at || !lp->work) {
    set_err("out of memory");
    return -1;
  }
  lp->res.status = MVB_UNSOLVED;
  return 0;
}

mvb_lp *mvb_lp_create(int m, int n, int row_cap) {
  if (m < 0 || n < 0) { set_err("bad dims"); return NULL; }
  mvb_lp *lp = (mvb_lp *)malloc(sizeof *lp);
  if (!lp) return NULL;
  if (lp_alloc(lp, m, n, row_cap)) { lp_release(lp); free(lp); return NULL; }
  return lp;
}

void mvb_lp_free(mvb_lp *lp) {
  if (!lp) return;
  lp_release(lp);
  free(lp);
}

void mvb_lp_free_many(mvb_lp *const *lps, int count) {
  for (int k = 0; k < count; k++) mvb_lp_free(lps[k]);
}

void *mvb_host_staging(size_t bytes) {
  static __thread void *buf;
  static __thread size_t cap;
  if (cap < bytes) {
    free(buf);
    buf = malloc(bytes);
    cap = buf ? bytes : 0;
  }
  return buf;
}

int mvb_lp_dims(const mvb_lp *lp, int *m, int *n) {
  if (m) *m = lp->m;
  if (n) *n = lp->n;
  return 0;
}

/* flag a non-basic variable takes for bounds [l,u], preferring `keep` */
static int nb_flag_for(double l, double u, int keep) {
  int hasl = l > -HUGE_VAL, hasu = u < HUGE_VAL;
  if (hasl && hasu && l == u) return MVB_AT_FX;
  if (keep == MVB_AT_UB && hasu) return MVB_AT_UB;
  if (keep == MVB_AT_LB && hasl) return MVB_AT_LB;
  if (hasl) return MVB_AT_LB;
  if (hasu) return MVB_AT_UB;
  return MVB_AT_FR;
}

int mvb_lp_load(mvb_lp *lp, const double *A, const double *cost,
                const double *lb, const double *ub) {
  int m = lp->m, n = lp->n, ld = lp->ld;
  memset(lp->T, 0, (size_t)(lp->mcap + 1) * ld * sizeof(double));
  for (int j = 0; j < n; j++) lp->T[j] = cost[j];
  for (int i = 0; i < m; i++)
    memcpy(lp->T + (size_t)(i + 1) * ld, A + (size_t)i * n, n * sizeof(double));
  memcpy(lp->cost, cost, n * sizeof(double));
  memcpy(lp->lb, lb, (n + m) * sizeof(double));
  memcpy(lp->ub, ub, (n + m) * sizeof(double));
  for (int i = 1; i <= m; i++) lp->head[i] = n + i - 1;
  for (int j = 0; j < n; j++) {
    lp->nb[j] = j;
    lp->at[j] = nb_flag_for(lb[j], ub[j], MVB_AT_LB);
  }
  lp->res.status = MVB_UNSOLVED;
  return 0;
}

int mvb_lp_assign(mvb_lp *dst, const mvb_lp *src) {
  if (dst->n != src->n || dst->mcap < src->m) { set_err("assign: shape"); return -1; }
  int m = src->m, n = src->n, ld = src->ld;
  dst->m = m;
  memcpy(dst->T, src->T, (size_t)(m + 1) * ld * sizeof(double));
  memcpy(dst->beta, src->beta, (m + 1) * sizeof(double));
  memcpy(dst->lb, src->lb, (n + m) * sizeof(double));
  memcpy(dst->ub, src->ub, (n + m) * sizeof(double));
  memcpy(dst->cost, src->cost, n * sizeof(double));
  memcpy(dst->head, src->head, (m + 1) * sizeof(int));
  memcpy(dst->nb, src->nb, n * sizeof(int));
  memcpy(dst->at, src->at, n * sizeof(int));
  dst->res = src->res;
  dst->pivlog_n = 0;
  return 0;
}

mvb_lp *mvb_lp_clone(const mvb_lp *src) {
  mvb_lp *lp = mvb_lp_create(src->m, src->n, src->mcap);
  if (!lp) return NULL;
  mvb_lp_assign(lp, src);
  return lp;
}

static int position_of(const mvb_lp *lp, int k) {
  for (int j = 0; j < lp->n; j++)
    if (lp->nb[j] == k) return j;
  return -1;
}

int mvb_lp_set_bounds(mvb_lp *lp, int k, double lb, double ub) {
  if (k < 0 || k >= lp->n + lp->m) { set_err("set_bounds: id"); return -1; }
  lp->lb[k] = lb;
  lp->ub[k] = ub;
  int j = position_of(lp, k);
  if (j >= 0) lp->at[j] = nb_flag_for(lb, ub, lp->at[j]);
  return 0;
}

static int lp_grow(mvb_lp *lp, int mcap) {
  mvb_lp nl;
  if (lp_alloc(&nl, lp->m, lp->n, mcap)) return -1;
  mvb_lp_assign(&nl, lp);
  nl.res = lp->res;
  lp_release(lp);
  *lp = nl;
  return 0;
}

/* new row over structurals a[0..n): tableau row = a_N + sum_i a_{head[i]} T[i,:]
 * accumulated in row order with one fma per term */
int mvb_lp_append_row(mvb_lp *lp, const double *a, double lb, double ub) {
  if (lp->m + 1 > lp->mcap && lp_grow(lp, lp->mcap + 16 + lp->mcap / 4)) return -1;
  int m = lp->m, n = lp->n, ld = lp->ld;
  double *row = lp->T + (size_t)(m + 1) * ld;
  for (int p = 0; p < ld; p++) row[p] = 0.0;
  for (int p = 0; p < n; p++) {
    int k = lp->nb[p];
    double acc = k < n ? a[k] : 0.0;
    for (int i = 1; i <= m; i++) {
      int h = lp->head[i];
      if (h < n && a[h] != 0.0) acc = fma(a[h], lp->T[(size_t)i * ld + p], acc);
    }
    row[p] = acc;
  }
  lp->head[m + 1] = n + m;
  lp->lb[n + m] = lb;
  lp->ub[n + m] = ub;
  lp->beta[m + 1] = 0.0;
  lp->m = m + 1;
  return 0;
}

/* ------------------------------------------------------------------ */
/* specification pieces                                                 */

static double nb_value(const mvb_lp *lp, int j) {
  int f = lp->at[j], k = lp->nb[j];
  if (f == MVB_AT_LB || f == MVB_AT_FX) return lp->lb[k];
  if (f == MVB_AT_UB) return lp->ub[k];
  return 0.0;
}

/* deterministic dot product: 32 lane-strided fma chains, halving tree */
static double det_dot(const double *a, const double *v, int n) {
  double p[MVB_DOT_LANES];
  for (int l = 0; l < MVB_DOT_LANES; l++) p[l] = 0.0;
  for (int j = 0; j < n; j++) p[j % MVB_DOT_LANES] = fma(a[j], v[j], p[j % MVB_DOT_LANES]);
  for (int off = MVB_DOT_LANES / 2; off >= 1; off /= 2)
    for (int l = 0; l < off; l++) p[l] = p[l] + p[l + off];
  return p[0];
}

static void compute_beta(mvb_lp *lp) {
  double *v = lp->work; /* n values */
  for (int j = 0; j < lp->n; j++) v[j] = nb_value(lp, j);
  for (int i = 1; i <= lp->m; i++)
    lp->beta[i] = det_dot(lp->T + (size_t)i * lp->ld, v, lp->n);
}

/* -1: below lb, +1: above ub, 0: within tolerance */
static int row_infeas(const mvb_lp *lp, int i) {
  int k = lp->head[i];
  double b = lp->beta[i], l = lp->lb[k], u = lp->ub[k];
  if (l - b > MVB_TOL_BND * (1.0 + fabs(l))) return -1;
  if (b - u > MVB_TOL_BND * (1.0 + fabs(u))) return 1;
  return 0;
}

static int primal_feasible(const mvb_lp *lp) {
  for (int i = 1; i <= lp->m; i++)
    if (row_infeas(lp, i)) return 0;
  return 1;
}

static int dj_eligible(int f, double dj) {
  if (f == MVB_AT_LB) return dj < -MVB_TOL_DJ;
  if (f == MVB_AT_UB) return dj > MVB_TOL_DJ;
  if (f == MVB_AT_FR) return dj < -MVB_TOL_DJ || dj > MVB_TOL_DJ;
  return 0;
}

static int dual_feasible(const mvb_lp *lp) {
  for (int j = 0; j < lp->n; j++)
    if (dj_eligible(lp->at[j], lp->T[j])) return 0;
  return 1;
}

/* entering column: largest |d_j| among eligible, lowest variable id on ties;
 * under Bland's rule the lowest eligible variable id */
static int price_primal(const mvb_lp *lp, const double *dr, int bland) {
  int best = -1, bestvar = 0;
  double besta = 0.0;
  for (int j = 0; j < lp->n; j++) {
    double dj = dr[j];
    if (!dj_eligible(lp->at[j], dj)) continue;
    double a = fabs(dj);
    int var = lp->nb[j];
    int better;
    if (best < 0) better = 1;
    else if (bland) better = var < bestvar;
    else better = a > besta || (a == besta && var < bestvar);
    if (better) { best = j; besta = a; bestvar = var; }
  }
  return best;
}

/* phase-1 reduced costs d1_j = sum_i w_i T_ij, w_i = infeasibility sign,
 * as chunked sequential sums (MVB_P1_CHUNK rows per partial) */
static int phase1_costs(mvb_lp *lp, double *d1, int *sgn) {
  int m = lp->m, n = lp->n, ld = lp->ld, ninf = 0;
  for (int i = 1; i <= m; i++) { sgn[i] = row_infeas(lp, i); ninf += sgn[i] != 0; }
  for (int j = 0; j < n; j++) d1[j] = 0.0;
  for (int c0 = 1; c0 <= m; c0 += MVB_P1_CHUNK) {
    int c1 = c0 + MVB_P1_CHUNK - 1 < m ? c0 + MVB_P1_CHUNK - 1 : m;
    for (int j = 0; j < n; j++) {
      double acc = 0.0;
      for (int i = c0; i <= c1; i++) {
        if (sgn[i] > 0) acc = acc + lp->T[(size_t)i * ld + j];
        else if (sgn[i] < 0) acc = acc - lp->T[(size_t)i * ld + j];
      }
      d1[j] = d1[j] + acc;
    }
  }
  return ninf;
}

typedef struct {
  int r;        /* leaving row, 0 = none */
  double t;     /* step length */
  int leave_at; /* flag the leaving variable takes */
} ratio_t;

/* primal ratio test.  sgn == NULL: phase 2.  sgn != NULL: phase 1, where an
 * infeasible basic variable blocks only when it reaches its violated bound */
static ratio_t ratio_primal(const mvb_lp *lp, int q, int dir, const int *sgn) {
  ratio_t out = {0, HUGE_VAL, MVB_AT_LB};
  int rvar = 0, ld = lp->ld;
  for (int i = 1; i <= lp->m; i++) {
    double alpha = lp->T[(size_t)i * ld + q];
    double a = dir > 0 ? alpha : -alpha;
    int var = lp->head[i];
    int s = sgn ? sgn[i] : 0;
    double t;
    int fl;
    if (a < -MVB_TOL_PIV) { /* basic variable decreases */
      if (s < 0) continue;
      double bound = s > 0 ? lp->ub[var] : lp->lb[var];
      if (!(bound > -HUGE_VAL)) continue;
      t = (lp->beta[i] - bound) / (-a);
      fl = s > 0 ? MVB_AT_UB : MVB_AT_LB;
    } else if (a > MVB_TOL_PIV) { /* basic variable increases */
      if (s > 0) continue;
      double bound = s < 0 ? lp->lb[var] : lp->ub[var];
      if (!(bound < HUGE_VAL)) continue;
      t = (bound - lp->beta[i]) / a;
      fl = s < 0 ? MVB_AT_LB : MVB_AT_UB;
    } else continue;
    if (!(t > 0.0)) t = 0.0;
    if (out.r == 0 || t < out.t || (t == out.t && var < rvar)) {
      out.r = i; out.t = t; out.leave_at = fl; rvar = var;
    }
  }
  return out;
}

/* rank-1 tableau update around (r,q) plus basis bookkeeping.
 * rho_q = 1/p, rho_j = -(T_rj * (1/p)); T_ij <- fma(T_iq, rho_j, T_ij) with
 * the old T_iq read as 0 in column q; row r <- rho */
static void pivot(mvb_lp *lp, int r, int q, int leave_at) {
  int m = lp->m, ld = lp->ld;
  double *T = lp->T;
  double *rho = lp->work + lp->ld;         /* ld */
  double *col = lp->work + 2 * (size_t)lp->ld; /* m+1 */
  double inv = 1.0 / T[(size_t)r * ld + q];
  for (int i = 0; i <= m; i++) col[i] = T[(size_t)i * ld + q];
  for (int j = 0; j < ld; j++) rho[j] = -(T[(size_t)r * ld + j] * inv);
  rho[q] = inv;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) if ((size_t)m * ld > 65536)
#endif
  for (int i = 0; i <= m; i++) {
    double *row = T + (size_t)i * ld;
    if (i == r) {
      for (int j = 0; j < ld; j++) row[j] = rho[j];
    } else {
      double ci = col[i];
      row[q] = 0.0;
      for (int j = 0; j < ld; j++) row[j] = fma(ci, rho[j], row[j]);
    }
  }
  int kin = lp->nb[q], kout = lp->head[r];
  lp->head[r] = kin;
  lp->nb[q] = kout;
  if (lp->lb[kout] == lp->ub[kout]) leave_at = MVB_AT_FX;
  lp->at[q] = leave_at;
}

static void log_iter(mvb_lp *lp, const mvb_opts *o, mvb_result *res, int mode,
                     int r, int q) {
  unsigned long long h = res->pivot_hash;
  h = (h ^ (unsigned long long)mode) * MVB_FNV_PRIME;
  h = (h ^ (unsigned long long)r) * MVB_FNV_PRIME;
  h = (h ^ (unsigned long long)q) * MVB_FNV_PRIME;
  res->pivot_hash = h;
  if (o && o->log_pivots) {
    if (lp->pivlog_n >= lp->pivlog_cap) {
      int nc = lp->pivlog_cap ? lp->pivlog_cap * 2 : 1024;
      lp->pivlog = (int *)realloc(lp->pivlog, (size_t)nc * 2 * sizeof(int));
      lp->pivlog_cap = nc;
    }
    lp->pivlog[2 * lp->pivlog_n] = r;
    lp->pivlog[2 * lp->pivlog_n + 1] = q;
    lp->pivlog_n++;
  }
  res->iters++;
  if (r) res->pivots++; else res->flips++;
  if (mode == MVB_MODE_PRIMAL1) res->iters_primal1++;
  else if (mode == MVB_MODE_PRIMAL2) res->iters_primal2++;
  else res->iters_dual++;
}

/* move entering variable q by delta: beta_i <- fma(T_iq, delta, beta_i) */
static void shift_beta(mvb_lp *lp, int q, double delta) {
  for (int i = 1; i <= lp->m; i++)
    lp->beta[i] = fma(lp->T[(size_t)i * lp->ld + q], delta, lp->beta[i]);
}

static void finish(mvb_lp *lp, mvb_result *res) {
  int n = lp->n;
  double *x = lp->work; /* n */
  for (int j = 0; j < n; j++) x[j] = 0.0;
  for (int j = 0; j < n; j++)
    if (lp->nb[j] < n) x[lp->nb[j]] = nb_value(lp, j);
  for (int i = 1; i <= lp->m; i++)
    if (lp->head[i] < n) x[lp->head[i]] = lp->beta[i];
  res->obj = det_dot(lp->cost, x, n);
  res->primal_feasible = primal_feasible(lp);
}

static void solve_one(mvb_lp *lp, const mvb_opts *o, mvb_result *res) {
  int m = lp->m, n = lp->n;
  memset(res, 0, sizeof *res);
  res->pivot_hash = MVB_FNV_OFFSET;
  res->status = MVB_ITLIM;
  lp->pivlog_n = 0;
  int max_iters = (o && o->max_iters > 0) ? o->max_iters : 50 * (m + n) + 10000;
  double *d1 = (double *)malloc((size_t)(n > 0 ? n : 1) * sizeof(double));
  int *sgn = (int *)malloc((size_t)(m + 2) * sizeof(int));

  compute_beta(lp);
  int mode;
  if (primal_feasible(lp)) mode = MVB_MODE_PRIMAL2;
  else if (dual_feasible(lp)) mode = MVB_MODE_DUAL;
  else mode = MVB_MODE_PRIMAL1;
  int degen = 0;

  while (1) {
    int bland = degen >= MVB_BLAND_AFTER;
    if (mode == MVB_MODE_DUAL) {
      /* leaving row: largest bound violation, lowest variable id on ties */
      int r = 0, rvar = 0, rs = 0;
      double rviol = 0.0;
      for (int i = 1; i <= m; i++) {
        int s = row_infeas(lp, i);
        if (!s) continue;
        int var = lp->head[i];
        double viol = s < 0 ? lp->lb[var] - lp->beta[i] : lp->beta[i] - lp->ub[var];
        int better;
        if (r == 0) better = 1;
        else if (bland) better = var < rvar;
        else better = viol > rviol || (viol == rviol && var < rvar);
        if (better) { r = i; rvar = var; rviol = viol; rs = s; }
      }
      if (r == 0) { res->status = MVB_OPT; break; }
      if (res->iters >= max_iters) break;
      /* entering column: min |d_j| / |T_rj| over columns that can move x_r
       * toward its violated bound, lowest variable id on ties */
      const double *row = lp->T + (size_t)r * lp->ld;
      int q = -1, qvar = 0;
      double qratio = 0.0;
      for (int j = 0; j < n; j++) {
        int f = lp->at[j];
        if (f == MVB_AT_FX) continue;
        double alpha = row[j];
        double a = rs < 0 ? alpha : -alpha;
        int ok = (f == MVB_AT_LB && a > MVB_TOL_PIV) ||
                 (f == MVB_AT_UB && a < -MVB_TOL_PIV) ||
                 (f == MVB_AT_FR && (a > MVB_TOL_PIV || a < -MVB_TOL_PIV));
        if (!ok) continue;
        double ratio = fabs(lp->T[j]) / fabs(alpha);
        int var = lp->nb[j];
        if (q < 0 || ratio < qratio || (ratio == qratio && var < qvar)) {
          q = j; qratio = ratio; qvar = var;
        }
      }
      if (q < 0) { res->status = MVB_NOFEAS; break; }
      double target = rs < 0 ? lp->lb[rvar] : lp->ub[rvar];
      double delta = (target - lp->beta[r]) / row[q];
      double vq = nb_value(lp, q);
      shift_beta(lp, q, delta);
      lp->beta[r] = vq + delta;
      if (bland) res->bland_iters++;
      log_iter(lp, o, res, mode, r, q);
      pivot(lp, r, q, rs < 0 ? MVB_AT_LB : MVB_AT_UB);
      degen = (qratio == 0.0) ? degen + 1 : 0;
      continue;
    }

    /* primal modes */
    const double *dr;
    const int *sg = NULL;
    if (mode == MVB_MODE_PRIMAL1) {
      int ninf = phase1_costs(lp, d1, sgn);
      if (ninf == 0) { mode = MVB_MODE_PRIMAL2; degen = 0; continue; }
      dr = d1;
      sg = sgn;
    } else {
      dr = lp->T;
    }
    int q = price_primal(lp, dr, bland);
    if (q < 0) {
      res->status = mode == MVB_MODE_PRIMAL1 ? MVB_NOFEAS : MVB_OPT;
      break;
    }
    if (res->iters >= max_iters) break;
    int dir = dr[q] < 0.0 ? 1 : -1;
    ratio_t rt = ratio_primal(lp, q, dir, sg);
    int kq = lp->nb[q];
    double tflip = lp->ub[kq] - lp->lb[kq];
    if (bland) res->bland_iters++;
    if (tflip <= rt.t) {
      if (!(tflip < HUGE_VAL)) {
        res->status = mode == MVB_MODE_PRIMAL1 ? MVB_NOFEAS : MVB_UNBND;
        break;
      }
      /* bound flip: no basis change */
      shift_beta(lp, q, dir > 0 ? tflip : -tflip);
      lp->at[q] = lp->at[q] == MVB_AT_LB ? MVB_AT_UB : MVB_AT_LB;
      log_iter(lp, o, res, mode, 0, q);
      degen = 0;
      continue;
    }
    double delta = dir > 0 ? rt.t : -rt.t;
    double vq = nb_value(lp, q);
    shift_beta(lp, q, delta);
    lp->beta[rt.r] = vq + delta;
    log_iter(lp, o, res, mode, rt.r, q);
    pivot(lp, rt.r, q, rt.leave_at);
    degen = (rt.t == 0.0) ? degen + 1 : 0;
  }
  finish(lp, res);
  lp->res = *res;
  free(d1);
  free(sgn);
}

int mvb_solve_batch(mvb_lp *const *lps, int count, const mvb_opts *opts,
                    mvb_result *out) {
#ifdef _OPENMP
  int nt = (opts && opts->threads > 0) ? opts->threads : omp_get_max_threads();
  omp_set_num_threads(nt);
  if (count > 1) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int k = 0; k < count; k++) solve_one(lps[k], opts, &out[k]);
    return 0;
  }
#endif
  for (int k = 0; k < count; k++) solve_one(lps[k], opts, &out[k]);
  return 0;
}

int mvb_lp_get_primal(mvb_lp *lp, double *x) {
  int n = lp->n, m = lp->m;
  for (int k = 0; k < n + m; k++) x[k] = 0.0;
  for (int j = 0; j < n; j++) x[lp->nb[j]] = nb_value(lp, j);
  for (int i = 1; i <= m; i++) x[lp->head[i]] = lp->beta[i];
  return 0;
}

int mvb_lp_get_basis(mvb_lp *lp, int *head, int *nb, int *at) {
  for (int i = 0; i < lp->m; i++) head[i] = lp->head[i + 1];
  for (int j = 0; j < lp->n; j++) { nb[j] = lp->nb[j]; at[j] = lp->at[j]; }
  return 0;
}

/* pointers valid until the next call on this thread (the shim consumes them at once) */
int mvb_lp_solution_view(mvb_lp *lp, const double **x, const int **head, const int **nb, const int **at) {
  static __thread double *vx;
  static __thread int *vh;
  static __thread int cap;
  if (cap < lp->n + lp->m + 1) {
    cap = 2 * (lp->n + lp->m) + 64;
    vx = (double *)realloc(vx, (size_t)cap * sizeof(double));
    vh = (int *)realloc(vh, (size_t)cap * sizeof(int));
  }
  mvb_lp_get_primal(lp, vx);
  for (int i = 0; i < lp->m; i++) vh[i] = lp->head[i + 1];
  *x = vx; *head = vh; *nb = lp->nb; *at = lp->at;
  return 0;
}

int mvb_lp_get_tab_rows(mvb_lp *lp, const int *rows, int nrows, double *out) {
  for (int t = 0; t < nrows; t++) {
    if (rows[t] < 0 || rows[t] > lp->m) { set_err("get_tab_rows: row"); return -1; }
    memcpy(out + (size_t)t * lp->n, lp->T + (size_t)rows[t] * lp->ld,
           lp->n * sizeof(double));
  }
  return 0;
}

/* the two-halves form of mvb_solve_batch: the CPU oracle solves inside begin */
static mvb_result *g_pending = NULL;
static int g_pending_n = -1;
int mvb_solve_batch_begin(mvb_lp *const *lps, int count, const mvb_opts *opts) {
  if (g_pending_n >= 0) { set_err("mvb_solve_batch_begin: a batch is already in flight"); return -1; }
  free(g_pending);
  g_pending = (mvb_result *)malloc(sizeof(mvb_result) * (size_t)(count > 0 ? count : 1));
  g_pending_n = count > 0 ? count : 0;
  return count > 0 ? mvb_solve_batch(lps, count, opts, g_pending) : 0;
}
int mvb_solve_batch_end(mvb_result *out) {
  if (g_pending_n < 0) { set_err("mvb_solve_batch_end: no batch in flight"); return -1; }
  if (g_pending_n > 0) memcpy(out, g_pending, sizeof(mvb_result) * (size_t)g_pending_n);
  g_pending_n = -1;
  return 0;
}

/* CUDA-only entry point (static device scratch); the CPU oracle has none: always fails */
int mvb_lp_get_tab_rows_noalloc(mvb_lp *lp, const int *rows, int nrows, double *out) {
  (void)lp; (void)rows; (void)nrows; (void)out;
  set_err("mvb_lp_get_tab_rows_noalloc: CUDA engine only");
  return -1;
}

int mvb_lp_get_tab_rows_batch(mvb_lp *const *lps, int count, const int *rows, const int *nrows,
                              double *out) {
  for (int k = 0; k < count; k++) {
    if (mvb_lp_get_tab_rows(lps[k], rows, nrows[k], out)) return -1;
    rows += nrows[k];
    out += (size_t)nrows[k] * lps[k]->n;
  }
  return 0;
}

int mvb_lp_get_tableau(mvb_lp *lp, double *T, double *beta) {
  for (int i = 0; i <= lp->m; i++)
    memcpy(T + (size_t)i * lp->n, lp->T + (size_t)i * lp->ld, lp->n * sizeof(double));
  if (beta) memcpy(beta, lp->beta, (lp->m + 1) * sizeof(double));
  return 0;
}

int mvb_lp_get_pivot_log(mvb_lp *lp, int *rowcol, int cap) {
  int c = lp->pivlog_n < cap ? lp->pivlog_n : cap;
  if (c > 0) memcpy(rowcol, lp->pivlog, (size_t)c * 2 * sizeof(int));
  return lp->pivlog_n;
}
