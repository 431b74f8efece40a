/* mvolps_engine.h -- thin C-ABI between the C++ host (glpk-subset shim,
 * B&B driver) and the LP-relaxation engine.
 *
 * Two libraries export exactly these symbols:
 *   product : mvolps_b200/lib/libglpk_b200.so      (CUDA sm_100a kernels)
 *   oracle  : oracle/_build/libglpk_oracle.so       (plain C, test-only)
 * The product never links or loads the oracle.
 *
 * The engine replaces what the reference reaches through GLPK:
 *   glp_simplex        bs.cpp:117, bs.cpp:279, bs.cpp:287
 *   glp_eval_tab_row   gmi.cpp:36
 *   glp_copy_prob      util.cpp:34, bs.cpp:116   (tableau clone for warm start)
 *   glp_set_col_bnds   bs.cpp:274, bs.cpp:282    (branching bound change)
 *   glp_add_rows/glp_set_mat_row/glp_set_row_bnds  cut.cpp:23,40-43 (cut row)
 *
 * Data model.  An LP has n structural variables
 * (internal id 0..n-1) and m auxiliary row variables (id n+i for 0-based row
 * i), every variable with bounds [lb, ub] (+-HUGE_VAL when absent).  The
 * dense tableau has m+1 rows and n columns, fp64 row-major:
 *   row 0      : reduced costs d_j of the internally MINIMISED objective
 *   row i>=1   : x_head[i] = sum_j T[i][j] * x_nb[j]
 * beta[i] holds the value of the basic variable of row i.
 */
#ifndef MVOLPS_ENGINE_H
#define MVOLPS_ENGINE_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mvb_lp mvb_lp;

/* termination status of a solve */
#define MVB_OPT 0     /* optimal basis found */
#define MVB_NOFEAS 1  /* primal infeasible */
#define MVB_UNBND 2   /* primal unbounded */
#define MVB_ITLIM 3   /* iteration limit reached */
#define MVB_UNSOLVED 4

/* non-basic position flags */
#define MVB_AT_LB 0
#define MVB_AT_UB 1
#define MVB_AT_FR 2 /* free variable, value 0 */
#define MVB_AT_FX 3 /* fixed variable, lb == ub */

/* simplex mode tags that enter the pivot hash / log */
#define MVB_MODE_PRIMAL2 1
#define MVB_MODE_DUAL 2
#define MVB_MODE_PRIMAL1 3

typedef struct mvb_result {
  int status;               /* MVB_* */
  int iters;                /* iterations = pivots + bound flips */
  int pivots;               /* basis changes (rank-1 tableau updates) */
  int flips;                /* bound flips without basis change */
  int iters_primal1;        /* iterations spent in each mode */
  int iters_primal2;
  int iters_dual;
  int bland_iters;          /* iterations taken under the anti-cycling rule */
  unsigned long long pivot_hash; /* FNV-1a over (mode,row,col) per iteration */
  double obj;               /* internal-min objective: sum_j cost_j x_j */
  int primal_feasible;      /* final point within bound tolerance */
  int reserved;
} mvb_result;

typedef struct mvb_opts {
  int max_iters;   /* <=0 : 50*(m+n)+10000 */
  int log_pivots;  /* !=0 : keep (row,col) of every iteration (see get_pivot_log) */
  int ctas_per_lp; /* CUDA: CTAs cooperating on one LP; <=0 auto. ignored by oracle */
  int threads;     /* oracle: OpenMP threads; <=0 all. ignored by CUDA */
} mvb_opts;

/* "cuda-sm100a" or "oracle-c" */
const char *mvb_backend(void);
/* last error text of the calling thread's most recent failing call */
const char *mvb_last_error(void);
/* CUDA: select device, create stream and memory pool. oracle: no-op. 0 = ok */
int mvb_device_init(int device);
/* CUDA: number of kernels this library launched since load (bench evidence) */
unsigned long long mvb_kernel_launches(void);

/* measurement hooks (bench.py): CUDA events on the engine's launch stream
 * (oracle: host clock).  slot in [0,16) */
int mvb_timer_record(int slot);
/* waits for slot b; milliseconds between the two records */
double mvb_timer_elapsed_ms(int a, int b);
/* CUDA: overwrite a buffer larger than L2 so the next step starts cold */
int mvb_flush_l2(void);
/* device time of the solve kernels of the last mvb_solve_batch alone (CUDA events around each
 * launch, summed: an LP that needs phase 1 takes a second launch for its phase 2) */
double mvb_last_kernel_ms(void);
/* kernel of the last solve launch: 0 general (simplex_kernel), 1 private-state lazy loop
 * (simplex_ps_kernel), 2 resident, tableau in one CTA's shared memory (simplex_res_kernel),
 * 3 cluster-resident, tableau in the distributed shared memory of a CTA cluster (simplex_cres_kernel) */
int mvb_last_kernel_mode(void);
/* A/B switches of the CUDA engine (tests, profiling; each also has an environment variable read at
 * start-up): "no_res" (MVOLPS_NO_RES) never use the resident kernel, "no_ps" (MVOLPS_NO_PS) general
 * kernel only, "no_handover" (MVOLPS_NO_HANDOVER) phase 2 after a phase 1 stays in the general kernel,
 * "cres" (MVOLPS_CRES) 1 = LPs too large for one CTA's shared memory that fit a cluster of 2/4/8 CTAs use the
 * cluster-resident kernel (default 0: the lazy loop is faster on full solves),
 * "lz_k" (MVOLPS_LZ_K) cap of the pending-update depth, "ps_nt" (MVOLPS_PS_NT) CTA width of wide
 * launches, "lz_ring" (MVOLPS_LZ_RING) 0 = the flush of HBM-streaming batches reads through registers
 * instead of the cp.async ring, "last_ring" / "last_lz_k" / "last_stream" (query only) ring slots per
 * thread, pending depth and kernel (1 = simplex_ps_stream_kernel) of the last lazy launch.
 * value < 0 only queries; returns the previous value, -1 on an unknown name.  oracle: -1 */
int mvb_engine_option(const char *name, int value);
/* shape of the last launch: CTAs per LP, groups, threads per CTA */
int mvb_last_launch(int *ctas_per_lp, int *groups, int *threads);

/* ---- LP objects ---- */
/* new empty LP with room for row_cap rows (grown on demand) */
mvb_lp *mvb_lp_create(int m, int n, int row_cap);
void mvb_lp_free(mvb_lp *lp);
/* the same for count LPs at once (the open queue of a finished B&B run: 10^5 handles) */
void mvb_lp_free_many(mvb_lp *const *lps, int count);
/* load host data and reset to the slack basis (all auxiliaries basic).
 * A: m*n row-major; cost: n (internal min form); lb/ub: n+m by variable id */
int mvb_lp_load(mvb_lp *lp, const double *A, const double *cost,
                const double *lb, const double *ub);
/* a host buffer of at least `bytes` bytes the caller may build A in before mvb_lp_load (CUDA: pinned
 * memory, so the upload is one DMA; reused by the next call, valid until then).  NULL: use your own */
void *mvb_host_staging(size_t bytes);
/* deep copy incl. tableau, basis and bounds (device-to-device on CUDA) */
mvb_lp *mvb_lp_clone(const mvb_lp *src);
/* same, into an existing object of equal (n, row capacity >= src m) */
int mvb_lp_assign(mvb_lp *dst, const mvb_lp *src);
int mvb_lp_dims(const mvb_lp *lp, int *m, int *n);
/* change bounds of variable id k; non-basic flags are repaired */
int mvb_lp_set_bounds(mvb_lp *lp, int k, double lb, double ub);
/* append constraint row a.x in [lb,ub]; a: n dense over structurals.
 * Its auxiliary enters the basis; the tableau row is formed on the engine. */
int mvb_lp_append_row(mvb_lp *lp, const double *a, double lb, double ub);

/* ---- solve ---- */
/* solve count LPs (one launch on CUDA); out[count] */
int mvb_solve_batch(mvb_lp *const *lps, int count, const mvb_opts *opts,
                    mvb_result *out);

/* the same solve in two halves, so that the host can work while the device does: begin enqueues the
 * launch (bounds patches up, kernel, results down) and returns; end waits, runs what the first launch
 * hands on (phase 1 / phase 2 hand-overs) and delivers out[count] in the order of begin's lps.  One
 * batch at a time; between the two calls its LPs must not be touched, other LPs may be created,
 * cloned, changed and freed (the batched B&B driver prepares the next round meanwhile, DESIGN.md
 * section 3).  A batch that needs more than one first launch is solved inside begin. */
int mvb_solve_batch_begin(mvb_lp *const *lps, int count, const mvb_opts *opts);
int mvb_solve_batch_end(mvb_result *out);

/* ---- read-back (only small vectors cross PCIe) ---- */
/* x[n+m] by variable id (raw, no rounding) */
int mvb_lp_get_primal(mvb_lp *lp, double *x);
/* head[m]: id basic in row i (0-based); nb[n]: id at position j; at[n]: MVB_AT_* */
int mvb_lp_get_basis(mvb_lp *lp, int *head, int *nb, int *at);
/* the same without a copy: pointers to the engine's own arrays of the last solve (x[n+m], head[m],
 * nb[n], at[n]), valid until the LP is solved again, changed or freed */
int mvb_lp_solution_view(mvb_lp *lp, const double **x, const int **head, const int **nb, const int **at);
/* gather tableau rows (1-based row numbers, 0 = reduced-cost row) into
 * out[nrows*n]; columns follow the non-basic positions (see get_basis) */
int mvb_lp_get_tab_rows(mvb_lp *lp, const int *rows, int nrows, double *out);
/* the same for `count` LPs in ONE device pass (all branching candidates of a B&B round,
 * bs.cpp:249-258 / gmi.cpp:36): LP k contributes nrows[k] rows, its row numbers follow
 * those of LP k-1 in `rows`, its rows (n_k doubles each) follow those of LP k-1 in `out` */
int mvb_lp_get_tab_rows_batch(mvb_lp *const *lps, int count, const int *rows, const int *nrows,
                              double *out);
/* mvb_lp_get_tab_rows for callers that may not allocate device memory (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
 * reference.  Same arguments and result as
 * mvb_lp_get_tab_rows; the row numbers and the gathered rows are staged in ONE static device array
 * of MVB_NOALLOC_SCRATCH_BYTES bytes (held by every GPU that loads the library until the process
 * ends), so the call makes no cudaMalloc / cudaFree / allocating library call.  It fails with a
 * message in mvb_last_error() when align256(4*nrows) + 8*nrows*n exceeds that size (e.g. more
 * than 255 rows of a 512-column LP), or while clones / cut rows are still queued on the engine
 * (carrying them out may allocate: mvb_solve_batch or mvb_lp_get_tab_rows does it).  All calls
 * share the array and it keeps the last call's contents: calls that overlap on the GPU (several
 * host threads) race -- serialise them. */
#define MVB_NOALLOC_SCRATCH_BYTES (1 << 20)
int mvb_lp_get_tab_rows_noalloc(mvb_lp *lp, const int *rows, int nrows, double *out);
/* whole current tableau ((m+1)*n) and beta (m+1, [0] unused); parity tests */
int mvb_lp_get_tableau(mvb_lp *lp, double *T, double *beta);
/* copies up to cap (row,col) pairs of the last solve; returns total logged */
int mvb_lp_get_pivot_log(mvb_lp *lp, int *rowcol, int cap);

#ifdef __cplusplus
}
#endif
#endif
