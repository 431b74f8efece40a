/* glpk.h -- the GLPK C-API subset that MVOLPS links against, served by the
 * B200 engine (product: mvolps_b200/lib/libglpk_b200.so) or, for tests only,
 * by the CPU oracle (oracle/_build/libglpk_oracle.so).
 *
 * This is the drop-in boundary of the hot path: the
 * reference's objects leave exactly these symbols undefined.  Every entry
 * point cites the reference call site(s) it serves (paths relative to the
 * MVOLPS tree).  Index conventions are GLPK's: rows 1..m, columns 1..n,
 * ind[1..len]/val[1..len] arrays with element 0 ignored, variable numbers
 * k = 1..m for auxiliary (row) variables and m+1..m+n for structural ones.
 *
 * Written from the public GLPK API documentation; no GLPK source is used.
 */
#ifndef MVOLPS_B200_GLPK_H
#define MVOLPS_B200_GLPK_H

#ifdef __cplusplus
extern "C" {
#endif

#define GLP_MAJOR_VERSION 4
#define GLP_MINOR_VERSION 65

typedef struct glp_prob glp_prob;
typedef struct glp_smcp glp_smcp; /* only NULL is ever passed (bs.cpp:117) */

/* optimisation direction */
#define GLP_MIN 1
#define GLP_MAX 2
/* kind of structural variable */
#define GLP_CV 1
#define GLP_IV 2
#define GLP_BV 3
/* type of auxiliary/structural variable */
#define GLP_FR 1
#define GLP_LO 2
#define GLP_UP 3
#define GLP_DB 4
#define GLP_FX 5
/* status of auxiliary/structural variable */
#define GLP_BS 1
#define GLP_NL 2
#define GLP_NU 3
#define GLP_NF 4
#define GLP_NS 5
/* solution status */
#define GLP_UNDEF 1
#define GLP_FEAS 2
#define GLP_INFEAS 3
#define GLP_NOFEAS 4
#define GLP_OPT 5
#define GLP_UNBND 6
/* enable/disable flag */
#define GLP_ON 1
#define GLP_OFF 0
/* MPS file format */
#define GLP_MPS_DECK 1
#define GLP_MPS_FILE 2
/* glp_simplex return codes used here */
#define GLP_EFAIL 0x05
#define GLP_EITLIM 0x08

/* ---- lifecycle (util.cpp:32-41, bs.cpp:88,114-116) ---- */
glp_prob *glp_create_prob(void);
void glp_delete_prob(glp_prob *P);
void glp_erase_prob(glp_prob *P);
/* copies rows, columns, bounds, kinds, objective, basis statuses, solution
 * and (engine extension) the device-resident tableau, so that a child node
 * warm-starts from its parent's final basis (util.cpp:34, bs.cpp:116) */
void glp_copy_prob(glp_prob *dest, glp_prob *prob, int names);

/* ---- I/O (util.cpp:278,284,290; 2test.cpp:45-62) ---- */
int glp_read_lp(glp_prob *P, const void *parm, const char *fname);
int glp_read_mps(glp_prob *P, int fmt, const void *parm, const char *fname);
const char *glp_version(void);
int glp_term_out(int flag);

/* ---- dimensions (util.cpp:45-46,299,305-306; gmi.cpp:15-16) ---- */
int glp_get_num_rows(glp_prob *P);
int glp_get_num_cols(glp_prob *P);
int glp_get_num_int(glp_prob *P);

/* ---- objective (util.cpp:51-58,436-455; bs.cpp:125-190) ---- */
int glp_get_obj_dir(glp_prob *P);
void glp_set_obj_dir(glp_prob *P, int dir);
double glp_get_obj_coef(glp_prob *P, int j); /* j = 0 is the constant term */
void glp_set_obj_coef(glp_prob *P, int j, double coef);
double glp_get_obj_val(glp_prob *P);

/* ---- rows (cut.cpp:23,40-43; gmi.cpp:45-47,84; util.cpp:65-130,355-375) ---- */
int glp_add_rows(glp_prob *P, int nrs);
int glp_get_row_type(glp_prob *P, int i);
double glp_get_row_lb(glp_prob *P, int i); /* -DBL_MAX when absent */
double glp_get_row_ub(glp_prob *P, int i); /* +DBL_MAX when absent */
void glp_set_row_bnds(glp_prob *P, int i, int type, double lb, double ub);
int glp_get_mat_row(glp_prob *P, int i, int ind[], double val[]);
void glp_set_mat_row(glp_prob *P, int i, int len, const int ind[],
                     const double val[]);
int glp_get_row_stat(glp_prob *P, int i);

/* ---- columns (bs.cpp:274,282; gmi.cpp:18,23,50-52; util.cpp:319-321,436-444) ---- */
int glp_get_col_type(glp_prob *P, int j);
double glp_get_col_lb(glp_prob *P, int j);
double glp_get_col_ub(glp_prob *P, int j);
void glp_set_col_bnds(glp_prob *P, int j, int type, double lb, double ub);
int glp_get_col_kind(glp_prob *P, int j);
double glp_get_col_prim(glp_prob *P, int j);
int glp_get_col_stat(glp_prob *P, int j);

/* ---- solve (bs.cpp:117,279,287; util.cpp:423; gmi.cpp:36) ---- */
int glp_simplex(glp_prob *P, const glp_smcp *parm);
int glp_get_status(glp_prob *P);
/* row of the current simplex tableau for basic variable k, expressed over
 * the non-basic variables: x_k = sum val[t] * x_{ind[t]}.  Entries are
 * returned in ascending variable number (documented engine ordering). */
int glp_eval_tab_row(glp_prob *P, int k, int ind[], double val[]);

/* ---- model-building extras (GLPK names; used by this repo's generators,
 *      tests and driver, not by the reference) ---- */
int glp_add_cols(glp_prob *P, int ncs);
void glp_set_col_kind(glp_prob *P, int j, int kind);
double glp_get_row_prim(glp_prob *P, int i);
int glp_get_it_cnt(glp_prob *P); /* simplex iterations of the last solve */
/* Not GLPK: the tolerance within which a primal value is reported as the integer next to it
 * (default 1e-9, environment MVOLPS_SNAP_TOL; 0 = report the raw tableau values).  It stands in
 * front of the reference's exact integrality test (util.cpp:435-451).  tol < 0 only queries;
 * returns the previous value. */
double mvb_glp_snap_tol(double tol);

#ifdef __cplusplus
}
#endif
#endif
