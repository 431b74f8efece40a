/* mvolps_dist.h -- multi-GPU layer of the LP-relaxation engine, C ABI.
 *
 * (1) One very large tableau, COLUMN-SHARDED across the GPUs of one box
 *     (north_star: "one very large tableau is column-sharded with an NCCL
 *     argmin allreduce per pivot"; BASELINE configs[4], m=16384 n=32768).
 *     Rank k owns the global non-basic positions [col0, col0+ncols) of every
 *     tableau row; beta, basis heads and bounds are replicated.  Per pivot:
 *       local pricing (K1) -> candidate (key, variable id, its column, m+1
 *       doubles) -> ONE exchange (one process per GPU: the pricing kernel
 *       pushes its record into every peer's CUDA-IPC mailbox over NVLink and
 *       the deciding kernel waits on arrival counters; ncclAllGather with
 *       MVOLPS_SHARD_EXCHANGE=nccl) -> every rank picks the same global
 *       entering column by the total order (value bits, variable id), runs the
 *       ratio test (K2) on it redundantly, and applies the rank-1 update (K3,
 *       K4) to its own columns; the pivot row is local to every shard.
 *     Pivot sequence, tableau and objective are bit-identical to the
 *     single-GPU engine (mvolps_engine.h) and therefore to the oracle.
 *     It stands behind the same reference call as the single-GPU path:
 *     glp_simplex (bs.cpp:117,279,287).
 *     Restrictions of this path: it starts from the slack basis with every
 *     non-basic variable at value 0 and runs primal phase 2 when that basis is
 *     primal feasible (configs[1]/[4]) or the dual simplex when it is dual
 *     feasible instead (covering rows: K5a on the replicated basic values, K5b
 *     over the local columns of the leaving row, the same record exchange),
 *     or the composite phase 1 followed by phase 2 when it is neither (the
 *     infeasibility costs of a rank's columns need no exchange: all rows of
 *     a column live on its rank).  Non-zero non-basic start values are
 *     rejected with an error (mvb_shard_start_mode tells in advance).
 *
 * (2) Open B&B nodes SHARDED across GPUs: node LPs are independent, so each
 *     rank solves its own nodes with mvb_solve_batch; the only exchange is
 *     the incumbent / bound and the open-node counts (max-reductions of a
 *     few doubles per round, `mvb_comm::allreduce_max`) plus the occasional
 *     node hand-over between a rich and a poor rank (work stealing,
 *     `mvb_comm::sendrecv_host`).  The driver that does this for one tree is
 *     mvb_bnb_run_dist (include/mvolps_bnb.h).
 *
 * The communicator is a small vtable so that the same host code runs over
 * NCCL (product; libnccl.so.2 is dlopen'ed, preferring the copy already loaded
 * by the process), over a single-process loop-back of several shards on one
 * GPU (tests), or over callbacks supplied by the caller (gloo in CPU tests).
 */
#ifndef MVOLPS_DIST_H
#define MVOLPS_DIST_H

#include <stddef.h>

#include "mvolps_comm.h"
#include "mvolps_engine.h"

#ifdef __cplusplus
extern "C" {
#endif

/* struct mvb_comm: include/mvolps_comm.h */

/* ---- NCCL communicator (one process per GPU) ---- */
#define MVB_NCCL_ID_BYTES 128
/* rank 0: fill id[128] (ncclGetUniqueId); the caller ships it to the other ranks
 * (bench.py: torch.distributed broadcast) */
int mvb_nccl_unique_id(void *id);
/* every rank, after mvb_device_init(local_rank) */
mvb_comm *mvb_comm_nccl(const void *id, int rank, int world);
/* member `rank` of a single-process group of `world` shards on ONE GPU (tests,
 * emulation of the ranks on one device): allgather_dev is NULL and
 * mvb_shard_solve is called once with all shards */
mvb_comm *mvb_comm_local(int rank, int world);
void mvb_comm_free(mvb_comm *c);

/* incumbent / bound broadcast of a node-sharded tree (bs.cpp:172 keeps the first
 * strictly better incumbent): on return *incumbent is the maximum over the ranks
 * and *owner the lowest rank that held it.  Uses comm->allreduce_max twice. */
int mvb_incumbent_exchange(const mvb_comm *c, double *incumbent, int *owner);

/* contiguous split of n positions over `world` ranks, multiples of 16 columns
 * (row stride alignment) except for the last rank */
void mvb_shard_partition(int n, int world, int rank, int *col0, int *ncols);

/* ---- column-sharded LP ---- */
typedef struct mvb_shard mvb_shard;

/* shard of an LP with m rows and n_global structural columns holding the
 * columns [col0, col0+ncols) = mvb_shard_partition(n_global, comm->world, comm->rank) */
mvb_shard *mvb_shard_create(int m, int n_global, int col0, int ncols, const mvb_comm *comm);
void mvb_shard_free(mvb_shard *s);
/* which simplex the column-sharded path would run from the slack basis of this LP (cost: n, internal
 * min form; lb/ub: n + m by variable id): MVB_MODE_PRIMAL2 (slack basis primal feasible), MVB_MODE_DUAL
 * (not primal but dual feasible: covering rows, B&B-style right-hand sides), MVB_MODE_PRIMAL1 (neither:
 * composite phase 1, then phase 2), 0 = not eligible (non-zero non-basic values: one GPU), by the rule of
 * the single-GPU engine's mode selection.  What glp_simplex asks before it routes an
 * LP to the shard group (bs.cpp:117). */
int mvb_shard_start_mode(int m, int n, const double *cost, const double *lb, const double *ub);
/* A_local: m x ncols row-major (the shard's columns of A); cost_local: ncols
 * (internal min form); lb/ub: n_global + m by variable id (replicated) */
int mvb_shard_load(mvb_shard *s, const double *A_local, const double *cost_local,
                   const double *lb, const double *ub);
/* Collective.  NCCL: every rank calls it with its one shard (nlocal = 1).
 * Single-process group: one call with all shards of the LP in rank order; the
 * all-gather is emulated with device-to-device copies.  out: one result,
 * identical on every rank. */
int mvb_shard_solve(mvb_shard *const *shards, int nlocal, const mvb_opts *opts, mvb_result *out);
/* x[n_global + m] by variable id of the last solve (replicated) */
int mvb_shard_get_primal(mvb_shard *s, double *x);
/* basis of the last solve (replicated): head[m] = id basic in row i, nb[n_global] / at[n_global] =
 * id and MVB_AT_* flag at every global non-basic position */
int mvb_shard_get_basis(mvb_shard *s, int *head, int *nb, int *at);
/* copies up to cap (row, global position) pairs of the last solve; returns total logged */
int mvb_shard_get_pivot_log(mvb_shard *s, int *rowcol, int cap);
/* device time of the last solve's kernels + exchanges (CUDA events), ms */
double mvb_shard_last_ms(mvb_shard *s);
/* how the per-pivot records travelled in the last solve: 0 device-to-device copies by the host
 * (single-process group), 1 ncclAllGather, 2 pushed into the peers' mailboxes by the price
 * kernel (CUDA IPC over NVLink; default for one process per GPU, MVOLPS_SHARD_EXCHANGE=nccl: 1) */
int mvb_shard_last_exchange(mvb_shard *s);

/* ---- the sharded path behind glp_simplex (bs.cpp:117) ----
 * Installs the group glp_simplex may route a large LP to: comm != NULL: the NCCL (or caller-supplied)
 * communicator of this process, glp_simplex is then COLLECTIVE for such LPs -- every rank calls it on
 * its own copy of the same problem; comm == NULL && local_world > 1: a single-process group of that
 * many shards on this GPU (tests); local_world <= 1 and comm == NULL: routing off (default).  An LP
 * takes the sharded path when it has no engine tableau yet, starts primal feasible from the slack
 * basis with every non-basic at value 0 (what that path solves), and its tableau has at least
 * min_bytes bytes (<= 0: 1 GiB).  Afterwards status, objective, primal values and row / column
 * statuses are those of the single-GPU solve, bit for bit; the tableau stays with the shards (a
 * later glp_eval_tab_row or warm start rebuilds on one GPU). */
int mvb_glp_use_shards(const mvb_comm *comm, int local_world, long long min_bytes);

#ifdef __cplusplus
}
#endif
#endif
