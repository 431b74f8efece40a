/* mvolps_bnb.h -- batched branch-and-bound driver, C ABI.
 *
 * Restates the reference driver `int branchAndBound(glp_prob*, ParameterObj&)`
 * (bs.h:7, bs.cpp:54-348) with identical node order, object ids, prune tags,
 * incumbent updates and printed tree, but solves the node LPs in device
 * batches: with FIFO node selection the first W open nodes are all going to
 * be popped and solved whatever the others return (bs.cpp:117 precedes the
 * bound test at bs.cpp:210), so they are solved as one engine launch and then
 * consumed on the host in FIFO order; the children created while consuming a
 * batch (bs.cpp:269-288) are solved as a second launch.
 */
#ifndef MVOLPS_BNB_H
#define MVOLPS_BNB_H

#include "glpk.h"
#include "mvolps_comm.h"

#ifdef __cplusplus
extern "C" {
#endif

/* strategies: values follow util.h:30-33 and the CLI flags (2test.cpp:25-37) */
#define MVB_NODE_FIFO 0 /* "-bs 0": front of the deque (util.cpp:157-167) */
#define MVB_NODE_BEST 1 /* "-bs 1": first maximum of upperBound (util.cpp:170-187) */
#define MVB_VAR_VO 0    /* "-vs 0": first violated (util.cpp:192-195) */
#define MVB_VAR_VFP 1   /* "-vs 1": as the reference: evaluated on the unsolved root object */
#define MVB_VAR_VGO 2   /* "-vs 2": largest positive objective coefficient */
#define MVB_VAR_VFP_NODE 3 /* not in the reference: VFP as util.cpp:197-213 intends it, the violated
                              variable whose fractional part in the NODE's LP solution is closest
                              to 1/2 (first on ties); "-vs 3" of mvolps_b200 */
#define MVB_CUT_NONE 0  /* "-cm 0" */
#define MVB_CUT_REF 1   /* "-cm 1": literal restatement of gmi.cpp:11-117 + cut.cpp:11-46 */
#define MVB_CUT_GMI 2   /* textbook Gomory mixed-integer cut from the tableau row */

/* modes of the node-sharded run (mvb_bnb_run_dist) */
#define MVB_DIST_STEAL 0 /* every rank walks its own share of the tree; exact optimum, per-rank oids */
#define MVB_DIST_SYNC 1  /* round-synchronous: the LPs of a round are dealt over the ranks, results are
                            exchanged and consumed in the serial order: oids, prune tags, tree text and
                            event stream are those of the serial run on every rank */

/* work stealing of the node-sharded run (mvb_bnb_run_dist) */
#define MVB_STEAL_LOW 4   /* a rank with fewer open nodes than this asks for more */
#define MVB_STEAL_MAX 256 /* nodes per hand-over */

typedef struct mvb_bnb_params {
  int node_strategy;
  int var_strategy;
  int cut_mode;
  int batch_width; /* FIFO nodes solved per launch; <=0: 256 */
  int max_iters;   /* node budget: stop after this many loop iterations; <=0: none
                      (the reference's own limit, bs.cpp:320, always applies) */
  int quiet;       /* !=0: do not print the tree / solution to stdout */
  int no_steal;    /* dist: !=0 disables the node hand-over between ranks (A/B runs, tests) */
  int no_speculation; /* MVB_NODE_BEST: !=0 solves the two children of one node per launch, as the
                         reference order suggests, instead of speculating over batch_width nodes */
  int timed_events;   /* !=0: every line of mvb_bnb_event_text starts with the milliseconds since the
                         start of the run (BaseMessagePOD::timeSpan, message.h:142-146) */
  /* ---- fields below: 0 = the reference's behaviour ---- */
  int total_budget;   /* dist: !=0: max_iters is the node budget of the WHOLE job (after the queue is
                         dealt out every rank gets an equal share of what is left), not per rank */
  int cut_select;     /* which cut of the node becomes its new row: 0 = the last one pooled, whichever
                         node produced it (cut.cpp:16-21, the reference's "TODO heuristic"); 1 = among the
                         cuts THIS node generated, the one its LP solution violates most (largest
                         (lb - a.x)/|a|, first on ties); none generated: no row */
  int root_heuristic; /* !=0: before branching on the root, round its LP solution down (integer columns)
                         and, when that point is feasible, take it as the first incumbent (event
                         "heuristic"); not in the reference */
  int dist_mode;      /* mvb_bnb_run_dist: MVB_DIST_STEAL (0) or MVB_DIST_SYNC (1), see below */
  double cut_skip;    /* "-cf x" (2test.cpp:135-149 parses it into _cutChance, util.cpp:259-261, which the
                         reference never reads): cuts are generated at a branched node with probability
                         1 - cut_skip, decided by a fixed hash of the node's oid (reproducible); 0 = at
                         every branched node, as the reference does */
  int no_event_text;  /* !=0: the event lines are not formatted and mvb_bnb_event_text is empty (the reference
                         formats and sends them only with --server, message.cpp:32-168); the event hash,
                         which covers (event, oid, pid, prune tag), is computed all the same */
} mvb_bnb_params;

typedef struct mvb_bnb_stats {
  int iterations;        /* `count` of bs.cpp:94 */
  int oids;              /* object ids handed out */
  int lp_solves;         /* glp_simplex-equivalent solves (3 per branched node) */
  int launches;          /* engine batch launches */
  long long simplex_iters;
  int has_incumbent;
  int incumbent_oid;
  double incumbent;      /* bestLower */
  int hit_limit;         /* stopped by max_iters or the reference's 200000 cap */
  unsigned long long event_hash; /* FNV-1a over (event, oid, pid, prune tag) */
  int split_at;          /* dist: iteration count at which the queue was dealt out (-1: never) */
  int incumbent_rank;    /* dist: rank that found the incumbent (0 in serial runs) */
  long long total_nodes; /* dist: nodes processed by all ranks (serial: = iterations) */
  int cached_solves;     /* of lp_solves: requests on an LP unchanged since its last optimal solve
                            (the re-solve of bs.cpp:117 on a copy), answered without a launch */
  int nodes_sent;        /* dist: open nodes this rank handed to poorer ranks */
  int nodes_received;    /* dist: open nodes this rank took over from richer ranks */
  int steal_rounds;      /* dist: rounds in which any pair of ranks exchanged nodes */
  int spec_nodes;        /* best-first: nodes whose children were solved ahead of their turn */
  int spec_adopted;      /* ... of which the serial order branched (the rest were fathomed first
                            or are still open) */
  int launches_aux;      /* batched tableau-row gathers of the cut step (one per round with cuts) */
  int solved_here;       /* LP solves this rank's device ran (MVB_DIST_SYNC: its share of lp_solves) */
} mvb_bnb_stats;

typedef struct mvb_bnb mvb_bnb;

/* run B&B on prob (not modified, like the reference, which works on copies) */
mvb_bnb *mvb_bnb_run(glp_prob *prob, const mvb_bnb_params *params);

/* Node-sharded B&B over comm->world ranks (one process per GPU).  Every rank
 * calls it with the same problem and parameters.  The ranks run the serial
 * driver in lock step (identical decisions) until the FIFO queue holds at
 * least 4 nodes per rank; the open nodes are then dealt round-robin and each
 * rank continues on its own share, one batch per round, exchanging only the
 * incumbent (comm->allreduce_max: the bound test of bs.cpp:210 uses the best
 * incumbent of ALL ranks) and the per-rank numbers of open nodes (termination
 * and balancing).  Work stealing: every round all ranks derive the same
 * matching from those counts -- the poorest rank with fewer than
 * MVB_STEAL_LOW open nodes is paired with the richest, which hands over half
 * of the difference (at most MVB_STEAL_MAX nodes, newest first) through
 * comm->sendrecv_host; a node travels as its column bounds + cut rows + bound
 * (the root problem is known everywhere) and is re-solved from the slack
 * basis by the receiver.  Exact search: the optimum equals the serial one;
 * node order, oids and tree text are per rank and NOT those of the serial
 * run (the non-sequence-identical mode); a received node hangs
 * under the root in the receiver's tree.
 * Results per rank: its own tree / events; stats.incumbent is the global
 * optimum on every rank, mvb_bnb_get_incumbent returns the winning x on every
 * rank, stats.iterations counts the nodes this rank processed. */
mvb_bnb *mvb_bnb_run_dist(glp_prob *prob, const mvb_bnb_params *params, const mvb_comm *comm);
/* The hand-over plan of one round of mvb_bnb_run_dist, a pure function of the per-rank open-node
 * counts (cnt[r] < 0: rank r has retired): fills donor[i] -> thief[i], nodes[i] for up to cap
 * pairs and returns the number of pairs.  Exposed for tests and for hosts that run their own
 * exchange. */
int mvb_bnb_steal_plan(const double *cnt, int world, int *donor, int *thief, int *nodes, int cap);
void mvb_bnb_free(mvb_bnb *h);
const mvb_bnb_stats *mvb_bnb_get_stats(const mvb_bnb *h);
/* tree exactly as PrettyPrintTree prints it (tree_print.h:12-22, bs.cpp:331-341) */
const char *mvb_bnb_tree_text(const mvb_bnb *h);
/* the "[oid] Solution is: ..." string of bs.cpp:179-190 ("" when none) */
const char *mvb_bnb_solution_text(const mvb_bnb *h);
/* one line per event, GrUMPy-like: "type oid pid dir [fields]" (message.h:141-226) */
const char *mvb_bnb_event_text(const mvb_bnb *h);
/* incumbent x[1..n] (x[0] unused); returns n or 0 when there is no incumbent */
int mvb_bnb_get_incumbent(const mvb_bnb *h, double *x, int cap);

#ifdef __cplusplus
}
#endif
#endif
