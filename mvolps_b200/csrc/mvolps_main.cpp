// mvolps_main.cpp -- command line of the B200 build; same flags and file-type
// rule as the reference's main (2test.cpp:13-157, InputParser.h:8-35), running
// the batched driver (bnb_driver.cpp) instead of the serial one.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>
#include <vector>

#include <sys/wait.h>
#include <unistd.h>

#include "../../include/glpk.h"
#include "../../include/mvolps_bnb.h"
#include "../../include/mvolps_dist.h"
#include "../../include/mvolps_engine.h"

namespace {
// presence + next-token lookup, as InputParser.h:16-31
struct Args {
  std::vector<std::string> tokens;
  Args(int argc, char **argv) { for (int i = 1; i < argc; i++) tokens.emplace_back(argv[i]); }
  bool has(const std::string &o) const { return std::find(tokens.begin(), tokens.end(), o) != tokens.end(); }
  std::string get(const std::string &o) const {
    auto it = std::find(tokens.begin(), tokens.end(), o);
    if (it != tokens.end() && ++it != tokens.end()) return *it;
    return "";
  }
};
}  // namespace

int main(int argc, char **argv) {
  Args in(argc, argv);
  if (in.has("-h") || in.has("--help")) {
    std::cout << "Usage: mvolps_b200 [OPTION]\n"
              << "  -f [FILENAME.{mps|lp}]\n"
              << "  -s/--silent  -v/--verbose  -d/--debug  -so/--solver-output\n"
              << "  -vs [{0|1|2|3}] variable choice: order | fractional part (as the reference: on the root object) |\n"
              << "                 objective impact | fractional part of the node's own solution\n"
              << "  -bs [{0|1}]    node choice: FIFO | best bound\n"
              << "  -cm [{0|1|2}]  cuts: off | reference gmi.cpp formula | textbook GMI\n"
              << "  -cf X          cut frequency in [0, 1]: share of the branched nodes that generate cuts (with -cm)\n"
              << "  -cs [{0|1}]    which cut becomes the node's new row: last pooled (reference) | most violated of the node's own\n"
              << "  -rh            root heuristic: the root LP solution rounded down as first incumbent\n"
              << "  -np N          one tree over N GPUs of this box: one process per GPU (forked here), NCCL over NVLink\n"
              << "  -dm [{0|1}]    with -np: node-sharded with work stealing (exact optimum, per-rank trees) |\n"
              << "                 round-synchronous (tree, ids and events of the serial run)\n"
              << "  -bw N          FIFO nodes solved per device launch (default 256)\n"
              << "  -nl N          node budget\n"
              << "  -ev FILE       write the event stream, timeSpan type oid pid dir ... (message.h:141-226)\n";
    return 0;
  }
  glp_term_out(GLP_OFF);
  if (in.has("-d") || in.has("--debug") || in.has("-so") || in.has("--solver-output")) glp_term_out(GLP_ON);
  if (!in.has("-f")) { std::cout << "see ./mvolps_b200 -h for usage\n"; return 0; }
  std::string fn = in.get("-f");
  std::string ext = fn.substr(fn.find_last_of('.') == std::string::npos ? 0 : fn.find_last_of('.'));
  glp_prob *prob = glp_create_prob();
  int rc;
  if (ext == ".lp") rc = glp_read_lp(prob, nullptr, fn.c_str());
  else if (ext == ".mps") rc = glp_read_mps(prob, GLP_MPS_FILE, nullptr, fn.c_str());
  else { std::cout << "Unrecognized filetype\n"; return -1; }
  if (rc) std::exit(-1);  // util.cpp:284-292
  mvb_bnb_params p{};
  auto ival = [&](const char *flag, int lo, int hi, int &dst) {
    if (!in.has(flag)) return true;
    int v = std::atoi(in.get(flag).c_str());
    if (v < lo || v > hi) { std::fprintf(stderr, "Unknown parameter value for %s\n", flag); return false; }
    dst = v;
    return true;
  };
  if (!ival("-bs", 0, 1, p.node_strategy) || !ival("-vs", 0, 3, p.var_strategy) ||
      !ival("-cm", 0, 2, p.cut_mode) || !ival("-bw", 1, 1 << 20, p.batch_width) ||
      !ival("-nl", 1, 1 << 30, p.max_iters) || !ival("-cs", 0, 1, p.cut_select))
    return -1;
  if (in.has("-cm") && in.has("-cf")) {  // 2test.cpp:135-149: read only together with -cm
    const double chance = std::atof(in.get("-cf").c_str());
    if (!(chance <= 1.0 && chance >= 0.0)) {
      std::fprintf(stderr, "Cut Frequency parameter must be in range [0.0, 1.0]\n");
      return -1;
    }
    p.cut_skip = 1.0 - chance;
  }
  p.root_heuristic = in.has("-rh");
  p.no_event_text = !in.has("-ev");  // the event lines are formatted only when they are written somewhere
  p.timed_events = in.has("-ev");
  int np = 1;
  if (!ival("-np", 1, 16, np) || !ival("-dm", 0, 1, p.dist_mode)) return -1;
  // -np N: fork one process per GPU BEFORE any CUDA call; rank 0's NCCL id travels through pipes
  int rank = 0;
  std::vector<int> id_pipe(2 * (size_t)np, -1);
  std::vector<pid_t> kids;
  if (np > 1) {
    for (int r = 1; r < np; r++)
      if (pipe(&id_pipe[2 * r])) { std::perror("pipe"); return 2; }
    std::fflush(stdout);
    for (int r = 1; r < np; r++) {
      pid_t c = fork();
      if (c < 0) { std::perror("fork"); return 2; }
      if (c == 0) { rank = r; kids.clear(); break; }
      kids.push_back(c);
    }
  }
  if (mvb_device_init(rank)) { std::fprintf(stderr, "rank %d: %s\n", rank, mvb_last_error()); return 2; }
  mvb_comm *comm = nullptr;
  if (np > 1) {
    unsigned char id[MVB_NCCL_ID_BYTES];
    if (rank == 0) {
      if (mvb_nccl_unique_id(id)) { std::fprintf(stderr, "%s\n", mvb_last_error()); return 2; }
      for (int r = 1; r < np; r++)
        if (write(id_pipe[2 * r + 1], id, sizeof id) != (ssize_t)sizeof id) { std::perror("write"); return 2; }
    } else if (read(id_pipe[2 * rank], id, sizeof id) != (ssize_t)sizeof id) {
      std::perror("read");
      return 2;
    }
    comm = mvb_comm_nccl(id, rank, np);
    if (!comm) { std::fprintf(stderr, "rank %d: %s\n", rank, mvb_last_error()); return 2; }
    if (rank != 0) p.quiet = 1;
  }
  mvb_bnb *h = comm ? mvb_bnb_run_dist(prob, &p, comm) : mvb_bnb_run(prob, &p);
  if (!h) return 2;
  if (rank != 0) {  // the other ranks have done their share; rank 0 reports
    mvb_bnb_free(h);
    mvb_comm_free(comm);
    return 0;
  }
  if (in.has("-ev")) {  // file sink of what the reference sends to its visualiser (message.cpp:32-191)
    std::FILE *f = std::fopen(in.get("-ev").c_str(), "w");
    if (!f) { std::fprintf(stderr, "cannot write %s\n", in.get("-ev").c_str()); return -1; }
    std::fputs(mvb_bnb_event_text(h), f);
    std::fputs("END", f);  // ~IPCDispatch closes its stream with this token (message.h:230-236)
    std::fclose(f);
  }
  if (!in.has("-s") && !in.has("--silent")) {
    const mvb_bnb_stats *st = mvb_bnb_get_stats(h);
    std::printf("[%s] %d iterations, %d LP solves in %d launches, %lld simplex iterations\n",
                mvb_backend(), st->iterations, st->lp_solves, st->launches, st->simplex_iters);
    if (np > 1)
      std::printf("[%d GPUs, %s] %lld nodes in all, incumbent %.12g (rank %d), rank 0 solved %d LPs, sent %d / received %d nodes\n",
                  np, p.dist_mode == MVB_DIST_SYNC ? "round-synchronous" : "node-sharded", st->total_nodes, st->incumbent,
                  st->incumbent_rank, st->solved_here, st->nodes_sent, st->nodes_received);
  }
  mvb_bnb_free(h);
  if (comm) mvb_comm_free(comm);
  glp_delete_prob(prob);
  int exit_code = 0;
  for (pid_t c : kids) {
    int stc = 0;
    if (waitpid(c, &stc, 0) < 0 || !WIFEXITED(stc) || WEXITSTATUS(stc) != 0) exit_code = 3;
  }
  return exit_code;
}
