// engine_shard.cuh -- column-sharded tableau simplex across the GPUs of one
// box (include/mvolps_dist.h).  Included at the end of engine_cuda.cu: it uses
// that file's device helpers (Key order, block_argmin, primal_ratio_row) so
// that the sharded path takes bit-for-bit the decisions of solve_lp / the oracle.
//
// Per pivot and rank:  shard_price_kernel  (K1 over the local columns + pack
// of the candidate column)  ->  ONE all-gather of (key, id, bounds, column)
// ->  shard_decide_kernel  (same global winner everywhere by the (value bits,
// variable id) order, replicated K2 ratio test, step, K3 scaled pivot row)
// on the decision stream, and  shard_update_kernel  (K4 on the local columns,
// tableau A -> B) on the update stream.  The decision chain of iteration t+1
// overlaps the update of iteration t: it reads the buffer that update reads
// and applies the one pending step on the fly (ShardView), bit for bit what
// the finished tableau holds.  Kernel boundaries and two events per iteration
// are the only synchronisation; the control block ping-pongs by iteration
// parity so a finished solve turns the remaining launches of a chunk into
// no-ops.
//
// Exchange step, one process per GPU: the price kernel PUSHES its record
// straight into every peer's mailbox over NVLink (peer pointers from CUDA IPC)
// and raises a per-sender arrival counter there; the decide kernel waits on
// the counters in its own memory.  No collective launch, no host in between:
// the pack, the transfer and the hand-shake are part of the two kernels that
// produce and consume the record.  The mailbox has SHARD_SLOTS records per
// sender (slot = iteration mod SHARD_SLOTS) because a fast rank may run up to
// two iterations ahead of a peer whose update kernels still read older
// columns.  ncclAllGather into the same slots is kept as the path of
// MVOLPS_SHARD_EXCHANGE=nccl (A/B runs) and when IPC is not available.
#pragma once
#include <dlfcn.h>

#include "../../include/mvolps_dist.h"

namespace {

struct ShardCtl {
  int status;  // -1 while running, then MVB_*
  int iters, pivots, flips, bland_iters, degen;
  int cur, cS;  // tableau buffer that holds the current tableau once the pending update landed; small-state parity
  int pend;     // 1: the update of the previous iteration (step[par ^ 1]) may still be running on the
                // update stream, readers take the other buffer and apply that step on the fly
  int mode;     // MVB_MODE_PRIMAL2 / MVB_MODE_DUAL (neither changes mode) or MVB_MODE_PRIMAL1, which becomes
                // PRIMAL2 when the decide kernel finds the NEXT basis primal feasible
  int it1;      // iterations spent in phase 1
  unsigned long long hash;
};

struct Cand {  // header of one exchange record; the column (m+1 doubles) follows at byte 64
  unsigned long long key;  // ~bits(|d_q|), 0 under Bland's rule, ~0: no eligible column
  int var, pos, flag, pad;
  double lq, uq;
  double dq;  // the cost the column was priced on (phase 1: the infeasibility cost d1_q; else T[0][q])
};
constexpr size_t SHARD_HDR = 64;

struct ShardStep {  // what iteration t leaves for its update kernel and for the readers of t+1
  int valid;   // 1: a basis change with a tableau update
  int r;       // pivot row
  int qloc;    // local position of the entering column, outside [0, nloc) when another rank owns it
  int wk;      // rank whose record holds the pivot column
  int src;     // tableau buffer the update reads (it writes src ^ 1)
  int slot;    // mailbox slot that holds the records of this iteration
};
constexpr int SHARD_SLOTS = 4;       // mailbox depth (iterations in flight across the ranks: 3)
constexpr int SHARD_MAX_PEERS = 16;
constexpr int SHARD_ST_TIMEOUT = 100;  // ctl.status: a peer's record did not arrive
constexpr unsigned long long SHARD_WAIT_NS = 10ULL * 1000 * 1000 * 1000;

struct ShardDev {
  int m, nloc, ld, col0, world, max_iters;
  double *T[2];
  double *beta[2];
  int *head[2];
  int *nb[2];
  int *at[2];
  double *lb, *ub;  // by variable id, n_global + m (replicated)
  ShardCtl *ctl;    // [2] by iteration parity
  ShardStep *step;  // [2] by iteration parity
  double *rho[2];   // scaled pivot row over the local columns, by iteration parity
  double *p1part;   // phase 1: partial column sums, ceil(m / MVB_P1_CHUNK) x ld (as LPDev::p1part)
  double *d1;       // phase 1: infeasibility costs of the local columns (ld)
  unsigned char *send, *recv[SHARD_SLOTS];  // recv[slot]: world records, by sender rank
  size_t rec_bytes;
  // peer exchange (nullptr / 0 when the host copies or NCCL moves the records)
  int rank, push;
  unsigned long long *flag;                    // [SHARD_SLOTS][world] arrival counters, in this rank's mailbox
  unsigned char *peer[SHARD_MAX_PEERS];        // base of every rank's mailbox as mapped here (own included)
  size_t flag_off;                             // offset of the counters inside a mailbox
  int *pivlog;
  int pivlog_cap;
};

// The pipeline (mvb_shard_solve): the decision chain of iteration t+1 (price -> all-gather ->
// decide, on the high-priority stream) runs WHILE the rank-1 update of iteration t streams the
// tableau on the update stream.  The chain therefore reads the buffer the running update reads
// (complete: the host makes it wait for update t-1) and applies step t on the fly -- one element
// costs the same fma the update applies, so every bit is the one the finished tableau will hold:
//     T_{t+1}[i][j] = i == r ? rho[j] : fma(col[i], rho[j], j == q ? 0 : T_t[i][j])
struct ShardView {
  const double *T;     // complete buffer
  const double *rho;   // pending step: scaled pivot row (local columns); nullptr: nothing pending
  const double *col;   // pending step: pivot column (m+1)
  int r, qloc, ld;
  __device__ __forceinline__ double at(int i, int j) const {
    const double v = ldcg(T + (size_t)i * ld + j);
    if (!rho) return v;
    const double rj = rho[j];
    return i == r ? rj : fma(col[i], rj, j == qloc ? 0.0 : v);
  }
};
__device__ __forceinline__ ShardView shard_view(const ShardDev &D, const ShardCtl &c, int par) {
  ShardView v;
  v.ld = D.ld;
  if (c.pend) {
    const ShardStep st = D.step[par ^ 1];
    v.T = D.T[c.cur ^ 1];
    v.rho = D.rho[par ^ 1];
    v.col = reinterpret_cast<const double *>(D.recv[st.slot] + (size_t)st.wk * D.rec_bytes + SHARD_HDR);
    v.r = st.r;
    v.qloc = st.qloc;
  } else {
    v.T = D.T[c.cur];
    v.rho = nullptr;
    v.col = nullptr;
    v.r = -1;
    v.qloc = -1;
  }
  return v;
}

__device__ void shard_emit_record(const ShardDev &D, const ShardView &V, int slot, const Key best, const int *at,
                                  double dq);

// K1 over the local columns + gather of the winning column into the send
// record.  Every CTA of the (small) grid repeats the cheap argmin over the
// local reduced costs and then gathers its own slice of the column rows, so the
// strided gather (one 32-byte sector per element) runs on several SMs.
__global__ void __launch_bounds__(1024, 1) shard_price_kernel(ShardDev D, int par, int slot) {
  __shared__ Key red[32];
  const ShardCtl c = D.ctl[par];
  if (c.status != -1) return;  // finished on every rank alike: nobody waits for a record
  const ShardView V = shard_view(D, c, par);
  const bool bland = c.degen >= MVB_BLAND_AFTER;
  const bool phase1 = c.mode == MVB_MODE_PRIMAL1;  // costs = D.d1 (shard_p1_* kernels of this iteration)
  const int *at = D.at[c.cS], *nb = D.nb[c.cS];
  Key best = key_none();
  for (int j = threadIdx.x; j < D.nloc; j += blockDim.x) {
    const int f = at[j];
    if (f == MVB_AT_FX) continue;
    const double dj = phase1 ? ldcg(D.d1 + j) : V.at(0, j);
    if (!dj_eligible(f, dj)) continue;
    const Key k{bland ? 0ULL : ~dbits(fabs(dj)), nb[j], j};
    if (key_less(k, best)) best = k;
  }
  best = block_argmin(best, red);  // result on every thread, identical in every CTA
  shard_emit_record(D, V, slot, best, at, best.idx >= 0 ? (phase1 ? ldcg(D.d1 + best.idx) : V.at(0, best.idx)) : 0.0);
}

// ---- composite phase 1 on the shards: the infeasibility costs d1_j = sum_i sign_i * T_ij of the
// LOCAL columns (all rows of a column live on its rank: no exchange), summed in the order of the
// specification: 32-row chunks in row order, then the chunk sums in order (p1_partials / p1_cost
// of the general loop).  Two small-state-free kernels in front of the price kernel, no-ops once
// the solve has left phase 1.
__global__ void __launch_bounds__(256) shard_p1_partials_kernel(ShardDev D, int par) {
  const ShardCtl c = D.ctl[par];
  if (c.status != -1 || c.mode != MVB_MODE_PRIMAL1) return;
  const ShardView V = shard_view(D, c, par);
  const double *b = D.beta[c.cS];
  const int *h = D.head[c.cS];
  const int m = D.m, n = D.nloc, ld = D.ld;
  const int nchunks = (m + MVB_P1_CHUNK - 1) / MVB_P1_CHUNK;
  const long long items = (long long)nchunks * n;
  for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < items;
       it += (long long)gridDim.x * blockDim.x) {
    const int ch = (int)(it / n), j = (int)(it - (long long)ch * n);
    const int i0 = 1 + ch * MVB_P1_CHUNK;
    const int i1 = i0 + MVB_P1_CHUNK - 1 < m ? i0 + MVB_P1_CHUNK - 1 : m;
    double acc = 0.0;
    for (int i = i0; i <= i1; i++) {
      const int k = h[i];
      const int sg = row_infeas(b[i], D.lb[k], D.ub[k]);
      if (sg > 0) acc = acc + V.at(i, j);
      else if (sg < 0) acc = acc - V.at(i, j);
    }
    D.p1part[(size_t)ch * ld + j] = acc;
  }
}
__global__ void __launch_bounds__(256) shard_p1_reduce_kernel(ShardDev D, int par) {
  const ShardCtl c = D.ctl[par];
  if (c.status != -1 || c.mode != MVB_MODE_PRIMAL1) return;
  const int nchunks = (D.m + MVB_P1_CHUNK - 1) / MVB_P1_CHUNK;
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < D.nloc; j += gridDim.x * blockDim.x) {
    double d = 0.0;
    for (int ch = 0; ch < nchunks; ch++) d = d + ldcg(D.p1part + (size_t)ch * D.ld + j);
    D.d1[j] = d;
  }
}

// The candidate `best` (idx = local position or -1, k, var) of this rank goes out as one record:
// header + its column through the view, into the send buffer (host copy / NCCL) or straight into
// this rank's place in every peer's mailbox, followed by the arrival counters (push mode).
__device__ void shard_emit_record(const ShardDev &D, const ShardView &V, int slot, const Key best, const int *at,
                                  double dq) {
  const int q = best.idx;
  const int ndst = D.push ? D.world : 1;
  const size_t place = (size_t)slot * D.world * D.rec_bytes + (size_t)D.rank * D.rec_bytes;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    Cand h{};
    h.key = q >= 0 ? best.k : ~0ULL;
    h.var = q >= 0 ? best.var : 0x7fffffff;
    h.pos = q >= 0 ? D.col0 + q : -1;
    h.flag = q >= 0 ? at[q] : 0;
    h.lq = q >= 0 ? D.lb[best.var] : 0.0;
    h.uq = q >= 0 ? D.ub[best.var] : 0.0;
    h.dq = dq;
    for (int k = 0; k < ndst; k++) *reinterpret_cast<Cand *>(D.push ? D.peer[k] + place : D.send) = h;
  }
  if (q >= 0)
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= D.m; i += gridDim.x * blockDim.x) {
      const double v = V.at(i, q);
      for (int k = 0; k < ndst; k++)
        reinterpret_cast<double *>((D.push ? D.peer[k] + place : D.send) + SHARD_HDR)[i] = v;
    }
  if (!D.push) return;
  // arrival: this CTA's stores are ordered before its increment of the sender's counter in
  // every mailbox (bar.sync, then fence.sys + atomic by one thread); the receiver expects
  // gridDim.x increments per use of the slot
  __syncthreads();
  if (threadIdx.x < D.world) {
    __threadfence_system();
    unsigned long long *f = reinterpret_cast<unsigned long long *>(D.peer[threadIdx.x] + D.flag_off) +
                            (size_t)slot * D.world + D.rank;
    atomicAdd_system(f, 1ULL);
  }
}

// the records of `slot` from every rank have landed (pushed by the peers' price kernels):
// one thread per sender spins on its counter with acquire loads; false on time-out
__device__ bool shard_wait_records(const ShardDev &D, int slot, unsigned long long expect, int *flagS) {
  if (threadIdx.x == 0) *flagS = 1;
  __syncthreads();
  if (threadIdx.x < D.world) {
    const unsigned long long *f = D.flag + (size_t)slot * D.world + threadIdx.x;
    unsigned long long t0 = 0, v;
    unsigned spins = 0;
    while (true) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
      if (v >= expect) break;
      if ((++spins & 1023u) == 0) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (!t0) t0 = now;
        else if (now - t0 > SHARD_WAIT_NS) { *flagS = 0; break; }
      }
    }
  }
  __syncthreads();
  return *flagS != 0;
}

// rank-1 update of this CTA's share of the local columns: pivot column from
// `colg` (m+1 doubles), scaled pivot row from `rhog` (both in global memory),
// same unit split and arithmetic as tile_update
__device__ void shard_tile_update(const ShardDev &D, const double *__restrict__ Ts,
                                  double *__restrict__ Td, int r, int qloc,
                                  const double *__restrict__ rhog, const double *__restrict__ colg,
                                  Smem &sm) {
  const int R = D.m + 1, ld = D.ld, g = blockIdx.x, G = gridDim.x;
  int tcl = 0;
  while ((1 << tcl) < (ld + 1) / 2 && (2 << tcl) <= (int)blockDim.x) tcl++;
  const int tcols = 1 << tcl, trows = blockDim.x >> tcl;
  const int tx = threadIdx.x & (tcols - 1), ty = threadIdx.x >> tcl;
  const int CW = 2 * tcols, nch = (ld + CW - 1) / CW;
  const long long nunits = (long long)nch * R;
  long long u0 = nunits * g / G;
  const long long u1 = nunits * (g + 1) / G;
  while (u0 < u1) {
    const int ch = (int)(u0 / R);
    const int ia = (int)(u0 - (long long)ch * R);
    const long long segend = u1 < (long long)(ch + 1) * R ? u1 : (long long)(ch + 1) * R;
    const int ib = ia + (int)(segend - u0);
    const int col = ch * CW + 2 * tx;
    const bool act = col < ld;
    const bool qx = col == qloc, qy = col + 1 == qloc;
    double2 rho = make_double2(0.0, 0.0);
    if (act) rho = ldcg2(rhog + col);
    for (int s0 = ia; s0 < ib; s0 += STAGE_ROWS) {
      const int s1 = ib < s0 + STAGE_ROWS ? ib : s0 + STAGE_ROWS;
      __syncthreads();
      for (int k = threadIdx.x; k < s1 - s0; k += blockDim.x) sm.colS[k] = colg[s0 + k];
      __syncthreads();
      if (!act) continue;
      for (int i = s0 + ty; i < s1; i += trows * UNROLL) {
        double2 t[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
          const int ii = i + u * trows;
          if (ii < s1) t[u] = ldcg2(Ts + (size_t)ii * ld + col);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
          const int ii = i + u * trows;
          if (ii < s1) {
            const double c = sm.colS[ii - s0];
            double2 o;
            o.x = fma(c, rho.x, qx ? 0.0 : t[u].x);
            o.y = fma(c, rho.y, qy ? 0.0 : t[u].y);
            if (ii == r) o = rho;
            stcg2(Td + (size_t)ii * ld + col, o);
          }
        }
      }
    }
    u0 = segend;
  }
}

}  // namespace

namespace {

// Decisions of one iteration on every rank (a few CTAs, each repeating the
// cheap part): global winner, replicated K2 ratio test, the step, the next
// small state, and the scaled pivot row rho over the local columns (K3) for
// the update kernel and for the readers of the next iteration.
__global__ void __launch_bounds__(512, 1) shard_decide_kernel(ShardDev D, int par, int slot,
                                                              unsigned long long expect) {
  __shared__ Smem sm;
  __shared__ int arrived;
  const ShardCtl c = D.ctl[par];
  ShardCtl *next = &D.ctl[par ^ 1];
  const bool lead = blockIdx.x == 0 && threadIdx.x == 0;
  ShardCtl n = c;
  n.pend = 0;
  ShardStep stp{};
  stp.slot = slot;
  if (c.status != -1) {
    if (lead) { *next = n; D.step[par] = stp; }
    return;
  }
  if (D.push && !shard_wait_records(D, slot, expect, &arrived)) {
    n.status = SHARD_ST_TIMEOUT;
    if (lead) { *next = n; D.step[par] = stp; }
    return;
  }
  const int m = D.m, tid = threadIdx.x, NT = blockDim.x;
  // same winner everywhere: minimum (key, variable id) over the ranks' candidates
  int wk = -1;
  unsigned long long bk = ~0ULL;
  int bvar = 0x7fffffff;
  for (int k = 0; k < D.world; k++) {
    const Cand *h = reinterpret_cast<const Cand *>(D.recv[slot] + (size_t)k * D.rec_bytes);
    if (h->pos < 0) continue;
    if (h->key < bk || (h->key == bk && h->var < bvar)) { bk = h->key; bvar = h->var; wk = k; }
  }
  const bool phase1 = c.mode == MVB_MODE_PRIMAL1;
  if (wk < 0) {
    n.status = phase1 ? MVB_NOFEAS : MVB_OPT;
    if (lead) { *next = n; D.step[par] = stp; }
    return;
  }
  if (c.iters >= D.max_iters) {
    n.status = MVB_ITLIM;
    if (lead) { *next = n; D.step[par] = stp; }
    return;
  }
  const Cand cand = *reinterpret_cast<const Cand *>(D.recv[slot] + (size_t)wk * D.rec_bytes);
  const double *col = reinterpret_cast<const double *>(D.recv[slot] + (size_t)wk * D.rec_bytes + SHARD_HDR);
  const bool bland = c.degen >= MVB_BLAND_AFTER;
  const double *beta = D.beta[c.cS];
  const int *head = D.head[c.cS];
  const int dir = (phase1 ? cand.dq : col[0]) < 0.0 ? 1 : -1;
  // K2 ratio test over the gathered pivot column, lowest variable id on ties (phase 1: an
  // infeasible basic variable blocks only where it reaches its violated bound)
  Key bi = key_none();
  for (int i = 1 + tid; i <= m; i += NT) {
    const int var = head[i];
    const double b = beta[i], l = D.lb[var], u = D.ub[var];
    double t;
    int fl;
    if (!primal_ratio_row(col[i], dir, phase1 ? row_infeas(b, l, u) : 0, b, l, u, t, fl)) continue;
    const Key k{dbits(t), var, i};
    if (key_less(k, bi)) bi = k;
  }
  bi = block_argmin(bi, sm.redK);
  const double lq = cand.lq, uq = cand.uq;
  const int fq = cand.flag;
  const double tflip = uq - lq;
  const double tmin = bi.idx >= 0 ? __longlong_as_double((long long)bi.k) : HUGE_VAL;
  int r = 0, newat = MVB_AT_LB;
  double delta, newbr = 0.0;
  if (tflip <= tmin) {
    if (!(tflip < HUGE_VAL)) {
      n.status = phase1 ? MVB_NOFEAS : MVB_UNBND;
      if (lead) { *next = n; D.step[par] = stp; }
      return;
    }
    delta = dir > 0 ? tflip : -tflip;
    newat = fq == MVB_AT_LB ? MVB_AT_UB : MVB_AT_LB;
    n.degen = 0;
  } else {
    r = bi.idx;
    double t;
    const double b = beta[r], l = D.lb[bi.var], u = D.ub[bi.var];
    primal_ratio_row(col[r], dir, phase1 ? row_infeas(b, l, u) : 0, b, l, u, t, newat);
    delta = dir > 0 ? tmin : -tmin;
    newbr = nb_value(fq, lq, uq) + delta;
    n.degen = tmin == 0.0 ? c.degen + 1 : 0;
  }
  const int kin = cand.var;
  int kout = kin;
  if (r) {
    kout = bi.var;
    if (D.lb[kout] == D.ub[kout]) newat = MVB_AT_FX;
  }
  const int qloc = cand.pos - D.col0;  // local position of the entering column, if it is ours
  const int gt = blockIdx.x * NT + tid, stride = gridDim.x * NT;
  // next small state (ping-pong), split over the grid
  {
    double *bd = D.beta[c.cS ^ 1];
    int *hd = D.head[c.cS ^ 1], *nd = D.nb[c.cS ^ 1], *ad = D.at[c.cS ^ 1];
    const int *ns = D.nb[c.cS], *as = D.at[c.cS];
    for (int i = 1 + gt; i <= m; i += stride) {
      bd[i] = i == r ? newbr : fma(col[i], delta, beta[i]);
      hd[i] = i == r ? kin : head[i];
    }
    for (int j = gt; j < D.nloc; j += stride) {
      nd[j] = j == qloc ? kout : ns[j];
      ad[j] = j == qloc ? newat : as[j];
    }
  }
  if (r) {  // K3: rho_q = 1/p, rho_j = -(T_rj * (1/p)) on the current row r (through the view)
    const ShardView V = shard_view(D, c, par);
    const double inv = 1.0 / col[r];
    double *rho = D.rho[par];
    for (int j = gt; j < D.ld; j += stride) rho[j] = j == qloc ? inv : -(V.at(r, j) * inv);
    stp.valid = 1;
    stp.r = r;
    stp.qloc = qloc;
    stp.wk = wk;
    stp.src = c.cur;
    n.cur = c.cur ^ 1;
    n.pend = 1;
  }
  n.cS = c.cS ^ 1;
  if (phase1) {
    // leave phase 1 as soon as the NEXT basis is primal feasible (the loop head of solve_lp: mode =
    // PRIMAL2, degen = 0): every CTA counts the infeasible rows of the state it has just derived
    int cnt = 0;
    for (int i = 1 + tid; i <= m; i += NT) {
      const double bn = i == r ? newbr : fma(col[i], delta, beta[i]);
      const int hv = i == r ? kin : head[i];
      cnt += row_infeas(bn, D.lb[hv], D.ub[hv]) != 0;
    }
    cnt = block_sum(cnt, sm.redI);
    if (cnt == 0) {
      n.mode = MVB_MODE_PRIMAL2;
      n.degen = 0;
    }
  }
  if (lead) {
    unsigned long long h = c.hash;
    h = (h ^ (unsigned long long)c.mode) * MVB_FNV_PRIME;
    h = (h ^ (unsigned long long)r) * MVB_FNV_PRIME;
    h = (h ^ (unsigned long long)cand.pos) * MVB_FNV_PRIME;
    n.hash = h;
    if (D.pivlog && c.iters < D.pivlog_cap) {
      D.pivlog[2 * c.iters] = r;
      D.pivlog[2 * c.iters + 1] = cand.pos;
    }
    n.iters = c.iters + 1;
    if (phase1) n.it1++;
    if (r) n.pivots++; else n.flips++;
    if (bland) n.bland_iters++;
    *next = n;
    D.step[par] = stp;
  }
}

// K4 of iteration t on the update stream: tableau src -> src ^ 1 with the step
// the decide kernel left (pivot column in the exchange record, rho in D.rho)
__global__ void __launch_bounds__(512, 1) shard_update_kernel(ShardDev D, int par) {
  __shared__ Smem sm;
  const ShardStep st = D.step[par];
  if (!st.valid) return;
  const double *col = reinterpret_cast<const double *>(D.recv[st.slot] + (size_t)st.wk * D.rec_bytes + SHARD_HDR);
  shard_tile_update(D, D.T[st.src], D.T[st.src ^ 1], st.r, st.qloc, D.rho[par], col, sm);
}

// ---- dual simplex on the column shards (B&B children / covering rows: the slack basis is dual
// feasible but not primal feasible).  Same pipeline and the same records as the primal pair:
//   shard_price_dual_kernel : K5a leaving row on the replicated beta (every rank finds the same r),
//                             K5b ratio test |d_j| / |T_rj| over the LOCAL columns of row r (through
//                             the view), record = (ratio bits, id, position, flag, bounds) + column
//   shard_decide_dual_kernel: same global winner everywhere by (ratio bits, variable id), step
//                             delta = (target - beta_r) / T_rq, next small state, K3 scaled row.
// Arithmetic and tie-breaks are those of the general loop (solve_lp, MVB_MODE_DUAL) and the oracle.

// K5a: largest bound violation, lowest variable id on ties (Bland: lowest id); idx = r or -1
__device__ Key shard_dual_leaving(const ShardDev &D, const ShardCtl &c, Key *red) {
  const bool bland = c.degen >= MVB_BLAND_AFTER;
  const double *beta = D.beta[c.cS];
  const int *head = D.head[c.cS];
  Key best = key_none();
  for (int i = 1 + threadIdx.x; i <= D.m; i += blockDim.x) {
    const int var = head[i];
    const double b = beta[i], lo = D.lb[var], up = D.ub[var];
    const int sg = row_infeas(b, lo, up);
    if (!sg) continue;
    const double viol = sg < 0 ? lo - b : b - up;
    const Key k{bland ? 0ULL : ~dbits(viol), var, i};
    if (key_less(k, best)) best = k;
  }
  return block_argmin(best, red);
}

__global__ void __launch_bounds__(1024, 1) shard_price_dual_kernel(ShardDev D, int par, int slot) {
  __shared__ Key red[32];
  const ShardCtl c = D.ctl[par];
  if (c.status != -1) return;
  const ShardView V = shard_view(D, c, par);
  const int *at = D.at[c.cS], *nb = D.nb[c.cS];
  const Key lv = shard_dual_leaving(D, c, red);
  Key best = key_none();
  if (lv.idx >= 0) {
    const int r = lv.idx;
    const int rs = row_infeas(D.beta[c.cS][r], D.lb[lv.var], D.ub[lv.var]);
    for (int j = threadIdx.x; j < D.nloc; j += blockDim.x) {
      const int f = at[j];
      if (f == MVB_AT_FX) continue;
      const double alpha = V.at(r, j);
      const double a = rs < 0 ? alpha : -alpha;
      const bool ok = (f == MVB_AT_LB && a > MVB_TOL_PIV) || (f == MVB_AT_UB && a < -MVB_TOL_PIV) ||
                      (f == MVB_AT_FR && (a > MVB_TOL_PIV || a < -MVB_TOL_PIV));
      if (!ok) continue;
      const double ratio = fabs(V.at(0, j)) / fabs(alpha);
      const Key k{dbits(ratio), nb[j], j};
      if (key_less(k, best)) best = k;
    }
  }
  best = block_argmin(best, red);  // primal feasible basis (no r): an empty record from every rank
  shard_emit_record(D, V, slot, best, at, best.idx >= 0 ? V.at(0, best.idx) : 0.0);
}

__global__ void __launch_bounds__(512, 1) shard_decide_dual_kernel(ShardDev D, int par, int slot,
                                                                   unsigned long long expect) {
  __shared__ Smem sm;
  __shared__ int arrived;
  const ShardCtl c = D.ctl[par];
  ShardCtl *next = &D.ctl[par ^ 1];
  const bool lead = blockIdx.x == 0 && threadIdx.x == 0;
  ShardCtl n = c;
  n.pend = 0;
  ShardStep stp{};
  stp.slot = slot;
  if (c.status != -1) {
    if (lead) { *next = n; D.step[par] = stp; }
    return;
  }
  if (D.push && !shard_wait_records(D, slot, expect, &arrived)) {
    n.status = SHARD_ST_TIMEOUT;
    if (lead) { *next = n; D.step[par] = stp; }
    return;
  }
  const int m = D.m, tid = threadIdx.x, NT = blockDim.x;
  const Key lv = shard_dual_leaving(D, c, sm.redK);
  if (lv.idx < 0) {
    n.status = MVB_OPT;
    if (lead) { *next = n; D.step[par] = stp; }
    return;
  }
  if (c.iters >= D.max_iters) {
    n.status = MVB_ITLIM;
    if (lead) { *next = n; D.step[par] = stp; }
    return;
  }
  // same winner everywhere: minimum (ratio bits, variable id) over the ranks' candidates
  int wk = -1;
  unsigned long long bk = ~0ULL;
  int bvar = 0x7fffffff;
  for (int k = 0; k < D.world; k++) {
    const Cand *h = reinterpret_cast<const Cand *>(D.recv[slot] + (size_t)k * D.rec_bytes);
    if (h->pos < 0) continue;
    if (wk < 0 || h->key < bk || (h->key == bk && h->var < bvar)) { bk = h->key; bvar = h->var; wk = k; }
  }
  if (wk < 0) {
    n.status = MVB_NOFEAS;
    if (lead) { *next = n; D.step[par] = stp; }
    return;
  }
  const Cand cand = *reinterpret_cast<const Cand *>(D.recv[slot] + (size_t)wk * D.rec_bytes);
  const double *col = reinterpret_cast<const double *>(D.recv[slot] + (size_t)wk * D.rec_bytes + SHARD_HDR);
  const bool bland = c.degen >= MVB_BLAND_AFTER;
  const double *beta = D.beta[c.cS];
  const int *head = D.head[c.cS];
  const int r = lv.idx, kout = lv.var, kin = cand.var;
  const double br = beta[r], lr = D.lb[kout], ur = D.ub[kout];
  const int rs = row_infeas(br, lr, ur);
  const double qratio = __longlong_as_double((long long)cand.key);
  const double target = rs < 0 ? lr : ur;
  const double delta = (target - br) / col[r];
  const double newbr = nb_value(cand.flag, cand.lq, cand.uq) + delta;
  const int newat = lr == ur ? MVB_AT_FX : (rs < 0 ? MVB_AT_LB : MVB_AT_UB);
  n.degen = qratio == 0.0 ? c.degen + 1 : 0;
  const int qloc = cand.pos - D.col0;
  const int gt = blockIdx.x * NT + tid, stride = gridDim.x * NT;
  {
    double *bd = D.beta[c.cS ^ 1];
    int *hd = D.head[c.cS ^ 1], *nd = D.nb[c.cS ^ 1], *ad = D.at[c.cS ^ 1];
    const int *ns = D.nb[c.cS], *as = D.at[c.cS];
    for (int i = 1 + gt; i <= m; i += stride) {
      bd[i] = i == r ? newbr : fma(col[i], delta, beta[i]);
      hd[i] = i == r ? kin : head[i];
    }
    for (int j = gt; j < D.nloc; j += stride) {
      nd[j] = j == qloc ? kout : ns[j];
      ad[j] = j == qloc ? newat : as[j];
    }
  }
  {  // K3 on the current row r (through the view)
    const ShardView V = shard_view(D, c, par);
    const double inv = 1.0 / col[r];
    double *rho = D.rho[par];
    for (int j = gt; j < D.ld; j += stride) rho[j] = j == qloc ? inv : -(V.at(r, j) * inv);
    stp.valid = 1;
    stp.r = r;
    stp.qloc = qloc;
    stp.wk = wk;
    stp.src = c.cur;
    n.cur = c.cur ^ 1;
    n.pend = 1;
  }
  n.cS = c.cS ^ 1;
  if (lead) {
    unsigned long long h = c.hash;
    h = (h ^ (unsigned long long)MVB_MODE_DUAL) * MVB_FNV_PRIME;
    h = (h ^ (unsigned long long)r) * MVB_FNV_PRIME;
    h = (h ^ (unsigned long long)cand.pos) * MVB_FNV_PRIME;
    n.hash = h;
    if (D.pivlog && c.iters < D.pivlog_cap) {
      D.pivlog[2 * c.iters] = r;
      D.pivlog[2 * c.iters + 1] = cand.pos;
    }
    n.iters = c.iters + 1;
    n.pivots++;
    if (bland) n.bland_iters++;
    *next = n;
    D.step[par] = stp;
  }
}

}  // namespace

// ------------------------------------------------------------------ host
namespace {

// NCCL through dlopen: prefer the copy the process has already loaded (torch's),
// then $MVOLPS_NCCL_LIB, then the system library
struct NcclId { char internal[MVB_NCCL_ID_BYTES]; };
struct NcclApi {
  void *lib = nullptr;
  int (*GetUniqueId)(void *) = nullptr;
  int (*CommInitRank)(void **, int, NcclId /* ncclUniqueId, by value */, int) = nullptr;
  int (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  int (*CommDestroy)(void *) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
  int (*Send)(const void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  int (*Recv)(void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*CommCount)(void *, int *) = nullptr;
};
NcclApi g_nccl;

bool nccl_load() {
  if (g_nccl.lib) return true;
  void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!h && getenv("MVOLPS_NCCL_LIB")) h = dlopen(getenv("MVOLPS_NCCL_LIB"), RTLD_NOW);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW);
  if (!h) return fail(std::string("cannot load NCCL: ") + dlerror());
  g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(h, "ncclGetUniqueId");
  g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(h, "ncclCommInitRank");
  g_nccl.AllGather = (decltype(g_nccl.AllGather))dlsym(h, "ncclAllGather");
  g_nccl.AllReduce = (decltype(g_nccl.AllReduce))dlsym(h, "ncclAllReduce");
  g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(h, "ncclCommDestroy");
  g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(h, "ncclGetErrorString");
  g_nccl.Send = (decltype(g_nccl.Send))dlsym(h, "ncclSend");
  g_nccl.Recv = (decltype(g_nccl.Recv))dlsym(h, "ncclRecv");
  g_nccl.GroupStart = (decltype(g_nccl.GroupStart))dlsym(h, "ncclGroupStart");
  g_nccl.GroupEnd = (decltype(g_nccl.GroupEnd))dlsym(h, "ncclGroupEnd");
  g_nccl.CommCount = (decltype(g_nccl.CommCount))dlsym(h, "ncclCommCount");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllGather || !g_nccl.AllReduce ||
      !g_nccl.CommDestroy)
    return fail("NCCL library lacks a required symbol");
  g_nccl.lib = h;
  return true;
}

struct NcclCtx {
  void *comm = nullptr;
  double *dbuf = nullptr;  // device scratch for the small host reductions
  char *xbuf = nullptr;    // device staging of the node hand-over (grown on demand)
  size_t xcap = 0;
};

int nccl_allgather(void *ctx, const void *send, void *recv, size_t bytes, void *stream) {
  NcclCtx *c = (NcclCtx *)ctx;
  const int rc = g_nccl.AllGather(send, recv, bytes, /*ncclChar*/ 0, c->comm, (cudaStream_t)stream);
  if (rc) { g_err = std::string("ncclAllGather: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error"); return -1; }
  return 0;
}
int nccl_allreduce_max(void *ctx, double *vals, int count) {
  NcclCtx *c = (NcclCtx *)ctx;
  if (count > 64) { g_err = "allreduce_max: at most 64 values"; return -1; }
  cudaStream_t st = g_ctx.stream;
  if (cudaMemcpyAsync(c->dbuf, vals, count * sizeof(double), cudaMemcpyHostToDevice, st) != cudaSuccess) return -1;
  const int rc = g_nccl.AllReduce(c->dbuf, c->dbuf, count, /*ncclDouble*/ 8, /*ncclMax*/ 2, c->comm, st);
  if (rc) { g_err = "ncclAllReduce failed"; return -1; }
  if (cudaMemcpyAsync(vals, c->dbuf, count * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess) return -1;
  return cudaStreamSynchronize(st) == cudaSuccess ? 0 : -1;
}
// node hand-over between the two ranks of a pair: host buffer -> device staging -> ncclSend /
// ncclRecv over NVLink -> host buffer (a hand-over is a few hundred KB once in a while)
int nccl_sendrecv_host(void *ctx, const void *send, size_t sbytes, int dst, void *recv, size_t rbytes, int src) {
  NcclCtx *c = (NcclCtx *)ctx;
  const size_t need = std::max(sbytes, rbytes);
  if (need > c->xcap) {
    if (c->xbuf) cudaFree(c->xbuf);
    c->xcap = 0;
    if (cudaMalloc((void **)&c->xbuf, need * 2) != cudaSuccess) { g_err = "sendrecv_host: cudaMalloc failed"; return -1; }
    c->xcap = need * 2;
  }
  cudaStream_t st = g_ctx.stream;
  if (dst >= 0 && src >= 0) { g_err = "sendrecv_host: a rank either sends or receives"; return -1; }
  if (dst >= 0) {
    if (cudaMemcpyAsync(c->xbuf, send, sbytes, cudaMemcpyHostToDevice, st) != cudaSuccess) return -1;
    if (g_nccl.Send(c->xbuf, sbytes, /*ncclChar*/ 0, dst, c->comm, st)) { g_err = "ncclSend failed"; return -1; }
  }
  if (src >= 0) {
    if (g_nccl.Recv(c->xbuf, rbytes, /*ncclChar*/ 0, src, c->comm, st)) { g_err = "ncclRecv failed"; return -1; }
    if (cudaMemcpyAsync(recv, c->xbuf, rbytes, cudaMemcpyDeviceToHost, st) != cudaSuccess) return -1;
  }
  return cudaStreamSynchronize(st) == cudaSuccess ? 0 : -1;
}
// result exchange of the round-synchronous tree: host buffers through device staging and
// ncclAllGather over NVLink (a round moves a few KB per solved LP)
int nccl_allgather_host(void *ctx, const void *send, void *recv, size_t bytes) {
  NcclCtx *c = (NcclCtx *)ctx;
  int world = 0;
  if (!g_nccl.CommCount || g_nccl.CommCount(c->comm, &world)) { g_err = "allgather_host: ncclCommCount failed"; return -1; }
  const size_t need = bytes * (size_t)(world + 1);
  if (need > c->xcap) {
    if (c->xbuf) cudaFree(c->xbuf);
    c->xcap = 0;
    if (cudaMalloc((void **)&c->xbuf, need * 2) != cudaSuccess) { g_err = "allgather_host: cudaMalloc failed"; return -1; }
    c->xcap = need * 2;
  }
  cudaStream_t st = g_ctx.stream;
  if (cudaMemcpyAsync(c->xbuf, send, bytes, cudaMemcpyHostToDevice, st) != cudaSuccess) return -1;
  if (g_nccl.AllGather(c->xbuf, c->xbuf + bytes, bytes, /*ncclChar*/ 0, c->comm, st)) { g_err = "ncclAllGather failed"; return -1; }
  if (cudaMemcpyAsync(recv, c->xbuf + bytes, bytes * world, cudaMemcpyDeviceToHost, st) != cudaSuccess) return -1;
  return cudaStreamSynchronize(st) == cudaSuccess ? 0 : -1;
}
void nccl_destroy(void *ctx) {
  NcclCtx *c = (NcclCtx *)ctx;
  if (c->comm) g_nccl.CommDestroy(c->comm);
  if (c->dbuf) cudaFree(c->dbuf);
  if (c->xbuf) cudaFree(c->xbuf);
  delete c;
}

}  // namespace

extern "C" {

int mvb_nccl_unique_id(void *id) {
  if (!nccl_load()) return -1;
  return g_nccl.GetUniqueId(id) ? (g_err = "ncclGetUniqueId failed", -1) : 0;
}

mvb_comm *mvb_comm_nccl(const void *id, int rank, int world) {
  if (!ctx_init(-1) || !nccl_load()) return nullptr;
  NcclId nid;
  std::memcpy(nid.internal, id, MVB_NCCL_ID_BYTES);
  NcclCtx *c = new NcclCtx();
  const int rc = g_nccl.CommInitRank(&c->comm, world, nid, rank);
  if (rc || cudaMalloc((void **)&c->dbuf, 64 * sizeof(double)) != cudaSuccess) {
    g_err = "ncclCommInitRank failed";
    delete c;
    return nullptr;
  }
  mvb_comm *m = new mvb_comm();
  m->ctx = c;
  m->rank = rank;
  m->world = world;
  m->allgather_dev = nccl_allgather;
  m->allreduce_max = nccl_allreduce_max;
  m->destroy = nccl_destroy;
  m->sendrecv_host = (g_nccl.Send && g_nccl.Recv) ? nccl_sendrecv_host : nullptr;
  m->allgather_host = nccl_allgather_host;
  return m;
}

// member of a single-process group of `world` shards on one GPU (tests, emulation):
// mvb_shard_solve then takes all shards in one call and copies device-to-device
mvb_comm *mvb_comm_local(int rank, int world) {
  mvb_comm *m = new mvb_comm();
  m->ctx = nullptr;
  m->rank = rank;
  m->world = world;
  m->allgather_dev = nullptr;
  m->allreduce_max = nullptr;
  m->destroy = nullptr;
  m->sendrecv_host = nullptr;
  m->allgather_host = nullptr;
  return m;
}

void mvb_comm_free(mvb_comm *c) {
  if (!c) return;
  if (c->destroy) c->destroy(c->ctx);
  delete c;
}

// incumbent / bound broadcast of a sharded tree: two max-reductions of one double
int mvb_incumbent_exchange(const mvb_comm *c, double *incumbent, int *owner) {
  if (!c || !c->allreduce_max) { g_err = "mvb_incumbent_exchange: communicator has no allreduce_max"; return -1; }
  const double mine = *incumbent;
  double v = mine;
  if (c->allreduce_max(c->ctx, &v, 1)) return -1;
  double who = (mine == v) ? -(double)c->rank : -HUGE_VAL;  // lowest rank holding the maximum
  if (c->allreduce_max(c->ctx, &who, 1)) return -1;
  *incumbent = v;
  if (owner) *owner = (int)(-who);
  return 0;
}

void mvb_shard_partition(int n, int world, int rank, int *col0, int *ncols) {
  int per = (n + world - 1) / world;
  per = (per + MVB_LD_ALIGN - 1) / MVB_LD_ALIGN * MVB_LD_ALIGN;
  int a = std::min(n, rank * per), b = std::min(n, (rank + 1) * per);
  if (col0) *col0 = a;
  if (ncols) *ncols = b - a;
}

}  // extern "C"

struct mvb_shard {
  int m = 0, nglob = 0, col0 = 0, nloc = 0, ld = 0, maxloc = 0;
  const mvb_comm *comm = nullptr;
  int world = 1, rank = 0;
  char *slab = nullptr;
  ShardDev dev{};
  size_t fin_bytes = 0;            // per-rank record of the final gather (nb | at | cost)
  unsigned char *fin_send = nullptr, *fin_recv = nullptr;
  double *cost = nullptr;          // local costs (device)
  std::vector<double> lb, ub;      // replicated bounds (host)
  std::vector<double> x;           // primal by id after the last solve
  std::vector<int> bhead, bnb, bat; // basis after the last solve: head[m], nb / at by GLOBAL position (n_global)
  int *pivlog = nullptr; int pivlog_cap = 0, pivlog_n = 0;
  cudaEvent_t ev[2] = {};
  cudaEvent_t ev_dec[2] = {}, ev_upd[2] = {};  // by iteration parity: decide done / update done
  float last_ms = 0.f;
  int last_exchange = 0;
  unsigned char *mail = nullptr;   // in-slab mailbox
  size_t mail_flag_off = 0;
  int start_mode = MVB_MODE_PRIMAL2;  // from the slack basis: primal phase 2, or the dual simplex
  bool dual_ok_local = true;          // dual start: no local column prices (the ranks agree at solve time)
};

namespace {

// records land in the shard's own slab (host copies, ncclAllGather, or pushes inside a single-process group)
void shard_use_slab_mailbox(mvb_shard *s) {
  ShardDev &d = s->dev;
  for (int k = 0; k < SHARD_SLOTS; k++) d.recv[k] = s->mail + (size_t)k * s->world * d.rec_bytes;
  d.flag = (unsigned long long *)(s->mail + s->mail_flag_off);
  d.flag_off = s->mail_flag_off;
  d.push = 0;
  for (int k = 0; k < SHARD_MAX_PEERS; k++) d.peer[k] = nullptr;
}

bool shard_alloc(mvb_shard *s) {
  const int m = s->m, nloc = s->nloc;
  s->ld = (int)align_up((size_t)std::max(nloc, 1), MVB_LD_ALIGN);
  int per = (s->nglob + s->world - 1) / s->world;
  s->maxloc = (per + MVB_LD_ALIGN - 1) / MVB_LD_ALIGN * MVB_LD_ALIGN;
  const size_t A = 256;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, A); return o; };
  const size_t tb = (size_t)(m + 1) * s->ld * sizeof(double);
  const size_t rec = align_up(SHARD_HDR + (size_t)(m + 1) * sizeof(double), 16);
  s->fin_bytes = align_up((size_t)s->maxloc * (2 * sizeof(int) + sizeof(double)), 16);
  size_t oT0 = take(tb), oT1 = take(tb);
  size_t ob0 = take((m + 1) * sizeof(double)), ob1 = take((m + 1) * sizeof(double));
  size_t oh0 = take((m + 1) * sizeof(int)), oh1 = take((m + 1) * sizeof(int));
  size_t on0 = take((size_t)s->ld * sizeof(int)), on1 = take((size_t)s->ld * sizeof(int));
  size_t oa0 = take((size_t)s->ld * sizeof(int)), oa1 = take((size_t)s->ld * sizeof(int));
  size_t olb = take((size_t)(s->nglob + m) * sizeof(double)), oub = take((size_t)(s->nglob + m) * sizeof(double));
  size_t oc = take((size_t)s->ld * sizeof(double));
  size_t octl = take(2 * sizeof(ShardCtl)), ostep = take(2 * sizeof(ShardStep));
  size_t orho0 = take((size_t)s->ld * sizeof(double)), orho1 = take((size_t)s->ld * sizeof(double));
  const size_t p1chunks = ((size_t)m + MVB_P1_CHUNK - 1) / MVB_P1_CHUNK;
  size_t op1 = take(p1chunks * s->ld * sizeof(double)), od1 = take((size_t)s->ld * sizeof(double));
  // in-slab mailbox, same layout as the IPC one (PeerMail): SHARD_SLOTS x world records, then the counters
  const size_t mail_flag_off = align_up((size_t)SHARD_SLOTS * s->world * rec, 256);
  size_t osend = take(rec), omail = take(mail_flag_off + (size_t)SHARD_SLOTS * s->world * sizeof(unsigned long long));
  size_t ofs = take(s->fin_bytes), ofr = take(s->fin_bytes * s->world);
  CK(cudaMallocAsync((void **)&s->slab, off, g_ctx.stream));
  CK(cudaMemsetAsync(s->slab, 0, off, g_ctx.stream));
  char *b = s->slab;
  ShardDev &d = s->dev;
  d.m = m; d.nloc = nloc; d.ld = s->ld; d.col0 = s->col0; d.world = s->world;
  d.T[0] = (double *)(b + oT0); d.T[1] = (double *)(b + oT1);
  d.beta[0] = (double *)(b + ob0); d.beta[1] = (double *)(b + ob1);
  d.head[0] = (int *)(b + oh0); d.head[1] = (int *)(b + oh1);
  d.nb[0] = (int *)(b + on0); d.nb[1] = (int *)(b + on1);
  d.at[0] = (int *)(b + oa0); d.at[1] = (int *)(b + oa1);
  d.lb = (double *)(b + olb); d.ub = (double *)(b + oub);
  s->cost = (double *)(b + oc);
  d.ctl = (ShardCtl *)(b + octl);
  d.step = (ShardStep *)(b + ostep);
  d.rho[0] = (double *)(b + orho0); d.rho[1] = (double *)(b + orho1);
  d.p1part = (double *)(b + op1); d.d1 = (double *)(b + od1);
  d.send = (unsigned char *)(b + osend);
  s->mail = (unsigned char *)(b + omail);
  s->mail_flag_off = mail_flag_off;
  d.rec_bytes = rec;
  d.rank = s->rank;
  shard_use_slab_mailbox(s);
  s->fin_send = (unsigned char *)(b + ofs); s->fin_recv = (unsigned char *)(b + ofr);
  return true;
}

// ---- peer mailbox: one per process, exported through CUDA IPC, mapped by every peer ----
// layout (identical on every rank): SHARD_SLOTS x world records, then SHARD_SLOTS x world counters
struct PeerMail {
  unsigned char *base = nullptr;
  size_t rec = 0, flag_off = 0, bytes = 0;
  int world = 0;
  unsigned char *peer[SHARD_MAX_PEERS] = {};
  bool failed = false;  // IPC is not available here: every rank stays on ncclAllGather
};
PeerMail g_mail;

// collective over the ranks of s->comm; 1: s->dev points into the mailboxes, 0: no peer path, -1: error
int shard_peer_setup(mvb_shard *s) {
  const mvb_comm *cm = s->comm;
  const char *mode = getenv("MVOLPS_SHARD_EXCHANGE");
  if (mode && !strcmp(mode, "nccl")) return 0;
  if (s->world < 2 || s->world > SHARD_MAX_PEERS || !cm->allgather_dev || !cm->allreduce_max) return 0;
  if (g_mail.failed) return 0;
  cudaStream_t st = g_ctx.stream;
  const size_t rec = s->dev.rec_bytes;
  if (!g_mail.base || g_mail.rec != rec || g_mail.world != s->world) {
    // (a mailbox of another shape stays mapped until the process ends: peers may hold it open)
    PeerMail nm;
    nm.rec = rec;
    nm.world = s->world;
    nm.flag_off = align_up((size_t)SHARD_SLOTS * s->world * rec, 256);
    nm.bytes = std::max<size_t>(nm.flag_off + (size_t)SHARD_SLOTS * s->world * sizeof(unsigned long long), 1 << 16);
    double bad = 0.0;
    cudaIpcMemHandle_t mine{};
    std::vector<cudaIpcMemHandle_t> all(s->world);
    if (cudaMalloc((void **)&nm.base, nm.bytes) != cudaSuccess || cudaIpcGetMemHandle(&mine, nm.base) != cudaSuccess) {
      cudaGetLastError();
      bad = 1.0;
    }
    if (nm.base) {  // the handles travel through the mailbox itself (not in use yet)
      if (cudaMemcpyAsync(nm.base, &mine, sizeof(mine), cudaMemcpyHostToDevice, st) != cudaSuccess) bad = 1.0;
      // (the collective runs even after a local failure: every rank must take part; `bad` is agreed below)
      if (cm->allgather_dev(cm->ctx, nm.base, nm.base + 4096, sizeof(mine), st)) { g_err = "peer set-up: all-gather failed"; return -1; }
      if (cudaMemcpyAsync(all.data(), nm.base + 4096, sizeof(mine) * s->world, cudaMemcpyDeviceToHost, st) != cudaSuccess) bad = 1.0;
      if (cudaStreamSynchronize(st) != cudaSuccess) bad = 1.0;
      if (bad != 0.0) cudaGetLastError();
    }
    if (bad == 0.0)
      for (int k = 0; k < s->world; k++) {
        if (k == s->rank) { nm.peer[k] = nm.base; continue; }
        void *p = nullptr;
        if (cudaIpcOpenMemHandle(&p, all[k], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
          cudaGetLastError();
          bad = 1.0;
          break;
        }
        nm.peer[k] = (unsigned char *)p;
      }
    if (cm->allreduce_max(cm->ctx, &bad, 1)) { g_err = "peer set-up: reduction failed"; return -1; }
    if (bad != 0.0) {
      if (getenv("MVOLPS_VERBOSE")) std::fprintf(stderr, "mvolps: CUDA IPC unavailable, exchange stays on ncclAllGather\n");
      g_mail.failed = true;
      return 0;
    }
    g_mail = nm;
  }
  ShardDev &d = s->dev;
  for (int k = 0; k < SHARD_SLOTS; k++) d.recv[k] = g_mail.base + (size_t)k * s->world * rec;
  for (int k = 0; k < s->world; k++) d.peer[k] = g_mail.peer[k];
  d.flag = (unsigned long long *)(g_mail.base + g_mail.flag_off);
  d.flag_off = g_mail.flag_off;
  d.push = 1;
  // counters back to zero, and nobody pushes before every rank has done so: the (stream-ordered)
  // all-gather below completes only after every rank has enqueued its memset
  const bool zeroed = cudaMemsetAsync(d.flag, 0, (size_t)SHARD_SLOTS * s->world * sizeof(unsigned long long), st) == cudaSuccess;
  if (cm->allgather_dev(cm->ctx, s->fin_send, s->fin_recv, 16, st)) { g_err = "peer set-up: barrier failed"; return -1; }
  if (!zeroed) { g_err = "peer set-up: cannot reset the arrival counters"; return -1; }
  return 1;
}

// final gather record: nb[maxloc] | at[maxloc] | cost[maxloc]
__global__ void shard_pack_final(ShardDev D, int par, const double *cost, int maxloc, unsigned char *out) {
  const ShardCtl c = D.ctl[par];
  int *nb = (int *)out, *at = nb + maxloc;
  double *co = (double *)(at + maxloc);
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < maxloc; j += gridDim.x * blockDim.x) {
    const bool in = j < D.nloc;
    nb[j] = in ? D.nb[c.cS][j] : -1;
    at[j] = in ? D.at[c.cS][j] : 0;
    co[j] = in ? cost[j] : 0.0;
  }
}

}  // namespace

extern "C" {

mvb_shard *mvb_shard_create(int m, int n_global, int col0, int ncols, const mvb_comm *comm) {
  if (m < 1 || n_global < 1 || col0 < 0 || ncols < 0 || col0 + ncols > n_global) {
    g_err = "mvb_shard_create: bad dimensions";
    return nullptr;
  }
  if (!ctx_init(-1)) return nullptr;
  mvb_shard *s = new mvb_shard();
  if (!comm) { g_err = "mvb_shard_create: a communicator is required (mvb_comm_nccl / mvb_comm_local)"; delete s; return nullptr; }
  s->m = m; s->nglob = n_global; s->col0 = col0; s->nloc = ncols; s->comm = comm;
  s->world = comm->world; s->rank = comm->rank;
  return s;  // device memory is allocated at load
}

void mvb_shard_free(mvb_shard *s) {
  if (!s) return;
  if (s->slab) cudaFreeAsync(s->slab, g_ctx.stream);
  if (s->pivlog) cudaFreeAsync(s->pivlog, g_ctx.stream);
  if (s->ev[0]) { cudaEventDestroy(s->ev[0]); cudaEventDestroy(s->ev[1]); }
  for (int k = 0; k < 2; k++) {
    if (s->ev_dec[k]) cudaEventDestroy(s->ev_dec[k]);
    if (s->ev_upd[k]) cudaEventDestroy(s->ev_upd[k]);
  }
  delete s;
}

// rows of the slack basis with every non-basic at value 0 (beta = 0), by the rule of the general
// loop (row_infeas): 1 = primal feasible (primal phase 2), 2 = some row violates its bound beyond
// the tolerance (the dual simplex if no column prices, else the composite phase 1)
static int shard_row_start(int m, const double *rlb, const double *rub) {
  for (int i = 0; i < m; i++) {
    const double l = rlb[i], u = rub[i];
    if (l - 0.0 > MVB_TOL_BND * (1.0 + std::fabs(l)) || 0.0 - u > MVB_TOL_BND * (1.0 + std::fabs(u))) return 2;
  }
  return 1;
}
// no column [0, n) prices at its slack-basis flag: the slack basis is dual feasible
static bool shard_cols_dual_ok(int n, const double *cost, const double *clb, const double *cub) {
  for (int j = 0; j < n; j++) {
    const int f = nb_flag_for(clb[j], cub[j], MVB_AT_LB);
    const double dj = cost[j];
    const bool elig = f == MVB_AT_LB ? dj < -MVB_TOL_DJ : f == MVB_AT_UB ? dj > MVB_TOL_DJ
                      : f == MVB_AT_FR ? (dj < -MVB_TOL_DJ || dj > MVB_TOL_DJ) : false;
    if (elig) return false;
  }
  return true;
}

int mvb_shard_start_mode(int m, int n, const double *cost, const double *lb, const double *ub) {
  for (int j = 0; j < n; j++) {
    const double l = lb[j], u = ub[j];
    const bool zero_ok = (l == 0.0) || (!(l > -HUGE_VAL) && (u == 0.0 || !(u < HUGE_VAL)));
    if (!zero_ok) return 0;
  }
  if (shard_row_start(m, lb + n, ub + n) == 1) return MVB_MODE_PRIMAL2;
  return shard_cols_dual_ok(n, cost, lb, ub) ? MVB_MODE_DUAL : MVB_MODE_PRIMAL1;
}

int mvb_shard_load(mvb_shard *s, const double *A_local, const double *cost_local, const double *lb,
                   const double *ub) {
  const int m = s->m, nloc = s->nloc, ng = s->nglob;
  cudaStream_t st = g_ctx.stream;
  // this path starts from the slack basis with every non-basic at value 0: primal phase 2 when
  // that basis is primal feasible, the dual simplex when it is dual feasible instead, the
  // composite phase 1 (then phase 2) when it is neither
  for (int j = 0; j < nloc; j++) {
    const double l = lb[s->col0 + j], u = ub[s->col0 + j];
    const bool zero_ok = (l == 0.0) || (!(l > -HUGE_VAL) && (u == 0.0 || !(u < HUGE_VAL)));
    if (!zero_ok) { g_err = "mvb_shard_load: sharded path needs non-basic values 0 (lb = 0 or free)"; return -1; }
  }
  const int rstart = shard_row_start(m, lb + ng, ub + ng);
  s->start_mode = rstart == 1 ? MVB_MODE_PRIMAL2 : MVB_MODE_DUAL;
  // (not primal feasible: the dual simplex if no column of ANY rank prices, else the composite
  // phase 1 -- agreed in mvb_shard_solve, where all ranks meet)
  s->dual_ok_local = rstart == 1 || shard_cols_dual_ok(nloc, cost_local, lb + s->col0, ub + s->col0);
  if (s->slab) { CKI(cudaFreeAsync(s->slab, st)); s->slab = nullptr; }
  if (!shard_alloc(s)) return -1;
  ShardDev &d = s->dev;
  if (nloc > 0) {
    CKI(cudaMemcpyAsync(d.T[0], cost_local, (size_t)nloc * sizeof(double), cudaMemcpyHostToDevice, st));
    CKI(cudaMemcpy2DAsync(d.T[0] + s->ld, (size_t)s->ld * sizeof(double), A_local, (size_t)nloc * sizeof(double),
                          (size_t)nloc * sizeof(double), m, cudaMemcpyHostToDevice, st));
    CKI(cudaMemcpyAsync(s->cost, cost_local, (size_t)nloc * sizeof(double), cudaMemcpyHostToDevice, st));
  }
  CKI(cudaMemcpyAsync(d.lb, lb, (size_t)(ng + m) * sizeof(double), cudaMemcpyHostToDevice, st));
  CKI(cudaMemcpyAsync(d.ub, ub, (size_t)(ng + m) * sizeof(double), cudaMemcpyHostToDevice, st));
  s->lb.assign(lb, lb + ng + m);
  s->ub.assign(ub, ub + ng + m);
  std::vector<int> head(m + 1, 0), nb(std::max(nloc, 1)), at(std::max(nloc, 1));
  for (int i = 1; i <= m; i++) head[i] = ng + i - 1;
  for (int j = 0; j < nloc; j++) {
    nb[j] = s->col0 + j;
    at[j] = nb_flag_for(lb[s->col0 + j], ub[s->col0 + j], MVB_AT_LB);
  }
  CKI(cudaMemcpyAsync(d.head[0], head.data(), (size_t)(m + 1) * sizeof(int), cudaMemcpyHostToDevice, st));
  if (nloc > 0) {
    CKI(cudaMemcpyAsync(d.nb[0], nb.data(), (size_t)nloc * sizeof(int), cudaMemcpyHostToDevice, st));
    CKI(cudaMemcpyAsync(d.at[0], at.data(), (size_t)nloc * sizeof(int), cudaMemcpyHostToDevice, st));
  }
  // beta = T v with v = 0: exactly +0.0 in every row (slab is zero-filled)
  CKI(cudaStreamSynchronize(st));
  s->x.assign(ng + m, 0.0);
  return 0;
}

int mvb_shard_get_primal(mvb_shard *s, double *x) {
  if ((int)s->x.size() != s->nglob + s->m) { g_err = "mvb_shard_get_primal: no solution"; return -1; }
  std::memcpy(x, s->x.data(), s->x.size() * sizeof(double));
  return 0;
}

int mvb_shard_get_pivot_log(mvb_shard *s, int *rowcol, int cap) {
  int c = std::min(s->pivlog_n, cap);
  if (c > 0) {
    CKI(cudaMemcpyAsync(rowcol, s->pivlog, (size_t)c * 2 * sizeof(int), cudaMemcpyDeviceToHost, g_ctx.stream));
    CKI(cudaStreamSynchronize(g_ctx.stream));
  }
  return s->pivlog_n;
}

int mvb_shard_get_basis(mvb_shard *s, int *head, int *nb, int *at) {
  if ((int)s->bhead.size() != s->m || (int)s->bnb.size() != s->nglob) { g_err = "mvb_shard_get_basis: no solution"; return -1; }
  std::memcpy(head, s->bhead.data(), (size_t)s->m * sizeof(int));
  std::memcpy(nb, s->bnb.data(), (size_t)s->nglob * sizeof(int));
  std::memcpy(at, s->bat.data(), (size_t)s->nglob * sizeof(int));
  return 0;
}

double mvb_shard_last_ms(mvb_shard *s) { return s->last_ms; }
int mvb_shard_last_exchange(mvb_shard *s) { return s->last_exchange; }

}  // extern "C"

extern "C" int mvb_shard_solve(mvb_shard *const *shards, int nlocal, const mvb_opts *opts,
                               mvb_result *out) {
  if (nlocal < 1) return 0;
  if (!ctx_init(-1)) return -1;
  cudaStream_t st = g_ctx.stream;
  mvb_shard *s0 = shards[0];
  const bool local = s0->comm->allgather_dev == nullptr;
  if (local && nlocal != s0->world) { g_err = "mvb_shard_solve: a local group needs all its shards"; return -1; }
  if (!local && nlocal != 1) { g_err = "mvb_shard_solve: one shard per process over NCCL"; return -1; }
  const int m = s0->m, ng = s0->nglob;
  const int mi = (opts && opts->max_iters > 0) ? opts->max_iters : 50 * (m + ng) + 10000;
  for (int k = 0; k < nlocal; k++)
    if (!shards[k]->slab) { g_err = "mvb_shard_solve: shard not loaded"; return -1; }
  int mode = s0->start_mode;
  if (mode == MVB_MODE_DUAL) {  // a column prices somewhere: neither primal nor dual feasible = composite phase 1
    double bad = 0.0;
    for (int k = 0; k < nlocal; k++) if (!shards[k]->dual_ok_local) bad = 1.0;
    if (!local) {
      if (!s0->comm->allreduce_max) { g_err = "mvb_shard_solve: the communicator has no reduction"; return -1; }
      if (s0->comm->allreduce_max(s0->comm->ctx, &bad, 1)) { g_err = "mvb_shard_solve: reduction failed"; return -1; }
    }
    if (bad != 0.0) mode = MVB_MODE_PRIMAL1;
  }
  const bool dual = mode == MVB_MODE_DUAL;
  bool phase1 = mode == MVB_MODE_PRIMAL1;  // the host stops launching the phase-1 kernels once it sees PRIMAL2
  ShardCtl init{};
  init.status = -1;
  init.mode = mode;
  init.hash = MVB_FNV_OFFSET;
  for (int k = 0; k < nlocal; k++) {
    mvb_shard *s = shards[k];
    if (opts && opts->log_pivots) {
      const int want = std::min(mi, 1 << 22);
      if (s->pivlog_cap < want) {
        if (s->pivlog) CKI(cudaFreeAsync(s->pivlog, st));
        CKI(cudaMallocAsync((void **)&s->pivlog, (size_t)want * 2 * sizeof(int), st));
        s->pivlog_cap = want;
      }
    }
    s->dev.max_iters = mi;
    s->dev.pivlog = (opts && opts->log_pivots) ? s->pivlog : nullptr;
    s->dev.pivlog_cap = s->dev.pivlog ? s->pivlog_cap : 0;
    CKI(cudaMemcpyAsync(&s->dev.ctl[0], &init, sizeof(init), cudaMemcpyHostToDevice, st));
    if (!s->ev[0]) { CKI(cudaEventCreate(&s->ev[0])); CKI(cudaEventCreate(&s->ev[1])); }
    for (int e = 0; e < 2; e++) {
      if (!s->ev_dec[e]) CKI(cudaEventCreateWithFlags(&s->ev_dec[e], cudaEventDisableTiming));
      if (!s->ev_upd[e]) CKI(cudaEventCreateWithFlags(&s->ev_upd[e], cudaEventDisableTiming));
    }
  }
  if (!g_ctx.stream_upd) CKI(cudaStreamCreateWithFlags(&g_ctx.stream_upd, cudaStreamNonBlocking));
  cudaStream_t su = g_ctx.stream_upd;  // the tableau updates; `st` carries the decision chain
  // one process per GPU: records are pushed into the peers' mailboxes by the price kernel itself
  for (int k = 0; k < nlocal; k++) shard_use_slab_mailbox(shards[k]);
  int pst = local ? 0 : shard_peer_setup(s0);
  if (pst < 0) return -1;
  if (local && nlocal <= SHARD_MAX_PEERS) {
    // single-process group (tests): MVOLPS_SHARD_EXCHANGE=push runs the same push / arrival-counter
    // kernels between the shards of one GPU (the price kernels of an iteration are all enqueued
    // before its decide kernels, so the waits cannot starve)
    const char *mode = getenv("MVOLPS_SHARD_EXCHANGE");
    if (mode && !strcmp(mode, "push")) {
      for (int k = 0; k < nlocal; k++) {
        ShardDev &d = shards[k]->dev;
        for (int t = 0; t < nlocal; t++) d.peer[shards[t]->rank] = shards[t]->mail;
        d.push = 1;
        CKI(cudaMemsetAsync(d.flag, 0, (size_t)SHARD_SLOTS * s0->world * sizeof(unsigned long long), st));
      }
      pst = 1;
    }
  }
  const bool push = pst == 1;
  for (int k = 0; k < nlocal; k++) shards[k]->last_exchange = push ? 2 : local ? 0 : 1;
  auto exchange = [&](bool fin, int slot) -> bool {
    if (push && !fin) return true;
    for (int k = 0; k < nlocal; k++) {
      mvb_shard *s = shards[k];
      const unsigned char *snd = fin ? s->fin_send : s->dev.send;
      const size_t bytes = fin ? s->fin_bytes : s->dev.rec_bytes;
      if (!local) {
        unsigned char *rcv = fin ? s->fin_recv : s->dev.recv[slot];
        if (s->comm->allgather_dev(s->comm->ctx, snd, rcv, bytes, st)) return false;
      } else {
        for (int t = 0; t < nlocal; t++) {
          unsigned char *rcv = (fin ? shards[t]->fin_recv : shards[t]->dev.recv[slot]) + (size_t)s->rank * bytes;
          CK(cudaMemcpyAsync(rcv, snd, bytes, cudaMemcpyDeviceToDevice, st));
        }
      }
    }
    return true;
  };
  // Two streams.  Decision chain of iteration t (st): price -> all-gather -> decide.  Update
  // stream (su): the rank-1 update of iteration t, after its decide.  The chain of t+1 reads
  // the buffer update t reads and applies step t on the fly (ShardView), so it needs update
  // t-1 finished, not update t: it overlaps the running update.  A few SMs are left to the
  // chain's small kernels (the update is HBM-bound long before it needs every SM).
  const int grid = std::max(1, g_ctx.max_ctas - 8);
  const int small_grid = 8;
  const int price_grid = std::max(1, std::min(small_grid, (m + 1 + 1023) / 1024));  // CTAs sharing the column gather
  const int CHUNK = 32;  // iterations enqueued between two looks at the control block
  int par = 0;
  long long t_it = 0;
  ShardCtl h{};
  CKI(cudaEventRecord(s0->ev[0], st));
  CKI(cudaStreamWaitEvent(su, s0->ev[0], 0));  // the update stream starts after the set-up on st
  while (true) {
    for (int it = 0; it < CHUNK; it++, t_it++) {
      // the view of iteration t is the output of update t-2 (+ step t-1 on the fly)
      if (t_it >= 2)
        for (int k = 0; k < nlocal; k++) CKI(cudaStreamWaitEvent(st, shards[k]->ev_upd[par], 0));
      const int slot = (int)(t_it % SHARD_SLOTS);
      const unsigned long long expect = (unsigned long long)price_grid * (unsigned long long)(t_it / SHARD_SLOTS + 1);
      if (phase1)
        for (int k = 0; k < nlocal; k++) {
          shard_p1_partials_kernel<<<grid, 256, 0, st>>>(shards[k]->dev, par);
          shard_p1_reduce_kernel<<<std::max(1, (shards[k]->nloc + 255) / 256), 256, 0, st>>>(shards[k]->dev, par);
          g_launches += 2;
        }
      for (int k = 0; k < nlocal; k++) {
        if (dual) shard_price_dual_kernel<<<price_grid, 1024, 0, st>>>(shards[k]->dev, par, slot);
        else shard_price_kernel<<<price_grid, 1024, 0, st>>>(shards[k]->dev, par, slot);
      }
      if (!exchange(false, slot)) return -1;
      for (int k = 0; k < nlocal; k++) {
        if (dual) shard_decide_dual_kernel<<<small_grid, 512, 0, st>>>(shards[k]->dev, par, slot, expect);
        else shard_decide_kernel<<<small_grid, 512, 0, st>>>(shards[k]->dev, par, slot, expect);
        CKI(cudaEventRecord(shards[k]->ev_dec[par], st));
      }
      for (int k = 0; k < nlocal; k++) {
        CKI(cudaStreamWaitEvent(su, shards[k]->ev_dec[par], 0));
        shard_update_kernel<<<grid, 512, 0, su>>>(shards[k]->dev, par);
        CKI(cudaEventRecord(shards[k]->ev_upd[par], su));
      }
      g_launches += 3 * nlocal;
      par ^= 1;
    }
    CKI(cudaGetLastError());
    CKI(cudaMemcpyAsync(&h, &s0->dev.ctl[par], sizeof(h), cudaMemcpyDeviceToHost, st));
    CKI(cudaStreamSynchronize(st));
    if (h.status != -1) break;
    if (h.mode != MVB_MODE_PRIMAL1) phase1 = false;
  }
  if (h.status == SHARD_ST_TIMEOUT) { g_err = "mvb_shard_solve: a peer's record did not arrive (peer exchange timed out)"; return -1; }
  // the last updates, then back to one stream
  for (int k = 0; k < nlocal; k++) {
    CKI(cudaStreamWaitEvent(st, shards[k]->ev_upd[0], 0));
    CKI(cudaStreamWaitEvent(st, shards[k]->ev_upd[1], 0));
  }
  // final gather: positions / flags / costs of every rank, then x and the objective on the host
  for (int k = 0; k < nlocal; k++)
    shard_pack_final<<<32, 256, 0, st>>>(shards[k]->dev, par, shards[k]->cost, shards[k]->maxloc,
                                         shards[k]->fin_send);
  g_launches += nlocal;
  if (!exchange(true, 0)) return -1;
  CKI(cudaEventRecord(s0->ev[1], st));
  for (int k = 0; k < nlocal; k++) {
    mvb_shard *s = shards[k];
    const int W = s->world, ML = s->maxloc;
    std::vector<unsigned char> fin(s->fin_bytes * W);
    std::vector<double> beta(m + 1);
    std::vector<int> head(m + 1);
    CKI(cudaMemcpyAsync(fin.data(), s->fin_recv, fin.size(), cudaMemcpyDeviceToHost, st));
    CKI(cudaMemcpyAsync(beta.data(), s->dev.beta[h.cS], (size_t)(m + 1) * sizeof(double), cudaMemcpyDeviceToHost, st));
    CKI(cudaMemcpyAsync(head.data(), s->dev.head[h.cS], (size_t)(m + 1) * sizeof(int), cudaMemcpyDeviceToHost, st));
    CKI(cudaStreamSynchronize(st));
    std::vector<double> cost(ng, 0.0);
    s->x.assign(ng + m, 0.0);
    s->bhead.assign(head.begin() + 1, head.end());
    s->bnb.assign(ng, 0);
    s->bat.assign(ng, 0);
    for (int r = 0; r < W; r++) {
      int c0, nc;
      mvb_shard_partition(ng, W, r, &c0, &nc);
      const int *nb = (const int *)(fin.data() + (size_t)r * s->fin_bytes), *at = nb + ML;
      const double *co = (const double *)(at + ML);
      for (int j = 0; j < nc; j++) {
        cost[c0 + j] = co[j];
        const int id = nb[j], f = at[j];
        s->bnb[c0 + j] = id;
        s->bat[c0 + j] = f;
        s->x[id] = (f == MVB_AT_LB || f == MVB_AT_FX) ? s->lb[id] : f == MVB_AT_UB ? s->ub[id] : 0.0;
      }
    }
    int ninf = 0;
    for (int i = 1; i <= m; i++) {
      const int id = head[i];
      s->x[id] = beta[i];
      const double l = s->lb[id], u = s->ub[id], b = beta[i];
      if (l - b > MVB_TOL_BND * (1.0 + std::fabs(l)) || b - u > MVB_TOL_BND * (1.0 + std::fabs(u))) ninf++;
    }
    // objective: the engine's 32-lane fma chains + halving tree (solve_lp epilogue)
    double lane[32];
    for (int l = 0; l < 32; l++) {
      double p = 0.0;
      for (int j = l; j < ng; j += 32) p = std::fma(cost[j], s->x[j], p);
      lane[l] = p;
    }
    for (int off = 16; off >= 1; off >>= 1)
      for (int l = 0; l < off; l++) lane[l] = lane[l] + lane[l + off];
    s->pivlog_n = s->dev.pivlog ? std::min(h.iters, s->pivlog_cap) : 0;
    if (k == 0) {
      mvb_result r{};
      r.status = h.status;
      r.iters = h.iters;
      r.pivots = h.pivots;
      r.flips = h.flips;
      if (dual) r.iters_dual = h.iters;
      else { r.iters_primal1 = h.it1; r.iters_primal2 = h.iters - h.it1; }
      r.bland_iters = h.bland_iters;
      r.pivot_hash = h.hash;
      r.obj = lane[0];
      r.primal_feasible = ninf == 0;
      *out = r;
    }
  }
  CKI(cudaEventElapsedTime(&s0->last_ms, s0->ev[0], s0->ev[1]));
  for (int k = 1; k < nlocal; k++) shards[k]->last_ms = s0->last_ms;
  return 0;
}
