// engine_ps.cuh -- "private state" simplex kernel: the fast path of
// engine_cuda.cu for LPs whose vectors fit in shared memory.
//
// Every CTA of the group keeps its own copy of all O(m)+O(n) state in shared
// memory (reduced costs d, basic values beta, per-row / per-position bounds,
// basis heads, flags) plus the current pivot column and scaled pivot row; the
// tableau stays in global memory (L2/HBM) as two ping-pong buffers.  Pricing
// and both ratio tests then run out of shared memory, and the data the NEXT
// pivot needs from the tableau -- its column and its row -- are derived from
// the OLD buffer (value' = fma(c_i, rho_j, value)), so their loads overlap the
// streaming rank-1 update and the group barrier instead of waiting for them:
//   apply state k -> price k+1 -> [column k+1 loads in flight] -> tile update k
//   -> barrier arrive -> ratio test k+1 -> [row k+1 by 1-D TMA bulk copy]
//   -> barrier wait.
// The pivot row travels global->shared by cp.async.bulk (TMA, mbarrier
// completion); the pivot column is a strided 8-byte gather (ld.global.cg).
// Arithmetic and decisions are those of solve_lp / the oracle, bit for bit.
// Phase 1 is not implemented here: such LPs are handed to simplex_kernel.
#pragma once

namespace {

constexpr int PS_CREG = 8;            // pivot-column values a thread can hold in flight
constexpr int MVB_PS_FALLBACK = 100;  // DevResult status: needs the general kernel

__device__ __forceinline__ Key warp_argmin(Key v) {
  const unsigned full = 0xffffffffu;
  const unsigned hi = (unsigned)(v.k >> 32), lo = (unsigned)v.k;
  const unsigned mh = __reduce_min_sync(full, hi);
  const unsigned ml = __reduce_min_sync(full, hi == mh ? lo : 0xffffffffu);
  const bool c = hi == mh && lo == ml;
  const unsigned mv = __reduce_min_sync(full, c ? (unsigned)v.var : 0xffffffffu);
  const unsigned b = __ballot_sync(full, c && (unsigned)v.var == mv);
  const int src = b ? __ffs(b) - 1 : 0;
  Key o;
  o.k = ((unsigned long long)mh << 32) | ml;
  o.var = (int)mv;
  o.idx = __shfl_sync(full, v.idx, src);
  return o;
}

// block argmin with ONE __syncthreads: the staging array alternates between
// two halves from call to call
struct FastRed {
  Key slot[2][32];
  int par;
};
__device__ __forceinline__ Key block_argmin_fast(Key v, Key (*slot)[32], int &par) {
  v = warp_argmin(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  if (lane == 0) slot[par][w] = v;
  __syncthreads();
  v = lane < nw ? slot[par][lane] : key_none();
  par ^= 1;
  return warp_argmin(v);
}

// ---- mbarrier / 1-D TMA bulk copy (global -> shared) ----
__device__ __forceinline__ unsigned smem_u32(const void *p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_row_load(double *dst, const double *src, unsigned bytes,
                                              unsigned long long *bar) {
  // order earlier generic-proxy accesses of dst before the async-proxy write
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@!p bra WAIT_%=;\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// split group barrier: arrive after the tile stores, wait before the next
// iteration touches the other buffer
__device__ __forceinline__ void group_arrive(unsigned *ctr, int G, unsigned &gen) {
  __syncthreads();
  gen++;
  if (G > 1 && threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
  }
}
__device__ __forceinline__ void group_wait(unsigned *ctr, int G, unsigned gen) {
  if (G > 1) {
    if (threadIdx.x == 0) {
      const unsigned target = gen * (unsigned)G;
      unsigned seen;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(ctr) : "memory");
      } while (seen < target);
    }
    __syncthreads();
  }
}

// shared-memory state of one CTA
struct PS {
  double *d, *rho[2], *lbN, *ubN;          // n-vectors (ld long)
  int *nbv, *atv;
  double *col[2], *beta, *lbB, *ubB;       // m-vectors (m+1 long, 1-based; col[.][0] = d_q)
  int *head;
  Key (*slot)[32];
  unsigned long long *mbar;
  double *scal;                            // broadcast scratch (8 doubles)
};

__host__ __device__ inline size_t ps_smem_bytes(int m, int ld) {
  size_t nv = (size_t)ld * (8 * 5 + 4 * 2);
  size_t mv = (size_t)(m + 2) * (8 * 5 + 4);
  return nv + mv + sizeof(Key) * 64 + 64 + 64 + 256;
}

__device__ inline void ps_carve(PS &S, unsigned char *base, int m, int ld) {
  size_t off = 0;
  auto take = [&](size_t bytes) { unsigned char *p = base + off; off += (bytes + 15) / 16 * 16; return p; };
  S.d = (double *)take((size_t)ld * 8);
  S.rho[0] = (double *)take((size_t)ld * 8);
  S.rho[1] = (double *)take((size_t)ld * 8);
  S.lbN = (double *)take((size_t)ld * 8);
  S.ubN = (double *)take((size_t)ld * 8);
  S.col[0] = (double *)take((size_t)(m + 2) * 8);
  S.col[1] = (double *)take((size_t)(m + 2) * 8);
  S.beta = (double *)take((size_t)(m + 2) * 8);
  S.lbB = (double *)take((size_t)(m + 2) * 8);
  S.ubB = (double *)take((size_t)(m + 2) * 8);
  S.slot = (Key(*)[32])take(sizeof(Key) * 64);
  S.mbar = (unsigned long long *)take(16);
  S.scal = (double *)take(64);
  S.nbv = (int *)take((size_t)ld * 4);
  S.atv = (int *)take((size_t)ld * 4);
  S.head = (int *)take((size_t)(m + 2) * 4);
}

}  // namespace

namespace {

// rank-1 update of this CTA's share, rho and the pivot column read from shared
// memory (same unit split and arithmetic as tile_update)
__device__ void tile_update_ps(const LPDev &L, const double *__restrict__ Ts,
                               double *__restrict__ Td, int r, int q, int g, int G, int tcl,
                               const double *rho_s, const double *col_s) {
  const int R = L.m + 1, ld = L.ld;
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
    if (col < ld) {
      const bool qx = col == q, qy = col + 1 == q;
      const double2 rho = *reinterpret_cast<const double2 *>(rho_s + col);
      for (int i = ia + ty; i < ib; i += trows * UNROLL) {
        double2 t[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
          const int ii = i + u * trows;
          if (ii < ib) t[u] = ldcg2(Ts + (size_t)ii * ld + col);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
          const int ii = i + u * trows;
          if (ii < ib) {
            const double c = col_s[ii];
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

// what one iteration decided
struct Dec {
  int status;   // -1: continue with (r,q); otherwise final MVB_* status
  int r, q;     // r == 0: bound flip
  double delta, newbr;
  int newat;
  bool degenerate;
};

// the pivot that is being applied while the next one is decided
struct Prev {
  bool piv;     // a rank-1 update is pending
  int r, q, p;  // p: which rho/col buffers hold its row and column
};

// value of T_{k+1}[i][j] given T_k[i][j] = raw and the pending pivot
__device__ __forceinline__ double derive(const Prev &pv, int i, int j, double raw, double ci,
                                         double rhoj) {
  if (!pv.piv) return raw;
  if (i == pv.r) return rhoj;
  return fma(ci, rhoj, j == pv.q ? 0.0 : raw);
}

}  // namespace

namespace {

struct PSCtx {
  const LPDev *L;
  PS *S;
  int g, G, tcl;
  unsigned *ctr;
  unsigned gen;
  int cT;  // tableau buffer that holds T_k (read side of the pending update)
  int par;          // block_argmin_fast slot parity (per thread, identical everywhere)
  unsigned mphase;  // mbarrier phase of the next row load
};

// run the pending rank-1 update (if any) and signal the group
__device__ __forceinline__ void run_pending(PSCtx &X, const Prev &pv) {
  if (!pv.piv) return;
  const LPDev &L = *X.L;
  tile_update_ps(L, L.T[X.cT], L.T[X.cT ^ 1], pv.r, pv.q, X.g, X.G, X.tcl, X.S->rho[pv.p],
                 X.S->col[pv.p]);
  group_arrive(X.ctr, X.G, X.gen);
}

// fetch row r of T_{k+1} into rho[pn] as raw row values (derive applied)
__device__ __forceinline__ void fetch_row(PSCtx &X, const Prev &pv, int r, int pn) {
  const LPDev &L = *X.L;
  PS &S = *X.S;
  __syncthreads();  // everyone is done with rho[pn]
  if (threadIdx.x == 0)
    bulk_row_load(S.rho[pn], L.T[X.cT] + (size_t)r * L.ld, (unsigned)(L.ld * sizeof(double)), S.mbar);
}
__device__ __forceinline__ void finish_row(PSCtx &X, const Prev &pv, int r, int pn) {
  PS &S = *X.S;
  mbar_wait(S.mbar, X.mphase);
  X.mphase ^= 1;
  if (pv.piv) {
    const double *rho = S.rho[pv.p];
    const double cr = S.col[pv.p][r];
    for (int j = threadIdx.x; j < X.L->ld; j += blockDim.x)
      S.rho[pn][j] = derive(pv, r, j, S.rho[pn][j], cr, rho[j]);
  }
  __syncthreads();
}

// rho[pn] <- scaled pivot row from raw row values: rho_q = inv, rho_j = -(v_j*inv)
__device__ __forceinline__ void scale_row(PSCtx &X, int q, double inv, int pn) {
  PS &S = *X.S;
  for (int j = threadIdx.x; j < X.L->ld; j += blockDim.x)
    S.rho[pn][j] = j == q ? inv : -(S.rho[pn][j] * inv);
  __syncthreads();
}

// issue / finish the strided gather of column q of T_{k+1} (rows 0..m)
__device__ __forceinline__ void issue_col(PSCtx &X, int q, double (&raw)[PS_CREG]) {
  const LPDev &L = *X.L;
  const double *T = L.T[X.cT];
#pragma unroll
  for (int u = 0; u < PS_CREG; u++) {
    const int i = threadIdx.x + u * blockDim.x;
    if (i <= L.m) raw[u] = ldcg(T + (size_t)i * L.ld + q);
  }
}
__device__ __forceinline__ void finish_col(PSCtx &X, const Prev &pv, int q, int pn,
                                           const double (&raw)[PS_CREG]) {
  PS &S = *X.S;
  const double rhoq = pv.piv ? S.rho[pv.p][q] : 0.0;
  const double *c = S.col[pv.p];
#pragma unroll
  for (int u = 0; u < PS_CREG; u++) {
    const int i = threadIdx.x + u * blockDim.x;
    if (i <= X.L->m) S.col[pn][i] = derive(pv, i, q, raw[u], pv.piv ? c[i] : 0.0, rhoq);
  }
  __syncthreads();
}

// PRIMAL2 decision for the next iteration, with the pending update folded in
__device__ Dec decide_primal(PSCtx &X, const Prev &pv, bool bland, int iters) {
  const LPDev &L = *X.L;
  PS &S = *X.S;
  const int m = L.m, n = L.n, tid = threadIdx.x, NT = blockDim.x, pn = pv.p ^ 1;
  Dec dc{};
  dc.status = -1;
  // K1 pricing out of shared memory
  Key bq = key_none();
  for (int j = tid; j < n; j += NT) {
    const double dj = S.d[j];
    if (!dj_eligible(S.atv[j], dj)) continue;
    const Key k{bland ? 0ULL : ~dbits(fabs(dj)), S.nbv[j], j};
    if (key_less(k, bq)) bq = k;
  }
  bq = block_argmin_fast(bq, S.slot, X.par);
  int q = bq.idx;
  if (q < 0) dc.status = MVB_OPT;
  else if (iters >= L.max_iters) { dc.status = MVB_ITLIM; q = -1; }
  double raw[PS_CREG];
  if (q >= 0) issue_col(X, q, raw);  // loads stay in flight across the update
  run_pending(X, pv);
  if (q >= 0) {
    finish_col(X, pv, q, pn, raw);
    const double dq = S.d[q];
    const int dir = dq < 0.0 ? 1 : -1;
    // K2 ratio test out of shared memory
    const double *col = S.col[pn];
    Key bi = key_none();
    for (int i = 1 + tid; i <= m; i += NT) {
      double t;
      int fl;
      if (!primal_ratio_row(col[i], dir, 0, S.beta[i], S.lbB[i], S.ubB[i], t, fl)) continue;
      const Key k{dbits(t), S.head[i], i};
      if (key_less(k, bi)) bi = k;
    }
    bi = block_argmin_fast(bi, S.slot, X.par);
    const double lq = S.lbN[q], uq = S.ubN[q];
    const int fq = S.atv[q];
    const double tflip = uq - lq;
    const double tmin = bi.idx >= 0 ? __longlong_as_double((long long)bi.k) : HUGE_VAL;
    dc.q = q;
    if (tflip <= tmin) {
      if (!(tflip < HUGE_VAL)) dc.status = MVB_UNBND;
      dc.r = 0;
      dc.delta = dir > 0 ? tflip : -tflip;
      dc.newat = fq == MVB_AT_LB ? MVB_AT_UB : MVB_AT_LB;
      dc.degenerate = false;
    } else {
      const int r = bi.idx;
      double t;
      primal_ratio_row(col[r], dir, 0, S.beta[r], S.lbB[r], S.ubB[r], t, dc.newat);
      dc.r = r;
      dc.delta = dir > 0 ? tmin : -tmin;
      dc.newbr = nb_value(fq, lq, uq) + dc.delta;
      dc.degenerate = tmin == 0.0;
      fetch_row(X, pv, r, pn);
      finish_row(X, pv, r, pn);
      scale_row(X, q, 1.0 / col[r], pn);
    }
  }
  if (pv.piv) group_wait(X.ctr, X.G, X.gen);
  return dc;
}

}  // namespace

namespace {

// DUAL decision for the next iteration, with the pending update folded in
__device__ Dec decide_dual(PSCtx &X, const Prev &pv, bool bland, int iters) {
  const LPDev &L = *X.L;
  PS &S = *X.S;
  const int m = L.m, n = L.n, tid = threadIdx.x, NT = blockDim.x, pn = pv.p ^ 1;
  Dec dc{};
  dc.status = -1;
  // K5a leaving row out of shared memory
  Key best = key_none();
  for (int i = 1 + tid; i <= m; i += NT) {
    const double b = S.beta[i], l = S.lbB[i], u = S.ubB[i];
    const int s = row_infeas(b, l, u);
    if (!s) continue;
    const double viol = s < 0 ? l - b : b - u;
    const Key k{bland ? 0ULL : ~dbits(viol), S.head[i], i};
    if (key_less(k, best)) best = k;
  }
  best = block_argmin_fast(best, S.slot, X.par);
  int r = best.idx;
  if (r < 0) dc.status = MVB_OPT;
  else if (iters >= L.max_iters) { dc.status = MVB_ITLIM; r = -1; }
  if (r >= 0) fetch_row(X, pv, r, pn);  // TMA bulk copy in flight across the update
  run_pending(X, pv);
  if (r >= 0) {
    finish_row(X, pv, r, pn);
    const double br = S.beta[r], lr = S.lbB[r], ur = S.ubB[r];
    const int rs = row_infeas(br, lr, ur);
    // K5b entering column: min |d_j| / |T_rj|
    const double *row = S.rho[pn];
    Key bq = key_none();
    for (int j = tid; j < n; j += NT) {
      const int f = S.atv[j];
      if (f == MVB_AT_FX) continue;
      const double alpha = row[j];
      const double a = rs < 0 ? alpha : -alpha;
      const bool ok = (f == MVB_AT_LB && a > MVB_TOL_PIV) || (f == MVB_AT_UB && a < -MVB_TOL_PIV) ||
                      (f == MVB_AT_FR && (a > MVB_TOL_PIV || a < -MVB_TOL_PIV));
      if (!ok) continue;
      const double ratio = fabs(S.d[j]) / fabs(alpha);
      const Key k{dbits(ratio), S.nbv[j], j};
      if (key_less(k, bq)) bq = k;
    }
    bq = block_argmin_fast(bq, S.slot, X.par);
    const int q = bq.idx;
    if (q < 0) {
      dc.status = MVB_NOFEAS;
    } else {
      double raw[PS_CREG];
      issue_col(X, q, raw);
      finish_col(X, pv, q, pn, raw);
      const double alpha_rq = row[q];
      const double target = rs < 0 ? lr : ur;
      dc.r = r;
      dc.q = q;
      dc.delta = (target - br) / alpha_rq;
      dc.newbr = nb_value(S.atv[q], S.lbN[q], S.ubN[q]) + dc.delta;
      dc.newat = rs < 0 ? MVB_AT_LB : MVB_AT_UB;
      dc.degenerate = __longlong_as_double((long long)bq.k) == 0.0;
      __syncthreads();  // all reads of the raw row done before it is rescaled
      scale_row(X, q, 1.0 / alpha_rq, pn);
    }
  }
  if (pv.piv) group_wait(X.ctr, X.G, X.gen);
  return dc;
}

// apply a decision to the private state (all in shared memory):
//   d'_j = fma(d_q, rho_j, j == q ? 0 : d_j)      (row 0 of the rank-1 update)
//   beta'_i = i == r ? newbr : fma(T_iq, delta, beta_i)
//   basis swap incl. the per-row / per-position bounds and flags
__device__ void apply_state(PSCtx &X, const Dec &dc, int p) {
  const LPDev &L = *X.L;
  PS &S = *X.S;
  const int tid = threadIdx.x, NT = blockDim.x, r = dc.r, q = dc.q;
  const double *col = S.col[p], *rho = S.rho[p];
  for (int i = 1 + tid; i <= L.m; i += NT) S.beta[i] = i == r ? dc.newbr : fma(col[i], dc.delta, S.beta[i]);
  if (r) {
    const double dq = col[0];
    for (int j = tid; j < L.ld; j += NT) S.d[j] = fma(dq, rho[j], j == q ? 0.0 : S.d[j]);
  }
  __syncthreads();
  if (tid == 0) {
    if (r) {
      const int kin = S.nbv[q], kout = S.head[r];
      const double lo = S.lbB[r], uo = S.ubB[r];
      S.head[r] = kin;
      S.lbB[r] = S.lbN[q];
      S.ubB[r] = S.ubN[q];
      S.nbv[q] = kout;
      S.lbN[q] = lo;
      S.ubN[q] = uo;
      S.atv[q] = lo == uo ? MVB_AT_FX : dc.newat;
    } else {
      S.atv[q] = dc.newat;
    }
  }
  __syncthreads();
}

}  // namespace

namespace {

// The private-state iteration loop of one LP (mode PRIMAL2 or DUAL; neither
// changes mode).  Entered after the common prologue (beta[cS] valid in global
// memory); leaves the final small state in the global arrays of parity cS and
// the final tableau in buffer cT, as the general loop does.  Returns the status.
__device__ int ps_loop(const LPDev &L, int g, int G, unsigned *ctr, unsigned &gen, int tcl,
                       unsigned char *dyn, int mode, int &cT, int cS, Tally &ty) {
  const int m = L.m, n = L.n, ld = L.ld, tid = threadIdx.x, NT = blockDim.x;
  __shared__ PS S;  // pointers into the carved dynamic shared memory
  if (tid == 0) {
    ps_carve(S, dyn, m, ld);
    mbar_init(S.mbar, 1);
  }
  __syncthreads();
  {
    const double *d0 = L.T[cT];
    const int *nb = L.nb[cS], *at = L.at[cS], *head = L.head[cS];
    const double *beta = L.beta[cS];
    for (int j = tid; j < ld; j += NT) {
      if (j < n) {
        const int k = ldcg(nb + j);
        S.d[j] = ldcg(d0 + j);
        S.nbv[j] = k;
        S.atv[j] = ldcg(at + j);
        S.lbN[j] = ldcg(L.lb + k);
        S.ubN[j] = ldcg(L.ub + k);
      } else {  // padding columns never price
        S.d[j] = 0.0;
        S.nbv[j] = 0x7fffffff;
        S.atv[j] = MVB_AT_FX;
        S.lbN[j] = S.ubN[j] = 0.0;
      }
    }
    for (int i = tid; i <= m + 1; i += NT) {
      if (i >= 1 && i <= m) {
        const int k = ldcg(head + i);
        S.beta[i] = ldcg(beta + i);
        S.head[i] = k;
        S.lbB[i] = ldcg(L.lb + k);
        S.ubB[i] = ldcg(L.ub + k);
      } else {
        S.beta[i] = 0.0;
        S.head[i] = 0x7fffffff;
        S.lbB[i] = -HUGE_VAL;
        S.ubB[i] = HUGE_VAL;
      }
    }
  }
  __syncthreads();

  PSCtx X{&L, &S, g, G, tcl, ctr, gen, cT, 0, 0u};
  Prev pv{false, 0, 0, 0};
  int status, degen = 0;
  while (true) {
    const bool bland = degen >= MVB_BLAND_AFTER;
    const Dec dc = mode == MVB_MODE_DUAL ? decide_dual(X, pv, bland, ty.iters)
                                         : decide_primal(X, pv, bland, ty.iters);
    if (pv.piv) X.cT ^= 1;  // the pending update has landed everywhere
    if (dc.status >= 0) { status = dc.status; break; }
    const int pn = pv.p ^ 1;
    tally_iter(ty, L, g, mode, dc.r, dc.q, bland);
    apply_state(X, dc, pn);
    degen = dc.degenerate ? degen + 1 : 0;
    pv.piv = dc.r != 0;
    pv.r = dc.r;
    pv.q = dc.q;
    pv.p = pn;
  }
  gen = X.gen;
  cT = X.cT;
  // hand the private state back (every CTA holds the same values)
  if (g == 0) {
    double *beta = L.beta[cS];
    int *head = L.head[cS], *nb = L.nb[cS], *at = L.at[cS];
    for (int i = 1 + tid; i <= m; i += NT) { beta[i] = S.beta[i]; head[i] = S.head[i]; }
    for (int j = tid; j < n; j += NT) { nb[j] = S.nbv[j]; at[j] = S.atv[j]; }
  }
  group_sync(ctr, G, gen);
  return status;
}

}  // namespace
