// engine_cres.cuh -- CLUSTER-RESIDENT simplex kernel (SURVEY.md K6 second half,
// north_star "one tableau per CTA-cluster"): a thread-block cluster of C = 2, 4
// or 8 CTAs owns one LP; the tableau is split by ROW BLOCKS over the cluster's
// shared memories (distributed shared memory) and stays there for the whole
// solve.  A configs[3]-shape node LP (257 x 512 doubles = 1.05 MB) takes a
// cluster of 8 (132 KB of tableau per CTA); it is read from HBM once, solved
// with eager rank-1 updates on chip, written back once.
//
// Who holds what.  CTA c of the cluster owns the global rows
// [1 + c*RB, 1 + (c+1)*RB) with beta, bounds and basis heads of exactly those
// rows; EVERY CTA keeps its own copy of the reduced-cost row (local row 0) and
// of the non-basic ids / flags, and updates them with the same operations, so
// pricing needs no exchange.  Per primal iteration:
//   local K2 ratio test over the CTA's rows
//   -> EXCHANGE 1: every CTA stores its candidate (key, row, pivot element) into
//      a slot of every peer (st.shared::cluster) and the cluster meets at the
//      hardware barrier; the same winner everywhere by the (value bits, id) order
//   -> EXCHANGE 2: the CTA that owns the pivot row scales it (K3) and sends it to
//      every CTA of the cluster -- itself included -- with one bulk copy each
//      (cp.async.bulk shared::cta -> shared::cluster, completion on the
//      receiver's mbarrier): no second cluster barrier
//   -> every CTA updates its row block and its reduced-cost copy (K4) and prices
//      the next iteration (K1) on the values it has just written.
// Dual iterations: EXCHANGE 1 selects the leaving row (K5a, local candidates);
// its owner runs the K5b ratio test alone (it holds the row) and EXCHANGE 2
// carries the scaled row plus a header (entering position, flag, id, pivot
// element, degeneracy); the others take the pivot column from their own rows.
// The arithmetic is the specification's, operation for operation; every
// reduction is the total order on (value bits, variable id): bit-identical to
// the oracle and to the other three kernels.
#pragma once

namespace {

constexpr int CRES_NT = 256;
constexpr int CRES_MAXC = 8;

struct XRec {  // one candidate of a cluster-wide argmin
  unsigned long long k;
  int var, idx;
  double a, b, c;
  double pad;
};

struct CresLay {
  unsigned lds, RB;  // row stride (doubles); rows per CTA
  unsigned oT, oCol, oBeta, oLb, oUb, oHead, oRho, oRhoS, oNb, oAt, oX, oSlot, oBar, total;
};
__host__ __device__ inline CresLay cres_layout(int m, int ld, int C) {
  CresLay l;
  const unsigned RB = ((unsigned)m + (unsigned)C - 1u) / (unsigned)C, L = (unsigned)ld;
  l.RB = RB;
  l.lds = L + 2u;
  unsigned off = 0;
  l.oT = off;    off = res_up16(off + (RB + 1u) * l.lds * 8u);
  l.oCol = off;  off = res_up16(off + (RB + 2u) * 8u);
  l.oBeta = off; off = res_up16(off + (RB + 2u) * 8u);
  l.oLb = off;   off = res_up16(off + (RB + 2u) * 8u);
  l.oUb = off;   off = res_up16(off + (RB + 2u) * 8u);
  l.oHead = off; off = res_up16(off + (RB + 2u) * 4u);
  l.oRho = off;  off = res_up16(off + (L + 8u) * 8u);   // received: rho[ld] + header[8]
  l.oRhoS = off; off = res_up16(off + (L + 8u) * 8u);   // staging of the owner
  l.oNb = off;   off = res_up16(off + L * 4u);
  l.oAt = off;   off = res_up16(off + L * 4u);
  l.oX = off;    off = res_up16(off + 2u * CRES_MAXC * (unsigned)sizeof(XRec));
  l.oSlot = off; off = res_up16(off + (unsigned)sizeof(Key) * 64u);
  l.oBar = off;  off = res_up16(off + 32u);
  l.total = off;
  return l;
}
__host__ __device__ inline size_t cres_smem_bytes(int m, int ld, int C) { return cres_layout(m, ld, C).total; }
// smallest cluster size (2, 4, 8) whose CTAs can hold the LP; 0: none
__host__ inline int cres_pick_c(int m, int ld, size_t avail) {
  for (int C = 2; C <= CRES_MAXC; C *= 2)
    if (cres_smem_bytes(m, ld, C) <= avail) return C;
  return 0;
}

__device__ __forceinline__ unsigned cres_rank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned cres_cluster_id() {  // clusters are 1-D: cluster id along x
  unsigned r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned cres_nclusters() {
  unsigned r;
  asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cres_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ unsigned cres_map(const void *p, unsigned rank) {
  unsigned r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(rank));
  return r;
}
__device__ __forceinline__ void cres_st64(unsigned addr, unsigned long long v) {
  asm volatile("st.shared::cluster.u64 [%0], %1;" ::"r"(addr), "l"(v) : "memory");
}

// cluster-wide argmin: every CTA contributes one record (identical in all its threads); on return
// every thread of every CTA holds the winner by (k, var).  One hardware cluster barrier.
__device__ XRec cres_argmin(const XRec &mine, XRec *xbuf /* [2][CRES_MAXC] */, int par, int C, unsigned rank) {
  if ((int)threadIdx.x < C) {
    const unsigned dst = cres_map(xbuf + par * CRES_MAXC + rank, threadIdx.x);
    cres_st64(dst, mine.k);
    cres_st64(dst + 8, ((unsigned long long)(unsigned)mine.idx << 32) | (unsigned)mine.var);
    cres_st64(dst + 16, (unsigned long long)__double_as_longlong(mine.a));
    cres_st64(dst + 24, (unsigned long long)__double_as_longlong(mine.b));
    cres_st64(dst + 32, (unsigned long long)__double_as_longlong(mine.c));
  }
  cres_sync();
  const XRec *r = xbuf + par * CRES_MAXC;
  XRec best = r[0];
  for (int c = 1; c < C; c++) {
    const XRec o = r[c];
    if (o.k < best.k || (o.k == best.k && o.var < best.var)) best = o;
  }
  return best;
}
// cluster-wide sum of one small integer per CTA (mode detection, final feasibility)
__device__ int cres_sum(int mine, XRec *xbuf, int par, int C, unsigned rank) {
  if ((int)threadIdx.x < C) {
    const unsigned dst = cres_map(xbuf + par * CRES_MAXC + rank, threadIdx.x);
    cres_st64(dst + 8, (unsigned long long)(unsigned)mine);  // var = mine, idx = 0
  }
  cres_sync();
  int s = 0;
  for (int c = 0; c < C; c++) s += xbuf[par * CRES_MAXC + c].var;
  return s;
}

struct CresCtx {
  Tally ty;
  int degen;
};

// K4 on this CTA's tile (local row 0 = its reduced-cost copy, local rows 1..nloc = its row block):
// same arithmetic and thread mapping as res_update; rloc = local index of the pivot row or -1.
// The entering position's new occupant (kout, atq) comes in registers: the basis swap by another
// thread is not ordered with the pricing reads.
template <bool PRICE>
__device__ __forceinline__ Key cres_update(const CresLay &l, int nloc, int n, int ld, int rloc, int q, bool bland,
                                           int tcl, int kout, int atq) {
  return smem_tile_update<PRICE>(reinterpret_cast<double *>(dyn_smem + l.oT), l.lds, nloc, n, ld,
                                 reinterpret_cast<const double *>(dyn_smem + l.oCol),
                                 reinterpret_cast<const double *>(dyn_smem + l.oRho),
                                 reinterpret_cast<const int *>(dyn_smem + l.oNb),
                                 reinterpret_cast<const int *>(dyn_smem + l.oAt), rloc, q, bland, tcl, kout, atq);
}

// the owner of the pivot row sends rho (+ header) from its staging buffer to every CTA of the
// cluster, itself included: one bulk copy each, completion counted on the receiver's mbarrier
__device__ __forceinline__ void cres_send_rho(const CresLay &l, int ld, int C) {
  // (the staging buffer was written through the generic proxy by all threads: fenced by them, then a
  // block barrier, before this single thread issues the async-proxy copies)
  const unsigned bytes = (unsigned)(ld + 8) * 8u;
  const void *src = dyn_smem + l.oRhoS;
  for (int c = 0; c < C; c++) {
    const unsigned dst = cres_map(dyn_smem + l.oRho, (unsigned)c);
    const unsigned bar = cres_map(dyn_smem + l.oBar + 8, (unsigned)c);
    asm volatile(
        "cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
        "r"(smem_u32(src)), "r"(bytes), "r"(bar)
        : "memory");
  }
}

__device__ void cres_solve(const LPDev &L, CresCtx *S, int C, unsigned rank) {
  const int m = L.m, n = L.n, ld = L.ld, tid = threadIdx.x, NT = blockDim.x;
  const CresLay l = cres_layout(m, ld, C);
  const unsigned lds = l.lds;
  const int RB = (int)l.RB;
  const int g0 = 1 + (int)rank * RB;                       // first global row of this CTA
  const int nloc = max(0, min(m + 1, g0 + RB) - g0);       // its rows: local 1..nloc <-> global g0..g0+nloc-1
  double *T = reinterpret_cast<double *>(dyn_smem + l.oT);
  double *col = reinterpret_cast<double *>(dyn_smem + l.oCol), *beta = reinterpret_cast<double *>(dyn_smem + l.oBeta);
  double *lbB = reinterpret_cast<double *>(dyn_smem + l.oLb), *ubB = reinterpret_cast<double *>(dyn_smem + l.oUb);
  int *head = reinterpret_cast<int *>(dyn_smem + l.oHead);
  double *rho = reinterpret_cast<double *>(dyn_smem + l.oRho), *rhoS = reinterpret_cast<double *>(dyn_smem + l.oRhoS);
  int *nbv = reinterpret_cast<int *>(dyn_smem + l.oNb), *atv = reinterpret_cast<int *>(dyn_smem + l.oAt);
  XRec *xbuf = reinterpret_cast<XRec *>(dyn_smem + l.oX);
  Key(*slot)[32] = reinterpret_cast<Key(*)[32]>(dyn_smem + l.oSlot);
  unsigned long long *barL = reinterpret_cast<unsigned long long *>(dyn_smem + l.oBar);       // load
  unsigned long long *barR = reinterpret_cast<unsigned long long *>(dyn_smem + l.oBar + 8);   // rho arrivals
  const int cT = L.curT, cS = L.curS;
  int tcl = 0;
  while ((1 << tcl) < (ld + 1) / 2 && (2 << tcl) <= NT) tcl++;
  const unsigned rho_bytes = (unsigned)(ld + 8) * 8u;
  unsigned rphase = 0;  // parity of the rho barrier
  int xpar = 0;         // exchange buffer in use

  // ---- load: the reduced-cost row and this CTA's row block, one bulk copy per row ----
  const double *srcT = L.srcT ? L.srcT : L.T[cT];
  if (tid == 0) {
    mbar_init(barL, 1);
    mbar_init(barR, 1);
    asm volatile("fence.proxy.async;" ::: "memory");
    const unsigned rowb = (unsigned)ld * 8u;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(barL)), "r"(rowb * (unsigned)(nloc + 1))
                 : "memory");
    for (int i = 0; i <= nloc; i++) {
      const int gi = i == 0 ? 0 : g0 + i - 1;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       smem_u32(T + (size_t)i * lds)),
                   "l"(srcT + (size_t)gi * ld), "r"(rowb), "r"(smem_u32(barL))
                   : "memory");
    }
    S->ty = Tally();
    S->degen = 0;
  }
  {
    const int *gnb = L.nb[cS], *gat = L.at[cS], *ghead = L.head[cS];
    for (int j = tid; j < ld; j += NT) {
      const bool in = j < n;
      nbv[j] = in ? ldcg(gnb + j) : 0x7fffffff;
      atv[j] = in ? ldcg(gat + j) : MVB_AT_FX;
      double v = 0.0;
      if (in) {
        const int k = ldcg(gnb + j);
        v = nb_value(ldcg(gat + j), ldcg(L.lb + k), ldcg(L.ub + k));
      }
      rhoS[j] = v;  // values of the non-basic variables (scratch) for the beta rebuild
    }
    for (int i = tid; i <= RB + 1; i += NT) {
      const bool in = i >= 1 && i <= nloc;
      const int k = in ? ldcg(ghead + g0 + i - 1) : 0;
      head[i] = in ? k : 0x7fffffff;
      lbB[i] = in ? ldcg(L.lb + k) : -HUGE_VAL;
      ubB[i] = in ? ldcg(L.ub + k) : HUGE_VAL;
      beta[i] = 0.0;
    }
  }
  // every CTA of the cluster has initialised its barriers before anybody may send to them
  cres_sync();
  mbar_wait(barL, 0);
  int mode;
  if (L.carry) {
    for (int i = 1 + tid; i <= nloc; i += NT) beta[i] = ldcg(L.beta[cS] + g0 + i - 1);
    if (tid == 0) {
      Tally &t = S->ty;
      t.iters = L.c_iters; t.pivots = L.c_pivots; t.flips = L.c_flips; t.it1 = L.c_it1;
      t.it2 = L.c_it2; t.itd = L.c_itd; t.bland = L.c_bland; t.hash = L.c_hash;
    }
    mode = MVB_MODE_PRIMAL2;
    __syncthreads();
  } else {
    const int lane = tid & 31, w = tid >> 5, nw = NT >> 5;
    for (int i = 1 + w; i <= nloc; i += nw) {
      const double *row = T + (size_t)i * lds;
      double p = 0.0;
      for (int j = lane; j < n; j += 32) p = fma(row[j], rhoS[j], p);
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) p = p + __shfl_down_sync(0xffffffffu, p, off);
      if (lane == 0) beta[i] = p;
    }
    __syncthreads();
    int pinf = 0, dinf = 0;
    for (int i = 1 + tid; i <= nloc; i += NT) pinf += row_infeas(beta[i], lbB[i], ubB[i]) != 0;
    for (int j = tid; j < n; j += NT) dinf += dj_eligible(atv[j], T[j]);
    int *red = reinterpret_cast<int *>(slot);
    pinf = block_sum(pinf, red);
    dinf = block_sum(dinf, red);
    pinf = cres_sum(pinf, xbuf, xpar, C, rank);
    xpar ^= 1;
    mode = pinf == 0 ? MVB_MODE_PRIMAL2 : dinf == 0 ? MVB_MODE_DUAL : MVB_MODE_PRIMAL1;
  }
  if (mode == MVB_MODE_PRIMAL1) {  // needs the general kernel: nothing has been changed
    if (rank == 0 && tid == 0) L.res->r.status = MVB_PS_FALLBACK;
    cres_sync();
    if (tid == 0) { mbar_inval(barL); mbar_inval(barR); }
    __syncthreads();
    return;
  }

  int status = MVB_ITLIM;
  int par = 0;
  if (mode == MVB_MODE_PRIMAL2) {
    // first pricing: every CTA on its own copy of the reduced costs
    int q;
    {
      Key best = key_none();
      for (int j = tid; j < n; j += NT) {
        const Key k = res_price_key(T[j], atv[j], nbv[j], j, false);
        if (key_less(k, best)) best = k;
      }
      best = block_argmin_fast(best, slot, par);
      par ^= 1;
      q = best.idx;
    }
    while (true) {
      if (q < 0) { status = MVB_OPT; break; }
      if (S->ty.iters >= L.max_iters) break;
      const int degen0 = S->degen;
      const bool bland = degen0 >= MVB_BLAND_AFTER;
      const double dq = T[q];
      const int dir = dq < 0.0 ? 1 : -1;
      const int fq = atv[q], kq = nbv[q];
      const double lq = ldcg(L.lb + kq), uq = ldcg(L.ub + kq);
      // local K2 over this CTA's rows (rule of lz_primal_iter), pivot column copied on the way
      double bt = HUGE_VAL;
      int bvar = 0x7fffffff, bidx = -1;
      if (tid == 0) col[0] = dq;
      for (int i = 1 + tid; i <= nloc; i += NT) {
        const double v = T[(size_t)i * lds + q];
        col[i] = v;
        const double b = beta[i], lo = lbB[i], up = ubB[i];
        const int hv = head[i];
        const double a = dir > 0 ? v : -v;
        const bool dec = a < -MVB_TOL_PIV, inc = a > MVB_TOL_PIV;
        const double bound = dec ? lo : up;
        const bool ok = dec ? bound > -HUGE_VAL : (inc && bound < HUGE_VAL);
        const double num0 = dec ? b - bound : bound - b;
        const bool live = ok & (num0 != 0.0);
        const double num = live ? num0 : 1.0;
        const double den = live ? fabs(a) : 1.0;
        double t = num / den;
        if (!live | !(t > 0.0)) t = 0.0;
        const bool better = ok & ((t < bt) | ((t == bt) & (hv < bvar)));
        bt = better ? t : bt;
        bvar = better ? hv : bvar;
        bidx = better ? (lo == up ? i | LZ_FXBIT : i) : bidx;
      }
      Key bi = bidx >= 0 ? Key{dbits(bt), bvar, bidx} : key_none();
      bi = block_argmin_fast(bi, slot, par);  // barrier: also publishes col
      par ^= 1;
      // EXCHANGE 1: this CTA's candidate -> the cluster's winner
      XRec mine{};
      mine.k = bi.k;
      mine.var = bi.var;
      mine.idx = -1;
      if (bi.idx >= 0) {
        const int li = bi.idx & (LZ_FXBIT - 1);
        mine.idx = (g0 + li - 1) | (bi.idx & LZ_FXBIT);
        mine.a = col[li];
      }
      const XRec win = cres_argmin(mine, xbuf, xpar, C, rank);
      xpar ^= 1;
      const double tflip = uq - lq;
      const double tmin = win.idx >= 0 ? __longlong_as_double((long long)win.k) : HUGE_VAL;
      if (tflip <= tmin) {
        if (!(tflip < HUGE_VAL)) { status = MVB_UNBND; break; }
        const double delta = dir > 0 ? tflip : -tflip;
        for (int i = 1 + tid; i <= nloc; i += NT) beta[i] = fma(col[i], delta, beta[i]);
        if (tid == 0) {
          atv[q] = fq == MVB_AT_LB ? MVB_AT_UB : MVB_AT_LB;
          res_tally_to(S->ty, L, mode, 0, q, bland, rank == 0);
          S->degen = 0;
        }
        __syncthreads();
        Key best = key_none();
        for (int j = tid; j < n; j += NT) {
          const Key k = res_price_key(T[j], atv[j], nbv[j], j, false);
          if (key_less(k, best)) best = k;
        }
        best = block_argmin_fast(best, slot, par);
        par ^= 1;
        q = best.idx;
        continue;
      }
      const int r = win.idx & (LZ_FXBIT - 1), kout = win.var;
      const double colr = win.a;
      const int atq = (win.idx & LZ_FXBIT) ? MVB_AT_FX : ((dir > 0 ? colr : -colr) < 0.0 ? MVB_AT_LB : MVB_AT_UB);
      const double inv = 1.0 / colr;
      const double delta = dir > 0 ? tmin : -tmin;
      const double newbr = nb_value(fq, lq, uq) + delta;
      const int owner = (r - 1) / RB;
      const int rloc = owner == (int)rank ? r - g0 + 1 : -1;
      const int degen1 = tmin == 0.0 ? degen0 + 1 : 0;
      for (int i = 1 + tid; i <= nloc; i += NT) {
        beta[i] = i == rloc ? newbr : fma(col[i], delta, beta[i]);
        if (i != rloc) T[(size_t)i * lds + q] = 0.0;  // column q enters the update as 0 (smem_tile_update)
      }
      if (tid == 0) {
        T[q] = 0.0;
        // arm the rho barrier for this iteration's arrival
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(barR)), "r"(rho_bytes) : "memory");
        res_tally_to(S->ty, L, mode, r, q, bland, rank == 0);
        S->degen = degen1;
      }
      if (tid == 32) {  // basis swap: the non-basic side everywhere, the row side at the owner
        nbv[q] = kout;
        atv[q] = atq;
        if (rloc > 0) { head[rloc] = kq; lbB[rloc] = lq; ubB[rloc] = uq; }
      }
      if (rloc > 0) {  // EXCHANGE 2, owner: K3 into the staging buffer, then one bulk copy per CTA
        const double *row = T + (size_t)rloc * lds;
        for (int j = tid; j < ld; j += NT) rhoS[j] = j == q ? inv : -(row[j] * inv);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) cres_send_rho(l, ld, C);
      }
      mbar_wait(barR, rphase);
      rphase ^= 1;
      Key bq = cres_update<true>(l, nloc, n, ld, rloc, q, degen1 >= MVB_BLAND_AFTER, tcl, kout, atq);
      bq = block_argmin_fast(bq, slot, par);  // barrier: end of the update, next entering column
      par ^= 1;
      q = bq.idx;
    }
  } else {
    bool blandsel = false;
    while (true) {
      // EXCHANGE 1: K5a, the leaving row = largest bound violation over all row blocks
      XRec mine{};
      {
        Key best = key_none();
        for (int i = 1 + tid; i <= nloc; i += NT) {
          const double b = beta[i], lo = lbB[i], up = ubB[i];
          const int s = row_infeas(b, lo, up);
          if (!s) continue;
          const double viol = s < 0 ? lo - b : b - up;
          const Key k{blandsel ? 0ULL : ~dbits(viol), head[i], i};
          if (key_less(k, best)) best = k;
        }
        best = block_argmin_fast(best, slot, par);  // barrier: also ends the previous update
        par ^= 1;
        mine.k = best.k;
        mine.var = best.var;
        mine.idx = -1;
        if (best.idx >= 0) {
          mine.idx = g0 + best.idx - 1;
          mine.a = beta[best.idx];
          mine.b = lbB[best.idx];
          mine.c = ubB[best.idx];
        }
      }
      const XRec win = cres_argmin(mine, xbuf, xpar, C, rank);
      xpar ^= 1;
      if (win.idx < 0) { status = MVB_OPT; break; }
      if (S->ty.iters >= L.max_iters) break;
      const int r = win.idx;
      const double br = win.a, lr = win.b, ur = win.c;
      const int rs = row_infeas(br, lr, ur);
      const int owner = (r - 1) / RB;
      const int rloc = owner == (int)rank ? r - g0 + 1 : -1;
      const int degen0 = S->degen;
      const bool bland0 = degen0 >= MVB_BLAND_AFTER;
      if (tid == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(barR)), "r"(rho_bytes) : "memory");
      if (rloc > 0) {
        // EXCHANGE 2, owner: K5b on the row it holds, then K3 + header into the staging buffer
        const double *row = T + (size_t)rloc * lds;
        double bt = HUGE_VAL;
        int bvar = 0x7fffffff, bidx = -1;
        for (int j = tid; j < n; j += NT) {
          const int f = atv[j], var = nbv[j];
          const double alpha = row[j];
          const double a = rs < 0 ? alpha : -alpha;
          const bool ok = ((f == MVB_AT_LB) & (a > MVB_TOL_PIV)) | ((f == MVB_AT_UB) & (a < -MVB_TOL_PIV)) |
                          ((f == MVB_AT_FR) & ((a > MVB_TOL_PIV) | (a < -MVB_TOL_PIV)));
          const double num0 = fabs(T[j]);
          const bool live = ok & (num0 != 0.0);
          const double t = live ? (num0 / (live ? fabs(alpha) : 1.0)) : 0.0;
          const bool better = ok & ((t < bt) | ((t == bt) & (var < bvar)));
          bt = better ? t : bt;
          bvar = better ? var : bvar;
          bidx = better ? (j | (f << 28)) : bidx;
        }
        Key bq = bidx >= 0 ? Key{dbits(bt), bvar, bidx} : key_none();
        bq = block_argmin_fast(bq, slot, par);
        par ^= 1;
        const int q = bq.idx < 0 ? -1 : (bq.idx & 0x0fffffff);
        const double alpha = q >= 0 ? row[q] : 1.0;
        const double inv = 1.0 / alpha;
        for (int j = tid; j < ld; j += NT) rhoS[j] = q < 0 ? 0.0 : (j == q ? inv : -(row[j] * inv));
        if (tid == 0) {
          rhoS[ld + 0] = __longlong_as_double((long long)bq.idx);  // -1: no entering column (primal infeasible)
          rhoS[ld + 1] = __longlong_as_double((long long)bq.var);
          rhoS[ld + 2] = alpha;
          rhoS[ld + 3] = __longlong_as_double((long long)bq.k);     // the ratio's bits (0.0: degenerate)
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) cres_send_rho(l, ld, C);
      }
      mbar_wait(barR, rphase);
      rphase ^= 1;
      const int qidx = (int)__double_as_longlong(rho[ld + 0]);
      if (qidx < 0) { status = MVB_NOFEAS; break; }
      const int q = qidx & 0x0fffffff, fq = (qidx >> 28) & 3, kin = (int)__double_as_longlong(rho[ld + 1]);
      const double alpha = rho[ld + 2];
      const bool degenerate = rho[ld + 3] == 0.0;
      const double lq = ldcg(L.lb + kin), uq = ldcg(L.ub + kin);
      const double target = rs < 0 ? lr : ur;
      const double delta = (target - br) / alpha;
      const double newbr = nb_value(fq, lq, uq) + delta;
      const int degen1 = degenerate ? degen0 + 1 : 0;
      const int kout = win.var;
      const int atq = lr == ur ? MVB_AT_FX : (rs < 0 ? MVB_AT_LB : MVB_AT_UB);
      // pivot column from this CTA's rows (not yet updated), beta
      if (tid == 0) { col[0] = T[q]; T[q] = 0.0; }
      for (int i = 1 + tid; i <= nloc; i += NT) {
        const double v = T[(size_t)i * lds + q];
        col[i] = v;
        beta[i] = i == rloc ? newbr : fma(v, delta, beta[i]);
        if (i != rloc) T[(size_t)i * lds + q] = 0.0;  // column q enters the update as 0 (smem_tile_update)
      }
      if (tid == 32) {
        nbv[q] = kout;
        atv[q] = atq;
        if (rloc > 0) { head[rloc] = kin; lbB[rloc] = lq; ubB[rloc] = uq; }
      }
      if (tid == 0) {
        res_tally_to(S->ty, L, mode, r, q, bland0, rank == 0);
        S->degen = degen1;
      }
      __syncthreads();
      cres_update<false>(l, nloc, n, ld, rloc, q, false, tcl, kout, atq);
      blandsel = degen1 >= MVB_BLAND_AFTER;
    }
  }
  // every CTA leaves the loop in the same iteration (all decisions are replicated)
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();

  // ---- store: this CTA's rows (CTA 0 also the reduced-cost row), small state, outbox ----
  double *dstT = L.T[cT];
  if (tid == 0) {
    const unsigned rowb = (unsigned)ld * 8u;
    for (int i = rank == 0 ? 0 : 1; i <= nloc; i++) {
      const int gi = i == 0 ? 0 : g0 + i - 1;
      bulk_store_row(dstT + (size_t)gi * ld, T + (size_t)i * lds, rowb);
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
  {
    double *gbeta = L.beta[cS];
    int *ghead = L.head[cS], *gnb = L.nb[cS], *gat = L.at[cS];
    for (int i = 1 + tid; i <= nloc; i += NT) {
      const int k = head[i], gi = g0 + i - 1;
      gbeta[gi] = beta[i];
      ghead[gi] = k;
      L.xval[k] = beta[i];
      L.ohead[gi - 1] = k;
    }
    if (rank == 0)
      for (int j = tid; j < n; j += NT) {
        const int k = nbv[j], f = atv[j];
        gnb[j] = k;
        gat[j] = f;
        L.xval[k] = nb_value(f, ldcg(L.lb + k), ldcg(L.ub + k));
        L.onb[j] = k;
        L.oat[j] = f;
      }
  }
  int ninf = 0;
  for (int i = 1 + tid; i <= nloc; i += NT) ninf += row_infeas(beta[i], lbB[i], ubB[i]) != 0;
  ninf = block_sum(ninf, reinterpret_cast<int *>(slot));
  __threadfence();  // xval of every CTA before CTA 0 reads it back for the objective
  ninf = cres_sum(ninf, xbuf, xpar, C, rank);
  xpar ^= 1;
  if (rank == 0 && tid < 32) {
    double p = 0.0;
    for (int j = tid; j < n; j += 32) p = fma(ldcg(L.cost + j), ldcg(L.xval + j), p);
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) p = p + __shfl_down_sync(0xffffffffu, p, off);
    if (tid == 0) {
      const Tally &t = S->ty;
      mvb_result &o = L.res->r;
      o.status = status;
      o.iters = t.iters;
      o.pivots = t.pivots;
      o.flips = t.flips;
      o.iters_primal1 = t.it1;
      o.iters_primal2 = t.it2;
      o.iters_dual = t.itd;
      o.bland_iters = t.bland;
      o.pivot_hash = t.hash;
      o.obj = p;
      o.primal_feasible = ninf == 0;
      o.reserved = 0;
      L.res->curT = cT;
      L.res->curS = cS;
    }
  }
  if (tid == 0) {
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    mbar_inval(barL);
    mbar_inval(barR);
  }
  // nobody may run ahead into the next LP (and re-initialise barriers peers could still send to)
  cres_sync();
}

// clusters stride over the LPs of the launch: cluster k solves LPs k, k + nclusters, ...
__global__ void __launch_bounds__(CRES_NT, 1) simplex_cres_kernel(const LPDev *lps, int nlp, int C) {
  __shared__ LPDev L;
  __shared__ CresCtx ctx;
  const unsigned rank = cres_rank();
  for (int k = (int)cres_cluster_id(); k < nlp; k += (int)cres_nclusters()) {
    __syncthreads();
    if (threadIdx.x == 0) L = lps[k];
    __syncthreads();
    if (L.npatch) {  // host-side bound changes since the last solve (bs.cpp:274,282): CTA 0 applies them
      if (rank == 0)
        for (int t = threadIdx.x; t < L.npatch; t += blockDim.x) {
          const BoundPatch p = L.patch[t];
          L.lb[p.var] = p.lb;
          L.ub[p.var] = p.ub;
          if (p.pos >= 0) L.at[L.curS][p.pos] = p.flag;
        }
      __threadfence();
      cres_sync();
    }
    cres_solve(L, &ctx, C, rank);
  }
}

}  // namespace
