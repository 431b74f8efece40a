// engine_res.cuh -- RESIDENT simplex kernel (north_star "a batched
// variant solves many branch-and-bound node LPs at once"): one CTA per LP, the
// whole tableau lives in that CTA's shared memory for the whole solve.
//
// A B&B node LP (configs[2]: 51 x 208 doubles = 85 KB; + one row per cut) is
// loaded from HBM once (one cp.async.bulk per row, all rows in flight at once,
// mbarrier completion), solved with EAGER rank-1 updates in shared memory --
// no pending updates, no flush passes, no group barrier: the CTA is alone with
// its LP -- and stored back once (cp.async.bulk shared -> global).  DRAM traffic
// of a solve is the tableau in + the tableau out, whatever the pivot count.
// The source of the load may be ANOTHER LP's buffer (LPDev::srcT: the parent's
// final tableau): the copy glp_copy_prob asks for (util.cpp:34, bs.cpp:116) is
// then this load and no clone kernel runs for that child.
//
// The arithmetic is the specification's (oracle/simplex_oracle.c), operation for
// operation, in the eager order: rho_q = 1/p, rho_j = -(T_rj * (1/p)),
// T'_ij = fma(T_iq, rho_j, j == q ? 0 : T_ij), T'_rj = rho_j,
// beta'_i = fma(T_iq, delta, beta_i); reductions are total orders on (value bits,
// variable id).  Modes: primal phase 2 and dual (what B&B children need); an LP
// that needs phase 1 reports MVB_PS_FALLBACK untouched and goes to
// simplex_kernel, as with the private-state kernel.
//
// Per primal iteration three block barriers: [argmin of the ratio test, which
// also publishes the pivot column] [rho + beta + basis swap done] [argmin of the
// NEXT pricing = end of the update: every thread prices the reduced costs it has
// just written].  Per dual iteration four.
#pragma once

namespace {

constexpr int RES_NT = 256;  // threads per CTA of the resident kernel

struct ResLay {
  unsigned lds;  // row stride of the tableau in shared memory, doubles (ld + 2: rows stay 16-byte aligned for
                 // the bulk copies, and a column walk is spread over the banks)
  unsigned oT, oCol, oBeta, oLb, oUb, oRho, oHead, oNb, oAt, oSlot, oBar, total;
};
__host__ __device__ inline unsigned res_up16(unsigned x) { return (x + 15u) & ~15u; }
__host__ __device__ inline ResLay res_layout(int m, int ld) {
  ResLay l;
  const unsigned R = (unsigned)m + 1u, L = (unsigned)ld;
  l.lds = L + 2u;
  unsigned off = 0;
  l.oT = off;    off = res_up16(off + R * l.lds * 8u);
  l.oCol = off;  off = res_up16(off + (R + 1u) * 8u);
  l.oBeta = off; off = res_up16(off + (R + 1u) * 8u);
  l.oLb = off;   off = res_up16(off + (R + 1u) * 8u);
  l.oUb = off;   off = res_up16(off + (R + 1u) * 8u);
  l.oRho = off;  off = res_up16(off + L * 8u);
  l.oHead = off; off = res_up16(off + (R + 1u) * 4u);
  l.oNb = off;   off = res_up16(off + L * 4u);
  l.oAt = off;   off = res_up16(off + L * 4u);
  l.oSlot = off; off = res_up16(off + (unsigned)sizeof(Key) * 64u);
  l.oBar = off;  off = res_up16(off + 16u);
  l.total = off;
  return l;
}
// 0 when the tableau is too large for one CTA's shared memory (or for 32-bit offsets)
__host__ __device__ inline size_t res_smem_bytes(int m, int ld) {
  if ((size_t)(m + 1) * (size_t)(ld + 2) * 8u > (size_t)1 << 20) return 0;
  return res_layout(m, ld).total;
}

__device__ __forceinline__ void bulk_store_row(double *dst, const double *src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)),
               "r"(bytes)
               : "memory");
}

// per-solve scalars every thread reads after a barrier
struct ResCtx {
  Tally ty;
  int degen, status, stop;
  double lq, uq;
};

// K1 on the reduced-cost row T[0][*] (rule of lz_price / price_primal)
__device__ __forceinline__ Key res_price_key(double dn, int f, int var, int j, bool bland) {
  const unsigned flip = f == MVB_AT_LB ? 0x80000000u : 0u;
  const unsigned keep = f == MVB_AT_FR ? 0x7fffffffu : 0xffffffffu;
  const double sv = __hiloint2double((int)(((unsigned)__double2hiint(dn) & keep) ^ flip), __double2loint(dn));
  const bool elig = (f != MVB_AT_FX) & (sv > MVB_TOL_DJ);
  return elig ? Key{bland ? 0ULL : ~dbits(sv), var, j} : key_none();
}

__device__ int res_price(const ResLay &l, int n, bool bland, int par) {
  const double *d = reinterpret_cast<const double *>(dyn_smem + l.oT);
  const int *nbv = reinterpret_cast<const int *>(dyn_smem + l.oNb), *atv = reinterpret_cast<const int *>(dyn_smem + l.oAt);
  Key best = key_none();
  for (int j = threadIdx.x; j < n; j += blockDim.x) {
    const Key k = res_price_key(d[j], atv[j], nbv[j], j, bland);
    if (key_less(k, best)) best = k;
  }
  best = block_argmin_fast(best, reinterpret_cast<Key(*)[32]>(dyn_smem + l.oSlot), par);
  return best.idx;
}

// K5a: leaving row of the dual simplex (rule of lz_dual_select)
__device__ int res_dual_select(const ResLay &l, int m, bool bland, int par) {
  const double *beta = reinterpret_cast<const double *>(dyn_smem + l.oBeta);
  const double *lbB = reinterpret_cast<const double *>(dyn_smem + l.oLb), *ubB = reinterpret_cast<const double *>(dyn_smem + l.oUb);
  const int *head = reinterpret_cast<const int *>(dyn_smem + l.oHead);
  Key best = key_none();
  for (int i = 1 + threadIdx.x; i <= m; i += blockDim.x) {
    const double b = beta[i], lo = lbB[i], up = ubB[i];
    const int s = row_infeas(b, lo, up);
    if (!s) continue;
    const double viol = s < 0 ? lo - b : b - up;
    const Key k{bland ? 0ULL : ~dbits(viol), head[i], i};
    if (key_less(k, best)) best = k;
  }
  best = block_argmin_fast(best, reinterpret_cast<Key(*)[32]>(dyn_smem + l.oSlot), par);
  return best.idx;
}

// K4 in shared memory on a tile of rows 0..nrow (row 0 = reduced costs): T'[i][j] = fma(col[i],
// rho[j], j == q ? 0 : T[i][j]), row rloc <- rho (rloc < 0: the pivot row is not in this tile).
// Thread (tx, ty) owns the column pairs {ch * CW + 2 tx} and the rows ty, ty + trows, ...  The
// main loop is the plain fma on every element (six instructions per row and column pair).  The two
// special cases of the specification are taken out of it: column q reads its old value as 0 --
// the CALLER has stored 0.0 into T[i][q] of every row but the pivot row after copying the column
// out (each row by the thread that copied it), so the plain fma yields exactly fma(c, rho_q, 0);
// the pivot row becomes rho -- rewritten after the loop by the thread that owns that element.
// PRICE: the threads that own row 0 price their columns for the next iteration (position q is
// taken from registers: kout / atq) and return their best key.
template <bool PRICE>
__device__ __forceinline__ Key smem_tile_update(double *T, unsigned lds, int nrow, int n, int ld, const double *col,
                                                const double *rhoS, const int *nbv, const int *atv, int rloc,
                                                int q, bool bland, int tcl, int kout, int atq) {
  const int tcols = 1 << tcl, trows = (int)blockDim.x >> tcl;
  const int tx = threadIdx.x & (tcols - 1), ty = threadIdx.x >> tcl;
  const int CW = 2 * tcols;
  const size_t step = (size_t)trows * lds;
  Key best = key_none();
  for (int c0 = 2 * tx; c0 < ld; c0 += CW) {
    const double2 rho = *reinterpret_cast<const double2 *>(rhoS + c0);
    double *p = T + (size_t)ty * lds + c0;
    int i = ty;
    for (; i + 3 * trows <= nrow; i += 4 * trows, p += 4 * step) {
      double2 t[4];
      double c[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        t[u] = *reinterpret_cast<const double2 *>(p + u * step);
        c[u] = col[i + u * trows];
      }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        double2 o;
        o.x = fma(c[u], rho.x, t[u].x);
        o.y = fma(c[u], rho.y, t[u].y);
        *reinterpret_cast<double2 *>(p + u * step) = o;
      }
    }
    for (; i <= nrow; i += trows, p += step) {
      const double2 t = *reinterpret_cast<const double2 *>(p);
      const double c = col[i];
      double2 o;
      o.x = fma(c, rho.x, t.x);
      o.y = fma(c, rho.y, t.y);
      *reinterpret_cast<double2 *>(p) = o;
    }
    // the pivot row becomes rho
    if (rloc >= 0 && (rloc - ty) % trows == 0 && rloc >= ty)
      *reinterpret_cast<double2 *>(T + (size_t)rloc * lds + c0) = rho;
    if (PRICE && ty == 0) {  // this thread wrote T[0][c0], T[0][c0 + 1]
      const double2 d = *reinterpret_cast<const double2 *>(T + c0);
      if (c0 < n) {
        const bool isq = c0 == q;
        const Key k = res_price_key(d.x, isq ? atq : atv[c0], isq ? kout : nbv[c0], c0, bland);
        if (key_less(k, best)) best = k;
      }
      if (c0 + 1 < n) {
        const bool isq = c0 + 1 == q;
        const Key k = res_price_key(d.y, isq ? atq : atv[c0 + 1], isq ? kout : nbv[c0 + 1], c0 + 1, bland);
        if (key_less(k, best)) best = k;
      }
    }
  }
  return best;
}

template <bool PRICE>
__device__ __forceinline__ Key res_update(const ResLay &l, int m, int n, int ld, int r, int q, bool bland,
                                          int tcl, int kout, int atq) {
  return smem_tile_update<PRICE>(reinterpret_cast<double *>(dyn_smem + l.oT), l.lds, m, n, ld,
                                 reinterpret_cast<const double *>(dyn_smem + l.oCol),
                                 reinterpret_cast<const double *>(dyn_smem + l.oRho),
                                 reinterpret_cast<const int *>(dyn_smem + l.oNb),
                                 reinterpret_cast<const int *>(dyn_smem + l.oAt), r, q, bland, tcl, kout, atq);
}

// bookkeeping of one iteration (as tally_iter); log: this thread also writes the pivot log
__device__ __forceinline__ void res_tally_to(Tally &t, const LPDev &L, int mode, int r, int q, bool bland, bool log) {
  unsigned long long h = t.hash;
  h = (h ^ (unsigned long long)mode) * MVB_FNV_PRIME;
  h = (h ^ (unsigned long long)r) * MVB_FNV_PRIME;
  h = (h ^ (unsigned long long)q) * MVB_FNV_PRIME;
  t.hash = h;
  if (log && L.pivlog && t.iters < L.pivlog_cap) {
    L.pivlog[2 * t.iters] = r;
    L.pivlog[2 * t.iters + 1] = q;
  }
  t.iters++;
  if (r) t.pivots++; else t.flips++;
  if (mode == MVB_MODE_PRIMAL2) t.it2++; else t.itd++;
  if (bland) t.bland++;
}
__device__ __forceinline__ void res_tally(ResCtx *S, const LPDev &L, int mode, int r, int q, bool bland) {
  res_tally_to(S->ty, L, mode, r, q, bland, true);
}

// One whole LP solve by ONE CTA with the tableau in shared memory.
__device__ void res_solve(const LPDev &L, ResCtx *S) {
  const int m = L.m, n = L.n, ld = L.ld, tid = threadIdx.x, NT = blockDim.x;
  const ResLay l = res_layout(m, ld);
  const unsigned lds = l.lds;
  double *T = reinterpret_cast<double *>(dyn_smem + l.oT);
  double *col = reinterpret_cast<double *>(dyn_smem + l.oCol), *beta = reinterpret_cast<double *>(dyn_smem + l.oBeta);
  double *lbB = reinterpret_cast<double *>(dyn_smem + l.oLb), *ubB = reinterpret_cast<double *>(dyn_smem + l.oUb);
  double *rho = reinterpret_cast<double *>(dyn_smem + l.oRho);
  int *head = reinterpret_cast<int *>(dyn_smem + l.oHead);
  int *nbv = reinterpret_cast<int *>(dyn_smem + l.oNb), *atv = reinterpret_cast<int *>(dyn_smem + l.oAt);
  Key(*slot)[32] = reinterpret_cast<Key(*)[32]>(dyn_smem + l.oSlot);
  unsigned long long *bar = reinterpret_cast<unsigned long long *>(dyn_smem + l.oBar);
  const int cT = L.curT, cS = L.curS;
  int tcl = 0;
  while ((1 << tcl) < (ld + 1) / 2 && (2 << tcl) <= NT) tcl++;

  // ---- load: every row of the tableau by one bulk copy, all in flight at once ----
  const double *srcT = L.srcT ? L.srcT : L.T[cT];
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.proxy.async;" ::: "memory");
    const unsigned rowb = (unsigned)ld * 8u;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(rowb * (unsigned)(m + 1))
                 : "memory");
    for (int i = 0; i <= m; i++)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       smem_u32(T + (size_t)i * lds)),
                   "l"(srcT + (size_t)i * ld), "r"(rowb), "r"(smem_u32(bar))
                   : "memory");
    S->ty = Tally();
    S->degen = 0;
    S->status = MVB_ITLIM;
    S->stop = 0;
  }
  // small state while the rows are in flight
  {
    const int *gnb = L.nb[cS], *gat = L.at[cS], *ghead = L.head[cS];
    for (int j = tid; j < ld; j += NT) {
      const bool in = j < n;
      nbv[j] = in ? ldcg(gnb + j) : 0x7fffffff;
      atv[j] = in ? ldcg(gat + j) : MVB_AT_FX;
    }
    for (int i = tid; i <= m + 1; i += NT) {
      const bool in = i >= 1 && i <= m;
      const int k = in ? ldcg(ghead + i) : 0;
      head[i] = in ? k : 0x7fffffff;
      lbB[i] = in ? ldcg(L.lb + k) : -HUGE_VAL;
      ubB[i] = in ? ldcg(L.ub + k) : HUGE_VAL;
      beta[i] = 0.0;
    }
    // values of the non-basic variables (scratch: rho)
    for (int j = tid; j < ld; j += NT) {
      double v = 0.0;
      if (j < n) {
        const int k = ldcg(gnb + j);
        v = nb_value(ldcg(gat + j), ldcg(L.lb + k), ldcg(L.ub + k));
      }
      rho[j] = v;
    }
  }
  __syncthreads();
  mbar_wait(bar, 0);
  int mode;
  if (L.carry) {
    // second leg of a solve whose phase 1 ran in the general kernel (see solve_lp)
    for (int i = 1 + tid; i <= m; i += NT) beta[i] = ldcg(L.beta[cS] + i);
    if (tid == 0) {
      Tally &t = S->ty;
      t.iters = L.c_iters; t.pivots = L.c_pivots; t.flips = L.c_flips; t.it1 = L.c_it1;
      t.it2 = L.c_it2; t.itd = L.c_itd; t.bland = L.c_bland; t.hash = L.c_hash;
    }
    mode = MVB_MODE_PRIMAL2;
    __syncthreads();
  } else {
    // beta_i = det_dot(T_i, v): one warp per row, lane-strided fma chains + halving tree
    const int lane = tid & 31, w = tid >> 5, nw = NT >> 5;
    for (int i = 1 + w; i <= m; i += nw) {
      const double *row = T + (size_t)i * lds;
      double p = 0.0;
      for (int j = lane; j < n; j += 32) p = fma(row[j], rho[j], p);
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) p = p + __shfl_down_sync(0xffffffffu, p, off);
      if (lane == 0) beta[i] = p;
    }
    __syncthreads();
    int pinf = 0, dinf = 0;
    for (int i = 1 + tid; i <= m; i += NT) pinf += row_infeas(beta[i], lbB[i], ubB[i]) != 0;
    for (int j = tid; j < n; j += NT) dinf += dj_eligible(atv[j], T[j]);
    // both counts in one reduction (red[] of block_sum: the argmin slots)
    int *red = reinterpret_cast<int *>(slot);
    pinf = block_sum(pinf, red);
    dinf = block_sum(dinf, red);
    mode = pinf == 0 ? MVB_MODE_PRIMAL2 : dinf == 0 ? MVB_MODE_DUAL : MVB_MODE_PRIMAL1;
    __syncthreads();
  }
  if (mode == MVB_MODE_PRIMAL1) {  // needs the general kernel: nothing has been changed
    if (tid == 0) {
      L.res->r.status = MVB_PS_FALLBACK;
      mbar_inval(bar);
    }
    __syncthreads();
    return;
  }

  int status = MVB_ITLIM;
  int par = 0;  // which half of the argmin slots the next reduction uses
  if (mode == MVB_MODE_PRIMAL2) {
    int q = res_price(l, n, false, par);
    par ^= 1;
    while (true) {
      if (q < 0) { status = MVB_OPT; break; }
      if (S->ty.iters >= L.max_iters) break;
      const int degen0 = S->degen;
      const bool bland = degen0 >= MVB_BLAND_AFTER;
      const double dq = T[q];
      const int dir = dq < 0.0 ? 1 : -1;
      const int fq = atv[q], kq = nbv[q];
      const double lq = ldcg(L.lb + kq), uq = ldcg(L.ub + kq);
      // pivot column copy + K2 ratio test (branch-free primal_ratio_row, phase 2), as lz_primal_iter
      double bt = HUGE_VAL;
      int bvar = 0x7fffffff, bidx = -1;
      if (tid == 0) col[0] = dq;
      for (int i = 1 + tid; i <= m; i += NT) {
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
      const double tflip = uq - lq;
      const double tmin = bi.idx >= 0 ? __longlong_as_double((long long)bi.k) : HUGE_VAL;
      if (tflip <= tmin) {
        if (!(tflip < HUGE_VAL)) { status = MVB_UNBND; break; }
        // bound flip: no basis change
        const double delta = dir > 0 ? tflip : -tflip;
        for (int i = 1 + tid; i <= m; i += NT) beta[i] = fma(col[i], delta, beta[i]);
        if (tid == 0) {
          atv[q] = fq == MVB_AT_LB ? MVB_AT_UB : MVB_AT_LB;
          res_tally(S, L, mode, 0, q, bland);
          S->degen = 0;
        }
        __syncthreads();
        q = res_price(l, n, false, par);
        par ^= 1;
        continue;
      }
      const int r = bi.idx & (LZ_FXBIT - 1), kout = bi.var;
      const double colr = col[r];
      const int atq = (bi.idx & LZ_FXBIT) ? MVB_AT_FX : ((dir > 0 ? colr : -colr) < 0.0 ? MVB_AT_LB : MVB_AT_UB);
      const double inv = 1.0 / colr;
      const double delta = dir > 0 ? tmin : -tmin;
      const double newbr = nb_value(fq, lq, uq) + delta;
      for (int i = 1 + tid; i <= m; i += NT) {
        beta[i] = i == r ? newbr : fma(col[i], delta, beta[i]);
        if (i != r) T[(size_t)i * lds + q] = 0.0;  // column q enters the update as 0 (smem_tile_update)
      }
      if (tid == 0) T[q] = 0.0;
      {  // K3: scaled pivot row
        const double *row = T + (size_t)r * lds;
        for (int j = tid; j < ld; j += NT) rho[j] = j == q ? inv : -(row[j] * inv);
      }
      const int degen1 = tmin == 0.0 ? degen0 + 1 : 0;
      if (tid == 32 % NT) {  // basis swap incl. the per-row bounds and the non-basic flag
        head[r] = kq;
        lbB[r] = lq;
        ubB[r] = uq;
        nbv[q] = kout;
        atv[q] = atq;
      }
      if (tid == 0) {
        res_tally(S, L, mode, r, q, bland);
        S->degen = degen1;
      }
      __syncthreads();
      Key bq = res_update<true>(l, m, n, ld, r, q, degen1 >= MVB_BLAND_AFTER, tcl, kout, atq);
      bq = block_argmin_fast(bq, slot, par);  // barrier: end of the update, next entering column
      par ^= 1;
      q = bq.idx;
    }
  } else {
    int r = res_dual_select(l, m, false, par);
    par ^= 1;
    while (true) {
      if (r < 0) { status = MVB_OPT; break; }
      if (S->ty.iters >= L.max_iters) break;
      const int degen0 = S->degen;
      const bool bland0 = degen0 >= MVB_BLAND_AFTER;
      const double br = beta[r], lr = lbB[r], ur = ubB[r];
      const int rs = row_infeas(br, lr, ur);
      // K5b: entering column, min |d_j| / |T_rj| over the columns that move x_r toward its
      // violated bound (rule of lz_dual_iter)
      const double *row = T + (size_t)r * lds;
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
      if (bq.idx < 0) { status = MVB_NOFEAS; break; }
      const int q = bq.idx & 0x0fffffff, fq = (bq.idx >> 28) & 3, kin = bq.var;
      const bool degenerate = __longlong_as_double((long long)bq.k) == 0.0;
      const double lq = ldcg(L.lb + kin), uq = ldcg(L.ub + kin);
      const double alpha = row[q];
      const double inv = 1.0 / alpha;
      const double target = rs < 0 ? lr : ur;
      const double delta = (target - br) / alpha;
      const double newbr = nb_value(fq, lq, uq) + delta;
      // pivot column copy + beta + K3; nothing of the tableau is written before the barrier
      if (tid == 0) { col[0] = T[q]; T[q] = 0.0; }
      for (int i = 1 + tid; i <= m; i += NT) {
        const double v = T[(size_t)i * lds + q];
        col[i] = v;
        beta[i] = i == r ? newbr : fma(v, delta, beta[i]);
        if (i != r) T[(size_t)i * lds + q] = 0.0;  // column q enters the update as 0 (smem_tile_update)
      }
      for (int j = tid; j < ld; j += NT) rho[j] = j == q ? inv : -(row[j] * inv);
      const int degen1 = degenerate ? degen0 + 1 : 0;
      if (tid == 32 % NT) {
        const int kout = head[r];
        head[r] = kin;
        lbB[r] = lq;
        ubB[r] = uq;
        nbv[q] = kout;
        atv[q] = lr == ur ? MVB_AT_FX : (rs < 0 ? MVB_AT_LB : MVB_AT_UB);
      }
      if (tid == 0) {
        res_tally(S, L, mode, r, q, bland0);
        S->degen = degen1;
      }
      __syncthreads();
      res_update<false>(l, m, n, ld, r, q, false, tcl, 0, 0);
      // K5a for the next iteration runs on the fresh beta; its barrier also ends the update
      r = res_dual_select(l, m, degen1 >= MVB_BLAND_AFTER, par);
      par ^= 1;
    }
  }
  // every thread's generic-proxy writes of the tableau before the async-proxy reads of the store
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();

  // ---- store: tableau rows by bulk copies, small state + outbox by the threads ----
  double *dstT = L.T[cT];
  if (tid == 0) {
    const unsigned rowb = (unsigned)ld * 8u;
    for (int i = 0; i <= m; i++) bulk_store_row(dstT + (size_t)i * ld, T + (size_t)i * lds, rowb);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
  {
    double *gbeta = L.beta[cS];
    int *ghead = L.head[cS], *gnb = L.nb[cS], *gat = L.at[cS];
    for (int i = 1 + tid; i <= m; i += NT) {
      const int k = head[i];
      gbeta[i] = beta[i];
      ghead[i] = k;
      L.xval[k] = beta[i];
      L.ohead[i - 1] = k;
    }
    for (int j = tid; j < n; j += NT) {
      const int k = nbv[j], f = atv[j];
      gnb[j] = k;
      gat[j] = f;
      const double v = nb_value(f, ldcg(L.lb + k), ldcg(L.ub + k));
      L.xval[k] = v;
      L.onb[j] = k;
      L.oat[j] = f;
      if (k < n) rho[k] = v;  // x by structural id, for the objective (scratch)
    }
    for (int i = 1 + tid; i <= m; i += NT)
      if (head[i] < n) rho[head[i]] = beta[i];
  }
  __syncthreads();
  int ninf = 0;
  for (int i = 1 + tid; i <= m; i += NT) ninf += row_infeas(beta[i], lbB[i], ubB[i]) != 0;
  ninf = block_sum(ninf, reinterpret_cast<int *>(slot));
  if (tid < 32) {
    double p = 0.0;
    for (int j = tid; j < n; j += 32) p = fma(ldcg(L.cost + j), rho[j], p);
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
      // the rows must have left shared memory before the next LP of this CTA lands in it
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      mbar_inval(bar);
    }
  }
  __syncthreads();
}

// grid-stride over the LPs of the launch: CTA b solves LPs b, b + gridDim.x, ...
__global__ void __launch_bounds__(RES_NT, 2) simplex_res_kernel(const LPDev *lps, int nlp) {
  __shared__ LPDev L;
  __shared__ ResCtx ctx;
  for (int k = blockIdx.x; k < nlp; k += gridDim.x) {
    __syncthreads();
    if (threadIdx.x == 0) L = lps[k];
    __syncthreads();
    if (L.npatch) {  // host-side bound changes since the last solve (bs.cpp:274,282)
      for (int t = threadIdx.x; t < L.npatch; t += blockDim.x) {
        const BoundPatch p = L.patch[t];
        L.lb[p.var] = p.lb;
        L.ub[p.var] = p.ub;
        if (p.pos >= 0) L.at[L.curS][p.pos] = p.flag;
      }
      __syncthreads();
    }
    res_solve(L, &ctx);
  }
}

}  // namespace
