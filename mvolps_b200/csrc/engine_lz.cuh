// engine_lz.cuh -- "lazy" private-state simplex loop: the fast path of
// engine_cuda.cu for LPs whose O(m)+O(n) state fits in shared memory.
//
// Every CTA of the group keeps its own copy of the small state in shared
// memory (reduced costs d, basic values beta, per-row bounds, basis heads,
// flags) and takes every pivot decision by itself, so the CTAs never exchange
// anything per pivot.  The rank-1 tableau updates are NOT applied one by one:
// the scaled pivot row rho_p and pivot column c_p of up to K pivots stay
// PENDING in shared memory, and what the next pivot needs from the tableau --
// one column, one row -- is read from the stale buffer and brought up to date
// element by element with the same operations the eager update would have
// applied, in the same order:
//     T_{p+1}[i][j] = i == r_p ? rho_p[j] : fma(c_p[i], rho_p[j], j == q_p ? 0 : T_p[i][j])
// so decisions and tableau are bit-identical to the eager algorithm (the
// oracle).  Every K pivots a FLUSH pass streams the tableau once (128-bit
// ld.cg/st.cg, buffer A -> buffer B), applying the K pending updates in
// registers, and the group meets at one barrier.  Tableau traffic per pivot
// drops from 16*R*C bytes to 16*R*C/K, and the per-pivot group barrier of the
// eager kernel becomes one barrier per K pivots.
// The pivot row travels global->shared by cp.async.bulk (TMA, mbarrier
// completion); the pivot column is a strided 8-byte gather (ld.global.cg).
// Phase 1 is not implemented here: such LPs are handed to simplex_kernel.
#pragma once

namespace {

constexpr int PS_CREG = 4;            // pivot-column values a thread can hold in flight
constexpr int LZ_KMAX = 8;            // most pending pivots between two flush passes
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


#ifdef MVB_PROF
// per-CTA phase clocks (thread 0): built only with -DMVB_PROF, read by mvb_debug_prof
__device__ long long g_prof[256 * 16];
#define PROF(i)                                                     \
  do {                                                              \
    if (threadIdx.x == 0) {                                         \
      const long long t_ = clock64();                               \
      g_prof[blockIdx.x * 16 + (i)] += t_ - Z.prof_last;            \
      Z.prof_last = t_;                                             \
    }                                                               \
  } while (0)
#else
#define PROF(i) do {} while (0)
#endif

}  // namespace

extern __shared__ __align__(128) unsigned char dyn_smem[];

namespace {

// Shared-memory layout of one CTA, as byte offsets from dyn_smem (all
// accesses go through the extern __shared__ symbol, so they compile to LDS/STS):
//   n-vectors (nv bytes = ld doubles) : d, rho[0..K)
//   int n-vectors                     : nbv, atv             (together nv bytes)
//   m-vectors (mv bytes = m+2 doubles, 1-based; col[p][0] = d_q) :
//                                       col[0..K], beta, lbB, ubB
//   head (m+2 ints), reduction slots, mbarrier
// col has K+1 slots: slot np receives the column of the pivot being decided.
struct Lay {
  unsigned nv, mv;
  int K;
};
__host__ __device__ inline unsigned lz_nv(int ld) { return (unsigned)ld * 8u; }
__host__ __device__ inline unsigned lz_mv(int m) { return ((unsigned)(m + 2) * 8u + 15u) & ~15u; }
__host__ __device__ inline size_t lz_smem_bytes(int m, int ld, int K) {
  return (size_t)lz_nv(ld) * (K + 2) + (size_t)lz_mv(m) * (K + 4) + (size_t)(m + 2) * 4 + 16 +
         sizeof(Key) * 64 + 64;
}
// largest K <= LZ_KMAX that fits `avail` bytes (0: does not fit at all)
__host__ __device__ inline int lz_pick_k(int m, int ld, size_t avail) {
  for (int K = LZ_KMAX; K >= 1; K--)
    if (lz_smem_bytes(m, ld, K) <= avail) return K;
  return 0;
}

__device__ __forceinline__ double *zD(const Lay &l) { return reinterpret_cast<double *>(dyn_smem); }
__device__ __forceinline__ double *zRho(const Lay &l, int p) {
  return reinterpret_cast<double *>(dyn_smem + (size_t)(1 + p) * l.nv);
}
__device__ __forceinline__ int *zNbv(const Lay &l) {
  return reinterpret_cast<int *>(dyn_smem + (size_t)(1 + l.K) * l.nv);
}
__device__ __forceinline__ int *zAtv(const Lay &l) {
  return reinterpret_cast<int *>(dyn_smem + (size_t)(1 + l.K) * l.nv + l.nv / 2);
}
__device__ __forceinline__ unsigned char *zM(const Lay &l) { return dyn_smem + (size_t)(2 + l.K) * l.nv; }
__device__ __forceinline__ double *zCol(const Lay &l, int p) {
  return reinterpret_cast<double *>(zM(l) + (size_t)p * l.mv);
}
__device__ __forceinline__ double *zBeta(const Lay &l) { return zCol(l, l.K + 1); }
__device__ __forceinline__ double *zLbB(const Lay &l) { return zCol(l, l.K + 2); }
__device__ __forceinline__ double *zUbB(const Lay &l) { return zCol(l, l.K + 3); }
__device__ __forceinline__ int *zHead(const Lay &l) { return reinterpret_cast<int *>(zCol(l, l.K + 4)); }
__device__ __forceinline__ unsigned char *zTail(const Lay &l, int m) {
  return zM(l) + (size_t)(l.K + 4) * l.mv + (((size_t)(m + 2) * 4 + 15) & ~(size_t)15);
}
__device__ __forceinline__ Key (*zSlot(const Lay &l, int m))[32] {
  return reinterpret_cast<Key(*)[32]>(zTail(l, m));
}
__device__ __forceinline__ unsigned long long *zMbar(const Lay &l, int m) {
  return reinterpret_cast<unsigned long long *>(zTail(l, m) + sizeof(Key) * 64);
}

// rows pr[p] = zPend[2p], columns pq[p] = zPend[2p+1] of the pending pivots
__device__ __forceinline__ int *zPend(const Lay &l, int m) {
  return reinterpret_cast<int *>(zTail(l, m) + sizeof(Key) * 64 + 16);
}

// per-thread context of the lazy loop (identical values in every thread)
struct LZ {
  Lay lay;
  int m, n, ld;
  int g, G, tcl;
  unsigned *ctr;
  unsigned gen;
  int cT;            // tableau buffer the pending updates apply to (the stale one)
  int np;            // pending pivots (their rows / columns: zPend)
  int par;           // block_argmin_fast slot parity
  unsigned mphase;   // mbarrier phase of the next row load
  bool must_wait;    // a flush barrier has been arrived at but not yet waited for
#ifdef MVB_PROF
  long long prof_last;
#endif
};

}  // namespace

namespace {

// one pending update applied to the element (i, j) whose current value is v
__device__ __forceinline__ double lz_step(double v, bool is_r, bool is_q, double ci, double rhoj) {
  const double t = fma(ci, rhoj, is_q ? 0.0 : v);
  return is_r ? rhoj : t;
}

// FLUSH: stream this CTA's share of the tableau Ts -> Td applying the np
// pending updates in order.  Work unit = (row, column chunk of 2*tcols
// doubles) as in tile_update; each thread owns two adjacent columns, keeps the
// rho pairs of all pending pivots in registers and UNROLL rows in flight.
template <int KC>
__device__ __noinline__ void lz_flush_k(const Lay l, const int m, const int ld, const int tcl,
                                       const int g, const int G, const double *__restrict__ Ts,
                                       double *__restrict__ Td) {
  const int R = m + 1;
  const int tcols = 1 << tcl, trows = blockDim.x >> tcl;
  const int tx = threadIdx.x & (tcols - 1), ty = threadIdx.x >> tcl;
  const int CW = 2 * tcols, nch = (ld + CW - 1) / CW;
  const int *pend = zPend(l, m);
  int pr[KC];
#pragma unroll
  for (int p = 0; p < KC; p++) pr[p] = pend[2 * p];
  const long long nunits = (long long)nch * R;
  long long u0 = nunits * g / G;
  const long long u1 = nunits * (g + 1) / G;
  constexpr int UN = KC <= 2 ? 8 : 4;
  while (u0 < u1) {
    const int ch = (int)(u0 / R);
    const int ia = (int)(u0 - (long long)ch * R);
    const long long segend = u1 < (long long)(ch + 1) * R ? u1 : (long long)(ch + 1) * R;
    const int ib = ia + (int)(segend - u0);
    const int col = ch * CW + 2 * tx;
    if (col < ld) {
      double2 rh[KC];
      unsigned qmask = 0;  // bit 2p: col == q_p, bit 2p+1: col + 1 == q_p
#pragma unroll
      for (int p = 0; p < KC; p++) {
        rh[p] = *reinterpret_cast<const double2 *>(zRho(l, p) + col);
        const int qp = pend[2 * p + 1];
        qmask |= (col == qp ? 1u : 0u) << (2 * p) | (col + 1 == qp ? 1u : 0u) << (2 * p + 1);
      }
      const double *src = Ts + col;
      double *dst = Td + col;
      for (int i = ia + ty; i < ib; i += trows * UN) {
        double2 t[UN];
#pragma unroll
        for (int u = 0; u < UN; u++) {
          const int ii = i + u * trows;
          if (ii < ib) t[u] = ldcg2(src + (size_t)ii * ld);
        }
#pragma unroll
        for (int u = 0; u < UN; u++) {
          const int ii = i + u * trows;
          if (ii < ib) {
            double2 o = t[u];
#pragma unroll
            for (int p = 0; p < KC; p++) {
              const double c = zCol(l, p)[ii];
              const bool isr = ii == pr[p];
              o.x = lz_step(o.x, isr, (qmask >> (2 * p)) & 1u, c, rh[p].x);
              o.y = lz_step(o.y, isr, (qmask >> (2 * p + 1)) & 1u, c, rh[p].y);
            }
            stcg2(dst + (size_t)ii * ld, o);
          }
        }
      }
    }
    u0 = segend;
  }
}

// flush + arrive at the group barrier; the wait is deferred to the next read
// of the tableau (lz_need_tableau)
__device__ __forceinline__ void lz_flush(LZ &Z, const LPDev &L) {
  if (Z.np == 0) return;
  const double *Ts = L.T[Z.cT];
  double *Td = L.T[Z.cT ^ 1];
  switch (Z.np) {
    case 1: lz_flush_k<1>(Z.lay, Z.m, Z.ld, Z.tcl, Z.g, Z.G, Ts, Td); break;
    case 2: lz_flush_k<2>(Z.lay, Z.m, Z.ld, Z.tcl, Z.g, Z.G, Ts, Td); break;
    case 3: lz_flush_k<3>(Z.lay, Z.m, Z.ld, Z.tcl, Z.g, Z.G, Ts, Td); break;
    case 4: lz_flush_k<4>(Z.lay, Z.m, Z.ld, Z.tcl, Z.g, Z.G, Ts, Td); break;
    case 5: lz_flush_k<5>(Z.lay, Z.m, Z.ld, Z.tcl, Z.g, Z.G, Ts, Td); break;
    case 6: lz_flush_k<6>(Z.lay, Z.m, Z.ld, Z.tcl, Z.g, Z.G, Ts, Td); break;
    case 7: lz_flush_k<7>(Z.lay, Z.m, Z.ld, Z.tcl, Z.g, Z.G, Ts, Td); break;
    default: lz_flush_k<8>(Z.lay, Z.m, Z.ld, Z.tcl, Z.g, Z.G, Ts, Td); break;
  }
  group_arrive(Z.ctr, Z.G, Z.gen);
  Z.cT ^= 1;
  Z.np = 0;
  Z.must_wait = true;
}
__device__ __forceinline__ void lz_need_tableau(LZ &Z) {
  if (Z.must_wait) {
    group_wait(Z.ctr, Z.G, Z.gen);
    Z.must_wait = false;
  }
}

}  // namespace

namespace {

// column q of the CURRENT tableau (rows 0..m) into col slot np: strided gather
// from the stale buffer, then the pending updates in order
__device__ __forceinline__ void lz_issue_col(const LZ &Z, const double *T, int q,
                                             double (&raw)[PS_CREG]) {
#pragma unroll
  for (int u = 0; u < PS_CREG; u++) {
    const int i = threadIdx.x + u * blockDim.x;
    if (i <= Z.m) raw[u] = ldcg(T + (size_t)i * Z.ld + q);
  }
}
__device__ __forceinline__ void lz_finish_col(const LZ &Z, int q, double (&v)[PS_CREG]) {
  const Lay &l = Z.lay;
  const int *pend = zPend(l, Z.m);
  for (int p = 0; p < Z.np; p++) {
    const double rq = zRho(l, p)[q];
    const int prp = pend[2 * p];
    const bool isq = pend[2 * p + 1] == q;
    const double *cp = zCol(l, p);
#pragma unroll
    for (int u = 0; u < PS_CREG; u++) {
      const int i = threadIdx.x + u * blockDim.x;
      if (i <= Z.m) v[u] = lz_step(v[u], i == prp, isq, cp[i], rq);
    }
  }
  double *dst = zCol(l, Z.np);
#pragma unroll
  for (int u = 0; u < PS_CREG; u++) {
    const int i = threadIdx.x + u * blockDim.x;
    if (i <= Z.m) dst[i] = v[u];
  }
  __syncthreads();
}

// row r of the CURRENT tableau into rho slot np.  The raw row lands by TMA;
// `finish` applies the pending updates in place and, when scale is set,
// turns it into the scaled pivot row rho_q = inv, rho_j = -(v_j * inv).
__device__ __forceinline__ void lz_issue_row(LZ &Z, const double *T, int r) {
  if (threadIdx.x == 0)
    bulk_row_load(zRho(Z.lay, Z.np), T + (size_t)r * Z.ld, (unsigned)(Z.ld * sizeof(double)),
                  zMbar(Z.lay, Z.m));
}
__device__ __forceinline__ void lz_finish_row(LZ &Z, int r, bool scale, int q, double inv) {
  const Lay &l = Z.lay;
  const int *pend = zPend(l, Z.m);
  mbar_wait(zMbar(l, Z.m), Z.mphase);
  Z.mphase ^= 1;
  constexpr int JU = 4;
  double *buf = zRho(l, Z.np);
  for (int j0 = threadIdx.x; j0 < Z.ld; j0 += JU * blockDim.x) {
    double v[JU];
#pragma unroll
    for (int u = 0; u < JU; u++) {
      const int j = j0 + u * blockDim.x;
      if (j < Z.ld) v[u] = buf[j];
    }
    for (int p = 0; p < Z.np; p++) {
      const double cr = zCol(l, p)[r];
      const bool isr = pend[2 * p] == r;
      const int pqp = pend[2 * p + 1];
      const double *rp = zRho(l, p);
#pragma unroll
      for (int u = 0; u < JU; u++) {
        const int j = j0 + u * blockDim.x;
        if (j < Z.ld) v[u] = lz_step(v[u], isr, j == pqp, cr, rp[j]);
      }
    }
#pragma unroll
    for (int u = 0; u < JU; u++) {
      const int j = j0 + u * blockDim.x;
      if (j < Z.ld) buf[j] = scale ? (j == q ? inv : -(v[u] * inv)) : v[u];
    }
  }
  __syncthreads();
}
__device__ __forceinline__ void lz_scale_row(LZ &Z, int q, double inv) {
  double *buf = zRho(Z.lay, Z.np);
  for (int j = threadIdx.x; j < Z.ld; j += blockDim.x) buf[j] = j == q ? inv : -(buf[j] * inv);
  __syncthreads();
}

}  // namespace

namespace {

// The lazy private-state iteration loop of one LP (mode PRIMAL2 or DUAL;
// neither changes mode).  Entered after the common prologue (beta[cS] valid in
// global memory); leaves the final small state in the global arrays of parity
// cS and the final tableau in buffer cT, as the general loop does.
__device__ int lz_loop(const LPDev &L, int g, int G, unsigned *ctr, unsigned &gen, int tcl,
                       int K, int mode, int &cT, int cS, Tally &ty) {
  const int m = L.m, n = L.n, ld = L.ld, tid = threadIdx.x, NT = blockDim.x;
  LZ Z;
  Z.lay = Lay{lz_nv(ld), lz_mv(m), K};
  Z.m = m; Z.n = n; Z.ld = ld; Z.g = g; Z.G = G; Z.tcl = tcl;
  Z.ctr = ctr; Z.gen = gen; Z.cT = cT; Z.np = 0; Z.par = 0; Z.mphase = 0; Z.must_wait = false;
  const Lay &l = Z.lay;
  double *d = zD(l), *beta = zBeta(l), *lbB = zLbB(l), *ubB = zUbB(l);
  int *nbv = zNbv(l), *atv = zAtv(l), *head = zHead(l), *pend = zPend(l, m);
  Key(*slot)[32] = zSlot(l, m);
  if (tid == 0) mbar_init(zMbar(l, m), 1);
  {
    const double *d0 = L.T[cT];
    const int *gnb = L.nb[cS], *gat = L.at[cS], *ghead = L.head[cS];
    const double *gbeta = L.beta[cS];
    for (int j = tid; j < ld; j += NT) {
      const bool in = j < n;  // padding columns never price
      d[j] = in ? ldcg(d0 + j) : 0.0;
      nbv[j] = in ? ldcg(gnb + j) : 0x7fffffff;
      atv[j] = in ? ldcg(gat + j) : MVB_AT_FX;
    }
    for (int i = tid; i <= m + 1; i += NT) {
      const bool in = i >= 1 && i <= m;
      const int k = in ? ldcg(ghead + i) : 0x7fffffff;
      beta[i] = in ? ldcg(gbeta + i) : 0.0;
      head[i] = k;
      lbB[i] = in ? ldcg(L.lb + k) : -HUGE_VAL;
      ubB[i] = in ? ldcg(L.ub + k) : HUGE_VAL;
    }
  }
  __syncthreads();
#ifdef MVB_PROF
  Z.prof_last = clock64();
#endif

  int status, degen = 0;
  while (true) {
    const bool bland = degen >= MVB_BLAND_AFTER;
    const double *T = L.T[Z.cT];
    double *colN = zCol(l, Z.np), *rhoN = zRho(l, Z.np);
    int r = 0, q = -1, newat = MVB_AT_LB;
    double delta = 0.0, newbr = 0.0, lq = 0.0, uq = 0.0;
    bool degenerate = false;
    double raw[PS_CREG];

    if (mode == MVB_MODE_DUAL) {
      // K5a leaving row: largest bound violation, lowest variable id on ties
      Key best = key_none();
      for (int i = 1 + tid; i <= m; i += NT) {
        const double b = beta[i], lo = lbB[i], up = ubB[i];
        const int s = row_infeas(b, lo, up);
        if (!s) continue;
        const double viol = s < 0 ? lo - b : b - up;
        const Key k{bland ? 0ULL : ~dbits(viol), head[i], i};
        if (key_less(k, best)) best = k;
      }
      best = block_argmin_fast(best, slot, Z.par);
      r = best.idx;
      if (r < 0) { status = MVB_OPT; break; }
      if (ty.iters >= L.max_iters) { status = MVB_ITLIM; break; }
      PROF(0);
      lz_need_tableau(Z);
      lz_issue_row(Z, T, r);
      lz_finish_row(Z, r, false, 0, 0.0);
      PROF(1);
      const double br = beta[r], lr = lbB[r], ur = ubB[r];
      const int rs = row_infeas(br, lr, ur);
      // K5b entering column: min |d_j| / |T_rj| over columns that move x_r
      // toward its violated bound, lowest variable id on ties
      Key bq = key_none();
      for (int j = tid; j < n; j += NT) {
        const int f = atv[j];
        if (f == MVB_AT_FX) continue;
        const double alpha = rhoN[j];
        const double a = rs < 0 ? alpha : -alpha;
        const bool ok = (f == MVB_AT_LB && a > MVB_TOL_PIV) || (f == MVB_AT_UB && a < -MVB_TOL_PIV) ||
                        (f == MVB_AT_FR && (a > MVB_TOL_PIV || a < -MVB_TOL_PIV));
        if (!ok) continue;
        const double ratio = fabs(d[j]) / fabs(alpha);
        const Key k{dbits(ratio), nbv[j], j};
        if (key_less(k, bq)) bq = k;
      }
      bq = block_argmin_fast(bq, slot, Z.par);
      q = bq.idx;
      if (q < 0) { status = MVB_NOFEAS; break; }
      PROF(2);
      const double alpha_rq = rhoN[q];
      lz_issue_col(Z, T, q, raw);
      const int kq = nbv[q];
      lq = ldcg(L.lb + kq);
      uq = ldcg(L.ub + kq);
      lz_finish_col(Z, q, raw);
      PROF(3);
      const double target = rs < 0 ? lr : ur;
      delta = (target - br) / alpha_rq;
      newbr = nb_value(atv[q], lq, uq) + delta;
      newat = rs < 0 ? MVB_AT_LB : MVB_AT_UB;
      degenerate = __longlong_as_double((long long)bq.k) == 0.0;
      lz_scale_row(Z, q, 1.0 / alpha_rq);
      PROF(4);
    } else {
      // K1 pricing: largest |d_j| among eligible columns, lowest variable id on
      // ties (Bland: lowest eligible id)
      Key bq = key_none();
      for (int j = tid; j < n; j += NT) {
        const double dj = d[j];
        if (!dj_eligible(atv[j], dj)) continue;
        const Key k{bland ? 0ULL : ~dbits(fabs(dj)), nbv[j], j};
        if (key_less(k, bq)) bq = k;
      }
      bq = block_argmin_fast(bq, slot, Z.par);
      q = bq.idx;
      if (q < 0) { status = MVB_OPT; break; }
      if (ty.iters >= L.max_iters) { status = MVB_ITLIM; break; }
      PROF(0);
      lz_need_tableau(Z);
      lz_issue_col(Z, T, q, raw);
      const int kq = nbv[q];
      lq = ldcg(L.lb + kq);
      uq = ldcg(L.ub + kq);
      lz_finish_col(Z, q, raw);
      PROF(1);
      const int dir = d[q] < 0.0 ? 1 : -1;
      // K2 ratio test over the pivot column, lowest variable id on ties
      Key bi = key_none();
      for (int i = 1 + tid; i <= m; i += NT) {
        double t;
        int fl;
        if (!primal_ratio_row(colN[i], dir, 0, beta[i], lbB[i], ubB[i], t, fl)) continue;
        const Key k{dbits(t), head[i], i};
        if (key_less(k, bi)) bi = k;
      }
      bi = block_argmin_fast(bi, slot, Z.par);
      PROF(2);
      const int fq = atv[q];
      const double tflip = uq - lq;
      const double tmin = bi.idx >= 0 ? __longlong_as_double((long long)bi.k) : HUGE_VAL;
      if (tflip <= tmin) {
        if (!(tflip < HUGE_VAL)) { status = MVB_UNBND; break; }
        r = 0;  // bound flip, no basis change
        delta = dir > 0 ? tflip : -tflip;
        newat = fq == MVB_AT_LB ? MVB_AT_UB : MVB_AT_LB;
      } else {
        r = bi.idx;
        lz_issue_row(Z, T, r);
        double t;
        primal_ratio_row(colN[r], dir, 0, beta[r], lbB[r], ubB[r], t, newat);
        delta = dir > 0 ? tmin : -tmin;
        newbr = nb_value(fq, lq, uq) + delta;
        degenerate = tmin == 0.0;
        lz_finish_row(Z, r, true, q, 1.0 / colN[r]);
      }
      PROF(3);
    }

    // ---- apply the iteration to the private state; the tableau update stays pending ----
    tally_iter(ty, L, g, mode, r, q, bland);
    for (int i = 1 + tid; i <= m; i += NT) beta[i] = i == r ? newbr : fma(colN[i], delta, beta[i]);
    if (r) {
      const double dq = colN[0];
      for (int j = tid; j < ld; j += NT) d[j] = fma(dq, rhoN[j], j == q ? 0.0 : d[j]);
    }
    if (tid == 0) {
      if (r) {
        const int kin = nbv[q], kout = head[r];
        const double lo = lbB[r], uo = ubB[r];
        head[r] = kin;
        lbB[r] = lq;
        ubB[r] = uq;
        nbv[q] = kout;
        atv[q] = lo == uo ? MVB_AT_FX : newat;
        pend[2 * Z.np] = r;
        pend[2 * Z.np + 1] = q;
      } else {
        atv[q] = newat;
      }
    }
    __syncthreads();
    if (r) Z.np++;
    degen = degenerate ? degen + 1 : 0;
    PROF(5);
    if (Z.np == K) {
      lz_flush(Z, L);
      PROF(6);
    }
  }
  lz_flush(Z, L);
  lz_need_tableau(Z);
  gen = Z.gen;
  cT = Z.cT;
  // hand the private state back (every CTA holds the same values)
  if (g == 0) {
    double *gbeta = L.beta[cS];
    int *ghead = L.head[cS], *gnb = L.nb[cS], *gat = L.at[cS];
    for (int i = 1 + tid; i <= m; i += NT) { gbeta[i] = beta[i]; ghead[i] = head[i]; }
    for (int j = tid; j < n; j += NT) { gnb[j] = nbv[j]; gat[j] = atv[j]; }
  }
  group_sync(ctr, G, gen);
  return status;
}

}  // namespace
