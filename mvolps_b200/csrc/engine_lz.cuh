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
// One iteration is ONE fused function with two block barriers (the two argmin
// reductions): lz_primal_iter (column + K2 ratio test | step, row, reduced
// costs, K1 pricing of the next iteration) and lz_dual_iter (row + K5b ratio
// test | column, step, K5a selection of the next leaving row).  Values a later
// phase needs travel in registers or in the argmin payload, never through a
// shared location another thread may already be overwriting.
// Phase 1 is not implemented here: such LPs are handed to simplex_kernel.
//
// Two instances of the loop (lz_loop<STREAM>): simplex_ps_kernel (a group of CTAs per LP or small
// L2-resident batches: the flush pass above, up to LZ_KMAX pending pivots) and
// simplex_ps_stream_kernel (one CTA per LP, tableaux streamed from HBM: up to LZ_KMAX_STREAM pending
// pivots, PEELED flush passes -- the threads that own the column pair of a pending entering column
// stay out of the streaming loop and lz_flush_qpairs re-derives those pairs afterwards, so no warp runs
// the slow chain next to the fast one -- with the rows ahead prefetched into L2, or staged through a
// per-thread cp.async ring where the launch has the shared memory: lz_flush_ring_k).
#pragma once

namespace {

constexpr int PS_CREG = 8;            // pivot-column values a thread can hold in flight
#ifndef MVB_LZ_KMAX
#define MVB_LZ_KMAX 8
#endif
#ifndef MVB_LZ_KMAX_STREAM
#define MVB_LZ_KMAX_STREAM 12
#endif
#ifndef MVB_LZ_UN_DEEP
#define MVB_LZ_UN_DEEP 4
#endif
#ifndef MVB_LZ_PF
#define MVB_LZ_PF 4  // streaming register-path flush: L2 prefetch distance in 4-row blocks (0: none); 2: 14.7, 4: 14.9, 8: 13.8 M LP-pivots/s
#endif
// most pending pivots between two flush passes: simplex_ps_kernel (12 measured there: no gain), and
// the streaming variant simplex_ps_stream_kernel (HBM-bound batches: every pending pivot more divides
// the flush traffic; two qmask bits per pivot limit both to 16)
constexpr int LZ_KMAX = MVB_LZ_KMAX;
constexpr int LZ_KMAX_STREAM = MVB_LZ_KMAX_STREAM;
constexpr int LZ_KCAP = LZ_KMAX > LZ_KMAX_STREAM ? LZ_KMAX : LZ_KMAX_STREAM;
static_assert(LZ_KMAX >= 1 && LZ_KMAX <= 8 && LZ_KMAX_STREAM >= LZ_KMAX && LZ_KMAX_STREAM <= 16, "LZ_KMAX");
constexpr int MVB_PS_FALLBACK = 100;  // DevResult status: needs the general kernel
constexpr int MVB_PS_HANDOVER = 101;  // DevResult status of the general kernel: phase 1 has reached a primal
                                      // feasible basis, phase 2 goes to the private-state kernel


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
__device__ __forceinline__ Key block_argmin_fast(Key v, Key (*slot)[32], int par) {
  v = warp_argmin(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  if (lane == 0) slot[par][w] = v;
  __syncthreads();
  v = lane < nw ? slot[par][lane] : key_none();
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
  // order earlier generic-proxy accesses of dst before the async-proxy write (the source rows in
  // global memory, written with st.cg by the last flush pass, are fenced once per flush: lz_need_tableau)
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_inval(unsigned long long *bar) {
  asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
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
#define PROF(i)                                             \
  do {                                                      \
    if (threadIdx.x == 0) {                                 \
      const long long t_ = clock64();                       \
      S->prof[(i)] += t_ - S->prof_last;                    \
      S->prof_last = t_;                                    \
    }                                                       \
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
  int ring;  // flush reads staged through a per-thread cp.async ring of this many 16-byte slots (0: registers)
};
__host__ __device__ inline unsigned lz_nv(int ld) { return (unsigned)ld * 8u; }
__host__ __device__ inline unsigned lz_mv(int m) { return ((unsigned)(m + 2) * 8u + 15u) & ~15u; }
__host__ __device__ inline size_t lz_smem_bytes(int m, int ld, int K) {
  return (size_t)lz_nv(ld) * (K + 2) + (size_t)lz_mv(m) * (K + 4) + (size_t)(m + 2) * 4 + 16 +
         sizeof(Key) * 64 + 16 /* mbarrier */ + 8 * LZ_KCAP /* (row, column) of the pending pivots */;
}
// the cp.async ring of the streaming flush follows the layout above (16-byte aligned)
#ifndef MVB_LZ_RING_UN
#define MVB_LZ_RING_UN 4
#endif
constexpr int LZ_RING_UN = MVB_LZ_RING_UN;  // rows per block of the ring (3, which fits next to depth 11: 14.5 against 14.9 M
                                            // LP-pivots/s of depth 12 without the ring, same file)
constexpr int LZ_RING_SLOTS = 2 * LZ_RING_UN;  // two blocks per thread: one being consumed, one in flight
__host__ __device__ inline size_t lz_ring_off(int m, int ld, int K) { return (lz_smem_bytes(m, ld, K) + 15) & ~(size_t)15; }
__host__ __device__ inline size_t lz_ring_bytes(int slots, int nt) { return (size_t)slots * nt * 16; }
// largest K <= kmax that fits `avail` bytes (0: does not fit at all)
__host__ __device__ inline int lz_pick_k(int m, int ld, size_t avail, int kmax = LZ_KMAX) {
  for (int K = kmax; K >= 1; K--)
    if (lz_smem_bytes(m, ld, K) <= avail) return K;
  return 0;
}

__device__ __forceinline__ double *zD(const Lay &l) { return reinterpret_cast<double *>(dyn_smem); }
__device__ __forceinline__ double *zRho(const Lay &l, int p) {
  return reinterpret_cast<double *>(dyn_smem + (unsigned)(1 + p) * l.nv);
}
__device__ __forceinline__ int *zNbv(const Lay &l) {
  return reinterpret_cast<int *>(dyn_smem + (unsigned)(1 + l.K) * l.nv);
}
__device__ __forceinline__ int *zAtv(const Lay &l) {
  return reinterpret_cast<int *>(dyn_smem + (unsigned)(1 + l.K) * l.nv + l.nv / 2);
}
__device__ __forceinline__ unsigned char *zM(const Lay &l) { return dyn_smem + (unsigned)(2 + l.K) * l.nv; }
__device__ __forceinline__ double *zCol(const Lay &l, int p) {
  return reinterpret_cast<double *>(zM(l) + (unsigned)p * l.mv);
}
__device__ __forceinline__ double *zBeta(const Lay &l) { return zCol(l, l.K + 1); }
__device__ __forceinline__ double *zLbB(const Lay &l) { return zCol(l, l.K + 2); }
__device__ __forceinline__ double *zUbB(const Lay &l) { return zCol(l, l.K + 3); }
__device__ __forceinline__ int *zHead(const Lay &l) { return reinterpret_cast<int *>(zCol(l, l.K + 4)); }
__device__ __forceinline__ unsigned char *zTail(const Lay &l, int m) {
  return zM(l) + (unsigned)(l.K + 4) * l.mv + (((unsigned)(m + 2) * 4u + 15u) & ~15u);
}
__device__ __forceinline__ Key (*zSlot(const Lay &l, int m))[32] {
  return reinterpret_cast<Key(*)[32]>(zTail(l, m));
}
__device__ __forceinline__ unsigned long long *zMbar(const Lay &l, int m) {
  return reinterpret_cast<unsigned long long *>(zTail(l, m) + sizeof(Key) * 64);
}

__device__ __forceinline__ unsigned char *zRing(const Lay &l, int m, int ld) { return dyn_smem + lz_ring_off(m, ld, l.K); }

// rows pr[p] = zPend[2p], columns pq[p] = zPend[2p+1] of the pending pivots
__device__ __forceinline__ int *zPend(const Lay &l, int m) {
  return reinterpret_cast<int *>(zTail(l, m) + sizeof(Key) * 64 + 16);
}

// Context of one solve, in static shared memory.  The first block is written
// once; the evolving fields are written by thread 0 between barriers and read
// by everyone after the next barrier.  The phases below are separate
// (noinline) functions so that each gets its own register allocation.
struct LZS {
  const double *lb, *ub;  // bounds by variable id (global)
  double *T[2];           // ping-pong tableau buffers (global)
  int *pivlog;
  unsigned *ctr;
  Lay lay;
  int m, n, ld, g, G, tcl, mode, max_iters, pivlog_cap;
  // evolving
  unsigned gen, mphase;
  int cT, np, must_wait;
  int r, q, newat, degenerate, degen;
  double delta, newbr, lq, uq, inv, alpha;
  Tally ty;
#ifdef MVB_PROF
  long long prof_last, prof[16];
#endif
};

}  // namespace

namespace {

// one pending update applied to the element (i, j) whose current value is v
__device__ __forceinline__ double lz_step(double v, bool is_r, bool is_q, double ci, double rhoj) {
  const double t = fma(ci, rhoj, is_q ? 0.0 : v);
  return is_r ? rhoj : t;
}

// FLUSH: stream this CTA's share of the tableau Ts -> Td applying the KC
// pending updates in order.  Work unit = (row, column chunk of 2*tcols
// doubles) as in tile_update; each thread owns two adjacent columns, keeps the
// rho pairs of all pending pivots in registers and UN rows in flight.
// one (row, double2) element through the KC pending updates
template <int KC, bool SPECIAL>
__device__ __forceinline__ double2 lz_chain(double2 o, const unsigned char *cb, unsigned mv,
                                            const double2 (&rh)[KC], const int (&pr)[KC],
                                            unsigned qmask, int ii) {
#pragma unroll
  for (int p = 0; p < KC; p++) {
    const double c = *reinterpret_cast<const double *>(cb + p * mv);
    if (SPECIAL) {
      const bool isr = ii == pr[p];
      o.x = lz_step(o.x, isr, (qmask >> (2 * p)) & 1u, c, rh[p].x);
      o.y = lz_step(o.y, isr, (qmask >> (2 * p + 1)) & 1u, c, rh[p].y);
    } else {
      o.x = fma(c, rh[p].x, o.x);
      o.y = fma(c, rh[p].y, o.y);
    }
  }
  return o;
}

// The column pairs that hold a pending entering column q_p need the special cases of lz_step in
// EVERY row.  Left to the threads that own them, they would drag their whole warp through the slow
// chain next to the fast one (divergence: with 8 pending columns two warps in three, measured as
// 3.3x the instructions of the pass), so the owners sit the streaming loop out and the CTA
// re-derives those few pairs afterwards, rows spread over all threads: rows [ia, ib) of the
// column chunk [c0, c1).
template <int KC>
__device__ __noinline__ void lz_flush_qpairs(const LZS *S, const Lay &l, const double *__restrict__ Ts,
                                                double *__restrict__ Td, int ia, int ib, int c0, int c1) {
  const int m = S->m, ld = S->ld;
  const int *pend = zPend(l, m);
  int pair[KC], npair = 0;  // distinct q pairs inside this chunk (the same list in every thread)
#pragma unroll
  for (int p = 0; p < KC; p++) {
    const int cq = pend[2 * p + 1] & ~1;
    bool take = cq >= c0 && cq < c1;
#pragma unroll
    for (int e = 0; e < p; e++) take &= (pend[2 * e + 1] & ~1) != cq;
#pragma unroll
    for (int e = 0; e < KC; e++)
      if (e == npair && take) pair[e] = cq;  // (static indexing: the list stays in registers)
    npair += take;
  }
  // rows of one pair after the other, one element per thread and trip.  Each element is a DRAM sector
  // of its own (8 % of the kernel's stall samples sit on these loads, profiles/r2_stream_kernel_summary.txt),
  // but flattening the (pair, row) items over the threads and requesting them into L2 first measured
  // no faster (14.85 against 14.91 M LP-pivots/s): the other CTA of the SM fills the gaps.
  for (int p = 0; p < npair; p++) {
    int cq = 0;
#pragma unroll
    for (int e = 0; e < KC; e++)
      if (e == p) cq = pair[e];
    for (int i = ia + (int)threadIdx.x; i < ib; i += (int)blockDim.x) {
      double2 o = ldcg2(Ts + (size_t)i * ld + cq);
      for (int e = 0; e < KC; e++) {
        const double c = zCol(l, e)[i];
        const double2 rh = *reinterpret_cast<const double2 *>(zRho(l, e) + cq);
        const int qe = pend[2 * e + 1];
        const bool isr = i == pend[2 * e];
        o.x = lz_step(o.x, isr, cq == qe, c, rh.x);
        o.y = lz_step(o.y, isr, cq + 1 == qe, c, rh.y);
      }
      stcg2(Td + (size_t)i * ld + cq, o);
    }
  }
}

// flush stores / loads of the streaming kernel.  A/B switch MVB_STREAM_CS: 1 = evict-first stores (the written
// tableau is not read again before 300 MB of other tableaux have passed through L2): no change, 14.90 against
// 14.91 M LP-pivots/s; 2 = evict-first loads as well: 14.28 M (profiles/r2_runs/r3n_evict_first_ab.txt).  Off.
template <bool STREAMING>
__device__ __forceinline__ void st_flush(double *p, double2 v) {
#ifdef MVB_STREAM_CS
  if (STREAMING) { __stcs(reinterpret_cast<double2 *>(p), v); return; }
#endif
  stcg2(p, v);
}
template <bool STREAMING>
__device__ __forceinline__ double2 ld_flush(const double *p) {
#if defined(MVB_STREAM_CS) && MVB_STREAM_CS > 1
  if (STREAMING) return __ldcs(reinterpret_cast<const double2 *>(p));
#endif
  return ldcg2(p);
}

// PEEL (one CTA per LP): the owners of a q pair sit the loop out, lz_flush_qpairs re-derives those
// pairs afterwards (see there).  A group of CTAs keeps them in the loop: with a dozen rows per CTA
// and chunk the extra pass would cost a dependent L2 round trip per flush.
template <int KC, bool PEEL>
__device__ __noinline__ void lz_flush_k(const LZS *S) {
  const Lay l = S->lay;
  const int m = S->m, ld = S->ld, tcl = S->tcl, g = S->g, G = S->G;
  const double *__restrict__ Ts = S->T[S->cT];
  double *__restrict__ Td = S->T[S->cT ^ 1];
  const int R = m + 1;
  const int tcols = 1 << tcl, trows = blockDim.x >> tcl;
  const int tx = threadIdx.x & (tcols - 1), ty = threadIdx.x >> tcl;
  const int CW = 2 * tcols, nch = (ld + CW - 1) / CW;
  const int *pend = zPend(l, m);
  const unsigned mv = l.mv;
  const unsigned char *colbase = zM(l);
  int pr[KC];
#pragma unroll
  for (int p = 0; p < KC; p++) pr[p] = pend[2 * p];
  // units = (column chunk, row), chunk-major; the group's units are split evenly
  const unsigned nunits = (unsigned)nch * (unsigned)R;  // < 2^31 (checked on the host)
  unsigned u0 = (unsigned)((unsigned long long)nunits * (unsigned)g / (unsigned)G);
  const unsigned u1 = (unsigned)((unsigned long long)nunits * (unsigned)(g + 1) / (unsigned)G);
  // rows in flight per thread: measured flat between 2 and 4 at K = 5, 6 and 8 spill (tools/ab.sh).
  // The L2 prefetch of the rows ahead below is compiled into the PEEL instances only (the streaming
  // kernel, +8.5 % there): in r1, inside the one kernel of that time, it cost the L2-resident path 4 %.
  // A 32-48 KB TMA (bulk) ring measured 6-9 % slower than this register path in r1; the per-thread
  // cp.async ring of lz_flush_ring_k needs no barrier and gains 12 % at depth 8.
  constexpr int UN = KC <= 4 ? 8 : KC <= 8 ? 4 : MVB_LZ_UN_DEEP;
  while (u0 < u1) {
    const int ch = (int)(u0 / (unsigned)R);
    const int ia = (int)(u0 - (unsigned)ch * (unsigned)R);
    const unsigned chend = (unsigned)(ch + 1) * (unsigned)R;
    const unsigned segend = u1 < chend ? u1 : chend;
    const int ib = ia + (int)(segend - u0);
    const int col = ch * CW + 2 * tx;
    bool mine = col < ld;
    if (PEEL) {
#pragma unroll
      for (int p = 0; p < KC; p++) mine &= (pend[2 * p + 1] & ~1) != col;
    }
    if (mine) {
      double2 rh[KC];
      unsigned qmask = 0;  // bit 2p: col == q_p, bit 2p+1: col + 1 == q_p (PEEL: never set)
#pragma unroll
      for (int p = 0; p < KC; p++) {
        rh[p] = *reinterpret_cast<const double2 *>(zRho(l, p) + col);
        if (!PEEL) {
          const int qp = pend[2 * p + 1];
          qmask |= (col == qp ? 1u : 0u) << (2 * p) | (col + 1 == qp ? 1u : 0u) << (2 * p + 1);
        }
      }
      const size_t rs = (size_t)trows * ld;  // row step of this thread, in doubles
      const double *src = Ts + (size_t)(ia + ty) * ld + col;
      double *dst = Td + (size_t)(ia + ty) * ld + col;
      const unsigned span = (unsigned)(trows * UN);
      int i = ia + ty;
      for (; i + (UN - 1) * trows < ib; i += trows * UN, src += rs * UN, dst += rs * UN) {
        double2 t[UN];
#if MVB_LZ_PF > 0
        if (PEEL) {  // streaming kernel: the rows MVB_LZ_PF blocks ahead on their way into L2 (costs no registers)
#pragma unroll
          for (int u = 0; u < UN; u++)
            if (i + (UN * MVB_LZ_PF + u) * trows < ib)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(src + rs * (UN * MVB_LZ_PF + u)));
        }
#endif
#pragma unroll
        for (int u = 0; u < UN; u++) t[u] = ld_flush<PEEL>(src + rs * u);
        // pivot rows / columns of the pending updates need the special cases
        // of lz_step; everything else is a plain fma chain
        bool special = qmask != 0;
#pragma unroll
        for (int p = 0; p < KC; p++) special |= (unsigned)(pr[p] - i) < span;
        const unsigned char *cb = colbase + (size_t)i * 8;
        if (!special) {
#pragma unroll
          for (int u = 0; u < UN; u++)
            st_flush<PEEL>(dst + rs * u, lz_chain<KC, false>(t[u], cb + (size_t)u * trows * 8, mv, rh, pr, qmask, 0));
        } else {
#pragma unroll
          for (int u = 0; u < UN; u++)
            st_flush<PEEL>(dst + rs * u,
                  lz_chain<KC, true>(t[u], cb + (size_t)u * trows * 8, mv, rh, pr, qmask, i + u * trows));
        }
      }
      for (; i < ib; i += trows, src += rs, dst += rs)
        st_flush<PEEL>(dst, lz_chain<KC, true>(ldcg2(src), colbase + (size_t)i * 8, mv, rh, pr, qmask, i));
    }
    if (PEEL) lz_flush_qpairs<KC>(S, l, Ts, Td, ia, ib, ch * CW, min(ch * CW + CW, ld));
    u0 = segend;
  }
}

// ---- the same pass with the reads staged through shared memory by cp.async (LDGSTS) ----
// HBM-streaming batches (one CTA per LP, tableaux far larger than L2): the register path above has
// 4 rows x 16 B per thread in flight and issues the next loads only after the fma chains of the
// current rows, so a CTA sustains ~15 GB/s against a loaded DRAM latency of ~1 us (phase clocks,
// profiles/r2_runs/r2m_prof_batch.txt).  Here every thread keeps a private ring of two 4-row
// blocks in shared memory: the block after next is requested the moment a block has been read out
// of the ring, so 8 x 16 B per thread stay in flight THROUGH the arithmetic and cost no registers.
// A thread reads only what it requested itself (cp.async.wait_group): no barrier, no mbarrier.
__device__ __forceinline__ void cp_async16(unsigned dst, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ double2 lds2(unsigned addr) {
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr) : "memory");
  return v;
}

template <int KC>
__device__ __noinline__ void lz_flush_ring_k(const LZS *S) {
  const Lay l = S->lay;
  const int m = S->m, ld = S->ld, tcl = S->tcl, g = S->g, G = S->G;
  const double *__restrict__ Ts = S->T[S->cT];
  double *__restrict__ Td = S->T[S->cT ^ 1];
  const int R = m + 1;
  const int tcols = 1 << tcl, trows = blockDim.x >> tcl;
  const int tx = threadIdx.x & (tcols - 1), ty = threadIdx.x >> tcl;
  const int CW = 2 * tcols, nch = (ld + CW - 1) / CW;
  const int *pend = zPend(l, m);
  const unsigned mv = l.mv;
  const unsigned char *colbase = zM(l);
  int pr[KC];
#pragma unroll
  for (int p = 0; p < KC; p++) pr[p] = pend[2 * p];
  const unsigned nunits = (unsigned)nch * (unsigned)R;
  unsigned u0 = (unsigned)((unsigned long long)nunits * (unsigned)g / (unsigned)G);
  const unsigned u1 = (unsigned)((unsigned long long)nunits * (unsigned)(g + 1) / (unsigned)G);
  constexpr int UN = LZ_RING_UN;
  // slot s of this thread: ring0 + s * sstep (consecutive threads, consecutive 16-byte units)
  const unsigned sstep = blockDim.x * 16u;
  const unsigned ring0 = smem_u32(zRing(l, m, ld)) + threadIdx.x * 16u;
  while (u0 < u1) {
    const int ch = (int)(u0 / (unsigned)R);
    const int ia = (int)(u0 - (unsigned)ch * (unsigned)R);
    const unsigned chend = (unsigned)(ch + 1) * (unsigned)R;
    const unsigned segend = u1 < chend ? u1 : chend;
    const int ib = ia + (int)(segend - u0);
    const int col = ch * CW + 2 * tx;
    bool mine = col < ld;
#pragma unroll
    for (int p = 0; p < KC; p++) mine &= (pend[2 * p + 1] & ~1) != col;  // q pairs: lz_flush_qpairs below
    if (mine) {
      double2 rh[KC];
      constexpr unsigned qmask = 0;
#pragma unroll
      for (int p = 0; p < KC; p++) rh[p] = *reinterpret_cast<const double2 *>(zRho(l, p) + col);
      const size_t rs = (size_t)trows * ld;
      const double *src = Ts + (size_t)(ia + ty) * ld + col;
      double *dst = Td + (size_t)(ia + ty) * ld + col;
      const unsigned span = (unsigned)(trows * UN);
      // prologue: the first two blocks (a commit per block even when it is empty: the group count
      // the waits rely on stays the same for every thread and every tail)
      int iload = ia + ty;
      const double *lsrc = src;
#pragma unroll
      for (int b = 0; b < 2; b++) {
#pragma unroll
        for (int u = 0; u < UN; u++)
          if (iload + u * trows < ib) cp_async16(ring0 + (unsigned)(b * UN + u) * sstep, lsrc + rs * u);
        cp_async_commit();
        iload += trows * UN;
        lsrc += rs * UN;
      }
      unsigned blk = 0;  // which half of the ring holds the block being consumed
      for (int i = ia + ty; i < ib; i += trows * UN, src += rs * UN, dst += rs * UN, blk ^= (unsigned)UN) {
        cp_async_wait<1>();  // all but the newest block have landed: the block of row i is in the ring
        double2 t[UN];
#pragma unroll
        for (int u = 0; u < UN; u++) t[u] = lds2(ring0 + (blk + (unsigned)u) * sstep);
        // its slots are free again: request the block after next (same thread, program order)
#pragma unroll
        for (int u = 0; u < UN; u++)
          if (iload + u * trows < ib) cp_async16(ring0 + (blk + (unsigned)u) * sstep, lsrc + rs * u);
        cp_async_commit();
#ifdef MVB_LZ_RING_PF  // A/B: the rows MVB_LZ_RING_PF blocks beyond the ring on their way into L2: no gain
                       // (13.8 M LP-pivots/s with and without, profiles/r2_runs/r3l_ring_prefetch_ab.txt)
#pragma unroll
        for (int u = 0; u < UN; u++)
          if (iload + (UN * MVB_LZ_RING_PF + u) * trows < ib)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(lsrc + rs * (UN * MVB_LZ_RING_PF + u)));
#endif
        iload += trows * UN;
        lsrc += rs * UN;
        bool special = qmask != 0 || i + (UN - 1) * trows >= ib;  // (a tail block takes the guarded path)
#pragma unroll
        for (int p = 0; p < KC; p++) special |= (unsigned)(pr[p] - i) < span;
        const unsigned char *cb = colbase + (size_t)i * 8;
        if (!special) {
#pragma unroll
          for (int u = 0; u < UN; u++)
            st_flush<true>(dst + rs * u, lz_chain<KC, false>(t[u], cb + (size_t)u * trows * 8, mv, rh, pr, qmask, 0));
        } else {
#pragma unroll
          for (int u = 0; u < UN; u++)
            if (i + u * trows < ib)
              st_flush<true>(dst + rs * u,
                    lz_chain<KC, true>(t[u], cb + (size_t)u * trows * 8, mv, rh, pr, qmask, i + u * trows));
        }
      }
      cp_async_wait<0>();  // (only empty groups are left)
    }
    lz_flush_qpairs<KC>(S, l, Ts, Td, ia, ib, ch * CW, min(ch * CW + CW, ld));
    u0 = segend;
  }
}

// STREAM (simplex_ps_stream_kernel, one CTA per LP, tableaux streamed from HBM): peeled passes, reads
// through the cp.async ring when the launch has room for it, up to LZ_KMAX_STREAM pending pivots.
// Otherwise exactly the pass simplex_ps_kernel has always run.
template <bool STREAM, int KC>
__device__ __forceinline__ void lz_flush_pick(LZS *S) {
  if (STREAM) {
    if (S->lay.ring) lz_flush_ring_k<KC>(S);
    else lz_flush_k<KC, true>(S);
  } else {
#ifdef MVB_PEEL_GROUPS  // A/B: peeled passes for groups of CTAs as well: configs[1] 5.33 against 5.17 us per pivot
                        // (profiles/r2_runs/r2y_peel_groups_ring6_ab.txt), so groups keep the in-loop handling
    lz_flush_k < KC <= LZ_KMAX ? KC : LZ_KMAX, true > (S);
#else
    lz_flush_k < KC <= LZ_KMAX ? KC : LZ_KMAX, false > (S);
#endif
  }
}

// flush + arrive at the group barrier; the wait is deferred to the next read
// of the tableau (lz_need_tableau)
template <bool STREAM>
__device__ __forceinline__ void lz_flush(LZS *S) {
  const int np = S->np;
  if (np == 0) return;
  if (!STREAM || np <= 8) {
    switch (np) {
      case 1: lz_flush_pick<STREAM, 1>(S); break;
      case 2: lz_flush_pick<STREAM, 2>(S); break;
      case 3: lz_flush_pick<STREAM, 3>(S); break;
      case 4: lz_flush_pick<STREAM, 4>(S); break;
      case 5: lz_flush_pick<STREAM, 5>(S); break;
      case 6: lz_flush_pick<STREAM, 6>(S); break;
      case 7: lz_flush_pick<STREAM, 7>(S); break;
      default: lz_flush_pick<STREAM, 8>(S); break;
    }
  } else {
    switch (np) {
      case 9: lz_flush_pick<STREAM, 9>(S); break;
      case 10: lz_flush_pick<STREAM, 10>(S); break;
      case 11: lz_flush_pick<STREAM, 11>(S); break;
#if MVB_LZ_KMAX_STREAM > 12
      case 12: lz_flush_pick<STREAM, 12>(S); break;
      case 13: lz_flush_pick<STREAM, 13>(S); break;
      case 14: lz_flush_pick<STREAM, 14>(S); break;
      case 15: lz_flush_pick<STREAM, 15>(S); break;
      default: lz_flush_pick<STREAM, 16>(S); break;
#else
      default: lz_flush_pick<STREAM, 12>(S); break;
#endif
    }
  }
  __syncthreads();  // all stores of this CTA issued
  if (threadIdx.x == 0) {
    if (S->G > 1) {
      __threadfence();
      atomicAdd(S->ctr, 1u);
    }
    S->gen++;
    S->cT ^= 1;
    S->np = 0;
    S->must_wait = 1;
  }
  __syncthreads();
  PROF(6);
}
// before the first tableau read after a flush: wait for the whole group
__device__ __forceinline__ void lz_need_tableau(LZS *S) {
  if (S->must_wait) {
    if (threadIdx.x == 0 && S->G > 1) {
      const unsigned target = S->gen * (unsigned)S->G;
      unsigned seen;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(S->ctr) : "memory");
      } while (seen < target);
    }
    // thread 0 issues the bulk (async-proxy) row copies: the rows the group has just written through
    // the generic proxy and this thread has acquired must be ordered before them (once per flush)
    if (threadIdx.x == 0) asm volatile("fence.proxy.async.global;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) S->must_wait = 0;
  }
}

}  // namespace

namespace {

// TMA fetch of row r of the stale tableau into rho slot np (thread 0)
// (np comes from the caller's register: S->np may already have been advanced by the tally thread)
__device__ __forceinline__ void lz_issue_row(LZS *S, int r, int np) {
  bulk_row_load(zRho(S->lay, np), S->T[S->cT] + (size_t)r * S->ld,
                (unsigned)(S->ld * sizeof(double)), zMbar(S->lay, S->m));
}

// K1 pricing: largest |d_j| among eligible columns, lowest variable id on
// ties (Bland: lowest eligible id).  Returns q or -1.
__device__ __noinline__ int lz_price(LZS *S) {
  const Lay l = S->lay;
  const int n = S->n, tid = threadIdx.x, NT = blockDim.x;
  const bool bland = S->degen >= MVB_BLAND_AFTER;
  const double *d = zD(l);
  const int *atv = zAtv(l), *nbv = zNbv(l);
  // per thread: best score s = |d_j| of an eligible column (FX / ineligible: -1), ties to the
  // lower variable id; branch-free, NJ columns in flight.  Same order as the (~bits, id) keys.
  double best = 0.0;  // eligible scores exceed MVB_TOL_DJ > 0
  int bj = -1, bvar = 0x7fffffff;
  constexpr int NJ = 4;
  for (int j0 = tid; j0 < n; j0 += NJ * NT) {
#pragma unroll
    for (int u = 0; u < NJ; u++) {
      const int jj = min(j0 + u * NT, n - 1);  // clamped repeats of column n-1 never win a tie
      const double dj = d[jj];
      const int f = atv[jj], v = nbv[jj];
      // s = -d_j at the lower bound, d_j at the upper bound, |d_j| when free: sign-bit arithmetic
      const unsigned flip = f == MVB_AT_LB ? 0x80000000u : 0u;
      const unsigned keep = f == MVB_AT_FR ? 0x7fffffffu : 0xffffffffu;
      const double s = __hiloint2double((int)(((unsigned)__double2hiint(dj) & keep) ^ flip), __double2loint(dj));
      const bool elig = f != MVB_AT_FX && s > MVB_TOL_DJ;
      const double sc = bland ? 1.0 : s;
      const bool better = elig & ((sc > best) | ((sc == best) & (v < bvar)));  // no short circuit: no branches
      best = better ? sc : best;
      bj = better ? jj : bj;
      bvar = better ? v : bvar;
    }
  }
  Key bq = bj >= 0 ? Key{bland ? 0ULL : ~dbits(best), bvar, bj} : key_none();
  PROF(8);
#ifdef MVB_PROF
  {
    Key v = warp_argmin(bq);
    PROF(9);
    Key(*slot)[32] = zSlot(l, S->m);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    if (lane == 0) slot[0][w] = v;
    __syncthreads();
    PROF(10);
    v = lane < nw ? slot[0][lane] : key_none();
    v = warp_argmin(v);
    PROF(11);
    return v.idx;
  }
#endif
  bq = block_argmin_fast(bq, zSlot(l, S->m), 0);
  return bq.idx;
}

// One primal iteration after pricing chose q, fused so that nothing but the
// two argmin barriers and one closing barrier separates the phases:
//   pivot column: rows 1..m gathered from the stale buffer + pending updates
//   (row 0 is the private d_q), the ratio test runs on the rows a thread just
//   derived; block argmin (its barrier also publishes the column); then for a
//   pivot every thread derives the step itself (no broadcast through S):
//   beta update and basis swap while the TMA row is in flight, row derive +
//   scale + reduced-cost update in one pass, tally by thread 0.
//   The same pass prices the NEXT iteration on the fresh reduced costs (rule of
//   lz_price), so its argmin barrier is also the closing barrier: two barriers
//   per pivot.
// Returns r in the low word: r > 0 (pivot done; high word = next entering
// position or -1), 0 (bound flip: S holds the step, lz_apply finishes it) or
// -1 (unbounded).
constexpr int LZ_TALLY_TID = 64;      // bookkeeping thread of the primal iteration (not thread 0: TMA issue)
constexpr int LZ_FXBIT = 0x40000000;  // ratio-test payload: the leaving variable is a fixed one
__device__ __forceinline__ long long lz_pack(int r, int qn) {
  return (long long)(((unsigned long long)(unsigned)qn << 32) | (unsigned)r);
}
__device__ __noinline__ long long lz_primal_iter(LZS *S, int q, int np) {
  const Lay l = S->lay;
  const int m = S->m, ld = S->ld, tid = threadIdx.x, NT = blockDim.x;
  lz_need_tableau(S);
  PROF(12);
  const double *T = S->T[S->cT] + q;
  const size_t lds = (size_t)ld;
  const int *pend = zPend(l, m);
  double *colN = zCol(l, np), *beta = zBeta(l), *lbB = zLbB(l), *ubB = zUbB(l);
  int *head = zHead(l);
  const double dq = zD(l)[q];  // row 0 of the current tableau is the private reduced-cost vector
  const int dir = dq < 0.0 ? 1 : -1;
  const int fq = zAtv(l)[q];
  const int degen0 = S->degen;
  if (tid == 0) {
    const int kq = zNbv(l)[q];
    S->lq = ldcg(S->lb + kq);
    S->uq = ldcg(S->ub + kq);
    colN[0] = dq;
  }
  double bt = HUGE_VAL;
  int bvar = 0x7fffffff, bidx = -1;
  constexpr int NU = 2;  // rows per thread and pass, all computed unconditionally
  for (int i0 = 1 + tid; i0 <= m; i0 += NU * NT) {
    int ii[NU];
    double v[NU];
#pragma unroll
    for (int u = 0; u < NU; u++) {
      ii[u] = min(i0 + u * NT, m);  // clamped: out-of-range lanes redo row m
      v[u] = ldcg(T + (size_t)ii[u] * lds);
    }
    for (int p = 0; p < np; p++) {
      const double rq = zRho(l, p)[q];
      const int prp = pend[2 * p];
      const bool isq = pend[2 * p + 1] == q;
      const double *cp = zCol(l, p);
#pragma unroll
      for (int u = 0; u < NU; u++) v[u] = lz_step(v[u], ii[u] == prp, isq, cp[ii[u]], rq);
    }
#pragma unroll
    for (int u = 0; u < NU; u++) {
      const int i = ii[u];
      if (i0 + u * NT <= m) colN[i] = v[u];
      // ratio test of row i (branch-free primal_ratio_row, phase 2), as the general loop
      const double b = beta[i], lo = lbB[i], up = ubB[i];
      const int hv = head[i];
      const double a = dir > 0 ? v[u] : -v[u];
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
  }
  PROF(1);
  Key bi = bidx >= 0 ? Key{dbits(bt), bvar, bidx} : key_none();
  bi = block_argmin_fast(bi, zSlot(l, m), 1);  // the barrier inside also publishes colN, lq, uq
  PROF(14);
  const double lq = S->lq, uq = S->uq;
  const double tflip = uq - lq;
  const double tmin = bi.idx >= 0 ? __longlong_as_double((long long)bi.k) : HUGE_VAL;
  if (tflip <= tmin) {
    if (!(tflip < HUGE_VAL)) return lz_pack(-1, -1);
    if (tid == 0) {  // bound flip, no basis change
      S->delta = dir > 0 ? tflip : -tflip;
      S->newat = fq == MVB_AT_LB ? MVB_AT_UB : MVB_AT_LB;
      S->degenerate = 0;
      S->r = 0;
      S->q = q;
    }
    __syncthreads();
    return lz_pack(0, -1);
  }
  // payload of the winner: row, the variable basic there, whether that variable is fixed
  const int r = bi.idx & (LZ_FXBIT - 1), kout = bi.var;
  // the leaving variable goes to the bound it runs into (as primal_ratio_row, phase 2)
  const int atq = (bi.idx & LZ_FXBIT) ? MVB_AT_FX
                                      : ((dir > 0 ? colN[r] : -colN[r]) < 0.0 ? MVB_AT_LB : MVB_AT_UB);
  if (tid == 0) lz_issue_row(S, r, np);  // TMA; per-thread ld.cg of the owned columns measured 2 % slower
  // the step, derived by every thread from the published column
  const double colr = colN[r];
  const double inv = 1.0 / colr;
  const double delta = dir > 0 ? tmin : -tmin;
  const double newbr = nb_value(fq, lq, uq) + delta;
  for (int i = 1 + tid; i <= m; i += NT) beta[i] = i == r ? newbr : fma(colN[i], delta, beta[i]);
  if (tid == 32) {  // basis swap incl. the per-row bounds and the non-basic flag
    int *nbv = zNbv(l), *atv = zAtv(l), *pd = zPend(l, m);
    const int kin = nbv[q];
    head[r] = kin;
    lbB[r] = lq;
    ubB[r] = uq;
    nbv[q] = kout;
    atv[q] = atq;
    pd[2 * np] = r;
    pd[2 * np + 1] = q;
  }
  if (tid == LZ_TALLY_TID) {  // bookkeeping (as tally_iter), off thread 0's TMA / division path
    const bool bland = S->degen >= MVB_BLAND_AFTER;
    Tally &t = S->ty;
    unsigned long long h = t.hash;
    h = (h ^ (unsigned long long)S->mode) * MVB_FNV_PRIME;
    h = (h ^ (unsigned long long)r) * MVB_FNV_PRIME;
    h = (h ^ (unsigned long long)q) * MVB_FNV_PRIME;
    t.hash = h;
    if (S->g == 0 && S->pivlog && t.iters < S->pivlog_cap) {
      S->pivlog[2 * t.iters] = r;
      S->pivlog[2 * t.iters + 1] = q;
    }
    t.iters++;
    t.pivots++;
    t.it2++;
    if (bland) t.bland++;
    S->degen = tmin == 0.0 ? S->degen + 1 : 0;
    S->r = r;
    S->q = q;
    S->np = np + 1;
  }
  PROF(2);
  mbar_wait(zMbar(l, m), S->mphase);
  PROF(7);
  constexpr int JU = 4;  // columns per thread and pass, all computed unconditionally
  double *buf = zRho(l, np), *d = zD(l);
  const int *atv = zAtv(l), *nbv = zNbv(l);
  const bool bland = (tmin == 0.0 ? degen0 + 1 : 0) >= MVB_BLAND_AFTER;
  double best = 0.0;  // next pricing: eligible scores exceed MVB_TOL_DJ > 0
  int bj = -1, bv = 0x7fffffff;
  for (int j0 = tid; j0 < ld; j0 += JU * NT) {
    int jj[JU];
    double v[JU];
#pragma unroll
    for (int u = 0; u < JU; u++) {
      jj[u] = min(j0 + u * NT, ld - 1);
      v[u] = buf[jj[u]];
    }
    for (int p = 0; p < np; p++) {
      const double cr = zCol(l, p)[r];
      const bool isr = pend[2 * p] == r;
      const int pqp = pend[2 * p + 1];
      const double *rp = zRho(l, p);
#pragma unroll
      for (int u = 0; u < JU; u++) v[u] = lz_step(v[u], isr, jj[u] == pqp, cr, rp[jj[u]]);
    }
#pragma unroll
    for (int u = 0; u < JU; u++) {
      const int j = jj[u];
      const bool isq = j == q;
      const double rho = isq ? inv : -(v[u] * inv);
      const double dn = fma(dq, rho, isq ? 0.0 : d[j]);
      const bool mine = j0 + u * NT < ld;  // clamped repeats may see d[j] already updated: not priced
      if (mine) {
        buf[j] = rho;
        d[j] = dn;
      }
      // price column j for the next iteration; position q now holds the leaving variable
      // (the swap by thread 32 is not ordered with this read, so it comes from registers)
      const int f = isq ? atq : atv[j], var = isq ? kout : nbv[j];
      const unsigned flip = f == MVB_AT_LB ? 0x80000000u : 0u;
      const unsigned keep = f == MVB_AT_FR ? 0x7fffffffu : 0xffffffffu;
      const double sv = __hiloint2double((int)(((unsigned)__double2hiint(dn) & keep) ^ flip), __double2loint(dn));
      const bool elig = mine & (f != MVB_AT_FX) & (sv > MVB_TOL_DJ);
      const double sc = bland ? 1.0 : sv;
      const bool better = elig & ((sc > best) | ((sc == best) & (var < bv)));
      best = better ? sc : best;
      bj = better ? j : bj;
      bv = better ? var : bv;
    }
  }
  PROF(3);
  Key bq = bj >= 0 ? Key{bland ? 0ULL : ~dbits(best), bv, bj} : key_none();
  bq = block_argmin_fast(bq, zSlot(l, m), 0);  // also the closing barrier of the iteration
  if (tid == 0) S->mphase ^= 1;
  PROF(4);
  return lz_pack(r, bq.idx);
}

// K5a leaving row of the dual simplex: largest bound violation, lowest
// variable id on ties.  Returns r >= 1 (its row load is issued), -1 when the
// basis is primal feasible (optimal), -2 at the iteration limit.
__device__ __noinline__ int lz_dual_select(LZS *S, int issue) {
  const Lay l = S->lay;
  const int m = S->m, tid = threadIdx.x, NT = blockDim.x;
  const bool bland = S->degen >= MVB_BLAND_AFTER;
  const double *beta = zBeta(l), *lbB = zLbB(l), *ubB = zUbB(l);
  const int *head = zHead(l);
  Key best = key_none();
  for (int i = 1 + tid; i <= m; i += NT) {
    const double b = beta[i], lo = lbB[i], up = ubB[i];
    const int s = row_infeas(b, lo, up);
    if (!s) continue;
    const double viol = s < 0 ? lo - b : b - up;
    const Key k{bland ? 0ULL : ~dbits(viol), head[i], i};
    if (key_less(k, best)) best = k;
  }
  best = block_argmin_fast(best, zSlot(l, m), 0);
  const int r = best.idx;
  if (r < 0) return -1;
  if (S->ty.iters >= S->max_iters) return -2;
  if (issue) {
    lz_need_tableau(S);
    if (tid == 0) lz_issue_row(S, r, S->np);
  }
  return r;
}

// One dual iteration for the leaving row r, fused like lz_primal_iter (two
// block barriers):
//   row r by TMA -> pending updates on the thread's own columns (kept in
//   registers) + K5b dual ratio test -> block argmin (payload: entering
//   variable and its flag) -> column q gathered on the thread's own rows; the
//   two tableau elements every thread needs, alpha = T[r][q] and d_q =
//   T[0][q], are derived redundantly by each thread (no broadcast, no race
//   with the in-place updates) -> step, beta update, scaled row + reduced
//   costs, basis swap / tally, and K5a selection of the NEXT leaving row on
//   the fresh beta -> block argmin = closing barrier.
// Returns q in the low word (-1: primal infeasible) and the next leaving row
// in the high word (-1: the basis is primal feasible, optimal).
__device__ __noinline__ long long lz_dual_iter(LZS *S, int r, int np) {
  const Lay l = S->lay;
  const int m = S->m, n = S->n, ld = S->ld, tid = threadIdx.x, NT = blockDim.x;
  lz_need_tableau(S);
  if (tid == 0) lz_issue_row(S, r, np);
  const int *pend = zPend(l, m);
  double *beta = zBeta(l), *lbB = zLbB(l), *ubB = zUbB(l), *d = zD(l);
  int *head = zHead(l);
  const int *atv = zAtv(l), *nbv = zNbv(l);
  const double br = beta[r], lr = lbB[r], ur = ubB[r];
  const int rs = row_infeas(br, lr, ur);
  const int degen0 = S->degen;
  const double *Tst = S->T[S->cT];
  const size_t lds = (size_t)ld;
  double *buf = zRho(l, np), *colN = zCol(l, np);
  mbar_wait(zMbar(l, m), S->mphase);
  PROF(1);
  // ---- row r on the thread's own columns + dual ratio test ----
  // (the kernel serves LPs up to ld = JU * NT columns in one pass; wider ones loop)
  constexpr int JU = 4;
  double bt = HUGE_VAL;
  int bvar = 0x7fffffff, bidx = -1;
  for (int j0 = tid; j0 < ld; j0 += JU * NT) {
    int jj[JU];
    double v[JU];
#pragma unroll
    for (int u = 0; u < JU; u++) {
      jj[u] = min(j0 + u * NT, ld - 1);
      v[u] = buf[jj[u]];
    }
    for (int p = 0; p < np; p++) {
      const double cr = zCol(l, p)[r];
      const bool isr = pend[2 * p] == r;
      const int pqp = pend[2 * p + 1];
      const double *rp = zRho(l, p);
#pragma unroll
      for (int u = 0; u < JU; u++) v[u] = lz_step(v[u], isr, jj[u] == pqp, cr, rp[jj[u]]);
    }
#pragma unroll
    for (int u = 0; u < JU; u++) {
      const int j = jj[u];
      const bool mine = j0 + u * NT < ld;
      if (mine) buf[j] = v[u];  // derived, unscaled: pass 2 reads it back
      const int f = atv[j], var = nbv[j];
      const double alpha = v[u];
      const double a = rs < 0 ? alpha : -alpha;
      const bool ok = mine & (j < n) &
                      (((f == MVB_AT_LB) & (a > MVB_TOL_PIV)) | ((f == MVB_AT_UB) & (a < -MVB_TOL_PIV)) |
                       ((f == MVB_AT_FR) & ((a > MVB_TOL_PIV) | (a < -MVB_TOL_PIV))));
      const double num0 = fabs(d[j]);
      const bool live = ok & (num0 != 0.0);  // 0 / x would take the slow path of the division
      const double t = live ? (num0 / (live ? fabs(alpha) : 1.0)) : 0.0;
      const bool better = ok & ((t < bt) | ((t == bt) & (var < bvar)));
      bt = better ? t : bt;
      bvar = better ? var : bvar;
      bidx = better ? (j | (f << 28)) : bidx;
    }
  }
  PROF(2);
  Key bq = bidx >= 0 ? Key{dbits(bt), bvar, bidx} : key_none();
  bq = block_argmin_fast(bq, zSlot(l, m), 1);
  PROF(14);
  if (bq.idx < 0) return lz_pack(-1, -1);
  const int q = bq.idx & 0x0fffffff, fq = (bq.idx >> 28) & 3, kin = bq.var;
  const bool degenerate = __longlong_as_double((long long)bq.k) == 0.0;
  // ---- column q on the thread's own rows; alpha and d_q derived by every thread ----
  const double lq = ldcg(S->lb + kin), uq = ldcg(S->ub + kin);
  constexpr int NU = 2;
  const double *Tq = Tst + q;
  double xa = ldcg(Tq + (size_t)r * lds), xd = ldcg(Tq);
  // first pass of the row loop below also carries the two extra elements through the chain
  for (int p = 0; p < np; p++) {
    const double rq = zRho(l, p)[q];
    const int prp = pend[2 * p];
    const bool isq = pend[2 * p + 1] == q;
    const double *cp = zCol(l, p);
    xa = lz_step(xa, r == prp, isq, cp[r], rq);
    xd = lz_step(xd, false, isq, cp[0], rq);
  }
  const double alpha = xa, dq = xd;
  const double inv = 1.0 / alpha;
  const double target = rs < 0 ? lr : ur;
  const double delta = (target - br) / alpha;
  const double newbr = nb_value(fq, lq, uq) + delta;
  const int degen1 = degenerate ? degen0 + 1 : 0;
  const bool bland = degen1 >= MVB_BLAND_AFTER;
  if (tid == 0) colN[0] = dq;
  Key best = key_none();
  for (int i0 = 1 + tid; i0 <= m; i0 += NU * NT) {
    int ii[NU];
    double v[NU];
#pragma unroll
    for (int u = 0; u < NU; u++) {
      ii[u] = min(i0 + u * NT, m);
      v[u] = ldcg(Tq + (size_t)ii[u] * lds);
    }
    for (int p = 0; p < np; p++) {
      const double rq = zRho(l, p)[q];
      const int prp = pend[2 * p];
      const bool isq = pend[2 * p + 1] == q;
      const double *cp = zCol(l, p);
#pragma unroll
      for (int u = 0; u < NU; u++) v[u] = lz_step(v[u], ii[u] == prp, isq, cp[ii[u]], rq);
    }
#pragma unroll
    for (int u = 0; u < NU; u++) {
      const int i = ii[u];
      const bool mine = i0 + u * NT <= m;
      const bool isr = i == r;
      const double bn = isr ? newbr : fma(v[u], delta, beta[i]);
      if (mine) {
        colN[i] = v[u];
        beta[i] = bn;
      }
      // K5a for the next iteration; row r now holds the entering variable (from registers:
      // the swap by thread 32 is not ordered with this read)
      const double lo = isr ? lq : lbB[i], up = isr ? uq : ubB[i];
      const int hv = isr ? kin : head[i];
      const int s = row_infeas(bn, lo, up);
      const double viol = s < 0 ? lo - bn : bn - up;
      const Key k{bland ? 0ULL : ~dbits(viol), hv, i};
      if (mine && s != 0 && key_less(k, best)) best = k;
    }
  }
  PROF(3);
  // ---- scaled pivot row + reduced costs on the thread's own columns ----
  for (int j0 = tid; j0 < ld; j0 += JU * NT) {
#pragma unroll
    for (int u = 0; u < JU; u++) {
      const int j = j0 + u * NT;
      if (j < ld) {
        const bool isq = j == q;
        const double rho = isq ? inv : -(buf[j] * inv);
        buf[j] = rho;
        d[j] = fma(dq, rho, isq ? 0.0 : d[j]);
      }
    }
  }
  if (tid == 32) {  // basis swap incl. the per-row bounds and the non-basic flag
    int *nbw = zNbv(l), *atw = zAtv(l), *pd = zPend(l, m);
    const int kout = head[r];
    head[r] = kin;
    lbB[r] = lq;
    ubB[r] = uq;
    nbw[q] = kout;
    atw[q] = lr == ur ? MVB_AT_FX : (rs < 0 ? MVB_AT_LB : MVB_AT_UB);
    pd[2 * np] = r;
    pd[2 * np + 1] = q;
  }
  if (tid == 0) {  // bookkeeping (as tally_iter)
    const bool bl0 = degen0 >= MVB_BLAND_AFTER;
    Tally &t = S->ty;
    unsigned long long h = t.hash;
    h = (h ^ (unsigned long long)S->mode) * MVB_FNV_PRIME;
    h = (h ^ (unsigned long long)r) * MVB_FNV_PRIME;
    h = (h ^ (unsigned long long)q) * MVB_FNV_PRIME;
    t.hash = h;
    if (S->g == 0 && S->pivlog && t.iters < S->pivlog_cap) {
      S->pivlog[2 * t.iters] = r;
      S->pivlog[2 * t.iters + 1] = q;
    }
    t.iters++;
    t.pivots++;
    t.itd++;
    if (bl0) t.bland++;
    S->degen = degen1;
    S->r = r;
    S->q = q;
    S->lq = lq;
    S->uq = uq;
    S->np = np + 1;
  }
  PROF(4);
  best = block_argmin_fast(best, zSlot(l, m), 0);  // also the closing barrier of the iteration
  if (tid == 0) S->mphase ^= 1;
  PROF(5);
  return lz_pack(q, best.idx);
}

// primal bound flip (lz_primal_iter returned 0, the step is in S): no basis change,
//   beta'_i = fma(T_iq, delta, beta_i), the entering variable moves to its other bound
__device__ __noinline__ void lz_flip(LZS *S, int np, int q) {
  const Lay l = S->lay;
  const int m = S->m, tid = threadIdx.x, NT = blockDim.x;
  const double delta = S->delta;
  const double *colN = zCol(l, np);
  double *beta = zBeta(l);
  for (int i = 1 + tid; i <= m; i += NT) beta[i] = fma(colN[i], delta, beta[i]);
  __syncthreads();  // everyone has read the step out of S
  PROF(15);
  if (tid == 32) zAtv(l)[q] = S->newat;
  if (tid == 0) {  // bookkeeping (as tally_iter)
    const bool bland = S->degen >= MVB_BLAND_AFTER;
    Tally &t = S->ty;
    unsigned long long h = t.hash;
    h = (h ^ (unsigned long long)S->mode) * MVB_FNV_PRIME;
    h = (h ^ 0ULL) * MVB_FNV_PRIME;
    h = (h ^ (unsigned long long)q) * MVB_FNV_PRIME;
    t.hash = h;
    if (S->g == 0 && S->pivlog && t.iters < S->pivlog_cap) {
      S->pivlog[2 * t.iters] = 0;
      S->pivlog[2 * t.iters + 1] = q;
    }
    t.iters++;
    t.flips++;
    t.it2++;
    if (bland) t.bland++;
    S->degen = 0;  // a flip is never degenerate
  }
  __syncthreads();
}

}  // namespace

namespace {

// The lazy private-state iteration loop of one LP (mode PRIMAL2 or DUAL;
// neither changes mode).  Entered after the common prologue (beta[cS] valid in
// global memory); leaves the final small state in the global arrays of parity
// cS and the final tableau in buffer cT, as the general loop does.
template <bool STREAM>
__device__ int lz_loop(const LPDev &L, int g, int G, unsigned *ctr, unsigned &gen, int tcl,
                       int K, int ring, int mode, int &cT, int cS, Tally &ty) {
  const int m = L.m, n = L.n, ld = L.ld, tid = threadIdx.x, NT = blockDim.x;
  __shared__ LZS ctx;
  LZS *S = &ctx;
  const Lay l{lz_nv(ld), lz_mv(m), K, ring};
  if (tid == 0) {
    S->lb = L.lb; S->ub = L.ub; S->T[0] = L.T[0]; S->T[1] = L.T[1];
    S->pivlog = L.pivlog; S->pivlog_cap = L.pivlog_cap; S->ctr = ctr; S->lay = l;
    S->m = m; S->n = n; S->ld = ld; S->g = g; S->G = G; S->tcl = tcl; S->mode = mode;
    S->max_iters = L.max_iters;
    S->gen = gen; S->mphase = 0; S->cT = cT; S->np = 0; S->must_wait = 0;
    S->r = 0; S->q = 0; S->newat = 0; S->degenerate = 0; S->degen = 0;
    S->ty = ty;
    mbar_init(zMbar(l, m), 1);
#ifdef MVB_PROF
    for (int i = 0; i < 16; i++) S->prof[i] = 0;
    S->prof_last = clock64();
#endif
  }
  {
    double *d = zD(l), *beta = zBeta(l), *lbB = zLbB(l), *ubB = zUbB(l);
    int *nbv = zNbv(l), *atv = zAtv(l), *head = zHead(l);
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
      const int k = in ? ldcg(ghead + i) : 0;
      beta[i] = in ? ldcg(gbeta + i) : 0.0;
      head[i] = in ? k : 0x7fffffff;
      lbB[i] = in ? ldcg(L.lb + k) : -HUGE_VAL;
      ubB[i] = in ? ldcg(L.ub + k) : HUGE_VAL;
    }
  }
  __syncthreads();

  int status;
  int qnext = -2;  // -2: not priced yet (first iteration, after a bound flip)
  int rnext = -2;  // dual: -2: leaving row not selected yet (first iteration)
  while (true) {
    int r, q;
    const int np = S->np;
    if (mode == MVB_MODE_DUAL) {
      r = rnext != -2 ? rnext : lz_dual_select(S, 0);
      if (r == -1) { status = MVB_OPT; break; }
      if (r == -2 || S->ty.iters >= S->max_iters) { status = MVB_ITLIM; break; }
      PROF(0);
      const long long qr = lz_dual_iter(S, r, np);
      q = (int)(unsigned)qr;
      if (q < 0) { status = MVB_NOFEAS; break; }
      rnext = (int)(qr >> 32);
      if (S->np == K) lz_flush<STREAM>(S);
      continue;
    } else {
      q = qnext != -2 ? qnext : lz_price(S);
      qnext = -2;
      if (q < 0) { status = MVB_OPT; break; }
      if (S->ty.iters >= S->max_iters) { status = MVB_ITLIM; break; }
      PROF(0);
      const long long rq = lz_primal_iter(S, q, np);
      r = (int)(unsigned)rq;
      if (r < 0) { status = MVB_UNBND; break; }
      if (r > 0) {  // pivot: the private state is updated and the next column priced
        qnext = (int)(rq >> 32);
        if (S->np == K) lz_flush<STREAM>(S);
        continue;
      }
    }
    lz_flip(S, np, q);
    PROF(5);
  }
  __syncthreads();
  lz_flush<STREAM>(S);
  lz_need_tableau(S);
  __syncthreads();
  // the next LP of this CTA lays its shared memory out afresh: the barrier object must not stay
  // valid (re-initialising a live mbarrier is undefined)
  if (tid == 0) mbar_inval(zMbar(l, m));
  gen = S->gen;
  cT = S->cT;
  ty = S->ty;
#ifdef MVB_PROF
  if (tid == 0 && blockIdx.x < 256)
    for (int i = 0; i < 16; i++) g_prof[blockIdx.x * 16 + i] += S->prof[i];
#endif
  // hand the private state back (every CTA holds the same values)
  if (g == 0) {
    const double *beta = zBeta(l);
    const int *head = zHead(l), *nbv = zNbv(l), *atv = zAtv(l);
    double *gbeta = L.beta[cS];
    int *ghead = L.head[cS], *gnb = L.nb[cS], *gat = L.at[cS];
    for (int i = 1 + tid; i <= m; i += NT) { gbeta[i] = beta[i]; ghead[i] = head[i]; }
    for (int j = tid; j < n; j += NT) { gnb[j] = nbv[j]; gat[j] = atv[j]; }
  }
  group_sync(ctr, G, gen);
  return status;
}

}  // namespace
