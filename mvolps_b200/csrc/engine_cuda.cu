// engine_cuda.cu -- the LP-relaxation engine on B200 (sm_100a): dense fp64
// bounded-variable tableau simplex, device resident, behind the C ABI of
// include/mvolps_engine.h.  Replaces what the reference reaches through GLPK:
// glp_simplex (bs.cpp:117,279,287), glp_eval_tab_row (gmi.cpp:36) and the
// tableau part of glp_copy_prob (util.cpp:34, bs.cpp:116).
//
// One kernel (`simplex_kernel`) runs whole solves: an LP is owned by a GROUP
// of G co-resident CTAs (G = 1 for batched node LPs, G = all SMs for one large
// tableau).  Per iteration every CTA of the group takes the pivot decisions
// redundantly from L2-resident vectors (pricing / ratio test as warp-shuffle +
// block reductions on a total (value, variable-id) order), then updates its
// share of the tableau from buffer A into buffer B (ping-pong, so that no CTA
// overwrites what another still reads) and the group meets at ONE barrier.
// The arithmetic is the specification of oracle/simplex_oracle.c, operation
// for operation (explicit fma, IEEE divide, -fmad=false), so pivot sequences
// and tableaus are bit-identical to the CPU oracle.
//
// There is no CPU fallback in this file: without a CUDA device every entry
// point fails with an error.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/mvolps_engine.h"
#include "../../include/mvolps_spec.h"
#include "cow_vec.hpp"
#include "par_for.hpp"

#ifndef MVB_PS_NT
#define MVB_PS_NT 512  // threads per CTA of the private-state kernel
#endif

// ------------------------------------------------------------------ device
struct DevResult {
  mvb_result r;
  int curT, curS;  // which ping-pong buffers hold the final state
};

// a bound change made on the host since the last solve (branching, bs.cpp:274,282),
// applied by the solve kernel itself so that it costs no transfer of its own
struct BoundPatch {
  int var, pos, flag, pad;  // pos >= 0: non-basic position whose flag becomes `flag`
  double lb, ub;
};

struct LPDev {
  int m, n, ld, mcap;
  double *T[2];     // (mcap+1) x ld; row 0 = reduced costs
  double *beta[2];  // mcap+1, 1-based
  int *head[2];     // mcap+1, 1-based
  int *nb[2];       // n
  int *at[2];       // n
  double *lb, *ub;  // n+mcap by variable id
  double *cost;     // n
  double *p1part;   // phase-1 partial sums: ceil(mcap/32) x ld
  double *vnb;      // n: values of the non-basic variables (beta rebuild)
  // outbox: one contiguous block copied back after the solve
  DevResult *res;
  double *xval;     // n+mcap
  int *ohead;       // mcap
  int *onb, *oat;   // n
  int *pivlog;      // 2*pivlog_cap or null
  int pivlog_cap;
  int curT, curS;   // parities at entry
  int max_iters;
  int lzK;          // pending-update depth of the lazy loop for this launch
  int lzRing;       // slots of the flush's cp.async ring per thread (0: flush reads through registers)
  int npatch;       // bound patches to apply before the solve
  const BoundPatch *patch;
  // hand-over from the general kernel to the private-state kernel once phase 1 has reached a
  // primal feasible basis: ps_ok = the LP fits the private-state kernel (general kernel: stop
  // there); carry = 1: continue a solve (tallies below, beta[curS] valid, mode primal phase 2)
  const double *srcT;  // resident kernel: load the tableau from here instead of T[curT] (fused clone), or null
  int ps_ok, carry;
  int c_iters, c_pivots, c_flips, c_it1, c_it2, c_itd, c_bland, c_pad;
  unsigned long long c_hash;
};

namespace {

constexpr int STAGE_ROWS = 128;  // pivot-column values staged in smem per pass
constexpr int UNROLL = 8;        // independent 128-bit loads in flight per thread

struct Key {
  unsigned long long k;
  int var;
  int idx;
};

__device__ __forceinline__ Key key_none() { return Key{~0ULL, 0x7fffffff, -1}; }
__device__ __forceinline__ bool key_less(const Key &a, const Key &b) {
  return a.k < b.k || (a.k == b.k && a.var < b.var);
}
__device__ __forceinline__ unsigned long long dbits(double x) {
  return (unsigned long long)__double_as_longlong(x);
}

__device__ __forceinline__ double ldcg(const double *p) { return __ldcg(p); }
__device__ __forceinline__ int ldcg(const int *p) { return __ldcg(p); }
__device__ __forceinline__ double2 ldcg2(const double *p) {
  return __ldcg(reinterpret_cast<const double2 *>(p));
}
__device__ __forceinline__ void stcg2(double *p, double2 v) {
  __stcg(reinterpret_cast<double2 *>(p), v);
}

// warp-shuffle + block reduction to the minimum key; result on every thread
__device__ Key block_argmin(Key v, Key *red) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    Key o;
    o.k = __shfl_xor_sync(0xffffffffu, v.k, off);
    o.var = __shfl_xor_sync(0xffffffffu, v.var, off);
    o.idx = __shfl_xor_sync(0xffffffffu, v.idx, off);
    if (key_less(o, v)) v = o;
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();  // red[] free again
  if (lane == 0) red[w] = v;
  __syncthreads();
  v = lane < nw ? red[lane] : key_none();
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    Key o;
    o.k = __shfl_xor_sync(0xffffffffu, v.k, off);
    o.var = __shfl_xor_sync(0xffffffffu, v.var, off);
    o.idx = __shfl_xor_sync(0xffffffffu, v.idx, off);
    if (key_less(o, v)) v = o;
  }
  return v;
}

__device__ int block_sum(int v, int *red) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  v = lane < nw ? red[lane] : 0;
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

// barrier of the G CTAs that own one LP.  *ctr counts arrivals monotonically
// (zeroed by the host before the launch); generation `gen` is tracked in
// registers.  Data written before the barrier is read after it with ld.cg.
__device__ __forceinline__ void group_sync(unsigned *ctr, int G, unsigned &gen) {
  __syncthreads();
  gen++;
  if (G > 1) {
    if (threadIdx.x == 0) {
      __threadfence();
      atomicAdd(ctr, 1u);
      const unsigned target = gen * (unsigned)G;
      unsigned seen;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(ctr) : "memory");
      } while (seen < target);
    }
    __syncthreads();
  }
}

__device__ __forceinline__ int row_infeas(double b, double l, double u) {
  if (l - b > MVB_TOL_BND * (1.0 + fabs(l))) return -1;
  if (b - u > MVB_TOL_BND * (1.0 + fabs(u))) return 1;
  return 0;
}

__device__ __forceinline__ bool dj_eligible(int f, double dj) {
  if (f == MVB_AT_LB) return dj < -MVB_TOL_DJ;
  if (f == MVB_AT_UB) return dj > MVB_TOL_DJ;
  if (f == MVB_AT_FR) return dj < -MVB_TOL_DJ || dj > MVB_TOL_DJ;
  return false;
}

__device__ __forceinline__ double nb_value(int f, double l, double u) {
  if (f == MVB_AT_LB || f == MVB_AT_FX) return l;
  if (f == MVB_AT_UB) return u;
  return 0.0;
}

}  // namespace

namespace {

struct Smem {
  Key redK[32];
  int redI[32];
  double colS[STAGE_ROWS];
};

// ---- K4: rank-1 tableau update, buffer Ts -> Td -------------------------
// Work unit = (row, column chunk of 2*tcols doubles); the group's units are
// split evenly over its G CTAs.  Each thread owns two adjacent columns
// (128-bit ld.cg/st.cg), keeps its rho pair in registers and streams UNROLL
// rows at a time; the pivot-column values of the rows in flight are staged
// in shared memory.   rho_q = 1/p, rho_j = -(T_rj * (1/p));
//   T'_ij = fma(T_iq, rho_j, j == q ? 0 : T_ij),  T'_rj = rho_j.
__device__ void tile_update(const LPDev &L, const double *__restrict__ Ts,
                            double *__restrict__ Td, int r, int q, double inv, int g, int G,
                            int tcl, Smem &sm) {
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
    const bool act = col < ld;
    const bool qx = col == q, qy = col + 1 == q;
    double2 rho = make_double2(0.0, 0.0);
    if (act) {
      const double2 t = ldcg2(Ts + (size_t)r * ld + col);
      rho.x = qx ? inv : -(t.x * inv);
      rho.y = qy ? inv : -(t.y * inv);
    }
    for (int s0 = ia; s0 < ib; s0 += STAGE_ROWS) {
      const int s1 = ib < s0 + STAGE_ROWS ? ib : s0 + STAGE_ROWS;
      __syncthreads();
      for (int k = threadIdx.x; k < s1 - s0; k += blockDim.x)
        sm.colS[k] = ldcg(Ts + (size_t)(s0 + k) * ld + q);
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

// beta, basis heads and non-basic flags of the next state (ping-pong), split
// over all threads of the group.  r == 0: bound flip (no basis change).
__device__ void state_update(const LPDev &L, int cT, int cS, int r, int q, double delta,
                             double newbr, int kin, int kout, int newat, int g, int G) {
  const double *Ts = L.T[cT];
  const double *bs = L.beta[cS];
  double *bd = L.beta[cS ^ 1];
  const int *hs = L.head[cS], *ns = L.nb[cS], *as = L.at[cS];
  int *hd = L.head[cS ^ 1], *nd = L.nb[cS ^ 1], *ad = L.at[cS ^ 1];
  const int gt = g * blockDim.x + threadIdx.x, stride = G * blockDim.x;
  for (int i = 1 + gt; i <= L.m; i += stride) {
    const double a = ldcg(Ts + (size_t)i * L.ld + q);
    bd[i] = i == r ? newbr : fma(a, delta, ldcg(bs + i));
    hd[i] = i == r ? kin : ldcg(hs + i);
  }
  for (int j = gt; j < L.n; j += stride) {
    nd[j] = j == q ? kout : ldcg(ns + j);
    ad[j] = j == q ? newat : ldcg(as + j);
  }
}

__device__ int count_primal_infeas(const LPDev &L, int cS, Smem &sm) {
  const double *b = L.beta[cS];
  const int *h = L.head[cS];
  int cnt = 0;
  for (int i = 1 + threadIdx.x; i <= L.m; i += blockDim.x) {
    const int k = ldcg(h + i);
    cnt += row_infeas(ldcg(b + i), ldcg(L.lb + k), ldcg(L.ub + k)) != 0;
  }
  return block_sum(cnt, sm.redI);
}

__device__ int count_dual_infeas(const LPDev &L, int cT, int cS, Smem &sm) {
  const double *d = L.T[cT];
  const int *at = L.at[cS];
  int cnt = 0;
  for (int j = threadIdx.x; j < L.n; j += blockDim.x) cnt += dj_eligible(ldcg(at + j), ldcg(d + j));
  return block_sum(cnt, sm.redI);
}

// phase-1 partial sums: p1part[c][j] = sum over the 32 rows of chunk c of
// sign_i * T_ij, rows in order
__device__ void p1_partials(const LPDev &L, int cT, int cS, int g, int G) {
  const double *T = L.T[cT], *b = L.beta[cS];
  const int *h = L.head[cS];
  const int m = L.m, n = L.n, ld = L.ld;
  const int nchunks = (m + MVB_P1_CHUNK - 1) / MVB_P1_CHUNK;
  const long long items = (long long)nchunks * n;
  for (long long it = (long long)g * blockDim.x + threadIdx.x; it < items;
       it += (long long)G * blockDim.x) {
    const int c = (int)(it / n), j = (int)(it - (long long)c * n);
    const int i0 = 1 + c * MVB_P1_CHUNK;
    const int i1 = i0 + MVB_P1_CHUNK - 1 < m ? i0 + MVB_P1_CHUNK - 1 : m;
    double acc = 0.0;
    for (int i = i0; i <= i1; i++) {
      const int k = ldcg(h + i);
      const int s = row_infeas(ldcg(b + i), ldcg(L.lb + k), ldcg(L.ub + k));
      if (s > 0) acc = acc + ldcg(T + (size_t)i * ld + j);
      else if (s < 0) acc = acc - ldcg(T + (size_t)i * ld + j);
    }
    L.p1part[(size_t)c * ld + j] = acc;
  }
}

__device__ __forceinline__ double p1_cost(const LPDev &L, int j) {
  const int nchunks = (L.m + MVB_P1_CHUNK - 1) / MVB_P1_CHUNK;
  double d = 0.0;
  for (int c = 0; c < nchunks; c++) d = d + ldcg(L.p1part + (size_t)c * L.ld + j);
  return d;
}

}  // namespace

namespace {

// per-solve bookkeeping replicated in every thread (identical everywhere)
struct Tally {
  int iters = 0, pivots = 0, flips = 0, it1 = 0, it2 = 0, itd = 0, bland = 0;
  unsigned long long hash = MVB_FNV_OFFSET;
};

__device__ __forceinline__ void tally_iter(Tally &t, const LPDev &L, int g, int mode, int r,
                                           int q, bool bland) {
  unsigned long long h = t.hash;
  h = (h ^ (unsigned long long)mode) * MVB_FNV_PRIME;
  h = (h ^ (unsigned long long)r) * MVB_FNV_PRIME;
  h = (h ^ (unsigned long long)q) * MVB_FNV_PRIME;
  t.hash = h;
  if (g == 0 && threadIdx.x == 0 && L.pivlog && t.iters < L.pivlog_cap) {
    L.pivlog[2 * t.iters] = r;
    L.pivlog[2 * t.iters + 1] = q;
  }
  t.iters++;
  if (r) t.pivots++; else t.flips++;
  if (mode == MVB_MODE_PRIMAL1) t.it1++;
  else if (mode == MVB_MODE_PRIMAL2) t.it2++;
  else t.itd++;
  if (bland) t.bland++;
}

// beta_i = det_dot(T_i, v): one warp per row, lane-strided fma chains and a
// halving tree (the order fixed by the specification)
__device__ void compute_beta(const LPDev &L, int cT, int cS, int g, int G) {
  const double *T = L.T[cT];
  double *beta = L.beta[cS];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int i = 1 + g * nw + w; i <= L.m; i += G * nw) {
    const double *row = T + (size_t)i * L.ld;
    double p = 0.0;
    for (int j = lane; j < L.n; j += 32) p = fma(ldcg(row + j), ldcg(L.vnb + j), p);
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) p = p + __shfl_down_sync(0xffffffffu, p, off);
    if (lane == 0) beta[i] = p;
  }
}

}  // namespace

namespace {

// one row of the primal ratio test (phase 2: s == 0; phase 1: an infeasible
// basic variable blocks only when it reaches its violated bound).  Returns
// whether the row blocks, its step t >= 0 and the flag the leaving variable takes
__device__ __forceinline__ bool primal_ratio_row(double alpha, int dir, int s, double b, double l,
                                                 double u, double &t, int &fl) {
  const double a = dir > 0 ? alpha : -alpha;
  if (a < -MVB_TOL_PIV) {  // basic variable decreases
    if (s < 0) return false;
    const double bound = s > 0 ? u : l;
    if (!(bound > -HUGE_VAL)) return false;
    t = (b - bound) / (-a);
    fl = s > 0 ? MVB_AT_UB : MVB_AT_LB;
  } else if (a > MVB_TOL_PIV) {  // basic variable increases
    if (s > 0) return false;
    const double bound = s < 0 ? l : u;
    if (!(bound < HUGE_VAL)) return false;
    t = (bound - b) / a;
    fl = s < 0 ? MVB_AT_LB : MVB_AT_UB;
  } else {
    return false;
  }
  if (!(t > 0.0)) t = 0.0;
  return true;
}

}  // namespace

#include "engine_lz.cuh"
#include "engine_res.cuh"
#include "engine_cres.cuh"

namespace {

// One whole LP solve by a group of G CTAs (this CTA is number g of it).
// dyn != nullptr: the launch carries enough dynamic shared memory for the
// private-state loop (engine_ps.cuh), used unless the LP needs phase 1.
// PSK: 0 = the general loop below; 1 = the lazy private-state loop (simplex_ps_kernel); 2 = the same
// loop with the streaming flush (simplex_ps_stream_kernel)
template <int PSK>
__device__ void solve_lp(const LPDev &L, int g, int G, unsigned *ctr, unsigned &gen, int tcl,
                         Smem &sm, unsigned char *dyn) {
  const int m = L.m, n = L.n, ld = L.ld, tid = threadIdx.x, NT = blockDim.x;
  int cT = L.curT, cS = L.curS;
  Tally ty;
  int status = MVB_ITLIM, degen = 0;

  int mode;
  if (L.carry) {
    // second leg of a solve whose phase 1 ran in the general kernel: the state is exactly what
    // that loop left (beta is NOT rebuilt: the oracle does not rebuild it at the phase switch)
    ty.iters = L.c_iters; ty.pivots = L.c_pivots; ty.flips = L.c_flips; ty.it1 = L.c_it1;
    ty.it2 = L.c_it2; ty.itd = L.c_itd; ty.bland = L.c_bland; ty.hash = L.c_hash;
    mode = MVB_MODE_PRIMAL2;
  } else {
    // values of the non-basic variables, then beta = T v
    for (int j = g * NT + tid; j < n; j += G * NT) {
      const int k = ldcg(L.nb[cS] + j);
      L.vnb[j] = nb_value(ldcg(L.at[cS] + j), ldcg(L.lb + k), ldcg(L.ub + k));
    }
    group_sync(ctr, G, gen);
    compute_beta(L, cT, cS, g, G);
    group_sync(ctr, G, gen);
    if (count_primal_infeas(L, cS, sm) == 0) mode = MVB_MODE_PRIMAL2;
    else if (count_dual_infeas(L, cT, cS, sm) == 0) mode = MVB_MODE_DUAL;
    else mode = MVB_MODE_PRIMAL1;
  }

  if (PSK) {
    if (mode == MVB_MODE_PRIMAL1) {  // needs the general kernel: nothing has been changed yet
      if (g == 0 && tid == 0) L.res->r.status = MVB_PS_FALLBACK;
      return;
    }
    status = lz_loop<PSK == 2>(L, g, G, ctr, gen, tcl, L.lzK, L.lzRing, mode, cT, cS, ty);
  }

  while (!PSK) {
    const bool bland = degen >= MVB_BLAND_AFTER;
    const double *T = L.T[cT];
    const double *beta = L.beta[cS];
    const int *head = L.head[cS], *nb = L.nb[cS], *at = L.at[cS];
    int r = 0, q = -1, newat = MVB_AT_LB;
    double delta = 0.0, newbr = 0.0;

    if (mode == MVB_MODE_DUAL) {
      // K5a leaving row: largest bound violation, lowest variable id on ties
      Key best = key_none();
      for (int i = 1 + tid; i <= m; i += NT) {
        const int var = ldcg(head + i);
        const double b = ldcg(beta + i), l = ldcg(L.lb + var), u = ldcg(L.ub + var);
        const int s = row_infeas(b, l, u);
        if (!s) continue;
        const double viol = s < 0 ? l - b : b - u;
        const Key k{bland ? 0ULL : ~dbits(viol), var, i};
        if (key_less(k, best)) best = k;
      }
      best = block_argmin(best, sm.redK);
      if (best.idx < 0) { status = MVB_OPT; break; }
      if (ty.iters >= L.max_iters) break;
      r = best.idx;
      const int rvar = best.var;
      const double br = ldcg(beta + r), lr = ldcg(L.lb + rvar), ur = ldcg(L.ub + rvar);
      const int rs = row_infeas(br, lr, ur);
      // K5b entering column: min |d_j| / |T_rj| over columns that move x_r
      // toward its violated bound, lowest variable id on ties
      const double *row = T + (size_t)r * ld;
      Key bq = key_none();
      for (int j = tid; j < n; j += NT) {
        const int f = ldcg(at + j);
        if (f == MVB_AT_FX) continue;
        const double alpha = ldcg(row + j);
        const double a = rs < 0 ? alpha : -alpha;
        const bool ok = (f == MVB_AT_LB && a > MVB_TOL_PIV) || (f == MVB_AT_UB && a < -MVB_TOL_PIV) ||
                        (f == MVB_AT_FR && (a > MVB_TOL_PIV || a < -MVB_TOL_PIV));
        if (!ok) continue;
        const double ratio = fabs(ldcg(T + j)) / fabs(alpha);
        const Key k{dbits(ratio), ldcg(nb + j), j};
        if (key_less(k, bq)) bq = k;
      }
      bq = block_argmin(bq, sm.redK);
      if (bq.idx < 0) { status = MVB_NOFEAS; break; }
      q = bq.idx;
      const double qratio = __longlong_as_double((long long)bq.k);
      const double target = rs < 0 ? lr : ur;
      delta = (target - br) / ldcg(row + q);
      const int kq = ldcg(nb + q);
      newbr = nb_value(ldcg(at + q), ldcg(L.lb + kq), ldcg(L.ub + kq)) + delta;
      newat = rs < 0 ? MVB_AT_LB : MVB_AT_UB;
      degen = qratio == 0.0 ? degen + 1 : 0;
    } else {
      const bool phase1 = mode == MVB_MODE_PRIMAL1;
      if (phase1) {
        if (count_primal_infeas(L, cS, sm) == 0) {
          mode = MVB_MODE_PRIMAL2;
          degen = 0;
          if (L.ps_ok) { status = MVB_PS_HANDOVER; break; }  // phase 2 runs in the private-state kernel
          continue;
        }
        p1_partials(L, cT, cS, g, G);
        group_sync(ctr, G, gen);
      }
      // K1 pricing: largest |d_j| among eligible columns, lowest variable id on
      // ties (Bland: lowest eligible id)
      Key bq = key_none();
      for (int j = tid; j < n; j += NT) {
        const int f = ldcg(at + j);
        if (f == MVB_AT_FX) continue;
        const double dj = phase1 ? p1_cost(L, j) : ldcg(T + j);
        if (!dj_eligible(f, dj)) continue;
        const Key k{bland ? 0ULL : ~dbits(fabs(dj)), ldcg(nb + j), j};
        if (key_less(k, bq)) bq = k;
      }
      bq = block_argmin(bq, sm.redK);
      if (bq.idx < 0) { status = phase1 ? MVB_NOFEAS : MVB_OPT; break; }
      if (ty.iters >= L.max_iters) break;
      q = bq.idx;
      const double dq = phase1 ? p1_cost(L, q) : ldcg(T + q);
      const int dir = dq < 0.0 ? 1 : -1;
      // K2 ratio test over the pivot column, lowest variable id on ties
      Key bi = key_none();
      for (int i = 1 + tid; i <= m; i += NT) {
        const double alpha = ldcg(T + (size_t)i * ld + q);
        const int var = ldcg(head + i);
        const double b = ldcg(beta + i), l = ldcg(L.lb + var), u = ldcg(L.ub + var);
        const int s = phase1 ? row_infeas(b, l, u) : 0;
        double t;
        int fl;
        if (!primal_ratio_row(alpha, dir, s, b, l, u, t, fl)) continue;
        const Key k{dbits(t), var, i};
        if (key_less(k, bi)) bi = k;
      }
      bi = block_argmin(bi, sm.redK);
      const int kq = ldcg(nb + q);
      const double lq = ldcg(L.lb + kq), uq = ldcg(L.ub + kq);
      const int fq = ldcg(at + q);
      const double tflip = uq - lq;
      const double tmin = bi.idx >= 0 ? __longlong_as_double((long long)bi.k) : HUGE_VAL;
      if (tflip <= tmin) {
        if (!(tflip < HUGE_VAL)) { status = phase1 ? MVB_NOFEAS : MVB_UNBND; break; }
        r = 0;  // bound flip, no basis change
        delta = dir > 0 ? tflip : -tflip;
        newat = fq == MVB_AT_LB ? MVB_AT_UB : MVB_AT_LB;
        degen = 0;
      } else {
        r = bi.idx;
        const double alpha = ldcg(T + (size_t)r * ld + q);
        const double b = ldcg(beta + r), l = ldcg(L.lb + bi.var), u = ldcg(L.ub + bi.var);
        double t;
        primal_ratio_row(alpha, dir, phase1 ? row_infeas(b, l, u) : 0, b, l, u, t, newat);
        delta = dir > 0 ? tmin : -tmin;
        newbr = nb_value(fq, lq, uq) + delta;
        degen = tmin == 0.0 ? degen + 1 : 0;
      }
    }

    // ---- apply the iteration: next small state, then the rank-1 update ----
    const int kin = ldcg(nb + q);
    int kout = kin;
    if (r) {
      kout = ldcg(head + r);
      if (ldcg(L.lb + kout) == ldcg(L.ub + kout)) newat = MVB_AT_FX;
    }
    tally_iter(ty, L, g, mode, r, q, bland);
    state_update(L, cT, cS, r, q, delta, newbr, kin, kout, newat, g, G);
    if (r) {
      const double inv = 1.0 / ldcg(T + (size_t)r * ld + q);
      tile_update(L, T, L.T[cT ^ 1], r, q, inv, g, G, tcl, sm);
      cT ^= 1;
    }
    cS ^= 1;
    group_sync(ctr, G, gen);
  }

  // ---- finish: primal values by variable id, basis copy, objective ----
  const double *beta = L.beta[cS];
  const int *head = L.head[cS], *nb = L.nb[cS], *at = L.at[cS];
  for (int j = g * NT + tid; j < n; j += G * NT) {
    const int k = ldcg(nb + j), f = ldcg(at + j);
    L.xval[k] = nb_value(f, ldcg(L.lb + k), ldcg(L.ub + k));
    L.onb[j] = k;
    L.oat[j] = f;
  }
  for (int i = 1 + g * NT + tid; i <= m; i += G * NT) {
    const int k = ldcg(head + i);
    L.xval[k] = ldcg(beta + i);
    L.ohead[i - 1] = k;
  }
  group_sync(ctr, G, gen);
  if (g == 0) {
    const int ninf = count_primal_infeas(L, cS, sm);
    if (tid < 32) {
      double p = 0.0;
      for (int j = tid; j < n; j += 32) p = fma(ldcg(L.cost + j), ldcg(L.xval + j), p);
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) p = p + __shfl_down_sync(0xffffffffu, p, off);
      if (tid == 0) {
        mvb_result &o = L.res->r;
        o.status = status;
        o.iters = ty.iters;
        o.pivots = ty.pivots;
        o.flips = ty.flips;
        o.iters_primal1 = ty.it1;
        o.iters_primal2 = ty.it2;
        o.iters_dual = ty.itd;
        o.bland_iters = ty.bland;
        o.pivot_hash = ty.hash;
        o.obj = p;
        o.primal_feasible = ninf == 0;
        o.reserved = 0;
        L.res->curT = cT;
        L.res->curS = cS;
      }
    }
  }
}

// grid = ngroups * G CTAs (all co-resident: cooperative launch); group k owns
// LPs k, k + ngroups, ...
template <int PSK>
__device__ __forceinline__ void kernel_body(const LPDev *lps, int nlp, int G, unsigned *ctrs,
                                            int tcl, unsigned char *dyn) {
  __shared__ Smem sm;
  __shared__ LPDev L;
  const int group = blockIdx.x / G, g = blockIdx.x - group * G, ngroups = gridDim.x / G;
  unsigned gen = 0;
  for (int k = group; k < nlp; k += ngroups) {
    __syncthreads();
    if (threadIdx.x == 0) L = lps[k];
    __syncthreads();
    int t = tcl;
    if (t < 0) {  // per-LP thread shape: columns = pow2 >= ld/2, capped at the block
      t = 0;
      while ((1 << t) < (L.ld + 1) / 2 && (2 << t) <= (int)blockDim.x) t++;
    }
    if (L.npatch) {  // host-side bound changes since the last solve
      if (g == 0)
        for (int k = threadIdx.x; k < L.npatch; k += blockDim.x) {
          const BoundPatch p = L.patch[k];
          L.lb[p.var] = p.lb;
          L.ub[p.var] = p.ub;
          if (p.pos >= 0) L.at[L.curS][p.pos] = p.flag;
        }
      group_sync(ctrs + group, G, gen);
    }
    solve_lp<PSK>(L, g, G, ctrs + group, gen, t, sm, dyn);
  }
}

// grid = ngroups * G CTAs (all co-resident: cooperative launch); group k owns
// LPs k, k + ngroups, ...
__global__ void __launch_bounds__(512, 1)
simplex_kernel(const LPDev *lps, int nlp, int G, unsigned *ctrs, int tcl) {
  kernel_body<0>(lps, nlp, G, ctrs, tcl, nullptr);
}

// private-state variant (engine_ps.cuh): every LP of the launch fits the
// dynamic shared memory; LPs that turn out to need phase 1 come back with
// status MVB_PS_FALLBACK and are re-run by simplex_kernel.
__global__ void __launch_bounds__(MVB_PS_NT, 1)
simplex_ps_kernel(const LPDev *lps, int nlp, int G, unsigned *ctrs, int tcl) {
  kernel_body<1>(lps, nlp, G, ctrs, tcl, dyn_smem);
}

// streaming variant: launches of one CTA per LP whose tableaux do not stay in L2 (configs[3]-shape
// batches, B&B rounds of large node LPs).  A kernel of its own so that the code of the
// L2-resident path above stays exactly what it was (its per-pivot time moved by 2-8 % with the
// placement of unrelated functions in the same kernel, profiles/r2_runs/r2q_ab*.txt).
__global__ void __launch_bounds__(MVB_PS_NT, 1)
simplex_ps_stream_kernel(const LPDev *lps, int nlp, int G, unsigned *ctrs, int tcl) {
  kernel_body<2>(lps, nlp, G, ctrs, tcl, dyn_smem);
}

}  // namespace

// ------------------------------------------------------------------ host
namespace {

thread_local std::string g_err;
std::atomic<unsigned long long> g_launches{0};

struct CloneJob {
  const char *src;
  char *dst;
  unsigned long long bytes;  // multiple of 256
};

// every queued clone in one launch: blockIdx.y = job, blockIdx.x strides the 16-byte units
__global__ void __launch_bounds__(256) clone_kernel(const CloneJob *jobs) {
  const CloneJob j = jobs[blockIdx.y];
  const uint4 *__restrict__ s = reinterpret_cast<const uint4 *>(j.src);
  uint4 *__restrict__ d = reinterpret_cast<uint4 *>(j.dst);
  const size_t n16 = j.bytes / 16, step = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * step < n16; i += 4 * step) {
    const uint4 a = s[i], b = s[i + step], c = s[i + 2 * step], e = s[i + 3 * step];
    d[i] = a; d[i + step] = b; d[i + 2 * step] = c; d[i + 3 * step] = e;
  }
  for (; i < n16; i += step) d[i] = s[i];
}

// a queued cut row: tableau row m+1 of L (L.m = rows before the append) from the dense
// coefficients at `a_off` doubles into the coefficient arena
struct AppendJob {
  LPDev L;
  int cT, cS;
  unsigned long long a_off;
  double lb, ub;  // bounds of the new auxiliary variable (id n + m)
};

// one row of one tableau into the staging buffer (K7 of a whole B&B round)
struct GatherJob {
  const double *src;
  double *dst;
  int n, pad;
};

struct DeviceCtx {
  bool ready = false;
  int device = -1, sms = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t stream_upd = nullptr;  // sharded path: the tableau updates (engine_shard.cuh)
  int max_ctas = 0;        // co-resident CTAs of simplex_kernel at NT threads
  int nt = 512;            // threads per CTA (general kernel)
  int nt_ps = MVB_PS_NT;   // threads per CTA (private-state kernel)
  unsigned *ctrs = nullptr; int ctrs_cap = 0;
  LPDev *dlps = nullptr; int dlps_cap = 0;
  char *stage = nullptr; size_t stage_cap = 0;   // device staging (row gathers, new rows)
  void *h_stage = nullptr; size_t h_stage_cap = 0;  // pinned host staging handed out by mvb_host_staging
  // per-launch arenas: bound patches go up and all outboxes come back in ONE copy each
  char *d_out = nullptr, *h_out = nullptr; size_t out_cap = 0;
  BoundPatch *d_patch = nullptr, *h_patch = nullptr; size_t patch_cap = 0;
  // recycled LP slabs by size (node LPs of one problem come and go in equal sizes, util.cpp:25-42;
  // a second problem in the same process must not have to search through the first one's)
  std::unordered_map<size_t, std::vector<char *>> free_slabs;
  size_t free_bytes = 0, free_count = 0;
  // small slabs are carved out of chunks (one driver allocation per chunk instead of one per open
  // B&B node); a carved slab is only ever recycled, its chunk stays reserved until the process ends
  std::vector<std::pair<char *, size_t>> chunks;  // (base, bytes)
  size_t chunk_bytes = 0;
  // deferred clones (glp_copy_prob, util.cpp:34 / bs.cpp:116): queued here and carried out
  // by ONE clone_kernel launch in front of the next device operation
  std::vector<CloneJob> clone_jobs;
  struct Zombie { size_t bytes; char *slab; bool carved; };
  std::vector<Zombie> zombies;  // slabs released while a queued clone / cut row may still touch them
  CloneJob *d_jobs = nullptr; size_t jobs_cap = 0;
  unsigned clone_epoch = 1;  // bumped by every flush: marks which handles take part in queued clones
  // deferred cut rows (cut.cpp:23-43 -> mvb_lp_append_row): queued like the clones and formed by ONE
  // append_rows_kernel launch, in front of the queued clones (a child is cloned from the parent
  // that has just received its cut row)
  std::vector<struct AppendJob> append_jobs;
  std::vector<double> append_a;                  // the rows' coefficients, n doubles per job
  char *d_append = nullptr; size_t append_cap = 0;  // device copy: jobs, then coefficients
  // pinned staging of the batched row gather (mvb_lp_get_tab_rows_batch)
  char *h_gather = nullptr; size_t gather_cap = 0;
  cudaEvent_t ev[16] = {};
  cudaEvent_t kev[2] = {};
  char *flush = nullptr; size_t flush_bytes = 0;
  int last_G = 0, last_groups = 0;
  bool last_ps = false;
  double batch_kernel_ms = 0.0;   // device time of the solve kernels of the last mvb_solve_batch (all its launches)
  int batch_solve_launches = 0;
  int max_ctas_ps = 0;     // co-resident CTAs of simplex_ps_kernel with max_dyn
  size_t max_dyn = 0;      // dynamic shared memory simplex_kernel may ask for
  bool no_ps = false;      // MVOLPS_NO_PS=1: general loop only (debugging / A-B timing)
  bool no_res = false;       // MVOLPS_NO_RES=1: never use the resident (tableau in shared memory) kernel
  size_t max_dyn_res = 0;    // dynamic shared memory simplex_res_kernel may ask for
  size_t max_dyn_cres = 0;   // dynamic shared memory of one CTA of simplex_cres_kernel
  int use_cres = 0;          // MVOLPS_CRES=1 / option "cres": LPs that fit no single CTA but a cluster of 2/4/8 go to the
                             // cluster-resident kernel (measured slower than the lazy loop on full solves: opt-in)
  int last_kind = 0;         // kernel of the last solve launch: 0 general, 1 private-state (lazy), 2 resident
  bool no_handover = false;  // MVOLPS_NO_HANDOVER=1: phase 2 after a phase 1 stays in the general loop (A-B timing)
  int force_k = 0;         // MVOLPS_LZ_K=k: cap the pending-update depth (A-B timing)
  int ring_slots = LZ_RING_SLOTS;  // option "lz_ring" / MVOLPS_LZ_RING=0: no cp.async ring in the streaming flush (A-B timing)
  // a launch whose results have not been read back yet (mvb_solve_batch_begin): the D2H of its outbox is
  // enqueued; launch_finish() waits for it and fills the host mirrors.  At most one at a time.
  struct InFlight {
    bool active = false;
    std::vector<mvb_lp *> lps;
    std::vector<size_t> ooff;
  } inflight;
  int last_ring = 0, last_lz_k = 0;
  bool last_stream = false;  // the last lazy launch ran simplex_ps_stream_kernel
  int force_nt = 0;        // MVOLPS_PS_NT=t: CTA width for launches with more LPs than SMs (A-B timing)
  int last_nt = 0;
};
DeviceCtx g_ctx;

bool fail(const std::string &m) { g_err = m; return false; }
bool launch_finish();  // read back a deferred solve launch (below)
#define CK(call)                                                                     \
  do {                                                                               \
    cudaError_t e_ = (call);                                                         \
    if (e_ != cudaSuccess) {                                                         \
      g_err = std::string(#call) + ": " + cudaGetErrorString(e_);                    \
      return false;                                                                  \
    }                                                                                \
  } while (0)

bool ctx_init(int device) {
  if (g_ctx.ready && (device < 0 || device == g_ctx.device)) return true;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
    return fail("no CUDA device: the mvolps_b200 engine has no CPU fallback");
  if (device < 0) device = 0;
  CK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) return fail("mvolps_b200 is built for sm_100a (B200) only");
  g_ctx.device = device;
  g_ctx.sms = prop.multiProcessorCount;
  CK(cudaStreamCreateWithFlags(&g_ctx.stream, cudaStreamNonBlocking));
  cudaMemPool_t pool;
  CK(cudaDeviceGetDefaultMemPool(&pool, device));
  unsigned long long keep = ~0ULL;  // keep freed slabs cached for reuse
  CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
  int optin = 0;
  CK(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
  cudaFuncAttributes fa;
  CK(cudaFuncGetAttributes(&fa, simplex_ps_kernel));
  g_ctx.max_dyn = (size_t)optin > fa.sharedSizeBytes + 1024 ? (size_t)optin - fa.sharedSizeBytes - 1024 : 0;
  CK(cudaFuncSetAttribute(simplex_ps_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_ctx.max_dyn));
  {
    // (the streaming variant takes the same dynamic shared memory; its static part may differ by a few bytes)
    cudaFuncAttributes fs;
    CK(cudaFuncGetAttributes(&fs, simplex_ps_stream_kernel));
    if ((size_t)optin < fs.sharedSizeBytes + 1024 + g_ctx.max_dyn)
      g_ctx.max_dyn = (size_t)optin > fs.sharedSizeBytes + 1024 ? (size_t)optin - fs.sharedSizeBytes - 1024 : 0;
    CK(cudaFuncSetAttribute(simplex_ps_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_ctx.max_dyn));
  }
  const char *nops = getenv("MVOLPS_NO_PS");
  g_ctx.no_ps = nops && nops[0] == '1';
  const char *nores = getenv("MVOLPS_NO_RES");
  g_ctx.no_res = nores && nores[0] == '1';
  {
    cudaFuncAttributes fr;
    CK(cudaFuncGetAttributes(&fr, simplex_res_kernel));
    g_ctx.max_dyn_res = (size_t)optin > fr.sharedSizeBytes + 1024 ? (size_t)optin - fr.sharedSizeBytes - 1024 : 0;
    CK(cudaFuncSetAttribute(simplex_res_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_ctx.max_dyn_res));
  }
  {
    cudaFuncAttributes fc;
    CK(cudaFuncGetAttributes(&fc, simplex_cres_kernel));
    g_ctx.max_dyn_cres = (size_t)optin > fc.sharedSizeBytes + 1024 ? (size_t)optin - fc.sharedSizeBytes - 1024 : 0;
    CK(cudaFuncSetAttribute(simplex_cres_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_ctx.max_dyn_cres));
    const char *cr = getenv("MVOLPS_CRES");
    if (cr) g_ctx.use_cres = atoi(cr);
  }
  const char *noh = getenv("MVOLPS_NO_HANDOVER");
  g_ctx.no_handover = noh && noh[0] == '1';
  const char *fk = getenv("MVOLPS_LZ_K");
  g_ctx.force_k = fk ? atoi(fk) : 0;
  const char *frg = getenv("MVOLPS_LZ_RING");
  if (frg) g_ctx.ring_slots = atoi(frg) > 0 ? LZ_RING_SLOTS : 0;
  const char *fnt = getenv("MVOLPS_PS_NT");
  g_ctx.force_nt = fnt ? atoi(fnt) : 0;
  int per_sm = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, simplex_kernel, g_ctx.nt, 0));
  if (per_sm < 1) return fail("simplex_kernel does not fit an SM");
  g_ctx.max_ctas = per_sm * g_ctx.sms;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, simplex_ps_kernel, g_ctx.nt_ps, g_ctx.max_dyn));
  if (per_sm < 1) return fail("simplex_ps_kernel does not fit an SM");
  g_ctx.max_ctas_ps = per_sm * g_ctx.sms;
  g_ctx.ready = true;
  return true;
}

bool ensure_stage(size_t bytes) {
  if (bytes <= g_ctx.stage_cap) return true;
  if (g_ctx.stage) CK(cudaFreeAsync(g_ctx.stage, g_ctx.stream));
  g_ctx.stage = nullptr;
  g_ctx.stage_cap = 0;
  size_t cap = std::max<size_t>(bytes, 1 << 20);
  CK(cudaMallocAsync((void **)&g_ctx.stage, cap, g_ctx.stream));
  g_ctx.stage_cap = cap;
  return true;
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

// Host handle: device slab + host mirror of the small state.
struct mvb_lp {
  int m = 0, n = 0, ld = 0, mcap = 0;
  char *slab = nullptr;
  bool slab_carved = false;         // the slab is part of a chunk allocation (never freed on its own)
  size_t slab_bytes = 0, t_bytes = 0, small_off = 0, small_bytes = 0, out_off = 0, out_bytes = 0;
  LPDev dev{};
  int curT = 0, curS = 0;
  // host mirror of the last solve's small state, copy-on-write: a clone (one per open B&B node,
  // util.cpp:25-42) shares it until either side is solved again.  Bounds have no host mirror:
  // they live on the device and change through patches / the append kernel only.
  CowVec<int> head, nb, at;         // basis
  CowVec<double> x;                 // primal by id
  bool at_dirty = false;            // `at` was set on the host (load) and is not on the device yet
  std::vector<BoundPatch> patches;  // set_bounds calls not yet on the device
  mvb_result res{};
  int *pivlog = nullptr; int pivlog_cap = 0, pivlog_n = 0;
  DevResult dres{};                 // result block of the last launch this LP took part in
  bool in_flight = false;           // part of a launch whose results have not been read back (mvb_solve_batch_begin)
  unsigned clone_epoch = 0;         // == g_ctx.clone_epoch: source (1) / destination (2) of a queued clone
  int clone_role = 0;
  unsigned append_epoch = 0;        // == g_ctx.clone_epoch: a cut row of this LP is queued
  bool res_ok = false;              // the last solve request fits the resident kernel
  int cres_c = 0;                   // ... or a cluster of this many CTAs (cluster-resident kernel); 0: neither
};

namespace {

// carve the slab; every sub-array 256-byte aligned
bool lp_alloc(mvb_lp *lp, int m, int n, int mcap) {
  if (mcap < m) mcap = m;
  lp->m = m; lp->n = n; lp->mcap = mcap;
  lp->ld = (int)align_up((size_t)std::max(n, 1), MVB_LD_ALIGN);
  const size_t A = 256;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, A); return o; };
  lp->t_bytes = align_up((size_t)(mcap + 1) * lp->ld * sizeof(double), A);
  // T[0] | small state | T[1]: the current tableau buffer and the small state are
  // contiguous whichever buffer is current, so a clone is ONE device-to-device copy
  size_t oT0 = take(lp->t_bytes);
  lp->small_off = off;
  size_t ob0 = take((mcap + 1) * sizeof(double)), ob1 = take((mcap + 1) * sizeof(double));
  size_t olb = take((size_t)(n + mcap) * sizeof(double)), oub = take((size_t)(n + mcap) * sizeof(double));
  size_t ocost = take((size_t)std::max(n, 1) * sizeof(double));
  size_t oh0 = take((mcap + 1) * sizeof(int)), oh1 = take((mcap + 1) * sizeof(int));
  size_t on0 = take((size_t)std::max(n, 1) * sizeof(int)), on1 = take((size_t)std::max(n, 1) * sizeof(int));
  size_t oa0 = take((size_t)std::max(n, 1) * sizeof(int)), oa1 = take((size_t)std::max(n, 1) * sizeof(int));
  lp->small_bytes = off - lp->small_off;
  size_t oT1 = take(lp->t_bytes);
  size_t ovnb = take((size_t)std::max(n, 1) * sizeof(double));
  size_t nchunks = (mcap + MVB_P1_CHUNK - 1) / MVB_P1_CHUNK + 1;
  size_t op1 = take(nchunks * lp->ld * sizeof(double));
  lp->out_off = off;
  size_t ores = take(sizeof(DevResult));
  size_t ox = take((size_t)(n + mcap) * sizeof(double));
  size_t ooh = take((size_t)std::max(mcap, 1) * sizeof(int));
  size_t oon = take((size_t)std::max(n, 1) * sizeof(int)), ooa = take((size_t)std::max(n, 1) * sizeof(int));
  lp->out_bytes = off - lp->out_off;
  lp->slab_bytes = off;
  lp->slab = nullptr;
  lp->slab_carved = false;
  {
    auto it = g_ctx.free_slabs.find(lp->slab_bytes);
    if (it != g_ctx.free_slabs.end() && !it->second.empty()) {  // recycle: no driver call at all
      char *s = it->second.back();
      it->second.pop_back();
      lp->slab = (char *)((uintptr_t)s & ~(uintptr_t)1);
      lp->slab_carved = ((uintptr_t)s & 1) != 0;  // (slabs are 256-byte aligned: bit 0 carries the flag)
      g_ctx.free_bytes -= lp->slab_bytes;
      g_ctx.free_count--;
    }
  }
  if (!lp->slab && lp->slab_bytes <= ((size_t)4 << 20) && g_ctx.chunk_bytes < ((size_t)16 << 30)) {
    // a growing tree needs one new slab per branched node (util.cpp:25-42): carve them
    const size_t per = std::min<size_t>(256, std::max<size_t>(8, ((size_t)32 << 20) / lp->slab_bytes));
    char *base = nullptr;
    CK(cudaMallocAsync((void **)&base, per * lp->slab_bytes, g_ctx.stream));
    g_ctx.chunks.emplace_back(base, per * lp->slab_bytes);
    g_ctx.chunk_bytes += per * lp->slab_bytes;
    std::vector<char *> &fl = g_ctx.free_slabs[lp->slab_bytes];
    for (size_t k = per; k-- > 1;) fl.push_back((char *)((uintptr_t)(base + k * lp->slab_bytes) | 1));
    g_ctx.free_bytes += (per - 1) * lp->slab_bytes;
    g_ctx.free_count += per - 1;
    lp->slab = base;
    lp->slab_carved = true;
  }
  if (!lp->slab) CK(cudaMallocAsync((void **)&lp->slab, lp->slab_bytes, g_ctx.stream));
  char *s = lp->slab;
  LPDev &d = lp->dev;
  d.m = m; d.n = n; d.ld = lp->ld; d.mcap = mcap;
  d.T[0] = (double *)(s + oT0); d.T[1] = (double *)(s + oT1);
  d.beta[0] = (double *)(s + ob0); d.beta[1] = (double *)(s + ob1);
  d.lb = (double *)(s + olb); d.ub = (double *)(s + oub); d.cost = (double *)(s + ocost);
  d.head[0] = (int *)(s + oh0); d.head[1] = (int *)(s + oh1);
  d.nb[0] = (int *)(s + on0); d.nb[1] = (int *)(s + on1);
  d.at[0] = (int *)(s + oa0); d.at[1] = (int *)(s + oa1);
  d.vnb = (double *)(s + ovnb); d.p1part = (double *)(s + op1);
  d.res = (DevResult *)(s + ores); d.xval = (double *)(s + ox);
  d.ohead = (int *)(s + ooh); d.onb = (int *)(s + oon); d.oat = (int *)(s + ooa);
  d.pivlog = nullptr; d.pivlog_cap = 0;
  lp->curT = lp->curS = 0;
  return true;
}

void slab_recycle(size_t bytes, char *slab, bool carved) {
  // keep up to 1 GiB of slabs for reuse (carved slabs always: they cannot be handed back one by one)
  if (carved || (bytes <= ((size_t)64 << 20) && g_ctx.free_bytes + bytes <= ((size_t)1 << 30) && g_ctx.free_count < 8192)) {
    g_ctx.free_slabs[bytes].push_back(carved ? (char *)((uintptr_t)slab | 1) : slab);
    g_ctx.free_bytes += bytes;
    g_ctx.free_count++;
  } else {
    cudaFreeAsync(slab, g_ctx.stream);
  }
}

void lp_release(mvb_lp *lp) {
  if (lp->slab) {
    // a queued clone / cut row may still have to read (or write) this slab: park it until the flush
    if (!g_ctx.clone_jobs.empty() || !g_ctx.append_jobs.empty())
      g_ctx.zombies.push_back(DeviceCtx::Zombie{lp->slab_bytes, lp->slab, lp->slab_carved});
    else slab_recycle(lp->slab_bytes, lp->slab, lp->slab_carved);
  }
  if (lp->pivlog) cudaFreeAsync(lp->pivlog, g_ctx.stream);
  lp->slab = nullptr;
  lp->pivlog = nullptr;
  lp->pivlog_cap = 0;
}

__global__ void append_rows_kernel(const AppendJob *jobs, const double *arena);

// carry out the queued cut-row appends, then the queued clones: one H2D of each job list + one
// kernel each, stream-ordered in front of whatever the caller enqueues next
bool flush_clones() {
  cudaStream_t st = g_ctx.stream;
  if (!g_ctx.append_jobs.empty()) {
    std::vector<AppendJob> &aj = g_ctx.append_jobs;
    const size_t jb = align_up(aj.size() * sizeof(AppendJob), 256), ab = g_ctx.append_a.size() * sizeof(double);
    if (g_ctx.append_cap < jb + ab) {
      if (g_ctx.d_append) CK(cudaFreeAsync(g_ctx.d_append, st));
      g_ctx.append_cap = std::max((jb + ab) * 2, (size_t)1 << 20);
      CK(cudaMallocAsync((void **)&g_ctx.d_append, g_ctx.append_cap, st));
    }
    // pageable sources: the driver stages them before returning, so the vectors may be reused at once
    CK(cudaMemcpyAsync(g_ctx.d_append, aj.data(), aj.size() * sizeof(AppendJob), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(g_ctx.d_append + jb, g_ctx.append_a.data(), ab, cudaMemcpyHostToDevice, st));
    int maxld = 0;
    for (const AppendJob &j : aj) maxld = std::max(maxld, j.L.ld);
    append_rows_kernel<<<dim3((unsigned)((maxld + 127) / 128), (unsigned)aj.size()), 128, 0, st>>>(
        (const AppendJob *)g_ctx.d_append, (const double *)(g_ctx.d_append + jb));
    CK(cudaGetLastError());
    g_launches++;
    aj.clear();
    g_ctx.append_a.clear();
  }
  std::vector<CloneJob> &jobs = g_ctx.clone_jobs;
  if (jobs.empty()) {
    for (auto &z : g_ctx.zombies) slab_recycle(z.bytes, z.slab, z.carved);
    g_ctx.zombies.clear();
    g_ctx.clone_epoch++;
    return true;
  }
  if (g_ctx.jobs_cap < jobs.size()) {
    if (g_ctx.d_jobs) CK(cudaFreeAsync(g_ctx.d_jobs, st));
    g_ctx.jobs_cap = std::max(jobs.size() * 2, (size_t)1024);
    CK(cudaMallocAsync((void **)&g_ctx.d_jobs, g_ctx.jobs_cap * sizeof(CloneJob), st));
  }
  // pageable source: the driver stages it before returning, so `jobs` may be reused at once
  CK(cudaMemcpyAsync(g_ctx.d_jobs, jobs.data(), jobs.size() * sizeof(CloneJob), cudaMemcpyHostToDevice, st));
  unsigned long long max_bytes = 0;
  for (const CloneJob &j : jobs) max_bytes = std::max(max_bytes, j.bytes);
  const size_t per_block = (size_t)256 * 16 * 8;  // eight 16-byte units per thread
  size_t gx = (max_bytes + per_block - 1) / per_block;
  gx = std::max<size_t>(1, std::min<size_t>(gx, std::max<size_t>(1, (size_t)g_ctx.sms * 32 / jobs.size())));
  clone_kernel<<<dim3((unsigned)gx, (unsigned)jobs.size()), 256, 0, st>>>(g_ctx.d_jobs);
  CK(cudaGetLastError());
  g_launches++;
  jobs.clear();
  g_ctx.clone_epoch++;
  for (auto &z : g_ctx.zombies) slab_recycle(z.bytes, z.slab, z.carved);
  g_ctx.zombies.clear();
  return true;
}

int nb_flag_for(double l, double u, int keep) {
  const bool hasl = l > -HUGE_VAL, hasu = u < HUGE_VAL;
  if (hasl && hasu && l == u) return MVB_AT_FX;
  if (keep == MVB_AT_UB && hasu) return MVB_AT_UB;
  if (keep == MVB_AT_LB && hasl) return MVB_AT_LB;
  if (hasl) return MVB_AT_LB;
  if (hasu) return MVB_AT_UB;
  return MVB_AT_FR;
}

// push the non-basic flags a load set on the host
bool lp_push_small(mvb_lp *lp) {
  if (lp->at_dirty) {
    CK(cudaMemcpyAsync(lp->dev.at[lp->curS], lp->at.data(), (size_t)lp->n * sizeof(int), cudaMemcpyHostToDevice, g_ctx.stream));
    lp->at_dirty = false;
  }
  return true;
}

}  // namespace

#define CKI(call)                                                                    \
  do {                                                                               \
    cudaError_t e_ = (call);                                                         \
    if (e_ != cudaSuccess) {                                                         \
      g_err = std::string(#call) + ": " + cudaGetErrorString(e_);                    \
      return -1;                                                                     \
    }                                                                                \
  } while (0)

extern "C" {

const char *mvb_backend(void) { return "cuda-sm100a"; }
const char *mvb_last_error(void) { return g_err.c_str(); }
int mvb_device_init(int device) { return ctx_init(device) ? 0 : -1; }
unsigned long long mvb_kernel_launches(void) { return g_launches.load(); }

mvb_lp *mvb_lp_create(int m, int n, int row_cap) {
  if (m < 0 || n < 1) { g_err = "mvb_lp_create: bad dimensions"; return nullptr; }
  if (!ctx_init(-1)) return nullptr;
  mvb_lp *lp = new mvb_lp();
  if (!lp_alloc(lp, m, n, row_cap)) { delete lp; return nullptr; }
  return lp;
}

void mvb_lp_free(mvb_lp *lp) {
  if (!lp) return;
  if (lp->in_flight) launch_finish();  // (the read-back of a deferred launch writes into this object)
  lp_release(lp);
  delete lp;
}

void mvb_lp_free_many(mvb_lp *const *lps, int count) {
  if (count <= 0) return;
  for (int k = 0; k < count; k++)
    if (lps[k] && lps[k]->in_flight) { launch_finish(); break; }
  for (int k = 0; k < count; k++)
    if (lps[k]) { lp_release(lps[k]); delete lps[k]; }
}

void *mvb_host_staging(size_t bytes) {
  if (!ctx_init(-1)) return nullptr;
  if (g_ctx.h_stage_cap < bytes) {
    if (g_ctx.h_stage) cudaFreeHost(g_ctx.h_stage);
    g_ctx.h_stage = nullptr;
    g_ctx.h_stage_cap = 0;
    if (cudaMallocHost(&g_ctx.h_stage, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    g_ctx.h_stage_cap = bytes;
  }
  return g_ctx.h_stage;
}

int mvb_lp_dims(const mvb_lp *lp, int *m, int *n) {
  if (m) *m = lp->m;
  if (n) *n = lp->n;
  return 0;
}

int mvb_lp_load(mvb_lp *lp, const double *A, const double *cost, const double *lb,
                const double *ub) {
  if (!flush_clones()) return -1;
  const int m = lp->m, n = lp->n, ld = lp->ld;
  cudaStream_t st = g_ctx.stream;
  LPDev &d = lp->dev;
  lp->curT = lp->curS = 0;
  CKI(cudaMemsetAsync(d.T[0], 0, lp->t_bytes, st));
  CKI(cudaMemcpyAsync(d.T[0], cost, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, st));
  if (m > 0)
    CKI(cudaMemcpy2DAsync(d.T[0] + ld, (size_t)ld * sizeof(double), A, (size_t)n * sizeof(double),
                          (size_t)n * sizeof(double), m, cudaMemcpyHostToDevice, st));
  CKI(cudaMemcpyAsync(d.cost, cost, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, st));
  CKI(cudaMemcpyAsync(d.lb, lb, (size_t)(n + m) * sizeof(double), cudaMemcpyHostToDevice, st));
  CKI(cudaMemcpyAsync(d.ub, ub, (size_t)(n + m) * sizeof(double), cudaMemcpyHostToDevice, st));
  {
    int *hd = lp->head.fresh(m), *nb = lp->nb.fresh(n), *at = lp->at.fresh(n);
    for (int i = 0; i < m; i++) hd[i] = n + i;
    for (int j = 0; j < n; j++) { nb[j] = j; at[j] = nb_flag_for(lb[j], ub[j], MVB_AT_LB); }
  }
  if (m > 0)
    CKI(cudaMemcpyAsync(d.head[0] + 1, lp->head.data(), (size_t)m * sizeof(int), cudaMemcpyHostToDevice, st));
  CKI(cudaMemcpyAsync(d.nb[0], lp->nb.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice, st));
  lp->patches.clear();
  lp->at_dirty = true;
  if (!lp_push_small(lp)) return -1;
  {
    double *xv = lp->x.fresh(n + m);
    std::fill(xv, xv + n + m, 0.0);
  }
  lp->res = mvb_result{};
  lp->res.status = MVB_UNSOLVED;
  CKI(cudaStreamSynchronize(st));
  return 0;
}

int mvb_lp_assign(mvb_lp *dst, const mvb_lp *src) {
  if ((dst->in_flight || src->in_flight) && !launch_finish()) return -1;
  if (dst->n != src->n || dst->mcap != src->mcap) { g_err = "mvb_lp_assign: shape mismatch"; return -1; }
  dst->m = src->m;
  dst->dev.m = src->m;
  dst->curT = src->curT;
  dst->curS = src->curS;
  // the current tableau buffer and the small state are adjacent (lp_alloc): one copy
  if (dst->t_bytes != src->t_bytes || dst->small_bytes != src->small_bytes) { g_err = "mvb_lp_assign: layout mismatch"; return -1; }
  const size_t from = src->curT == 0 ? 0 : src->small_off;
  // queued, not issued: all clones of a B&B round go out as one kernel (flush_clones).
  // A handle may be read by several queued clones, but a chain (clone of a queued
  // destination, or overwriting a queued source/destination) is flushed first.
  const unsigned ep = g_ctx.clone_epoch;
  const bool src_is_dst = src->clone_epoch == ep && (src->clone_role & 2);
  const bool dst_busy = (dst->clone_epoch == ep && dst->clone_role != 0) || dst->append_epoch == ep;
  if (src_is_dst || dst_busy || g_ctx.clone_jobs.size() >= 32768) {
    if (!flush_clones()) return -1;
  }
  mvb_lp *msrc = const_cast<mvb_lp *>(src);
  if (msrc->clone_epoch != g_ctx.clone_epoch) { msrc->clone_epoch = g_ctx.clone_epoch; msrc->clone_role = 0; }
  msrc->clone_role |= 1;
  dst->clone_epoch = g_ctx.clone_epoch;
  dst->clone_role = 2;
  g_ctx.clone_jobs.push_back(CloneJob{src->slab + from, dst->slab + from,
                                      (unsigned long long)(src->t_bytes + src->small_bytes)});
  dst->head = src->head; dst->nb = src->nb; dst->at = src->at; dst->x = src->x;  // shared, copy-on-write
  dst->at_dirty = src->at_dirty;
  dst->patches = src->patches;
  dst->res = src->res;
  dst->pivlog_n = 0;
  return 0;
}

mvb_lp *mvb_lp_clone(const mvb_lp *src) {
  mvb_lp *lp = mvb_lp_create(src->m, src->n, src->mcap);
  if (!lp) return nullptr;
  if (mvb_lp_assign(lp, src)) { mvb_lp_free(lp); return nullptr; }
  return lp;
}

int mvb_lp_set_bounds(mvb_lp *lp, int k, double lb, double ub) {
  if (lp->in_flight && !launch_finish()) return -1;
  if (k < 0 || k >= lp->n + lp->m) { g_err = "mvb_lp_set_bounds: variable id"; return -1; }
  // the change travels as a patch the next solve kernel applies (bounds have no host mirror)
  int pos = -1, flag = 0;
  const int *nb = lp->nb.data();
  for (int j = 0; j < lp->n; j++)
    if (nb[j] == k) {
      pos = j;
      flag = nb_flag_for(lb, ub, lp->at[j]);
      if (flag != lp->at[j]) lp->at.w(j) = flag;
      break;
    }
  for (auto &p : lp->patches)
    if (p.var == k) { p.lb = lb; p.ub = ub; p.pos = pos; p.flag = flag; return 0; }
  lp->patches.push_back(BoundPatch{k, pos, flag, 0, lb, ub});
  return 0;
}

}  // extern "C"

namespace {

// tableau row of a new constraint a.x: a_N + sum_i a_{head[i]} T[i,:], terms in
// row order, one fma each (specification: oracle mvb_lp_append_row); blockIdx.y = job
__global__ void append_rows_kernel(const AppendJob *jobs, const double *arena) {
  const AppendJob &J = jobs[blockIdx.y];
  const LPDev &L = J.L;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = L.m, n = L.n, ld = L.ld;
  if (p >= ld) return;
  const double *a = arena + J.a_off;
  const double *T = L.T[J.cT];
  double acc = 0.0;
  if (p < n) {
    const int k = L.nb[J.cS][p];
    acc = k < n ? a[k] : 0.0;
    for (int i = 1; i <= m; i++) {
      const int h = L.head[J.cS][i];
      if (h < n && a[h] != 0.0) acc = fma(a[h], T[(size_t)i * ld + p], acc);
    }
  }
  L.T[J.cT][(size_t)(m + 1) * ld + p] = acc;
  if (p == 0) {
    L.head[J.cS][m + 1] = n + m;
    L.beta[J.cS][m + 1] = 0.0;
    L.lb[n + m] = J.lb;
    L.ub[n + m] = J.ub;
  }
}

// K7: gather tableau rows into a contiguous staging buffer; blockIdx.x = row job
__global__ void gather_rows_kernel(const GatherJob *jobs) {
  const GatherJob j = jobs[blockIdx.x];
  for (int c = threadIdx.x; c < j.n; c += blockDim.x) j.dst[c] = j.src[c];
}

bool lp_grow(mvb_lp *lp, int mcap) {
  mvb_lp nl;
  if (!lp_alloc(&nl, lp->m, lp->n, mcap)) return false;
  cudaStream_t st = g_ctx.stream;
  const int m = lp->m, n = lp->n;
  const LPDev &s = lp->dev;
  LPDev &d = nl.dev;
  CK(cudaMemcpyAsync(d.T[0], s.T[lp->curT], (size_t)(m + 1) * lp->ld * sizeof(double), cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(d.beta[0], s.beta[lp->curS], (size_t)(m + 1) * sizeof(double), cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(d.head[0], s.head[lp->curS], (size_t)(m + 1) * sizeof(int), cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(d.nb[0], s.nb[lp->curS], (size_t)n * sizeof(int), cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(d.at[0], s.at[lp->curS], (size_t)n * sizeof(int), cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(d.cost, s.cost, (size_t)n * sizeof(double), cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(d.lb, s.lb, (size_t)(n + m) * sizeof(double), cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(d.ub, s.ub, (size_t)(n + m) * sizeof(double), cudaMemcpyDeviceToDevice, st));
  lp_release(lp);
  lp->slab = nl.slab; nl.slab = nullptr;
  lp->slab_carved = nl.slab_carved;
  lp->slab_bytes = nl.slab_bytes; lp->t_bytes = nl.t_bytes;
  lp->small_off = nl.small_off; lp->small_bytes = nl.small_bytes;
  lp->out_off = nl.out_off; lp->out_bytes = nl.out_bytes;
  lp->mcap = nl.mcap;
  lp->dev = nl.dev;
  lp->curT = lp->curS = 0;
  return true;
}

}  // namespace

extern "C" int mvb_lp_append_row(mvb_lp *lp, const double *a, double lb, double ub) {
  if (lp->in_flight && !launch_finish()) return -1;
  // queued: the rows of a B&B round are formed by one kernel (flush_clones).  An LP that takes
  // part in a queued clone, already has a queued row, or must grow is brought up to date first.
  const bool in_clone = lp->clone_epoch == g_ctx.clone_epoch && lp->clone_role != 0;
  const bool grows = lp->m + 1 > lp->mcap;
  if (in_clone || grows || lp->append_epoch == g_ctx.clone_epoch || g_ctx.append_jobs.size() >= 16384) {
    if (!flush_clones()) return -1;
  }
  if (grows && !lp_grow(lp, lp->mcap + 16 + lp->mcap / 4)) return -1;
  const int n = lp->n, m = lp->m;
  lp->dev.m = m;
  AppendJob j;
  j.L = lp->dev;
  j.cT = lp->curT;
  j.cS = lp->curS;
  j.a_off = g_ctx.append_a.size();
  j.lb = lb;
  j.ub = ub;
  g_ctx.append_a.insert(g_ctx.append_a.end(), a, a + n);
  g_ctx.append_jobs.push_back(j);
  lp->append_epoch = g_ctx.clone_epoch;
  lp->m = m + 1;
  lp->dev.m = m + 1;
  lp->head.push_back(n + m);
  lp->x.push_back(0.0);
  return 0;
}

namespace {

// one cooperative launch over the LPs `idx` (indices into lps), results into the outboxes
// read back a deferred launch: wait for its D2H, fill the host mirrors of its LPs
bool launch_finish() {
  if (!g_ctx.inflight.active) return true;
  g_ctx.inflight.active = false;
  CK(cudaStreamSynchronize(g_ctx.stream));
  {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, g_ctx.kev[0], g_ctx.kev[1]) == cudaSuccess) g_ctx.batch_kernel_ms += ms;
    g_ctx.batch_solve_launches++;
  }
  // results of this launch into the host mirrors (an LP that reports MVB_PS_FALLBACK /
  // MVB_PS_HANDOVER is launched again and overwritten)
  const std::vector<mvb_lp *> &lps = g_ctx.inflight.lps;
  par_for(lps.size(), [&](size_t k0, size_t k1) {  // (every LP on its own: host worker threads)
    for (size_t k = k0; k < k1; k++) {
      mvb_lp *lp = lps[k];
      lp->in_flight = false;
      const char *ob = g_ctx.h_out + g_ctx.inflight.ooff[k];
      const char *base = lp->slab + lp->out_off;
      const DevResult *dr = (const DevResult *)(ob + ((const char *)lp->dev.res - base));
      lp->dres = *dr;
      if (dr->r.status == MVB_PS_FALLBACK) continue;
      const double *x = (const double *)(ob + ((const char *)lp->dev.xval - base));
      const int *oh = (const int *)(ob + ((const char *)lp->dev.ohead - base));
      const int *on = (const int *)(ob + ((const char *)lp->dev.onb - base));
      const int *oa = (const int *)(ob + ((const char *)lp->dev.oat - base));
      lp->x.assign(x, x + lp->n + lp->m);
      lp->head.assign(oh, oh + lp->m);
      lp->nb.assign(on, on + lp->n);
      lp->at.assign(oa, oa + lp->n);
    }
  });
  g_ctx.inflight.lps.clear();
  return true;
}

// defer: return as soon as the launch and the D2H of its results are enqueued (launch_finish reads them)
bool launch_solve(mvb_lp *const *lps, const std::vector<int> &idx, const mvb_opts *opts, int kind, int csize = 0,
                  bool defer = false) {
  const bool psk = kind == 1, res = kind == 2 || kind == 3, cres = kind == 3;
  const int count = (int)idx.size();
  if (!count) return true;
  if (!launch_finish()) return false;  // (its staging buffers are about to be reused)
  cudaStream_t st = g_ctx.stream;
  std::vector<LPDev> h(count);
  size_t max_bytes = 0, dyn = 0, tot_bytes = 0;
  for (int k = 0; k < count; k++) {
    mvb_lp *lp = lps[idx[k]];
    h[k] = lp->dev;
    max_bytes = std::max(max_bytes, (size_t)(lp->m + 1) * lp->ld * sizeof(double));
    tot_bytes += 2 * (size_t)(lp->m + 1) * lp->ld * sizeof(double);
  }
  int max_ctas = psk ? g_ctx.max_ctas_ps : g_ctx.max_ctas;
  int nt = psk ? g_ctx.nt_ps : g_ctx.nt;
  // the launch's tableaux do not stay in L2: flush passes stream from HBM (deeper pending chains,
  // reads staged through the cp.async ring when one CTA owns an LP)
  const bool streams = tot_bytes > ((size_t)96 << 20);
  int ring = psk && streams ? g_ctx.ring_slots : 0;
  if (psk && count > g_ctx.sms && !(opts && opts->ctas_per_lp > 0)) {
    // more small LPs than SMs (a B&B round): narrower CTAs, several resident per SM, so more
    // LPs are in flight; the vectors of one LP do not fill 512 threads anyway
    int span = 0;
    for (int k = 0; k < count; k++) span = std::max(span, std::max(lps[idx[k]]->ld, lps[idx[k]]->m + 1));
    const int want = g_ctx.force_nt > 0 ? g_ctx.force_nt : span <= 256 ? 128 : span <= 512 ? 256 : nt;
    if (want < nt) {
      // (more LPs than SMs: one CTA per LP, so a streaming launch runs the streaming kernel)
      const int kcap = g_ctx.force_k > 0 ? g_ctx.force_k : streams ? LZ_KMAX_STREAM : LZ_KMAX;
      const void *kern = streams ? (const void *)simplex_ps_stream_kernel : (const void *)simplex_ps_kernel;
      size_t dyn_max = 0;
      for (int k = 0; k < count; k++)
        dyn_max = std::max(dyn_max, lz_smem_bytes(lps[idx[k]]->m, lps[idx[k]]->ld,
                                                  lz_pick_k(lps[idx[k]]->m, lps[idx[k]]->ld, g_ctx.max_dyn, std::min(kcap, LZ_KCAP))));
      int per_sm = 0;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, want, dyn_max));
      if (ring) {  // the ring must not cost a resident CTA
        int per_sm_ring = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_ring, kern, want, dyn_max + 16 + lz_ring_bytes(ring, want)));
        if (per_sm_ring < per_sm) ring = 0;
      }
      if (per_sm > 1) {
        nt = want;
        max_ctas = per_sm * g_ctx.sms;
      }
    }
  }
  // CTAs per LP: by tableau size, bounded by what is co-resident
  const size_t per_cta = psk ? (96 << 10) : (128 << 10);
  int G = (opts && opts->ctas_per_lp > 0) ? opts->ctas_per_lp
                                          : (int)std::min<size_t>((max_bytes + per_cta - 1) / per_cta, (size_t)max_ctas);
  if (!(opts && opts->ctas_per_lp > 0) && count > 1) G = std::max(1, std::min(G, max_ctas / count));
  G = std::max(1, std::min(G, max_ctas));
  int ngroups = std::max(1, std::min(count, max_ctas / G));
  const bool stream_kernel = psk && streams && G == 1;
  if (psk) {
    // pending-update depth: 5 is the measured sweet spot while the tableaux stay in L2
    // (flush passes are cheap, deeper chains cost decisions); a wide batch of one-CTA LPs
    // streams its tableaux from HBM, where every pending pivot more divides the flush traffic
    // (depth 12 in the streaming kernel: 13.8 against 12.5 M LP-pivots/s at depth 8 with the ring)
    const int kmax = stream_kernel ? LZ_KMAX_STREAM : LZ_KMAX;
    const int cap = std::min(kmax, g_ctx.force_k > 0 ? g_ctx.force_k : stream_kernel ? LZ_KMAX_STREAM : 5);
    if (!stream_kernel) ring = 0;
    for (int pass = 0; pass < 2; pass++) {  // (second pass: without the ring, when it does not fit every LP)
      dyn = 0;
      bool fits = true;
      for (int k = 0; k < count; k++) {
        mvb_lp *lp = lps[idx[k]];
        const int K = std::min(lz_pick_k(lp->m, lp->ld, g_ctx.max_dyn, kmax), cap);
        h[k].lzK = K;
        h[k].lzRing = ring;
        g_ctx.last_lz_k = K;
        const size_t need = ring ? lz_ring_off(lp->m, lp->ld, K) + lz_ring_bytes(ring, nt) : lz_smem_bytes(lp->m, lp->ld, K);
        if (need > g_ctx.max_dyn) fits = false;
        dyn = std::max(dyn, need);
      }
      if (fits || !ring) break;
      ring = 0;
    }
    g_ctx.last_ring = ring;
    g_ctx.last_stream = stream_kernel;
  }
  int res_grid = 0;
  if (res) {
    // one CTA per LP, as many CTAs as are co-resident with the largest tableau of the launch
    // (grid-stride over the LPs: a CTA takes the next LP when it has stored the last one)
    dyn = 0;
    if (!cres) {
      for (int k = 0; k < count; k++) dyn = std::max(dyn, res_smem_bytes(lps[idx[k]]->m, lps[idx[k]]->ld));
      int per_sm = 0;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, simplex_res_kernel, RES_NT, dyn));
      if (per_sm < 1) return fail("simplex_res_kernel does not fit an SM");
      res_grid = std::min(count, per_sm * g_ctx.sms);
      G = 1;
      nt = RES_NT;
    } else {
      // clusters of `csize` CTAs, one LP per cluster at a time; as many clusters as are co-resident
      for (int k = 0; k < count; k++) dyn = std::max(dyn, cres_smem_bytes(lps[idx[k]]->m, lps[idx[k]]->ld, csize));
      cudaLaunchConfig_t qc{};
      qc.gridDim = dim3((unsigned)(csize * count));
      qc.blockDim = dim3(CRES_NT);
      qc.dynamicSmemBytes = dyn;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = (unsigned)csize; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
      qc.attrs = qa;
      qc.numAttrs = 1;
      int nclusters = 0;
      CK(cudaOccupancyMaxActiveClusters(&nclusters, simplex_cres_kernel, &qc));
      if (nclusters < 1) return fail("simplex_cres_kernel: no cluster of this size fits the device");
      res_grid = std::min(count, nclusters) * csize;
      G = csize;
      nt = CRES_NT;
    }
    ngroups = res_grid / G;
  }
  if (g_ctx.ctrs_cap < ngroups) {
    if (g_ctx.ctrs) CK(cudaFreeAsync(g_ctx.ctrs, st));
    CK(cudaMallocAsync((void **)&g_ctx.ctrs, (size_t)ngroups * sizeof(unsigned), st));
    g_ctx.ctrs_cap = ngroups;
  }
  if (g_ctx.dlps_cap < count) {
    if (g_ctx.dlps) CK(cudaFreeAsync(g_ctx.dlps, st));
    CK(cudaMallocAsync((void **)&g_ctx.dlps, (size_t)count * sizeof(LPDev), st));
    g_ctx.dlps_cap = count;
  }
  // outbox arena: every LP's result block at its offset, one D2H for the launch
  std::vector<size_t> ooff(count);
  size_t otot = 0, ptot = 0;
  for (int k = 0; k < count; k++) {
    ooff[k] = otot;
    otot += lps[idx[k]]->out_bytes;
    ptot += lps[idx[k]]->patches.size();
  }
  if (g_ctx.out_cap < otot) {
    if (g_ctx.d_out) CK(cudaFreeAsync(g_ctx.d_out, st));
    if (g_ctx.h_out) CK(cudaFreeHost(g_ctx.h_out));
    g_ctx.out_cap = std::max(otot * 2, (size_t)1 << 20);
    CK(cudaMallocAsync((void **)&g_ctx.d_out, g_ctx.out_cap, st));
    CK(cudaMallocHost((void **)&g_ctx.h_out, g_ctx.out_cap));
  }
  if (g_ctx.patch_cap < ptot) {
    if (g_ctx.d_patch) CK(cudaFreeAsync(g_ctx.d_patch, st));
    if (g_ctx.h_patch) CK(cudaFreeHost(g_ctx.h_patch));
    g_ctx.patch_cap = std::max(ptot * 2, (size_t)4096);
    CK(cudaMallocAsync((void **)&g_ctx.d_patch, g_ctx.patch_cap * sizeof(BoundPatch), st));
    CK(cudaMallocHost((void **)&g_ctx.h_patch, g_ctx.patch_cap * sizeof(BoundPatch)));
  }
  size_t pat = 0;
  for (int k = 0; k < count; k++) {
    mvb_lp *lp = lps[idx[k]];
    const ptrdiff_t rel = (g_ctx.d_out + ooff[k]) - (lp->slab + lp->out_off);
    h[k].res = (DevResult *)((char *)lp->dev.res + rel);
    h[k].xval = (double *)((char *)lp->dev.xval + rel);
    h[k].ohead = (int *)((char *)lp->dev.ohead + rel);
    h[k].onb = (int *)((char *)lp->dev.onb + rel);
    h[k].oat = (int *)((char *)lp->dev.oat + rel);
    h[k].npatch = (int)lp->patches.size();
    h[k].patch = g_ctx.d_patch + pat;
    for (const BoundPatch &bp : lp->patches) g_ctx.h_patch[pat++] = bp;
    lp->patches.clear();
  }
  if (ptot)
    CK(cudaMemcpyAsync(g_ctx.d_patch, g_ctx.h_patch, ptot * sizeof(BoundPatch), cudaMemcpyHostToDevice, st));
  CK(cudaMemsetAsync(g_ctx.ctrs, 0, (size_t)ngroups * sizeof(unsigned), st));
  CK(cudaMemcpyAsync(g_ctx.dlps, h.data(), (size_t)count * sizeof(LPDev), cudaMemcpyHostToDevice, st));
  const LPDev *dl = g_ctx.dlps;
  unsigned *ctrs = g_ctx.ctrs;
  int tcl = -1, cnt = count;
  void *args[] = {(void *)&dl, (void *)&cnt, (void *)&G, (void *)&ctrs, (void *)&tcl};
  if (!g_ctx.kev[0]) { CK(cudaEventCreate(&g_ctx.kev[0])); CK(cudaEventCreate(&g_ctx.kev[1])); }
  CK(cudaEventRecord(g_ctx.kev[0], st));
  if (cres) {
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3((unsigned)res_grid);
    lc.blockDim = dim3(CRES_NT);
    lc.dynamicSmemBytes = dyn;
    lc.stream = st;
    cudaLaunchAttribute la[1];
    la[0].id = cudaLaunchAttributeClusterDimension;
    la[0].val.clusterDim.x = (unsigned)csize; la[0].val.clusterDim.y = 1; la[0].val.clusterDim.z = 1;
    lc.attrs = la;
    lc.numAttrs = 1;
    CK(cudaLaunchKernelEx(&lc, simplex_cres_kernel, dl, cnt, csize));
  } else if (res) {
    simplex_res_kernel<<<res_grid, RES_NT, dyn, st>>>(dl, cnt);
    CK(cudaGetLastError());
  } else {
    CK(cudaLaunchCooperativeKernel(stream_kernel ? (const void *)simplex_ps_stream_kernel
                                   : psk         ? (const void *)simplex_ps_kernel
                                                 : (const void *)simplex_kernel,
                                   dim3(ngroups * G), dim3(nt), args, dyn, st));
  }
  CK(cudaEventRecord(g_ctx.kev[1], st));
  g_launches++;
  g_ctx.last_G = G;
  g_ctx.last_nt = nt;
  g_ctx.last_groups = ngroups;
  g_ctx.last_ps = psk;
  g_ctx.last_kind = kind;
  CK(cudaMemcpyAsync(g_ctx.h_out, g_ctx.d_out, otot, cudaMemcpyDeviceToHost, st));
  g_ctx.inflight.active = true;
  g_ctx.inflight.lps.resize(count);
  for (int k = 0; k < count; k++) {
    g_ctx.inflight.lps[k] = lps[idx[k]];
    lps[idx[k]]->in_flight = true;
  }
  g_ctx.inflight.ooff = ooff;
  return defer ? true : launch_finish();
}

}  // namespace

namespace {

// One mvb_solve_batch call: which kernel every LP goes to first (plan), the first launches, then
// what they hand on (phase 1 to the general kernel, phase 2 back to the fast ones) and the results.
struct BatchPlan {
  std::vector<mvb_lp *> lps;
  mvb_opts opts{};
  bool have_opts = false;
  std::vector<int> ps_idx, gen_idx, res_idx, cres_idx[3];
  bool deferred = false;  // the one first launch is in flight (mvb_solve_batch_begin)
  bool active = false;
  bool done = false;                // brought home early (another solve came in between): results held here
  std::vector<mvb_result> results;
};
BatchPlan g_plan;  // the batch between mvb_solve_batch_begin and mvb_solve_batch_end

int batch_plan(BatchPlan &P) {
  cudaStream_t st = g_ctx.stream;
  mvb_lp *const *lps = P.lps.data();
  const int count = (int)P.lps.size();
  const mvb_opts *opts = P.have_opts ? &P.opts : nullptr;
  std::vector<int> &ps_idx = P.ps_idx, &gen_idx = P.gen_idx, &res_idx = P.res_idx;
  std::vector<int> *cres_idx = P.cres_idx;
  for (int k = 0; k < count; k++) {
    mvb_lp *lp = lps[k];
    if (!lp_push_small(lp)) return -1;
    int mi = (opts && opts->max_iters > 0) ? opts->max_iters : 50 * (lp->m + lp->n) + 10000;
    if (opts && opts->log_pivots) {
      int want = std::min(mi, 1 << 22);
      if (lp->pivlog_cap < want) {
        if (lp->pivlog) CKI(cudaFreeAsync(lp->pivlog, st));
        CKI(cudaMallocAsync((void **)&lp->pivlog, (size_t)want * 2 * sizeof(int), st));
        lp->pivlog_cap = want;
      }
    }
    lp->dev.m = lp->m;
    lp->dev.curT = lp->curT;
    lp->dev.curS = lp->curS;
    lp->dev.max_iters = mi;
    lp->dev.pivlog = (opts && opts->log_pivots) ? lp->pivlog : nullptr;
    lp->dev.pivlog_cap = lp->dev.pivlog ? lp->pivlog_cap : 0;
    const bool fits = !g_ctx.no_ps && lp->m >= 1 && lz_pick_k(lp->m, lp->ld, g_ctx.max_dyn) >= 1 &&
                      lp->m + 1 <= PS_CREG * g_ctx.nt_ps;
    lp->dev.ps_ok = fits && !g_ctx.no_handover;
    lp->dev.carry = 0;
    lp->dev.srcT = nullptr;
    // resident kernel: the whole tableau fits one CTA's shared memory (a B&B node LP)
    const size_t rb = g_ctx.no_res || lp->m < 1 ? 0 : res_smem_bytes(lp->m, lp->ld);
    lp->res_ok = rb > 0 && rb <= g_ctx.max_dyn_res && !(opts && opts->ctas_per_lp > 0);  // an explicit group size asks for the lazy loop
    if (lp->res_ok) lp->dev.ps_ok = !g_ctx.no_handover;
    // cluster-resident kernel: too large for one CTA, but a cluster of 2 / 4 / 8 holds it
    lp->cres_c = 0;
    if (!lp->res_ok && g_ctx.use_cres && !g_ctx.no_res && lp->m >= 1 && !(opts && opts->ctas_per_lp > 0))
      lp->cres_c = cres_pick_c(lp->m, lp->ld, g_ctx.max_dyn_cres);
    if (lp->cres_c) lp->dev.ps_ok = !g_ctx.no_handover;
    if (lp->cres_c) cres_idx[lp->cres_c == 2 ? 0 : lp->cres_c == 4 ? 1 : 2].push_back(k);
    else (lp->res_ok ? res_idx : fits ? ps_idx : gen_idx).push_back(k);
  }
  return 0;
}

// resident and private-state kernels first; what needs phase 1 (or fits neither) goes to the general one
int batch_first(BatchPlan &P, bool defer) {
  mvb_lp *const *lps = P.lps.data();
  const mvb_opts *opts = P.have_opts ? &P.opts : nullptr;
  // a single first launch can stay in flight while the host goes on (its results are read by batch_rest)
  const int kinds = !P.res_idx.empty() + !P.ps_idx.empty() + !P.cres_idx[0].empty() + !P.cres_idx[1].empty() +
                    !P.cres_idx[2].empty();
  const bool d = defer && kinds == 1 && P.gen_idx.empty() && (!P.res_idx.empty() || !P.ps_idx.empty());
  if (!launch_solve(lps, P.res_idx, opts, 2, 0, d)) return -1;
  for (int c = 0; c < 3; c++)
    if (!launch_solve(lps, P.cres_idx[c], opts, 3, 2 << c)) return -1;
  if (!launch_solve(lps, P.ps_idx, opts, 1, 0, d)) return -1;
  P.deferred = d;
  return 0;
}

int batch_rest(BatchPlan &P, mvb_result *out) {
  if (!launch_finish()) return -1;
  P.deferred = false;
  mvb_lp *const *lps = P.lps.data();
  const int count = (int)P.lps.size();
  const mvb_opts *opts = P.have_opts ? &P.opts : nullptr;
  std::vector<int> &ps_idx = P.ps_idx, &gen_idx = P.gen_idx, &res_idx = P.res_idx;
  std::vector<int> *cres_idx = P.cres_idx;
  auto dev_result = [](mvb_lp *lp) { return (const DevResult *)&lp->dres; };
  for (int k : res_idx)
    if (dev_result(lps[k])->r.status == MVB_PS_FALLBACK) gen_idx.push_back(k);
  for (int c = 0; c < 3; c++)
    for (int k : cres_idx[c])
      if (dev_result(lps[k])->r.status == MVB_PS_FALLBACK) gen_idx.push_back(k);
  for (int k : ps_idx)
    if (dev_result(lps[k])->r.status == MVB_PS_FALLBACK) gen_idx.push_back(k);
  std::sort(gen_idx.begin(), gen_idx.end());
  if (!launch_solve(lps, gen_idx, opts, 0)) return -1;
  // ... which hands an LP back as soon as its phase 1 has found a primal feasible basis: phase 2
  // (the bulk of the iterations) then runs in the private-state kernel, carrying the tallies on
  std::vector<int> hand_idx, hand_res, hand_cres[3];
  for (int k : gen_idx) {
    mvb_lp *lp = lps[k];
    const DevResult *dr = dev_result(lp);
    if (dr->r.status != MVB_PS_HANDOVER) continue;
    lp->curT = lp->dev.curT = dr->curT;
    lp->curS = lp->dev.curS = dr->curS;
    lp->dev.carry = 1;
    lp->dev.c_iters = dr->r.iters; lp->dev.c_pivots = dr->r.pivots; lp->dev.c_flips = dr->r.flips;
    lp->dev.c_it1 = dr->r.iters_primal1; lp->dev.c_it2 = dr->r.iters_primal2; lp->dev.c_itd = dr->r.iters_dual;
    lp->dev.c_bland = dr->r.bland_iters; lp->dev.c_hash = dr->r.pivot_hash;
    if (lp->cres_c) hand_cres[lp->cres_c == 2 ? 0 : lp->cres_c == 4 ? 1 : 2].push_back(k);
    else (lp->res_ok ? hand_res : hand_idx).push_back(k);
  }
  if (!launch_solve(lps, hand_res, opts, 2)) return -1;
  for (int c = 0; c < 3; c++) {
    if (!launch_solve(lps, hand_cres[c], opts, 3, 2 << c)) return -1;
    for (int k : hand_cres[c]) lps[k]->dev.carry = 0;
  }
  if (!launch_solve(lps, hand_idx, opts, 1)) return -1;
  for (int k : hand_res) lps[k]->dev.carry = 0;
  for (int k : hand_idx) lps[k]->dev.carry = 0;
  for (int k = 0; k < count; k++) {
    mvb_lp *lp = lps[k];
    const DevResult *dr = &lp->dres;
    lp->res = dr->r;
    lp->curT = dr->curT;
    lp->curS = dr->curS;
    lp->pivlog_n = lp->dev.pivlog ? std::min(dr->r.iters, lp->pivlog_cap) : 0;
    out[k] = dr->r;
  }
  return 0;
}

int batch_setup(BatchPlan &P, mvb_lp *const *lps, int count, const mvb_opts *opts) {
  if (!ctx_init(-1)) return -1;
  if (!launch_finish()) return -1;
  if (!flush_clones()) return -1;
  g_ctx.batch_kernel_ms = 0.0;
  g_ctx.batch_solve_launches = 0;
  P = BatchPlan{};
  P.lps.assign(lps, lps + count);
  if (opts) { P.opts = *opts; P.have_opts = true; }
  return batch_plan(P);
}

}  // namespace

extern "C" int mvb_solve_batch(mvb_lp *const *lps, int count, const mvb_opts *opts,
                               mvb_result *out) {
  if (count <= 0) return 0;
  if (g_plan.active && !g_plan.done && !g_plan.lps.empty()) {
    // a batch is in flight: it comes home now (mvb_solve_batch_end will hand its results out)
    g_plan.results.resize(g_plan.lps.size());
    if (batch_rest(g_plan, g_plan.results.data())) return -1;
    g_plan.done = true;
  }
  BatchPlan P;
  if (batch_setup(P, lps, count, opts)) return -1;
  if (batch_first(P, false)) return -1;
  return batch_rest(P, out);
}

// The same solve in two halves: begin enqueues the (first) launch and returns while the device works,
// end waits, runs whatever the first launch hands on, and delivers the results.  One batch at a
// time; its LPs must not be touched in between (other LPs may be created, cloned, changed, freed).
extern "C" int mvb_solve_batch_begin(mvb_lp *const *lps, int count, const mvb_opts *opts) {
  if (g_plan.active) { g_err = "mvb_solve_batch_begin: a batch is already in flight"; return -1; }
  if (count <= 0) { g_plan = BatchPlan{}; g_plan.active = true; return 0; }
  if (batch_setup(g_plan, lps, count, opts)) return -1;
  if (batch_first(g_plan, true)) return -1;
  g_plan.active = true;
  return 0;
}

extern "C" int mvb_solve_batch_end(mvb_result *out) {
  if (!g_plan.active) { g_err = "mvb_solve_batch_end: no batch in flight"; return -1; }
  // (may be called from another host thread than begin -- the B&B driver reads a launch back on a
  // thread of its own -- and the current device is per thread)
  if (g_ctx.device >= 0) {
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess || cur != g_ctx.device) CKI(cudaSetDevice(g_ctx.device));
  }
  g_plan.active = false;
  if (g_plan.lps.empty()) return 0;
  if (g_plan.done) {
    std::copy(g_plan.results.begin(), g_plan.results.end(), out);
    return 0;
  }
  return batch_rest(g_plan, out);
}

extern "C" {

int mvb_lp_get_primal(mvb_lp *lp, double *x) {
  if ((int)lp->x.size() != lp->n + lp->m) { g_err = "mvb_lp_get_primal: no solution"; return -1; }
  std::memcpy(x, lp->x.data(), lp->x.size() * sizeof(double));
  return 0;
}

int mvb_lp_solution_view(mvb_lp *lp, const double **x, const int **head, const int **nb, const int **at) {
  if ((int)lp->x.size() != lp->n + lp->m) { g_err = "mvb_lp_solution_view: no solution"; return -1; }
  *x = lp->x.data();
  *head = lp->head.data();
  *nb = lp->nb.data();
  *at = lp->at.data();
  return 0;
}

int mvb_lp_get_basis(mvb_lp *lp, int *head, int *nb, int *at) {
  std::memcpy(head, lp->head.data(), (size_t)lp->m * sizeof(int));
  std::memcpy(nb, lp->nb.data(), (size_t)lp->n * sizeof(int));
  std::memcpy(at, lp->at.data(), (size_t)lp->n * sizeof(int));
  return 0;
}

int mvb_lp_get_tab_rows_batch(mvb_lp *const *lps, int count, const int *rows, const int *nrows, double *out) {
  if (!flush_clones()) return -1;
  size_t njobs = 0, odoubles = 0;
  for (int k = 0; k < count; k++) {
    if (nrows[k] < 0) { g_err = "mvb_lp_get_tab_rows: negative row count"; return -1; }
    njobs += (size_t)nrows[k];
    odoubles += (size_t)nrows[k] * lps[k]->n;
  }
  if (!njobs) return 0;
  cudaStream_t st = g_ctx.stream;
  const size_t jb = align_up(njobs * sizeof(GatherJob), 256), ob = odoubles * sizeof(double);
  if (!ensure_stage(jb + ob)) return -1;
  if (g_ctx.gather_cap < jb + ob) {
    if (g_ctx.h_gather) CKI(cudaFreeHost(g_ctx.h_gather));
    g_ctx.gather_cap = std::max((jb + ob) * 2, (size_t)1 << 20);
    CKI(cudaMallocHost((void **)&g_ctx.h_gather, g_ctx.gather_cap));
  }
  GatherJob *hj = (GatherJob *)g_ctx.h_gather;
  double *dout = (double *)(g_ctx.stage + jb);
  size_t t = 0, o = 0;
  for (int k = 0; k < count; k++) {
    mvb_lp *lp = lps[k];
    for (int r = 0; r < nrows[k]; r++, t++) {
      const int row = rows[t];
      if (row < 0 || row > lp->m) { g_err = "mvb_lp_get_tab_rows: row out of range"; return -1; }
      hj[t] = GatherJob{lp->dev.T[lp->curT] + (size_t)row * lp->ld, dout + o, lp->n, 0};
      o += lp->n;
    }
  }
  CKI(cudaMemcpyAsync(g_ctx.stage, hj, njobs * sizeof(GatherJob), cudaMemcpyHostToDevice, st));
  gather_rows_kernel<<<(unsigned)njobs, 128, 0, st>>>((const GatherJob *)g_ctx.stage);
  CKI(cudaGetLastError());
  g_launches++;
  CKI(cudaMemcpyAsync(g_ctx.h_gather + jb, dout, ob, cudaMemcpyDeviceToHost, st));
  CKI(cudaStreamSynchronize(st));
  std::memcpy(out, g_ctx.h_gather + jb, ob);
  return 0;
}

int mvb_lp_get_tab_rows(mvb_lp *lp, const int *rows, int nrows, double *out) {
  return mvb_lp_get_tab_rows_batch(&lp, 1, rows, &nrows, out);
}

}  // extern "C"

// ---- mvb_lp_get_tab_rows_noalloc: K7 with its staging in a static device array --------------------
// For callers that may not allocate (CUDA graph capture around them, tight serving loops, harnesses
// that forbid cudaMalloc).  gather_rows_kernel stages its job list and the gathered rows in a
// device buffer the engine grows on demand (ensure_stage -> cudaMallocAsync) and in pinned host
// memory (cudaMallocHost); this variant keeps both in ONE file-scope device array of a fixed size:
// no cudaMalloc*, cudaFree* or allocating library routine on its path.  Every GPU that loads the
// library holds the array until the process ends.  All calls share it and it keeps the previous
// call's contents (nothing in it has to start at zero: the row numbers and every gathered double
// of a call are written before they are read), so calls that overlap on the GPU race.
static __device__ __align__(256) unsigned char g_noalloc_scratch[MVB_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int MVB_NOALLOC_MAX_DEVICES = 64;
// device addresses of g_noalloc_scratch, one per device ordinal.  In host code the array's NAME is
// a host-side stand-in, not a device address: the kernel gets what cudaGetSymbolAddress returns
// with that device current.
unsigned char *g_noalloc_ptr[MVB_NOALLOC_MAX_DEVICES] = {};

// row job blockIdx.x: row rows[blockIdx.x] of the tableau T (row stride ld) -> dst[blockIdx.x * n ...]
__global__ void gather_rows_noalloc_kernel(const double *T, int ld, int n, const int *rows, double *dst) {
  const double *src = T + (size_t)rows[blockIdx.x] * ld;
  double *d = dst + (size_t)blockIdx.x * n;
  for (int c = threadIdx.x; c < n; c += blockDim.x) d[c] = src[c];
}

}  // namespace

extern "C" int mvb_lp_get_tab_rows_noalloc(mvb_lp *lp, const int *rows, int nrows, double *out) {
  if (!lp || !g_ctx.ready) { g_err = "mvb_lp_get_tab_rows_noalloc: engine not initialised"; return -1; }
  if (nrows < 0) { g_err = "mvb_lp_get_tab_rows_noalloc: negative row count"; return -1; }
  if (nrows == 0) return 0;
  if (!rows || !out) { g_err = "mvb_lp_get_tab_rows_noalloc: null rows / out"; return -1; }
  if (!g_ctx.append_jobs.empty() || !g_ctx.clone_jobs.empty()) {
    // carrying out queued cut rows / clones may grow the engine's job buffers (an allocation)
    g_err = "mvb_lp_get_tab_rows_noalloc: queued clones or cut rows are pending; run mvb_solve_batch or "
            "mvb_lp_get_tab_rows first";
    return -1;
  }
  const size_t rb = align_up((size_t)nrows * sizeof(int), 256), ob = (size_t)nrows * lp->n * sizeof(double);
  if (rb + ob > (size_t)MVB_NOALLOC_SCRATCH_BYTES) {
    g_err = "mvb_lp_get_tab_rows_noalloc: " + std::to_string(nrows) + " rows of " + std::to_string(lp->n) +
            " columns need " + std::to_string(rb + ob) + " bytes of staging, the static scratch holds " +
            std::to_string((size_t)MVB_NOALLOC_SCRATCH_BYTES) + " (MVB_NOALLOC_SCRATCH_BYTES); use mvb_lp_get_tab_rows";
    return -1;
  }
  for (int r = 0; r < nrows; r++)
    if (rows[r] < 0 || rows[r] > lp->m) { g_err = "mvb_lp_get_tab_rows_noalloc: row out of range"; return -1; }
  const int dev = g_ctx.device;
  if (dev < 0 || dev >= MVB_NOALLOC_MAX_DEVICES) { g_err = "mvb_lp_get_tab_rows_noalloc: device ordinal out of range"; return -1; }
  if (!g_noalloc_ptr[dev]) {  // set-up, once per device, with that device current
    int cur = -1;
    CKI(cudaGetDevice(&cur));
    if (cur != dev) CKI(cudaSetDevice(dev));
    unsigned char *ptr = nullptr;
    CKI(cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch));
    if (!ptr) { g_err = "mvb_lp_get_tab_rows_noalloc: cudaGetSymbolAddress returned null"; return -1; }
    g_noalloc_ptr[dev] = ptr;
  }
  unsigned char *scratch = g_noalloc_ptr[dev];
  cudaStream_t st = g_ctx.stream;
  int *drows = (int *)scratch;
  double *dout = (double *)(scratch + rb);
  // pageable source and destination: the driver stages them itself, nothing is allocated here
  CKI(cudaMemcpyAsync(drows, rows, (size_t)nrows * sizeof(int), cudaMemcpyHostToDevice, st));
  gather_rows_noalloc_kernel<<<(unsigned)nrows, 128, 0, st>>>(lp->dev.T[lp->curT], lp->ld, lp->n, drows, dout);
  CKI(cudaGetLastError());
  g_launches++;
  CKI(cudaMemcpyAsync(out, dout, ob, cudaMemcpyDeviceToHost, st));
  CKI(cudaStreamSynchronize(st));
  return 0;
}

extern "C" {

int mvb_lp_get_tableau(mvb_lp *lp, double *T, double *beta) {
  if (!flush_clones()) return -1;
  cudaStream_t st = g_ctx.stream;
  CKI(cudaMemcpy2DAsync(T, (size_t)lp->n * sizeof(double), lp->dev.T[lp->curT],
                        (size_t)lp->ld * sizeof(double), (size_t)lp->n * sizeof(double), lp->m + 1,
                        cudaMemcpyDeviceToHost, st));
  if (beta)
    CKI(cudaMemcpyAsync(beta, lp->dev.beta[lp->curS], (size_t)(lp->m + 1) * sizeof(double),
                        cudaMemcpyDeviceToHost, st));
  CKI(cudaStreamSynchronize(st));
  if (beta) beta[0] = 0.0;
  return 0;
}

int mvb_lp_get_pivot_log(mvb_lp *lp, int *rowcol, int cap) {
  if (!flush_clones()) return -1;
  int c = std::min(lp->pivlog_n, cap);
  if (c > 0) {
    CKI(cudaMemcpyAsync(rowcol, lp->pivlog, (size_t)c * 2 * sizeof(int), cudaMemcpyDeviceToHost, g_ctx.stream));
    CKI(cudaStreamSynchronize(g_ctx.stream));
  }
  return lp->pivlog_n;
}

}  // extern "C"

extern "C" {

int mvb_timer_record(int slot) {
  if (!ctx_init(-1)) return -1;
  slot &= 15;
  if (!g_ctx.ev[slot]) CKI(cudaEventCreate(&g_ctx.ev[slot]));
  CKI(cudaEventRecord(g_ctx.ev[slot], g_ctx.stream));
  return 0;
}

double mvb_timer_elapsed_ms(int a, int b) {
  float ms = -1.0f;
  if (cudaEventSynchronize(g_ctx.ev[b & 15]) != cudaSuccess) return -1.0;
  if (cudaEventElapsedTime(&ms, g_ctx.ev[a & 15], g_ctx.ev[b & 15]) != cudaSuccess) return -1.0;
  return ms;
}

double mvb_last_kernel_ms(void) {
  if (!g_ctx.kev[1]) return -1.0;
  return g_ctx.batch_kernel_ms;  // every launch of the last mvb_solve_batch has been waited for
}

int mvb_flush_l2(void) {
  if (!ctx_init(-1)) return -1;
  if (!g_ctx.flush) {
    g_ctx.flush_bytes = (size_t)256 << 20;  // > 126 MB L2
    CKI(cudaMalloc((void **)&g_ctx.flush, g_ctx.flush_bytes));
  }
  static int v = 0;
  CKI(cudaMemsetAsync(g_ctx.flush, ++v & 0xff, g_ctx.flush_bytes, g_ctx.stream));
  return 0;
}

int mvb_last_kernel_mode(void) { return g_ctx.last_kind; }

int mvb_engine_option(const char *name, int value) {
  if (!ctx_init(-1) || !name) return -1;
  int old = -1;
  auto flag = [&](bool &f) { old = f; if (value >= 0) f = value != 0; };
  auto num = [&](int &f) { old = f; if (value >= 0) f = value; };
  if (!strcmp(name, "no_res")) flag(g_ctx.no_res);
  else if (!strcmp(name, "no_ps")) flag(g_ctx.no_ps);
  else if (!strcmp(name, "no_handover")) flag(g_ctx.no_handover);
  else if (!strcmp(name, "cres")) num(g_ctx.use_cres);
  else if (!strcmp(name, "lz_k")) num(g_ctx.force_k);
  else if (!strcmp(name, "ps_nt")) num(g_ctx.force_nt);
  else if (!strcmp(name, "lz_ring")) { old = g_ctx.ring_slots; if (value >= 0) g_ctx.ring_slots = value ? LZ_RING_SLOTS : 0; }
  else if (!strcmp(name, "last_ring")) old = g_ctx.last_ring;  // read-only: ring slots of the last lazy launch
  else if (!strcmp(name, "last_lz_k")) old = g_ctx.last_lz_k;  // read-only: pending depth of the last lazy launch
  else if (!strcmp(name, "last_stream")) old = g_ctx.last_stream;  // read-only: it ran the streaming kernel
  else { g_err = std::string("mvb_engine_option: unknown option ") + name; return -1; }
  return old;
}

int mvb_last_launch(int *c, int *g, int *t) {
  if (c) *c = g_ctx.last_G;
  if (g) *g = g_ctx.last_groups;
  if (t) *t = g_ctx.last_nt;
  return 0;
}

}  // extern "C"

#ifdef MVB_PROF
// profiling builds only (make PROF=1): per-CTA phase clocks of the private-state loop
extern "C" int mvb_debug_prof(long long *out, int reset) {
  if (out) cudaMemcpyFromSymbol(out, g_prof, sizeof(long long) * 256 * 16);
  if (reset) {
    static long long z[256 * 16];
    cudaMemcpyToSymbol(g_prof, z, sizeof(z));
  }
  return 0;
}
#endif
#include "engine_shard.cuh"
